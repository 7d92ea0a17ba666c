# GPU-box helper: A/B of library variants / environment knobs on the bench workload.
#   tools/gpu_ab.sh "name|ENV=.. ENV=..|lib path or -" ...
cd $GRAFT_REPO_ROOT
for spec in "$@"; do
  name=$(echo "$spec" | cut -d'|' -f1); envs=$(echo "$spec" | cut -d'|' -f2); lib=$(echo "$spec" | cut -d'|' -f3)
  if [ "$lib" != "-" ] && [ -n "$lib" ]; then envs="$envs BCU_LIB=$lib"; fi
  env $envs timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-extra > gpurun_out/ab_$name.json 2> gpurun_out/ab_$name.err || { echo "$name FAILED"; tail -3 gpurun_out/ab_$name.err; continue; }
  python - <<P
import json
d=json.load(open("gpurun_out/ab_$name.json"))
print("%-22s strikes/s %.3fM exact %.3fM | ms/step %.3f exact %.3f | kernel_ms %.3f | e2e %.3fM | best %s" % ("$name", d["value"]/1e6, d["value_exact"]/1e6, d["ms_per_step"], d["ms_per_step_exact"], d["roofline"]["kernel_ms"], d["e2e"]["value"]/1e6, d["best_strike"]))
P
done

set -x
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_size.py -x -q > gpurun_out/r02_gputests_c.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02_gputests_c.log
tail -15 gpurun_out/r02_gputests_c.log
for cfg in "1 8" "1 4" "1 16" "0 8"; do set -- $cfg
  BCU_LAYOUT=$1 BCU_STEP_LANES=$2 timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-extra > gpurun_out/r02_c_layout$1_lanes$2.json 2> gpurun_out/r02_c_layout$1_lanes$2.err || tail -5 gpurun_out/r02_c_layout$1_lanes$2.err
  python - <<P
import json
d=json.load(open("gpurun_out/r02_c_layout$1_lanes$2.json"))
print("layout $1 lanes $2: strikes/s %.3fM exact %.3fM ms/step %.3f exact %.3f kernel_ms %.3f e2e %.3fM best %s" % (d["value"]/1e6, d["value_exact"]/1e6, d["ms_per_step"], d["ms_per_step_exact"], d["roofline"]["kernel_ms"], d["e2e"]["value"]/1e6, d["best_strike"]))
P
done
BCU_ROUND_TIMING=1 BCU_STREAMS=1 BCU_PRUNE=1 timeout 120 python tools/level_once.py 1 2> gpurun_out/r02_c_round_timing.txt | tail -1
grep "bcu timing" gpurun_out/r02_c_round_timing.txt | tail -36

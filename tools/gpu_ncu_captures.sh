cd $GRAFT_REPO_ROOT
export BCU_NO_GRAPH=1 BCU_STREAMS=1
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum --clock-control none --cache-control none --csv --log-file gpurun_out/r02_launch_list.csv python tools/level_once.py 1 > gpurun_out/ncu_ll.log 2>&1
for k in items quartic apply; do
  timeout 600 ncu --set full --import-source on --clock-control none -k regex:${k}_kernel --launch-skip 166 --launch-count 1 -o gpurun_out/r02_${k}_r11 python tools/level_once.py 1 > gpurun_out/ncu_$k.log 2>&1
  python tools/ncu_summary.py gpurun_out/r02_${k}_r11.ncu-rep > gpurun_out/r02_${k}_kernel_ncu_full.txt 2>&1
done
timeout 600 ncu --set full --import-source on --clock-control none -k regex:plan_count_kernel --launch-skip 1 --launch-count 1 -o gpurun_out/r02_plan_count python tools/plan_once.py > gpurun_out/ncu_plan.log 2>&1
python tools/ncu_summary.py gpurun_out/r02_plan_count.ncu-rep > gpurun_out/r02_plan_count_kernel_ncu_full.txt 2>&1
unset BCU_NO_GRAPH BCU_STREAMS
timeout 600 python -m pytest tests/test_gpu_prune_regressions.py -x -q > gpurun_out/r02_gputests_h.log 2>&1; tail -3 gpurun_out/r02_gputests_h.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/r02_bench_h.json 2> gpurun_out/r02_bench_h.err; echo "bench exit $?"
python - <<'P'
import json
d=json.load(open("gpurun_out/r02_bench_h.json"))
print({k:d[k] for k in ("value","value_exact","ms_per_step")}, d["e2e"]["value"])
for k,v in d["extra"].items(): print(k, json.dumps(v)[:700])
P
ls -la gpurun_out | head -50

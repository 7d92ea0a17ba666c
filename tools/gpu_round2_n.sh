cd $GRAFT_REPO_ROOT
BCU_NO_GRAPH=1 BCU_STREAMS=1 timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum --clock-control none --cache-control none --csv --log-file gpurun_out/r02_launch_list.csv python tools/level_once.py 1 > gpurun_out/ncu_ll.log 2>&1
BCU_NO_GRAPH=1 BCU_STREAMS=1 timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum --clock-control none --cache-control none --csv --log-file gpurun_out/r02_launch_list_prune0.csv python tools/level_once.py 0 > gpurun_out/ncu_ll0.log 2>&1
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "two_engines" 2>&1 | tail -2
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/r02_bench_n.json 2> gpurun_out/r02_bench_n.err; echo "bench exit $?"
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r02_bench_ref_n.json 2> /dev/null
python - <<'P'
import json
d=json.load(open("gpurun_out/r02_bench_n.json"))
print({k:d[k] for k in ("value","value_exact","ms_per_step")}, d["e2e"]["value"], d["roofline"]["traffic"], d["roofline"]["frac_issue_slots"])
print(json.dumps(d["extra"]["config4"])[:900])
P

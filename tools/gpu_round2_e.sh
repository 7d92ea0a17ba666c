cd $GRAFT_REPO_ROOT
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r02_gputests_e.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02_gputests_e.log
tail -6 gpurun_out/r02_gputests_e.log
BCU_LAYOUT=1 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_size.py tests/test_gpu_search_levels.py -x -q > gpurun_out/r02_gputests_e_fused.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02_gputests_e_fused.log
tail -3 gpurun_out/r02_gputests_e_fused.log
timeout 900 python bench.py > gpurun_out/r02_bench_e.json 2> gpurun_out/r02_bench_e.err; echo "bench exit $?"
python - <<'P'
import json
d=json.load(open("gpurun_out/r02_bench_e.json"))
print({k:d[k] for k in ("value","value_exact","ms_per_step")}, d["e2e"], d["parity"])
for k,v in d["extra"].items(): print(k, {a:b for a,b in v.items() if a!="workload"})
P

cd $GRAFT_REPO_ROOT
timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/r02_gputests_final.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02_gputests_final.log
tail -4 gpurun_out/r02_gputests_final.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/r02_bench_final.json 2> gpurun_out/r02_bench_final.err; echo "bench exit $?"
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r02_bench_final_ref.json 2>/dev/null
python - <<'P'
import json
d=json.load(open("gpurun_out/r02_bench_final.json"))
print({k:d[k] for k in ("value","value_exact","ms_per_step","gpu_launches")}, d["e2e"]["value"], d["parity"]["mismatches"], d["roofline"]["traffic"], d["roofline"]["frac"], d["roofline"]["frac_issue_slots"])
print(json.load(open("gpurun_out/r02_bench_final_ref.json"))["value"])
P

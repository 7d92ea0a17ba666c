#!/bin/bash
# GPU-box helper: parity spot check + short bench, prints one summary line
timeout 300 python tools/debug_mismatch.py 16 20000 3 2>&1 | head -1
timeout 300 python bench.py --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/qb.json 2> gpurun_out/qb.err || tail -5 gpurun_out/qb.err
python - <<'P'
import json
d=json.load(open("gpurun_out/qb.json"))
print("strikes/s %.3fM  ms/step %.2f  kernel_ms %.2f  e2e %.3fM  best %s" % (d["value"]/1e6, d["ms_per_step"], d["roofline"]["kernel_ms"], d["e2e"]["value"]/1e6, d["best_strike"]))
P

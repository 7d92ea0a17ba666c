cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_planner.py -x -q > gpurun_out/r02_gputests_i.log 2>&1; tail -3 gpurun_out/r02_gputests_i.log
python - <<'P'
import sys, time, numpy as np
sys.path.insert(0, '.')
from billiard_ai_b200 import Engine, workloads as W
tbl = W.default_table()
eng = Engine.from_table(tbl, W.SNOOKER_SCORES, cue_type=W.TYPE_CODE["WHITE"])
eng.set_planner(tbl, velocity_multiply=True)
for nst, nb in ((1, 16), (64, 16), (1024, 16), (1024, 22)):
    pos, types = W.random_tables(nst, nb, seed0=5000)
    eng.set_states(pos, types, 1 << W.TYPE_CODE["RED"])
    n = eng.plan()
    ms = []
    for _ in range(3):
        eng.timer_begin(); eng.plan(); ms.append(eng.timer_end())
    print(nst, nb, "strikes", n, "plan ms", np.median(ms))
P
BCU_NO_GRAPH=1 BCU_STREAMS=1 timeout 600 ncu --set full --import-source on --clock-control none -k regex:plan_count_kernel --launch-skip 1 --launch-count 1 -o gpurun_out/r02_plan_count_v2 python tools/plan_once.py > gpurun_out/ncu_plan.log 2>&1
python tools/ncu_summary.py gpurun_out/r02_plan_count_v2.ncu-rep > gpurun_out/r02_plan_count_v2_kernel_ncu_full.txt 2>&1
head -48 gpurun_out/r02_plan_count_v2_kernel_ncu_full.txt | tail -16

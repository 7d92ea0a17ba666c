# GPU-box helper (diagnosis): per-kernel times of a small and of the benched level with every quartic kernel launched twice --
# the second launch solves the same jobs with the solver's code still in the instruction caches (cold against warm).
cd $GRAFT_REPO_ROOT
export BCU_ROUND_TIMING=1 BCU_QUARTIC_TWICE=1 BCU_STREAMS=1
for n in 8 128 2048; do
  echo "== level of $((n * 32)) strikes"
  timeout 120 python tools/level_small.py $n 2>&1 | grep "bcu timing" | tail -34 | grep "round  [0-9] \|round 1[05] \|round 2[05] \|round 30\|twice\|finish"
done

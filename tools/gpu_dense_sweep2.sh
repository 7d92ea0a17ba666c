# GPU-box helper: items-dense and quartic-dense separately on small parts (32 768 and 65 536 strikes per part).
cd $GRAFT_REPO_ROOT
BIG=1000000000
for nt in 16 32; do
  for c in "ss $BIG $BIG" "sd $BIG 0" "ds 0 $BIG" "dd 0 0"; do
    set -- $c
    BCU_DENSE_MIN_ITEMS=$2 BCU_DENSE_MIN_QUARTIC=$3 timeout 300 python tools/big_launch.py ${1}_$nt $nt
  done
done
bash tools/gpu_ab.sh "level_ss|BCU_DENSE_MIN=$BIG|-" "level_sd|BCU_DENSE_MIN_ITEMS=$BIG BCU_DENSE_MIN_QUARTIC=0|-" "level_ds|BCU_DENSE_MIN_ITEMS=0 BCU_DENSE_MIN_QUARTIC=$BIG|-" "level_ss2|BCU_DENSE_MIN=$BIG|-" "level_sd2|BCU_DENSE_MIN_ITEMS=$BIG BCU_DENSE_MIN_QUARTIC=0|-"

# GPU-box helper: the build with quartic_kernel_wide: parity tests, launch sizes, the bench level, and the source-level ncu capture of
# the wide kernel on a large round (the same round as profiles/r02_big_round_instruction_fetch.txt had for the 128-thread kernel).
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_size.py tests/test_gpu_search_levels.py -m gpu -x -q 2>&1 | tail -3
for nt in 16 32 64 256 1000 2000; do timeout 300 python tools/big_launch.py default_$nt $nt; done
bash tools/gpu_ab.sh "level_default||-"
bash tools/gpu_ncu_quartic_big.sh 2>&1 | head -24

# GPU-box helper: the dense + per-SM-class quartic kernel forced on every launch size through the parity tests, then the
# default build on large launches.
cd $GRAFT_REPO_ROOT
BCU_SM_CLASS_MIN=0 BCU_DENSE_MIN=0 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_size.py tests/test_gpu_search_levels.py -m gpu -x -q 2>&1 | tail -4
for nt in 256 1000 2000; do timeout 300 python tools/big_launch.py default_$nt $nt; done

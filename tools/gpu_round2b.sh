# GPU-box helper (this round's second half): the no-allocation entry point's test alone in a fresh process, the dense
# crossover sweep, then the parity tests that cover small and large launches (dense instantiations included).
cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -15
bash tools/gpu_dense_sweep.sh 2>&1 | grep -v "^$" | tee gpurun_out/dense_sweep.txt
BCU_DENSE_MIN=0 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_size.py -m gpu -x -q 2>&1 | tail -5

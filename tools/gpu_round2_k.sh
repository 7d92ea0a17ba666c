cd $GRAFT_REPO_ROOT
BCU_NO_GRAPH=1 timeout 900 python -m pytest tests/test_gpu_parity.py -x -q > gpurun_out/k_nograph.log 2>&1; echo "nograph exit $?"; tail -2 gpurun_out/k_nograph.log
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "not filmed and not thesis" > gpurun_out/k_subset.log 2>&1; echo "subset exit $?"; tail -2 gpurun_out/k_subset.log
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q > gpurun_out/k_graph.log 2>&1; echo "graph exit $?"; grep -m3 "engine.py\|test_gpu_parity.py\", line" gpurun_out/k_graph.log; tail -2 gpurun_out/k_graph.log | cut -c1-200
nvidia-smi --query-gpu=name,memory.used --format=csv

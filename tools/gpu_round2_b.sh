set -x
cd $GRAFT_REPO_ROOT
timeout 1500 python -m pytest tests/test_gpu_search_dropin.py tests/test_gpu_search_levels.py -x -q -s > gpurun_out/r02_gputests_b.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02_gputests_b.log
tail -30 gpurun_out/r02_gputests_b.log
timeout 300 python tools/config4_search.py > gpurun_out/r02_config4_1gpu.json 2> gpurun_out/r02_config4_1gpu.err; cat gpurun_out/r02_config4_1gpu.json

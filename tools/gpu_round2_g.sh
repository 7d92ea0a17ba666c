cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_planner.py tests/test_gpu_search_dropin.py tests/test_gpu_prune_regressions.py -x -q -s > gpurun_out/r02_gputests_g.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02_gputests_g.log
tail -25 gpurun_out/r02_gputests_g.log

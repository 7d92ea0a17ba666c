cd $GRAFT_REPO_ROOT
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/r02_gputests_j.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02_gputests_j.log
tail -6 gpurun_out/r02_gputests_j.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3

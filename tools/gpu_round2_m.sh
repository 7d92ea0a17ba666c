cd $GRAFT_REPO_ROOT
for i in 1 2 3 4 5 6; do
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -k "not filmed and not thesis" > gpurun_out/m_$i.log 2>&1; echo "run $i exit $? $(tail -1 gpurun_out/m_$i.log | cut -c1-60)"
done

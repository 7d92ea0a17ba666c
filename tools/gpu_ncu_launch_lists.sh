# GPU-box helper: DRAM bytes / warp instructions / time of one level per kernel (ncu launch list), reach bound on and off, with the
# kernel instantiations the 2-part bench level runs (one part of 65 536 strikes under ncu would otherwise take the dense ones).
cd $GRAFT_REPO_ROOT
export BCU_NO_GRAPH=1 BCU_STREAMS=1 BCU_DENSE_MIN=1000000000
for p in 1 0; do
  timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum --clock-control none --cache-control none --csv --log-file gpurun_out/launch_list_prune$p.csv python tools/level_once.py $p > gpurun_out/ncu_ll$p.log 2>&1; tail -1 gpurun_out/ncu_ll$p.log
done

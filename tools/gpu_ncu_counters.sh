# GPU-box helper: the counter passes behind bench.py's roofline block, for the build as it is: per-opcode FP instruction counts and
# DRAM bytes / warp instructions / time of one level, with the reach bound on (1) and off (0).  The JSONs are made from the
# CSVs with tools/ncu_fp_ops.py and tools/ncu_level_json.py (see their headers).
cd $GRAFT_REPO_ROOT
export BCU_NO_GRAPH=1 BCU_STREAMS=1
FP=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_fadd_pred_on.sum,smsp__sass_thread_inst_executed_op_fmul_pred_on.sum,smsp__sass_thread_inst_executed_op_ffma_pred_on.sum
for p in 1 0; do
  timeout 900 ncu --metrics $FP --clock-control none --csv --log-file gpurun_out/fp_ops_prune$p.csv python tools/level_once.py $p > gpurun_out/ncu_fp$p.log 2>&1; tail -1 gpurun_out/ncu_fp$p.log
  timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum --clock-control none --cache-control none --csv --log-file gpurun_out/launch_list_prune$p.csv python tools/level_once.py $p > gpurun_out/ncu_ll$p.log 2>&1; tail -1 gpurun_out/ncu_ll$p.log
done
ls -la gpurun_out/*.csv gpurun_out/level_counters*

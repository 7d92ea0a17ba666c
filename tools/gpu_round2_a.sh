set -x
cd $GRAFT_REPO_ROOT
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02_gputests.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02_gputests.log
timeout 600 python bench.py > gpurun_out/r02_bench_a.json 2> gpurun_out/r02_bench_a.err; echo "bench exit $?"
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r02_bench_ref_a.json 2> gpurun_out/r02_bench_ref_a.err
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_fadd_pred_on.sum,smsp__sass_thread_inst_executed_op_fmul_pred_on.sum,smsp__sass_thread_inst_executed_op_ffma_pred_on.sum
timeout 900 ncu --metrics $M --clock-control none --csv --log-file gpurun_out/fp_ops_p1.csv python tools/level_once.py 1 > gpurun_out/level_once_p1.log 2>&1
timeout 900 ncu --metrics $M --clock-control none --csv --log-file gpurun_out/fp_ops_p0.csv python tools/level_once.py 0 > gpurun_out/level_once_p0.log 2>&1
tail -3 gpurun_out/r02_gputests.log; cat gpurun_out/r02_bench_a.json | cut -c1-1500

cd $GRAFT_REPO_ROOT
export BCU_NO_GRAPH=1 BCU_STREAMS=1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:step_kernel --launch-skip 138 --launch-count 1 -o gpurun_out/r02_step_peak python tools/level_once.py 1 > gpurun_out/ncu_step.log 2>&1
tail -3 gpurun_out/ncu_step.log
python tools/ncu_summary.py gpurun_out/r02_step_peak.ncu-rep > gpurun_out/r02_step_peak_summary.txt 2>&1
head -60 gpurun_out/r02_step_peak_summary.txt

# GPU-box helper (one call, several things): new parity test, csqrt A/B, source-level capture of items_kernel_dense with the reach bound on
cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "large_part or knobs or reach_bound" 2>&1 | tail -3
bash tools/gpu_variant_ab.sh base csq 2>&1 | grep -v "^$"
export BCU_NO_GRAPH=1 BCU_STREAMS=1 BCU_PRUNE=1
timeout 900 ncu --set full --import-source on --clock-control none -k regex:items_kernel --launch-skip 6 --launch-count 1 -o gpurun_out/r02_items_big_prune1 python tools/big_launch.py ncu 128 > gpurun_out/ncu_items_p1.log 2>&1
tail -2 gpurun_out/ncu_items_p1.log
python tools/ncu_lines.py gpurun_out/r02_items_big_prune1.ncu-rep > gpurun_out/r02_items_big_prune1_lines.txt 2>&1
head -40 gpurun_out/r02_items_big_prune1_lines.txt

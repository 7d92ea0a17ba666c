# GPU-box helper: ncu --set full capture (with source) of items_kernel_dense and apply_kernel on a LARGE round (128 tables x 4096
# strikes in one part, reach bound off, round 6) and the per-line breakdown of executed instructions and active lanes.
cd $GRAFT_REPO_ROOT
export BCU_NO_GRAPH=1 BCU_STREAMS=1
for k in items apply; do
  timeout 900 ncu --set full --import-source on --clock-control none -k regex:${k}_kernel --launch-skip 6 --launch-count 1 -o gpurun_out/r02_${k}_big python tools/big_launch.py ncu 128 > gpurun_out/ncu_${k}big.log 2>&1
  tail -2 gpurun_out/ncu_${k}big.log
  python tools/ncu_summary.py gpurun_out/r02_${k}_big.ncu-rep > gpurun_out/r02_${k}_dense_big_round_ncu_full.txt 2>&1
  python tools/ncu_lines.py gpurun_out/r02_${k}_big.ncu-rep > gpurun_out/r02_${k}_big_lines.txt 2>&1
  head -45 gpurun_out/r02_${k}_big_lines.txt
  grep -i "duration\|registers\|achieved occupancy\|executed ipc\|issue slots busy\|warp cycles per issued\|avg. active threads" gpurun_out/r02_${k}_dense_big_round_ncu_full.txt | head -12
done

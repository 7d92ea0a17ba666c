cd $GRAFT_REPO_ROOT
export BCU_NO_GRAPH=1 BCU_STREAMS=1
for k in items quartic apply; do
timeout 600 ncu --set full --import-source on --clock-control none -k regex:${k}_kernel --launch-skip 320 --launch-count 1 -o gpurun_out/r02_small_${k} python tools/level_small.py 32 > gpurun_out/ncu_small_$k.log 2>&1
python tools/ncu_summary.py gpurun_out/r02_small_${k}.ncu-rep > gpurun_out/r02_small_${k}_summary.txt 2>&1
done
head -40 gpurun_out/r02_small_items_summary.txt

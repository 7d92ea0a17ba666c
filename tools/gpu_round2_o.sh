cd $GRAFT_REPO_ROOT
python tools/level_small.py 32
python tools/level_small.py 8
BCU_PDL=0 python tools/level_small.py 32
BCU_NO_GRAPH=1 python tools/level_small.py 32
BCU_NO_GRAPH=1 BCU_STREAMS=1 timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none --cache-control none --csv --log-file gpurun_out/small_launch_list.csv python tools/level_small.py 32 > gpurun_out/ncu_small.log 2>&1
python tools/launch_rounds.py gpurun_out/small_launch_list.csv | head -40

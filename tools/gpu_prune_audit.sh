cd $GRAFT_REPO_ROOT
( timeout 420 python tools/prune_audit.py 16 262144 400 72000 | tail -2 ) 2>&1 | tee gpurun_out/r02_audit_16.txt | tail -1
( timeout 420 python tools/prune_audit.py 22 262144 250 71000 | tail -2 ) 2>&1 | tee gpurun_out/r02_audit_22.txt | tail -1
( timeout 200 python tools/prune_audit.py 8 262144 200 73000 | tail -2 ) 2>&1 | tee gpurun_out/r02_audit_8.txt | tail -1

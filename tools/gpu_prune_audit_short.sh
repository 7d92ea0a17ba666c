# GPU-box helper: audit of the reach bound on fresh seeds with the final build (tools/prune_audit.py: mode 2 counts marked jobs that
# would have mattered; mode 0 against mode 1 bit for bit).  usage: gpu_prune_audit_short.sh [tables16 tables22 tables8]
cd $GRAFT_REPO_ROOT
A=${1:-120}; B=${2:-80}; C=${3:-0}
( timeout 200 python tools/prune_audit.py 16 262144 $A 76000 | tail -2 ) 2>&1 | tee gpurun_out/r02_audit_final_16.txt | tail -1
( timeout 200 python tools/prune_audit.py 22 262144 $B 77000 | tail -2 ) 2>&1 | tee gpurun_out/r02_audit_final_22.txt | tail -1
[ "$C" -gt 0 ] && ( timeout 100 python tools/prune_audit.py 8 262144 $C 78000 | tail -2 ) 2>&1 | tee gpurun_out/r02_audit_final_8.txt | tail -1
true

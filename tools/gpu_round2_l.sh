cd $GRAFT_REPO_ROOT
for i in 1 2 3; do
timeout 600 cuda-gdb -batch -ex "set pagination off" -ex run -ex bt -ex "info registers rip" --args python -m pytest tests/test_gpu_parity.py -x -q -k "not filmed and not thesis" > gpurun_out/l_gdb_$i.log 2>&1
grep -n "SIGSEGV\|^#[0-9]" gpurun_out/l_gdb_$i.log | head -30
grep -q SIGSEGV gpurun_out/l_gdb_$i.log && break
done

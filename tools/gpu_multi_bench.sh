# multi-GPU: bench under torchrun (N = number of visible GPUs), one 2-rank merge test
cd $GRAFT_REPO_ROOT
N=$(nvidia-smi -L | wc -l)
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/r02_bench_${N}gpu.json 2> gpurun_out/r02_bench_${N}gpu.err; echo "bench exit $?"
python - <<P
import json
d=json.load(open("gpurun_out/r02_bench_${N}gpu.json"))
print({k:d[k] for k in ("n_gpus","value","value_exact","ms_per_step")}, d["e2e"], d["parity"])
for k,v in (d["extra"] or {}).items(): print(k, {a:b for a,b in v.items() if a!="workload"})
P
tail -3 gpurun_out/r02_bench_${N}gpu.err

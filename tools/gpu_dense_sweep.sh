# GPU-box helper: where does the dense instantiation of items / quartic (bcu_pipeline.cu launch_pipeline) start to pay?
# One launch of T random 16-ball tables x 4096 strikes, never dense against always dense, then the bench level both ways.
cd $GRAFT_REPO_ROOT
for nt in 16 32 64 128 256 512 1000; do
  BCU_DENSE_MIN=1000000000 timeout 300 python tools/big_launch.py sparse_$nt $nt
  BCU_DENSE_MIN=0 timeout 300 python tools/big_launch.py dense_$nt $nt
done
timeout 300 python tools/big_launch.py default_1000 1000
bash tools/gpu_ab.sh "level_default||-" "level_dense|BCU_DENSE_MIN=0|-"

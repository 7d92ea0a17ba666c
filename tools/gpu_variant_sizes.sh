# GPU-box helper: library variants (billiard_ai_b200/_variants/lib_<name>.so) across launch sizes: T random 16-ball tables x
# 4096 strikes in one launch, T = 16 .. 1000, then the bench level.   tools/gpu_variant_sizes.sh name ...
cd $GRAFT_REPO_ROOT
V=billiard_ai_b200/_variants
for nt in 16 32 64 128 256 1000; do
  for v in "$@"; do BCU_LIB=$V/lib_$v.so timeout 300 python tools/big_launch.py ${v}_$nt $nt; done
done
specs=""
for v in "$@"; do specs="$specs level_$v||$V/lib_$v.so"; done
bash tools/gpu_ab.sh $specs

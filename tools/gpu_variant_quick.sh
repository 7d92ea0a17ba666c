# GPU-box helper: library variants on one large launch (1000 tables x 4096 strikes) and on the bench level.  tools/gpu_variant_quick.sh name ...
cd $GRAFT_REPO_ROOT
V=billiard_ai_b200/_variants
for v in "$@"; do BCU_LIB=$V/lib_$v.so timeout 300 python tools/big_launch.py ${v}_1000 1000; done
specs=""
for v in "$@"; do specs="$specs level_$v||$V/lib_$v.so"; done
bash tools/gpu_ab.sh $specs

# GPU-box helper: A/B of library variants (billiard_ai_b200/_variants/lib_<name>.so) on a large launch, a small one and the bench level;
# the parity tests run with the last one.
cd $GRAFT_REPO_ROOT
V=billiard_ai_b200/_variants
for v in "$@"; do
  BCU_LIB=$V/lib_$v.so timeout 300 python tools/big_launch.py ${v}_1000 1000
  BCU_LIB=$V/lib_$v.so timeout 300 python tools/big_launch.py ${v}_16 16
done
specs=""
for v in "$@"; do specs="$specs level_$v||$V/lib_$v.so"; done
bash tools/gpu_ab.sh $specs
last="${@: -1}"
BCU_LIB=$V/lib_$last.so timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_size.py -m gpu -x -q --deselect tests/test_gpu_parity.py::test_native_library_is_loaded 2>&1 | tail -4

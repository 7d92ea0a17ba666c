# GPU-box helper: one ncu --set full capture (with source) of the large-part quartic kernel (quartic_kernel_wide) on a LARGE round (128 tables x 4096 strikes in
# one part, reach bound off, round 2) and the per-function / per-line breakdown of executed instructions and active lanes.
cd $GRAFT_REPO_ROOT
export BCU_NO_GRAPH=1 BCU_STREAMS=1
timeout 900 ncu --set full --import-source on --clock-control none -k regex:quartic_kernel --launch-skip 2 --launch-count 1 -o gpurun_out/r02_quartic_big python tools/big_launch.py ncu 128 > gpurun_out/ncu_qbig.log 2>&1
tail -3 gpurun_out/ncu_qbig.log
python tools/ncu_summary.py gpurun_out/r02_quartic_big.ncu-rep > gpurun_out/r02_quartic_dense_big_round_ncu_full.txt 2>&1
R="bcu_math.cuh:38-41:c_mul bcu_math.cuh:47-83:c_div bcu_math.cuh:87-114:hypot bcu_math.cuh:117-167:csqrt bcu_math.cuh:172-206:x2y2m1 bcu_math.cuh:209-249:clog bcu_math.cuh:252-260:c_pow_real"
R="$R bcu_libm.cuh:32-57:fma_bits bcu_libm.cuh:58-93:bcu_div bcu_libm.cuh:116-164:exp bcu_libm.cuh:165-232:log bcu_libm.cuh:233-321:pow bcu_libm.cuh:322-349:cubes3 bcu_libm.cuh:350-430:log1p bcu_libm.cuh:431-514:sincos bcu_libm.cuh:515-613:atan2"
R="$R bcu_quartic.cuh:1-97:quartic"
python tools/ncu_lines.py gpurun_out/r02_quartic_big.ncu-rep $R > gpurun_out/r02_quartic_big_functions.txt 2>&1
python tools/ncu_lines.py gpurun_out/r02_quartic_big.ncu-rep > gpurun_out/r02_quartic_big_lines.txt 2>&1
cat gpurun_out/r02_quartic_big_functions.txt; head -70 gpurun_out/r02_quartic_big_lines.txt

#!/bin/bash
# builds billiard_ai_b200/_variants/lib_<name>.so with extra -D flags: tools/build_variant.sh name -DITEM_MIN_BLOCKS=6 ...
set -e
name=$1; shift
mkdir -p billiard_ai_b200/_variants
cd billiard_ai_b200/csrc
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false -prec-div=true -prec-sqrt=true -ftz=false \
  -Xcompiler -fPIC,-ffp-contract=off,-O2 -shared "$@" bcu_kernels.cu bcu_pipeline.cu bcu_topk.cu bcu_merge.cu bcu_plan.cu bcu_api.cu bcu_rules.cpp bcu_search.cpp -o ../../billiard_ai_b200/_variants/lib_$name.so -ldl 2>&1 | grep -i "error" || true

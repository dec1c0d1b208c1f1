#!/bin/bash
# Build libb200attn.so with extra -D flags on one source file into build_variants/<name>.so (A/B runs on the GPU box:
# `cp build_variants/<name>.so ht_vllm_omni_b200/libb200attn.so` inside the gpurun command; the tree's own .so is untouched).
#   tests/tools/build_variant.sh <name> <source.cu> <-Dflag> [...]
set -e
name=$1; src=$2; shift 2
root=$(cd "$(dirname "$0")/../.." && pwd)
cd "$root/ht_vllm_omni_b200/csrc"
mkdir -p "$root/build_variants"
obj="$root/build_variants/${name}_$(basename "$src" .cu).o"
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -fvisibility=hidden \
     --expt-relaxed-constexpr -cudart static -Xptxas -v "$@" -c "$src" -o "$obj" 2> "$root/build_variants/${name}.ptxas.log"
objs=""
for o in abi kv_write rope_kv_write kv_gather decode decode_tc decode_simt decode_static prefill_tc plan_merge peer_exchange; do
  if [ "$o" = "$(basename "$src" .cu)" ]; then objs="$objs $obj"; else objs="$objs $o.o"; fi
done
nvcc -gencode arch=compute_100a,code=sm_100a -shared -cudart static -o "$root/build_variants/${name}.so" $objs
grep -A1 "prefill_tc_kernelILi1ELb0" "$root/build_variants/${name}.ptxas.log" | grep -i "spill\|registers" || true
echo "built build_variants/${name}.so"

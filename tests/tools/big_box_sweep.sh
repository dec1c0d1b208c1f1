set -x
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_fused_write.py tests/test_gpu_backend.py -m gpu -q -x 2>&1 | tail -5
for cfg in "--B 64 --L 4096 --Hq 16 --Hkv 2" "--B 32 --L 32768 --Hq 32 --Hkv 4 --layers 4"; do
for bs in 16 32 64 128; do for bb in 0 1; do
echo "== $cfg bs=$bs big_box=$bb"
B200ATTN_OPT_big_box=$bb timeout 120 python tests/tools/kernel_bench.py $cfg --bs $bs --graph 1 --pdl 1 2>&1 | grep "pdl=1: FUSED"
done; done; done

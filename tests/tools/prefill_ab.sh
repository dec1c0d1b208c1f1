for v in c1h1 c1h2 c2h1 c2h2; do
  cp build_variants/$v.so ht_vllm_omni_b200/libb200attn.so
  echo "=== $v"
  timeout 200 python -m pytest tests/test_gpu_prefill.py -m gpu -q -x 2>&1 | tail -2
  timeout 200 python tests/tools/prefill_bench.py --ours-only 2>&1 | sed -E 's/"requests.*"ours": //'
done
cp build_variants/c2h1.so ht_vllm_omni_b200/libb200attn.so; timeout 60 python tests/tools/prefill_profile.py prefill_8k 0 | tail -16

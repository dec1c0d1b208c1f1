/*
 * oracle.c — CPU ORACLE (test infrastructure + timed CPU baseline; never linked into the product).
 *
 * Plain-C / OpenMP restatement of the same algorithm as oracle/oracle.py, for sizes numpy is too slow for and as
 * the `cpu_baseline` / `--impl reference` arm of bench.py.  The reference's own CPU backend for this path
 * (vLLM CPUAttentionBackend, $VLLM/v1/attention/backends/cpu_attn.py:292-398 -> torch.ops._C.cpu_attention_with_kv_cache)
 * is NOT compiled into the CUDA wheel shipped in this image, so this port follows the published
 * algorithm of the GPU reference kernels instead:
 *   kv write  : $VLLM/v1/attention/ops/triton_reshape_and_cache_flash.py:44-124
 *   decode    : $VLLM/v1/attention/ops/triton_unified_attention.py:38-55, 395-545 (+ helpers softmax_step)
 * Semantics: fp8 tile -> fp32 * scale -> activation dtype (bf16/fp16 rounding), fp32 dot products, online softmax
 * over 16-token tiles, P rounded to the activation dtype before P.V, fp32 accumulate, final 1/L, output rounded.
 *
 * Build: make -C oracle   (gcc -O3 -march=x86-64-v3 -fopenmp -shared)
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

static float g_e4m3[256];
static int g_init = 0;

static void init_tables(void) {
  if (g_init) return;
  for (int b = 0; b < 256; ++b) {
    int e = (b >> 3) & 15, m = b & 7;
    float v;
    if (e == 15 && m == 7) v = NAN;
    else if (e == 0) v = ldexpf((float)m / 8.0f, -6);
    else v = ldexpf(1.0f + (float)m / 8.0f, e - 7);
    g_e4m3[b] = (b & 0x80) ? -v : v;
  }
  g_init = 1;
}

static inline float bf16_bits_to_f32(uint16_t u) {
  uint32_t w = (uint32_t)u << 16;
  float f;
  memcpy(&f, &w, 4);
  return f;
}
static inline float round_bf16(float x) {
  uint32_t u;
  memcpy(&u, &x, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return x; /* NaN */
  u = (u + 0x7fffu + ((u >> 16) & 1u)) & 0xffff0000u;
  float y;
  memcpy(&y, &u, 4);
  return y;
}
static inline float round_f16(float x) { return (float)(_Float16)x; }
static inline float act_round(float x, int act) { return act == 0 ? round_bf16(x) : round_f16(x); }

/* RN-even, saturate-to-finite float -> e4m3 (same results as oracle.py:e4m3_encode_satfinite) */
static inline uint8_t e4m3_encode(float x) {
  uint8_t sign = signbit(x) ? 0x80 : 0;
  if (isnan(x)) return sign | 0x7f;
  float a = fabsf(x);
  if (a >= 448.0f) return sign | 0x7e;
  /* binary search over the 127 non-negative codes */
  int lo = 0, hi = 126;
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (g_e4m3[mid] <= a) lo = mid; else hi = mid;
  }
  double dl = (double)a - g_e4m3[lo], dh = (double)g_e4m3[hi] - a;
  int code = dl < dh ? lo : (dh < dl ? hi : ((lo & 1) ? hi : lo));
  return sign | (uint8_t)code;
}

/* torchrun exports OMP_NUM_THREADS=1 to its workers; the timed CPU arm asks for every core it is allowed to run on */
void oracle_set_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

int oracle_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* key/value: 16-bit activations [T][Hkv][D] (act: 0 bf16, 1 fp16), caches [nb][bs][Hkv][D] contiguous (NHD).
 * kv_fp8: cache bytes are e4m3; else 16-bit copies.  Scales: scale_stride 0 (per tensor) or 1 (per head). */
void oracle_kv_write(const uint16_t* key, const uint16_t* value, void* k_cache, void* v_cache, const int64_t* slots,
                     int T, int Hkv, int D, int bs, int act, int kv_fp8, const float* k_scale, const float* v_scale,
                     int scale_stride) {
  init_tables();
#pragma omp parallel for schedule(static)
  for (int t = 0; t < T; ++t) {
    int64_t slot = slots[t];
    if (slot < 0) continue;
    size_t base = (size_t)slot * Hkv * D; /* (blk*bs + off) * Hkv * D */
    for (int h = 0; h < Hkv; ++h) {
      const uint16_t* ks = key + ((size_t)t * Hkv + h) * D;
      const uint16_t* vs = value + ((size_t)t * Hkv + h) * D;
      if (kv_fp8) {
        float sk = k_scale[h * scale_stride], sv = v_scale[h * scale_stride];
        uint8_t* kd = (uint8_t*)k_cache + base + (size_t)h * D;
        uint8_t* vd = (uint8_t*)v_cache + base + (size_t)h * D;
        for (int d = 0; d < D; ++d) {
          float kf = act == 0 ? bf16_bits_to_f32(ks[d]) : (float)(*(const _Float16*)&ks[d]);
          float vf = act == 0 ? bf16_bits_to_f32(vs[d]) : (float)(*(const _Float16*)&vs[d]);
          kd[d] = e4m3_encode(kf / sk);
          vd[d] = e4m3_encode(vf / sv);
        }
      } else {
        memcpy((uint16_t*)k_cache + base + (size_t)h * D, ks, (size_t)D * 2);
        memcpy((uint16_t*)v_cache + base + (size_t)h * D, vs, (size_t)D * 2);
      }
    }
  }
}

/* q [B][Hq][D] 16-bit, caches NHD contiguous, out [B][Hq][D] float32 (already rounded to the activation dtype). */
void oracle_paged_decode(const uint16_t* q, const void* k_cache, const void* v_cache, const int32_t* block_table,
                         int64_t bt_stride, const int32_t* seq_lens, int B, int Hq, int Hkv, int D, int bs, int act,
                         int kv_fp8, float sm_scale, const float* k_scale, const float* v_scale, int scale_stride,
                         float* out) {
  init_tables();
  const int G = Hq / Hkv;
  const int TILE = 16;
#pragma omp parallel
  {
    float* Kt = (float*)malloc(sizeof(float) * TILE * D);
    float* Vt = (float*)malloc(sizeof(float) * TILE * D);
    float* acc = (float*)malloc(sizeof(float) * G * D);
    float* qf = (float*)malloc(sizeof(float) * G * D);
    float* S = (float*)malloc(sizeof(float) * G * TILE);
    float* M = (float*)malloc(sizeof(float) * G);
    float* Ls = (float*)malloc(sizeof(float) * G);
#pragma omp for schedule(dynamic, 1) collapse(2)
    for (int b = 0; b < B; ++b) {
      for (int h = 0; h < Hkv; ++h) {
        const int L = seq_lens[b];
        float* o = out + ((size_t)b * Hq + (size_t)h * G) * D;
        if (L <= 0) { memset(o, 0, sizeof(float) * G * D); continue; }
        const float sk = kv_fp8 ? k_scale[h * scale_stride] : 1.0f, sv = kv_fp8 ? v_scale[h * scale_stride] : 1.0f;
        for (int g = 0; g < G; ++g) {
          M[g] = -INFINITY; Ls[g] = 1.0f;
          const uint16_t* qs = q + ((size_t)b * Hq + (size_t)h * G + g) * D;
          for (int d = 0; d < D; ++d) {
            qf[g * D + d] = act == 0 ? bf16_bits_to_f32(qs[d]) : (float)(*(const _Float16*)&qs[d]);
            acc[g * D + d] = 0.0f;
          }
        }
        for (int j = 0; j < L; j += TILE) {
          const int n = L - j < TILE ? L - j : TILE;
          for (int i = 0; i < n; ++i) {
            const int tok = j + i;
            const size_t blk = (size_t)block_table[(size_t)b * bt_stride + tok / bs];
            const size_t off = ((blk * bs + (size_t)(tok % bs)) * Hkv + h) * D;
            if (kv_fp8) {
              const uint8_t* kp = (const uint8_t*)k_cache + off;
              const uint8_t* vp = (const uint8_t*)v_cache + off;
              for (int d = 0; d < D; ++d) {
                Kt[i * D + d] = act_round(g_e4m3[kp[d]] * sk, act);
                Vt[i * D + d] = act_round(g_e4m3[vp[d]] * sv, act);
              }
            } else {
              const uint16_t* kp = (const uint16_t*)k_cache + off;
              const uint16_t* vp = (const uint16_t*)v_cache + off;
              for (int d = 0; d < D; ++d) {
                Kt[i * D + d] = act == 0 ? bf16_bits_to_f32(kp[d]) : (float)(*(const _Float16*)&kp[d]);
                Vt[i * D + d] = act == 0 ? bf16_bits_to_f32(vp[d]) : (float)(*(const _Float16*)&vp[d]);
              }
            }
          }
          for (int g = 0; g < G; ++g) {
            float mx = M[g];
            for (int i = 0; i < n; ++i) {
              float s = 0.0f;
              const float* qr = qf + g * D;
              const float* kr = Kt + i * D;
              for (int d = 0; d < D; ++d) s += qr[d] * kr[d];
              s *= sm_scale;
              S[g * TILE + i] = s;
              if (s > mx) mx = s;
            }
            const float alpha = expf(M[g] - mx);
            float lsum = 0.0f;
            float* ar = acc + g * D;
            for (int d = 0; d < D; ++d) ar[d] *= alpha;
            for (int i = 0; i < n; ++i) {
              const float p = expf(S[g * TILE + i] - mx);
              lsum += p;
              const float pr = act_round(p, act);
              const float* vr = Vt + i * D;
              for (int d = 0; d < D; ++d) ar[d] += pr * vr[d];
            }
            Ls[g] = Ls[g] * alpha + lsum;
            M[g] = mx;
          }
        }
        for (int g = 0; g < G; ++g)
          for (int d = 0; d < D; ++d) o[g * D + d] = act_round(acc[g * D + d] / Ls[g], act);
      }
    }
    free(Kt); free(Vt); free(acc); free(qf); free(S); free(M); free(Ls);
  }
}

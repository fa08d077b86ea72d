/*
 * TEST INFRASTRUCTURE ONLY -- plain-C restatement of the reference's hot-path algorithms.
 * Second, independent statement of oracle/paged_kv_oracle.py (the two are cross-checked in
 * tests/test_oracle_golden.py) and the multi-threaded CPU baseline that bench.py can time.
 * Built by oracle/build_oracle.sh into oracle/_build/liboracle_c.so; never linked into or
 * loaded by the product (vllm_splitwise_b200/).
 *
 * Cited reference lines are relative to the reference tree.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* float32 -> fp8 e5m2, RNE, saturate-to-finite: what __nv_cvt_float_to_fp8(x, __NV_SATFINITE,
 * __NV_E5M2) returns on sm_89+ (csrc/quantization/fp8_e5m2_kvcache/quant_utils.cuh:164-192). */
uint8_t oracle_e5m2_from_f32(float x) {
  uint32_t bits;
  memcpy(&bits, &x, 4);
  const uint8_t sign = (uint8_t)((bits >> 31) << 7);
  const uint32_t a = bits & 0x7FFFFFFFu;
  if (a > 0x7F800000u) return (uint8_t)0x7F;                /* NaN -> canonical +NaN, sign dropped */
  if (a <= 0x37000000u) return sign;                        /* |x| <= 2^-17 -> 0 */
  const int exp = (int)(a >> 23) - 127 + 15;
  const uint32_t man = a & 0x7FFFFFu;
  uint32_t res;
  if (exp >= 1) {
    res = ((uint32_t)exp << 2) | (man >> 21);
    const uint32_t rem = man & 0x1FFFFFu;
    if (rem > 0x100000u || (rem == 0x100000u && (res & 1u))) res += 1;
  } else {
    const int drop = 21 + (1 - exp);
    const uint64_t full = (uint64_t)(man | 0x800000u);
    res = (uint32_t)(full >> drop);
    const uint64_t rem = full & (((uint64_t)1 << drop) - 1);
    const uint64_t half = (uint64_t)1 << (drop - 1);
    if (rem > half || (rem == half && (res & 1u))) res += 1;
  }
  if (res > 0x7Bu) res = 0x7Bu;                             /* overflow / Inf -> 57344 */
  return (uint8_t)(sign | res);
}

/* fp8 e5m2 -> float32 (exact): the byte is the high byte of an IEEE half (quant_utils.cuh:23-29). */
float oracle_e5m2_to_f32(uint8_t b) {
  const uint32_t sign = (uint32_t)(b & 0x80) << 24;
  const int e = (b >> 2) & 0x1F, m = b & 3;
  float v;
  if (e == 0) v = ldexpf((float)m, -16);
  else if (e == 31) v = m ? NAN : INFINITY;
  else v = ldexpf(1.0f + (float)m * 0.25f, e - 15);
  uint32_t bits;
  memcpy(&bits, &v, 4);
  bits |= sign;
  memcpy(&v, &bits, 4);
  return v;
}

void oracle_e5m2_encode(const float* src, uint8_t* dst, int64_t n) {
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i) dst[i] = oracle_e5m2_from_f32(src[i]);
}
void oracle_e5m2_decode(const uint8_t* src, float* dst, int64_t n) {
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i) dst[i] = oracle_e5m2_to_f32(src[i]);
}

/* reshape_and_cache on raw elements of `elt` bytes (csrc/cache_kernels.cu:154-210):
 * key_cache[blk][h][d/x][off][d%x] = key[i][h][d], value_cache[blk][h][d][off] = value[i][h][d]. */
void oracle_reshape_and_cache(const uint8_t* key, const uint8_t* value, uint8_t* key_cache,
                              uint8_t* value_cache, const int64_t* slot_mapping, int num_tokens,
                              int num_heads, int head_size, int block_size, int x,
                              int64_t key_stride, int64_t value_stride, int elt) {
  for (int i = 0; i < num_tokens; ++i) {
    const int64_t slot = slot_mapping[i];
    if (slot < 0) continue; /* padding, cache_kernels.cu:169-172 */
    const int64_t blk = slot / block_size, off = slot % block_size;
    for (int h = 0; h < num_heads; ++h)
      for (int d = 0; d < head_size; ++d) {
        const int64_t ksrc = (int64_t)i * key_stride + (int64_t)h * head_size + d;
        const int64_t vsrc = (int64_t)i * value_stride + (int64_t)h * head_size + d;
        const int64_t kdst = (((blk * num_heads + h) * (head_size / x) + d / x) * block_size + off) * x + d % x;
        const int64_t vdst = ((blk * num_heads + h) * head_size + d) * block_size + off;
        memcpy(key_cache + kdst * elt, key + ksrc * elt, (size_t)elt);
        memcpy(value_cache + vdst * elt, value + vsrc * elt, (size_t)elt);
      }
  }
}

/* copy_blocks for one buffer (csrc/cache_kernels.cu:65-91). */
void oracle_copy_blocks(uint8_t* cache, const int64_t* pairs, int num_pairs, int64_t block_bytes) {
  for (int p = 0; p < num_pairs; ++p)
    memmove(cache + pairs[2 * p + 1] * block_bytes, cache + pairs[2 * p] * block_bytes, (size_t)block_bytes);
}

/* Decode attention, all fp32 (inputs widened by the caller), one (seq, head) per task.
 * Formula of paged_attention_kernel (csrc/attention/attention_kernels.cu:83-417) /
 * ref_single_query_cached_kv_attention (tests/kernels/test_attention.py:58-111):
 *   out[s,h,:] = softmax_t(scale*q.K_t + alibi_h*(t-ctx+1)) . V_t ,  t < ctx.
 * key_cache [NB,Hkv,D/x,bs,x], value_cache [NB,Hkv,D,bs], x given by the caller. */
void oracle_paged_attention_f32(float* out, const float* q, const float* key_cache, const float* value_cache,
                                const int32_t* block_tables, const int32_t* context_lens,
                                const float* alibi_slopes, float scale, int num_seqs, int num_heads,
                                int num_kv_heads, int head_size, int block_size, int x,
                                int max_num_blocks_per_seq) {
  const int group = num_heads / num_kv_heads;
#pragma omp parallel for collapse(2) schedule(dynamic)
  for (int s = 0; s < num_seqs; ++s)
    for (int h = 0; h < num_heads; ++h) {
      const int ctx = context_lens[s], kvh = h / group;
      const float* qv = q + ((int64_t)s * num_heads + h) * head_size;
      float* o = out + ((int64_t)s * num_heads + h) * head_size;
      float* logits = (float*)malloc(sizeof(float) * (size_t)(ctx > 0 ? ctx : 1));
      float mx = -INFINITY;
      for (int t = 0; t < ctx; ++t) {
        const int64_t blk = block_tables[(int64_t)s * max_num_blocks_per_seq + t / block_size];
        const int off = t % block_size;
        const float* kb = key_cache + (blk * num_kv_heads + kvh) * (int64_t)head_size * block_size;
        float acc = 0.f;
        for (int d = 0; d < head_size; ++d) acc += qv[d] * kb[((int64_t)(d / x) * block_size + off) * x + d % x];
        acc *= scale;
        if (alibi_slopes) acc += alibi_slopes[h] * (float)(t - ctx + 1);
        logits[t] = acc;
        if (acc > mx) mx = acc;
      }
      float sum = 0.f;
      for (int t = 0; t < ctx; ++t) { logits[t] = expf(logits[t] - mx); sum += logits[t]; }
      for (int d = 0; d < head_size; ++d) o[d] = 0.f;
      for (int t = 0; t < ctx; ++t) {
        const int64_t blk = block_tables[(int64_t)s * max_num_blocks_per_seq + t / block_size];
        const int off = t % block_size;
        const float* vb = value_cache + (blk * num_kv_heads + kvh) * (int64_t)head_size * block_size;
        const float p = logits[t] / sum;
        for (int d = 0; d < head_size; ++d) o[d] += p * vb[(int64_t)d * block_size + off];
      }
      free(logits);
    }
}

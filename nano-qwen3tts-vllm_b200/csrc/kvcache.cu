// kvcache.cu - a1: scatter new K/V rows into the paged cache at slot_mapping.
// Replaces store_kvcache_kernel / store_kvcache (layers/attention.py:10-54).
//
// HBM-bound bit copy: per token 2*kv_dim*2 bytes read + the same written (8196 B for
// kv_dim = 1024).  One CTA per KV_TOK tokens; every thread moves 16-byte vectors with
// streaming (no-L1-allocate) loads and plain stores and has the K and V vectors of all KV_TOK
// tokens in flight before the first store (8 independent loads per thread: with one CTA per
// (token, K|V) and one load per thread the launch was a chain of two dependent round trips
// per tiny CTA, 0.60 of the HBM peak at 4096 tokens).
#include "common.cuh"

namespace b2 {

constexpr int KV_TOK = 4;

__global__ void __launch_bounds__(128)
store_kvcache_kernel(const __nv_bfloat16* __restrict__ key, const __nv_bfloat16* __restrict__ value,
                     int64_t k_stride, int64_t v_stride,
                     __nv_bfloat16* __restrict__ k_cache, __nv_bfloat16* __restrict__ v_cache,
                     const int32_t* __restrict__ slot_mapping, int kv_dim, int n_tokens) {
  const int tok0 = blockIdx.x * KV_TOK;
  int slot[KV_TOK];
#pragma unroll
  for (int t = 0; t < KV_TOK; ++t) slot[t] = tok0 + t < n_tokens ? __ldg(slot_mapping + tok0 + t) : -1;   // attention.py:23 (slot == -1)
  const int nvec = kv_dim >> 3;                             // 8 bf16 per 16 bytes
  for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
    uint4 kv[KV_TOK][2];
#pragma unroll
    for (int t = 0; t < KV_TOK; ++t)
      if (slot[t] >= 0) {
        kv[t][0] = ldg_nc_v4(key + (int64_t)(tok0 + t) * k_stride + 8 * i);
        kv[t][1] = ldg_nc_v4(value + (int64_t)(tok0 + t) * v_stride + 8 * i);
      }
#pragma unroll
    for (int t = 0; t < KV_TOK; ++t)
      if (slot[t] >= 0) {
        stg_v4(k_cache + (int64_t)slot[t] * kv_dim + 8 * i, kv[t][0]);
        stg_v4(v_cache + (int64_t)slot[t] * kv_dim + 8 * i, kv[t][1]);
      }
  }
}

}  // namespace b2

extern "C" int b2tts_store_kvcache(const void* key, const void* value,
                                   int64_t key_row_stride, int64_t value_row_stride,
                                   void* k_cache, void* v_cache, const int32_t* slot_mapping,
                                   int n_tokens, int kv_dim, void* stream) {
  using namespace b2;
  B2_REQUIRE(n_tokens >= 0 && kv_dim > 0, B2TTS_EINVAL, "store_kvcache: n_tokens=%d kv_dim=%d", n_tokens, kv_dim);
  if (n_tokens == 0) return B2TTS_OK;
  B2_REQUIRE(key && value && k_cache && v_cache && slot_mapping, B2TTS_EINVAL, "store_kvcache: null pointer");
  B2_REQUIRE(kv_dim % 8 == 0 && key_row_stride % 8 == 0 && value_row_stride % 8 == 0, B2TTS_EALIGN,
             "store_kvcache: kv_dim/strides must be multiples of 8 elements (16 B), got %d/%lld/%lld", kv_dim,
             (long long)key_row_stride, (long long)value_row_stride);
  B2_REQUIRE(aligned16(key) && aligned16(value) && aligned16(k_cache) && aligned16(v_cache), B2TTS_EALIGN,
             "store_kvcache: pointers must be 16-byte aligned");
  dim3 grid((n_tokens + KV_TOK - 1) / KV_TOK);
  store_kvcache_kernel<<<grid, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(key), static_cast<const __nv_bfloat16*>(value), key_row_stride,
      value_row_stride, static_cast<__nv_bfloat16*>(k_cache), static_cast<__nv_bfloat16*>(v_cache),
      slot_mapping, kv_dim, n_tokens);
  B2_CHECK_LAUNCH("store_kvcache");
  return B2TTS_OK;
}

// kvcache.cu - a1: scatter new K/V rows into the paged cache at slot_mapping.
// Replaces store_kvcache_kernel / store_kvcache (layers/attention.py:10-54).
//
// HBM-bound bit copy: per token 2*kv_dim*2 bytes read + the same written (8196 B for
// kv_dim = 1024).  One CTA per (token, K|V); every thread moves 16-byte vectors with
// streaming (no-L1-allocate) loads and plain stores, so a 2 KiB row is one fully
// coalesced 128-thread x 16 B transaction pair.
#include "common.cuh"

namespace b2 {

__global__ void __launch_bounds__(128)
store_kvcache_kernel(const __nv_bfloat16* __restrict__ key, const __nv_bfloat16* __restrict__ value,
                     int64_t k_stride, int64_t v_stride,
                     __nv_bfloat16* __restrict__ k_cache, __nv_bfloat16* __restrict__ v_cache,
                     const int32_t* __restrict__ slot_mapping, int kv_dim) {
  const int tok = blockIdx.x;
  const int slot = __ldg(slot_mapping + tok);
  if (slot < 0) return;                                     // attention.py:23 (slot == -1)
  const bool is_v = blockIdx.y != 0;
  const __nv_bfloat16* src = is_v ? value + (int64_t)tok * v_stride : key + (int64_t)tok * k_stride;
  __nv_bfloat16* dst = (is_v ? v_cache : k_cache) + (int64_t)slot * kv_dim;
  const int nvec = kv_dim >> 3;                             // 8 bf16 per 16 bytes
  for (int i = threadIdx.x; i < nvec; i += blockDim.x)
    stg_v4(dst + 8 * i, ldg_nc_v4(src + 8 * i));
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
  dim3 grid(n_tokens, 2);
  store_kvcache_kernel<<<grid, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(key), static_cast<const __nv_bfloat16*>(value), key_row_stride,
      value_row_stride, static_cast<__nv_bfloat16*>(k_cache), static_cast<__nv_bfloat16*>(v_cache),
      slot_mapping, kv_dim);
  B2_CHECK_LAUNCH("store_kvcache");
  return B2TTS_OK;
}

// sampler.cu - a8: temperature -> top-k threshold (ties kept) -> softmax -> argmax(p / noise).
// Replaces Sampler.forward (layers/sampler.py:15-24) and the torch.multinomial single-sample
// fast path it ends in (argmax(probs / q), q ~ Exp(1) supplied by the caller so the Philox
// stream is the reference's own).
//
// One CTA per row; the row (V <= 16384 fp32) lives in shared memory for the whole call, so
// HBM sees each logit and each noise value exactly once: 2V (bf16) + 4V (noise) + 12 bytes.
// The k-th largest value is found exactly with a 4-pass 8-bit radix select on the
// order-preserving integer image of the fp32 values - no sort, no approximation, so the
// `logits < kth` tie rule of sampler.py:19-20 is reproduced bit for bit.
#include "common.cuh"

namespace b2 {

constexpr int SMP_THREADS = 256;
constexpr int SMP_MAXV = 16384;

__device__ __forceinline__ uint32_t f2key(float f) {       // larger float -> larger key
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key2f(uint32_t k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

template <bool FP32_IN>
__global__ void __launch_bounds__(SMP_THREADS)
topk_sample_kernel(const void* __restrict__ logits, int64_t row_stride, const float* __restrict__ temperatures,
                   const float* __restrict__ noise, int64_t* __restrict__ out_index, float* __restrict__ probs_out,
                   int vocab, int top_k) {
  extern __shared__ float vals[];                  // [vocab]
  __shared__ uint32_t hist[256];
  __shared__ uint32_t sel_prefix, sel_remaining;
  __shared__ float redf[SMP_THREADS / 32];
  __shared__ int redi[SMP_THREADS / 32];
  __shared__ float bcast;

  const int row = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const float T = __ldg(temperatures + row);

  // ---- l = float(logit) / T                                              sampler.py:16
  float lmax = -INFINITY;
  for (int i = tid; i < vocab; i += SMP_THREADS) {
    float x;
    if (FP32_IN) x = __ldg(static_cast<const float*>(logits) + (int64_t)row * row_stride + i);
    else x = __bfloat162float(static_cast<const __nv_bfloat16*>(logits)[(int64_t)row * row_stride + i]);
    x = x / T;                                      // IEEE division (no fast-math)
    vals[i] = x;
    lmax = fmaxf(lmax, x);
  }
  if (tid == 0) { sel_prefix = 0; sel_remaining = (uint32_t)top_k; }
  lmax = warp_max(lmax);
  if (lane == 0) redf[warp] = lmax;
  __syncthreads();
  lmax = redf[0];
#pragma unroll
  for (int w = 1; w < SMP_THREADS / 32; ++w) lmax = fmaxf(lmax, redf[w]);

  // ---- kth = topk(l, k).values[:, -1] by radix select                    sampler.py:18-19
  float kth = -INFINITY;
  if (top_k < vocab) {
    for (int pass = 0; pass < 4; ++pass) {
      const int shift = 24 - 8 * pass;
      hist[tid] = 0;                                // SMP_THREADS == 256 bins
      __syncthreads();
      const uint32_t prefix = sel_prefix;
      const uint32_t himask = pass == 0 ? 0u : (0xffffffffu << (shift + 8));
      // logits of one row share a handful of sign/exponent bytes, so plain atomics would serialise on 2-8 bins:
      // lanes with the same bin are grouped (match.any) and their leader adds the group's count once
      for (int i0 = 0; i0 < vocab; i0 += SMP_THREADS) {
        const int i = i0 + tid;
        uint32_t bin = 256u;                         // 256 = not a candidate
        if (i < vocab) {
          const uint32_t key = f2key(vals[i]);
          if ((key & himask) == prefix) bin = (key >> shift) & 255u;
        }
        const uint32_t peers = __match_any_sync(0xffffffffu, bin);
        if (bin < 256u && lane == __ffs(peers) - 1) atomicAdd(&hist[bin], (uint32_t)__popc(peers));
      }
      __syncthreads();
      if (warp == 0) {
        // bins walked from the largest digit down: lane l owns bins [255-8l-7, 255-8l]; an exclusive
        // warp scan of the per-lane totals finds the lane, 8 sequential steps find the bin.
        const uint32_t rem = sel_remaining;
        uint32_t c[8], tot = 0;
#pragma unroll
        for (int e = 0; e < 8; ++e) { c[e] = hist[255 - (8 * lane + e)]; tot += c[e]; }
        uint32_t incl = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
          if (lane >= o) incl += t;
        }
        const uint32_t excl = incl - tot;
        // the owning lane: first lane whose inclusive count reaches rem (lane 31 if none: digit 0)
        const uint32_t hit = __ballot_sync(0xffffffffu, incl >= rem);
        const int owner = hit ? __ffs(hit) - 1 : 31;
        if (lane == owner) {
          uint32_t left = rem - excl, d = 255 - 8 * lane;
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            if (c[e] >= left || (lane == 31 && e == 7)) { d = 255 - (8 * lane + e); break; }
            left -= c[e];
          }
          sel_prefix = prefix | (d << shift);
          sel_remaining = left;
        }
      }
      __syncthreads();
    }
    kth = key2f(sel_prefix);
  }

  // ---- p = softmax(masked l)                                             sampler.py:20-21
  float sum = 0.f;
  for (int i = tid; i < vocab; i += SMP_THREADS) {
    const float x = vals[i];
    const float e = x < kth ? 0.f : expf(x - lmax);
    vals[i] = e;
    sum += e;
  }
  sum = warp_sum(sum);
  __syncthreads();                                  // redf reuse
  if (lane == 0) redf[warp] = sum;
  __syncthreads();
  if (tid == 0) {
    float t = 0.f;
    for (int w = 0; w < SMP_THREADS / 32; ++w) t += redf[w];
    bcast = t;
  }
  __syncthreads();
  sum = bcast;

  // ---- index = argmax(p / q)                    torch.multinomial(probs, 1), sampler.py:22
  float best = -INFINITY;
  int besti = 0x7fffffff;
  for (int i = tid; i < vocab; i += SMP_THREADS) {
    const float pr = vals[i] / sum;
    if (probs_out) probs_out[(int64_t)row * vocab + i] = pr;
    const float sc = pr / __ldg(noise + (int64_t)row * vocab + i);
    if (sc > best) { best = sc; besti = i; }        // strict: first index wins inside a thread
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ob = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, besti, o);
    if (ob > best || (ob == best && oi < besti)) { best = ob; besti = oi; }
  }
  __syncthreads();
  if (lane == 0) { redf[warp] = best; redi[warp] = besti; }
  __syncthreads();
  if (tid == 0) {
    for (int w = 1; w < SMP_THREADS / 32; ++w)
      if (redf[w] > best || (redf[w] == best && redi[w] < besti)) { best = redf[w]; besti = redi[w]; }
    out_index[row] = besti == 0x7fffffff ? 0 : besti;
  }
}

}  // namespace b2

extern "C" int b2tts_topk_sample(const void* logits, int64_t logits_row_stride, int logits_fp32,
                                 const float* temperatures, const float* noise,
                                 int64_t* out_index, float* probs_out,
                                 int batch, int vocab, int top_k, float top_p, int apply_top_p,
                                 void* stream) {
  using namespace b2;
  (void)top_p;
  B2_REQUIRE(batch >= 0 && vocab > 0 && top_k > 0, B2TTS_EINVAL, "topk_sample: batch=%d vocab=%d top_k=%d", batch, vocab, top_k);
  if (batch == 0) return B2TTS_OK;
  B2_REQUIRE(logits && temperatures && noise && out_index, B2TTS_EINVAL, "topk_sample: null pointer");
  B2_REQUIRE(vocab <= SMP_MAXV, B2TTS_EUNSUPPORTED, "topk_sample: vocab %d > %d", vocab, SMP_MAXV);
  B2_REQUIRE(!apply_top_p, B2TTS_EUNSUPPORTED,
             "topk_sample: top_p filtering is not implemented (the reference accepts and ignores top_p, sampler.py:15)");
  const size_t smem = (size_t)vocab * sizeof(float);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (smem > 48 * 1024) {
    static bool configured = false;
    if (!configured) {
      cudaError_t e = cudaFuncSetAttribute(topk_sample_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMP_MAXV * 4);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(topk_sample_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMP_MAXV * 4);
      if (e != cudaSuccess) return cuda_fail(e, "topk_sample");
      configured = true;
    }
  }
  if (logits_fp32) topk_sample_kernel<true><<<batch, SMP_THREADS, smem, st>>>(logits, logits_row_stride, temperatures, noise, out_index, probs_out, vocab, top_k);
  else topk_sample_kernel<false><<<batch, SMP_THREADS, smem, st>>>(logits, logits_row_stride, temperatures, noise, out_index, probs_out, vocab, top_k);
  B2_CHECK_LAUNCH("topk_sample");
  return B2TTS_OK;
}

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

__device__ __forceinline__ uint32_t f2key(float f) {       // larger float -> larger key; -0.0 and +0.0 share one key,
  const uint32_t u = __float_as_uint(f + 0.0f);            // as they compare equal in torch.topk and in `l < kth`
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

// ---------------------------------------------------------------------------------------------------
// Register-resident sampler for vocab <= 4096 (the codec heads: 3072 talker, 2048 predictor).  The kernel above
// is a chain of ~20 block barriers, shared-memory atomics and 4 histogram passes over a shared-memory copy of
// the row: ~18 us whatever the batch - pure latency, 16 times per codec frame.  Here one 256-thread CTA owns a
// row and keeps it in REGISTERS as order-preserving integer keys (8 contiguous values per thread per 16-byte
// load, every load in flight at once):
//   * max: REDUX.MAX per warp + one barrier;
//   * k-th largest key, exactly: greedy bitwise search from the MSB - "count keys >= candidate" is 8-16 register
//     compares, one REDUX.ADD and one barrier per bit; the largest key K with count(key >= K) >= k is the k-th
//     largest, so `logits < kth` (sampler.py:19-20, ties kept) is reproduced bit for bit.  The search stops as
//     soon as exactly k keys are left above the prefix (then K is their minimum): ~12-16 of the 32 steps;
//   * only the survivors (>= K, about k of them) need exp, a noise value and two divisions: their Exp(1) draws are
//     prefetched while the partition function is summed - HBM sees 2V bytes of logits and ~k noise values instead
//     of 4V of them;  masked entries have p = 0 and can never be the arg-max (the row maximum has p > 0);
//   * argmax(p / q) with the first index on ties, as torch.argmax / the kernel above;
//   * optional fused gather: the CTA copies embed_table[index] to embed_out[row] - the next predictor pass's
//     input embedding (models/qwen3_tts_predictor.py:94-95) without another launch.
struct RowSampleParams {
  const void* logits;
  int64_t row_stride;
  const float* temperatures;
  const float* noise;
  int64_t* out_index;
  float* probs_out;
  int vocab, top_k;
  const __nv_bfloat16* embed_table;
  __nv_bfloat16* embed_out;
  int64_t embed_out_stride;
  int embed_cols;
};

constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;

template <int NCH, bool FP32_IN>
__global__ void __launch_bounds__(RS_THREADS)
topk_sample_row_kernel(const RowSampleParams p) {
  __shared__ int red_cnt[2][RS_WARPS];
  __shared__ uint32_t red_key[RS_WARPS];
  __shared__ float red_f[RS_WARPS];
  __shared__ int red_i[RS_WARPS];
  __shared__ int bcast_i;
  const int row = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  pdl_launch();
  pdl_wait();                                         // logits come from the previous kernel on the stream
  const float T = __ldg(p.temperatures + row);
  uint32_t key[NCH * 8];
  uint32_t k16[FP32_IN ? 1 : NCH * 8];                // bf16 input: the 16-bit order keys of the logits (0 = padding)
#pragma unroll
  for (int j = 0; j < (FP32_IN ? 1 : NCH * 8); ++j) k16[j] = 0u;
  // ---- l = float(logit) / T as keys                                      sampler.py:16
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    const int i0 = (c * RS_THREADS + tid) * 8;
    float x[8];
    if (i0 < p.vocab) {                               // vocab % 8 == 0: whole chunks
      if (FP32_IN) {
        const float4* src = reinterpret_cast<const float4*>(static_cast<const float*>(p.logits) + (int64_t)row * p.row_stride + i0);
        const float4 a = __ldg(src), b = __ldg(src + 1);
        x[0] = a.x; x[1] = a.y; x[2] = a.z; x[3] = a.w; x[4] = b.x; x[5] = b.y; x[6] = b.z; x[7] = b.w;
      } else {
        const uint4 r = ldg_v4(static_cast<const __nv_bfloat16*>(p.logits) + (int64_t)row * p.row_stride + i0);
        x[0] = bf16lo(r.x); x[1] = bf16hi(r.x); x[2] = bf16lo(r.y); x[3] = bf16hi(r.y);
        x[4] = bf16lo(r.z); x[5] = bf16hi(r.z); x[6] = bf16lo(r.w); x[7] = bf16hi(r.w);
#pragma unroll
        for (int e = 0; e < 8; ++e) k16[c * 8 + e] = f2key(x[e]) >> 16;      // order key of the bf16 logit itself
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) key[c * 8 + e] = f2key(__fdiv_rn(x[e], T));    // IEEE division (no fast-math)
    } else {
#pragma unroll
      for (int e = 0; e < 8; ++e) key[c * 8 + e] = 0u;   // below every real key (f2key(-inf) = 0x007fffff)
    }
  }
  uint32_t kmax = 0u;
#pragma unroll
  for (int j = 0; j < NCH * 8; ++j) kmax = max(kmax, key[j]);
  kmax = __reduce_max_sync(0xffffffffu, kmax);
  if (lane == 0) red_key[warp] = kmax;
  __syncthreads();
#pragma unroll
  for (int w = 0; w < RS_WARPS; ++w) kmax = max(kmax, red_key[w]);
  const float lmax = key2f(kmax);

  // ---- kth = topk(l, k).values[:, -1]                                     sampler.py:18-19
  uint32_t kth = 1u;                                   // top_k >= vocab: only the padding keys (0) are masked
  if (!FP32_IN && T > 0.f && p.top_k < p.vocab) {
    // bf16 logits and T > 0: l = x / T is non-decreasing in x, so the k-th largest l is (k-th largest x) / T, and x has
    // only 16 bits: radix select over the bf16 order keys, TWO bits per round (the counts for digits 1..3 packed in one
    // word: one redux, one shared-memory word per warp, one block barrier) = 8 rounds instead of 32
    uint32_t prefix = 0u;
    int above = p.vocab, buf = 0;
    for (int shift = 14; shift >= 0 && above != p.top_k; shift -= 2) {
      uint32_t pk = 0u;
#pragma unroll
      for (int d = 1; d < 4; ++d) {
        const uint32_t cand = prefix | ((uint32_t)d << shift);
        uint32_t cnt = 0u;
#pragma unroll
        for (int j = 0; j < NCH * 8; ++j) cnt += k16[j] >= cand ? 1u : 0u;
        pk |= cnt << (10 * (d - 1));                   // a warp's field sum <= 32 * 16 = 512: 10 bits
      }
      pk = __reduce_add_sync(0xffffffffu, pk);
      if (lane == 0) red_cnt[buf][warp] = (int)pk;
      __syncthreads();
      int tot[3] = {0, 0, 0};
#pragma unroll
      for (int w = 0; w < RS_WARPS; ++w) {
        const uint32_t v = (uint32_t)red_cnt[buf][w];
        tot[0] += (int)(v & 0x3ffu); tot[1] += (int)((v >> 10) & 0x3ffu); tot[2] += (int)((v >> 20) & 0x3ffu);
      }
      buf ^= 1;
#pragma unroll
      for (int d = 3; d >= 1; --d)
        if (tot[d - 1] >= p.top_k) { prefix |= (uint32_t)d << shift; above = tot[d - 1]; break; }
    }
    uint32_t kx = prefix;                              // all 16 bits decided: ties straddle the k-th place
    if (above == p.top_k) {                            // exactly k logits at or above the prefix: their minimum
      uint32_t kmin = 0xffffu;
#pragma unroll
      for (int j = 0; j < NCH * 8; ++j) if (k16[j] >= prefix) kmin = min(kmin, k16[j]);
      kmin = __reduce_min_sync(0xffffffffu, kmin);
      __syncthreads();                                 // red_key was read for the maximum above
      if (lane == 0) red_key[warp] = kmin;
      __syncthreads();
#pragma unroll
      for (int w = 0; w < RS_WARPS; ++w) kmin = min(kmin, red_key[w]);
      kx = kmin;
    }
    // back to the bf16 value (negative values carry ones in the dropped low half of their key), then the same division
    // the keys went through
    const uint32_t kfull = (kx << 16) | ((kx & 0x8000u) ? 0u : 0xffffu);
    kth = f2key(__fdiv_rn(key2f(kfull), T));
    if (kth == 0u) kth = 1u;
  } else if (p.top_k < p.vocab) {
    uint32_t prefix = 0u;
    int above = p.vocab, buf = 0;                      // keys >= prefix
    for (int bit = 31; bit >= 0 && above != p.top_k; --bit) {
      const uint32_t cand = prefix | (1u << bit);
      int cnt = 0;
#pragma unroll
      for (int j = 0; j < NCH * 8; ++j) cnt += key[j] >= cand ? 1 : 0;
      cnt = __reduce_add_sync(0xffffffffu, cnt);
      if (lane == 0) red_cnt[buf][warp] = cnt;
      __syncthreads();
      int total = 0;
#pragma unroll
      for (int w = 0; w < RS_WARPS; ++w) total += red_cnt[buf][w];
      buf ^= 1;
      if (total >= p.top_k) { prefix = cand; above = total; }
    }
    if (above == p.top_k) {                            // exactly k keys at or above the prefix: K is their minimum
      uint32_t kmin = 0xffffffffu;
#pragma unroll
      for (int j = 0; j < NCH * 8; ++j) if (key[j] >= prefix) kmin = min(kmin, key[j]);
      kmin = __reduce_min_sync(0xffffffffu, kmin);
      __syncthreads();                                 // red_key was read for the maximum above
      if (lane == 0) red_key[warp] = kmin;
      __syncthreads();
#pragma unroll
      for (int w = 0; w < RS_WARPS; ++w) kmin = min(kmin, red_key[w]);
      kth = kmin;
    } else {
      kth = prefix;                                    // all 32 bits decided (ties straddle the k-th place)
    }
    if (kth == 0u) kth = 1u;
  }

  // ---- partition function over the survivors; their noise values are requested meanwhile     sampler.py:20-21
  const float* nrow = p.noise + (int64_t)row * p.vocab;
  float lsum = 0.f;
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    const int i0 = (c * RS_THREADS + tid) * 8;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const uint32_t k = key[c * 8 + e];
      if (k >= kth) {
        asm volatile("prefetch.global.L1 [%0];" :: "l"(nrow + i0 + e));
        lsum += expf(key2f(k) - lmax);
      }
    }
  }
  lsum = warp_sum(lsum);
  if (lane == 0) red_f[warp] = lsum;
  __syncthreads();
  float sum = 0.f;
#pragma unroll
  for (int w = 0; w < RS_WARPS; ++w) sum += red_f[w];  // fixed order: deterministic

  // ---- index = argmax(p / q)                     torch.multinomial(probs, 1), sampler.py:22
  float best = -INFINITY;
  int besti = 0x7fffffff;
  float* prow = p.probs_out ? p.probs_out + (int64_t)row * p.vocab : nullptr;
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    const int i0 = (c * RS_THREADS + tid) * 8;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const uint32_t k = key[c * 8 + e];
      float pr = 0.f;
      if (k >= kth) {
        pr = __fdiv_rn(expf(key2f(k) - lmax), sum);
        const float sc = __fdiv_rn(pr, __ldg(nrow + i0 + e));
        if (sc > best) { best = sc; besti = i0 + e; }      // strict: first index wins inside a thread (ascending)
      }
      if (prow && i0 < p.vocab) prow[i0 + e] = pr;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ob = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, besti, o);
    if (ob > best || (ob == best && oi < besti)) { best = ob; besti = oi; }
  }
  __syncthreads();                                     // red_f was read for the sum above
  if (lane == 0) { red_f[warp] = best; red_i[warp] = besti; }
  __syncthreads();
  if (tid == 0) {
    for (int w = 1; w < RS_WARPS; ++w)
      if (red_f[w] > best || (red_f[w] == best && red_i[w] < besti)) { best = red_f[w]; besti = red_i[w]; }
    if (besti == 0x7fffffff) besti = 0;
    p.out_index[row] = besti;
    bcast_i = besti;
  }
  if (p.embed_table) {
    __syncthreads();
    const __nv_bfloat16* src = p.embed_table + (int64_t)bcast_i * p.embed_cols;
    __nv_bfloat16* dst = p.embed_out + (int64_t)row * p.embed_out_stride;
    for (int v = tid; v < (p.embed_cols >> 3); v += RS_THREADS) stg_v4(dst + 8 * v, ldg_v4(src + 8 * v));
  }
}

template <int NCH>
static int launch_row_sampler(const RowSampleParams& p, int batch, bool fp32, cudaStream_t st) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(batch, 1, 1);
  cfg.blockDim = dim3(RS_THREADS, 1, 1);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e = fp32 ? cudaLaunchKernelEx(&cfg, topk_sample_row_kernel<NCH, true>, p)
                       : cudaLaunchKernelEx(&cfg, topk_sample_row_kernel<NCH, false>, p);
  count_launch();
  if (e != cudaSuccess) return cuda_fail(e, "topk_sample");
  return B2TTS_OK;
}

}  // namespace b2

namespace b2 {
static int topk_sample_impl(const void* logits, int64_t logits_row_stride, int logits_fp32,
                            const float* temperatures, const float* noise,
                            int64_t* out_index, float* probs_out,
                            int batch, int vocab, int top_k, float top_p, int apply_top_p,
                            const void* embed_table, int embed_cols, void* embed_out, int64_t embed_out_row_stride,
                            void* stream) {
  (void)top_p;
  B2_REQUIRE(batch >= 0 && vocab > 0 && top_k > 0, B2TTS_EINVAL, "topk_sample: batch=%d vocab=%d top_k=%d", batch, vocab, top_k);
  if (batch == 0) return B2TTS_OK;
  B2_REQUIRE(logits && temperatures && noise && out_index, B2TTS_EINVAL, "topk_sample: null pointer");
  B2_REQUIRE(vocab <= SMP_MAXV, B2TTS_EUNSUPPORTED, "topk_sample: vocab %d > %d", vocab, SMP_MAXV);
  B2_REQUIRE(!apply_top_p, B2TTS_EUNSUPPORTED,
             "topk_sample: top_p filtering is not implemented (the reference accepts and ignores top_p, sampler.py:15)");
  B2_REQUIRE(!embed_table || (embed_out && embed_cols > 0 && embed_cols % 8 == 0 && embed_out_row_stride % 8 == 0 &&
                              aligned16(embed_table) && aligned16(embed_out)),
             B2TTS_EALIGN, "topk_sample: embedding gather needs 16-byte aligned rows (cols %d)", embed_cols);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int esz = logits_fp32 ? 4 : 2;
  const bool warp_ok = vocab <= 4096 && vocab % 8 == 0 && aligned16(logits) && (logits_row_stride * esz) % 16 == 0;
  B2_REQUIRE(!embed_table || warp_ok, B2TTS_EUNSUPPORTED,
             "topk_sample: the fused embedding gather needs vocab <= 4096, a multiple of 8, 16-byte aligned rows");
  if (warp_ok) {
    RowSampleParams p{};
    p.logits = logits; p.row_stride = logits_row_stride; p.temperatures = temperatures; p.noise = noise;
    p.out_index = out_index; p.probs_out = probs_out; p.vocab = vocab; p.top_k = top_k;
    p.embed_table = static_cast<const __nv_bfloat16*>(embed_table);
    p.embed_out = static_cast<__nv_bfloat16*>(embed_out);
    p.embed_out_stride = embed_out_row_stride; p.embed_cols = embed_cols;
    if (vocab <= RS_THREADS * 8) return launch_row_sampler<1>(p, batch, logits_fp32 != 0, st);
    return launch_row_sampler<2>(p, batch, logits_fp32 != 0, st);
  }
  const size_t smem = (size_t)vocab * sizeof(float);
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
}  // namespace b2

extern "C" int b2tts_topk_sample(const void* logits, int64_t logits_row_stride, int logits_fp32,
                                 const float* temperatures, const float* noise,
                                 int64_t* out_index, float* probs_out,
                                 int batch, int vocab, int top_k, float top_p, int apply_top_p,
                                 void* stream) {
  return b2::topk_sample_impl(logits, logits_row_stride, logits_fp32, temperatures, noise, out_index, probs_out, batch,
                              vocab, top_k, top_p, apply_top_p, nullptr, 0, nullptr, 0, stream);
}

extern "C" int b2tts_topk_sample_embed(const void* logits, int64_t logits_row_stride, int logits_fp32,
                                       const float* temperatures, const float* noise,
                                       int64_t* out_index, int batch, int vocab, int top_k,
                                       const void* embed_table, int embed_cols, void* embed_out,
                                       int64_t embed_out_row_stride, void* stream) {
  B2_REQUIRE(embed_table && embed_out, B2TTS_EINVAL, "topk_sample_embed: null embedding table / output");
  return b2::topk_sample_impl(logits, logits_row_stride, logits_fp32, temperatures, noise, out_index, nullptr, batch,
                              vocab, top_k, 1.0f, 0, embed_table, embed_cols, embed_out, embed_out_row_stride, stream);
}

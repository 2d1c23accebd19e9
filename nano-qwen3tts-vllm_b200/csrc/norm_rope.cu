// norm_rope.cu - a4/a5 (+a1 fused): RMSNorm, add+RMSNorm, and the fused per-head q/k RMSNorm +
// RoPE + paged KV write that replaces q_norm, k_norm, rotary_emb and store_kvcache as called
// from Qwen3TTSAttention.forward (models/qwen3_tts_share.py:96-106).
//
// All of these are HBM/launch-latency bound element-wise passes: 16-byte (hidden rows) or
// 4-byte-per-lane fully coalesced accesses, fp32 math, rounding points chosen by `rounding`:
//   B2TTS_ROUND_EAGER  layernorm.py:22-27 evaluated op by op: bf16(x*rstd) then bf16(. * w),
//                      rstd = 1/sqrt (correctly rounded, what ATen's CPU rsqrt does);
//   B2TTS_ROUND_FUSED  what torch-inductor emits for the same source on the GPU: fp32 through
//                      to one rounding at the store, rstd = rsqrtf.
// RoPE is never compiled in the reference, so it always keeps the three bf16 roundings of
// `(x*cos) + (rotate_half(x)*sin)` (rotary_embedding.py:33-34) and bf16 cos/sin (:138-139).
#include <stdlib.h>

#include "common.cuh"

namespace b2 {

template <bool EAGER>
__device__ __forceinline__ float rstd_of(float sumsq, int cols, float eps) {
  const float var = sumsq / (float)cols + eps;
  return EAGER ? 1.0f / sqrtf(var) : rsqrtf(var);
}

template <bool EAGER>
__device__ __forceinline__ float norm_out(float x, float rstd, float w) {
  return EAGER ? rbf(rbf(x * rstd) * w) : (x * rstd) * w;   // caller rounds once more (no-op for EAGER)
}

__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = (blockDim.x + 31) >> 5;
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float t = lane < nw ? red[lane] : 0.f;
  t = warp_sum(t);
  return t;
}

// One CTA per row; every thread keeps up to MAXV 16-byte vectors of the row in registers.
constexpr int RMS_THREADS = 256;
constexpr int RMS_MAXV = 4;       // cols <= 8 * 256 * 4 = 8192

template <bool ADD, bool EAGER>
__global__ void __launch_bounds__(RMS_THREADS)
rmsnorm_kernel(const __nv_bfloat16* __restrict__ x, int64_t x_stride, const __nv_bfloat16* res,
               const __nv_bfloat16* __restrict__ w, __nv_bfloat16* y, __nv_bfloat16* res_out, int cols, float eps,
               int round_sum_first) {
  __shared__ float red[32];
  // programmatic dependent launch: between two GEMMs of a wide batch this kernel must neither drain the chain (the next
  // GEMM's CTAs become resident and stream their first weight tiles while it runs) nor start late itself
  pdl_launch();
  pdl_wait();
  const int row = blockIdx.x;
  const int nvec = cols >> 3;
  const __nv_bfloat16* xr = x + (int64_t)row * x_stride;
  float f[RMS_MAXV][8];
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < RMS_MAXV; ++i) {
    const int v = threadIdx.x + i * RMS_THREADS;
    if (v < nvec) {
      uint4 r = ldg_v4(xr + 8 * v);
      f[i][0] = bf16lo(r.x); f[i][1] = bf16hi(r.x); f[i][2] = bf16lo(r.y); f[i][3] = bf16hi(r.y);
      f[i][4] = bf16lo(r.z); f[i][5] = bf16hi(r.z); f[i][6] = bf16lo(r.w); f[i][7] = bf16hi(r.w);
      if (ADD) {
        uint4 q = ldg_v4(res + (int64_t)row * cols + 8 * v);
        f[i][0] += bf16lo(q.x); f[i][1] += bf16hi(q.x); f[i][2] += bf16lo(q.y); f[i][3] += bf16hi(q.y);
        f[i][4] += bf16lo(q.z); f[i][5] += bf16hi(q.z); f[i][6] += bf16lo(q.w); f[i][7] += bf16hi(q.w);
        uint4 o;
        o.x = pack_bf16(f[i][0], f[i][1]); o.y = pack_bf16(f[i][2], f[i][3]);
        o.z = pack_bf16(f[i][4], f[i][5]); o.w = pack_bf16(f[i][6], f[i][7]);
        stg_v4(res_out + (int64_t)row * cols + 8 * v, o);
        if (round_sum_first) {
          f[i][0] = bf16lo(o.x); f[i][1] = bf16hi(o.x); f[i][2] = bf16lo(o.y); f[i][3] = bf16hi(o.y);
          f[i][4] = bf16lo(o.z); f[i][5] = bf16hi(o.z); f[i][6] = bf16lo(o.w); f[i][7] = bf16hi(o.w);
        }
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) ss = fmaf(f[i][e], f[i][e], ss);
    }
  }
  const float rstd = rstd_of<EAGER>(block_sum(ss, red), cols, eps);
#pragma unroll
  for (int i = 0; i < RMS_MAXV; ++i) {
    const int v = threadIdx.x + i * RMS_THREADS;
    if (v < nvec) {
      uint4 wr = ldg_v4(w + 8 * v);
      float o[8];
      o[0] = norm_out<EAGER>(f[i][0], rstd, bf16lo(wr.x)); o[1] = norm_out<EAGER>(f[i][1], rstd, bf16hi(wr.x));
      o[2] = norm_out<EAGER>(f[i][2], rstd, bf16lo(wr.y)); o[3] = norm_out<EAGER>(f[i][3], rstd, bf16hi(wr.y));
      o[4] = norm_out<EAGER>(f[i][4], rstd, bf16lo(wr.z)); o[5] = norm_out<EAGER>(f[i][5], rstd, bf16hi(wr.z));
      o[6] = norm_out<EAGER>(f[i][6], rstd, bf16lo(wr.w)); o[7] = norm_out<EAGER>(f[i][7], rstd, bf16hi(wr.w));
      uint4 ov;
      ov.x = pack_bf16(o[0], o[1]); ov.y = pack_bf16(o[2], o[3]);
      ov.z = pack_bf16(o[4], o[5]); ov.w = pack_bf16(o[6], o[7]);
      stg_v4(y + (int64_t)row * cols + 8 * v, ov);
    }
  }
}

// Narrow rows (cols <= 256, e.g. the per-head norm over 128): one warp per row, 8 rows per CTA.
template <bool EAGER>
__global__ void __launch_bounds__(256)
rmsnorm_narrow_kernel(const __nv_bfloat16* __restrict__ x, int64_t x_stride, const __nv_bfloat16* __restrict__ w,
                      __nv_bfloat16* y, int rows, int cols, float eps) {
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  const int nvec = cols >> 3;                       // <= 32
  const __nv_bfloat16* xr = x + (int64_t)row * x_stride;
  float f[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  float ss = 0.f;
  if (lane < nvec) {
    uint4 r = ldg_v4(xr + 8 * lane);
    f[0] = bf16lo(r.x); f[1] = bf16hi(r.x); f[2] = bf16lo(r.y); f[3] = bf16hi(r.y);
    f[4] = bf16lo(r.z); f[5] = bf16hi(r.z); f[6] = bf16lo(r.w); f[7] = bf16hi(r.w);
#pragma unroll
    for (int e = 0; e < 8; ++e) ss = fmaf(f[e], f[e], ss);
  }
  const float rstd = rstd_of<EAGER>(warp_sum(ss), cols, eps);
  if (lane < nvec) {
    uint4 wr = ldg_v4(w + 8 * lane);
    uint4 ov;
    ov.x = pack_bf16(norm_out<EAGER>(f[0], rstd, bf16lo(wr.x)), norm_out<EAGER>(f[1], rstd, bf16hi(wr.x)));
    ov.y = pack_bf16(norm_out<EAGER>(f[2], rstd, bf16lo(wr.y)), norm_out<EAGER>(f[3], rstd, bf16hi(wr.y)));
    ov.z = pack_bf16(norm_out<EAGER>(f[4], rstd, bf16lo(wr.z)), norm_out<EAGER>(f[5], rstd, bf16hi(wr.z)));
    ov.w = pack_bf16(norm_out<EAGER>(f[6], rstd, bf16lo(wr.w)), norm_out<EAGER>(f[7], rstd, bf16hi(wr.w)));
    stg_v4(y + (int64_t)row * cols + 8 * lane, ov);
  }
}

// ------------------------------------------------------------------------------------------
// Fused q/k head-norm + RoPE + KV write.  One CTA per token, 8 warps; a warp owns one head at a
// time (lane l holds dims {2l, 2l+1, 64+2l, 64+2l+1} -> two fully coalesced 128-byte accesses),
// computes cos/sin for its two frequencies once and walks heads warp, warp+8, ...
struct QkRopeParams {
  const __nv_bfloat16 *q, *k, *v;
  int64_t q_stride, k_stride, v_stride;
  __nv_bfloat16 *q_out, *k_out;
  const __nv_bfloat16 *q_w, *k_w;
  float eps;
  const int64_t* positions;
  int64_t pos_row_stride;
  const float* inv_freq;
  const int32_t* modality;
  float attention_scaling;
  const int32_t* slot_mapping;
  __nv_bfloat16 *k_cache, *v_cache;
  int num_heads, num_kv_heads;
};

template <bool EAGER>
__global__ void __launch_bounds__(256)
qknorm_rope_kvwrite_kernel(const QkRopeParams p) {
  constexpr int HD = 128;
  const int tok = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kv_dim = p.num_kv_heads * HD;
  const int slot = p.slot_mapping ? __ldg(p.slot_mapping + tok) : -1;

  // cos/sin of this lane's columns {2l, 2l+1} (c1/s1) and {64+2l, 64+2l+1} (c2/s2): fp32 angle,
  // fp32 cos/sin, bf16 result.  Columns f and 64+f share frequency f but - in the sectioned M-RoPE
  // branch (rotary_embedding.py:73-80) - not necessarily the position row, so the two halves keep
  // their own cos/sin; with 1-D positions (modality NULL) they are computed once.
  float c1[2], s1[2], c2[2], s2[2];
#pragma unroll
  for (int e = 0; e < 2; ++e) {
    const int fi = 2 * lane + e;
    const float f = __ldg(p.inv_freq + fi);
    const int m1 = p.modality ? __ldg(p.modality + fi) : 0;
    const float a1 = (float)__ldg(p.positions + (int64_t)m1 * p.pos_row_stride + tok) * f;
    c1[e] = rbf(cosf(a1) * p.attention_scaling);
    s1[e] = rbf(sinf(a1) * p.attention_scaling);
    const int m2 = p.modality ? __ldg(p.modality + HD / 2 + fi) : 0;
    if (m2 != m1) {
      const float a2 = (float)__ldg(p.positions + (int64_t)m2 * p.pos_row_stride + tok) * f;
      c2[e] = rbf(cosf(a2) * p.attention_scaling);
      s2[e] = rbf(sinf(a2) * p.attention_scaling);
    } else {
      c2[e] = c1[e]; s2[e] = s1[e];
    }
  }

  const int total = p.num_heads + p.num_kv_heads;
  for (int hh = warp; hh < total; hh += 8) {
    const bool is_k = hh >= p.num_heads;
    const int h = is_k ? hh - p.num_heads : hh;
    const __nv_bfloat16* src = is_k ? p.k + (int64_t)tok * p.k_stride + h * HD : p.q + (int64_t)tok * p.q_stride + h * HD;
    const __nv_bfloat16* w = is_k ? p.k_w : p.q_w;
    const uint32_t r1 = *reinterpret_cast<const uint32_t*>(src + 2 * lane);
    const uint32_t r2 = *reinterpret_cast<const uint32_t*>(src + 64 + 2 * lane);
    float x1[2] = {bf16lo(r1), bf16hi(r1)};
    float x2[2] = {bf16lo(r2), bf16hi(r2)};
    if (w) {
      float ss = x1[0] * x1[0];
      ss = fmaf(x1[1], x1[1], ss); ss = fmaf(x2[0], x2[0], ss); ss = fmaf(x2[1], x2[1], ss);
      const float rstd = rstd_of<EAGER>(warp_sum(ss), HD, p.eps);
      const uint32_t w1 = *reinterpret_cast<const uint32_t*>(w + 2 * lane);
      const uint32_t w2 = *reinterpret_cast<const uint32_t*>(w + 64 + 2 * lane);
      x1[0] = rbf(norm_out<EAGER>(x1[0], rstd, bf16lo(w1))); x1[1] = rbf(norm_out<EAGER>(x1[1], rstd, bf16hi(w1)));
      x2[0] = rbf(norm_out<EAGER>(x2[0], rstd, bf16lo(w2))); x2[1] = rbf(norm_out<EAGER>(x2[1], rstd, bf16hi(w2)));
    }
    float o1[2], o2[2];
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      o1[e] = rbf(x1[e] * c1[e]) + rbf(-x2[e] * s1[e]);   // rotate_half: first half pairs with -x2
      o2[e] = rbf(x2[e] * c2[e]) + rbf(x1[e] * s2[e]);
    }
    const uint32_t y1 = pack_bf16(o1[0], o1[1]), y2 = pack_bf16(o2[0], o2[1]);
    __nv_bfloat16* dst = nullptr;
    if (is_k) {
      if (p.k_out) dst = p.k_out + (int64_t)tok * kv_dim + h * HD;
      if (slot >= 0) {
        __nv_bfloat16* cd = p.k_cache + (int64_t)slot * kv_dim + h * HD;
        *reinterpret_cast<uint32_t*>(cd + 2 * lane) = y1;
        *reinterpret_cast<uint32_t*>(cd + 64 + 2 * lane) = y2;
      }
    } else {
      dst = p.q_out + ((int64_t)tok * p.num_heads + h) * HD;
    }
    if (dst) {
      *reinterpret_cast<uint32_t*>(dst + 2 * lane) = y1;
      *reinterpret_cast<uint32_t*>(dst + 64 + 2 * lane) = y2;
    }
  }
  if (slot >= 0 && p.v) {                                  // V goes to the cache untouched
    const __nv_bfloat16* vs = p.v + (int64_t)tok * p.v_stride;
    __nv_bfloat16* vd = p.v_cache + (int64_t)slot * kv_dim;
    for (int i = threadIdx.x; i < (kv_dim >> 3); i += blockDim.x) stg_v4(vd + 8 * i, ldg_nc_v4(vs + 8 * i));
  }
}

}  // namespace b2

extern "C" int b2tts_rmsnorm(const void* x, int64_t x_row_stride, const void* w, void* y,
                             int rows, int cols, float eps, int rounding, void* stream) {
  using namespace b2;
  B2_REQUIRE(rows >= 0 && cols > 0, B2TTS_EINVAL, "rmsnorm: rows=%d cols=%d", rows, cols);
  if (rows == 0) return B2TTS_OK;
  B2_REQUIRE(x && w && y, B2TTS_EINVAL, "rmsnorm: null pointer");
  B2_REQUIRE(cols % 8 == 0 && x_row_stride % 8 == 0 && aligned16(x) && aligned16(w) && aligned16(y), B2TTS_EALIGN,
             "rmsnorm: cols/stride must be multiples of 8 and pointers 16-byte aligned");
  B2_REQUIRE(cols <= 8 * RMS_THREADS * RMS_MAXV, B2TTS_EUNSUPPORTED, "rmsnorm: cols %d > %d", cols,
             8 * RMS_THREADS * RMS_MAXV);
  B2_REQUIRE(rounding == B2TTS_ROUND_EAGER || rounding == B2TTS_ROUND_FUSED, B2TTS_EINVAL, "rmsnorm: rounding %d", rounding);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  auto xp = static_cast<const __nv_bfloat16*>(x);
  auto wp = static_cast<const __nv_bfloat16*>(w);
  auto yp = static_cast<__nv_bfloat16*>(y);
  if (cols <= 256) {
    const int grid = (rows + 7) / 8;
    if (rounding == B2TTS_ROUND_EAGER) rmsnorm_narrow_kernel<true><<<grid, 256, 0, st>>>(xp, x_row_stride, wp, yp, rows, cols, eps);
    else rmsnorm_narrow_kernel<false><<<grid, 256, 0, st>>>(xp, x_row_stride, wp, yp, rows, cols, eps);
  } else {
    // launched as a programmatic dependent (B2TTS_PDL_NORM=0: a plain launch); correct after any kernel - one that never
    // executes griddepcontrol.launch_dependents releases its dependents when it completes
    static const bool pdl = [] { const char* e = getenv("B2TTS_PDL_NORM"); return !(e && e[0] == '0'); }();
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(rows, 1, 1);
    cfg.blockDim = dim3(RMS_THREADS, 1, 1);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    const __nv_bfloat16* nores = nullptr;
    __nv_bfloat16* noout = nullptr;
    cudaError_t e = rounding == B2TTS_ROUND_EAGER
        ? cudaLaunchKernelEx(&cfg, rmsnorm_kernel<false, true>, xp, x_row_stride, nores, wp, yp, noout, cols, eps, 0)
        : cudaLaunchKernelEx(&cfg, rmsnorm_kernel<false, false>, xp, x_row_stride, nores, wp, yp, noout, cols, eps, 0);
    if (e != cudaSuccess) return cuda_fail(e, "rmsnorm");
  }
  B2_CHECK_LAUNCH("rmsnorm");
  return B2TTS_OK;
}

extern "C" int b2tts_add_rmsnorm(const void* x, const void* res, const void* w, void* y, void* res_out,
                                 int rows, int cols, float eps, int rounding, int round_sum_first,
                                 void* stream) {
  using namespace b2;
  B2_REQUIRE(rows >= 0 && cols > 0, B2TTS_EINVAL, "add_rmsnorm: rows=%d cols=%d", rows, cols);
  if (rows == 0) return B2TTS_OK;
  B2_REQUIRE(x && res && w && y && res_out, B2TTS_EINVAL, "add_rmsnorm: null pointer");
  B2_REQUIRE(cols % 8 == 0 && aligned16(x) && aligned16(res) && aligned16(w) && aligned16(y) && aligned16(res_out),
             B2TTS_EALIGN, "add_rmsnorm: cols must be a multiple of 8 and pointers 16-byte aligned");
  B2_REQUIRE(cols <= 8 * RMS_THREADS * RMS_MAXV, B2TTS_EUNSUPPORTED, "add_rmsnorm: cols %d > %d", cols,
             8 * RMS_THREADS * RMS_MAXV);
  B2_REQUIRE(rounding == B2TTS_ROUND_EAGER || rounding == B2TTS_ROUND_FUSED, B2TTS_EINVAL, "add_rmsnorm: rounding %d", rounding);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  auto xp = static_cast<const __nv_bfloat16*>(x);
  auto rp = static_cast<const __nv_bfloat16*>(res);
  auto wp = static_cast<const __nv_bfloat16*>(w);
  auto yp = static_cast<__nv_bfloat16*>(y);
  auto op = static_cast<__nv_bfloat16*>(res_out);
  if (rounding == B2TTS_ROUND_EAGER)
    rmsnorm_kernel<true, true><<<rows, RMS_THREADS, 0, st>>>(xp, cols, rp, wp, yp, op, cols, eps, round_sum_first);
  else
    rmsnorm_kernel<true, false><<<rows, RMS_THREADS, 0, st>>>(xp, cols, rp, wp, yp, op, cols, eps, round_sum_first);
  B2_CHECK_LAUNCH("add_rmsnorm");
  return B2TTS_OK;
}

extern "C" int b2tts_qknorm_rope_kvwrite(const void* q, int64_t q_row_stride,
                                         const void* k, int64_t k_row_stride,
                                         const void* v, int64_t v_row_stride,
                                         void* q_out, void* k_out,
                                         const void* q_weight, const void* k_weight, float eps,
                                         const int64_t* positions, int64_t pos_row_stride,
                                         const float* inv_freq, const int32_t* modality,
                                         float attention_scaling,
                                         const int32_t* slot_mapping, void* k_cache, void* v_cache,
                                         int n_tokens, int num_heads, int num_kv_heads, int head_dim,
                                         int rounding, void* stream) {
  using namespace b2;
  const char* what = "qknorm_rope_kvwrite";
  B2_REQUIRE(n_tokens >= 0 && num_heads > 0 && num_kv_heads > 0, B2TTS_EINVAL, "%s: n=%d heads=%d kv=%d", what,
             n_tokens, num_heads, num_kv_heads);
  if (n_tokens == 0) return B2TTS_OK;
  B2_REQUIRE(head_dim == 128, B2TTS_EUNSUPPORTED, "%s: head_dim %d unsupported (128)", what, head_dim);
  B2_REQUIRE(q && k && q_out && positions && inv_freq, B2TTS_EINVAL, "%s: null pointer", what);
  B2_REQUIRE((q_weight == nullptr) == (k_weight == nullptr), B2TTS_EINVAL, "%s: q/k weights must both be set or both NULL", what);
  if (slot_mapping) B2_REQUIRE(k_cache && v_cache && v, B2TTS_EINVAL, "%s: slot_mapping needs v, k_cache, v_cache", what);
  B2_REQUIRE(q_row_stride % 8 == 0 && k_row_stride % 8 == 0 && v_row_stride % 8 == 0 && aligned16(q) && aligned16(k) &&
                 aligned16(v) && aligned16(q_out) && aligned16(k_out) && aligned16(k_cache) && aligned16(v_cache),
             B2TTS_EALIGN, "%s: strides must be multiples of 8 and pointers 16-byte aligned", what);
  B2_REQUIRE(rounding == B2TTS_ROUND_EAGER || rounding == B2TTS_ROUND_FUSED, B2TTS_EINVAL, "%s: rounding %d", what, rounding);
  QkRopeParams p{};
  p.q = static_cast<const __nv_bfloat16*>(q); p.k = static_cast<const __nv_bfloat16*>(k);
  p.v = static_cast<const __nv_bfloat16*>(v);
  p.q_stride = q_row_stride; p.k_stride = k_row_stride; p.v_stride = v_row_stride;
  p.q_out = static_cast<__nv_bfloat16*>(q_out); p.k_out = static_cast<__nv_bfloat16*>(k_out);
  p.q_w = static_cast<const __nv_bfloat16*>(q_weight); p.k_w = static_cast<const __nv_bfloat16*>(k_weight);
  p.eps = eps; p.positions = positions; p.pos_row_stride = pos_row_stride;
  p.inv_freq = inv_freq; p.modality = modality; p.attention_scaling = attention_scaling;
  p.slot_mapping = slot_mapping;
  p.k_cache = static_cast<__nv_bfloat16*>(k_cache); p.v_cache = static_cast<__nv_bfloat16*>(v_cache);
  p.num_heads = num_heads; p.num_kv_heads = num_kv_heads;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (rounding == B2TTS_ROUND_EAGER) qknorm_rope_kvwrite_kernel<true><<<n_tokens, 256, 0, st>>>(p);
  else qknorm_rope_kvwrite_kernel<false><<<n_tokens, 256, 0, st>>>(p);
  B2_CHECK_LAUNCH(what);
  return B2TTS_OK;
}

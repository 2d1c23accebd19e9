// decoder_stack.cu - a10 in ONE launch: the whole decode pass of a small decoder stack (the code predictor's
// 5 layers, models/qwen3_tts_predictor.py:30-44 over models/qwen3_tts_share.py:79-176 of the reference) as a
// persistent DATAFLOW kernel.
//
// Why: one predictor pass is 5 x (qkv GEMM, attention, o GEMM, gate_up GEMM, down GEMM) = 25 dependent phases over
// 157 MB of weights; as separate launches every phase costs 8-12 us of launch boundary, barrier/TMEM set-up,
// operand fetch and split-K exchange for < 2 us of weight streaming, 15 times per codec frame.  The stack is small
// enough that ONE (projection, 128-feature tile, K split) job per layer fits on its own SM: 32 qkv tiles + 8 x 4 o
// splits + 48 gate_up tiles + 8 x 4 down splits = 144 <= 148 SMs.  So the assignment is STATIC:
//   * every CTA owns the same job in every layer; its producer warp streams the job's packed weight tiles
//     (b2tts_gemm_pack_weight images, 16 KiB 1-D bulk copies) into a 9-stage ring and runs AHEAD: while layer l
//     waits for its activations, layer l's (and then l+1's) weights are already in shared memory, so HBM never
//     waits for a dependency and no phase pays a weight-fetch latency;
//   * phases are linked by monotonic arrival counters in L2 (red.release / ld.acquire, epoch-scaled so that a CUDA
//     graph can replay the launch without a memset) instead of kernel boundaries; the activations (<= 64 rows) are
//     fetched by 2-D tensor-map copies as the UMMA B operand once their counter is complete;
//   * tcgen05.mma (weights = 128-row A operand, batch = N, fp32 accumulator in TMEM) exactly as gemm_stream_kernel;
//   * the o / down splits reduce through an L2 scratch in fixed split order (deterministic), apply the residual
//     add, and the same CTAs then normalise the finished rows (RMSNorm, arithmetic of rmsnorm_kernel) for the next
//     projection;
//   * attention (q/k head norm + RoPE + paged KV write + GQA attention over the cached tokens, arithmetic and
//     summation order of fused_decode_attn_kernel) runs on the 4 worker warps of ALL CTAs, one (sequence, kv head)
//     per warp, between the qkv and o phases.
// Rounding points are those of the separate launches; with the same K splits the result is
// bit-identical to rmsnorm -> gemm -> fused_decode_attn -> gemm(residual) -> rmsnorm -> gemm(SiLU*mul) -> gemm(residual).
//
// Every wait is bounded: a dependency that does not resolve within DS_TIMEOUT_NS traps (launch failure) instead of
// hanging the device.
#include <stdlib.h>

#include "umma.cuh"

namespace b2 {

constexpr int DS_MAX_LAYERS = 8;
constexpr int DS_THREADS = 224;                  // warp 0: W producer, 1: MMA issuer, 2: x producer, 3-6: workers
constexpr int DS_W0 = 96;                        // first worker thread
constexpr int DS_WORKERS = 128;
constexpr int DS_MAX_M = 64;                     // batch rows (UMMA N)
constexpr int DS_STAGES = 9;
constexpr int DS_STAGE_BYTES = G_WTILE_BYTES + DS_MAX_M * 128;   // 24 KiB: W tile + x tile
constexpr int DS_HD = 128;
constexpr int DS_GROUP = 2;                      // q heads per kv head
constexpr int DS_ATT_TILE = 8;                   // tokens per online-softmax step (as fused_decode_attn_kernel)
constexpr int DS_SCR_BYTES = 1024;               // per worker warp: q [2][128], k [128], v [128] bf16
constexpr int DS_MAX_SPLIT_TILES = 16;
constexpr int DS_CTRS = 8 + 2 * DS_MAX_SPLIT_TILES;   // counters per layer
constexpr int DS_CTR_STRIDE = 8;                 // words between counters (one 32-byte sector each)
constexpr int DS_CTR_BASE = 64;                  // words 0..63: epoch, exit counter, error
constexpr size_t DS_SYNC_BYTES = 32768;
constexpr unsigned long long DS_TIMEOUT_NS = 4000000000ull;

enum { DS_XN_A = 0, DS_QKV = 1, DS_ATT = 2, DS_O = 3, DS_XN_B = 4, DS_ACT = 5, DS_DOWN = 6, DS_T_O = 8,
       DS_T_D = 8 + DS_MAX_SPLIT_TILES };
enum { OP_QKV = 0, OP_O = 1, OP_GU = 2, OP_DOWN = 3, OP_NONE = 4 };

struct DsLayer {
  const uint8_t *w_qkv, *w_o, *w_gu, *w_down;            // packed weights
  const __nv_bfloat16 *ln_in, *ln_post, *qn_w, *kn_w;    // norm weights
  __nv_bfloat16 *k_cache, *v_cache;                      // [pages, page, kvh, 128]
};

struct DsParams {
  DsLayer layer[DS_MAX_LAYERS];
  int num_layers;
  const __nv_bfloat16* x_in;
  int64_t x_stride;
  __nv_bfloat16* out;
  int64_t out_stride;
  const __nv_bfloat16* final_w;
  int m, mpad, tmem_cols, rounding;
  float eps;
  int H, Q, KV, I, heads, kvh;
  int nq_tiles, no_tiles, so, ng_tiles, nd_tiles, sd;
  const int64_t* positions;
  const float* inv_freq;
  float attention_scaling, scale_log2;
  const int32_t* slot_mapping;
  const int32_t* block_table;
  int bt_stride, page_size;
  const int32_t* context_lens;
  unsigned* sync;
  __nv_bfloat16 *xn, *resid, *qkv, *attn, *act;
  float *part_o, *part_d;
  unsigned long long* trace;                     // debug: 16 globaltimer stamps per (CTA, layer) (null = off)
};
#define DS_TRACE(l, slot) do { if (p.trace) p.trace[((int64_t)cta * DS_MAX_LAYERS + (l)) * 16 + (slot)] = gtimer(); } while (0)

// ---------------------------------------------------------------- bounded waits
__device__ __forceinline__ unsigned ds_ld_acquire(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void ds_arrive(unsigned* ctr) {
  asm volatile("red.release.gpu.global.add.u32 [%0], 1;" :: "l"(ctr) : "memory");
}
__device__ __noinline__ void ds_fail(unsigned* sync, unsigned code) {
  atomicExch(sync + 2, code);
  __threadfence_system();
  __trap();
}
// one thread: spin until the counter has reached `target` (wrap-safe)
__device__ __forceinline__ void ds_wait(unsigned* sync, const unsigned* ctr, unsigned target, unsigned code) {
  unsigned long long t0 = 0;
  unsigned n = 0;
  while ((int)(ds_ld_acquire(ctr) - target) < 0) {
    if ((++n & 255u) == 0) {
      const unsigned long long t = gtimer();
      if (t0 == 0) t0 = t;
      else if (t - t0 > DS_TIMEOUT_NS) ds_fail(sync, code);
    }
  }
}
__device__ __forceinline__ bool ds_mbar_try(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void ds_mbar_wait(unsigned* sync, uint64_t* bar, uint32_t parity, unsigned code) {
  unsigned long long t0 = 0;
  unsigned n = 0;
  while (!ds_mbar_try(bar, parity)) {
    if ((++n & 63u) == 0) {
      const unsigned long long t = gtimer();
      if (t0 == 0) t0 = t;
      else if (t - t0 > DS_TIMEOUT_NS) ds_fail(sync, code);
    }
  }
}
__device__ __forceinline__ void ds_fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }
__device__ __forceinline__ void ds_workers_bar() { asm volatile("bar.sync 1, 128;" ::: "memory"); }
__device__ __forceinline__ unsigned* ds_ctr(const DsParams& p, int l, int k) {
  return p.sync + DS_CTR_BASE + (l * DS_CTRS + k) * DS_CTR_STRIDE;
}
__device__ __forceinline__ float ds_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void ds_unpack8(const uint4& r, float* f) {
  f[0] = bf16lo(r.x); f[1] = bf16hi(r.x); f[2] = bf16lo(r.y); f[3] = bf16hi(r.y);
  f[4] = bf16lo(r.z); f[5] = bf16hi(r.z); f[6] = bf16lo(r.w); f[7] = bf16hi(r.w);
}
__device__ __forceinline__ uint4 ds_pack8(const float* f) {
  uint4 r;
  r.x = pack_bf16(f[0], f[1]); r.y = pack_bf16(f[2], f[3]); r.z = pack_bf16(f[4], f[5]); r.w = pack_bf16(f[6], f[7]);
  return r;
}
__device__ __forceinline__ uint4 ds_ldcg_v4(const void* p) { return __ldcg(reinterpret_cast<const uint4*>(p)); }
__device__ __forceinline__ uint32_t ds_ldcg_u32(const void* p) { return __ldcg(reinterpret_cast<const unsigned*>(p)); }

// ---------------------------------------------------------------- one attention item: (row, kv head) on one warp.
// Arithmetic, lane mapping and summation order of fused_decode_attn_kernel (attention_fused.cu) with nsplit == 1;
// the cached tokens are read straight from the paged cache (L2), the new token from the warp's scratch.
template <bool EAGER>
__device__ __forceinline__ void ds_attn_item(const DsParams& p, const DsLayer& ly, int row, int hk, int lane,
                                             __nv_bfloat16* scr, const float* frq) {
  const int kv_dim = p.kvh * DS_HD;
  const int qkv_w = p.Q + 2 * p.KV;
  const int head0 = hk * DS_GROUP;
  const int g = lane >> 3, j = lane & 7;
  int L = __ldg(p.context_lens + row);
  if (L <= 0) {                                   // CUDA-graph padding row: zeros
    if (g == 0) {
#pragma unroll
      for (int h = 0; h < DS_GROUP; ++h) {
        __nv_bfloat16* o = p.attn + (int64_t)row * p.Q + (head0 + h) * DS_HD;
        stg_v4(o + 8 * j, make_uint4(0, 0, 0, 0));
        stg_v4(o + 64 + 8 * j, make_uint4(0, 0, 0, 0));
      }
    }
    return;
  }
  const int slot = __ldg(p.slot_mapping + row);
  const int64_t pos = __ldg(p.positions + row);
  const __nv_bfloat16* qkv_row = p.qkv + (int64_t)row * qkv_w;
  uint32_t raw1[DS_GROUP + 1], raw2[DS_GROUP + 1];
#pragma unroll
  for (int i = 0; i < DS_GROUP + 1; ++i) {
    const __nv_bfloat16* src = i == DS_GROUP ? qkv_row + p.Q + hk * DS_HD : qkv_row + (head0 + i) * DS_HD;
    raw1[i] = ds_ldcg_u32(src + 2 * lane);
    raw2[i] = ds_ldcg_u32(src + 64 + 2 * lane);
  }
  uint4 vraw = make_uint4(0, 0, 0, 0);
  if (lane < 16) vraw = ds_ldcg_v4(qkv_row + p.Q + p.KV + hk * DS_HD + 8 * lane);
  const uint32_t qw1 = *reinterpret_cast<const uint32_t*>(ly.qn_w + 2 * lane);
  const uint32_t qw2 = *reinterpret_cast<const uint32_t*>(ly.qn_w + 64 + 2 * lane);
  const uint32_t kw1 = *reinterpret_cast<const uint32_t*>(ly.kn_w + 2 * lane);
  const uint32_t kw2 = *reinterpret_cast<const uint32_t*>(ly.kn_w + 64 + 2 * lane);

  __nv_bfloat16* q_s = scr;
  __nv_bfloat16* kn_s = scr + DS_GROUP * DS_HD;
  __nv_bfloat16* vn_s = kn_s + DS_HD;
  __syncwarp();                                   // the previous item's readers are done with the scratch
  {
    float c1[2], s1[2];
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const float a1 = (float)pos * frq[e];
      c1[e] = rbf(cosf(a1) * p.attention_scaling);
      s1[e] = rbf(sinf(a1) * p.attention_scaling);
    }
#pragma unroll
    for (int i = 0; i < DS_GROUP + 1; ++i) {
      const bool is_k = i == DS_GROUP;
      float x1[2] = {bf16lo(raw1[i]), bf16hi(raw1[i])};
      float x2[2] = {bf16lo(raw2[i]), bf16hi(raw2[i])};
      {
        float ss = x1[0] * x1[0];
        ss = fmaf(x1[1], x1[1], ss); ss = fmaf(x2[0], x2[0], ss); ss = fmaf(x2[1], x2[1], ss);
        const float var = warp_sum(ss) / (float)DS_HD + p.eps;
        const float rstd = EAGER ? 1.0f / sqrtf(var) : rsqrtf(var);
        const uint32_t w1 = is_k ? kw1 : qw1, w2 = is_k ? kw2 : qw2;
        const float wf[4] = {bf16lo(w1), bf16hi(w1), bf16lo(w2), bf16hi(w2)};
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          x1[e] = rbf(EAGER ? rbf(x1[e] * rstd) * wf[e] : (x1[e] * rstd) * wf[e]);
          x2[e] = rbf(EAGER ? rbf(x2[e] * rstd) * wf[2 + e] : (x2[e] * rstd) * wf[2 + e]);
        }
      }
      float o1[2], o2[2];
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        o1[e] = rbf(x1[e] * c1[e]) + rbf(-x2[e] * s1[e]);
        o2[e] = rbf(x2[e] * c1[e]) + rbf(x1[e] * s1[e]);
      }
      const uint32_t y1 = pack_bf16(o1[0], o1[1]), y2 = pack_bf16(o2[0], o2[1]);
      __nv_bfloat16* dst = is_k ? kn_s : q_s + i * DS_HD;
      *reinterpret_cast<uint32_t*>(dst + 2 * lane) = y1;
      *reinterpret_cast<uint32_t*>(dst + 64 + 2 * lane) = y2;
      if (is_k && slot >= 0) {
        __nv_bfloat16* cd = ly.k_cache + (int64_t)slot * kv_dim + hk * DS_HD;
        *reinterpret_cast<uint32_t*>(cd + 2 * lane) = y1;
        *reinterpret_cast<uint32_t*>(cd + 64 + 2 * lane) = y2;
      }
    }
    if (lane < 16) {
      *reinterpret_cast<uint4*>(vn_s + 8 * lane) = vraw;
      if (slot >= 0) stg_v4(ly.v_cache + (int64_t)slot * kv_dim + hk * DS_HD + 8 * lane, vraw);
    }
  }
  __syncwarp();

  float qv[DS_GROUP][16];
#pragma unroll
  for (int h = 0; h < DS_GROUP; ++h) {
    const __nv_bfloat16* qp = q_s + h * DS_HD;
    ds_unpack8(*reinterpret_cast<const uint4*>(qp + 8 * j), qv[h]);
    ds_unpack8(*reinterpret_cast<const uint4*>(qp + 64 + 8 * j), qv[h] + 8);
#pragma unroll
    for (int d = 0; d < 16; ++d) qv[h][d] *= p.scale_log2;
  }
  float mx[DS_GROUP], l[DS_GROUP], acc[DS_GROUP][16];
#pragma unroll
  for (int h = 0; h < DS_GROUP; ++h) {
    mx[h] = -INFINITY; l[h] = 0.f;
#pragma unroll
    for (int d = 0; d < 16; ++d) acc[h][d] = 0.f;
  }
  const int n_tiles = (L + DS_ATT_TILE - 1) / DS_ATT_TILE;
  const int32_t* bt = p.block_table + (int64_t)row * p.bt_stride;
  for (int i = 0; i < n_tiles; ++i) {
    const int tok0 = i * DS_ATT_TILE;
    const int ntok = min(DS_ATT_TILE, L - tok0);
    // both token slots of this lane: K and V rows requested together (one L2 round trip per tile)
    uint4 kraw[2][2], vr[2][2];
    bool valid[2];
#pragma unroll
    for (int it = 0; it < DS_ATT_TILE / 4; ++it) {
      const int tok = it * 4 + g;
      valid[it] = tok < ntok;
      const int t = tok0 + (valid[it] ? tok : 0);
      if (t == L - 1) {
        kraw[it][0] = *reinterpret_cast<const uint4*>(kn_s + 8 * j);
        kraw[it][1] = *reinterpret_cast<const uint4*>(kn_s + 64 + 8 * j);
        vr[it][0] = *reinterpret_cast<const uint4*>(vn_s + 8 * j);
        vr[it][1] = *reinterpret_cast<const uint4*>(vn_s + 64 + 8 * j);
      } else {
        int page = __ldg(bt + t / p.page_size);
        if (page < 0) page = 0;
        const int64_t off = ((int64_t)page * p.page_size + t % p.page_size) * kv_dim + hk * DS_HD;
        kraw[it][0] = ds_ldcg_v4(ly.k_cache + off + 8 * j);
        kraw[it][1] = ds_ldcg_v4(ly.k_cache + off + 64 + 8 * j);
        vr[it][0] = ds_ldcg_v4(ly.v_cache + off + 8 * j);
        vr[it][1] = ds_ldcg_v4(ly.v_cache + off + 64 + 8 * j);
      }
    }
    float s[DS_ATT_TILE / 4][DS_GROUP];
#pragma unroll
    for (int it = 0; it < DS_ATT_TILE / 4; ++it) {
      float kf[16];
      ds_unpack8(kraw[it][0], kf);
      ds_unpack8(kraw[it][1], kf + 8);
#pragma unroll
      for (int h = 0; h < DS_GROUP; ++h) {
        float d0 = 0.f, d1 = 0.f;
#pragma unroll
        for (int d = 0; d < 16; d += 2) {
          d0 = fmaf(qv[h][d], kf[d], d0);
          d1 = fmaf(qv[h][d + 1], kf[d + 1], d1);
        }
        float dot = d0 + d1;
        dot += __shfl_xor_sync(0xffffffffu, dot, 1);
        dot += __shfl_xor_sync(0xffffffffu, dot, 2);
        dot += __shfl_xor_sync(0xffffffffu, dot, 4);
        s[it][h] = valid[it] ? dot : -INFINITY;
      }
    }
    float alpha[DS_GROUP];
#pragma unroll
    for (int h = 0; h < DS_GROUP; ++h) {
      float mt = s[0][h];
#pragma unroll
      for (int it = 1; it < DS_ATT_TILE / 4; ++it) mt = fmaxf(mt, s[it][h]);
      mt = fmaxf(mt, __shfl_xor_sync(0xffffffffu, mt, 8));
      mt = fmaxf(mt, __shfl_xor_sync(0xffffffffu, mt, 16));
      const float mn = fmaxf(mx[h], mt);
      alpha[h] = ds_exp2(mx[h] - mn);
      mx[h] = mn;
      float ps = 0.f;
#pragma unroll
      for (int it = 0; it < DS_ATT_TILE / 4; ++it) {
        s[it][h] = ds_exp2(s[it][h] - mn);
        ps += s[it][h];
      }
      l[h] = fmaf(l[h], alpha[h], ps);
#pragma unroll
      for (int d = 0; d < 16; ++d) acc[h][d] *= alpha[h];
    }
#pragma unroll
    for (int it = 0; it < DS_ATT_TILE / 4; ++it) {
      float vf[16];
      ds_unpack8(vr[it][0], vf);
      ds_unpack8(vr[it][1], vf + 8);
#pragma unroll
      for (int h = 0; h < DS_GROUP; ++h)
#pragma unroll
        for (int d = 0; d < 16; ++d) acc[h][d] = fmaf(s[it][h], vf[d], acc[h][d]);
    }
  }
#pragma unroll
  for (int h = 0; h < DS_GROUP; ++h) {
    l[h] += __shfl_xor_sync(0xffffffffu, l[h], 8);
    l[h] += __shfl_xor_sync(0xffffffffu, l[h], 16);
#pragma unroll
    for (int d = 0; d < 16; ++d) {
      acc[h][d] += __shfl_xor_sync(0xffffffffu, acc[h][d], 8);
      acc[h][d] += __shfl_xor_sync(0xffffffffu, acc[h][d], 16);
    }
  }
  if (g == 0) {
#pragma unroll
    for (int h = 0; h < DS_GROUP; ++h) {
      const float inv = 1.f / l[h];
      float o[16];
#pragma unroll
      for (int d = 0; d < 16; ++d) o[d] = acc[h][d] * inv;
      __nv_bfloat16* op = p.attn + (int64_t)row * p.Q + (head0 + h) * DS_HD;
      stg_v4(op + 8 * j, ds_pack8(o));
      stg_v4(op + 64 + 8 * j, ds_pack8(o + 8));
    }
  }
}

// ---------------------------------------------------------------- RMSNorm of one row on one warp: the arithmetic and
// reduction tree of rmsnorm_kernel (norm_rope.cu: 256 threads, thread t owns the 16-byte vector t; warp sums,
// then one butterfly over the 8 warp totals) with the 8 warps of that kernel walked in sequence.
template <bool EAGER>
__device__ __forceinline__ void ds_norm_row(const __nv_bfloat16* x, const __nv_bfloat16* w, __nv_bfloat16* y, int cols,
                                            float eps, int lane) {
  const int nvec = cols >> 3;                     // <= 256
  uint4 raw[8];
  float tot[8];
#pragma unroll
  for (int vw = 0; vw < 8; ++vw) {
    const int v = vw * 32 + lane;
    float ss = 0.f;
    raw[vw] = make_uint4(0, 0, 0, 0);
    if (v < nvec) {
      raw[vw] = ds_ldcg_v4(x + 8 * v);
      float f[8];
      ds_unpack8(raw[vw], f);
#pragma unroll
      for (int e = 0; e < 8; ++e) ss = fmaf(f[e], f[e], ss);
    }
    tot[vw] = warp_sum(ss);
  }
  float t = 0.f;
#pragma unroll
  for (int vw = 0; vw < 8; ++vw) t = lane == vw ? tot[vw] : t;
  const float var = warp_sum(t) / (float)cols + eps;
  const float rstd = EAGER ? 1.0f / sqrtf(var) : rsqrtf(var);
#pragma unroll
  for (int vw = 0; vw < 8; ++vw) {
    const int v = vw * 32 + lane;
    if (v < nvec) {
      float f[8], wf[8], o[8];
      ds_unpack8(raw[vw], f);
      ds_unpack8(ldg_v4(w + 8 * v), wf);
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] = EAGER ? rbf(rbf(f[e] * rstd) * wf[e]) : (f[e] * rstd) * wf[e];
      stg_v4(y + 8 * v, ds_pack8(o));
    }
  }
}

template <bool EAGER>
__global__ void __launch_bounds__(DS_THREADS, 1)
decoder_stack_kernel(const DsParams p, const __grid_constant__ CUtensorMap tm_xn, const __grid_constant__ CUtensorMap tm_attn,
                     const __grid_constant__ CUtensorMap tm_act) {
  extern __shared__ uint8_t smem_dyn[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~uintptr_t(1023));
  uint8_t* scr_base = smem + DS_STAGES * DS_STAGE_BYTES;
  uint64_t* wfull = reinterpret_cast<uint64_t*>(scr_base + 4 * DS_SCR_BYTES);
  uint64_t* xfull = wfull + DS_STAGES;
  uint64_t* empty = xfull + DS_STAGES;
  uint64_t* accfull = empty + DS_STAGES;
  uint64_t* accempty = accfull + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accempty + 1);

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const int G = gridDim.x;
  const int cta = blockIdx.x;
  const int L = p.num_layers;

  // ---- this CTA's job (the same in every layer)
  int op = OP_NONE, tile = 0, split = 0, S = 1, kblocks = 0;
  {
    int c = cta;
    if (c < p.nq_tiles) { op = OP_QKV; tile = c; kblocks = p.H / G_BK; }
    else if ((c -= p.nq_tiles) < p.no_tiles * p.so) { op = OP_O; S = p.so; tile = c / S; split = c - tile * S; kblocks = p.Q / G_BK; }
    else if ((c -= p.no_tiles * p.so) < p.ng_tiles) { op = OP_GU; tile = c; kblocks = p.H / G_BK; }
    else if ((c -= p.ng_tiles) < p.nd_tiles * p.sd) { op = OP_DOWN; S = p.sd; tile = c / S; split = c - tile * S; kblocks = p.I / G_BK; }
  }
  const int kb0 = (split * kblocks) / S, kb1 = ((split + 1) * kblocks) / S;
  const int n_o = p.no_tiles * p.so, n_d = p.nd_tiles * p.sd;

  if (threadIdx.x == 0) {
    for (int s = 0; s < DS_STAGES; ++s) { mbar_init(wfull + s, 1); mbar_init(xfull + s, 1); mbar_init(empty + s, 1); }
    mbar_init(accfull, 1);
    mbar_init(accempty, 4);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, (uint32_t)p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const unsigned E = __ldcg(p.sync);               // launch epoch: counter targets are (E + 1) * arrivals
  const unsigned E1 = E + 1u;
  if (threadIdx.x == 0) DS_TRACE(0, 13);

  if (warp == 0) {
    // =========================== W producer: runs ahead through all layers, bounded only by the ring
    if (op != OP_NONE) {
      uint32_t slot = 0, ph = 1;
      for (int l = 0; l < L; ++l) {
        const DsLayer& ly = p.layer[l];
        const uint8_t* wp = op == OP_QKV ? ly.w_qkv : op == OP_O ? ly.w_o : op == OP_GU ? ly.w_gu : ly.w_down;
        const uint8_t* src = wp + ((int64_t)tile * kblocks + kb0) * G_WTILE_BYTES;
        for (int kb = kb0; kb < kb1; ++kb, src += G_WTILE_BYTES) {
          ds_mbar_wait(p.sync, empty + slot, ph, 0x100u + l);
          if (elect_one()) {
            mbar_expect_tx(wfull + slot, G_WTILE_BYTES);
            bulk_g2s(smem + slot * DS_STAGE_BYTES, src, G_WTILE_BYTES, wfull + slot);
          }
          __syncwarp();
          if (++slot == (uint32_t)DS_STAGES) { slot = 0; ph ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // =========================== MMA issuer
    if (op != OP_NONE) {
      const uint32_t idesc = umma_idesc_bf16(p.mpad);
      uint32_t slot = 0, ph = 0;
      for (int l = 0; l < L; ++l) {
        ds_mbar_wait(p.sync, accempty, (uint32_t)((l & 1) ^ 1), 0x200u + l);     // the previous layer's accumulator is drained
        tc_fence_after();
        for (int kb = kb0; kb < kb1; ++kb) {
          ds_mbar_wait(p.sync, wfull + slot, ph, 0x300u + l);
          ds_mbar_wait(p.sync, xfull + slot, ph, 0x400u + l);
          if (kb == kb0 && lane == 0) DS_TRACE(l, 11);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem + slot * DS_STAGE_BYTES);
          const uint32_t b_addr = a_addr + G_WTILE_BYTES;
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < G_BK / 16; ++k)
              umma_bf16(tmem_base, umma_desc_sw128(a_addr + k * 32), umma_desc_sw128(b_addr + k * 32), idesc,
                        (uint32_t)((kb > kb0) || (k > 0)));
            umma_commit(empty + slot);
            if (kb == kb1 - 1) umma_commit(accfull);
          }
          __syncwarp();
          if (kb == kb1 - 1 && lane == 0) DS_TRACE(l, 12);
          if (++slot == (uint32_t)DS_STAGES) { slot = 0; ph ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 2) {
    // =========================== x producer: waits for the phase that writes this job's input rows
    if (op != OP_NONE) {
      uint32_t slot = 0, ph = 1;
      const uint32_t xbytes = (uint32_t)p.mpad * 128u;
      for (int l = 0; l < L; ++l) {
        if (lane == 0) {
          if (op == OP_QKV) ds_wait(p.sync, ds_ctr(p, l, DS_XN_A), E1 * (unsigned)n_d, 0x500u + l);
          else if (op == OP_O) ds_wait(p.sync, ds_ctr(p, l, DS_ATT), E1 * (unsigned)G, 0x600u + l);
          else if (op == OP_GU) ds_wait(p.sync, ds_ctr(p, l, DS_XN_B), E1 * (unsigned)n_o, 0x700u + l);
          else ds_wait(p.sync, ds_ctr(p, l, DS_ACT), E1 * (unsigned)p.ng_tiles, 0x800u + l);
        }
        __syncwarp();
        if (lane == 0) DS_TRACE(l, 9);
        ds_fence_proxy_async_all();
        for (int kb = kb0; kb < kb1; ++kb) {
          ds_mbar_wait(p.sync, empty + slot, ph, 0x900u + l);
          if (elect_one()) {
            mbar_expect_tx(xfull + slot, xbytes);
            uint8_t* dst = smem + slot * DS_STAGE_BYTES + G_WTILE_BYTES;
            // tensor maps addressed in param space (a run-time pointer select would copy them to local memory)
            if (op == OP_O) tma_load_2d(dst, &tm_attn, kb * G_BK, 0, xfull + slot);
            else if (op == OP_DOWN) tma_load_2d(dst, &tm_act, kb * G_BK, 0, xfull + slot);
            else tma_load_2d(dst, &tm_xn, kb * G_BK, 0, xfull + slot);
          }
          __syncwarp();
          if (++slot == (uint32_t)DS_STAGES) { slot = 0; ph ^= 1; }
        }
        if (lane == 0) DS_TRACE(l, 10);
      }
    }
    __syncwarp();
  } else {
    // =========================== workers: epilogues, split reduce, row norms, attention
    const int tw = threadIdx.x - DS_W0;             // 0..127
    const int ww = warp - 3;                        // 0..3
    const int wq = warp & 3;                        // TMEM lane quarter this warp may read
    const int r = wq * 32 + lane;                   // row of the 128-feature tile held by this thread
    const int M = p.m;
    __nv_bfloat16* scr = reinterpret_cast<__nv_bfloat16*>(scr_base + ww * DS_SCR_BYTES);
    float frq[2];
    frq[0] = __ldg(p.inv_freq + 2 * lane);
    frq[1] = __ldg(p.inv_freq + 2 * lane + 1);
    pdl_wait();

    // rows this CTA normalises after its split reduce: CTA j of the n jobs of the op takes rows j, j + n, ...
    auto norm_rows = [&](const __nv_bfloat16* src, int64_t src_stride, const __nv_bfloat16* w, __nv_bfloat16* dst,
                         int64_t dst_stride, int j, int n) {
      for (int row = j + ww * n; row < M; row += 4 * n)
        ds_norm_row<EAGER>(src + (int64_t)row * src_stride, w, dst + (int64_t)row * dst_stride, p.H, p.eps, lane);
    };
    auto publish = [&](unsigned* ctr) {             // this CTA's stores -> visible, one arrival
      __threadfence();
      ds_workers_bar();
      if (tw == 0) ds_arrive(ctr);
    };
    auto await = [&](unsigned* ctr, unsigned target, unsigned code) {
      if (tw == 0) ds_wait(p.sync, ctr, target, code);
      ds_workers_bar();
    };
    auto load_acc = [&](uint32_t* v) {
      const int nq = p.mpad >> 4;
#pragma unroll
      for (int qq = 0; qq < 4; ++qq)
        if (qq < nq) tmem_ld16_nowait(tmem_base + ((uint32_t)(wq * 32) << 16) + (uint32_t)(16 * qq), v + 16 * qq);
      tmem_ld_wait();
    };
    // split-K job: accumulator -> L2 scratch, per-tile counter, fixed-order reduce of this CTA's feature slice + residual
    auto split_job = [&](int l, float* part, int N, int tctr, const __nv_bfloat16* res, int64_t res_stride) {
      ds_mbar_wait(p.sync, accfull, (uint32_t)(l & 1), 0xa00u + l);
      if (tw == 0) DS_TRACE(l, 1);
      tc_fence_after();
      uint32_t v[64];
      load_acc(v);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(accempty);
      float* dst = part + (int64_t)split * DS_MAX_M * N + (int64_t)tile * G_BN + r;
#pragma unroll
      for (int c = 0; c < 64; ++c)
        if (c < M) __stcg(dst + (int64_t)c * N, __uint_as_float(v[c]));
      publish(ds_ctr(p, l, tctr + tile));
      if (tw == 0) DS_TRACE(l, 2);
      await(ds_ctr(p, l, tctr + tile), E1 * (unsigned)S, 0xb00u + l);
      if (tw == 0) DS_TRACE(l, 3);
      const int rows_own = G_BN / S;
      const int upr = rows_own >> 3;                // (row, 8 features) units per batch row
      const int n0 = tile * G_BN + split * rows_own;
      for (int u = tw; u < M * upr; u += DS_WORKERS) {
        const int m = u / upr, n = n0 + (u - m * upr) * 8;
        float acc[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = 0.f;
        const float* src = part + (int64_t)m * N + n;
        for (int s = 0; s < S; ++s, src += (int64_t)DS_MAX_M * N) {
          const float4 a = __ldcg(reinterpret_cast<const float4*>(src));
          const float4 b = __ldcg(reinterpret_cast<const float4*>(src) + 1);
          acc[0] += a.x; acc[1] += a.y; acc[2] += a.z; acc[3] += a.w;
          acc[4] += b.x; acc[5] += b.y; acc[6] += b.z; acc[7] += b.w;
        }
        const uint4 rw = ds_ldcg_v4(res + (int64_t)m * res_stride + n);
        acc[0] = rbf(acc[0]) + bf16lo(rw.x); acc[1] = rbf(acc[1]) + bf16hi(rw.x);
        acc[2] = rbf(acc[2]) + bf16lo(rw.y); acc[3] = rbf(acc[3]) + bf16hi(rw.y);
        acc[4] = rbf(acc[4]) + bf16lo(rw.z); acc[5] = rbf(acc[5]) + bf16hi(rw.z);
        acc[6] = rbf(acc[6]) + bf16lo(rw.w); acc[7] = rbf(acc[7]) + bf16hi(rw.w);
        stg_v4(p.resid + (int64_t)m * p.H + n, ds_pack8(acc));
      }
    };

    // the first layer's input norm: done by the down jobs (they are idle until the first MLP)
    if (op == OP_DOWN) {
      norm_rows(p.x_in, p.x_stride, p.layer[0].ln_in, p.xn, p.H, cta - (p.nq_tiles + n_o + p.ng_tiles), n_d);
      publish(ds_ctr(p, 0, DS_XN_A));
    }
    for (int l = 0; l < L; ++l) {
      const DsLayer& ly = p.layer[l];
      if (tw == 0) DS_TRACE(l, 0);
      if (op == OP_QKV) {
        ds_mbar_wait(p.sync, accfull, (uint32_t)(l & 1), 0xc00u + l);
        if (tw == 0) DS_TRACE(l, 1);
        tc_fence_after();
        uint32_t v[64];
        load_acc(v);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(accempty);
        __nv_bfloat16* dst = p.qkv + tile * G_BN + r;
        const int qkv_w = p.Q + 2 * p.KV;
#pragma unroll
        for (int c = 0; c < 64; ++c)
          if (c < M) dst[(int64_t)c * qkv_w] = __float2bfloat16_rn(__uint_as_float(v[c]));
        publish(ds_ctr(p, l, DS_QKV));
        if (tw == 0) DS_TRACE(l, 2);
      }
      // ---- attention: every CTA's workers, one (row, kv head) item per warp
      await(ds_ctr(p, l, DS_QKV), E1 * (unsigned)p.nq_tiles, 0xd00u + l);
      if (tw == 0) DS_TRACE(l, 7);
      {
        const int items = M * p.kvh;
        for (int it = ww * G + cta; it < items; it += 4 * G)
          ds_attn_item<EAGER>(p, ly, it / p.kvh, it % p.kvh, lane, scr, frq);
      }
      publish(ds_ctr(p, l, DS_ATT));
      if (tw == 0) DS_TRACE(l, 8);
      if (op == OP_O) {
        split_job(l, p.part_o, p.H, DS_T_O, l == 0 ? p.x_in : p.resid, l == 0 ? p.x_stride : (int64_t)p.H);
        publish(ds_ctr(p, l, DS_O));
        if (tw == 0) DS_TRACE(l, 4);
        await(ds_ctr(p, l, DS_O), E1 * (unsigned)n_o, 0xe00u + l);
        if (tw == 0) DS_TRACE(l, 5);
        norm_rows(p.resid, p.H, ly.ln_post, p.xn, p.H, cta - p.nq_tiles, n_o);
        publish(ds_ctr(p, l, DS_XN_B));
        if (tw == 0) DS_TRACE(l, 6);
      } else if (op == OP_GU) {
        ds_mbar_wait(p.sync, accfull, (uint32_t)(l & 1), 0xf00u + l);
        if (tw == 0) DS_TRACE(l, 1);
        tc_fence_after();
        uint32_t v[64];
        load_acc(v);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(accempty);
        // packed gate_up rows (b2tts_gemm_pack_weight, B2TTS_PACK_GATE_UP): rows r with (r & 4) == 0 are gate features,
        // r + 4 the matching up features; activation feature = tile * 64 + (r >> 3) * 4 + (r & 3)
        __nv_bfloat16* dst = p.act + tile * 64 + (r >> 3) * 4 + (r & 3);
#pragma unroll
        for (int c = 0; c < 64; ++c) {
          const float mine = __uint_as_float(v[c]);
          const float other = __shfl_down_sync(0xffffffffu, mine, 4);
          if (c < M && (r & 4) == 0) {
            const float gt = rbf(mine), up = rbf(other);
            const float o = EAGER ? rbf(gsilu_f(gt)) * up : gsilu_f(gt) * up;
            dst[(int64_t)c * p.I] = __float2bfloat16_rn(o);
          }
        }
        publish(ds_ctr(p, l, DS_ACT));
        if (tw == 0) DS_TRACE(l, 2);
      } else if (op == OP_DOWN) {
        split_job(l, p.part_d, p.H, DS_T_D, p.resid, p.H);
        publish(ds_ctr(p, l, DS_DOWN));
        if (tw == 0) DS_TRACE(l, 4);
        await(ds_ctr(p, l, DS_DOWN), E1 * (unsigned)n_d, 0x1000u + l);
        if (tw == 0) DS_TRACE(l, 5);
        const int j = cta - (p.nq_tiles + n_o + p.ng_tiles);
        if (l + 1 < L) norm_rows(p.resid, p.H, p.layer[l + 1].ln_in, p.xn, p.H, j, n_d);
        else norm_rows(p.resid, p.H, p.final_w, p.out, p.out_stride, j, n_d);
        publish(ds_ctr(p, l + 1, DS_XN_A));
        if (tw == 0) DS_TRACE(l, 6);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) DS_TRACE(0, 14);
  if (warp == 1) tmem_dealloc(tmem_base, (uint32_t)p.tmem_cols);
  // the last CTA to leave opens the next epoch (every CTA has read this one long ago)
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned old = atomicAdd(p.sync + 1, 1u);
    if (old == E1 * (unsigned)G - 1u) {
      p.sync[0] = E1;
      __threadfence();
    }
  }
}

static unsigned long long* g_ds_trace = nullptr;   // debug only (b2tts_debug_stack_trace)
struct DsPlan { int nq, no, so, ng, nd, sd; };
static bool ds_plan(int hidden, int q_size, int kv_size, int inter, DsPlan* out) {
  if (hidden <= 0 || q_size <= 0 || kv_size <= 0 || inter <= 0) return false;
  if (hidden % G_BN || (q_size + 2 * kv_size) % G_BN || (2 * inter) % G_BN) return false;
  if (hidden % G_BK || q_size % G_BK || inter % G_BK || hidden > 2048) return false;
  DsPlan pl{(q_size + 2 * kv_size) / G_BN, hidden / G_BN, 1, 2 * inter / G_BN, hidden / G_BN, 1};
  if (pl.no > DS_MAX_SPLIT_TILES) return false;
  const int G = sm_count();
  if (pl.nq + pl.no + pl.ng + pl.nd > G) return false;
  // double the K split of whichever of o / down has more k-blocks per job while the jobs still fit one per SM
  for (;;) {
    const int kb_o = q_size / G_BK / pl.so, kb_d = inter / G_BK / pl.sd;
    const bool pick_d = kb_d >= kb_o;
    int so = pl.so, sd = pl.sd;
    if (pick_d) sd *= 2; else so *= 2;
    const bool ok = so <= 8 && sd <= 8 && (q_size / G_BK) % so == 0 && (inter / G_BK) % sd == 0 &&
                    pl.nq + pl.no * so + pl.ng + pl.nd * sd <= G;
    if (!ok) {
      // try the other one once
      so = pl.so; sd = pl.sd;
      if (pick_d) so *= 2; else sd *= 2;
      const bool ok2 = so <= 8 && sd <= 8 && (q_size / G_BK) % so == 0 && (inter / G_BK) % sd == 0 &&
                       pl.nq + pl.no * so + pl.ng + pl.nd * sd <= G && (pick_d ? kb_o : kb_d) > DS_STAGES;
      if (!ok2) break;
    }
    pl.so = so; pl.sd = sd;
  }
  *out = pl;
  return true;
}

static size_t ds_align(size_t v) { return (v + 255) & ~(size_t)255; }

}  // namespace b2

extern "C" B2TTS_API int b2tts_decoder_stack_plan(int hidden, int q_size, int kv_size, int intermediate,
                                                  int* o_splits, int* down_splits) {
  b2::DsPlan pl;
  if (!b2::ds_plan(hidden, q_size, kv_size, intermediate, &pl)) return 0;
  if (o_splits) *o_splits = pl.so;
  if (down_splits) *down_splits = pl.sd;
  return 1;
}

extern "C" B2TTS_API size_t b2tts_decoder_stack_workspace_bytes(int hidden, int q_size, int kv_size, int intermediate) {
  using namespace b2;
  DsPlan pl;
  if (!ds_plan(hidden, q_size, kv_size, intermediate, &pl)) return 0;
  const size_t rows = DS_MAX_M;
  return DS_SYNC_BYTES + ds_align(rows * hidden * 2) * 2 + ds_align(rows * (q_size + 2 * kv_size) * 2) +
         ds_align(rows * q_size * 2) + ds_align(rows * intermediate * 2) + ds_align((size_t)pl.so * rows * hidden * 4) +
         ds_align((size_t)pl.sd * rows * hidden * 4) + 256;
}

extern "C" B2TTS_API int b2tts_decoder_stack_decode(
    const void* x, int64_t x_row_stride, void* out, int64_t out_row_stride, const b2tts_stack_layer* layers, int num_layers,
    const void* final_norm_weight, int m, int hidden, int num_heads, int num_kv_heads, int head_dim, int intermediate,
    float eps, int rounding, const int64_t* positions, const float* inv_freq, float attention_scaling,
    const int32_t* slot_mapping, const int32_t* block_table, int bt_stride, const int32_t* context_lens, int page_size,
    float scale, void* workspace, size_t workspace_bytes, void* stream) {
  using namespace b2;
  const char* what = "decoder_stack_decode";
  B2_REQUIRE(m >= 0 && m <= DS_MAX_M, B2TTS_EUNSUPPORTED, "%s: m=%d (0..%d)", what, m, DS_MAX_M);
  if (m == 0) return B2TTS_OK;
  B2_REQUIRE(layers && num_layers >= 1 && num_layers <= DS_MAX_LAYERS, B2TTS_EUNSUPPORTED, "%s: 1..%d layers", what, DS_MAX_LAYERS);
  B2_REQUIRE(head_dim == DS_HD && num_kv_heads >= 1 && num_heads == DS_GROUP * num_kv_heads, B2TTS_EUNSUPPORTED,
             "%s: head_dim %d / heads %d / kv heads %d unsupported (128, heads = 2 x kv heads)", what, head_dim, num_heads,
             num_kv_heads);
  const int Q = num_heads * head_dim, KV = num_kv_heads * head_dim;
  DsPlan pl;
  B2_REQUIRE(ds_plan(hidden, Q, KV, intermediate, &pl), B2TTS_EUNSUPPORTED,
             "%s: hidden %d / intermediate %d do not fit one job per SM", what, hidden, intermediate);
  B2_REQUIRE(x && out && final_norm_weight && positions && inv_freq && slot_mapping && block_table && context_lens && workspace,
             B2TTS_EINVAL, "%s: null pointer", what);
  B2_REQUIRE(rounding == B2TTS_ROUND_EAGER || rounding == B2TTS_ROUND_FUSED, B2TTS_EINVAL, "%s: rounding %d", what, rounding);
  B2_REQUIRE(page_size > 0 && bt_stride > 0, B2TTS_EINVAL, "%s: page_size=%d bt_stride=%d", what, page_size, bt_stride);
  B2_REQUIRE(aligned16(x) && aligned16(out) && aligned16(final_norm_weight) && x_row_stride % 8 == 0 && out_row_stride % 8 == 0 &&
                 (reinterpret_cast<uintptr_t>(workspace) & 255u) == 0,
             B2TTS_EALIGN, "%s: pointers / strides must be 16-byte aligned (workspace 256)", what);
  B2_REQUIRE(workspace_bytes >= b2tts_decoder_stack_workspace_bytes(hidden, Q, KV, intermediate), B2TTS_EWORKSPACE,
             "%s: workspace %zu B too small", what, workspace_bytes);
  DsParams p{};
  for (int l = 0; l < num_layers; ++l) {
    const b2tts_stack_layer& s = layers[l];
    B2_REQUIRE(s.qkv_w && s.o_w && s.gate_up_w && s.down_w && s.input_norm_w && s.post_norm_w && s.q_norm_w && s.k_norm_w &&
                   s.k_cache && s.v_cache, B2TTS_EINVAL, "%s: layer %d null pointer", what, l);
    B2_REQUIRE(aligned16(s.qkv_w) && aligned16(s.o_w) && aligned16(s.gate_up_w) && aligned16(s.down_w) &&
                   aligned16(s.input_norm_w) && aligned16(s.post_norm_w) && aligned16(s.k_cache) && aligned16(s.v_cache) &&
                   (reinterpret_cast<uintptr_t>(s.q_norm_w) & 3u) == 0 && (reinterpret_cast<uintptr_t>(s.k_norm_w) & 3u) == 0,
               B2TTS_EALIGN, "%s: layer %d alignment", what, l);
    DsLayer& d = p.layer[l];
    d.w_qkv = static_cast<const uint8_t*>(s.qkv_w); d.w_o = static_cast<const uint8_t*>(s.o_w);
    d.w_gu = static_cast<const uint8_t*>(s.gate_up_w); d.w_down = static_cast<const uint8_t*>(s.down_w);
    d.ln_in = static_cast<const __nv_bfloat16*>(s.input_norm_w); d.ln_post = static_cast<const __nv_bfloat16*>(s.post_norm_w);
    d.qn_w = static_cast<const __nv_bfloat16*>(s.q_norm_w); d.kn_w = static_cast<const __nv_bfloat16*>(s.k_norm_w);
    d.k_cache = static_cast<__nv_bfloat16*>(s.k_cache); d.v_cache = static_cast<__nv_bfloat16*>(s.v_cache);
  }
  p.num_layers = num_layers;
  p.x_in = static_cast<const __nv_bfloat16*>(x); p.x_stride = x_row_stride;
  p.out = static_cast<__nv_bfloat16*>(out); p.out_stride = out_row_stride;
  p.final_w = static_cast<const __nv_bfloat16*>(final_norm_weight);
  p.m = m; p.mpad = (m + 15) & ~15; p.rounding = rounding; p.eps = eps;
  p.tmem_cols = p.mpad <= 32 ? 32 : 64;
  p.H = hidden; p.Q = Q; p.KV = KV; p.I = intermediate; p.heads = num_heads; p.kvh = num_kv_heads;
  p.nq_tiles = pl.nq; p.no_tiles = pl.no; p.so = pl.so; p.ng_tiles = pl.ng; p.nd_tiles = pl.nd; p.sd = pl.sd;
  p.positions = positions; p.inv_freq = inv_freq; p.attention_scaling = attention_scaling;
  p.scale_log2 = scale * 1.4426950408889634f;
  p.slot_mapping = slot_mapping; p.block_table = block_table; p.bt_stride = bt_stride; p.page_size = page_size;
  p.context_lens = context_lens;
  uint8_t* w = static_cast<uint8_t*>(workspace);
  p.sync = reinterpret_cast<unsigned*>(w); w += DS_SYNC_BYTES;
  const size_t rows = DS_MAX_M;
  p.xn = reinterpret_cast<__nv_bfloat16*>(w); w += ds_align(rows * hidden * 2);
  p.resid = reinterpret_cast<__nv_bfloat16*>(w); w += ds_align(rows * hidden * 2);
  p.qkv = reinterpret_cast<__nv_bfloat16*>(w); w += ds_align(rows * (Q + 2 * KV) * 2);
  p.attn = reinterpret_cast<__nv_bfloat16*>(w); w += ds_align(rows * Q * 2);
  p.act = reinterpret_cast<__nv_bfloat16*>(w); w += ds_align(rows * intermediate * 2);
  p.part_o = reinterpret_cast<float*>(w); w += ds_align((size_t)pl.so * rows * hidden * 4);
  p.part_d = reinterpret_cast<float*>(w);
  p.trace = g_ds_trace;
  CUtensorMap tm_xn, tm_attn, tm_act;
  int rc = make_x_tensor_map(&tm_xn, p.xn, hidden, DS_MAX_M, hidden, p.mpad, what);
  if (rc) return rc;
  rc = make_x_tensor_map(&tm_attn, p.attn, Q, DS_MAX_M, Q, p.mpad, what);
  if (rc) return rc;
  rc = make_x_tensor_map(&tm_act, p.act, intermediate, DS_MAX_M, intermediate, p.mpad, what);
  if (rc) return rc;

  auto kern = rounding == B2TTS_ROUND_EAGER ? decoder_stack_kernel<true> : decoder_stack_kernel<false>;
  const size_t smem_bytes = (size_t)DS_STAGES * DS_STAGE_BYTES + 4 * DS_SCR_BYTES + (3 * DS_STAGES + 2) * 8 + 16 + 1024;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
  if (e != cudaSuccess) return cuda_fail(e, "decoder_stack_decode(smem attr)");
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(sm_count(), 1, 1);
  cfg.blockDim = dim3(DS_THREADS, 1, 1);
  cfg.dynamicSmemBytes = smem_bytes;
  cfg.stream = static_cast<cudaStream_t>(stream);
  cudaLaunchAttribute attrs[1];
  int na = 0;
  static const bool coop = [] { const char* v = getenv("B2TTS_DS_COOP"); return !(v && v[0] == '0'); }();
  if (coop) {                                    // every CTA spins on its peers: all of them must be resident
    attrs[na].id = cudaLaunchAttributeCooperative;
    attrs[na].val.cooperative = 1;
    ++na;
  }
  cfg.attrs = attrs;
  cfg.numAttrs = na;
  e = cudaLaunchKernelEx(&cfg, kern, p, tm_xn, tm_attn, tm_act);
  count_launch();
  if (e != cudaSuccess) return cuda_fail(e, what);
  return B2TTS_OK;
}

// Debug aid (tools/stack_trace.py): following launches record 16 globaltimer stamps per (CTA, layer) into
// buffer[(cta * 8 + layer) * 16 + slot] (uint64).  NULL switches tracing off.
extern "C" B2TTS_API int b2tts_debug_stack_trace(void* buffer) {
  b2::g_ds_trace = static_cast<unsigned long long*>(buffer);
  return B2TTS_OK;
}

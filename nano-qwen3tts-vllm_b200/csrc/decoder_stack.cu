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
//     (b2tts_gemm_pack_weight images, 16 KiB 1-D bulk copies) into a 7-stage ring and runs AHEAD: while layer l
//     waits for its activations, layer l's (and then l+1's) weights are already in shared memory, so HBM never
//     waits for a dependency and no phase pays a weight-fetch latency;
//   * phases are linked by monotonic arrival counters in L2 (red.release / ld.acquire, epoch-scaled so that a CUDA
//     graph can replay the launch without a memset) instead of kernel boundaries; the activations (<= 64 rows) are
//     fetched by 2-D tensor-map copies as the UMMA B operand once their counter is complete;
//   * tcgen05.mma (weights = 128-row A operand, batch = N, fp32 accumulator in TMEM) exactly as gemm_stream_kernel;
//   * the o / down splits leave fp32 partials in an L2 scratch; the same CTAs then each FINISH whole rows: fixed-order
//     sum over the splits (deterministic), residual add, and the RMSNorm of the finished row for the next projection
//     (arithmetic and reduction tree of rmsnorm_kernel) - one hand-over instead of reduce -> barrier -> norm;
//   * attention (q/k head norm + RoPE + paged KV write + GQA attention over the cached tokens, arithmetic and
//     summation order of fused_decode_attn_kernel) runs on 4 worker warps of ALL CTAs, one (sequence, kv head) per
//     warp, between the qkv and o phases; the cached K / V rows of the item (written by earlier launches) are staged
//     in shared memory by cp.async while the QKV projection is still running.
// Rounding points are those of the separate launches; with the same K splits the result is
// bit-identical to rmsnorm -> gemm -> fused_decode_attn -> gemm(residual) -> rmsnorm -> gemm(SiLU*mul) -> gemm(residual).
//
// Every wait is bounded: a dependency that does not resolve within DS_TIMEOUT_NS traps (launch failure) instead of
// hanging the device.
#include <stdlib.h>

#include "umma.cuh"

namespace b2 {

constexpr int DS_MAX_LAYERS = 8;
constexpr int DS_THREADS = 352;                  // warp 0: W producer, 1: MMA issuer, 2: x producer, 3-10: workers
constexpr int DS_W0 = 96;                        // first worker thread
constexpr int DS_WORKERS = 256;
constexpr int DS_MAX_M = 64;                     // batch rows (UMMA N)
constexpr int DS_STAGES = 7;
constexpr int DS_STAGE_BYTES = G_WTILE_BYTES + DS_MAX_M * 128;   // 24 KiB: W tile + x tile
constexpr int DS_HD = 128;
constexpr int DS_GROUP = 2;                      // q heads per kv head
constexpr int DS_ATT_TILE = 8;                   // tokens per online-softmax step (as fused_decode_attn_kernel)
constexpr int DS_ATT_WARPS = 4;                  // worker warps that own an attention item
constexpr int DS_SCR_BYTES = 1024;               // per attention warp: q [2][128], k [128], v [128] bf16 of the new token
constexpr int DS_KV_PRE = 16;                    // cached tokens staged in shared memory per item
constexpr int DS_KV_BYTES = 2 * DS_KV_PRE * DS_HD * 2;   // K and V rows of one head
constexpr int DS_MAX_SPLIT_TILES = 16;
constexpr int DS_CTRS = 8;                       // counters per layer
constexpr int DS_CTR_STRIDE = 8;                 // words between counters (one 32-byte sector each)
constexpr int DS_CTR_BASE = 64;                  // words 0..63: epoch, exit counter, error
constexpr size_t DS_SYNC_BYTES = 32768;
constexpr unsigned long long DS_TIMEOUT_NS = 4000000000ull;

enum { DS_XN_A = 0, DS_QKV = 1, DS_ATT = 2, DS_PO = 3, DS_XN_B = 4, DS_ACT = 5, DS_PD = 6 };
enum { OP_QKV = 0, OP_O = 1, OP_GU = 2, OP_DOWN = 3, OP_NONE = 4 };

struct DsLayer {
  const uint8_t *w_qkv, *w_o, *w_gu, *w_down;            // packed weights
  const __nv_bfloat16 *ln_in, *ln_post, *qn_w, *kn_w;    // norm weights
  __nv_bfloat16 *k_cache, *v_cache;                      // [pages, page, kvh, 128]
};

struct DsCore {
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
  unsigned long long* trace;                     // debug: 32 globaltimer stamps per (CTA, layer) (null = off)
};
struct DsParams : DsCore {
  DsLayer layer[DS_MAX_LAYERS];
};
#define DS_TRACE(l, slot) do { if (p.trace) p.trace[((int64_t)blockIdx.x * DS_MAX_LAYERS + (l)) * 32 + (slot)] = gtimer(); } while (0)

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
__device__ __forceinline__ void ds_workers_bar() { asm volatile("bar.sync 1, 256;" ::: "memory"); }
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

__device__ __forceinline__ void ds_cp_async16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void ds_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void ds_cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// The (row, kv head) item an attention warp owns for the whole launch: everything that does not depend on the
// layer (context length, cache slot, RoPE cos/sin, where the cached tokens live) is worked out once.
struct DsAttnItem {
  int row, hk, L, slot, npre;                    // npre: cached tokens staged in shared memory (<= DS_KV_PRE)
  float c1[2], s1[2];
  uint32_t tokslot[DS_KV_PRE / 2];               // cache slot of the token copy i of this lane stages (or ~0u)
};

// stage the cached K / V rows of this item's head (tokens 0 .. npre-1) in shared memory: lanes 0-15 copy the 16
// 16-byte pieces of token 2i, lanes 16-31 those of token 2i+1.  They were written by earlier launches, so the copies
// run while the QKV projection of the layer is still being computed.
__device__ __forceinline__ void ds_kv_prefetch(const DsLayer& ly, const DsAttnItem& it, int kv_dim, int lane,
                                               __nv_bfloat16* ksm, __nv_bfloat16* vsm) {
  const int part = lane & 15;
#pragma unroll
  for (int i = 0; i < DS_KV_PRE / 2; ++i) {
    if (it.tokslot[i] != 0xffffffffu) {
      const int tok = 2 * i + (lane >> 4);
      const int64_t off = (int64_t)it.tokslot[i] * kv_dim + it.hk * DS_HD + part * 8;
      ds_cp_async16(ksm + tok * DS_HD + part * 8, ly.k_cache + off);
      ds_cp_async16(vsm + tok * DS_HD + part * 8, ly.v_cache + off);
    }
  }
  ds_cp_async_commit();
}

// ---------------------------------------------------------------- one attention item: (row, kv head) on one warp.
// Arithmetic, lane mapping and summation order of fused_decode_attn_kernel (attention_fused.cu) with nsplit == 1;
// cached tokens come from the shared-memory copy (or, beyond DS_KV_PRE tokens, straight from the paged cache),
// the new token from the warp's scratch.
template <bool EAGER>
__device__ __forceinline__ void ds_attn_item(const DsParams& p, const DsLayer& ly, int row, int hk, int L, int slot, int npre,
                                          float c10, float c11, float s10, float s11, int lane,
                                          __nv_bfloat16* scr, const __nv_bfloat16* ksm, const __nv_bfloat16* vsm,
                                          unsigned long long* tr) {
  const int kv_dim = p.kvh * DS_HD;
  const int qkv_w = p.Q + 2 * p.KV;
  const float c1[2] = {c10, c11}, s1[2] = {s10, s11};
  const int head0 = hk * DS_GROUP;
  const int g = lane >> 3, j = lane & 7;
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
  {
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
  if (tr && lane == 0) tr[4] = gtimer();
  ds_cp_async_wait_all();                         // this lane's staged K / V pieces have landed ...
  __syncwarp();                                   // ... and so have everybody else's, and the scratch rows

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
    // K rows of this lane's two token slots, then (after the softmax) the V rows: staged tokens and the new token come
    // from shared memory, anything beyond the staged window straight from the paged cache
    const __nv_bfloat16* vp[DS_ATT_TILE / 4];
    bool vsh[DS_ATT_TILE / 4];
    float sc[DS_ATT_TILE / 4][DS_GROUP];
#pragma unroll
    for (int s = 0; s < DS_ATT_TILE / 4; ++s) {
      const int tok = s * 4 + g;
      const bool valid = tok < ntok;
      const int t = tok0 + (valid ? tok : 0);
      uint4 k0, k1;
      if (t == L - 1) {
        k0 = *reinterpret_cast<const uint4*>(kn_s + 8 * j);
        k1 = *reinterpret_cast<const uint4*>(kn_s + 64 + 8 * j);
        vp[s] = vn_s; vsh[s] = true;
      } else if (t < npre) {
        k0 = *reinterpret_cast<const uint4*>(ksm + t * DS_HD + 8 * j);
        k1 = *reinterpret_cast<const uint4*>(ksm + t * DS_HD + 64 + 8 * j);
        vp[s] = vsm + t * DS_HD; vsh[s] = true;
      } else {
        int page = __ldg(bt + t / p.page_size);
        if (page < 0) page = 0;
        const int64_t off = ((int64_t)page * p.page_size + t % p.page_size) * kv_dim + hk * DS_HD;
        k0 = ds_ldcg_v4(ly.k_cache + off + 8 * j);
        k1 = ds_ldcg_v4(ly.k_cache + off + 64 + 8 * j);
        vp[s] = ly.v_cache + off; vsh[s] = false;
      }
      float kf[16];
      ds_unpack8(k0, kf);
      ds_unpack8(k1, kf + 8);
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
        sc[s][h] = valid ? dot : -INFINITY;
      }
    }
    float alpha[DS_GROUP];
#pragma unroll
    for (int h = 0; h < DS_GROUP; ++h) {
      float mt = sc[0][h];
#pragma unroll
      for (int s = 1; s < DS_ATT_TILE / 4; ++s) mt = fmaxf(mt, sc[s][h]);
      mt = fmaxf(mt, __shfl_xor_sync(0xffffffffu, mt, 8));
      mt = fmaxf(mt, __shfl_xor_sync(0xffffffffu, mt, 16));
      const float mn = fmaxf(mx[h], mt);
      alpha[h] = ds_exp2(mx[h] - mn);
      mx[h] = mn;
      float ps = 0.f;
#pragma unroll
      for (int s = 0; s < DS_ATT_TILE / 4; ++s) {
        sc[s][h] = ds_exp2(sc[s][h] - mn);
        ps += sc[s][h];
      }
      l[h] = fmaf(l[h], alpha[h], ps);
#pragma unroll
      for (int d = 0; d < 16; ++d) acc[h][d] *= alpha[h];
    }
#pragma unroll
    for (int s = 0; s < DS_ATT_TILE / 4; ++s) {
      uint4 v0, v1;
      if (vsh[s]) {
        v0 = *reinterpret_cast<const uint4*>(vp[s] + 8 * j);
        v1 = *reinterpret_cast<const uint4*>(vp[s] + 64 + 8 * j);
      } else {
        v0 = ds_ldcg_v4(vp[s] + 8 * j);
        v1 = ds_ldcg_v4(vp[s] + 64 + 8 * j);
      }
      float vf[16];
      ds_unpack8(v0, vf);
      ds_unpack8(v1, vf + 8);
#pragma unroll
      for (int h = 0; h < DS_GROUP; ++h)
#pragma unroll
        for (int d = 0; d < 16; ++d) acc[h][d] = fmaf(sc[s][h], vf[d], acc[h][d]);
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
  if (tr && lane == 0) tr[5] = gtimer();
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
  __syncwarp();                                   // the scratch / staged rows may be overwritten from here on
}

// ---------------------------------------------------------------- one finished row on the 256 worker threads:
// fixed-order sum of the split-K partials + residual add (bf16(residual + bf16(acc)), written back to the residual
// stream), then the RMSNorm of the rounded row for the next projection.  Thread t owns the 16-byte vector t, warp
// sums, then one butterfly over the 8 warp totals: the mapping and reduction tree of rmsnorm_kernel (norm_rope.cu).
template <bool EAGER>
__device__ __forceinline__ void ds_row_phase(const DsParams& p, int row, const float* part, int S, const __nv_bfloat16* res,
                                             int64_t res_stride, const __nv_bfloat16* w, __nv_bfloat16* dst, int tw, float* red) {
  const int nvec = p.H >> 3;                      // <= 256
  const int lane = tw & 31, ww = tw >> 5;
  const bool active = tw < nvec;
  uint4 rowv = make_uint4(0, 0, 0, 0), wv = make_uint4(0, 0, 0, 0);
  float ss = 0.f;
  if (active) {
    wv = ldg_v4(w + 8 * tw);
    const uint4 rw = ds_ldcg_v4(res + (int64_t)row * res_stride + 8 * tw);
    if (part) {
      const float* src = part + (int64_t)row * p.H + 8 * tw;
      const int64_t sstride = (int64_t)DS_MAX_M * p.H;
      float acc[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] = 0.f;
      for (int s0 = 0; s0 < S; s0 += 4) {           // 4 splits in flight; summed in split order
        float4 a[4], b[4];
#pragma unroll
        for (int s = 0; s < 4; ++s)
          if (s0 + s < S) {
            a[s] = __ldcg(reinterpret_cast<const float4*>(src + (s0 + s) * sstride));
            b[s] = __ldcg(reinterpret_cast<const float4*>(src + (s0 + s) * sstride) + 1);
          }
#pragma unroll
        for (int s = 0; s < 4; ++s)
          if (s0 + s < S) {
            acc[0] += a[s].x; acc[1] += a[s].y; acc[2] += a[s].z; acc[3] += a[s].w;
            acc[4] += b[s].x; acc[5] += b[s].y; acc[6] += b[s].z; acc[7] += b[s].w;
          }
      }
      acc[0] = rbf(acc[0]) + bf16lo(rw.x); acc[1] = rbf(acc[1]) + bf16hi(rw.x);
      acc[2] = rbf(acc[2]) + bf16lo(rw.y); acc[3] = rbf(acc[3]) + bf16hi(rw.y);
      acc[4] = rbf(acc[4]) + bf16lo(rw.z); acc[5] = rbf(acc[5]) + bf16hi(rw.z);
      acc[6] = rbf(acc[6]) + bf16lo(rw.w); acc[7] = rbf(acc[7]) + bf16hi(rw.w);
      rowv = ds_pack8(acc);
      stg_v4(p.resid + (int64_t)row * p.H + 8 * tw, rowv);
    } else {
      rowv = rw;
    }
    float f[8];
    ds_unpack8(rowv, f);
#pragma unroll
    for (int e = 0; e < 8; ++e) ss = fmaf(f[e], f[e], ss);
  }
  ss = warp_sum(ss);
  if (lane == 0) red[ww] = ss;
  ds_workers_bar();
  float t = lane < 8 ? red[lane] : 0.f;
  t = warp_sum(t);
  const float var = t / (float)p.H + p.eps;
  const float rstd = EAGER ? 1.0f / sqrtf(var) : rsqrtf(var);
  if (active) {
    float f[8], wf[8], o[8];
    ds_unpack8(rowv, f);
    ds_unpack8(wv, wf);
#pragma unroll
    for (int e = 0; e < 8; ++e) o[e] = EAGER ? rbf(rbf(f[e] * rstd) * wf[e]) : (f[e] * rstd) * wf[e];
    stg_v4(dst + 8 * tw, ds_pack8(o));
  }
}

// Everything a worker warp needs about its CTA's job (scalarised by the compiler: every user is inlined).
struct DsCtx {
  uint32_t tmem_base;
  uint64_t *accfull, *accempty;
  unsigned E1;
  int G, cta, L, tile, split, job, n_o, n_d;
  DsLayer* layers;
  uint8_t *scr_base, *kv_base;
  float* red_s;
};

// One worker warp for the whole launch, specialised by the CTA's job so that each job's layer loop carries only its
// own epilogue (one merged loop made the compiler keep every job's hoisted addresses live at once, and spill).
template <bool EAGER, int OP>
__device__ __forceinline__ void ds_worker(const DsParams& p, const DsCtx& c, int warp, int lane) {
  const int cta = c.cta, G = c.G, L = c.L;
  const int tw = threadIdx.x - DS_W0;             // 0..255
  const int ww = warp - 3;                        // 0..7
  const int wq = warp & 3;                        // TMEM lane quarter this warp may read
  const int ch = ww >> 2;                         // which 32 accumulator columns (batch rows) this warp drains
  const int r = wq * 32 + lane;                   // row of the 128-feature tile held by this thread
  const int M = p.m;
  const int kv_dim = p.kvh * DS_HD;
  const unsigned E1 = c.E1;
  __nv_bfloat16* scr = reinterpret_cast<__nv_bfloat16*>(c.scr_base + (ww & (DS_ATT_WARPS - 1)) * DS_SCR_BYTES);
  __nv_bfloat16* ksm = reinterpret_cast<__nv_bfloat16*>(c.kv_base + (ww & (DS_ATT_WARPS - 1)) * DS_KV_BYTES);
  __nv_bfloat16* vsm = ksm + DS_KV_PRE * DS_HD;

  // ---- this warp's attention item (warps 0-3: item = warp * G + cta), fixed for the launch
  DsAttnItem item;
  const int item_id = ww * G + cta;
  const bool has_item = ww < DS_ATT_WARPS && item_id < M * p.kvh;
  item.row = has_item ? item_id / p.kvh : 0;
  item.hk = has_item ? item_id % p.kvh : 0;
  item.L = 0; item.slot = -1; item.npre = 0;
  item.c1[0] = item.c1[1] = item.s1[0] = item.s1[1] = 0.f;
#pragma unroll
  for (int i = 0; i < DS_KV_PRE / 2; ++i) item.tokslot[i] = 0xffffffffu;
  if (has_item) {
    item.L = __ldg(p.context_lens + item.row);
    item.slot = __ldg(p.slot_mapping + item.row);
    const int64_t pos = __ldg(p.positions + item.row);
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const float a1 = (float)pos * __ldg(p.inv_freq + 2 * lane + e);
      item.c1[e] = rbf(cosf(a1) * p.attention_scaling);
      item.s1[e] = rbf(sinf(a1) * p.attention_scaling);
    }
    item.npre = max(0, min(item.L - 1, DS_KV_PRE));
    const int32_t* bt = p.block_table + (int64_t)item.row * p.bt_stride;
#pragma unroll
    for (int i = 0; i < DS_KV_PRE / 2; ++i) {
      const int tok = 2 * i + (lane >> 4);
      if (tok < item.npre) {
        int page = __ldg(bt + tok / p.page_size);
        if (page < 0) page = 0;
        item.tokslot[i] = (uint32_t)(page * p.page_size + tok % p.page_size);
      }
    }
  }

  auto publish = [&](unsigned* ctr) {             // bar.sync orders every worker's stores before the one release
    ds_workers_bar();
    if (tw == 0) ds_arrive(ctr);
  };
  auto await = [&](unsigned* ctr, unsigned target, unsigned code) {
    if (tw == 0) ds_wait(p.sync, ctr, target, code);
    ds_workers_bar();
  };
  auto load_acc = [&](int l, uint32_t* v, unsigned code) {      // this warp's 32 accumulator columns of row r
    ds_mbar_wait(p.sync, c.accfull, (uint32_t)(l & 1), code);
    if (tw == 0) DS_TRACE(l, 1);
    tc_fence_after();
    const int nq = p.mpad >> 4;
#pragma unroll
    for (int qq = 0; qq < 2; ++qq)
      if (ch * 2 + qq < nq)
        tmem_ld16_nowait(c.tmem_base + ((uint32_t)(wq * 32) << 16) + (uint32_t)(ch * 32 + 16 * qq), v + 16 * qq);
    tmem_ld_wait();
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(c.accempty);
  };
  // rows this CTA finishes after a split-K projection: job j of the n jobs of the op owns rows j, j + n, ...
  int red_flip = 0;
  auto rows_phase = [&](const float* part, int nsplit, const __nv_bfloat16* res, int64_t res_stride, const __nv_bfloat16* w,
                        __nv_bfloat16* dst, int64_t dst_stride, int n) {
    for (int row = c.job; row < M; row += n) {
      ds_row_phase<EAGER>(p, row, part, nsplit, res, res_stride, w, dst + (int64_t)row * dst_stride, tw, c.red_s + 8 * red_flip);
      red_flip ^= 1;
    }
  };
  auto drain_partial = [&](int l, float* part, unsigned code) {
    uint32_t v[32];
    load_acc(l, v, code);
    float* dst = part + (int64_t)c.split * DS_MAX_M * p.H + (int64_t)c.tile * G_BN + r;
    asm volatile("" : "+l"(dst));                 // keeps the 32 store addresses from being hoisted out of the layer loop
#pragma unroll
    for (int cc = 0; cc < 32; ++cc)
      if (ch * 32 + cc < M) __stcg(dst + (int64_t)(ch * 32 + cc) * p.H, __uint_as_float(v[cc]));
  };

  // the cached K / V rows of an item do not depend on this launch: layer 0's are staged now, layer l + 1's as soon as
  // the attention of layer l has released the staging area
  if (has_item && item.L > 0) ds_kv_prefetch(c.layers[0], item, kv_dim, lane, ksm, vsm);
  // the first layer's input norm: done by the down jobs (idle until the first MLP)
  if (OP == OP_DOWN) {
    rows_phase(nullptr, 0, p.x_in, p.x_stride, c.layers[0].ln_in, p.xn, p.H, c.n_d);
    publish(ds_ctr(p, 0, DS_XN_A));
  }
  for (int l = 0; l < L; ++l) {
    const DsLayer& ly = c.layers[l];
    if (tw == 0) DS_TRACE(l, 0);
    if (OP == OP_QKV) {
      uint32_t v[32];
      load_acc(l, v, 0xc00u + l);
      __nv_bfloat16* dst = p.qkv + c.tile * G_BN + r;
      asm volatile("" : "+l"(dst));
      const int qkv_w = p.Q + 2 * p.KV;
#pragma unroll
      for (int cc = 0; cc < 32; ++cc)
        if (ch * 32 + cc < M) dst[(int64_t)(ch * 32 + cc) * qkv_w] = __float2bfloat16_rn(__uint_as_float(v[cc]));
      publish(ds_ctr(p, l, DS_QKV));
      if (tw == 0) DS_TRACE(l, 2);
    }
    // ---- attention: the first 4 worker warps of every CTA, one (row, kv head) item each
    await(ds_ctr(p, l, DS_QKV), E1 * (unsigned)p.nq_tiles, 0xd00u + l);
    if (tw == 0) DS_TRACE(l, 7);
    if (has_item)
      ds_attn_item<EAGER>(p, ly, item.row, item.hk, item.L, item.slot, item.npre, item.c1[0], item.c1[1], item.s1[0],
                          item.s1[1], lane, scr, ksm, vsm,
                          (p.trace && ww == 0) ? p.trace + ((int64_t)cta * DS_MAX_LAYERS + l) * 32 : nullptr);
    if (tw == 0) DS_TRACE(l, 15);
    if (lane == 0 && ww < 4) DS_TRACE(l, 20 + ww);
    ds_workers_bar();
    if (tw == 0) DS_TRACE(l, 16);
    if (tw == 0) ds_arrive(ds_ctr(p, l, DS_ATT));
    if (tw == 0) DS_TRACE(l, 17);
    if (has_item && item.L > 0 && l + 1 < L) ds_kv_prefetch(c.layers[l + 1], item, kv_dim, lane, ksm, vsm);
    if (tw == 0) DS_TRACE(l, 8);
    if (OP == OP_O) {
      drain_partial(l, p.part_o, 0xa00u + l);
      publish(ds_ctr(p, l, DS_PO));
      if (tw == 0) DS_TRACE(l, 2);
      await(ds_ctr(p, l, DS_PO), E1 * (unsigned)c.n_o, 0xe00u + l);
      if (tw == 0) DS_TRACE(l, 3);
      rows_phase(p.part_o, p.so, l == 0 ? p.x_in : p.resid, l == 0 ? p.x_stride : (int64_t)p.H, ly.ln_post, p.xn, p.H, c.n_o);
      publish(ds_ctr(p, l, DS_XN_B));
      if (tw == 0) DS_TRACE(l, 6);
    } else if (OP == OP_GU) {
      uint32_t v[32];
      load_acc(l, v, 0xf00u + l);
      // packed gate_up rows (b2tts_gemm_pack_weight, B2TTS_PACK_GATE_UP): rows r with (r & 4) == 0 are gate features,
      // r + 4 the matching up features; activation feature = tile * 64 + (r >> 3) * 4 + (r & 3).  The two lanes of a
      // pair exchange their values and share the work: the gate lane finishes the even batch rows, the up lane the odd.
      const bool is_gate = (r & 4) == 0;
      __nv_bfloat16* dst = p.act + c.tile * 64 + (r >> 3) * 4 + (r & 3);
      asm volatile("" : "+l"(dst));
#pragma unroll
      for (int c2 = 0; c2 < 16; ++c2) {
        const float a = __uint_as_float(v[2 * c2]), b = __uint_as_float(v[2 * c2 + 1]);
        const float oa = __shfl_xor_sync(0xffffffffu, a, 4), ob = __shfl_xor_sync(0xffffffffu, b, 4);
        const int col = ch * 32 + 2 * c2 + (is_gate ? 0 : 1);
        if (col < M) {
          const float gt = rbf(is_gate ? a : ob), up = rbf(is_gate ? oa : b);
          const float o = EAGER ? rbf(gsilu_f(gt)) * up : gsilu_f(gt) * up;
          dst[(int64_t)col * p.I] = __float2bfloat16_rn(o);
        }
      }
      publish(ds_ctr(p, l, DS_ACT));
      if (tw == 0) DS_TRACE(l, 2);
    } else if (OP == OP_DOWN) {
      drain_partial(l, p.part_d, 0xb00u + l);
      publish(ds_ctr(p, l, DS_PD));
      if (tw == 0) DS_TRACE(l, 2);
      await(ds_ctr(p, l, DS_PD), E1 * (unsigned)c.n_d, 0x1000u + l);
      if (tw == 0) DS_TRACE(l, 3);
      const bool last = l + 1 == L;               // the last layer's rows get the final norm and leave the kernel
      rows_phase(p.part_d, p.sd, p.resid, p.H, last ? p.final_w : c.layers[l + 1].ln_in, last ? p.out : p.xn,
                 last ? p.out_stride : (int64_t)p.H, c.n_d);
      publish(ds_ctr(p, l + 1, DS_XN_A));
      if (tw == 0) DS_TRACE(l, 6);
    }
  }
}


template <bool EAGER>
__global__ void __launch_bounds__(DS_THREADS, 1)
decoder_stack_kernel(const DsParams kp, const __grid_constant__ CUtensorMap tm_xn, const __grid_constant__ CUtensorMap tm_attn,
                     const __grid_constant__ CUtensorMap tm_act) {
  extern __shared__ uint8_t smem_dyn[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~uintptr_t(1023));
  uint8_t* kv_base = smem + DS_STAGES * DS_STAGE_BYTES;
  uint8_t* scr_base = kv_base + DS_ATT_WARPS * DS_KV_BYTES;
  float* red_s = reinterpret_cast<float*>(scr_base + DS_ATT_WARPS * DS_SCR_BYTES);      // [2][8]
  uint64_t* wfull = reinterpret_cast<uint64_t*>(red_s + 16);
  uint64_t* xfull = wfull + DS_STAGES;
  uint64_t* empty = xfull + DS_STAGES;
  uint64_t* accfull = empty + DS_STAGES;
  uint64_t* accempty = accfull + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accempty + 1);
  // the per-layer pointer table, copied out of param space once: indexing p.layer[] with a run-time layer would make
  // the compiler keep a copy in LOCAL memory, and every acquire / release invalidates L1 (each pointer read = L2 trip)
  DsParams* ps = reinterpret_cast<DsParams*>(tmem_slot + 4);
  DsLayer* layers_s = ps->layer;
  const DsParams& p = *ps;                          // the shared-memory copy (valid after the first __syncthreads)

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const int G = gridDim.x;
  const int cta = blockIdx.x;
  const int L = kp.num_layers;

  // ---- this CTA's job (the same in every layer)
  int op = OP_NONE, tile = 0, split = 0, S = 1, kblocks = 0, job = 0;
  {
    int c = cta;
    if (c < kp.nq_tiles) { op = OP_QKV; tile = c; kblocks = kp.H / G_BK; }
    else if ((c -= kp.nq_tiles) < kp.no_tiles * kp.so) { op = OP_O; S = kp.so; tile = c / S; split = c - tile * S; kblocks = kp.Q / G_BK; }
    else if ((c -= kp.no_tiles * kp.so) < kp.ng_tiles) { op = OP_GU; tile = c; kblocks = kp.H / G_BK; }
    else if ((c -= kp.ng_tiles) < kp.nd_tiles * kp.sd) { op = OP_DOWN; S = kp.sd; tile = c / S; split = c - tile * S; kblocks = kp.I / G_BK; }
    job = c;
  }
  const int kb0 = (split * kblocks) / S, kb1 = ((split + 1) * kblocks) / S;
  const int n_o = kp.no_tiles * kp.so, n_d = kp.nd_tiles * kp.sd;

#pragma unroll
  for (int i = 0; i < DS_MAX_LAYERS; ++i)
    if ((int)threadIdx.x == 32 + i) layers_s[i] = kp.layer[i];
  if (threadIdx.x == 64) static_cast<DsCore&>(*ps) = static_cast<const DsCore&>(kp);
  if (threadIdx.x == 0) {
    for (int s = 0; s < DS_STAGES; ++s) { mbar_init(wfull + s, 1); mbar_init(xfull + s, 1); mbar_init(empty + s, 1); }
    mbar_init(accfull, 1);
    mbar_init(accempty, DS_WORKERS / 32);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, (uint32_t)kp.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const unsigned E = __ldcg(kp.sync);               // launch epoch: counter targets are (E + 1) * arrivals
  const unsigned E1 = E + 1u;
  if (threadIdx.x == 0) DS_TRACE(0, 13);

  if (warp == 0) {
    // =========================== W producer: runs ahead through all layers, bounded only by the ring
    if (op != OP_NONE) {
      uint32_t slot = 0, ph = 1;
      for (int l = 0; l < L; ++l) {
        const DsLayer& ly = layers_s[l];
        const uint8_t* wp = op == OP_QKV ? ly.w_qkv : op == OP_O ? ly.w_o : op == OP_GU ? ly.w_gu : ly.w_down;
        const uint8_t* src = wp + ((int64_t)tile * kblocks + kb0) * G_WTILE_BYTES;
        // the k-blocks that do not fit the ring are at least pulled into L2 now, a layer ahead of their use
        if (elect_one())
          for (int kb = kb0 + DS_STAGES; kb < kb1; ++kb)
            asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;"
                         :: "l"(src + (int64_t)(kb - kb0) * G_WTILE_BYTES), "r"(G_WTILE_BYTES) : "memory");
        __syncwarp();
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
    // =========================== workers (8 warps): epilogues, split reduce + row norms, attention
    pdl_wait();
    DsCtx c;
    c.tmem_base = tmem_base; c.accfull = accfull; c.accempty = accempty; c.E1 = E1;
    c.G = G; c.cta = cta; c.L = L; c.tile = tile; c.split = split; c.job = job; c.n_o = n_o; c.n_d = n_d;
    c.layers = layers_s; c.scr_base = scr_base; c.kv_base = kv_base; c.red_s = red_s;
    switch (op) {
      case OP_QKV: ds_worker<EAGER, OP_QKV>(p, c, warp, lane); break;
      case OP_O: ds_worker<EAGER, OP_O>(p, c, warp, lane); break;
      case OP_GU: ds_worker<EAGER, OP_GU>(p, c, warp, lane); break;
      case OP_DOWN: ds_worker<EAGER, OP_DOWN>(p, c, warp, lane); break;
      default: ds_worker<EAGER, OP_NONE>(p, c, warp, lane); break;
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
  B2_REQUIRE(m * num_kv_heads <= DS_ATT_WARPS * sm_count(), B2TTS_EUNSUPPORTED,
             "%s: %d attention items exceed %d warps", what, m * num_kv_heads, DS_ATT_WARPS * sm_count());
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
  const size_t smem_bytes = (size_t)DS_STAGES * DS_STAGE_BYTES + DS_ATT_WARPS * (DS_KV_BYTES + DS_SCR_BYTES) + 64 +
                            (3 * DS_STAGES + 2) * 8 + 16 + sizeof(DsParams) + 1024;
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

// Debug aid (tools/stack_trace.py): following launches record 32 globaltimer stamps per (CTA, layer) into
// buffer[(cta * 8 + layer) * 16 + slot] (uint64).  NULL switches tracing off.
extern "C" B2TTS_API int b2tts_debug_stack_trace(void* buffer) {
  b2::g_ds_trace = static_cast<unsigned long long*>(buffer);
  return B2TTS_OK;
}

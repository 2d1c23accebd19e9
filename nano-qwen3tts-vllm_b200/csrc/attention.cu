// attention.cu - a2/a3: paged GQA attention for single-query decode and varlen causal prefill.
// Replaces flash_attn_with_kvcache / flash_attn_varlen_func as called from
// Attention.forward (layers/attention.py:73-96).
//
// HBM-bound design (algorithmic bytes per (sequence, layer): 4096*L + 8192 for 8 kv heads x 128):
//   * the cache layout [page, 256, kv_heads, 128] makes the K (or V) rows of ALL kv heads of
//     TILE consecutive tokens one contiguous TILE*kv_dim*2-byte span, so one CTA owns a token
//     range of one sequence for every kv head and streams it with 1-D bulk async copies
//     (cp.async.bulk -> SASS UBLKCP, the TMA unit) into an NSTAGE-deep shared-memory ring that
//     completes on mbarriers; a dedicated producer warp issues, kv_heads consumer warps compute;
//   * consumer warp w owns kv head w and both (GROUP) query heads that share it, so each K/V
//     byte is read from HBM once and from shared memory once;
//   * inside a warp 8 lanes share one token (16 head-dims each, two conflict-free LDS.128),
//     4 tokens in flight; scores are reduced with 3 shuffles, the online softmax is per tile;
//   * long contexts / small batches are split over `nsplit` CTAs per sequence (split-KV); the
//     fp32 partials (acc, m, l) go to a caller-provided workspace and a small combine kernel
//     folds them.  nsplit == 1 writes bf16 output directly.
// Prefill runs the same kernel with one CTA per query token (causal length = lk-lq+i+1);
// K/V rows come either from the packed projections (per-row bulk copies, any row stride) or
// from the paged cache when a block table is given (attention.py:79-80).
#include <stdlib.h>

#include "common.cuh"

namespace b2 {

// attention_prefill_tc.cu
bool prefill_tc_applicable(int max_seqlen_q, int head_dim, int num_heads, int num_kv_heads, const void* block_table);
int launch_prefill_tc(const char* what, const void* q, int64_t q_row_stride, const void* k, int64_t k_row_stride,
                      const void* v, int64_t v_row_stride, const int32_t* cu_q, const int32_t* cu_k, void* out,
                      int num_seqs, int total_q, int total_k, int max_seqlen_q, int num_heads, int num_kv_heads,
                      float scale, cudaStream_t st);

constexpr int ATT_HD = 128;
constexpr int ATT_TILE = 8;      // tokens per pipeline stage
constexpr int ATT_NSTAGE = 3;

struct AttnParams {
  const __nv_bfloat16* q;
  int64_t q_stride;
  const __nv_bfloat16* k;        // paged: cache base; packed: [total_k, kv_dim] with row stride
  const __nv_bfloat16* v;
  int64_t k_stride, v_stride;    // packed only (elements)
  const int32_t* block_table;    // NULL => packed
  int bt_stride;
  int page_size;
  const int32_t* context_lens;   // decode
  const int32_t* cu_q;           // prefill
  const int32_t* cu_k;
  int num_seqs;
  __nv_bfloat16* out;            // [rows, heads*128]
  float* part_acc;               // [rows, nsplit, heads, 128]
  float* part_ml;                // [rows, nsplit, heads, 2]
  int nsplit;
  int num_heads;
  int num_kv_heads;
  float scale_log2;              // softmax scale * log2(e)
};

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__device__ __forceinline__ void unpack8(const uint4& r, float* f) {
  f[0] = bf16lo(r.x); f[1] = bf16hi(r.x);
  f[2] = bf16lo(r.y); f[3] = bf16hi(r.y);
  f[4] = bf16lo(r.z); f[5] = bf16hi(r.z);
  f[6] = bf16lo(r.w); f[7] = bf16hi(r.w);
}

template <int GROUP, bool PREFILL>
__global__ void __launch_bounds__(288, 2)
paged_attn_kernel(const AttnParams p) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  const int kvh = p.num_kv_heads;
  const int kv_dim = kvh * ATT_HD;
  const int stage_elems = ATT_TILE * kv_dim;
  __nv_bfloat16* Ks = reinterpret_cast<__nv_bfloat16*>(smem_raw);
  __nv_bfloat16* Vs = Ks + ATT_NSTAGE * stage_elems;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(Vs + ATT_NSTAGE * stage_elems);
  uint64_t* empty_bar = full_bar + ATT_NSTAGE;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x;
  const int split = blockIdx.y;

  // ---- which sequence, how many keys this query sees
  int seq, L, k0 = 0;
  if (PREFILL) {
    int lo = 0, hi = p.num_seqs - 1;               // smallest s with cu_q[s+1] > row
    while (lo < hi) {
      int mid = (lo + hi) >> 1;
      if (__ldg(p.cu_q + mid + 1) > row) hi = mid; else lo = mid + 1;
    }
    seq = lo;
    const int q0 = __ldg(p.cu_q + seq), q1 = __ldg(p.cu_q + seq + 1);
    k0 = __ldg(p.cu_k + seq);
    const int lk = __ldg(p.cu_k + seq + 1) - k0;
    L = lk - (q1 - q0) + (row - q0) + 1;           // bottom-right aligned causal mask
    if (L > lk) L = lk;
  } else {
    seq = row;
    L = __ldg(p.context_lens + row);
  }
  if (L < 0) L = 0;
  const int n_tiles = (L + ATT_TILE - 1) / ATT_TILE;
  const int t_begin = (int)(((int64_t)split * n_tiles) / p.nsplit);
  const int t_end = (int)(((int64_t)(split + 1) * n_tiles) / p.nsplit);
  const int my_tiles = t_end - t_begin;

  if (threadIdx.x == 0) {
    for (int s = 0; s < ATT_NSTAGE; ++s) {
      mbar_init(full_bar + s, 1);
      mbar_init(empty_bar + s, kvh);
    }
    mbar_fence_init();
  }
  __syncthreads();

  if (warp == kvh) {
    // =========================== producer warp: bulk async copies into the ring
    const bool paged = p.block_table != nullptr;
    const bool contiguous = paged || (p.k_stride == kv_dim && p.v_stride == kv_dim);
    for (int i = 0; i < my_tiles; ++i) {
      const int stage = i % ATT_NSTAGE;
      const uint32_t ph = (i / ATT_NSTAGE) & 1;
      mbar_wait(empty_bar + stage, ph ^ 1);
      const int tok0 = (t_begin + i) * ATT_TILE;
      const int ntok = min(ATT_TILE, L - tok0);
      const uint32_t row_bytes = kv_dim * 2;
      const __nv_bfloat16 *ksrc, *vsrc;
      int64_t ks = kv_dim, vs = kv_dim;
      if (paged) {
        int page = __ldg(p.block_table + (int64_t)seq * p.bt_stride + tok0 / p.page_size);
        if (page < 0) page = 0;                    // never faults on a stale (-1) table entry
        const int64_t slot = (int64_t)page * p.page_size + tok0 % p.page_size;
        ksrc = p.k + slot * kv_dim;
        vsrc = p.v + slot * kv_dim;
      } else {
        ks = p.k_stride; vs = p.v_stride;
        ksrc = p.k + (int64_t)(k0 + tok0) * ks;
        vsrc = p.v + (int64_t)(k0 + tok0) * vs;
      }
      __nv_bfloat16* kd = Ks + stage * stage_elems;
      __nv_bfloat16* vd = Vs + stage * stage_elems;
      if (lane == 0) mbar_expect_tx(full_bar + stage, 2u * ntok * row_bytes);
      __syncwarp();
      if (contiguous) {
        if (lane == 0) {
          bulk_g2s(kd, ksrc, ntok * row_bytes, full_bar + stage);
          bulk_g2s(vd, vsrc, ntok * row_bytes, full_bar + stage);
        }
      } else if (lane < ntok) {
        bulk_g2s(kd + lane * kv_dim, ksrc + lane * ks, row_bytes, full_bar + stage);
        bulk_g2s(vd + lane * kv_dim, vsrc + lane * vs, row_bytes, full_bar + stage);
      }
    }
    return;
  }

  // =========================== consumer warp `warp` owns kv head `warp`
  const int g = lane >> 3;        // token slot within a 4-token step
  const int j = lane & 7;         // which 16 head-dims: [8j,8j+8) and [64+8j,64+8j+8)
  const int head0 = warp * GROUP;

  if (L == 0) {                   // CUDA-graph padding row (talker_mode_runner.py:97-100): zeros
    if (split == 0 && g == 0) {
#pragma unroll
      for (int h = 0; h < GROUP; ++h) {
        __nv_bfloat16* o = p.out + ((int64_t)row * p.num_heads + head0 + h) * ATT_HD;
        stg_v4(o + 8 * j, make_uint4(0, 0, 0, 0));
        stg_v4(o + 64 + 8 * j, make_uint4(0, 0, 0, 0));
      }
    }
    if (p.nsplit > 1 && lane < GROUP) {
      float* ml = p.part_ml + (((int64_t)row * p.nsplit + split) * p.num_heads + head0 + lane) * 2;
      ml[0] = -INFINITY; ml[1] = 0.f;
    }
    return;
  }

  float qv[GROUP][16];
#pragma unroll
  for (int h = 0; h < GROUP; ++h) {
    const __nv_bfloat16* qp = p.q + (int64_t)row * p.q_stride + (head0 + h) * ATT_HD;
    unpack8(ldg_v4(qp + 8 * j), qv[h]);
    unpack8(ldg_v4(qp + 64 + 8 * j), qv[h] + 8);
#pragma unroll
    for (int d = 0; d < 16; ++d) qv[h][d] *= p.scale_log2;
  }

  float m[GROUP], l[GROUP], acc[GROUP][16];
#pragma unroll
  for (int h = 0; h < GROUP; ++h) {
    m[h] = -INFINITY; l[h] = 0.f;
#pragma unroll
    for (int d = 0; d < 16; ++d) acc[h][d] = 0.f;
  }

  for (int i = 0; i < my_tiles; ++i) {
    const int stage = i % ATT_NSTAGE;
    const uint32_t ph = (i / ATT_NSTAGE) & 1;
    const int ntok = min(ATT_TILE, L - (t_begin + i) * ATT_TILE);
    mbar_wait(full_bar + stage, ph);
    const __nv_bfloat16* kt = Ks + stage * stage_elems + warp * ATT_HD;
    const __nv_bfloat16* vt = Vs + stage * stage_elems + warp * ATT_HD;

    float s[ATT_TILE / 4][GROUP];
    int tokc[ATT_TILE / 4];
#pragma unroll
    for (int it = 0; it < ATT_TILE / 4; ++it) {
      const int tok = it * 4 + g;
      const bool valid = tok < ntok;
      tokc[it] = valid ? tok : 0;                  // token 0 of a tile is always loaded
      const __nv_bfloat16* kr = kt + tokc[it] * kv_dim;
      float kf[16];
      unpack8(*reinterpret_cast<const uint4*>(kr + 8 * j), kf);
      unpack8(*reinterpret_cast<const uint4*>(kr + 64 + 8 * j), kf + 8);
#pragma unroll
      for (int h = 0; h < GROUP; ++h) {
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
        s[it][h] = valid ? dot : -INFINITY;
      }
    }
    float alpha[GROUP];
#pragma unroll
    for (int h = 0; h < GROUP; ++h) {
      float mt = s[0][h];
#pragma unroll
      for (int it = 1; it < ATT_TILE / 4; ++it) mt = fmaxf(mt, s[it][h]);
      mt = fmaxf(mt, __shfl_xor_sync(0xffffffffu, mt, 8));
      mt = fmaxf(mt, __shfl_xor_sync(0xffffffffu, mt, 16));
      const float mn = fmaxf(m[h], mt);            // finite: every processed tile has a valid token
      alpha[h] = fast_exp2(m[h] - mn);
      m[h] = mn;
      float ps = 0.f;
#pragma unroll
      for (int it = 0; it < ATT_TILE / 4; ++it) {
        s[it][h] = fast_exp2(s[it][h] - mn);       // s now holds p
        ps += s[it][h];
      }
      l[h] = fmaf(l[h], alpha[h], ps);
#pragma unroll
      for (int d = 0; d < 16; ++d) acc[h][d] *= alpha[h];
    }
#pragma unroll
    for (int it = 0; it < ATT_TILE / 4; ++it) {
      const __nv_bfloat16* vr = vt + tokc[it] * kv_dim;
      float vf[16];
      unpack8(*reinterpret_cast<const uint4*>(vr + 8 * j), vf);
      unpack8(*reinterpret_cast<const uint4*>(vr + 64 + 8 * j), vf + 8);
#pragma unroll
      for (int h = 0; h < GROUP; ++h)
#pragma unroll
        for (int d = 0; d < 16; ++d) acc[h][d] = fmaf(s[it][h], vf[d], acc[h][d]);
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(empty_bar + stage);
  }

  // ---- fold the 4 token slots of the warp
#pragma unroll
  for (int h = 0; h < GROUP; ++h) {
    l[h] += __shfl_xor_sync(0xffffffffu, l[h], 8);
    l[h] += __shfl_xor_sync(0xffffffffu, l[h], 16);
#pragma unroll
    for (int d = 0; d < 16; ++d) {
      acc[h][d] += __shfl_xor_sync(0xffffffffu, acc[h][d], 8);
      acc[h][d] += __shfl_xor_sync(0xffffffffu, acc[h][d], 16);
    }
  }

  if (p.nsplit == 1) {
    if (g == 0) {
#pragma unroll
      for (int h = 0; h < GROUP; ++h) {
        const float inv = 1.f / l[h];
        __nv_bfloat16* o = p.out + ((int64_t)row * p.num_heads + head0 + h) * ATT_HD;
        uint4 a, b;
        a.x = pack_bf16(acc[h][0] * inv, acc[h][1] * inv);   a.y = pack_bf16(acc[h][2] * inv, acc[h][3] * inv);
        a.z = pack_bf16(acc[h][4] * inv, acc[h][5] * inv);   a.w = pack_bf16(acc[h][6] * inv, acc[h][7] * inv);
        b.x = pack_bf16(acc[h][8] * inv, acc[h][9] * inv);   b.y = pack_bf16(acc[h][10] * inv, acc[h][11] * inv);
        b.z = pack_bf16(acc[h][12] * inv, acc[h][13] * inv); b.w = pack_bf16(acc[h][14] * inv, acc[h][15] * inv);
        stg_v4(o + 8 * j, a);
        stg_v4(o + 64 + 8 * j, b);
      }
    }
  } else {
    if (g == 0) {
#pragma unroll
      for (int h = 0; h < GROUP; ++h) {
        const int64_t idx = ((int64_t)row * p.nsplit + split) * p.num_heads + head0 + h;
        if (my_tiles > 0) {
          float* pa = p.part_acc + idx * ATT_HD;
          *reinterpret_cast<float4*>(pa + 8 * j) = make_float4(acc[h][0], acc[h][1], acc[h][2], acc[h][3]);
          *reinterpret_cast<float4*>(pa + 8 * j + 4) = make_float4(acc[h][4], acc[h][5], acc[h][6], acc[h][7]);
          *reinterpret_cast<float4*>(pa + 64 + 8 * j) = make_float4(acc[h][8], acc[h][9], acc[h][10], acc[h][11]);
          *reinterpret_cast<float4*>(pa + 64 + 8 * j + 4) = make_float4(acc[h][12], acc[h][13], acc[h][14], acc[h][15]);
        }
        if (j == 0) {
          p.part_ml[idx * 2] = my_tiles > 0 ? m[h] : -INFINITY;
          p.part_ml[idx * 2 + 1] = my_tiles > 0 ? l[h] : 0.f;
        }
      }
    }
  }
}

// Fold the split-KV partials of one (row, head): 128 threads, one head-dim each.
__global__ void __launch_bounds__(128)
attn_combine_kernel(const float* __restrict__ part_acc, const float* __restrict__ part_ml,
                    __nv_bfloat16* __restrict__ out, const int32_t* __restrict__ context_lens,
                    int nsplit, int num_heads) {
  const int row = blockIdx.x / num_heads, head = blockIdx.x % num_heads, d = threadIdx.x;
  if (context_lens && __ldg(context_lens + row) <= 0) return;   // zeros already written
  const int64_t base = ((int64_t)row * nsplit) * num_heads + head;
  float M = -INFINITY;
  for (int s = 0; s < nsplit; ++s) M = fmaxf(M, __ldg(part_ml + (base + (int64_t)s * num_heads) * 2));
  float num = 0.f, den = 0.f;
  for (int s = 0; s < nsplit; ++s) {
    const int64_t idx = base + (int64_t)s * num_heads;
    const float ls = __ldg(part_ml + idx * 2 + 1);
    if (ls > 0.f) {
      const float w = fast_exp2(__ldg(part_ml + idx * 2) - M);
      num = fmaf(w, __ldg(part_acc + idx * ATT_HD + d), num);
      den = fmaf(w, ls, den);
    }
  }
  out[((int64_t)row * num_heads + head) * ATT_HD + d] = __float2bfloat16_rn(den > 0.f ? num / den : 0.f);
}

static size_t attn_smem_bytes(int kv_heads) {
  return (size_t)2 * ATT_NSTAGE * ATT_TILE * kv_heads * ATT_HD * 2 + 2 * ATT_NSTAGE * sizeof(uint64_t);
}

// split-KV factor: fill 2 CTAs/SM for two waves when the batch alone cannot, but keep >= 4
// tiles per split so the ring has something to pipeline.
static int choose_nsplit(int rows, int max_ctx) {
  const int max_tiles = (max_ctx + ATT_TILE - 1) / ATT_TILE;
  const int target = 4 * sm_count();
  int n = (target + rows - 1) / rows;
  const int cap = (max_tiles + 3) / 4;
  if (n > cap) n = cap;
  if (n < 1) n = 1;
  if (n > 64) n = 64;
  return n;
}

template <bool PREFILL>
static int launch_attn(const AttnParams& p, int rows, cudaStream_t stream, const char* what) {
  const int group = p.num_heads / p.num_kv_heads;
  const size_t smem = attn_smem_bytes(p.num_kv_heads);
  dim3 grid(rows, p.nsplit), block(32 * (p.num_kv_heads + 1));
#define B2_LAUNCH_ATTN(G)                                                                             \
  do {                                                                                                \
    static bool configured = false;                                                                   \
    if (!configured) {                                                                                \
      cudaError_t e = cudaFuncSetAttribute(paged_attn_kernel<G, PREFILL>,                             \
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024); \
      if (e != cudaSuccess) return cuda_fail(e, what);                                                \
      configured = true;                                                                              \
    }                                                                                                 \
    paged_attn_kernel<G, PREFILL><<<grid, block, smem, stream>>>(p);                                  \
  } while (0)
  switch (group) {
    case 1: B2_LAUNCH_ATTN(1); break;
    case 2: B2_LAUNCH_ATTN(2); break;
    case 4: B2_LAUNCH_ATTN(4); break;
    default:
      set_error("%s: GQA group %d unsupported (1, 2, 4)", what, group);
      return B2TTS_EUNSUPPORTED;
  }
#undef B2_LAUNCH_ATTN
  B2_CHECK_LAUNCH(what);
  return B2TTS_OK;
}

static int check_heads(const char* what, int num_heads, int num_kv_heads, int head_dim) {
  B2_REQUIRE(head_dim == ATT_HD, B2TTS_EUNSUPPORTED, "%s: head_dim %d unsupported (128)", what, head_dim);
  B2_REQUIRE(num_kv_heads >= 1 && num_kv_heads <= 8 && num_heads % num_kv_heads == 0, B2TTS_EUNSUPPORTED,
             "%s: heads %d / kv heads %d unsupported (kv heads 1..8 dividing heads)", what, num_heads, num_kv_heads);
  return B2TTS_OK;
}

}  // namespace b2

extern "C" size_t b2tts_paged_attn_decode_workspace_bytes(int batch, int num_heads, int head_dim, int max_ctx) {
  if (batch <= 0 || num_heads <= 0 || head_dim <= 0 || max_ctx <= 0) return 0;
  const int nsplit = b2::choose_nsplit(batch, max_ctx);
  if (nsplit == 1) return 0;
  return (size_t)batch * nsplit * num_heads * (head_dim + 2) * sizeof(float);
}

extern "C" int b2tts_paged_attn_decode(const void* q, int64_t q_row_stride,
                                       const void* k_cache, const void* v_cache,
                                       const int32_t* block_table, int bt_stride,
                                       const int32_t* context_lens,
                                       void* out, void* workspace, size_t workspace_bytes,
                                       int batch, int num_heads, int num_kv_heads, int head_dim,
                                       int page_size, float scale, int max_ctx, void* stream) {
  using namespace b2;
  const char* what = "paged_attn_decode";
  B2_REQUIRE(batch >= 0 && max_ctx > 0 && page_size > 0 && bt_stride > 0, B2TTS_EINVAL,
             "%s: batch=%d max_ctx=%d page_size=%d bt_stride=%d", what, batch, max_ctx, page_size, bt_stride);
  if (batch == 0) return B2TTS_OK;
  B2_REQUIRE(q && k_cache && v_cache && block_table && context_lens && out, B2TTS_EINVAL, "%s: null pointer", what);
  int rc = check_heads(what, num_heads, num_kv_heads, head_dim);
  if (rc) return rc;
  B2_REQUIRE(page_size % ATT_TILE == 0, B2TTS_EUNSUPPORTED, "%s: page_size %d must be a multiple of %d", what,
             page_size, ATT_TILE);
  B2_REQUIRE(aligned16(q) && aligned16(k_cache) && aligned16(v_cache) && aligned16(out) && q_row_stride % 8 == 0,
             B2TTS_EALIGN, "%s: pointers / q stride must be 16-byte aligned", what);
  AttnParams p{};
  p.q = static_cast<const __nv_bfloat16*>(q);
  p.q_stride = q_row_stride;
  p.k = static_cast<const __nv_bfloat16*>(k_cache);
  p.v = static_cast<const __nv_bfloat16*>(v_cache);
  p.block_table = block_table;
  p.bt_stride = bt_stride;
  p.page_size = page_size;
  p.context_lens = context_lens;
  p.out = static_cast<__nv_bfloat16*>(out);
  p.nsplit = choose_nsplit(batch, max_ctx);
  p.num_heads = num_heads;
  p.num_kv_heads = num_kv_heads;
  p.scale_log2 = scale * 1.4426950408889634f;
  if (p.nsplit > 1) {
    const size_t need = (size_t)batch * p.nsplit * num_heads * (head_dim + 2) * sizeof(float);
    B2_REQUIRE(workspace && workspace_bytes >= need && aligned16(workspace), B2TTS_EWORKSPACE,
               "%s: workspace %zu B < %zu B needed (or unaligned)", what, workspace_bytes, need);
    p.part_acc = static_cast<float*>(workspace);
    p.part_ml = p.part_acc + (size_t)batch * p.nsplit * num_heads * head_dim;
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  rc = launch_attn<false>(p, batch, st, what);
  if (rc) return rc;
  if (p.nsplit > 1) {
    attn_combine_kernel<<<batch * num_heads, 128, 0, st>>>(p.part_acc, p.part_ml, p.out, context_lens,
                                                                 p.nsplit, num_heads);
    B2_CHECK_LAUNCH("paged_attn_decode(combine)");
  }
  return B2TTS_OK;
}

extern "C" int b2tts_varlen_attn_prefill(const void* q, int64_t q_row_stride,
                                         const void* k, int64_t k_row_stride,
                                         const void* v, int64_t v_row_stride,
                                         const int32_t* cu_seqlens_q, const int32_t* cu_seqlens_k,
                                         void* out, int num_seqs, int total_q, int total_k,
                                         int max_seqlen_q, int max_seqlen_k,
                                         int num_heads, int num_kv_heads, int head_dim, float scale,
                                         const int32_t* block_table, int bt_stride, int page_size,
                                         void* stream) {
  using namespace b2;
  const char* what = "varlen_attn_prefill";
  (void)max_seqlen_k;
  B2_REQUIRE(num_seqs >= 0 && total_q >= 0 && total_k >= 0, B2TTS_EINVAL, "%s: num_seqs=%d total_q=%d total_k=%d", what,
             num_seqs, total_q, total_k);
  if (num_seqs == 0 || total_q == 0) return B2TTS_OK;
  B2_REQUIRE(q && k && v && cu_seqlens_q && cu_seqlens_k && out, B2TTS_EINVAL, "%s: null pointer", what);
  int rc = check_heads(what, num_heads, num_kv_heads, head_dim);
  if (rc) return rc;
  B2_REQUIRE(aligned16(q) && aligned16(k) && aligned16(v) && aligned16(out) && q_row_stride % 8 == 0 &&
                 k_row_stride % 8 == 0 && v_row_stride % 8 == 0,
             B2TTS_EALIGN, "%s: pointers / strides must be 16-byte aligned", what);
  if (block_table)
    B2_REQUIRE(page_size > 0 && page_size % ATT_TILE == 0 && bt_stride > 0, B2TTS_EUNSUPPORTED,
               "%s: page_size %d must be a positive multiple of %d", what, page_size, ATT_TILE);
  // long prompts: the tcgen05 tile kernel (attention_prefill_tc.cu); short ones: one CTA per query token below
  static const bool tc_enabled = [] { const char* e = getenv("B2TTS_PREFILL_TC"); return !(e && e[0] == '0'); }();
  if (tc_enabled && max_seqlen_q > 0 && prefill_tc_applicable(max_seqlen_q, head_dim, num_heads, num_kv_heads, block_table))
    return launch_prefill_tc(what, q, q_row_stride, k, k_row_stride, v, v_row_stride, cu_seqlens_q, cu_seqlens_k, out,
                             num_seqs, total_q, total_k, max_seqlen_q, num_heads, num_kv_heads, scale,
                             static_cast<cudaStream_t>(stream));
  AttnParams p{};
  p.q = static_cast<const __nv_bfloat16*>(q);
  p.q_stride = q_row_stride;
  p.k = static_cast<const __nv_bfloat16*>(k);
  p.v = static_cast<const __nv_bfloat16*>(v);
  p.k_stride = k_row_stride;
  p.v_stride = v_row_stride;
  p.block_table = block_table;
  p.bt_stride = bt_stride;
  p.page_size = block_table ? page_size : ATT_TILE;
  p.cu_q = cu_seqlens_q;
  p.cu_k = cu_seqlens_k;
  p.num_seqs = num_seqs;
  p.out = static_cast<__nv_bfloat16*>(out);
  p.nsplit = 1;
  p.num_heads = num_heads;
  p.num_kv_heads = num_kv_heads;
  p.scale_log2 = scale * 1.4426950408889634f;
  return launch_attn<true>(p, total_q, static_cast<cudaStream_t>(stream), what);
}

// ---------------------------------------------------------------------------------------------------------------------
// b2tts_paged_attn_decode_static: the same launch(es) as b2tts_paged_attn_decode with the split-KV partials in a
// file-scope device array instead of a caller-provided workspace, for callers that cannot allocate (code under CUDA-graph
// capture, serving loops and harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Every GPU that loads the library holds the 8 MiB until the process exits.  All calls share
// the array: two calls that can overlap on the GPU (different streams, or graph branches) race on it.  Nothing in it has
// to start at zero - every (row, split, head) partial, empty splits included, is written before the combine reads it.
namespace b2 {
constexpr size_t ATT_STATIC_SCRATCH_BYTES = 8u << 20;
static __device__ __align__(256) unsigned char g_attn_static_scratch[ATT_STATIC_SCRATCH_BYTES];
constexpr int ATT_STATIC_MAX_DEVICES = 64;
static void* g_attn_static_ptr[ATT_STATIC_MAX_DEVICES];       // device address of the array, per device (host code only
                                                               // knows the symbol: the address is asked for once per device)
}  // namespace b2

extern "C" size_t b2tts_paged_attn_decode_static_max_bytes(void) { return b2::ATT_STATIC_SCRATCH_BYTES; }

extern "C" int b2tts_paged_attn_decode_static(const void* q, int64_t q_row_stride,
                                              const void* k_cache, const void* v_cache,
                                              const int32_t* block_table, int bt_stride,
                                              const int32_t* context_lens, void* out,
                                              int batch, int num_heads, int num_kv_heads, int head_dim,
                                              int page_size, float scale, int max_ctx, void* stream) {
  using namespace b2;
  const char* what = "paged_attn_decode_static";
  const size_t need = b2tts_paged_attn_decode_workspace_bytes(batch, num_heads, head_dim, max_ctx);
  B2_REQUIRE(need <= ATT_STATIC_SCRATCH_BYTES, B2TTS_EWORKSPACE,
             "%s: batch=%d heads=%d max_ctx=%d needs %zu B of split-KV scratch, the built-in array holds %zu B; "
             "use b2tts_paged_attn_decode with a workspace of your own", what, batch, num_heads, max_ctx, need,
             (size_t)ATT_STATIC_SCRATCH_BYTES);
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return cuda_fail(e, what);
  B2_REQUIRE(dev >= 0 && dev < ATT_STATIC_MAX_DEVICES, B2TTS_EUNSUPPORTED, "%s: device ordinal %d", what, dev);
  void* scratch = g_attn_static_ptr[dev];
  if (!scratch) {                                 // set-up, once per device, with that device current
    e = cudaGetSymbolAddress(&scratch, g_attn_static_scratch);
    if (e != cudaSuccess) return cuda_fail(e, "paged_attn_decode_static(cudaGetSymbolAddress)");
    B2_REQUIRE(scratch && aligned16(scratch), B2TTS_ECUDA, "%s: scratch symbol has no usable device address", what);
    g_attn_static_ptr[dev] = scratch;
  }
  return b2tts_paged_attn_decode(q, q_row_stride, k_cache, v_cache, block_table, bt_stride, context_lens, out, scratch,
                                 ATT_STATIC_SCRATCH_BYTES, batch, num_heads, num_kv_heads, head_dim, page_size, scale,
                                 max_ctx, stream);
}

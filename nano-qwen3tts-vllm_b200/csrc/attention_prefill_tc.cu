// attention_prefill_tc.cu - a3 at its real size: varlen causal prefill attention on the 5th-gen tensor cores.
// Replaces flash_attn_varlen_func(q, k, v, cu_seqlens_q/k, max_seqlen_q/k, softmax_scale, causal=True) as called from
// Attention.forward's prefill branch (layers/attention.py:78-91) when the prompt is long enough for the work to be a
// real contraction - the voice-clone ICL prompt of BASELINE configs[3] is ~4096 tokens: 2*T^2*128*16 / 2 = 69 GFLOP
// per layer, tensor-bound.  (Short prompts and the paged prefix-cache variant stay on paged_attn_kernel<G,true>.)
//
// One CTA = 128 queries of one (sequence, q head); K/V are walked in 128-key tiles up to the causal diagonal:
//   S = Q K^T      tcgen05.mma kind::f16, M = 128 queries, N = 128 keys, K = 128 dims: A = Q and B = K tile, both
//                  K-major SWIZZLE_128B shared-memory operands, each a pair of [128 rows x 64 dims] boxes written by
//                  the TMA unit (cp.async.bulk.tensor.2d) straight from the packed projections (any row stride);
//                  the fp32 scores land in TMEM (two 128-column buffers, so S of tile j+1 is computed while the
//                  softmax of tile j runs);
//   softmax        4 warps, thread r owns query row r = TMEM lane r: tcgen05.ld 32 columns at a time, running max
//                  and sum in the exp2 domain (scale * log2 e folded in), causal / length mask only on the tiles
//                  that straddle the diagonal or the sequence end, P = exp2(S - m) rounded to bf16 and stored to
//                  shared memory in the K-major SWIZZLE_128B image of an MMA A operand;
//   O += P V       tcgen05.mma, A = P (shared), B = the V tile exactly as TMA delivered it: [128 keys x 64 dims] boxes
//                  are the MN-major SWIZZLE_128B canonical layout (keys = K, dims = N; LBO = 16 KiB between the two
//                  64-dim halves, SBO = 1 KiB between 8-key groups), so V needs no transpose;  each tile's P V goes to
//                  its own TMEM buffer and the softmax threads fold it into the fp32 running output they keep in
//                  registers (O = O * alpha + PV): the accumulator never needs a read-modify-write through TMEM.
// Warp roles: warp 0 TMA producer (Q once, then a 2-stage ring of K/V tiles), warp 1 MMA issuer + TMEM owner,
// warps 2-5 softmax / epilogue.  Everything is signalled through mbarriers (tcgen05.commit for the MMA side).
// Causal tile skipping: a q tile only visits the key tiles at or below its diagonal (bottom-right aligned when
// lk > lq); CTAs are ordered heaviest-first.
#include <cuda.h>

#include "common.cuh"

namespace b2 {

constexpr int PF_BM = 128;                 // queries per CTA (UMMA M)
constexpr int PF_BN = 128;                 // keys per tile (UMMA N of S, K extent of P V)
constexpr int PF_HD = 128;
constexpr int PF_STAGES = 2;
constexpr int PF_THREADS = 192;
constexpr int PF_BOX_BYTES = 128 * 64 * 2; // one [128 rows x 64 columns] bf16 box = 16 KiB
constexpr int PF_TILE_BYTES = 2 * PF_BOX_BYTES;
constexpr int PF_MIN_SEQLEN = 128;         // below this the per-query kernel wins (launcher's choice)

struct PrefillTcParams {
  const int32_t* cu_q;
  const int32_t* cu_k;
  __nv_bfloat16* out;                      // [total_q, heads * 128]
  int num_heads, num_kv_heads;
  float scale_log2;
};

// ---- tcgen05 / TMA PTX (same forms as gemm.cu)
__device__ __forceinline__ void pf_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void pf_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void pf_fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ bool pf_elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void pf_tma_2d(void* smem_dst, const CUtensorMap* tmap, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      :: "r"(smem_u32(smem_dst)), "l"(tmap), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
// K-major SWIZZLE_128B operand (rows of 128 bytes, 8-row groups 1 KiB apart)
__device__ __forceinline__ uint64_t pf_desc_kmajor(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3ffffu) >> 4);
  d |= (uint64_t)1 << 16;                                // LBO unused for swizzled K-major
  d |= (uint64_t)(1024 >> 4) << 32;                      // SBO: next 8-row group
  d |= (uint64_t)1 << 46;                                // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;                                // SWIZZLE_128B
  return d;
}
// MN-major SWIZZLE_128B operand: each K index (key) is a 128-byte row of 64 contiguous N elements (dims);
// LBO = distance to the next 64 N elements (the second box), SBO = distance between groups of 8 K rows.
__device__ __forceinline__ uint64_t pf_desc_mnmajor(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3ffffu) >> 4);
  d |= (uint64_t)(PF_BOX_BYTES >> 4) << 16;              // LBO: 16 KiB
  d |= (uint64_t)(1024 >> 4) << 32;                      // SBO: 1 KiB
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// kind::f16 instruction descriptor: bf16 x bf16 -> fp32, M = 128, N = 128; b_mn = 1: B operand is MN-major
__device__ __forceinline__ uint32_t pf_idesc(int b_mn) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)b_mn << 16) | ((uint32_t)(PF_BN >> 3) << 17) |
         ((uint32_t)(PF_BM >> 4) << 24);
}
__device__ __forceinline__ void pf_mma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}"
      :: "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void pf_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void pf_tmem_ld32(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ float pf_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__global__ void __launch_bounds__(PF_THREADS, 1)
prefill_attn_tc_kernel(const PrefillTcParams p, const __grid_constant__ CUtensorMap tmap_q,
                       const __grid_constant__ CUtensorMap tmap_k, const __grid_constant__ CUtensorMap tmap_v) {
  extern __shared__ uint8_t pf_smem_dyn[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(pf_smem_dyn) + 1023) & ~uintptr_t(1023));
  uint8_t* q_s = smem;                                             // 2 boxes: dims 0-63 | 64-127
  uint8_t* k_s = q_s + PF_TILE_BYTES;                              // [stage][2 boxes]
  uint8_t* v_s = k_s + PF_STAGES * PF_TILE_BYTES;                  // [stage][2 boxes]
  uint8_t* p_s = v_s + PF_STAGES * PF_TILE_BYTES;                  // 2 boxes: keys 0-63 | 64-127
  uint64_t* bars = reinterpret_cast<uint64_t*>(p_s + PF_TILE_BYTES);
  uint64_t* q_full = bars;
  uint64_t* kv_full = bars + 1;                                    // [PF_STAGES]
  uint64_t* kv_empty = kv_full + PF_STAGES;
  uint64_t* s_full = kv_empty + PF_STAGES;                         // [2]
  uint64_t* s_empty = s_full + 2;
  uint64_t* p_full = s_empty + 2;
  uint64_t* pv_full = p_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(pv_full + 1);

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const int seq = blockIdx.z;
  const int head = blockIdx.y;
  const int tile = (int)gridDim.x - 1 - (int)blockIdx.x;           // heaviest (latest) q tiles first
  const int kv_head = head / (p.num_heads / p.num_kv_heads);

  const int q0 = __ldg(p.cu_q + seq), lq = __ldg(p.cu_q + seq + 1) - q0;
  const int k0 = __ldg(p.cu_k + seq), lk = __ldg(p.cu_k + seq + 1) - k0;
  if (tile * PF_BM >= lq || lk <= 0) return;                        // uniform over the CTA
  const int offset = lk - lq;                                       // bottom-right aligned causal mask
  const int rows_valid = min(PF_BM, lq - tile * PF_BM);
  // last key any valid row of this tile may see
  int last_key = tile * PF_BM + rows_valid - 1 + offset;
  if (last_key > lk - 1) last_key = lk - 1;
  const int n_tiles = last_key / PF_BN + 1;

  if (threadIdx.x == 0) {
    mbar_init(q_full, 1);
    for (int s = 0; s < PF_STAGES; ++s) { mbar_init(kv_full + s, 1); mbar_init(kv_empty + s, 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(s_full + b, 1); mbar_init(s_empty + b, 4); }
    mbar_init(p_full, 4);
    mbar_init(pv_full, 1);
    mbar_fence_init();
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(smem_u32(tmem_slot)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  pf_fence_before();
  __syncthreads();
  pf_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_s0 = tmem_base, tmem_pv = tmem_base + 256;   // S buffers at columns 0 / 128, P V at 256

  if (warp == 0) {
    // =========================== TMA producer
    if (pf_elect_one()) {
      mbar_expect_tx(q_full, PF_TILE_BYTES);
      pf_tma_2d(q_s, &tmap_q, head * PF_HD, q0 + tile * PF_BM, q_full);
      pf_tma_2d(q_s + PF_BOX_BYTES, &tmap_q, head * PF_HD + 64, q0 + tile * PF_BM, q_full);
    }
    __syncwarp();
    for (int j = 0; j < n_tiles; ++j) {
      const int stage = j % PF_STAGES;
      const uint32_t ph = (j / PF_STAGES) & 1;
      mbar_wait(kv_empty + stage, ph ^ 1);                          // P V of tile j - PF_STAGES has read the stage
      if (pf_elect_one()) {
        const int row = k0 + j * PF_BN, col = kv_head * PF_HD;
        mbar_expect_tx(kv_full + stage, 2 * PF_TILE_BYTES);
        pf_tma_2d(k_s + stage * PF_TILE_BYTES, &tmap_k, col, row, kv_full + stage);
        pf_tma_2d(k_s + stage * PF_TILE_BYTES + PF_BOX_BYTES, &tmap_k, col + 64, row, kv_full + stage);
        pf_tma_2d(v_s + stage * PF_TILE_BYTES, &tmap_v, col, row, kv_full + stage);
        pf_tma_2d(v_s + stage * PF_TILE_BYTES + PF_BOX_BYTES, &tmap_v, col + 64, row, kv_full + stage);
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    // =========================== MMA issuer: S_j = Q K_j^T, then (one tile behind) PV_{j-1} = P_{j-1} V_{j-1}
    const uint32_t idesc_s = pf_idesc(0), idesc_pv = pf_idesc(1);
    const uint32_t q_addr = smem_u32(q_s), p_addr = smem_u32(p_s);
    auto issue_pv = [&](int i) {
      const int stage = i % PF_STAGES;
      mbar_wait(p_full, (uint32_t)(i & 1));                         // P_i is in shared memory, PV buffer consumed
      pf_fence_after();
      const uint32_t v_addr = smem_u32(v_s + stage * PF_TILE_BYTES);
      if (pf_elect_one()) {
#pragma unroll
        for (int s = 0; s < PF_BN / 16; ++s)                        // 16 keys per MMA
          pf_mma(tmem_pv, pf_desc_kmajor(p_addr + (s >> 2) * PF_BOX_BYTES + (s & 3) * 32),
                 pf_desc_mnmajor(v_addr + s * 2048), idesc_pv, (uint32_t)(s != 0));
        pf_commit(pv_full);
        pf_commit(kv_empty + stage);
      }
      __syncwarp();
    };
    mbar_wait(q_full, 0);
    for (int j = 0; j < n_tiles; ++j) {
      const int stage = j % PF_STAGES, b = j & 1;
      mbar_wait(kv_full + stage, (uint32_t)((j / PF_STAGES) & 1));
      mbar_wait(s_empty + b, (uint32_t)(((j >> 1) & 1) ^ 1));       // softmax of tile j - 2 has drained this buffer
      pf_fence_after();
      const uint32_t k_addr = smem_u32(k_s + stage * PF_TILE_BYTES);
      if (pf_elect_one()) {
#pragma unroll
        for (int s = 0; s < PF_HD / 16; ++s)                        // 16 dims per MMA
          pf_mma(tmem_s0 + b * PF_BN, pf_desc_kmajor(q_addr + (s >> 2) * PF_BOX_BYTES + (s & 3) * 32),
                 pf_desc_kmajor(k_addr + (s >> 2) * PF_BOX_BYTES + (s & 3) * 32), idesc_s, (uint32_t)(s != 0));
        pf_commit(s_full + b);
      }
      __syncwarp();
      if (j > 0) issue_pv(j - 1);
    }
    issue_pv(n_tiles - 1);
  } else {
    // =========================== softmax + output: thread r owns query row r (TMEM lane r)
    const int quarter = warp & 3;                                   // TMEM lane quarter this warp may access
    const int r = quarter * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(quarter * 32) << 16;
    const int q_idx = tile * PF_BM + r;
    int kmax = q_idx + offset;                                      // last key this row may see
    if (kmax > lk - 1) kmax = lk - 1;
    float m = -INFINITY, l = 0.f, alpha_prev = 1.f;
    float o[PF_HD];
#pragma unroll
    for (int d = 0; d < PF_HD; ++d) o[d] = 0.f;
    uint8_t* p_row = p_s + r * 128;
    const int sw = r & 7;

    for (int j = 0; j < n_tiles; ++j) {
      const int b = j & 1;
      const uint32_t s_addr = tmem_s0 + b * PF_BN + lane_addr;
      const int lim = kmax - j * PF_BN;                             // columns <= lim are visible
      const bool need_mask = lim < PF_BN - 1;
      mbar_wait(s_full + b, (uint32_t)((j >> 1) & 1));
      pf_fence_after();
      // ---- pass 1: row maximum
      float mx = -INFINITY;
#pragma unroll 1
      for (int c = 0; c < PF_BN / 32; ++c) {
        uint32_t v[32];
        pf_tmem_ld32(s_addr + c * 32, v);
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          float x = __uint_as_float(v[i]);
          if (need_mask && c * 32 + i > lim) x = -INFINITY;
          mx = fmaxf(mx, x);
        }
      }
      const float m_new = fmaxf(m, mx * p.scale_log2);
      const float m_safe = m_new == -INFINITY ? 0.f : m_new;        // a row with no visible key yet
      const float alpha = pf_exp2(m - m_safe);
      // ---- fold the previous tile's P V into the running output (also frees the P buffer and the PV accumulator)
      if (j > 0) {
        mbar_wait(pv_full, (uint32_t)((j - 1) & 1));
        pf_fence_after();
#pragma unroll
        for (int c = 0; c < PF_HD / 32; ++c) {
          uint32_t v[32];
          pf_tmem_ld32(tmem_pv + lane_addr + c * 32, v);
#pragma unroll
          for (int i = 0; i < 32; ++i) o[c * 32 + i] = fmaf(o[c * 32 + i], alpha_prev, __uint_as_float(v[i]));
        }
      }
      // ---- pass 2: P = exp2(S * scale - m) -> bf16 -> shared memory (K-major SWIZZLE_128B A operand), row sum
      float rs = 0.f;
#pragma unroll 1
      for (int c = 0; c < PF_BN / 32; ++c) {
        uint32_t v[32];
        pf_tmem_ld32(s_addr + c * 32, v);
        uint32_t pk[16];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float x0 = __uint_as_float(v[i]) * p.scale_log2 - m_safe;
          float x1 = __uint_as_float(v[i + 1]) * p.scale_log2 - m_safe;
          if (need_mask && c * 32 + i > lim) x0 = -INFINITY;
          if (need_mask && c * 32 + i + 1 > lim) x1 = -INFINITY;
          const float e0 = pf_exp2(x0), e1 = pf_exp2(x1);
          pk[i >> 1] = pack_bf16(e0, e1);
          // the row sum uses the ROUNDED probabilities, the values the tensor core multiplies with V
          rs += bf16lo(pk[i >> 1]) + bf16hi(pk[i >> 1]);
        }
        uint8_t* box = p_row + (c >> 1) * PF_BOX_BYTES;
#pragma unroll
        for (int q4 = 0; q4 < 4; ++q4) {
          const int chunk = (c & 1) * 4 + q4;                       // 16-byte chunk (8 keys) within the 64-key row
          *reinterpret_cast<uint4*>(box + ((chunk ^ sw) << 4)) =
              make_uint4(pk[4 * q4], pk[4 * q4 + 1], pk[4 * q4 + 2], pk[4 * q4 + 3]);
        }
      }
      l = fmaf(l, alpha, rs);
      m = m_new;
      alpha_prev = alpha;
      pf_fence_before();                                            // TMEM reads of S_j / PV_{j-1} are complete
      pf_fence_proxy_async();                                       // generic-proxy stores of P -> visible to the MMA
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(s_empty + b);
        mbar_arrive(p_full);
      }
    }
    // ---- the last tile's P V, normalise, store
    mbar_wait(pv_full, (uint32_t)((n_tiles - 1) & 1));
    pf_fence_after();
    const float inv = l > 0.f ? 1.f / l : 0.f;
    __nv_bfloat16* orow = p.out + ((int64_t)(q0 + q_idx) * p.num_heads + head) * PF_HD;
#pragma unroll
    for (int c = 0; c < PF_HD / 32; ++c) {
      uint32_t v[32];
      pf_tmem_ld32(tmem_pv + lane_addr + c * 32, v);
      if (r < rows_valid) {
#pragma unroll
        for (int i = 0; i < 32; i += 8) {
          float f[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) f[e] = fmaf(o[c * 32 + i + e], alpha_prev, __uint_as_float(v[i + e])) * inv;
          uint4 w;
          w.x = pack_bf16(f[0], f[1]); w.y = pack_bf16(f[2], f[3]); w.z = pack_bf16(f[4], f[5]); w.w = pack_bf16(f[6], f[7]);
          stg_v4(orow + c * 32 + i, w);
        }
      }
    }
    pf_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    pf_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem_base), "r"(512u) : "memory");
  }
}

// ---- host side
typedef CUresult (*PfEncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
// [rows, cols] bf16 with a row stride, box = 64 columns x 128 rows, SWIZZLE_128B, rows beyond `rows` zero-filled
static int pf_make_map(CUtensorMap* tm, const void* base, int64_t row_stride, int rows, int cols, const char* what) {
  static PfEncodeTiledFn fn = nullptr;
  if (!fn) {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !sym) {
      set_error("%s: cuTensorMapEncodeTiled unavailable", what);
      return B2TTS_ECUDA;
    }
    fn = reinterpret_cast<PfEncodeTiledFn>(sym);
  }
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)row_stride * 2};
  const cuuint32_t box[2] = {64, 128};
  const cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("%s: cuTensorMapEncodeTiled failed (%d) for [%d, %d] stride %lld", what, (int)r, rows, cols, (long long)row_stride);
    return B2TTS_ECUDA;
  }
  return B2TTS_OK;
}

bool prefill_tc_applicable(int max_seqlen_q, int head_dim, int num_heads, int num_kv_heads, const void* block_table) {
  return !block_table && head_dim == PF_HD && max_seqlen_q >= PF_MIN_SEQLEN && num_kv_heads > 0 &&
         num_heads % num_kv_heads == 0;
}

int launch_prefill_tc(const char* what, const void* q, int64_t q_row_stride, const void* k, int64_t k_row_stride,
                      const void* v, int64_t v_row_stride, const int32_t* cu_q, const int32_t* cu_k, void* out,
                      int num_seqs, int total_q, int total_k, int max_seqlen_q, int num_heads, int num_kv_heads,
                      float scale, cudaStream_t st) {
  CUtensorMap mq, mk, mv;
  int rc = pf_make_map(&mq, q, q_row_stride, total_q, num_heads * PF_HD, what);
  if (!rc) rc = pf_make_map(&mk, k, k_row_stride, total_k, num_kv_heads * PF_HD, what);
  if (!rc) rc = pf_make_map(&mv, v, v_row_stride, total_k, num_kv_heads * PF_HD, what);
  if (rc) return rc;
  PrefillTcParams p{};
  p.cu_q = cu_q; p.cu_k = cu_k;
  p.out = static_cast<__nv_bfloat16*>(out);
  p.num_heads = num_heads; p.num_kv_heads = num_kv_heads;
  p.scale_log2 = scale * 1.4426950408889634f;
  const size_t smem = (size_t)PF_TILE_BYTES * (2 + 2 * PF_STAGES) + 256 + 1024;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(prefill_attn_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return cuda_fail(e, what);
    configured = true;
  }
  dim3 grid((max_seqlen_q + PF_BM - 1) / PF_BM, num_heads, num_seqs);
  prefill_attn_tc_kernel<<<grid, PF_THREADS, smem, st>>>(p, mq, mk, mv);
  B2_CHECK_LAUNCH(what);
  return B2TTS_OK;
}

}  // namespace b2

// gemm.cu - a7: the projections of the decode path as ONE weight-streaming tcgen05 GEMM.
// Replaces ColumnParallelLinear / RowParallelLinear forward = F.linear (layers/linear.py:72-73,
// 149-153), the three q/k/v and two gate/up launches of models/qwen3_tts_share.py:85-87,126-134,
// nn.Linear of small_to_mtp_projection / lm_head (models/qwen3_tts_predictor.py:54-61) and the
// codec_head GEMM (layers/embed_head.py:61).
//
//   y[M, N] = x[M, K] @ W[N, K]^T      M = streams in the step (1..256), W bf16, fp32 accumulate
//
// In decode M is small, so the GEMM is a stream of W through the SM: HBM-bound, K*N*2 bytes.
// The roles are therefore swapped with respect to a training GEMM:
//   * W is the 128-row UMMA "A" operand (128 output features per CTA), the batch is the UMMA
//     "N" dimension (16..256), the accumulator D[128 features, Mpad] lives in TMEM;
//   * W is re-packed ONCE at load time (b2tts_gemm_pack_weight) into the exact shared-memory
//     image of a K-major SWIZZLE_128B operand tile, [n_tile][k_block] x 16 KiB, so one 1-D bulk
//     async copy (cp.async.bulk -> UBLKCP, the TMA unit) per pipeline stage streams it, fully
//     contiguous in HBM, with no tensor map;
//   * x (L2-resident, written by the previous kernel) is staged by 4 producer warps into the
//     same swizzled layout (generic-proxy stores + fence.proxy.async);
//   * one thread issues tcgen05.mma (kind::f16, bf16 x bf16 -> fp32), tcgen05.commit releases
//     ring stages and signals the epilogue through mbarriers;
//   * N/128 tiles alone cannot fill 148 SMs (N = 1024 gives 8), so K is split over a thread-block
//     CLUSTER of S CTAs (S = 1,2,4,8,16); the S partial accumulators are exchanged through
//     distributed shared memory (each CTA owns 128/S features of the tile: push with
//     st.shared::cluster, one cluster barrier, fixed-order fp32 sum => deterministic);
//   * the epilogue fuses what follows the projection in the decoder layer: bias, the residual add
//     (qwen3_tts_share.py:171,175) or SiLU(gate)*up (qwen3_tts_share.py:131-135);
//   * programmatic dependent launch: the W stream starts before the previous kernel has finished
//     (weights do not depend on it); only the x producers / epilogue execute griddepcontrol.wait.
#include "common.cuh"

namespace b2 {

constexpr int G_BN = 128;                       // output features per CTA = UMMA M
constexpr int G_BK = 64;                        // K per stage = one 128-byte swizzle row of bf16
constexpr int G_STAGES = 4;
constexpr int G_THREADS = 192;                  // warp 0: W producer, warp 1: MMA issuer + TMEM owner, warps 2-5: x producers + epilogue
constexpr int G_WTILE_BYTES = G_BN * G_BK * 2;  // 16 KiB
constexpr int G_MAX_M = 256;

struct GemmParams {
  const __nv_bfloat16* x;
  int64_t x_stride;
  const uint8_t* wp;            // packed weight
  const __nv_bfloat16* bias;    // [N] or null (epilogue 0)
  const __nv_bfloat16* res;     // [M, N] (epilogue 1)
  int64_t res_stride;
  __nv_bfloat16* out;
  int64_t out_stride;
  int m, n, k;
  int mpad;                     // M rounded up to 16 (UMMA N)
  int kblocks;                  // K / 64
  int splits;                   // cluster size along K
  int epilogue;                 // B2TTS_GEMM_*
  int rounding;                 // B2TTS_ROUND_* of the SiLU*mul epilogue
  int tmem_cols;                // power of two >= 32, >= mpad
};

// ---------------------------------------------------------------- tcgen05 / cluster PTX
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(smem_u32(smem_dst)), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(taddr), "r"(cols) : "memory");
}
// K-major SWIZZLE_128B shared-memory matrix descriptor: 8-row groups 1024 B apart.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3ffffu) >> 4);          // start address, bits [0,14)
  d |= (uint64_t)1 << 16;                                // leading byte offset (unused for swizzled K-major)
  d |= (uint64_t)(1024 >> 4) << 32;                      // stride byte offset, bits [32,46)
  d |= (uint64_t)1 << 46;                                // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;                                // layout type SWIZZLE_128B
  return d;
}
// kind::f16 instruction descriptor: bf16 x bf16 -> fp32, both operands K-major, M=128, N=n.
__device__ __forceinline__ uint32_t umma_idesc_bf16(int n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(G_BN >> 4) << 24);
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}"
      :: "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float* v) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ uint32_t mapa_u32(uint32_t local_addr, uint32_t cta_rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(cta_rank));
  return r;
}
__device__ __forceinline__ void st_cluster_f32(uint32_t addr, float v) {
  asm volatile("st.shared::cluster.f32 [%0], %1;" :: "r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

__device__ __forceinline__ float gsilu_f(float g) { return g / (1.0f + expf(-g)); }   // as elementwise.cu

// One GEMM CTA: 128 features x Mpad batch columns over its K range.  NCH = 16-byte x chunks each
// producer thread stages per k-block (Mpad * 8 / 128).
template <int NCH>
__global__ void __launch_bounds__(G_THREADS, 1)
gemm_stream_kernel(const GemmParams p) {
  extern __shared__ uint8_t smem_dyn[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~uintptr_t(1023));
  const int mpad = p.mpad;
  const int stage_bytes = G_WTILE_BYTES + mpad * 128;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + G_STAGES * stage_bytes);
  uint64_t* empty_bar = full_bar + G_STAGES;
  uint64_t* accum_bar = empty_bar + G_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int S = p.splits;
  const int rank = blockIdx.x;                  // gridDim.x == S == cluster size: rank in cluster
  const int tile = blockIdx.y;
  const int kb_begin = (int)(((int64_t)rank * p.kblocks) / S);
  const int kb_end = (int)(((int64_t)(rank + 1) * p.kblocks) / S);
  const int nkb = kb_end - kb_begin;

  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

  if (threadIdx.x == 0) {
    for (int s = 0; s < G_STAGES; ++s) {
      mbar_init(full_bar + s, 1 + 4);           // W producer (expect_tx arrive) + 4 x-producer warps
      mbar_init(empty_bar + s, 1);              // tcgen05.commit
    }
    mbar_init(accum_bar, 1);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, (uint32_t)p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // =========================== W producer: one 16 KiB bulk copy per k-block, no dependency on
    // the previous kernel, so it runs ahead of griddepcontrol.wait.
    if (lane == 0) {
      const uint8_t* src = p.wp + ((int64_t)tile * p.kblocks + kb_begin) * G_WTILE_BYTES;
      for (int i = 0; i < nkb; ++i) {
        const int stage = i % G_STAGES;
        const uint32_t ph = (i / G_STAGES) & 1;
        mbar_wait(empty_bar + stage, ph ^ 1);
        mbar_expect_tx(full_bar + stage, G_WTILE_BYTES);
        bulk_g2s(smem + stage * stage_bytes, src + (int64_t)i * G_WTILE_BYTES, G_WTILE_BYTES, full_bar + stage);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // =========================== MMA issuer
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(mpad);
      for (int i = 0; i < nkb; ++i) {
        const int stage = i % G_STAGES;
        const uint32_t ph = (i / G_STAGES) & 1;
        mbar_wait(full_bar + stage, ph);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(smem + stage * stage_bytes);
        const uint32_t b_addr = a_addr + G_WTILE_BYTES;
#pragma unroll
        for (int k = 0; k < G_BK / 16; ++k)
          umma_bf16(tmem_base, umma_desc_sw128(a_addr + k * 32), umma_desc_sw128(b_addr + k * 32), idesc,
                    (uint32_t)((i | k) != 0));
        umma_commit(empty_bar + stage);         // stage reusable once these MMAs have read it
      }
      umma_commit(accum_bar);                   // accumulator complete
    }
    __syncwarp();
  } else {
    // =========================== x producers (128 threads): global (L2) -> swizzled smem
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int tp = threadIdx.x - 64;
    uint4 cur[NCH], nxt[NCH];
    auto load = [&](uint4* dst, int kb) {
      const __nv_bfloat16* base = p.x + (int64_t)(kb_begin + kb) * G_BK;
#pragma unroll
      for (int j = 0; j < NCH; ++j) {
        const int idx = tp + 128 * j;
        const int r = idx >> 3, c = idx & 7;
        dst[j] = (r < p.m) ? ldg_v4(base + (int64_t)r * p.x_stride + c * 8) : make_uint4(0, 0, 0, 0);
      }
    };
    if (nkb > 0) load(cur, 0);
    for (int i = 0; i < nkb; ++i) {
      const int stage = i % G_STAGES;
      const uint32_t ph = (i / G_STAGES) & 1;
      if (i + 1 < nkb) load(nxt, i + 1);
      mbar_wait(empty_bar + stage, ph ^ 1);
      uint8_t* xs = smem + stage * stage_bytes + G_WTILE_BYTES;
#pragma unroll
      for (int j = 0; j < NCH; ++j) {
        const int idx = tp + 128 * j;
        const int r = idx >> 3, c = idx & 7;
        if (r < mpad) *reinterpret_cast<uint4*>(xs + r * 128 + ((c ^ (r & 7)) << 4)) = cur[j];
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(full_bar + stage);
#pragma unroll
      for (int j = 0; j < NCH; ++j) cur[j] = nxt[j];
    }
  }

  // =========================== all threads: accumulator complete => this CTA's ring is free
  mbar_wait(accum_bar, 0);
  tc_fence_after();
  if (warp < 2) asm volatile("griddepcontrol.wait;" ::: "memory");   // the epilogue below reads res / writes out
  if (S > 1) cluster_sync_all();                // every CTA of the cluster is past its MMAs: rings are reusable

  // reduction buffer red[src_rank][m][feature within owner slice] (fp32) aliases the ring
  const int rows_own = G_BN / S;
  float* red = reinterpret_cast<float*>(smem);
  if (warp >= 2) {
    const int q = warp & 3;                     // TMEM lane quarter this warp may read
    const int f = q * 32 + lane;                // feature row of the tile held by this thread
    const int owner = f / rows_own, fo = f - owner * rows_own;
    const uint32_t local = smem_u32(red + ((int64_t)rank * mpad) * rows_own + fo);
    const uint32_t dst = (S > 1) ? mapa_u32(local, (uint32_t)owner) : local;
    for (int c0 = 0; c0 < mpad; c0 += 16) {
      float v[16];
      if (nkb > 0) {
        tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, v);
      } else {
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = 0.f;
      }
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const uint32_t a = dst + (uint32_t)((c0 + j) * rows_own * 4);
        if (S > 1) st_cluster_f32(a, v[j]);
        else asm volatile("st.shared.f32 [%0], %1;" :: "r"(a), "f"(v[j]) : "memory");
      }
    }
  }
  tc_fence_before();
  if (S > 1) cluster_sync_all(); else __syncthreads();

  // =========================== fixed-order sum of the S partials + fused epilogue.
  // One item = (batch row m, 8 consecutive features of this CTA's slice).
  {
    const int groups = rows_own >> 3;
    const int items = p.m * groups;
    const int n0 = tile * G_BN + rank * rows_own;
    for (int it = threadIdx.x; it < items; it += G_THREADS) {
      const int m = it / groups, g8 = it - m * groups;
      float acc[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = 0.f;
      for (int s = 0; s < S; ++s) {
        const float4* src = reinterpret_cast<const float4*>(red + ((int64_t)s * mpad + m) * rows_own + g8 * 8);
        const float4 a = src[0], b = src[1];
        acc[0] += a.x; acc[1] += a.y; acc[2] += a.z; acc[3] += a.w;
        acc[4] += b.x; acc[5] += b.y; acc[6] += b.z; acc[7] += b.w;
      }
      const int n = n0 + g8 * 8;
      if (p.epilogue == B2TTS_GEMM_SILU_MUL) {
        // packed rows of this group: 4 gate features then the 4 matching up features
        float o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float g = rbf(acc[j]), u = rbf(acc[4 + j]);       // the gate_up GEMM output is bf16
          o[j] = (p.rounding == B2TTS_ROUND_EAGER) ? rbf(gsilu_f(g)) * u : gsilu_f(g) * u;
        }
        uint2 w;
        w.x = pack_bf16(o[0], o[1]); w.y = pack_bf16(o[2], o[3]);
        stg_v2(p.out + (int64_t)m * p.out_stride + (n >> 1), w);
      } else {
        if (p.epilogue == B2TTS_GEMM_BIAS && p.bias) {
          const uint4 bw = ldg_v4(p.bias + n);
          acc[0] += bf16lo(bw.x); acc[1] += bf16hi(bw.x); acc[2] += bf16lo(bw.y); acc[3] += bf16hi(bw.y);
          acc[4] += bf16lo(bw.z); acc[5] += bf16hi(bw.z); acc[6] += bf16lo(bw.w); acc[7] += bf16hi(bw.w);
        } else if (p.epilogue == B2TTS_GEMM_RESIDUAL) {
          // bf16(residual + bf16(x W^T)): the projection output is rounded before the add
          const uint4 rw = ldg_v4(p.res + (int64_t)m * p.res_stride + n);
          acc[0] = rbf(acc[0]) + bf16lo(rw.x); acc[1] = rbf(acc[1]) + bf16hi(rw.x);
          acc[2] = rbf(acc[2]) + bf16lo(rw.y); acc[3] = rbf(acc[3]) + bf16hi(rw.y);
          acc[4] = rbf(acc[4]) + bf16lo(rw.z); acc[5] = rbf(acc[5]) + bf16hi(rw.z);
          acc[6] = rbf(acc[6]) + bf16lo(rw.w); acc[7] = rbf(acc[7]) + bf16hi(rw.w);
        }
        uint4 w;
        w.x = pack_bf16(acc[0], acc[1]); w.y = pack_bf16(acc[2], acc[3]);
        w.z = pack_bf16(acc[4], acc[5]); w.w = pack_bf16(acc[6], acc[7]);
        stg_v4(p.out + (int64_t)m * p.out_stride + n, w);
      }
    }
  }
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, (uint32_t)p.tmem_cols);
}

// Re-pack W[N, K] (row-major) into [N/128][K/64] tiles of 16 KiB, each the K-major SWIZZLE_128B
// shared-memory image of 128 rows x 64 k.  One thread moves one 16-byte chunk.
__global__ void __launch_bounds__(256)
gemm_pack_kernel(const __nv_bfloat16* __restrict__ w, int64_t w_stride, uint8_t* __restrict__ packed,
                 int n, int k, int mode) {
  const int kblocks = k / G_BK;
  const int64_t total = (int64_t)n * k / 8;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i & 7);
    const int r = (int)((i >> 3) & 127);
    const int64_t tk = i >> 10;                 // tile * kblocks + kb
    const int kb = (int)(tk % kblocks);
    const int tile = (int)(tk / kblocks);
    int feature;
    if (mode == B2TTS_PACK_GATE_UP) {
      const int half = n >> 1;
      const int j = tile * 64 + (r >> 3) * 4 + (r & 3);
      feature = (r & 4) ? half + j : j;
    } else {
      feature = tile * G_BN + r;
    }
    const uint4 v = ldg_nc_v4(w + (int64_t)feature * w_stride + (int64_t)kb * G_BK + c * 8);
    *reinterpret_cast<uint4*>(packed + tk * G_WTILE_BYTES + r * 128 + ((c ^ (r & 7)) << 4)) = v;
  }
}

static int choose_splits(int n_tiles, int kblocks) {
  int s = 1;
  while (s < 16 && n_tiles * s < 120 && s * 2 <= kblocks) s *= 2;
  return s;
}

static size_t gemm_smem_bytes(int mpad) {
  return (size_t)G_STAGES * (G_WTILE_BYTES + mpad * 128) + 1024 + 128;
}

template <int NCH>
static int launch_gemm(const GemmParams& p, cudaStream_t st, int pdl) {
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(gemm_stream_kernel<NCH>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return cuda_fail(e, "gemm_bf16(smem attr)");
    e = cudaFuncSetAttribute(gemm_stream_kernel<NCH>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    if (e != cudaSuccess) return cuda_fail(e, "gemm_bf16(cluster attr)");
    configured = true;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(p.splits, p.n / G_BN, 1);
  cfg.blockDim = dim3(G_THREADS, 1, 1);
  cfg.dynamicSmemBytes = gemm_smem_bytes(p.mpad);
  cfg.stream = st;
  cudaLaunchAttribute attrs[2];
  int na = 0;
  if (p.splits > 1) {
    attrs[na].id = cudaLaunchAttributeClusterDimension;
    attrs[na].val.clusterDim.x = p.splits;
    attrs[na].val.clusterDim.y = 1;
    attrs[na].val.clusterDim.z = 1;
    ++na;
  }
  if (pdl) {
    attrs[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attrs[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attrs;
  cfg.numAttrs = na;
  cudaError_t e = cudaLaunchKernelEx(&cfg, gemm_stream_kernel<NCH>, p);
  count_launch();
  if (e != cudaSuccess) return cuda_fail(e, "gemm_bf16");
  return B2TTS_OK;
}

}  // namespace b2

extern "C" int b2tts_gemm_bf16_supported(int m, int n, int k) {
  return m >= 1 && m <= b2::G_MAX_M && n >= b2::G_BN && n % b2::G_BN == 0 && k >= b2::G_BK && k % b2::G_BK == 0;
}

extern "C" size_t b2tts_gemm_packed_bytes(int n, int k) {
  if (n <= 0 || k <= 0 || n % b2::G_BN || k % b2::G_BK) return 0;
  return (size_t)n * k * 2;
}

extern "C" int b2tts_gemm_pack_weight(const void* weight, int64_t weight_row_stride, void* packed,
                                      int n, int k, int mode, void* stream) {
  using namespace b2;
  const char* what = "gemm_pack_weight";
  B2_REQUIRE(weight && packed, B2TTS_EINVAL, "%s: null pointer", what);
  B2_REQUIRE(n > 0 && k > 0 && n % G_BN == 0 && k % G_BK == 0, B2TTS_EUNSUPPORTED,
             "%s: n=%d must be a multiple of %d and k=%d of %d", what, n, G_BN, k, G_BK);
  B2_REQUIRE(mode == B2TTS_PACK_PLAIN || mode == B2TTS_PACK_GATE_UP, B2TTS_EINVAL, "%s: mode %d", what, mode);
  B2_REQUIRE(aligned16(weight) && aligned16(packed) && weight_row_stride % 8 == 0, B2TTS_EALIGN,
             "%s: pointers / row stride must be 16-byte aligned", what);
  const int64_t total = (int64_t)n * k / 8;
  const int blocks = (int)((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
  gemm_pack_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(weight), weight_row_stride, static_cast<uint8_t*>(packed), n, k, mode);
  B2_CHECK_LAUNCH(what);
  return B2TTS_OK;
}

extern "C" int b2tts_gemm_bf16(const void* x, int64_t x_row_stride, const void* w_packed,
                               const void* bias, const void* residual, int64_t residual_row_stride,
                               void* out, int64_t out_row_stride, int m, int n, int k,
                               int epilogue, int rounding, int splits, int pdl, void* stream) {
  using namespace b2;
  const char* what = "gemm_bf16";
  B2_REQUIRE(m >= 0 && n > 0 && k > 0, B2TTS_EINVAL, "%s: m=%d n=%d k=%d", what, m, n, k);
  if (m == 0) return B2TTS_OK;
  B2_REQUIRE(b2tts_gemm_bf16_supported(m, n, k), B2TTS_EUNSUPPORTED,
             "%s: m=%d (1..%d), n=%d (multiple of %d), k=%d (multiple of %d)", what, m, G_MAX_M, n, G_BN, k, G_BK);
  B2_REQUIRE(x && w_packed && out, B2TTS_EINVAL, "%s: null pointer", what);
  B2_REQUIRE(epilogue == B2TTS_GEMM_BIAS || epilogue == B2TTS_GEMM_RESIDUAL || epilogue == B2TTS_GEMM_SILU_MUL,
             B2TTS_EINVAL, "%s: epilogue %d", what, epilogue);
  B2_REQUIRE(epilogue != B2TTS_GEMM_RESIDUAL || residual, B2TTS_EINVAL, "%s: residual epilogue without residual", what);
  B2_REQUIRE(aligned16(x) && aligned16(w_packed) && aligned16(out) && x_row_stride % 8 == 0 && out_row_stride % 4 == 0 &&
                 (epilogue == B2TTS_GEMM_SILU_MUL || out_row_stride % 8 == 0) &&
                 (!bias || aligned16(bias)) && (!residual || (aligned16(residual) && residual_row_stride % 8 == 0)),
             B2TTS_EALIGN, "%s: pointers / strides must be 16-byte aligned", what);
  GemmParams p{};
  p.x = static_cast<const __nv_bfloat16*>(x);
  p.x_stride = x_row_stride;
  p.wp = static_cast<const uint8_t*>(w_packed);
  p.bias = static_cast<const __nv_bfloat16*>(bias);
  p.res = static_cast<const __nv_bfloat16*>(residual);
  p.res_stride = residual_row_stride;
  p.out = static_cast<__nv_bfloat16*>(out);
  p.out_stride = out_row_stride;
  p.m = m; p.n = n; p.k = k;
  p.mpad = (m + 15) & ~15;
  p.kblocks = k / G_BK;
  p.epilogue = epilogue;
  p.rounding = rounding;
  int s = splits > 0 ? splits : choose_splits(n / G_BN, p.kblocks);
  B2_REQUIRE((s == 1 || s == 2 || s == 4 || s == 8 || s == 16) && s <= p.kblocks, B2TTS_EINVAL,
             "%s: splits %d must be 1,2,4,8,16 and <= k/64", what, s);
  p.splits = s;
  p.tmem_cols = 32;
  while (p.tmem_cols < p.mpad) p.tmem_cols <<= 1;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int nch = p.mpad / 16;                  // 16-byte chunks per producer thread per k-block
  if (nch <= 1) return launch_gemm<1>(p, st, pdl);
  if (nch <= 2) return launch_gemm<2>(p, st, pdl);
  if (nch <= 4) return launch_gemm<4>(p, st, pdl);
  if (nch <= 8) return launch_gemm<8>(p, st, pdl);
  return launch_gemm<16>(p, st, pdl);
}

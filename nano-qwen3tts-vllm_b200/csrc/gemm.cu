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
#include <cuda.h>
#include <stdio.h>
#include <stdlib.h>

#include "umma.cuh"

namespace b2 {

constexpr int G_STAGES = 4;                    // ring depth the split heuristic assumes
constexpr int G_MAX_STAGES = 12;
constexpr int G_THREADS = 192;                  // warp 0: W producer, warp 1: MMA issuer + TMEM owner, warps 2-5: x producers + epilogue
constexpr int G_MAX_M = 128;                    // batch rows per CTA (UMMA N); larger batches are cut into row chunks
constexpr int G_MAX_ROWS = 1 << 16;

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
  int stages;                   // ring depth (2..4)
  // fused input RMSNorm: x is the RESIDUAL stream; the x producers stage norm(x) = f(x * rstd[m], w[k])
  const __nv_bfloat16* norm_w;  // [K] or null
  const float* ssq_in;          // [M, ssq_in_slots] partial sums of squares of the x rows (fixed-order sum)
  int ssq_in_slots;
  float eps;
  // per-row partial sum of squares of the OUTPUT rows (bias / residual epilogues): [M, n / (128 / splits)]
  float* ssq_out;
  unsigned long long* trace;    // debug: 8 globaltimer stamps per CTA (null = off)
  int ssq_lanes;                // lanes per row when summing ssq_in (power of two, divides ssq_in_slots)
  // L2 prefetch of the NEXT GEMM's packed weights (or null): weights never depend on activations, and a decode
  // step is a chain of latency-bound launches during which HBM idles ~75 % of the time; every CTA asks the TMA
  // unit to pull its slice of the next weight image into L2 (cp.async.bulk.prefetch.L2) once its own stream is fully
  // requested, so the next launch streams from L2.
  const uint8_t* pf_ptr;
  unsigned long long pf_bytes;
  int bsplit;                   // > 1: the cluster splits the BATCH, not K (then splits == 1): CTA r owns rows [r * mpad, (r + 1) * mpad)
                                // of the chunk over the whole K range, and every weight tile is fetched once per cluster - each
                                // CTA requests 1 / bsplit of it and the copy is multicast into all of them.  No partial sums, no
                                // exchange, a TMEM drain of mpad = 16 / 32 columns: what K <= 1024 projections want, where a
                                // CTA's whole K range (16 k-blocks of MMA issue, ~3 us) costs less than drain + exchange.
  int chunk_rows;               // rows per chunk when the batch is cut (<= G_MAX_M)
  int nchunks;                  // M > 128: the batch is cut into ceil(M / 128) row chunks, one CTA (cluster) per
                                // (feature tile, chunk); blockIdx.y = tile * nchunks + chunk so that the chunks of one
                                // weight tile are launched next to each other and share it through L2
};

#define G_TRACE(slot) do { if (p.trace) p.trace[(blockIdx.y * gridDim.x + blockIdx.x) * 8 + (slot)] = gtimer(); } while (0)

// One GEMM CTA: 128 features x Mpad batch columns over its K range.  NCH = 16-byte x chunks each
// producer thread stages per k-block (Mpad * 8 / 128).
template <int NCH>
__global__ void __launch_bounds__(G_THREADS, 1)
gemm_stream_kernel(const GemmParams p, const __grid_constant__ CUtensorMap tmap_x) {
  extern __shared__ uint8_t smem_dyn[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~uintptr_t(1023));
  const int mpad = p.mpad;
  const int NS = p.stages;
  const int stage_bytes = G_WTILE_BYTES + mpad * 128;
  uint8_t* ring_end = smem + NS * stage_bytes;
  // receive buffer of the split-K exchange (S > 1): recv[src_rank][feature within my slice][mpad + 4] fp32
  float* recv = reinterpret_cast<float*>(ring_end);
  uint8_t* tail = ring_end + (p.splits > 1 ? (mpad + 4) * 512 : 0);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(tail);
  uint64_t* empty_bar = full_bar + G_MAX_STAGES;
  uint64_t* raw_bar = empty_bar + G_MAX_STAGES;    // x tile landed (TMA), before the optional in-place norm
  uint64_t* accum_bar = raw_bar + G_MAX_STAGES;
  uint64_t* recv_bar = accum_bar + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(recv_bar + 1);
  float* rstd_s = reinterpret_cast<float*>(tmem_slot + 2);        // [mpad]
  float* qx_s = rstd_s + mpad;                                    // [S][mpad]: the cluster's per-row ssq partials (rank 0's copy is used)

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);       // provably warp-uniform
  const int lane = threadIdx.x & 31;
  const int S = p.splits;
  const int BS = p.bsplit;                      // batch-split cluster (then S == 1)
  const int rank = BS > 1 ? 0 : (int)blockIdx.x;                // gridDim.x == cluster size: K-split rank ...
  const int brank = BS > 1 ? (int)blockIdx.x : 0;               // ... or batch-split rank
  const uint16_t bmask = (uint16_t)((1u << BS) - 1u);
  const int tile = blockIdx.y / p.nchunks;
  const int row0 = (blockIdx.y % p.nchunks) * p.chunk_rows + brank * mpad;   // first batch row of this CTA
  const int m_loc = max(0, min(p.m - row0, BS > 1 ? mpad : p.chunk_rows));   // its rows (the last CTA may be short, or empty)
  const int kb_begin = (int)(((int64_t)rank * p.kblocks) / S);
  const int kb_end = (int)(((int64_t)(rank + 1) * p.kblocks) / S);
  const int nkb = kb_end - kb_begin;

  if (threadIdx.x == 0) G_TRACE(0);
  pdl_launch();
  // the x tensor map's descriptor is fetched now, under the previous kernel's tail, not by the first copy after the wait
  if (threadIdx.x == 32) asm volatile("prefetch.tensormap [%0];" :: "l"(reinterpret_cast<uint64_t>(&tmap_x)) : "memory");

  if (threadIdx.x == 0) {
    for (int s = 0; s < NS; ++s) {
      // fused norm: W producer (expect_tx arrive) + the 4 x warps that transform the landed tile; otherwise the x copy
      // completes straight on this barrier (two expect_tx arrivals: W's and x's) - no relay through the x warps
      mbar_init(full_bar + s, p.norm_w ? 1 + 4 : 2);
      mbar_init(empty_bar + s, BS > 1 ? BS : 1);          // tcgen05.commit (of every CTA that received the multicast tile)
      mbar_init(raw_bar + s, 1);
    }
    mbar_init(accum_bar, 1);
    mbar_init(recv_bar, 1);
    mbar_fence_init();
    if (S > 1) mbar_expect_tx(recv_bar, (uint32_t)((mpad + 4) * 512));   // S slices of (128/S) rows
  }
  if (warp == 1) tmem_alloc(tmem_slot, (uint32_t)p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (S > 1 || BS > 1) cluster_arrive();        // my barriers are initialised (waited on before the exchange / the first multicast)
  if (threadIdx.x == 0) G_TRACE(1);

  if (warp == 0) {
    // =========================== producer thread.  W: one 16 KiB bulk copy per k-block; it does not
    // depend on the previous kernel, so the first ring-full is requested BEFORE griddepcontrol.wait.
    // x: one 2-D tensor-map copy per k-block (SWIZZLE_128B: lands as the UMMA B tile, rows >= M
    // zero-filled), requested after the wait.  Both go through the TMA unit, so neither the HBM nor
    // the L2 round trip is ever paid per block by a waiting thread.
    {
      // the whole warp walks the ring (uniform addresses and barriers), one elected lane issues each copy
      const uint8_t* src = p.wp + ((int64_t)tile * p.kblocks + kb_begin) * G_WTILE_BYTES;
      const uint32_t xbytes = (uint32_t)mpad * 128u;
      const int first = min(nkb, NS);
      const uint32_t wslice = G_WTILE_BYTES / (uint32_t)BS, woff = (uint32_t)brank * wslice;
      if (BS > 1) cluster_wait();                 // every CTA of the cluster can receive (this warp's wait; the others' is below)
      if (elect_one()) {
        for (int i = 0; i < first; ++i) {
          mbar_expect_tx(full_bar + i, G_WTILE_BYTES);
          if (BS > 1) bulk_g2s_multicast(smem + i * stage_bytes + woff, src + (int64_t)i * G_WTILE_BYTES + woff, wslice, full_bar + i, bmask);
          else bulk_g2s(smem + i * stage_bytes, src + (int64_t)i * G_WTILE_BYTES, G_WTILE_BYTES, full_bar + i);
        }
      }
      __syncwarp();
      pdl_wait();
      if (elect_one()) {
        for (int i = 0; i < first; ++i) {
          uint64_t* landed = p.norm_w ? raw_bar + i : full_bar + i;
          mbar_expect_tx(landed, xbytes);
          tma_load_2d(smem + i * stage_bytes + G_WTILE_BYTES, &tmap_x, (kb_begin + i) * G_BK, row0, landed);
        }
      }
      __syncwarp();
      int stage = first == NS ? 0 : first;
      uint32_t ph = 1;                              // parity of the "empty" wait: first reuse of a stage waits on phase 0
      for (int i = first; i < nkb; ++i) {
        mbar_wait(empty_bar + stage, ph ^ 1);
        if (elect_one()) {
          mbar_expect_tx(full_bar + stage, G_WTILE_BYTES);
          if (BS > 1) bulk_g2s_multicast(smem + stage * stage_bytes + woff, src + (int64_t)i * G_WTILE_BYTES + woff, wslice, full_bar + stage, bmask);
          else bulk_g2s(smem + stage * stage_bytes, src + (int64_t)i * G_WTILE_BYTES, G_WTILE_BYTES, full_bar + stage);
          uint64_t* landed = p.norm_w ? raw_bar + stage : full_bar + stage;
          mbar_expect_tx(landed, xbytes);
          tma_load_2d(smem + stage * stage_bytes + G_WTILE_BYTES, &tmap_x, (kb_begin + i) * G_BK, row0, landed);
        }
        __syncwarp();
        if (++stage == NS) { stage = 0; ph ^= 1; }
      }
      // my own stream is fully requested: from here to the next launch's first copy HBM would idle (MMA tail, split-K
      // exchange, epilogue, kernel boundary), so this is when the next weight image is pulled towards L2
      if (p.pf_ptr && elect_one()) {
        const unsigned long long ncta = (unsigned long long)gridDim.x * gridDim.y;
        const unsigned long long me = (unsigned long long)blockIdx.y * gridDim.x + blockIdx.x;
        const unsigned long long pieces = p.pf_bytes / G_WTILE_BYTES;
        for (unsigned long long t = me; t < pieces; t += ncta)
          asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(p.pf_ptr + t * G_WTILE_BYTES), "r"(G_WTILE_BYTES) : "memory");
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // =========================== MMA issuer: the whole warp walks the ring (uniform control flow and operands),
    // one elected lane issues
    {
      const uint32_t idesc = umma_idesc_bf16(mpad);
      int stage = 0;
      uint32_t ph = 0;
      for (int i = 0; i < nkb; ++i) {
        mbar_wait(full_bar + stage, ph);
        tc_fence_after();
        const uint32_t a_lo = umma_desc_lo_sw128(smem_u32(smem + stage * stage_bytes));
        const uint32_t b_lo = a_lo + (G_WTILE_BYTES >> 4);
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < G_BK / 16; ++k)
            umma_bf16_lo(tmem_base, a_lo + 2 * k, b_lo + 2 * k, idesc, (uint32_t)((i | k) != 0));
          // stage reusable once these MMAs have read it - but only a k-block NS later ever waits for that: the
          // commits of the last NS k-blocks (ALL of them for the split-K decode shapes, whose K range fits the ring)
          // would be ~0.2 us of issue time each for a barrier nobody reads
          if (i + NS < nkb) { if (BS > 1) umma_commit_multicast(empty_bar + stage, bmask); else umma_commit(empty_bar + stage); }
          else if (i == nkb - 1) umma_commit(accum_bar);      // accumulator complete
        }
        __syncwarp();
        if (++stage == NS) { stage = 0; ph ^= 1; }
      }
      if (nkb == 0 && elect_one()) umma_commit(accum_bar);
    }
    __syncwarp();
  } else {
    // =========================== x tiles land by TMA; with the fused norm these 4 warps transform the
    // landed tile in place (shared-memory latency only), then release it to the MMA issuer.
    const int tp = threadIdx.x - 64;
    const bool normed = p.norm_w != nullptr;
    const bool eager = p.rounding == B2TTS_ROUND_EAGER;
    // the norm weights of the first k-block are constants: requested before the wait, not after the rstd prologue
    // (where their ~0.7 us of load latency sat in front of the first transform)
    uint4 wv_next = make_uint4(0, 0, 0, 0);
    if (normed && nkb > 0) wv_next = ldg_v4(p.norm_w + (int64_t)kb_begin * G_BK + (tp & 7) * 8);
    pdl_wait();
    if (threadIdx.x == 64) G_TRACE(2);
    if (normed) {
      // rstd[m] from the partial sums of squares the producing kernel left: G lanes per row, each a
      // contiguous chunk summed in order, then a fixed shuffle tree (deterministic).
      const int G = p.ssq_lanes, rows_per_pass = 128 / G;
      const int gl = tp % G, chunk = p.ssq_in_slots / G;
      for (int r0 = 0; r0 < mpad; r0 += rows_per_pass) {
        const int r = r0 + tp / G;
        float sum = 0.f;
        if (r < m_loc) {
          const float* src = p.ssq_in + (int64_t)(row0 + r) * p.ssq_in_slots + gl * chunk;
          if ((chunk & 3) == 0) {
            // 8 loads in flight, added in slot order: with one load per trip the 64-slot chunk of a row (16 tiles x 8
            // splits of the producing GEMM, two lanes per row) was ~16 dependent L2 round trips, 2 us before the first
            // x tile could be transformed
            int i = 0;
            for (; i + 32 <= chunk; i += 32) {
              float4 v[8];
#pragma unroll
              for (int u = 0; u < 8; ++u) v[u] = *reinterpret_cast<const float4*>(src + i + 4 * u);
#pragma unroll
              for (int u = 0; u < 8; ++u) { sum += v[u].x; sum += v[u].y; sum += v[u].z; sum += v[u].w; }
            }
            for (; i < chunk; i += 4) {
              const float4 v = *reinterpret_cast<const float4*>(src + i);
              sum += v.x; sum += v.y; sum += v.z; sum += v.w;
            }
          } else {
            for (int i = 0; i < chunk; ++i) sum += src[i];
          }
        }
        for (int o = 1; o < G; o <<= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        if (r < mpad && gl == 0) {
          const float var = sum / (float)p.k + p.eps;
          rstd_s[r] = eager ? 1.0f / sqrtf(var) : rsqrtf(var);
        }
      }
      bar_sync_producers();
    }
    // the norm weights of the NEXT k-block are requested before waiting for this one's x tile: the global-load
    // latency (~0.7 us) would otherwise sit on the critical path of every k-block
    int stage = 0;
    uint32_t ph = 0;
    for (int i = 0; normed && i < nkb; ++i) {       // plain GEMMs: the x copies complete on full_bar themselves
      const uint4 wv = wv_next;
      if (normed && i + 1 < nkb) wv_next = ldg_v4(p.norm_w + (int64_t)(kb_begin + i + 1) * G_BK + (tp & 7) * 8);
      mbar_wait(raw_bar + stage, ph);
      if (normed) {
        uint8_t* xs = smem + stage * stage_bytes + G_WTILE_BYTES;
#pragma unroll
        for (int j = 0; j < NCH; ++j) {
          const int idx = tp + 128 * j;
          const int r = idx >> 3, c = idx & 7;
          if (r < m_loc) {
            uint4* cell = reinterpret_cast<uint4*>(xs + r * 128 + ((c ^ (r & 7)) << 4));
            *cell = norm8(*cell, wv, rstd_s[r], eager);
          }
        }
        fence_proxy_async();
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(full_bar + stage);
      if (threadIdx.x == 64 && i == 0) G_TRACE(3);
      if (++stage == NS) { stage = 0; ph ^= 1; }
    }
  }

  // =========================== all threads.  Everything that does not need the accumulator happens BEFORE the wait
  // for it, i.e. while the last MMAs are still running: the (long complete) cluster barrier that says every peer has
  // initialised its receive barrier (~0.25 us of instruction latency), and the first residual / bias vector.
  if (warp < 2) pdl_wait();                     // the epilogue below reads res / writes out
  if (S > 1 || (BS > 1 && warp != 0)) cluster_wait();
  // batch split: nobody leaves while a peer may still signal one of its barriers (arrival now, wait at the very end)
  if (BS > 1) cluster_arrive();
  const int rows_own = G_BN / S;
  const int rstride = mpad + 4;
  const int groups = rows_own >> 3;
  const int n0 = tile * G_BN + rank * rows_own;
  const int items = m_loc * groups;
  uint4 pre = make_uint4(0, 0, 0, 0);
  if ((int)threadIdx.x < items) {
    const int g8 = threadIdx.x / m_loc, m = threadIdx.x - g8 * m_loc;
    if (p.epilogue == B2TTS_GEMM_RESIDUAL) pre = ldg_v4(p.res + (int64_t)(row0 + m) * p.res_stride + n0 + g8 * 8);
    else if (p.epilogue == B2TTS_GEMM_BIAS && p.bias) pre = ldg_v4(p.bias + n0 + g8 * 8);
  }
  // accumulator complete => this CTA's ring is free
  mbar_wait(accum_bar, 0);
  tc_fence_after();
  if (threadIdx.x == 0) G_TRACE(4);

  // Split-K exchange.  Every CTA spills its 128 x mpad fp32 partial to a staging area that aliases
  // its own (now idle) ring, rows = features, padded by 4 floats (conflict-free 16-byte stores, and
  // the reduce reads it conflict-free with lanes along m).  Feature rows of one owner are contiguous,
  // so ONE bulk async copy per peer (shared::cta -> shared::cluster through the TMA unit) moves the
  // slice into the owner's receive buffer recv[src][row][mpad+4] and completes on the OWNER's
  // mbarrier: no cluster barrier on the critical path, no per-thread remote stores.  (Tried and removed:
  // st.async straight from the TMEM registers into the peers' buffers - no staging, no proxy fence, no closing
  // cluster barrier, yet 1.2 us slower per launch: 2048 16-byte remote stores per CTA cost more than 8 bulk copies.)
  float* stg = reinterpret_cast<float*>(smem);
  const float* red = (S > 1) ? recv : stg;
  if (warp >= 2) {
    const int q = warp & 3;                     // TMEM lane quarter this warp may read
    const int f = q * 32 + lane;                // feature row of the tile held by this thread
    float* row = stg + (int64_t)f * rstride;
    for (int c0 = 0; c0 < mpad; c0 += 64) {
      uint32_t v[64];
      const int nq = min(4, (mpad - c0) >> 4);  // 16-column quarters in this pass
      if (nkb > 0) {
#pragma unroll
        for (int qq = 0; qq < 4; ++qq)
          if (qq < nq) tmem_ld16_nowait(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(c0 + 16 * qq), v + 16 * qq);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int j = 0; j < 64; ++j) v[j] = 0u;
      }
#pragma unroll
      for (int qq = 0; qq < 4; ++qq)
        if (qq < nq) {
#pragma unroll
          for (int j = 0; j < 16; j += 4) {
            const int jj = 16 * qq + j;
            *reinterpret_cast<uint4*>(row + c0 + jj) = make_uint4(v[jj], v[jj + 1], v[jj + 2], v[jj + 3]);
          }
        }
    }
    if (S > 1) {
      fence_proxy_async();                      // generic-proxy stores -> visible to the bulk copies
      bar_sync_producers();
      const int tp = (int)threadIdx.x - 64;
      if (tp < S) {
        // one lane per peer: the S copies are issued side by side (one thread issuing them in turn spent ~0.1 us each)
        const uint32_t chunk = (uint32_t)(rows_own * rstride * 4);
        const uint32_t recv_local = smem_u32(recv) + (uint32_t)rank * chunk;
        const uint32_t bar_local = smem_u32(recv_bar);
        const int peer = (rank + tp) % S;       // spread the traffic
        bulk_s2c(mapa_u32(recv_local, (uint32_t)peer), smem_u32(stg) + (uint32_t)peer * chunk, chunk,
                 mapa_u32(bar_local, (uint32_t)peer));
      }
    }
    if (threadIdx.x == 64) G_TRACE(5);
  }
  tc_fence_before();
  if (S > 1) {
    mbar_wait(recv_bar, 0);                     // all S slices (own included) have landed
    // my receive buffer is complete; peers may retire their staging.  With ssq_out the arrival comes after the epilogue
    // instead: it then also publishes this CTA's per-row sums of squares to rank 0 (below)
    if (!p.ssq_out) cluster_arrive();
  } else {
    __syncthreads();
  }
  if (threadIdx.x == 0) G_TRACE(6);

  // =========================== fixed-order sum of the S partials + fused epilogue.
  // One item = (batch row m, 8 consecutive features of this CTA's slice); returns the sum of squares
  // of the 8 bf16 values it stored.
  {
    auto item = [&](int m, int g8, bool have_pre) -> float {
      float acc[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = 0.f;
      const int rstride = mpad + 4;
      for (int s = 0; s < S; ++s) {
        const float* src = red + ((int64_t)s * rows_own + g8 * 8) * rstride + m;
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += src[j * rstride];
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
        stg_v2(p.out + (int64_t)(row0 + m) * p.out_stride + (n >> 1), w);
        return 0.f;
      }
      if (p.epilogue == B2TTS_GEMM_BIAS && p.bias) {
        const uint4 bw = have_pre ? pre : ldg_v4(p.bias + n);
        acc[0] += bf16lo(bw.x); acc[1] += bf16hi(bw.x); acc[2] += bf16lo(bw.y); acc[3] += bf16hi(bw.y);
        acc[4] += bf16lo(bw.z); acc[5] += bf16hi(bw.z); acc[6] += bf16lo(bw.w); acc[7] += bf16hi(bw.w);
      } else if (p.epilogue == B2TTS_GEMM_RESIDUAL) {
        // bf16(residual + bf16(x W^T)): the projection output is rounded before the add
        const uint4 rw = have_pre ? pre : ldg_v4(p.res + (int64_t)(row0 + m) * p.res_stride + n);
        acc[0] = rbf(acc[0]) + bf16lo(rw.x); acc[1] = rbf(acc[1]) + bf16hi(rw.x);
        acc[2] = rbf(acc[2]) + bf16lo(rw.y); acc[3] = rbf(acc[3]) + bf16hi(rw.y);
        acc[4] = rbf(acc[4]) + bf16lo(rw.z); acc[5] = rbf(acc[5]) + bf16hi(rw.z);
        acc[6] = rbf(acc[6]) + bf16lo(rw.w); acc[7] = rbf(acc[7]) + bf16hi(rw.w);
      }
      uint4 w;
      w.x = pack_bf16(acc[0], acc[1]); w.y = pack_bf16(acc[2], acc[3]);
      w.z = pack_bf16(acc[4], acc[5]); w.w = pack_bf16(acc[6], acc[7]);
      stg_v4(p.out + (int64_t)(row0 + m) * p.out_stride + n, w);
      float q = 0.f;
      q = fmaf(bf16lo(w.x), bf16lo(w.x), q); q = fmaf(bf16hi(w.x), bf16hi(w.x), q);
      q = fmaf(bf16lo(w.y), bf16lo(w.y), q); q = fmaf(bf16hi(w.y), bf16hi(w.y), q);
      q = fmaf(bf16lo(w.z), bf16lo(w.z), q); q = fmaf(bf16hi(w.z), bf16hi(w.z), q);
      q = fmaf(bf16lo(w.w), bf16lo(w.w), q); q = fmaf(bf16hi(w.w), bf16hi(w.w), q);
      return q;
    };
    if (p.ssq_out) {
      // the items are spread over all threads as below; every item leaves the sum of squares of its 8 stored values
      // in the (now dead) first cell it read, and one thread per batch row then adds the row's feature groups IN ORDER:
      // a deterministic partial sum of squares of the stored row slice, slot = (tile, rank).  (One thread per row
      // walking its groups itself left 2/3 of the CTA idle: 2.3 us of epilogue against 1.1 us without ssq.)
      // The S CTAs of a tile then combine their per-row sums inside the cluster (remote store into rank 0's shared
      // memory, the closing cluster barrier, rank 0 adds in rank order): ONE slot per (row, tile).  With a slot per (tile,
      // rank) every CTA of the consuming GEMM summed tiles x S partials per row - 256 CTAs reading the same 32 KB
      // at once made a hot spot in L2 that held the first fused-norm tile back by ~3 us.
      const int slots = p.n / G_BN;
      float* qcell = const_cast<float*>(red);
      for (int it = threadIdx.x; it < items; it += G_THREADS) {
        const int g8 = it / m_loc, m = it - g8 * m_loc;
        const float q = item(m, g8, it == (int)threadIdx.x);
        qcell[(int64_t)(g8 * 8) * (mpad + 4) + m] = q;
      }
      __syncthreads();
      for (int m = threadIdx.x; m < m_loc; m += G_THREADS) {
        float q = 0.f;
        for (int g8 = 0; g8 < groups; ++g8) q += qcell[(int64_t)(g8 * 8) * (mpad + 4) + m];
        if (S == 1) p.ssq_out[(int64_t)(row0 + m) * slots + tile] = q;
        else st_cluster_f32(mapa_u32(smem_u32(qx_s + rank * mpad + m), 0u), q);
      }
      if (S > 1) {
        cluster_arrive();                       // release: the remote stores above; also "my receive buffer is complete"
        cluster_wait();
        if (rank == 0)
          for (int m = threadIdx.x; m < m_loc; m += G_THREADS) {
            float q = 0.f;
            for (int r = 0; r < S; ++r) q += qx_s[r * mpad + m];
            p.ssq_out[(int64_t)(row0 + m) * slots + tile] = q;
          }
      }
    } else {
      // item it = (feature group g8 = it / m_loc, batch row m = it % m_loc): lanes along m read red conflict-free
      for (int it = threadIdx.x; it < items; it += G_THREADS) {
        const int g8 = it / m_loc;
        item(it - g8 * m_loc, g8, it == (int)threadIdx.x);
      }
    }
  }
  if (S > 1 && !p.ssq_out) cluster_wait();      // peers' copies out of my staging area have completed
  if (BS > 1) cluster_wait();
  __syncthreads();
  if (threadIdx.x == 0) G_TRACE(7);
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

static size_t gemm_smem_bytes(int mpad, int stages, int splits) {
  return (size_t)stages * (G_WTILE_BYTES + mpad * 128) + (splits > 1 ? (size_t)(mpad + 4) * 512 : 0) + 384 + (size_t)mpad * 4 +
         (splits > 1 ? (size_t)splits * mpad * 4 : 0) + 1024;
}

// ring depth: the deepest ring that still lets two CTAs share an SM (the next GEMM's weight prefetch
// then overlaps this one's epilogue under programmatic dependent launch), else the deepest that fits.
static int choose_stages(int mpad, int splits) {
  const size_t two_per_sm = 113 * 1024, one_per_sm = 227 * 1024;
  if (const char* e = getenv("B2TTS_GEMM_STAGES")) {            // experiments only
    const int st = atoi(e);
    if (st >= 2 && st <= G_MAX_STAGES && gemm_smem_bytes(mpad, st, splits) <= one_per_sm) return st;
  }
  for (int st : {4, 3})
    if (gemm_smem_bytes(mpad, st, splits) <= two_per_sm) return st;
  for (int st : {4, 3, 2})
    if (gemm_smem_bytes(mpad, st, splits) <= one_per_sm) return st;
  return 0;
}

static int choose_splits(int n_tiles, int kblocks, int mpad) {
  // the largest K-split (cluster size <= 8: 16-CTA clusters schedule badly) whose CTAs are all
  // resident at once given the shared-memory footprint, with at least two k-blocks per CTA
  int best = 1;
  for (int s = 2; s <= 8; s *= 2) {
    if (s * 2 > kblocks) break;
    const int st = choose_stages(mpad, s);
    if (st < 2) break;
    const int per_sm = gemm_smem_bytes(mpad, st, s) <= 113 * 1024 ? 2 : 1;
    if (n_tiles * s <= sm_count() * per_sm) best = s;
  }
  return best;
}

// Batch split instead of K split (GemmParams::bsplit): for K ranges short enough that one CTA issuing all of their MMAs
// (~0.16 us per k-block) is cheaper than the drain + DSMEM exchange + closing barrier of a K split (~3 us).
static int choose_bsplit(int n_tiles, int kblocks, int mpad, int nchunks) {
  static int mode = -1, max_kb = 16;
  if (mode < 0) {
    // Off unless asked for: measured in the frame (64 streams) the batch split LOSES - 6.01 ms per frame against 5.26 -
    // because one issuing thread needs ~0.2 us per k-block (0.4 us with the fused-norm hand-over), so a CTA's 16 k-blocks
    // of K = 1024 take 6.4 us where 4 k-blocks + drain + exchange take 4.5 us.
    const char* e = getenv("B2TTS_GEMM_BSPLIT");
    mode = e ? atoi(e) : 0;
    if (const char* k = getenv("B2TTS_GEMM_BSPLIT_MAX_KB")) max_kb = atoi(k);
  }
  if (!mode || nchunks != 1 || kblocks > max_kb || kblocks < 2) return 1;
  for (int bs : {4, 2})
    if (mpad % (16 * bs) == 0 && n_tiles * bs <= 2 * sm_count()) return bs;
  return 1;
}

template <int NCH>
static int launch_gemm(const GemmParams& p, const CUtensorMap& tmap, cudaStream_t st, int pdl) {
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(gemm_stream_kernel<NCH>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return cuda_fail(e, "gemm_bf16(smem attr)");
    e = cudaFuncSetAttribute(gemm_stream_kernel<NCH>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    if (e != cudaSuccess) return cuda_fail(e, "gemm_bf16(cluster attr)");
    configured = true;
  }
  cudaLaunchConfig_t cfg{};
  const int cluster = p.bsplit > 1 ? p.bsplit : p.splits;
  cfg.gridDim = dim3(cluster, (p.n / G_BN) * p.nchunks, 1);
  cfg.blockDim = dim3(G_THREADS, 1, 1);
  cfg.dynamicSmemBytes = gemm_smem_bytes(p.mpad, p.stages, p.splits);
  cfg.stream = st;
  cudaLaunchAttribute attrs[2];
  int na = 0;
  if (cluster > 1) {
    attrs[na].id = cudaLaunchAttributeClusterDimension;
    attrs[na].val.clusterDim.x = cluster;
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
  cudaError_t e = cudaLaunchKernelEx(&cfg, gemm_stream_kernel<NCH>, p, tmap);
  count_launch();
  if (e != cudaSuccess) return cuda_fail(e, "gemm_bf16");
  return B2TTS_OK;
}

}  // namespace b2

extern "C" int b2tts_gemm_bf16_supported(int m, int n, int k) {
  return m >= 1 && m <= b2::G_MAX_ROWS && n >= b2::G_BN && n % b2::G_BN == 0 && k >= b2::G_BK && k % b2::G_BK == 0 &&
         (int64_t)(n / b2::G_BN) * ((m + b2::G_MAX_M - 1) / b2::G_MAX_M) <= 65535;
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

namespace b2 {
// x[M, K] (bf16, row stride in elements) as a 2-D tensor map: box = 64 k x mpad rows, SWIZZLE_128B, rows
// beyond M zero-filled.  The driver entry point is resolved through the runtime (no libcuda link).
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int make_x_tensor_map(CUtensorMap* tm, const void* x, int64_t row_stride, int m, int k, int mpad, const char* what) {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !sym) {
      set_error("%s: cuTensorMapEncodeTiled unavailable", what);
      return B2TTS_ECUDA;
    }
    fn = reinterpret_cast<EncodeTiledFn>(sym);
  }
  const cuuint64_t dims[2] = {(cuuint64_t)k, (cuuint64_t)m};
  const cuuint64_t strides[1] = {(cuuint64_t)row_stride * 2};
  const cuuint32_t box[2] = {(cuuint32_t)G_BK, (cuuint32_t)mpad};
  const cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(x), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("%s: cuTensorMapEncodeTiled failed (%d) for x [%d, %d] stride %lld", what, (int)r, m, k, (long long)row_stride);
    return B2TTS_ECUDA;
  }
  return B2TTS_OK;
}

static unsigned long long* g_trace_base = nullptr;   // debug only (b2tts_debug_gemm_trace)
static size_t g_trace_slots = 0, g_trace_next = 0;
static int gemm_impl(const char* what, const void* x, int64_t x_row_stride, const void* w_packed,
                     const void* bias, const void* residual, int64_t residual_row_stride,
                     void* out, int64_t out_row_stride, int m, int n, int k,
                     int epilogue, int rounding, int splits, int pdl,
                     const void* norm_weight, const float* ssq_in, int ssq_in_slots, float eps, float* ssq_out,
                     const void* prefetch_next, size_t prefetch_bytes, void* stream) {
  B2_REQUIRE(m >= 0 && n > 0 && k > 0, B2TTS_EINVAL, "%s: m=%d n=%d k=%d", what, m, n, k);
  B2_REQUIRE(!prefetch_next || (aligned16(prefetch_next) && prefetch_bytes % G_WTILE_BYTES == 0), B2TTS_EALIGN,
             "%s: prefetch_next must be a packed weight image (16-byte aligned, a multiple of %d bytes)", what, G_WTILE_BYTES);
  if (m == 0) return B2TTS_OK;
  B2_REQUIRE(b2tts_gemm_bf16_supported(m, n, k), B2TTS_EUNSUPPORTED,
             "%s: m=%d (1..%d), n=%d (multiple of %d), k=%d (multiple of %d)", what, m, G_MAX_ROWS, n, G_BN, k, G_BK);
  B2_REQUIRE(x && w_packed && out, B2TTS_EINVAL, "%s: null pointer", what);
  B2_REQUIRE(epilogue == B2TTS_GEMM_BIAS || epilogue == B2TTS_GEMM_RESIDUAL || epilogue == B2TTS_GEMM_SILU_MUL,
             B2TTS_EINVAL, "%s: epilogue %d", what, epilogue);
  B2_REQUIRE(epilogue != B2TTS_GEMM_RESIDUAL || residual, B2TTS_EINVAL, "%s: residual epilogue without residual", what);
  B2_REQUIRE(aligned16(x) && aligned16(w_packed) && aligned16(out) && x_row_stride % 8 == 0 && out_row_stride % 4 == 0 &&
                 (epilogue == B2TTS_GEMM_SILU_MUL || out_row_stride % 8 == 0) &&
                 (!bias || aligned16(bias)) && (!residual || (aligned16(residual) && residual_row_stride % 8 == 0)),
             B2TTS_EALIGN, "%s: pointers / strides must be 16-byte aligned", what);
  B2_REQUIRE(!norm_weight || (ssq_in && ssq_in_slots > 0 && aligned16(norm_weight) && aligned16(ssq_in)), B2TTS_EINVAL,
             "%s: fused norm needs ssq_in with >= 1 slot per row (16-byte aligned)", what);
  B2_REQUIRE(!ssq_out || epilogue != B2TTS_GEMM_SILU_MUL, B2TTS_EINVAL, "%s: ssq_out with the SiLU*mul epilogue", what);
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
  // batches above the chunk size are cut into row chunks (one cluster per (feature tile, chunk)).  The per-CTA costs that
  // grow with the rows - x tile, in-place norm transform, TMEM drain, split-K exchange - make 64-row chunks the better cut
  // (B2TTS_GEMM_CHUNK_ROWS overrides: 16 .. 128, a multiple of 16)
  static int chunk_cfg = -1;
  if (chunk_cfg < 0) {
    const char* e = getenv("B2TTS_GEMM_CHUNK_ROWS");
    chunk_cfg = e ? atoi(e) : 64;                 // measured: 128 streams 7.81 -> 6.39 ms per frame, 256: 10.3 -> 8.7, 64: 5.24 -> 5.20
    if (chunk_cfg < 16 || chunk_cfg > G_MAX_M || chunk_cfg % 16) chunk_cfg = G_MAX_M;
  }
  p.chunk_rows = (splits >= 0 && m > chunk_cfg) ? chunk_cfg : G_MAX_M;      // an explicit batch split takes the whole batch
  // Plain x (no fused norm: above 64 rows the norms are separate launches) cut into chunks: a k-block costs ~0.2 us
  // there (0.45 with the in-place norm transform), so (measured at 256 streams, tools/plan_sweep.py)
  //   * K <= 1024 is not split at all: 16 k-blocks per CTA cost less than drain + DSMEM exchange of 64-row partials
  //     (predictor qkv: -0.16 ms per frame at 256 streams; 128 streams with separate norms 5.94 -> 5.70 ms);
  //   * an unsplit launch takes the chunk size that fills ONE wave of 2 CTAs per SM best (predictor gate_up: 48 tiles x
  //     6 chunks of 48 rows = 288 CTAs, -0.19 ms; talker gate_up: 96 tiles x 3 chunks of 96 rows = 288 CTAs instead of
  //     384 in two waves, -0.23 ms); when nothing fits in one wave, the largest chunk (fewest waves, fewest L2 re-reads).
  const bool auto_plan = splits == 0;
  static const bool plain_rule = [] { const char* e = getenv("B2TTS_GEMM_PLAIN_RULE"); return !(e && e[0] == '0'); }();
  static const bool chunk_forced = getenv("B2TTS_GEMM_CHUNK_ROWS") != nullptr;
  if (plain_rule && splits == 0 && !chunk_forced && m > chunk_cfg) {
    // (the fill rule also holds with the fused norm: talker gate_up at 128 streams, 3 chunks of 48 rows, -0.12 ms)
    const int tiles = n / G_BN, kb = k / G_BK;
    const int units64 = tiles * ((m + 63) / 64);
    if ((!norm_weight && kb <= 16) || choose_splits(units64, kb, 64) == 1) {
      int best_chunk = G_MAX_M, best_ctas = 0;
      for (int ch : {128, 96, 64, 48}) {
        const int ctas = tiles * ((m + ch - 1) / ch);
        if (ctas <= 2 * sm_count() && ctas > best_ctas) { best_ctas = ctas; best_chunk = ch; }
      }
      p.chunk_rows = best_chunk;
      splits = 1;
    }
  }
  // experiments: B2TTS_GEMM_PLAN="n,k,m_lo,m_hi,chunk,splits;..." overrides the cut and the K split of matching launches
  {
    struct PlanEntry { int n, k, mlo, mhi, chunk, splits; };
    static PlanEntry plan[32];
    static int nplan = -1;
    if (nplan < 0) {
      nplan = 0;
      if (const char* e = getenv("B2TTS_GEMM_PLAN")) {
        while (*e && nplan < 32) {
          PlanEntry pe{};
          int used = 0;
          if (sscanf(e, "%d,%d,%d,%d,%d,%d%n", &pe.n, &pe.k, &pe.mlo, &pe.mhi, &pe.chunk, &pe.splits, &used) == 6) plan[nplan++] = pe;
          else break;
          e += used;
          if (*e == ';') ++e;
        }
      }
    }
    for (int i = 0; i < nplan; ++i)
      if (plan[i].n == n && plan[i].k == k && m >= plan[i].mlo && m <= plan[i].mhi && auto_plan) {
        if (plan[i].chunk >= 16 && plan[i].chunk <= G_MAX_M && plan[i].chunk % 16 == 0) p.chunk_rows = m > plan[i].chunk ? plan[i].chunk : G_MAX_M;
        splits = plan[i].splits;
        break;
      }
  }
  p.nchunks = (m + p.chunk_rows - 1) / p.chunk_rows;
  B2_REQUIRE((int64_t)(n / G_BN) * p.nchunks <= 65535, B2TTS_EUNSUPPORTED, "%s: m=%d n=%d: too many tiles", what, m, n);
  p.mpad = p.nchunks > 1 ? p.chunk_rows : (m + 15) & ~15;
  p.kblocks = k / G_BK;
  p.epilogue = epilogue;
  p.rounding = rounding;
  // splits < 0: a batch split of -splits CTAs (see GemmParams::bsplit); 0: chosen here
  int bs = 1;
  if (splits < 0) {
    bs = -splits;
    B2_REQUIRE((bs == 2 || bs == 4) && p.nchunks == 1 && p.mpad % (16 * bs) == 0, B2TTS_EINVAL,
               "%s: batch split %d needs 2 or 4 and a padded batch (%d rows) that is a multiple of %d", what, bs, p.mpad, 16 * bs);
  } else if (splits == 0) {
    bs = choose_bsplit((n / G_BN), p.kblocks, p.mpad, p.nchunks);
  }
  int s = bs > 1 ? 1 : (splits > 0 ? splits : choose_splits((n / G_BN) * p.nchunks, p.kblocks, p.mpad));
  B2_REQUIRE((s == 1 || s == 2 || s == 4 || s == 8 || s == 16) && s <= p.kblocks, B2TTS_EINVAL,
             "%s: splits %d must be 1,2,4,8,16 and <= k/64", what, s);
  p.splits = s;
  p.bsplit = bs;
  if (bs > 1) {
    p.mpad /= bs;                                 // rows per CTA: 16 or 32
    // the deepest ring that still lets two CTAs share an SM, at most the whole K range
    int st = G_MAX_STAGES;
    static int deep = -1;
    if (deep < 0) { const char* e = getenv("B2TTS_GEMM_BSPLIT_DEEP"); deep = e ? atoi(e) : 0; }     // experiments: one CTA per SM
    while (st > 2 && gemm_smem_bytes(p.mpad, st, 1) > (deep ? 227 : 113) * 1024) --st;
    p.stages = st < p.kblocks ? st : (p.kblocks < 2 ? 2 : p.kblocks);
  } else {
    p.stages = choose_stages(p.mpad, s);
  }
  B2_REQUIRE(p.stages >= 2, B2TTS_EUNSUPPORTED, "%s: no shared-memory plan for m=%d", what, m);
  p.tmem_cols = 32;
  while (p.tmem_cols < p.mpad) p.tmem_cols <<= 1;
  p.norm_w = static_cast<const __nv_bfloat16*>(norm_weight);
  p.ssq_in = ssq_in;
  p.ssq_in_slots = ssq_in_slots;
  p.eps = eps;
  p.ssq_out = ssq_out;
  p.ssq_lanes = 1;
  p.pf_ptr = prefetch_bytes ? static_cast<const uint8_t*>(prefetch_next) : nullptr;
  p.pf_bytes = prefetch_bytes;
  p.trace = nullptr;
  if (g_trace_base && g_trace_next < g_trace_slots) p.trace = g_trace_base + (g_trace_next++) * 4096 * 8;
  if (norm_weight) {
    int g = 128 / p.mpad;
    if (g > 8) g = 8;
    if (g < 1) g = 1;
    while (g > 1 && (ssq_in_slots % g)) g >>= 1;
    // float4 loads need 16-byte aligned chunks
    while (g > 1 && ((ssq_in_slots / g) % 4)) g >>= 1;
    p.ssq_lanes = g;
    B2_REQUIRE(ssq_in_slots % 4 == 0 || g == 1, B2TTS_EALIGN, "%s: ssq_in_slots %d", what, ssq_in_slots);
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int nch = p.mpad / 16;                  // 16-byte chunks per producer thread per k-block
  CUtensorMap tmap;
  int rc = make_x_tensor_map(&tmap, x, x_row_stride, m, k, p.mpad, what);
  if (rc) return rc;
  if (nch <= 1) return launch_gemm<1>(p, tmap, st, pdl);
  if (nch <= 2) return launch_gemm<2>(p, tmap, st, pdl);
  if (nch <= 4) return launch_gemm<4>(p, tmap, st, pdl);
  return launch_gemm<8>(p, tmap, st, pdl);
}
}  // namespace b2

extern "C" int b2tts_gemm_bf16(const void* x, int64_t x_row_stride, const void* w_packed,
                               const void* bias, const void* residual, int64_t residual_row_stride,
                               void* out, int64_t out_row_stride, int m, int n, int k,
                               int epilogue, int rounding, int splits, int pdl, void* stream) {
  return b2::gemm_impl("gemm_bf16", x, x_row_stride, w_packed, bias, residual, residual_row_stride, out, out_row_stride,
                       m, n, k, epilogue, rounding, splits, pdl, nullptr, nullptr, 0, 0.f, nullptr, nullptr, 0, stream);
}

extern "C" int b2tts_gemm_ssq_slots(int m, int n, int k, int splits) {
  if (n <= 0 || k <= 0 || n % b2::G_BN || k % b2::G_BK) return 0;
  (void)m; (void)splits;                        // one slot per 128-feature tile, whatever the K split or the row chunks
  return n / b2::G_BN;
}

extern "C" int b2tts_gemm_norm_bf16(const void* x, int64_t x_row_stride, const void* w_packed,
                                    const void* bias, const void* residual, int64_t residual_row_stride,
                                    void* out, int64_t out_row_stride, int m, int n, int k,
                                    int epilogue, int rounding, int splits, int pdl,
                                    const void* norm_weight, const float* ssq_in, int ssq_in_slots, float eps,
                                    float* ssq_out, void* stream) {
  return b2::gemm_impl("gemm_norm_bf16", x, x_row_stride, w_packed, bias, residual, residual_row_stride, out,
                       out_row_stride, m, n, k, epilogue, rounding, splits, pdl, norm_weight, ssq_in, ssq_in_slots, eps,
                       ssq_out, nullptr, 0, stream);
}

extern "C" int b2tts_gemm_prefetch_bf16(const void* x, int64_t x_row_stride, const void* w_packed,
                                        const void* bias, const void* residual, int64_t residual_row_stride,
                                        void* out, int64_t out_row_stride, int m, int n, int k,
                                        int epilogue, int rounding, int splits, int pdl,
                                        const void* norm_weight, const float* ssq_in, int ssq_in_slots, float eps,
                                        float* ssq_out, const void* prefetch_next, size_t prefetch_bytes, void* stream) {
  return b2::gemm_impl("gemm_prefetch_bf16", x, x_row_stride, w_packed, bias, residual, residual_row_stride, out,
                       out_row_stride, m, n, k, epilogue, rounding, splits, pdl, norm_weight, ssq_in, ssq_in_slots, eps,
                       ssq_out, prefetch_next, prefetch_bytes, stream);
}

// =====================================================================================================
// Persistent GEMM CHAIN: up to 4 dependent projections of the decoder layer in ONE launch
// (o_proj + residual -> gate_up (+ post-attention RMSNorm in, SiLU*mul out) -> down_proj + residual ->
// the NEXT layer's qkv_proj (+ input RMSNorm in); models/qwen3_tts_share.py:163-176 of the reference).
//
// Why: at decode batch sizes every projection is a few microseconds of weight streaming wrapped in
// ~5 us of dependent latencies (kernel boundary, x fetch, accumulator drain, split-K exchange,
// epilogue), during which HBM idles: a per-GEMM launch reaches < 30 % of the HBM roofline however
// well its inner loop streams.  Weights do not depend on activations, so here ONE persistent CTA per SM
// runs a static schedule over all GEMMs of the chain and its producer thread keeps a deep ring of
// weight tiles (10 x 16 KiB per SM, 23 MB chip-wide) full ACROSS GEMM boundaries: while the CTAs
// exchange split-K partials and cross the grid barrier of GEMM i, HBM is already streaming GEMM i+1.
//
//   warp 0   W producer: 16 KiB bulk copies of packed weight tiles, runs ahead through the whole chain
//   warp 1   MMA issuer: tcgen05.mma into one of two TMEM accumulators (the drain of item j overlaps
//            the MMAs of item j+1)
//   warp 2   x producer: 2-D tensor-map copies of the activation tile (SWIZZLE_128B = UMMA B operand);
//            starts GEMM i once grid barrier B(i-1) says its input rows are complete
//   warps 3-6  fused input RMSNorm on the landed x tile; accumulator drain (tcgen05.ld) to the fp32
//            split-K scratch in L2; after grid barrier A(i) the REDUCE phase: every (row, 16 features)
//            unit of the output is summed over the splits in a fixed order (deterministic), gets the
//            fused epilogue (residual add / SiLU*mul / plain) and its sum of squares, then barrier B(i).
// Work item = (128-feature tile, K split); items are dealt round-robin to the CTAs; the split count
// per GEMM balances k-blocks per CTA.  Grid barriers are monotonic counters in the workspace, reset by
// the last CTA to leave, so a CUDA graph can replay the launch without a memset.
// Rounding points are those of gemm_stream_kernel (bf16(residual + bf16(acc)), norm of the rounded sum).
namespace b2 {
constexpr int C_MAX_OPS = 4;
constexpr int C_THREADS = 480;                   // warps 0-2 producers/MMA, 3-6 epilogue, 7-14 fused-norm warps
constexpr int C_NORM0 = 224;                     // first norm-warp thread
constexpr int C_WORKERS = 384;                   // epilogue + norm threads: the reduce phase uses all of them
constexpr int C_EPI0 = 96;                       // first epilogue thread
constexpr int C_SSQ_FEATS = 128;                 // output features per ssq slot (16 lanes x 8 features)
constexpr int C_NORMW_MAX = 4096;                // largest K whose norm weight is staged in shared memory
constexpr int C_SYNC_WORDS = 16;                 // 2 * C_MAX_OPS barrier counters + exit counter, padded

struct ChainOp {
  const uint8_t* wp;
  const __nv_bfloat16* norm_w;
  const float* ssq_in;
  const __nv_bfloat16* res;
  __nv_bfloat16* out;
  float* ssq_out;
  int64_t res_stride, out_stride;
  int n, kblocks, n_tiles, splits, epilogue, ssq_in_slots, ssq_lanes, k;
};

struct ChainParams {
  ChainOp op[C_MAX_OPS];
  int nops, m, mpad, tmem_cols, nw, nx, cg, rounding;   // nw == nx ring slots, released in groups of cg k-blocks
  float eps;
  float* partial;                                // [splits][n][mpad] fp32, reused by every op
  unsigned* sync;                                // C_SYNC_WORDS counters, zero between launches
  unsigned long long* trace;                     // debug: 8 globaltimer stamps per (CTA, op) (null = off)
};
// debug: per-k-block clock64 stamps of CTA 0, op 1: stream e (0 TMA x issue, 1 x landed at the hop, 2 hop done, 3 MMA has W,
// 4 MMA has x) x 64 k-blocks at trace[8192 + e * 64 + idx]
#define C_KB_TRACE(e, idx) do { if (p.trace && cta == 0 && i == 2 && (idx) < 64) p.trace[8192 + (e) * 64 + (idx)] = (unsigned long long)clock64(); } while (0)
#define C_TRACE(op_i, slot) do { if (p.trace && te == 0) p.trace[((int64_t)cta * C_MAX_OPS + (op_i)) * 8 + (slot)] = gtimer(); } while (0)

__device__ __forceinline__ unsigned ld_acquire_gpu(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void grid_wait(const unsigned* ctr, unsigned target) {
  while (ld_acquire_gpu(ctr) < target) { }
}
__device__ __forceinline__ void grid_arrive(unsigned* ctr) {
  asm volatile("red.release.gpu.global.add.u32 [%0], 1;" :: "l"(ctr) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }
__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 2, 128;" ::: "memory"); }
__device__ __forceinline__ void workers_bar() { asm volatile("bar.sync 3, 384;" ::: "memory"); }

// Reduce phase of one chain op (all C_WORKERS worker threads of every CTA): unit = (row m, 8 consecutive features), lanes
// run along the features: 32-byte scratch reads per split and 16-byte stores, all coalesced; the sums of squares of
// 16 lanes (128 features) fold in a fixed shuffle tree into one ssq slot.  Splits are added in order: deterministic.
__device__ __forceinline__ void chain_reduce(const ChainParams& p, const ChainOp& op, int S, int cta, int G, int tr, int lane,
                                             bool eager) {
    const int M = p.m;
    const int N = op.n;
    const int per_row = N >> 3;                       // units per row (a multiple of 16)
    const int units = M * per_row;
    const int slots = N / C_SSQ_FEATS;
    for (int u0 = cta * C_WORKERS + (tr & ~31); u0 < units; u0 += G * C_WORKERS) {
      const int u = u0 + lane;
      const bool valid = u < units;
      const int m = valid ? u / per_row : 0;
      const int n = valid ? (u - m * per_row) * 8 : 0;
      float acc[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] = 0.f;
      if (valid) {
        const float* src = p.partial + (int64_t)m * N + n;
        for (int sp = 0; sp < S; ++sp, src += (int64_t)M * N) {
          const float4 a = __ldcg(reinterpret_cast<const float4*>(src));
          const float4 c = __ldcg(reinterpret_cast<const float4*>(src) + 1);
          acc[0] += a.x; acc[1] += a.y; acc[2] += a.z; acc[3] += a.w;
          acc[4] += c.x; acc[5] += c.y; acc[6] += c.z; acc[7] += c.w;
        }
      }
      float qsum = 0.f;
      if (op.epilogue == B2TTS_GEMM_SILU_MUL) {
        if (valid) {
          float o[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float gt = rbf(acc[e]), up = rbf(acc[4 + e]);
            o[e] = eager ? rbf(gsilu_f(gt)) * up : gsilu_f(gt) * up;
          }
          uint2 w;
          w.x = pack_bf16(o[0], o[1]); w.y = pack_bf16(o[2], o[3]);
          stg_v2(op.out + (int64_t)m * op.out_stride + (n >> 1), w);
        }
        continue;
      }
      if (valid) {
        if (op.epilogue == B2TTS_GEMM_RESIDUAL) {
          const uint4 rw = __ldcg(reinterpret_cast<const uint4*>(op.res + (int64_t)m * op.res_stride + n));
          acc[0] = rbf(acc[0]) + bf16lo(rw.x); acc[1] = rbf(acc[1]) + bf16hi(rw.x);
          acc[2] = rbf(acc[2]) + bf16lo(rw.y); acc[3] = rbf(acc[3]) + bf16hi(rw.y);
          acc[4] = rbf(acc[4]) + bf16lo(rw.z); acc[5] = rbf(acc[5]) + bf16hi(rw.z);
          acc[6] = rbf(acc[6]) + bf16lo(rw.w); acc[7] = rbf(acc[7]) + bf16hi(rw.w);
        }
        uint4 w;
        w.x = pack_bf16(acc[0], acc[1]); w.y = pack_bf16(acc[2], acc[3]);
        w.z = pack_bf16(acc[4], acc[5]); w.w = pack_bf16(acc[6], acc[7]);
        stg_v4(op.out + (int64_t)m * op.out_stride + n, w);
        qsum = fmaf(bf16lo(w.x), bf16lo(w.x), qsum); qsum = fmaf(bf16hi(w.x), bf16hi(w.x), qsum);
        qsum = fmaf(bf16lo(w.y), bf16lo(w.y), qsum); qsum = fmaf(bf16hi(w.y), bf16hi(w.y), qsum);
        qsum = fmaf(bf16lo(w.z), bf16lo(w.z), qsum); qsum = fmaf(bf16hi(w.z), bf16hi(w.z), qsum);
        qsum = fmaf(bf16lo(w.w), bf16lo(w.w), qsum); qsum = fmaf(bf16hi(w.w), bf16hi(w.w), qsum);
      }
      if (op.ssq_out) {                               // warp-uniform
#pragma unroll
        for (int o = 1; o < 16; o <<= 1) qsum += __shfl_xor_sync(0xffffffffu, qsum, o);
        if (valid && (lane & 15) == 0) op.ssq_out[(int64_t)m * slots + n / C_SSQ_FEATS] = qsum;
      }
    }
}

__global__ void __launch_bounds__(C_THREADS, 1)
gemm_chain_kernel(const ChainParams p, const __grid_constant__ CUtensorMap tm0, const __grid_constant__ CUtensorMap tm1,
                  const __grid_constant__ CUtensorMap tm2, const __grid_constant__ CUtensorMap tm3) {
  extern __shared__ uint8_t smem_dyn[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~uintptr_t(1023));
  const int mpad = p.mpad, NW = p.nw, NX = p.nx, CG = p.cg;          // NW == NX in {4, 8}; CG a power of two
  const uint32_t CGL = 31 - __clz(CG);
  const int xstage = mpad * 128;
  uint8_t* wring = smem;
  uint8_t* xring = wring + NW * G_WTILE_BYTES;
  uint64_t* wfull = reinterpret_cast<uint64_t*>(xring + NX * xstage);
  uint64_t* wempty = wfull + G_MAX_STAGES;
  uint64_t* xraw = wempty + G_MAX_STAGES;
  uint64_t* xfull = xraw + G_MAX_STAGES;
  uint64_t* xempty = xfull + G_MAX_STAGES;
  uint64_t* accfull = xempty + G_MAX_STAGES;
  uint64_t* accempty = accfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accempty + 2);
  float* rstd_s = reinterpret_cast<float*>(tmem_slot + 4);                         // keeps normw_s 16-byte aligned
  __nv_bfloat16* normw_s = reinterpret_cast<__nv_bfloat16*>(rstd_s + G_MAX_M);     // [C_NORMW_MAX], 16-byte aligned

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);       // provably warp-uniform
  const int lane = threadIdx.x & 31;
  const int G = gridDim.x;
  const int cta = blockIdx.x;

  if (threadIdx.x == 0) {
    // wempty[g]: ring slots g*CG .. g*CG+CG-1 of BOTH rings are free again (one tcgen05.commit per CG k-blocks:
    // a commit costs the issuing thread ~0.2 us, so one per k-block per ring would pace the whole stream)
    for (int s = 0; s < NW; ++s) { mbar_init(wfull + s, 1); mbar_init(wempty + s, 1); }
    for (int s = 0; s < NX; ++s) { mbar_init(xraw + s, 1); mbar_init(xfull + s, 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(accfull + b, 1); mbar_init(accempty + b, 4); }
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, (uint32_t)p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // =========================== W producer: never waits for activations (whole warp walks the schedule, one
    // elected lane issues: addresses and barriers stay in uniform registers)
    {
      uint32_t slot = 0, ph = 1;                 // ring slot and "empty" parity as counters: no div/mod per k-block
      for (int i = 0; i < p.nops; ++i) {
        const ChainOp& op = p.op[i];
        const int S = op.splits, items = op.n_tiles * S;
        for (int t = cta; t < items; t += G) {
          const int tile = t / S, s = t - tile * S;
          const int kb0 = (s * op.kblocks) / S, kb1 = ((s + 1) * op.kblocks) / S;     // 32-bit: splits <= 16, kblocks < 2^20
          const uint8_t* src = op.wp + ((int64_t)tile * op.kblocks + kb0) * G_WTILE_BYTES;
          for (int kb = kb0; kb < kb1; ++kb, src += G_WTILE_BYTES) {
            mbar_wait(wempty + (slot >> CGL), ph);
            if (elect_one()) {
              mbar_expect_tx(wfull + slot, G_WTILE_BYTES);
              bulk_g2s(wring + slot * G_WTILE_BYTES, src, G_WTILE_BYTES, wfull + slot);
            }
            __syncwarp();
            if (++slot == (uint32_t)NW) { slot = 0; ph ^= 1; }
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // =========================== MMA issuer: the whole warp walks the schedule (uniform control flow, operand
    // descriptors in uniform registers), one elected lane issues; ring slot / phase / commit group are counters
    {
      const uint32_t idesc = umma_idesc_bf16(mpad);
      uint32_t it = 0, sw = 0, ph = 0, cgc = 0;
      for (int i = 0; i < p.nops; ++i) {
        const ChainOp& op = p.op[i];
        const int S = op.splits, items = op.n_tiles * S;
        unsigned long long wait_w = 0, wait_x = 0;
        int kbi = 0;
        for (int t = cta; t < items; t += G, ++it) {
          const int tile = t / S, s = t - tile * S;
          const int kb0 = (s * op.kblocks) / S, kb1 = ((s + 1) * op.kblocks) / S;     // 32-bit: splits <= 16, kblocks < 2^20
          const uint32_t b = it & 1;
          mbar_wait(accempty + b, ((it >> 1) & 1) ^ 1);
          tc_fence_after();
          const uint32_t d_addr = tmem_base + b * (uint32_t)mpad;
          for (int kb = kb0; kb < kb1; ++kb) {
            const long long c0 = p.trace ? clock64() : 0;
            mbar_wait(wfull + sw, ph);
            const long long c1 = p.trace ? clock64() : 0;
            if (lane == 0) C_KB_TRACE(3, kbi);
            mbar_wait(xfull + sw, ph);
            if (lane == 0) C_KB_TRACE(4, kbi);
            ++kbi;
            if (p.trace) { wait_w += (unsigned long long)(c1 - c0); wait_x += (unsigned long long)(clock64() - c1); }
            tc_fence_after();
            const uint32_t a_addr = smem_u32(wring + sw * G_WTILE_BYTES);
            const uint32_t b_addr = smem_u32(xring + sw * xstage);
            const bool release = ++cgc == (uint32_t)CG;
            if (elect_one()) {
#pragma unroll
              for (int k = 0; k < G_BK / 16; ++k)
                umma_bf16(d_addr, umma_desc_sw128(a_addr + k * 32), umma_desc_sw128(b_addr + k * 32), idesc,
                          (uint32_t)((kb > kb0) || (k > 0)));
              if (release) umma_commit(wempty + (sw >> CGL));
              if (kb == kb1 - 1) umma_commit(accfull + b);
            }
            __syncwarp();
            if (release) cgc = 0;
            if (++sw == (uint32_t)NW) { sw = 0; ph ^= 1; }
          }
        }
        if (p.trace && lane == 0) p.trace[((int64_t)cta * C_MAX_OPS + i) * 8 + 7] = (wait_w << 32) | (wait_x & 0xffffffffull);
      }
    }
    __syncwarp();
  } else if (warp == 2) {
    // =========================== x producer (warp-uniform as well)
    {
      uint32_t slot = 0, ph = 1;
      const uint32_t xbytes = (uint32_t)xstage;
      for (int i = 0; i < p.nops; ++i) {
        const ChainOp& op = p.op[i];
        if (i == 0) {
          pdl_wait();
        } else {
          if (lane == 0) grid_wait(p.sync + 2 * (i - 1) + 1, (unsigned)G);
          __syncwarp();
          fence_proxy_async_all();
        }
        const int S = op.splits, items = op.n_tiles * S;
        const bool normed = op.norm_w != nullptr;
        int kbi = 0;
        for (int t = cta; t < items; t += G) {
          const int tile = t / S, s = t - tile * S;
          const int kb0 = (s * op.kblocks) / S, kb1 = ((s + 1) * op.kblocks) / S;     // 32-bit: splits <= 16, kblocks < 2^20
          for (int kb = kb0; kb < kb1; ++kb) {
            mbar_wait(wempty + (slot >> CGL), ph);
            if (lane == 0) C_KB_TRACE(0, kbi);
            ++kbi;
            // with a fused norm the tile lands "raw" and an epilogue warp releases it after the transform;
            // otherwise the copy completes straight on the barrier the MMA issuer waits on
            uint64_t* landed = normed ? xraw + slot : xfull + slot;
            if (elect_one()) {
            mbar_expect_tx(landed, xbytes);
            // the tensor map must be addressed in param space: selecting a pointer at run time would make the
            // compiler copy the maps to local memory, and every copy would then fetch its descriptor from DRAM
            switch (i) {
              case 0: tma_load_2d(xring + slot * xstage, &tm0, kb * G_BK, 0, landed); break;
              case 1: tma_load_2d(xring + slot * xstage, &tm1, kb * G_BK, 0, landed); break;
              case 2: tma_load_2d(xring + slot * xstage, &tm2, kb * G_BK, 0, landed); break;
              default: tma_load_2d(xring + slot * xstage, &tm3, kb * G_BK, 0, landed); break;
            }
            }
            __syncwarp();
            if (++slot == (uint32_t)NX) { slot = 0; ph ^= 1; }
          }
        }
      }
    }
    __syncwarp();
  } else if (warp >= 7) {
    // =========================== fused-norm warps: ring slot sx belongs to norm warp sx & 7 (or & 3 with a 4-slot
    // ring), which tracks the phase of its slots' xraw barriers; 8 k-blocks are transformed concurrently.  They also
    // take their share of every reduce phase.
    const int wn = warp - 7;
    const int tr = threadIdx.x - C_EPI0;        // 128..383 among the workers
    const bool eager = p.rounding == B2TTS_ROUND_EAGER;
    const int M = p.m;
    const uint32_t own_mask = (uint32_t)(NX >= 8 ? 7 : 3);
    uint32_t sx = 0, xraw_ph = 0;
    for (int i = 0; i < p.nops; ++i) {
      const ChainOp& op = p.op[i];
      const bool normed = op.norm_w != nullptr;
      const bool w_in_smem = normed && op.k <= C_NORMW_MAX;
      const int S = op.splits, items = op.n_tiles * S;
      if (normed) workers_bar();                // rstd_s / normw_s of this op are ready
      for (int t = cta; t < items; t += G) {
        const int tile = t / S, s = t - tile * S;
        const int kb0 = (s * op.kblocks) / S, kb1 = ((s + 1) * op.kblocks) / S;     // 32-bit: splits <= 16, kblocks < 2^20
        if (!normed) { sx = (sx + (uint32_t)(kb1 - kb0)) & (uint32_t)(NX - 1); continue; }   // NX is a power of two
        for (int kb = kb0; kb < kb1; ++kb, sx = (sx + 1) & (uint32_t)(NX - 1)) {
          if ((int)(sx & own_mask) != wn) continue;
          mbar_wait(xraw + sx, (xraw_ph >> sx) & 1);
          xraw_ph ^= 1u << sx;
          uint8_t* xs = xring + sx * xstage;
          const int c = lane & 7;
          const uint4 wv = w_in_smem ? reinterpret_cast<const uint4*>(normw_s)[kb * 8 + c]
                                     : ldg_v4(op.norm_w + (int64_t)kb * G_BK + c * 8);
#pragma unroll 4
          for (int r = lane >> 3; r < M; r += 4) {
            uint4* cell = reinterpret_cast<uint4*>(xs + r * 128 + ((c ^ (r & 7)) << 4));
            *cell = norm8(*cell, wv, rstd_s[r], eager);
          }
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) mbar_arrive(xfull + sx);
        }
      }
      workers_bar();                            // barrier A(i) passed (the epilogue warps waited on it)
      chain_reduce(p, op, S, cta, G, tr, lane, eager);
      workers_bar();                            // this CTA's share of the reduce is complete
    }
  } else {
    // =========================== epilogue warps: accumulator drain, grid barriers, reduce phase
    const int te = threadIdx.x - C_EPI0;
    const int q = warp & 3;                     // TMEM lane quarter this warp may read
    const bool eager = p.rounding == B2TTS_ROUND_EAGER;
    const int M = p.m;
    uint32_t it = 0;
    C_TRACE(0, 0);
    pdl_wait();
    for (int i = 0; i < p.nops; ++i) {
      const ChainOp& op = p.op[i];
      if (i > 0) {
        if (te == 0) grid_wait(p.sync + 2 * (i - 1) + 1, (unsigned)G);
        epi_bar();
      }
      C_TRACE(i, 1);
      const bool normed = op.norm_w != nullptr;
      const bool w_in_smem = normed && op.k <= C_NORMW_MAX;
      if (normed) {
        // rstd[m]: one thread per row sums the row's partial sums of squares in slot order (deterministic)
        if (te < mpad) {
          float sum = 0.f;
          if (te < M) {
            const float* src = op.ssq_in + (int64_t)te * op.ssq_in_slots;
            if ((op.ssq_in_slots & 3) == 0) {
#pragma unroll 4
              for (int c = 0; c < op.ssq_in_slots; c += 4) {
                const float4 v = __ldcg(reinterpret_cast<const float4*>(src + c));
                sum += v.x; sum += v.y; sum += v.z; sum += v.w;
              }
            } else {
#pragma unroll 4
              for (int c = 0; c < op.ssq_in_slots; ++c) sum += __ldcg(src + c);
            }
          }
          const float var = sum / (float)op.k + p.eps;
          rstd_s[te] = eager ? 1.0f / sqrtf(var) : rsqrtf(var);
        }
        if (w_in_smem)                                    // the norm weight once per op, not once per k-block
          for (int c = te; c < op.k / 8; c += 128)
            reinterpret_cast<uint4*>(normw_s)[c] = ldg_v4(op.norm_w + c * 8);
        workers_bar();                            // releases the norm warps into this op
      }
      C_TRACE(i, 2);
      const int S = op.splits, items = op.n_tiles * S;
      int kbi = 0;
      for (int t = cta; t < items; t += G, ++it) {
        const int tile = t / S, s = t - tile * S;
        const int kb0 = (s * op.kblocks) / S, kb1 = ((s + 1) * op.kblocks) / S;     // 32-bit: splits <= 16, kblocks < 2^20
        // drain the accumulator of this item to the split-K scratch [split][row m][feature n]: thread = feature,
        // so for each batch row a warp stores 32 consecutive floats
        const uint32_t b = it & 1;
        mbar_wait(accfull + b, (it >> 1) & 1);
        tc_fence_after();
        float* dst = p.partial + (int64_t)s * M * op.n + (int64_t)tile * G_BN + q * 32 + lane;
        for (int c0 = 0; c0 < M; c0 += 32) {
          uint32_t v[32];
          tmem_ld16_nowait(tmem_base + ((uint32_t)(q * 32) << 16) + b * (uint32_t)mpad + (uint32_t)c0, v);
          if (c0 + 16 < mpad)
            tmem_ld16_nowait(tmem_base + ((uint32_t)(q * 32) << 16) + b * (uint32_t)mpad + (uint32_t)(c0 + 16), v + 16);
          tmem_ld_wait();
#pragma unroll
          for (int c = 0; c < 32; ++c)
            if (c0 + c < M) __stcg(dst + (int64_t)(c0 + c) * op.n, __uint_as_float(v[c]));
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(accempty + b);
      }
      // ---- grid barrier A(i): every partial of op i is in L2 (release by one thread after the CTA-wide
      // barrier is cumulative over the other threads' stores, as in a cooperative-groups grid sync)
      C_TRACE(i, 4);
      epi_bar();
      if (te == 0) {
        grid_arrive(p.sync + 2 * i);
        grid_wait(p.sync + 2 * i, (unsigned)G);
      }
      epi_bar();
      C_TRACE(i, 5);
      if (i == p.nops - 1) pdl_launch();
      workers_bar();
      chain_reduce(p, op, S, cta, G, te, lane, eager);
      workers_bar();
      // ---- barrier B(i): the rows op i+1 reads are complete (arrive only; the readers wait)
      C_TRACE(i, 6);
      if (i + 1 < p.nops && te == 0) grid_arrive(p.sync + 2 * i + 1);
    }
  }
  pdl_launch();
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, (uint32_t)p.tmem_cols);
  // the last CTA to leave resets the barrier counters: every role of every CTA is past every wait
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned old = atomicAdd(p.sync + 2 * C_MAX_OPS, 1u);
    if (old == (unsigned)G - 1) {
      for (int k = 0; k < 2 * C_MAX_OPS + 1; ++k) p.sync[k] = 0u;
      __threadfence();
    }
  }
}

static size_t chain_smem_bytes(int mpad, int nw, int nx) {
  return (size_t)nw * G_WTILE_BYTES + (size_t)nx * mpad * 128 + (5 * G_MAX_STAGES + 4) * 8 + 16 + (size_t)G_MAX_M * 4 +
         (size_t)C_NORMW_MAX * 2 + 1024;
}

// K split of one chain op: minimise k-blocks on the busiest CTA (+ a term for the reduce cost of many splits)
static int chain_choose_splits(int n_tiles, int kblocks, int ctas) {
  if (const char* e = getenv("B2TTS_CHAIN_SPLITS")) {            // experiments only
    const int v = atoi(e);
    if (v >= 1 && v <= 16 && v <= kblocks) return v;
  }
  int best = 1;
  float best_cost = 1e30f;
  for (int s = 1; s <= 16 && s <= kblocks; ++s) {
    const int rounds = (n_tiles * s + ctas - 1) / ctas;
    const float cost = (float)(rounds * ((kblocks + s - 1) / s)) + 0.25f * (float)s;
    if (cost < best_cost - 1e-6f) { best_cost = cost; best = s; }
  }
  return best;
}
}  // namespace b2

extern "C" B2TTS_API int b2tts_gemm_chain_ssq_slots(int n) { return n > 0 && n % b2::C_SSQ_FEATS == 0 ? n / b2::C_SSQ_FEATS : 0; }

extern "C" B2TTS_API size_t b2tts_gemm_chain_workspace_bytes(int m, const int* n, const int* k, int nops) {
  using namespace b2;
  if (m < 1 || m > G_MAX_M || nops < 1 || nops > C_MAX_OPS || !n || !k) return 0;
  const int mpad = (m + 15) & ~15;
  size_t part = 0;
  for (int i = 0; i < nops; ++i) {
    if (n[i] <= 0 || k[i] <= 0 || n[i] % G_BN || k[i] % G_BK) return 0;
    const int s = 16 < k[i] / G_BK ? 16 : k[i] / G_BK;           // upper bound of any split choice
    const size_t b = (size_t)s * n[i] * mpad * 4;
    if (b > part) part = b;
  }
  return C_SYNC_WORDS * 4 + 192 + part;
}

extern "C" B2TTS_API int b2tts_gemm_chain(const b2tts_gemm_op* ops, int nops, int m, float eps, int rounding,
                                          void* workspace, size_t workspace_bytes, int pdl, void* stream) {
  using namespace b2;
  const char* what = "gemm_chain";
  B2_REQUIRE(ops && nops >= 1 && nops <= C_MAX_OPS, B2TTS_EINVAL, "%s: 1..%d ops", what, C_MAX_OPS);
  B2_REQUIRE(m >= 0 && m <= G_MAX_M, B2TTS_EUNSUPPORTED, "%s: m=%d (0..%d)", what, m, G_MAX_M);
  if (m == 0) return B2TTS_OK;
  B2_REQUIRE(workspace && aligned16(workspace), B2TTS_EINVAL, "%s: workspace", what);
  ChainParams p{};
  p.nops = nops; p.m = m; p.mpad = (m + 15) & ~15; p.rounding = rounding; p.eps = eps;
  p.tmem_cols = 32;
  while (p.tmem_cols < 2 * p.mpad) p.tmem_cols <<= 1;
  p.nx = p.nw = p.mpad <= 64 ? 8 : 4;           // 192 KiB / 128 KiB of ring; both rings share slot indices
  if (const char* e = getenv("B2TTS_CHAIN_NW")) {                // experiments only
    const int v = atoi(e);
    if ((v == 4 || v == 8) && chain_smem_bytes(p.mpad, v, v) <= 227 * 1024) p.nw = p.nx = v;
  }
  p.cg = p.nw >= 8 ? 4 : 2;
  if (const char* e = getenv("B2TTS_CHAIN_CG")) {                // experiments only
    const int v = atoi(e);
    if (v >= 1 && v <= p.nw / 2 && p.nw % v == 0) p.cg = v;
  }
  const int ctas = sm_count();
  CUtensorMap tmaps[C_MAX_OPS];
  size_t part = 0;
  for (int i = 0; i < nops; ++i) {
    const b2tts_gemm_op& o = ops[i];
    B2_REQUIRE(o.x && o.w_packed && o.out, B2TTS_EINVAL, "%s: op %d null pointer", what, i);
    B2_REQUIRE(o.n > 0 && o.k > 0 && o.n % G_BN == 0 && o.k % G_BK == 0, B2TTS_EUNSUPPORTED,
               "%s: op %d n=%d (multiple of %d), k=%d (multiple of %d)", what, i, o.n, G_BN, o.k, G_BK);
    B2_REQUIRE(o.epilogue == B2TTS_GEMM_BIAS || o.epilogue == B2TTS_GEMM_RESIDUAL || o.epilogue == B2TTS_GEMM_SILU_MUL,
               B2TTS_EINVAL, "%s: op %d epilogue %d", what, i, o.epilogue);
    B2_REQUIRE(o.epilogue != B2TTS_GEMM_RESIDUAL || o.residual, B2TTS_EINVAL, "%s: op %d residual epilogue without residual", what, i);
    B2_REQUIRE(aligned16(o.x) && aligned16(o.w_packed) && aligned16(o.out) && o.x_row_stride % 8 == 0 &&
                   o.out_row_stride % 4 == 0 && (o.epilogue == B2TTS_GEMM_SILU_MUL || o.out_row_stride % 8 == 0) &&
                   (!o.residual || (aligned16(o.residual) && o.residual_row_stride % 8 == 0)),
               B2TTS_EALIGN, "%s: op %d pointers / strides must be 16-byte aligned", what, i);
    B2_REQUIRE(!o.norm_weight || (o.ssq_in && o.ssq_in_slots > 0 && aligned16(o.norm_weight) && aligned16(o.ssq_in)),
               B2TTS_EINVAL, "%s: op %d fused norm needs ssq_in", what, i);
    B2_REQUIRE(!o.ssq_out || o.epilogue != B2TTS_GEMM_SILU_MUL, B2TTS_EINVAL, "%s: op %d ssq_out with SiLU*mul", what, i);
    ChainOp& c = p.op[i];
    c.wp = static_cast<const uint8_t*>(o.w_packed);
    c.norm_w = static_cast<const __nv_bfloat16*>(o.norm_weight);
    c.ssq_in = o.ssq_in;
    c.res = static_cast<const __nv_bfloat16*>(o.residual);
    c.out = static_cast<__nv_bfloat16*>(o.out);
    c.ssq_out = o.ssq_out;
    c.res_stride = o.residual_row_stride; c.out_stride = o.out_row_stride;
    c.n = o.n; c.k = o.k; c.kblocks = o.k / G_BK; c.n_tiles = o.n / G_BN;
    c.splits = chain_choose_splits(c.n_tiles, c.kblocks, ctas);
    c.epilogue = o.epilogue;
    c.ssq_in_slots = o.ssq_in_slots;
    c.ssq_lanes = 1;
    if (o.norm_weight) {
      int g = 128 / p.mpad;
      if (g > 8) g = 8;
      if (g < 1) g = 1;
      while (g > 1 && (o.ssq_in_slots % g)) g >>= 1;
      while (g > 1 && ((o.ssq_in_slots / g) % 4)) g >>= 1;
      c.ssq_lanes = g;
      B2_REQUIRE(o.ssq_in_slots % 4 == 0 || g == 1, B2TTS_EALIGN, "%s: op %d ssq_in_slots %d", what, i, o.ssq_in_slots);
    }
    const size_t b = (size_t)c.splits * c.n * m * 4;
    if (b > part) part = b;
    int rc = make_x_tensor_map(&tmaps[i], o.x, o.x_row_stride, m, o.k, p.mpad, what);
    if (rc) return rc;
  }
  for (int i = nops; i < C_MAX_OPS; ++i) tmaps[i] = tmaps[0];
  B2_REQUIRE(workspace_bytes >= C_SYNC_WORDS * 4 + 192 + part, B2TTS_EINVAL, "%s: workspace %zu < %zu bytes", what,
             workspace_bytes, (size_t)(C_SYNC_WORDS * 4 + 192 + part));
  p.trace = nullptr;
  if (g_trace_base && g_trace_next < g_trace_slots) p.trace = g_trace_base + (g_trace_next++) * 4096 * 8;
  p.sync = static_cast<unsigned*>(workspace);
  p.partial = reinterpret_cast<float*>(static_cast<uint8_t*>(workspace) + 256);
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(gemm_chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return cuda_fail(e, "gemm_chain(smem attr)");
    configured = true;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(ctas, 1, 1);
  cfg.blockDim = dim3(C_THREADS, 1, 1);
  cfg.dynamicSmemBytes = chain_smem_bytes(p.mpad, p.nw, p.nx);
  cfg.stream = static_cast<cudaStream_t>(stream);
  cudaLaunchAttribute attrs[1];
  int na = 0;
  if (pdl) {
    attrs[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attrs[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attrs;
  cfg.numAttrs = na;
  cudaError_t e = cudaLaunchKernelEx(&cfg, gemm_chain_kernel, p, tmaps[0], tmaps[1], tmaps[2], tmaps[3]);
  count_launch();
  if (e != cudaSuccess) return cuda_fail(e, what);
  return B2TTS_OK;
}

// Debug aid (tools/gemm_trace.py): every following GEMM launch records 8 globaltimer stamps per CTA into
// buffer + launch_index * 4096 * 8 (uint64), for `launches` launches.  NULL switches tracing off.
extern "C" B2TTS_API int b2tts_debug_gemm_trace(void* buffer, int launches) {
  b2::g_trace_base = static_cast<unsigned long long*>(buffer);
  b2::g_trace_slots = buffer ? (size_t)launches : 0;
  b2::g_trace_next = 0;
  return B2TTS_OK;
}

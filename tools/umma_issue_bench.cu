// umma_issue_bench.cu - what ONE issuing thread pays per k-block of tcgen05.mma at the decode GEMM's tile shape
// (A = 128 x 64 bf16 weight tile, B = N x 64 activations, SWIZZLE_128B, accumulator in TMEM), with every operand
// already resident in shared memory.  Separates: the MMA issue itself, tcgen05.commit per k-block, the mbarrier wait on
// an already complete barrier, the commit -> mbarrier -> waiting-thread latency, and a ring fill by bulk copies.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_bin/umma_issue_bench tools/umma_issue_bench.cu
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../nano-qwen3tts-vllm_b200/csrc/umma.cuh"

using namespace b2;

namespace b2 {                                  // host plumbing common.cuh declares; unused here
void set_error(const char*, ...) {}
int cuda_fail(cudaError_t, const char*) { return -1; }
void count_launch() {}
int sm_count() { return 148; }
}

constexpr int NS = 12;                          // resident k-blocks (16 KiB A + up to 32 KiB B each would not fit: B shared)

struct Result { long long t[8]; };

// mode bits: 1 = commit per k-block, 2 = wait a (complete) full barrier per k-block, 4 = whole warp walks + elect_one,
// 8 = the 4 K-steps of a k-block alternate between 2 accumulators, 24 = each K-step has its own accumulator (4)
template <int N>
__global__ void __launch_bounds__(128, 1) issue_kernel(Result* out, int nkb, int mode, const uint8_t* gsrc) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* a_tiles = smem;                                  // NS x 16 KiB
  uint8_t* b_tile = smem + NS * G_WTILE_BYTES;              // one N x 64 tile, reused (contents do not matter)
  uint64_t* bars = reinterpret_cast<uint64_t*>(b_tile + 256 * 128);
  uint64_t* full_bar = bars;                                // [64]
  uint64_t* empty_bar = bars + 64;                          // [64]
  uint64_t* accum_bar = bars + 128;
  uint64_t* fill_bar = bars + 129;                          // [NS]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 160);
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  for (int i = threadIdx.x; i < (NS * G_WTILE_BYTES + 256 * 128) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 64; ++i) { mbar_init(full_bar + i, 1); mbar_init(empty_bar + i, 1); }
    mbar_init(accum_bar, 1);
    for (int i = 0; i < NS; ++i) mbar_init(fill_bar + i, 1);
    mbar_fence_init();
    for (int i = 0; i < 64; ++i) mbar_arrive(full_bar + i);      // "already landed"
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  long long t0 = 0, t1 = 0, t2 = 0, t3 = 0;
  if (warp == 1) {
    // production form (gemm.cu): the whole warp walks the ring, elect_one() right around the MMAs => uniform datapath
    const int nacc = ((mode >> 3) & 3) + 1;     // 1, 2 or 4 accumulators (K-steps alternate)
    t0 = clock64();
    for (int i = 0; i < nkb; ++i) {
      if (mode & 2) mbar_wait(full_bar + (i & 63), 0);
      tc_fence_after();
      const uint32_t a_addr = smem_u32(a_tiles + (i % NS) * G_WTILE_BYTES);
      const uint32_t b_addr = smem_u32(b_tile);
      if (mode & 32) {
        // lean form: the 64-bit descriptors differ from the stage's base only in the low word, by 2 per K step
        const uint64_t a0 = umma_desc_sw128(a_addr), b0 = umma_desc_sw128(b_addr);
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16(tmem_base, a0 + 2 * k, b0 + 2 * k, idesc, (uint32_t)((i | k) != 0));
          if (mode & 1) umma_commit(empty_bar + (i & 63));
          if (i == nkb - 1) umma_commit(accum_bar);
        }
      } else if (elect_one()) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_bf16(tmem_base + (uint32_t)((k & (nacc - 1)) * N), umma_desc_sw128(a_addr + k * 32),
                    umma_desc_sw128(b_addr + k * 32), idesc, (uint32_t)(nacc > 1 ? (i != 0 || k >= nacc) : ((i | k) != 0)));
        if (mode & 1) umma_commit(empty_bar + (i & 63));
        if (i == nkb - 1) umma_commit(accum_bar);
      }
      __syncwarp();
    }
    t1 = clock64();                            // everything issued
    mbar_wait(accum_bar, 0);
    t2 = clock64();                            // accumulator complete seen by the issuing warp
    __syncwarp();
  }
  if (mode & 64) __syncthreads();
  if ((mode & 64) && (warp == 1 || warp == 2)) {
    const int w = warp - 1;
    if (threadIdx.x == 32) { /* accum_bar was consumed by the single-warp pass above: use fill_bar[NS-1]/[NS-2] */ }
    const long long u0 = clock64();
    int last = -1;
    for (int i = w; i < nkb; i += 2) last = i;
    for (int i = w; i < nkb; i += 2) {
      mbar_wait(full_bar + (i & 63), 0);
      tc_fence_after();
      const uint64_t a0 = umma_desc_sw128(smem_u32(a_tiles + (i % NS) * G_WTILE_BYTES)), b0 = umma_desc_sw128(smem_u32(b_tile));
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_bf16(tmem_base + 256u + (uint32_t)(w * N), a0 + 2 * k, b0 + 2 * k, idesc, (uint32_t)((i >= 2) || k));
        if (i == last) umma_commit(empty_bar + 60 + w);
      }
      __syncwarp();
    }
    if (last >= 0) mbar_wait(empty_bar + 60 + w, 0);
    const long long u1 = clock64();
    if ((threadIdx.x & 31) == 0) out[blockIdx.x].t[7] = u1 - u0;      // either warp: they finish together
  }
  if (warp == 0) {                               // another warp waiting on the same barrier
    mbar_wait(accum_bar, 0);
    t3 = clock64();
  }
  __syncthreads();
  // ring fill: NS bulk copies of 16 KiB from global (L2-resident after the first launch), issued by one thread
  long long f0 = 0, f1 = 0, f2 = 0;
  if (threadIdx.x == 0 && gsrc) {
    f0 = clock64();
    for (int i = 0; i < NS; ++i) {
      mbar_expect_tx(fill_bar + i, G_WTILE_BYTES);
      bulk_g2s(a_tiles + i * G_WTILE_BYTES, gsrc + ((size_t)blockIdx.x * NS + i) * G_WTILE_BYTES, G_WTILE_BYTES, fill_bar + i);
    }
    f1 = clock64();
    mbar_wait(fill_bar + 0, 0);
    f2 = clock64();
    for (int i = 1; i < NS; ++i) mbar_wait(fill_bar + i, 0);
    long long f3 = clock64();
    out[blockIdx.x].t[6] = f3 - f0;
  }
  __syncthreads();
  if (threadIdx.x == 32) {
    out[blockIdx.x].t[0] = t1 - t0;
    out[blockIdx.x].t[1] = t2 - t0;
  }
  if (threadIdx.x == 0) {
    out[blockIdx.x].t[2] = t3;
    out[blockIdx.x].t[4] = f1 - f0;
    out[blockIdx.x].t[5] = f2 - f0;
  }
  if (threadIdx.x == 32) out[blockIdx.x].t[3] = t2;
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

template <int N>
static void run(int grid, const uint8_t* gsrc, Result* dres) {
  const int smem = NS * G_WTILE_BYTES + 256 * 128 + 2048;
  cudaFuncSetAttribute(issue_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  std::vector<Result> h(grid);
  const int modes[] = {2, 3, 34, 35, 98};
  for (int mode : modes) {
    for (int nkb : {1, 4, 8, 16, 32}) {
      long long issue = 0, done = 0, other = 0, fi = 0, ff = 0, fa = 0;
      for (int rep = 0; rep < 3; ++rep) {
        issue_kernel<N><<<grid, 128, smem>>>(dres, nkb, mode, gsrc);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("launch failed: %s\n", cudaGetErrorString(e)); exit(1); }
      }
      cudaMemcpy(h.data(), dres, grid * sizeof(Result), cudaMemcpyDeviceToHost);
      for (int b = 0; b < grid; ++b) {
        issue += h[b].t[0]; done += h[b].t[1]; other += h[b].t[2] - h[b].t[3];
        fi += h[b].t[4]; ff += h[b].t[5]; fa += h[b].t[6];
      }
      if (mode & 64) printf("   two issuing warps, alternate k-blocks, two accumulators: %lld cyc (%.0f per k-block)\n", h[0].t[7], (double)h[0].t[7] / nkb);
      printf("N=%3d grid=%3d mode=%d nkb=%2d: issued after %6lld cyc, accumulator seen after %6lld cyc (%.0f per k-block), "
             "other warp %+lld cyc | fill %d x 16 KiB: issue %lld, first landed %lld, all %lld cyc\n",
             N, grid, mode, nkb, issue / grid, done / grid, (double)done / grid / nkb, other / grid, NS, fi / grid, ff / grid, fa / grid);
    }
  }
}


// second experiment: is ~125 cycles per tcgen05.mma a property of the instruction or of the issuing thread's code?
// W issuing warps (lane 0 each), constant descriptors (nothing but the MMA in the loop), accumulator w; M = 128 or 64.
template <int N, int M>
__global__ void __launch_bounds__(256, 1) issue2_kernel(Result* out, int nmma, int nwarps) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 64 * 1024);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 64 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 8; ++i) mbar_init(bars + i, 1);
    mbar_fence_init();
  }
  if (warp == 0) tmem_alloc(tmem_slot, 512);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
  if (warp < nwarps && (threadIdx.x & 31) == 0) {
    const uint64_t ad = umma_desc_sw128(smem_u32(smem + warp * 1024 * 0));
    const uint64_t bd = umma_desc_sw128(smem_u32(smem + 32 * 1024));
    const uint32_t d = tmem_base + (uint32_t)(warp * N);
    const long long t0 = clock64();
    umma_bf16(d, ad, bd, idesc, 0u);
    for (int i = 1; i < nmma; ++i) umma_bf16(d, ad, bd, idesc, 1u);
    const long long t1 = clock64();
    umma_commit(bars + warp);
    mbar_wait(bars + warp, 0);
    const long long t2 = clock64();
    out[blockIdx.x].t[warp & 3] = t1 - t0;
    out[blockIdx.x].t[4 + (warp & 3)] = t2 - t0;
  }
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, 512);
}

template <int N, int M>
static void run2(Result* dres) {
  const int smem = 64 * 1024 + 1024;
  cudaFuncSetAttribute(issue2_kernel<N, M>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  Result h;
  for (int nwarps : {1, 2, 4})
    for (int nmma : {4, 16, 64}) {
      for (int rep = 0; rep < 3; ++rep) {
        issue2_kernel<N, M><<<1, 256, smem>>>(dres, nmma, nwarps);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("launch failed: %s\n", cudaGetErrorString(e)); exit(1); }
      }
      cudaMemcpy(&h, dres, sizeof(Result), cudaMemcpyDeviceToHost);
      printf("M=%3d N=%3d warps=%d mma/warp=%2d: issue %lld %lld %lld %lld cyc, complete %lld %lld %lld %lld cyc (%.0f cyc per MMA per warp)\n",
             M, N, nwarps, nmma, h.t[0], nwarps > 1 ? h.t[1] : 0, nwarps > 2 ? h.t[2] : 0, nwarps > 3 ? h.t[3] : 0,
             h.t[4], nwarps > 1 ? h.t[5] : 0, nwarps > 2 ? h.t[6] : 0, nwarps > 3 ? h.t[7] : 0, (double)h.t[4] / nmma);
    }
}

int main() {
  uint8_t* gsrc;
  Result* dres;
  cudaMalloc(&gsrc, (size_t)296 * NS * G_WTILE_BYTES);
  cudaMemset(gsrc, 0, (size_t)296 * NS * G_WTILE_BYTES);
  cudaMalloc(&dres, 296 * sizeof(Result));
  run<64>(1, gsrc, dres);
  run<16>(1, gsrc, dres);
  if (getenv("UMMA_RUN2")) {
    run2<64, 128>(dres);
    run2<64, 64>(dres);
    run2<128, 128>(dres);
  }
  return 0;
}

// attention_fused.cu - the decode step of Qwen3TTSAttention.forward (models/qwen3_tts_share.py:96-109)
// as ONE kernel: per-head q/k RMSNorm + RoPE (layers/rotary_embedding.py:11-212), the paged KV write of
// the new token (store_kvcache, layers/attention.py:10-54), the block_table paged GQA attention
// (flash_attn_with_kvcache, attention.py:92-95) and the split-KV combine.  The separate
// qknorm_rope_kvwrite / paged_attn / attn_combine launches of the decode path collapse into this one.
//
// Same HBM design as attention.cu: one CTA per (sequence, KV split) covers all kv heads; a producer
// warp streams 8-token K and V tiles (all heads of a token are contiguous in the page) with 1-D bulk
// async copies into a 3-stage mbarrier ring; consumer warp w owns kv head w and its GROUP q heads.
// Fusion:
//   * prologue: the consumer warps run the arithmetic of qknorm_rope_kvwrite_kernel (norm_rope.cu; same
//     lane mapping, rounding points and summation order, so K cache bytes are identical to the unfused
//     path) on the one new token: q heads go to shared memory; the split that owns the sequence's last
//     tile also normalises + rotates k, writes it and the raw v to cache[slot] and keeps a copy in
//     shared memory: when the last tile lands, each warp overwrites its head's slice of the new token's
//     row with that copy, so it never depends on reading its own global write;
//   * combine: split partials go to the workspace; the LAST CTA of a sequence to finish (atomic
//     ticket) folds them in split order - deterministic - and resets the ticket for the next launch.
#include <stdlib.h>

#include "common.cuh"

namespace b2 {

constexpr int FA_HD = 128;
constexpr int FA_TILE = 8;
constexpr int FA_NSTAGE = 3;
// The per-row split-KV tickets live in a FIXED span at the start of the workspace, whatever the batch size of the
// launch: one workspace serves every batch size / graph of a runner, and the partial accumulators of one launch
// must never land on the (always-zero-between-launches) tickets another launch will read.
constexpr int FA_MAX_ROWS = 4096;
constexpr size_t FA_TICKET_BYTES = (size_t)FA_MAX_ROWS * sizeof(int);

struct FusedAttnParams {
  const __nv_bfloat16 *q, *k, *v;          // rows of the QKV projection
  int64_t q_stride, k_stride, v_stride;
  const __nv_bfloat16 *q_w, *k_w;          // head-norm weights [128]
  float eps;
  const int64_t* positions;                // [rows] or [3][rows]
  int64_t pos_row_stride;
  const float* inv_freq;                   // [64]
  const int32_t* modality;                 // [128] or null
  float attention_scaling;
  const int32_t* slot_mapping;             // [rows]
  __nv_bfloat16 *k_cache, *v_cache;        // [pages, page, kvh, 128]
  const int32_t* block_table;
  int bt_stride, page_size;
  const int32_t* context_lens;
  __nv_bfloat16* out;                      // [rows, heads*128]
  float* part_acc;                         // [rows, nsplit, heads, 128]
  float* part_ml;                          // [rows, nsplit, heads, 2]
  int* tickets;                            // [rows], zero between launches
  int nsplit, num_heads, num_kv_heads;
  float scale_log2;
};

__device__ __forceinline__ float fa_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void fa_unpack8(const uint4& r, float* f) {
  f[0] = bf16lo(r.x); f[1] = bf16hi(r.x); f[2] = bf16lo(r.y); f[3] = bf16hi(r.y);
  f[4] = bf16lo(r.z); f[5] = bf16hi(r.z); f[6] = bf16lo(r.w); f[7] = bf16hi(r.w);
}
__device__ __forceinline__ uint4 fa_pack8(const float* f) {
  uint4 r;
  r.x = pack_bf16(f[0], f[1]); r.y = pack_bf16(f[2], f[3]); r.z = pack_bf16(f[4], f[5]); r.w = pack_bf16(f[6], f[7]);
  return r;
}

template <int GROUP, bool EAGER>
__global__ void __launch_bounds__(288, 2)
fused_decode_attn_kernel(const FusedAttnParams p) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  const int kvh = p.num_kv_heads;
  const int kv_dim = kvh * FA_HD;
  const int stage_elems = FA_TILE * kv_dim;
  __nv_bfloat16* Ks = reinterpret_cast<__nv_bfloat16*>(smem_raw);
  __nv_bfloat16* Vs = Ks + FA_NSTAGE * stage_elems;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(Vs + FA_NSTAGE * stage_elems);
  uint64_t* empty_bar = full_bar + FA_NSTAGE;
  int* last_flag = reinterpret_cast<int*>(empty_bar + FA_NSTAGE);
  __nv_bfloat16* q_s = reinterpret_cast<__nv_bfloat16*>(last_flag + 4);     // [heads][128] normed + rotated q
  __nv_bfloat16* kn_s = q_s + p.num_heads * FA_HD;                          // [kvh][128] the new token's K
  __nv_bfloat16* vn_s = kn_s + kv_dim;                                      // [kvh][128] the new token's V

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x;
  const int split = blockIdx.y;

  pdl_launch();
  if (threadIdx.x == 0) {
    for (int s = 0; s < FA_NSTAGE; ++s) {
      mbar_init(full_bar + s, 1);
      mbar_init(empty_bar + s, kvh);
    }
    mbar_fence_init();
  }
  // Latency discipline: this kernel runs 103 times per codec frame on contexts of 3 .. a few hundred tokens, so
  // its cost is the chain of dependent global-memory round trips, not bytes.  Everything that does not depend on
  // the previous kernel (rotary frequencies, M-RoPE row selectors, head-norm weights) is requested BEFORE
  // griddepcontrol.wait; everything that does (length, slot, positions, the q/k/v rows, the block table) is
  // requested in ONE batch right after it, so the prologue costs one round trip instead of six.
  const int g = lane >> 3;        // token slot within a 4-token step
  const int j = lane & 7;         // which 16 head-dims: [8j,8j+8) and [64+8j,64+8j+8)
  const bool consumer = warp < kvh;
  float frq[2] = {0.f, 0.f};
  int md1[2] = {0, 0}, md2[2] = {0, 0};
  uint32_t qw1 = 0, qw2 = 0, kw1 = 0, kw2 = 0;
  if (consumer) {
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int fi = 2 * lane + e;
      frq[e] = __ldg(p.inv_freq + fi);
      if (p.modality) { md1[e] = __ldg(p.modality + fi); md2[e] = __ldg(p.modality + FA_HD / 2 + fi); }
    }
    if (p.q_w) { qw1 = *reinterpret_cast<const uint32_t*>(p.q_w + 2 * lane); qw2 = *reinterpret_cast<const uint32_t*>(p.q_w + 64 + 2 * lane); }
    if (p.k_w) { kw1 = *reinterpret_cast<const uint32_t*>(p.k_w + 2 * lane); kw2 = *reinterpret_cast<const uint32_t*>(p.k_w + 64 + 2 * lane); }
  }
  __syncthreads();
  pdl_wait();                                   // q/k/v rows, slots and lengths come from earlier kernels

  const bool writer = (split == p.nsplit - 1);  // owns the last tile, i.e. the token appended this step
  int L = __ldg(p.context_lens + row);
  // ---- one batch of independent requests (consumers: this token's rows; producer: the block-table row)
  int tbl = 0;                                  // producer: lane l holds block_table[row][l]
  int slot = -1;
  int64_t pos[3] = {0, 0, 0};
  uint32_t raw1[GROUP + 1], raw2[GROUP + 1];
  uint4 vraw = make_uint4(0, 0, 0, 0);
  const int nh_q = p.num_heads / kvh;           // q heads this warp normalises: warp, warp + kvh, ...  (== GROUP)
  if (consumer) {
    if (writer) slot = __ldg(p.slot_mapping + row);
    pos[0] = __ldg(p.positions + row);
    if (p.modality && p.pos_row_stride) {
      pos[1] = __ldg(p.positions + p.pos_row_stride + row);
      pos[2] = __ldg(p.positions + 2 * p.pos_row_stride + row);
    } else {
      pos[1] = pos[2] = pos[0];
    }
#pragma unroll
    for (int i = 0; i < GROUP + 1; ++i) {
      const bool is_k = i == GROUP;
      raw1[i] = raw2[i] = 0u;
      if (i < nh_q || (is_k && writer)) {
        const __nv_bfloat16* src = is_k ? p.k + (int64_t)row * p.k_stride + warp * FA_HD
                                        : p.q + (int64_t)row * p.q_stride + (warp + i * kvh) * FA_HD;
        raw1[i] = *reinterpret_cast<const uint32_t*>(src + 2 * lane);
        raw2[i] = *reinterpret_cast<const uint32_t*>(src + 64 + 2 * lane);
      }
    }
    if (writer && (int)threadIdx.x < (kv_dim >> 3)) vraw = ldg_v4(p.v + (int64_t)row * p.v_stride + 8 * threadIdx.x);
  } else if (lane < p.bt_stride) {
    tbl = __ldg(p.block_table + (int64_t)row * p.bt_stride + lane);
  }
  if (L < 0) L = 0;
  const int n_tiles = (L + FA_TILE - 1) / FA_TILE;
  const int t_begin = (int)(((int64_t)split * n_tiles) / p.nsplit);
  const int t_end = (int)(((int64_t)(split + 1) * n_tiles) / p.nsplit);
  const int my_tiles = t_end - t_begin;

  if (warp == kvh) {
    // =========================== producer warp: the whole warp walks the ring (the page of a tile comes from the
    // lanes' copy of the block-table row), lane 0 issues the copies
    for (int i = 0; i < my_tiles; ++i) {
      const int stage = i % FA_NSTAGE;
      const uint32_t ph = (i / FA_NSTAGE) & 1;
      const int tok0 = (t_begin + i) * FA_TILE;
      const int pidx = tok0 / p.page_size;
      int page = __shfl_sync(0xffffffffu, tbl, pidx & 31);
      if (pidx >= 32) page = __ldg(p.block_table + (int64_t)row * p.bt_stride + pidx);
      if (lane == 0) {
        mbar_wait(empty_bar + stage, ph ^ 1);
        const int ntok = min(FA_TILE, L - tok0);
        if (page < 0) page = 0;                  // never faults on a stale (-1) table entry
        const int64_t slot0 = (int64_t)page * p.page_size + tok0 % p.page_size;
        const uint32_t bytes = (uint32_t)(ntok * kv_dim * 2);
        mbar_expect_tx(full_bar + stage, 2u * bytes);
        bulk_g2s(Ks + stage * stage_elems, p.k_cache + slot0 * kv_dim, bytes, full_bar + stage);
        bulk_g2s(Vs + stage * stage_elems, p.v_cache + slot0 * kv_dim, bytes, full_bar + stage);
      }
      __syncwarp();
    }
    return;
  }

  // =========================== consumer warps
  const int head0 = warp * GROUP;

  if (L == 0) {                   // CUDA-graph padding row (talker_mode_runner.py:97-100): zeros, no ticket
    if (split == 0 && g == 0) {
#pragma unroll
      for (int h = 0; h < GROUP; ++h) {
        __nv_bfloat16* o = p.out + ((int64_t)row * p.num_heads + head0 + h) * FA_HD;
        stg_v4(o + 8 * j, make_uint4(0, 0, 0, 0));
        stg_v4(o + 64 + 8 * j, make_uint4(0, 0, 0, 0));
      }
    }
    return;
  }

  // ---- prologue: the body of qknorm_rope_kvwrite_kernel (norm_rope.cu) for this one token, same lane
  // mapping and arithmetic (lane l holds dims {2l, 2l+1, 64+2l, 64+2l+1}; a warp walks heads warp,
  // warp+kvh, ...): q heads -> shared memory, and on the writer split k (normed + rotated) and v (raw)
  // -> cache[slot] and shared memory.
  {
    float c1[2], s1[2], c2[2], s2[2];
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const float a1 = (float)(md1[e] == 0 ? pos[0] : md1[e] == 1 ? pos[1] : pos[2]) * frq[e];
      c1[e] = rbf(cosf(a1) * p.attention_scaling);
      s1[e] = rbf(sinf(a1) * p.attention_scaling);
      if (md2[e] != md1[e]) {
        const float a2 = (float)(md2[e] == 0 ? pos[0] : md2[e] == 1 ? pos[1] : pos[2]) * frq[e];
        c2[e] = rbf(cosf(a2) * p.attention_scaling);
        s2[e] = rbf(sinf(a2) * p.attention_scaling);
      } else {
        c2[e] = c1[e]; s2[e] = s1[e];
      }
    }
#pragma unroll
    for (int i = 0; i < GROUP + 1; ++i) {
      const bool is_k = i == GROUP;
      if (!(i < nh_q || (is_k && writer))) continue;
      const int h = is_k ? warp : warp + i * kvh;
      const bool has_w = is_k ? (p.k_w != nullptr) : (p.q_w != nullptr);
      float x1[2] = {bf16lo(raw1[i]), bf16hi(raw1[i])};
      float x2[2] = {bf16lo(raw2[i]), bf16hi(raw2[i])};
      if (has_w) {
        float ss = x1[0] * x1[0];
        ss = fmaf(x1[1], x1[1], ss); ss = fmaf(x2[0], x2[0], ss); ss = fmaf(x2[1], x2[1], ss);
        const float var = warp_sum(ss) / (float)FA_HD + p.eps;
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
        o1[e] = rbf(x1[e] * c1[e]) + rbf(-x2[e] * s1[e]);   // rotate_half: first half pairs with -x2
        o2[e] = rbf(x2[e] * c2[e]) + rbf(x1[e] * s2[e]);
      }
      const uint32_t y1 = pack_bf16(o1[0], o1[1]), y2 = pack_bf16(o2[0], o2[1]);
      __nv_bfloat16* dst = is_k ? kn_s + h * FA_HD : q_s + h * FA_HD;
      *reinterpret_cast<uint32_t*>(dst + 2 * lane) = y1;
      *reinterpret_cast<uint32_t*>(dst + 64 + 2 * lane) = y2;
      if (is_k && slot >= 0) {
        __nv_bfloat16* cd = p.k_cache + (int64_t)slot * kv_dim + h * FA_HD;
        *reinterpret_cast<uint32_t*>(cd + 2 * lane) = y1;
        *reinterpret_cast<uint32_t*>(cd + 64 + 2 * lane) = y2;
      }
    }
    if (writer && (int)threadIdx.x < (kv_dim >> 3)) {      // V goes to the cache untouched
      *reinterpret_cast<uint4*>(vn_s + 8 * threadIdx.x) = vraw;
      if (slot >= 0) stg_v4(p.v_cache + (int64_t)slot * kv_dim + 8 * threadIdx.x, vraw);
    }
  }
  asm volatile("bar.sync 1, %0;" :: "r"(kvh * 32) : "memory");      // the consumer warps only

  // ---- this warp's q heads from shared memory, pre-scaled for exp2
  float qv[GROUP][16];
#pragma unroll
  for (int h = 0; h < GROUP; ++h) {
    const __nv_bfloat16* qp = q_s + (head0 + h) * FA_HD;
    fa_unpack8(*reinterpret_cast<const uint4*>(qp + 8 * j), qv[h]);
    fa_unpack8(*reinterpret_cast<const uint4*>(qp + 64 + 8 * j), qv[h] + 8);
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
    const int stage = i % FA_NSTAGE;
    const uint32_t ph = (i / FA_NSTAGE) & 1;
    const int tok0 = (t_begin + i) * FA_TILE;
    const int ntok = min(FA_TILE, L - tok0);
    mbar_wait(full_bar + stage, ph);
    __nv_bfloat16* kt = Ks + stage * stage_elems + warp * FA_HD;
    __nv_bfloat16* vt = Vs + stage * stage_elems + warp * FA_HD;
    if (writer && i == my_tiles - 1) {
      // the bulk copy may have read the new token's row before or after the global store above:
      // replace this head's slice of that row with the register copy
      if (g == 0) {
        __nv_bfloat16* kr = kt + (L - 1 - tok0) * kv_dim;
        __nv_bfloat16* vr = vt + (L - 1 - tok0) * kv_dim;
        const __nv_bfloat16* kn = kn_s + warp * FA_HD;
        const __nv_bfloat16* vn = vn_s + warp * FA_HD;
        *reinterpret_cast<uint4*>(kr + 8 * j) = *reinterpret_cast<const uint4*>(kn + 8 * j);
        *reinterpret_cast<uint4*>(kr + 64 + 8 * j) = *reinterpret_cast<const uint4*>(kn + 64 + 8 * j);
        *reinterpret_cast<uint4*>(vr + 8 * j) = *reinterpret_cast<const uint4*>(vn + 8 * j);
        *reinterpret_cast<uint4*>(vr + 64 + 8 * j) = *reinterpret_cast<const uint4*>(vn + 64 + 8 * j);
      }
      __syncwarp();
    }

    float s[FA_TILE / 4][GROUP];
    int tokc[FA_TILE / 4];
#pragma unroll
    for (int it = 0; it < FA_TILE / 4; ++it) {
      const int tok = it * 4 + g;
      const bool valid = tok < ntok;
      tokc[it] = valid ? tok : 0;
      const __nv_bfloat16* kr = kt + tokc[it] * kv_dim;
      float kf[16];
      fa_unpack8(*reinterpret_cast<const uint4*>(kr + 8 * j), kf);
      fa_unpack8(*reinterpret_cast<const uint4*>(kr + 64 + 8 * j), kf + 8);
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
      for (int it = 1; it < FA_TILE / 4; ++it) mt = fmaxf(mt, s[it][h]);
      mt = fmaxf(mt, __shfl_xor_sync(0xffffffffu, mt, 8));
      mt = fmaxf(mt, __shfl_xor_sync(0xffffffffu, mt, 16));
      const float mn = fmaxf(m[h], mt);
      alpha[h] = fa_exp2(m[h] - mn);
      m[h] = mn;
      float ps = 0.f;
#pragma unroll
      for (int it = 0; it < FA_TILE / 4; ++it) {
        s[it][h] = fa_exp2(s[it][h] - mn);
        ps += s[it][h];
      }
      l[h] = fmaf(l[h], alpha[h], ps);
#pragma unroll
      for (int d = 0; d < 16; ++d) acc[h][d] *= alpha[h];
    }
#pragma unroll
    for (int it = 0; it < FA_TILE / 4; ++it) {
      const __nv_bfloat16* vr = vt + tokc[it] * kv_dim;
      float vf[16];
      fa_unpack8(*reinterpret_cast<const uint4*>(vr + 8 * j), vf);
      fa_unpack8(*reinterpret_cast<const uint4*>(vr + 64 + 8 * j), vf + 8);
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
        float o[16];
#pragma unroll
        for (int d = 0; d < 16; ++d) o[d] = acc[h][d] * inv;
        __nv_bfloat16* op = p.out + ((int64_t)row * p.num_heads + head0 + h) * FA_HD;
        stg_v4(op + 8 * j, fa_pack8(o));
        stg_v4(op + 64 + 8 * j, fa_pack8(o + 8));
      }
    }
    return;
  }

  // ---- split-KV: partials out, then the last CTA of this sequence folds all of them
  if (g == 0) {
#pragma unroll
    for (int h = 0; h < GROUP; ++h) {
      const int64_t idx = ((int64_t)row * p.nsplit + split) * p.num_heads + head0 + h;
      float* pa = p.part_acc + idx * FA_HD;
      *reinterpret_cast<float4*>(pa + 8 * j) = make_float4(acc[h][0], acc[h][1], acc[h][2], acc[h][3]);
      *reinterpret_cast<float4*>(pa + 8 * j + 4) = make_float4(acc[h][4], acc[h][5], acc[h][6], acc[h][7]);
      *reinterpret_cast<float4*>(pa + 64 + 8 * j) = make_float4(acc[h][8], acc[h][9], acc[h][10], acc[h][11]);
      *reinterpret_cast<float4*>(pa + 64 + 8 * j + 4) = make_float4(acc[h][12], acc[h][13], acc[h][14], acc[h][15]);
      if (j == 0) {
        p.part_ml[idx * 2] = my_tiles > 0 ? m[h] : -INFINITY;
        p.part_ml[idx * 2 + 1] = my_tiles > 0 ? l[h] : 0.f;
      }
    }
  }
  // One thread releases the CTA's partials after the CTA barrier (cumulative over the other threads' stores, as in a
  // cooperative-groups grid sync) and acquires for the CTA when it turns out to be the last: a __threadfence() per
  // thread is a MEMBAR.SC.GPU + L1 invalidate each (~1-2 us with 256 threads storing).  The readers below use ld.cg.
  asm volatile("bar.sync 1, %0;" :: "r"(kvh * 32) : "memory");      // the consumer warps only
  if (threadIdx.x == 0) {
    asm volatile("fence.acq_rel.gpu;" ::: "memory");
    const int t = atomicAdd(p.tickets + row, 1);
    asm volatile("fence.acq_rel.gpu;" ::: "memory");
    *last_flag = (t == p.nsplit - 1);
    if (t == p.nsplit - 1) p.tickets[row] = 0;                     // ready for the next launch
  }
  asm volatile("bar.sync 1, %0;" :: "r"(kvh * 32) : "memory");
  if (!*last_flag) return;
  // warp w folds heads head0 .. head0+GROUP-1; lane owns 4 consecutive dims
#pragma unroll
  for (int h = 0; h < GROUP; ++h) {
    const int head = head0 + h;
    const int64_t base = ((int64_t)row * p.nsplit) * p.num_heads + head;
    float M = -INFINITY;
    for (int s = 0; s < p.nsplit; ++s) M = fmaxf(M, __ldcg(p.part_ml + (base + (int64_t)s * p.num_heads) * 2));
    float4 num = make_float4(0.f, 0.f, 0.f, 0.f);
    float den = 0.f;
    for (int s = 0; s < p.nsplit; ++s) {
      const int64_t idx = base + (int64_t)s * p.num_heads;
      const float ls = __ldcg(p.part_ml + idx * 2 + 1);
      if (ls > 0.f) {
        const float w = fa_exp2(__ldcg(p.part_ml + idx * 2) - M);
        const float4 a = __ldcg(reinterpret_cast<const float4*>(p.part_acc + idx * FA_HD) + lane);
        num.x = fmaf(w, a.x, num.x); num.y = fmaf(w, a.y, num.y);
        num.z = fmaf(w, a.z, num.z); num.w = fmaf(w, a.w, num.w);
        den = fmaf(w, ls, den);
      }
    }
    const float inv = den > 0.f ? 1.f / den : 0.f;
    uint2 o;
    o.x = pack_bf16(num.x * inv, num.y * inv);
    o.y = pack_bf16(num.z * inv, num.w * inv);
    stg_v2(p.out + ((int64_t)row * p.num_heads + head) * FA_HD + 4 * lane, o);
  }
}

// ---------------------------------------------------------------------------------------------------------
// Short contexts (the code predictor: 3..17 tokens, 75 launches per codec frame).  The kernel above is built to stream
// long contexts (bulk-copy ring, producer warp, split-KV tickets); for a handful of tokens all of that is latency.
// Here one WARP owns one (sequence, kv head) item: no shared-memory ring, no block barrier, no split - the cached
// rows are read straight from the paged cache, two 8-token tiles per round trip, and the new token never leaves
// the warp.  Arithmetic, lane mapping and summation order are those of fused_decode_attn_kernel with nsplit == 1
// (GROUP 2, 1-D RoPE), so outputs and K / V cache bytes are bit-identical.
constexpr int FS_WARPS = 4;
constexpr int FS_MAX_CTX = 64;

template <bool EAGER>
__global__ void __launch_bounds__(FS_WARPS * 32)
fused_decode_attn_small_kernel(const FusedAttnParams p, int rows) {
  constexpr int GROUP = 2;
  __shared__ __align__(16) __nv_bfloat16 scr_all[FS_WARPS][(GROUP + 2) * FA_HD];   // q [2][128], k [128], v [128] of the new token
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kvh = p.num_kv_heads, kv_dim = kvh * FA_HD;
  const int item = blockIdx.x * FS_WARPS + warp;
  pdl_launch();
  if (item >= rows * kvh) return;
  const int row = item / kvh, hk = item - row * kvh;
  const int head0 = hk * GROUP;
  const int g = lane >> 3, j = lane & 7;
  __nv_bfloat16* q_s = scr_all[warp];
  __nv_bfloat16* kn_s = q_s + GROUP * FA_HD;
  __nv_bfloat16* vn_s = kn_s + FA_HD;
  // independent of the previous kernel: rotary frequencies, head-norm weights
  float frq[2];
  frq[0] = __ldg(p.inv_freq + 2 * lane);
  frq[1] = __ldg(p.inv_freq + 2 * lane + 1);
  uint32_t qw1 = 0, qw2 = 0, kw1 = 0, kw2 = 0;
  if (p.q_w) { qw1 = *reinterpret_cast<const uint32_t*>(p.q_w + 2 * lane); qw2 = *reinterpret_cast<const uint32_t*>(p.q_w + 64 + 2 * lane); }
  if (p.k_w) { kw1 = *reinterpret_cast<const uint32_t*>(p.k_w + 2 * lane); kw2 = *reinterpret_cast<const uint32_t*>(p.k_w + 64 + 2 * lane); }
  pdl_wait();
  // ---- one batch of independent requests
  int L = __ldg(p.context_lens + row);
  const int slot = __ldg(p.slot_mapping + row);
  const int64_t pos = __ldg(p.positions + row);
  const int32_t* bt = p.block_table + (int64_t)row * p.bt_stride;
  int page0 = __ldg(bt);
  uint32_t raw1[GROUP + 1], raw2[GROUP + 1];
#pragma unroll
  for (int i = 0; i < GROUP + 1; ++i) {
    const __nv_bfloat16* src = i == GROUP ? p.k + (int64_t)row * p.k_stride + hk * FA_HD
                                          : p.q + (int64_t)row * p.q_stride + (head0 + i) * FA_HD;
    raw1[i] = *reinterpret_cast<const uint32_t*>(src + 2 * lane);
    raw2[i] = *reinterpret_cast<const uint32_t*>(src + 64 + 2 * lane);
  }
  uint4 vraw = make_uint4(0, 0, 0, 0);
  if (lane < 16) vraw = ldg_v4(p.v + (int64_t)row * p.v_stride + hk * FA_HD + 8 * lane);
  if (L <= 0) {                   // CUDA-graph padding row: zeros
    if (g == 0) {
#pragma unroll
      for (int h = 0; h < GROUP; ++h) {
        __nv_bfloat16* o = p.out + ((int64_t)row * p.num_heads + head0 + h) * FA_HD;
        stg_v4(o + 8 * j, make_uint4(0, 0, 0, 0));
        stg_v4(o + 64 + 8 * j, make_uint4(0, 0, 0, 0));
      }
    }
    return;
  }
  if (page0 < 0) page0 = 0;
  // a token's K / V row of this head in the paged cache
  auto tok_off = [&](int t) -> int64_t {
    int page = page0;
    if (t >= p.page_size) { page = __ldg(bt + t / p.page_size); if (page < 0) page = 0; }
    return ((int64_t)page * p.page_size + t % p.page_size) * kv_dim + hk * FA_HD;
  };
  // the cached rows of the first two tiles are requested NOW, under the prologue arithmetic
  const int n_tiles = (L + FA_TILE - 1) / FA_TILE;
  uint4 kbuf[2][2][2], vbuf[2][2][2];            // [tile of the pair][token slot][half]
  auto load_pair = [&](int i0) {
#pragma unroll
    for (int u = 0; u < 2; ++u)
#pragma unroll
      for (int s = 0; s < 2; ++s) {
        const int tok0 = (i0 + u) * FA_TILE;
        const int ntok = min(FA_TILE, L - tok0);
        const int tok = s * 4 + g;
        const int t = tok0 + (tok < ntok ? tok : 0);
        if (i0 + u < n_tiles && t != L - 1) {
          const int64_t off = tok_off(t);
          kbuf[u][s][0] = ldg_v4(p.k_cache + off + 8 * j);
          kbuf[u][s][1] = ldg_v4(p.k_cache + off + 64 + 8 * j);
          vbuf[u][s][0] = ldg_v4(p.v_cache + off + 8 * j);
          vbuf[u][s][1] = ldg_v4(p.v_cache + off + 64 + 8 * j);
        }
      }
  };
  load_pair(0);
  // ---- prologue: q/k head norm + RoPE of the new token (the body of qknorm_rope_kvwrite_kernel), K / V to the cache
  {
    float c1[2], s1[2];
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const float a1 = (float)pos * frq[e];
      c1[e] = rbf(cosf(a1) * p.attention_scaling);
      s1[e] = rbf(sinf(a1) * p.attention_scaling);
    }
#pragma unroll
    for (int i = 0; i < GROUP + 1; ++i) {
      const bool is_k = i == GROUP;
      const bool has_w = is_k ? (p.k_w != nullptr) : (p.q_w != nullptr);
      float x1[2] = {bf16lo(raw1[i]), bf16hi(raw1[i])};
      float x2[2] = {bf16lo(raw2[i]), bf16hi(raw2[i])};
      if (has_w) {
        float ss = x1[0] * x1[0];
        ss = fmaf(x1[1], x1[1], ss); ss = fmaf(x2[0], x2[0], ss); ss = fmaf(x2[1], x2[1], ss);
        const float var = warp_sum(ss) / (float)FA_HD + p.eps;
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
      __nv_bfloat16* dst = is_k ? kn_s : q_s + i * FA_HD;
      *reinterpret_cast<uint32_t*>(dst + 2 * lane) = y1;
      *reinterpret_cast<uint32_t*>(dst + 64 + 2 * lane) = y2;
      if (is_k && slot >= 0) {
        __nv_bfloat16* cd = p.k_cache + (int64_t)slot * kv_dim + hk * FA_HD;
        *reinterpret_cast<uint32_t*>(cd + 2 * lane) = y1;
        *reinterpret_cast<uint32_t*>(cd + 64 + 2 * lane) = y2;
      }
    }
    if (lane < 16) {
      *reinterpret_cast<uint4*>(vn_s + 8 * lane) = vraw;
      if (slot >= 0) stg_v4(p.v_cache + (int64_t)slot * kv_dim + hk * FA_HD + 8 * lane, vraw);
    }
  }
  __syncwarp();
  float qv[GROUP][16];
#pragma unroll
  for (int h = 0; h < GROUP; ++h) {
    const __nv_bfloat16* qp = q_s + h * FA_HD;
    fa_unpack8(*reinterpret_cast<const uint4*>(qp + 8 * j), qv[h]);
    fa_unpack8(*reinterpret_cast<const uint4*>(qp + 64 + 8 * j), qv[h] + 8);
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
  for (int i0 = 0; i0 < n_tiles; i0 += 2) {
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (i0 + u >= n_tiles) break;
      const int tok0 = (i0 + u) * FA_TILE;
      const int ntok = min(FA_TILE, L - tok0);
      float s[FA_TILE / 4][GROUP];
      bool isnew[FA_TILE / 4];
#pragma unroll
      for (int it = 0; it < FA_TILE / 4; ++it) {
        const int tok = it * 4 + g;
        const bool valid = tok < ntok;
        isnew[it] = tok0 + (valid ? tok : 0) == L - 1;
        float kf[16];
        if (isnew[it]) {
          fa_unpack8(*reinterpret_cast<const uint4*>(kn_s + 8 * j), kf);
          fa_unpack8(*reinterpret_cast<const uint4*>(kn_s + 64 + 8 * j), kf + 8);
        } else {
          fa_unpack8(kbuf[u][it][0], kf);
          fa_unpack8(kbuf[u][it][1], kf + 8);
        }
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
        for (int it = 1; it < FA_TILE / 4; ++it) mt = fmaxf(mt, s[it][h]);
        mt = fmaxf(mt, __shfl_xor_sync(0xffffffffu, mt, 8));
        mt = fmaxf(mt, __shfl_xor_sync(0xffffffffu, mt, 16));
        const float mn = fmaxf(m[h], mt);
        alpha[h] = fa_exp2(m[h] - mn);
        m[h] = mn;
        float ps = 0.f;
#pragma unroll
        for (int it = 0; it < FA_TILE / 4; ++it) {
          s[it][h] = fa_exp2(s[it][h] - mn);
          ps += s[it][h];
        }
        l[h] = fmaf(l[h], alpha[h], ps);
#pragma unroll
        for (int d = 0; d < 16; ++d) acc[h][d] *= alpha[h];
      }
#pragma unroll
      for (int it = 0; it < FA_TILE / 4; ++it) {
        float vf[16];
        if (isnew[it]) {
          fa_unpack8(*reinterpret_cast<const uint4*>(vn_s + 8 * j), vf);
          fa_unpack8(*reinterpret_cast<const uint4*>(vn_s + 64 + 8 * j), vf + 8);
        } else {
          fa_unpack8(vbuf[u][it][0], vf);
          fa_unpack8(vbuf[u][it][1], vf + 8);
        }
#pragma unroll
        for (int h = 0; h < GROUP; ++h)
#pragma unroll
          for (int d = 0; d < 16; ++d) acc[h][d] = fmaf(s[it][h], vf[d], acc[h][d]);
      }
    }
    if (i0 + 2 < n_tiles) load_pair(i0 + 2);
  }
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
  if (g == 0) {
#pragma unroll
    for (int h = 0; h < GROUP; ++h) {
      const float inv = 1.f / l[h];
      float o[16];
#pragma unroll
      for (int d = 0; d < 16; ++d) o[d] = acc[h][d] * inv;
      __nv_bfloat16* op = p.out + ((int64_t)row * p.num_heads + head0 + h) * FA_HD;
      stg_v4(op + 8 * j, fa_pack8(o));
      stg_v4(op + 64 + 8 * j, fa_pack8(o + 8));
    }
  }
}

static size_t fa_smem_bytes(int kv_heads, int num_heads) {
  return (size_t)2 * FA_NSTAGE * FA_TILE * kv_heads * FA_HD * 2 + 2 * FA_NSTAGE * sizeof(uint64_t) + 16 +
         (size_t)(num_heads + 2 * kv_heads) * FA_HD * 2;
}

// split-KV factor: one wave of 2 CTAs per SM, at least 8 tiles (64 tokens) of table capacity per split
static int fa_choose_nsplit(int rows, int max_ctx) {
  if (const char* e = getenv("B2TTS_FA_NSPLIT")) {             // experiments only
    const int v = atoi(e);
    if (v >= 1 && v <= 32) return v;
  }
  const int max_tiles = (max_ctx + FA_TILE - 1) / FA_TILE;
  int n = (2 * sm_count()) / (rows > 0 ? rows : 1);
  const int cap = max_tiles / 8;
  if (n > cap) n = cap;
  if (n < 1) n = 1;
  if (n > 32) n = 32;
  return n;
}

}  // namespace b2

extern "C" size_t b2tts_fused_decode_attn_workspace_bytes(int batch, int num_heads, int head_dim, int max_ctx) {
  if (batch <= 0 || num_heads <= 0 || head_dim <= 0 || max_ctx <= 0) return 0;
  const int nsplit = b2::fa_choose_nsplit(batch, max_ctx);
  // fixed ticket span (zero before the first launch, left zero by every launch) + partials + (m, l)
  return b2::FA_TICKET_BYTES + (size_t)batch * nsplit * num_heads * (head_dim + 2) * sizeof(float) + 16;
}

extern "C" int b2tts_fused_decode_attn(const void* q, int64_t q_row_stride, const void* k, int64_t k_row_stride,
                                       const void* v, int64_t v_row_stride,
                                       const void* q_weight, const void* k_weight, float eps,
                                       const int64_t* positions, int64_t pos_row_stride,
                                       const float* inv_freq, const int32_t* modality, float attention_scaling,
                                       const int32_t* slot_mapping, void* k_cache, void* v_cache,
                                       const int32_t* block_table, int bt_stride, const int32_t* context_lens,
                                       void* out, void* workspace, size_t workspace_bytes,
                                       int batch, int num_heads, int num_kv_heads, int head_dim, int page_size,
                                       float scale, int max_ctx, int rounding, int pdl, void* stream) {
  using namespace b2;
  const char* what = "fused_decode_attn";
  B2_REQUIRE(batch >= 0 && max_ctx > 0 && page_size > 0 && bt_stride > 0, B2TTS_EINVAL,
             "%s: batch=%d max_ctx=%d page_size=%d bt_stride=%d", what, batch, max_ctx, page_size, bt_stride);
  if (batch == 0) return B2TTS_OK;
  B2_REQUIRE(batch <= FA_MAX_ROWS, B2TTS_EUNSUPPORTED, "%s: batch %d > %d rows", what, batch, FA_MAX_ROWS);
  B2_REQUIRE(q && k && v && positions && inv_freq && slot_mapping && k_cache && v_cache && block_table && context_lens &&
                 out && workspace, B2TTS_EINVAL, "%s: null pointer", what);
  B2_REQUIRE(head_dim == FA_HD, B2TTS_EUNSUPPORTED, "%s: head_dim %d unsupported (128)", what, head_dim);
  B2_REQUIRE(num_kv_heads >= 1 && num_kv_heads <= 8 && num_heads % num_kv_heads == 0, B2TTS_EUNSUPPORTED,
             "%s: heads %d / kv heads %d unsupported", what, num_heads, num_kv_heads);
  B2_REQUIRE(page_size % FA_TILE == 0, B2TTS_EUNSUPPORTED, "%s: page_size %d must be a multiple of %d", what, page_size,
             FA_TILE);
  B2_REQUIRE(rounding == B2TTS_ROUND_EAGER || rounding == B2TTS_ROUND_FUSED, B2TTS_EINVAL, "%s: rounding %d", what, rounding);
  B2_REQUIRE(aligned16(q) && aligned16(k) && aligned16(v) && aligned16(k_cache) && aligned16(v_cache) && aligned16(out) &&
                 aligned16(workspace) && q_row_stride % 8 == 0 && k_row_stride % 8 == 0 && v_row_stride % 8 == 0 &&
                 (!q_weight || aligned16(q_weight)) && (!k_weight || aligned16(k_weight)),
             B2TTS_EALIGN, "%s: pointers / strides must be 16-byte aligned", what);
  FusedAttnParams p{};
  p.q = static_cast<const __nv_bfloat16*>(q);
  p.k = static_cast<const __nv_bfloat16*>(k);
  p.v = static_cast<const __nv_bfloat16*>(v);
  p.q_stride = q_row_stride; p.k_stride = k_row_stride; p.v_stride = v_row_stride;
  p.q_w = static_cast<const __nv_bfloat16*>(q_weight);
  p.k_w = static_cast<const __nv_bfloat16*>(k_weight);
  p.eps = eps;
  p.positions = positions;
  p.pos_row_stride = pos_row_stride;
  p.inv_freq = inv_freq;
  p.modality = modality;
  p.attention_scaling = attention_scaling;
  p.slot_mapping = slot_mapping;
  p.k_cache = static_cast<__nv_bfloat16*>(k_cache);
  p.v_cache = static_cast<__nv_bfloat16*>(v_cache);
  p.block_table = block_table;
  p.bt_stride = bt_stride;
  p.page_size = page_size;
  p.context_lens = context_lens;
  p.out = static_cast<__nv_bfloat16*>(out);
  p.nsplit = fa_choose_nsplit(batch, max_ctx);
  p.num_heads = num_heads;
  p.num_kv_heads = num_kv_heads;
  p.scale_log2 = scale * 1.4426950408889634f;
  const size_t acc_elems = (size_t)batch * p.nsplit * num_heads * head_dim;
  const size_t need = b2tts_fused_decode_attn_workspace_bytes(batch, num_heads, head_dim, max_ctx);
  B2_REQUIRE(workspace_bytes >= need, B2TTS_EWORKSPACE, "%s: workspace %zu B < %zu B needed", what, workspace_bytes, need);
  p.tickets = static_cast<int*>(workspace);
  p.part_acc = reinterpret_cast<float*>(static_cast<uint8_t*>(workspace) + FA_TICKET_BYTES);
  p.part_ml = p.part_acc + acc_elems;

  const int group = num_heads / num_kv_heads;
  const bool eager = rounding == B2TTS_ROUND_EAGER;
  static const bool small_ok = [] { const char* e = getenv("B2TTS_FA_SMALL"); return !(e && e[0] == '0'); }();
  if (small_ok && max_ctx <= FS_MAX_CTX && group == 2 && !modality && p.nsplit == 1) {
    // a handful of cached tokens: one warp per (sequence, kv head), no ring, no split
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((batch * num_kv_heads + FS_WARPS - 1) / FS_WARPS, 1, 1);
    cfg.blockDim = dim3(FS_WARPS * 32, 1, 1);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = static_cast<cudaStream_t>(stream);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    cudaError_t e = eager ? cudaLaunchKernelEx(&cfg, fused_decode_attn_small_kernel<true>, p, batch)
                          : cudaLaunchKernelEx(&cfg, fused_decode_attn_small_kernel<false>, p, batch);
    count_launch();
    if (e != cudaSuccess) return cuda_fail(e, what);
    return B2TTS_OK;
  }
  void (*kern)(const FusedAttnParams) = nullptr;
  if (group == 1) kern = eager ? fused_decode_attn_kernel<1, true> : fused_decode_attn_kernel<1, false>;
  else if (group == 2) kern = eager ? fused_decode_attn_kernel<2, true> : fused_decode_attn_kernel<2, false>;
  else if (group == 4) kern = eager ? fused_decode_attn_kernel<4, true> : fused_decode_attn_kernel<4, false>;
  B2_REQUIRE(kern, B2TTS_EUNSUPPORTED, "%s: GQA group %d unsupported (1, 2, 4)", what, group);
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  if (e != cudaSuccess) return cuda_fail(e, what);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(batch, p.nsplit, 1);
  cfg.blockDim = dim3(32 * (num_kv_heads + 1), 1, 1);
  cfg.dynamicSmemBytes = fa_smem_bytes(num_kv_heads, num_heads);
  cfg.stream = static_cast<cudaStream_t>(stream);
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  e = cudaLaunchKernelEx(&cfg, kern, p);
  count_launch();
  if (e != cudaSuccess) return cuda_fail(e, what);
  return B2TTS_OK;
}

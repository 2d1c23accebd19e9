// elementwise.cu - a6 SiLU / SiLU*mul, a9 row gathers and the frame-loop codebook embedding sum.
// Streaming 16-byte vector passes; grids are sized to cover the data once (no reuse to tile for).
#include "common.cuh"

namespace b2 {

__device__ __forceinline__ float silu_f(float g) { return g / (1.0f + expf(-g)); }

template <bool EAGER, bool MUL>
__global__ void __launch_bounds__(256)
silu_mul_kernel(const __nv_bfloat16* __restrict__ gate, const __nv_bfloat16* __restrict__ up, int64_t in_stride,
                __nv_bfloat16* __restrict__ out, int rows, int inter) {
  const int nvec = inter >> 3;
  const int64_t total = (int64_t)rows * nvec;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int row = (int)(i / nvec), v = (int)(i % nvec);
    const uint4 g = ldg_nc_v4(gate + (int64_t)row * in_stride + 8 * v);
    uint4 u = make_uint4(0, 0, 0, 0);
    if (MUL) u = ldg_nc_v4(up + (int64_t)row * in_stride + 8 * v);
    const uint32_t gw[4] = {g.x, g.y, g.z, g.w};
    const uint32_t uw[4] = {u.x, u.y, u.z, u.w};
    uint32_t ow[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      float a = silu_f(bf16lo(gw[e])), b = silu_f(bf16hi(gw[e]));
      if (MUL) {
        if (EAGER) { a = rbf(a); b = rbf(b); }       // eager: silu rounds, then the product rounds
        a *= bf16lo(uw[e]);
        b *= bf16hi(uw[e]);
      }
      ow[e] = pack_bf16(a, b);
    }
    stg_v4(out + (int64_t)row * inter + 8 * v, make_uint4(ow[0], ow[1], ow[2], ow[3]));
  }
}

// out[i] = tables[table_index[i]][row_index[i]]; one warp per output row.
__global__ void __launch_bounds__(256)
gather_rows_kernel(const __nv_bfloat16* __restrict__ tables, int64_t table_stride, int rows_per_table,
                   const int64_t* __restrict__ row_index, const int32_t* __restrict__ table_index,
                   __nv_bfloat16* __restrict__ out, int64_t out_stride, int n, int cols) {
  const int i = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (i >= n) return;
  const int64_t r = __ldg(row_index + i);
  const int t = table_index ? __ldg(table_index + i) : 0;
  const bool ok = r >= 0 && r < rows_per_table && t >= 0;
  const __nv_bfloat16* src = tables + (int64_t)t * table_stride + r * cols;
  __nv_bfloat16* dst = out + (int64_t)i * out_stride;
  for (int v = lane; v < (cols >> 3); v += 32) stg_v4(dst + 8 * v, ok ? ldg_v4(src + 8 * v) : make_uint4(0, 0, 0, 0));
}

// out[i] = bf16( bf16(sum_c f32(E_c[codes[i,c]])) + extra[i] ): interface.py:1276-1286 - the
// 16-codebook sum is one fp32 accumulate rounded once (`codec_hiddens.sum(1)` on bf16), then a
// bf16 add of the trailing-text / pad embedding.  One CTA per row, a 16-byte vector per thread.
__global__ void __launch_bounds__(256)
codec_embed_sum_kernel(const __nv_bfloat16* __restrict__ table0, const __nv_bfloat16* __restrict__ tables_rest,
                       int64_t table_stride, const int64_t* __restrict__ codes, int64_t codes_stride,
                       const __nv_bfloat16* __restrict__ extra, int64_t extra_stride,
                       const int64_t* __restrict__ extra_index,
                       __nv_bfloat16* __restrict__ out, int ncb, int cols) {
  const int i = blockIdx.x;
  const int64_t erow = extra_index ? __ldg(extra_index + i) : (int64_t)i;
  for (int v = threadIdx.x; v < (cols >> 3); v += blockDim.x) {
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int c = 0; c < ncb; ++c) {
      const int64_t code = __ldg(codes + (int64_t)i * codes_stride + c);
      const __nv_bfloat16* row = (c == 0 ? table0 : tables_rest + (int64_t)(c - 1) * table_stride) + code * cols;
      const uint4 r = ldg_v4(row + 8 * v);
      acc[0] += bf16lo(r.x); acc[1] += bf16hi(r.x); acc[2] += bf16lo(r.y); acc[3] += bf16hi(r.y);
      acc[4] += bf16lo(r.z); acc[5] += bf16hi(r.z); acc[6] += bf16lo(r.w); acc[7] += bf16hi(r.w);
    }
    if (extra) {
      const uint4 e = ldg_v4(extra + erow * extra_stride + 8 * v);
      acc[0] = rbf(acc[0]) + bf16lo(e.x); acc[1] = rbf(acc[1]) + bf16hi(e.x);
      acc[2] = rbf(acc[2]) + bf16lo(e.y); acc[3] = rbf(acc[3]) + bf16hi(e.y);
      acc[4] = rbf(acc[4]) + bf16lo(e.z); acc[5] = rbf(acc[5]) + bf16hi(e.z);
      acc[6] = rbf(acc[6]) + bf16lo(e.w); acc[7] = rbf(acc[7]) + bf16hi(e.w);
    }
    uint4 o;
    o.x = pack_bf16(acc[0], acc[1]); o.y = pack_bf16(acc[2], acc[3]);
    o.z = pack_bf16(acc[4], acc[5]); o.w = pack_bf16(acc[6], acc[7]);
    stg_v4(out + (int64_t)i * cols + 8 * v, o);
  }
}

}  // namespace b2

extern "C" int b2tts_silu_and_mul(const void* gate, const void* up, int64_t in_row_stride,
                                  void* out, int rows, int inter, int rounding, void* stream) {
  using namespace b2;
  B2_REQUIRE(rows >= 0 && inter > 0, B2TTS_EINVAL, "silu_and_mul: rows=%d inter=%d", rows, inter);
  if (rows == 0) return B2TTS_OK;
  B2_REQUIRE(gate && out, B2TTS_EINVAL, "silu_and_mul: null pointer");
  B2_REQUIRE(inter % 8 == 0 && in_row_stride % 8 == 0 && aligned16(gate) && aligned16(up) && aligned16(out),
             B2TTS_EALIGN, "silu_and_mul: inter/stride must be multiples of 8 and pointers 16-byte aligned");
  B2_REQUIRE(rounding == B2TTS_ROUND_EAGER || rounding == B2TTS_ROUND_FUSED, B2TTS_EINVAL, "silu_and_mul: rounding %d", rounding);
  const int64_t total = (int64_t)rows * (inter >> 3);
  int64_t blocks = (total + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  auto g = static_cast<const __nv_bfloat16*>(gate);
  auto u = static_cast<const __nv_bfloat16*>(up);
  auto o = static_cast<__nv_bfloat16*>(out);
  if (!up) silu_mul_kernel<true, false><<<(int)blocks, 256, 0, st>>>(g, u, in_row_stride, o, rows, inter);
  else if (rounding == B2TTS_ROUND_EAGER) silu_mul_kernel<true, true><<<(int)blocks, 256, 0, st>>>(g, u, in_row_stride, o, rows, inter);
  else silu_mul_kernel<false, true><<<(int)blocks, 256, 0, st>>>(g, u, in_row_stride, o, rows, inter);
  B2_CHECK_LAUNCH("silu_and_mul");
  return B2TTS_OK;
}

extern "C" int b2tts_gather_rows(const void* tables, int64_t table_stride, int rows_per_table,
                                 const int64_t* row_index, const int32_t* table_index,
                                 void* out, int64_t out_row_stride, int n, int cols, void* stream) {
  using namespace b2;
  B2_REQUIRE(n >= 0 && cols > 0 && rows_per_table > 0, B2TTS_EINVAL, "gather_rows: n=%d cols=%d rows=%d", n, cols, rows_per_table);
  if (n == 0) return B2TTS_OK;
  B2_REQUIRE(tables && row_index && out, B2TTS_EINVAL, "gather_rows: null pointer");
  B2_REQUIRE(cols % 8 == 0 && table_stride % 8 == 0 && out_row_stride % 8 == 0 && aligned16(tables) && aligned16(out),
             B2TTS_EALIGN, "gather_rows: cols/strides must be multiples of 8 and pointers 16-byte aligned");
  gather_rows_kernel<<<(n + 7) / 8, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(tables), table_stride, rows_per_table, row_index, table_index,
      static_cast<__nv_bfloat16*>(out), out_row_stride, n, cols);
  B2_CHECK_LAUNCH("gather_rows");
  return B2TTS_OK;
}

extern "C" int b2tts_codec_embed_sum(const void* table0, const void* tables_rest, int64_t table_stride,
                                     const int64_t* codes, int64_t codes_row_stride,
                                     const void* extra, int64_t extra_row_stride, const int64_t* extra_index,
                                     void* out, int n, int num_codebooks, int cols, void* stream) {
  using namespace b2;
  B2_REQUIRE(n >= 0 && cols > 0 && num_codebooks >= 1, B2TTS_EINVAL, "codec_embed_sum: n=%d cols=%d ncb=%d", n, cols, num_codebooks);
  if (n == 0) return B2TTS_OK;
  B2_REQUIRE(table0 && codes && out && (num_codebooks == 1 || tables_rest), B2TTS_EINVAL, "codec_embed_sum: null pointer");
  B2_REQUIRE(cols % 8 == 0 && table_stride % 8 == 0 && extra_row_stride % 8 == 0 && aligned16(table0) &&
                 aligned16(tables_rest) && aligned16(extra) && aligned16(out),
             B2TTS_EALIGN, "codec_embed_sum: cols/strides must be multiples of 8 and pointers 16-byte aligned");
  codec_embed_sum_kernel<<<n, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(table0), static_cast<const __nv_bfloat16*>(tables_rest), table_stride, codes,
      codes_row_stride, static_cast<const __nv_bfloat16*>(extra), extra_row_stride, extra ? extra_index : nullptr,
      static_cast<__nv_bfloat16*>(out), num_codebooks, cols);
  B2_CHECK_LAUNCH("codec_embed_sum");
  return B2TTS_OK;
}

// ---- a12 on the device: the decode layout of prepare_decode / prepare_decode_talker
// (engine/model_runner/base.py:218-234, talker_mode_runner.py:116-140) computed from a device-resident
// sequence-length vector, so that a frame loop can run without the host touching the layout.
// seq_lens[i] = tokens of stream i INCLUDING the one decoded this step (0 = inactive row):
//   context_lens[i] = len;  slot[i] = block_table[i][(len-1)/page]*page + (len-1)%page;
//   positions[i]    = len - 1 + position_offset   (offset 1 for the talker, 0 for the predictor);
// then, if advance != 0, seq_lens[i] += advance for active rows (the next step's length).
namespace b2 {
__global__ void decode_layout_kernel(int32_t* __restrict__ seq_lens, const int32_t* __restrict__ block_table,
                                     int bt_stride, int bt_width, int page, int position_offset, int advance,
                                     int64_t* __restrict__ positions, int32_t* __restrict__ slot_mapping,
                                     int32_t* __restrict__ context_lens, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int len = seq_lens[i];
  if (len <= 0) {
    positions[i] = 0; slot_mapping[i] = -1; context_lens[i] = 0;
    return;
  }
  const int t = len - 1;
  if (t / page >= bt_width) {            // beyond the table: never index another stream's row (the row goes idle)
    positions[i] = 0; slot_mapping[i] = -1; context_lens[i] = 0;
    return;
  }
  const int pg = block_table[(int64_t)i * bt_stride + t / page];
  positions[i] = (int64_t)t + position_offset;
  slot_mapping[i] = pg < 0 ? -1 : pg * page + t % page;
  context_lens[i] = len;
  if (advance) seq_lens[i] = len + advance;
}
}  // namespace b2

extern "C" int b2tts_decode_layout(int32_t* seq_lens, const int32_t* block_table, int bt_stride, int bt_width,
                                   int page_size, int position_offset, int advance, int64_t* positions,
                                   int32_t* slot_mapping, int32_t* context_lens, int n, void* stream) {
  using namespace b2;
  B2_REQUIRE(n >= 0 && page_size > 0 && bt_stride > 0 && bt_width > 0 && bt_width <= bt_stride, B2TTS_EINVAL,
             "decode_layout: n=%d page=%d bt_stride=%d bt_width=%d", n, page_size, bt_stride, bt_width);
  if (n == 0) return B2TTS_OK;
  B2_REQUIRE(seq_lens && block_table && positions && slot_mapping && context_lens, B2TTS_EINVAL, "decode_layout: null pointer");
  decode_layout_kernel<<<(n + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      seq_lens, block_table, bt_stride, bt_width, page_size, position_offset, advance, positions, slot_mapping,
      context_lens, n);
  B2_CHECK_LAUNCH("decode_layout");
  return B2TTS_OK;
}

// ---- f1: per-frame stream bookkeeping of the frame loop (interface.py:1233-1239, 1282-1285) on the device.
// For stream i, with code0[i] the talker code sampled for this frame:
//   * code0 == eos_id (2150) ends the stream BEFORE the frame is produced (`send_clear_request; break`): active = 0,
//     seq_lens = 0, so the row becomes CUDA-graph padding for the rest of the run; so does a stream that has
//     produced max_frames frames;
//   * extra_row[i] = i * (trail_rows + 1) + min(gen_step, trail_rows): the row of a [n, trail_rows + 1, H] buffer
//     holding trailing_text_hiddens[:, generation_step] for generation_step < trail_rows and tts_pad_embed after;
//   * active streams count the frame: gen_step += 1, frames_done += 1.
namespace b2 {
__global__ void frame_advance_kernel(const int64_t* __restrict__ code0, int eos_id, int32_t* __restrict__ active,
                                     int32_t* __restrict__ seq_lens, int32_t* __restrict__ gen_step,
                                     int32_t* __restrict__ frames_done, int64_t* __restrict__ extra_row,
                                     int trail_rows, int max_frames, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int a = active[i];
  if (a && ((eos_id >= 0 && code0[i] == (int64_t)eos_id) || frames_done[i] >= max_frames)) {
    a = 0;
    active[i] = 0;
    seq_lens[i] = 0;
  }
  const int g = gen_step[i];
  extra_row[i] = (int64_t)i * (trail_rows + 1) + (g < trail_rows ? g : trail_rows);
  if (a) {
    gen_step[i] = g + 1;
    frames_done[i] += 1;
  }
}
}  // namespace b2

extern "C" int b2tts_frame_advance(const int64_t* code0, int eos_id, int32_t* active, int32_t* seq_lens,
                                   int32_t* gen_step, int32_t* frames_done, int64_t* extra_row, int trail_rows,
                                   int max_frames, int n, void* stream) {
  using namespace b2;
  B2_REQUIRE(n >= 0 && trail_rows >= 0 && max_frames >= 0, B2TTS_EINVAL, "frame_advance: n=%d trail_rows=%d max_frames=%d",
             n, trail_rows, max_frames);
  if (n == 0) return B2TTS_OK;
  B2_REQUIRE(code0 && active && seq_lens && gen_step && frames_done && extra_row, B2TTS_EINVAL, "frame_advance: null pointer");
  frame_advance_kernel<<<(n + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      code0, eos_id, active, seq_lens, gen_step, frames_done, extra_row, trail_rows, max_frames, n);
  B2_CHECK_LAUNCH("frame_advance");
  return B2TTS_OK;
}

// ------------------------------------------------------------------------------------------
// f3 - prompt assembly (utils/generation.py:18-379, 412-515 of the reference).  Every row of a talker prompt (and of
// the trailing text hiddens) is  text side (+) codec side : a projected text-embedding row, a codec / speaker / ICL
// reference-frame row, or their bf16 sum.  The host plans the rows of ALL requests of a batch as two index vectors
// and this one launch gathers, adds and scatters them into the left-padded [B, Tmax, H] batch the reference builds
// with ~10 tiny tensor ops per request: out[r] = a[a_idx[r]] + b[b_idx[r]], an index < 0 = absent side (both absent =
// zero padding row).  The add is torch's bf16 tensor add: bf16(f32(a) + f32(b)).
namespace b2 {
__global__ void __launch_bounds__(256)
prompt_assemble_kernel(const __nv_bfloat16* __restrict__ a, int64_t a_stride, const __nv_bfloat16* __restrict__ b,
                       int64_t b_stride, const int32_t* __restrict__ a_idx, const int32_t* __restrict__ b_idx,
                       __nv_bfloat16* __restrict__ out, int64_t out_stride, int cols) {
  const int r = blockIdx.x;
  const int ia = a_idx[r], ib = b_idx[r];
  const __nv_bfloat16* pa = ia >= 0 ? a + (int64_t)ia * a_stride : nullptr;
  const __nv_bfloat16* pb = ib >= 0 ? b + (int64_t)ib * b_stride : nullptr;
  __nv_bfloat16* po = out + (int64_t)r * out_stride;
  for (int v = threadIdx.x; v < (cols >> 3); v += blockDim.x) {
    uint4 o = make_uint4(0, 0, 0, 0);
    if (pa && pb) {
      const uint4 x = ldg_nc_v4(pa + 8 * v), y = ldg_nc_v4(pb + 8 * v);
      o.x = pack_bf16(bf16lo(x.x) + bf16lo(y.x), bf16hi(x.x) + bf16hi(y.x));
      o.y = pack_bf16(bf16lo(x.y) + bf16lo(y.y), bf16hi(x.y) + bf16hi(y.y));
      o.z = pack_bf16(bf16lo(x.z) + bf16lo(y.z), bf16hi(x.z) + bf16hi(y.z));
      o.w = pack_bf16(bf16lo(x.w) + bf16lo(y.w), bf16hi(x.w) + bf16hi(y.w));
    } else if (pa) {
      o = ldg_nc_v4(pa + 8 * v);
    } else if (pb) {
      o = ldg_nc_v4(pb + 8 * v);
    }
    stg_na_v4(po + 8 * v, o);
  }
}
}  // namespace b2

extern "C" int b2tts_prompt_assemble(const void* a, int64_t a_row_stride, int a_rows, const void* b, int64_t b_row_stride,
                                     int b_rows, const int32_t* a_index, const int32_t* b_index, void* out,
                                     int64_t out_row_stride, int rows, int cols, void* stream) {
  using namespace b2;
  const char* what = "prompt_assemble";
  B2_REQUIRE(rows >= 0 && cols > 0 && cols % 8 == 0 && a_rows >= 0 && b_rows >= 0, B2TTS_EINVAL,
             "%s: rows=%d cols=%d (multiple of 8)", what, rows, cols);
  if (rows == 0) return B2TTS_OK;
  B2_REQUIRE(a_index && b_index && out && (a || a_rows == 0) && (b || b_rows == 0), B2TTS_EINVAL, "%s: null pointer", what);
  B2_REQUIRE(aligned16(out) && (!a || aligned16(a)) && (!b || aligned16(b)) && a_row_stride % 8 == 0 && b_row_stride % 8 == 0 &&
                 out_row_stride % 8 == 0, B2TTS_EALIGN, "%s: pointers / strides must be 16-byte aligned", what);
  prompt_assemble_kernel<<<rows, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(a), a_row_stride, static_cast<const __nv_bfloat16*>(b), b_row_stride, a_index,
      b_index, static_cast<__nv_bfloat16*>(out), out_row_stride, cols);
  B2_CHECK_LAUNCH(what);
  return B2TTS_OK;
}

// ------------------------------------------------------------------------------------------
// f4 - the tail of the codec decode window (examples/server.py:338-349 of the reference): float waveform -> PCM16,
//   pcm = (int16) trunc(clip(x, -1, 1) * 32767)                       _float_to_pcm16   (numpy astype: toward zero)
//   pcm = (int16) clip(rint((float)pcm * volume), -32768, 32767)      _apply_volume_pcm16 (numpy round: half to even)
// on the device, so that a decoded window leaves the GPU as 2 bytes per sample instead of 4.
namespace b2 {
__global__ void __launch_bounds__(256)
pcm16_kernel(const float* __restrict__ wav, int16_t* __restrict__ out, int64_t n, float volume) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float x = wav[i];
    x = fminf(fmaxf(x, -1.0f), 1.0f);                  // NaN -> -1 like np.clip? np.clip keeps NaN; the cast is then UB there
    int v = (int)(x * 32767.0f);                       // truncation toward zero
    if (volume != 1.0f) {
      const float s = rintf((float)v * volume);        // round half to even
      v = (int)fminf(fmaxf(s, -32768.0f), 32767.0f);
    }
    out[i] = (int16_t)v;
  }
}
}  // namespace b2

extern "C" int b2tts_float_to_pcm16(const float* wav, int16_t* out, int64_t n, float volume, void* stream) {
  using namespace b2;
  B2_REQUIRE(n >= 0, B2TTS_EINVAL, "float_to_pcm16: n=%lld", (long long)n);
  if (n == 0) return B2TTS_OK;
  B2_REQUIRE(wav && out, B2TTS_EINVAL, "float_to_pcm16: null pointer");
  const int64_t want = (n + 255) / 256;
  const int blocks = (int)(want < 148 * 8 ? want : 148 * 8);
  pcm16_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(wav, out, n, volume);
  B2_CHECK_LAUNCH("float_to_pcm16");
  return B2TTS_OK;
}

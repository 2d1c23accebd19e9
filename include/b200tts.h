/*
 * b200tts.h - C ABI of libb200tts.so: the Qwen3-TTS decode hot path as hand-written
 * sm_100a CUDA kernels (paged KV write, paged GQA attention for decode and varlen
 * prefill, RMSNorm, q/k-norm + RoPE, SiLU*mul, top-k sampler, frame-loop gathers).
 *
 * The reference (tsdocode/nano-qwen3tts-vllm) has no FFI: its "operator API" for this
 * path is a set of Python callables.  Each entry point below replaces the arithmetic
 * of one of them; citations are file:line under /root/reference/nano-qwen3tts-vllm/.
 * The Python modules in nano-qwen3tts-vllm_b200/layers/ keep the reference's names and
 * signatures and call these through ctypes.
 *
 * Contract common to every entry point
 *   - plain pointers and sizes only; every pointer is a DEVICE pointer unless named host_*;
 *   - bf16 tensors are `const void*` / `void*` to 2-byte bfloat16, row-major, innermost
 *     stride 1; `*_stride` arguments are in ELEMENTS;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 *   - no allocation, no synchronisation, no host read of device memory: every call is
 *     capturable in a CUDA graph; scratch comes from the caller (`*_workspace_bytes`);
 *   - returns B2TTS_OK or a negative B2TTS_E* code; b2tts_last_error() gives a
 *     thread-local message.  Nothing throws, nothing calls exit().
 *   - rows with slot -1 / context length 0 are CUDA-graph padding and are skipped
 *     (talker_mode_runner.py:97-100, base.py:166-170).
 */
#ifndef B200TTS_H
#define B200TTS_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define B2TTS_API __attribute__((visibility("default")))
#else
#define B2TTS_API
#endif

#define B2TTS_OK            0
#define B2TTS_EINVAL       -1   /* bad size / null pointer / inconsistent arguments      */
#define B2TTS_EALIGN       -2   /* pointer or stride not 16-byte aligned                 */
#define B2TTS_EUNSUPPORTED -3   /* shape outside what the sm_100a kernels are built for  */
#define B2TTS_ECUDA        -4   /* CUDA launch error (message carries cudaGetErrorString) */
#define B2TTS_EWORKSPACE   -5   /* workspace too small                                   */

/* rounding modes of the element-wise kernels (see DESIGN.md "rounding points") */
#define B2TTS_ROUND_EAGER 0   /* the reference's Python source evaluated op by op in bf16 */
#define B2TTS_ROUND_FUSED 1   /* what torch-inductor emits for the same source on a GPU:
                                  fp32 throughout, one rounding at the store            */

B2TTS_API int         b2tts_version(void);                 /* ABI version, bumped on any signature change */
B2TTS_API const char* b2tts_last_error(void);              /* thread-local, never NULL */
/* Kernels launched (or recorded into a CUDA graph under capture) by this library in this process. */
B2TTS_API unsigned long long b2tts_launch_count(void);
/* SM count / max dynamic smem of the current device (host query, cached). */
B2TTS_API int         b2tts_device_info(int* sm_count, int* max_smem_bytes, int* cc_major, int* cc_minor);

/* ---- a1: paged KV write --------------------------------------------------------------
 * replaces store_kvcache_kernel / store_kvcache          layers/attention.py:10-30, 33-54
 * key,value [n_tokens, kv_dim] with row strides; caches [num_slots, kv_dim] contiguous
 * (the reference asserts k_cache.stride(1) == D, :52); slot_mapping[i] == -1 skips row i. */
B2TTS_API int b2tts_store_kvcache(const void* key, const void* value,
                        int64_t key_row_stride, int64_t value_row_stride,
                        void* k_cache, void* v_cache, const int32_t* slot_mapping,
                        int n_tokens, int kv_dim, void* stream);

/* ---- a2: paged GQA decode attention ---------------------------------------------------
 * replaces flash_attn_with_kvcache(q.unsqueeze(1), k_cache, v_cache, cache_seqlens,
 *          block_table, softmax_scale, causal=True)        layers/attention.py:92-95
 * q [batch, num_heads, head_dim] (row stride q_row_stride), caches
 * [num_pages, page_size, num_kv_heads, head_dim], block_table [batch, bt_stride] int32
 * (entries beyond ceil(len/page) are never read, -1 allowed), context_lens [batch] int32
 * INCLUDING the token written this step.  out [batch, num_heads, head_dim] bf16.
 * max_ctx bounds the split-KV grid (capacity of the block table in tokens). */
B2TTS_API size_t b2tts_paged_attn_decode_workspace_bytes(int batch, int num_heads, int head_dim, int max_ctx);
B2TTS_API int b2tts_paged_attn_decode(const void* q, int64_t q_row_stride,
                            const void* k_cache, const void* v_cache,
                            const int32_t* block_table, int bt_stride,
                            const int32_t* context_lens,
                            void* out, void* workspace, size_t workspace_bytes,
                            int batch, int num_heads, int num_kv_heads, int head_dim,
                            int page_size, float scale, int max_ctx, void* stream);

/* The same operation with the split-KV partial sums in a file-scope device array inside the library instead of a
 * caller-provided workspace: the call allocates nothing (no cudaMalloc / cudaFree, no allocating library routine), for
 * callers that cannot allocate - code under CUDA-graph capture, tight serving loops, harnesses that forbid it.
 * NO COUNTERPART IN THE REFERENCE.
 * Limit: the array holds b2tts_paged_attn_decode_static_max_bytes() = 8 MiB, held by every GPU that loads the library until
 * the process exits; shapes whose b2tts_paged_attn_decode_workspace_bytes(...) exceeds it are rejected with
 * B2TTS_EWORKSPACE (use b2tts_paged_attn_decode with your own workspace).  All calls share the array, which keeps the last
 * call's contents (nothing in it needs clearing: every partial is written before it is read): calls that can overlap on
 * the GPU - different streams, parallel graph branches - race on it.  Arguments and results as b2tts_paged_attn_decode. */
B2TTS_API size_t b2tts_paged_attn_decode_static_max_bytes(void);
B2TTS_API int b2tts_paged_attn_decode_static(const void* q, int64_t q_row_stride,
                            const void* k_cache, const void* v_cache,
                            const int32_t* block_table, int bt_stride,
                            const int32_t* context_lens, void* out,
                            int batch, int num_heads, int num_kv_heads, int head_dim,
                            int page_size, float scale, int max_ctx, void* stream);

/* ---- a3: varlen causal prefill attention ----------------------------------------------
 * replaces flash_attn_varlen_func(q, k, v, cu_seqlens_q/k, max_seqlen_q/k, softmax_scale,
 *          causal=True, block_table)                       layers/attention.py:78-91
 * q [total_q, num_heads, head_dim], k,v [total_k, num_kv_heads, head_dim] packed, row
 * strides in elements; cu_seqlens_* [num_seqs+1] int32.  Causal mask is bottom-right
 * aligned (query i of a sequence sees keys 0 .. len_k-len_q+i).  When block_table is
 * non-NULL (prefix-cache hit, attention.py:79-80) k,v are the paged caches
 * [num_pages, page_size, kv_heads, head_dim] and are read through block_table.
 * total_k = rows of the packed k/v (ignored with a block table).  max_seqlen_q >= 128 (packed k/v) runs the
 * tcgen05 tile kernel: 128-query x 128-key tiles, QK^T and PV on the tensor cores with TMEM accumulators, K/V by
 * TMA, causal tile skipping; shorter prompts run one CTA per query token. */
B2TTS_API int b2tts_varlen_attn_prefill(const void* q, int64_t q_row_stride,
                              const void* k, int64_t k_row_stride,
                              const void* v, int64_t v_row_stride,
                              const int32_t* cu_seqlens_q, const int32_t* cu_seqlens_k,
                              void* out, int num_seqs, int total_q, int total_k,
                              int max_seqlen_q, int max_seqlen_k,
                              int num_heads, int num_kv_heads, int head_dim, float scale,
                              const int32_t* block_table, int bt_stride, int page_size,
                              void* stream);

/* ---- a4: RMSNorm ----------------------------------------------------------------------
 * replaces Qwen3TTSRMSNorm.rms_forward / add_rms_forward    layers/layernorm.py:18-27, 30-41
 * x [rows, cols] (row stride x_row_stride), w [cols]; y [rows, cols] contiguous.
 * add variant: sum = f32(x)+f32(res); res_out = bf16(sum); norm uses the fp32 sum
 * (layernorm.py:35-38) or, with round_sum_first != 0, the bf16-rounded sum - which is
 * what the decoder layer's `residual + hidden` followed by rms_forward computes
 * (models/qwen3_tts_share.py:171-173).  y/res_out may alias x/res. */
B2TTS_API int b2tts_rmsnorm(const void* x, int64_t x_row_stride, const void* w, void* y,
                  int rows, int cols, float eps, int rounding, void* stream);
B2TTS_API int b2tts_add_rmsnorm(const void* x, const void* res, const void* w, void* y, void* res_out,
                      int rows, int cols, float eps, int rounding, int round_sum_first,
                      void* stream);

/* ---- a4+a5+a1 fused: per-head q/k RMSNorm + RoPE (+ paged KV write) ---------------------
 * replaces q_norm, k_norm, rotary_emb and store_kvcache as called from
 * Qwen3TTSAttention.forward                       models/qwen3_tts_share.py:96-106
 *   RotaryEmbedding.forward / TalkerRotaryEmbedding.forward / apply_*_rotary_pos_emb
 *                                                 layers/rotary_embedding.py:11-212
 * q [n, heads, 128], k,v [n, kv_heads, 128] with row strides (they may be slices of one
 * fused QKV GEMM output).  q_out [n, heads*128] and k_out [n, kv_heads*128] contiguous
 * (may alias q / k when the strides match).  q_weight/k_weight NULL skips the norm
 * (plain RoPE, the rotary_emb.forward surface).  positions int64 [3 or 1][n]
 * (pos_row_stride = 0 for 1-D); inv_freq fp32 [head_dim/2]; modality int32 [head_dim]
 * selects the position row per cos/sin column (NULL = row 0; columns f and head_dim/2+f share
 * frequency f but may read different rows, rotary_embedding.py:73-80).  cos/sin are computed in fp32,
 * scaled by attention_scaling, rounded to bf16, and the rotation is three bf16-rounded
 * ops exactly as the reference's eager expression (rotary_embedding.py:33-34).
 * slot_mapping NULL skips the cache write; otherwise K (normed+rotated) and V (raw) rows
 * go to k_cache/v_cache[slot] and slot -1 is skipped. */
B2TTS_API int b2tts_qknorm_rope_kvwrite(const void* q, int64_t q_row_stride,
                              const void* k, int64_t k_row_stride,
                              const void* v, int64_t v_row_stride,
                              void* q_out, void* k_out,
                              const void* q_weight, const void* k_weight, float eps,
                              const int64_t* positions, int64_t pos_row_stride,
                              const float* inv_freq, const int32_t* modality,
                              float attention_scaling,
                              const int32_t* slot_mapping, void* k_cache, void* v_cache,
                              int n_tokens, int num_heads, int num_kv_heads, int head_dim,
                              int rounding, void* stream);

/* ---- a6: SiLU and the gated product ------------------------------------------------------
 * replaces Silu.forward (layers/activation.py:12-13) and `act_fn(gate) * up`
 * (models/qwen3_tts_share.py:131-135).  gate, up [rows, inter] with a common row stride
 * (two halves of a fused gate_up GEMM output); up == NULL computes silu(gate) only. */
B2TTS_API int b2tts_silu_and_mul(const void* gate, const void* up, int64_t in_row_stride,
                       void* out, int rows, int inter, int rounding, void* stream);

/* ---- a8: sampler ---------------------------------------------------------------------------
 * replaces Sampler.forward                                  layers/sampler.py:15-24
 * logits [batch, vocab] bf16 (or fp32 when logits_fp32 != 0), temperatures [batch] fp32,
 * noise [batch, vocab] fp32 = the Exp(1) draws torch.multinomial makes
 * (`empty_like(probs).exponential_(1)`), out_index [batch] int64 = argmax(softmax(topk-masked
 * logits / T) / noise).  Values tied with the k-th largest are kept (`logits < kth`).
 * top_p is accepted and, like the reference, ignored unless apply_top_p != 0.
 * probs_out (optional, fp32 [batch, vocab]) receives the softmax for inspection. */
B2TTS_API int b2tts_topk_sample(const void* logits, int64_t logits_row_stride, int logits_fp32,
                      const float* temperatures, const float* noise,
                      int64_t* out_index, float* probs_out,
                      int batch, int vocab, int top_k, float top_p, int apply_top_p,
                      void* stream);
/* The same sampler followed, in the same launch, by the gather of the next predictor pass's input embedding:
 * embed_out[i, :] = embed_table[out_index[i], :]   (bf16 rows of embed_cols)
 * replaces Sampler.forward + codec_embedding[g-1](id)   layers/sampler.py:15-24, models/qwen3_tts_predictor.py:94-95
 * vocab <= 4096 and a multiple of 8. */
B2TTS_API int b2tts_topk_sample_embed(const void* logits, int64_t logits_row_stride, int logits_fp32,
                            const float* temperatures, const float* noise,
                            int64_t* out_index, int batch, int vocab, int top_k,
                            const void* embed_table, int embed_cols, void* embed_out,
                            int64_t embed_out_row_stride, void* stream);

/* ---- a9 / frame-loop gathers ------------------------------------------------------------------
 * b2tts_gather_rows: out[i] = tables[table_index[i]][row_index[i]]  (bf16 rows of `cols`)
 *   replaces the per-sequence loop codec_embedding[g_i-1](id_i)   models/qwen3_tts_predictor.py:94-98
 *   tables = base pointer of a stacked [num_tables, rows_per_table, cols] tensor;
 *   table_index NULL = table 0; an index < 0 writes a zero row.
 * b2tts_codec_embed_sum: out[i] = bf16(sum_{c<ncb} f32(E_c[codes[i,c]])) + extra[extra_index ? extra_index[i] : i]
 *   replaces the 16-codebook gather-sum + trailing-text add        interface.py:1276-1286
 *   (extra_index selects trailing_text_hiddens[:, generation_step] or the tts_pad_embed row per stream) */
B2TTS_API int b2tts_gather_rows(const void* tables, int64_t table_stride, int rows_per_table,
                      const int64_t* row_index, const int32_t* table_index,
                      void* out, int64_t out_row_stride, int n, int cols, void* stream);
B2TTS_API int b2tts_codec_embed_sum(const void* table0, const void* tables_rest, int64_t table_stride,
                          const int64_t* codes, int64_t codes_row_stride,
                          const void* extra, int64_t extra_row_stride, const int64_t* extra_index,
                          void* out, int n, int num_codebooks, int cols, void* stream);

/* ---- a12 on the device: decode batch layout -------------------------------------------------------
 * replaces the host loops of prepare_decode / prepare_decode_talker
 *                          engine/model_runner/base.py:218-234, talker_mode_runner.py:116-140
 * seq_lens [n] int32 (device) = tokens of each stream INCLUDING the one decoded this step, 0 =
 * inactive row.  Writes context_lens = len, slot = block_table[i][(len-1)/page]*page + (len-1)%page,
 * positions = len-1+position_offset (1 = talker quirk, 0 = predictor); inactive rows - and rows whose
 * token would fall beyond the bt_width pages of their table row - get slot -1, context 0.
 * advance != 0 then adds `advance` to the active seq_lens (next step's length). */
B2TTS_API int b2tts_decode_layout(int32_t* seq_lens, const int32_t* block_table, int bt_stride, int bt_width,
                        int page_size, int position_offset, int advance, int64_t* positions,
                        int32_t* slot_mapping, int32_t* context_lens, int n, void* stream);

/* ---- f1: per-frame stream state of the on-device frame loop ---------------------------------------
 * replaces the host bookkeeping of generate_async                      interface.py:1233-1239, 1282-1285
 * code0 [n] int64 = the talker code sampled for this frame.  A stream whose code0 == eos_id (2150; eos_id < 0
 * ignores EOS) or that has produced max_frames frames goes inactive BEFORE the frame is produced: active[i] = 0,
 * seq_lens[i] = 0 (a padding row for b2tts_decode_layout from now on).  extra_row[i] = i*(trail_rows+1) +
 * min(gen_step[i], trail_rows) indexes a [n, trail_rows+1, H] buffer whose row trail_rows holds tts_pad_embed
 * (b2tts_codec_embed_sum's extra_index).  Active streams then count the frame: gen_step += 1, frames_done += 1. */
B2TTS_API int b2tts_frame_advance(const int64_t* code0, int eos_id, int32_t* active, int32_t* seq_lens,
                        int32_t* gen_step, int32_t* frames_done, int64_t* extra_row, int trail_rows,
                        int max_frames, int n, void* stream);

/* ---- a7: the projections (weight-streaming tcgen05 GEMM) -------------------------------------------
 * replaces F.linear in ColumnParallelLinear / RowParallelLinear.forward   layers/linear.py:72-73, 149-153
 *   (q/k/v and gate/up as ONE GEMM each over a merged weight: models/qwen3_tts_share.py:85-87, 126-134),
 *   nn.Linear small_to_mtp_projection / lm_head[g]                 models/qwen3_tts_predictor.py:54-61,
 *   ParallelLMHead's F.linear                                      layers/embed_head.py:61
 * y[m, n] = x[m, k] @ W[n, k]^T, bf16 in/out, fp32 accumulate in TMEM, m >= 1 (rows beyond 128 run as
 * 128-row chunks inside the same launch), n % 128 == 0, k % 64 == 0.  W is passed PRE-PACKED: b2tts_gemm_pack_weight() re-tiles a row-major [n, k] weight
 * once at load time into [n/128][k/64] tiles of 16 KiB (the K-major SWIZZLE_128B shared-memory image
 * of a UMMA operand), b2tts_gemm_packed_bytes(n, k) bytes.  B2TTS_PACK_GATE_UP packs a merged
 * [gate; up] weight (n = 2 * intermediate) so that every 8 consecutive tile rows hold 4 gate features
 * and the 4 matching up features - required by the SiLU*mul epilogue.
 * epilogue: B2TTS_GEMM_BIAS      out[m, n]   = bf16(acc + bias)              (bias may be NULL)
 *           B2TTS_GEMM_RESIDUAL  out[m, n]   = bf16(bf16(acc) + residual)    (qwen3_tts_share.py:171,175;
 *                                              out may alias residual)
 *           B2TTS_GEMM_SILU_MUL  out[m, n/2] = silu(bf16(gate)) * bf16(up)   (qwen3_tts_share.py:131-135;
 *                                              `rounding` as b2tts_silu_and_mul)
 * splits: K-split = cluster size (1,2,4,8,16; 0 = choose so that tiles * splits fill the SMs); the
 * partial sums are exchanged through distributed shared memory and added in a fixed order.
 * splits = -2 / -4: the cluster splits the BATCH instead (m padded to 16 must be a multiple of 32 / 64, m <= 128):
 * every CTA runs the whole K range for its 16 - 48 rows and each weight tile is fetched once per cluster (1 / |splits|
 * per CTA, multicast to all) - no partial sums, no exchange.  splits = 0 never picks it unless B2TTS_GEMM_BSPLIT=1
 * (then for K <= 1024): in the decode frame it is slower than the K split.
 * pdl != 0 launches with programmatic stream serialization: the weight stream starts while the
 * previous kernel on the stream is still running (every kernel of this library that is launched
 * this way executes griddepcontrol.wait before touching its inputs/outputs). */
#define B2TTS_GEMM_BIAS     0
#define B2TTS_GEMM_RESIDUAL 1
#define B2TTS_GEMM_SILU_MUL 2
#define B2TTS_PACK_PLAIN    0
#define B2TTS_PACK_GATE_UP  1
B2TTS_API int    b2tts_gemm_bf16_supported(int m, int n, int k);
B2TTS_API size_t b2tts_gemm_packed_bytes(int n, int k);
B2TTS_API int    b2tts_gemm_pack_weight(const void* weight, int64_t weight_row_stride, void* packed,
                                        int n, int k, int mode, void* stream);
B2TTS_API int    b2tts_gemm_bf16(const void* x, int64_t x_row_stride, const void* w_packed,
                                 const void* bias, const void* residual, int64_t residual_row_stride,
                                 void* out, int64_t out_row_stride, int m, int n, int k,
                                 int epilogue, int rounding, int splits, int pdl, void* stream);


/* ---- a4 fused into a7: input RMSNorm in the GEMM's x producers, row sums of squares out of its epilogue ----
 * replaces `hidden = layernorm(residual)` followed by the projection that reads it
 *                                          models/qwen3_tts_share.py:166-167,172-174, layers/layernorm.py:18-27
 * As b2tts_gemm_bf16, plus:
 *   norm_weight != NULL: x is the residual stream; the GEMM multiplies norm(x)[m, k] =
 *     rounding(x[m, k] * rstd[m], norm_weight[k]) (same rounding points as b2tts_rmsnorm) with
 *     rstd[m] = 1/sqrt(sum_j ssq_in[m, j] / k + eps), j < ssq_in_slots summed in a fixed order.
 *   ssq_out != NULL (bias / residual epilogues): ssq_out[m, slot] = sum of squares of the bf16 values this
 *     call stored in out[m, :] over slot's 128-feature tile; slots per row = b2tts_gemm_ssq_slots(m, n, k, splits)
 *     = n / 128 (the K-split CTAs of a tile combine their sums inside the cluster, in rank order).
 *     The next b2tts_gemm_norm_bf16 over `out` takes it as ssq_in: the decoder layer's two norm launches
 *     disappear, and the normalised activations never reach HBM. */
B2TTS_API int    b2tts_gemm_ssq_slots(int m, int n, int k, int splits);
B2TTS_API int    b2tts_gemm_norm_bf16(const void* x, int64_t x_row_stride, const void* w_packed,
                                      const void* bias, const void* residual, int64_t residual_row_stride,
                                      void* out, int64_t out_row_stride, int m, int n, int k,
                                      int epilogue, int rounding, int splits, int pdl,
                                      const void* norm_weight, const float* ssq_in, int ssq_in_slots, float eps,
                                      float* ssq_out, void* stream);

/* b2tts_gemm_norm_bf16 plus an L2 prefetch of the NEXT projection's packed weights (prefetch_next, prefetch_bytes =
 * b2tts_gemm_packed_bytes of that weight, or NULL / 0): weights do not depend on activations, and a decode step is a
 * chain of latency-bound launches during which HBM idles most of the time - every CTA pulls its slice of the next
 * weight image into the 126 MB L2 (cp.async.bulk.prefetch.L2) so the next launch streams from L2 while HBM already
 * fetches the launch after it.  Results are identical to b2tts_gemm_norm_bf16. */
B2TTS_API int    b2tts_gemm_prefetch_bf16(const void* x, int64_t x_row_stride, const void* w_packed,
                                          const void* bias, const void* residual, int64_t residual_row_stride,
                                          void* out, int64_t out_row_stride, int m, int n, int k,
                                          int epilogue, int rounding, int splits, int pdl,
                                          const void* norm_weight, const float* ssq_in, int ssq_in_slots, float eps,
                                          float* ssq_out, const void* prefetch_next, size_t prefetch_bytes, void* stream);

/* ---- a7 x 4 in one launch: the persistent GEMM chain of the decoder layer ------------------------------------
 * replaces the four dependent projections between two attention calls
 *     o_proj (+ residual add) -> gate_up_proj (post_attention_layernorm in, SiLU*mul out) -> down_proj (+ residual add)
 *     -> the next layer's qkv_proj (input_layernorm in)                      models/qwen3_tts_share.py:163-176
 * with ONE persistent kernel (one CTA per SM) whose weight stream never stops at a GEMM boundary: a producer
 * thread keeps a deep shared-memory ring of packed weight tiles full across the whole chain while the CTAs
 * reduce split-K partials through L2 and cross a grid barrier between GEMMs.  Every op is a b2tts_gemm_norm_bf16
 * (same packed weights, epilogues, fused input norm and rounding points); op i+1 may read what op i wrote
 * (x, residual, ssq_in).  ssq_out of an op has b2tts_gemm_chain_ssq_slots(n) = n/128 slots per row.
 * 1 <= nops <= 4, 1 <= m <= 128.  workspace: b2tts_gemm_chain_workspace_bytes(...) bytes, ZERO-FILLED once by the
 * caller (the kernel leaves its barrier counters zero); one workspace must not be used by two launches that
 * can run concurrently.  The grid needs every SM of the device: do not run two chains concurrently. */
typedef struct b2tts_gemm_op {
  const void* x;            int64_t x_row_stride;         /* [m, k] bf16 */
  const void* w_packed;                                   /* b2tts_gemm_pack_weight image of W[n, k] */
  const void* residual;     int64_t residual_row_stride;  /* B2TTS_GEMM_RESIDUAL */
  void* out;                int64_t out_row_stride;       /* [m, n] (or [m, n/2] for SiLU*mul); may alias residual */
  const void* norm_weight;  const float* ssq_in;          /* fused input RMSNorm (or NULL) */
  float* ssq_out;                                         /* [m, n/128] partial row sums of squares (or NULL) */
  int n, k, epilogue, ssq_in_slots;
} b2tts_gemm_op;
B2TTS_API int    b2tts_gemm_chain_ssq_slots(int n);
B2TTS_API size_t b2tts_gemm_chain_workspace_bytes(int m, const int* n, const int* k, int nops);
B2TTS_API int    b2tts_gemm_chain(const b2tts_gemm_op* ops, int nops, int m, float eps, int rounding,
                                  void* workspace, size_t workspace_bytes, int pdl, void* stream);

/* ---- a5 + a4 + a1 + a2 in one launch: the decode step of Qwen3TTSAttention.forward ------------------------
 * replaces q_norm, k_norm, rotary_emb, store_kvcache and flash_attn_with_kvcache as called per decode token
 *                                       models/qwen3_tts_share.py:96-109, layers/attention.py:33-54, 92-95
 * q/k/v are the rows of the QKV projection (one new token per sequence).  The kernel normalises + rotates q
 * and k (arguments as b2tts_qknorm_rope_kvwrite), writes K (normed + rotated) and V (raw) of the new token
 * to cache[slot_mapping[i]], attends over context_lens[i] cached tokens INCLUDING the new one
 * (arguments as b2tts_paged_attn_decode) and folds the split-KV partials itself (the last CTA of a
 * sequence to finish does it, in split order).  K cache bytes and attention outputs equal those of the
 * separate calls.  workspace: b2tts_fused_decode_attn_workspace_bytes(...) bytes, ZERO-FILLED once by the
 * caller before the first launch (the kernel leaves its tickets zero).  The tickets occupy a fixed span at the
 * start of the workspace, so ONE workspace sized for the largest batch serves launches of every smaller batch
 * (batch <= 4096).  pdl as b2tts_gemm_bf16. */
B2TTS_API size_t b2tts_fused_decode_attn_workspace_bytes(int batch, int num_heads, int head_dim, int max_ctx);
B2TTS_API int b2tts_fused_decode_attn(const void* q, int64_t q_row_stride, const void* k, int64_t k_row_stride,
                            const void* v, int64_t v_row_stride,
                            const void* q_weight, const void* k_weight, float eps,
                            const int64_t* positions, int64_t pos_row_stride,
                            const float* inv_freq, const int32_t* modality, float attention_scaling,
                            const int32_t* slot_mapping, void* k_cache, void* v_cache,
                            const int32_t* block_table, int bt_stride, const int32_t* context_lens,
                            void* out, void* workspace, size_t workspace_bytes,
                            int batch, int num_heads, int num_kv_heads, int head_dim, int page_size,
                            float scale, int max_ctx, int rounding, int pdl, void* stream);

/* ---- f3: prompt-prep embedding assembly -------------------------------------------------------------------
 * replaces the per-request tensor ops of prepare_inputs / generate_icl_prompt   utils/generation.py:18-379, 412-515
 * (torch.cat / expand / `+` of projected text rows, codec-tag rows, the speaker row and the ICL reference frames,
 * then pad_sequence into the left-padded batch, :329-377).  The host plans every row of every request of a batch;
 * one launch writes  out[r] = a[a_index[r]] (+) b[b_index[r]]  (bf16 rows of `cols`; the add is bf16(f32 + f32),
 * torch's bf16 tensor add; an index < 0 = that side absent, both absent = a zero padding row).
 * a = the projected text rows text_projection(text_embedding(ids)), b = codec-side rows; indices are NOT range-checked
 * on the device (a_rows / b_rows are validated by the host wrapper). */
B2TTS_API int b2tts_prompt_assemble(const void* a, int64_t a_row_stride, int a_rows, const void* b, int64_t b_row_stride,
                                    int b_rows, const int32_t* a_index, const int32_t* b_index, void* out,
                                    int64_t out_row_stride, int rows, int cols, void* stream);

/* ---- f4 (tail): decoded waveform -> PCM16 ----------------------------------------------------------------------
 * replaces _float_to_pcm16 / _apply_volume_pcm16                                    examples/server.py:338-349
 *   out[i] = (int16) trunc(clip(wav[i], -1, 1) * 32767);  volume != 1: out[i] = clip(rint(out[i] * volume), -32768, 32767)
 * (the vocoder itself - qwen_tts 0.0.5, not part of the reference tree - is out of reach). */
B2TTS_API int b2tts_float_to_pcm16(const float* wav, int16_t* out, int64_t n, float volume, void* stream);

/* ---- a10 in one launch: the decode pass of a whole (small) decoder stack ----------------------------------------
 * replaces, for one decode step of up to 64 sequences, the layer loop of Qwen3TTSCodePredictorModel.forward /
 * Qwen3TTSTalkerModel.forward                          models/qwen3_tts_predictor.py:30-44, qwen3_tts_talker.py:53-62
 * over Qwen3TTSDecoderLayer.forward                    models/qwen3_tts_share.py:163-176 (and :79-113, :131-135)
 * i.e. per layer: input RMSNorm, QKV projection, q/k head norm + RoPE + paged KV write + paged GQA attention,
 * o projection + residual add, post-attention RMSNorm, gate_up projection + SiLU*mul, down projection + residual
 * add; then the final RMSNorm.  ONE persistent dataflow kernel (one CTA per SM): every CTA owns one (projection,
 * 128-feature tile, K split) job in every layer and prefetches its packed weights ahead of the activations; phases
 * are linked by arrival counters in the workspace instead of kernel boundaries (csrc/decoder_stack.cu).  Same packed
 * weights (b2tts_gemm_pack_weight; gate_up in B2TTS_PACK_GATE_UP mode), rounding points and summation orders as
 * b2tts_rmsnorm -> b2tts_gemm_bf16 -> b2tts_fused_decode_attn -> b2tts_gemm_bf16(RESIDUAL, splits = o_splits) ->
 * b2tts_rmsnorm -> b2tts_gemm_bf16(SILU_MUL) -> b2tts_gemm_bf16(RESIDUAL, splits = down_splits): bit-identical.
 * x [m, hidden] is the residual stream entering layer 0 (left intact), out [m, hidden] the normalised final hidden
 * states.  positions [m] int64 (1-D RoPE), slot_mapping / block_table / context_lens as b2tts_fused_decode_attn.
 * 1 <= m <= 64, 1 <= num_layers <= 8, head_dim 128, num_heads == 2 * num_kv_heads, hidden <= 2048, and the jobs must
 * fit one per SM (b2tts_decoder_stack_plan returns 0 otherwise).  workspace: b2tts_decoder_stack_workspace_bytes(...)
 * bytes, 256-byte aligned, ZERO-FILLED once by the caller; one workspace must not be used by two launches that can
 * run concurrently, and the grid needs every SM of the device.  A dependency that fails to resolve within 4 s traps
 * (CUDA launch failure) rather than hanging the device. */
typedef struct b2tts_stack_layer {
  const void *qkv_w, *o_w, *gate_up_w, *down_w;                      /* packed weights */
  const void *input_norm_w, *post_norm_w, *q_norm_w, *k_norm_w;       /* bf16 norm weights [hidden] / [128] */
  void *k_cache, *v_cache;                                            /* [pages, page, kv_heads, 128] bf16 */
} b2tts_stack_layer;
B2TTS_API int    b2tts_decoder_stack_plan(int hidden, int q_size, int kv_size, int intermediate,
                                          int* o_splits, int* down_splits);
B2TTS_API size_t b2tts_decoder_stack_workspace_bytes(int hidden, int q_size, int kv_size, int intermediate);
B2TTS_API int    b2tts_decoder_stack_decode(const void* x, int64_t x_row_stride, void* out, int64_t out_row_stride,
                                            const b2tts_stack_layer* layers, int num_layers, const void* final_norm_weight,
                                            int m, int hidden, int num_heads, int num_kv_heads, int head_dim,
                                            int intermediate, float eps, int rounding,
                                            const int64_t* positions, const float* inv_freq, float attention_scaling,
                                            const int32_t* slot_mapping, const int32_t* block_table, int bt_stride,
                                            const int32_t* context_lens, int page_size, float scale,
                                            void* workspace, size_t workspace_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200TTS_H */

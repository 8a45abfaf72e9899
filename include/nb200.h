/* nb200.h — C ABI of libnemob200.so: the B200 (sm_100a) backend for nemotron-asr.cpp's streaming hot path
 *   s16le PCM -> log-mel -> cache-aware FastConformer encoder -> RNNT LSTM/joint -> greedy tokens.
 *
 * This is the drop-in boundary. Every entry point names the reference interface it replaces
 * (paths relative to the reference repo m1el/nemotron-asr.cpp @ 4a863ff). The reference API is C++ over public
 * structs; this ABI is plain C (opaque handles, pointers and sizes, int status codes) so it can be bound from the
 * reference's own sources, from ctypes (nemotron-asr.cpp_b200/nb200.py) or any FFI.
 *
 * Threading contract = the reference's (src/nemo-server.cpp:6-10): one thread owns a model and its streams.
 * Errors: functions return NULL / a negative status and nb200_last_error() holds a message (the reference
 * returns nullptr/false/"" and prints to stderr). There is NO CPU fallback: without a CUDA device or with a
 * failed kernel every call fails loudly.
 */
#ifndef NB200_H
#define NB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NB200_OK 0
#define NB200_ERR (-1)

typedef struct nb200_model nb200_model;   /* replaces struct nemo_context   (src/nemo-ggml.h:236-252) */
typedef struct nb200_stream nb200_stream; /* replaces struct nemo_stream_context (src/nemo-stream.h:177-262) */

/* model-load flags */
#define NB200_FLAG_ACT_QUANT_AUTO 0 /* emulate ggml-CPU's Q8_0 activation re-quantisation iff the file has Q8_0 weights */
#define NB200_FLAG_ACT_QUANT_OFF 1  /* never (weights are still the dequantised Q8_0 values)                       */
#define NB200_FLAG_ACT_QUANT_ON 2   /* always                                                                      */
#define NB200_FLAG_GEMM_SIMT 16     /* debug: CUDA-core GEMM instead of the tcgen05 kernel (same operands/epilogues) */
#define NB200_FLAG_NO_GRAPHS 32     /* debug: launch every kernel eagerly instead of replaying per-shape CUDA graphs */
#define NB200_FLAG_BF16 64          /* bf16 instead of fp16 operands in the 24 layers' GEMMs (weights converted at load, activations stored as bf16
                                     * by the producing kernels; SURVEY section 8(d) "F32 -> bf16 at load"). For F32 / F16 files: not together with
                                     * the Q8_0 activation re-quantisation. */

typedef struct nb200_info {
    int32_t n_mels, d_model, n_heads, d_head, d_ff, n_layers, kernel_size;
    int32_t vocab_size, blank_token, decoder_dim, joint_dim, subsampling_factor, att_left_context, num_prompts;
    int32_t weights_quantized; /* 1 if the GGUF holds Q8_0 encoder matrices */
    int32_t act_quant;         /* 1 if activation re-quantisation is emulated */
    int32_t max_streams;       /* stream slots allocated in HBM */
    int32_t device;
} nb200_info;

typedef struct nb200_stats {
    int64_t steps;            /* batched steps run                                   */
    int64_t chunks;           /* stream-chunks processed (sum of batch sizes)        */
    int64_t frames;           /* encoder frames decoded                              */
    int64_t tokens;           /* non-blank tokens emitted                            */
    int64_t decode_rounds;    /* batched joint evaluations                           */
    int64_t kernel_launches;  /* kernels of this library launched so far             */
    int64_t h2d_bytes;        /* bytes copied host->device (audio + per-step metadata)          */
    int64_t d2h_bytes;        /* bytes copied device->host (tokens, counts, the decode loop's flag) */
    double audio_seconds;     /* audio consumed by processed chunks                  */
    double gpu_seconds;       /* device time of all steps (CUDA events on the library's stream) */
    double last_step_gpu_ms;  /* device time of the last nb200_step                  */
    double last_frontend_ms, last_encoder_ms, last_decoder_ms; /* split of the last step (events) */
} nb200_stats;

const char* nb200_last_error(void);
int nb200_version(void);

/* nemo_init_with_backend(path, backend) — src/nemo-ggml.h:259, src/nemo-ggml.cpp:448-495 (GGUF load :83-442).
 * device: CUDA ordinal. max_streams: stream slots to reserve (K/V/conv caches live in HBM per slot). */
nb200_model* nb200_model_load(const char* gguf_path, int device, int max_streams, int flags);
/* nb200_model_load through the NATIVE WEIGHT FILE (the reference's on-disk formats are chosen by its converter,
 * scripts/convert_to_gguf.py:322-336, and ggml dequantises on the fly): the big operands in the device-ready form this backend
 * lays them out in (fp16 [N][K] with q|k|v stacked, GLU rows interleaved, pre_encode.out columns permuted, the split-precision
 * triples and hi/lo pairs of joint / LSTM). cache_path NULL = "<gguf_path>.nb200w". If the file exists and was made from this
 * GGUF (size + mtime stamp, layout version) the operands are copied straight from its mmap; otherwise the model is built from the
 * GGUF as usual and the file is (re)written, atomically. The GGUF itself is always opened (hyper-parameters, vocabulary, the small
 * fp32 tensors). nb200_model_cache_stats: how many big operands came from the cache / were built from the GGUF. */
nb200_model* nb200_model_load_cached(const char* gguf_path, const char* cache_path, int device, int max_streams, int flags);
int nb200_model_cache_stats(const nb200_model* m, int64_t* from_cache, int64_t* from_gguf);
/* Host-only round trip of the native weight file format (no GPU needed): write, mmap-read, reject foreign stamps and truncated files.
 * Uses `path` (and path + ".trunc") as scratch files and removes them. NB200_OK or <0 with nb200_last_error(). */
int nb200_weight_cache_selftest(const char* path);
/* Host-only check of a GGUF file (no GPU needed): parses the container exactly like nb200_model_load — header, KVs, tensor
 * table with the bounds of every tensor inside the file, the vocabulary (tokenizer.vocab_list, or the legacy 8-byte-piece blob
 * tokenizer.vocab, src/nemo-ggml.cpp:149-169) — and writes a JSON summary {"tensors", "tensor_bytes", "vocab_source",
 * "vocab_pieces", "n_layers", "first_pieces"} into json_out. Returns its length, or <0 with nb200_last_error() on a malformed file. */
int nb200_gguf_check(const char* gguf_path, char* json_out, int cap);
/* nemo_free — src/nemo-ggml.h:261 */
void nb200_model_free(nb200_model* m);
int nb200_model_info(const nb200_model* m, nb200_info* out);
/* vocabulary piece of a token id (model.vocab, src/nemo-ggml.cpp:149-169); NULL if out of range */
const char* nb200_model_token(const nb200_model* m, int32_t id);
/* tokens_to_text — src/nemo-ggml.cpp:1556-1583 (U+2581 -> ' '); returns bytes written (excluding NUL) or <0 */
int nb200_tokens_to_text(const nb200_model* m, const int32_t* tokens, int n, char* out, int cap);
/* tokens_to_text(timed tokens, vocab, timestamp_words = true) — src/nemo-ggml.cpp:1556-1583: a piece that starts with
 * U+2581 becomes " {%.2f}" + rest, the time being timed_token::to_seconds (src/nemo-ggml.h:383-395): frame * 1280 /
 * 16000 in fp32. frames[i] is the encoder frame token i was emitted at (nb200_stream_poll_timed). */
int nb200_tokens_to_text_timed(const nb200_model* m, const int32_t* tokens, const int64_t* frames, int n, char* out, int cap);

/* nemo_stream_init(ctx, cfg) with cfg.att_right_context in {0,1,6,13} — src/nemo-stream.h:271, .cpp:696-733 */
nb200_stream* nb200_stream_create(nb200_model* m, int right_context);
/* nemo_set_language(ctx, lang) — src/nemo-ggml.h:266, .cpp:497-511: the model-wide language prompt (multilingual files):
 * used by the offline path (nb200_encode_audio) and inherited by streams created afterwards (nemo_stream_init copies
 * ctx->prompt_index). Unknown code or a monolingual model: error, nothing changes. */
int nb200_model_set_language(nb200_model* m, const char* lang);
/* nemo_stream_set_language — src/nemo-stream.h:279, src/nemo-stream.cpp:735-749 (multilingual checkpoints: the language
 * code is looked up in the GGUF's nemo.prompt_langs / nemo.prompt_ids; it may change between chunks). Fails like the
 * reference (which returns false) when the model has no prompts or the code is unknown. New streams start with the
 * loader's default prompt 101 = "auto" (src/nemo-ggml.cpp:459-462). _set_prompt_index sets the field the server writes
 * directly (sctx->prompt_index, src/nemo-server.cpp:244). */
int nb200_stream_set_language(nb200_stream* s, const char* lang);
int nb200_stream_set_prompt_index(nb200_stream* s, int prompt_index);
/* nemo_stream_free — src/nemo-stream.h:324 */
void nb200_stream_free(nb200_stream* s);
/* nemo_stream_reset — src/nemo-stream.h:319 (this backend also re-zeroes the conv cache; the reference does not) */
int nb200_stream_reset(nb200_stream* s);

/* First half of nemo_stream_process_incremental (src/nemo-stream.cpp:1145-1162): buffer PCM. The samples are
 * copied into a pinned staging arena; the next nb200_step moves everything fed since the last one with ONE
 * host->device copy and one scatter kernel into the streams' device-resident PCM rings (a cudaMemcpyAsync per stream
 * cost 64 x 5 us of host time per step). Enqueue only: no kernel runs here unless the arena fills up. */
int nb200_stream_feed(nb200_stream* s, const int16_t* pcm, int n_samples);
/* nemo_stream_finalize (src/nemo-stream.cpp:1217-1293): mark the stream so the next nb200_step flushes its
 * zero-padded tail chunk (keeps floor(new_mel/8) frames). */
int nb200_stream_finalize(nb200_stream* s);
/* chunks this stream could run right now (full chunks buffered, +1 if a finalize tail is pending) */
int nb200_stream_ready_chunks(const nb200_stream* s);

/* Second half of nemo_stream_process_incremental (:1174-1199) for ALL streams of the model at once: takes at
 * most one ready chunk from every stream, runs them as batched kernels (mel, conv subsampling, 24 conformer
 * layers, joint/LSTM greedy decode) and appends the emitted tokens to each stream. This is what gives
 * cross-stream batching; the reference runs one stream at a time (src/nemo-server.cpp:192-271).
 * Returns the number of stream-chunks processed (0 = nothing ready) or <0 on error. */
int nb200_step(nb200_model* m);
/* nb200_step in two halves, for a server loop that keeps receiving audio while the GPU works: _begin collects the
 * ready chunks and launches the batched step without waiting; _end waits for it and appends the tokens. Between the
 * two, nb200_stream_feed / nb200_feed_batch / polls are allowed (host-side staging only); create / free / reset
 * complete the running step first. nb200_step == _begin followed by _end. The reference has no counterpart: its worker
 * is one thread doing one stream at a time (src/nemo-server.cpp:192-271). Both return the stream-chunk count or <0. */
int nb200_step_begin(nb200_model* m);
int nb200_step_end(nb200_model* m);
/* nb200_stream_feed for `count` streams of one model in one call (the server's reader threads hand the worker a list
 * of (stream, bytes) events, src/nemo-server.cpp:201-262). */
int nb200_feed_batch(nb200_model* m, nb200_stream* const* streams, const int16_t* const* pcm, const int* n_samples, int count);
/* nb200_stream_poll_tokens for `count` streams: new tokens concatenated into tokens_out (capacity cap), per-stream
 * counts into counts_out[count]. Returns the total number of tokens written; when cap is too small nothing is
 * consumed, counts_out holds the waiting counts and the call returns <0. */
int nb200_poll_batch(nb200_model* m, nb200_stream* const* streams, int count, int32_t* tokens_out, int cap, int32_t* counts_out);
/* Offline path — nemo_encode_audio(ctx, pcm, n) -> std::vector<timed_token> (src/nemo-ggml.cpp:1690-1715, nemo_encode
 * :1592-1670): whole-utterance log-mel, subsampling, 24 layers with FULL-sequence relative-position attention, greedy decode
 * from a fresh decoder state. Writes min(n, cap) (token, encoder frame) pairs and returns the number of tokens n, or <0.
 * enc_out (may be NULL) receives the encoder output [T][1024] fp32 (parity tests), *n_frames_out = T. Utterances up to
 * min(512, 14 * max_streams) encoder frames (the reference's positional table ends at 512, :1061); needs one free stream
 * slot; synchronous; must not be called between nb200_step_begin and nb200_step_end of other streams' steps. */
int nb200_encode_audio(nb200_model* m, const int16_t* pcm, int n_samples, int32_t* tokens_out, int64_t* frames_out, int cap,
                       float* enc_out, int64_t enc_cap_floats, int* n_frames_out);
/* Convenience: nb200_step until no stream has a ready chunk. Returns total stream-chunks processed. */
int nb200_drain(nb200_model* m);

/* Tokens emitted since the last poll (sctx->tokens, src/nemo-stream.h:231). Returns count copied. */
int nb200_stream_poll_tokens(nb200_stream* s, int32_t* out, int cap);
/* Same, with the encoder frame (80 ms units, counted over the frames the stream has decoded so far) at which each
 * token was emitted: the timed_token{token_id, frame_idx} pairs of the reference's offline decoder
 * (greedy_decode, src/nemo-ggml.cpp:1233-1380; struct at src/nemo-ggml.h:383-395), which its streaming path drops
 * (tokens_to_text_simple fills frame 0, src/nemo-stream.cpp:755-762). frames_out may be NULL. */
int nb200_stream_poll_timed(nb200_stream* s, int32_t* tokens_out, int64_t* frames_out, int cap);
/* frame index of every token so far; returns total count, copies min(total, cap) */
int nb200_stream_get_token_frames(const nb200_stream* s, int64_t* out, int cap);
/* nemo_stream_get_tokens — all tokens so far; returns total count, copies min(total, cap) */
int nb200_stream_get_tokens(const nb200_stream* s, int32_t* out, int cap);
/* per-chunk token counts (for parity tests); returns number of chunks processed */
int nb200_stream_chunk_token_counts(const nb200_stream* s, int32_t* out, int cap);
/* per-chunk smallest top1-top2 logit gap over the chunk's symbol steps (near-tie tracing of token mismatches) */
int nb200_stream_chunk_margins(const nb200_stream* s, float* out, int cap);

/* Parity taps (the reference reads sctx->encoder_graph.encoder_out, src/nemo-stream.h:156). When enabled the
 * stream keeps the log-mel frames and every chunk's encoder output [T,1024] on the host. */
#define NB200_TAP_MEL 1
#define NB200_TAP_ENC 2
int nb200_stream_set_taps(nb200_stream* s, int taps);
/* returns number of floats available; copies min(available, cap) */
int64_t nb200_stream_fetch_tap(const nb200_stream* s, int tap, float* out, int64_t cap);

int nb200_get_stats(const nb200_model* m, nb200_stats* out);

/* ---- measurement (the reference has only wall-clock accumulators, src/nemo-stream.h:237-244) ------------- */
/* CUDA-event stopwatch on the library's own stream: start records an event, stop records a second one, waits for
 * it and returns the elapsed DEVICE time in ms (covers everything queued on that stream in between). */
int nb200_timer_start(nb200_model* m);
int nb200_timer_stop(nb200_model* m, double* ms_out);
/* Per-kernel-family device timing: while enabled every launch is bracketed by CUDA events on the launching stream.
 * report() writes a JSON object {"family": {"launches","ms","flops","bytes"}} (algorithmic FLOPs / bytes of the
 * launches) into json_out, clears the records and returns the string length, or <0 if cap is too small. */
int nb200_profile_enable(nb200_model* m, int on);
/* In-graph marginal cost of a kernel family: while set, every launch of that family ("gemm", "layernorm", "attention",
 * "dwconv") is issued TWICE inside the step's graph, so (step time with - step time without) / launches is what one launch of
 * the family costs where it actually runs (between its producer and consumer, PDL and graph replay included). The GEMM
 * epilogues are pure stores, so "gemm" leaves the results intact; the other families' results are garbage while set.
 * NULL or "" switches it off. Drops the model's captured graphs (they are re-captured on the next steps). */
int nb200_set_dup_family(nb200_model* m, const char* family);
int nb200_profile_report(nb200_model* m, char* json_out, int cap);

/* ---- kernel-level entry points (used by the parity tests and bench; same kernels the step uses) -------- */
/* nemo_preprocessor_process (src/preprocessor.cpp:330-395) for a batch of independent signals, each starting
 * from a fresh preprocessor: pcm [n_streams][n_samples] int16 (host) -> mel [n_streams][n_frames][128] (host),
 * n_frames = (256 + n_samples - 512)/160 + 1. Returns n_frames or <0. */
int nb200_mel_batch(nb200_model* m, const int16_t* pcm, int n_streams, int n_samples, float* mel_out);
/* Y[M,N] = X[M,K] * W^T with W one of the model's encoder matrices (by GGUF tensor name), using exactly the
 * operand preparation (activation re-quantisation, fp16 operands) and GEMM kernel of the step. Host buffers. */
int nb200_debug_matmul(nb200_model* m, const char* tensor_name, const float* x, int M, float* y);
/* The same, through the FUSED EPILOGUE the step uses for that matrix class and with the tile configuration the step would pick
 * for this M (what src/nemo-stream.cpp:593-603,631-690 does as separate ggml nodes): epi_kind 0 = fp32 store, 1 = SiLU (+ the
 * Q8_0 re-quantisation when the model emulates it) -> fp16 (linear1), 4 = GLU on pointwise_conv1 -> [M][N/2], 5 = split-K partial
 * planes added to a zero residual stream by the step's LayerNorm kernel (N = 1024 matrices: linear2, linear_out, pointwise_conv2):
 * y = the sum, y_ln (may be NULL) = the LayerNorm output (unit gain) as the next GEMM's fp16 operand, 6 = plain fp16 store
 * ("encoder.layers.<l>.self_attn.qkv" names the stacked q|k|v matrix). fp16 results are returned widened to fp32.
 * cfg_out (may be NULL) receives {tile width BN, K splits, CTAs per tile (2 = cta_group::2 pair), CTAs launched}. */
int nb200_debug_gemm(nb200_model* m, const char* tensor_name, const float* x, int M, int epi_kind, float* y, float* y_ln, int* cfg_out);
/* Times the tcgen05 GEMM alone: C[M,N] = A[M,K] * W[N,K]^T on synthetic fp16 operands, cycling through > 2x L2 worth of
 * weight matrices so that W streams from HBM as it does inside a step. epi_kind: 0 store f32, 1 SiLU + re-quantise ->
 * f16, 4 GLU, 5 split-K partials, 6 plain f16 store. bn / splits force a tile configuration (bn = 0: the library's heuristic). Device time
 * per launch (CUDA events on the library's stream) in microseconds. */
int nb200_bench_gemm(nb200_model* m, int M, int N, int K, int epi_kind, int bn, int splits, int iters, double* us_per_launch,
                     int* bn_used, int* splits_used);

/* Phase trace of `launches` back-to-back launches of that GEMM: 8 uint64 per CTA and launch (globaltimer ns):
 * [0] kernel entry, [1] prologue done, [2] previous kernel's results visible (griddepcontrol.wait returned), [3] first operand
 * stage landed, [4] accumulator complete, [5] epilogue stored. Returns the CTA count per launch or <0. */
int nb200_trace_gemm(nb200_model* m, int M, int N, int K, int epi_kind, int bn, int splits, int launches, uint64_t* out,
                     int cap_words);

/* Tensor-pipe peak of one tcgen05.mma kind on this device (measurement only: the int8 question of SURVEY section 8d; ggml's Q8_0
 * arithmetic is ggml/src/ggml-cpu/quants.c:451-478). Every CTA keeps a 128 x 256 operand pair resident in shared memory and
 * issues iters k-blocks of 128 bytes of K (4 MMAs) into one TMEM accumulator. kind 0 = kind::f16 (fp16 x fp16 -> fp32), 1 =
 * kind::i8 (s8 x s8 -> s32), 2 = kind::i8 with ggml's Q8_0 block scales: ONE K = 32 MMA per iteration into one of two TMEM
 * accumulators, read out by 16 epilogue warps and added as d_x[m][blk] * d_w[n][blk] * acc into fp32 running sums (synthetic scale
 * tables; check_out then holds fp32 sums over 8 blocks). *tops_out = 1e-12 * 2 * 128 * 256 * K_total * ctas / device time (TFLOP/s or TOP/s).
 * a_bytes [128][128] / b_bytes [256][128] (host, K-major rows of 128 bytes) are the operands; check_out (host, [128][256] 32-bit
 * words, may be NULL) receives CTA 0's accumulator of a SEPARATE one-k-block launch, for an arithmetic check. */
int nb200_bench_mma_peak(nb200_model* m, int kind, int ctas, int iters, const uint8_t* a_bytes, const uint8_t* b_bytes,
                         uint32_t* check_out, double* tops_out);

/* nb200_mel_batch without any allocation. No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it). Same
 * arguments and results as nb200_mel_batch (host pcm [n_streams][n_samples] in, host mel [n_streams][n_frames][128] out, returns
 * n_frames), same kernel; the device scratch (PCM rings, metadata, the mel output before its copy to the host) is a fixed
 * file-scope device array of 5.2 MiB that every GPU holds while the library is loaded. Limits: n_streams <=
 * NB200_NOALLOC_MAX_STREAMS and n_samples <= NB200_NOALLOC_MAX_SAMPLES, anything larger is rejected with an error (use
 * nb200_mel_batch). Calls no cudaMalloc / cudaFree and no allocating library routine. All calls share the scratch: calls on one
 * model are ordered by its stream; calls that overlap on the GPU (two models on one device, two threads) race. */
#define NB200_NOALLOC_MAX_STREAMS 16
#define NB200_NOALLOC_MAX_SAMPLES 65536
int nb200_mel_batch_noalloc(nb200_model* m, const int16_t* pcm, int n_streams, int n_samples, float* mel_out);

#ifdef __cplusplus
}
#endif
#endif /* NB200_H */

// kernels.cuh — launchers of the bandwidth-bound kernels (frontend.cu, encoder.cu, decoder.cu).
#pragma once
#include "common.cuh"

namespace nb {

// ---- frontend.cu ---------------------------------------------------------------------------------
struct MelTables {           // device pointers, built once per model from the GGUF filterbank/window
    const float* win512;     // hann-400 centred in 512 (src/preprocessor.cpp:297-299)
    const float* cos_tbl;    // host libm cosf/sinf of 2*pi*i/512 (src/preprocessor.cpp:86-89)
    const float* sin_tbl;
    const float* fbT;        // filterbank transposed [257][128]
    const int2* fb_range;    // per mel bin: [first, last+1) FFT bins with non-zero weight
};
// Audio lives in per-slot device rings (int16, ring_cap a power of two): absolute sample a of slot s is
// ring[s*ring_cap + (a & (ring_cap-1))]. Item b computes n_frames[b] frames; frame f covers absolute samples
// a0[b] + 160 f + [0, 513): element 0 is the sample BEFORE the frame's first sample (pre-emphasis needs x[i-1]).
// Samples with a < 0 (the 256-sample centre pad) or a >= n_avail[b] (finalize tail) read as zero.
// out: item b, frame f -> mel_out + out_off[b] + f*128 floats (out_off in floats). n_frames[b] <= max_frames.
struct MelBatch {
    const int16_t* ring;
    int ring_cap;
    const int* slot;            // [B]
    const int* n_frames;        // [B]
    const long long* a0;        // [B]
    const long long* n_avail;   // [B]
    const long long* out_off;   // [B]
};
void launch_mel(cudaStream_t st, const MelTables& tb, const MelBatch& mb, float* mel_out, int B, int max_frames);

struct SubWeights {          // encoder.pre_encode.* (src/nemo-ggml.cpp:296-307)
    const float* w0;         // [256][9] fp16-rounded values (conv0 runs through an F16 im2col in the reference)
    const float* b0;         // [256]
    const float* w2;         // [256][9] fp32 depthwise
    const float* b2;
    const float* w5;
    const float* b5;
};
// mel chunk [B][chunk_mel][128] -> relu(conv0) -> depthwise conv2 (+bias) -> fp16 [B][t2][33][256]
void launch_sub_stage1(cudaStream_t st, const SubWeights& w, const float* mel, __half* out, int B, int chunk_mel, int t1, int t2);
// relu(pw3) fp32 [B][t2][33][256] -> depthwise conv5 (+bias), rows h3 >= drop only -> fp16 [B][t3-drop][17][256]
void launch_sub_stage2(cudaStream_t st, const SubWeights& w, const float* a3, __half* out, int B, int t2, int t3, int drop = 2);

// ---- encoder.cu ----------------------------------------------------------------------------------
// Optional residual input: x += alpha * sum_{s<nsplit} partials[s*stride + ...] first (the split-K partial outputs of the
// GEMM that precedes this norm; src/nemo-stream.cpp:631-634,654,677,682-685 are the adds this replaces), written back to x.
struct ResidualIn {
    const float* partials = nullptr;
    int nsplit = 0;
    size_t stride = 0;
    float alpha = 1.0f;
};
// y = LN(x; g1,b1); if out_f32: out_f32 = y; z = g2 ? LN(y; g2,b2) : y; out_f16 = fp16(fq ? fake_quant(z) : z)
// h[m][n] = relu(h[m][n] + w_lang[lang[m / T]][n]) over [M][2048]: the one-hot half of prompt_kernel.0 (src/nemo-ggml.cpp:1097-1101)
void launch_prompt_add_relu(cudaStream_t st, float* h, const float* w_lang, const int* lang, int T, int M);
void launch_layernorm(cudaStream_t st, float* x, const float* g1, const float* b1, float* out_f32, const float* g2,
                      const float* b2, __half* out_f16, int act_quant, int M, const ResidualIn& rin = ResidualIn(), int out_bf16 = 0);

struct AttnArgs {
    const __half* qkv;     // [M][3072] fp16: q | k | v of this chunk's frames (row m = b*T + i), written by the QKV GEMM
    __half* k_ring;        // this layer's [slots][RING][1024] fp16: the last 70 key rows of every stream; this chunk's rows are appended
    __half* v_ring;
    const __half* pos;     // this layer's projected positions [npos][1024] fp16, row p + (T-1), p in [-(T-1), 70+T-1]
    const float* bias_u;   // [8][128]
    const float* bias_v;
    const int* slot;       // [B]
    const int* wpos;       // [B] ring index of this chunk's frame 0
    const int* valid;      // [B] cache_valid_len before this chunk
    __half* out;           // [M][1024] fp16 context (fq per 32 features when act_quant)
    int B, T, act_quant;
    int out_bf16 = 0;      // store the context as bf16 (NB200_FLAG_BF16 models)
};
void launch_attention(cudaStream_t st, const AttnArgs& a);

// Offline path: full-sequence relative-position attention over one utterance (build_rel_pos_mha, src/nemo-ggml.cpp:668-757):
// no cache, no mask; query i, key j use position i - j (the rel-shift, :625-662).
struct AttnFullArgs {
    const __half* qkv;     // [T][3072] fp16: q | k | v
    const __half* pos;     // this layer's projected positions [2T-1][1024] fp16, row p + (T-1), p = i - j in [-(T-1), T-1]
    const float* bias_u;   // [8][128]
    const float* bias_v;
    __half* out;           // [T][1024] fp16 context (fq per 32 features when act_quant)
    int T, act_quant;
    int out_bf16 = 0;
};
constexpr int ATTN_FULL_MAX_T = 512;  // the reference's positional table covers 512 frames (src/nemo-ggml.cpp:233,1061)
void launch_attention_full(cudaStream_t st, const AttnFullArgs& a);

struct ConvArgs {
    const float* g;        // GLU output [M][1024]
    float* conv_cache;     // this layer's [slots][2][8][1024]: two buffers per stream, read parity[b], write the other
    const int* parity;     // [B] chunks processed so far (its low bit selects the buffer to read)
    const float* w_dw;     // [9][1024]
    const float* ln_g;     // conv.batch_norm (a LayerNorm)
    const float* ln_b;
    const int* slot;
    __half* out;           // [M][1024] fp16 = fq(silu(LN(dwconv)))
    int B, T, act_quant;
    int out_bf16 = 0;
};
void launch_dwconv(cudaStream_t st, const ConvArgs& a);  // out = fq(silu(LN(dwconv([cache ; g])))), other cache buffer = last 8 of [cache ; g]

// fp32 [rows][1024] -> fp16 with optional fake-quant per 32 (used to prepare the positional table GEMM input)
// out[m] = (fp16(x), fp16(x - fp16(x)), fp16(x)) over [M][3 * cols]: A operand of the split-precision GEMMs (joint.enc, prompt kernel)
void launch_split_f16x3(cudaStream_t st, const float* x, __half* out, int M, int cols = D_MODEL);
void launch_cast_f16(cudaStream_t st, const float* x, __half* out, int act_quant, int rows, int cols, int out_bf16 = 0);
// sinusoid rows for relative positions p0 .. p0+n-1 (src/nemo-ggml.cpp:17-32) -> fp32 [n][1024]
void launch_pos_emb(cudaStream_t st, float* out, int p0, int n);

}  // namespace nb

// ---- decoder.cu ----------------------------------------------------------------------------------
namespace nb {

// C[M,N] = A[M,K] * W[N,K]^T + bias, fp32 on CUDA cores (the reference keeps LSTM/joint weights and math in fp32,
// src/nemo-ggml.cpp:580-619,1186-1224). N % 64 == 0, K % 16 == 0. Row m of C is stored at row out_rows[m] if given.
void launch_sgemm(cudaStream_t st, const float* A, const float* W, const float* bias, float* C, int M, int N, int K,
                  int ldc, const int* out_rows, int* zero3 = nullptr);

struct DecodeState {           // per stream slot, persistent in HBM
    float* h;                  // [slots][2][640] committed LSTM hidden state
    float* c;                  // [slots][2][640]
    float* pend_h;             // [slots][2][640] LSTM(embed[prev], h, c) — what the next symbol step would commit
    float* pend_c;
    float* pend_p;             // [slots][640]    joint.pred(pend_h[1]) + bias
    int* prev;                 // [slots] previous non-blank token (blank at start)
};
struct DecodeBatch {           // per batched step
    const float* e;            // [B*T][640] joint.enc(enc_out) + bias
    const int* slot;           // [B]
    const int* nvalid;         // [B] encoder frames to decode for this chunk
    int* fp;                   // [B] frame pointer
    int* sym;                  // [B] symbols emitted at the current frame
    int* tok_out;              // [B][T*MAX_SYMBOLS]
    int* tok_cnt;              // [B]
    float* min_margin;         // [B] smallest top1-top2 logit gap over this chunk's symbol steps (near-tie tracing)
    int* counters;             // [16]: unfinished[3], emit_count[3] (rotating per round), [6] streams left, [7] rounds run, [8] = bar,
                               //       [10..13] the kernel's start / end globaltimer stamps
    unsigned* bar;             // = counters + 8: arrival counter of the grid-wide barriers
    int4* emit_list;           // [B] (batch index, slot, token, -) of the streams that emitted in the current round
    float* partial;            // [B][G][4] per-CTA (max, runner-up, index) of the current round's logits
    float* hid;                // [B*W][640] joint hidden vector of every stream's next symbol step(s)
    int B, T, max_rounds;
    int W;                     // frames evaluated per stream and round (1 = one symbol step per round; > 1: window mode, B * W <= 64 items)
};
struct DecodeWeights {
    const float* embed;        // [1025][640]
    const float* wcat[2];      // [2560][1280] = [W_ih | W_hh]
    const float* bcat[2];      // [2560] = b_ih + b_hh
    const float* wpred;        // [640][640]
    const float* bpred;
    const float* wout;         // [1088][640], rows >= 1025 are zero
    const float* bout;         // [1088], entries >= 1025 are -inf
    // The same matrices pre-split for the decode kernel's fp16 tensor-core dot products: every 4 consecutive floats become 16
    // bytes {hi0 hi1 hi2 hi3 | lo0 lo1 lo2 lo3} of halves, hi = fp16(w), lo = fp16((w - hi) * 2048) (decoder.cu). Same geometry
    // as the fp32 matrix (4 bytes per element), so rows are addressed alike.
    const float* wcat_s[2];
    const float* wpred_s;
    const float* wout_s;
    // Layer 0 of the LSTM sees x = embed[prev]: its input half W_ih0 . embed[tok] + (b_ih0 + b_hh0) only depends on the token, so
    // it is a table [1025][2560] built at load (fp32 SGEMM); the kernel then multiplies only the recurrent half W_hh0 . h0
    // (K = 640 instead of 1280: half the staging and half the MMAs of that phase) and adds the token's row.
    const float* t0;           // [1025][2560]
    const float* whh0_s;       // [2560][640] W_hh of layer 0, pre-split like wcat_s
};
// all greedy rounds of one batched chunk (src/nemo-stream.cpp:840-930, 1106-1118) as one persistent launch
void launch_decode_chunk(cudaStream_t st, const DecodeWeights& w, const DecodeState& s, const DecodeBatch& b);
int decode_partial_floats(int B);  // size of DecodeBatch::partial
// recompute the pending LSTM/pred state of the given slots from their committed state (after create/reset)
void launch_decode_refresh(cudaStream_t st, const DecodeWeights& w, const DecodeState& s, const DecodeBatch& b);

}  // namespace nb

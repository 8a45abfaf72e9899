// frontend.cu — log-mel preprocessor and the direct-convolution stages of the dw-striding subsampling.
//
// Mel (reference: src/preprocessor.cpp:330-395): one CTA per (frame, stream). The operation ORDER of the reference is
// mirrored so that fp32 results agree to an ulp of logf: pre-emphasis (:349-356), window (:184-194), bit-reversed
// radix-2 DIT FFT with the host libm twiddle table (:113-161), sqrtf then re-square (:200-203,:363-368), sequential
// mel sum over FFT bins (:374-383). Every multiply/add is an explicit round-to-nearest intrinsic so nvcc cannot
// contract them into FMAs (the reference binary is built without FMA, Makefile:7). Zero filterbank weights are skipped:
// adding +0 never changes an fp32 sum, so the result is bit-identical to the dense 257-term loop.
#include <cstdlib>

#include "kernels.cuh"

namespace nb {

__global__ void __launch_bounds__(256) mel_kernel(MelTables tb, MelBatch mb, float* __restrict__ mel_out) {
    pdl_prologue();
    const int f = blockIdx.x, b = blockIdx.y;
    if (f >= mb.n_frames[b]) return;
    __shared__ float xs[N_FFT + 1];
    __shared__ float re[N_FFT], im[N_FFT];
    __shared__ float pw[N_BINS];
    const int t = threadIdx.x;
    // element i of this frame is absolute sample a0 + 160 f + i; samples before the stream start are the centre
    // padding (zeros), samples at or beyond n_avail are the zero padding of a finalize tail.
    const int16_t* ring = mb.ring + (size_t)mb.slot[b] * mb.ring_cap;
    const long long a_first = mb.a0[b] + (long long)f * HOP, n_avail = mb.n_avail[b];
    for (int i = t; i < N_FFT + 1; i += 256) {
        const long long a = a_first + i;
        const int16_t s = (a >= 0 && a < n_avail) ? ring[a & (long long)(mb.ring_cap - 1)] : (int16_t)0;
        xs[i] = __fmul_rn((float)s, 1.0f / 32768.0f);
    }
    __syncthreads();
    for (int i = t; i < N_FFT; i += 256) {
        float y = __fsub_rn(xs[i + 1], __fmul_rn(0.97f, xs[i]));
        y = __fmul_rn(y, tb.win512[i]);
        int r = __brev((unsigned)i) >> 23;  // 9-bit reversal
        re[r] = y;
        im[r] = 0.0f;
    }
    __syncthreads();
#pragma unroll 1
    for (int lg = 1; lg <= 9; lg++) {
        const int m2 = 1 << (lg - 1);
        const int j = t & (m2 - 1);
        const int i1 = ((t >> (lg - 1)) << lg) + j;
        const int i2 = i1 + m2;
        const int idx = j << (9 - lg);
        const float wr = tb.cos_tbl[idx];
        const float wi = -tb.sin_tbl[idx];
        const float r2 = re[i2], q2 = im[i2], r1 = re[i1], q1 = im[i1];
        const float tr = __fsub_rn(__fmul_rn(wr, r2), __fmul_rn(wi, q2));
        const float ti = __fadd_rn(__fmul_rn(wr, q2), __fmul_rn(wi, r2));
        re[i2] = __fsub_rn(r1, tr);
        im[i2] = __fsub_rn(q1, ti);
        re[i1] = __fadd_rn(r1, tr);
        im[i1] = __fadd_rn(q1, ti);
        __syncthreads();
    }
    for (int k = t; k < N_BINS; k += 256) {
        float mag = __fsqrt_rn(__fadd_rn(__fmul_rn(re[k], re[k]), __fmul_rn(im[k], im[k])));
        pw[k] = __fmul_rn(mag, mag);
    }
    __syncthreads();
    if (t < N_MELS) {
        const int2 rg = tb.fb_range[t];
        float sum = 0.0f;
        for (int k = rg.x; k < rg.y; k++) sum = __fadd_rn(sum, __fmul_rn(tb.fbT[k * N_MELS + t], pw[k]));
        mel_out[mb.out_off[b] + (long long)f * N_MELS + t] = logf(__fadd_rn(sum, 5.960464477539063e-8f));
    }
}

// ---- radix-8 register version -------------------------------------------------------------------------------------
// Same butterflies, same twiddle table entries, same operation order inside every butterfly as mel_kernel above (hence
// bit-identical: a radix-2 DIT stage only orders independent butterflies), but three consecutive stages are done on 8
// complex values held in registers: 3 shared-memory round trips and 5 barriers per frame instead of 9 and 13. 64
// threads per frame, 4 frames per CTA. Shared-memory index i lives at i + 8 * (i >> 6): the pad makes the stride-8
// (stages 4-6) and stride-64 (stages 7-9) accesses conflict-free.
constexpr int MEL_FPB = 4;  // frames per CTA
__device__ __forceinline__ int mel_pos(int i) { return i + ((i >> 6) << 3); }

// One radix-2 butterfly exactly as in the reference's fft_frame (src/preprocessor.cpp:140-156)
__device__ __forceinline__ void mel_bfly(float& r1, float& q1, float& r2, float& q2, float wr, float wi) {
    const float tr = __fsub_rn(__fmul_rn(wr, r2), __fmul_rn(wi, q2));
    const float ti = __fadd_rn(__fmul_rn(wr, q2), __fmul_rn(wi, r2));
    const float a = r1, c = q1;
    r2 = __fsub_rn(a, tr);
    q2 = __fsub_rn(c, ti);
    r1 = __fadd_rn(a, tr);
    q1 = __fadd_rn(c, ti);
}

// Stages 3*PASS+1 .. 3*PASS+3 on the thread's 8 values. Element k of thread g is global index
//   PASS 0: 8 g + k      PASS 1: 64 (g >> 3) + (g & 7) + 8 k      PASS 2: g + 64 k
// j0 = the part of the butterfly's twiddle index j = i1 & (m2 - 1) that does not depend on k.
template <int PASS>
__device__ __forceinline__ void mel_fft_pass(float (&r)[8], float (&q)[8], int g, const float* __restrict__ cs, const float* __restrict__ sn) {
    const int j0 = PASS == 0 ? 0 : (PASS == 1 ? (g & 7) : g);
    constexpr int unit = PASS == 0 ? 1 : (PASS == 1 ? 8 : 64);  // index distance between consecutive k
#pragma unroll
    for (int sp = 1; sp <= 3; sp++) {
        const int lg = 3 * PASS + sp, half = 1 << (sp - 1);
#pragma unroll
        for (int k = 0; k < 8; k++) {
            if (k & half) continue;
            const int j = j0 + unit * (k & (half - 1));
            const int idx = j << (9 - lg);
            mel_bfly(r[k], q[k], r[k + half], q[k + half], cs[idx], -sn[idx]);
        }
    }
}

__global__ void __launch_bounds__(64 * MEL_FPB) mel_r8_kernel(MelTables tb, MelBatch mb, float* __restrict__ mel_out) {
    pdl_prologue();
    constexpr int PADN = N_FFT + 8 * (N_FFT / 64);  // 576
    __shared__ float xs[MEL_FPB][N_FFT + 1];
    __shared__ float re[MEL_FPB][PADN], im[MEL_FPB][PADN];
    __shared__ float pw[MEL_FPB][N_BINS + 3];
    __shared__ float cs[N_FFT / 2], sn[N_FFT / 2];
    const int t = threadIdx.x, fs = t >> 6, g = t & 63, b = blockIdx.y;
    const int f = blockIdx.x * MEL_FPB + fs;
    const bool active = f < mb.n_frames[b];
    for (int i = t; i < N_FFT / 2; i += 64 * MEL_FPB) {
        cs[i] = tb.cos_tbl[i];
        sn[i] = tb.sin_tbl[i];
    }
    if (active) {
        const int16_t* ring = mb.ring + (size_t)mb.slot[b] * mb.ring_cap;
        const long long a_first = mb.a0[b] + (long long)f * HOP, n_avail = mb.n_avail[b];
        for (int i = g; i < N_FFT + 1; i += 64) {
            const long long a = a_first + i;
            const int16_t s = (a >= 0 && a < n_avail) ? ring[a & (long long)(mb.ring_cap - 1)] : (int16_t)0;
            xs[fs][i] = __fmul_rn((float)s, 1.0f / 32768.0f);
        }
    }
    __syncthreads();
    float r[8], q[8];
    if (active) {
        // pass 0 reads its inputs straight from the windowed signal: element k of thread g is bit-reversed index 8 g + k,
        // i.e. time sample brev9(8 g + k) (src/preprocessor.cpp:120-131)
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const int i = __brev((unsigned)(8 * g + k)) >> 23;
            float y = __fsub_rn(xs[fs][i + 1], __fmul_rn(0.97f, xs[fs][i]));
            r[k] = __fmul_rn(y, tb.win512[i]);
            q[k] = 0.0f;
        }
        mel_fft_pass<0>(r, q, g, cs, sn);
#pragma unroll
        for (int k = 0; k < 8; k++) {
            re[fs][mel_pos(8 * g + k)] = r[k];
            im[fs][mel_pos(8 * g + k)] = q[k];
        }
    }
    __syncthreads();
    if (active) {
        const int base = 64 * (g >> 3) + (g & 7);
#pragma unroll
        for (int k = 0; k < 8; k++) {
            r[k] = re[fs][mel_pos(base + 8 * k)];
            q[k] = im[fs][mel_pos(base + 8 * k)];
        }
        mel_fft_pass<1>(r, q, g, cs, sn);
#pragma unroll
        for (int k = 0; k < 8; k++) {
            re[fs][mel_pos(base + 8 * k)] = r[k];
            im[fs][mel_pos(base + 8 * k)] = q[k];
        }
    }
    __syncthreads();
    if (active) {
#pragma unroll
        for (int k = 0; k < 8; k++) {
            r[k] = re[fs][mel_pos(g + 64 * k)];
            q[k] = im[fs][mel_pos(g + 64 * k)];
        }
        mel_fft_pass<2>(r, q, g, cs, sn);
        // bins 0..256: elements k = 0..3 of every thread, plus element 4 of thread 0
#pragma unroll
        for (int k = 0; k < 5; k++) {
            if (k == 4 && g != 0) break;
            const float mag = __fsqrt_rn(__fadd_rn(__fmul_rn(r[k], r[k]), __fmul_rn(q[k], q[k])));
            pw[fs][g + 64 * k] = __fmul_rn(mag, mag);
        }
    }
    __syncthreads();
    if (active) {
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const int mbin = g + 64 * h;
            const int2 rg = tb.fb_range[mbin];
            float sum = 0.0f;
            for (int k = rg.x; k < rg.y; k++) sum = __fadd_rn(sum, __fmul_rn(tb.fbT[k * N_MELS + mbin], pw[fs][k]));
            mel_out[mb.out_off[b] + (long long)f * N_MELS + mbin] = logf(__fadd_rn(sum, 5.960464477539063e-8f));
        }
    }
}

void launch_mel(cudaStream_t st, const MelTables& tb, const MelBatch& mb, float* mel_out, int B, int max_frames) {
    if (B <= 0 || max_frames <= 0) return;
    static const bool radix2 = getenv("NB200_MEL_RADIX2") != nullptr;  // the one-stage-per-barrier kernel, kept as the checker
    ProfScope ps(st, "mel", (double)B * max_frames * 2.0e5, (double)B * max_frames * (HOP * 2 + N_MELS * 4));
    if (radix2)
        launch_k(mel_kernel, dim3(dim3(max_frames, B)), dim3(256), 0, st, tb, mb, mel_out);
    else
        for (int rep = dup_family("mel") ? 2 : 1; rep > 0; rep--) launch_k(mel_r8_kernel, dim3(dim3(cdiv(max_frames, MEL_FPB), B)), dim3(64 * MEL_FPB), 0, st, tb, mb, mel_out);
    NB_KERNEL_CHECK();
}

// ---- subsampling stage 1: conv0 (1->256, 3x3 s2, fp16-rounded operands) + ReLU + depthwise conv2 (3x3 s2) ----------
// Causal zero padding (2 before, 1 after) on both axes at every stage (src/nemo-ggml.cpp:905-913). One CTA per
// (h2, stream): thread = channel; the 7 mel rows feeding this output row sit in shared memory, fp16-rounded.
// The thread walks along the frequency axis: per output column w2 it reads 4 new mel bins of each row (one 128-bit
// broadcast load per row), evaluates the two new conv0 columns (w1 = 2 w2 - 1, 2 w2) of the three conv0 rows and
// reuses column 2 w2 - 2 from the previous step, so every conv0 value is computed once per CTA instead of 9/4 times.
// Summation orders (conv0: taps row-major; conv2: kh outer, kw inner) are those of the plain triple loop.
constexpr int S1_PAD = 7;                 // smem column of mel bin 0: the 4 new bins of step w2 start at column 4 w2 + 4
constexpr int S1_COLS = N_MELS + 1 + S1_PAD;  // bins -7 .. 128

__global__ void __launch_bounds__(256) sub_stage1_kernel(SubWeights w, const float* __restrict__ mel, __half* __restrict__ out,
                                                         int chunk_mel, int t1, int t2) {
    pdl_prologue();
    const int h2 = blockIdx.x, b = blockIdx.y, c = threadIdx.x;
    __shared__ __align__(16) float ms[7][S1_COLS];
    const float* msrc = mel + (size_t)b * chunk_mel * N_MELS;
    for (int i = c; i < 7 * S1_COLS; i += 256) {
        const int r = i / S1_COLS, col = i - r * S1_COLS;
        const int tt = 4 * h2 - 6 + r, ff = col - S1_PAD;
        float v = 0.0f;
        if (tt >= 0 && tt < chunk_mel && ff >= 0 && ff < N_MELS) v = __half2float(__float2half_rn(msrc[(size_t)tt * N_MELS + ff]));
        ms[r][col] = v;
    }
    float k0[9], k2[9];
#pragma unroll
    for (int i = 0; i < 9; i++) {
        k0[i] = w.w0[c * 9 + i];
        k2[i] = w.w2[c * 9 + i];
    }
    const float bias0 = w.b0[c], bias2 = w.b2[c];
    bool row_ok[3];
#pragma unroll
    for (int kh = 0; kh < 3; kh++) {
        const int h1 = 2 * h2 + kh - 2;
        row_ok[kh] = h1 >= 0 && h1 < t1;  // otherwise a zero padding row of the conv0 output
    }
    __syncthreads();
    __half* orow = out + ((size_t)(b * t2 + h2) * SUB_F2) * SUB_CH + c;
    float carry[7], prev_e[3] = {0.f, 0.f, 0.f};
#pragma unroll
    for (int r = 0; r < 7; r++) carry[r] = 0.0f;  // bin -4: only feeds the (invalid) column w1 = -1
#pragma unroll 1
    for (int w2 = 0; w2 < SUB_F2; w2++) {
        float4 v[7];
#pragma unroll
        for (int r = 0; r < 7; r++) v[r] = *reinterpret_cast<const float4*>(&ms[r][4 * w2 + 4]);  // bins 4 w2 - 3 .. 4 w2
        float acc = bias2;
#pragma unroll
        for (int kh = 0; kh < 3; kh++) {
            if (!row_ok[kh]) continue;
            float ao = 0.0f, ae = 0.0f;
#pragma unroll
            for (int i = 0; i < 3; i++) {
                const int r = 2 * kh + i;
                ao = fmaf(k0[i * 3 + 0], carry[r], ao);
                ao = fmaf(k0[i * 3 + 1], v[r].x, ao);
                ao = fmaf(k0[i * 3 + 2], v[r].y, ao);
                ae = fmaf(k0[i * 3 + 0], v[r].y, ae);
                ae = fmaf(k0[i * 3 + 1], v[r].z, ae);
                ae = fmaf(k0[i * 3 + 2], v[r].w, ae);
            }
            ao = fmaxf(ao + bias0, 0.0f);
            ae = fmaxf(ae + bias0, 0.0f);
            if (w2 > 0) {  // columns w1 = 2 w2 - 2 and 2 w2 - 1 exist
                acc = fmaf(k2[kh * 3 + 0], prev_e[kh], acc);
                acc = fmaf(k2[kh * 3 + 1], ao, acc);
            }
            acc = fmaf(k2[kh * 3 + 2], ae, acc);
            prev_e[kh] = ae;
        }
#pragma unroll
        for (int r = 0; r < 7; r++) carry[r] = v[r].w;
        orow[(size_t)w2 * SUB_CH] = __float2half_rn(acc);
    }
}

void launch_sub_stage1(cudaStream_t st, const SubWeights& w, const float* mel, __half* out, int B, int chunk_mel, int t1, int t2) {
    ProfScope ps(st, "sub_conv0_dw2", (double)B * t2 * SUB_F2 * SUB_CH * 2.0 * 9 * 10,
                 (double)B * (chunk_mel * N_MELS * 4.0 + (double)t2 * SUB_F2 * SUB_CH * 2));
    for (int rep = dup_family("sub1") ? 2 : 1; rep > 0; rep--) launch_k(sub_stage1_kernel, dim3(dim3(t2, B)), dim3(256), 0, st, w, mel, out, chunk_mel, t1, t2);
    NB_KERNEL_CHECK();
}

// ---- subsampling stage 2: depthwise conv5 over relu(pw3), only the rows that survive the drop of `drop` frames (2 in the
// streaming path, src/nemo-stream.cpp:154-162; 0 in the offline path) -------------------------------------------------
__global__ void __launch_bounds__(256) sub_stage2_kernel(SubWeights w, const float* __restrict__ a3, __half* __restrict__ out,
                                                         int t2, int t3, int drop) {
    pdl_prologue();
    const int w3 = blockIdx.x, hh = blockIdx.y, b = blockIdx.z, c = threadIdx.x;
    const int h3 = hh + drop;
    float acc = w.b5[c];
#pragma unroll
    for (int kh = 0; kh < 3; kh++) {
        const int h2 = 2 * h3 + kh - 2;
        if (h2 < 0 || h2 >= t2) continue;
#pragma unroll
        for (int kw = 0; kw < 3; kw++) {
            const int w2 = 2 * w3 + kw - 2;
            if (w2 < 0 || w2 >= SUB_F2) continue;
            acc = fmaf(w.w5[c * 9 + kh * 3 + kw], a3[((size_t)(b * t2 + h2) * SUB_F2 + w2) * SUB_CH + c], acc);
        }
    }
    out[((size_t)(b * (t3 - drop) + hh) * SUB_F3 + w3) * SUB_CH + c] = __float2half_rn(acc);
}

void launch_sub_stage2(cudaStream_t st, const SubWeights& w, const float* a3, __half* out, int B, int t2, int t3, int drop) {
    ProfScope ps(st, "sub_dw5", (double)B * (t3 - drop) * SUB_F3 * SUB_CH * 18.0,
                 (double)B * ((double)t2 * SUB_F2 * SUB_CH * 4 + (double)(t3 - drop) * SUB_F3 * SUB_CH * 2));
    for (int rep = dup_family("sub2") ? 2 : 1; rep > 0; rep--) launch_k(sub_stage2_kernel, dim3(dim3(SUB_F3, t3 - drop, B)), dim3(256), 0, st, w, a3, out, t2, t3, drop);
    NB_KERNEL_CHECK();
}

}  // namespace nb

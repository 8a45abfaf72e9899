// decoder.cu — RNNT prediction network (2-layer LSTM), joint network and the greedy symbol loop, batched over streams.
//
// Reference semantics (src/nemo-stream.cpp:840-930): per encoder frame up to 10 symbol steps; each step runs
// LSTM(embed[prev], h, c) and the joint, takes the FIRST maximum; blank -> next frame with the state untouched;
// non-blank -> emit, prev = token, commit the LSTM state. Because a blank leaves (prev, h, c) untouched, the LSTM
// output of a step only changes after an emission: it is kept per stream as "pending" state and recomputed after
// each emission, so a round costs one joint evaluation plus one batched LSTM update. Streams advance through their
// frames independently (no lock-step over frames).
#include <algorithm>

#include "kernels.cuh"

namespace nb {

constexpr int VOCAB_PAD = 1088;

// ---- fp32 SGEMM (NT), 64x64x16 tiles, 4x4 per thread -------------------------------------------------
__global__ void __launch_bounds__(256) sgemm_nt_kernel(const float* __restrict__ A, const float* __restrict__ W,
                                                       const float* __restrict__ bias, float* __restrict__ C, int M, int N,
                                                       int K, int ldc, const int* __restrict__ out_rows) {
    pdl_prologue();
    __shared__ float As[16][64 + 4];
    __shared__ float Ws[16][64 + 4];
    const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
    const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
    const int lr = t >> 2, lk = (t & 3) * 4;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] = 0.0f;
    for (int k0 = 0; k0 < K; k0 += 16) {
        float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
        if (m0 + lr < M) a = *reinterpret_cast<const float4*>(A + (size_t)(m0 + lr) * K + k0 + lk);
        const float4 w = *reinterpret_cast<const float4*>(W + (size_t)(n0 + lr) * K + k0 + lk);
        As[lk + 0][lr] = a.x; As[lk + 1][lr] = a.y; As[lk + 2][lr] = a.z; As[lk + 3][lr] = a.w;
        Ws[lk + 0][lr] = w.x; Ws[lk + 1][lr] = w.y; Ws[lk + 2][lr] = w.z; Ws[lk + 3][lr] = w.w;
        __syncthreads();
#pragma unroll
        for (int k = 0; k < 16; k++) {
            const float4 av = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
            const float4 wv = *reinterpret_cast<const float4*>(&Ws[k][tx * 4]);
            const float aa[4] = {av.x, av.y, av.z, av.w};
            const float ww[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
            for (int i = 0; i < 4; i++)
#pragma unroll
                for (int j = 0; j < 4; j++) acc[i][j] = fmaf(aa[i], ww[j], acc[i][j]);
        }
        __syncthreads();
    }
    const int n = n0 + tx * 4;
    float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
    if (bias) bv = *reinterpret_cast<const float4*>(bias + n);
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int m = m0 + ty * 4 + i;
        if (m >= M) continue;
        const int orow = out_rows ? out_rows[m] : m;
        *reinterpret_cast<float4*>(C + (size_t)orow * ldc + n) =
            make_float4(acc[i][0] + bv.x, acc[i][1] + bv.y, acc[i][2] + bv.z, acc[i][3] + bv.w);
    }
}

// Skinny variant for the decode loop (M <= 64 rows, weights streamed from L2): 64 x 16 tiles so that N/16 CTAs
// (160 for the LSTM gates, 68 for the joint output) spread over the SMs; each thread owns 4 rows x 1 column.
// `gate`: when non-null and *gate == 0 the whole launch is a no-op (no stream emitted this round -> LSTM unchanged).
__global__ void __launch_bounds__(256) sgemm_skinny_kernel(const float* __restrict__ A, const float* __restrict__ W,
                                                           const float* __restrict__ bias, float* __restrict__ C, int M, int N,
                                                           int K, int ldc, const int* __restrict__ out_rows,
                                                           const int* __restrict__ gate) {
    pdl_prologue();
    if (gate != nullptr && *gate == 0) return;
    __shared__ float As[32][64 + 4];
    __shared__ float Ws[32][16 + 1];
    const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
    const int n0 = blockIdx.x * 16;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    const int ar = t >> 2, ak = (t & 3) * 8;   // A tile 64 rows x 32 k: thread loads 8 consecutive k of one row
    const int wr = t >> 4, wk = (t & 15) * 2;  // W tile 16 rows x 32 k: thread loads 2 consecutive k of one row
    for (int k0 = 0; k0 < K; k0 += 32) {
        float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0;
        if (ar < M) {
            const float4* src = reinterpret_cast<const float4*>(A + (size_t)ar * K + k0 + ak);
            a0 = src[0];
            a1 = src[1];
        }
        const float2 w = *reinterpret_cast<const float2*>(W + (size_t)(n0 + wr) * K + k0 + wk);
        As[ak + 0][ar] = a0.x; As[ak + 1][ar] = a0.y; As[ak + 2][ar] = a0.z; As[ak + 3][ar] = a0.w;
        As[ak + 4][ar] = a1.x; As[ak + 5][ar] = a1.y; As[ak + 6][ar] = a1.z; As[ak + 7][ar] = a1.w;
        Ws[wk + 0][wr] = w.x;
        Ws[wk + 1][wr] = w.y;
        __syncthreads();
#pragma unroll
        for (int k = 0; k < 32; k++) {
            const float4 av = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
            const float wv = Ws[k][tx];
            acc[0] = fmaf(av.x, wv, acc[0]);
            acc[1] = fmaf(av.y, wv, acc[1]);
            acc[2] = fmaf(av.z, wv, acc[2]);
            acc[3] = fmaf(av.w, wv, acc[3]);
        }
        __syncthreads();
    }
    const int n = n0 + tx;
    const float bv = bias ? bias[n] : 0.0f;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int m = ty * 4 + i;
        if (m >= M) continue;
        const int orow = out_rows ? out_rows[m] : m;
        C[(size_t)orow * ldc + n] = acc[i] + bv;
    }
}

void launch_sgemm(cudaStream_t st, const float* A, const float* W, const float* bias, float* C, int M, int N, int K,
                  int ldc, const int* out_rows, const int* gate) {
    if (M <= 0) return;
    if (N % 64 != 0 || K % 32 != 0) throw Error("sgemm: N must be a multiple of 64 and K of 32");
    const double flops = 2.0 * M * (double)N * K, bytes = 4.0 * ((double)M * K + (double)N * K + (double)M * N);
    if (M <= 64) {
        ProfScope ps(st, "sgemm_f32_skinny", flops, bytes);
        launch_k(sgemm_skinny_kernel, dim3(N / 16), dim3(256), 0, st, A, W, bias, C, M, N, K, ldc, out_rows, gate);
        NB_KERNEL_CHECK();
        return;
    }
    if (gate) throw Error("sgemm: gated launches are only supported for M <= 64");
    ProfScope ps(st, "sgemm_f32", flops, bytes);
    launch_k(sgemm_nt_kernel, dim3(dim3(N / 64, cdiv(M, 64))), dim3(256), 0, st, A, W, bias, C, M, N, K, ldc, out_rows);
    NB_KERNEL_CHECK();
}

// ---- round kernels -------------------------------------------------------------------------------------
// hid[b] = relu(e[b, fp[b]] + pend_p[slot[b]])   (src/nemo-ggml.cpp:1215-1217); zeros for finished streams
__global__ void __launch_bounds__(160) hid_kernel(DecodeState s, DecodeBatch d) {
    pdl_prologue();
    const int b = blockIdx.x, i = threadIdx.x * 4;
    if (b == 0 && threadIdx.x == 0) {  // first kernel of a round: reset the round's flags
        *d.unfinished = 0;
        *d.any_emit = 0;
    }
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    const int fp = d.fp[b];
    if (fp < d.nvalid[b]) {
        const float4 e = *reinterpret_cast<const float4*>(d.e + ((size_t)(b * d.T + fp)) * JOINT + i);
        const float4 p = *reinterpret_cast<const float4*>(s.pend_p + (size_t)d.slot[b] * JOINT + i);
        v = make_float4(fmaxf(e.x + p.x, 0.f), fmaxf(e.y + p.y, 0.f), fmaxf(e.z + p.z, 0.f), fmaxf(e.w + p.w, 0.f));
    }
    *reinterpret_cast<float4*>(d.hid + (size_t)b * JOINT + i) = v;
}

// first-max argmax over the 1025 logits, then the reference's state machine (src/nemo-stream.cpp:899-926);
// also stages the LSTM layer-0 input row x[b] = [embed[prev] | h0].
__global__ void __launch_bounds__(256) argmax_commit_kernel(DecodeWeights w, DecodeState s, DecodeBatch d) {
    pdl_prologue();
    const int b = blockIdx.x, t = threadIdx.x, slot = d.slot[b];
    __shared__ float bv[8], bs[8];
    __shared__ int bi[8];
    __shared__ int s_emit;
    const int fp = d.fp[b];
    const bool active = fp < d.nvalid[b];
    if (t == 0) s_emit = 0;
    if (active) {
        const float* lg = d.logits + (size_t)b * VOCAB_PAD;
        float best = -INFINITY, second = -INFINITY;
        int besti = 0x7fffffff;
        for (int v = t; v < VOCAB; v += 256) {
            const float x = lg[v];
            if (x > best) { second = best; best = x; besti = v; }  // increasing v per thread: strict '>' keeps the first maximum
            else if (x > second) second = x;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
            const float os = __shfl_xor_sync(0xffffffffu, second, o);
            const int oi = __shfl_xor_sync(0xffffffffu, besti, o);
            if (ob > best || (ob == best && oi < besti)) { second = fmaxf(best, os); best = ob; besti = oi; }
            else second = fmaxf(second, ob);
        }
        if ((t & 31) == 0) { bv[t >> 5] = best; bi[t >> 5] = besti; bs[t >> 5] = second; }
        __syncthreads();
        if (t == 0) {
            for (int i = 1; i < 8; i++) {
                if (bv[i] > best || (bv[i] == best && bi[i] < besti)) { second = fmaxf(best, bs[i]); best = bv[i]; besti = bi[i]; }
                else second = fmaxf(second, bv[i]);
            }
            d.min_margin[b] = fminf(d.min_margin[b], best - second);
            int f = fp, sy = d.sym[b];
            if (besti == BLANK || besti >= VOCAB) {
                f++;
                sy = 0;
            } else {
                d.tok_out[(size_t)b * d.T * MAX_SYMBOLS + d.tok_cnt[b]] = besti;
                d.tok_cnt[b]++;
                s.prev[slot] = besti;
                s_emit = 1;
                *d.any_emit = 1;  // benign race: every writer stores 1
                sy++;
                if (sy >= MAX_SYMBOLS) { f++; sy = 0; }
            }
            d.fp[b] = f;
            d.sym[b] = sy;
            if (f < d.nvalid[b]) atomicAdd(d.unfinished, 1);
        }
        __syncthreads();
        if (s_emit) {  // commit the pending LSTM state
            for (int i = t; i < 2 * DEC; i += 256) {
                s.h[(size_t)slot * 2 * DEC + i] = s.pend_h[(size_t)slot * 2 * DEC + i];
                s.c[(size_t)slot * 2 * DEC + i] = s.pend_c[(size_t)slot * 2 * DEC + i];
            }
        }
    }
    __syncthreads();
    const int prev = s.prev[slot];
    for (int i = t; i < DEC; i += 256) {
        d.x[(size_t)b * 2 * DEC + i] = w.embed[(size_t)prev * DEC + i];
        d.x[(size_t)b * 2 * DEC + DEC + i] = s.h[(size_t)slot * 2 * DEC + i];
    }
}

// stage the layer-0 input without touching the state machine (used by refresh)
__global__ void __launch_bounds__(256) stage_x0_kernel(DecodeWeights w, DecodeState s, DecodeBatch d) {
    pdl_prologue();
    const int b = blockIdx.x, t = threadIdx.x, slot = d.slot[b];
    const int prev = s.prev[slot];
    for (int i = t; i < DEC; i += 256) {
        d.x[(size_t)b * 2 * DEC + i] = w.embed[(size_t)prev * DEC + i];
        d.x[(size_t)b * 2 * DEC + DEC + i] = s.h[(size_t)slot * 2 * DEC + i];
    }
}

// LSTM cell, gate order i,f,g,o (src/nemo-ggml.cpp:601-618). layer 0 also stages x[b] = [h0' | h1]; layer 1 stages hp.
__global__ void __launch_bounds__(160) cell_kernel(DecodeState s, DecodeBatch d, int layer, const int* __restrict__ gate) {
    pdl_prologue();
    if (gate != nullptr && *gate == 0) return;
    const int b = blockIdx.x, slot = d.slot[b];
    const float* g = d.gates + (size_t)b * 4 * DEC;
    const size_t so = (size_t)slot * 2 * DEC + (size_t)layer * DEC;
    for (int i = threadIdx.x; i < DEC; i += 160) {
        const float ig = sigmoid_f(g[i]), fg = sigmoid_f(g[DEC + i]), gg = tanhf(g[2 * DEC + i]), og = sigmoid_f(g[3 * DEC + i]);
        const float c2 = fg * s.c[so + i] + ig * gg;
        const float h2 = og * tanhf(c2);
        s.pend_c[so + i] = c2;
        s.pend_h[so + i] = h2;
        if (layer == 0) {
            d.x[(size_t)b * 2 * DEC + i] = h2;
            d.x[(size_t)b * 2 * DEC + DEC + i] = s.h[(size_t)slot * 2 * DEC + DEC + i];
        } else {
            d.hp[(size_t)b * DEC + i] = h2;
        }
    }
}

// Recomputes the pending LSTM / joint.pred state of every stream of the batch from its committed state. In the
// greedy loop it is gated on the round's any_emit flag: a blank leaves (prev, h, c) untouched (src/nemo-stream.cpp:
// 907-911), so when no stream emitted the pending state is still valid and the five launches are no-ops.
static void lstm_update(cudaStream_t st, const DecodeWeights& w, const DecodeState& s, const DecodeBatch& d, const int* gate) {
    for (int b0 = 0; b0 < d.B; b0 += 64) {  // the skinny GEMM takes up to 64 rows
        const int nb = std::min(64, d.B - b0);
        DecodeBatch e = d;
        e.slot = d.slot + b0;
        float* x = d.x + (size_t)b0 * 2 * DEC;
        float* gates = d.gates + (size_t)b0 * 4 * DEC;
        float* hp = d.hp + (size_t)b0 * DEC;
        e.x = x; e.gates = gates; e.hp = hp;
        launch_sgemm(st, x, w.wcat[0], w.bcat[0], gates, nb, 4 * DEC, 2 * DEC, 4 * DEC, nullptr, gate);
        launch_k(cell_kernel, dim3(nb), dim3(160), 0, st, s, e, 0, gate);
        NB_KERNEL_CHECK();
        launch_sgemm(st, x, w.wcat[1], w.bcat[1], gates, nb, 4 * DEC, 2 * DEC, 4 * DEC, nullptr, gate);
        launch_k(cell_kernel, dim3(nb), dim3(160), 0, st, s, e, 1, gate);
        NB_KERNEL_CHECK();
        launch_sgemm(st, hp, w.wpred, w.bpred, s.pend_p, nb, JOINT, DEC, JOINT, e.slot, gate);
    }
}

void launch_decode_round(cudaStream_t st, const DecodeWeights& w, const DecodeState& s, const DecodeBatch& d) {
    launch_k(hid_kernel, dim3(d.B), dim3(160), 0, st, s, d);
    NB_KERNEL_CHECK();
    for (int b0 = 0; b0 < d.B; b0 += 64)
        launch_sgemm(st, d.hid + (size_t)b0 * JOINT, w.wout, w.bout, d.logits + (size_t)b0 * VOCAB_PAD, std::min(64, d.B - b0), VOCAB_PAD,
                     JOINT, VOCAB_PAD, nullptr, nullptr);
    launch_k(argmax_commit_kernel, dim3(d.B), dim3(256), 0, st, w, s, d);
    NB_KERNEL_CHECK();
    lstm_update(st, w, s, d, d.any_emit);
}

void launch_decode_refresh(cudaStream_t st, const DecodeWeights& w, const DecodeState& s, const DecodeBatch& d) {
    launch_k(stage_x0_kernel, dim3(d.B), dim3(256), 0, st, w, s, d);
    NB_KERNEL_CHECK();
    lstm_update(st, w, s, d, nullptr);
}

}  // namespace nb

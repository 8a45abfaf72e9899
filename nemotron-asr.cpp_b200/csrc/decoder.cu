// decoder.cu — RNNT prediction network (2-layer LSTM), joint network and the greedy symbol loop, batched over streams.
//
// Reference semantics (src/nemo-stream.cpp:840-930): per encoder frame up to 10 symbol steps; each step runs
// LSTM(embed[prev], h, c) and the joint, takes the FIRST maximum; blank -> next frame with the state untouched;
// non-blank -> emit, prev = token, commit the LSTM state. Because a blank leaves (prev, h, c) untouched, the LSTM
// output of a step only changes after an emission: it is kept per stream as "pending" state and recomputed after
// each emission, so a round costs one joint evaluation plus one batched LSTM update. Streams advance through their
// frames independently (no lock-step over frames).
#include <algorithm>
#include <cstdlib>

#include "kernels.cuh"

namespace nb {



// ---- fp32 SGEMM (NT), 64x64x16 tiles, 4x4 per thread; the next k-chunk is fetched into registers while the current one
// is multiplied (the K loop is otherwise a chain of exposed L2 round trips) --------------------------------------------
// `zero3`: when non-null, thread 0 of block (0,0) clears three ints there (the decode round's counters, see below).
__global__ void __launch_bounds__(256) sgemm_nt_kernel(const float* __restrict__ A, const float* __restrict__ W,
                                                       const float* __restrict__ bias, float* __restrict__ C, int M, int N,
                                                       int K, int ldc, const int* __restrict__ out_rows, int* zero3) {
    pdl_prologue();
    if (zero3 != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) zero3[0] = zero3[1] = zero3[2] = 0;
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
    const bool a_ok = m0 + lr < M;
    const float* ap = A + (size_t)(m0 + lr) * K + lk;
    const float* wp = W + (size_t)(n0 + lr) * K + lk;
    float4 a = a_ok ? *reinterpret_cast<const float4*>(ap) : make_float4(0.f, 0.f, 0.f, 0.f);
    float4 w = *reinterpret_cast<const float4*>(wp);
    for (int k0 = 0; k0 < K; k0 += 16) {
        As[lk + 0][lr] = a.x; As[lk + 1][lr] = a.y; As[lk + 2][lr] = a.z; As[lk + 3][lr] = a.w;
        Ws[lk + 0][lr] = w.x; Ws[lk + 1][lr] = w.y; Ws[lk + 2][lr] = w.z; Ws[lk + 3][lr] = w.w;
        __syncthreads();
        if (k0 + 16 < K) {
            if (a_ok) a = *reinterpret_cast<const float4*>(ap + k0 + 16);
            w = *reinterpret_cast<const float4*>(wp + k0 + 16);
        }
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
// (68 for the joint output) spread over the SMs; each thread owns 4 rows x 1 column. Same register prefetch.
__global__ void __launch_bounds__(256) sgemm_skinny_kernel(const float* __restrict__ A, const float* __restrict__ W,
                                                           const float* __restrict__ bias, float* __restrict__ C, int M, int N,
                                                           int K, int ldc, const int* __restrict__ out_rows, int* zero3) {
    pdl_prologue();
    if (zero3 != nullptr && blockIdx.x == 0 && threadIdx.x == 0) zero3[0] = zero3[1] = zero3[2] = 0;
    __shared__ float As[32][64 + 4];
    __shared__ float Ws[32][16 + 1];
    const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
    const int n0 = blockIdx.x * 16;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    const int ar = t >> 2, ak = (t & 3) * 8;   // A tile 64 rows x 32 k: thread loads 8 consecutive k of one row
    const int wr = t >> 4, wk = (t & 15) * 2;  // W tile 16 rows x 32 k: thread loads 2 consecutive k of one row
    const bool a_ok = ar < M;
    const float* ap = A + (size_t)ar * K + ak;
    const float* wp = W + (size_t)(n0 + wr) * K + wk;
    float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0;
    if (a_ok) {
        a0 = *reinterpret_cast<const float4*>(ap);
        a1 = *reinterpret_cast<const float4*>(ap + 4);
    }
    float2 w = *reinterpret_cast<const float2*>(wp);
    for (int k0 = 0; k0 < K; k0 += 32) {
        As[ak + 0][ar] = a0.x; As[ak + 1][ar] = a0.y; As[ak + 2][ar] = a0.z; As[ak + 3][ar] = a0.w;
        As[ak + 4][ar] = a1.x; As[ak + 5][ar] = a1.y; As[ak + 6][ar] = a1.z; As[ak + 7][ar] = a1.w;
        Ws[wk + 0][wr] = w.x;
        Ws[wk + 1][wr] = w.y;
        __syncthreads();
        if (k0 + 32 < K) {
            if (a_ok) {
                a0 = *reinterpret_cast<const float4*>(ap + k0 + 32);
                a1 = *reinterpret_cast<const float4*>(ap + k0 + 36);
            }
            w = *reinterpret_cast<const float2*>(wp + k0 + 32);
        }
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
                  int ldc, const int* out_rows, int* zero3) {
    if (M <= 0) return;
    if (N % 64 != 0 || K % 32 != 0) throw Error("sgemm: N must be a multiple of 64 and K of 32");
    const double flops = 2.0 * M * (double)N * K, bytes = 4.0 * ((double)M * K + (double)N * K + (double)M * N);
    if (M <= 64) {
        ProfScope ps(st, "sgemm_f32_skinny", flops, bytes);
        launch_k(sgemm_skinny_kernel, dim3(N / 16), dim3(256), 0, st, A, W, bias, C, M, N, K, ldc, out_rows, zero3);
        NB_KERNEL_CHECK();
        return;
    }
    ProfScope ps(st, "sgemm_f32", flops, bytes);
    launch_k(sgemm_nt_kernel, dim3(dim3(N / 64, cdiv(M, 64))), dim3(256), 0, st, A, W, bias, C, M, N, K, ldc, out_rows, zero3);
    NB_KERNEL_CHECK();
}

// ---- the greedy loop: one persistent launch per batched chunk -------------------------------------------------------
// G CTAs (one per SM) stay resident for all symbol rounds of the chunk and meet at grid-wide barriers. A round:
//   L  logits: CTA j owns vocabulary rows [j*R, j*R+R) of joint.out (kept in shared memory for the whole launch). For
//      every tile of 8 streams it stages hid = relu(joint.enc(frame fp) + pending pred) (src/nemo-ggml.cpp:1215-1217),
//      one warp per row forms the 8 dot products, and the slice's (first max, runner-up, index) goes to partial[b][j].
//   -- barrier --
//   A  argmax: one warp per stream merges the G partials (ties -> lowest index = the reference's strict '>' scan,
//      src/nemo-stream.cpp:899-906) and runs its state machine (:907-926): blank -> next frame, state untouched;
//      token -> emit, prev = token, commit the pending LSTM state, join emit_list.
//   -- barrier --
//   E  only if some stream emitted (a blank leaves (prev, h, c) untouched, so the pending state stays valid): the
//      pending LSTM / joint.pred state of the streams in emit_list is recomputed: layer 0, layer 1, joint.pred, each
//      followed by a barrier (every hidden unit of the previous one is needed); CTA j owns 4-5 hidden units (= the i,f,
//      g,o gate rows of those units, src/nemo-ggml.cpp:601-618) and loops over the emitting streams in tiles of 8.
// Window mode (DecodeBatch::W > 1, used by the offline path where one utterance is the whole batch): a blank leaves the
// decoder state untouched, so the logits of the next W frames can all be formed with the same pending prediction; a round
// evaluates W consecutive frames of the stream as W items, the argmax phase walks them in order up to the first non-blank
// one (everything before it is a blank step, everything after it is discarded), and a round advances up to W frames instead
// of one. Same steps, same arithmetic per step, same tokens and frame stamps as W = 1.
// The loop ends when no stream has frames left. Counters rotate over three slots so that clearing the next round's slot
// cannot race with a CTA still reading the previous one.
constexpr int DK_TS = 8, DK_MAXU = 8, DK_MAXR = 8, DK_GROUP = 64;
struct DkSmem {
    float xs[DK_GROUP * JOINT];        // L: hid of up to 64 streams [q][640]; E: input vectors of 8 streams [q][1280]
    float wrow[DK_MAXR][JOINT];        // this CTA's joint.out rows
    float gs[DK_TS][4 * DK_MAXU];      // gate pre-activations / logits of the tile
    int sl[DK_TS], bq[DK_TS], rows[4 * DK_MAXU];
    float bias[DK_MAXR];               // joint.out bias of my rows
    int act[DK_GROUP];                 // L: item indices (stream * W + window offset) of the group's pending symbol steps
    int n_act;
    float wbest[DK_GROUP], wsecond[DK_GROUP];  // A, window mode: merged (max, runner-up, index) of a stream's window items
    int widx[DK_GROUP], wemit, wf;
};
static_assert(DK_GROUP * JOINT >= DK_TS * 2 * DEC, "the LSTM tile reuses the hid buffer");

__device__ __forceinline__ void grid_barrier(unsigned* bar, unsigned& phase) {
    __syncthreads();
    if (threadIdx.x == 0) {
        phase++;
        __threadfence();
        atomicAdd(bar, 1u);
        const unsigned target = phase * gridDim.x;
        unsigned v;
        do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(bar) : "memory");
        } while (v < target);
        __threadfence();
    }
    __syncthreads();
}

// One warp: n <= R weight rows (wr[i] -> K floats, global or shared memory) times the 8 staged vectors xs[q * LD + k].
// Lane = (kg = lane & 7: interleaved K slice, sg = lane >> 3: stream pair). Every vector element a warp touches is read
// once per call, whatever the number of rows (one row per warp would re-read the whole stream tile from shared memory for
// every row: measured, that made shared-memory bandwidth the limit of the round). On return the lanes with kg == 0 hold
// v[i][j] = wr[i] . xs[2 sg + j].
// UNROLL k-steps are in flight at once: with the weights in L2 (the LSTM / joint.pred rows of phase E) the loop is bound by
// that latency, not by bandwidth, so E asks for 10 (30 independent 16-byte loads per lane); the logits phase reads its rows
// from shared memory and keeps 4.
template <int K, int LD, int R, int UNROLL = 4>
__device__ __forceinline__ void dk_block(const float* const (&wr)[R], int n, const float* xs, float (&v)[R][2]) {
    const int lane = threadIdx.x & 31, kg = lane & 7, sg = lane >> 3;
    const float* x0 = xs + (size_t)(2 * sg) * LD + 4 * kg;
    const float* x1 = x0 + LD;
#pragma unroll
    for (int i = 0; i < R; i++) v[i][0] = v[i][1] = 0.0f;
#pragma unroll UNROLL
    for (int k = 0; k < K; k += 32) {
        const float4 a = *reinterpret_cast<const float4*>(x0 + k), c = *reinterpret_cast<const float4*>(x1 + k);
#pragma unroll
        for (int i = 0; i < R; i++) {
            if (i < n) {
                const float4 wv = *reinterpret_cast<const float4*>(wr[i] + 4 * kg + k);
                v[i][0] = fmaf(wv.x, a.x, v[i][0]);
                v[i][0] = fmaf(wv.y, a.y, v[i][0]);
                v[i][0] = fmaf(wv.z, a.z, v[i][0]);
                v[i][0] = fmaf(wv.w, a.w, v[i][0]);
                v[i][1] = fmaf(wv.x, c.x, v[i][1]);
                v[i][1] = fmaf(wv.y, c.y, v[i][1]);
                v[i][1] = fmaf(wv.z, c.z, v[i][1]);
                v[i][1] = fmaf(wv.w, c.w, v[i][1]);
            }
        }
    }
#pragma unroll
    for (int i = 0; i < R; i++) {
#pragma unroll
        for (int j = 0; j < 2; j++) {
            float x = v[i][j];
            x += __shfl_xor_sync(0xffffffffu, x, 1);
            x += __shfl_xor_sync(0xffffffffu, x, 2);
            x += __shfl_xor_sync(0xffffffffu, x, 4);
            v[i][j] = x;
        }
    }
}

// rows[0..nr) of W (K floats each, row-major) times the DK_TS staged vectors xs[q][0..K): warp w takes rows w, w+8, w+16;
// out(q, r, value + bias[row]). wsm (optional): the same rows already staged in shared memory, row r at wsm + r * K.
template <int K, int LD, typename Out>
__device__ __forceinline__ void dk_rows(const float* __restrict__ W, const float* __restrict__ bias, const float* xs, int nr,
                                        const int* rows, Out&& out, const float* wsm = nullptr) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, sg = lane >> 3;
    const float* wr[3];
    int n = 0;
#pragma unroll
    for (int i = 0; i < 3; i++) {
        const int r = warp + 8 * i;
        wr[i] = wsm ? wsm + (size_t)(r < nr ? r : 0) * K : W + (size_t)rows[r < nr ? r : 0] * K;
        if (r < nr) n = i + 1;
    }
    if (n == 0) return;
    float v[3][2];
    dk_block<K, LD, 3, 10>(wr, n, xs, v);
    if ((lane & 7) == 0) {
#pragma unroll
        for (int i = 0; i < 3; i++) {
            if (i < n) {
                const int r = warp + 8 * i;
                const float bv = bias[rows[r]];
                out(2 * sg, r, v[i][0] + bv);
                out(2 * sg + 1, r, v[i][1] + bv);
            }
        }
    }
}

// cp.async copy of rows[0..nrows) of W (K floats each) to dst[r * K ..] by the whole CTA, one commit group
__device__ __forceinline__ void dk_stage_rows(float* dst, const float* __restrict__ W, int K, const int* rows, int nrows) {
    const int per_row = K / 4;
    for (int i = threadIdx.x; i < nrows * per_row; i += 256) {
        const int r = i / per_row, c = i - r * per_row;
        const uint32_t d = (uint32_t)__cvta_generic_to_shared(dst + (size_t)r * K + 4 * c);
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(W + (size_t)rows[r] * K + 4 * c) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
}
__device__ __forceinline__ void dk_stage_wait() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// Recomputes pend_h / pend_c / pend_p of the E streams listed in `list` (batch indices; nullptr = 0..E-1).
// The CTA's weight rows of a phase (16-20 gate rows x 1280 floats = 100 KB per LSTM layer, 4-5 x 640 for joint.pred) are brought
// into the idle part of the hid buffer with cp.async by all 256 threads (bandwidth- instead of latency-bound), the next
// phase's rows while this phase finishes its gate math and waits at the grid barrier.
__device__ void dk_lstm_update(const DecodeWeights& w, const DecodeState& s, const DecodeBatch& d, DkSmem& sm, const int* list, int E,
                               unsigned& phase) {
    const int cta = blockIdx.x, G = gridDim.x, t = threadIdx.x;
    const int base = DEC / G, rem = DEC % G;
    const int nu = base + (cta < rem ? 1 : 0), u0 = cta * base + min(cta, rem);  // my hidden units [u0, u0 + nu)
    // behind the input tile: room for 24 rows x 1280 floats (nu <= 5 on 148 SMs needs 20); otherwise the rows are read from L2
    float* wsm = 4 * nu * 2 * DEC <= DK_GROUP * JOINT - DK_TS * 2 * DEC ? sm.xs + DK_TS * 2 * DEC : nullptr;
    __syncthreads();
    if (t < 4 * nu) sm.rows[t] = (t / nu) * DEC + u0 + (t % nu);  // the same gate rows in both layers; rows[0..nu) = u0.. are also joint.pred's
    __syncthreads();
    if (wsm) dk_stage_rows(wsm, w.wcat[0], 2 * DEC, sm.rows, 4 * nu);
    for (int layer = 0; layer < 2; layer++) {
        for (int e0 = 0; e0 < E; e0 += DK_TS) {
            __syncthreads();
            if (t < DK_TS) {
                const int b = (e0 + t < E) ? (list ? __ldcg(list + e0 + t) : e0 + t) : -1;
                sm.bq[t] = b;
                sm.sl[t] = b >= 0 ? d.slot[b] : -1;
            }
            __syncthreads();
            // x = [embed[prev] | h0] for layer 0, [h0' | h1] for layer 1 (h0' = this round's new layer-0 output)
            for (int i = t; i < DK_TS * (2 * DEC / 4); i += 256) {
                const int q = i / (2 * DEC / 4), k = (i - q * (2 * DEC / 4)) * 4, slot = sm.sl[q];
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (slot >= 0) {
                    const float* src;
                    if (k < DEC) src = layer == 0 ? w.embed + (size_t)__ldcg(s.prev + slot) * DEC + k : s.pend_h + (size_t)slot * 2 * DEC + k;
                    else src = s.h + (size_t)slot * 2 * DEC + (size_t)layer * DEC + (k - DEC);
                    v = __ldcg(reinterpret_cast<const float4*>(src));  // state written by other CTAs of this launch
                }
                *reinterpret_cast<float4*>(&sm.xs[q * 2 * DEC + k]) = v;
            }
            dk_stage_wait();  // this layer's rows (requested before the previous barrier), my copies ...
            __syncthreads();  // ... and everyone's, and the input tile
            dk_rows<2 * DEC, 2 * DEC>(w.wcat[layer], w.bcat[layer], sm.xs, 4 * nu, sm.rows, [&](int q, int r, float v) { sm.gs[q][r] = v; }, wsm);
            __syncthreads();
            if (wsm && e0 + DK_TS >= E) {  // last tile: the rows are dead, fetch the next phase's while the gates and the barrier run
                if (layer == 0) dk_stage_rows(wsm, w.wcat[1], 2 * DEC, sm.rows, 4 * nu);
                else dk_stage_rows(wsm, w.wpred, DEC, sm.rows, nu);
            }
            if (t < DK_TS * nu) {
                const int q = t / nu, ui = t % nu, slot = sm.sl[q];
                if (slot >= 0) {
                    const size_t so = (size_t)slot * 2 * DEC + (size_t)layer * DEC + u0 + ui;
                    const float ig = sigmoid_f(sm.gs[q][ui]), fg = sigmoid_f(sm.gs[q][nu + ui]);
                    const float gg = tanhf(sm.gs[q][2 * nu + ui]), og = sigmoid_f(sm.gs[q][3 * nu + ui]);
                    const float c2 = fg * __ldcg(s.c + so) + ig * gg;
                    s.pend_c[so] = c2;
                    s.pend_h[so] = og * tanhf(c2);
                }
            }
        }
        grid_barrier(d.bar, phase);
    }
    // joint.pred (src/nemo-ggml.cpp:1207-1211): pend_p = W_pred h1' + b_pred, rows u0 .. u0+nu-1
    if (t < nu) sm.rows[t] = u0 + t;
    for (int e0 = 0; e0 < E; e0 += DK_TS) {
        __syncthreads();
        if (t < DK_TS) {
            const int b = (e0 + t < E) ? (list ? __ldcg(list + e0 + t) : e0 + t) : -1;
            sm.bq[t] = b;
            sm.sl[t] = b >= 0 ? d.slot[b] : -1;
        }
        __syncthreads();
        for (int i = t; i < DK_TS * (DEC / 4); i += 256) {
            const int q = i / (DEC / 4), k = (i - q * (DEC / 4)) * 4, slot = sm.sl[q];
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (slot >= 0) v = __ldcg(reinterpret_cast<const float4*>(s.pend_h + (size_t)slot * 2 * DEC + DEC + k));
            *reinterpret_cast<float4*>(&sm.xs[q * 2 * DEC + k]) = v;
        }
        dk_stage_wait();
        __syncthreads();
        dk_rows<DEC, 2 * DEC>(w.wpred, w.bpred, sm.xs, nu, sm.rows, [&](int q, int r, float v) {
            if (sm.sl[q] < 0) return;
            s.pend_p[(size_t)sm.sl[q] * JOINT + u0 + r] = v;
            if (list) {  // the stream's next symbol step: hid = relu(joint.enc(frame fp) + new pending pred), my rows
                const int b = sm.bq[q], fp = __ldcg(d.fp + b), nv = d.nvalid[b];
                for (int wi = 0; wi < d.W && fp + wi < nv; wi++)
                    d.hid[(size_t)(b * d.W + wi) * JOINT + u0 + r] = fmaxf(d.e[((size_t)(b * d.T + fp + wi)) * JOINT + u0 + r] + v, 0.0f);
            }
        }, wsm);
    }
    grid_barrier(d.bar, phase);
}

// hid[b] = relu(joint.enc(frame fp) + pend_p[slot]) for one stream, by one warp (src/nemo-ggml.cpp:1215-1217)
// (item = b * W + window offset; W = 1: item = b)
__device__ __forceinline__ void dk_stage_hid(const DecodeState& s, const DecodeBatch& d, int item, int b, int fp, int slot, int lane) {
    for (int k = 4 * lane; k < JOINT; k += 128) {
        const float4 e = *reinterpret_cast<const float4*>(d.e + ((size_t)(b * d.T + fp)) * JOINT + k);
        const float4 p = __ldcg(reinterpret_cast<const float4*>(s.pend_p + (size_t)slot * JOINT + k));
        *reinterpret_cast<float4*>(d.hid + (size_t)item * JOINT + k) =
            make_float4(fmaxf(e.x + p.x, 0.f), fmaxf(e.y + p.y, 0.f), fmaxf(e.z + p.z, 0.f), fmaxf(e.w + p.w, 0.f));
    }
}

__global__ void __launch_bounds__(256) decode_persistent_kernel(DecodeWeights w, DecodeState s, DecodeBatch d, int refresh) {
    pdl_prologue();
    extern __shared__ __align__(16) uint8_t dk_raw[];
    DkSmem& sm = *reinterpret_cast<DkSmem*>(dk_raw);
    const int cta = blockIdx.x, G = gridDim.x, t = threadIdx.x, warp = t >> 5, lane = t & 31;
    unsigned phase = 0;
    unsigned long long* stamps = reinterpret_cast<unsigned long long*>(d.counters + 10);  // [10..11] start, [12..13] end (globaltimer ns)
    if (cta == 0 && t == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(stamps[0])::"memory");
    if (refresh) {  // stream create / reset: pending state of every batch item from its committed state
        dk_lstm_update(w, s, d, sm, nullptr, d.B, phase);
        return;
    }
    // my vocabulary rows, resident in shared memory for the whole launch
    const int R = (VOCAB + G - 1) / G, v0 = cta * R, nr = max(0, min(R, VOCAB - v0));
    const int n_part = (VOCAB + R - 1) / R;  // CTAs that own rows
    for (int i = t; i < nr * (JOINT / 4); i += 256) {
        const int r = i / (JOINT / 4), k = (i - r * (JOINT / 4)) * 4;
        *reinterpret_cast<float4*>(&sm.wrow[r][k]) = *reinterpret_cast<const float4*>(w.wout + (size_t)(v0 + r) * JOINT + k);
    }
    if (t < nr) sm.bias[t] = w.bout[v0 + t];
    // per-chunk decode state (was four memset nodes in front of this kernel): frame pointers, symbol counters, token counts,
    // margins = 3.39e38. Nothing reads them before the grid barrier below.
    for (int b = cta * 256 + t; b < d.B; b += G * 256) {
        d.fp[b] = 0;
        d.sym[b] = 0;
        d.tok_cnt[b] = 0;
        d.min_margin[b] = __int_as_float(0x7f7f7f7f);
    }
    // hid of every stream's first symbol step (fp = 0, pending pred from the previous chunk); the owner warp of stream b
    const int W = d.W;
    for (int b = cta * 8 + warp; b < d.B; b += G * 8)
        for (int wi = 0; wi < W && wi < d.nvalid[b]; wi++) dk_stage_hid(s, d, b * W + wi, b, wi, d.slot[b], lane);
    grid_barrier(d.bar, phase);

    float4* partial = reinterpret_cast<float4*>(d.partial);
    int* unfinished = d.counters;      // [3]
    int* emit_count = d.counters + 3;  // [3]
    int round = 0;
    unsigned long long* acc = reinterpret_cast<unsigned long long*>(d.counters + 16);
    auto now = [] { unsigned long long x; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(x)::"memory"); return x; };
    const bool tr = cta == 0 && t == 0;
    for (;; round++) {
        const int cs = round % 3;
        unsigned long long t0 = tr ? now() : 0;
        if (cta == 0 && t == 0) {  // clear the next round's counter slot (last read two barriers ago)
            unfinished[(round + 1) % 3] = 0;
            emit_count[(round + 1) % 3] = 0;
        }
        // ---- L: logits of my rows; the hid vectors of up to 64 unfinished streams are staged at once ----
        if (nr > 0) {
            for (int g0 = 0; g0 < d.B * W; g0 += DK_GROUP) {
                __syncthreads();
                if (warp == 0) {  // compact the group's pending items (ballot scan)
                    int n = 0;
                    for (int c = 0; c < DK_GROUP; c += 32) {
                        const int it = g0 + c + lane, b = it / W, wi = it - b * W;
                        const bool on = it < d.B * W && __ldcg(d.fp + b) + wi < d.nvalid[b];
                        const unsigned m = __ballot_sync(0xffffffffu, on);
                        if (on) sm.act[n + __popc(m & ((1u << lane) - 1u))] = it;
                        n += __popc(m);
                    }
                    if (lane == 0) sm.n_act = n;
                }
                __syncthreads();
                const int n_act = sm.n_act;
                unsigned long long u0 = tr ? now() : 0;
                if (n_act == 0) continue;  // uniform
                const int n_pad = (n_act + DK_TS - 1) / DK_TS * DK_TS;
                for (int i0 = 0; i0 < n_pad * (JOINT / 4); i0 += 256 * 8) {  // 8 independent 16-byte loads in flight per thread
                    float4 v[8];
#pragma unroll
                    for (int u = 0; u < 8; u++) {
                        const int i = i0 + u * 256 + t, q = i / (JOINT / 4), k = (i - q * (JOINT / 4)) * 4;
                        v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (q < n_act) v[u] = __ldcg(reinterpret_cast<const float4*>(d.hid + (size_t)sm.act[q] * JOINT + k));
                    }
#pragma unroll
                    for (int u = 0; u < 8; u++) {
                        const int i = i0 + u * 256 + t;
                        if (i < n_pad * (JOINT / 4)) *reinterpret_cast<float4*>(&sm.xs[(size_t)i * 4]) = v[u];
                    }
                }
                __syncthreads();
                unsigned long long u1 = tr ? now() : 0;
                if (n_act <= DK_TS) {
                    // few pending items (the tail of a chunk, or a batch of a few streams): a warp per row instead of one warp for
                    // all rows; same per-row arithmetic, rows merged in index order like below
                    if (warp < nr) {
                        const float* w1[1] = {&sm.wrow[warp][0]};
                        float v1[1][2];
                        dk_block<JOINT, JOINT, 1>(w1, 1, sm.xs, v1);
                        if ((lane & 7) == 0) {
                            sm.gs[2 * (lane >> 3)][warp] = v1[0][0] + sm.bias[warp];
                            sm.gs[2 * (lane >> 3) + 1][warp] = v1[0][1] + sm.bias[warp];
                        }
                    }
                    __syncthreads();
                    if (t < n_act) {
                        float best = -INFINITY, second = -INFINITY;
                        int besti = 0;
                        for (int i = 0; i < nr; i++) {
                            const float x = sm.gs[t][i];
                            if (x > best) { second = best; best = x; besti = v0 + i; }
                            else if (x > second) second = x;
                        }
                        partial[(size_t)sm.act[t] * G + cta] = make_float4(best, second, __int_as_float(besti), 0.0f);
                    }
                } else if (8 * warp < n_act) {  // warp w: streams 8w .. 8w+7 of the group, all of my rows
                    const float* wr[DK_MAXR];
#pragma unroll
                    for (int i = 0; i < DK_MAXR; i++) wr[i] = &sm.wrow[i < nr ? i : 0][0];
                    float v[DK_MAXR][2];
                    dk_block<JOINT, JOINT, DK_MAXR>(wr, nr, sm.xs + (size_t)(8 * warp) * JOINT, v);
                    if ((lane & 7) == 0) {
#pragma unroll
                        for (int j = 0; j < 2; j++) {
                            const int q = 8 * warp + 2 * (lane >> 3) + j;
                            if (q < n_act) {
                                float best = -INFINITY, second = -INFINITY;
                                int besti = 0;
#pragma unroll
                                for (int i = 0; i < DK_MAXR; i++) {
                                    if (i < nr) {
                                        const float x = v[i][j] + sm.bias[i];
                                        if (x > best) { second = best; best = x; besti = v0 + i; }
                                        else if (x > second) second = x;
                                    }
                                }
                                partial[(size_t)sm.act[q] * G + cta] = make_float4(best, second, __int_as_float(besti), 0.0f);
                            }
                        }
                    }
                }
                if (tr) { acc[4] += u0 - t0; acc[5] += u1 - u0; acc[6] += now() - u1; }
            }
        }
        unsigned long long t1 = tr ? now() : 0;
        grid_barrier(d.bar, phase);
        unsigned long long t2 = tr ? now() : 0;
        // ---- A: merge the partial maxima, state machine, next hid; one warp per stream ----
        if (W > 1) {
            // window mode: a CTA per stream; its 8 warps merge the window's items, thread 0 runs the state machine over them in frame order
            for (int b = cta; b < d.B; b += G) {
                const int nvalid = d.nvalid[b], slot = d.slot[b];
                const int fp = __ldcg(d.fp + b);
                if (fp >= nvalid) continue;  // uniform
                const int nw = min(W, nvalid - fp);
                for (int wi = warp; wi < nw; wi += 8) {
                    float best = -INFINITY, second = -INFINITY;
                    int besti = 0x7fffffff;
                    for (int c = lane; c < n_part; c += 32) {
                        const float4 p = __ldcg(partial + (size_t)(b * W + wi) * G + c);
                        const int pi = __float_as_int(p.z);
                        if (p.x > best || (p.x == best && pi < besti)) { second = fmaxf(best, p.y); best = p.x; besti = pi; }
                        else second = fmaxf(second, p.x);
                    }
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) {
                        const float ob = __shfl_xor_sync(0xffffffffu, best, o);
                        const float os = __shfl_xor_sync(0xffffffffu, second, o);
                        const int oi = __shfl_xor_sync(0xffffffffu, besti, o);
                        if (ob > best || (ob == best && oi < besti)) { second = fmaxf(best, os); best = ob; besti = oi; }
                        else second = fmaxf(second, ob);
                    }
                    if (lane == 0) { sm.wbest[wi] = best; sm.wsecond[wi] = second; sm.widx[wi] = besti; }
                }
                __syncthreads();
                if (t == 0) {
                    int f = fp, sy = d.sym[b], emit = 0;
                    float mg = d.min_margin[b];
                    for (int wi = 0; wi < nw; wi++) {
                        mg = fminf(mg, sm.wbest[wi] - sm.wsecond[wi]);
                        const int besti = sm.widx[wi];
                        if (besti == BLANK) {
                            f++;
                            sy = 0;
                        } else {
                            const int n = d.tok_cnt[b];
                            d.tok_out[(size_t)b * d.T * MAX_SYMBOLS + n] = besti | (f << 16);
                            d.tok_cnt[b] = n + 1;
                            s.prev[slot] = besti;
                            emit = 1;
                            d.emit_list[atomicAdd(emit_count + cs, 1)] = b;
                            sy++;
                            if (sy >= MAX_SYMBOLS) { f++; sy = 0; }
                            break;  // the later items of the window were formed with the old prediction: discarded
                        }
                    }
                    d.min_margin[b] = mg;
                    d.fp[b] = f;
                    d.sym[b] = sy;
                    if (f < nvalid) atomicAdd(unfinished + cs, 1);
                    sm.wemit = emit;
                    sm.wf = f;
                }
                __syncthreads();
                const int emit = sm.wemit, f = sm.wf;
                if (emit) {  // commit the pending LSTM state; the new pending state and the next window's hid follow in phase E
                    for (int i = t; i < 2 * DEC; i += 256) {
                        s.h[(size_t)slot * 2 * DEC + i] = __ldcg(s.pend_h + (size_t)slot * 2 * DEC + i);
                        s.c[(size_t)slot * 2 * DEC + i] = __ldcg(s.pend_c + (size_t)slot * 2 * DEC + i);
                    }
                } else {  // all blank: same pending pred, next window
                    for (int wi = warp; wi < W && f + wi < nvalid; wi += 8) dk_stage_hid(s, d, b * W + wi, b, f + wi, slot, lane);
                }
                __syncthreads();
            }
        } else
        for (int b = cta * 8 + warp; b < d.B; b += G * 8) {
            const int nvalid = d.nvalid[b], slot = d.slot[b];
            const int fp = __ldcg(d.fp + b);
            if (fp >= nvalid) continue;
            float best = -INFINITY, second = -INFINITY;
            int besti = 0x7fffffff;
            for (int c = lane; c < n_part; c += 32) {
                const float4 p = __ldcg(partial + (size_t)b * G + c);
                const int pi = __float_as_int(p.z);
                if (p.x > best || (p.x == best && pi < besti)) { second = fmaxf(best, p.y); best = p.x; besti = pi; }
                else second = fmaxf(second, p.x);
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float ob = __shfl_xor_sync(0xffffffffu, best, o);
                const float os = __shfl_xor_sync(0xffffffffu, second, o);
                const int oi = __shfl_xor_sync(0xffffffffu, besti, o);
                if (ob > best || (ob == best && oi < besti)) { second = fmaxf(best, os); best = ob; besti = oi; }
                else second = fmaxf(second, ob);
            }
            int f = fp, emit = 0;
            if (lane == 0) {
                d.min_margin[b] = fminf(d.min_margin[b], best - second);
                int sy = d.sym[b];
                if (besti == BLANK) {
                    f++;
                    sy = 0;
                } else {
                    const int n = d.tok_cnt[b];
                    d.tok_out[(size_t)b * d.T * MAX_SYMBOLS + n] = besti | (fp << 16);  // token id | chunk-local frame it was emitted at
                    d.tok_cnt[b] = n + 1;
                    s.prev[slot] = besti;
                    emit = 1;
                    d.emit_list[atomicAdd(emit_count + cs, 1)] = b;  // order is irrelevant: the streams are independent
                    sy++;
                    if (sy >= MAX_SYMBOLS) { f++; sy = 0; }
                }
                d.fp[b] = f;
                d.sym[b] = sy;
                if (f < nvalid) atomicAdd(unfinished + cs, 1);
            }
            emit = __shfl_sync(0xffffffffu, emit, 0);
            f = __shfl_sync(0xffffffffu, f, 0);
            if (emit) {  // commit the pending LSTM state; the new pending state and hid follow in phase E
                for (int i = lane; i < 2 * DEC; i += 32) {
                    s.h[(size_t)slot * 2 * DEC + i] = __ldcg(s.pend_h + (size_t)slot * 2 * DEC + i);
                    s.c[(size_t)slot * 2 * DEC + i] = __ldcg(s.pend_c + (size_t)slot * 2 * DEC + i);
                }
            } else if (f < nvalid) {
                dk_stage_hid(s, d, b, b, f, slot, lane);  // blank: same pending pred, next frame
            }
        }
        unsigned long long t3 = tr ? now() : 0;
        grid_barrier(d.bar, phase);
        if (tr) { const unsigned long long t4 = now(); acc[0] += t1 - t0; acc[1] += t2 - t1; acc[2] += t3 - t2; acc[3] += t4 - t3; }
        const int E = __ldcg(emit_count + cs), U = __ldcg(unfinished + cs);
        if (E > 0) dk_lstm_update(w, s, d, sm, d.emit_list, E, phase);
        if (U == 0 || round + 1 >= d.max_rounds) {
            if (cta == 0 && t == 0) {
                d.counters[6] = U;          // streams left unfinished (non-zero only if the round cap was hit)
                d.counters[7] = round + 1;  // rounds run
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(stamps[1])::"memory");
            }
            break;
        }
    }
}

// The persistent kernel's CTAs meet at hand-rolled grid barriers, so all of them must be resident at once: it is launched
// cooperatively (the driver refuses a grid it cannot place whole, e.g. while a second model's decode kernel holds SMs of the
// same device, instead of letting it hang), and the grid is checked against the occupancy once per device.
// NB200_DECODE_NO_COOP=1 falls back to a plain programmatic-dependent launch (debugging only).
static int decode_grid() {
    static int g = 0;
    static PerDeviceOnce once;
    if (!g) {
        int dev = 0, sms = 0;
        CUDA_CHECK(cudaGetDevice(&dev));
        CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        g = std::min(sms, 148);
        if ((VOCAB + g - 1) / g > DK_MAXR || (DEC + g - 1) / g > DK_MAXU) throw Error("decode kernel: too few SMs for its row split");
    }
    if (once.first()) {
        CUDA_CHECK(cudaFuncSetAttribute(decode_persistent_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(DkSmem)));
        int dev = 0, sms = 0, per_sm = 0, coop = 0;
        CUDA_CHECK(cudaGetDevice(&dev));
        CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        CUDA_CHECK(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
        CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, decode_persistent_kernel, 256, sizeof(DkSmem)));
        if (!coop) throw Error("decode kernel: the device does not support cooperative launches");
        if (per_sm * sms < g) throw Error("decode kernel: its grid cannot be co-resident on this device");
    }
    return g;
}
static void launch_decode(cudaStream_t st, int g, const DecodeWeights& w, const DecodeState& s, const DecodeBatch& d, int refresh) {
    static const bool no_coop = getenv("NB200_DECODE_NO_COOP") != nullptr;
    if (no_coop) launch_k(decode_persistent_kernel, dim3(g), dim3(256), sizeof(DkSmem), st, w, s, d, refresh);
    else launch_coop(decode_persistent_kernel, dim3(g), dim3(256), sizeof(DkSmem), st, w, s, d, refresh);
}

int decode_partial_floats(int B) { return 4 * B * 148; }

// All greedy rounds of the batch's chunk in one launch (src/nemo-stream.cpp:1106-1118, 840-930).
void launch_decode_chunk(cudaStream_t st, const DecodeWeights& w, const DecodeState& s, const DecodeBatch& d) {
    const int g = decode_grid();
    CUDA_CHECK(cudaMemsetAsync(d.counters, 0, 32 * sizeof(int), st));  // counters, the barrier word, time stamps
    ProfScope ps(st, "decode_chunk", 2.0 * d.B * d.T * (double)VOCAB * JOINT, 0.0);
    launch_decode(st, g, w, s, d, 0);
    NB_KERNEL_CHECK();
}

// Recomputes the pending LSTM / joint.pred state of every stream of the batch from its committed state (create / reset).
void launch_decode_refresh(cudaStream_t st, const DecodeWeights& w, const DecodeState& s, const DecodeBatch& d) {
    const int g = decode_grid();
    CUDA_CHECK(cudaMemsetAsync(d.counters, 0, 9 * sizeof(int), st));
    launch_decode(st, g, w, s, d, 1);
    NB_KERNEL_CHECK();
}

}  // namespace nb

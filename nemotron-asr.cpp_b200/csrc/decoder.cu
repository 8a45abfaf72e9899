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
//   L  logits: CTA j owns vocabulary rows [j*R, j*R+R) of joint.out (kept in shared memory for the whole launch, pre-split).
//      The pending items (stream, window offset) are compacted once per round; for every group of up to 64 of them it stages
//      hid = relu(joint.enc(frame fp) + pending pred) (src/nemo-ggml.cpp:1215-1217) with cp.async, the 8 warps form the
//      [items x my rows] products as MMA tiles of 16 items (K split over the warps), and the slice's (first max, runner-up,
//      index) goes to partial[item][j].
//   -- barrier --
//   A  argmax: one warp per stream merges the G partials (ties -> lowest index = the reference's strict '>' scan,
//      src/nemo-stream.cpp:899-906) and runs its state machine (:907-926): blank -> next frame, state untouched;
//      token -> emit, prev = token, commit the pending LSTM state, join emit_list.
//   -- barrier --
//   E  only if some stream emitted (a blank leaves (prev, h, c) untouched, so the pending state stays valid): the
//      pending LSTM / joint.pred state of the streams in emit_list is recomputed: layer 0, layer 1, joint.pred, each
//      followed by a barrier (every hidden unit of the previous one is needed); CTA j owns 4-5 hidden units (= the i,f,
//      g,o gate rows of those units, src/nemo-ggml.cpp:601-618) and loops over the emitting streams in tiles of 16. Layer 0
//      multiplies only its recurrent half: W_ih0 . embed[prev] + b is a per-token table built at load (DecodeWeights::t0).
// Window mode (DecodeBatch::W > 1: 3 in the streaming step, 32 in the offline path where one utterance is the whole batch): a blank leaves the
// decoder state untouched, so the logits of the next W frames can all be formed with the same pending prediction; a round
// evaluates W consecutive frames of the stream as W items, the argmax phase walks them in order up to the first non-blank
// one (everything before it is a blank step, everything after it is discarded), and a round advances up to W frames instead
// of one. Same steps, same arithmetic per step, same tokens and frame stamps as W = 1.
// The loop ends when no stream has frames left. Counters rotate over three slots so that clearing the next round's slot
// cannot race with a CTA still reading the previous one.
// The dot products of both phases run on the tensor cores in split precision (mma.sync.m16n8k16, fp16 operands, fp32
// accumulation): every fp32 operand is hi + lo / 2048 with hi = fp16(x), lo = fp16((x - hi) * 2048), and hi*hi + (hi*lo + lo*hi) /
// 2048 is accumulated in fp32 (weights pre-split at load, activations split in the loop); the dropped lo*lo term and the
// rounding of the lo parts are ~2^-22 relative per product, the same class as the fp32 summation-order noise of a CUDA-core
// dot product (3xTF32 on m16n8k8 gave the same results at twice the MMA count: the legacy tensor path is the limiter). The
// reference keeps these weights and this math in F32 (src/nemo-ggml.cpp:580-619,1186-1224). M = 16 items (streams / symbol steps), N = 8 weight rows per tile; a CTA's 8 warps
// split K and their partial tiles are summed in warp order through shared memory (fixed order: deterministic).
// The CUDA-core version this replaces was bound by shared-memory bandwidth (every warp re-read the whole item tile for 3
// weight rows): 9.7 us of dot products per round in the logits phase and ~1.5 us per 8 emitting streams and LSTM layer.
constexpr int DK_TS = 16, DK_MAXU = 5, DK_MAXR = 8, DK_GROUP = 64, DK_SCAN = 256;
constexpr int DK_LDH = JOINT + 16;     // row stride (floats) of hid vectors / joint.out rows in shared memory
constexpr int DK_LDX = 2 * DEC + 16;   // row stride of LSTM input vectors / gate rows; both = 16 mod 32: the 128-bit fragment
                                       // loads of a quarter warp (2 rows x 4 x 16 B) then cover all 32 banks once
struct DkSmem {
    float xs[(DK_TS + 4 * DK_MAXU) * DK_LDX];  // L: hid of up to 64 items [q][DK_LDH]; E: inputs of 16 streams [q][DK_LDX], then this CTA's gate rows [r][DK_LDX]
    float wrow[DK_MAXR][DK_LDH];       // this CTA's joint.out rows (rows >= nr are zero)
    float red[8][16][24];              // per-warp partial tiles [item][row]
    float gs[DK_TS][4 * DK_MAXU];      // gate pre-activations of the tile
    int sl[DK_TS], bq[DK_TS], tok[DK_TS], fpq[DK_TS], nvq[DK_TS], rows[4 * DK_MAXU];
    float bias[DK_MAXR];               // joint.out bias of my rows
    int act[DK_SCAN];                  // L: item indices (stream * W + window offset) of the pending symbol steps of up to 256 item slots
    unsigned cmask[8];                 // L: per-warp ballot of the scan
    float wbest[DK_GROUP], wsecond[DK_GROUP];  // A, window mode: merged (max, runner-up, index) of a stream's window items
    int widx[DK_GROUP], wemit, wf;
};
static_assert((DK_TS + 4 * DK_MAXU) * DK_LDX >= DK_GROUP * DK_LDH, "the hid group shares the LSTM tile buffer");
static_assert(sizeof(DkSmem) <= 227 * 1024, "decode kernel shared memory");

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

// hi / lo halves of 4 consecutive fp32 values: hi = fp16(x), lo = fp16((x - hi) * 2048). x = hi + lo / 2048 to ~2^-22 relative
// (the scale keeps lo a normal fp16 number wherever hi is one).
__device__ __forceinline__ void f16_split4(const float4& v, uint32_t& h01, uint32_t& h23, uint32_t& l01, uint32_t& l23) {
    const __half2 a = __floats2half2_rn(v.x, v.y), b = __floats2half2_rn(v.z, v.w);
    const float2 fa = __half22float2(a), fb = __half22float2(b);
    const __half2 la = __floats2half2_rn((v.x - fa.x) * 2048.0f, (v.y - fa.y) * 2048.0f);
    const __half2 lb = __floats2half2_rn((v.z - fb.x) * 2048.0f, (v.w - fb.y) * 2048.0f);
    h01 = *reinterpret_cast<const uint32_t*>(&a);
    h23 = *reinterpret_cast<const uint32_t*>(&b);
    l01 = *reinterpret_cast<const uint32_t*>(&la);
    l23 = *reinterpret_cast<const uint32_t*>(&lb);
}
__device__ __forceinline__ void mma_f16(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// One warp: out[m][n] = sum_{k in [k0, k1)} xs[m * ldx + k] * W[row(n)][k] for m < 16, n < 8 NT, row(n) = n < nrows ? n : 0
// (the padding columns repeat row 0 and are never read back). xs: fp32; ws: the rows in the pre-split format, row r at
// ws + r * ldw floats. (k1 - k0) % 16 == 0; both strides = 16 mod 32 floats.
// K runs through the MMAs in a permuted order: per 16 consecutive k the lane with tig = lane & 3 takes k0 + 4 tig .. + 3 of its
// two A rows and of its B row (one 128-bit load each) as the fragment's logical k = 2 tig, 2 tig + 1, 2 tig + 8, 2 tig + 9 —
// the same permutation for both operands, so the products pair up correctly.
// x . w = hi hi + (hi lo' + lo' hi) / 2048: the main and the correction terms have their own fp32 accumulators.
template <int NT>
__device__ __forceinline__ void dk_mma_tile(const float* xs, int ldx, const float* ws, int ldw, int nrows, int k0, int k1, float (*out)[24]) {
    const int lane = threadIdx.x & 31, g = lane >> 2, tig = lane & 3;
    const float* xa = xs + (size_t)g * ldx + 4 * tig;
    const float* xb = xa + (size_t)8 * ldx;
    const float* wp[NT];
#pragma unroll
    for (int nt = 0; nt < NT; nt++) {
        const int n = nt * 8 + g;
        wp[nt] = ws + (size_t)(n < nrows ? n : 0) * ldw + 4 * tig;
    }
    float acc[NT][4], cor[NT][4];
#pragma unroll
    for (int nt = 0; nt < NT; nt++)
#pragma unroll
        for (int i = 0; i < 4; i++) acc[nt][i] = cor[nt][i] = 0.0f;
#pragma unroll 2
    for (int k = k0; k < k1; k += 16) {
        const float4 A0 = *reinterpret_cast<const float4*>(xa + k), A1 = *reinterpret_cast<const float4*>(xb + k);
        uint32_t ah[4], al[4];  // (a0, a1, a2, a3) = (row g k-lo, row g+8 k-lo, row g k-hi, row g+8 k-hi)
        f16_split4(A0, ah[0], ah[2], al[0], al[2]);
        f16_split4(A1, ah[1], ah[3], al[1], al[3]);
#pragma unroll
        for (int nt = 0; nt < NT; nt++) {
            const uint4 Wv = *reinterpret_cast<const uint4*>(wp[nt] + k);  // {hi b0, hi b1, lo b0, lo b1}
            mma_f16(cor[nt], al[0], al[1], al[2], al[3], Wv.x, Wv.y);
            mma_f16(cor[nt], ah[0], ah[1], ah[2], ah[3], Wv.z, Wv.w);
            mma_f16(acc[nt], ah[0], ah[1], ah[2], ah[3], Wv.x, Wv.y);
        }
    }
#pragma unroll
    for (int nt = 0; nt < NT; nt++) {
        *reinterpret_cast<float2*>(&out[g][nt * 8 + 2 * tig]) =
            make_float2(fmaf(cor[nt][0], 1.0f / 2048.0f, acc[nt][0]), fmaf(cor[nt][1], 1.0f / 2048.0f, acc[nt][1]));
        *reinterpret_cast<float2*>(&out[g + 8][nt * 8 + 2 * tig]) =
            make_float2(fmaf(cor[nt][2], 1.0f / 2048.0f, acc[nt][2]), fmaf(cor[nt][3], 1.0f / 2048.0f, acc[nt][3]));
    }
}

// cp.async copy of rows[0..nrows) of W (K floats each) to dst[r * ld ..] by the whole CTA, one commit group
__device__ __forceinline__ void dk_stage_rows(float* dst, int ld, const float* __restrict__ W, int K, const int* rows, int nrows) {
    const int per_row = K / 4;
    for (int i = threadIdx.x; i < nrows * per_row; i += 256) {
        const int r = i / per_row, c = i - r * per_row;
        const uint32_t d = (uint32_t)__cvta_generic_to_shared(dst + (size_t)r * ld + 4 * c);
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(W + (size_t)rows[r] * K + 4 * c) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
}
__device__ __forceinline__ void cp_async16(float* dst_smem, const float* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void dk_stage_wait() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// The tile's 16 x nrows outputs: all 8 warps split K (K / 8 each), then thread (q, r) sums the 8 partial tiles in warp order.
template <int K, typename Out>
__device__ __forceinline__ void dk_tile_rows(DkSmem& sm, const float* xs, int ldx, const float* ws, int ldw, int nrows, Out&& out) {
    const int warp = threadIdx.x >> 5, t = threadIdx.x;
    static_assert(K % 128 == 0, "8 warps x 16-wide k steps");
    if (nrows > 16) dk_mma_tile<3>(xs, ldx, ws, ldw, nrows, warp * (K / 8), (warp + 1) * (K / 8), sm.red[warp]);
    else if (nrows > 8) dk_mma_tile<2>(xs, ldx, ws, ldw, nrows, warp * (K / 8), (warp + 1) * (K / 8), sm.red[warp]);
    else dk_mma_tile<1>(xs, ldx, ws, ldw, nrows, warp * (K / 8), (warp + 1) * (K / 8), sm.red[warp]);
    __syncthreads();
    for (int i = t; i < DK_TS * nrows; i += 256) {
        const int q = i / nrows, r = i - q * nrows;
        float v = 0.0f;
#pragma unroll
        for (int w = 0; w < 8; w++) v += sm.red[w][q][r];
        out(q, r, v);
    }
}

// Item q of an LSTM tile: (batch index, slot, previous token) of emitting stream e0 + q — one 16-byte record written by the
// argmax phase: (batch index, slot, token | next frame << 16, frames of the chunk) (no dependent slot / prev / frame look-ups) — or, for a refresh of streams 0..E-1, from the batch arrays.
__device__ __forceinline__ void dk_tile_item(const DecodeState& s, const DecodeBatch& d, DkSmem& sm, const int4* list, int e0, int E, int q) {
    int b = -1, slot = -1, tok = 0, fp = 0, nv = 0;
    if (e0 + q < E) {
        if (list) {
            const int4 r = __ldcg(list + e0 + q);
            b = r.x; slot = r.y; tok = r.z & 0xffff; fp = r.z >> 16; nv = r.w;
        } else {
            b = e0 + q;
            slot = d.slot[b];
            tok = __ldcg(s.prev + slot);
        }
    }
    sm.bq[q] = b;
    sm.sl[q] = slot;
    sm.tok[q] = tok;
    sm.fpq[q] = fp;  // the stream's frame pointer after the emission and its frame count: joint.pred stages the next hid from them
    sm.nvq[q] = nv;
}

// Recomputes pend_h / pend_c / pend_p of the E streams listed in `list` (batch indices; nullptr = 0..E-1), 16 streams per tile.
// The CTA's weight rows of a phase (16-20 gate rows x 1280 floats = 100 KB per LSTM layer, 4-5 x 640 for joint.pred) are brought
// into shared memory behind the input tile with cp.async by all 256 threads, the next phase's rows while this phase finishes
// its gate math and waits at the grid barrier.
__device__ void dk_lstm_update(const DecodeWeights& w, const DecodeState& s, const DecodeBatch& d, DkSmem& sm, const int4* list, int E,
                               unsigned& phase, unsigned long long* acc = nullptr) {
    const int cta = blockIdx.x, G = gridDim.x, t = threadIdx.x;
    auto now = [] { unsigned long long x; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(x)::"memory"); return x; };
    const bool tr = acc != nullptr && cta == 0 && t == 0;
    unsigned long long q0 = tr ? now() : 0, q1;
    const int base = DEC / G, rem = DEC % G;
    const int nu = base + (cta < rem ? 1 : 0), u0 = cta * base + min(cta, rem);  // my hidden units [u0, u0 + nu)
    float* wsm = sm.xs + DK_TS * DK_LDX;
    __syncthreads();
    if (t < 4 * nu) sm.rows[t] = (t / nu) * DEC + u0 + (t % nu);  // the same gate rows in both layers; rows[0..nu) = u0.. are also joint.pred's
    __syncthreads();
    dk_stage_rows(wsm, DK_LDH, w.whh0_s, DEC, sm.rows, 4 * nu);
    for (int layer = 0; layer < 2; layer++) {
        for (int e0 = 0; e0 < E; e0 += DK_TS) {
            __syncthreads();
            if (t < DK_TS) dk_tile_item(s, d, sm, list, e0, E, t);
            __syncthreads();
            // layer 0: x = h0 (the input half W_ih0 . embed[prev] + b is the token's row of the table w.t0); layer 1: x = [h0' | h1]
            // (h0' = this round's new layer-0 output). Every 16-byte piece of a thread is in flight at once (cp.async; L2 only, the
            // state was written by other CTAs of this launch)
            const int kx = layer == 0 ? DEC : 2 * DEC;
            for (int i = t; i < DK_TS * (kx / 4); i += 256) {
                const int q = i / (kx / 4), k = (i - q * (kx / 4)) * 4, slot = sm.sl[q];
                float* dst = &sm.xs[q * DK_LDX + k];
                if (slot >= 0) {
                    const float* src;
                    if (layer == 0) src = s.h + (size_t)slot * 2 * DEC + k;
                    else if (k < DEC) src = s.pend_h + (size_t)slot * 2 * DEC + k;
                    else src = s.h + (size_t)slot * 2 * DEC + DEC + (k - DEC);
                    cp_async16(dst, src);
                }  // absent streams' rows are left as they are (their output rows are not read)
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
            dk_stage_wait();  // this layer's rows (requested before the previous barrier) and the input tile, my copies ...
            __syncthreads();  // ... and everyone's, and the input tile
            if (tr) { q1 = now(); acc[7] += q1 - q0; q0 = q1; }
            float c_old = 0.0f;  // the gate threads' cell state, requested now so that it is there when the MMAs are done
            if (t < DK_TS * nu && sm.sl[t / nu] >= 0) c_old = __ldcg(s.c + (size_t)sm.sl[t / nu] * 2 * DEC + (size_t)layer * DEC + u0 + t % nu);
            if (layer == 0)
                dk_tile_rows<DEC>(sm, sm.xs, DK_LDX, wsm, DK_LDH, 4 * nu,
                                  [&](int q, int r, float v) { sm.gs[q][r] = sm.sl[q] >= 0 ? v + __ldg(w.t0 + (size_t)sm.tok[q] * 4 * DEC + sm.rows[r]) : 0.0f; });
            else
                dk_tile_rows<2 * DEC>(sm, sm.xs, DK_LDX, wsm, DK_LDX, 4 * nu, [&](int q, int r, float v) { sm.gs[q][r] = v + w.bcat[1][sm.rows[r]]; });
            __syncthreads();
            if (tr) { q1 = now(); acc[8] += q1 - q0; q0 = q1; }
            if (e0 + DK_TS >= E) {  // last tile: the rows are dead, fetch the next phase's while the gates and the barrier run
                if (layer == 0) dk_stage_rows(wsm, DK_LDX, w.wcat_s[1], 2 * DEC, sm.rows, 4 * nu);
                else dk_stage_rows(wsm, DK_LDH, w.wpred_s, DEC, sm.rows, nu);
            }
            if (t < DK_TS * nu) {
                const int q = t / nu, ui = t % nu, slot = sm.sl[q];
                if (slot >= 0) {
                    const size_t so = (size_t)slot * 2 * DEC + (size_t)layer * DEC + u0 + ui;
                    const float ig = sigmoid_f(sm.gs[q][ui]), fg = sigmoid_f(sm.gs[q][nu + ui]);
                    const float gg = tanhf(sm.gs[q][2 * nu + ui]), og = sigmoid_f(sm.gs[q][3 * nu + ui]);
                    const float c2 = fg * c_old + ig * gg;
                    s.pend_c[so] = c2;
                    s.pend_h[so] = og * tanhf(c2);
                }
            }
            if (tr) { q1 = now(); acc[9] += q1 - q0; q0 = q1; }
        }
        grid_barrier(d.bar, phase);
        if (tr) { q1 = now(); acc[10] += q1 - q0; q0 = q1; }
    }
    // joint.pred (src/nemo-ggml.cpp:1207-1211): pend_p = W_pred h1' + b_pred, rows u0 .. u0+nu-1
    if (t < nu) sm.rows[t] = u0 + t;
    for (int e0 = 0; e0 < E; e0 += DK_TS) {
        __syncthreads();
        if (t < DK_TS) dk_tile_item(s, d, sm, list, e0, E, t);
        __syncthreads();
        for (int i = t; i < DK_TS * (DEC / 4); i += 256) {
            const int q = i / (DEC / 4), k = (i - q * (DEC / 4)) * 4, slot = sm.sl[q];
            float* dst = &sm.xs[q * DK_LDX + k];
            if (slot >= 0) cp_async16(dst, s.pend_h + (size_t)slot * 2 * DEC + DEC + k);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        dk_stage_wait();
        __syncthreads();
        if (tr) { q1 = now(); acc[11] += q1 - q0; q0 = q1; }
        dk_tile_rows<DEC>(sm, sm.xs, DK_LDX, wsm, DK_LDH, nu, [&](int q, int r, float acc) {
            if (sm.sl[q] < 0) return;
            const float v = acc + w.bpred[u0 + r];
            s.pend_p[(size_t)sm.sl[q] * JOINT + u0 + r] = v;
            if (list) {  // the stream's next symbol step: hid = relu(joint.enc(frame fp) + new pending pred), my rows
                const int b = sm.bq[q], fp = sm.fpq[q], n_w = min(d.W, sm.nvq[q] - fp);
                for (int w0 = 0; w0 < n_w; w0 += 4) {  // loads before stores (the pointers may alias as far as the compiler knows)
                    float ev[4];
#pragma unroll
                    for (int j = 0; j < 4; j++)
                        if (w0 + j < n_w) ev[j] = d.e[((size_t)(b * d.T + fp + w0 + j)) * JOINT + u0 + r];
#pragma unroll
                    for (int j = 0; j < 4; j++)
                        if (w0 + j < n_w) d.hid[(size_t)(b * d.W + w0 + j) * JOINT + u0 + r] = fmaxf(ev[j] + v, 0.0f);
                }
            }
        });
        if (tr) { q1 = now(); acc[12] += q1 - q0; q0 = q1; }
    }
    grid_barrier(d.bar, phase);
    if (tr) { q1 = now(); acc[13] += q1 - q0; acc[14] += 1; acc[15] += (unsigned long long)E; }
}

// Commit of an emission: (h, c) = (pend_h, pend_c) of one slot (2 x 1280 floats) by NT threads. All of a thread's loads are
// issued before its stores: a load / store pair per iteration serialises on the L2 round trip (the pointers may alias as far as
// the compiler knows) — 40 dependent round trips per warp, which made the streams that emit hold the round's barrier for ~20 us.
template <int NT>
__device__ __forceinline__ void dk_commit_state(const DecodeState& s, int slot, int tid) {
    constexpr int N4 = 2 * DEC / 4, IT = (N4 + NT - 1) / NT;
    const float4* src[2] = {reinterpret_cast<const float4*>(s.pend_h + (size_t)slot * 2 * DEC), reinterpret_cast<const float4*>(s.pend_c + (size_t)slot * 2 * DEC)};
    float4* dst[2] = {reinterpret_cast<float4*>(s.h + (size_t)slot * 2 * DEC), reinterpret_cast<float4*>(s.c + (size_t)slot * 2 * DEC)};
#pragma unroll
    for (int a = 0; a < 2; a++) {
        float4 v[IT];
#pragma unroll
        for (int i = 0; i < IT; i++)
            if (tid + i * NT < N4) v[i] = __ldcg(src[a] + tid + i * NT);
#pragma unroll
        for (int i = 0; i < IT; i++)
            if (tid + i * NT < N4) dst[a][tid + i * NT] = v[i];
    }
}

// hid[b] = relu(joint.enc(frame fp) + pend_p[slot]) for one stream, by one warp (src/nemo-ggml.cpp:1215-1217)
// (item = b * W + window offset; W = 1: item = b)
__device__ __forceinline__ void dk_stage_hid(const DecodeState& s, const DecodeBatch& d, int item, int b, int fp, int slot, int lane) {
    constexpr int IT = JOINT / 128;  // 5 x (32 lanes x 4 floats); all loads are issued before the stores (see dk_commit_state)
    float4 e[IT], p[IT];
#pragma unroll
    for (int i = 0; i < IT; i++) {
        const int k = 4 * lane + 128 * i;
        e[i] = *reinterpret_cast<const float4*>(d.e + ((size_t)(b * d.T + fp)) * JOINT + k);
        p[i] = __ldcg(reinterpret_cast<const float4*>(s.pend_p + (size_t)slot * JOINT + k));
    }
#pragma unroll
    for (int i = 0; i < IT; i++) {
        const int k = 4 * lane + 128 * i;
        *reinterpret_cast<float4*>(d.hid + (size_t)item * JOINT + k) =
            make_float4(fmaxf(e[i].x + p[i].x, 0.f), fmaxf(e[i].y + p[i].y, 0.f), fmaxf(e[i].z + p[i].z, 0.f), fmaxf(e[i].w + p[i].w, 0.f));
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
    for (int i = t; i < DK_MAXR * (JOINT / 4); i += 256) {  // rows >= nr: zero (padding columns of the MMA tile)
        const int r = i / (JOINT / 4), k = (i - r * (JOINT / 4)) * 4;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (r < nr) v = *reinterpret_cast<const float4*>(w.wout_s + (size_t)(v0 + r) * JOINT + k);  // pre-split halves; zero bits = zero
        *reinterpret_cast<float4*>(&sm.wrow[r][k]) = v;
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
    for (int b = cta + G * warp; b < d.B; b += G * 8)
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
        unsigned long long tg = t0;
        if (nr > 0) {
            // The pending items (stream, window offset) of up to 256 consecutive item slots are compacted at once — warp w scans 32 of
            // them, one L2 round trip for the whole round — and then processed 64 at a time, so late rounds with a few unfinished
            // streams scattered over the batch are one pass, not one per 64 slots.
            for (int s0 = 0; s0 < d.B * W; s0 += DK_SCAN)
            {
                __syncthreads();
                const int it_ = s0 + 32 * warp + lane, b_ = it_ / W, wi_ = it_ - b_ * W;
                const bool on_ = it_ < d.B * W && __ldcg(d.fp + b_) + wi_ < d.nvalid[b_];
                const unsigned m_ = __ballot_sync(0xffffffffu, on_);
                if (lane == 0) sm.cmask[warp] = m_;
                __syncthreads();
                int base_ = 0, total = 0;
#pragma unroll
                for (int w2 = 0; w2 < 8; w2++) {
                    const int c2 = __popc(sm.cmask[w2]);
                    if (w2 < warp) base_ += c2;
                    total += c2;
                }
                if (on_) sm.act[base_ + __popc(m_ & ((1u << lane) - 1u))] = it_;
            for (int a0 = 0; a0 < total; a0 += DK_GROUP) {
                __syncthreads();  // the list is complete; the previous group's tiles and partial sums are consumed
                const int n_act = min(DK_GROUP, total - a0);
                const int* act = sm.act + a0;
                unsigned long long u0 = tr ? now() : 0;
                const int n_mt = (n_act + 15) >> 4;                      // 16-item MMA tiles in use
                const int mtp = n_mt <= 1 ? 1 : (n_mt <= 2 ? 2 : 4);      // warps are split (item tile, K slice): 1x8, 2x4 or 4x2
                for (int i = t; i < n_act * (JOINT / 4); i += 256) {  // all of a thread's 16-byte pieces in flight at once (L2 only)
                    const int q = i / (JOINT / 4), k = (i - q * (JOINT / 4)) * 4;
                    float* dst = &sm.xs[(size_t)q * DK_LDH + k];
                    if (q < n_act) cp_async16(dst, d.hid + (size_t)act[q] * JOINT + k);  // rows >= n_act keep whatever the buffer holds: an MMA
                                                                                         // output row depends on its own A row only, and those rows are never read back
                }
                asm volatile("cp.async.commit_group;" ::: "memory");
                dk_stage_wait();
                __syncthreads();
                unsigned long long u1 = tr ? now() : 0;
                {
                    const int mt = warp % mtp, ks = warp / mtp, klen = JOINT / (8 / mtp);  // 80 / 160 / 320: multiples of 16
                    if (mt < n_mt) dk_mma_tile<1>(sm.xs + (size_t)(16 * mt) * DK_LDH, DK_LDH, &sm.wrow[0][0], DK_LDH, DK_MAXR, ks * klen, (ks + 1) * klen, sm.red[warp]);
                    __syncthreads();
                    if (t < n_act) {  // my item's logits on this CTA's rows: K slices summed in order, then the slice's (first max, runner-up)
                        const int q = t & 15, m_t = t >> 4;
                        float best = -INFINITY, second = -INFINITY;
                        int besti = 0;
                        for (int i = 0; i < nr; i++) {
                            float x = 0.0f;
                            for (int k2 = 0; k2 < 8 / mtp; k2++) x += sm.red[k2 * mtp + m_t][q][i];
                            x += sm.bias[i];
                            if (x > best) { second = best; best = x; besti = v0 + i; }
                            else if (x > second) second = x;
                        }
                        partial[(size_t)act[t] * G + cta] = make_float4(best, second, __int_as_float(besti), 0.0f);
                    }
                }
                if (tr) { const unsigned long long u2 = now(); acc[4] += u0 - tg; acc[5] += u1 - u0; acc[6] += u2 - u1; tg = u2; }
            }
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
                int sy0 = 0, tc0 = 0;  // the state machine's counters, requested beside the partial maxima (one round trip, not three)
                float mg0 = 0.0f;
                if (t == 0) { sy0 = d.sym[b]; mg0 = d.min_margin[b]; tc0 = d.tok_cnt[b]; }
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
                    int f = fp, sy = sy0, emit = 0;
                    float mg = mg0;
                    for (int wi = 0; wi < nw; wi++) {
                        mg = fminf(mg, sm.wbest[wi] - sm.wsecond[wi]);
                        const int besti = sm.widx[wi];
                        if (besti == BLANK) {
                            f++;
                            sy = 0;
                        } else {
                            const int n = tc0;
                            d.tok_out[(size_t)b * d.T * MAX_SYMBOLS + n] = besti | (f << 16);
                            d.tok_cnt[b] = n + 1;
                            s.prev[slot] = besti;
                            emit = 1;
                            sy++;
                            if (sy >= MAX_SYMBOLS) { f++; sy = 0; }
                            d.emit_list[atomicAdd(emit_count + cs, 1)] = make_int4(b, slot, besti | (f << 16), nvalid);
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
                    dk_commit_state<256>(s, slot, t);
                } else {  // all blank: same pending pred, next window
                    for (int wi = warp; wi < W && f + wi < nvalid; wi += 8) dk_stage_hid(s, d, b * W + wi, b, f + wi, slot, lane);
                }
                __syncthreads();
            }
        } else
        for (int b = cta + G * warp; b < d.B; b += G * 8) {  // a warp per stream, the streams spread over all CTAs
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
                    sy++;
                    if (sy >= MAX_SYMBOLS) { f++; sy = 0; }
                    d.emit_list[atomicAdd(emit_count + cs, 1)] = make_int4(b, slot, besti | (f << 16), nvalid);  // order is irrelevant: the streams are independent
                }
                d.fp[b] = f;
                d.sym[b] = sy;
                if (f < nvalid) atomicAdd(unfinished + cs, 1);
            }
            emit = __shfl_sync(0xffffffffu, emit, 0);
            f = __shfl_sync(0xffffffffu, f, 0);
            if (emit) {  // commit the pending LSTM state; the new pending state and hid follow in phase E
                dk_commit_state<32>(s, slot, lane);
            } else if (f < nvalid) {
                dk_stage_hid(s, d, b, b, f, slot, lane);  // blank: same pending pred, next frame
            }
        }
        unsigned long long t3 = tr ? now() : 0;
        grid_barrier(d.bar, phase);
        if (tr) { const unsigned long long t4 = now(); acc[0] += t1 - t0; acc[1] += t2 - t1; acc[2] += t3 - t2; acc[3] += t4 - t3; }
        const int E = __ldcg(emit_count + cs), U = __ldcg(unfinished + cs);
        if (E > 0) dk_lstm_update(w, s, d, sm, d.emit_list, E, phase, acc);
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
    CUDA_CHECK(cudaMemsetAsync(d.counters, 0, 48 * sizeof(int), st));  // counters, the barrier word, time stamps, phase trace
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

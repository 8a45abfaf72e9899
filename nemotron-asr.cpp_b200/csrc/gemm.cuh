// gemm.cuh — C[M,N] = A[M,K] * W[N,K]^T for the encoder's dense contractions, fp16 operands, fp32 accumulation.
//
// Two kernels share one set of fused epilogues:
//   gemm_tc_kernel   tcgen05.mma (kind::f16, M=128 x N=BN x K=16 per instruction) with the accumulator in TMEM,
//                    operands staged by TMA (cp.async.bulk.tensor, 128-byte swizzle) through an mbarrier ring;
//                    warp 0 = TMA producer, warp 1 = MMA issuer + TMEM allocator, then 4 / 8 / 16 epilogue warps (tcgen05.ld) by tile width and epilogue class.
//   gemm_simt_kernel CUDA-core checker/debug path (NB200_FLAG_GEMM_SIMT), identical operands and epilogues.
//
// Epilogues (what the reference does as separate ggml nodes, src/nemo-stream.cpp:593-690):
//   EPI_STORE      out_f32 = act(acc + bias)                      (pre_encode pw3, pre_encode.out, joint.enc)
//   EPI_STORE_F16  out_f16 = fq(act(acc + bias))                  (FFN linear1 + SiLU, pre_encode pw6 + ReLU, q|k|v, linear_pos)
//   EPI_RESID      x += alpha * acc                               (FFN linear2 *0.5, attention out, pointwise_conv2)
//   EPI_GLU        out_f32[j] = a_j * sigmoid(b_j) with (a_j,b_j) interleaved as columns (2j,2j+1) at load
//   EPI_PARTIAL    split-K: CTA z of gridDim.z sums its K range and stores the raw partial tile to
//                  out_f32 + z*split_stride; the LayerNorm that follows every residual GEMM adds alpha * sum_z partial_z
//                  to x in a fixed order (deterministic, unlike atomics). Used for the N=1024 GEMMs, which would
//                  otherwise run on 28-112 CTAs with tiles too small for the L2 -> SM bandwidth.
// fq = ggml-CPU's Q8_0 activation re-quantisation emulated per 32 consecutive features (common.cuh fake_quant).
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace nb {

enum EpiKind : int { EPI_STORE = 0, EPI_STORE_F16 = 1, EPI_RESID = 2, EPI_GLU = 4, EPI_PARTIAL = 5 };
enum ActKind : int { ACT_NONE = 0, ACT_RELU = 1, ACT_SILU = 2 };

struct EpiParams {
    int kind = EPI_STORE;
    int act = ACT_NONE;
    int act_quant = 0;          // EPI_STORE_F16: emulate Q8_0 re-quantisation of the output row blocks
    int out_bf16 = 0;           // EPI_STORE_F16: store bf16 (the output is the A operand of a bf16 GEMM)
    int ab_bf16 = 0;            // the operands A and W hold bf16 (kind::f16 with A = B = bf16; NB200_FLAG_BF16 models)
    float alpha = 1.0f;         // EPI_RESID
    const float* bias = nullptr;  // [N] or null
    float* out_f32 = nullptr;   // EPI_STORE / EPI_RESID / EPI_GLU / EPI_QKV(q)
    __half* out_f16 = nullptr;  // EPI_STORE_F16
    int ldo = 0;                // leading dimension of the output (elements)
    size_t split_stride = 0;    // EPI_PARTIAL: elements between the partial outputs of consecutive K splits (a multiple of ldo)
    unsigned long long* trace = nullptr;  // debug: per-CTA phase timestamps (globaltimer ns), 8 words per CTA
    // L2 prefetch of the NEXT GEMM's weights (they do not depend on anything this step computes): every CTA asks for its slice
    // while its own main loop runs — the tensor pipe is the limiter then and HBM is ~10 % busy — so the next kernel's operand
    // ring fills from L2 instead of ramping up an HBM stream behind a dependency edge. pf_slice is set by gemm().
    const void* pf_ptr = nullptr;
    size_t pf_bytes = 0;
    unsigned pf_slice = 0;
    // ... and of per-stream state the kernels behind this GEMM will read from HBM (the K / V ring slabs of the batch's streams for the
    // attention two kernels later): slab i = pf2_base + pf2_slot[i] * pf2_stride, pf2_bytes each, i < pf2_count; CTA c takes slabs
    // c, c + #CTAs, ...
    const void* pf2_base = nullptr;
    size_t pf2_stride = 0;
    unsigned pf2_bytes = 0;
    const int* pf2_slot = nullptr;
    int pf2_count = 0;
};

// ---- epilogue, stage 1 (registers): bias / activation / re-quantisation / GLU on 32 consecutive accumulator columns
// [n0, n0+32) of row m. Leaves the row's output words in w[] and returns how many there are: 32 (fp32 outputs) or 16
// (32 fp16 values, or 16 GLU outputs). `dst` receives the address of the row's first output byte of this chunk.
__device__ __forceinline__ int epilogue_transform(const EpiParams& p, int m, int n0, float (&acc)[32], uint32_t (&w)[32], char*& dst) {
    if (p.kind != EPI_RESID && p.kind != EPI_PARTIAL) {
        if (p.bias != nullptr) {
#pragma unroll
            for (int j = 0; j < 32; j++) acc[j] += __ldg(p.bias + n0 + j);
        }
        if (p.act == ACT_RELU) {
#pragma unroll
            for (int j = 0; j < 32; j++) acc[j] = fmaxf(acc[j], 0.0f);
        } else if (p.act == ACT_SILU) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) silu_fast4(acc[j], acc[j + 1], acc[j + 2], acc[j + 3]);
        }
    }
    if (p.kind == EPI_STORE_F16) {
        if (p.act_quant) {
            float amax = 0.0f;
#pragma unroll
            for (int j = 0; j < 32; j++) amax = fmaxf(amax, fabsf(acc[j]));
            float d, id;
            fq_scales(amax, d, id);
#pragma unroll
            // round-to-nearest-even of |v| <= 127 by the 1.5 * 2^23 trick (two FADDs on the FMA pipe; rintf is an FRND on the
            // quarter-rate conversion pipe, which this epilogue already saturates with ex2/rcp)
            for (int j = 0; j < 32; j++) acc[j] = __fsub_rn(__fadd_rn(__fmul_rn(acc[j], id), 12582912.0f), 12582912.0f) * d;
        }
#pragma unroll
        for (int j = 0; j < 16; j++) {
            w[j] = pack_operand2(acc[2 * j], acc[2 * j + 1], p.out_bf16);
        }
        dst = reinterpret_cast<char*>(p.out_f16 + (size_t)m * p.ldo + n0);
        return 16;
    }
    if (p.kind == EPI_GLU) {
#pragma unroll
        for (int j = 0; j < 16; j += 4) {
            float o[4];
            glu_fast4(o, acc[2 * j], acc[2 * j + 1], acc[2 * j + 2], acc[2 * j + 3], acc[2 * j + 4], acc[2 * j + 5], acc[2 * j + 6], acc[2 * j + 7]);
#pragma unroll
            for (int e = 0; e < 4; e++) w[j + e] = __float_as_uint(o[e]);
        }
        dst = reinterpret_cast<char*>(p.out_f32 + (size_t)m * p.ldo + (n0 >> 1));
        return 16;
    }
#pragma unroll
    for (int j = 0; j < 32; j++) w[j] = __float_as_uint(acc[j]);
    dst = reinterpret_cast<char*>(p.out_f32 + (size_t)m * p.ldo + n0);
    return 32;
}

// Epilogue classes the tcgen05 kernel is compiled for: with the kind / activation / bias fixed at compile time the other
// branches of epilogue_transform (predicated-off instructions that still take issue slots) disappear. EC_GENERIC keeps
// everything a run-time switch (subsampling GEMMs with bias / ReLU, anything else).
enum EpiClass : int { EC_GENERIC = 0, EC_SILU_F16 = 1, EC_PARTIAL = 2, EC_GLU = 3, EC_PLAIN_F16 = 4 };
__host__ __device__ inline int epi_class_of(const EpiParams& p) {
    if (p.kind == EPI_PARTIAL) return EC_PARTIAL;
    if (p.bias != nullptr) return EC_GENERIC;
    if (p.kind == EPI_STORE_F16 && p.act == ACT_SILU) return EC_SILU_F16;
    if (p.kind == EPI_STORE_F16 && p.act == ACT_NONE) return EC_PLAIN_F16;
    if (p.kind == EPI_GLU && p.act == ACT_NONE) return EC_GLU;
    return EC_GENERIC;
}
template <int EC>
__device__ __forceinline__ EpiParams epi_specialize(const EpiParams& in) {
    EpiParams p = in;
    if constexpr (EC == EC_SILU_F16) { p.kind = EPI_STORE_F16; p.act = ACT_SILU; p.bias = nullptr; }
    if constexpr (EC == EC_PARTIAL) { p.kind = EPI_PARTIAL; p.act = ACT_NONE; p.bias = nullptr; p.act_quant = 0; }
    if constexpr (EC == EC_GLU) { p.kind = EPI_GLU; p.act = ACT_NONE; p.bias = nullptr; p.act_quant = 0; }
    if constexpr (EC == EC_PLAIN_F16) { p.kind = EPI_STORE_F16; p.act = ACT_NONE; p.bias = nullptr; }
    return p;
}

// Direct (one thread = one row) version used by the CUDA-core checker kernel: same transform, per-thread stores.
__device__ __forceinline__ void epilogue_apply(const EpiParams& p, int m, int n0, float (&acc)[32]) {
    uint32_t w[32];
    char* dst;
    const int nw = epilogue_transform(p, m, n0, acc, w, dst);
    uint4* d4 = reinterpret_cast<uint4*>(dst);
    if (p.kind == EPI_RESID) {
#pragma unroll
        for (int j = 0; j < 8; j++) {
            float4 v = reinterpret_cast<float4*>(dst)[j];
            v.x += p.alpha * __uint_as_float(w[4 * j + 0]);
            v.y += p.alpha * __uint_as_float(w[4 * j + 1]);
            v.z += p.alpha * __uint_as_float(w[4 * j + 2]);
            v.w += p.alpha * __uint_as_float(w[4 * j + 3]);
            reinterpret_cast<float4*>(dst)[j] = v;
        }
        return;
    }
#pragma unroll
    for (int j = 0; j < 8; j++)
        if (4 * j < nw) d4[j] = make_uint4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
}

// One fp16 operand matrix [rows, K] (K contiguous) in HBM, with the TMA descriptors of its two tile heights.
struct Operand {
    const __half* ptr = nullptr;
    int rows = 0, K = 0;
    CUtensorMap maps[5];  // boxes of 64 (K) x {32, 64, 96, 128, 256} rows, 128B swizzle
    const CUtensorMap& box(int rows) const { return maps[rows == 32 ? 0 : (rows == 64 ? 1 : (rows == 96 ? 2 : (rows == 128 ? 3 : 4)))]; }
};
void operand_init(Operand& op, const __half* ptr, int rows, int K);  // builds the tensor maps

// C = A * W^T (+ epilogue). M valid rows of A; N = W.rows; K = A.K = W.K. use_simt selects the checker kernel.
// Returns the number of K splits used (1 unless epi.kind == EPI_PARTIAL; never more than max_splits).
int gemm(cudaStream_t st, const Operand& A, const Operand& W, int M, const EpiParams& epi, bool use_simt, int max_splits = 1);
// Tile / split configuration the tcgen05 path would pick (exposed for the benchmarks and DESIGN.md)
struct GemmConfig {
    int bn, splits, ctas;
    int cg = 1;  // 1: one CTA per 128 x bn tile; 2: a CTA pair (cta_group::2) per 256 x bn tile
};
GemmConfig gemm_pick_config(int M, int N, int K, int max_splits, int min_bn = 32);
// Forces a configuration for tuning runs: bn in {32,64,128,256} (single CTA) or 1000 + {128,192,256} (CTA pair), splits >= 1;
// bn = 0 restores the heuristic.
void gemm_force_config(int bn, int splits);


// Tensor-pipe peak and arithmetic check of one tcgen05.mma kind (measurement only; SURVEY section 8d asks for the int8 peak
// beside the fp16 one): every CTA keeps one 128 x 256 operand pair of K = 128 bytes resident in shared memory and issues
// `iters` k-blocks of 4 MMAs into the same TMEM accumulator. kind 0 = kind::f16 (fp16 operands, fp32 accumulate, K = 16 per
// MMA), 1 = kind::i8 (signed int8 operands, int32 accumulate, K = 32 per MMA). Returns the device time of the launch in ms;
// check_out (device, [128][256] 32-bit words, may be null) receives CTA 0's accumulator after ONE k-block of the operands
// a_bytes / b_bytes (device, [128][128] and [256][128] bytes, K-major rows of 128 bytes).
float mma_peak_run(cudaStream_t st, int kind, int ctas, int iters, const uint8_t* a_bytes, const uint8_t* b_bytes, uint32_t* check_out);

}  // namespace nb

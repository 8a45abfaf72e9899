// gemm.cuh — C[M,N] = A[M,K] * W[N,K]^T for the encoder's dense contractions, fp16 operands, fp32 accumulation.
//
// Two kernels share one set of fused epilogues:
//   gemm_tc_kernel   tcgen05.mma (kind::f16, M=128 x N=BN x K=16 per instruction) with the accumulator in TMEM,
//                    operands staged by TMA (cp.async.bulk.tensor, 128-byte swizzle) through an mbarrier ring;
//                    warp 0 = TMA producer, warp 1 = MMA issuer + TMEM allocator, warps 2-5 = epilogue (tcgen05.ld).
//   gemm_simt_kernel CUDA-core checker/debug path (NB200_FLAG_GEMM_SIMT), identical operands and epilogues.
//
// Epilogues (what the reference does as separate ggml nodes, src/nemo-stream.cpp:593-690):
//   EPI_STORE      out_f32 = act(acc + bias)                      (pre_encode pw3, pre_encode.out, joint.enc)
//   EPI_STORE_F16  out_f16 = fq(act(acc + bias))                  (FFN linear1 + SiLU, pre_encode pw6 + ReLU)
//   EPI_RESID      x += alpha * acc                               (FFN linear2 *0.5, attention out, pointwise_conv2)
//   EPI_QKV        q -> fp32 buffer, k/v -> the streams' K/V ring caches
//   EPI_GLU        out_f32[j] = a_j * sigmoid(b_j) with (a_j,b_j) interleaved as columns (2j,2j+1) at load
// fq = ggml-CPU's Q8_0 activation re-quantisation emulated per 32 consecutive features (common.cuh fake_quant).
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace nb {

enum EpiKind : int { EPI_STORE = 0, EPI_STORE_F16 = 1, EPI_RESID = 2, EPI_QKV = 3, EPI_GLU = 4 };
enum ActKind : int { ACT_NONE = 0, ACT_RELU = 1, ACT_SILU = 2 };

struct EpiParams {
    int kind = EPI_STORE;
    int act = ACT_NONE;
    int act_quant = 0;          // EPI_STORE_F16: emulate Q8_0 re-quantisation of the output row blocks
    float alpha = 1.0f;         // EPI_RESID
    const float* bias = nullptr;  // [N] or null
    float* out_f32 = nullptr;   // EPI_STORE / EPI_RESID / EPI_GLU / EPI_QKV(q)
    __half* out_f16 = nullptr;  // EPI_STORE_F16
    int ldo = 0;                // leading dimension of the output (elements)
    // EPI_QKV: row m = b*T + i belongs to batch item b; its stream slot and ring write position
    float* k_ring = nullptr;    // this layer's [slots][RING][1024]
    float* v_ring = nullptr;
    const int* slot = nullptr;  // [B]
    const int* wpos = nullptr;  // [B] ring index where frame 0 of this chunk goes
    int T = 1;
};

// Applies the epilogue to 32 consecutive accumulator columns [n0, n0+32) of row m.
__device__ __forceinline__ void epilogue_apply(const EpiParams& p, int m, int n0, float (&acc)[32]) {
    if (p.kind == EPI_RESID) {
        float4* dst = reinterpret_cast<float4*>(p.out_f32 + (size_t)m * p.ldo + n0);
#pragma unroll
        for (int j = 0; j < 8; j++) {
            float4 v = dst[j];
            v.x += p.alpha * acc[4 * j + 0];
            v.y += p.alpha * acc[4 * j + 1];
            v.z += p.alpha * acc[4 * j + 2];
            v.w += p.alpha * acc[4 * j + 3];
            dst[j] = v;
        }
        return;
    }
    if (p.bias != nullptr) {
#pragma unroll
        for (int j = 0; j < 32; j++) acc[j] += __ldg(p.bias + n0 + j);
    }
    if (p.act == ACT_RELU) {
#pragma unroll
        for (int j = 0; j < 32; j++) acc[j] = fmaxf(acc[j], 0.0f);
    } else if (p.act == ACT_SILU) {
#pragma unroll
        for (int j = 0; j < 32; j++) acc[j] = silu_f(acc[j]);
    }
    if (p.kind == EPI_STORE) {
        float4* dst = reinterpret_cast<float4*>(p.out_f32 + (size_t)m * p.ldo + n0);
#pragma unroll
        for (int j = 0; j < 8; j++) dst[j] = make_float4(acc[4 * j], acc[4 * j + 1], acc[4 * j + 2], acc[4 * j + 3]);
    } else if (p.kind == EPI_STORE_F16) {
        if (p.act_quant) {
            float amax = 0.0f;
#pragma unroll
            for (int j = 0; j < 32; j++) amax = fmaxf(amax, fabsf(acc[j]));
#pragma unroll
            for (int j = 0; j < 32; j++) acc[j] = fake_quant(acc[j], amax);
        }
        uint4* dst = reinterpret_cast<uint4*>(p.out_f16 + (size_t)m * p.ldo + n0);
#pragma unroll
        for (int j = 0; j < 4; j++) {
            __half2 h0 = __floats2half2_rn(acc[8 * j + 0], acc[8 * j + 1]);
            __half2 h1 = __floats2half2_rn(acc[8 * j + 2], acc[8 * j + 3]);
            __half2 h2 = __floats2half2_rn(acc[8 * j + 4], acc[8 * j + 5]);
            __half2 h3 = __floats2half2_rn(acc[8 * j + 6], acc[8 * j + 7]);
            uint4 u;
            u.x = *reinterpret_cast<uint32_t*>(&h0);
            u.y = *reinterpret_cast<uint32_t*>(&h1);
            u.z = *reinterpret_cast<uint32_t*>(&h2);
            u.w = *reinterpret_cast<uint32_t*>(&h3);
            dst[j] = u;
        }
    } else if (p.kind == EPI_QKV) {
        float* base;
        if (n0 < D_MODEL) {
            base = p.out_f32 + (size_t)m * D_MODEL + n0;
        } else {
            int b = m / p.T, i = m - b * p.T;
            int pos = p.wpos[b] + i;
            if (pos >= RING) pos -= RING;
            float* ring = (n0 < 2 * D_MODEL) ? p.k_ring : p.v_ring;
            base = ring + ((size_t)p.slot[b] * RING + pos) * D_MODEL + (n0 & (D_MODEL - 1));
        }
        float4* dst = reinterpret_cast<float4*>(base);
#pragma unroll
        for (int j = 0; j < 8; j++) dst[j] = make_float4(acc[4 * j], acc[4 * j + 1], acc[4 * j + 2], acc[4 * j + 3]);
    } else if (p.kind == EPI_GLU) {
        float4* dst = reinterpret_cast<float4*>(p.out_f32 + (size_t)m * p.ldo + (n0 >> 1));
#pragma unroll
        for (int j = 0; j < 4; j++) {
            float4 v;
            v.x = acc[8 * j + 0] * sigmoid_f(acc[8 * j + 1]);
            v.y = acc[8 * j + 2] * sigmoid_f(acc[8 * j + 3]);
            v.z = acc[8 * j + 4] * sigmoid_f(acc[8 * j + 5]);
            v.w = acc[8 * j + 6] * sigmoid_f(acc[8 * j + 7]);
            dst[j] = v;
        }
    }
}

// One fp16 operand matrix [rows, K] (K contiguous) in HBM, with the TMA descriptors of its two tile heights.
struct Operand {
    const __half* ptr = nullptr;
    int rows = 0, K = 0;
    CUtensorMap map128;  // box 64(K) x 128 rows, 128B swizzle
    CUtensorMap map64;   // box 64(K) x 64 rows
};
void operand_init(Operand& op, const __half* ptr, int rows, int K);  // builds the tensor maps

// C = A * W^T (+ epilogue). M valid rows of A; N = W.rows; K = A.K = W.K. use_simt selects the checker kernel.
void gemm(cudaStream_t st, const Operand& A, const Operand& W, int M, const EpiParams& epi, bool use_simt);

}  // namespace nb

// encoder.cu — the bandwidth-bound pieces of the cache-aware conformer layer (reference: src/nemo-stream.cpp:336-690).
#include <cstdlib>

#include "kernels.cuh"

namespace nb {

// ---- block reductions over 256 threads ------------------------------------------------------------
// `red` must be a different 8-float array for consecutive reductions of one kernel: that removes the barrier that
// would otherwise protect the scratch from the next reduction's writes (one __syncthreads per reduction, not two).
__device__ __forceinline__ float block_sum_256(float v, float* red) {
    v = warp_sum(v);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) red[w] = v;
    __syncthreads();
    float s = 0.0f;
#pragma unroll
    for (int i = 0; i < 8; i++) s += red[i];
    return s;
}

// amax over each group of 8 consecutive lanes (= 32 consecutive features when every lane holds 4)
__device__ __forceinline__ float group8_max(float v) {
    v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 1));
    v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 2));
    v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 4));
    return v;
}

__device__ __forceinline__ void store_f16x4(__half* dst, float4 v, int act_quant) {
    if (act_quant) {
        float amax = group8_max(fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w))));
        v.x = fake_quant(v.x, amax);
        v.y = fake_quant(v.y, amax);
        v.z = fake_quant(v.z, amax);
        v.w = fake_quant(v.w, amax);
    }
    __half2 a = __floats2half2_rn(v.x, v.y), b = __floats2half2_rn(v.z, v.w);
    uint2 u;
    u.x = *reinterpret_cast<uint32_t*>(&a);
    u.y = *reinterpret_cast<uint32_t*>(&b);
    *reinterpret_cast<uint2*>(dst) = u;
}

// LayerNorm over 1024 features the way ggml_norm does it (ggml/src/ggml-cpu/ops.cpp:3698-3741): mean, then the
// variance of the centred values, eps 1e-5; affine as separate mul/add (src/nemo-stream.cpp:580-591).
__device__ __forceinline__ float4 ln_1024(float4 v, const float* g, const float* b, float* red) {
    const int c = threadIdx.x * 4;
    const float mean = block_sum_256(v.x + v.y + v.z + v.w, red) * (1.0f / D_MODEL);
    v.x -= mean; v.y -= mean; v.z -= mean; v.w -= mean;
    const float var = block_sum_256(v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w, red + 8) * (1.0f / D_MODEL);
    const float sc = 1.0f / sqrtf(var + 1e-5f);
    const float4 gg = *reinterpret_cast<const float4*>(g + c), bb = *reinterpret_cast<const float4*>(b + c);
    v.x = v.x * sc * gg.x + bb.x;
    v.y = v.y * sc * gg.y + bb.y;
    v.z = v.z * sc * gg.z + bb.z;
    v.w = v.w * sc * gg.w + bb.w;
    return v;
}

__global__ void __launch_bounds__(256) layernorm_kernel(float* __restrict__ x, const float* __restrict__ g1,
                                                        const float* __restrict__ b1, float* __restrict__ out_f32,
                                                        const float* __restrict__ g2, const float* __restrict__ b2,
                                                        __half* __restrict__ out_f16, int act_quant, ResidualIn rin) {
    pdl_prologue();
    __shared__ float red[32];
    const size_t off = (size_t)blockIdx.x * D_MODEL + threadIdx.x * 4;
    float4 v = *reinterpret_cast<const float4*>(x + off);
    if (rin.partials != nullptr) {  // x += alpha * (sum of the preceding GEMM's split-K partials), in split order
        float4 acc = *reinterpret_cast<const float4*>(rin.partials + off);
        for (int s = 1; s < rin.nsplit; s++) {
            const float4 p = *reinterpret_cast<const float4*>(rin.partials + (size_t)s * rin.stride + off);
            acc.x += p.x; acc.y += p.y; acc.z += p.z; acc.w += p.w;
        }
        v.x += rin.alpha * acc.x; v.y += rin.alpha * acc.y; v.z += rin.alpha * acc.z; v.w += rin.alpha * acc.w;
        if (out_f32 != x) *reinterpret_cast<float4*>(x + off) = v;  // the residual stream (unless the LN output replaces it)
    }
    v = ln_1024(v, g1, b1, red);
    if (out_f32) *reinterpret_cast<float4*>(out_f32 + off) = v;
    if (g2) v = ln_1024(v, g2, b2, red + 16);
    if (out_f16) store_f16x4(out_f16 + off, v, act_quant);
}

void launch_layernorm(cudaStream_t st, float* x, const float* g1, const float* b1, float* out_f32, const float* g2,
                      const float* b2, __half* out_f16, int act_quant, int M, const ResidualIn& rin) {
    if (M <= 0) return;
    ProfScope ps(st, "layernorm", (double)M * D_MODEL * 8,
                 (double)M * D_MODEL * (4 + (out_f32 ? 4 : 0) + (out_f16 ? 2 : 0) + (rin.partials ? 4.0 * rin.nsplit + (out_f32 != x ? 4 : 0) : 0)));
    launch_k(layernorm_kernel, dim3(M), dim3(256), 0, st, x, g1, b1, out_f32, g2, b2, out_f16, act_quant, rin);
    NB_KERNEL_CHECK();
}

// ---- cached relative-position attention -----------------------------------------------------------
// One CTA per (head, stream). S[i][j] = ((q_i+u).K_j + (q_i+v).P[70+i-j]) / sqrt(128) + mask_j, softmax over j,
// ctx_i = sum_j A[i][j] V_j  (src/nemo-stream.cpp:463-573; the pad/reshape rel-shift :419-461 is the index 70+i-j).
// K_j/V_j for j < 70 are the ring-cache rows before this chunk, j >= 70 this chunk's rows (already written by the
// QKV GEMM epilogue). mask_j = -1e9 for j < 70 - valid (:1035-1045).
constexpr int ATT_LD = D_HEAD + 1;  // padded row: conflict-free when a warp reads one column of 32 rows

__global__ void __launch_bounds__(256) attention_kernel(AttnArgs a) {
    pdl_prologue();
    extern __shared__ float sm[];
    const int h = blockIdx.x, b = blockIdx.y, T = a.T;
    const int kv = LEFT + T, npos = LEFT + 2 * T - 1;
    float* Ks = sm;                       // [kv][ATT_LD]   (re-used for V)
    float* Ps = Ks + kv * ATT_LD;         // [npos][ATT_LD]
    float* Qu = Ps + npos * ATT_LD;       // [T][128]
    float* Qv = Qu + T * D_HEAD;          // [T][128]
    float* Sc = Qv + T * D_HEAD;          // [T][kv]
    const int t = threadIdx.x;
    const int slot = a.slot[b];
    int start = a.wpos[b] - LEFT;         // ring index of logical key 0
    if (start < 0) start += RING;
    const float* kbase = a.k_ring + (size_t)slot * RING * D_MODEL + h * D_HEAD;
    const float* vbase = a.v_ring + (size_t)slot * RING * D_MODEL + h * D_HEAD;
    // load K rows, P rows, q (+u, +v)
    for (int i = t; i < kv * (D_HEAD / 4); i += 256) {
        int j = i >> 5, d = (i & 31) * 4;
        int pos = start + j;
        if (pos >= RING) pos -= RING;
        float4 v = *reinterpret_cast<const float4*>(kbase + (size_t)pos * D_MODEL + d);
        float* dst = Ks + j * ATT_LD + d;
        dst[0] = v.x; dst[1] = v.y; dst[2] = v.z; dst[3] = v.w;
    }
    for (int i = t; i < npos * (D_HEAD / 4); i += 256) {
        int j = i >> 5, d = (i & 31) * 4;
        float4 v = *reinterpret_cast<const float4*>(a.pos + (size_t)j * D_MODEL + h * D_HEAD + d);
        float* dst = Ps + j * ATT_LD + d;
        dst[0] = v.x; dst[1] = v.y; dst[2] = v.z; dst[3] = v.w;
    }
    for (int i = t; i < T * D_HEAD; i += 256) {
        int r = i >> 7, d = i & 127;
        float qv = a.q[((size_t)(b * T + r)) * D_MODEL + h * D_HEAD + d];
        Qu[i] = qv + a.bias_u[h * D_HEAD + d];
        Qv[i] = qv + a.bias_v[h * D_HEAD + d];
    }
    __syncthreads();
    const int masked = LEFT - a.valid[b];
    const float scale = 1.0f / sqrtf((float)D_HEAD);
    for (int idx = t; idx < T * kv; idx += 256) {
        const int i = idx / kv, j = idx - i * kv;
        const float* kr = Ks + j * ATT_LD;
        const float* pr = Ps + (LEFT + i - j + T - 1) * ATT_LD;
        const float* qu = Qu + i * D_HEAD;
        const float* qv = Qv + i * D_HEAD;
        float s1 = 0.0f, s2 = 0.0f;
#pragma unroll 8
        for (int d = 0; d < D_HEAD; d++) {
            s1 = fmaf(qu[d], kr[d], s1);
            s2 = fmaf(qv[d], pr[d], s2);
        }
        float s = (s1 + s2) * scale;
        if (j < masked) s += -1e9f;
        Sc[idx] = s;
    }
    __syncthreads();
    // V over K's storage
    for (int i = t; i < kv * (D_HEAD / 4); i += 256) {
        int j = i >> 5, d = (i & 31) * 4;
        int pos = start + j;
        if (pos >= RING) pos -= RING;
        float4 v = *reinterpret_cast<const float4*>(vbase + (size_t)pos * D_MODEL + d);
        float* dst = Ks + j * ATT_LD + d;
        dst[0] = v.x; dst[1] = v.y; dst[2] = v.z; dst[3] = v.w;
    }
    // softmax per query row: one warp per row
    const int warp = t >> 5, lane = t & 31;
    for (int i = warp; i < T; i += 8) {
        float* row = Sc + i * kv;
        float mx = -INFINITY;
        for (int j = lane; j < kv; j += 32) mx = fmaxf(mx, row[j]);
        mx = warp_max(mx);
        float sum = 0.0f;
        for (int j = lane; j < kv; j += 32) {
            float e = expf(row[j] - mx);
            row[j] = e;
            sum += e;
        }
        sum = warp_sum(sum);
        const float inv = 1.0f / sum;
        for (int j = lane; j < kv; j += 32) row[j] *= inv;
    }
    __syncthreads();
    // ctx[i][d]: each thread 4 consecutive d of one row i
    for (int idx = t; idx < T * (D_HEAD / 4); idx += 256) {
        const int i = idx >> 5, d = (idx & 31) * 4;
        const float* row = Sc + i * kv;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int j = 0; j < kv; j++) {
            const float p = row[j];
            const float* vr = Ks + j * ATT_LD + d;
            acc.x = fmaf(p, vr[0], acc.x);
            acc.y = fmaf(p, vr[1], acc.y);
            acc.z = fmaf(p, vr[2], acc.z);
            acc.w = fmaf(p, vr[3], acc.w);
        }
        // 8 consecutive lanes hold 32 consecutive features of the same row -> fake-quant group
        store_f16x4(a.out + ((size_t)(b * T + i)) * D_MODEL + h * D_HEAD + d, acc, a.act_quant);
    }
}

static size_t attn_smem_bytes(int T) {
    const int kv = LEFT + T, npos = LEFT + 2 * T - 1;
    return sizeof(float) * ((size_t)kv * ATT_LD + (size_t)npos * ATT_LD + 2 * T * D_HEAD + (size_t)T * kv);
}

// ---- tensor-core version (the one the step uses) ------------------------------------------------------------
// Same math, restructured as three small matrix products per (stream, head) on mma.sync.m16n8k8 with 3xTF32 operand
// splitting (hi + lo, three MMAs: fp32-level accuracy):  C1 = (Q+u) K^T  [16 x kv],  C2 = (Q+v) P^T  [16 x npos],
// ctx = softmax(...) V  [16 x 128].  T <= 14 query rows fit one 16-row MMA tile. K, V and P rows are read straight
// from global memory as MMA B-fragments (each element is used once per CTA, so staging them in shared memory buys
// nothing): the k-slot <-> feature mapping is permuted so that every lane loads 16 contiguous bytes (K, P) or 8
// (V) and every 32-byte sector fetched is fully used. Shared memory holds only Q(+u,+v), the scores and the context
// (39 KB -> 5 CTAs per SM instead of 1), so the K/V streams of different CTAs overlap.
constexpr int QS = 144;   // Q row stride (floats): 128-bit A-fragment loads are conflict-free per quarter-warp
constexpr int S1S = 92;   // scores/probabilities row stride (>= 88)
constexpr int S2S = 108;  // positional scores row stride (>= 104)
constexpr int CXS = 132;

__device__ __forceinline__ void tf32_split(float x, uint32_t& hi, uint32_t& lo) {
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hi) : "f"(x));
    const float r = x - __uint_as_float(hi);
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(lo) : "f"(r));
}
__device__ __forceinline__ void mma_tf32(float (&c)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
// c += A(16x8) * B(8x8) with fp32 inputs split into tf32 hi/lo: a_hi*b_hi + a_lo*b_hi + a_hi*b_lo
__device__ __forceinline__ void mma_3xtf32(float (&c)[4], const float (&a)[4], const float (&b)[2]) {
    uint32_t ah[4], al[4], bh[2], bl[2];
#pragma unroll
    for (int i = 0; i < 4; i++) tf32_split(a[i], ah[i], al[i]);
#pragma unroll
    for (int i = 0; i < 2; i++) tf32_split(b[i], bh[i], bl[i]);
    mma_tf32(c, al, bh);
    mma_tf32(c, ah, bl);
    mma_tf32(c, ah, bh);
}

__global__ void __launch_bounds__(256) attention_tc_kernel(AttnArgs a) {
    pdl_prologue();
    __shared__ __align__(16) float Qu[16 * QS];
    __shared__ __align__(16) float Qv[16 * QS];
    __shared__ __align__(16) float S1[16 * S1S];
    __shared__ __align__(16) float S2[16 * S2S];
    __shared__ __align__(16) float Cx[16 * CXS];
    const int h = blockIdx.x, b = blockIdx.y, T = a.T;
    const int kv = LEFT + T, npos = LEFT + 2 * T - 1;
    const int kvt = (kv + 7) >> 3, npt = (npos + 7) >> 3;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const int slot = a.slot[b];
    int start = a.wpos[b] - LEFT;  // ring index of logical key 0
    if (start < 0) start += RING;
    const float* kbase = a.k_ring + (size_t)slot * RING * D_MODEL + h * D_HEAD;
    const float* vbase = a.v_ring + (size_t)slot * RING * D_MODEL + h * D_HEAD;
    const float* pbase = a.pos + h * D_HEAD;

    for (int idx = tid; idx < 16 * (D_HEAD / 4); idx += 256) {
        const int r = idx >> 5, d = (idx & 31) * 4;
        float4 qu = make_float4(0.f, 0.f, 0.f, 0.f), qv = qu;
        if (r < T) {
            const float4 q = *reinterpret_cast<const float4*>(a.q + ((size_t)(b * T + r)) * D_MODEL + h * D_HEAD + d);
            const float4 u = *reinterpret_cast<const float4*>(a.bias_u + h * D_HEAD + d);
            const float4 v = *reinterpret_cast<const float4*>(a.bias_v + h * D_HEAD + d);
            qu = make_float4(q.x + u.x, q.y + u.y, q.z + u.z, q.w + u.w);
            qv = make_float4(q.x + v.x, q.y + v.y, q.z + v.z, q.w + v.w);
        }
        *reinterpret_cast<float4*>(Qu + r * QS + d) = qu;
        *reinterpret_cast<float4*>(Qv + r * QS + d) = qv;
    }
    __syncthreads();

    // ---- scores: 8-column tiles of C1 (keys) then C2 (relative positions), round-robin over the warps ----
    for (int tile = warp; tile < kvt + npt; tile += 8) {
        const bool is_key = tile < kvt;
        const int n0 = 8 * (is_key ? tile : tile - kvt);
        const int lim = is_key ? kv : npos;
        const int r = min(n0 + g, lim - 1);  // padding columns re-read the last valid row; their scores are never used
        const float* src;
        if (is_key) {
            int pos = start + r;
            if (pos >= RING) pos -= RING;
            src = kbase + (size_t)pos * D_MODEL;
        } else {
            src = pbase + (size_t)r * D_MODEL;
        }
        const float* Qs = is_key ? Qu : Qv;
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
        for (int kk = 0; kk < 8; kk++) {
            const float4 bv = __ldg(reinterpret_cast<const float4*>(src + 16 * kk + 4 * t));
            const float4 qa = *reinterpret_cast<const float4*>(Qs + g * QS + 16 * kk + 4 * t);
            const float4 qb = *reinterpret_cast<const float4*>(Qs + (g + 8) * QS + 16 * kk + 4 * t);
            {   // k-slot t <-> feature 16kk+4t, k-slot t+4 <-> feature 16kk+4t+1
                const float af[4] = {qa.x, qb.x, qa.y, qb.y};
                const float bf[2] = {bv.x, bv.y};
                mma_3xtf32(acc, af, bf);
            }
            {   // features 16kk+4t+2 and +3
                const float af[4] = {qa.z, qb.z, qa.w, qb.w};
                const float bf[2] = {bv.z, bv.w};
                mma_3xtf32(acc, af, bf);
            }
        }
        float* S = is_key ? S1 : S2;
        const int ld = is_key ? S1S : S2S;
        *reinterpret_cast<float2*>(S + g * ld + n0 + 2 * t) = make_float2(acc[0], acc[1]);
        *reinterpret_cast<float2*>(S + (g + 8) * ld + n0 + 2 * t) = make_float2(acc[2], acc[3]);
    }
    __syncthreads();

    // ---- rel-shift by indexing, scale, mask, softmax: one warp per query row; probabilities overwrite S1 ----
    {
        const int masked = LEFT - a.valid[b];
        const float scale = 1.0f / sqrtf((float)D_HEAD);
        for (int i = warp; i < 16; i += 8) {
            float* row = S1 + i * S1S;
            if (i >= T) {
                for (int j = lane; j < kvt * 8; j += 32) row[j] = 0.0f;
                continue;
            }
            const float* prow = S2 + i * S2S + (LEFT + i + T - 1);
            float sc[3];
            float mx = -INFINITY;
#pragma unroll
            for (int q = 0; q < 3; q++) {
                const int j = lane + 32 * q;
                float s = -INFINITY;
                if (j < kv) {
                    s = (row[j] + prow[-j]) * scale;
                    if (j < masked) s += -1e9f;
                }
                sc[q] = s;
                mx = fmaxf(mx, s);
            }
            mx = warp_max(mx);
            float sum = 0.0f;
#pragma unroll
            for (int q = 0; q < 3; q++) {
                const int j = lane + 32 * q;
                sc[q] = (j < kv) ? expf(sc[q] - mx) : 0.0f;
                sum += sc[q];
            }
            sum = warp_sum(sum);
            const float inv = 1.0f / sum;
#pragma unroll
            for (int q = 0; q < 3; q++) {
                const int j = lane + 32 * q;
                if (j < kvt * 8) row[j] = sc[q] * inv;
            }
        }
    }
    __syncthreads();

    // ---- context: warp w owns features [16w, 16w+16) as two 8-column tiles (feature = 16w + 2n + q) ----
    {
        float acc0[4] = {0.f, 0.f, 0.f, 0.f}, acc1[4] = {0.f, 0.f, 0.f, 0.f};
        const float* vcol = vbase + 16 * warp + 2 * g;
        for (int ks = 0; ks < kvt; ks++) {
            const int j0 = 8 * ks + t, j1 = j0 + 4;
            int p0 = start + min(j0, kv - 1), p1 = start + min(j1, kv - 1);  // padded keys have probability 0
            if (p0 >= RING) p0 -= RING;
            if (p1 >= RING) p1 -= RING;
            const float2 v0 = __ldg(reinterpret_cast<const float2*>(vcol + (size_t)p0 * D_MODEL));
            const float2 v1 = __ldg(reinterpret_cast<const float2*>(vcol + (size_t)p1 * D_MODEL));
            const float af[4] = {S1[g * S1S + j0], S1[(g + 8) * S1S + j0], S1[g * S1S + j1], S1[(g + 8) * S1S + j1]};
            const float b0[2] = {v0.x, v1.x};
            const float b1[2] = {v0.y, v1.y};
            mma_3xtf32(acc0, af, b0);
            mma_3xtf32(acc1, af, b1);
        }
        *reinterpret_cast<float4*>(Cx + g * CXS + 16 * warp + 4 * t) = make_float4(acc0[0], acc1[0], acc0[1], acc1[1]);
        *reinterpret_cast<float4*>(Cx + (g + 8) * CXS + 16 * warp + 4 * t) = make_float4(acc0[2], acc1[2], acc0[3], acc1[3]);
    }
    __syncthreads();
    // ---- fp16 store (+ per-32-feature re-quantisation): a warp covers one row, 8 lanes = one 32-feature block ----
    for (int idx = tid; idx < T * (D_HEAD / 4); idx += 256) {
        const int i = idx >> 5, d = (idx & 31) * 4;
        const float4 v = *reinterpret_cast<const float4*>(Cx + i * CXS + d);
        store_f16x4(a.out + ((size_t)(b * T + i)) * D_MODEL + h * D_HEAD + d, v, a.act_quant);
    }
}

void launch_attention(cudaStream_t st, const AttnArgs& a) {
    static const bool use_simt = getenv("NB200_ATTN_SIMT") != nullptr;
    const double kv = LEFT + a.T;
    const double flops = 6.0 * a.B * a.T * kv * D_MODEL;
    const double bytes = (double)a.B * (2.0 * kv * D_MODEL * 4 + (double)a.T * D_MODEL * 6) + (LEFT + 2.0 * a.T - 1) * D_MODEL * 4;
    if (use_simt) {
        static bool configured = false;
        if (!configured) {
            CUDA_CHECK(cudaFuncSetAttribute(attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)attn_smem_bytes(MAX_T)));
            configured = true;
        }
        ProfScope ps(st, "attention_simt", flops, bytes);
        launch_k(attention_kernel, dim3(dim3(N_HEADS, a.B)), dim3(256), attn_smem_bytes(a.T), st, a);
        NB_KERNEL_CHECK();
        return;
    }
    if (a.T > 14) throw Error("attention: more than 14 frames per chunk");
    ProfScope ps(st, "attention_tc", flops, bytes);
    launch_k(attention_tc_kernel, dim3(dim3(N_HEADS, a.B)), dim3(256), 0, st, a);
    NB_KERNEL_CHECK();
}

// ---- cached causal depthwise conv (9 taps) + LayerNorm + SiLU --------------------------------------
// One CTA per (frame t, stream): d[t][ch] = sum_k G[t+k][ch] * w[k][ch], G = [cache(8) ; g(T)] (src/nemo-stream.cpp:336-412),
// products accumulated in tap order like the reference's mul/add chain (:370-388); then LN (conv.batch_norm), SiLU (:671-674).
__global__ void __launch_bounds__(256) dwconv_kernel(ConvArgs a) {
    pdl_prologue();
    __shared__ float red[16];
    const int t = blockIdx.x, b = blockIdx.y, c = threadIdx.x * 4;
    const float* cache = a.conv_cache + (size_t)a.slot[b] * CONV_CACHE * D_MODEL;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int k = 0; k < KERNEL; k++) {
        const int r = t + k;  // row of G
        const float* src = (r < CONV_CACHE) ? cache + (size_t)r * D_MODEL + c
                                            : a.g + ((size_t)(b * a.T + (r - CONV_CACHE))) * D_MODEL + c;
        const float4 x = *reinterpret_cast<const float4*>(src);
        const float4 w = *reinterpret_cast<const float4*>(a.w_dw + k * D_MODEL + c);
        if (k == 0) {
            acc.x = x.x * w.x; acc.y = x.y * w.y; acc.z = x.z * w.z; acc.w = x.w * w.w;
        } else {
            acc.x += x.x * w.x; acc.y += x.y * w.y; acc.z += x.z * w.z; acc.w += x.w * w.w;
        }
    }
    float4 v = ln_1024(acc, a.ln_g, a.ln_b, red);
    v.x = silu_f(v.x); v.y = silu_f(v.y); v.z = silu_f(v.z); v.w = silu_f(v.w);
    store_f16x4(a.out + ((size_t)(b * a.T + t)) * D_MODEL + c, v, a.act_quant);
}

void launch_dwconv(cudaStream_t st, const ConvArgs& a) {
    ProfScope ps(st, "dwconv_ln_silu", (double)a.B * a.T * D_MODEL * 30, (double)a.B * ((a.T + 8.0) * D_MODEL * 4 + (double)a.T * D_MODEL * 2));
    launch_k(dwconv_kernel, dim3(dim3(a.T, a.B)), dim3(256), 0, st, a);
    NB_KERNEL_CHECK();
}

// new cache = last 8 rows of [cache ; g]; one CTA per stream, thread = 4 channels, rows staged in registers
__global__ void __launch_bounds__(256) conv_cache_update_kernel(ConvArgs a) {
    pdl_prologue();
    const int b = blockIdx.x, c = threadIdx.x * 4, T = a.T;
    float* cache = a.conv_cache + (size_t)a.slot[b] * CONV_CACHE * D_MODEL;
    float4 rows[CONV_CACHE];
#pragma unroll
    for (int r = 0; r < CONV_CACHE; r++) {
        const int src = T + r;  // row of G = [cache ; g]
        rows[r] = (src < CONV_CACHE) ? *reinterpret_cast<const float4*>(cache + (size_t)src * D_MODEL + c)
                                     : *reinterpret_cast<const float4*>(a.g + ((size_t)(b * T + (src - CONV_CACHE))) * D_MODEL + c);
    }
#pragma unroll
    for (int r = 0; r < CONV_CACHE; r++) *reinterpret_cast<float4*>(cache + (size_t)r * D_MODEL + c) = rows[r];
}

void launch_conv_cache_update(cudaStream_t st, const ConvArgs& a) {
    ProfScope ps(st, "conv_cache_update", 0.0, (double)a.B * 16.0 * D_MODEL * 4);
    launch_k(conv_cache_update_kernel, dim3(a.B), dim3(256), 0, st, a);
    NB_KERNEL_CHECK();
}

// ---- helpers for the positional table ---------------------------------------------------------------
__global__ void __launch_bounds__(256) cast_f16_kernel(const float* __restrict__ x, __half* __restrict__ out, int act_quant, int cols) {
    pdl_prologue();
    const size_t row = blockIdx.x;
    for (int c = threadIdx.x * 4; c < cols; c += 1024) {
        float4 v = *reinterpret_cast<const float4*>(x + row * cols + c);
        store_f16x4(out + row * cols + c, v, act_quant);
    }
}
void launch_cast_f16(cudaStream_t st, const float* x, __half* out, int act_quant, int rows, int cols) {
    if (cols % 1024 != 0) throw Error("cast_f16: cols must be a multiple of 1024");
    launch_k(cast_f16_kernel, dim3(rows), dim3(256), 0, st, x, out, act_quant, cols);
    NB_KERNEL_CHECK();
}

// PE(p)[2i] = sin(p * w_i), PE(p)[2i+1] = cos(p * w_i), w_i = exp(-(2i) * ln(10000) / 1024)  (src/nemo-ggml.cpp:17-32)
__global__ void pos_emb_kernel(float* out, int p0) {
    pdl_prologue();
    const int r = blockIdx.x, i2 = threadIdx.x * 2;  // 512 threads
    const float p = (float)(p0 + r);
    const float div = expf(-(float)i2 * logf(10000.0f) / (float)D_MODEL);
    out[(size_t)r * D_MODEL + i2] = sinf(p * div);
    out[(size_t)r * D_MODEL + i2 + 1] = cosf(p * div);
}
void launch_pos_emb(cudaStream_t st, float* out, int p0, int n) {
    launch_k(pos_emb_kernel, dim3(n), dim3(512), 0, st, out, p0);
    NB_KERNEL_CHECK();
}

}  // namespace nb

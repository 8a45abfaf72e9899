// encoder.cu — the bandwidth-bound pieces of the cache-aware conformer layer (reference: src/nemo-stream.cpp:336-690).
#include "kernels.cuh"

namespace nb {

// ---- block reductions over 256 threads ------------------------------------------------------------
__device__ __forceinline__ float block_sum_256(float v, float* red) {
    v = warp_sum(v);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
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
    const float var = block_sum_256(v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w, red) * (1.0f / D_MODEL);
    const float sc = 1.0f / sqrtf(var + 1e-5f);
    const float4 gg = *reinterpret_cast<const float4*>(g + c), bb = *reinterpret_cast<const float4*>(b + c);
    v.x = v.x * sc * gg.x + bb.x;
    v.y = v.y * sc * gg.y + bb.y;
    v.z = v.z * sc * gg.z + bb.z;
    v.w = v.w * sc * gg.w + bb.w;
    return v;
}

__global__ void __launch_bounds__(256) layernorm_kernel(const float* __restrict__ x, const float* __restrict__ g1,
                                                        const float* __restrict__ b1, float* __restrict__ out_f32,
                                                        const float* __restrict__ g2, const float* __restrict__ b2,
                                                        __half* __restrict__ out_f16, int act_quant) {
    __shared__ float red[8];
    const size_t off = (size_t)blockIdx.x * D_MODEL + threadIdx.x * 4;
    float4 v = *reinterpret_cast<const float4*>(x + off);
    v = ln_1024(v, g1, b1, red);
    if (out_f32) *reinterpret_cast<float4*>(out_f32 + off) = v;
    if (g2) v = ln_1024(v, g2, b2, red);
    if (out_f16) store_f16x4(out_f16 + off, v, act_quant);
}

void launch_layernorm(cudaStream_t st, const float* x, const float* g1, const float* b1, float* out_f32, const float* g2,
                      const float* b2, __half* out_f16, int act_quant, int M) {
    if (M <= 0) return;
    ProfScope ps(st, "layernorm", (double)M * D_MODEL * 8, (double)M * D_MODEL * (4 + (out_f32 ? 4 : 0) + (out_f16 ? 2 : 0)));
    layernorm_kernel<<<M, 256, 0, st>>>(x, g1, b1, out_f32, g2, b2, out_f16, act_quant);
    NB_KERNEL_CHECK();
}

// ---- cached relative-position attention -----------------------------------------------------------
// One CTA per (head, stream). S[i][j] = ((q_i+u).K_j + (q_i+v).P[70+i-j]) / sqrt(128) + mask_j, softmax over j,
// ctx_i = sum_j A[i][j] V_j  (src/nemo-stream.cpp:463-573; the pad/reshape rel-shift :419-461 is the index 70+i-j).
// K_j/V_j for j < 70 are the ring-cache rows before this chunk, j >= 70 this chunk's rows (already written by the
// QKV GEMM epilogue). mask_j = -1e9 for j < 70 - valid (:1035-1045).
constexpr int ATT_LD = D_HEAD + 1;  // padded row: conflict-free when a warp reads one column of 32 rows

__global__ void __launch_bounds__(256) attention_kernel(AttnArgs a) {
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

void launch_attention(cudaStream_t st, const AttnArgs& a) {
    static bool configured = false;
    if (!configured) {
        CUDA_CHECK(cudaFuncSetAttribute(attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)attn_smem_bytes(MAX_T)));
        configured = true;
    }
    {
        const double kv = LEFT + a.T;
        ProfScope ps(st, "attention", 6.0 * a.B * a.T * kv * D_MODEL,
                     (double)a.B * (2.0 * kv * D_MODEL * 4 + (double)a.T * D_MODEL * 6) + (LEFT + 2.0 * a.T - 1) * D_MODEL * 4);
        attention_kernel<<<dim3(N_HEADS, a.B), 256, attn_smem_bytes(a.T), st>>>(a);
        NB_KERNEL_CHECK();
    }
}

// ---- cached causal depthwise conv (9 taps) + LayerNorm + SiLU --------------------------------------
// One CTA per (frame t, stream): d[t][ch] = sum_k G[t+k][ch] * w[k][ch], G = [cache(8) ; g(T)] (src/nemo-stream.cpp:336-412),
// products accumulated in tap order like the reference's mul/add chain (:370-388); then LN (conv.batch_norm), SiLU (:671-674).
__global__ void __launch_bounds__(256) dwconv_kernel(ConvArgs a) {
    __shared__ float red[8];
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
    dwconv_kernel<<<dim3(a.T, a.B), 256, 0, st>>>(a);
    NB_KERNEL_CHECK();
}

// new cache = last 8 rows of [cache ; g]; one CTA per stream, thread = 4 channels, rows staged in registers
__global__ void __launch_bounds__(256) conv_cache_update_kernel(ConvArgs a) {
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
    conv_cache_update_kernel<<<a.B, 256, 0, st>>>(a);
    NB_KERNEL_CHECK();
}

// ---- helpers for the positional table ---------------------------------------------------------------
__global__ void __launch_bounds__(256) cast_f16_kernel(const float* __restrict__ x, __half* __restrict__ out, int act_quant, int cols) {
    const size_t row = blockIdx.x;
    for (int c = threadIdx.x * 4; c < cols; c += 1024) {
        float4 v = *reinterpret_cast<const float4*>(x + row * cols + c);
        store_f16x4(out + row * cols + c, v, act_quant);
    }
}
void launch_cast_f16(cudaStream_t st, const float* x, __half* out, int act_quant, int rows, int cols) {
    if (cols % 1024 != 0) throw Error("cast_f16: cols must be a multiple of 1024");
    cast_f16_kernel<<<rows, 256, 0, st>>>(x, out, act_quant, cols);
    NB_KERNEL_CHECK();
}

// PE(p)[2i] = sin(p * w_i), PE(p)[2i+1] = cos(p * w_i), w_i = exp(-(2i) * ln(10000) / 1024)  (src/nemo-ggml.cpp:17-32)
__global__ void pos_emb_kernel(float* out, int p0) {
    const int r = blockIdx.x, i2 = threadIdx.x * 2;  // 512 threads
    const float p = (float)(p0 + r);
    const float div = expf(-(float)i2 * logf(10000.0f) / (float)D_MODEL);
    out[(size_t)r * D_MODEL + i2] = sinf(p * div);
    out[(size_t)r * D_MODEL + i2 + 1] = cosf(p * div);
}
void launch_pos_emb(cudaStream_t st, float* out, int p0, int n) {
    pos_emb_kernel<<<n, 512, 0, st>>>(out, p0);
    NB_KERNEL_CHECK();
}

}  // namespace nb

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

__device__ __forceinline__ void store_f16x4(__half* dst, float4 v, int act_quant, int bf16 = 0) {
    if (act_quant) {
        float amax = group8_max(fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w))));
        v.x = fake_quant(v.x, amax);
        v.y = fake_quant(v.y, amax);
        v.z = fake_quant(v.z, amax);
        v.w = fake_quant(v.w, amax);
    }
    uint2 u;
    u.x = pack_operand2(v.x, v.y, bf16);
    u.y = pack_operand2(v.z, v.w, bf16);
    *reinterpret_cast<uint2*>(dst) = u;
}

// LayerNorm over 1024 features the way ggml_norm does it (ggml/src/ggml-cpu/ops.cpp:3698-3741): mean, then the
// variance of the centred values, eps 1e-5; affine as separate mul/add (src/nemo-stream.cpp:580-591).
// Block-wide version (256 threads = one row), used by the depthwise-conv kernel.
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

// NS = number of split-K partials to add (0: plain LayerNorm of x). A compile-time count lets all NS + 1 loads of a thread
// be in flight together (with a run-time loop they were issued one round trip after the other); the sums keep split order.
template <int NS>
__global__ void __launch_bounds__(256) layernorm_kernel(float* __restrict__ x, const float* __restrict__ g1,
                                                        const float* __restrict__ b1, float* __restrict__ out_f32,
                                                        const float* __restrict__ g2, const float* __restrict__ b2,
                                                        __half* __restrict__ out_f16, int act_quant, ResidualIn rin, int out_bf16) {
    pdl_prologue();
    __shared__ float red[32];
    const size_t off = (size_t)blockIdx.x * D_MODEL + threadIdx.x * 4;
    float4 v = *reinterpret_cast<const float4*>(x + off);
    if constexpr (NS > 0) {  // x += alpha * (sum of the preceding GEMM's split-K partials), in split order
        float4 p[NS];
#pragma unroll
        for (int s = 0; s < NS; s++) p[s] = *reinterpret_cast<const float4*>(rin.partials + (size_t)s * rin.stride + off);
        float4 acc = p[0];
#pragma unroll
        for (int s = 1; s < NS; s++) {
            acc.x += p[s].x; acc.y += p[s].y; acc.z += p[s].z; acc.w += p[s].w;
        }
        v.x += rin.alpha * acc.x; v.y += rin.alpha * acc.y; v.z += rin.alpha * acc.z; v.w += rin.alpha * acc.w;
        if (out_f32 != x) *reinterpret_cast<float4*>(x + off) = v;  // the residual stream (unless the LN output replaces it)
    }
    v = ln_1024(v, g1, b1, red);
    if (out_f32) *reinterpret_cast<float4*>(out_f32 + off) = v;
    if (g2) v = ln_1024(v, g2, b2, red + 16);
    if (out_f16) store_f16x4(out_f16 + off, v, act_quant, out_bf16);
}

template <int NS>
static void launch_ln_ns(cudaStream_t st, float* x, const float* g1, const float* b1, float* out_f32, const float* g2, const float* b2,
                         __half* out_f16, int act_quant, int M, const ResidualIn& rin, int out_bf16) {
    launch_k(layernorm_kernel<NS>, dim3(M), dim3(256), 0, st, x, g1, b1, out_f32, g2, b2, out_f16, act_quant, rin, out_bf16);
}

void launch_layernorm(cudaStream_t st, float* x, const float* g1, const float* b1, float* out_f32, const float* g2,
                      const float* b2, __half* out_f16, int act_quant, int M, const ResidualIn& rin, int out_bf16) {
    if (M <= 0) return;
    ProfScope ps(st, "layernorm", (double)M * D_MODEL * 8,
                 (double)M * D_MODEL * (4 + (out_f32 ? 4 : 0) + (out_f16 ? 2 : 0) + (rin.partials ? 4.0 * rin.nsplit + (out_f32 != x ? 4 : 0) : 0)));
    const int ns = rin.partials ? rin.nsplit : 0;
    if (ns < 0 || ns > 8) throw Error("layernorm: at most 8 split-K partials");
    for (int rep = dup_family("layernorm") ? 2 : 1; rep > 0; rep--) {
        switch (ns) {
            case 0: launch_ln_ns<0>(st, x, g1, b1, out_f32, g2, b2, out_f16, act_quant, M, rin, out_bf16); break;
            case 1: launch_ln_ns<1>(st, x, g1, b1, out_f32, g2, b2, out_f16, act_quant, M, rin, out_bf16); break;
            case 2: launch_ln_ns<2>(st, x, g1, b1, out_f32, g2, b2, out_f16, act_quant, M, rin, out_bf16); break;
            case 3: launch_ln_ns<3>(st, x, g1, b1, out_f32, g2, b2, out_f16, act_quant, M, rin, out_bf16); break;
            case 4: launch_ln_ns<4>(st, x, g1, b1, out_f32, g2, b2, out_f16, act_quant, M, rin, out_bf16); break;
            case 5: launch_ln_ns<5>(st, x, g1, b1, out_f32, g2, b2, out_f16, act_quant, M, rin, out_bf16); break;
            case 6: launch_ln_ns<6>(st, x, g1, b1, out_f32, g2, b2, out_f16, act_quant, M, rin, out_bf16); break;
            case 7: launch_ln_ns<7>(st, x, g1, b1, out_f32, g2, b2, out_f16, act_quant, M, rin, out_bf16); break;
            default: launch_ln_ns<8>(st, x, g1, b1, out_f32, g2, b2, out_f16, act_quant, M, rin, out_bf16); break;
        }
    }
    NB_KERNEL_CHECK();
}

// ---- cached causal depthwise conv (9 taps) + LayerNorm + SiLU + cache roll-forward --------------------------------
// One CTA per (frame t, stream), thread = 4 channels: d[t][ch] = sum_k G[t+k][ch] * w[k][ch], G = [cache(8) ; g(T)] (src/nemo-stream.cpp:336-412),
// products accumulated in tap order like the reference's mul/add chain (:370-388); then LN (conv.batch_norm), SiLU (:671-674).
// The new cache (last 8 rows of G, :399-409) goes to the stream's OTHER cache buffer (two buffers per stream and layer,
// flipped every chunk), so the CTAs of a stream can write it while others still read the old one: row t of g is new cache
// row t + 8 - T, written by the CTA that owns frame t; when T < 8 the CTA of frame 0 also moves old rows T.. up.
__global__ void __launch_bounds__(256) dwconv_kernel(ConvArgs a) {
    pdl_prologue();
    __shared__ float red[16];
    const int t = blockIdx.x, b = blockIdx.y, c = threadIdx.x * 4, T = a.T;
    const int par = a.parity[b] & 1;
    const float* rd = a.conv_cache + ((size_t)a.slot[b] * 2 + par) * CONV_CACHE * D_MODEL;
    float* wr = a.conv_cache + ((size_t)a.slot[b] * 2 + (par ^ 1)) * CONV_CACHE * D_MODEL;
    const float* g = a.g + (size_t)b * T * D_MODEL;
    float4 x[KERNEL];
#pragma unroll
    for (int k = 0; k < KERNEL; k++) {
        const int r = t + k;  // row of G
        x[k] = *reinterpret_cast<const float4*>(r < CONV_CACHE ? rd + (size_t)r * D_MODEL + c : g + (size_t)(r - CONV_CACHE) * D_MODEL + c);
    }
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int k = 0; k < KERNEL; k++) {
        const float4 w = *reinterpret_cast<const float4*>(a.w_dw + k * D_MODEL + c);
        if (k == 0) {
            acc.x = x[k].x * w.x; acc.y = x[k].y * w.y; acc.z = x[k].z * w.z; acc.w = x[k].w * w.w;
        } else {
            acc.x += x[k].x * w.x; acc.y += x[k].y * w.y; acc.z += x[k].z * w.z; acc.w += x[k].w * w.w;
        }
    }
    if (t + CONV_CACHE - T >= 0) *reinterpret_cast<float4*>(wr + (size_t)(t + CONV_CACHE - T) * D_MODEL + c) = x[KERNEL - 1];  // g row t
    if (t == 0) {
        for (int r = 0; r < CONV_CACHE - T; r++)
            *reinterpret_cast<float4*>(wr + (size_t)r * D_MODEL + c) = *reinterpret_cast<const float4*>(rd + (size_t)(T + r) * D_MODEL + c);
    }
    float4 v = ln_1024(acc, a.ln_g, a.ln_b, red);
    v.x = silu_f(v.x); v.y = silu_f(v.y); v.z = silu_f(v.z); v.w = silu_f(v.w);
    store_f16x4(a.out + ((size_t)(b * T + t)) * D_MODEL + c, v, a.act_quant, a.out_bf16);
}

void launch_dwconv(cudaStream_t st, const ConvArgs& a) {
    ProfScope ps(st, "dwconv_ln_silu", (double)a.B * a.T * D_MODEL * 30, (double)a.B * ((a.T + 16.0) * D_MODEL * 4 + (double)a.T * D_MODEL * 2));
    for (int rep = dup_family("dwconv") ? 2 : 1; rep > 0; rep--) launch_k(dwconv_kernel, dim3(dim3(a.T, a.B)), dim3(256), 0, st, a);
    NB_KERNEL_CHECK();
}

// ---- helpers for the positional table ---------------------------------------------------------------
// ---- prompt fusion, one-hot half of fc1 + ReLU: concat([enc ; onehot(lang)]) . W1^T = enc . W1[:, :1024]^T + W1[:, 1024 + lang] ----
__global__ void __launch_bounds__(256) prompt_add_relu_kernel(float* __restrict__ h, const float* __restrict__ w_lang, const int* __restrict__ lang, int T) {
    pdl_prologue();
    const int m = blockIdx.x;
    const float* wl = w_lang + (size_t)lang[m / T] * PROMPT_HIDDEN;
    float* row = h + (size_t)m * PROMPT_HIDDEN;
    for (int c = threadIdx.x * 4; c < PROMPT_HIDDEN; c += 1024) {
        float4 v = *reinterpret_cast<float4*>(row + c);
        const float4 w = *reinterpret_cast<const float4*>(wl + c);
        v.x = fmaxf(v.x + w.x, 0.0f); v.y = fmaxf(v.y + w.y, 0.0f); v.z = fmaxf(v.z + w.z, 0.0f); v.w = fmaxf(v.w + w.w, 0.0f);
        *reinterpret_cast<float4*>(row + c) = v;
    }
}

void launch_prompt_add_relu(cudaStream_t st, float* h, const float* w_lang, const int* lang, int T, int M) {
    if (M <= 0) return;
    ProfScope ps(st, "prompt_add_relu", (double)M * PROMPT_HIDDEN * 2, (double)M * PROMPT_HIDDEN * 8);
    launch_k(prompt_add_relu_kernel, dim3(M), dim3(256), 0, st, h, w_lang, lang, T);
    NB_KERNEL_CHECK();
}

// ---- fp32 -> (hi, lo, hi) fp16 triple for a tensor-core GEMM of fp32 quality ---------------------------------------------
// x = hi + lo + O(2^-24 |x|) with hi = fp16(x), lo = fp16(x - hi). With weights split the same way and laid out as
// [Wh | Wh | Wl] along K, one fp16 GEMM over K' = 3 * 1024 gives hi.Wh + lo.Wh + hi.Wl = x.W up to ~1e-6 relative (the lo.Wl
// term, 2^-24, is dropped; fp32 accumulation in TMEM). Used for joint.enc, whose weights the reference keeps in F32.
__global__ void __launch_bounds__(256) split_f16x3_kernel(const float* __restrict__ x, __half* __restrict__ out, int cols) {
    pdl_prologue();
    const size_t m = blockIdx.x;
    const int c = blockIdx.y * 1024 + threadIdx.x * 4;
    const float4 v = *reinterpret_cast<const float4*>(x + m * cols + c);
    const float f[4] = {v.x, v.y, v.z, v.w};
    __half hi[4], lo[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        hi[i] = __float2half_rn(f[i]);
        lo[i] = __float2half_rn(f[i] - __half2float(hi[i]));
    }
    __half* row = out + m * (3 * (size_t)cols) + c;
    *reinterpret_cast<uint2*>(row) = *reinterpret_cast<const uint2*>(hi);
    *reinterpret_cast<uint2*>(row + cols) = *reinterpret_cast<const uint2*>(lo);
    *reinterpret_cast<uint2*>(row + 2 * cols) = *reinterpret_cast<const uint2*>(hi);
}

void launch_split_f16x3(cudaStream_t st, const float* x, __half* out, int M, int cols) {
    if (M <= 0) return;
    if (cols % 1024 != 0) throw Error("split_f16x3: the row length must be a multiple of 1024");
    ProfScope ps(st, "split_f16x3", (double)M * cols * 3, (double)M * cols * 10);
    launch_k(split_f16x3_kernel, dim3(dim3(M, cols / 1024)), dim3(256), 0, st, x, out, cols);
    NB_KERNEL_CHECK();
}

__global__ void __launch_bounds__(256) cast_f16_kernel(const float* __restrict__ x, __half* __restrict__ out, int act_quant, int cols, int out_bf16) {
    pdl_prologue();
    const size_t row = blockIdx.x;
    for (int c = threadIdx.x * 4; c < cols; c += 1024) {
        float4 v = *reinterpret_cast<const float4*>(x + row * cols + c);
        store_f16x4(out + row * cols + c, v, act_quant, out_bf16);
    }
}
void launch_cast_f16(cudaStream_t st, const float* x, __half* out, int act_quant, int rows, int cols, int out_bf16) {
    if (cols % 1024 != 0) throw Error("cast_f16: cols must be a multiple of 1024");
    launch_k(cast_f16_kernel, dim3(rows), dim3(256), 0, st, x, out, act_quant, cols, out_bf16);
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

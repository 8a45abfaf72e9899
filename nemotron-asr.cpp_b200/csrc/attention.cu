// attention.cu — cached relative-position multi-head attention (reference: src/nemo-stream.cpp:463-573, rel-shift
// :419-461, mask fill :1035-1045, cache write-back :251-266).
//
// One CTA per (head, stream). For the chunk's T query frames i and kv = 70 + T keys j:
//   S[i][j] = ((q_i + u_h) . K_j + (q_i + v_h) . P[70 + i - j]) / sqrt(128) + mask_j,  mask_j = -1e9 for j < 70 - valid
//   ctx_i   = sum_j softmax_j(S[i][.]) V_j
// (the reference's pad/reshape/drop "rel-shift" is exactly the index 70 + i - j into the projected position table).
// Keys 0..69 are the stream's cached rows (fp16 ring in HBM, oldest first), keys 70.. are this chunk's rows, which the
// QKV GEMM has just written densely to qkv[M][3072]; this kernel also appends them to the ring (the reference copies
// the last 70 rows of [cache ; new] back into the cache tensors after every layer of every chunk).
//
// Data movement is the cost (2 x 84 rows x 256 B per CTA from HBM, the rest from L2), so every operand row is brought
// into shared memory with 16-byte cp.async issued up front (V in a second group that lands while the scores are
// computed), and the three small products run on mma.sync.m16n8k16 (fp16 in, fp32 accumulate) fed by ldmatrix from
// conflict-free padded rows. Softmax is fp32.
#include <cstdlib>

#include "kernels.cuh"

namespace nb {

namespace {

constexpr int AT_THREADS = 256;
constexpr int ROWB = 272;      // bytes per staged operand row: 128 halves + 16 B pad (ldmatrix rows hit 8 distinct 16-B bank groups)
constexpr int K_ROWS = 88, V_ROWS = 96, P_ROWS = 104;  // padded to MMA tiles (kv <= 84 -> 11 n-tiles / 6 k-steps; npos <= 97 -> 13 n-tiles)
constexpr int S1_LD = 92, S2_LD = 108, PR_LD = 104 /* halves: 208 B rows -> conflict-free ldmatrix */, CX_LD = 132;

struct AttSmem {
    static constexpr int K_OFF = 0;
    static constexpr int V_OFF = K_OFF + K_ROWS * ROWB;
    static constexpr int P_OFF = V_OFF + V_ROWS * ROWB;
    static constexpr int QU_OFF = P_OFF + P_ROWS * ROWB;
    static constexpr int QV_OFF = QU_OFF + 16 * ROWB;
    static constexpr int S1_OFF = QV_OFF + 16 * ROWB;
    static constexpr int S2_OFF = S1_OFF + 16 * S1_LD * 4;
    static constexpr int PR_OFF = S2_OFF + 16 * S2_LD * 4;
    static constexpr int TOTAL = PR_OFF + 16 * PR_LD * 2;
    // the fp32 context [16][CX_LD] is staged over the K rows (dead after the scores)
    static_assert(16 * CX_LD * 4 <= K_ROWS * ROWB, "context staging must fit the K area");
};

__device__ __forceinline__ uint32_t s_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma_f16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ __forceinline__ float group8_max(float v) {
    v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 1));
    v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 2));
    v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 4));
    return v;
}

}  // namespace

__global__ void __launch_bounds__(AT_THREADS, 2) attention_f16_kernel(AttnArgs a) {
    // PDL: the cached ring rows and the position table were written by earlier steps / at load, so they are fetched
    // before griddepcontrol.wait (overlapping the tail of the QKV GEMM); everything that reads qkv comes after it.
    pdl_launch();
    extern __shared__ __align__(16) uint8_t smem[];
    using L = AttSmem;
    const int h = blockIdx.x, b = blockIdx.y, T = a.T;
    const int kv = LEFT + T, npos = LEFT + 2 * T - 1;
    const int kvt = (kv + 7) >> 3, npt = (npos + 7) >> 3, kvs = (kv + 15) >> 4;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const int slot = a.slot[b], wpos = a.wpos[b];
    int start = wpos - LEFT;  // ring index of cached key 0
    if (start < 0) start += RING;
    __half* kring = a.k_ring + (size_t)slot * RING * D_MODEL + h * D_HEAD;
    __half* vring = a.v_ring + (size_t)slot * RING * D_MODEL + h * D_HEAD;
    const __half* qkv = a.qkv + (size_t)b * T * (3 * D_MODEL) + h * D_HEAD;  // row i: q at +0, k at +1024, v at +2048
    const __half* pbase = a.pos + h * D_HEAD;
    const uint32_t sb = s_u32(smem);

    // ---- stage K (cached, then new), P, then V; 16 B per cp.async ----
    for (int idx = tid; idx < LEFT * 16; idx += AT_THREADS) {
        const int j = idx >> 4, c = idx & 15;
        int pos = start + j;
        if (pos >= RING) pos -= RING;
        cp_async16(sb + L::K_OFF + j * ROWB + c * 16, kring + (size_t)pos * D_MODEL + c * 8);
    }
    for (int idx = tid; idx < npos * 16; idx += AT_THREADS) {
        const int j = idx >> 4, c = idx & 15;
        cp_async16(sb + L::P_OFF + j * ROWB + c * 16, pbase + (size_t)j * D_MODEL + c * 8);
    }
    pdl_wait();
    for (int idx = tid; idx < T * 16; idx += AT_THREADS) {
        const int i = idx >> 4, c = idx & 15;
        cp_async16(sb + L::K_OFF + (LEFT + i) * ROWB + c * 16, qkv + (size_t)i * (3 * D_MODEL) + D_MODEL + c * 8);
    }
    cp_commit();
    for (int idx = tid; idx < kv * 16; idx += AT_THREADS) {
        const int j = idx >> 4, c = idx & 15;
        const __half* src;
        if (j < LEFT) {
            int pos = start + j;
            if (pos >= RING) pos -= RING;
            src = vring + (size_t)pos * D_MODEL + c * 8;
        } else {
            src = qkv + (size_t)(j - LEFT) * (3 * D_MODEL) + 2 * D_MODEL + c * 8;
        }
        cp_async16(sb + L::V_OFF + j * ROWB + c * 16, src);
    }
    cp_commit();
    // zero the tile padding rows (their scores are never read, but 0 * garbage must stay finite in P.V)
    {
        const uint4 z = make_uint4(0, 0, 0, 0);
        for (int idx = tid; idx < (K_ROWS - kv) * 16; idx += AT_THREADS)
            *reinterpret_cast<uint4*>(smem + L::K_OFF + (kv + (idx >> 4)) * ROWB + (idx & 15) * 16) = z;
        for (int idx = tid; idx < (V_ROWS - kv) * 16; idx += AT_THREADS)
            *reinterpret_cast<uint4*>(smem + L::V_OFF + (kv + (idx >> 4)) * ROWB + (idx & 15) * 16) = z;
        for (int idx = tid; idx < (P_ROWS - npos) * 16; idx += AT_THREADS)
            *reinterpret_cast<uint4*>(smem + L::P_OFF + (npos + (idx >> 4)) * ROWB + (idx & 15) * 16) = z;
    }
    // Q + u, Q + v (fp32 add, one rounding to fp16); rows >= T are zero
    for (int idx = tid; idx < 16 * 16; idx += AT_THREADS) {
        const int r = idx >> 4, c = idx & 15;
        uint4 ou = make_uint4(0, 0, 0, 0), ov = ou;
        if (r < T) {
            const uint4 q8 = *reinterpret_cast<const uint4*>(qkv + (size_t)r * (3 * D_MODEL) + c * 8);
            const __half2* qh = reinterpret_cast<const __half2*>(&q8);
            const float4 u0 = *reinterpret_cast<const float4*>(a.bias_u + h * D_HEAD + c * 8), u1 = *reinterpret_cast<const float4*>(a.bias_u + h * D_HEAD + c * 8 + 4);
            const float4 v0 = *reinterpret_cast<const float4*>(a.bias_v + h * D_HEAD + c * 8), v1 = *reinterpret_cast<const float4*>(a.bias_v + h * D_HEAD + c * 8 + 4);
            const float uu[8] = {u0.x, u0.y, u0.z, u0.w, u1.x, u1.y, u1.z, u1.w};
            const float vv[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
            __half2 hu[4], hv[4];
#pragma unroll
            for (int e = 0; e < 4; e++) {
                const float2 q2 = __half22float2(qh[e]);
                hu[e] = __floats2half2_rn(q2.x + uu[2 * e], q2.y + uu[2 * e + 1]);
                hv[e] = __floats2half2_rn(q2.x + vv[2 * e], q2.y + vv[2 * e + 1]);
            }
            ou = *reinterpret_cast<uint4*>(hu);
            ov = *reinterpret_cast<uint4*>(hv);
        }
        *reinterpret_cast<uint4*>(smem + L::QU_OFF + r * ROWB + c * 16) = ou;
        *reinterpret_cast<uint4*>(smem + L::QV_OFF + r * ROWB + c * 16) = ov;
    }
    cp_wait<1>();  // K and P have landed (this thread's copies); the barrier makes everyone's visible
    __syncthreads();

    // ---- append this chunk's K rows to the ring (V follows after its group lands) ----
    for (int idx = tid; idx < T * 16; idx += AT_THREADS) {
        const int i = idx >> 4, c = idx & 15;
        int pos = wpos + i;
        if (pos >= RING) pos -= RING;
        *reinterpret_cast<uint4*>(kring + (size_t)pos * D_MODEL + c * 8) = *reinterpret_cast<const uint4*>(smem + L::K_OFF + (LEFT + i) * ROWB + c * 16);
    }

    // ---- scores: 8-key tiles of (Q+u) K^T, then 8-position tiles of (Q+v) P^T, round-robin over the warps ----
    float* S1 = reinterpret_cast<float*>(smem + L::S1_OFF);
    float* S2 = reinterpret_cast<float*>(smem + L::S2_OFF);
    {
        const int a_row = (lane & 7) + ((lane >> 3) & 1) * 8, a_col = (lane >> 4) * 16;  // ldmatrix address roles (A operand)
        const int b_row = lane & 7, b_col = (lane >> 3) * 16;                              // B operand: 4 k-chunks of one 8-row tile
        for (int tile = warp; tile < kvt + npt; tile += 8) {
            const bool is_key = tile < kvt;
            const int n0 = 8 * (is_key ? tile : tile - kvt);
            const uint32_t qa = sb + (is_key ? L::QU_OFF : L::QV_OFF) + a_row * ROWB + a_col;
            const uint32_t kb = sb + (is_key ? L::K_OFF : L::P_OFF) + (n0 + b_row) * ROWB + b_col;
            float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int ks = 0; ks < 4; ks++) {  // 32 features per iteration
                uint32_t bf[4], a0[4], a1[4];
                ldsm_x4(bf, kb + ks * 64);
                ldsm_x4(a0, qa + ks * 64);
                ldsm_x4(a1, qa + ks * 64 + 32);
                mma_f16(acc, a0, bf[0], bf[1]);
                mma_f16(acc, a1, bf[2], bf[3]);
            }
            float* S = is_key ? S1 : S2;
            const int ld = is_key ? S1_LD : S2_LD;
            *reinterpret_cast<float2*>(S + g * ld + n0 + 2 * t) = make_float2(acc[0], acc[1]);
            *reinterpret_cast<float2*>(S + (g + 8) * ld + n0 + 2 * t) = make_float2(acc[2], acc[3]);
        }
    }
    __syncthreads();

    // ---- rel-shift by indexing, scale, mask, softmax (fp32): one warp per query row -> fp16 probabilities ----
    {
        __half* Pr = reinterpret_cast<__half*>(smem + L::PR_OFF);
        const int masked = LEFT - a.valid[b];
        const float scale = 1.0f / sqrtf((float)D_HEAD);
        for (int i = warp; i < 16; i += 8) {
            __half* prow_out = Pr + i * PR_LD;
            if (i >= T) {
                for (int j = lane; j < PR_LD; j += 32) prow_out[j] = __float2half_rn(0.0f);
                continue;
            }
            const float* row = S1 + i * S1_LD;
            const float* prow = S2 + i * S2_LD + (LEFT + i + T - 1);
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
                prow_out[j] = __float2half_rn(sc[q] * inv);
            }
            if (lane < PR_LD - 96) prow_out[96 + lane] = __float2half_rn(0.0f);
        }
    }
    cp_wait<0>();  // V has landed
    __syncthreads();

    for (int idx = tid; idx < T * 16; idx += AT_THREADS) {
        const int i = idx >> 4, c = idx & 15;
        int pos = wpos + i;
        if (pos >= RING) pos -= RING;
        *reinterpret_cast<uint4*>(vring + (size_t)pos * D_MODEL + c * 8) = *reinterpret_cast<const uint4*>(smem + L::V_OFF + (LEFT + i) * ROWB + c * 16);
    }

    // ---- context: warp w owns features [16w, 16w+16); k runs over the keys in steps of 16 ----
    float* Cx = reinterpret_cast<float*>(smem + L::K_OFF);
    {
        float acc0[4] = {0.f, 0.f, 0.f, 0.f}, acc1[4] = {0.f, 0.f, 0.f, 0.f};
        const int a_row = (lane & 7) + ((lane >> 3) & 1) * 8, a_col = (lane >> 4) * 16;
        const uint32_t pa = sb + L::PR_OFF + a_row * (PR_LD * 2) + a_col;
        // V^T fragments via ldmatrix.trans: matrices (keys k0..+7 | k0+8..+15) x (features n0..+7 | n0+8..+15)
        const uint32_t vb = sb + L::V_OFF + ((lane & 7) + ((lane >> 3) & 1) * 8) * ROWB + (16 * warp + (lane >> 4) * 8) * 2;
        for (int ks = 0; ks < kvs; ks++) {
            uint32_t af[4], bf[4];
            ldsm_x4(af, pa + ks * 32);
            ldsm_x4_t(bf, vb + ks * 16 * ROWB);
            mma_f16(acc0, af, bf[0], bf[1]);
            mma_f16(acc1, af, bf[2], bf[3]);
        }
        // K rows are dead (every warp passed the barrier after the scores); stage the fp32 context over them
        *reinterpret_cast<float2*>(Cx + g * CX_LD + 16 * warp + 2 * t) = make_float2(acc0[0], acc0[1]);
        *reinterpret_cast<float2*>(Cx + (g + 8) * CX_LD + 16 * warp + 2 * t) = make_float2(acc0[2], acc0[3]);
        *reinterpret_cast<float2*>(Cx + g * CX_LD + 16 * warp + 8 + 2 * t) = make_float2(acc1[0], acc1[1]);
        *reinterpret_cast<float2*>(Cx + (g + 8) * CX_LD + 16 * warp + 8 + 2 * t) = make_float2(acc1[2], acc1[3]);
    }
    __syncthreads();
    // ---- fp16 store (+ per-32-feature re-quantisation): a warp covers one row, 8 lanes = one 32-feature block ----
    for (int idx = tid; idx < T * (D_HEAD / 4); idx += AT_THREADS) {
        const int i = idx >> 5, d = (idx & 31) * 4;
        float4 v = *reinterpret_cast<const float4*>(Cx + i * CX_LD + d);
        if (a.act_quant) {
            const float amax = group8_max(fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w))));
            v.x = fake_quant(v.x, amax);
            v.y = fake_quant(v.y, amax);
            v.z = fake_quant(v.z, amax);
            v.w = fake_quant(v.w, amax);
        }
        uint2 u;
        u.x = pack_operand2(v.x, v.y, a.out_bf16);
        u.y = pack_operand2(v.z, v.w, a.out_bf16);
        *reinterpret_cast<uint2*>(a.out + ((size_t)(b * T + i)) * D_MODEL + h * D_HEAD + d) = u;
    }
}

// ---- single-wave variant (default for T > 2) --------------------------------------------------------------------------
// Same math as attention_f16_kernel with 51 KB of shared memory instead of 103 KB, so that four CTAs fit on an SM and the
// 512 (head, stream) items of the 64-stream step are ONE wave instead of 1.7:
//   * the projected positions are not staged: their MMA B fragments come straight from global memory (the table is shared
//     by all streams and stays in L2). Lane (g, t) loads the 16 contiguous bytes P[n0 + g][32 ks + 8 t .. + 8): a dot product
//     does not care in which order k is summed, so the 16 k-slots of an MMA are mapped to features 8 t + {0..3} (first MMA
//     of a 32-feature block) and 8 t + {4..7} (second), and the A fragments of Q + v are read with the same map;
//   * K and V share one buffer: V is fetched (after an L2 prefetch issued at the start) once every warp is done with K,
//     under the softmax.
struct AttSmemS {
    static constexpr int KV_ROWS = 96;
    static constexpr int KV_OFF = 0;
    static constexpr int QU_OFF = KV_OFF + KV_ROWS * ROWB;
    static constexpr int QV_OFF = QU_OFF + 16 * ROWB;
    static constexpr int S1_OFF = QV_OFF + 16 * ROWB;
    static constexpr int S2_OFF = S1_OFF + 16 * S1_LD * 4;
    static constexpr int PR_OFF = S2_OFF + 16 * S2_LD * 4;
    static constexpr int TOTAL = PR_OFF + 16 * PR_LD * 2;
    static_assert(16 * CX_LD * 4 <= 16 * S1_LD * 4 + 16 * S2_LD * 4, "context staging must fit the score area");
    static_assert(TOTAL <= 52 * 1024, "four CTAs per SM");
};

__global__ void __launch_bounds__(AT_THREADS, 4) attention_f16s_kernel(AttnArgs a) {
    pdl_launch();
    extern __shared__ __align__(16) uint8_t smem[];
    using L = AttSmemS;
    const int h = blockIdx.x, b = blockIdx.y, T = a.T;
    const int kv = LEFT + T, npos = LEFT + 2 * T - 1;
    const int kvt = (kv + 7) >> 3, npt = (npos + 7) >> 3, kvs = (kv + 15) >> 4;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const int slot = a.slot[b], wpos = a.wpos[b];
    int start = wpos - LEFT;
    if (start < 0) start += RING;
    __half* kring = a.k_ring + (size_t)slot * RING * D_MODEL + h * D_HEAD;
    __half* vring = a.v_ring + (size_t)slot * RING * D_MODEL + h * D_HEAD;
    const __half* qkv = a.qkv + (size_t)b * T * (3 * D_MODEL) + h * D_HEAD;
    const __half* pbase = a.pos + h * D_HEAD;
    const uint32_t sb = s_u32(smem);

    // cached K rows (written by earlier steps): requested before griddepcontrol.wait; cached V rows: pulled into L2 now
    for (int idx = tid; idx < LEFT * 16; idx += AT_THREADS) {
        const int j = idx >> 4, c = idx & 15;
        int pos = start + j;
        if (pos >= RING) pos -= RING;
        cp_async16(sb + L::KV_OFF + j * ROWB + c * 16, kring + (size_t)pos * D_MODEL + c * 8);
    }
    for (int idx = tid; idx < LEFT * 2; idx += AT_THREADS) {
        int pos = start + (idx >> 1);
        if (pos >= RING) pos -= RING;
        asm volatile("prefetch.global.L2 [%0];" ::"l"(vring + (size_t)pos * D_MODEL + (idx & 1) * 64));
    }
    {   // rows kv .. 95 stay zero for both operands (0 * garbage must stay finite in P.V)
        const uint4 z = make_uint4(0, 0, 0, 0);
        for (int idx = tid; idx < (L::KV_ROWS - kv) * 16; idx += AT_THREADS)
            *reinterpret_cast<uint4*>(smem + L::KV_OFF + (kv + (idx >> 4)) * ROWB + (idx & 15) * 16) = z;
    }
    pdl_wait();
    for (int idx = tid; idx < T * 16; idx += AT_THREADS) {
        const int i = idx >> 4, c = idx & 15;
        cp_async16(sb + L::KV_OFF + (LEFT + i) * ROWB + c * 16, qkv + (size_t)i * (3 * D_MODEL) + D_MODEL + c * 8);
    }
    cp_commit();
    // Q + u, Q + v (fp32 add, one rounding to fp16); rows >= T are zero
    for (int idx = tid; idx < 16 * 16; idx += AT_THREADS) {
        const int r = idx >> 4, c = idx & 15;
        uint4 ou = make_uint4(0, 0, 0, 0), ov = ou;
        if (r < T) {
            const uint4 q8 = *reinterpret_cast<const uint4*>(qkv + (size_t)r * (3 * D_MODEL) + c * 8);
            const __half2* qh = reinterpret_cast<const __half2*>(&q8);
            const float4 u0 = *reinterpret_cast<const float4*>(a.bias_u + h * D_HEAD + c * 8), u1 = *reinterpret_cast<const float4*>(a.bias_u + h * D_HEAD + c * 8 + 4);
            const float4 v0 = *reinterpret_cast<const float4*>(a.bias_v + h * D_HEAD + c * 8), v1 = *reinterpret_cast<const float4*>(a.bias_v + h * D_HEAD + c * 8 + 4);
            const float uu[8] = {u0.x, u0.y, u0.z, u0.w, u1.x, u1.y, u1.z, u1.w};
            const float vv[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
            __half2 hu[4], hv[4];
#pragma unroll
            for (int e = 0; e < 4; e++) {
                const float2 q2 = __half22float2(qh[e]);
                hu[e] = __floats2half2_rn(q2.x + uu[2 * e], q2.y + uu[2 * e + 1]);
                hv[e] = __floats2half2_rn(q2.x + vv[2 * e], q2.y + vv[2 * e + 1]);
            }
            ou = *reinterpret_cast<uint4*>(hu);
            ov = *reinterpret_cast<uint4*>(hv);
        }
        *reinterpret_cast<uint4*>(smem + L::QU_OFF + r * ROWB + c * 16) = ou;
        *reinterpret_cast<uint4*>(smem + L::QV_OFF + r * ROWB + c * 16) = ov;
    }
    __syncthreads();  // Q + u, Q + v visible

    float* S1 = reinterpret_cast<float*>(smem + L::S1_OFF);
    float* S2 = reinterpret_cast<float*>(smem + L::S2_OFF);
    // ---- position scores (Q + v) P^T, 8 positions per tile, B fragments from global with the permuted k map ----
    for (int tile = warp; tile < npt; tile += 8) {
        const int n0 = 8 * tile;
        const int prow = min(n0 + g, npos - 1);  // rows beyond the table are never read back: any finite row will do
        const __half* pp = pbase + (size_t)prow * D_MODEL + 8 * t;
        uint4 pf[4];
#pragma unroll
        for (int ks = 0; ks < 4; ks++) pf[ks] = __ldg(reinterpret_cast<const uint4*>(pp + 32 * ks));
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int ks = 0; ks < 4; ks++) {
            const uint4 q0 = *reinterpret_cast<const uint4*>(smem + L::QV_OFF + g * ROWB + (32 * ks + 8 * t) * 2);
            const uint4 q1 = *reinterpret_cast<const uint4*>(smem + L::QV_OFF + (g + 8) * ROWB + (32 * ks + 8 * t) * 2);
            const uint32_t a0[4] = {q0.x, q1.x, q0.y, q1.y}, a1[4] = {q0.z, q1.z, q0.w, q1.w};
            mma_f16(acc, a0, pf[ks].x, pf[ks].y);
            mma_f16(acc, a1, pf[ks].z, pf[ks].w);
        }
        *reinterpret_cast<float2*>(S2 + g * S2_LD + n0 + 2 * t) = make_float2(acc[0], acc[1]);
        *reinterpret_cast<float2*>(S2 + (g + 8) * S2_LD + n0 + 2 * t) = make_float2(acc[2], acc[3]);
    }
    cp_wait<0>();  // K has landed (this thread's copies); the barrier makes everyone's visible
    __syncthreads();

    // ---- append this chunk's K rows to the ring ----
    for (int idx = tid; idx < T * 16; idx += AT_THREADS) {
        const int i = idx >> 4, c = idx & 15;
        int pos = wpos + i;
        if (pos >= RING) pos -= RING;
        *reinterpret_cast<uint4*>(kring + (size_t)pos * D_MODEL + c * 8) = *reinterpret_cast<const uint4*>(smem + L::KV_OFF + (LEFT + i) * ROWB + c * 16);
    }
    // ---- key scores (Q + u) K^T ----
    {
        const int a_row = (lane & 7) + ((lane >> 3) & 1) * 8, a_col = (lane >> 4) * 16;
        const int b_row = lane & 7, b_col = (lane >> 3) * 16;
        for (int tile = warp; tile < kvt; tile += 8) {
            const int n0 = 8 * tile;
            const uint32_t qa = sb + L::QU_OFF + a_row * ROWB + a_col;
            const uint32_t kb = sb + L::KV_OFF + (n0 + b_row) * ROWB + b_col;
            float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int ks = 0; ks < 4; ks++) {
                uint32_t bf[4], a0[4], a1[4];
                ldsm_x4(bf, kb + ks * 64);
                ldsm_x4(a0, qa + ks * 64);
                ldsm_x4(a1, qa + ks * 64 + 32);
                mma_f16(acc, a0, bf[0], bf[1]);
                mma_f16(acc, a1, bf[2], bf[3]);
            }
            *reinterpret_cast<float2*>(S1 + g * S1_LD + n0 + 2 * t) = make_float2(acc[0], acc[1]);
            *reinterpret_cast<float2*>(S1 + (g + 8) * S1_LD + n0 + 2 * t) = make_float2(acc[2], acc[3]);
        }
    }
    __syncthreads();  // every warp is done with K: the buffer now receives V

    for (int idx = tid; idx < kv * 16; idx += AT_THREADS) {
        const int j = idx >> 4, c = idx & 15;
        const __half* src;
        if (j < LEFT) {
            int pos = start + j;
            if (pos >= RING) pos -= RING;
            src = vring + (size_t)pos * D_MODEL + c * 8;
        } else {
            src = qkv + (size_t)(j - LEFT) * (3 * D_MODEL) + 2 * D_MODEL + c * 8;
        }
        cp_async16(sb + L::KV_OFF + j * ROWB + c * 16, src);
    }
    cp_commit();

    // ---- rel-shift by indexing, scale, mask, softmax (fp32): one warp per query row -> fp16 probabilities ----
    {
        __half* Pr = reinterpret_cast<__half*>(smem + L::PR_OFF);
        const int masked = LEFT - a.valid[b];
        const float scale = 1.0f / sqrtf((float)D_HEAD);
        for (int i = warp; i < 16; i += 8) {
            __half* prow_out = Pr + i * PR_LD;
            if (i >= T) {
                for (int j = lane; j < PR_LD; j += 32) prow_out[j] = __float2half_rn(0.0f);
                continue;
            }
            const float* row = S1 + i * S1_LD;
            const float* prow = S2 + i * S2_LD + (LEFT + i + T - 1);
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
                prow_out[j] = __float2half_rn(sc[q] * inv);
            }
            if (lane < PR_LD - 96) prow_out[96 + lane] = __float2half_rn(0.0f);
        }
    }
    cp_wait<0>();  // V has landed
    __syncthreads();  // ... and the probabilities are complete; S1 / S2 are dead from here on

    for (int idx = tid; idx < T * 16; idx += AT_THREADS) {
        const int i = idx >> 4, c = idx & 15;
        int pos = wpos + i;
        if (pos >= RING) pos -= RING;
        *reinterpret_cast<uint4*>(vring + (size_t)pos * D_MODEL + c * 8) = *reinterpret_cast<const uint4*>(smem + L::KV_OFF + (LEFT + i) * ROWB + c * 16);
    }
    // ---- context: warp w owns features [16w, 16w+16); k runs over the keys in steps of 16 ----
    float* Cx = reinterpret_cast<float*>(smem + L::S1_OFF);
    {
        float acc0[4] = {0.f, 0.f, 0.f, 0.f}, acc1[4] = {0.f, 0.f, 0.f, 0.f};
        const int a_row = (lane & 7) + ((lane >> 3) & 1) * 8, a_col = (lane >> 4) * 16;
        const uint32_t pa = sb + L::PR_OFF + a_row * (PR_LD * 2) + a_col;
        const uint32_t vb = sb + L::KV_OFF + ((lane & 7) + ((lane >> 3) & 1) * 8) * ROWB + (16 * warp + (lane >> 4) * 8) * 2;
        for (int ks = 0; ks < kvs; ks++) {
            uint32_t af[4], bf[4];
            ldsm_x4(af, pa + ks * 32);
            ldsm_x4_t(bf, vb + ks * 16 * ROWB);
            mma_f16(acc0, af, bf[0], bf[1]);
            mma_f16(acc1, af, bf[2], bf[3]);
        }
        *reinterpret_cast<float2*>(Cx + g * CX_LD + 16 * warp + 2 * t) = make_float2(acc0[0], acc0[1]);
        *reinterpret_cast<float2*>(Cx + (g + 8) * CX_LD + 16 * warp + 2 * t) = make_float2(acc0[2], acc0[3]);
        *reinterpret_cast<float2*>(Cx + g * CX_LD + 16 * warp + 8 + 2 * t) = make_float2(acc1[0], acc1[1]);
        *reinterpret_cast<float2*>(Cx + (g + 8) * CX_LD + 16 * warp + 8 + 2 * t) = make_float2(acc1[2], acc1[3]);
    }
    __syncthreads();
    for (int idx = tid; idx < T * (D_HEAD / 4); idx += AT_THREADS) {
        const int i = idx >> 5, d = (idx & 31) * 4;
        float4 v = *reinterpret_cast<const float4*>(Cx + i * CX_LD + d);
        if (a.act_quant) {
            const float amax = group8_max(fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w))));
            v.x = fake_quant(v.x, amax);
            v.y = fake_quant(v.y, amax);
            v.z = fake_quant(v.z, amax);
            v.w = fake_quant(v.w, amax);
        }
        uint2 u;
        u.x = pack_operand2(v.x, v.y, a.out_bf16);
        u.y = pack_operand2(v.z, v.w, a.out_bf16);
        *reinterpret_cast<uint2*>(a.out + ((size_t)(b * T + i)) * D_MODEL + h * D_HEAD + d) = u;
    }
}

// ---- streaming kernel for the low-latency modes (T <= 2: right_context 0 and 1) -----------------------------------
// With one or two query frames there is nothing for a tensor core to do (the MMA kernel pads to 16 query rows and
// stages 100 KB per CTA, two CTAs per SM): the work is reading 70 K rows, 70 V rows and ~71 position rows per (stream,
// head) exactly once. Here one WARP owns a (stream, head): every load is a fully coalesced 512-byte warp access (two
// 256-byte rows, 16 lanes each, 16 B per lane) straight from global memory into registers, no shared-memory staging
// of operands, ~24 warps resident per SM for latency hiding. Lane c of a half-warp holds features [8c, 8c+8) of q+u /
// q+v in fp32; a row's dot product is 8 FMAs + a 4-step shuffle reduction over the half-warp. Probabilities stay fp32.
// UN = rows in flight per half-warp and phase (UN x 512 B per warp). Each phase is a chain of ceil(71 / 2 UN) dependent memory round
// trips: 6 at UN = 6. With up to ~256 streams there are only a few warps per SM to hide them (7 at 128 streams), so the small-batch
// instantiation keeps 12 rows in flight (3 round trips per phase; 48 more registers, irrelevant at that occupancy) and requests
// the first V rows before the softmax; the large-batch one (1024 streams: 55 warps per SM, 73 % of the HBM peak) keeps 6.
template <int NQ, int UN>
__global__ void __launch_bounds__(128) attention_stream_kernel(AttnArgs a) {
    pdl_launch();
    constexpr int WARPS = 4, SLD = 96;
    __shared__ float S[WARPS][NQ][SLD];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, hw = lane >> 4, c = lane & 15;
    const int item = blockIdx.x * WARPS + warp;
    const int h = item & (N_HEADS - 1), b = item >> 3;  // the 4 warps of a CTA read 4 adjacent heads of the same rows
    if (b >= a.B) return;
    const int T = a.T, kv = LEFT + T, npos = LEFT + 2 * T - 1;
    const int slot = a.slot[b], wpos = a.wpos[b];
    int start = wpos - LEFT;
    if (start < 0) start += RING;
    __half* kring = a.k_ring + (size_t)slot * RING * D_MODEL + h * D_HEAD + c * 8;
    __half* vring = a.v_ring + (size_t)slot * RING * D_MODEL + h * D_HEAD + c * 8;
    const __half* pbase = a.pos + h * D_HEAD + c * 8;
    float (*Sw)[SLD] = S[warp];

    // The position table was written at load time: its first batch of rows is requested before the dependency wait (it lands
    // while the QKV GEMM finishes); everything that depends on qkv follows the wait.
    uint4 rp[UN];
#pragma unroll
    for (int u = 0; u < UN; u++) {
        const int rr = 2 * u + hw;
        rp[u] = rr < npos ? __ldg(reinterpret_cast<const uint4*>(pbase + (size_t)rr * D_MODEL)) : make_uint4(0, 0, 0, 0);
    }
    pdl_wait();
    const __half* qkv = a.qkv + (size_t)b * T * (3 * D_MODEL) + h * D_HEAD + c * 8;
    float qu[NQ][8], qv[NQ][8];
    {
        const float4 u0 = *reinterpret_cast<const float4*>(a.bias_u + h * D_HEAD + c * 8), u1 = *reinterpret_cast<const float4*>(a.bias_u + h * D_HEAD + c * 8 + 4);
        const float4 v0 = *reinterpret_cast<const float4*>(a.bias_v + h * D_HEAD + c * 8), v1 = *reinterpret_cast<const float4*>(a.bias_v + h * D_HEAD + c * 8 + 4);
        const float uu[8] = {u0.x, u0.y, u0.z, u0.w, u1.x, u1.y, u1.z, u1.w};
        const float vv[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
        for (int i = 0; i < NQ; i++) {
            uint4 q8 = make_uint4(0, 0, 0, 0);
            if (i < T) q8 = *reinterpret_cast<const uint4*>(qkv + (size_t)i * (3 * D_MODEL));
            const __half2* qh = reinterpret_cast<const __half2*>(&q8);
#pragma unroll
            for (int e = 0; e < 4; e++) {
                const float2 f = __half22float2(qh[e]);
                qu[i][2 * e] = f.x + uu[2 * e]; qu[i][2 * e + 1] = f.y + uu[2 * e + 1];
                qv[i][2 * e] = f.x + vv[2 * e]; qv[i][2 * e + 1] = f.y + vv[2 * e + 1];
            }
        }
    }
    auto dot8 = [](const uint4& r, const float (&q)[8]) {
        const __half2* rh = reinterpret_cast<const __half2*>(&r);
        float s = 0.0f;
#pragma unroll
        for (int e = 0; e < 4; e++) {
            const float2 f = __half22float2(rh[e]);
            s = fmaf(f.x, q[2 * e], s);
            s = fmaf(f.y, q[2 * e + 1], s);
        }
        return s;
    };
    auto half_sum = [](float s) {  // sum over the 16 lanes of a half-warp
        s += __shfl_xor_sync(0xffffffffu, s, 8);
        s += __shfl_xor_sync(0xffffffffu, s, 4);
        s += __shfl_xor_sync(0xffffffffu, s, 2);
        s += __shfl_xor_sync(0xffffffffu, s, 1);
        return s;
    };
    const uint4 z4 = make_uint4(0, 0, 0, 0);

    // ---- key scores (Q+u).K_j; this chunk's key rows come from qkv and are appended to the ring on the way ----
    for (int j0 = 0; j0 < kv; j0 += 2 * UN) {
        uint4 r[UN];
#pragma unroll
        for (int u = 0; u < UN; u++) {
            const int j = j0 + 2 * u + hw;
            r[u] = z4;
            if (j < LEFT) {
                int pos = start + j;
                if (pos >= RING) pos -= RING;
                r[u] = *reinterpret_cast<const uint4*>(kring + (size_t)pos * D_MODEL);
            } else if (j < kv) {
                r[u] = *reinterpret_cast<const uint4*>(qkv + (size_t)(j - LEFT) * (3 * D_MODEL) + D_MODEL);
                int pos = wpos + (j - LEFT);
                if (pos >= RING) pos -= RING;
                *reinterpret_cast<uint4*>(kring + (size_t)pos * D_MODEL) = r[u];
            }
        }
#pragma unroll
        for (int u = 0; u < UN; u++) {
            const int j = j0 + 2 * u + hw;
#pragma unroll
            for (int i = 0; i < NQ; i++) {
                const float d = half_sum(dot8(r[u], qu[i]));
                if (c == 0 && j < kv) Sw[i][j] = d;
            }
        }
    }
    __syncwarp();
    // ---- position scores (Q+v).P_r, added at key j = LEFT + i + T - 1 - r (the reference's rel-shift) ----
    for (int r0 = 0; r0 < npos; r0 += 2 * UN) {
        if (r0 > 0) {
#pragma unroll
            for (int u = 0; u < UN; u++) {
                const int rr = r0 + 2 * u + hw;
                rp[u] = rr < npos ? __ldg(reinterpret_cast<const uint4*>(pbase + (size_t)rr * D_MODEL)) : z4;
            }
        }
#pragma unroll
        for (int u = 0; u < UN; u++) {
            const int rr = r0 + 2 * u + hw;
#pragma unroll
            for (int i = 0; i < NQ; i++) {
                const float d = half_sum(dot8(rp[u], qv[i]));
                const int j = LEFT + i + T - 1 - rr;
                if (c == 0 && i < T && rr < npos && j >= 0 && j < kv) Sw[i][j] += d;
            }
        }
    }
    __syncwarp();
    // the first batch of V rows is requested now: it is in flight while the softmax runs
    auto load_v = [&](int j0, uint4 (&r)[UN]) {
#pragma unroll
        for (int u = 0; u < UN; u++) {
            const int j = j0 + 2 * u + hw;
            r[u] = z4;
            if (j < LEFT) {
                int pos = start + j;
                if (pos >= RING) pos -= RING;
                r[u] = *reinterpret_cast<const uint4*>(vring + (size_t)pos * D_MODEL);
            } else if (j < kv) {
                r[u] = *reinterpret_cast<const uint4*>(qkv + (size_t)(j - LEFT) * (3 * D_MODEL) + 2 * D_MODEL);
                int pos = wpos + (j - LEFT);
                if (pos >= RING) pos -= RING;
                *reinterpret_cast<uint4*>(vring + (size_t)pos * D_MODEL) = r[u];
            }
        }
    };
    uint4 rv[UN];
    load_v(0, rv);
    // ---- scale, mask, softmax (fp32) ----
    {
        const int masked = LEFT - a.valid[b];
        const float scale = 1.0f / sqrtf((float)D_HEAD);
#pragma unroll
        for (int i = 0; i < NQ; i++) {
            if (i >= T) break;
            float sc[3];
            float mx = -INFINITY;
#pragma unroll
            for (int q = 0; q < 3; q++) {
                const int j = lane + 32 * q;
                float s = -INFINITY;
                if (j < kv) {
                    s = Sw[i][j] * scale;
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
                if (j < SLD) Sw[i][j] = sc[q] * inv;
            }
        }
    }
    __syncwarp();
    // ---- context: each half-warp accumulates every other key row, lane c owns features [8c, 8c+8) ----
    float acc[NQ][8];
#pragma unroll
    for (int i = 0; i < NQ; i++)
#pragma unroll
        for (int e = 0; e < 8; e++) acc[i][e] = 0.0f;
    for (int j0 = 0; j0 < kv; j0 += 2 * UN) {
        if (j0 > 0) load_v(j0, rv);
#pragma unroll
        for (int u = 0; u < UN; u++) {
            const int j = j0 + 2 * u + hw;
            const __half2* rh = reinterpret_cast<const __half2*>(&rv[u]);
#pragma unroll
            for (int i = 0; i < NQ; i++) {
                const float p = (j < kv) ? Sw[i][j] : 0.0f;
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    const float2 f = __half22float2(rh[e]);
                    acc[i][2 * e] = fmaf(p, f.x, acc[i][2 * e]);
                    acc[i][2 * e + 1] = fmaf(p, f.y, acc[i][2 * e + 1]);
                }
            }
        }
    }
#pragma unroll
    for (int i = 0; i < NQ; i++) {
#pragma unroll
        for (int e = 0; e < 8; e++) acc[i][e] += __shfl_xor_sync(0xffffffffu, acc[i][e], 16);
        if (a.act_quant) {  // 32 consecutive features = 4 adjacent lanes
            float amax = 0.0f;
#pragma unroll
            for (int e = 0; e < 8; e++) amax = fmaxf(amax, fabsf(acc[i][e]));
            amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
            amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
#pragma unroll
            for (int e = 0; e < 8; e++) acc[i][e] = fake_quant(acc[i][e], amax);
        }
        if (hw == 0 && i < T) {
            uint32_t o[4];
#pragma unroll
            for (int e = 0; e < 4; e++) o[e] = pack_operand2(acc[i][2 * e], acc[i][2 * e + 1], a.out_bf16);
            *reinterpret_cast<uint4*>(a.out + ((size_t)(b * T + i)) * D_MODEL + h * D_HEAD + c * 8) = *reinterpret_cast<const uint4*>(o);
        }
    }
}

// ---- fp32 CUDA-core checker (NB200_ATTN_SIMT=1): same inputs and outputs, no tensor cores, no staging tricks ----
__global__ void __launch_bounds__(256) attention_simt_kernel(AttnArgs a) {
    pdl_prologue();
    __shared__ float Sc[MAX_T * (LEFT + MAX_T)];
    const int h = blockIdx.x, b = blockIdx.y, T = a.T, kv = LEFT + T, tid = threadIdx.x;
    const int slot = a.slot[b], wpos = a.wpos[b];
    int start = wpos - LEFT;
    if (start < 0) start += RING;
    __half* kring = a.k_ring + (size_t)slot * RING * D_MODEL + h * D_HEAD;
    __half* vring = a.v_ring + (size_t)slot * RING * D_MODEL + h * D_HEAD;
    const __half* qkv = a.qkv + (size_t)b * T * (3 * D_MODEL) + h * D_HEAD;
    for (int idx = tid; idx < T * D_HEAD; idx += 256) {  // append first: keys >= 70 are then read from the ring as well
        const int i = idx >> 7, d = idx & 127;
        int pos = wpos + i;
        if (pos >= RING) pos -= RING;
        kring[(size_t)pos * D_MODEL + d] = qkv[(size_t)i * 3 * D_MODEL + D_MODEL + d];
        vring[(size_t)pos * D_MODEL + d] = qkv[(size_t)i * 3 * D_MODEL + 2 * D_MODEL + d];
    }
    __syncthreads();
    const int masked = LEFT - a.valid[b];
    const float scale = 1.0f / sqrtf((float)D_HEAD);
    for (int idx = tid; idx < T * kv; idx += 256) {
        const int i = idx / kv, j = idx - i * kv;
        int pos = start + j;
        if (pos >= RING) pos -= RING;
        const __half* kr = kring + (size_t)pos * D_MODEL;
        const __half* pr = a.pos + (size_t)(LEFT + i - j + T - 1) * D_MODEL + h * D_HEAD;
        const __half* q = qkv + (size_t)i * 3 * D_MODEL;
        float s1 = 0.0f, s2 = 0.0f;
        for (int d = 0; d < D_HEAD; d++) {
            const float qf = __half2float(q[d]);
            const float qu = __half2float(__float2half_rn(qf + a.bias_u[h * D_HEAD + d]));
            const float qv = __half2float(__float2half_rn(qf + a.bias_v[h * D_HEAD + d]));
            s1 = fmaf(qu, __half2float(kr[d]), s1);
            s2 = fmaf(qv, __half2float(pr[d]), s2);
        }
        float s = (s1 + s2) * scale;
        if (j < masked) s += -1e9f;
        Sc[idx] = s;
    }
    __syncthreads();
    const int warp = tid >> 5, lane = tid & 31;
    for (int i = warp; i < T; i += 8) {
        float* row = Sc + i * kv;
        float mx = -INFINITY;
        for (int j = lane; j < kv; j += 32) mx = fmaxf(mx, row[j]);
        mx = warp_max(mx);
        float sum = 0.0f;
        for (int j = lane; j < kv; j += 32) {
            const float e = expf(row[j] - mx);
            row[j] = e;
            sum += e;
        }
        sum = warp_sum(sum);
        const float inv = 1.0f / sum;
        for (int j = lane; j < kv; j += 32) row[j] = __half2float(__float2half_rn(row[j] * inv));
    }
    __syncthreads();
    for (int idx = tid; idx < T * (D_HEAD / 4); idx += 256) {
        const int i = idx >> 5, d = (idx & 31) * 4;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int j = 0; j < kv; j++) {
            int pos = start + j;
            if (pos >= RING) pos -= RING;
            const __half* vr = vring + (size_t)pos * D_MODEL + d;
            const float p = Sc[i * kv + j];
            acc.x = fmaf(p, __half2float(vr[0]), acc.x);
            acc.y = fmaf(p, __half2float(vr[1]), acc.y);
            acc.z = fmaf(p, __half2float(vr[2]), acc.z);
            acc.w = fmaf(p, __half2float(vr[3]), acc.w);
        }
        if (a.act_quant) {
            const float amax = group8_max(fmaxf(fmaxf(fabsf(acc.x), fabsf(acc.y)), fmaxf(fabsf(acc.z), fabsf(acc.w))));
            acc.x = fake_quant(acc.x, amax);
            acc.y = fake_quant(acc.y, amax);
            acc.z = fake_quant(acc.z, amax);
            acc.w = fake_quant(acc.w, amax);
        }
        uint2 u;
        u.x = pack_operand2(acc.x, acc.y, a.out_bf16);
        u.y = pack_operand2(acc.z, acc.w, a.out_bf16);
        *reinterpret_cast<uint2*>(a.out + ((size_t)(b * T + i)) * D_MODEL + h * D_HEAD + d) = u;
    }
}

// ---- offline path: full-sequence relative-position attention over one utterance -------------------------------------
// build_rel_pos_mha (src/nemo-ggml.cpp:668-757): scores[i][j] = ((q_i + u) . k_j + (q_i + v) . pos[i - j]) / sqrt(128), softmax over
// all T keys (no cache, no mask), context = P V. One CTA = 8 query rows of one head (a warp each); keys go through shared
// memory in tiles of 32 (lane = key) together with the 39 position rows the 8 x 32 (query, key) pairs of the tile need; the
// score rows (<= 512 keys, fp32) stay in shared memory for the softmax and the P V pass (lane = 4 output features).
// K, V and the projected positions are fp16 like in the streaming kernels; q + u, q + v, scores and probabilities are fp32.
constexpr int AF_QB = 8, AF_KT = 32, AF_PITCH = D_HEAD + 8;  // 272-byte rows: 16-byte row pieces of consecutive rows fall in different bank groups

__global__ void __launch_bounds__(256) attention_full_kernel(AttnFullArgs a) {
    pdl_prologue();
    __shared__ float S[AF_QB][ATTN_FULL_MAX_T];
    __shared__ __align__(16) __half Kt[AF_KT][AF_PITCH];
    __shared__ __align__(16) __half Pt[AF_KT + AF_QB - 1][AF_PITCH];
    __shared__ __align__(16) float qu[AF_QB][D_HEAD];
    __shared__ __align__(16) float qv[AF_QB][D_HEAD];
    const int T = a.T, h = blockIdx.y, i0 = blockIdx.x * AF_QB;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int i = i0 + warp;
    const __half* qkv = a.qkv + h * D_HEAD;
    const __half* pbase = a.pos + h * D_HEAD;
    {   // q + u, q + v of the CTA's 8 query rows (rows >= T: zero)
        float4 u4 = make_float4(0.f, 0.f, 0.f, 0.f), v4 = u4;
        if (i < T) {
            const uint2 raw = *reinterpret_cast<const uint2*>(qkv + (size_t)i * (3 * D_MODEL) + 4 * lane);
            const float2 q01 = __half22float2(*reinterpret_cast<const __half2*>(&raw.x)), q23 = __half22float2(*reinterpret_cast<const __half2*>(&raw.y));
            const float4 bu = *reinterpret_cast<const float4*>(a.bias_u + h * D_HEAD + 4 * lane);
            const float4 bv = *reinterpret_cast<const float4*>(a.bias_v + h * D_HEAD + 4 * lane);
            u4 = make_float4(q01.x + bu.x, q01.y + bu.y, q23.x + bu.z, q23.y + bu.w);
            v4 = make_float4(q01.x + bv.x, q01.y + bv.y, q23.x + bv.z, q23.y + bv.w);
        }
        *reinterpret_cast<float4*>(&qu[warp][4 * lane]) = u4;
        *reinterpret_cast<float4*>(&qv[warp][4 * lane]) = v4;
    }
    const float scale = 1.0f / sqrtf((float)D_HEAD);
    const uint4 z4 = make_uint4(0, 0, 0, 0);
    for (int j0 = 0; j0 < T; j0 += AF_KT) {
        __syncthreads();  // the previous tile has been consumed (first pass: q + u, q + v are visible after the barrier below)
        for (int idx = tid; idx < AF_KT * 16; idx += 256) {
            const int r = idx >> 4, c = idx & 15, j = j0 + r;
            *reinterpret_cast<uint4*>(&Kt[r][c * 8]) = j < T ? *reinterpret_cast<const uint4*>(qkv + (size_t)j * (3 * D_MODEL) + D_MODEL + c * 8) : z4;
        }
        // pair (query i0 + w, key j0 + l) uses position p = (i0 + w) - (j0 + l): table row p + T - 1 = rlo + (w - l + 31)
        const int rlo = i0 - j0 - (AF_KT - 1) + T - 1;
        for (int idx = tid; idx < (AF_KT + AF_QB - 1) * 16; idx += 256) {
            const int r = idx >> 4, c = idx & 15, row = rlo + r;
            *reinterpret_cast<uint4*>(&Pt[r][c * 8]) = (row >= 0 && row < 2 * T - 1) ? *reinterpret_cast<const uint4*>(pbase + (size_t)row * D_MODEL + c * 8) : z4;
        }
        __syncthreads();
        const int j = j0 + lane;
        if (i < T && j < T) {
            const __half* kr = &Kt[lane][0];
            const __half* pr = &Pt[warp - lane + AF_KT - 1][0];
            float sc = 0.0f, sp = 0.0f;
#pragma unroll 4
            for (int d = 0; d < D_HEAD; d += 8) {
                const uint4 k8 = *reinterpret_cast<const uint4*>(kr + d), p8 = *reinterpret_cast<const uint4*>(pr + d);
                const float4 u0 = *reinterpret_cast<const float4*>(&qu[warp][d]), u1 = *reinterpret_cast<const float4*>(&qu[warp][d + 4]);
                const float4 v0 = *reinterpret_cast<const float4*>(&qv[warp][d]), v1 = *reinterpret_cast<const float4*>(&qv[warp][d + 4]);
                const __half2* kh = reinterpret_cast<const __half2*>(&k8);
                const __half2* ph = reinterpret_cast<const __half2*>(&p8);
                const float2 k0 = __half22float2(kh[0]), k1 = __half22float2(kh[1]), k2 = __half22float2(kh[2]), k3 = __half22float2(kh[3]);
                const float2 p0 = __half22float2(ph[0]), p1 = __half22float2(ph[1]), p2 = __half22float2(ph[2]), p3 = __half22float2(ph[3]);
                sc = fmaf(u0.x, k0.x, sc); sc = fmaf(u0.y, k0.y, sc); sc = fmaf(u0.z, k1.x, sc); sc = fmaf(u0.w, k1.y, sc);
                sc = fmaf(u1.x, k2.x, sc); sc = fmaf(u1.y, k2.y, sc); sc = fmaf(u1.z, k3.x, sc); sc = fmaf(u1.w, k3.y, sc);
                sp = fmaf(v0.x, p0.x, sp); sp = fmaf(v0.y, p0.y, sp); sp = fmaf(v0.z, p1.x, sp); sp = fmaf(v0.w, p1.y, sp);
                sp = fmaf(v1.x, p2.x, sp); sp = fmaf(v1.y, p2.y, sp); sp = fmaf(v1.z, p3.x, sp); sp = fmaf(v1.w, p3.y, sp);
            }
            S[warp][j] = (sc + sp) * scale;
        }
    }
    // softmax over the T keys of the warp's row (ggml_soft_max: max-subtracted exp, normalised), fp32
    float inv = 0.0f;
    if (i < T) {
        float mx = -INFINITY;
        for (int j = lane; j < T; j += 32) mx = fmaxf(mx, S[warp][j]);
        mx = warp_max(mx);
        float sum = 0.0f;
        for (int j = lane; j < T; j += 32) {
            const float e = expf(S[warp][j] - mx);
            S[warp][j] = e;
            sum += e;
        }
        inv = 1.0f / warp_sum(sum);
    }
    // context: lane owns features 4 lane .. 4 lane + 3; V goes through the K tile buffer
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int j0 = 0; j0 < T; j0 += AF_KT) {
        __syncthreads();
        for (int idx = tid; idx < AF_KT * 16; idx += 256) {
            const int r = idx >> 4, c = idx & 15, j = j0 + r;
            *reinterpret_cast<uint4*>(&Kt[r][c * 8]) = j < T ? *reinterpret_cast<const uint4*>(qkv + (size_t)j * (3 * D_MODEL) + 2 * D_MODEL + c * 8) : z4;
        }
        __syncthreads();
        if (i < T) {
            const int n = min(AF_KT, T - j0);
            for (int jj = 0; jj < n; jj++) {
                const float p = S[warp][j0 + jj];
                const uint2 raw = *reinterpret_cast<const uint2*>(&Kt[jj][4 * lane]);
                const float2 v01 = __half22float2(*reinterpret_cast<const __half2*>(&raw.x)), v23 = __half22float2(*reinterpret_cast<const __half2*>(&raw.y));
                acc.x = fmaf(p, v01.x, acc.x); acc.y = fmaf(p, v01.y, acc.y); acc.z = fmaf(p, v23.x, acc.z); acc.w = fmaf(p, v23.y, acc.w);
            }
        }
    }
    if (i < T) {
        float4 v = make_float4(acc.x * inv, acc.y * inv, acc.z * inv, acc.w * inv);
        if (a.act_quant) {  // the warp is converged here: i < T is warp-uniform
            const float amax = group8_max(fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w))));
            v.x = fake_quant(v.x, amax);
            v.y = fake_quant(v.y, amax);
            v.z = fake_quant(v.z, amax);
            v.w = fake_quant(v.w, amax);
        }
        uint2 u;
        u.x = pack_operand2(v.x, v.y, a.out_bf16);
        u.y = pack_operand2(v.z, v.w, a.out_bf16);
        *reinterpret_cast<uint2*>(a.out + (size_t)i * D_MODEL + h * D_HEAD + 4 * lane) = u;
    }
}

// ---- offline path, tensor-core version: 16 query rows of one head per CTA, mma.sync.m16n8k16 for the three products ---------
// Per tile of 64 keys: K rows and the 79 position rows the 16 x 64 (query, key) pairs need are staged with cp.async; content
// scores (Q + u) K^T and position scores (Q + v) P^T run on the tensor cores (fp16 operands, fp32 accumulate, the fragments
// of the streaming kernel above); the rel-shift is the index i - jj + 63 into the tile's position scores. The whole score row
// (<= 512 keys, fp32) stays in shared memory for the softmax; P (fp16, unnormalised) . V follows tile by tile with the
// accumulators in registers. The CUDA-core kernel above is kept as the checker (NB200_ATTN_FULL_SIMT=1).
struct AfSmem {
    static constexpr int KT = 64, PROWS = 80;
    static constexpr int S_LD = 516, P2_LD = 84, PR_PITCH = 1040;  // floats, floats, bytes (1040 / 16 odd: conflict-free ldmatrix)
    static constexpr int QU_OFF = 0;
    static constexpr int QV_OFF = QU_OFF + 16 * ROWB;
    static constexpr int KV_OFF = QV_OFF + 16 * ROWB;
    static constexpr int POS_OFF = KV_OFF + KT * ROWB;
    static constexpr int S_OFF = POS_OFF + PROWS * ROWB;
    static constexpr int P2_OFF = S_OFF + 16 * S_LD * 4;
    static constexpr int INV_OFF = P2_OFF + 16 * P2_LD * 4;
    static constexpr int TOTAL = INV_OFF + 16 * 4;
    static constexpr int PR_OFF = POS_OFF;  // probabilities (fp16) over the position rows, dead after the scores
    static constexpr int CX_OFF = S_OFF;    // fp32 context over the score rows, dead after the softmax
    static_assert(16 * PR_PITCH <= PROWS * ROWB, "probabilities must fit the position area");
    static_assert(16 * CX_LD * 4 <= 16 * S_LD * 4, "context staging must fit the score area");
};

__global__ void __launch_bounds__(128) attention_full_mma_kernel(AttnFullArgs a) {
    pdl_prologue();
    extern __shared__ __align__(16) uint8_t smem[];
    using L = AfSmem;
    const int T = a.T, h = blockIdx.y, i0 = blockIdx.x * 16;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const __half* qkv = a.qkv + h * D_HEAD;
    const __half* pbase = a.pos + h * D_HEAD;
    const uint32_t sb = s_u32(smem);
    const uint4 z4 = make_uint4(0, 0, 0, 0);
    float* S = reinterpret_cast<float*>(smem + L::S_OFF);
    float* P2 = reinterpret_cast<float*>(smem + L::P2_OFF);
    float* inv = reinterpret_cast<float*>(smem + L::INV_OFF);

    // Q + u, Q + v (fp32 add, one rounding to fp16); rows >= T are zero
    for (int idx = tid; idx < 16 * 16; idx += 128) {
        const int r = idx >> 4, c = idx & 15;
        uint4 ou = z4, ov = z4;
        if (i0 + r < T) {
            const uint4 q8 = *reinterpret_cast<const uint4*>(qkv + (size_t)(i0 + r) * (3 * D_MODEL) + c * 8);
            const __half2* qh = reinterpret_cast<const __half2*>(&q8);
            const float4 u0 = *reinterpret_cast<const float4*>(a.bias_u + h * D_HEAD + c * 8), u1 = *reinterpret_cast<const float4*>(a.bias_u + h * D_HEAD + c * 8 + 4);
            const float4 v0 = *reinterpret_cast<const float4*>(a.bias_v + h * D_HEAD + c * 8), v1 = *reinterpret_cast<const float4*>(a.bias_v + h * D_HEAD + c * 8 + 4);
            const float uu[8] = {u0.x, u0.y, u0.z, u0.w, u1.x, u1.y, u1.z, u1.w};
            const float vv[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
            __half2 hu[4], hv[4];
#pragma unroll
            for (int e = 0; e < 4; e++) {
                const float2 q2 = __half22float2(qh[e]);
                hu[e] = __floats2half2_rn(q2.x + uu[2 * e], q2.y + uu[2 * e + 1]);
                hv[e] = __floats2half2_rn(q2.x + vv[2 * e], q2.y + vv[2 * e + 1]);
            }
            ou = *reinterpret_cast<uint4*>(hu);
            ov = *reinterpret_cast<uint4*>(hv);
        }
        *reinterpret_cast<uint4*>(smem + L::QU_OFF + r * ROWB + c * 16) = ou;
        *reinterpret_cast<uint4*>(smem + L::QV_OFF + r * ROWB + c * 16) = ov;
    }
    const float scale = 1.0f / sqrtf((float)D_HEAD);
    const int a_row = (lane & 7) + ((lane >> 3) & 1) * 8, a_col = (lane >> 4) * 16;
    const int b_row = lane & 7, b_col = (lane >> 3) * 16;
    const int n_tiles = (T + L::KT - 1) / L::KT;

    // ---- scores ----
    for (int kt = 0; kt < n_tiles; kt++) {
        const int j0 = kt * L::KT;
        __syncthreads();  // the previous tile's K / position rows and P2 have been consumed
        for (int idx = tid; idx < L::KT * 16; idx += 128) {
            const int r = idx >> 4, c = idx & 15, j = j0 + r;
            if (j < T) cp_async16(sb + L::KV_OFF + r * ROWB + c * 16, qkv + (size_t)j * (3 * D_MODEL) + D_MODEL + c * 8);
            else *reinterpret_cast<uint4*>(smem + L::KV_OFF + r * ROWB + c * 16) = z4;
        }
        // pair (query i0 + i, key j0 + jj) uses position (i0 + i) - (j0 + jj): table row rlo + (i - jj + 63)
        const int rlo = i0 - j0 - (L::KT - 1) + T - 1;
        for (int idx = tid; idx < L::PROWS * 16; idx += 128) {
            const int r = idx >> 4, c = idx & 15, row = rlo + r;
            if (row >= 0 && row < 2 * T - 1) cp_async16(sb + L::POS_OFF + r * ROWB + c * 16, pbase + (size_t)row * D_MODEL + c * 8);
            else *reinterpret_cast<uint4*>(smem + L::POS_OFF + r * ROWB + c * 16) = z4;
        }
        cp_commit();
        cp_wait<0>();
        __syncthreads();
        // content scores: 8 key tiles of 8, two per warp -> S[i][j0 + jj] (raw)
        for (int tile = warp; tile < L::KT / 8; tile += 4) {
            const int n0 = 8 * tile;
            const uint32_t qa = sb + L::QU_OFF + a_row * ROWB + a_col;
            const uint32_t kb = sb + L::KV_OFF + (n0 + b_row) * ROWB + b_col;
            float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int ks = 0; ks < 4; ks++) {
                uint32_t bf[4], a0[4], a1[4];
                ldsm_x4(bf, kb + ks * 64);
                ldsm_x4(a0, qa + ks * 64);
                ldsm_x4(a1, qa + ks * 64 + 32);
                mma_f16(acc, a0, bf[0], bf[1]);
                mma_f16(acc, a1, bf[2], bf[3]);
            }
            *reinterpret_cast<float2*>(S + g * L::S_LD + j0 + n0 + 2 * t) = make_float2(acc[0], acc[1]);
            *reinterpret_cast<float2*>(S + (g + 8) * L::S_LD + j0 + n0 + 2 * t) = make_float2(acc[2], acc[3]);
        }
        // position scores of the tile's 80 rows -> P2[i][r]
        for (int tile = warp; tile < L::PROWS / 8; tile += 4) {
            const int n0 = 8 * tile;
            const uint32_t qa = sb + L::QV_OFF + a_row * ROWB + a_col;
            const uint32_t kb = sb + L::POS_OFF + (n0 + b_row) * ROWB + b_col;
            float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int ks = 0; ks < 4; ks++) {
                uint32_t bf[4], a0[4], a1[4];
                ldsm_x4(bf, kb + ks * 64);
                ldsm_x4(a0, qa + ks * 64);
                ldsm_x4(a1, qa + ks * 64 + 32);
                mma_f16(acc, a0, bf[0], bf[1]);
                mma_f16(acc, a1, bf[2], bf[3]);
            }
            *reinterpret_cast<float2*>(P2 + g * L::P2_LD + n0 + 2 * t) = make_float2(acc[0], acc[1]);
            *reinterpret_cast<float2*>(P2 + (g + 8) * L::P2_LD + n0 + 2 * t) = make_float2(acc[2], acc[3]);
        }
        __syncthreads();
        // rel-shift by indexing, scale
        for (int idx = tid; idx < 16 * L::KT; idx += 128) {
            const int i = idx >> 6, jj = idx & 63;
            S[i * L::S_LD + j0 + jj] = (S[i * L::S_LD + j0 + jj] + P2[i * L::P2_LD + i - jj + (L::KT - 1)]) * scale;
        }
    }
    __syncthreads();
    // ---- softmax over the T keys (fp32), one warp per 4 query rows -> fp16 probabilities (unnormalised), 1 / sum kept per row ----
    const int kpad = n_tiles * L::KT;
    for (int r = warp * 4; r < warp * 4 + 4; r++) {
        __half* prow = reinterpret_cast<__half*>(smem + L::PR_OFF + r * L::PR_PITCH);
        float mx = -INFINITY;
        for (int j = lane; j < T; j += 32) mx = fmaxf(mx, S[r * L::S_LD + j]);
        mx = warp_max(mx);
        float sum = 0.0f;
        for (int j = lane; j < kpad; j += 32) {
            float e = 0.0f;
            if (j < T) {
                e = expf(S[r * L::S_LD + j] - mx);
                sum += e;
            }
            prow[j] = __float2half_rn(e);
        }
        sum = warp_sum(sum);
        if (lane == 0) inv[r] = 1.0f / sum;
    }
    // ---- context: warp w owns features [32 w, 32 w + 32) ----
    float acc[2][2][4];
#pragma unroll
    for (int x = 0; x < 2; x++)
#pragma unroll
        for (int y = 0; y < 2; y++)
#pragma unroll
            for (int e = 0; e < 4; e++) acc[x][y][e] = 0.0f;
    for (int kt = 0; kt < n_tiles; kt++) {
        const int j0 = kt * L::KT;
        __syncthreads();  // first pass: probabilities complete, K tile consumed; later: previous V tile consumed
        for (int idx = tid; idx < L::KT * 16; idx += 128) {
            const int r = idx >> 4, c = idx & 15, j = j0 + r;
            if (j < T) cp_async16(sb + L::KV_OFF + r * ROWB + c * 16, qkv + (size_t)j * (3 * D_MODEL) + 2 * D_MODEL + c * 8);
            else *reinterpret_cast<uint4*>(smem + L::KV_OFF + r * ROWB + c * 16) = z4;
        }
        cp_commit();
        cp_wait<0>();
        __syncthreads();
        const uint32_t pa = sb + L::PR_OFF + a_row * L::PR_PITCH + a_col + j0 * 2;
#pragma unroll
        for (int ks = 0; ks < L::KT / 16; ks++) {
            uint32_t af[4];
            ldsm_x4(af, pa + ks * 32);
#pragma unroll
            for (int x = 0; x < 2; x++) {
                const uint32_t vb = sb + L::KV_OFF + ((lane & 7) + ((lane >> 3) & 1) * 8 + ks * 16) * ROWB + (32 * warp + 16 * x + (lane >> 4) * 8) * 2;
                uint32_t bf[4];
                ldsm_x4_t(bf, vb);
                mma_f16(acc[x][0], af, bf[0], bf[1]);
                mma_f16(acc[x][1], af, bf[2], bf[3]);
            }
        }
    }
    __syncthreads();  // every warp is done with the score rows' neighbours (probabilities alias the position area, not S): S is free
    float* Cx = reinterpret_cast<float*>(smem + L::CX_OFF);
#pragma unroll
    for (int x = 0; x < 2; x++) {
        const int f0 = 32 * warp + 16 * x;
        *reinterpret_cast<float2*>(Cx + g * CX_LD + f0 + 2 * t) = make_float2(acc[x][0][0], acc[x][0][1]);
        *reinterpret_cast<float2*>(Cx + (g + 8) * CX_LD + f0 + 2 * t) = make_float2(acc[x][0][2], acc[x][0][3]);
        *reinterpret_cast<float2*>(Cx + g * CX_LD + f0 + 8 + 2 * t) = make_float2(acc[x][1][0], acc[x][1][1]);
        *reinterpret_cast<float2*>(Cx + (g + 8) * CX_LD + f0 + 8 + 2 * t) = make_float2(acc[x][1][2], acc[x][1][3]);
    }
    __syncthreads();
    for (int idx = tid; idx < 16 * (D_HEAD / 4); idx += 128) {  // a warp = one row: the 8-lane groups of group8_max are 32 features
        const int i = idx >> 5, d = (idx & 31) * 4;
        const float s1 = inv[i];
        float4 v = *reinterpret_cast<const float4*>(Cx + i * CX_LD + d);
        v = make_float4(v.x * s1, v.y * s1, v.z * s1, v.w * s1);
        if (a.act_quant) {
            const float amax = group8_max(fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w))));
            v.x = fake_quant(v.x, amax);
            v.y = fake_quant(v.y, amax);
            v.z = fake_quant(v.z, amax);
            v.w = fake_quant(v.w, amax);
        }
        if (i0 + i < T) {
            uint2 u;
            u.x = pack_operand2(v.x, v.y, a.out_bf16);
            u.y = pack_operand2(v.z, v.w, a.out_bf16);
            *reinterpret_cast<uint2*>(a.out + (size_t)(i0 + i) * D_MODEL + h * D_HEAD + d) = u;
        }
    }
}

void launch_attention_full(cudaStream_t st, const AttnFullArgs& a) {
    if (a.T <= 0) return;
    if (a.T > ATTN_FULL_MAX_T) throw Error("offline attention: more than 512 encoder frames");
    static const bool simt = getenv("NB200_ATTN_FULL_SIMT") != nullptr;
    ProfScope ps(st, "attention_full", 6.0 * N_HEADS * (double)a.T * a.T * D_HEAD, (double)a.T * D_MODEL * 2 * 6);
    if (simt) {
        launch_k(attention_full_kernel, dim3(dim3(cdiv(a.T, AF_QB), N_HEADS)), dim3(256), 0, st, a);
    } else {
        static PerDeviceOnce once;
        if (once.first()) CUDA_CHECK(cudaFuncSetAttribute(attention_full_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AfSmem::TOTAL));
        launch_k(attention_full_mma_kernel, dim3(dim3(cdiv(a.T, 16), N_HEADS)), dim3(128), (size_t)AfSmem::TOTAL, st, a);
    }
    NB_KERNEL_CHECK();
}

void launch_attention(cudaStream_t st, const AttnArgs& a) {
    static const bool use_simt = getenv("NB200_ATTN_SIMT") != nullptr;
    if (a.T > MAX_T) throw Error("attention: more than 14 frames per chunk");
    const double kv = LEFT + a.T;
    const double flops = 6.0 * a.B * a.T * kv * D_MODEL;
    // algorithmic bytes: cached K+V rows read, this chunk's q/k/v read, K/V rows appended, context written, positions once
    const double bytes = (double)a.B * (2.0 * LEFT * D_MODEL * 2 + (double)a.T * D_MODEL * (6 + 4 + 2)) + (LEFT + 2.0 * a.T - 1) * D_MODEL * 2;
    if (use_simt) {
        ProfScope ps(st, "attention_simt", flops, bytes);
        launch_k(attention_simt_kernel, dim3(dim3(N_HEADS, a.B)), dim3(256), 0, st, a);
        NB_KERNEL_CHECK();
        return;
    }
    static const bool no_stream_kernel = getenv("NB200_ATTN_MMA_ONLY") != nullptr;
    if (a.T <= 2 && !no_stream_kernel) {
        ProfScope ps(st, "attention_stream", flops, bytes);
        const dim3 grid(cdiv(a.B * N_HEADS, 4));
        for (int rep = dup_family("attention") ? 2 : 1; rep > 0; rep--) {
            const bool small = a.B <= 320;  // <= ~17 warps per SM: latency-bound, keep more rows in flight (see the kernel's comment)
            if (a.T == 1) {
                if (small) launch_k(attention_stream_kernel<1, 12>, grid, dim3(128), 0, st, a);
                else launch_k(attention_stream_kernel<1, 6>, grid, dim3(128), 0, st, a);
            } else {
                if (small) launch_k(attention_stream_kernel<2, 12>, grid, dim3(128), 0, st, a);
                else launch_k(attention_stream_kernel<2, 6>, grid, dim3(128), 0, st, a);
            }
        }
        NB_KERNEL_CHECK();
        return;
    }
    static const bool two_wave = getenv("NB200_ATTN_STAGED") != nullptr;  // the 103 KB version that stages K, V and P at once
    static PerDeviceOnce once;
    if (once.first()) {
        CUDA_CHECK(cudaFuncSetAttribute(attention_f16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AttSmem::TOTAL));
        CUDA_CHECK(cudaFuncSetAttribute(attention_f16s_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AttSmemS::TOTAL));
    }
    ProfScope ps(st, "attention_f16", flops, bytes);
    for (int rep = dup_family("attention") ? 2 : 1; rep > 0; rep--) {
        if (two_wave)
            launch_k(attention_f16_kernel, dim3(dim3(N_HEADS, a.B)), dim3(AT_THREADS), (size_t)AttSmem::TOTAL, st, a);
        else
            launch_k(attention_f16s_kernel, dim3(dim3(N_HEADS, a.B)), dim3(AT_THREADS), (size_t)AttSmemS::TOTAL, st, a);
    }
    NB_KERNEL_CHECK();
}

}  // namespace nb

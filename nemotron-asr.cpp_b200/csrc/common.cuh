// common.cuh — shared definitions for libnemob200 (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <stdexcept>
#include <string>
#include <utility>

namespace nb {

// Architecture constants of nemotron-speech-streaming-0.6B (src/nemo-ggml.h:37-55 in the reference); only
// n_layers is read from the file, everything else is checked against these at load.
constexpr int N_MELS = 128, D_MODEL = 1024, N_HEADS = 8, D_HEAD = 128, D_FF = 4096, KERNEL = 9;
constexpr int VOCAB = 1025, BLANK = 1024, DEC = 640, JOINT = 640, LEFT = 70, CONV_CACHE = 8;
constexpr int SUB_CH = 256, SUB_F1 = 65, SUB_F2 = 33, SUB_F3 = 17, SUB_FEAT = SUB_CH * SUB_F3;  // 4352
constexpr int N_FFT = 512, HOP = 160, N_BINS = 257, MEL_OVERLAP = 9;
constexpr int RING = 84;          // K/V ring capacity: 70 cached + up to 14 new frames
constexpr int PROMPT_HIDDEN = 2048;  // prompt_kernel hidden width (multilingual checkpoints)
constexpr int MAX_T = 14;         // encoder frames per chunk at right_context 13
constexpr int MAX_SYMBOLS = 10;   // src/nemo-stream.cpp:849

struct Error : std::runtime_error {
    using std::runtime_error::runtime_error;
};

inline void cuda_check(cudaError_t e, const char* what, const char* file, int line) {
    if (e != cudaSuccess) {
        char buf[512];
        snprintf(buf, sizeof buf, "CUDA error %s at %s:%d: %s", what, file, line, cudaGetErrorString(e));
        throw Error(buf);
    }
}
#define CUDA_CHECK(x) ::nb::cuda_check((x), #x, __FILE__, __LINE__)

// Kernels launched by this library on the calling thread (one thread owns a model, include/nb200.h: the engine adds the
// per-entry-point difference to the model's own total, which is what nb200_get_stats reports as kernel_launches).
extern thread_local long long g_launches;
#define NB_LAUNCHED() (++::nb::g_launches)
#define NB_KERNEL_CHECK() do { NB_LAUNCHED(); CUDA_CHECK(cudaGetLastError()); } while (0)

// Per-kernel-family device timing (nb200_profile_*): when enabled, every launcher brackets its launch with CUDA events
// on the launching stream and reports the algorithmic FLOPs / bytes of that launch; prof.cu aggregates per family.
extern thread_local bool g_prof_on;  // per thread, like the records it fills
void prof_begin(cudaStream_t st, const char* family, double flops, double bytes);
void prof_end(cudaStream_t st);
std::string prof_report();  // JSON of per-family totals since the last report; clears the records
struct ProfScope {
    cudaStream_t st;
    bool on;
    ProfScope(cudaStream_t s, const char* family, double flops, double bytes) : st(s), on(g_prof_on) {
        if (on) prof_begin(st, family, flops, bytes);
    }
    ~ProfScope() {
        if (on) prof_end(st);
    }
};

// ---- programmatic dependent launch (PDL) -----------------------------------------------------------------------
// Every kernel of the library starts with pdl_prologue(): "launch_dependents" lets the NEXT kernel in the stream be
// scheduled as soon as all CTAs of this one have started (its CTAs then sit in their own prologue / at their wait),
// and "wait" blocks until the PREVIOUS kernel has completed and its writes are visible. All global-memory reads of
// produced data and all global writes come after the wait, so the only thing that overlaps is launch latency and
// per-CTA setup (barrier init, TMEM allocation, descriptor prefetch). Kernels are launched through launch_k(), which
// sets the programmatic-stream-serialization attribute (also inside CUDA-graph capture). NB200_NO_PDL=1 disables it.
extern bool g_pdl;
// Timing experiment: NB200_DUP=<family> makes the launcher of that kernel family launch it twice, so its marginal in-graph cost is
// (step time with - without) / launches. Results are garbage in that mode.
bool dup_family(const char* family);
void set_dup_family(const char* family);  // nullptr / "" = off (nb200_set_dup_family; initialised from NB200_DUP)
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_prologue() {
    pdl_launch();
    pdl_wait();
}
// cluster_x > 1 launches thread-block clusters of (cluster_x, 1, 1); grid.x must be a multiple of it.
template <typename... KArgs, typename... Args>
void launch_kc(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, int cluster_x, Args&&... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = g_pdl ? 1 : 0;
    cfg.numAttrs = 1;
    if (cluster_x > 1) {
        at[1].id = cudaLaunchAttributeClusterDimension;
        at[1].val.clusterDim.x = (unsigned)cluster_x;
        at[1].val.clusterDim.y = 1;
        at[1].val.clusterDim.z = 1;
        cfg.numAttrs = 2;
    }
    cfg.attrs = at;
    CUDA_CHECK(cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...));
}
// Cooperative launch (no programmatic overlap): for kernels whose CTAs wait for each other at hand-rolled grid barriers. The
// driver then places the whole grid at once or fails the launch (cudaErrorCooperativeLaunchTooLarge) instead of letting a
// partially resident grid spin forever when another kernel holds some of the SMs. Capturable into CUDA graphs.
template <typename... KArgs, typename... Args>
void launch_coop(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeCooperative;
    at[0].val.cooperative = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    CUDA_CHECK(cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...));
}
template <typename... KArgs, typename... Args>
void launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
    launch_kc(kernel, grid, block, smem, st, 1, std::forward<Args>(args)...);
}
#endif

// cudaFuncSetAttribute is per device: "done once" flags must be kept per device, or a second model on another GPU of
// the same process (server_loop.h ShardedWorkers) would launch with the default 48 KB dynamic shared-memory limit.
struct PerDeviceOnce {
    bool done[64] = {};
    bool first() {  // true the first time it is called for the current device
        int dev = 0;
        cudaGetDevice(&dev);
        dev &= 63;
        if (done[dev]) return false;
        done[dev] = true;
        return true;
    }
};

template <typename T>
T* dalloc(size_t n) {
    T* p = nullptr;
    CUDA_CHECK(cudaMalloc(&p, n * sizeof(T)));
    return p;
}

__host__ __device__ inline int cdiv(int a, int b) { return (a + b - 1) / b; }

// Latency-mode arithmetic (src/nemo-stream.h:65-100)
struct Mode {
    int rc;         // att_right_context
    int T;          // encoder frames per chunk = 1 + rc
    int shift;      // new mel frames per chunk = 8 * (1 + rc)
    int chunk_mel;  // 9 + shift
    int t1, t2, t3; // time extents after each stride-2 stage
    int npos;       // relative positions used: p in [-(T-1), 70+T-1]
    explicit Mode(int rc_) : rc(rc_) {
        T = 1 + rc;
        shift = 8 * (1 + rc);
        chunk_mel = MEL_OVERLAP + shift;
        t1 = chunk_mel / 2 + 1;
        t2 = t1 / 2 + 1;
        t3 = t2 / 2 + 1;
        npos = LEFT + 2 * T - 1;
    }
};

// ---- device helpers -------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
// Two fp32 values -> one 32-bit word of the GEMM operand type: fp16 (default) or bf16 (models loaded with NB200_FLAG_BF16: the
// encoder GEMMs then run kind::f16 with A = B = bf16, so every kernel that produces an A operand stores bf16).
__device__ __forceinline__ uint32_t pack_operand2(float a, float b, int bf16) {
    if (bf16) {
        const __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
        return *reinterpret_cast<const uint32_t*>(&v);
    }
    const __half2 v = __floats2half2_rn(a, b);
    return *reinterpret_cast<const uint32_t*>(&v);
}
__device__ __forceinline__ float silu_f(float x) { return x / (1.0f + expf(-x)); }
__device__ __forceinline__ float sigmoid_f(float x) { return 1.0f / (1.0f + expf(-x)); }
// GEMM-epilogue versions. The epilogue of a 128 x 256 tile evaluates 32768 sigmoids; it is bound by instruction issue
// (ncu: ~20 instructions per element with the IEEE-safe __expf / division sequences, IPC ~0.5), so these are the
// shortest forms: ex2.approx.ftz and rcp.approx.ftz (1 ulp each) -> FMUL, MUFU.EX2, FADD, MUFU.RCP, FMUL.
// x -> -inf gives e = +inf, 1/inf = 0, x*0 = -0: no NaN for finite x.
__device__ __forceinline__ float ex2_ftz(float t) {
    float e;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(t));
    return e;
}
__device__ __forceinline__ float rcp_ftz(float y) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(y));
    return r;
}
__device__ __forceinline__ float sigmoid_fast(float x) { return rcp_ftz(1.0f + ex2_ftz(x * -1.4426950408889634f)); }
__device__ __forceinline__ float silu_fast(float x) { return x * sigmoid_fast(x); }
// SiLU of four values with ONE reciprocal: 1/(1+a), 1/(1+b), ... = r * (product of the other three), r = 1 / product of all.
// MUFU work drops from 2 to 1.25 per element (4 ex2 + 1 rcp); the extra multiplies go to the FMA pipe. The exponent is
// clamped at 2^30 so the product of four stays finite: below x = -20.8 sigmoid is then 2^-30 instead of e^x, i.e. |silu|
// < 2e-8 either way, which rounds to zero in the fp16 output. Each result carries <= 3 extra fp32 roundings.
__device__ __forceinline__ void silu_fast4(float& x0, float& x1, float& x2, float& x3) {
    const float d0 = 1.0f + ex2_ftz(fminf(x0 * -1.4426950408889634f, 30.0f));
    const float d1 = 1.0f + ex2_ftz(fminf(x1 * -1.4426950408889634f, 30.0f));
    const float d2 = 1.0f + ex2_ftz(fminf(x2 * -1.4426950408889634f, 30.0f));
    const float d3 = 1.0f + ex2_ftz(fminf(x3 * -1.4426950408889634f, 30.0f));
    const float p01 = d0 * d1, p23 = d2 * d3;
    const float r = rcp_ftz(p01 * p23);
    const float r01 = r * p23, r23 = r * p01;  // 1/(d0 d1), 1/(d2 d3)
    x0 *= r01 * d1;
    x1 *= r01 * d0;
    x2 *= r23 * d3;
    x3 *= r23 * d2;
}
// a_i * sigmoid(b_i) for four pairs with one reciprocal (GLU epilogue), same scheme as silu_fast4
__device__ __forceinline__ void glu_fast4(float (&o)[4], float a0, float b0, float a1, float b1, float a2, float b2, float a3, float b3) {
    const float d0 = 1.0f + ex2_ftz(fminf(b0 * -1.4426950408889634f, 30.0f));
    const float d1 = 1.0f + ex2_ftz(fminf(b1 * -1.4426950408889634f, 30.0f));
    const float d2 = 1.0f + ex2_ftz(fminf(b2 * -1.4426950408889634f, 30.0f));
    const float d3 = 1.0f + ex2_ftz(fminf(b3 * -1.4426950408889634f, 30.0f));
    const float p01 = d0 * d1, p23 = d2 * d3;
    const float r = rcp_ftz(p01 * p23);
    const float r01 = r * p23, r23 = r * p01;
    o[0] = a0 * (r01 * d1);
    o[1] = a1 * (r01 * d0);
    o[2] = a2 * (r23 * d3);
    o[3] = a3 * (r23 * d2);
}
// Scale pair of ggml-CPU's Q8_0 quantiser for a 32-block with the given amax: d = fp16(amax/127), id = 127/amax
__device__ __forceinline__ void fq_scales(float amax, float& d, float& id) {
    d = amax == 0.0f ? 0.0f : __half2float(__float2half_rn(__fdiv_rn(amax, 127.0f)));
    id = amax == 0.0f ? 0.0f : __fdiv_rn(127.0f, amax);
}

// ggml-CPU's Q8_0 activation quantiser applied to a value given its block's amax, returned dequantised:
// d = fp16(amax/127), q = rne(x * (127/amax))  (ggml/src/ggml-cpu/arch/x86/quants.c:331-346), value = q*d.
__device__ __forceinline__ float fake_quant(float x, float amax) {
    if (amax == 0.0f) return 0.0f;
    float d = __half2float(__float2half_rn(__fdiv_rn(amax, 127.0f)));
    float id = __fdiv_rn(127.0f, amax);
    return rintf(__fmul_rn(x, id)) * d;
}

}  // namespace nb

// engine.cu — model load, stream sessions, the batched step scheduler and the C ABI (include/nb200.h).
//
// Replaces, for the streaming hot path, the reference's nemo_model_load (src/nemo-ggml.cpp:83-442),
// nemo_stream_context (src/nemo-stream.h:177-262), process_mel_chunk_streaming / nemo_stream_process_incremental /
// nemo_stream_finalize (src/nemo-stream.cpp:1013-1293). One step runs ONE chunk of EVERY ready stream as batched
// kernels; per-stream state (K/V ring caches, conv cache, mel overlap, LSTM state) lives in HBM slots.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/nb200.h"
#include "gemm.cuh"
#include "gguf_reader.h"
#include "kernels.cuh"
#include "weight_cache.h"

using namespace nb;

namespace {

thread_local std::string g_error;
void set_error(const std::string& e) {
    g_error = e;
    fprintf(stderr, "nb200: %s\n", e.c_str());
}

struct F16Matrix {  // fp16 [rows][K] in HBM + TMA descriptors
    __half* ptr = nullptr;
    Operand op;
};

struct LayerW {
    float *ln_ff1_g, *ln_ff1_b, *ln_att_g, *ln_att_b, *ln_conv_g, *ln_conv_b, *ln_ff2_g, *ln_ff2_b, *ln_out_g, *ln_out_b;
    float *bn_g, *bn_b, *pos_u, *pos_v, *w_dw;
    F16Matrix ff1_w1, ff1_w2, wqkv, wpos, wout, pw1, pw2, ff2_w1, ff2_w2;
};

template <typename T>
struct PinnedBuf {
    T* p = nullptr;
    size_t n = 0;
    void reserve(size_t m) {
        if (m <= n) return;
        if (p) cudaFreeHost(p);
        CUDA_CHECK(cudaMallocHost(&p, m * sizeof(T)));
        n = m;
    }
    ~PinnedBuf() {
        if (p) cudaFreeHost(p);
    }
};

// Per latency mode: the precomputed positional table and the host side of a batched step of that mode (pinned metadata and
// result buffers, events), so that the groups of a step with streams in several modes are all launched before any is waited
// for (round 1 completed every group but the last synchronously inside nb200_step_begin).
struct ModeCtx {
    Mode mode;
    __half* pos_tab = nullptr;  // [L][npos][1024] fp16 linear_pos(PE(p)), p = -(T-1) .. 70+T-1
    PinnedBuf<int> h_meta;
    PinnedBuf<long long> h_meta64;
    PinnedBuf<int> h_tok;
    cudaEvent_t ev[4]{};
    explicit ModeCtx(int rc) : mode(rc) {}
    ~ModeCtx() {
        for (auto& e : ev)
            if (e) cudaEventDestroy(e);
    }
};

}  // namespace

namespace {
// A captured kernel sequence (CUDA graph) plus the number of kernels it launches per replay.
struct GraphExec {
    cudaGraphExec_t exec = nullptr;
    long long launches = 0;
    int seen = 0;  // times this shape was requested; the first request runs eagerly (lazy cudaFuncSetAttribute calls)
};

// One contiguous run of fed samples: arena offset -> position in a slot's device PCM ring
struct UploadSeg {
    long long dst;  // sample index into pcm_ring
    int src;        // sample offset in the staging arena
    int n;
};
constexpr int MAX_UPLOAD_SEGS = 8192;
struct UploadArena {
    PinnedBuf<int16_t> pcm;
    PinnedBuf<UploadSeg> segs;
    size_t off = 0;
    int nseg = 0, max_n = 0;
    cudaEvent_t done = nullptr;
    bool busy = false;
};
struct BatchItem {
    nb200_stream* s;
    int n_frames;  // new mel frames this chunk really has (shift for a full chunk, fewer for a finalize tail)
    int nvalid;    // encoder frames to decode
};

}  // namespace

struct nb200_stream {
    nb200_model* model = nullptr;
    int slot = -1;
    int rc = 0;
    // audio: samples [0, fed) have been uploaded into this slot's device PCM ring; `pending` holds samples that did
    // not fit yet (the ring keeps what unprocessed chunks still need) and is flushed as chunks are consumed.
    std::vector<int16_t> pending;
    long long fed = 0;
    long long samples_total = 0;  // fed + pending.size()
    long long chunks_done = 0;
    bool finalize_requested = false, finalized = false;
    int wpos = 0;   // K/V ring write index
    int valid = 0;  // cache_valid_len
    std::vector<int32_t> tokens;
    std::vector<int64_t> token_frames;  // encoder frame each token was emitted at (timed_token::frame_idx, src/nemo-ggml.h:383-395)
    long long frames_done = 0;          // encoder frames decoded so far
    size_t poll_cursor = 0;
    std::vector<int32_t> chunk_tokens;
    std::vector<float> chunk_margins;  // per chunk: smallest top1-top2 logit gap of its symbol steps
    int prompt_index = -1;  // multilingual checkpoints: language prompt id (src/nemo-stream.h prompt_index); -1 = model has none
    int taps = 0;
    std::vector<float> tap_mel, tap_enc;
};

struct nb200_model {
    int device = 0;
    cudaStream_t st = nullptr;
    nb200_info info{};
    std::vector<std::string> vocab;
    bool act_quant = false, simt = false;
    int n_layers = 0, max_streams = 0;

    std::vector<void*> allocs;  // everything cudaMalloc'ed
    // weights
    MelTables mel_tb{};
    SubWeights sub{};
    float *sub_b3 = nullptr, *sub_b6 = nullptr, *sub_bout = nullptr;
    F16Matrix sub_w3, sub_w6, sub_wout;
    float *joint_enc_w = nullptr, *joint_enc_b = nullptr;  // fp32: the decoder side stays in fp32 like the reference
    F16Matrix pk_w1_3, pk_w2_3;  // prompt kernel fc1 (encoder part) / fc2 as fp16 triples
    __half* ws_pk16 = nullptr;   // (hi, lo, hi) of the prompt kernel's hidden layer [rows][3 * 2048]
    Operand op_pk3;
    F16Matrix joint_enc_w3;    // joint.enc as [Wh | Wh | Wl] fp16 [640][3072] (split-precision tensor-core GEMM, see split_f16x3_kernel)
    Operand op_x3;             // [rows][3072] view of ws_h16: (hi, lo, hi) of the encoder output
    bool joint_enc_tc = true;  // NB200_JOINT_SGEMM=1: fp32 CUDA-core SGEMM instead
    // multilingual prompt fusion (src/nemo-ggml.cpp:1087-1105), fp32 like the file: fc1 split into its encoder part
    // [2048][1024] and its one-hot part stored per language [num_prompts][2048]; fc2 [1024][2048]
    int num_prompts = 0, default_prompt = -1;
    // offline path (nb200_encode_audio): positional table of the last utterance length, allocated on first use
    float* off_posf = nullptr;
    __half *off_pos16 = nullptr, *off_pos_tab = nullptr;
    Operand op_off_pos16;
    int off_max_T = 0, off_pos_T = 0;
    int dec_window = 1;
    float *pk_w1 = nullptr, *pk_w1_lang = nullptr, *pk_b1 = nullptr, *pk_w2 = nullptr, *pk_b2 = nullptr;
    std::map<std::string, int> prompt_dict;
    std::vector<LayerW> layers;
    DecodeWeights dec{};
    struct NamedMatrix {   // GGUF tensor name -> where its rows live in our (possibly fused / interleaved) fp16 storage
        const __half* ptr;
        int rows, K;
        bool glu_interleaved;  // pointwise_conv1: row 2j = value j, row 2j+1 = gate j
    };
    std::map<std::string, NamedMatrix> by_name;
    // per-slot state
    __half *k_ring = nullptr, *v_ring = nullptr;  // [L][slots][RING][1024] fp16
    bool bf16 = false;           // NB200_FLAG_BF16: the layers' GEMM operands are bf16
    int16_t* off_pcm = nullptr;  // offline path: the utterance's PCM (persistent, sized for 512 encoder frames)
    float* conv_cache = nullptr;                  // [L][slots][2][8][1024] (two buffers, flipped every chunk)
    float* mel_overlap = nullptr;                                      // [slots][9][128]
    DecodeState dstate{};
    std::vector<int> free_slots;
    std::vector<nb200_stream*> streams;
    std::map<int, std::unique_ptr<ModeCtx>> modes;
    bool use_graphs = true;                        // NB200_NO_GRAPHS=1 or NB200_FLAG_NO_GRAPHS disables CUDA graphs
    std::map<long long, GraphExec> enc_graphs, dec_graphs;
    // workspace (capacity: cap_B streams x MAX_T frames)
    int cap_B = 0;
    float *ws_mel = nullptr, *ws_a3 = nullptr, *ws_x = nullptr, *ws_g = nullptr, *ws_e = nullptr;
    __half *ws_a2h = nullptr, *ws_a5h = nullptr, *ws_a6h = nullptr, *ws_a16 = nullptr, *ws_h16 = nullptr, *ws_pos16 = nullptr, *ws_qkv16 = nullptr;
    float* ws_posf = nullptr;
    float* ws_pk = nullptr;        // prompt fusion hidden layer [rows][2048] fp32 (multilingual checkpoints only)
    float* ws_part = nullptr;      // split-K partial outputs of the residual GEMMs: [splits][part_rows][1024]
    size_t part_rows = 0;
    int part_max_splits = 1;
    Operand op_a2h, op_a5h, op_a6h, op_a16, op_h16, op_pos16;
    int16_t* pcm_ring = nullptr;   // [slots][ring_cap] device-resident audio (nb200_stream_feed uploads into it)
    int ring_cap = 0;              // samples per slot, power of two
    int* ws_meta = nullptr;        // int arrays [8][cap_B]
    long long* ws_meta64 = nullptr;  // long long arrays [3][cap_B]: out_off, a0, n_avail
    int *ws_fp = nullptr, *ws_sym = nullptr, *ws_tok = nullptr, *ws_tokcnt = nullptr, *ws_unfinished = nullptr;
    float *ws_partial = nullptr, *ws_hid = nullptr;
    int4* ws_emit = nullptr;
    // Fed audio is gathered on the host (pinned arenas) and moved at the next step by ONE copy + one scatter kernel
    // instead of a cudaMemcpyAsync per stream; two arenas so feeding can go on while a flushed one is still in flight.
    UploadArena arena[2];
    int cur_arena = 0;
    cudaStream_t st_copy = nullptr;  // audio uploads (H2D + scatter) run here, beside the step's stream
    cudaEvent_t upload_ev = nullptr; // recorded after every flush; the next step's stream waits for it
    bool upload_wait = false;
    int16_t* d_stage = nullptr;    // device copy of the arena being flushed
    UploadSeg* d_segs = nullptr;
    // a step launched by nb200_step_begin and not yet completed by nb200_step_end
    struct Inflight {
        bool active = false;
        int total = 0;
        std::vector<std::pair<ModeCtx*, std::vector<BatchItem>>> groups;  // one batched chunk per latency mode, in launch order
    } inflight;
    // slots waiting for their state reset (apply_resets), the pinned / device copies of that list
    std::vector<int> reset_queue;
    PinnedBuf<int> h_reset;
    int* d_reset = nullptr;
    cudaEvent_t reset_ev = nullptr;
    bool reset_busy = false;
    long long launches = 0;  // kernels this model has launched (the per-thread counter g_launches, accumulated per entry point)
    PinnedBuf<int> h_meta;
    PinnedBuf<long long> h_meta64;
    PinnedBuf<int> h_tok;
    PinnedBuf<float> h_tap;
    cudaEvent_t ev[4]{};
    cudaEvent_t tev[2]{};
    nb200_stats stats{};

    template <typename T>
    T* alloc(size_t n, bool zero = true) {
        T* p = dalloc<T>(n);
        allocs.push_back(p);
        if (zero) CUDA_CHECK(cudaMemset(p, 0, n * sizeof(T)));
        return p;
    }
    template <typename T>
    T* upload(const std::vector<T>& v) {
        T* p = alloc<T>(v.size(), false);
        CUDA_CHECK(cudaMemcpy(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
        return p;
    }
    // The native weight file (weight_cache.h): a big operand is taken from the cache when it is there (mmap -> one copy to the
    // device; `produce`, the dequantise-and-repack on the host, is not run), otherwise produced, uploaded and appended to the cache
    // being written. n = element count the layout expects.
    std::unique_ptr<WeightCacheReader> wc_in;
    std::unique_ptr<WeightCacheWriter> wc_out;
    long long wc_hits = 0, wc_misses = 0;
    template <typename T, typename F>
    T* upload_cached(const std::string& key, size_t n, F&& produce) {
        if (wc_in) {
            if (const WeightCacheBlob* b = wc_in->find(key, n * sizeof(T))) {
                T* p = alloc<T>(n, false);
                CUDA_CHECK(cudaMemcpy(p, b->data, n * sizeof(T), cudaMemcpyHostToDevice));
                wc_hits++;
                return p;
            }
        }
        const std::vector<T> v = produce();
        if (v.size() != n) throw Error("weight layout: " + key + " has " + std::to_string(v.size()) + " elements, expected " + std::to_string(n));
        if (wc_out) wc_out->add(key, v.data(), n * sizeof(T));
        wc_misses++;
        return upload(v);
    }
    ~nb200_model() {
        for (auto* s : streams) delete s;
        for (void* p : allocs) cudaFree(p);
        for (auto& e : ev)
            if (e) cudaEventDestroy(e);
        for (auto& e : tev)
            if (e) cudaEventDestroy(e);
        for (auto& A : arena)
            if (A.done) cudaEventDestroy(A.done);
        for (auto* gm : {&enc_graphs, &dec_graphs})
            for (auto& kv : *gm)
                if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
        if (upload_ev) cudaEventDestroy(upload_ev);
        if (reset_ev) cudaEventDestroy(reset_ev);
        if (st_copy) cudaStreamDestroy(st_copy);
        if (st) cudaStreamDestroy(st);
    }
};

namespace {

std::vector<float> load_f32(const GgufFile& g, const std::string& name, int64_t expect = -1) {
    const GgufTensor& t = g.tensor(name);
    if (expect >= 0 && t.nelements() != expect)
        throw Error("tensor " + name + " has " + std::to_string(t.nelements()) + " elements, expected " + std::to_string(expect));
    std::vector<float> v((size_t)t.nelements());
    dequantize_to_f32(t, v.data());
    return v;
}

// fp32 [rows][K] (from `produce`) -> the fp16 GEMM operand on the device, through the weight cache under `key`
template <typename F>
void make_f16(nb200_model* m, F16Matrix& dst, const std::string& key, int rows, int K, F&& produce, const std::string& name = "", bool as_bf16 = false) {
    dst.ptr = m->upload_cached<__half>(key + (as_bf16 ? "#bf16" : ""), (size_t)rows * K, [&] {
        const std::vector<float> w = produce();
        if (w.size() != (size_t)rows * K) throw Error("weight layout: " + key + " has the wrong size");
        std::vector<__half> h(w.size());
        if (as_bf16) {  // bf16 bit patterns in the 16-bit container (NB200_FLAG_BF16)
            uint16_t* bits = reinterpret_cast<uint16_t*>(h.data());
            for (size_t i = 0; i < h.size(); i++) bits[i] = __bfloat16_as_ushort(__float2bfloat16_rn(w[i]));
        } else {
            for (size_t i = 0; i < h.size(); i++) h[i] = __float2half_rn(w[i]);
        }
        return h;
    });
    operand_init(dst.op, dst.ptr, rows, K);
    if (!name.empty()) m->by_name[name] = {dst.ptr, rows, K, false};
}

// Decode-kernel operand format (kernels.cuh DecodeWeights::wcat_s): per 4 consecutive floats 16 bytes of halves
// {hi0..hi3 | lo0..lo3}, hi = fp16(w), lo = fp16((w - hi) * 2048); returned as floats (bit containers) of the same geometry.
static std::vector<float> split_f16_pairs(const std::vector<float>& w) {
    std::vector<float> out(w.size());
    uint16_t* o = reinterpret_cast<uint16_t*>(out.data());
    for (size_t i = 0; i + 3 < w.size(); i += 4) {
        for (int j = 0; j < 4; j++) {
            const __half h = __float2half_rn(w[i + j]);
            const __half l = __float2half_rn((w[i + j] - __half2float(h)) * 2048.0f);
            o[2 * i + j] = __half_as_ushort(h);
            o[2 * i + 4 + j] = __half_as_ushort(l);
        }
    }
    return out;
}

// fp32 weights [rows][K] as the fp16 triple [Wh | Wh | Wl] along K (see split_f16x3_kernel): the B operand of a GEMM of fp32 quality
template <typename F>
void make_split3(nb200_model* m, F16Matrix& dst, const std::string& key, int rows, int K, F&& produce) {
    make_f16(m, dst, key + "#split3", rows, 3 * K, [&] {
        const std::vector<float> w = produce();
        if (w.size() != (size_t)rows * K) throw Error("weight layout: " + key + " has the wrong size");
        std::vector<float> w3((size_t)rows * 3 * K);
        for (int n = 0; n < rows; n++)
            for (int k = 0; k < K; k++) {
                const float v = w[(size_t)n * K + k];
                const float hi = __half2float(__float2half_rn(v));
                w3[(size_t)n * 3 * K + k] = hi;
                w3[(size_t)n * 3 * K + K + k] = hi;
                w3[(size_t)n * 3 * K + 2 * K + k] = v - hi;  // rounded to fp16 by make_f16
            }
        return w3;
    });
}

void load_matrix(nb200_model* m, const GgufFile& g, F16Matrix& dst, const std::string& name, int rows, int K, bool as_bf16 = false) {
    make_f16(m, dst, name, rows, K, [&] { return load_f32(g, name, (int64_t)rows * K); }, name, as_bf16);
}

uint32_t kv_u32(const GgufFile& g, const char* key, uint32_t dflt) {
    auto it = g.u32.find(key);
    return it == g.u32.end() ? dflt : it->second;
}

// Vocabulary (src/nemo-ggml.cpp:149-169): the string array `tokenizer.vocab_list`, else the legacy `tokenizer.vocab` blob of
// fixed 8-byte, NUL-padded pieces (vocab_size - 1 of them: the blank has no piece). Returns which of the two was read.
const char* load_vocab(GgufFile& g, int vocab_size, std::vector<std::string>& vocab) {
    vocab.clear();
    if (g.str_arr.count("tokenizer.vocab_list")) {
        vocab = g.str_arr["tokenizer.vocab_list"];
        return "tokenizer.vocab_list";
    }
    if (g.str.count("tokenizer.vocab")) {
        const std::string& blob = g.str["tokenizer.vocab"];
        for (size_t i = 0; i + 8 <= blob.size() && (int)(i / 8) < vocab_size - 1; i += 8)
            vocab.emplace_back(blob.data() + i, strnlen(blob.data() + i, 8));
        return "tokenizer.vocab";
    }
    throw Error("no vocabulary in GGUF (need tokenizer.vocab_list or tokenizer.vocab)");
}

void build_model(nb200_model* m, const std::string& path, int flags, const char* cache_path) {
    GgufFile g;
    g.open(path);
    if (cache_path) {  // the native weight file: use it if it was made from this GGUF, else (re)write it while loading
        uint64_t src_size = 0, src_mtime = 0;
        file_stamp(path, src_size, src_mtime);
        const std::string cp = *cache_path ? std::string(cache_path) : path + ".nb200w";
        m->wc_in.reset(new WeightCacheReader());
        if (!m->wc_in->open(cp, src_size, src_mtime)) {
            m->wc_in.reset();
            m->wc_out.reset(new WeightCacheWriter());
            try {
                m->wc_out->open(cp, src_size, src_mtime);
            } catch (const Error&) {  // read-only directory, disk full ...: the model still loads from the GGUF, just without writing the cache
                m->wc_out.reset();
            }
        }
    }
    nb200_info& hp = m->info;
    hp.n_mels = kv_u32(g, "nemo.n_mels", N_MELS);
    hp.d_model = kv_u32(g, "nemo.d_model", D_MODEL);
    hp.n_heads = kv_u32(g, "nemo.n_heads", N_HEADS);
    hp.d_head = kv_u32(g, "nemo.d_head", D_HEAD);
    hp.d_ff = kv_u32(g, "nemo.d_ff", D_FF);
    hp.n_layers = kv_u32(g, "nemo.n_layers", 24);
    hp.vocab_size = kv_u32(g, "nemo.vocab_size", VOCAB);
    hp.decoder_dim = kv_u32(g, "nemo.decoder_dim", DEC);
    hp.joint_dim = kv_u32(g, "nemo.joint_dim", JOINT);
    hp.subsampling_factor = kv_u32(g, "nemo.subsampling_factor", 8);
    hp.att_left_context = kv_u32(g, "nemo.att_left_context", LEFT);
    hp.num_prompts = kv_u32(g, "nemo.num_prompts", 0);
    hp.blank_token = hp.vocab_size - 1;
    hp.device = m->device;
    hp.max_streams = m->max_streams;
    if (hp.n_mels != N_MELS || hp.d_model != D_MODEL || hp.n_heads != N_HEADS || hp.d_head != D_HEAD || hp.d_ff != D_FF ||
        hp.vocab_size != VOCAB || hp.decoder_dim != DEC || hp.joint_dim != JOINT || hp.subsampling_factor != 8 ||
        hp.att_left_context != LEFT)
        throw Error("this backend is built for the nemotron-speech-streaming 0.6B architecture; the GGUF header differs");
    if (hp.num_prompts < 0 || hp.num_prompts > 4096) throw Error("nemo.num_prompts out of range");
    m->n_layers = hp.n_layers;
    load_vocab(g, hp.vocab_size, m->vocab);
    const std::string l0 = "encoder.layers.0.";
    const GgufTensor& dw0 = g.tensor(l0 + "conv.depthwise_conv.weight");
    hp.kernel_size = dw0.dims.size() > 1 ? (int)dw0.dims[1] : 0;  // inferred like src/nemo-ggml.cpp:357-360
    if (hp.kernel_size != KERNEL) throw Error("depthwise kernel size must be 9");
    hp.weights_quantized = g.tensor(l0 + "feed_forward1.linear1.weight").type != GGML_F32 &&
                           g.tensor(l0 + "feed_forward1.linear1.weight").type != GGML_F16;
    const int aq = flags & 3;
    m->act_quant = aq == NB200_FLAG_ACT_QUANT_ON || (aq == NB200_FLAG_ACT_QUANT_AUTO && hp.weights_quantized);
    hp.act_quant = m->act_quant;
    m->simt = (flags & NB200_FLAG_GEMM_SIMT) != 0;
    m->bf16 = (flags & NB200_FLAG_BF16) != 0;
    if (m->bf16 && m->act_quant) throw Error("NB200_FLAG_BF16 is for F32 / F16 files (or NB200_FLAG_ACT_QUANT_OFF): the int8 activation grid does not fit bf16's 8 mantissa bits");
    if (m->bf16 && m->simt) throw Error("NB200_FLAG_BF16 cannot be combined with the CUDA-core GEMM checker");
    m->use_graphs = !(flags & NB200_FLAG_NO_GRAPHS) && !getenv("NB200_NO_GRAPHS");

    // ---- mel tables (src/preprocessor.cpp:80-110, 270-307) ----
    {
        std::vector<float> fb = load_f32(g, "preprocessor.featurizer.fb", (int64_t)N_MELS * N_BINS);
        std::vector<float> win = load_f32(g, "preprocessor.featurizer.window", 400);
        std::vector<float> win512(N_FFT, 0.0f), cs(N_FFT), sn(N_FFT), fbT((size_t)N_BINS * N_MELS);
        std::copy(win.begin(), win.end(), win512.begin() + (N_FFT - 400) / 2);
        for (int i = 0; i < N_FFT; i++) {
            float theta = (2.0f * (float)M_PI * i) / N_FFT;
            sn[i] = sinf(theta);
            cs[i] = cosf(theta);
        }
        std::vector<int2> range(N_MELS);
        for (int mm = 0; mm < N_MELS; mm++) {
            int lo = N_BINS, hi = 0;
            for (int k = 0; k < N_BINS; k++) {
                float v = fb[(size_t)mm * N_BINS + k];
                fbT[(size_t)k * N_MELS + mm] = v;
                if (v != 0.0f) {
                    lo = std::min(lo, k);
                    hi = std::max(hi, k + 1);
                }
            }
            if (lo >= hi) lo = hi = 0;
            range[mm] = make_int2(lo, hi);
        }
        m->mel_tb.win512 = m->upload(win512);
        m->mel_tb.cos_tbl = m->upload(cs);
        m->mel_tb.sin_tbl = m->upload(sn);
        m->mel_tb.fbT = m->upload(fbT);
        m->mel_tb.fb_range = m->upload(range);
    }
    // ---- pre_encode (src/nemo-ggml.cpp:296-307) ----
    {
        const std::string pe = "encoder.pre_encode.";
        std::vector<float> w0 = load_f32(g, pe + "conv.0.weight", 256 * 9);
        for (auto& v : w0) v = __half2float(__float2half_rn(v));  // conv0 runs through an F16 im2col (ggml.c:4658-4680)
        m->sub.w0 = m->upload(w0);
        m->sub.b0 = m->upload(load_f32(g, pe + "conv.0.bias", 256));
        m->sub.w2 = m->upload(load_f32(g, pe + "conv.2.weight", 256 * 9));
        m->sub.b2 = m->upload(load_f32(g, pe + "conv.2.bias", 256));
        m->sub.w5 = m->upload(load_f32(g, pe + "conv.5.weight", 256 * 9));
        m->sub.b5 = m->upload(load_f32(g, pe + "conv.5.bias", 256));
        load_matrix(m, g, m->sub_w3, pe + "conv.3.weight", 256, 256);
        m->sub_b3 = m->upload(load_f32(g, pe + "conv.3.bias", 256));
        load_matrix(m, g, m->sub_w6, pe + "conv.6.weight", 256, 256);
        m->sub_b6 = m->upload(load_f32(g, pe + "conv.6.bias", 256));
        // out.weight [1024][c*17+w] -> [1024][w*256+c]: our activations are position-major, channel fastest
        make_f16(m, m->sub_wout, pe + "out.weight#w256c", D_MODEL, SUB_FEAT, [&] {
            std::vector<float> wo = load_f32(g, pe + "out.weight", (int64_t)D_MODEL * SUB_FEAT), wp(wo.size());
            for (int n = 0; n < D_MODEL; n++)
                for (int c = 0; c < SUB_CH; c++)
                    for (int w = 0; w < SUB_F3; w++) wp[(size_t)n * SUB_FEAT + w * SUB_CH + c] = wo[(size_t)n * SUB_FEAT + c * SUB_F3 + w];
            return wp;
        });
        m->sub_bout = m->upload(load_f32(g, pe + "out.bias", D_MODEL));
    }
    // ---- conformer layers (src/nemo-ggml.cpp:310-361) ----
    m->layers.resize(m->n_layers);
    for (int l = 0; l < m->n_layers; l++) {
        LayerW& L = m->layers[l];
        const std::string p = "encoder.layers." + std::to_string(l) + ".";
        auto vec = [&](const std::string& n, int64_t cnt) { return m->upload(load_f32(g, p + n, cnt)); };
        L.ln_ff1_g = vec("norm_feed_forward1.weight", D_MODEL); L.ln_ff1_b = vec("norm_feed_forward1.bias", D_MODEL);
        L.ln_att_g = vec("norm_self_att.weight", D_MODEL);      L.ln_att_b = vec("norm_self_att.bias", D_MODEL);
        L.ln_conv_g = vec("norm_conv.weight", D_MODEL);         L.ln_conv_b = vec("norm_conv.bias", D_MODEL);
        L.ln_ff2_g = vec("norm_feed_forward2.weight", D_MODEL); L.ln_ff2_b = vec("norm_feed_forward2.bias", D_MODEL);
        L.ln_out_g = vec("norm_out.weight", D_MODEL);           L.ln_out_b = vec("norm_out.bias", D_MODEL);
        L.bn_g = vec("conv.batch_norm.weight", D_MODEL);        L.bn_b = vec("conv.batch_norm.bias", D_MODEL);
        L.pos_u = vec("self_attn.pos_bias_u", D_MODEL);         L.pos_v = vec("self_attn.pos_bias_v", D_MODEL);
        L.w_dw = vec("conv.depthwise_conv.weight", (int64_t)KERNEL * D_MODEL);
        load_matrix(m, g, L.ff1_w1, p + "feed_forward1.linear1.weight", D_FF, D_MODEL, m->bf16);
        load_matrix(m, g, L.ff1_w2, p + "feed_forward1.linear2.weight", D_MODEL, D_FF, m->bf16);
        load_matrix(m, g, L.ff2_w1, p + "feed_forward2.linear1.weight", D_FF, D_MODEL, m->bf16);
        load_matrix(m, g, L.ff2_w2, p + "feed_forward2.linear2.weight", D_MODEL, D_FF, m->bf16);
        load_matrix(m, g, L.wpos, p + "self_attn.linear_pos.weight", D_MODEL, D_MODEL, m->bf16);
        load_matrix(m, g, L.wout, p + "self_attn.linear_out.weight", D_MODEL, D_MODEL, m->bf16);
        load_matrix(m, g, L.pw2, p + "conv.pointwise_conv2.weight", D_MODEL, D_MODEL, m->bf16);
        {   // q|k|v stacked: one N=3072 GEMM over the same (re-quantised) LN output
            const char* nm[3] = {"self_attn.linear_q.weight", "self_attn.linear_k.weight", "self_attn.linear_v.weight"};
            make_f16(m, L.wqkv, p + "self_attn.qkv", 3 * D_MODEL, D_MODEL, [&] {
                std::vector<float> w((size_t)3 * D_MODEL * D_MODEL);
                for (int i = 0; i < 3; i++) {
                    std::vector<float> part = load_f32(g, p + nm[i], (int64_t)D_MODEL * D_MODEL);
                    std::copy(part.begin(), part.end(), w.begin() + (size_t)i * D_MODEL * D_MODEL);
                }
                return w;
            }, "", m->bf16);
            for (int i = 0; i < 3; i++) m->by_name[p + nm[i]] = {L.wqkv.ptr + (size_t)i * D_MODEL * D_MODEL, D_MODEL, D_MODEL, false};
            m->by_name[p + "self_attn.qkv"] = {L.wqkv.ptr, 3 * D_MODEL, D_MODEL, false};  // the stacked matrix the step multiplies by (debug entry points)
        }
        {   // pointwise_conv1: rows 0..1023 = GLU value, 1024..2047 = gate -> interleave (value_j, gate_j)
            make_f16(m, L.pw1, p + "conv.pointwise_conv1.weight#glu", 2 * D_MODEL, D_MODEL, [&] {
                std::vector<float> w = load_f32(g, p + "conv.pointwise_conv1.weight", (int64_t)2 * D_MODEL * D_MODEL), wi(w.size());
                for (int j = 0; j < D_MODEL; j++) {
                    std::copy(w.begin() + (size_t)j * D_MODEL, w.begin() + (size_t)(j + 1) * D_MODEL, wi.begin() + (size_t)(2 * j) * D_MODEL);
                    std::copy(w.begin() + (size_t)(D_MODEL + j) * D_MODEL, w.begin() + (size_t)(D_MODEL + j + 1) * D_MODEL,
                              wi.begin() + (size_t)(2 * j + 1) * D_MODEL);
                }
                return wi;
            }, "", m->bf16);
            m->by_name[p + "conv.pointwise_conv1.weight"] = {L.pw1.ptr, 2 * D_MODEL, D_MODEL, true};
        }
    }
    // ---- decoder + joint (src/nemo-ggml.cpp:369-388), kept fp32 ----
    {
        const std::string d = "decoder.prediction.";
        m->dec.embed = m->upload(load_f32(g, d + "embed.weight", (int64_t)VOCAB * DEC));
        for (int l = 0; l < 2; l++) {
            const std::string s = std::to_string(l);
            auto wcat = [&] {  // [W_ih | W_hh] per gate row
                std::vector<float> wih = load_f32(g, d + "dec_rnn.lstm.weight_ih_l" + s, (int64_t)4 * DEC * DEC);
                std::vector<float> whh = load_f32(g, d + "dec_rnn.lstm.weight_hh_l" + s, (int64_t)4 * DEC * DEC);
                std::vector<float> wc((size_t)4 * DEC * 2 * DEC);
                for (int r = 0; r < 4 * DEC; r++) {
                    std::copy(wih.begin() + (size_t)r * DEC, wih.begin() + (size_t)(r + 1) * DEC, wc.begin() + (size_t)r * 2 * DEC);
                    std::copy(whh.begin() + (size_t)r * DEC, whh.begin() + (size_t)(r + 1) * DEC, wc.begin() + (size_t)r * 2 * DEC + DEC);
                }
                return wc;
            };
            m->dec.wcat[l] = m->upload_cached<float>(d + "dec_rnn.lstm.l" + s + "#cat", (size_t)4 * DEC * 2 * DEC, wcat);
            m->dec.wcat_s[l] = m->upload_cached<float>(d + "dec_rnn.lstm.l" + s + "#cat#hilo", (size_t)4 * DEC * 2 * DEC, [&] { return split_f16_pairs(wcat()); });
            std::vector<float> bih = load_f32(g, d + "dec_rnn.lstm.bias_ih_l" + s, 4 * DEC);
            std::vector<float> bhh = load_f32(g, d + "dec_rnn.lstm.bias_hh_l" + s, 4 * DEC);
            std::vector<float> bc(4 * DEC);
            for (int r = 0; r < 4 * DEC; r++) bc[r] = bih[r] + bhh[r];
            m->dec.bcat[l] = m->upload(bc);
        }
        {   // layer 0: recurrent half for the kernel, input half folded into the per-token table t0 (kernels.cuh DecodeWeights::t0)
            m->dec.whh0_s = m->upload_cached<float>(d + "dec_rnn.lstm.l0#hh#hilo", (size_t)4 * DEC * DEC,
                                                    [&] { return split_f16_pairs(load_f32(g, d + "dec_rnn.lstm.weight_hh_l0", (int64_t)4 * DEC * DEC)); });
            float* wih = m->alloc<float>((size_t)4 * DEC * DEC, false);  // dense [2560][640] copy of the left half of wcat[0]
            CUDA_CHECK(cudaMemcpy2D(wih, DEC * sizeof(float), m->dec.wcat[0], 2 * DEC * sizeof(float), DEC * sizeof(float), 4 * DEC, cudaMemcpyDeviceToDevice));
            float* t0 = m->alloc<float>((size_t)VOCAB * 4 * DEC);
            launch_sgemm(m->st, m->dec.embed, wih, m->dec.bcat[0], t0, VOCAB, 4 * DEC, DEC, 4 * DEC, nullptr);
            CUDA_CHECK(cudaStreamSynchronize(m->st));
            m->dec.t0 = t0;
        }
        m->dec.wpred = m->upload(load_f32(g, "joint.pred.weight", (int64_t)JOINT * DEC));
        m->dec.wpred_s = m->upload_cached<float>("joint.pred.weight#hilo", (size_t)JOINT * DEC, [&] { return split_f16_pairs(load_f32(g, "joint.pred.weight", (int64_t)JOINT * DEC)); });
        m->dec.bpred = m->upload(load_f32(g, "joint.pred.bias", JOINT));
        auto wout = [&] {
            std::vector<float> wo = load_f32(g, "joint.joint_net.2.weight", (int64_t)VOCAB * JOINT);
            wo.resize((size_t)1088 * JOINT, 0.0f);
            return wo;
        };
        std::vector<float> bo = load_f32(g, "joint.joint_net.2.bias", VOCAB);
        bo.resize(1088, -INFINITY);
        m->dec.wout = m->upload_cached<float>("joint.joint_net.2.weight#pad1088", (size_t)1088 * JOINT, wout);
        m->dec.wout_s = m->upload_cached<float>("joint.joint_net.2.weight#pad1088#hilo", (size_t)1088 * JOINT, [&] { return split_f16_pairs(wout()); });
        m->dec.bout = m->upload(bo);
        m->joint_enc_w = m->upload(load_f32(g, "joint.enc.weight", (int64_t)JOINT * D_MODEL));
        m->joint_enc_b = m->upload(load_f32(g, "joint.enc.bias", JOINT));
        {
            make_split3(m, m->joint_enc_w3, "joint.enc.weight", JOINT, D_MODEL, [&] { return load_f32(g, "joint.enc.weight", (int64_t)JOINT * D_MODEL); });
            m->joint_enc_tc = getenv("NB200_JOINT_SGEMM") == nullptr;
        }
    }
    m->num_prompts = hp.num_prompts;
    if (hp.num_prompts > 0) {  // the prompt kernel must be present exactly when the header advertises prompts (src/nemo-ggml.cpp:424-434)
        const int P = hp.num_prompts, H = PROMPT_HIDDEN;
        std::vector<float> w1 = load_f32(g, "prompt_kernel.0.weight", (int64_t)H * (D_MODEL + P));
        std::vector<float> a((size_t)H * D_MODEL), l((size_t)P * H);
        for (int n = 0; n < H; n++) {
            memcpy(&a[(size_t)n * D_MODEL], &w1[(size_t)n * (D_MODEL + P)], D_MODEL * sizeof(float));
            for (int q = 0; q < P; q++) l[(size_t)q * H + n] = w1[(size_t)n * (D_MODEL + P) + D_MODEL + q];
        }
        m->pk_w1 = m->upload(a);
        make_split3(m, m->pk_w1_3, "prompt_kernel.0.weight#enc", H, D_MODEL, [&] { return a; });
        m->pk_w1_lang = m->upload(l);
        m->pk_b1 = m->upload(load_f32(g, "prompt_kernel.0.bias", H));
        {
            const std::vector<float> w2 = load_f32(g, "prompt_kernel.2.weight", (int64_t)D_MODEL * H);
            m->pk_w2 = m->upload(w2);
            make_split3(m, m->pk_w2_3, "prompt_kernel.2.weight", D_MODEL, H, [&] { return w2; });
        }
        m->pk_b2 = m->upload(load_f32(g, "prompt_kernel.2.bias", D_MODEL));
        m->default_prompt = 101 < P ? 101 : 0;  // 101 = "auto" (src/nemo-ggml.cpp:459-462)
        auto lk = g.str_arr.find("nemo.prompt_langs");
        auto vk = g.i32_arr.find("nemo.prompt_ids");
        if (lk != g.str_arr.end() && vk != g.i32_arr.end())
            for (size_t i = 0; i < lk->second.size() && i < vk->second.size(); i++) m->prompt_dict[lk->second[i]] = vk->second[i];
    } else if (g.has("prompt_kernel.0.weight")) {
        throw Error("prompt_kernel weights present but nemo.num_prompts is 0");
    }
    if (m->wc_out) {
        try {
            m->wc_out->commit();
        } catch (const Error&) {  // the partial file is removed by the writer; loading does not depend on it
        }
        m->wc_out.reset();
    }
    m->wc_in.reset();  // unmap: everything is on the device now
}

}  // namespace

namespace {

constexpr int META_ROWS = 8;  // int arrays per batch: slot, wpos, valid, nvalid, n_frames, conv-cache parity, (2 spare)

void alloc_state(nb200_model* m) {
    const size_t S = (size_t)m->max_streams, L = (size_t)m->n_layers;
    m->k_ring = m->alloc<__half>(L * S * RING * D_MODEL);
    m->v_ring = m->alloc<__half>(L * S * RING * D_MODEL);
    m->conv_cache = m->alloc<float>(L * S * 2 * CONV_CACHE * D_MODEL);
    m->mel_overlap = m->alloc<float>(S * MEL_OVERLAP * N_MELS);
    m->dstate.h = m->alloc<float>(S * 2 * DEC);
    m->dstate.c = m->alloc<float>(S * 2 * DEC);
    m->dstate.pend_h = m->alloc<float>(S * 2 * DEC);
    m->dstate.pend_c = m->alloc<float>(S * 2 * DEC);
    m->dstate.pend_p = m->alloc<float>(S * JOINT);
    m->dstate.prev = m->alloc<int>(S);
    for (int i = m->max_streams - 1; i >= 0; i--) m->free_slots.push_back(i);

    const int B = m->cap_B = m->max_streams;
    const Mode big(13);
    const size_t rows = (size_t)B * MAX_T;
    const size_t rows_pad = (rows + 127) / 128 * 128;
    m->ws_mel = m->alloc<float>((size_t)B * big.chunk_mel * N_MELS);
    m->ws_a2h = m->alloc<__half>((size_t)B * big.t2 * SUB_F2 * SUB_CH + 128 * SUB_CH);
    m->ws_a3 = m->alloc<float>((size_t)B * big.t2 * SUB_F2 * SUB_CH);
    m->ws_a5h = m->alloc<__half>((rows_pad + 128) * SUB_F3 * SUB_CH);
    m->ws_x = m->alloc<float>(rows_pad * D_MODEL);
    m->ws_qkv16 = m->alloc<__half>(rows_pad * 3 * D_MODEL);
    m->ws_g = m->alloc<float>(rows_pad * D_MODEL);
    m->ws_e = m->alloc<float>(rows_pad * JOINT);
    m->ws_a16 = m->alloc<__half>(rows_pad * D_MODEL);
    m->ws_h16 = m->alloc<__half>(rows_pad * D_FF);
    m->ws_posf = m->alloc<float>((size_t)128 * D_MODEL);
    m->part_rows = rows_pad;
    m->part_max_splits = rows_pad <= 4096 ? 8 : (rows_pad <= 8192 ? 4 : (rows_pad <= 16384 ? 2 : 1));
    m->ws_part = m->alloc<float>((size_t)m->part_max_splits * rows_pad * D_MODEL);
    if (m->num_prompts > 0) {
        m->ws_pk = m->alloc<float>((size_t)rows_pad * PROMPT_HIDDEN);
        m->ws_pk16 = m->alloc<__half>((size_t)rows_pad * 3 * PROMPT_HIDDEN);
        operand_init(m->op_pk3, m->ws_pk16, (int)rows_pad, 3 * PROMPT_HIDDEN);
    }
    m->ws_pos16 = m->alloc<__half>((size_t)128 * D_MODEL);
    operand_init(m->op_a2h, m->ws_a2h, B * big.t2 * SUB_F2, SUB_CH);
    operand_init(m->op_a5h, m->ws_a5h, (int)rows * SUB_F3, SUB_CH);
    m->ws_a6h = m->alloc<__half>((rows_pad + 128) * SUB_FEAT);
    operand_init(m->op_a6h, m->ws_a6h, (int)rows_pad, SUB_FEAT);
    operand_init(m->op_a16, m->ws_a16, (int)rows_pad, D_MODEL);
    operand_init(m->op_h16, m->ws_h16, (int)rows_pad, D_FF);
    operand_init(m->op_x3, m->ws_h16, (int)rows_pad, 3 * D_MODEL);  // same memory, free once the last layer has run
    operand_init(m->op_pos16, m->ws_pos16, 128, D_MODEL);
    {   // device PCM ring per slot: default 2^17 samples (8.2 s); NB200_PCM_RING_LOG2 overrides (15..24)
        int lg = 17;
        if (const char* e = getenv("NB200_PCM_RING_LOG2")) lg = std::min(24, std::max(15, atoi(e)));
        m->ring_cap = 1 << lg;
        m->pcm_ring = m->alloc<int16_t>(S * (size_t)m->ring_cap);
    }
    m->ws_meta = m->alloc<int>((size_t)META_ROWS * B);
    m->ws_meta64 = m->alloc<long long>((size_t)3 * B);
    m->ws_fp = m->alloc<int>(B);
    m->ws_sym = m->alloc<int>(B);
    m->ws_tok = m->alloc<int>((size_t)B * MAX_T * MAX_SYMBOLS);
    m->ws_tokcnt = m->alloc<int>((size_t)2 * B);  // [B] counts + [B] margins (as float bits)
    m->ws_unfinished = m->alloc<int>(64);  // decode kernel counters (kernels.cuh DecodeBatch::counters)
    // decode window (frames evaluated per stream and round, decoder.cu window mode): 2 for the streaming step. Measured at 64
    // streams x 14 frames, all-blank decoding: 258 / 211 / 197 us per chunk at 1 / 2 / 4; 2 keeps the speculative extra work small
    // when tokens are frequent. NB200_DECODE_WINDOW overrides (1 = one symbol step per round, the checker).
    m->dec_window = 3;  // frames of a stream evaluated per decode round (decoder.cu window mode; measured 2 / 3 / 4: 587 / 560 / 577 us per chunk)
    if (const char* e = getenv("NB200_DECODE_WINDOW")) m->dec_window = std::max(1, std::min(8, atoi(e)));
    m->ws_partial = m->alloc<float>((size_t)decode_partial_floats(B) * m->dec_window);
    m->ws_hid = m->alloc<float>((size_t)B * JOINT * m->dec_window);
    m->ws_emit = m->alloc<int4>(B);
    m->d_reset = m->alloc<int>(B);
    CUDA_CHECK(cudaEventCreateWithFlags(&m->reset_ev, cudaEventDisableTiming));
    {   // upload arenas: >= two rc=13 chunks per slot each
        const size_t cap = std::max<size_t>((size_t)1 << 22, (size_t)B * 2 * (HOP * big.shift + 1024));
        for (auto& A : m->arena) {
            A.pcm.reserve(cap);
            A.segs.reserve(MAX_UPLOAD_SEGS);
            CUDA_CHECK(cudaEventCreateWithFlags(&A.done, cudaEventDisableTiming));
        }
        m->d_stage = m->alloc<int16_t>(cap, false);
        m->d_segs = m->alloc<UploadSeg>(MAX_UPLOAD_SEGS, false);
    }
    for (auto& e : m->ev) CUDA_CHECK(cudaEventCreate(&e));
    for (auto& e : m->tev) CUDA_CHECK(cudaEventCreate(&e));
}

// linear_pos(PE(p)) for every layer: constant per latency mode, so it is computed once here instead of in every layer
// of every chunk as the reference does (src/nemo-stream.cpp:516). Same arithmetic as any other encoder matmul.
ModeCtx* get_mode(nb200_model* m, int rc) {
    auto it = m->modes.find(rc);
    if (it != m->modes.end()) return it->second.get();
    auto ctx = std::make_unique<ModeCtx>(rc);
    const Mode& md = ctx->mode;
    ctx->h_meta.reserve((size_t)META_ROWS * m->cap_B);
    ctx->h_meta64.reserve((size_t)3 * m->cap_B);
    ctx->h_tok.reserve((size_t)m->cap_B * (MAX_T * MAX_SYMBOLS + 2) + 8);
    for (auto& e : ctx->ev) CUDA_CHECK(cudaEventCreate(&e));
    ctx->pos_tab = m->alloc<__half>((size_t)m->n_layers * md.npos * D_MODEL);
    launch_pos_emb(m->st, m->ws_posf, -(md.T - 1), md.npos);
    launch_cast_f16(m->st, m->ws_posf, m->ws_pos16, m->act_quant ? 1 : 0, md.npos, D_MODEL, m->bf16 ? 1 : 0);
    for (int l = 0; l < m->n_layers; l++) {
        EpiParams ep;
        ep.kind = EPI_STORE_F16;
        ep.out_f16 = ctx->pos_tab + (size_t)l * md.npos * D_MODEL;  // fp16 also for bf16 models: the attention kernels read it
        ep.ldo = D_MODEL;
        ep.ab_bf16 = m->bf16 ? 1 : 0;
        gemm(m->st, m->op_pos16, m->layers[l].wpos.op, md.npos, ep, m->simt);
    }
    CUDA_CHECK(cudaStreamSynchronize(m->st));
    ModeCtx* r = ctx.get();
    m->modes[rc] = std::move(ctx);
    return r;
}

DecodeBatch decode_batch(nb200_model* m, int B, int T, int W = 1) {
    DecodeBatch d{};
    d.e = m->ws_e;
    d.slot = m->ws_meta + 0 * m->cap_B;
    d.nvalid = m->ws_meta + 3 * m->cap_B;
    d.fp = m->ws_fp;
    d.sym = m->ws_sym;
    d.tok_out = m->ws_tok;
    d.tok_cnt = m->ws_tokcnt;
    d.min_margin = reinterpret_cast<float*>(m->ws_tokcnt + m->cap_B);
    d.counters = m->ws_unfinished;
    d.bar = reinterpret_cast<unsigned*>(m->ws_unfinished + 8);
    d.emit_list = m->ws_emit;
    d.partial = m->ws_partial;
    d.hid = m->ws_hid;
    d.max_rounds = T * (MAX_SYMBOLS + 1) + 2;
    d.B = B;
    d.T = T;
    d.W = W;
    return d;
}

// ---- stream create / reset: zero the slots' state slabs and recompute their pending decoder state --------------------------
// One kernel over all queued slots instead of 75 memsets + a host synchronisation per stream (1024 creates = 77 k calls):
// block (x = slot index in the list, y = layer, z = part). Parts 0 / 1 / 2 = the layer's K ring, V ring, conv cache of the slot;
// the blocks of layer 0 / part 0 also clear the slot's mel overlap and committed LSTM state and set prev = blank.
struct ResetArgs {
    const int* slots;
    __half *k_ring, *v_ring;
    float *conv_cache, *mel_overlap, *h, *c;
    int* prev;
    size_t S;  // slots per layer
};
__device__ __forceinline__ void zero_bytes(void* p, size_t bytes) {  // bytes % 16 == 0, p 16-byte aligned
    uint4* q = reinterpret_cast<uint4*>(p);
    const uint4 z = make_uint4(0, 0, 0, 0);
    for (size_t i = threadIdx.x; i < bytes / 16; i += blockDim.x) q[i] = z;
}
__global__ void __launch_bounds__(256) reset_slots_kernel(ResetArgs a) {
    pdl_prologue();
    const size_t slot = (size_t)a.slots[blockIdx.x], l = blockIdx.y;
    if (blockIdx.z == 0) zero_bytes(a.k_ring + (l * a.S + slot) * RING * D_MODEL, sizeof(__half) * RING * D_MODEL);
    else if (blockIdx.z == 1) zero_bytes(a.v_ring + (l * a.S + slot) * RING * D_MODEL, sizeof(__half) * RING * D_MODEL);
    else zero_bytes(a.conv_cache + (l * a.S + slot) * 2 * CONV_CACHE * D_MODEL, sizeof(float) * 2 * CONV_CACHE * D_MODEL);
    if (l == 0 && blockIdx.z == 0) {
        zero_bytes(a.mel_overlap + slot * MEL_OVERLAP * N_MELS, sizeof(float) * MEL_OVERLAP * N_MELS);
        zero_bytes(a.h + slot * 2 * DEC, sizeof(float) * 2 * DEC);
        zero_bytes(a.c + slot * 2 * DEC, sizeof(float) * 2 * DEC);
        if (threadIdx.x == 0) a.prev[slot] = BLANK;
    }
}

// Resets are queued (nb200_stream_create / _reset) and applied here, stream-ordered on the step's stream, by everything that
// touches slot state next (the next step, the offline path): no host synchronisation, and a create between nb200_step_begin and
// nb200_step_end neither waits for nor disturbs the step in flight (its kernels precede these in stream order).
void apply_resets(nb200_model* m) {
    if (m->reset_queue.empty()) return;
    const int n = (int)m->reset_queue.size();
    if (m->reset_busy) CUDA_CHECK(cudaEventSynchronize(m->reset_ev));  // the pinned list of the previous batch of resets is still being read
    m->h_reset.reserve((size_t)m->cap_B);
    std::copy(m->reset_queue.begin(), m->reset_queue.end(), m->h_reset.p);
    m->reset_queue.clear();
    cudaStream_t st = m->st;
    CUDA_CHECK(cudaMemcpyAsync(m->d_reset, m->h_reset.p, (size_t)n * sizeof(int), cudaMemcpyHostToDevice, st));
    CUDA_CHECK(cudaEventRecord(m->reset_ev, st));
    m->reset_busy = true;
    ResetArgs ra{m->d_reset, m->k_ring, m->v_ring, m->conv_cache, m->mel_overlap, m->dstate.h, m->dstate.c, m->dstate.prev, (size_t)m->max_streams};
    launch_k(reset_slots_kernel, dim3(n, m->n_layers, 3), dim3(256), 0, st, ra);
    NB_KERNEL_CHECK();
    DecodeBatch d = decode_batch(m, n, 1);
    d.slot = m->d_reset;
    launch_decode_refresh(st, m->dec, m->dstate, d);
}

void reset_slot(nb200_model* m, int slot) {
    if (std::find(m->reset_queue.begin(), m->reset_queue.end(), slot) == m->reset_queue.end()) m->reset_queue.push_back(slot);
}

// ---- small copy kernels of the step ---------------------------------------------------------------------
// chunk[b][0:9] = overlap[slot[b]]; chunk[b][9+n_frames[b] : chunk_mel] = 0 (tail chunk of a finalize)
__global__ void mel_assemble_kernel(float* chunk, const float* overlap, const int* slot, const int* n_frames, int chunk_mel) {
    pdl_prologue();
    const int b = blockIdx.y, f = blockIdx.x, t = threadIdx.x;  // 128 threads
    float* dst = chunk + ((size_t)b * chunk_mel + f) * N_MELS + t;
    if (f < MEL_OVERLAP) *dst = overlap[((size_t)slot[b] * MEL_OVERLAP + f) * N_MELS + t];
    else if (f >= MEL_OVERLAP + n_frames[b]) *dst = 0.0f;
}
// overlap[slot[b]] = chunk[b][shift : shift+9]
__global__ void mel_save_overlap_kernel(const float* chunk, float* overlap, const int* slot, int chunk_mel, int shift) {
    pdl_prologue();
    const int b = blockIdx.y, f = blockIdx.x, t = threadIdx.x;
    overlap[((size_t)slot[b] * MEL_OVERLAP + f) * N_MELS + t] = chunk[((size_t)b * chunk_mel + shift + f) * N_MELS + t];
}

long long frames_available(long long n_samples) {  // src/preprocessor.cpp:320-328 with the 256-sample centre pad
    return n_samples + 256 < N_FFT ? 0 : (n_samples + 256 - N_FFT) / HOP + 1;
}

// oldest absolute sample the stream's next chunk still needs (everything before it may be overwritten in the ring)
long long keep_from(const nb200_stream* s) {
    const Mode md(s->rc);
    return std::max<long long>(0, (long long)HOP * md.shift * s->chunks_done - 257);
}

// Scatter of one flushed arena into the slots' PCM rings: block (x = segment, y = 4096-sample piece).
__global__ void __launch_bounds__(256) pcm_scatter_kernel(const int16_t* __restrict__ stage, const UploadSeg* __restrict__ segs, int16_t* __restrict__ ring) {
    pdl_prologue();
    const UploadSeg sg = segs[blockIdx.x];
    const int16_t* src = stage + sg.src;
    int16_t* dst = ring + sg.dst;
    const int end = min(sg.n, (int)(blockIdx.y + 1) * 4096);
    for (int i = blockIdx.y * 4096 + threadIdx.x; i < end; i += 256) dst[i] = src[i];
}

// Sends the current arena to the device (one H2D copy of the samples, one of the segment table, one scatter kernel) and
// switches to the other arena. No host synchronisation unless that one is still in flight. The work goes to the COPY
// stream, so audio fed while a step is running (nb200_feed_batch between nb200_step_begin and _end) is already in the
// rings when the next step starts; that step's stream waits for `upload_ev`. Ring positions written here are never ones
// a launched-but-unfinished chunk still reads (upload_pcm only takes what fits before keep_from()).
void flush_uploads(nb200_model* m) {
    UploadArena& A = m->arena[m->cur_arena];
    if (A.nseg == 0) return;
    cudaStream_t st = m->st_copy;
    CUDA_CHECK(cudaMemcpyAsync(m->d_stage, A.pcm.p, A.off * sizeof(int16_t), cudaMemcpyHostToDevice, st));
    CUDA_CHECK(cudaMemcpyAsync(m->d_segs, A.segs.p, (size_t)A.nseg * sizeof(UploadSeg), cudaMemcpyHostToDevice, st));
    launch_k(pcm_scatter_kernel, dim3(A.nseg, cdiv(A.max_n, 4096)), dim3(256), 0, st, (const int16_t*)m->d_stage, (const UploadSeg*)m->d_segs, m->pcm_ring);
    NB_KERNEL_CHECK();
    CUDA_CHECK(cudaEventRecord(A.done, st));
    A.busy = true;
    CUDA_CHECK(cudaEventRecord(m->upload_ev, st));
    m->upload_wait = true;
    m->stats.h2d_bytes += (long long)(A.off * sizeof(int16_t) + (size_t)A.nseg * sizeof(UploadSeg));
    m->cur_arena ^= 1;
    UploadArena& N = m->arena[m->cur_arena];
    if (N.busy) {
        CUDA_CHECK(cudaEventSynchronize(N.done));
        N.busy = false;
    }
    N.off = 0;
    N.nseg = 0;
    N.max_n = 0;
}

// Stage up to n samples for the slot's device ring (at most two segments when the ring wraps). Returns the number of
// samples taken (less than n when the ring is full). The data reaches the device at the next flush_uploads().
size_t upload_pcm(nb200_model* m, nb200_stream* s, const int16_t* pcm, size_t n) {
    const long long space = (long long)m->ring_cap - (s->fed - keep_from(s));
    n = (size_t)std::max<long long>(0, std::min<long long>((long long)n, space));
    size_t done = 0;
    while (done < n) {
        UploadArena* A = &m->arena[m->cur_arena];
        if (A->off >= A->pcm.n || A->nseg >= MAX_UPLOAD_SEGS) {  // arena full: send it and continue in the other one
            flush_uploads(m);
            A = &m->arena[m->cur_arena];
        }
        const size_t pos = (size_t)(s->fed & (long long)(m->ring_cap - 1));
        const size_t k = std::min({n - done, A->pcm.n - A->off, (size_t)m->ring_cap - pos});
        memcpy(A->pcm.p + A->off, pcm + done, k * sizeof(int16_t));
        A->segs.p[A->nseg++] = UploadSeg{(long long)s->slot * m->ring_cap + (long long)pos, (int)A->off, (int)k};
        A->max_n = std::max(A->max_n, (int)k);
        A->off += k;
        s->fed += (long long)k;
        done += k;
    }
    return n;
}

void flush_pending(nb200_model* m, nb200_stream* s) {
    if (s->pending.empty()) return;
    const size_t k = upload_pcm(m, s, s->pending.data(), s->pending.size());
    s->pending.erase(s->pending.begin(), s->pending.begin() + k);
}

void run_batch_launch(nb200_model* m, ModeCtx* mc, std::vector<BatchItem>& items);
void run_batch_finish(nb200_model* m, ModeCtx* mc, std::vector<BatchItem>& items);

}  // namespace

namespace {

// LayerNorm + 24 conformer layers (+ the multilingual prompt fusion) over the M = B * T rows in ws_x; `attn(l)` turns layer l's
// q | k | v (ws_qkv16) into the attention context (ws_a16). Shared by the streaming step and the offline path. Returns the
// encoder output ([M][1024] fp32, ws_x or ws_g).
template <typename AttnFn>
const float* enqueue_layers(nb200_model* m, int B, int T, AttnFn&& attn, bool kv_prefetch = false) {
    const int M = B * T, capB = m->cap_B;
    const int aq = m->act_quant ? 1 : 0;
    const bool simt = m->simt;
    cudaStream_t st = m->st;
    const int* d_slot = m->ws_meta + 0 * capB;
    const int* d_parity = m->ws_meta + 5 * capB;
    // 24 cache-aware conformer layers (src/nemo-stream.cpp:605-690)
    const size_t S = (size_t)m->max_streams;
    const int bf = m->bf16 ? 1 : 0;  // bf16 GEMM operands: every kernel that writes an A operand stores bf16
    launch_layernorm(st, m->ws_x, m->layers[0].ln_ff1_g, m->layers[0].ln_ff1_b, nullptr, nullptr, nullptr, m->ws_a16, aq, M, ResidualIn(), bf);
    for (int l = 0; l < m->n_layers; l++) {
        LayerW& L = m->layers[l];
        EpiParams up;  // linear1 + SiLU (+ re-quantise) -> fp16 [M,4096]
        up.kind = EPI_STORE_F16; up.act = ACT_SILU; up.act_quant = aq; up.out_f16 = m->ws_h16; up.ldo = D_FF; up.out_bf16 = bf; up.ab_bf16 = bf;
        // The four residual GEMMs (N = 1024) write split-K partials; the LayerNorm that follows adds them to x.
        EpiParams part;
        part.kind = EPI_PARTIAL; part.out_f32 = m->ws_part; part.ldo = D_MODEL; part.split_stride = m->part_rows * D_MODEL; part.ab_bf16 = bf;
        ResidualIn rin;
        rin.partials = m->ws_part; rin.stride = part.split_stride;
        // every GEMM asks for the NEXT GEMM's weights to be brought into L2 while its own main loop runs (EpiParams::pf_ptr)
        auto next_w = [](EpiParams& e, const F16Matrix* w) {
            e.pf_ptr = w ? w->ptr : nullptr;
            e.pf_bytes = w ? (size_t)w->op.rows * w->op.K * sizeof(__half) : 0;
        };
        // ... and, during FFN1, for this layer's K / V ring slabs of the batch's streams (read by the attention kernel 4 launches later)
        auto kv_slabs = [&](EpiParams& e, const __half* ring) {
            e.pf2_base = kv_prefetch ? ring + (size_t)l * S * RING * D_MODEL : nullptr;
            e.pf2_stride = (size_t)RING * D_MODEL * sizeof(__half);
            e.pf2_bytes = (unsigned)(RING * D_MODEL * sizeof(__half));
            e.pf2_slot = d_slot;
            e.pf2_count = B;
        };
        next_w(up, &L.ff1_w2);
        kv_slabs(up, m->k_ring);
        gemm(st, m->op_a16, L.ff1_w1.op, M, up, simt);
        up.pf2_base = nullptr;
        next_w(part, &L.wqkv);
        kv_slabs(part, m->v_ring);
        rin.nsplit = gemm(st, m->op_h16, L.ff1_w2.op, M, part, simt, m->part_max_splits);
        part.pf2_base = nullptr;
        rin.alpha = 0.5f;  // x += 0.5 * FFN1
        // self-attention
        launch_layernorm(st, m->ws_x, L.ln_att_g, L.ln_att_b, nullptr, nullptr, nullptr, m->ws_a16, aq, M, rin, bf);
        EpiParams qkv;  // q | k | v of the chunk's frames, dense fp16; the attention kernel appends k, v to the stream rings
        qkv.kind = EPI_STORE_F16; qkv.out_f16 = m->ws_qkv16; qkv.ldo = 3 * D_MODEL; qkv.ab_bf16 = bf;  // q | k | v stay fp16: the attention kernels' type
        next_w(qkv, &L.wout);
        gemm(st, m->op_a16, L.wqkv.op, M, qkv, simt);
        attn(l);  // q | k | v in ws_qkv16 -> context in ws_a16 (cached streaming attention, or full-sequence attention offline)
        next_w(part, &L.pw1);
        rin.nsplit = gemm(st, m->op_a16, L.wout.op, M, part, simt, m->part_max_splits);
        rin.alpha = 1.0f;  // x += attention
        // convolution module
        launch_layernorm(st, m->ws_x, L.ln_conv_g, L.ln_conv_b, nullptr, nullptr, nullptr, m->ws_a16, aq, M, rin, bf);
        EpiParams glu;
        glu.kind = EPI_GLU; glu.out_f32 = m->ws_g; glu.ldo = D_MODEL; glu.ab_bf16 = bf;
        next_w(glu, &L.pw2);
        gemm(st, m->op_a16, L.pw1.op, M, glu, simt);
        ConvArgs cv{};
        cv.g = m->ws_g; cv.conv_cache = m->conv_cache + (size_t)l * S * 2 * CONV_CACHE * D_MODEL; cv.parity = d_parity; cv.w_dw = L.w_dw;
        cv.ln_g = L.bn_g; cv.ln_b = L.bn_b; cv.slot = d_slot; cv.out = m->ws_a16; cv.B = B; cv.T = T; cv.act_quant = aq; cv.out_bf16 = bf;
        launch_dwconv(st, cv);
        next_w(part, &L.ff2_w1);
        rin.nsplit = gemm(st, m->op_a16, L.pw2.op, M, part, simt, m->part_max_splits);
        // FFN2 and the layer's output norm (fused with the next layer's first norm)
        launch_layernorm(st, m->ws_x, L.ln_ff2_g, L.ln_ff2_b, nullptr, nullptr, nullptr, m->ws_a16, aq, M, rin, bf);  // x += conv module
        next_w(up, &L.ff2_w2);
        gemm(st, m->op_a16, L.ff2_w1.op, M, up, simt);
        next_w(part, l + 1 < m->n_layers ? &m->layers[l + 1].ff1_w1 : nullptr);
        rin.nsplit = gemm(st, m->op_h16, L.ff2_w2.op, M, part, simt, m->part_max_splits);
        rin.alpha = 0.5f;  // x += 0.5 * FFN2, then x = LN_out(x)
        if (l + 1 < m->n_layers)
            launch_layernorm(st, m->ws_x, L.ln_out_g, L.ln_out_b, m->ws_x, m->layers[l + 1].ln_ff1_g, m->layers[l + 1].ln_ff1_b, m->ws_a16, aq, M, rin, bf);
        else
            launch_layernorm(st, m->ws_x, L.ln_out_g, L.ln_out_b, m->ws_x, nullptr, nullptr, nullptr, 0, M, rin);
    }
    // multilingual checkpoints: language-prompt fusion on the encoder output (src/nemo-ggml.cpp:1087-1105, src/nemo-stream.cpp:236-242),
    // fp32 like the file: relu(fc1 . [enc ; onehot]) = relu(enc . W1a^T + b1 + W1[:, 1024 + lang]), then fc2
    const float* enc_out = m->ws_x;
    if (m->num_prompts > 0) {
        if (m->joint_enc_tc && !simt) {  // split-precision tensor-core GEMMs (fp32 quality), like joint.enc below
            launch_split_f16x3(st, m->ws_x, m->ws_h16, M);
            EpiParams e1;
            e1.kind = EPI_STORE; e1.bias = m->pk_b1; e1.out_f32 = m->ws_pk; e1.ldo = PROMPT_HIDDEN;
            gemm(st, m->op_x3, m->pk_w1_3.op, M, e1, false);
            launch_prompt_add_relu(st, m->ws_pk, m->pk_w1_lang, m->ws_meta + 6 * capB, T, M);
            launch_split_f16x3(st, m->ws_pk, m->ws_pk16, M, PROMPT_HIDDEN);
            EpiParams e2;
            e2.kind = EPI_STORE; e2.bias = m->pk_b2; e2.out_f32 = m->ws_g; e2.ldo = D_MODEL;
            gemm(st, m->op_pk3, m->pk_w2_3.op, M, e2, false);
        } else {
            launch_sgemm(st, m->ws_x, m->pk_w1, m->pk_b1, m->ws_pk, M, PROMPT_HIDDEN, D_MODEL, PROMPT_HIDDEN, nullptr);
            launch_prompt_add_relu(st, m->ws_pk, m->pk_w1_lang, m->ws_meta + 6 * capB, T, M);
            launch_sgemm(st, m->ws_pk, m->pk_w2, m->pk_b2, m->ws_g, M, D_MODEL, PROMPT_HIDDEN, D_MODEL, nullptr);
        }
        enc_out = m->ws_g;
    }
    return enc_out;
}

// joint.enc over all frames (fp32 quality: split-precision tensor-core GEMM, or the fp32 CUDA-core checker)
void enqueue_joint_enc(nb200_model* m, const float* enc_out, int M) {
    cudaStream_t st = m->st;
    if (m->joint_enc_tc && !m->simt) {
        launch_split_f16x3(st, enc_out, m->ws_h16, M);
        EpiParams ej;
        ej.kind = EPI_STORE; ej.bias = m->joint_enc_b; ej.out_f32 = m->ws_e; ej.ldo = JOINT;
        gemm(st, m->op_x3, m->joint_enc_w3.op, M, ej, false);
    } else {
        launch_sgemm(st, enc_out, m->joint_enc_w, m->joint_enc_b, m->ws_e, M, JOINT, D_MODEL, JOINT, nullptr);
    }
}

// Adds the kernels launched inside a scope (this thread's counter) to the model's own total: g_launches is per thread, so
// several models stepping on their own threads (server_loop.h ShardedWorkers) do not race on it.
// Temporary device buffers of the synchronous entry points, freed on every path (including a throw)
struct DevBufs {
    std::vector<void*> v;
    DevBufs() = default;
    DevBufs(const DevBufs&) = delete;
    ~DevBufs() { for (void* p : v) cudaFree(p); }
    template <typename T> T* get(size_t n) { T* p = dalloc<T>(n); v.push_back(p); return p; }
};

struct LaunchCount {
    nb200_model* m;
    long long t0;
    explicit LaunchCount(nb200_model* mm) : m(mm), t0(g_launches) {}
    ~LaunchCount() { m->launches += g_launches - t0; }
};

struct StepCtx {
    nb200_model* m;
    ModeCtx* mc;
    int B;
    bool taps;
    std::vector<BatchItem>* items;
};

// ---- frontend + 24 conformer layers + joint.enc: no host synchronisation inside unless parity taps are on ----
void enqueue_encoder(const StepCtx& c) {
    nb200_model* m = c.m;
    ModeCtx* mc = c.mc;
    const Mode& md = mc->mode;
    const int B = c.B, T = md.T, M = B * T, capB = m->cap_B;
    const int aq = m->act_quant ? 1 : 0;
    const bool simt = m->simt;
    cudaStream_t st = m->st;
    const int* d_slot = m->ws_meta + 0 * capB;
    const int* d_wpos = m->ws_meta + 1 * capB;
    const int* d_valid = m->ws_meta + 2 * capB;
    const int* d_nframes = m->ws_meta + 4 * capB;

    // frontend: log-mel of the new samples, then dw-striding subsampling (src/nemo-ggml.cpp:954-1029)
    launch_k(mel_assemble_kernel, dim3(dim3(md.chunk_mel, B)), dim3(N_MELS), 0, st, m->ws_mel, m->mel_overlap, d_slot, d_nframes, md.chunk_mel);
    NB_KERNEL_CHECK();
    {
        MelBatch mb{m->pcm_ring, m->ring_cap, d_slot, d_nframes, m->ws_meta64 + 1 * capB, m->ws_meta64 + 2 * capB, m->ws_meta64 + 0 * capB};
        launch_mel(st, m->mel_tb, mb, m->ws_mel, B, md.shift);
    }
    if (c.taps) {
        for (int b = 0; b < B; b++) {
            nb200_stream* s = (*c.items)[b].s;
            const int nfr = (*c.items)[b].n_frames;
            if ((s->taps & NB200_TAP_MEL) && nfr > 0) {
                const size_t n = (size_t)nfr * N_MELS, old = s->tap_mel.size();
                s->tap_mel.resize(old + n);
                CUDA_CHECK(cudaMemcpyAsync(s->tap_mel.data() + old, m->ws_mel + ((size_t)b * md.chunk_mel + MEL_OVERLAP) * N_MELS,
                                           n * sizeof(float), cudaMemcpyDeviceToHost, st));
                CUDA_CHECK(cudaStreamSynchronize(st));
            }
        }
    }
    launch_sub_stage1(st, m->sub, m->ws_mel, m->ws_a2h, B, md.chunk_mel, md.t1, md.t2);
    launch_k(mel_save_overlap_kernel, dim3(dim3(MEL_OVERLAP, B)), dim3(N_MELS), 0, st, m->ws_mel, m->mel_overlap, d_slot, md.chunk_mel, md.shift);
    NB_KERNEL_CHECK();
    {
        EpiParams ep;  // pw3 + bias + ReLU -> fp32 (input of the fp32 depthwise conv5)
        ep.kind = EPI_STORE; ep.act = ACT_RELU; ep.bias = m->sub_b3; ep.out_f32 = m->ws_a3; ep.ldo = SUB_CH;
        gemm(st, m->op_a2h, m->sub_w3.op, B * md.t2 * SUB_F2, ep, simt);
        launch_sub_stage2(st, m->sub, m->ws_a3, m->ws_a5h, B, md.t2, md.t3);
        EpiParams e6;  // pw6 + bias + ReLU -> fp16 [B*T*17][256], whose rows (b,t) are the K=4352 operand of pre_encode.out
        e6.kind = EPI_STORE_F16; e6.act = ACT_RELU; e6.bias = m->sub_b6; e6.out_f16 = m->ws_a6h; e6.ldo = SUB_CH; e6.act_quant = 0;
        gemm(st, m->op_a5h, m->sub_w6.op, M * SUB_F3, e6, simt);
        EpiParams eo;
        eo.kind = EPI_STORE; eo.bias = m->sub_bout; eo.out_f32 = m->ws_x; eo.ldo = D_MODEL;
        gemm(st, m->op_a6h, m->sub_wout.op, M, eo, simt);
    }

    // 24 cache-aware conformer layers (src/nemo-stream.cpp:605-690)
    const size_t S = (size_t)m->max_streams;
    const float* enc_out = enqueue_layers(m, B, T, [&](int l) {
        LayerW& L = m->layers[l];
        AttnArgs at{};
        at.qkv = m->ws_qkv16; at.k_ring = m->k_ring + (size_t)l * S * RING * D_MODEL; at.v_ring = m->v_ring + (size_t)l * S * RING * D_MODEL;
        at.pos = mc->pos_tab + (size_t)l * md.npos * D_MODEL;
        at.bias_u = L.pos_u; at.bias_v = L.pos_v; at.slot = d_slot; at.wpos = d_wpos; at.valid = d_valid;
        at.out = m->ws_a16; at.B = B; at.T = T; at.act_quant = aq; at.out_bf16 = m->bf16 ? 1 : 0;
        launch_attention(st, at);
    }, /*kv_prefetch=*/true);
    if (c.taps) {
        for (int b = 0; b < B; b++) {
            nb200_stream* s = (*c.items)[b].s;
            if (s->taps & NB200_TAP_ENC) {
                const size_t n = (size_t)T * D_MODEL, old = s->tap_enc.size();
                s->tap_enc.resize(old + n);
                CUDA_CHECK(cudaMemcpyAsync(s->tap_enc.data() + old, enc_out + (size_t)b * T * D_MODEL, n * sizeof(float), cudaMemcpyDeviceToHost, st));
                CUDA_CHECK(cudaStreamSynchronize(st));
            }
        }
    }
    // joint.enc over all frames of the batch, and the decode loop's per-step state
    enqueue_joint_enc(m, enc_out, M);
    launch_decode_chunk(st, m->dec, m->dstate, decode_batch(m, B, T, std::min(m->dec_window, T)));  // every greedy round of the chunk, one persistent launch
}

// Runs `fn` through a CUDA graph keyed by `key`: eagerly on the first request, captured on the second, replayed after.
template <typename F>
void run_graphed(nb200_model* m, std::map<long long, GraphExec>& cache, long long key, bool allow_graph, F&& fn) {
    if (!allow_graph) {
        fn();
        return;
    }
    GraphExec& g = cache[key];
    g.seen++;
    if (g.seen == 1) {
        fn();
        return;
    }
    if (!g.exec) {
        const long long before = g_launches;
        cudaGraph_t graph = nullptr;
        CUDA_CHECK(cudaStreamBeginCapture(m->st, cudaStreamCaptureModeThreadLocal));
        try {
            fn();
        } catch (...) {
            cudaStreamEndCapture(m->st, &graph);
            if (graph) cudaGraphDestroy(graph);
            throw;
        }
        CUDA_CHECK(cudaStreamEndCapture(m->st, &graph));
        g.launches = g_launches - before;
        g_launches = before;  // nothing ran yet; replays are counted below
        CUDA_CHECK(cudaGraphInstantiate(&g.exec, graph, 0));
        CUDA_CHECK(cudaGraphDestroy(graph));
    }
    CUDA_CHECK(cudaGraphLaunch(g.exec, m->st));
    g_launches += g.launches;
}


// Enqueues one batched chunk (metadata upload, encoder + decode graph, token copies) without waiting for it.
void run_batch_launch(nb200_model* m, ModeCtx* mc, std::vector<BatchItem>& items) {
    const Mode& md = mc->mode;
    const int B = (int)items.size(), T = md.T, capB = m->cap_B;
    cudaStream_t st = m->st;
    LaunchCount lc(m);

    // ---- per-item metadata (the audio itself is already resident in the slots' device PCM rings) ----
    int* meta = mc->h_meta.p;
    long long* meta64 = mc->h_meta64.p;
    bool taps = false;
    for (int b = 0; b < B; b++) {
        nb200_stream* s = items[b].s;
        taps = taps || s->taps != 0;
        meta[0 * capB + b] = s->slot;
        meta[1 * capB + b] = s->wpos;
        meta[2 * capB + b] = s->valid;
        meta[3 * capB + b] = items[b].nvalid;
        meta[4 * capB + b] = items[b].n_frames;
        meta[5 * capB + b] = (int)(s->chunks_done & 1);
        meta[6 * capB + b] = (s->prompt_index >= 0 && s->prompt_index < m->num_prompts) ? s->prompt_index : 0;  // src/nemo-stream.cpp:1052-1053
        meta64[0 * capB + b] = ((long long)b * md.chunk_mel + MEL_OVERLAP) * N_MELS;
        // frame f of this chunk covers absolute samples a0 + 160 f + [0, 513): a0 = first new frame's first sample - 1
        meta64[1 * capB + b] = (long long)HOP * md.shift * s->chunks_done - 257;
        meta64[2 * capB + b] = s->fed;
    }
    CUDA_CHECK(cudaEventRecord(mc->ev[0], st));
    CUDA_CHECK(cudaMemcpyAsync(m->ws_meta, meta, (size_t)META_ROWS * capB * sizeof(int), cudaMemcpyHostToDevice, st));
    CUDA_CHECK(cudaMemcpyAsync(m->ws_meta64, meta64, (size_t)3 * capB * sizeof(long long), cudaMemcpyHostToDevice, st));
    m->stats.h2d_bytes += (long long)META_ROWS * capB * sizeof(int) + (long long)3 * capB * sizeof(long long);

    // ---- encoder: one CUDA graph per (latency mode, batch size); eager when taps / profiling need host access ----
    const bool allow_graph = m->use_graphs && !taps && !g_prof_on;
    StepCtx ctx{m, mc, B, taps, &items};
    run_graphed(m, m->enc_graphs, (long long)md.rc * 1000003 + B, allow_graph, [&] { enqueue_encoder(ctx); });
    CUDA_CHECK(cudaEventRecord(mc->ev[2], st));

    // The greedy rounds (src/nemo-stream.cpp:1106-1118, 840-930) ran inside the same graph as one persistent launch.
    // pinned result buffer: [0..7] = decode counters 6..13 (streams left unfinished, rounds run, barrier word, -, the kernel's
    // start / end globaltimer stamps), then token counts [capB], margins [capB], tokens [B][T*MAX_SYMBOLS]: three copies
    int* h_unf = mc->h_tok.p;
    int* h_cnt = mc->h_tok.p + 8;
    int* h_toks = h_cnt + 2 * capB;
    CUDA_CHECK(cudaMemcpyAsync(h_unf, m->ws_unfinished + 6, 8 * sizeof(int), cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaMemcpyAsync(h_cnt, m->ws_tokcnt, sizeof(int) * 2 * capB, cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaMemcpyAsync(h_toks, m->ws_tok, sizeof(int) * (size_t)B * T * MAX_SYMBOLS, cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaEventRecord(mc->ev[3], st));
}

// Waits for the batch launched by run_batch_launch and does the host bookkeeping (tokens, cache positions, statistics).
void run_batch_finish(nb200_model* m, ModeCtx* mc, std::vector<BatchItem>& items) {
    const Mode& md = mc->mode;
    const int B = (int)items.size(), T = md.T, capB = m->cap_B;
    int* h_unf = mc->h_tok.p;
    int* h_cnt = mc->h_tok.p + 8;
    int* h_margin = h_cnt + capB;
    int* h_toks = h_cnt + 2 * capB;
    CUDA_CHECK(cudaEventSynchronize(mc->ev[3]));
    m->stats.d2h_bytes += (long long)sizeof(int) * ((size_t)2 * capB + (size_t)B * T * MAX_SYMBOLS + 8);
    if (h_unf[0] != 0) throw Error("decode loop did not terminate");
    if (getenv("NB200_DK_TRACE")) {
        unsigned long long a[16];
        CUDA_CHECK(cudaMemcpy(a, m->ws_unfinished + 16, sizeof a, cudaMemcpyDeviceToHost));
        fprintf(stderr,
                "decode phases (us): L %.1f (compact %.1f stage %.1f rows %.1f) bar1 %.1f A %.1f bar2 %.1f | E: lstm stage %.1f mma %.1f gates %.1f bar %.1f, pred stage %.1f "
                "mma %.1f bar %.1f, %llu rounds with %llu emissions | rounds %d\n",
                a[0] * 1e-3, a[4] * 1e-3, a[5] * 1e-3, a[6] * 1e-3, a[1] * 1e-3, a[2] * 1e-3, a[3] * 1e-3, a[7] * 1e-3, a[8] * 1e-3, a[9] * 1e-3, a[10] * 1e-3,
                a[11] * 1e-3, a[12] * 1e-3, a[13] * 1e-3, a[14], a[15], h_unf[1]);
    }
    const long long rounds = h_unf[1];

    // ---- host bookkeeping ----
    long long frames = 0, toks = 0;
    for (int b = 0; b < B; b++) {
        nb200_stream* s = items[b].s;
        const int n = h_cnt[b];
        for (int i = 0; i < n; i++) {  // word = token id | chunk-local frame << 16 (decoder.cu, phase A)
            const int w = h_toks[(size_t)b * T * MAX_SYMBOLS + i];
            s->tokens.push_back(w & 0xffff);
            s->token_frames.push_back(s->frames_done + (w >> 16));
        }
        s->frames_done += items[b].nvalid;
        s->chunk_tokens.push_back(n);
        float mg;
        memcpy(&mg, &h_margin[b], sizeof(float));
        s->chunk_margins.push_back(mg);
        toks += n;
        frames += items[b].nvalid;
        s->chunks_done++;
        s->wpos = (s->wpos + T) % RING;
        s->valid = std::min(s->valid + T, LEFT);  // src/nemo-stream.cpp:1085
        m->stats.audio_seconds += (double)items[b].n_frames * HOP / 16000.0;
    }
    float ms02 = 0, ms23 = 0;
    CUDA_CHECK(cudaEventElapsedTime(&ms02, mc->ev[0], mc->ev[2]));
    CUDA_CHECK(cudaEventElapsedTime(&ms23, mc->ev[2], mc->ev[3]));
    m->stats.steps++;
    m->stats.chunks += B;
    m->stats.frames += frames;
    m->stats.tokens += toks;
    m->stats.decode_rounds += rounds;
    m->stats.last_frontend_ms = 0.0;  // the frontend is part of the encoder graph
    // the decode kernel stamps its own start / end (globaltimer ns) into counters[10..13]
    unsigned long long dk0, dk1;
    memcpy(&dk0, h_unf + 4, 8);
    memcpy(&dk1, h_unf + 6, 8);
    const double dec_ms = dk1 > dk0 ? (double)(dk1 - dk0) * 1e-6 : 0.0;
    m->stats.last_encoder_ms = ms02 + ms23 - dec_ms;
    m->stats.last_decoder_ms = dec_ms;
    m->stats.last_step_gpu_ms = ms02 + ms23;
    m->stats.gpu_seconds += (ms02 + ms23) * 1e-3;
    m->stats.kernel_launches = m->launches;
}

}  // namespace

// =====================================================================================================
// C ABI
// =====================================================================================================
#define NB_TRY try {
#define NB_CATCH(ret)                         \
    }                                         \
    catch (const std::exception& e) {         \
        set_error(e.what());                  \
        return ret;                           \
    }

extern "C" {

static int finish_inflight(nb200_model* m);

const char* nb200_last_error(void) { return g_error.c_str(); }
int nb200_version(void) { return 100; }

// Host-only: parses the GGUF container (header, KVs, tensor table, bounds of every tensor) and the vocabulary the way
// nb200_model_load does, without touching a GPU.
int nb200_gguf_check(const char* gguf_path, char* json_out, int cap) {
    NB_TRY
    if (!gguf_path || !json_out || cap <= 0) throw Error("bad argument");
    GgufFile g;
    g.open(gguf_path);
    std::vector<std::string> vocab;
    const int vocab_size = (int)kv_u32(g, "nemo.vocab_size", VOCAB);
    const char* src = load_vocab(g, vocab_size, vocab);
    size_t bytes = 0;
    for (auto& kv : g.tensors) bytes += kv.second.nbytes;
    auto esc = [](const std::string& t) {
        std::string o;
        for (unsigned char ch : t) {
            if (ch == '"' || ch == '\\') { o += '\\'; o += (char)ch; }
            else if (ch < 0x20) { char b[8]; snprintf(b, sizeof b, "\\u%04x", ch); o += b; }
            else o += (char)ch;
        }
        return o;
    };
    std::string r = "{\"tensors\": " + std::to_string(g.tensors.size()) + ", \"tensor_bytes\": " + std::to_string(bytes) +
                    ", \"vocab_source\": \"" + src + "\", \"vocab_pieces\": " + std::to_string(vocab.size()) + ", \"n_layers\": " +
                    std::to_string(kv_u32(g, "nemo.n_layers", 24)) + ", \"first_pieces\": [";
    for (size_t i = 0; i < vocab.size() && i < 4; i++) r += std::string(i ? ", " : "") + "\"" + esc(vocab[i]) + "\"";
    r += "]}";
    if ((int)r.size() + 1 > cap) throw Error("gguf_check: buffer too small");
    memcpy(json_out, r.c_str(), r.size() + 1);
    return (int)r.size();
    NB_CATCH(NB200_ERR)
}

// Host-only round trip of the native weight file format (no GPU needed): writes a few blobs, reads them back through the mmap
// reader, and checks that a wrong source stamp, a wrong size request and a truncated file are all rejected.
int nb200_weight_cache_selftest(const char* path) {
    NB_TRY
    if (!path) throw Error("null path");
    const std::string p = path;
    std::vector<float> a(1000), b(77);
    std::vector<uint8_t> c(5);
    for (size_t i = 0; i < a.size(); i++) a[i] = (float)i * 0.5f;
    for (size_t i = 0; i < b.size(); i++) b[i] = -(float)i;
    for (size_t i = 0; i < c.size(); i++) c[i] = (uint8_t)(i + 1);
    {
        WeightCacheWriter w;
        w.open(p, 1234, 5678);
        w.add("first", a.data(), a.size() * sizeof(float));
        w.add("second#tag", b.data(), b.size() * sizeof(float));
        w.add("", c.data(), c.size());
        w.commit();
    }
    {
        WeightCacheReader r;
        if (!r.open(p, 1234, 5678) || r.entries() != 3) throw Error("selftest: cache did not open");
        const WeightCacheBlob* x = r.find("first", a.size() * sizeof(float));
        const WeightCacheBlob* y = r.find("second#tag", b.size() * sizeof(float));
        const WeightCacheBlob* z = r.find("", c.size());
        if (!x || !y || !z || memcmp(x->data, a.data(), x->bytes) || memcmp(y->data, b.data(), y->bytes) || memcmp(z->data, c.data(), z->bytes))
            throw Error("selftest: blobs differ");
        if (((uintptr_t)x->data & 255) || ((uintptr_t)y->data & 255)) throw Error("selftest: blobs are not 256-byte aligned");
        if (r.find("first", 4) || r.find("missing", 4)) throw Error("selftest: a wrong size / key was accepted");
    }
    {
        WeightCacheReader r;
        if (r.open(p, 1234, 5679) || r.open(p, 1235, 5678)) throw Error("selftest: a foreign source stamp was accepted");
    }
    {   // truncated at every interesting length: must be rejected, never read out of bounds
        FILE* f = fopen(p.c_str(), "rb");
        if (!f) throw Error("selftest: cannot reopen");
        std::vector<uint8_t> all;
        uint8_t buf[4096];
        size_t n;
        while ((n = fread(buf, 1, sizeof buf, f)) > 0) all.insert(all.end(), buf, buf + n);
        fclose(f);
        const std::string t = p + ".trunc";
        for (size_t len : {(size_t)0, (size_t)7, (size_t)39, (size_t)40, all.size() / 2, all.size() - 9, all.size() - 1}) {
            FILE* g = fopen(t.c_str(), "wb");
            if (!g) throw Error("selftest: cannot write");
            fwrite(all.data(), 1, len, g);
            fclose(g);
            WeightCacheReader r;
            if (r.open(t, 1234, 5678)) throw Error("selftest: a truncated cache (" + std::to_string(len) + " bytes) was accepted");
        }
        remove(t.c_str());
    }
    remove(p.c_str());
    return NB200_OK;
    NB_CATCH(NB200_ERR)
}

static nb200_model* model_load_impl(const char* gguf_path, const char* cache_path, int device, int max_streams, int flags) {
    nb200_model* m = nullptr;
    try {
        if (!gguf_path) throw Error("null path");
        if (max_streams <= 0) max_streams = 64;
        int n_dev = 0;
        cudaError_t e = cudaGetDeviceCount(&n_dev);
        if (e != cudaSuccess || n_dev == 0)
            throw Error(std::string("no CUDA device: this backend has no CPU fallback (") + cudaGetErrorString(e) + ")");
        if (device < 0 || device >= n_dev) throw Error("CUDA device ordinal out of range");
        CUDA_CHECK(cudaSetDevice(device));
        cudaDeviceProp prop;
        CUDA_CHECK(cudaGetDeviceProperties(&prop, device));
        if (prop.major != 10) throw Error(std::string("this library is built for sm_100a (B200); found ") + prop.name);
        m = new nb200_model();
        m->device = device;
        m->max_streams = max_streams;
        CUDA_CHECK(cudaStreamCreateWithFlags(&m->st, cudaStreamNonBlocking));
        CUDA_CHECK(cudaStreamCreateWithFlags(&m->st_copy, cudaStreamNonBlocking));
        CUDA_CHECK(cudaEventCreateWithFlags(&m->upload_ev, cudaEventDisableTiming));
        build_model(m, gguf_path, flags, cache_path);
        alloc_state(m);
        CUDA_CHECK(cudaStreamSynchronize(m->st));
        return m;
    } catch (const std::exception& e) {
        set_error(e.what());
        delete m;
        return nullptr;
    }
}

nb200_model* nb200_model_load(const char* gguf_path, int device, int max_streams, int flags) {
    return model_load_impl(gguf_path, nullptr, device, max_streams, flags);
}

nb200_model* nb200_model_load_cached(const char* gguf_path, const char* cache_path, int device, int max_streams, int flags) {
    return model_load_impl(gguf_path, cache_path ? cache_path : "", device, max_streams, flags);
}

int nb200_model_cache_stats(const nb200_model* m, int64_t* from_cache, int64_t* from_gguf) {
    if (!m) return NB200_ERR;
    if (from_cache) *from_cache = m->wc_hits;
    if (from_gguf) *from_gguf = m->wc_misses;
    return NB200_OK;
}

void nb200_model_free(nb200_model* m) {
    if (!m) return;
    cudaSetDevice(m->device);
    cudaDeviceSynchronize();
    delete m;
}

int nb200_model_info(const nb200_model* m, nb200_info* out) {
    if (!m || !out) return NB200_ERR;
    *out = m->info;
    return NB200_OK;
}

const char* nb200_model_token(const nb200_model* m, int32_t id) {
    if (!m || id < 0 || id >= (int)m->vocab.size()) return nullptr;
    return m->vocab[id].c_str();
}

int nb200_tokens_to_text(const nb200_model* m, const int32_t* tokens, int n, char* out, int cap) {
    if (!m || (!tokens && n > 0) || !out || cap <= 0) return NB200_ERR;
    std::string text;
    for (int i = 0; i < n; i++) {
        if (tokens[i] < 0 || tokens[i] >= (int)m->vocab.size()) continue;
        text += m->vocab[tokens[i]];
    }
    std::string res;  // U+2581 (E2 96 81) -> space, src/nemo-ggml.cpp:1556-1583
    for (size_t i = 0; i < text.size();) {
        if (i + 2 < text.size() && (unsigned char)text[i] == 0xE2 && (unsigned char)text[i + 1] == 0x96 && (unsigned char)text[i + 2] == 0x81) {
            res += ' ';
            i += 3;
        } else {
            res += text[i++];
        }
    }
    const int w = std::min((int)res.size(), cap - 1);
    memcpy(out, res.data(), w);
    out[w] = 0;
    return w;
}

int nb200_tokens_to_text_timed(const nb200_model* m, const int32_t* tokens, const int64_t* frames, int n, char* out, int cap) {
    if (!m || ((!tokens || !frames) && n > 0) || !out || cap <= 0) return NB200_ERR;
    std::string res;
    for (int i = 0; i < n; i++) {
        if (tokens[i] < 0 || tokens[i] >= (int)m->vocab.size()) continue;
        const std::string& piece = m->vocab[tokens[i]];
        if (piece.compare(0, 3, "\xe2\x96\x81") == 0) {  // word start: " {seconds}" + the rest of the piece
            const float sec = (float)frames[i] * 1280 / 16000;  // timed_token::to_seconds, src/nemo-ggml.h:393-395
            char buf[32];
            snprintf(buf, sizeof(buf), "{%.2f}", sec);
            res += ' ';
            res += buf;
            res.append(piece, 3, std::string::npos);
        } else {
            res += piece;
        }
    }
    const int w = std::min((int)res.size(), cap - 1);
    memcpy(out, res.data(), w);
    out[w] = 0;
    return w;
}

nb200_stream* nb200_stream_create(nb200_model* m, int right_context) {
    NB_TRY
    if (!m) throw Error("null model");
    if (right_context != 0 && right_context != 1 && right_context != 6 && right_context != 13)
        throw Error("right_context must be one of 0, 1, 6, 13 (src/nemo-stream.h:15-20)");
    if (m->free_slots.empty()) throw Error("no free stream slot (raise max_streams at model load)");
    CUDA_CHECK(cudaSetDevice(m->device));
    get_mode(m, right_context);
    nb200_stream* s = new nb200_stream();
    s->model = m;
    s->rc = right_context;
    s->prompt_index = m->default_prompt;
    s->slot = m->free_slots.back();
    m->free_slots.pop_back();
    reset_slot(m, s->slot);
    m->streams.push_back(s);
    return s;
    NB_CATCH(nullptr)
}

void nb200_stream_free(nb200_stream* s) {
    if (!s) return;
    nb200_model* m = s->model;
    try {  // the step in flight may reference this stream; audio staged for its slot must land before the slot is reused
        cudaSetDevice(m->device);
        finish_inflight(m);
        flush_uploads(m);
    } catch (const std::exception& e) {
        set_error(e.what());
    }
    m->free_slots.push_back(s->slot);
    m->streams.erase(std::remove(m->streams.begin(), m->streams.end(), s), m->streams.end());
    delete s;
}

int nb200_stream_reset(nb200_stream* s) {
    NB_TRY
    if (!s) throw Error("null stream");
    CUDA_CHECK(cudaSetDevice(s->model->device));
    finish_inflight(s->model);
    flush_uploads(s->model);  // audio staged before the reset lands (stream-ordered) before anything fed after it
    reset_slot(s->model, s->slot);
    s->pending.clear();
    s->fed = s->samples_total = s->chunks_done = 0;
    s->finalize_requested = s->finalized = false;
    s->wpos = s->valid = 0;
    s->tokens.clear();
    s->token_frames.clear();
    s->frames_done = 0;
    s->poll_cursor = 0;
    s->chunk_tokens.clear();
    s->chunk_margins.clear();
    s->tap_mel.clear();
    s->tap_enc.clear();
    return NB200_OK;
    NB_CATCH(NB200_ERR)
}

int nb200_stream_feed(nb200_stream* s, const int16_t* pcm, int n_samples) {
    NB_TRY
    if (!s) throw Error("null stream");
    if (n_samples <= 0 || !pcm) return NB200_OK;  // like the reference: nothing to do (src/nemo-stream.cpp:1150)
    if (s->finalized || s->finalize_requested) throw Error("stream already finalized; reset it first");
    CUDA_CHECK(cudaSetDevice(s->model->device));
    flush_pending(s->model, s);
    size_t took = 0;
    if (s->pending.empty()) took = upload_pcm(s->model, s, pcm, (size_t)n_samples);
    s->pending.insert(s->pending.end(), pcm + took, pcm + n_samples);
    s->samples_total += n_samples;
    return NB200_OK;
    NB_CATCH(NB200_ERR)
}

static int full_chunks_ready(const nb200_stream* s) {  // counts device-resident audio only
    const Mode md(s->rc);
    const long long frames = frames_available(s->fed);
    const long long have = frames - (long long)md.shift * s->chunks_done;
    return have < md.shift ? 0 : (int)(have / md.shift);
}

static int tail_frames(const nb200_stream* s) {  // new mel frames left for a finalize tail (src/nemo-stream.cpp:1240-1244)
    const Mode md(s->rc);
    return (int)(frames_available(s->fed) - (long long)md.shift * s->chunks_done);
}

int nb200_stream_finalize(nb200_stream* s) {
    if (!s) return NB200_ERR;
    if (!s->finalized) s->finalize_requested = true;
    return NB200_OK;
}

int nb200_stream_ready_chunks(const nb200_stream* s) {
    if (!s || s->finalized) return 0;
    int n = full_chunks_ready(s);
    if (n == 0 && !s->pending.empty()) return 1;  // the next step uploads the rest of the fed audio
    if (s->finalize_requested) {
        const Mode md(s->rc);
        const int rem = tail_frames(s) - n * md.shift;
        if (rem / 8 > 0) n++;
    }
    return n;
}

// Completes the step nb200_step_begin left running (no-op when there is none). Returns its stream-chunk count.
static int finish_inflight(nb200_model* m) {
    if (!m->inflight.active) return 0;
    m->inflight.active = false;  // cleared first: a failure below must not leave a half-finished step behind
    std::vector<std::pair<ModeCtx*, std::vector<BatchItem>>> groups = std::move(m->inflight.groups);
    m->inflight.groups.clear();
    double gpu_ms = 0.0, enc_ms = 0.0, dec_ms = 0.0;
    for (auto& g : groups) {
        run_batch_finish(m, g.first, g.second);
        gpu_ms += m->stats.last_step_gpu_ms;
        enc_ms += m->stats.last_encoder_ms;
        dec_ms += m->stats.last_decoder_ms;
    }
    if (groups.size() > 1) {  // the "last step" figures cover all of the step's groups
        m->stats.last_step_gpu_ms = gpu_ms;
        m->stats.last_encoder_ms = enc_ms;
        m->stats.last_decoder_ms = dec_ms;
        m->stats.steps -= (long long)groups.size() - 1;
    }
    return m->inflight.total;
}

int nb200_step_begin(nb200_model* m) {
    NB_TRY
    if (!m) throw Error("null model");
    if (m->inflight.active) throw Error("nb200_step_begin: the previous step has not been completed by nb200_step_end");
    CUDA_CHECK(cudaSetDevice(m->device));
    std::map<int, std::vector<BatchItem>> groups;
    for (nb200_stream* s : m->streams) {
        if (s->finalized) continue;
        flush_pending(m, s);
        const Mode md(s->rc);
        if (full_chunks_ready(s) > 0) {
            groups[s->rc].push_back({s, md.shift, md.T});
        } else if (s->finalize_requested && s->pending.empty()) {
            const int rem = tail_frames(s);
            const int n_valid = rem / 8;  // floor(real_new / subsampling_factor), src/nemo-stream.cpp:1243
            if (n_valid > 0) groups[s->rc].push_back({s, rem, n_valid});
            s->finalized = true;
            s->finalize_requested = false;
        }
    }
    flush_uploads(m);  // whatever was fed since the last flush: one copy + one scatter kernel on the copy stream
    apply_resets(m);   // state of the streams created / reset since the last step (one kernel + one decode refresh, no host sync)
    if (m->upload_wait) {
        CUDA_CHECK(cudaStreamWaitEvent(m->st, m->upload_ev, 0));
        m->upload_wait = false;
    }
    // One batched chunk per latency mode, all enqueued before any is waited for: the modes share the device workspace, which the
    // stream orders; metadata and results have their own pinned buffers per mode (ModeCtx).
    int total = 0;
    m->inflight.groups.clear();
    for (auto& kv : groups) {
        ModeCtx* mc = get_mode(m, kv.first);
        total += (int)kv.second.size();
        run_batch_launch(m, mc, kv.second);
        m->inflight.active = true;
        m->inflight.groups.emplace_back(mc, std::move(kv.second));
    }
    m->inflight.total = total;
    return total;
    NB_CATCH(NB200_ERR)
}

int nb200_step_end(nb200_model* m) {
    NB_TRY
    if (!m) throw Error("null model");
    CUDA_CHECK(cudaSetDevice(m->device));
    return finish_inflight(m);
    NB_CATCH(NB200_ERR)
}

int nb200_step(nb200_model* m) {
    const int n = nb200_step_begin(m);
    if (n <= 0) return n;
    return nb200_step_end(m);
}

int nb200_feed_batch(nb200_model* m, nb200_stream* const* streams, const int16_t* const* pcm, const int* n_samples, int count) {
    if (!m || !streams || !pcm || !n_samples) return NB200_ERR;
    for (int i = 0; i < count; i++) {
        if (!streams[i] || streams[i]->model != m) {
            set_error("nb200_feed_batch: stream " + std::to_string(i) + " does not belong to this model");
            return NB200_ERR;
        }
        const int r = nb200_stream_feed(streams[i], pcm[i], n_samples[i]);
        if (r < 0) return r;
    }
    try {  // start the transfer now: it overlaps the step that may be running
        CUDA_CHECK(cudaSetDevice(m->device));
        flush_uploads(m);
    } catch (const std::exception& e) {
        set_error(e.what());
        return NB200_ERR;
    }
    return NB200_OK;
}

int nb200_poll_batch(nb200_model* m, nb200_stream* const* streams, int count, int32_t* tokens_out, int cap, int32_t* counts_out) {
    if (!m || !streams || !counts_out || (!tokens_out && cap > 0)) return NB200_ERR;
    long long need = 0;
    for (int i = 0; i < count; i++) {
        counts_out[i] = streams[i] ? (int)(streams[i]->tokens.size() - streams[i]->poll_cursor) : 0;
        need += counts_out[i];
    }
    if (need > cap) {  // nothing is consumed: counts_out holds what is waiting, the caller retries with a larger buffer
        set_error("nb200_poll_batch: " + std::to_string(need) + " tokens waiting, buffer holds " + std::to_string(cap));
        return NB200_ERR;
    }
    int total = 0;
    for (int i = 0; i < count; i++) {
        if (counts_out[i] == 0) continue;
        total += nb200_stream_poll_tokens(streams[i], tokens_out + total, counts_out[i]);
    }
    return total;
}

int nb200_drain(nb200_model* m) {
    int total = 0;
    for (;;) {
        const int n = nb200_step(m);
        if (n < 0) return n;
        if (n == 0) return total;
        total += n;
    }
}

int nb200_stream_poll_tokens(nb200_stream* s, int32_t* out, int cap) {
    if (!s || (!out && cap > 0)) return NB200_ERR;
    const int n = std::min<int>(cap, (int)(s->tokens.size() - s->poll_cursor));
    for (int i = 0; i < n; i++) out[i] = s->tokens[s->poll_cursor + i];
    s->poll_cursor += n;
    return n;
}

int nb200_stream_poll_timed(nb200_stream* s, int32_t* tokens_out, int64_t* frames_out, int cap) {
    if (!s || (!tokens_out && cap > 0)) return NB200_ERR;
    const int n = std::min<int>(cap, (int)(s->tokens.size() - s->poll_cursor));
    for (int i = 0; i < n; i++) {
        tokens_out[i] = s->tokens[s->poll_cursor + i];
        if (frames_out) frames_out[i] = s->token_frames[s->poll_cursor + i];
    }
    s->poll_cursor += n;
    return n;
}

int nb200_stream_get_token_frames(const nb200_stream* s, int64_t* out, int cap) {
    if (!s || (!out && cap > 0)) return NB200_ERR;
    const int n = std::min<int>(cap, (int)s->token_frames.size());
    for (int i = 0; i < n; i++) out[i] = s->token_frames[i];
    return (int)s->token_frames.size();
}

int nb200_stream_get_tokens(const nb200_stream* s, int32_t* out, int cap) {
    if (!s) return NB200_ERR;
    const int n = std::min<int>(cap, (int)s->tokens.size());
    for (int i = 0; i < n; i++) out[i] = s->tokens[i];
    return (int)s->tokens.size();
}

int nb200_stream_chunk_token_counts(const nb200_stream* s, int32_t* out, int cap) {
    if (!s) return NB200_ERR;
    const int n = std::min<int>(cap, (int)s->chunk_tokens.size());
    for (int i = 0; i < n; i++) out[i] = s->chunk_tokens[i];
    return (int)s->chunk_tokens.size();
}

int nb200_stream_chunk_margins(const nb200_stream* s, float* out, int cap) {
    if (!s) return NB200_ERR;
    const int n = std::min<int>(cap, (int)s->chunk_margins.size());
    for (int i = 0; i < n; i++) out[i] = s->chunk_margins[i];
    return (int)s->chunk_margins.size();
}

int nb200_model_set_language(nb200_model* m, const char* lang) {
    NB_TRY
    if (!m || !lang) throw Error("null argument");
    if (m->num_prompts <= 0) throw Error("model is not multilingual (num_prompts=0)");
    auto it = m->prompt_dict.find(lang);
    if (it == m->prompt_dict.end()) throw Error(std::string("unknown language code '") + lang + "'");
    m->default_prompt = it->second;
    return NB200_OK;
    NB_CATCH(NB200_ERR)
}

int nb200_stream_set_language(nb200_stream* s, const char* lang) {
    NB_TRY
    if (!s || !lang) throw Error("null argument");
    nb200_model* m = s->model;
    if (m->num_prompts <= 0) throw Error("model is not multilingual (num_prompts=0)");
    auto it = m->prompt_dict.find(lang);
    if (it == m->prompt_dict.end()) throw Error(std::string("unknown language code '") + lang + "'");
    s->prompt_index = it->second;
    return NB200_OK;
    NB_CATCH(NB200_ERR)
}

int nb200_stream_set_prompt_index(nb200_stream* s, int prompt_index) {
    NB_TRY
    if (!s) throw Error("null stream");
    if (s->model->num_prompts <= 0) throw Error("model is not multilingual (num_prompts=0)");
    s->prompt_index = prompt_index;  // out-of-range ids fall back to 0 when the chunk runs, like src/nemo-stream.cpp:1052-1053
    return NB200_OK;
    NB_CATCH(NB200_ERR)
}

int nb200_stream_set_taps(nb200_stream* s, int taps) {
    if (!s) return NB200_ERR;
    s->taps = taps;
    return NB200_OK;
}

int64_t nb200_stream_fetch_tap(const nb200_stream* s, int tap, float* out, int64_t cap) {
    if (!s) return NB200_ERR;
    const std::vector<float>& v = (tap == NB200_TAP_MEL) ? s->tap_mel : s->tap_enc;
    const int64_t n = std::min<int64_t>(cap, (int64_t)v.size());
    if (out && n > 0) memcpy(out, v.data(), (size_t)n * sizeof(float));
    return (int64_t)v.size();
}

int nb200_get_stats(const nb200_model* m, nb200_stats* out) {
    if (!m || !out) return NB200_ERR;
    *out = m->stats;
    out->kernel_launches = m->launches;
    return NB200_OK;
}

int nb200_mel_batch(nb200_model* m, const int16_t* pcm, int n_streams, int n_samples, float* mel_out) {
    NB_TRY
    if (!m || !pcm || !mel_out) throw Error("null argument");
    CUDA_CHECK(cudaSetDevice(m->device));
    const int nf = (int)frames_available(n_samples);
    if (nf <= 0 || n_streams <= 0) return 0;
    int cap = 1;
    while (cap < n_samples) cap <<= 1;  // a private "ring" that holds each whole signal
    std::vector<int> slot(n_streams), nfr(n_streams, nf);
    std::vector<long long> m64((size_t)3 * n_streams);
    for (int b = 0; b < n_streams; b++) {
        slot[b] = b;
        m64[0 * n_streams + b] = (long long)b * nf * N_MELS;  // out_off
        m64[1 * n_streams + b] = -257;                        // a0: 256 zeros of centre padding + the pre-emphasis sample
        m64[2 * n_streams + b] = n_samples;                   // n_avail
    }
    DevBufs bufs;
    LaunchCount lc(m);
    int16_t* d_pcm = bufs.get<int16_t>((size_t)n_streams * cap);
    int* d_i = bufs.get<int>((size_t)2 * n_streams);
    long long* d_l = bufs.get<long long>(m64.size());
    float* d_mel = bufs.get<float>((size_t)n_streams * nf * N_MELS);
    CUDA_CHECK(cudaMemcpy2DAsync(d_pcm, (size_t)cap * sizeof(int16_t), pcm, (size_t)n_samples * sizeof(int16_t),
                                 (size_t)n_samples * sizeof(int16_t), n_streams, cudaMemcpyHostToDevice, m->st));
    CUDA_CHECK(cudaMemcpyAsync(d_i, slot.data(), n_streams * sizeof(int), cudaMemcpyHostToDevice, m->st));
    CUDA_CHECK(cudaMemcpyAsync(d_i + n_streams, nfr.data(), n_streams * sizeof(int), cudaMemcpyHostToDevice, m->st));
    CUDA_CHECK(cudaMemcpyAsync(d_l, m64.data(), m64.size() * sizeof(long long), cudaMemcpyHostToDevice, m->st));
    MelBatch mb{d_pcm, cap, d_i, d_i + n_streams, d_l + n_streams, d_l + 2 * n_streams, d_l};
    launch_mel(m->st, m->mel_tb, mb, d_mel, n_streams, nf);
    CUDA_CHECK(cudaMemcpyAsync(mel_out, d_mel, (size_t)n_streams * nf * N_MELS * sizeof(float), cudaMemcpyDeviceToHost, m->st));
    CUDA_CHECK(cudaStreamSynchronize(m->st));
    return nf;
    NB_CATCH(NB200_ERR)
}

// ---- offline path: nemo_encode_audio (src/nemo-ggml.cpp:1690-1715 -> nemo_encode :1592-1670) ---------------------------
// Whole-utterance log-mel -> dw-striding subsampling with its causal zero padding, nothing dropped -> 24 layers with
// full-sequence relative-position attention (no cache, no mask) and the causal depthwise conv over zero left padding ->
// (prompt fusion) -> greedy decode from a zero LSTM state with timed tokens. Same kernels as the streaming step (one
// "chunk" of T frames on a scratch slot with zeroed caches) except for the attention. Eager launches, synchronous.
int nb200_encode_audio(nb200_model* m, const int16_t* pcm, int n_samples, int32_t* tokens_out, int64_t* frames_out, int cap,
                       float* enc_out, int64_t enc_cap_floats, int* n_frames_out) {
    NB_TRY
    if (!m || (!pcm && n_samples > 0) || (!tokens_out && cap > 0)) throw Error("null argument");
    CUDA_CHECK(cudaSetDevice(m->device));
    if (finish_inflight(m) < 0) return NB200_ERR;
    if (n_frames_out) *n_frames_out = 0;
    const int nf = (int)frames_available(n_samples);
    if (nf <= 0) return 0;  // nemo_encode_audio returns no tokens for an utterance without a single frame
    const int t1 = nf / 2 + 1, t2 = t1 / 2 + 1, T = t2 / 2 + 1;  // causal stride-2 convs: pad 2 before, 1 after (src/nemo-ggml.cpp:897-952)
    const int capB = m->cap_B;
    const Mode big(13);
    if (T > ATTN_FULL_MAX_T) throw Error("offline path: utterance longer than 512 encoder frames (the reference's positional table, src/nemo-ggml.cpp:1061)");
    if (nf > capB * big.chunk_mel || t2 > capB * big.t2 || T > capB * MAX_T)
        throw Error("offline path: utterance too long for this model's workspaces (" + std::to_string(T) + " frames; max_streams * 14 = " +
                    std::to_string(capB * MAX_T) + "): raise max_streams at model load");
    if (m->free_slots.empty()) throw Error("offline path needs one free stream slot");
    const int slot = m->free_slots.back();  // scratch: zero conv caches and decoder state, left free afterwards
    cudaStream_t st = m->st;
    LaunchCount lc(m);
    flush_uploads(m);
    reset_slot(m, slot);
    apply_resets(m);
    const int aq = m->act_quant ? 1 : 0;

    // positional table for this length: rows p + (T - 1), p = i - j in [-(T-1), T-1] (ascending p; the reference stores the
    // same rows in descending order and its rel-shift reads row j - i + T - 1, src/nemo-ggml.cpp:17-32,625-662)
    if (!m->off_pos_tab) {
        m->off_max_T = std::min(ATTN_FULL_MAX_T, capB * MAX_T);
        const size_t rows = (size_t)2 * m->off_max_T;
        m->off_posf = m->alloc<float>(rows * D_MODEL);
        m->off_pos16 = m->alloc<__half>(rows * D_MODEL);
        m->off_pos_tab = m->alloc<__half>((size_t)m->n_layers * rows * D_MODEL);
        operand_init(m->op_off_pos16, m->off_pos16, (int)rows, D_MODEL);
    }
    const int npos = 2 * T - 1;
    if (m->off_pos_T != T) {
        launch_pos_emb(st, m->off_posf, -(T - 1), npos);
        launch_cast_f16(st, m->off_posf, m->off_pos16, aq, npos, D_MODEL, m->bf16 ? 1 : 0);
        for (int l = 0; l < m->n_layers; l++) {
            EpiParams ep;
            ep.kind = EPI_STORE_F16;
            ep.out_f16 = m->off_pos_tab + (size_t)l * npos * D_MODEL;
            ep.ldo = D_MODEL;
            ep.ab_bf16 = m->bf16 ? 1 : 0;
            gemm(st, m->op_off_pos16, m->layers[l].wpos.op, npos, ep, m->simt);
        }
        m->off_pos_T = T;
    }

    // audio + metadata of the one "stream" (b = 0)
    const int ring = 1 << 30;  // the mel kernel indexes a ring: make it larger than any utterance
    // the utterance's PCM lives in a persistent device buffer (no cudaMalloc per call; a stable pointer for the per-length graphs below)
    const size_t pcm_cap = ((size_t)ATTN_FULL_MAX_T * 8 + 32) * HOP + 1024;
    if ((size_t)n_samples + 8 > pcm_cap) throw Error("offline path: utterance longer than 512 encoder frames");
    if (!m->off_pcm) m->off_pcm = m->alloc<int16_t>(pcm_cap, false);
    int16_t* d_pcm = m->off_pcm;
    std::vector<int> meta((size_t)META_ROWS * capB, 0);
    std::vector<long long> meta64((size_t)3 * capB, 0);
    meta[0 * capB] = slot;      // slot (scratch: zeroed caches)
    meta[3 * capB] = T;         // frames to decode
    meta[4 * capB] = nf;
    meta[6 * capB] = (m->default_prompt >= 0 && m->default_prompt < m->num_prompts) ? m->default_prompt : 0;
    meta64[0 * capB] = 0;       // mel rows start at ws_mel
    meta64[1 * capB] = -257;    // 256 zeros of centre padding + the pre-emphasis sample (src/preprocessor.cpp:300-328)
    meta64[2 * capB] = n_samples;
    int zero = 0;
    CUDA_CHECK(cudaMemcpyAsync(d_pcm, pcm, (size_t)n_samples * sizeof(int16_t), cudaMemcpyHostToDevice, st));
    CUDA_CHECK(cudaMemcpyAsync(m->ws_meta, meta.data(), meta.size() * sizeof(int), cudaMemcpyHostToDevice, st));
    CUDA_CHECK(cudaMemcpyAsync(m->ws_meta64, meta64.data(), meta64.size() * sizeof(long long), cudaMemcpyHostToDevice, st));
    int* d_zero_slot = m->ws_meta + 7 * capB;  // spare meta row: the mel kernel's ring slot index (0: d_pcm is the whole "ring")
    CUDA_CHECK(cudaMemcpyAsync(d_zero_slot, &zero, sizeof(int), cudaMemcpyHostToDevice, st));
    // Everything from the log-mel to the decode kernel depends on the utterance only through its frame count (the audio, its length
    // and the scratch slot travel in device buffers): one CUDA graph per mel-frame count, captured on the second utterance of that
    // length and replayed afterwards (~470 launches; eager launches cost ~2 ms of host time per call). A small cache: cleared when full.
    const float* enc = nullptr;
    if (m->dec_graphs.size() > 24) {
        for (auto& kv : m->dec_graphs)
            if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
        m->dec_graphs.clear();
    }
    const int win_off = [&] {
        static const int win_env = getenv("NB200_OFFLINE_WINDOW") ? atoi(getenv("NB200_OFFLINE_WINDOW")) : 0;
        return std::max(1, std::min(win_env > 0 ? win_env : 32, std::min(capB, 64)));
    }();
    run_graphed(m, m->dec_graphs, (long long)nf, m->use_graphs && !g_prof_on, [&] {
    {
        MelBatch mb{d_pcm, ring, d_zero_slot, m->ws_meta + 4 * capB, m->ws_meta64 + 1 * capB, m->ws_meta64 + 2 * capB, m->ws_meta64 + 0 * capB};
        launch_mel(st, m->mel_tb, mb, m->ws_mel, 1, nf);
    }
    launch_sub_stage1(st, m->sub, m->ws_mel, m->ws_a2h, 1, nf, t1, t2);
    {
        EpiParams ep;
        ep.kind = EPI_STORE; ep.act = ACT_RELU; ep.bias = m->sub_b3; ep.out_f32 = m->ws_a3; ep.ldo = SUB_CH;
        gemm(st, m->op_a2h, m->sub_w3.op, t2 * SUB_F2, ep, m->simt);
        launch_sub_stage2(st, m->sub, m->ws_a3, m->ws_a5h, 1, t2, T, 0);
        EpiParams e6;
        e6.kind = EPI_STORE_F16; e6.act = ACT_RELU; e6.bias = m->sub_b6; e6.out_f16 = m->ws_a6h; e6.ldo = SUB_CH; e6.act_quant = 0;
        gemm(st, m->op_a5h, m->sub_w6.op, T * SUB_F3, e6, m->simt);
        EpiParams eo;
        eo.kind = EPI_STORE; eo.bias = m->sub_bout; eo.out_f32 = m->ws_x; eo.ldo = D_MODEL;
        gemm(st, m->op_a6h, m->sub_wout.op, T, eo, m->simt);
    }
    enc = enqueue_layers(m, 1, T, [&](int l) {
        LayerW& L = m->layers[l];
        AttnFullArgs at{};
        at.qkv = m->ws_qkv16; at.pos = m->off_pos_tab + (size_t)l * npos * D_MODEL;
        at.bias_u = L.pos_u; at.bias_v = L.pos_v; at.out = m->ws_a16; at.T = T; at.act_quant = aq; at.out_bf16 = m->bf16 ? 1 : 0;
        launch_attention_full(st, at);
    });
    enqueue_joint_enc(m, enc, T);
    // greedy_decode (src/nemo-ggml.cpp:1233-1380) from the fresh state; one utterance is the whole batch, so a round evaluates a
    // window of consecutive frames (decoder.cu, window mode): blank runs cost one round per window instead of one per frame
    launch_decode_chunk(st, m->dec, m->dstate, decode_batch(m, 1, T, win_off));
    });
    if (enc == nullptr) enc = m->num_prompts > 0 ? m->ws_g : m->ws_x;  // a replayed graph: enqueue_layers' (fixed) output buffer
    if (enc_out) {
        if (enc_cap_floats < (int64_t)T * D_MODEL) throw Error("encoder output buffer too small");
        CUDA_CHECK(cudaMemcpyAsync(enc_out, enc, (size_t)T * D_MODEL * sizeof(float), cudaMemcpyDeviceToHost, st));
    }
    int counters[8], n_tok = 0;
    CUDA_CHECK(cudaMemcpyAsync(counters, m->ws_unfinished + 6, sizeof(counters), cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaMemcpyAsync(&n_tok, m->ws_tokcnt, sizeof(int), cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    if (counters[0] != 0) throw Error("decode loop did not terminate");
    std::vector<int> words((size_t)std::max(n_tok, 1));
    CUDA_CHECK(cudaMemcpy(words.data(), m->ws_tok, (size_t)n_tok * sizeof(int), cudaMemcpyDeviceToHost));
    for (int i = 0; i < std::min(n_tok, cap); i++) {
        tokens_out[i] = words[i] & 0xffff;
        if (frames_out) frames_out[i] = words[i] >> 16;
    }
    if (n_frames_out) *n_frames_out = T;
    m->stats.audio_seconds += (double)n_samples / 16000.0;
    m->stats.frames += T;
    m->stats.tokens += n_tok;
    return n_tok;
    NB_CATCH(NB200_ERR)
}

int nb200_timer_start(nb200_model* m) {
    NB_TRY
    if (!m) throw Error("null model");
    CUDA_CHECK(cudaSetDevice(m->device));
    CUDA_CHECK(cudaEventRecord(m->tev[0], m->st));
    return NB200_OK;
    NB_CATCH(NB200_ERR)
}

int nb200_timer_stop(nb200_model* m, double* ms_out) {
    NB_TRY
    if (!m || !ms_out) throw Error("null argument");
    CUDA_CHECK(cudaSetDevice(m->device));
    CUDA_CHECK(cudaEventRecord(m->tev[1], m->st));
    CUDA_CHECK(cudaEventSynchronize(m->tev[1]));
    float ms = 0;
    CUDA_CHECK(cudaEventElapsedTime(&ms, m->tev[0], m->tev[1]));
    *ms_out = ms;
    return NB200_OK;
    NB_CATCH(NB200_ERR)
}

int nb200_profile_enable(nb200_model* m, int on) {
    NB_TRY
    if (!m) throw Error("null model");
    CUDA_CHECK(cudaSetDevice(m->device));
    CUDA_CHECK(cudaStreamSynchronize(m->st));
    if (on) prof_report();  // drop stale records
    g_prof_on = on != 0;
    return NB200_OK;
    NB_CATCH(NB200_ERR)
}

int nb200_set_dup_family(nb200_model* m, const char* family) {
    NB_TRY
    if (!m) throw Error("null model");
    CUDA_CHECK(cudaSetDevice(m->device));
    if (finish_inflight(m) < 0) return NB200_ERR;
    CUDA_CHECK(cudaStreamSynchronize(m->st));
    set_dup_family(family);
    for (auto* gm : {&m->enc_graphs, &m->dec_graphs}) {  // the captured graphs hold the old launch sequence
        for (auto& kv : *gm)
            if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
        gm->clear();
    }
    return NB200_OK;
    NB_CATCH(NB200_ERR)
}

int nb200_profile_report(nb200_model* m, char* json_out, int cap) {
    NB_TRY
    if (!m || !json_out || cap <= 0) throw Error("bad argument");
    CUDA_CHECK(cudaSetDevice(m->device));
    CUDA_CHECK(cudaStreamSynchronize(m->st));
    const std::string r = prof_report();
    if ((int)r.size() + 1 > cap) throw Error("profile report buffer too small");
    memcpy(json_out, r.c_str(), r.size() + 1);
    return (int)r.size();
    NB_CATCH(NB200_ERR)
}

namespace {
// Synthetic operands for the GEMM microbenchmarks: one activation matrix and enough weight matrices to exceed 2x L2.
struct GemmBench {
    int M, N, K, Mp, n_w;
    __half *a = nullptr, *w = nullptr, *o16 = nullptr;
    float* o32 = nullptr;
    Operand A;
    std::vector<Operand> W;
    EpiParams ep;
    GemmBench(int M_, int N_, int K_, int epi_kind) : M(M_), N(N_), K(K_) {
        Mp = (M + 127) / 128 * 128;
        n_w = std::max(2, (int)(((size_t)300 << 20) / ((size_t)N * K * 2)));
        a = dalloc<__half>((size_t)Mp * K);
        w = dalloc<__half>((size_t)n_w * N * K);
        o32 = dalloc<float>((size_t)8 * Mp * N);
        o16 = dalloc<__half>((size_t)Mp * N);
        std::vector<__half> h((size_t)Mp * K);
        for (size_t i = 0; i < h.size(); i++) h[i] = __float2half_rn((float)((int)((i * 2654435761u) >> 20 & 255) - 128) / 64.0f);
        CUDA_CHECK(cudaMemcpy(a, h.data(), h.size() * 2, cudaMemcpyHostToDevice));
        std::vector<__half> hw((size_t)N * K);
        for (size_t i = 0; i < hw.size(); i++) hw[i] = __float2half_rn((float)((int)((i * 40503u) >> 9 & 255) - 128) / 2048.0f);
        for (int j = 0; j < n_w; j++) CUDA_CHECK(cudaMemcpy(w + (size_t)j * N * K, hw.data(), hw.size() * 2, cudaMemcpyHostToDevice));
        operand_init(A, a, Mp, K);
        W.resize(n_w);
        for (int j = 0; j < n_w; j++) operand_init(W[j], w + (size_t)j * N * K, N, K);
        ep.kind = epi_kind == 6 ? (int)EPI_STORE_F16 : epi_kind;  // 6 = fp16 store without activation (q|k|v)
        ep.ldo = N;
        ep.out_f32 = o32;
        ep.out_f16 = o16;
        ep.split_stride = (size_t)Mp * N;
        if (epi_kind == EPI_STORE_F16) { ep.act = ACT_SILU; ep.act_quant = 1; }
        if (epi_kind == EPI_GLU) ep.ldo = N / 2;
    }
    ~GemmBench() { cudaFree(a); cudaFree(w); cudaFree(o32); cudaFree(o16); }
};
}  // namespace

int nb200_bench_gemm(nb200_model* m, int M, int N, int K, int epi_kind, int bn, int splits, int iters, double* us_per_launch,
                     int* bn_used, int* splits_used) {
    NB_TRY
    if (!m || M <= 0 || N % 64 || K % 64 || iters <= 0 || !us_per_launch) throw Error("bad argument");
    CUDA_CHECK(cudaSetDevice(m->device));
    GemmBench gb(M, N, K, epi_kind);
    gemm_force_config(bn, splits);
    const GemmConfig c = gemm_pick_config(M, N, K, epi_kind == EPI_PARTIAL ? 8 : 1);
    if (bn_used) *bn_used = c.bn;
    if (splits_used) *splits_used = c.splits;
    for (int i = 0; i < 3; i++) gemm(m->st, gb.A, gb.W[i % gb.n_w], M, gb.ep, false, 8);
    CUDA_CHECK(cudaEventRecord(m->tev[0], m->st));
    for (int i = 0; i < iters; i++) gemm(m->st, gb.A, gb.W[i % gb.n_w], M, gb.ep, false, 8);
    CUDA_CHECK(cudaEventRecord(m->tev[1], m->st));
    CUDA_CHECK(cudaEventSynchronize(m->tev[1]));
    gemm_force_config(0, 1);
    float ms = 0;
    CUDA_CHECK(cudaEventElapsedTime(&ms, m->tev[0], m->tev[1]));
    *us_per_launch = ms * 1e3 / iters;
    return NB200_OK;
    NB_CATCH(NB200_ERR)
}

int nb200_trace_gemm(nb200_model* m, int M, int N, int K, int epi_kind, int bn, int splits, int launches, uint64_t* out, int cap_words) {
    NB_TRY
    if (!m || M <= 0 || N % 64 || K % 64 || launches <= 0 || !out) throw Error("bad argument");
    CUDA_CHECK(cudaSetDevice(m->device));
    GemmBench gb(M, N, K, epi_kind);
    gemm_force_config(bn, splits);
    const GemmConfig c = gemm_pick_config(M, N, K, epi_kind == EPI_PARTIAL ? 8 : 1);
    const size_t words = (size_t)launches * c.ctas * 8;
    if ((size_t)cap_words < words) { gemm_force_config(0, 1); throw Error("trace buffer too small"); }
    unsigned long long* d = dalloc<unsigned long long>(words);
    CUDA_CHECK(cudaMemset(d, 0, words * 8));
    for (int i = 0; i < 3; i++) gemm(m->st, gb.A, gb.W[i % gb.n_w], M, gb.ep, false, 8);
    for (int i = 0; i < launches; i++) {
        EpiParams ep = gb.ep;
        ep.trace = d + (size_t)i * c.ctas * 8;
        gemm(m->st, gb.A, gb.W[i % gb.n_w], M, ep, false, 8);
    }
    CUDA_CHECK(cudaStreamSynchronize(m->st));
    gemm_force_config(0, 1);
    CUDA_CHECK(cudaMemcpy(out, d, words * 8, cudaMemcpyDeviceToHost));
    cudaFree(d);
    return c.ctas;
    NB_CATCH(NB200_ERR)
}

int nb200_debug_matmul(nb200_model* m, const char* tensor_name, const float* x, int M, float* y) {
    NB_TRY
    if (!m || !tensor_name || !x || !y || M <= 0) throw Error("bad argument");
    CUDA_CHECK(cudaSetDevice(m->device));
    auto it = m->by_name.find(tensor_name);
    if (it == m->by_name.end()) throw Error(std::string("no such encoder matrix: ") + tensor_name);
    Operand W;
    operand_init(W, it->second.ptr, it->second.rows, it->second.K);
    const int K = W.K, N = W.rows;
    if (K % 1024 != 0) throw Error("debug_matmul supports K multiple of 1024");
    const int Mp = (M + 127) / 128 * 128;
    DevBufs bufs;
    float* dx = bufs.get<float>((size_t)M * K);
    __half* dx16 = bufs.get<__half>((size_t)Mp * K);
    float* dy = bufs.get<float>((size_t)M * N);
    CUDA_CHECK(cudaMemsetAsync(dx16, 0, (size_t)Mp * K * sizeof(__half), m->st));
    CUDA_CHECK(cudaMemcpyAsync(dx, x, (size_t)M * K * sizeof(float), cudaMemcpyHostToDevice, m->st));
    launch_cast_f16(m->st, dx, dx16, m->act_quant ? 1 : 0, M, K, m->bf16 ? 1 : 0);
    Operand A;
    operand_init(A, dx16, Mp, K);
    EpiParams ep;
    ep.kind = EPI_STORE;
    ep.out_f32 = dy;
    ep.ldo = N;
    ep.ab_bf16 = m->bf16 ? 1 : 0;
    gemm(m->st, A, W, M, ep, m->simt);
    if (it->second.glu_interleaved) {  // undo the load-time row interleave so y is in the tensor's own row order
        std::vector<float> tmp((size_t)M * N);
        CUDA_CHECK(cudaMemcpyAsync(tmp.data(), dy, tmp.size() * sizeof(float), cudaMemcpyDeviceToHost, m->st));
        CUDA_CHECK(cudaStreamSynchronize(m->st));
        for (int r = 0; r < M; r++)
            for (int j = 0; j < N / 2; j++) {
                y[(size_t)r * N + j] = tmp[(size_t)r * N + 2 * j];
                y[(size_t)r * N + N / 2 + j] = tmp[(size_t)r * N + 2 * j + 1];
            }
    } else {
        CUDA_CHECK(cudaMemcpyAsync(y, dy, (size_t)M * N * sizeof(float), cudaMemcpyDeviceToHost, m->st));
    }
    CUDA_CHECK(cudaStreamSynchronize(m->st));
    return NB200_OK;
    NB_CATCH(NB200_ERR)
}

// The step's GEMM + fused epilogue on one of the model's matrices, exactly as enqueue_layers launches it (same operand preparation,
// same tile selection for this M, same epilogue class), for the parity tests of the shapes the benchmark runs.
int nb200_debug_gemm(nb200_model* m, const char* tensor_name, const float* x, int M, int epi_kind, float* y, float* y_ln, int* cfg_out) {
    NB_TRY
    if (!m || !tensor_name || !x || !y || M <= 0) throw Error("bad argument");
    CUDA_CHECK(cudaSetDevice(m->device));
    if (finish_inflight(m) < 0) return NB200_ERR;
    if (m->bf16) throw Error("debug_gemm reads its fp16 results back as fp16: not available for NB200_FLAG_BF16 models (use debug_matmul)");
    auto it = m->by_name.find(tensor_name);
    if (it == m->by_name.end()) throw Error(std::string("no such encoder matrix: ") + tensor_name);
    Operand W;
    operand_init(W, it->second.ptr, it->second.rows, it->second.K);
    const int K = W.K, N = W.rows, aq = m->act_quant ? 1 : 0;
    if (K % 1024 != 0) throw Error("debug_gemm supports K multiple of 1024");
    if (epi_kind == EPI_GLU && !it->second.glu_interleaved) throw Error("debug_gemm: the GLU epilogue needs pointwise_conv1 (value / gate rows interleaved)");
    if (epi_kind == EPI_PARTIAL && N != D_MODEL) throw Error("debug_gemm: the split-K + LayerNorm class needs N = 1024");
    const int Mp = (M + 127) / 128 * 128;
    const int n_out = epi_kind == EPI_GLU ? N / 2 : N;
    DevBufs bufs;
    float* dx = bufs.get<float>((size_t)M * K);
    __half* dx16 = bufs.get<__half>((size_t)Mp * K);
    cudaStream_t st = m->st;
    CUDA_CHECK(cudaMemsetAsync(dx16, 0, (size_t)Mp * K * sizeof(__half), st));
    CUDA_CHECK(cudaMemcpyAsync(dx, x, (size_t)M * K * sizeof(float), cudaMemcpyHostToDevice, st));
    launch_cast_f16(st, dx, dx16, aq, M, K);
    Operand A;
    operand_init(A, dx16, Mp, K);
    EpiParams ep;
    ep.ldo = n_out;
    const int max_splits = epi_kind == EPI_PARTIAL ? m->part_max_splits : 1;
    const bool packed = epi_kind == EPI_STORE_F16 || epi_kind == 6 || epi_kind == EPI_GLU;
    const GemmConfig c = gemm_pick_config(M, N, K, max_splits, packed ? 64 : 32);
    if (cfg_out) { cfg_out[0] = c.bn; cfg_out[1] = c.splits; cfg_out[2] = c.cg; cfg_out[3] = c.ctas; }
    std::vector<float> host((size_t)M * n_out);
    if (epi_kind == EPI_STORE_F16 || epi_kind == 6) {  // linear1 + SiLU (+ re-quantise) -> fp16, or a plain fp16 store (q | k | v)
        __half* d16 = bufs.get<__half>((size_t)Mp * N);
        ep.kind = EPI_STORE_F16; ep.out_f16 = d16;
        if (epi_kind == EPI_STORE_F16) { ep.act = ACT_SILU; ep.act_quant = aq; }
        gemm(st, A, W, M, ep, m->simt);
        std::vector<__half> h16((size_t)M * N);
        CUDA_CHECK(cudaMemcpyAsync(h16.data(), d16, h16.size() * sizeof(__half), cudaMemcpyDeviceToHost, st));
        CUDA_CHECK(cudaStreamSynchronize(st));
        for (size_t i = 0; i < h16.size(); i++) y[i] = __half2float(h16[i]);
        return NB200_OK;
    }
    if (epi_kind == EPI_PARTIAL) {  // split-K partial planes, then the step's LayerNorm adds them to a zero residual stream: y = sum, y_ln = fp16(fq(LN(sum)))
        const size_t plane = (size_t)Mp * N;
        float* dpart = bufs.get<float>((size_t)max_splits * plane);
        float* dres = bufs.get<float>(plane);
        __half* dln = bufs.get<__half>(plane);
        std::vector<float> ones(N, 1.0f), zeros(N, 0.0f);
        float* dg = bufs.get<float>(N);
        float* db = bufs.get<float>(N);
        CUDA_CHECK(cudaMemcpyAsync(dg, ones.data(), N * sizeof(float), cudaMemcpyHostToDevice, st));
        CUDA_CHECK(cudaMemcpyAsync(db, zeros.data(), N * sizeof(float), cudaMemcpyHostToDevice, st));
        CUDA_CHECK(cudaMemsetAsync(dres, 0, plane * sizeof(float), st));
        ep.kind = EPI_PARTIAL; ep.out_f32 = dpart; ep.split_stride = plane;
        ResidualIn rin;
        rin.partials = dpart; rin.stride = plane; rin.alpha = 1.0f;
        rin.nsplit = gemm(st, A, W, M, ep, m->simt, max_splits);
        launch_layernorm(st, dres, dg, db, nullptr, nullptr, nullptr, dln, aq, M, rin);
        CUDA_CHECK(cudaMemcpyAsync(y, dres, (size_t)M * N * sizeof(float), cudaMemcpyDeviceToHost, st));
        std::vector<__half> h16((size_t)M * N);
        CUDA_CHECK(cudaMemcpyAsync(h16.data(), dln, h16.size() * sizeof(__half), cudaMemcpyDeviceToHost, st));
        CUDA_CHECK(cudaStreamSynchronize(st));
        if (y_ln) for (size_t i = 0; i < h16.size(); i++) y_ln[i] = __half2float(h16[i]);
        return NB200_OK;
    }
    float* dy = bufs.get<float>((size_t)Mp * n_out);
    ep.out_f32 = dy;
    if (epi_kind == EPI_GLU) ep.kind = EPI_GLU;
    else if (epi_kind == EPI_STORE) ep.kind = EPI_STORE;
    else throw Error("debug_gemm: epilogue kind must be 0 (f32), 1 (SiLU -> f16), 4 (GLU), 5 (split-K + LayerNorm) or 6 (f16)");
    gemm(st, A, W, M, ep, m->simt);
    CUDA_CHECK(cudaMemcpyAsync(y, dy, (size_t)M * n_out * sizeof(float), cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    return NB200_OK;
    NB_CATCH(NB200_ERR)
}

int nb200_bench_mma_peak(nb200_model* m, int kind, int ctas, int iters, const uint8_t* a_bytes, const uint8_t* b_bytes, uint32_t* check_out,
                         double* tops_out) {
    NB_TRY
    if (!m || !a_bytes || !b_bytes || !tops_out || (kind < 0 || kind > 2) || ctas <= 0 || iters <= 0) throw Error("bad argument");
    CUDA_CHECK(cudaSetDevice(m->device));
    if (finish_inflight(m) < 0) return NB200_ERR;
    DevBufs bufs;
    uint8_t* da = bufs.get<uint8_t>(128 * 128);
    uint8_t* db = bufs.get<uint8_t>(256 * 128);
    uint32_t* dc = bufs.get<uint32_t>(128 * 256);
    cudaStream_t st = m->st;
    CUDA_CHECK(cudaMemcpyAsync(da, a_bytes, 128 * 128, cudaMemcpyHostToDevice, st));
    CUDA_CHECK(cudaMemcpyAsync(db, b_bytes, 256 * 128, cudaMemcpyHostToDevice, st));
    if (check_out) {
        mma_peak_run(st, kind, 1, kind == 2 ? 8 : 1, da, db, dc);  // kind 2: 8 blocks (4 distinct K slices of the operands, twice)
        CUDA_CHECK(cudaMemcpyAsync(check_out, dc, 128 * 256 * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
        CUDA_CHECK(cudaStreamSynchronize(st));
    }
    mma_peak_run(st, kind, ctas, std::min(iters, 64), da, db, nullptr);  // warm-up
    const float ms = mma_peak_run(st, kind, ctas, iters, da, db, nullptr);
    const double k_total = (double)iters * (kind == 0 ? 64.0 : (kind == 1 ? 128.0 : 32.0));
    *tops_out = 2.0 * 128.0 * 256.0 * k_total * ctas / (ms * 1e-3) * 1e-12;
    return NB200_OK;
    NB_CATCH(NB200_ERR)
}

// ---- allocation-free variant of nb200_mel_batch (no counterpart in the reference) ----
// For callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it):
// the device-side scratch nb200_mel_batch allocates per call (PCM rings, per-stream metadata, the log-mel output before it is
// copied to the host) lives in ONE file-scope device array of fixed size instead. Every GPU that loads the library holds the
// array until exit. Limits: NB200_NOALLOC_MAX_STREAMS signals of at most NB200_NOALLOC_MAX_SAMPLES samples each.
// All calls share the array and it keeps the last call's contents. Nothing has to start at zero: every word the kernel reads is
// written by the same call first (samples beyond n_samples are masked by the n_avail bound, all n_streams * n_frames * 128 outputs
// are written before they are copied out). Calls that overlap on the GPU (two models on one device stepping from two threads)
// race on it; calls on one model are ordered by the model's stream.
}  // extern "C" (the array below has file scope)

constexpr int NA_MAX_STREAMS = NB200_NOALLOC_MAX_STREAMS, NA_MAX_SAMPLES = NB200_NOALLOC_MAX_SAMPLES;  // 16 x 65536 (4.096 s at 16 kHz)
constexpr int NA_MAX_FRAMES = (256 + NA_MAX_SAMPLES - 512) / 160 + 1;                                   // 409
constexpr size_t NA_PCM_BYTES = (size_t)NA_MAX_STREAMS * NA_MAX_SAMPLES * sizeof(int16_t);              // 2 MiB: one 65536-sample "ring" per signal
constexpr size_t NA_MEL_BYTES = (size_t)NA_MAX_STREAMS * NA_MAX_FRAMES * nb::N_MELS * sizeof(float);    // 3.2 MiB
constexpr size_t NA_META_BYTES = 1024;                                                                  // 2 x 16 ints + 3 x 16 long longs
constexpr size_t NA_TOTAL_BYTES = NA_PCM_BYTES + NA_MEL_BYTES + NA_META_BYTES;                          // 5.2 MiB per GPU
static_assert(NA_PCM_BYTES % 256 == 0 && NA_MEL_BYTES % 256 == 0, "sub-buffers keep the 256-byte alignment");
static_assert((NA_MAX_SAMPLES & (NA_MAX_SAMPLES - 1)) == 0, "the mel kernel indexes the PCM buffer as a power-of-two ring");
static_assert(2 * NA_MAX_STREAMS * sizeof(int) + 3 * NA_MAX_STREAMS * sizeof(long long) <= NA_META_BYTES, "metadata region");

static __device__ __align__(256) unsigned char nb200_noalloc_scratch[NA_TOTAL_BYTES];

// In host code the array's name is only a stand-in for the symbol: its device address comes from cudaGetSymbolAddress, once per
// device, with that device current.
static unsigned char* noalloc_scratch_ptr() {
    static unsigned char* ptr[64] = {};
    static nb::PerDeviceOnce once;
    int dev = 0;
    CUDA_CHECK(cudaGetDevice(&dev));
    dev &= 63;
    if (once.first() || ptr[dev] == nullptr) {
        void* p = nullptr;
        CUDA_CHECK(cudaGetSymbolAddress(&p, nb200_noalloc_scratch));
        if (p == nullptr) throw nb::Error("cudaGetSymbolAddress returned a null address for the no-alloc scratch");
        ptr[dev] = static_cast<unsigned char*>(p);
    }
    return ptr[dev];
}

extern "C" {

int nb200_mel_batch_noalloc(nb200_model* m, const int16_t* pcm, int n_streams, int n_samples, float* mel_out) {
    NB_TRY
    if (!m || !pcm || !mel_out) throw Error("null argument");
    if (n_streams > NA_MAX_STREAMS || n_samples > NA_MAX_SAMPLES)
        throw Error("nb200_mel_batch_noalloc: at most " + std::to_string(NA_MAX_STREAMS) + " signals of at most " + std::to_string(NA_MAX_SAMPLES) +
                    " samples fit its fixed device scratch (got " + std::to_string(n_streams) + " x " + std::to_string(n_samples) +
                    "); use nb200_mel_batch for larger batches");
    CUDA_CHECK(cudaSetDevice(m->device));
    const int nf = (int)frames_available(n_samples);
    if (nf <= 0 || n_streams <= 0) return 0;
    unsigned char* base = noalloc_scratch_ptr();
    int16_t* d_pcm = reinterpret_cast<int16_t*>(base);
    float* d_mel = reinterpret_cast<float*>(base + NA_PCM_BYTES);
    int* d_i = reinterpret_cast<int*>(base + NA_PCM_BYTES + NA_MEL_BYTES);
    long long* d_l = reinterpret_cast<long long*>(base + NA_PCM_BYTES + NA_MEL_BYTES + 2 * NA_MAX_STREAMS * sizeof(int));
    int h_i[2 * NA_MAX_STREAMS];          // slot | n_frames (host metadata on the stack: no host allocation either)
    long long h_l[3 * NA_MAX_STREAMS];    // out_off | a0 | n_avail
    for (int b = 0; b < n_streams; b++) {
        h_i[b] = b;
        h_i[NA_MAX_STREAMS + b] = nf;
        h_l[0 * NA_MAX_STREAMS + b] = (long long)b * nf * N_MELS;
        h_l[1 * NA_MAX_STREAMS + b] = -257;  // 256 zeros of centre padding + the pre-emphasis sample
        h_l[2 * NA_MAX_STREAMS + b] = n_samples;
    }
    LaunchCount lc(m);
    cudaStream_t st = m->st;
    CUDA_CHECK(cudaMemcpy2DAsync(d_pcm, (size_t)NA_MAX_SAMPLES * sizeof(int16_t), pcm, (size_t)n_samples * sizeof(int16_t),
                                 (size_t)n_samples * sizeof(int16_t), n_streams, cudaMemcpyHostToDevice, st));
    CUDA_CHECK(cudaMemcpyAsync(d_i, h_i, sizeof h_i, cudaMemcpyHostToDevice, st));
    CUDA_CHECK(cudaMemcpyAsync(d_l, h_l, sizeof h_l, cudaMemcpyHostToDevice, st));
    MelBatch mb{d_pcm, NA_MAX_SAMPLES, d_i, d_i + NA_MAX_STREAMS, d_l + NA_MAX_STREAMS, d_l + 2 * NA_MAX_STREAMS, d_l};
    launch_mel(st, m->mel_tb, mb, d_mel, n_streams, nf);
    CUDA_CHECK(cudaMemcpyAsync(mel_out, d_mel, (size_t)n_streams * nf * N_MELS * sizeof(float), cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));  // h_i / h_l / the caller's buffers are in use until here
    return nf;
    NB_CATCH(NB200_ERR)
}

}  // extern "C"

// oracle/ref_harness.cpp — TEST INFRASTRUCTURE, not product.
//
// Driver around the UNMODIFIED reference (linked from oracle/_ref/<variant>/libnemoref.so, built by
// oracle/Makefile from /root/reference). It calls only the reference's public API
//   nemo_init_with_backend            (/root/reference/src/nemo-ggml.h:259)
//   nemo_stream_init / _process_incremental / _finalize / _free   (src/nemo-stream.h:271-326)
//   nemo_preprocessor_init_from_data / _process                   (src/preprocessor.h:16-49)
// and reads the public struct fields sctx->encoder_graph.encoder_out, sctx->tokens,
// sctx->{encoder,decoder,transfer}_seconds to dump what the parity tests compare against.
//
//   ref_harness stream <gguf> <pcm.s16le> <right_context> <threads> <out_prefix|-> [max_seconds] [language|-]
//       feeds shift*160 samples per call (=> at most one encoder chunk per call), then finalize.
//       writes <prefix>.mel.f32 [frames,128], <prefix>.enc.f32 [chunks,T,d_model] (all T frames of
//       every chunk, before the valid_out truncation), <prefix>.tokens.i32, <prefix>.chunk_tokens.i32
//       (tokens emitted per chunk) and prints one JSON line with timings.
//   ref_harness mel <gguf> <pcm.s16le> <slice_samples> <out.f32>
//       preprocessor only, PCM fed in slices of the given size.
//   ref_harness matmul <gguf> <tensor_name> <x.f32> <M> <out.f32> [threads]
//       y[M,N] = ggml_mul_mat(W, x) on the reference's CPU backend for one weight tensor of the file (Q8_0 or F32),
//       i.e. exactly the kernel the encoder graph runs for that tensor (src/nemo-stream.cpp:593-603).
//   ref_harness offline <gguf> <pcm.s16le> <threads> <out_prefix> [max_seconds] [language|-]
//       the reference's offline path: nemo_encode_audio (src/nemo-ggml.cpp:1690-1715 = nemo_preprocessor_process over the
//       whole utterance + build_encoder with full-sequence attention + greedy_decode) -> <prefix>.tokens.i32,
//       <prefix>.frames.i64 (timed_token::frame_idx); and, through the public build_encoder (src/nemo-ggml.h:342) on the
//       same mel (+ build_prompt_fusion :351 for multilingual files, like nemo_encode), <prefix>.enc.f32 [T, d_model] and
//       <prefix>.mel.f32. language: nemo_set_language (src/nemo-ggml.cpp:497-511).
//   ref_harness text <gguf> <tokens.i32> <frames.i64> <timestamp_words 0|1> <out.txt>
//       tokens_to_text (src/nemo-ggml.cpp:1556-1583) over timed_token{token, frame} pairs with the file's vocabulary.
#include "nemo-ggml.h"
#include "nemo-stream.h"
#include "preprocessor.h"

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

static std::vector<int16_t> read_pcm(const char* path) {
    std::vector<int16_t> pcm;
    FILE* f = fopen(path, "rb");
    if (!f) { fprintf(stderr, "cannot open %s\n", path); exit(2); }
    fseek(f, 0, SEEK_END);
    long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    pcm.resize(n / 2);
    if (fread(pcm.data(), 2, pcm.size(), f) != pcm.size()) { fprintf(stderr, "short read\n"); exit(2); }
    fclose(f);
    return pcm;
}

static void write_file(const std::string& path, const void* data, size_t bytes) {
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) { fprintf(stderr, "cannot write %s\n", path.c_str()); exit(2); }
    fwrite(data, 1, bytes, f);
    fclose(f);
}

static double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

static int run_mel(int argc, char** argv) {
    if (argc < 6) { fprintf(stderr, "usage: mel <gguf> <pcm> <slice> <out.f32>\n"); return 2; }
    nemo_context* ctx = nemo_init_with_backend(argv[2], NEMO_BACKEND_CPU);
    if (!ctx) return 1;
    std::vector<int16_t> pcm = read_pcm(argv[3]);
    size_t slice = (size_t)atol(argv[4]);
    nemo_preprocessor* pp = nemo_preprocessor_init_from_data(
        ctx->model.preproc_fb_host.data(), ctx->model.preproc_fb_host.size(),
        ctx->model.preproc_win_host.data(), ctx->model.preproc_win_host.size());
    std::vector<float> all, mel;
    for (size_t i = 0; i < pcm.size(); i += slice) {
        size_t n = std::min(slice, pcm.size() - i);
        nemo_preprocessor_process(pp, pcm.data() + i, n, mel);
        all.insert(all.end(), mel.begin(), mel.end());
    }
    write_file(argv[5], all.data(), all.size() * sizeof(float));
    printf("{\"mode\":\"mel\",\"frames\":%zu}\n", all.size() / 128);
    nemo_preprocessor_free(pp);
    nemo_free(ctx);
    return 0;
}

static int run_stream(int argc, char** argv) {
    if (argc < 7) { fprintf(stderr, "usage: stream <gguf> <pcm> <rc> <threads> <prefix|-> [max_seconds]\n"); return 2; }
    const char* gguf = argv[2];
    std::vector<int16_t> pcm = read_pcm(argv[3]);
    int rc = atoi(argv[4]);
    int threads = atoi(argv[5]);
    std::string prefix = argv[6];
    bool dump = prefix != "-";
    double max_seconds = argc > 7 ? atof(argv[7]) : 0.0;
    if (max_seconds > 0 && pcm.size() > (size_t)(max_seconds * 16000)) pcm.resize((size_t)(max_seconds * 16000));

    double t0 = now_s();
    nemo_context* ctx = nemo_init_with_backend(gguf, NEMO_BACKEND_CPU);
    if (!ctx) return 1;
    if (threads > 0) ggml_backend_cpu_set_n_threads(ctx->model.backend, threads);
    double t_load = now_s() - t0;

    nemo_cache_config cfg = nemo_cache_config::default_config();
    cfg.att_right_context = rc;
    nemo_stream_context* s = nemo_stream_init(ctx, &cfg);
    if (!s) return 1;
    // multilingual checkpoints: optional language code (nemo_stream_set_language, src/nemo-stream.cpp:735-749); "-" keeps
    // the loader's default prompt ("auto", src/nemo-ggml.cpp:459-462)
    if (argc > 8 && strcmp(argv[8], "-") != 0 && !nemo_stream_set_language(s, argv[8])) return 3;

    const size_t feed = s->config.get_shift_mel_frames() * s->config.hop_length;
    const int d_model = s->config.d_model;

    // mel dump: a second, independent preprocessor instance over the same PCM
    std::vector<float> mel_all;
    if (dump) {
        nemo_preprocessor* pp = nemo_preprocessor_init_from_data(
            ctx->model.preproc_fb_host.data(), ctx->model.preproc_fb_host.size(),
            ctx->model.preproc_win_host.data(), ctx->model.preproc_win_host.size());
        std::vector<float> mel;
        for (size_t i = 0; i < pcm.size(); i += feed) {
            nemo_preprocessor_process(pp, pcm.data() + i, std::min(feed, pcm.size() - i), mel);
            mel_all.insert(mel_all.end(), mel.begin(), mel.end());
        }
        nemo_preprocessor_free(pp);
    }

    std::vector<float> enc_all;
    std::vector<int32_t> chunk_tokens;
    int chunks_seen = 0;
    size_t tokens_seen = 0;
    int enc_T = 0;
    auto after_call = [&]() {
        if (s->total_chunks_processed > chunks_seen) {
            // by construction at most one chunk per call
            if (dump) {
                ggml_tensor* e = s->encoder_graph.encoder_out;
                enc_T = (int)e->ne[1];
                std::vector<float> buf(ggml_nelements(e));
                ggml_backend_tensor_get(e, buf.data(), 0, buf.size() * sizeof(float));
                enc_all.insert(enc_all.end(), buf.begin(), buf.end());
            }
            chunk_tokens.push_back((int32_t)(s->tokens.size() - tokens_seen));
            tokens_seen = s->tokens.size();
            chunks_seen = s->total_chunks_processed;
        }
    };

    double t1 = now_s();
    for (size_t i = 0; i < pcm.size(); i += feed) {
        nemo_stream_process_incremental(s, pcm.data() + i, (int)std::min(feed, pcm.size() - i));
        after_call();
    }
    nemo_stream_finalize(s);
    after_call();
    double t_proc = now_s() - t1;

    if (dump) {
        write_file(prefix + ".mel.f32", mel_all.data(), mel_all.size() * sizeof(float));
        write_file(prefix + ".enc.f32", enc_all.data(), enc_all.size() * sizeof(float));
        std::vector<int32_t> tok(s->tokens.begin(), s->tokens.end());
        write_file(prefix + ".tokens.i32", tok.data(), tok.size() * sizeof(int32_t));
        write_file(prefix + ".chunk_tokens.i32", chunk_tokens.data(), chunk_tokens.size() * sizeof(int32_t));
    }
    double audio_s = (double)pcm.size() / 16000.0;
    printf("{\"mode\":\"stream\",\"rc\":%d,\"threads\":%d,\"audio_s\":%.4f,\"proc_s\":%.4f,\"load_s\":%.3f,"
           "\"rtfx\":%.4f,\"chunks\":%d,\"enc_T\":%d,\"d_model\":%d,\"tokens\":%zu,\"decode_iters\":%d,"
           "\"encoder_s\":%.4f,\"decoder_s\":%.4f,\"transfer_s\":%.4f}\n",
           rc, threads, audio_s, t_proc, t_load, audio_s / t_proc, chunks_seen, enc_T, d_model,
           s->tokens.size(), s->total_decode_iterations, s->encoder_seconds, s->decoder_seconds,
           s->transfer_seconds);
    nemo_stream_free(s);
    nemo_free(ctx);
    return 0;
}

static int run_matmul(int argc, char** argv) {
    if (argc < 7) { fprintf(stderr, "usage: matmul <gguf> <tensor> <x.f32> <M> <out.f32> [threads]\n"); return 2; }
    nemo_context* ctx = nemo_init_with_backend(argv[2], NEMO_BACKEND_CPU);
    if (!ctx) return 1;
    if (argc > 7) ggml_backend_cpu_set_n_threads(ctx->model.backend, atoi(argv[7]));
    auto it = ctx->model.tensors.find(argv[3]);
    if (it == ctx->model.tensors.end()) { fprintf(stderr, "no tensor %s\n", argv[3]); return 1; }
    ggml_tensor* W = it->second;
    const int M = atoi(argv[5]);
    const int64_t K = W->ne[0], N = W->ne[1];
    std::vector<float> x((size_t)M * K);
    FILE* f = fopen(argv[4], "rb");
    if (!f || fread(x.data(), sizeof(float), x.size(), f) != x.size()) { fprintf(stderr, "cannot read %s\n", argv[4]); return 2; }
    fclose(f);
    ggml_init_params ip = {ggml_tensor_overhead() * 8 + ggml_graph_overhead(), nullptr, true};
    ggml_context* c = ggml_init(ip);
    ggml_tensor* X = ggml_new_tensor_2d(c, GGML_TYPE_F32, K, M);
    ggml_tensor* Y = ggml_mul_mat(c, W, X);
    ggml_cgraph* g = ggml_new_graph(c);
    ggml_build_forward_expand(g, Y);
    ggml_gallocr_t ga = ggml_gallocr_new(ggml_backend_get_default_buffer_type(ctx->model.backend));
    ggml_gallocr_alloc_graph(ga, g);
    ggml_backend_tensor_set(X, x.data(), 0, x.size() * sizeof(float));
    ggml_backend_graph_compute(ctx->model.backend, g);
    std::vector<float> y((size_t)M * N);
    ggml_backend_tensor_get(Y, y.data(), 0, y.size() * sizeof(float));
    write_file(argv[6], y.data(), y.size() * sizeof(float));
    printf("{\"mode\":\"matmul\",\"M\":%d,\"N\":%lld,\"K\":%lld,\"type\":\"%s\"}\n", M, (long long)N, (long long)K, ggml_type_name(W->type));
    ggml_gallocr_free(ga);
    ggml_free(c);
    nemo_free(ctx);
    return 0;
}

static int run_offline(int argc, char** argv) {
    if (argc < 6) { fprintf(stderr, "usage: offline <gguf> <pcm> <threads> <prefix> [max_seconds]\n"); return 2; }
    std::vector<int16_t> pcm = read_pcm(argv[3]);
    const int threads = atoi(argv[4]);
    const std::string prefix = argv[5];
    const double max_seconds = argc > 6 ? atof(argv[6]) : 0.0;
    if (max_seconds > 0 && pcm.size() > (size_t)(max_seconds * 16000)) pcm.resize((size_t)(max_seconds * 16000));
    nemo_context* ctx = nemo_init_with_backend(argv[2], NEMO_BACKEND_CPU);
    if (!ctx) return 1;
    if (threads > 0) ggml_backend_cpu_set_n_threads(ctx->model.backend, threads);
    if (argc > 7 && strcmp(argv[7], "-") != 0 && !nemo_set_language(ctx, argv[7])) return 3;

    // nemo_encode_audio's body (src/nemo-ggml.cpp:1690-1715): preprocessor over the whole utterance, then nemo_encode. (The
    // header declares it with a std::vector& parameter that has no definition, src/nemo-ggml.h:370-373, so the two calls
    // are made here.)
    double t0 = now_s();
    std::vector<float> mel;
    const size_t n_frames = nemo_preprocessor_process(ctx->preprocessor, pcm.data(), pcm.size(), mel);
    std::vector<timed_token> tt = nemo_encode(ctx, mel.data(), (int)n_frames);
    const double t_proc = now_s() - t0;
    std::vector<int32_t> tok;
    std::vector<int64_t> frames;
    for (const timed_token& t : tt) { tok.push_back(t.token_id); frames.push_back(t.frame_idx); }
    write_file(prefix + ".tokens.i32", tok.data(), tok.size() * sizeof(int32_t));
    write_file(prefix + ".frames.i64", frames.data(), frames.size() * sizeof(int64_t));

    // encoder output of the same utterance: the graph nemo_encode builds (src/nemo-ggml.cpp:1592-1658), minus the decode
    write_file(prefix + ".mel.f32", mel.data(), mel.size() * sizeof(float));
    const int n_mels = ctx->model.hparams.n_mels;
    size_t buf_size = ggml_tensor_overhead() * 8000 + ggml_graph_overhead_custom(16384, false);
    std::vector<uint8_t> buf(buf_size);
    ggml_init_params ip = {buf_size, buf.data(), true};
    ggml_context* c0 = ggml_init(ip);
    ggml_tensor* inp = ggml_new_tensor_3d(c0, GGML_TYPE_F32, n_mels, (int64_t)n_frames, 1);
    ggml_set_input(inp);
    ggml_tensor* enc = build_encoder(c0, inp, &ctx->model);
    ggml_tensor* prompt = nullptr;
    const int num_prompts = ctx->model.hparams.num_prompts;
    if (num_prompts > 0) {  // what nemo_encode does at src/nemo-ggml.cpp:1630-1636 (its helpers are file-static)
        prompt = ggml_new_tensor_3d(c0, GGML_TYPE_F32, num_prompts, enc->ne[1], enc->ne[2]);
        ggml_set_input(prompt);
        enc = build_prompt_fusion(c0, enc, prompt, &ctx->model);
    }
    ggml_set_output(enc);
    ggml_cgraph* gf = ggml_new_graph_custom(c0, 16384, false);
    ggml_build_forward_expand(gf, enc);
    ggml_gallocr_t ga = ggml_gallocr_new(ggml_backend_get_default_buffer_type(ctx->model.backend));
    if (!ggml_gallocr_alloc_graph(ga, gf)) return 1;
    ggml_backend_tensor_set(inp, mel.data(), 0, mel.size() * sizeof(float));
    if (prompt) {  // one-hot rows, fill_prompt_onehot src/nemo-ggml.cpp:1118-1130
        std::vector<float> oh((size_t)ggml_nelements(prompt), 0.0f);
        const int idx = (ctx->prompt_index >= 0 && ctx->prompt_index < num_prompts) ? ctx->prompt_index : 0;
        for (int64_t r = 0; r < prompt->ne[1] * prompt->ne[2]; r++) oh[(size_t)r * num_prompts + idx] = 1.0f;
        ggml_backend_tensor_set(prompt, oh.data(), 0, oh.size() * sizeof(float));
    }
    ggml_backend_graph_compute(ctx->model.backend, gf);
    std::vector<float> out(ggml_nelements(enc));
    ggml_backend_tensor_get(enc, out.data(), 0, out.size() * sizeof(float));
    write_file(prefix + ".enc.f32", out.data(), out.size() * sizeof(float));
    printf("{\"mode\":\"offline\",\"threads\":%d,\"audio_s\":%.4f,\"proc_s\":%.4f,\"mel_frames\":%zu,\"enc_T\":%lld,\"tokens\":%zu}\n",
           threads, (double)pcm.size() / 16000.0, t_proc, n_frames, (long long)enc->ne[1], tok.size());
    ggml_gallocr_free(ga);
    ggml_free(c0);
    nemo_free(ctx);
    return 0;
}

// declared in the reference's translation units only (src/nemo-stream.cpp:752), defined at src/nemo-ggml.cpp:1556
std::string tokens_to_text(const std::vector<timed_token>& tokens, const std::vector<std::string>& vocab, bool timestamp_words);

static int run_text(int argc, char** argv) {
    if (argc < 7) { fprintf(stderr, "usage: text <gguf> <tokens.i32> <frames.i64> <timestamp_words> <out.txt>\n"); return 2; }
    nemo_context* ctx = nemo_init_with_backend(argv[2], NEMO_BACKEND_CPU);
    if (!ctx) return 1;
    std::vector<int32_t> tok;
    std::vector<int64_t> frames;
    {
        FILE* f = fopen(argv[3], "rb");
        if (!f) return 2;
        int32_t v;
        while (fread(&v, 4, 1, f) == 1) tok.push_back(v);
        fclose(f);
        f = fopen(argv[4], "rb");
        if (!f) return 2;
        int64_t w;
        while (fread(&w, 8, 1, f) == 1) frames.push_back(w);
        fclose(f);
    }
    if (tok.size() != frames.size()) { fprintf(stderr, "tokens / frames length mismatch\n"); return 2; }
    std::vector<timed_token> tt;
    for (size_t i = 0; i < tok.size(); i++) tt.push_back(timed_token(tok[i], frames[i]));
    const std::string text = tokens_to_text(tt, ctx->model.vocab, atoi(argv[5]) != 0);
    write_file(argv[6], text.data(), text.size());
    nemo_free(ctx);
    return 0;
}

int main(int argc, char** argv) {
    if (argc < 2) { fprintf(stderr, "usage: ref_harness stream|mel|matmul|text|offline ...\n"); return 2; }
    if (!strcmp(argv[1], "mel")) return run_mel(argc, argv);
    if (!strcmp(argv[1], "stream")) return run_stream(argc, argv);
    if (!strcmp(argv[1], "matmul")) return run_matmul(argc, argv);
    if (!strcmp(argv[1], "text")) return run_text(argc, argv);
    if (!strcmp(argv[1], "offline")) return run_offline(argc, argv);
    fprintf(stderr, "unknown mode %s\n", argv[1]);
    return 2;
}

// tests/adapter/nemo_stream_b200.cpp — TEST INFRASTRUCTURE: the reference-side binding of INTEGRATION.md section 2, compiled for real.
//
// Implements the free functions the reference's streaming front-ends call (declared in the reference's own headers,
// src/nemo-ggml.h:259-269 and src/nemo-stream.h:271-326, which this file includes from the reference tree) on top of the C ABI
// of libnemob200.so. oracle/Makefile links it with the UNMODIFIED src/transcribe_stream.cpp (the CLI's read loop, :229-254):
// the resulting binary is the reference's CLI running on the B200 backend, with no ggml object linked at all.
// The contexts are opaque to the CLI (it only passes the pointers around), so nemo_context / nemo_stream_context pointers carry
// our own wrapper structs; the reference's structs are never instantiated.
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "nemo-ggml.h"
#include "nemo-stream.h"

#include "nb200.h"

namespace {

struct B200Context {   // travels as struct nemo_context*
    nb200_model* model = nullptr;
};
struct B200Stream {    // travels as struct nemo_stream_context*
    B200Context* ctx = nullptr;
    nb200_stream* stream = nullptr;
    std::vector<int> tokens;   // sctx->tokens (src/nemo-stream.h:231)
    std::string transcript;    // sctx->transcript
};

B200Context* unwrap(struct nemo_context* c) { return reinterpret_cast<B200Context*>(c); }
B200Stream* unwrap(struct nemo_stream_context* s) { return reinterpret_cast<B200Stream*>(s); }

bool fail(const char* what) {
    fprintf(stderr, "%s: %s\n", what, nb200_last_error());
    return false;
}

// new tokens of the stream -> text of exactly those tokens (what process_mel_chunk_streaming returns per chunk, concatenated)
std::string collect(B200Stream* s) {
    std::string text;
    for (;;) {
        int32_t tok[1024];
        const int n = nb200_stream_poll_tokens(s->stream, tok, 1024);
        if (n <= 0) break;
        s->tokens.insert(s->tokens.end(), tok, tok + n);
        std::vector<char> buf(64 * (size_t)n + 16);
        const int len = nb200_tokens_to_text(s->ctx->model, tok, n, buf.data(), (int)buf.size());
        if (len > 0) text.append(buf.data(), (size_t)len);
        if (n < 1024) break;
    }
    s->transcript += text;
    return text;
}

}  // namespace

// ---- src/nemo-ggml.h:259-269 ----------------------------------------------------------------------------------
struct nemo_context* nemo_init_with_backend(const char* model_path, nemo_backend_type backend) {
    if (backend == NEMO_BACKEND_CPU || backend == NEMO_BACKEND_METAL) {
        fprintf(stderr, "this build has only the B200 backend (no CPU fallback)\n");
        return nullptr;
    }
    const char* dev = getenv("NB200_DEVICE");
    auto* c = new B200Context();
    c->model = nb200_model_load(model_path, dev ? atoi(dev) : 0, /*max_streams*/ 4, 0);
    if (!c->model) {
        fail("nb200_model_load");
        delete c;
        return nullptr;
    }
    return reinterpret_cast<struct nemo_context*>(c);
}

void nemo_free(struct nemo_context* ctx) {
    if (!ctx) return;
    nb200_model_free(unwrap(ctx)->model);
    delete unwrap(ctx);
}

bool nemo_set_language(struct nemo_context* ctx, const char* lang) {
    return ctx && nb200_model_set_language(unwrap(ctx)->model, lang) == NB200_OK;
}

const char* nemo_get_backend_name(struct nemo_context*) { return "B200 (sm_100a)"; }

// ---- src/nemo-stream.h:271-326 --------------------------------------------------------------------------------
struct nemo_stream_context* nemo_stream_init(struct nemo_context* ctx, const nemo_cache_config* config) {
    if (!ctx) return nullptr;
    const nemo_cache_config cfg = config ? *config : nemo_cache_config::default_config();
    auto* s = new B200Stream();
    s->ctx = unwrap(ctx);
    s->stream = nb200_stream_create(s->ctx->model, cfg.att_right_context);
    if (!s->stream) {
        fail("nb200_stream_create");
        delete s;
        return nullptr;
    }
    return reinterpret_cast<struct nemo_stream_context*>(s);
}

bool nemo_stream_set_language(struct nemo_stream_context* sctx, const char* lang) {
    return sctx && nb200_stream_set_language(unwrap(sctx)->stream, lang) == NB200_OK;
}

std::string nemo_stream_process_incremental(struct nemo_stream_context* sctx, const int16_t* audio, int n_samples) {
    if (!sctx || !audio || n_samples <= 0) return "";
    B200Stream* s = unwrap(sctx);
    if (nb200_stream_feed(s->stream, audio, n_samples) != NB200_OK) { fail("nb200_stream_feed"); return ""; }
    while (nb200_stream_ready_chunks(s->stream) > 0)
        if (nb200_step(s->ctx->model) < 0) { fail("nb200_step"); return ""; }
    return collect(s);
}

std::string nemo_stream_process(struct nemo_stream_context* sctx, const int16_t* audio, int n_samples) {
    return nemo_stream_process_incremental(sctx, audio, n_samples);
}

std::string nemo_stream_finalize(struct nemo_stream_context* sctx) {
    if (!sctx) return "";
    B200Stream* s = unwrap(sctx);
    if (nb200_stream_finalize(s->stream) != NB200_OK) { fail("nb200_stream_finalize"); return ""; }
    if (nb200_drain(s->ctx->model) < 0) { fail("nb200_drain"); return ""; }
    return collect(s);
}

std::string nemo_stream_get_transcript(struct nemo_stream_context* sctx) { return sctx ? unwrap(sctx)->transcript : std::string(); }

const std::vector<int>& nemo_stream_get_tokens(struct nemo_stream_context* sctx) {
    static const std::vector<int> empty;
    return sctx ? unwrap(sctx)->tokens : empty;
}

void nemo_stream_reset(struct nemo_stream_context* sctx) {
    if (!sctx) return;
    B200Stream* s = unwrap(sctx);
    if (nb200_stream_reset(s->stream) != NB200_OK) fail("nb200_stream_reset");
    s->tokens.clear();
    s->transcript.clear();
}

void nemo_stream_free(struct nemo_stream_context* sctx) {
    if (!sctx) return;
    nb200_stream_free(unwrap(sctx)->stream);
    delete unwrap(sctx);
}

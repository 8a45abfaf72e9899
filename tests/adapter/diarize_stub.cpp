// tests/adapter/diarize_stub.cpp — TEST INFRASTRUCTURE: link-time stand-ins for the diarization pipeline the reference's CLI
// can optionally drive (src/diarize_pipeline.h:64-88). Diarization is out of scope; the CLI fixture is run
// without --diarize, so none of these is ever called: init fails loudly if someone tries.
#include <cstdio>

#include "diarize_pipeline.h"

diarize_pipeline* diarize_pipeline_init(const diarize_pipeline_cfg&) {
    fprintf(stderr, "diarization is not part of this test fixture\n");
    return nullptr;
}
void diarize_pipeline_free(diarize_pipeline*) {}
size_t diarize_pipeline_push_audio(diarize_pipeline*, const float*, size_t) { return 0; }
void diarize_pipeline_push_text(diarize_pipeline*, const std::string&, double) {}
std::string diarize_pipeline_drain_json(diarize_pipeline*) { return ""; }
std::string diarize_pipeline_finalize(diarize_pipeline*) { return ""; }
size_t diarize_pipeline_n_embeddings(const diarize_pipeline*) { return 0; }
size_t diarize_pipeline_n_segments(const diarize_pipeline*) { return 0; }
size_t diarize_pipeline_n_words(const diarize_pipeline*) { return 0; }

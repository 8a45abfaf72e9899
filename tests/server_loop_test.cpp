// Test driver for nemotron-asr.cpp_b200/server_loop.h (built and run by tests/test_gpu_parity.py::test_server_loop_batches).
// usage: server_loop_test <gguf> <n_streams> <pcm file>...
// Plays the reference server's event sequence (CREATE, DATA in 100 ms packets from two reader threads, END) for
// n_streams sessions into a BatchWorker, then runs the same sessions one at a time through the plain C ABI
// (feed / step / poll, the reference's serial order) and compares the texts. Prints one JSON line.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "server_loop.h"

static std::vector<int16_t> read_pcm(const char* path) {
    std::ifstream f(path, std::ios::binary);
    std::vector<char> raw((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
    std::vector<int16_t> v(raw.size() / 2);
    memcpy(v.data(), raw.data(), v.size() * 2);
    return v;
}

static std::string run_serial(nb200_model* m, const std::vector<int16_t>& pcm, int rc) {
    nb200_stream* s = nb200_stream_create(m, rc);
    std::string text;
    std::vector<int32_t> toks(4096);
    std::vector<char> buf(1 << 16);
    auto collect = [&] {
        for (;;) {
            const int n = nb200_stream_poll_tokens(s, toks.data(), (int)toks.size());
            if (n <= 0) break;
            const int len = nb200_tokens_to_text(m, toks.data(), n, buf.data(), (int)buf.size());
            text.append(buf.data(), (size_t)len);
        }
    };
    for (size_t i = 0; i < pcm.size(); i += 1600) {
        nb200_stream_feed(s, pcm.data() + i, (int)std::min<size_t>(1600, pcm.size() - i));
        while (nb200_stream_ready_chunks(s) > 0) nb200_step(m);
        collect();
    }
    nb200_stream_finalize(s);
    while (nb200_stream_ready_chunks(s) > 0) nb200_step(m);
    nb200_step(m);
    collect();
    nb200_stream_free(s);
    return text;
}

int main(int argc, char** argv) {
    if (argc < 4) return 2;
    const int n_streams = atoi(argv[2]);
    std::vector<std::vector<int16_t>> pcms;
    for (int i = 3; i < argc; i++) pcms.push_back(read_pcm(argv[i]));
    // NB200_TEST_DEVICES=n: one model replica per GPU in this process, sessions pinned by stream_id % n (ShardedWorkers)
    const int n_dev = getenv("NB200_TEST_DEVICES") ? atoi(getenv("NB200_TEST_DEVICES")) : 1;
    std::vector<nb200_model*> models;
    for (int d = 0; d < n_dev; d++) {
        nb200_model* md = nb200_model_load(argv[1], d, n_streams + 1, 0);
        if (!md) {
            fprintf(stderr, "load failed on device %d: %s\n", d, nb200_last_error());
            return 1;
        }
        models.push_back(md);
    }
    nb200_model* m = models[0];
    const int modes[4] = {13, 0, 6, 13};
    std::mutex mu;
    std::map<uint64_t, std::string> texts;
    std::map<uint64_t, int> finals;
    nb200::WorkerStats st;
    {
        nb200::ShardedWorkers w(models, [&](uint64_t id, const std::string& t, bool fin) {
            std::lock_guard<std::mutex> g(mu);
            texts[id] += t;
            finals[id] += fin ? 1 : 0;
        }, 3000);
        for (int i = 0; i < n_streams; i++) w.post({nb200::Event::CREATE, (uint64_t)i, modes[i % 4], {}});
        auto reader = [&](int first) {  // two "connections threads", each serving every other session, 100 ms packets
            size_t longest = 0;
            for (auto& p : pcms) longest = std::max(longest, p.size());
            for (size_t off = 0; off < longest; off += 1600) {
                for (int i = first; i < n_streams; i += 2) {
                    const auto& p = pcms[i % pcms.size()];
                    if (off >= p.size()) continue;
                    const size_t k = std::min<size_t>(1600, p.size() - off);
                    w.post({nb200::Event::DATA, (uint64_t)i, 0, std::vector<int16_t>(p.begin() + off, p.begin() + off + k)});
                }
                std::this_thread::sleep_for(std::chrono::microseconds(500));
            }
            for (int i = first; i < n_streams; i += 2) w.post({nb200::Event::END, (uint64_t)i, 0, {}});
        };
        std::thread r0(reader, 0), r1(reader, 1);
        r0.join();
        r1.join();
        w.stop();
        for (size_t d = 0; d < w.size(); d++) {
            const nb200::WorkerStats sd = w.stats(d);
            st.steps += sd.steps; st.stream_chunks += sd.stream_chunks; st.events += sd.events; st.errors += sd.errors;
            st.max_batch = std::max(st.max_batch, sd.max_batch);
        }
    }
    int mismatches = 0, not_closed = 0;
    for (int i = 0; i < n_streams; i++) {
        const std::string ref = run_serial(m, pcms[i % pcms.size()], modes[i % 4]);
        if (ref != texts[(uint64_t)i]) {
            mismatches++;
            fprintf(stderr, "stream %d differs:\n  serial : %s\n  batched: %s\n", i, ref.c_str(), texts[(uint64_t)i].c_str());
        }
        if (finals[(uint64_t)i] != 1) not_closed++;
    }
    nb200_stats gs;
    nb200_get_stats(m, &gs);
    printf("{\"devices\": %d, \"streams\": %d, \"mismatches\": %d, \"not_closed\": %d, \"steps\": %llu, \"stream_chunks\": %llu, \"max_batch\": %d, \"events\": %llu, \"errors\": %llu}\n",
           n_dev, n_streams, mismatches, not_closed, (unsigned long long)st.steps, (unsigned long long)st.stream_chunks, st.max_batch,
           (unsigned long long)st.events, (unsigned long long)st.errors);
    for (nb200_model* md : models) nb200_model_free(md);
    return mismatches || not_closed || st.errors ? 1 : 0;
}

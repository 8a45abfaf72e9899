// server_loop.h — cross-stream batching worker for a streaming ASR server, on top of the C ABI (include/nb200.h) only.
//
// What it replaces: the reference server's worker thread (src/nemo-server.cpp:192-271) pops ONE event at a time and
// runs that stream's chunk alone (CREATE -> nemo_stream_init, DATA -> nemo_stream_process_incremental, END ->
// nemo_stream_finalize), so N sessions cost N encoder runs per chunk period. Here the reader threads post the same
// three event kinds; the worker drains the queue, stages all audio (nb200_stream_feed is a host memcpy), and runs ONE
// batched step over every session that has a full chunk. While the GPU works (nb200_step_begin .. nb200_step_end) it
// keeps draining the queue, so ingest overlaps compute. A short batching window (deadline after the first pending
// audio) lets sessions whose packets arrive a little apart share a step; it bounds the added latency.
//
// The wire protocol, the readers and the text frames stay the reference's (server-protocol.h); `TextSink` is where
// send_text_frame(conn, OP_TEXT, id, text) goes. One BatchWorker per GPU; ShardedWorkers routes stream_id % n_gpu.
#pragma once
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <deque>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "nb200.h"

namespace nb200 {

struct Event {
    enum Kind { CREATE, DATA, END } kind;
    uint64_t stream_id;
    int right_context = 0;       // CREATE: latency mode (att_right_context 0 / 1 / 6 / 13)
    std::vector<int16_t> pcm;    // DATA: s16le 16 kHz samples
};

struct WorkerStats {
    uint64_t steps = 0, stream_chunks = 0, events = 0, errors = 0;
    int max_batch = 0;
};

class BatchWorker {
  public:
    // text: newly decoded text of the session (may be empty on the final call); final: session closed after END
    using TextSink = std::function<void(uint64_t stream_id, const std::string& text, bool final)>;

    BatchWorker(nb200_model* model, TextSink sink, int window_us = 2000)
        : model_(model), sink_(std::move(sink)), window_(std::chrono::microseconds(window_us)), thread_([this] { loop(); }) {}
    ~BatchWorker() { stop(); }
    BatchWorker(const BatchWorker&) = delete;
    BatchWorker& operator=(const BatchWorker&) = delete;

    // thread-safe; called by the connection reader threads
    void post(Event ev) {
        {
            std::lock_guard<std::mutex> g(mu_);
            queue_.push_back(std::move(ev));
        }
        cv_.notify_one();
    }
    // processes everything posted so far (including END flushes), then stops the thread
    void stop() {
        bool was = false;
        if (!stopping_.compare_exchange_strong(was, true)) {
            if (thread_.joinable()) thread_.join();
            return;
        }
        cv_.notify_one();
        if (thread_.joinable()) thread_.join();
    }
    WorkerStats stats() const {
        std::lock_guard<std::mutex> g(mu_);
        return stats_;
    }

  private:
    struct Session {
        nb200_stream* s = nullptr;
        bool ending = false;
    };

    // moves the queued events out; returns false when there were none
    bool take(std::deque<Event>& out) {
        std::lock_guard<std::mutex> g(mu_);
        if (queue_.empty()) return false;
        out.swap(queue_);
        return true;
    }
    void apply(std::deque<Event>& evs) {
        for (Event& e : evs) {
            stats_local_.events++;
            auto it = sessions_.find(e.stream_id);
            switch (e.kind) {
                case Event::CREATE: {
                    if (it != sessions_.end()) break;
                    nb200_stream* s = nb200_stream_create(model_, e.right_context);
                    if (!s) {
                        stats_local_.errors++;
                        sink_(e.stream_id, std::string("error: ") + nb200_last_error(), true);
                        break;
                    }
                    sessions_[e.stream_id].s = s;
                    break;
                }
                case Event::DATA:
                    if (it == sessions_.end() || it->second.ending) break;
                    if (nb200_stream_feed(it->second.s, e.pcm.data(), (int)e.pcm.size()) != NB200_OK) stats_local_.errors++;
                    if (!have_audio_) {
                        have_audio_ = true;
                        first_audio_ = std::chrono::steady_clock::now();
                    }
                    break;
                case Event::END:
                    if (it == sessions_.end() || it->second.ending) break;
                    nb200_stream_finalize(it->second.s);
                    it->second.ending = true;
                    n_ending_++;
                    break;
            }
        }
        evs.clear();
    }
    int ready_streams() const {
        int n = 0;
        for (auto& kv : sessions_) n += nb200_stream_ready_chunks(kv.second.s) > 0 ? 1 : 0;
        return n;
    }
    void emit(bool force_close = false) {
        std::vector<int32_t> toks(4096);
        std::string text;
        for (auto it = sessions_.begin(); it != sessions_.end();) {
            Session& ss = it->second;
            text.clear();
            for (;;) {
                const int n = nb200_stream_poll_tokens(ss.s, toks.data(), (int)toks.size());
                if (n <= 0) break;
                std::vector<char> buf((size_t)n * 16 + 16);
                const int len = nb200_tokens_to_text(model_, toks.data(), n, buf.data(), (int)buf.size());
                if (len > 0) text.append(buf.data(), (size_t)len);
            }
            // after END: once no chunk and no finalize tail is left, every token of the session has been polled above
            const bool closed = ss.ending && (force_close || nb200_stream_ready_chunks(ss.s) == 0);
            if (!text.empty() || closed) sink_(it->first, text, closed);
            if (closed) {
                nb200_stream_free(ss.s);
                n_ending_--;
                it = sessions_.erase(it);
            } else {
                ++it;
            }
        }
    }
    void loop() {
        std::deque<Event> evs;
        for (;;) {
            if (!take(evs)) {
                std::unique_lock<std::mutex> g(mu_);
                if (queue_.empty()) {
                    if (stopping_ && !have_audio_ && n_ending_ == 0) break;
                    // idle: sleep until an event arrives, or until the batching window of already staged audio closes
                    if (have_audio_ || n_ending_ > 0)
                        cv_.wait_until(g, first_audio_ + window_);
                    else
                        cv_.wait(g, [this] { return !queue_.empty() || stopping_.load(); });
                }
                g.unlock();
                take(evs);
            }
            apply(evs);
            const bool window_closed = have_audio_ && std::chrono::steady_clock::now() >= first_audio_ + window_;
            if (!(window_closed || n_ending_ > 0 || stopping_)) continue;
            // one batched step over every session with a full chunk (or a finalize tail); ingest continues meanwhile
            const int ready = ready_streams();
            if (ready == 0 && n_ending_ == 0) {
                have_audio_ = false;  // nothing complete yet: wait for more audio
                continue;
            }
            const int n = nb200_step_begin(model_);
            if (n < 0) {  // report, drop the sessions that were closing, wait for new audio
                stats_local_.errors++;
                have_audio_ = false;
                emit(true);
                continue;
            }
            while (take(evs)) apply(evs);  // staged on the host while the GPU runs the step
            if (nb200_step_end(model_) < 0) stats_local_.errors++;
            stats_local_.steps++;
            stats_local_.stream_chunks += (uint64_t)n;
            if (n > stats_local_.max_batch) stats_local_.max_batch = n;
            emit();
            // more complete chunks buffered (a session sent several at once)? keep stepping without a new window
            have_audio_ = ready_streams() > 0;
            if (have_audio_) first_audio_ = std::chrono::steady_clock::now() - window_;
            {
                std::lock_guard<std::mutex> g(mu_);
                stats_ = stats_local_;
            }
        }
        for (auto& kv : sessions_) nb200_stream_free(kv.second.s);
        sessions_.clear();
        std::lock_guard<std::mutex> g(mu_);
        stats_ = stats_local_;
    }

    nb200_model* model_;
    TextSink sink_;
    std::chrono::microseconds window_;
    mutable std::mutex mu_;
    std::condition_variable cv_;
    std::deque<Event> queue_;
    std::atomic<bool> stopping_{false};
    std::map<uint64_t, Session> sessions_;
    bool have_audio_ = false;
    int n_ending_ = 0;
    std::chrono::steady_clock::time_point first_audio_{};
    WorkerStats stats_, stats_local_;
    std::thread thread_;  // last member: starts after everything else is initialised
};

// One worker (and one full model replica) per GPU; sessions are pinned to GPU stream_id % n (no cross-GPU traffic).
class ShardedWorkers {
  public:
    ShardedWorkers(const std::vector<nb200_model*>& models, BatchWorker::TextSink sink, int window_us = 2000) {
        for (nb200_model* m : models) workers_.emplace_back(new BatchWorker(m, sink, window_us));
    }
    void post(Event ev) { workers_[ev.stream_id % workers_.size()]->post(std::move(ev)); }
    void stop() {
        for (auto& w : workers_) w->stop();
    }
    size_t size() const { return workers_.size(); }
    WorkerStats stats(size_t i) const { return workers_[i]->stats(); }

  private:
    std::vector<std::unique_ptr<BatchWorker>> workers_;
};

}  // namespace nb200

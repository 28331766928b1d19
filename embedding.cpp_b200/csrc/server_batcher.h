// server_batcher.h -- dynamic batching across connections for the embedding server.
//
// The reference server (examples/server.cpp:100-118) serves one connection at a time and calls bert_encode once per
// received text, so the GPU engine would see batches of one.  Here every connection thread submits its text to one
// queue; a single worker drains the queue into ONE bert_encode_batch call (length-sorted, packed, sharded over the
// GPUs by the engine) and hands each caller its own embedding back.  While a batch is on the GPU new requests
// queue up, so the batch size adapts to the load by itself; a short collection window (window_us) additionally lets
// near-simultaneous requests share a launch when the server is idle.
//
// The backend is a callable  void(const std::vector<const char *> &texts, const std::vector<float *> &outs)  that
// writes n_embd floats per text: bert_encode_batch in the server, a stub in the host-side unit test.
#pragma once
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <functional>
#include <future>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace bk {

class DynamicBatcher {
public:
    using Backend = std::function<void(const std::vector<const char *> &, const std::vector<float *> &)>;

    DynamicBatcher(Backend backend, int n_embd, int max_batch, int window_us)
        : backend_(std::move(backend)), n_embd_(n_embd), max_batch_(max_batch < 1 ? 1 : max_batch), window_us_(window_us < 0 ? 0 : window_us),
          worker_([this] { run(); }) {}

    ~DynamicBatcher() {
        {
            std::lock_guard<std::mutex> lk(mu_);
            stop_ = true;
        }
        cv_.notify_all();
        worker_.join();
    }

    // thread-safe; blocks the calling connection thread until its embedding is ready
    std::vector<float> embed(std::string text) {
        auto req = std::make_shared<Request>();
        req->text = std::move(text);
        req->out.resize((size_t)n_embd_);
        std::future<void> done = req->done.get_future();
        {
            std::lock_guard<std::mutex> lk(mu_);
            queue_.push_back(req);
        }
        cv_.notify_all();
        done.get();
        return std::move(req->out);
    }

    // statistics (for logs and tests)
    uint64_t batches() const { return n_batches_.load(); }
    uint64_t requests() const { return n_requests_.load(); }
    uint64_t largest_batch() const { return largest_.load(); }

private:
    struct Request {
        std::string text;
        std::vector<float> out;
        std::promise<void> done;
    };

    void run() {
        std::vector<std::shared_ptr<Request>> batch;
        for (;;) {
            batch.clear();
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_.wait(lk, [this] { return stop_ || !queue_.empty(); });
                if (queue_.empty()) return;  // stop requested and nothing left to serve
                if (window_us_ > 0 && (int)queue_.size() < max_batch_) {
                    // idle server: give requests that are about to arrive a chance to share this launch
                    const auto deadline = std::chrono::steady_clock::now() + std::chrono::microseconds(window_us_);
                    cv_.wait_until(lk, deadline, [this] { return stop_ || (int)queue_.size() >= max_batch_; });
                }
                const size_t n = queue_.size() < (size_t)max_batch_ ? queue_.size() : (size_t)max_batch_;
                batch.assign(queue_.begin(), queue_.begin() + (long)n);
                queue_.erase(queue_.begin(), queue_.begin() + (long)n);
            }
            std::vector<const char *> texts(batch.size());
            std::vector<float *> outs(batch.size());
            for (size_t i = 0; i < batch.size(); i++) { texts[i] = batch[i]->text.c_str(); outs[i] = batch[i]->out.data(); }
            n_batches_++;
            n_requests_ += batch.size();
            if (batch.size() > largest_.load()) largest_ = batch.size();
            try {
                backend_(texts, outs);
                for (auto &r : batch) r->done.set_value();
            } catch (...) {
                for (auto &r : batch) r->done.set_exception(std::current_exception());
            }
        }
    }

    Backend backend_;
    const int n_embd_, max_batch_, window_us_;
    std::mutex mu_;
    std::condition_variable cv_;
    std::vector<std::shared_ptr<Request>> queue_;
    bool stop_ = false;
    std::atomic<uint64_t> n_batches_{0}, n_requests_{0}, largest_{0};
    std::thread worker_;  // last member: starts after everything above is initialised
};

}  // namespace bk

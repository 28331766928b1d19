// Host-side unit test of embedding.cpp_b200/csrc/server_batcher.h with a stub backend (no GPU, no model):
// many "connection" threads submit texts concurrently; every caller must get the embedding of ITS text, requests must
// have been merged into fewer backend calls than requests, and no call may exceed max_batch.
#include <chrono>
#include <cstdio>
#include <cstring>
#include <functional>
#include <thread>

#include "../../embedding.cpp_b200/csrc/server_batcher.h"

static float text_hash(const char *s) {
    uint32_t h = 2166136261u;
    for (; *s; s++) h = (h ^ (uint8_t)*s) * 16777619u;
    return (float)(h % 100003);
}

int main(int argc, char **argv) {
    const int n_threads = argc > 1 ? atoi(argv[1]) : 48, per_thread = argc > 2 ? atoi(argv[2]) : 25;
    const int max_batch = argc > 3 ? atoi(argv[3]) : 16, window_us = argc > 4 ? atoi(argv[4]) : 200;
    const int n_embd = 8;
    std::atomic<int> over{0};
    bk::DynamicBatcher batcher(
        [&](const std::vector<const char *> &texts, const std::vector<float *> &outs) {
            if ((int)texts.size() > max_batch || texts.empty()) over++;
            std::this_thread::sleep_for(std::chrono::milliseconds(2));  // "the GPU is busy": requests pile up meanwhile
            for (size_t i = 0; i < texts.size(); i++) {
                for (int k = 0; k < n_embd; k++) outs[i][k] = 0.f;
                outs[i][0] = text_hash(texts[i]);
                outs[i][1] = (float)texts.size();
            }
        },
        n_embd, max_batch, window_us);
    std::atomic<int> wrong{0};
    std::vector<std::thread> th;
    for (int t = 0; t < n_threads; t++)
        th.emplace_back([&, t] {
            for (int i = 0; i < per_thread; i++) {
                char text[64];
                snprintf(text, sizeof text, "connection %d request %d", t, i);
                const std::vector<float> e = batcher.embed(text);
                if ((int)e.size() != n_embd || e[0] != text_hash(text) || e[1] < 1.f || e[1] > (float)max_batch) wrong++;
            }
        });
    for (auto &x : th) x.join();
    printf("requests %llu batches %llu largest %llu wrong %d over %d\n", (unsigned long long)batcher.requests(),
           (unsigned long long)batcher.batches(), (unsigned long long)batcher.largest_batch(), wrong.load(), over.load());
    return wrong.load() || over.load() ? 1 : 0;
}

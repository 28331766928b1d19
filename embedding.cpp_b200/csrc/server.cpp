// server.cpp -- build/bert-server: the reference's TCP embedding server (examples/server.cpp) on the B200 engine.
//
// Wire protocol unchanged (examples/server.cpp:100-118, examples/sample_client.py): on connect the server sends
// bert_n_embd() as a 4-byte int; then every message the client sends (one read() of up to 32 KiB = one text) is
// answered with n_embd raw floats; the client closes the connection when done.  Same command line
// (bert_params_parse: -m model, --port, -t).  What changes: connections are served concurrently (one thread each,
// the reference handles one client at a time) and their requests are batched into bert_encode_batch calls by
// server_batcher.h.  Tuning through the environment: BERT_SERVER_MAX_BATCH (default 512 texts per launch),
// BERT_SERVER_WINDOW_US (default 200: how long an idle server waits for more requests before launching).
#include <arpa/inet.h>
#include <netinet/tcp.h>
#include <signal.h>
#include <sys/socket.h>
#include <unistd.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../include/bert.h"
#include "server_batcher.h"

static int env_int(const char *name, int dflt) {
    const char *v = getenv(name);
    return v ? atoi(v) : dflt;
}

static bool send_all(int fd, const void *buf, size_t n) {
    const char *p = static_cast<const char *>(buf);
    while (n > 0) {
        const ssize_t k = send(fd, p, n, MSG_NOSIGNAL);
        if (k <= 0) return false;
        p += k;
        n -= (size_t)k;
    }
    return true;
}

static void serve_connection(int fd, bk::DynamicBatcher *batcher, int n_embd) {
    const int one = 1;
    setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
    if (send_all(fd, &n_embd, sizeof(int))) {       // examples/server.cpp:108
        std::vector<char> buf(1 << 15);             // the reference's receive buffer size (examples/server.cpp:26)
        for (;;) {
            const ssize_t n = read(fd, buf.data(), buf.size());
            if (n <= 0) break;                       // empty read: the client is done (examples/server.cpp:111-113)
            std::vector<float> emb;
            try {
                emb = batcher->embed(std::string(buf.data(), (size_t)n));
            } catch (...) {
                break;
            }
            if (!send_all(fd, emb.data(), emb.size() * sizeof(float))) break;
        }
    }
    close(fd);
}

int main(int argc, char **argv) {
    bert_params params;
    if (!bert_params_parse(argc, argv, params)) return 1;
    bert_ctx *ctx = bert_load_from_file(params.model);
    if (!ctx) {
        fprintf(stderr, "%s: failed to load model from '%s'\n", argv[0], params.model);
        return 1;
    }
    const int n_embd = bert_n_embd(ctx);
    const int n_threads = params.n_threads;
    bk::DynamicBatcher batcher(
        [ctx, n_threads](const std::vector<const char *> &texts, const std::vector<float *> &outs) {
            bert_encode_batch(ctx, n_threads, (int32_t)texts.size(), (int32_t)texts.size(), const_cast<const char **>(texts.data()),
                              const_cast<float **>(outs.data()));
        },
        n_embd, env_int("BERT_SERVER_MAX_BATCH", 512), env_int("BERT_SERVER_WINDOW_US", 200));

    const int server_fd = socket(AF_INET, SOCK_STREAM, 0);
    if (server_fd < 0) { perror("socket"); return 1; }
    const int one = 1;
    setsockopt(server_fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
    sockaddr_in address{};
    address.sin_family = AF_INET;
    address.sin_addr.s_addr = INADDR_ANY;
    address.sin_port = htons((uint16_t)params.port);
    if (bind(server_fd, reinterpret_cast<sockaddr *>(&address), sizeof address) < 0) { perror("bind"); return 1; }
    if (listen(server_fd, 128) < 0) { perror("listen"); return 1; }
    socklen_t alen = sizeof address;
    getsockname(server_fd, reinterpret_cast<sockaddr *>(&address), &alen);   // --port 0: report the port the kernel chose
    printf("Server running on port %d with dynamic batching (max %d texts per launch)\n", (int)ntohs(address.sin_port),
           env_int("BERT_SERVER_MAX_BATCH", 512));
    fflush(stdout);
    for (;;) {
        const int fd = accept(server_fd, nullptr, nullptr);
        if (fd < 0) {
            if (errno == EINTR) continue;
            perror("accept");
            break;
        }
        std::thread(serve_connection, fd, &batcher, n_embd).detach();
    }
    close(server_fd);
    bert_free(ctx);
    return 0;
}

// CPU ring all-reduce / all-gather over TCP for host tensors (part of libhetu_ps.so).
//
// reference: src/communication/c_communication_nthread.cc (ZeroMQ REQ/REP ring: DL_Connect_Init,
// DL_Communicate_Init, RingAllreduce_{Server,Client}[_nthread], RingAllgather_*, average, close) -
// the legacy CPU data-parallel path.  Same algorithm (reduce-scatter then all-gather around the
// ring, segments of len / n, result averaged); each step sends to the next rank and receives from
// the previous one concurrently on two threads, so no acknowledgement round-trips are needed and
// large segments cannot deadlock on socket buffers.
#include "net.h"

#include <algorithm>
#include <thread>
#include <vector>

#define HRING_API extern "C" __attribute__((visibility("default")))

namespace {

struct Ring {
    int n = 1, id = 0;
    int listen_fd = -1, next_fd = -1, prev_fd = -1;
    std::vector<float> recv_buf;
};
Ring g_ring;
thread_local std::string g_ring_err;

void seg(int64_t len, int n, int s, int64_t* b, int64_t* e) { *b = len * s / n; *e = len * (s + 1) / n; }

// one ring step: send [sb, se) of data to next while receiving `rn` floats from prev into recv_buf
void exchange(const float* send, int64_t sn, float* recv, int64_t rn) {
    std::string err;
    std::thread tx([&] { try { if (sn > 0) hps::write_all(g_ring.next_fd, send, (size_t)sn * sizeof(float)); } catch (const std::exception& e) { err = e.what(); } });
    bool ok = rn == 0 || hps::read_all(g_ring.prev_fd, recv, (size_t)rn * sizeof(float));
    tx.join();
    if (!err.empty()) throw std::runtime_error(err);
    if (!ok) throw std::runtime_error("ring neighbour closed the connection");
}

}  // namespace

#define HRING_TRY(...)                                   \
    try { __VA_ARGS__; return 0; }                       \
    catch (const std::exception& e) { g_ring_err = e.what(); return -1; }

HRING_API const char* hring_last_error() { return g_ring_err.c_str(); }

// Rank `id` of `n` listens on `listen_port`, connects to the next rank at next_host:next_port and
// accepts the previous rank.  Every rank must call this (order does not matter).
HRING_API int hring_connect_init(int n, int id, int listen_port, const char* next_host, int next_port) {
    HRING_TRY({
        g_ring.n = n; g_ring.id = id;
        if (n <= 1) return 0;
        g_ring.listen_fd = hps::listen_on(listen_port);
        std::string host(next_host);
        int fd = -1;
        std::string err;
        std::thread conn([&] { try { fd = hps::connect_to(host, next_port); } catch (const std::exception& e) { err = e.what(); } });
        g_ring.prev_fd = ::accept(g_ring.listen_fd, nullptr, nullptr);
        conn.join();
        if (!err.empty()) throw std::runtime_error(err);
        if (g_ring.prev_fd < 0) throw std::runtime_error("accept from the previous ring rank failed");
        int one = 1;
        setsockopt(g_ring.prev_fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
        g_ring.next_fd = fd;
    });
}

static void ring_allreduce(float* data, int64_t len, int average) {
    int n = g_ring.n, id = g_ring.id;
    if (n > 1) {
        g_ring.recv_buf.resize((size_t)(len / n + 2));
        float* rb = g_ring.recv_buf.data();
        // reduce-scatter: after n-1 steps rank id holds the full sum of segment (id + 1) % n
        for (int i = 0; i < n - 1; ++i) {
            int ss = ((id - i) % n + n) % n, rs = ((id - i - 1) % n + n) % n;
            int64_t sb, se, rbeg, rend;
            seg(len, n, ss, &sb, &se); seg(len, n, rs, &rbeg, &rend);
            exchange(data + sb, se - sb, rb, rend - rbeg);
#pragma omp parallel for schedule(static)
            for (int64_t k = 0; k < rend - rbeg; ++k) data[rbeg + k] += rb[k];
        }
        // all-gather of the reduced segments
        for (int i = 0; i < n - 1; ++i) {
            int ss = ((id + 1 - i) % n + n) % n, rs = ((id - i) % n + n) % n;
            int64_t sb, se, rbeg, rend;
            seg(len, n, ss, &sb, &se); seg(len, n, rs, &rbeg, &rend);
            exchange(data + sb, se - sb, data + rbeg, rend - rbeg);
        }
    }
    if (average && n > 1) {
        float inv = 1.f / (float)n;
#pragma omp parallel for schedule(static)
        for (int64_t k = 0; k < len; ++k) data[k] *= inv;
    }
}

// In-place sum (average != 0: mean) of `data[len]` over the ring.
HRING_API int hring_allreduce(float* data, int64_t len, int average) { HRING_TRY(ring_allreduce(data, len, average)); }

// `data[len]` with this rank's segment (len * id / n .. len * (id + 1) / n) filled in; on return
// every segment is present everywhere.
HRING_API int hring_allgather(float* data, int64_t len) {
    HRING_TRY({
        int n = g_ring.n, id = g_ring.id;
        for (int i = 0; i < n - 1; ++i) {
            int ss = ((id - i) % n + n) % n, rs = ((id - i - 1) % n + n) % n;
            int64_t sb, se, rbeg, rend;
            seg(len, n, ss, &sb, &se); seg(len, n, rs, &rbeg, &rend);
            exchange(data + sb, se - sb, data + rbeg, rend - rbeg);
        }
    });
}

HRING_API int hring_close() {
    for (int* fd : {&g_ring.next_fd, &g_ring.prev_fd, &g_ring.listen_fd})
        if (*fd >= 0) { ::close(*fd); *fd = -1; }
    g_ring.n = 1; g_ring.id = 0;
    return 0;
}

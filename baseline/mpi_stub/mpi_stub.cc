// Socket-backed implementation of the handful of MPI calls declared in mpi.h (byte payloads only).
// Full mesh: rank r listens on port base + r, connects to every lower rank and accepts from every higher
// one.  Point-to-point messages carry (tag, length); a receive that meets another tag first parks that
// message in a per-peer queue.  Collectives are built from the point-to-point layer through rank 0 / root.
#include "mpi.h"
#include <arpa/inet.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <sys/socket.h>
#include <unistd.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <string>
#include <vector>

namespace {
int g_rank = 0, g_size = 1;
bool g_up = false;
std::vector<int> g_fd;
struct Parked { int tag; std::vector<char> data; };
std::vector<std::deque<Parked>> g_parked;
const int kCollTag = -7777;

int env_int(const char *a, const char *b, int dflt) {
    const char *v = getenv(a);
    if (!v && b) v = getenv(b);
    return v ? atoi(v) : dflt;
}
void die(const char *what) { perror(what); fprintf(stderr, "mpi_stub: rank %d fatal: %s\n", g_rank, what); exit(1); }
void write_all(int fd, const void *p, size_t n) {
    const char *c = (const char *)p;
    while (n) { ssize_t k = ::send(fd, c, n, MSG_NOSIGNAL); if (k <= 0) die("send"); c += k; n -= k; }
}
void read_all(int fd, void *p, size_t n) {
    char *c = (char *)p;
    while (n) { ssize_t k = ::recv(fd, c, n, 0); if (k <= 0) die("recv"); c += k; n -= k; }
}
void p2p_send(int dst, int tag, const void *buf, size_t n) {
    long long hdr[2] = {tag, (long long)n};
    write_all(g_fd[dst], hdr, sizeof(hdr));
    if (n) write_all(g_fd[dst], buf, n);
}
void p2p_recv(int src, int tag, void *buf, size_t n) {
    auto &q = g_parked[src];
    for (auto it = q.begin(); it != q.end(); ++it)
        if (it->tag == tag) { memcpy(buf, it->data.data(), n < it->data.size() ? n : it->data.size()); q.erase(it); return; }
    for (;;) {
        long long hdr[2];
        read_all(g_fd[src], hdr, sizeof(hdr));
        if ((int)hdr[0] == tag) {
            std::vector<char> tmp;
            if ((size_t)hdr[1] <= n) { read_all(g_fd[src], buf, (size_t)hdr[1]); }
            else { tmp.resize(hdr[1]); read_all(g_fd[src], tmp.data(), tmp.size()); memcpy(buf, tmp.data(), n); }
            return;
        }
        Parked p; p.tag = (int)hdr[0]; p.data.resize(hdr[1]);
        if (hdr[1]) read_all(g_fd[src], p.data.data(), p.data.size());
        q.push_back(std::move(p));
    }
}
}  // namespace

extern "C" {
int MPI_Init(int *, char ***) {
    if (g_up) return MPI_SUCCESS;
    g_rank = env_int("RANK", "OMPI_COMM_WORLD_RANK", 0);
    g_size = env_int("WORLD_SIZE", "OMPI_COMM_WORLD_SIZE", 1);
    g_fd.assign(g_size, -1);
    g_parked.resize(g_size);
    g_up = true;
    if (g_size == 1) return MPI_SUCCESS;
    const char *addr = getenv("MASTER_ADDR");
    if (!addr) addr = "127.0.0.1";
    int base = env_int("MPISTUB_PORT", nullptr, env_int("MASTER_PORT", nullptr, 29500) + 1000);
    int lfd = socket(AF_INET, SOCK_STREAM, 0), one = 1;
    setsockopt(lfd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
    sockaddr_in sa{}; sa.sin_family = AF_INET; sa.sin_addr.s_addr = htonl(INADDR_LOOPBACK); sa.sin_port = htons(base + g_rank);
    if (strcmp(addr, "127.0.0.1") != 0 && strcmp(addr, "localhost") != 0) sa.sin_addr.s_addr = htonl(INADDR_ANY);
    if (bind(lfd, (sockaddr *)&sa, sizeof(sa)) != 0) die("bind");
    if (listen(lfd, g_size) != 0) die("listen");
    for (int peer = 0; peer < g_rank; ++peer) {          // connect downwards
        sockaddr_in pa{}; pa.sin_family = AF_INET; pa.sin_port = htons(base + peer);
        if (inet_pton(AF_INET, strcmp(addr, "localhost") == 0 ? "127.0.0.1" : addr, &pa.sin_addr) != 1) die("inet_pton");
        int fd = -1;
        for (int tries = 0; tries < 6000; ++tries) {
            fd = socket(AF_INET, SOCK_STREAM, 0);
            if (connect(fd, (sockaddr *)&pa, sizeof(pa)) == 0) break;
            close(fd); fd = -1; usleep(10000);
        }
        if (fd < 0) die("connect");
        setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
        int me = g_rank; write_all(fd, &me, sizeof(me));
        g_fd[peer] = fd;
    }
    for (int k = g_rank + 1; k < g_size; ++k) {          // accept from above
        int fd = accept(lfd, nullptr, nullptr);
        if (fd < 0) die("accept");
        setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
        int who = -1; read_all(fd, &who, sizeof(who));
        if (who <= g_rank || who >= g_size) die("bad hello");
        g_fd[who] = fd;
    }
    close(lfd);
    return MPI_SUCCESS;
}
int MPI_Finalize(void) {
    for (int &fd : g_fd) if (fd >= 0) { close(fd); fd = -1; }
    return MPI_SUCCESS;
}
int MPI_Comm_rank(MPI_Comm, int *rank) { *rank = g_rank; return MPI_SUCCESS; }
int MPI_Comm_size(MPI_Comm, int *size) { *size = g_size; return MPI_SUCCESS; }
int MPI_Send(const void *buf, int count, MPI_Datatype, int dest, int tag, MPI_Comm) {
    p2p_send(dest, tag, buf, (size_t)count); return MPI_SUCCESS;
}
int MPI_Recv(void *buf, int count, MPI_Datatype, int source, int tag, MPI_Comm, MPI_Status *st) {
    p2p_recv(source, tag, buf, (size_t)count);
    if (st) { st->MPI_SOURCE = source; st->MPI_TAG = tag; st->MPI_ERROR = 0; }
    return MPI_SUCCESS;
}
int MPI_Bcast(void *buf, int count, MPI_Datatype, int root, MPI_Comm) {
    if (g_size == 1) return MPI_SUCCESS;
    if (g_rank == root) { for (int r = 0; r < g_size; ++r) if (r != root) p2p_send(r, kCollTag, buf, (size_t)count); }
    else p2p_recv(root, kCollTag, buf, (size_t)count);
    return MPI_SUCCESS;
}
int MPI_Allgather(const void *sendbuf, int sendcount, MPI_Datatype, void *recvbuf, int recvcount, MPI_Datatype, MPI_Comm) {
    char *out = (char *)recvbuf;
    size_t n = (size_t)recvcount;
    if (sendbuf != MPI_IN_PLACE) memcpy(out + n * g_rank, sendbuf, (size_t)sendcount < n ? (size_t)sendcount : n);
    if (g_size == 1) return MPI_SUCCESS;
    if (g_rank == 0) {
        for (int r = 1; r < g_size; ++r) p2p_recv(r, kCollTag - 1, out + n * r, n);
        for (int r = 1; r < g_size; ++r) p2p_send(r, kCollTag - 2, out, n * g_size);
    } else {
        p2p_send(0, kCollTag - 1, out + n * g_rank, n);
        p2p_recv(0, kCollTag - 2, out, n * g_size);
    }
    return MPI_SUCCESS;
}
int MPI_Barrier(MPI_Comm c) { char z = 0; std::vector<char> all(g_size); return MPI_Allgather(&z, 1, MPI_BYTE, all.data(), 1, MPI_BYTE, c); }
}

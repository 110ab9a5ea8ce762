// Minimal blocking TCP transport + message framing for the parameter server
// (replaces ps-lite's ZeroMQ "Van": ps-lite/src/zmq_van.h, p3_van.h).  One message =
//   [u32 magic][u32 type][u64 key][i64 arg0][i64 arg1][u64 sig][u32 narrays] { [u64 nbytes][bytes] }*
// `sig` = (sender uid << 32 | sequence number); 0 when the sender does not use the resend protocol
// (ps-lite's Resender keys its duplicate filter on the same kind of signature, resender.h:15-148).
#pragma once
#include <arpa/inet.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <sys/socket.h>
#include <sys/time.h>
#include <unistd.h>

#include <cerrno>
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

namespace hps {

constexpr uint32_t kMagic = 0x48455455;  // "HETU"

struct Message {
    uint32_t type = 0;
    uint64_t key = 0;
    int64_t arg0 = 0, arg1 = 0;
    uint64_t sig = 0;
    std::vector<std::vector<char>> arrays;

    template <typename T> void add(const T* p, size_t n) {
        arrays.emplace_back(reinterpret_cast<const char*>(p), reinterpret_cast<const char*>(p) + n * sizeof(T));
    }
    void add_str(const std::string& s) { arrays.emplace_back(s.begin(), s.end()); }
    template <typename T> const T* data(size_t i) const { return reinterpret_cast<const T*>(arrays[i].data()); }
    template <typename T> T* mdata(size_t i) { return reinterpret_cast<T*>(arrays[i].data()); }
    template <typename T> size_t count(size_t i) const { return arrays[i].size() / sizeof(T); }
    std::string str(size_t i) const { return std::string(arrays[i].begin(), arrays[i].end()); }
    size_t bytes() const { size_t b = 44; for (auto& a : arrays) b += 8 + a.size(); return b; }
};

inline void write_all(int fd, const void* buf, size_t n) {
    const char* p = static_cast<const char*>(buf);
    while (n > 0) {
        ssize_t w = ::send(fd, p, n, MSG_NOSIGNAL);
        if (w < 0) { if (errno == EINTR) continue; throw std::runtime_error(std::string("send: ") + strerror(errno)); }
        p += w; n -= (size_t)w;
    }
}
inline bool read_all(int fd, void* buf, size_t n) {
    char* p = static_cast<char*>(buf);
    while (n > 0) {
        ssize_t r = ::recv(fd, p, n, 0);
        if (r == 0) return false;
        if (r < 0) { if (errno == EINTR) continue; return false; }
        p += r; n -= (size_t)r;
    }
    return true;
}
inline void send_msg(int fd, const Message& m) {
    char hdr[44];
    uint32_t na = (uint32_t)m.arrays.size();
    memcpy(hdr, &kMagic, 4); memcpy(hdr + 4, &m.type, 4); memcpy(hdr + 8, &m.key, 8);
    memcpy(hdr + 16, &m.arg0, 8); memcpy(hdr + 24, &m.arg1, 8); memcpy(hdr + 32, &m.sig, 8); memcpy(hdr + 40, &na, 4);
    write_all(fd, hdr, 44);
    for (auto& a : m.arrays) {
        uint64_t nb = a.size();
        write_all(fd, &nb, 8);
        if (nb) write_all(fd, a.data(), nb);
    }
}
inline bool recv_msg(int fd, Message* m) {
    char hdr[44];
    if (!read_all(fd, hdr, 44)) return false;
    uint32_t magic, na;
    memcpy(&magic, hdr, 4);
    if (magic != kMagic) return false;
    memcpy(&m->type, hdr + 4, 4); memcpy(&m->key, hdr + 8, 8);
    memcpy(&m->arg0, hdr + 16, 8); memcpy(&m->arg1, hdr + 24, 8); memcpy(&m->sig, hdr + 32, 8); memcpy(&na, hdr + 40, 4);
    // a peer is not trusted with our memory: bounded array count and sizes (HETU_PS_MAX_ARRAY_MB, default 16 GiB)
    static const uint64_t max_bytes = [] {
        const char* e = getenv("HETU_PS_MAX_ARRAY_MB");
        return (uint64_t)(e ? strtoull(e, nullptr, 10) : 16384ull) << 20;
    }();
    if (na > 16) return false;
    m->arrays.clear();
    m->arrays.resize(na);
    for (uint32_t i = 0; i < na; ++i) {
        uint64_t nb;
        if (!read_all(fd, &nb, 8)) return false;
        if (nb > max_bytes) return false;
        m->arrays[i].resize(nb);
        if (nb && !read_all(fd, m->arrays[i].data(), nb)) return false;
    }
    return true;
}
// receive timeout in milliseconds (0 = block forever); a timed-out recv makes read_all return false
inline void set_recv_timeout(int fd, int ms) {
    timeval tv{};
    tv.tv_sec = ms / 1000; tv.tv_usec = (ms % 1000) * 1000;
    setsockopt(fd, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof(tv));
}
inline int listen_on(int port, int* bound_port = nullptr) {
    int fd = ::socket(AF_INET, SOCK_STREAM, 0);
    int one = 1;
    setsockopt(fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
    sockaddr_in addr{};
    addr.sin_family = AF_INET;
    // Listen where the job lives: HETU_PS_BIND names the interface address; a job whose scheduler is on the loopback
    // address (single-node runs, the tests) is not reachable from outside at all; otherwise every interface.
    addr.sin_addr.s_addr = htonl(INADDR_ANY);
    const char* bind_to = getenv("HETU_PS_BIND");
    const char* root = getenv("DMLC_PS_ROOT_URI");
    if (bind_to != nullptr && bind_to[0]) {
        if (inet_pton(AF_INET, bind_to, &addr.sin_addr) != 1) { ::close(fd); throw std::runtime_error(std::string("HETU_PS_BIND: bad address ") + bind_to); }
    } else if (root != nullptr && (strcmp(root, "127.0.0.1") == 0 || strcmp(root, "localhost") == 0)) {
        addr.sin_addr.s_addr = htonl(INADDR_LOOPBACK);
    }
    addr.sin_port = htons((uint16_t)port);
    if (::bind(fd, (sockaddr*)&addr, sizeof(addr)) != 0) { ::close(fd); throw std::runtime_error(std::string("bind: ") + strerror(errno)); }
    if (::listen(fd, 256) != 0) { ::close(fd); throw std::runtime_error("listen failed"); }
    if (bound_port) {
        socklen_t len = sizeof(addr);
        getsockname(fd, (sockaddr*)&addr, &len);
        *bound_port = ntohs(addr.sin_port);
    }
    return fd;
}
inline int connect_to(const std::string& host, int port, int retries = 600) {
    for (int attempt = 0; attempt < retries; ++attempt) {
        addrinfo hints{}, *res = nullptr;
        hints.ai_family = AF_INET; hints.ai_socktype = SOCK_STREAM;
        if (getaddrinfo(host.c_str(), std::to_string(port).c_str(), &hints, &res) == 0 && res) {
            int fd = ::socket(res->ai_family, res->ai_socktype, res->ai_protocol);
            if (fd >= 0 && ::connect(fd, res->ai_addr, res->ai_addrlen) == 0) {
                int one = 1;
                setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
                freeaddrinfo(res);
                return fd;
            }
            if (fd >= 0) ::close(fd);
            freeaddrinfo(res);
        }
        usleep(100 * 1000);
    }
    throw std::runtime_error("connect to " + host + ":" + std::to_string(port) + " failed");
}

}  // namespace hps

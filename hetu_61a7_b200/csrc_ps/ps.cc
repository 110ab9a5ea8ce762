// Parameter server: scheduler / server / worker roles in one library (libhetu_ps.so).
//
// reference: ps-lite/ (9.2 kLoC: ZeroMQ/IB vans, protobuf meta, Postoffice, KVWorker/KVServer,
// PSAgent partitioning, PS function set in include/ps/psf/PSFunc.h:33-58, server handlers in
// include/ps/server/PSFHandle.h and src/PSFhandle_embedding.cc, SSP in ssp_handler.h, partial
// reduce matchmaking in src/preduce_handler.cc, C API src/python_binding.cc:8-149).
//
// Same roles, same PS-function set and semantics ("server does +=", versioned cache rows,
// SSP, partial reduce, .dat save format <dir>/<node>_<server>.dat), new implementation: plain
// TCP with length-prefixed frames (net.h), one handler thread per connection, contiguous
// per-server shards (AveragePartitioner semantics), async requests on the worker through
// std::async with per-tensor Wait().  This is the host-DRAM tier; the B200 fast path for
// embeddings is the NVLink-sharded table (parallel/sharded_embedding.py).
//
// Fault handling (reference: ps-lite/src/van.cc:23-27,132-162,334-336,567-579, resender.h:15-148,
// postoffice.cc:201-221), same environment knobs:
//   PS_HEARTBEAT_INTERVAL=<s>  workers and servers beat to the scheduler on a side connection;
//   PS_HEARTBEAT_TIMEOUT=<s>   a node silent for longer is "dead" (hps_query_dead_nodes) and a node
//                              of the same role that registers late takes over its rank (recovery;
//                              parameter state is not restored - use LoadParam, as in the reference);
//   PS_RESEND=1, PS_RESEND_TIMEOUT=<ms>  worker->server requests carry a signature; after a timeout
//                              the worker reconnects and retransmits (up to 10 times), the server
//                              answers a retransmitted signature from its reply cache instead of
//                              applying the update twice;
//   PS_DROP_MSG=<percent>      fault injection: the server loses that share of data-plane messages,
//                              half before handling them (lost request), half after (lost reply).
#include "net.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <deque>
#include <functional>
#include <future>
#include <map>
#include <memory>
#include <mutex>
#include <random>
#include <shared_mutex>
#include <thread>
#include <unordered_map>

#define HPS_API extern "C" __attribute__((visibility("default")))

namespace hps {

enum PsfType : uint32_t {
    kDensePull = 0, kDensePush, kDDPushPull, kSparsePull, kSparsePush, kSDPushPull, kSSPushPull, kParamInit,
    kParamClear, kParamSave, kParamLoad, kSyncEmbedding, kPushEmbedding, kPushSyncEmbedding, kSSPInit, kSSPSync,
    kPReduceGetPartner,
    kRegister = 100, kBarrier, kTerminate, kHeartbeat, kOk, kError, kGetDeadNodes, kGetServers
};

// ------------------------------------------------------------------------------- server state
struct Param {
    int type = 0;                 // 0 dense, 1 2-D sparse-addressable, 2 cache table (versioned rows)
    size_t len = 0, width = 1;    // len = elements (dense) or rows (2-D)
    std::vector<float> data;
    std::vector<int64_t> ver;
    std::shared_mutex mu;
    // optional server-side optimizer (reference include/ps/server/optimizer.h:36-357): when set, pushed
    // values are gradients.  0 none (server does +=), 1 SGD, 2 Momentum, 3 Nesterov, 4 AdaGrad, 5 Adam.
    int opt = 0;
    double lr = 0, h0 = 0, h1 = 0, eps = 1e-7;
    std::vector<float> s1, s2;          // velocity / accumulator / m, v
    std::vector<int64_t> step;          // Adam step per element block (one per row for 2-D, one for dense)
    void apply(size_t off, const float* g, size_t n, size_t row) {
        float* w = data.data() + off;
        switch (opt) {
            case 0: for (size_t i = 0; i < n; ++i) w[i] += g[i]; break;
            case 1: for (size_t i = 0; i < n; ++i) w[i] -= (float)lr * g[i]; break;
            case 2: { float* v = s1.data() + off; for (size_t i = 0; i < n; ++i) { v[i] = (float)h0 * v[i] - (float)lr * g[i]; w[i] += v[i]; } break; }
            case 3: { float* v = s1.data() + off; for (size_t i = 0; i < n; ++i) { float nv = (float)h0 * v[i] - (float)lr * g[i]; w[i] += (float)h0 * nv - (float)lr * g[i]; v[i] = nv; } break; }
            case 4: { float* a = s1.data() + off; for (size_t i = 0; i < n; ++i) { a[i] += g[i] * g[i]; w[i] -= (float)lr * g[i] / (std::sqrt(a[i]) + (float)eps); } break; }
            case 5: {
                float* m = s1.data() + off; float* v = s2.data() + off;
                int64_t t = ++step[row];
                double c1 = 1.0 - std::pow(h0, (double)t), c2 = 1.0 - std::pow(h1, (double)t);
                for (size_t i = 0; i < n; ++i) {
                    m[i] = (float)h0 * m[i] + (1.f - (float)h0) * g[i];
                    v[i] = (float)h1 * v[i] + (1.f - (float)h1) * g[i] * g[i];
                    w[i] -= (float)(lr * (m[i] / c1) / (std::sqrt(v[i] / c2) + eps));
                }
                break;
            }
        }
    }
};

class Server {
  public:
    Server(int id, int nservers) : id_(id), nservers_(nservers) {}
    void set_id(int id) { id_ = id; }

    void handle(const Message& req, Message* rsp) {
        while (id_.load() < 0) usleep(1000);        // rank not yet assigned by the scheduler
        rsp->type = kOk;
        rsp->key = req.key;
        switch (req.type) {
            case kParamInit: param_init(req); break;
            case kParamClear: { std::unique_lock<std::mutex> l(map_mu_); params_.erase(req.key); break; }
            case kDensePull: dense_pull(req, rsp); break;
            case kDensePush: dense_push(req); break;
            case kDDPushPull: dense_push(req); dense_pull(req, rsp); break;
            case kSparsePull: sparse_pull(req, 0, rsp); break;
            case kSparsePush: sparse_push(req, 0, 1); break;
            case kSDPushPull: sparse_push(req, 0, 1); dense_pull(req, rsp); break;
            case kSSPushPull: sparse_push(req, 0, 1); sparse_pull(req, 2, rsp); break;
            case kParamSave: save(req); break;
            case kParamLoad: load(req); break;
            case kSyncEmbedding: sync_embedding(req, 0, rsp); break;
            case kPushEmbedding: push_embedding(req, 0); break;
            case kPushSyncEmbedding: push_embedding(req, 2); sync_embedding(req, 0, rsp); break;
            default: rsp->type = kError; rsp->add_str("unknown PS function");
        }
    }

  private:
    std::shared_ptr<Param> get(uint64_t key) {
        std::unique_lock<std::mutex> l(map_mu_);
        auto it = params_.find(key);
        if (it == params_.end()) throw std::runtime_error("parameter " + std::to_string(key) + " not initialised on server");
        return it->second;
    }
    // every handler checks what it is about to index: a malformed or mismatching request gets kError, not a wild read
    static void need(const Message& req, size_t i, size_t min_bytes, const char* what) {
        if (req.arrays.size() <= i || req.arrays[i].size() < min_bytes)
            throw std::runtime_error(std::string("malformed request: ") + what);
    }
    // arrays: [0] = {type, len, width, init_type, opt_type} as int64, [1] = {a, b} double, [2] = seed u64
    void param_init(const Message& req) {
        need(req, 0, 5 * sizeof(int64_t), "ParamInit header");
        need(req, 1, 2 * sizeof(double), "ParamInit range");
        need(req, 2, sizeof(uint64_t), "ParamInit seed");
        const int64_t* h = req.data<int64_t>(0);
        if (h[0] < 0 || h[0] > 2 || h[1] < 0 || h[2] < 0 || (h[0] != 0 && h[2] == 0) ||
            (double)h[1] * (double)(h[0] == 0 ? 1 : h[2]) > 1e12)
            throw std::runtime_error("malformed request: ParamInit shape");
        const double* ab = req.data<double>(1);
        uint64_t seed = *req.data<uint64_t>(2);
        auto p = std::make_shared<Param>();
        p->type = (int)h[0]; p->len = (size_t)h[1]; p->width = (size_t)h[2];
        size_t n = p->type == 0 ? p->len : p->len * p->width;
        p->data.resize(n);
        if (p->type == 2) p->ver.assign(p->len, 0);
        int init_type = (int)h[3];
        p->opt = (int)h[4];
        if (p->opt != 0 && req.arrays.size() > 3) {
            const double* o = req.data<double>(3);
            size_t no = req.count<double>(3);
            p->lr = no > 0 ? o[0] : 0; p->h0 = no > 1 ? o[1] : 0; p->h1 = no > 2 ? o[2] : 0;
            if (no > 3 && o[3] > 0) p->eps = o[3];
            if (p->opt >= 2) p->s1.assign(n, 0.f);
            if (p->opt == 5) { p->s2.assign(n, 0.f); p->step.assign(p->type == 0 ? 1 : p->len, 0); }
        }
        std::mt19937_64 rng(seed * 1000003ull + (uint64_t)id_.load());
        if (init_type == 0) {
            std::fill(p->data.begin(), p->data.end(), (float)ab[0]);
        } else if (init_type == 1) {
            std::uniform_real_distribution<float> d((float)ab[0], (float)ab[1]);
            for (auto& v : p->data) v = d(rng);
        } else {
            std::normal_distribution<float> d(0.f, 1.f);
            for (auto& v : p->data) {
                float z = d(rng);
                if (init_type == 3) while (z < -2.f || z > 2.f) z = d(rng);
                v = (float)ab[0] + (float)ab[1] * z;
            }
        }
        std::unique_lock<std::mutex> l(map_mu_);
        if (!params_.count(req.key)) params_[req.key] = p;      // first initialiser wins (all workers call it)
    }
    void dense_pull(const Message& req, Message* rsp) {
        auto p = get(req.key);
        std::shared_lock<std::shared_mutex> l(p->mu);
        rsp->add(p->data.data(), p->data.size());
    }
    void dense_push(const Message& req) {
        auto p = get(req.key);
        std::unique_lock<std::shared_mutex> l(p->mu);
        need(req, 0, 0, "dense push data");
        const float* g = req.data<float>(0);
        size_t n = std::min(p->data.size(), req.count<float>(0));
        p->apply(0, g, n, 0);
    }
    void sparse_pull(const Message& req, size_t off_idx, Message* rsp) {
        auto p = get(req.key);
        std::shared_lock<std::shared_mutex> l(p->mu);
        need(req, off_idx, 0, "sparse pull rows");
        const int64_t* rows = req.data<int64_t>(off_idx);
        size_t n = req.count<int64_t>(off_idx), w = p->width;
        std::vector<float> out(n * w);
        for (size_t i = 0; i < n; ++i) {
            int64_t r = rows[i];
            if (r >= 0 && (size_t)r < p->len) memcpy(&out[i * w], &p->data[(size_t)r * w], w * sizeof(float));
        }
        rsp->add(out.data(), out.size());
    }
    void sparse_push(const Message& req, size_t off_idx, size_t data_idx) {
        auto p = get(req.key);
        std::unique_lock<std::shared_mutex> l(p->mu);
        need(req, off_idx, 0, "sparse push rows");
        need(req, data_idx, 0, "sparse push data");
        const int64_t* rows = req.data<int64_t>(off_idx);
        const float* g = req.data<float>(data_idx);
        size_t n = req.count<int64_t>(off_idx), w = p->width;
        if (req.count<float>(data_idx) != n * w) throw std::runtime_error("malformed request: sparse push carries " +
            std::to_string(req.count<float>(data_idx)) + " floats for " + std::to_string(n) + " rows of width " + std::to_string(w));
        for (size_t i = 0; i < n; ++i) {
            int64_t r = rows[i];
            if (r < 0 || (size_t)r >= p->len) continue;
            p->apply((size_t)r * w, g + i * w, w, (size_t)r);
        }
    }
    // arrays: rows(i64), versions(i64); arg0 = bound.  response: idx(i64), versions(i64), data(f32)
    void sync_embedding(const Message& req, size_t base, Message* rsp) {
        auto p = get(req.key);
        std::shared_lock<std::shared_mutex> l(p->mu);
        need(req, base, 0, "sync rows");
        need(req, base + 1, 0, "sync versions");
        const int64_t* rows = req.data<int64_t>(base);
        const int64_t* cver = req.data<int64_t>(base + 1);
        size_t n = req.count<int64_t>(base), w = p->width;
        if (req.count<int64_t>(base + 1) != n) throw std::runtime_error("malformed request: sync rows / versions differ in length");
        int64_t bound = req.arg0;
        std::vector<int64_t> idx, ver;
        std::vector<float> data;
        for (size_t i = 0; i < n; ++i) {
            int64_t r = rows[i];
            if (r < 0 || (size_t)r >= p->len) continue;
            int64_t sv = p->ver.empty() ? 0 : p->ver[(size_t)r];
            if (cver[i] == -1 || sv - cver[i] > bound) {
                idx.push_back((int64_t)i);
                ver.push_back(sv);
                data.insert(data.end(), &p->data[(size_t)r * w], &p->data[(size_t)r * w] + w);
            }
        }
        rsp->add(idx.data(), idx.size());
        rsp->add(ver.data(), ver.size());
        rsp->add(data.data(), data.size());
    }
    // arrays at base: rows(i64), data(f32), updates(i64)
    void push_embedding(const Message& req, size_t base) {
        auto p = get(req.key);
        std::unique_lock<std::shared_mutex> l(p->mu);
        need(req, base, 0, "push rows");
        need(req, base + 1, 0, "push data");
        need(req, base + 2, 0, "push update counts");
        const int64_t* rows = req.data<int64_t>(base);
        const float* g = req.data<float>(base + 1);
        const int64_t* upd = req.data<int64_t>(base + 2);
        size_t n = req.count<int64_t>(base), w = p->width;
        if (req.count<float>(base + 1) != n * w || req.count<int64_t>(base + 2) != n)
            throw std::runtime_error("malformed request: embedding push arrays do not match " + std::to_string(n) + " rows of width " + std::to_string(w));
        for (size_t i = 0; i < n; ++i) {
            int64_t r = rows[i];
            if (r < 0 || (size_t)r >= p->len) continue;
            if (!p->ver.empty()) p->ver[(size_t)r] += upd[i];
            float* dst = &p->data[(size_t)r * w];
            for (size_t j = 0; j < w; ++j) dst[j] += g[i * w + j];
        }
    }
    // Checkpoint files go under the directory the request names - which must not climb (".."), and, when the operator
    // sets HETU_PS_CKPT_ROOT, must lie under that root.
    std::string path_of(const Message& req) const {
        need(req, 0, 1, "checkpoint directory");
        const std::string dir = req.str(0);
        if (dir.find("..") != std::string::npos || dir.find('\0') != std::string::npos)
            throw std::runtime_error("checkpoint directory may not contain '..': " + dir);
        const char* root = getenv("HETU_PS_CKPT_ROOT");
        if (root != nullptr && root[0] && dir.compare(0, strlen(root), root) != 0)
            throw std::runtime_error("checkpoint directory " + dir + " is outside HETU_PS_CKPT_ROOT");
        return dir + "/" + std::to_string(req.key) + "_" + std::to_string(id_.load()) + ".dat";
    }
    void save(const Message& req) {
        auto p = get(req.key);
        std::shared_lock<std::shared_mutex> l(p->mu);
        FILE* f = fopen(path_of(req).c_str(), "wb");
        if (!f) throw std::runtime_error("cannot open " + path_of(req));
        fwrite(p->data.data(), sizeof(float), p->data.size(), f);
        fclose(f);
    }
    void load(const Message& req) {
        auto p = get(req.key);
        std::unique_lock<std::shared_mutex> l(p->mu);
        FILE* f = fopen(path_of(req).c_str(), "rb");
        if (!f) throw std::runtime_error("cannot open " + path_of(req));
        size_t got = fread(p->data.data(), sizeof(float), p->data.size(), f);
        fclose(f);
        if (got != p->data.size()) throw std::runtime_error("short read from " + path_of(req));
    }

    std::atomic<int> id_;
    int nservers_;
    std::mutex map_mu_;
    std::unordered_map<uint64_t, std::shared_ptr<Param>> params_;
};

// ---------------------------------------------------------------------------- scheduler state
struct NodeInfo { int role; std::string host; int port; };

static const char* env_or(const char* k, const char* d) { const char* v = getenv(k); return v ? v : d; }
// ps-lite node ids: scheduler 1, server rank r -> 8 + 2r, worker rank r -> 9 + 2r
inline int node_id_of(int role, int rank) { return (role == 1 ? 8 : 9) + 2 * rank; }

class Scheduler {
  public:
    Scheduler(int nworkers, int nservers) : nworkers_(nworkers), nservers_(nservers) {}

    // register: arg0 = role (0 worker, 1 server), arg1 = port, array0 = host.  Blocks until the
    // whole cluster is present; response: arg0 = rank, arrays = server "host:port" strings.
    void handle(const Message& req, Message* rsp) {
        rsp->type = kOk;
        switch (req.type) {
            case kRegister: do_register(req, rsp); break;
            case kBarrier: barrier((int)req.arg0); break;
            case kSSPInit: ssp_init(req); break;
            case kSSPSync: ssp_sync(req, rsp); break;
            case kPReduceGetPartner: preduce(req, rsp); break;
            case kHeartbeat: { std::unique_lock<std::mutex> l(mu_); heartbeat_[req.arg0] = std::chrono::steady_clock::now(); break; }
            case kGetDeadNodes: {
                std::unique_lock<std::mutex> l(mu_);
                std::vector<int> dead = dead_nodes((double)req.arg0 / 1000.0);
                rsp->add(dead.data(), dead.size());
                break;
            }
            case kGetServers: {
                std::unique_lock<std::mutex> l(mu_);
                for (auto& sv : servers_) rsp->add_str(sv.host + ":" + std::to_string(sv.port));
                break;
            }
            default: rsp->type = kError; rsp->add_str("unknown scheduler request");
        }
    }

  private:
    // ids registered so far whose last heartbeat is older than `timeout_s` (caller holds mu_)
    std::vector<int> dead_nodes(double timeout_s) {
        std::vector<int> dead;
        auto now = std::chrono::steady_clock::now();
        for (auto& kv : heartbeat_)
            if (std::chrono::duration<double>(now - kv.second).count() > timeout_s) dead.push_back((int)kv.first);
        std::sort(dead.begin(), dead.end());
        return dead;
    }
    void do_register(const Message& req, Message* rsp) {
        std::unique_lock<std::mutex> l(mu_);
        int role = (int)req.arg0;
        int rank = -1;
        bool full = (int)servers_.size() == nservers_ && nworkers_seen_ >= nworkers_;
        if (full) {
            // late registration = a restarted node: give it the rank of a dead node of its role
            double timeout = atof(env_or("PS_HEARTBEAT_TIMEOUT", "0"));
            if (timeout <= 0) throw std::runtime_error("cluster is complete and recovery is off (PS_HEARTBEAT_TIMEOUT unset)");
            for (int id : dead_nodes(timeout))
                if ((id % 2 == 0) == (role == 1)) { rank = (id - (role == 1 ? 8 : 9)) / 2; break; }
            if (rank < 0) throw std::runtime_error("cluster is complete and no node of this role is dead");
            if (role == 1) servers_[(size_t)rank] = {role, req.str(0), (int)req.arg1};
            rsp->key = 1;                                        // is_recovery
        } else {
            if (role == 1) { rank = (int)servers_.size(); servers_.push_back({role, req.str(0), (int)req.arg1}); }
            else { rank = nworkers_seen_++; }
            cv_.notify_all();
            cv_.wait(l, [&] { return (int)servers_.size() == nservers_ && nworkers_seen_ >= nworkers_; });
        }
        heartbeat_[node_id_of(role, rank)] = std::chrono::steady_clock::now();
        rsp->arg0 = rank;
        rsp->arg1 = nworkers_;
        for (auto& s : servers_) rsp->add_str(s.host + ":" + std::to_string(s.port));
    }
    void barrier(int group_size) {
        std::unique_lock<std::mutex> l(mu_);
        int gen = barrier_gen_;
        if (++barrier_count_ >= (group_size > 0 ? group_size : nworkers_)) {
            barrier_count_ = 0;
            ++barrier_gen_;
            cv_.notify_all();
        } else {
            cv_.wait(l, [&] { return barrier_gen_ != gen; });
        }
    }
    struct Ssp { std::vector<int64_t> clock; int64_t tol = 0; };
    void ssp_init(const Message& req) {
        std::unique_lock<std::mutex> l(mu_);
        auto& s = ssp_[req.key];
        if (s.clock.empty()) { s.clock.assign((size_t)req.arg1, 0); s.tol = *req.data<int64_t>(0); }
    }
    // arg0 = rank, arg1 = version: blocks while this worker is more than `tol` clocks ahead of the slowest
    void ssp_sync(const Message& req, Message* rsp) {
        std::unique_lock<std::mutex> l(mu_);
        auto& s = ssp_[req.key];
        if (s.clock.empty()) return;
        s.clock[(size_t)req.arg0] = req.arg1;
        cv_.notify_all();
        cv_.wait(l, [&] { return req.arg1 - *std::min_element(s.clock.begin(), s.clock.end()) <= s.tol; });
        rsp->arg0 = 1;
    }
    struct PGroup { std::vector<int> ranks; bool closed = false; std::chrono::steady_clock::time_point deadline; };
    // arg0 = rank, arg1 = wanted group size, array0 = {wait_ms} float.  Whoever is ready inside the
    // window forms the group (heterogeneity-aware partial reduce).
    void preduce(const Message& req, Message* rsp) {
        std::unique_lock<std::mutex> l(mu_);
        float wait_ms = *req.data<float>(0);
        auto& cur = pgroups_[req.key];
        if (!cur || cur->closed) {
            cur = std::make_shared<PGroup>();
            cur->deadline = std::chrono::steady_clock::now() + std::chrono::microseconds((long long)(wait_ms * 1000));
        }
        auto g = cur;
        g->ranks.push_back((int)req.arg0);
        if ((int)g->ranks.size() >= (int)req.arg1) { g->closed = true; cv_.notify_all(); }
        while (!g->closed) {
            if (cv_.wait_until(l, g->deadline) == std::cv_status::timeout) { g->closed = true; cv_.notify_all(); }
        }
        std::vector<int> out = g->ranks;
        std::sort(out.begin(), out.end());
        rsp->add(out.data(), out.size());
    }

    int nworkers_, nservers_, nworkers_seen_ = 0, barrier_count_ = 0, barrier_gen_ = 0;
    std::mutex mu_;
    std::condition_variable cv_;
    std::vector<NodeInfo> servers_;
    std::unordered_map<uint64_t, Ssp> ssp_;
    std::unordered_map<uint64_t, std::shared_ptr<PGroup>> pgroups_;
    std::unordered_map<int64_t, std::chrono::steady_clock::time_point> heartbeat_;
};

// ------------------------------------------------------------------------------ accept loop
template <class H>
class Listener {
  public:
    Listener(std::shared_ptr<H> h, int port) : st_(std::make_shared<Shared>()) {
        st_->h = h;
        st_->drop_pct = atoi(env_or("PS_DROP_MSG", "0"));
        st_->fd = listen_on(port, &port_);
        th_ = std::thread([st = st_] { loop(st); });
    }
    ~Listener() { stop(); }
    int port() const { return port_; }
    void stop() {
        if (st_->stopped.exchange(true)) { if (th_.joinable()) th_.join(); return; }
        ::shutdown(st_->fd, SHUT_RDWR);
        ::close(st_->fd);
        if (th_.joinable()) th_.join();
    }
    void wait() { if (th_.joinable()) th_.join(); }

  private:
    // Everything a connection handler touches lives in this shared block, not in the Listener: handlers run
    // detached and may sit in a blocking request (barrier, registration) after the Listener is gone.
    struct Slot { uint32_t seq = 0; bool done = false; Message rsp; };
    struct Shared {
        std::shared_ptr<H> h;
        int fd = -1, drop_pct = 0;
        std::atomic<bool> stopped{false};
        std::mutex dd_mu;
        std::condition_variable dd_cv;
        std::unordered_map<uint32_t, Slot> slots;
        // Reply cache for the resend protocol: one slot per sender uid (a sender has one request in
        // flight).  replay() returns true with the cached reply when `sig` was already handled; when
        // the first copy is still being handled it waits for it.  Otherwise it claims the slot.
        bool replay(uint64_t sig, Message* rsp) {
            uint32_t uid = (uint32_t)(sig >> 32), seq = (uint32_t)sig;
            std::unique_lock<std::mutex> l(dd_mu);
            Slot& s = slots[uid];
            if (s.seq == seq) {
                dd_cv.wait(l, [&] { return slots[uid].done || slots[uid].seq != seq; });
                if (slots[uid].seq != seq) return false;
                *rsp = slots[uid].rsp;
                return true;
            }
            s.seq = seq; s.done = false; s.rsp = Message();
            return false;
        }
        void remember(uint64_t sig, const Message& rsp) {
            uint32_t uid = (uint32_t)(sig >> 32), seq = (uint32_t)sig;
            std::unique_lock<std::mutex> l(dd_mu);
            Slot& s = slots[uid];
            if (s.seq == seq) { s.rsp = rsp; s.done = true; }
            dd_cv.notify_all();
        }
    };
    static void loop(std::shared_ptr<Shared> st) {
        while (!st->stopped) {
            int c = ::accept(st->fd, nullptr, nullptr);
            if (c < 0) { if (st->stopped) break; continue; }
            int one = 1;
            setsockopt(c, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
            std::thread([st, c] {
                Message req, rsp;
                std::minstd_rand rng((unsigned)c * 2654435761u + 12345u);
                while (recv_msg(c, &req)) {
                    rsp = Message();
                    if (req.type == kTerminate) {
                        rsp.type = kOk; send_msg(c, rsp);
                        if (!st->stopped.exchange(true)) ::shutdown(st->fd, SHUT_RDWR);      // wait() returns; the owner closes the socket
                        break;
                    }
                    int fate = 0;                        // 1 = lose the request, 2 = lose the reply
                    if (st->drop_pct > 0 && req.type < kRegister && (int)(rng() % 100) < st->drop_pct) fate = 1 + (int)(rng() % 2);
                    if (fate == 1) continue;
                    if (req.sig == 0 || !st->replay(req.sig, &rsp)) {
                        try { st->h->handle(req, &rsp); }
                        catch (const std::exception& e) { rsp = Message(); rsp.type = kError; rsp.add_str(e.what()); }
                        if (req.sig != 0) st->remember(req.sig, rsp);
                    }
                    if (fate == 2) continue;
                    try { send_msg(c, rsp); } catch (...) { break; }
                }
                ::close(c);
            }).detach();
        }
    }
    std::shared_ptr<Shared> st_;
    int port_ = 0;
    std::thread th_;
};

// ------------------------------------------------------------------------------------ worker
struct Conn {
    int fd = -1;
    std::mutex mu;
    // resend protocol (PS_RESEND=1): only on worker->server connections, whose handlers never block
    bool resend = false;
    int timeout_ms = 1000;
    uint32_t uid = 0, seq = 0;
    std::string host;
    int port = 0;
    std::function<bool(Conn*)> relocate;          // asks the scheduler where this server lives now
    std::atomic<uint64_t>* retransmits = nullptr;

    static Message& check(Message& rsp) {
        if (rsp.type == kError) throw std::runtime_error("PS error: " + (rsp.arrays.empty() ? std::string("?") : rsp.str(0)));
        return rsp;
    }
    void open(const std::string& h, int p, int retries = 600) {
        host = h; port = p;
        fd = connect_to(h, p, retries);
        if (resend) set_recv_timeout(fd, timeout_ms);
    }
    Message call(Message& req) {
        std::unique_lock<std::mutex> l(mu);
        Message rsp;
        if (!resend) {
            send_msg(fd, req);
            if (!recv_msg(fd, &rsp)) throw std::runtime_error("connection to PS lost");
            return std::move(check(rsp));
        }
        if (++seq == 0) ++seq;
        req.sig = ((uint64_t)uid << 32) | seq;
        for (int attempt = 0; attempt <= 10; ++attempt) {
            if (attempt > 0 && retransmits) ++*retransmits;
            try {
                if (fd >= 0) { send_msg(fd, req); if (recv_msg(fd, &rsp)) return std::move(check(rsp)); }
            } catch (const std::runtime_error& e) {
                if (strncmp(e.what(), "PS error", 8) == 0) throw;
            }
            // timed out, or the peer went away (possibly mid-frame): start over on a fresh connection
            if (fd >= 0) ::close(fd);
            fd = -1;
            try { open(host, port, 5); }
            catch (...) { if (relocate && relocate(this)) { try { open(host, port, 20); } catch (...) {} } }
        }
        throw std::runtime_error("PS request not acknowledged after 10 retransmissions");
    }
};

struct TensorMeta { int type; size_t len, width; };

// Beat to the scheduler every PS_HEARTBEAT_INTERVAL seconds on a connection of its own (the main
// scheduler connection may sit in a blocking barrier).  Returns an empty thread when disabled.
inline std::thread start_heartbeat(const std::string& host, int port, int node_id, std::atomic<bool>* stop) {
    double interval = atof(env_or("PS_HEARTBEAT_INTERVAL", "0"));
    if (interval <= 0) return std::thread();
    return std::thread([=] {
        int fd = -1;
        try { fd = connect_to(host, port, 50); } catch (...) { return; }
        auto us = (long long)(interval * 1e6);
        while (!*stop) {
            Message m, r;
            m.type = kHeartbeat; m.arg0 = node_id;
            try { send_msg(fd, m); if (!recv_msg(fd, &r)) break; } catch (...) { break; }
            for (long long t = 0; t < us && !*stop; t += 10000) usleep(10000);
        }
        ::close(fd);
    });
}

// Worker-side I/O pool for the *_async calls.  reference: ps-lite/src/worker.cc:25-34 + thread_pool.cc:4 (5 threads)
// and the P3 van's priority queue (ps-lite/src/p3_van.h:12-68).  Requests of one key run in submission order
// (a pull queued after a push of the same tensor sees that push); across keys the highest priority goes first
// (hps_set_priority; default 0 = first come first served).  PS_WORKER_THREADS sets the pool size.
class IoPool {
  public:
    ~IoPool() { shutdown(); }
    std::future<void> submit(uint64_t key, std::function<void()> fn) {
        std::unique_lock<std::mutex> l(mu_);
        if (threads_.empty()) {
            int n = std::max(1, atoi(env_or("PS_WORKER_THREADS", "5")));
            for (int i = 0; i < n; ++i) threads_.emplace_back([this] { loop(); });
        }
        auto t = std::make_unique<Task>();
        t->fn = std::move(fn);
        t->seq = ++seq_;
        std::future<void> fut = t->done.get_future();
        keys_[key].q.push_back(std::move(t));
        cv_.notify_one();
        return fut;
    }
    void set_priority(uint64_t key, int prio) { std::unique_lock<std::mutex> l(mu_); keys_[key].prio = prio; }
    void shutdown() {
        { std::unique_lock<std::mutex> l(mu_); stop_ = true; cv_.notify_all(); }
        for (auto& t : threads_) if (t.joinable()) t.join();
        threads_.clear();
    }

  private:
    struct Task { std::function<void()> fn; std::promise<void> done; uint64_t seq = 0; };
    struct KeyQ { std::deque<std::unique_ptr<Task>> q; bool running = false; int prio = 0; };
    KeyQ* pick() {
        KeyQ* best = nullptr;
        for (auto& kv : keys_) {
            KeyQ& k = kv.second;
            if (k.running || k.q.empty()) continue;
            if (!best || k.prio > best->prio || (k.prio == best->prio && k.q.front()->seq < best->q.front()->seq)) best = &k;
        }
        return best;
    }
    void loop() {
        std::unique_lock<std::mutex> l(mu_);
        for (;;) {
            KeyQ* k = nullptr;
            cv_.wait(l, [&] { return (k = pick()) != nullptr || stop_; });
            if (!k) return;                                     // stopping and nothing runnable
            std::unique_ptr<Task> t = std::move(k->q.front());
            k->q.pop_front();
            k->running = true;
            l.unlock();
            try { t->fn(); t->done.set_value(); }
            catch (...) { t->done.set_exception(std::current_exception()); }
            l.lock();
            k->running = false;
            cv_.notify_all();
        }
    }
    std::mutex mu_;
    std::condition_variable cv_;
    std::unordered_map<uint64_t, KeyQ> keys_;
    std::vector<std::thread> threads_;
    uint64_t seq_ = 0;
    bool stop_ = false;
};

class Worker {
  public:
    Worker(const std::string& sched_host, int sched_port) : sched_host_(sched_host), sched_port_(sched_port) {
        sched_.open(sched_host, sched_port);
        Message req;
        req.type = kRegister; req.arg0 = 0; req.arg1 = 0; req.add_str("worker");
        Message rsp = sched_.call(req);
        rank_ = (int)rsp.arg0;
        nrank_ = (int)rsp.arg1;
        is_recovery_ = rsp.key == 1;
        bool resend = atoi(env_or("PS_RESEND", "0")) != 0;
        int timeout_ms = std::max(1, atoi(env_or("PS_RESEND_TIMEOUT", "1000")));
        std::random_device rd;
        for (size_t i = 0; i < rsp.arrays.size(); ++i) {
            std::string hp = rsp.str(i);
            size_t c = hp.rfind(':');
            auto conn = std::make_unique<Conn>();
            conn->resend = resend; conn->timeout_ms = timeout_ms; conn->uid = (uint32_t)rd() | 1u;
            conn->retransmits = &retransmits_;
            conn->relocate = [this, i](Conn* cn) {
                Message q; q.type = kGetServers;
                Message a = sched_.call(q);
                if (i >= a.arrays.size()) return false;
                std::string now = a.str(i);
                size_t k = now.rfind(':');
                std::string h = now.substr(0, k); int p = atoi(now.substr(k + 1).c_str());
                if (h == cn->host && p == cn->port) return false;
                cn->host = h; cn->port = p;
                return true;
            };
            conn->open(hp.substr(0, c), atoi(hp.substr(c + 1).c_str()));
            servers_.push_back(std::move(conn));
        }
        hb_ = start_heartbeat(sched_host, sched_port, node_id_of(0, rank_), &hb_stop_);
    }
    ~Worker() {
        pool_.shutdown();
        hb_stop_ = true;
        if (hb_.joinable()) hb_.join();
        for (auto& s : servers_) if (s->fd >= 0) ::close(s->fd);
        ::close(sched_.fd);
    }
    bool is_recovery() const { return is_recovery_; }
    std::atomic<uint64_t> retransmits_{0};
    int rank() const { return rank_; }
    int nrank() const { return nrank_; }
    int nservers() const { return (int)servers_.size(); }

    // contiguous shard of `total` items owned by server s (ps-lite AveragePartitioner)
    void shard(size_t total, int s, size_t* beg, size_t* end) const {
        size_t S = servers_.size();
        *beg = total * (size_t)s / S;
        *end = total * (size_t)(s + 1) / S;
    }
    void init_tensor(uint64_t key, int type, size_t len, size_t width, int init_type, double a, double b, uint64_t seed,
                     int opt_type = 0, const double* opt_args = nullptr, int n_opt = 0) {
        { std::unique_lock<std::mutex> l(mu_); meta_[key] = {type, len, width}; }
        for (int s = 0; s < nservers(); ++s) {
            size_t beg, end;
            shard(len, s, &beg, &end);
            Message req;
            req.type = kParamInit; req.key = key;
            int64_t h[5] = {type, (int64_t)(end - beg), (int64_t)width, init_type, opt_type};
            double ab[2] = {a, b};
            req.add(h, 5); req.add(ab, 2); req.add(&seed, 1);
            if (opt_type != 0 && opt_args && n_opt > 0) req.add(opt_args, (size_t)n_opt);
            count(req, servers_[s]->call(req));
        }
        barrier(0);
    }
    TensorMeta meta(uint64_t key) {
        std::unique_lock<std::mutex> l(mu_);
        auto it = meta_.find(key);
        if (it == meta_.end()) throw std::runtime_error("tensor " + std::to_string(key) + " was not InitTensor'ed");
        return it->second;
    }
    void dense(uint32_t type, uint64_t key, const float* push, float* pull) {
        TensorMeta m = meta(key);
        size_t total = m.type == 0 ? m.len : m.len;             // shards are by element (dense) / row (2-D)
        size_t unit = m.type == 0 ? 1 : m.width;
        std::vector<std::future<void>> fs;
        for (int s = 0; s < nservers(); ++s)
            fs.push_back(std::async(std::launch::async, [=] {
                size_t beg, end;
                shard(total, s, &beg, &end);
                Message req;
                req.type = type; req.key = key;
                if (push) req.add(push + beg * unit, (end - beg) * unit);
                Message rsp = servers_[s]->call(req);
                count(req, rsp);
                if (pull && !rsp.arrays.empty()) memcpy(pull + beg * unit, rsp.arrays[0].data(), rsp.arrays[0].size());
            }));
        for (auto& f : fs) f.get();
    }
    // rows are global ids; grads [n, width] (may be null); out [n_pull, width] (may be null)
    void sparse(uint32_t type, uint64_t key, const int64_t* rows, size_t n, const float* grads, const int64_t* pull_rows,
                size_t n_pull, float* out, float* dense_out) {
        TensorMeta m = meta(key);
        size_t w = m.width;
        std::vector<std::future<void>> fs;
        for (int s = 0; s < nservers(); ++s)
            fs.push_back(std::async(std::launch::async, [=] {
                size_t beg, end;
                shard(m.len, s, &beg, &end);
                std::vector<int64_t> local, pos, plocal, ppos;
                std::vector<float> g;
                for (size_t i = 0; i < n; ++i)
                    if (rows[i] >= (int64_t)beg && rows[i] < (int64_t)end) {
                        local.push_back(rows[i] - (int64_t)beg);
                        pos.push_back((int64_t)i);
                        if (grads) g.insert(g.end(), grads + i * w, grads + (i + 1) * w);
                    }
                const int64_t* pr = pull_rows ? pull_rows : rows;
                size_t pn = pull_rows ? n_pull : n;
                if (type == kSSPushPull)
                    for (size_t i = 0; i < pn; ++i)
                        if (pr[i] >= (int64_t)beg && pr[i] < (int64_t)end) { plocal.push_back(pr[i] - (int64_t)beg); ppos.push_back((int64_t)i); }
                Message req;
                req.type = type; req.key = key;
                req.add(local.data(), local.size());
                if (type == kSparsePush || type == kSDPushPull || type == kSSPushPull) req.add(g.data(), g.size());
                if (type == kSSPushPull) req.add(plocal.data(), plocal.size());
                if (local.empty() && plocal.empty() && type != kSDPushPull) return;
                Message rsp = servers_[s]->call(req);
                count(req, rsp);
                if (type == kSparsePull && out) {
                    const float* d = rsp.data<float>(0);
                    for (size_t i = 0; i < pos.size(); ++i) memcpy(out + (size_t)pos[i] * w, d + i * w, w * sizeof(float));
                } else if (type == kSSPushPull && out) {
                    const float* d = rsp.data<float>(0);
                    for (size_t i = 0; i < ppos.size(); ++i) memcpy(out + (size_t)ppos[i] * w, d + i * w, w * sizeof(float));
                } else if (type == kSDPushPull && dense_out) {
                    memcpy(dense_out + beg * w, rsp.arrays[0].data(), rsp.arrays[0].size());
                }
            }));
        for (auto& f : fs) f.get();
    }
    // HET cache RPCs: returns (idx into request, versions, data) merged over servers
    void sync_embedding(uint32_t type, uint64_t key, const int64_t* rows, const int64_t* vers, size_t n, int64_t bound,
                        const int64_t* prow, const float* pdata, const int64_t* pupd, size_t pn,
                        std::vector<int64_t>* idx, std::vector<int64_t>* over, std::vector<float>* odata) {
        TensorMeta m = meta(key);
        size_t w = m.width;
        std::mutex out_mu;
        std::vector<std::future<void>> fs;
        for (int s = 0; s < nservers(); ++s)
            fs.push_back(std::async(std::launch::async, [&, s] {
                size_t beg, end;
                shard(m.len, s, &beg, &end);
                std::vector<int64_t> local, lver, pos, pl, pu;
                std::vector<float> pd;
                for (size_t i = 0; i < n; ++i)
                    if (rows[i] >= (int64_t)beg && rows[i] < (int64_t)end) { local.push_back(rows[i] - (int64_t)beg); lver.push_back(vers[i]); pos.push_back((int64_t)i); }
                for (size_t i = 0; i < pn; ++i)
                    if (prow[i] >= (int64_t)beg && prow[i] < (int64_t)end) {
                        pl.push_back(prow[i] - (int64_t)beg); pu.push_back(pupd[i]);
                        pd.insert(pd.end(), pdata + i * w, pdata + (i + 1) * w);
                    }
                if (local.empty() && pl.empty()) return;
                Message req;
                req.type = type; req.key = key; req.arg0 = bound;
                if (type == kPushEmbedding) { req.add(pl.data(), pl.size()); req.add(pd.data(), pd.size()); req.add(pu.data(), pu.size()); }
                else { req.add(local.data(), local.size()); req.add(lver.data(), lver.size());
                       if (type == kPushSyncEmbedding) { req.add(pl.data(), pl.size()); req.add(pd.data(), pd.size()); req.add(pu.data(), pu.size()); } }
                Message rsp = servers_[s]->call(req);
                count(req, rsp);
                if (type == kPushEmbedding || !idx) return;
                std::unique_lock<std::mutex> l(out_mu);
                const int64_t* ri = rsp.data<int64_t>(0);
                size_t k = rsp.count<int64_t>(0);
                for (size_t i = 0; i < k; ++i) idx->push_back(pos[(size_t)ri[i]]);
                over->insert(over->end(), rsp.data<int64_t>(1), rsp.data<int64_t>(1) + k);
                odata->insert(odata->end(), rsp.data<float>(2), rsp.data<float>(2) + k * w);
            }));
        for (auto& f : fs) f.get();
    }
    void simple(uint32_t type, uint64_t key, const std::string& path) {
        for (int s = 0; s < nservers(); ++s) {
            Message req;
            req.type = type; req.key = key;
            if (!path.empty()) req.add_str(path);
            servers_[s]->call(req);
        }
    }
    void barrier(int group) { Message r; r.type = kBarrier; r.arg0 = group; sched_.call(r); }
    void ssp_init(uint64_t key, int group_size, int64_t tol) { Message r; r.type = kSSPInit; r.key = key; r.arg0 = rank_; r.arg1 = group_size; r.add(&tol, 1); sched_.call(r); }
    void ssp_sync(uint64_t key, int64_t version) { Message r; r.type = kSSPSync; r.key = key; r.arg0 = rank_; r.arg1 = version; sched_.call(r); }
    std::vector<int> preduce(uint64_t key, int want, float wait_ms) {
        Message r; r.type = kPReduceGetPartner; r.key = key; r.arg0 = rank_; r.arg1 = want; r.add(&wait_ms, 1);
        Message rsp = sched_.call(r);
        return std::vector<int>(rsp.data<int>(0), rsp.data<int>(0) + rsp.count<int>(0));
    }
    // async bookkeeping -----------------------------------------------------------------
    void submit(uint64_t key, std::function<void()> fn) {
        std::unique_lock<std::mutex> l(mu_);
        pending_[key].push_back(pool_.submit(key, std::move(fn)));
    }
    void set_priority(uint64_t key, int prio) { pool_.set_priority(key, prio); }
    void wait(uint64_t key) {
        std::vector<std::future<void>> fs;
        { std::unique_lock<std::mutex> l(mu_); fs.swap(pending_[key]); }
        for (auto& f : fs) f.get();
    }
    void count(const Message& req, const Message& rsp) {
        if (!record_) return;
        sent_ += req.bytes(); recv_ += rsp.bytes();
        if (req.type < kNumPsf) { sent_by_[req.type] += req.bytes(); recv_by_[req.type] += rsp.bytes(); calls_by_[req.type] += 1; }
    }
    static constexpr uint32_t kNumPsf = 17;                    // data-plane PS functions (PsfType 0..16)
    std::atomic<uint64_t> sent_by_[17] = {}, recv_by_[17] = {}, calls_by_[17] = {};
    std::atomic<bool> record_{false};
    std::atomic<uint64_t> sent_{0}, recv_{0};

  private:
    Conn sched_;
    std::vector<std::unique_ptr<Conn>> servers_;
    std::string sched_host_;
    int sched_port_ = 0;
    bool is_recovery_ = false;
    std::atomic<bool> hb_stop_{false};
    std::thread hb_;
    int rank_ = 0, nrank_ = 1;
    std::mutex mu_;
    std::unordered_map<uint64_t, TensorMeta> meta_;
    std::unordered_map<uint64_t, std::vector<std::future<void>>> pending_;
    IoPool pool_;
};

static std::unique_ptr<Worker> g_worker;
static std::unique_ptr<Listener<Scheduler>> g_sched;
static std::vector<std::unique_ptr<Listener<Server>>> g_servers;
static thread_local std::string g_err;
static std::atomic<bool> g_server_hb_stop{false};
static std::vector<std::thread> g_server_hb;

}  // namespace hps

using namespace hps;

#define HPS_TRY(body)                                  \
    try { body; return 0; }                            \
    catch (const std::exception& e) { g_err = e.what(); return -1; }

HPS_API const char* hps_last_error() { return g_err.c_str(); }

// ---- roles (env: DMLC_PS_ROOT_URI, DMLC_PS_ROOT_PORT, DMLC_NUM_WORKER, DMLC_NUM_SERVER) ----------
HPS_API int hps_start_scheduler(int port, int nworkers, int nservers) {
    HPS_TRY(g_sched = std::make_unique<Listener<Scheduler>>(std::make_shared<Scheduler>(nworkers, nservers), port));
}
HPS_API int hps_scheduler_port() { return g_sched ? g_sched->port() : -1; }
HPS_API int hps_stop_scheduler() { if (g_sched) g_sched->stop(); g_sched.reset(); return 0; }
HPS_API int hps_wait_scheduler() { if (g_sched) g_sched->wait(); return 0; }

// start a server in this process (listens on an ephemeral port, registers with the scheduler)
HPS_API int hps_start_server(const char* sched_host, int sched_port, const char* my_host, int nservers) {
    HPS_TRY({
        int fd = connect_to(sched_host, sched_port);
        // the listener must exist before registering so that workers can connect once released
        auto srv = std::make_shared<Server>(-1, nservers);
        auto lst = std::make_unique<Listener<Server>>(srv, 0);
        Message req;
        req.type = kRegister; req.arg0 = 1; req.arg1 = lst->port(); req.add_str(my_host);
        g_servers.push_back(std::move(lst));
        std::string sh = sched_host;
        g_server_hb_stop = false;
        g_server_hb.emplace_back([fd, req, srv, sh, sched_port] {
            int rank = -1;
            try { Message rsp; send_msg(fd, req); if (recv_msg(fd, &rsp) && rsp.type == kOk) rank = (int)rsp.arg0; } catch (...) {}
            ::close(fd);
            if (rank < 0) return;
            srv->set_id(rank);
            std::thread hb = start_heartbeat(sh, sched_port, node_id_of(1, rank), &g_server_hb_stop);
            if (hb.joinable()) hb.join();
        });
    });
}
HPS_API int hps_wait_servers() { for (auto& s : g_servers) s->wait(); return 0; }
HPS_API int hps_stop_servers() {
    g_server_hb_stop = true;
    for (auto& s : g_servers) s->stop();
    g_servers.clear();
    // a registration that never completed (scheduler gone) would block join forever: detach those
    for (auto& t : g_server_hb) if (t.joinable()) t.detach();
    g_server_hb.clear();
    return 0;
}

HPS_API int hps_worker_init(const char* sched_host, int sched_port) { HPS_TRY(g_worker = std::make_unique<Worker>(sched_host, sched_port)); }
HPS_API int hps_worker_finalize() { g_worker.reset(); return 0; }
HPS_API int hps_rank() { return g_worker ? g_worker->rank() : -1; }
HPS_API int hps_nrank() { return g_worker ? g_worker->nrank() : 0; }
HPS_API int hps_nservers() { return g_worker ? g_worker->nservers() : 0; }

// ---- PS functions (synchronous; *_async variants enqueue and are joined by hps_wait(key)) -------
HPS_API int hps_init_tensor(uint64_t key, int type, uint64_t len, uint64_t width, int init_type, double a, double b, uint64_t seed) {
    HPS_TRY(g_worker->init_tensor(key, type, len, width, init_type, a, b, seed));
}
// same, with a server-side optimizer: pushes are then gradients.  opt_type 1 SGD, 2 Momentum, 3 Nesterov,
// 4 AdaGrad, 5 Adam; opt_args = {lr, h0 (momentum / beta1), h1 (beta2), eps}
HPS_API int hps_init_tensor_opt(uint64_t key, int type, uint64_t len, uint64_t width, int init_type, double a, double b, uint64_t seed,
                                int opt_type, const double* opt_args, int n_opt) {
    HPS_TRY(g_worker->init_tensor(key, type, len, width, init_type, a, b, seed, opt_type, opt_args, n_opt));
}
HPS_API int hps_pull(uint64_t key, float* out) { HPS_TRY(g_worker->dense(kDensePull, key, nullptr, out)); }
HPS_API int hps_push(uint64_t key, const float* grad) { HPS_TRY(g_worker->dense(kDensePush, key, grad, nullptr)); }
HPS_API int hps_dd_pushpull(uint64_t key, const float* grad, float* out) { HPS_TRY(g_worker->dense(kDDPushPull, key, grad, out)); }
HPS_API int hps_sparse_pull(uint64_t key, const int64_t* rows, uint64_t n, float* out) {
    HPS_TRY(g_worker->sparse(kSparsePull, key, rows, n, nullptr, nullptr, 0, out, nullptr));
}
HPS_API int hps_sparse_push(uint64_t key, const int64_t* rows, uint64_t n, const float* grads) {
    HPS_TRY(g_worker->sparse(kSparsePush, key, rows, n, grads, nullptr, 0, nullptr, nullptr));
}
HPS_API int hps_sd_pushpull(uint64_t key, const int64_t* rows, uint64_t n, const float* grads, float* dense_out) {
    HPS_TRY(g_worker->sparse(kSDPushPull, key, rows, n, grads, nullptr, 0, nullptr, dense_out));
}
HPS_API int hps_ss_pushpull(uint64_t key, const int64_t* rows, uint64_t n, const float* grads, const int64_t* pull_rows,
                            uint64_t n_pull, float* out) {
    HPS_TRY(g_worker->sparse(kSSPushPull, key, rows, n, grads, pull_rows, n_pull, out, nullptr));
}
HPS_API int hps_push_async(uint64_t key, const float* grad) { HPS_TRY(g_worker->submit(key, [=] { g_worker->dense(kDensePush, key, grad, nullptr); })); }
HPS_API int hps_pull_async(uint64_t key, float* out) { HPS_TRY(g_worker->submit(key, [=] { g_worker->dense(kDensePull, key, nullptr, out); })); }
HPS_API int hps_dd_pushpull_async(uint64_t key, const float* grad, float* out) {
    HPS_TRY(g_worker->submit(key, [=] { g_worker->dense(kDDPushPull, key, grad, out); }));
}
HPS_API int hps_sparse_push_async(uint64_t key, const int64_t* rows, uint64_t n, const float* grads) {
    HPS_TRY(g_worker->submit(key, [=] { g_worker->sparse(kSparsePush, key, rows, n, grads, nullptr, 0, nullptr, nullptr); }));
}
HPS_API int hps_sparse_pull_async(uint64_t key, const int64_t* rows, uint64_t n, float* out) {
    HPS_TRY(g_worker->submit(key, [=] { g_worker->sparse(kSparsePull, key, rows, n, nullptr, nullptr, 0, out, nullptr); }));
}
HPS_API int hps_wait(uint64_t key) { HPS_TRY(g_worker->wait(key)); }
// higher priority keys are served first by the async I/O pool (P3-style: front layers' parameters before back layers')
HPS_API int hps_set_priority(uint64_t key, int prio) { HPS_TRY(g_worker->set_priority(key, prio)); }
HPS_API int hps_clear(uint64_t key) { HPS_TRY(g_worker->simple(kParamClear, key, "")); }
HPS_API int hps_save(uint64_t key, const char* dir) { HPS_TRY(g_worker->simple(kParamSave, key, dir)); }
HPS_API int hps_load(uint64_t key, const char* dir) { HPS_TRY(g_worker->simple(kParamLoad, key, dir)); }
HPS_API int hps_barrier(int group) { HPS_TRY(g_worker->barrier(group)); }
HPS_API int hps_ssp_init(uint64_t key, int group_size, int64_t tolerance) { HPS_TRY(g_worker->ssp_init(key, group_size, tolerance)); }
HPS_API int hps_ssp_sync(uint64_t key, int64_t version) { HPS_TRY(g_worker->ssp_sync(key, version)); }
HPS_API int hps_preduce_get_partner(uint64_t key, int want, float wait_ms, int* out, int* n_out) {
    HPS_TRY({ auto v = g_worker->preduce(key, want, wait_ms); *n_out = (int)v.size(); for (size_t i = 0; i < v.size(); ++i) out[i] = v[i]; });
}
HPS_API int hps_start_record() { if (g_worker) g_worker->record_ = true; return 0; }
// per PS function (index = PsfType 0..16): bytes sent, bytes received, number of requests; arrays of 17
HPS_API int hps_get_loads_by_function(uint64_t* sent, uint64_t* recv, uint64_t* calls) {
    if (!g_worker) return -1;
    for (int i = 0; i < 17; ++i) { sent[i] = g_worker->sent_by_[i]; recv[i] = g_worker->recv_by_[i]; calls[i] = g_worker->calls_by_[i]; }
    return 0;
}
HPS_API int hps_get_loads(uint64_t* sent, uint64_t* recv) { if (!g_worker) return -1; *sent = g_worker->sent_; *recv = g_worker->recv_; return 0; }

// ---- failure detection -------------------------------------------------------------------------
// Node ids (ps-lite convention: server r -> 8+2r, worker r -> 9+2r) that registered with the
// scheduler at host:port and have not sent a heartbeat for `timeout_s` seconds.  Callable from any
// process (opens its own connection).  reference: Postoffice::GetDeadNodes, postoffice.cc:201-221.
HPS_API int hps_query_dead_nodes(const char* sched_host, int sched_port, double timeout_s, int* out, int cap, int* n_out) {
    HPS_TRY({
        Conn c;
        c.open(sched_host, sched_port, 50);
        Message q; q.type = kGetDeadNodes; q.arg0 = (int64_t)(timeout_s * 1000.0);
        Message r = c.call(q);
        ::close(c.fd);
        size_t k = r.count<int>(0);
        *n_out = (int)k;
        for (size_t i = 0; i < k && (int)i < cap; ++i) out[i] = r.data<int>(0)[i];
    });
}
HPS_API int hps_is_recovery() { return g_worker && g_worker->is_recovery() ? 1 : 0; }
HPS_API uint64_t hps_num_retransmits() { return g_worker ? g_worker->retransmits_.load() : 0; }

// HET-cache RPCs for a cache that lives elsewhere (the GPU tier, parallel/gpu_cache.py): the staleness check / row fetch and
// the push of kept gradients, exactly the two requests of the host cache (cache.cc).
//   sync: rows[n] with the versions the caller holds (-1 = not cached); returns in out_* the k <= n rows that are missing or
//         more than `bound` versions behind: out_idx[k] = position in rows, out_ver[k], out_rows[k, width].
HPS_API int hps_sync_embedding(uint64_t key, const int64_t* rows, const int64_t* vers, uint64_t n, int64_t bound, int64_t* out_idx,
                               int64_t* out_ver, float* out_rows, uint64_t* n_out) {
    try {
        if (!g_worker) throw std::runtime_error("PS worker not initialised");
        std::vector<int64_t> idx, ver;
        std::vector<float> data;
        g_worker->sync_embedding(kSyncEmbedding, key, rows, vers, n, bound, nullptr, nullptr, nullptr, 0, &idx, &ver, &data);
        *n_out = idx.size();
        if (!idx.empty()) {
            memcpy(out_idx, idx.data(), idx.size() * sizeof(int64_t));
            memcpy(out_ver, ver.data(), ver.size() * sizeof(int64_t));
            memcpy(out_rows, data.data(), data.size() * sizeof(float));
        }
        return 0;
    } catch (const std::exception& e) { g_err = e.what(); return -1; }
}
HPS_API int hps_push_embedding(uint64_t key, const int64_t* rows, const float* grads, const int64_t* updates, uint64_t n) {
    try {
        if (!g_worker) throw std::runtime_error("PS worker not initialised");
        if (n == 0) return 0;
        g_worker->sync_embedding(kPushEmbedding, key, nullptr, nullptr, 0, 0, rows, grads, updates, n, nullptr, nullptr, nullptr);
        return 0;
    } catch (const std::exception& e) { g_err = e.what(); return -1; }
}

// HET-cache RPCs used by cache.cc ---------------------------------------------------------------
namespace hps {
void worker_sync_embedding(uint32_t type, uint64_t key, const int64_t* rows, const int64_t* vers, size_t n, int64_t bound,
                           const int64_t* prow, const float* pdata, const int64_t* pupd, size_t pn,
                           std::vector<int64_t>* idx, std::vector<int64_t>* over, std::vector<float>* odata) {
    if (!g_worker) throw std::runtime_error("PS worker not initialised");
    g_worker->sync_embedding(type, key, rows, vers, n, bound, prow, pdata, pupd, pn, idx, over, odata);
}
}  // namespace hps

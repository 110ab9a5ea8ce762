// Distributed graph store + mini-batch samplers for GNN training (libhetu_graph.so).
//
// reference: third_party/GraphMix (graphmix/src/graph/{graph_handle,remote_handle,sampler,
// graph_client}.cc, include/graph/*.h; Python client python/graphmix/client.py): the graph is
// partitioned over graph servers, each server owns a contiguous range of (renumbered) node ids
// with their float / int features and out-edges, runs sampler threads (LocalNode, GlobalNode,
// RandomWalk, GraphSage) that keep a queue of ready mini-batch sub-graphs, pulls the node data it
// does not own from its peers through a cache (LRU / LFU / LFUOpt), and hands batches to workers
// on request.
//
// Same roles and sampler semantics; transport is the parameter server's framed TCP (net.h)
// instead of a forked ps-lite, server 0 doubles as the barrier point (no scheduler process), and
// a sub-graph travels as one message of flat arrays.
#include "../csrc_ps/net.h"

#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <deque>
#include <future>
#include <list>
#include <map>
#include <memory>
#include <mutex>
#include <random>
#include <set>
#include <thread>
#include <unordered_map>
#include <unordered_set>

#define HGM_API extern "C" __attribute__((visibility("default")))

namespace hgm {

using hps::Message;

enum MsgType : uint32_t { kMeta = 1, kPullNode, kPullLocal, kPullGraph, kBarrier, kStop, kStats, kOk = 200, kError };
enum SamplerType : int { kGraphSage = 0, kRandomWalk = 1, kGlobalNode = 2, kLocalNode = 3, kNumSampler = 4 };
enum CachePolicy : int { kLRU = 0, kLFU = 1, kLFUOpt = 2 };

struct NodePack {
    std::vector<float> f;
    std::vector<int32_t> i;
    std::vector<int64_t> nbr;
};

// ---------------------------------------------------------------------------------- cache
// Remote node data cache.  LRU: recency list.  LFU: evict the least-used (oldest among ties).
// LFUOpt: LFU whose entries become permanent once used `kPromote` times (reference
// hetu_cache/src/lfuopt_cache.cc idea applied to node packs).
class NodeCache {
  public:
    NodeCache(size_t cap, int policy) : cap_(cap), policy_(policy) {}
    bool lookup(int64_t id, NodePack* out) {
        std::unique_lock<std::mutex> l(mu_);
        ++lookups_;
        auto it = map_.find(id);
        if (it == map_.end()) return false;
        ++hits_;
        touch(it);
        *out = it->second.pack;
        return true;
    }
    void insert(int64_t id, const NodePack& p) {
        std::unique_lock<std::mutex> l(mu_);
        if (cap_ == 0 || map_.count(id)) return;
        while (map_.size() - pinned_ >= cap_) { if (!evict()) return; }
        Entry e;
        e.pack = p; e.cnt = 1; e.tick = ++tick_;
        if (policy_ == kLRU) { lru_.push_front(id); e.it = lru_.begin(); }
        else order_.insert({e.cnt, e.tick, id});
        map_.emplace(id, std::move(e));
    }
    void stats(int64_t* lookups, int64_t* hits, int64_t* size) {
        std::unique_lock<std::mutex> l(mu_);
        *lookups = lookups_; *hits = hits_; *size = (int64_t)map_.size();
    }

  private:
    static constexpr uint64_t kPromote = 16;
    struct Entry { NodePack pack; uint64_t cnt = 0, tick = 0; bool pinned = false; std::list<int64_t>::iterator it; };
    using Map = std::unordered_map<int64_t, Entry>;
    void touch(Map::iterator it) {
        Entry& e = it->second;
        if (policy_ == kLRU) { lru_.erase(e.it); lru_.push_front(it->first); e.it = lru_.begin(); return; }
        if (e.pinned) return;
        order_.erase({e.cnt, e.tick, it->first});
        ++e.cnt;
        if (policy_ == kLFUOpt && e.cnt >= kPromote) { e.pinned = true; ++pinned_; return; }
        order_.insert({e.cnt, e.tick, it->first});
    }
    bool evict() {
        if (policy_ == kLRU) {
            if (lru_.empty()) return false;
            map_.erase(lru_.back()); lru_.pop_back();
            return true;
        }
        if (order_.empty()) return false;
        auto victim = *order_.begin();
        order_.erase(order_.begin());
        map_.erase(std::get<2>(victim));
        return true;
    }
    size_t cap_, pinned_ = 0;
    int policy_;
    std::mutex mu_;
    Map map_;
    std::list<int64_t> lru_;
    std::set<std::tuple<uint64_t, uint64_t, int64_t>> order_;
    uint64_t tick_ = 0;
    int64_t lookups_ = 0, hits_ = 0;
};

// ---------------------------------------------------------------------------- sampled graph
struct SubGraph {
    int type = 0;
    int64_t n = 0, f_len = 0, i_len = 0, extra_len = 0;
    std::vector<float> f;             // [n, f_len]
    std::vector<int32_t> i;           // [n, i_len + 1]; last column low 31 bits of the global id (full id in `gid`)
    std::vector<int64_t> gid;         // [n]
    std::vector<int64_t> edges;       // [2, E] local indices (row 0 = source, row 1 = destination)
    std::vector<int32_t> extra;       // [n, extra_len]
    void to_msg(Message* m) const {
        int64_t hdr[5] = {type, n, f_len, i_len, extra_len};
        m->add(hdr, 5); m->add(f.data(), f.size()); m->add(i.data(), i.size()); m->add(gid.data(), gid.size());
        m->add(edges.data(), edges.size()); m->add(extra.data(), extra.size());
    }
    void from_msg(const Message& m, size_t base = 0) {
        const int64_t* hdr = m.data<int64_t>(base);
        type = (int)hdr[0]; n = hdr[1]; f_len = hdr[2]; i_len = hdr[3]; extra_len = hdr[4];
        f.assign(m.data<float>(base + 1), m.data<float>(base + 1) + m.count<float>(base + 1));
        i.assign(m.data<int32_t>(base + 2), m.data<int32_t>(base + 2) + m.count<int32_t>(base + 2));
        gid.assign(m.data<int64_t>(base + 3), m.data<int64_t>(base + 3) + m.count<int64_t>(base + 3));
        edges.assign(m.data<int64_t>(base + 4), m.data<int64_t>(base + 4) + m.count<int64_t>(base + 4));
        extra.assign(m.data<int32_t>(base + 5), m.data<int32_t>(base + 5) + m.count<int32_t>(base + 5));
    }
};

struct Peer {
    std::string host;
    int port = 0, fd = -1;
    std::mutex mu;
    Message call(Message& req) {
        std::unique_lock<std::mutex> l(mu);
        if (fd < 0) fd = hps::connect_to(host, port);
        hps::send_msg(fd, req);
        Message rsp;
        if (!hps::recv_msg(fd, &rsp)) { ::close(fd); fd = -1; throw std::runtime_error("graph server " + host + ":" + std::to_string(port) + " went away"); }
        if (rsp.type == kError) throw std::runtime_error("graph server error: " + (rsp.arrays.empty() ? std::string("?") : rsp.str(0)));
        return rsp;
    }
    ~Peer() { if (fd >= 0) ::close(fd); }
};

struct SamplerConf { int type = 0, batch_size = 128, depth = 2, width = 2, rw_head = 128, rw_length = 2, queue = 4; uint64_t seed = 1; };

// ----------------------------------------------------------------------------------- server
class Server {
  public:
    Server(int rank, int nparts, const int64_t* offsets, int64_t f_len, int64_t i_len, const float* f, const int32_t* ifeat,
           const int64_t* indptr, const int64_t* nbr, const std::string& meta, int num_workers)
        : rank_(rank), nparts_(nparts), f_len_(f_len), i_len_(i_len), meta_(meta), num_workers_(num_workers) {
        offset_.assign(offsets, offsets + nparts + 1);
        n_local_ = offset_[(size_t)rank + 1] - offset_[(size_t)rank];
        f_.assign(f, f + n_local_ * f_len);
        i_.assign(ifeat, ifeat + n_local_ * i_len);
        indptr_.assign(indptr, indptr + n_local_ + 1);
        nbr_.assign(nbr, nbr + indptr_[(size_t)n_local_]);
        peers_.resize((size_t)nparts);
    }
    ~Server() { stop(); }
    void set_peers(const std::vector<std::pair<std::string, int>>& addr) {
        for (int p = 0; p < nparts_; ++p) { peers_[(size_t)p] = std::make_unique<Peer>(); peers_[(size_t)p]->host = addr[(size_t)p].first; peers_[(size_t)p]->port = addr[(size_t)p].second; }
    }
    void init_cache(double ratio, int policy) {
        int64_t remote = offset_[(size_t)nparts_] - n_local_;
        cache_ = std::make_unique<NodeCache>((size_t)std::max<int64_t>(0, (int64_t)(ratio * (double)remote)), policy);
    }
    void add_sampler(const SamplerConf& c) { confs_.push_back(c); }
    int rank() const { return rank_; }

    void listen(int port) {
        fd_ = hps::listen_on(port, &port_);
        accept_th_ = std::thread([this] { accept_loop(); });
    }
    int port() const { return port_; }
    // start the sampler threads; serving pull_graph blocks until then
    void ready() {
        for (size_t s = 0; s < confs_.size(); ++s) {
            queues_.push_back(std::make_unique<Queue>());
            queues_.back()->conf = confs_[s];
        }
        for (size_t s = 0; s < queues_.size(); ++s)
            threads_.emplace_back([this, s] { produce(*queues_[s], s); });
        { std::unique_lock<std::mutex> l(ready_mu_); ready_ = true; }
        ready_cv_.notify_all();
    }
    void wait() { if (accept_th_.joinable()) accept_th_.join(); }
    // a client asked the server to stop: only flag it and close the listener; the owner's stop() / destructor does the joins
    void stop_async() {
        stop_flag_only_ = true;
        if (fd_ >= 0) ::shutdown(fd_, SHUT_RDWR);
    }
    void stop() {
        if (stopped_.exchange(true)) return;
        for (auto& q : queues_) { std::unique_lock<std::mutex> l(q->mu); q->cv.notify_all(); }
        ready_cv_.notify_all();
        barrier_cv_.notify_all();
        if (fd_ >= 0) { ::shutdown(fd_, SHUT_RDWR); ::close(fd_); }
        if (accept_th_.joinable() && std::this_thread::get_id() != accept_th_.get_id()) accept_th_.join();
        for (auto& t : threads_) if (t.joinable()) t.join();
        // connection handlers run detached and use this object: wake them (their sockets are shut down) and wait
        // until the last one has left before the caller may destroy the server
        std::unique_lock<std::mutex> l(conn_mu_);
        for (int c : conn_fds_) ::shutdown(c, SHUT_RDWR);
        conn_cv_.wait(l, [&] { return active_conns_ == 0; });
    }
    void cache_stats(int64_t* lookups, int64_t* hits, int64_t* size) {
        *lookups = *hits = *size = 0;
        if (cache_) cache_->stats(lookups, hits, size);
    }

    int owner(int64_t gid) const {
        return (int)(std::upper_bound(offset_.begin(), offset_.end(), gid) - offset_.begin()) - 1;
    }
    void local_pack(int64_t gid, NodePack* p) const {
        int64_t l = gid - offset_[(size_t)rank_];
        if (l < 0 || l >= n_local_) throw std::runtime_error("node " + std::to_string(gid) + " is not on server " + std::to_string(rank_));
        p->f.assign(f_.begin() + l * f_len_, f_.begin() + (l + 1) * f_len_);
        p->i.assign(i_.begin() + l * i_len_, i_.begin() + (l + 1) * i_len_);
        p->nbr.assign(nbr_.begin() + indptr_[(size_t)l], nbr_.begin() + indptr_[(size_t)l + 1]);
    }
    // node data for arbitrary global ids: local rows, cache hits, then one request per remote owner
    void get_packs(const std::vector<int64_t>& ids, std::vector<NodePack>* out) {
        out->assign(ids.size(), NodePack());
        std::map<int, std::vector<size_t>> need;
        int64_t total = offset_[(size_t)nparts_];
        for (size_t k = 0; k < ids.size(); ++k) {
            int64_t g = ids[k];
            if (g < 0 || g >= total) throw std::runtime_error("node id " + std::to_string(g) + " out of range");
            int o = owner(g);
            if (o == rank_) local_pack(g, &(*out)[k]);
            else if (!(cache_ && cache_->lookup(g, &(*out)[k]))) need[o].push_back(k);
        }
        for (auto& kv : need) {
            Message req;
            req.type = kPullLocal;
            std::vector<int64_t> want;
            for (size_t k : kv.second) want.push_back(ids[k]);
            req.add(want.data(), want.size());
            Message rsp = peers_[(size_t)kv.first]->call(req);
            unpack_nodes(rsp, kv.second, out);
            if (cache_) for (size_t k : kv.second) cache_->insert(ids[k], (*out)[k]);
        }
    }

  private:
    struct Queue { SamplerConf conf; std::mutex mu; std::condition_variable cv; std::deque<std::shared_ptr<SubGraph>> q; };

    // wire format of node packs: f [n*f_len], i [n*i_len], nbr_ptr [n+1], nbr
    void pack_nodes(const std::vector<NodePack>& packs, Message* m) const {
        std::vector<float> f; std::vector<int32_t> iv; std::vector<int64_t> ptr{0}, nb;
        for (auto& p : packs) {
            f.insert(f.end(), p.f.begin(), p.f.end());
            iv.insert(iv.end(), p.i.begin(), p.i.end());
            nb.insert(nb.end(), p.nbr.begin(), p.nbr.end());
            ptr.push_back((int64_t)nb.size());
        }
        m->add(f.data(), f.size()); m->add(iv.data(), iv.size()); m->add(ptr.data(), ptr.size()); m->add(nb.data(), nb.size());
    }
    void unpack_nodes(const Message& m, const std::vector<size_t>& slots, std::vector<NodePack>* out) const {
        const float* f = m.data<float>(0); const int32_t* iv = m.data<int32_t>(1);
        const int64_t* ptr = m.data<int64_t>(2); const int64_t* nb = m.data<int64_t>(3);
        for (size_t k = 0; k < slots.size(); ++k) {
            NodePack& p = (*out)[slots[k]];
            p.f.assign(f + k * f_len_, f + (k + 1) * f_len_);
            p.i.assign(iv + k * i_len_, iv + (k + 1) * i_len_);
            p.nbr.assign(nb + ptr[k], nb + ptr[k + 1]);
        }
    }

    // sub-graph induced by `ids` (unique, order kept): features + edges among them
    std::shared_ptr<SubGraph> induce(int type, const std::vector<int64_t>& ids, const std::vector<int32_t>& extra, int extra_len) {
        std::vector<NodePack> packs;
        get_packs(ids, &packs);
        auto g = std::make_shared<SubGraph>();
        g->type = type; g->n = (int64_t)ids.size(); g->f_len = f_len_; g->i_len = i_len_ + 1; g->extra_len = extra_len;
        g->gid = ids; g->extra = extra;
        std::unordered_map<int64_t, int64_t> index;
        index.reserve(ids.size() * 2);
        for (size_t k = 0; k < ids.size(); ++k) index[ids[k]] = (int64_t)k;
        std::vector<int64_t> src, dst;
        for (size_t k = 0; k < ids.size(); ++k) {
            g->f.insert(g->f.end(), packs[k].f.begin(), packs[k].f.end());
            g->i.insert(g->i.end(), packs[k].i.begin(), packs[k].i.end());
            g->i.push_back((int32_t)(ids[k] & 0x7fffffff));
            for (int64_t v : packs[k].nbr) {
                auto it = index.find(v);
                if (it != index.end()) { src.push_back((int64_t)k); dst.push_back(it->second); }
            }
        }
        g->edges = src;
        g->edges.insert(g->edges.end(), dst.begin(), dst.end());
        return g;
    }

    std::shared_ptr<SubGraph> sample(const SamplerConf& c, std::mt19937_64& rng) {
        int64_t total = offset_[(size_t)nparts_], base = offset_[(size_t)rank_];
        auto pick_local = [&](int64_t count) {
            std::vector<int64_t> ids;
            count = std::min(count, n_local_);
            std::unordered_set<int64_t> seen;
            while ((int64_t)ids.size() < count) {
                int64_t g = base + (int64_t)(rng() % (uint64_t)n_local_);
                if (seen.insert(g).second) ids.push_back(g);
            }
            return ids;
        };
        switch (c.type) {
            case kLocalNode: return induce(c.type, pick_local(c.batch_size), {}, 0);
            case kGlobalNode: {
                std::vector<int64_t> ids;
                std::unordered_set<int64_t> seen;
                int64_t count = std::min<int64_t>(c.batch_size, total);
                while ((int64_t)ids.size() < count) {
                    int64_t g = (int64_t)(rng() % (uint64_t)total);
                    if (seen.insert(g).second) ids.push_back(g);
                }
                return induce(c.type, ids, {}, 0);
            }
            case kRandomWalk: {
                // rw_head walkers start on local nodes and take rw_length uniform steps; the batch is every visited node
                std::vector<int64_t> ids = pick_local(c.rw_head), frontier = ids;
                std::unordered_set<int64_t> seen(ids.begin(), ids.end());
                for (int step = 0; step < c.rw_length; ++step) {
                    std::vector<NodePack> packs;
                    get_packs(frontier, &packs);
                    for (size_t k = 0; k < frontier.size(); ++k) {
                        if (packs[k].nbr.empty()) continue;
                        int64_t nx = packs[k].nbr[rng() % packs[k].nbr.size()];
                        frontier[k] = nx;
                        if (seen.insert(nx).second) ids.push_back(nx);
                    }
                }
                return induce(c.type, ids, {}, 0);
            }
            default: {
                // GraphSage: batch_size local seeds, `depth` hops, at most `width` sampled neighbours per node and hop.
                // extra = [is_seed, hop at which the node was first reached]
                std::vector<int64_t> ids = pick_local(c.batch_size), frontier = ids;
                std::vector<int32_t> extra;
                for (size_t k = 0; k < ids.size(); ++k) { extra.push_back(1); extra.push_back(0); }
                std::unordered_set<int64_t> seen(ids.begin(), ids.end());
                for (int hop = 1; hop <= c.depth && !frontier.empty(); ++hop) {
                    std::vector<NodePack> packs;
                    get_packs(frontier, &packs);
                    std::vector<int64_t> next;
                    for (size_t k = 0; k < frontier.size(); ++k) {
                        std::vector<int64_t>& nb = packs[k].nbr;
                        size_t take = std::min<size_t>((size_t)c.width, nb.size());
                        for (size_t t = 0; t < take; ++t) {                 // partial Fisher-Yates
                            size_t j = t + (size_t)(rng() % (nb.size() - t));
                            std::swap(nb[t], nb[j]);
                            if (seen.insert(nb[t]).second) { ids.push_back(nb[t]); next.push_back(nb[t]); extra.push_back(0); extra.push_back(hop); }
                        }
                    }
                    frontier.swap(next);
                }
                return induce(kGraphSage, ids, extra, 2);
            }
        }
    }

    void produce(Queue& q, size_t idx) {
        std::mt19937_64 rng(q.conf.seed * 1315423911ull + (uint64_t)rank_ * 2654435761ull + idx);
        while (!stopped_) {
            std::shared_ptr<SubGraph> g;
            try { g = sample(q.conf, rng); }
            catch (const std::exception& e) { if (stopped_) break; fprintf(stderr, "[graph server %d] sampler failed: %s\n", rank_, e.what()); usleep(200 * 1000); continue; }
            std::unique_lock<std::mutex> l(q.mu);
            q.cv.wait(l, [&] { return stopped_ || (int)q.q.size() < q.conf.queue; });
            if (stopped_) break;
            q.q.push_back(g);
            q.cv.notify_all();
        }
    }

    std::shared_ptr<SubGraph> next_graph(const int32_t* types, size_t ntypes) {
        { std::unique_lock<std::mutex> l(ready_mu_); ready_cv_.wait(l, [&] { return ready_ || stopped_.load(); }); }
        if (stopped_) throw std::runtime_error("graph server is shutting down");
        std::vector<Queue*> cands;
        for (auto& q : queues_) {
            bool ok = ntypes == 0;
            for (size_t t = 0; t < ntypes; ++t) ok = ok || types[t] == q->conf.type;
            if (ok) cands.push_back(q.get());
        }
        if (cands.empty()) throw std::runtime_error("no sampler of the requested type was added on this server");
        Queue* q = cands[(size_t)(rr_++ % cands.size())];
        std::unique_lock<std::mutex> l(q->mu);
        q->cv.wait(l, [&] { return stopped_ || !q->q.empty(); });
        if (stopped_) throw std::runtime_error("graph server is shutting down");
        auto g = q->q.front();
        q->q.pop_front();
        ++served_graphs_;
        q->cv.notify_all();
        return g;
    }

    void handle(const Message& req, Message* rsp) {
        rsp->type = kOk;
        switch (req.type) {
            case kMeta: {
                rsp->add_str(meta_);
                int64_t info[4] = {rank_, nparts_, num_workers_, worker_seq_++};
                rsp->add(info, 4);
                break;
            }
            case kPullLocal:
            case kPullNode: {
                std::vector<int64_t> ids(req.data<int64_t>(0), req.data<int64_t>(0) + req.count<int64_t>(0));
                std::vector<NodePack> packs;
                if (req.type == kPullLocal) { packs.resize(ids.size()); for (size_t k = 0; k < ids.size(); ++k) local_pack(ids[k], &packs[k]); }
                else get_packs(ids, &packs);
                pack_nodes(packs, rsp);
                break;
            }
            case kPullGraph: next_graph(req.arrays.empty() ? nullptr : req.data<int32_t>(0), req.arrays.empty() ? 0 : req.count<int32_t>(0))->to_msg(rsp); break;
            case kBarrier: {
                std::unique_lock<std::mutex> l(barrier_mu_);
                int64_t want = req.arg0 > 0 ? req.arg0 : num_workers_;
                int64_t gen = barrier_gen_;
                if (++barrier_cnt_ >= want) { barrier_cnt_ = 0; ++barrier_gen_; barrier_cv_.notify_all(); }
                else barrier_cv_.wait(l, [&] { return barrier_gen_ != gen || stopped_.load(); });
                break;
            }
            case kStats: {
                int64_t st[4] = {0, 0, 0, 0};
                cache_stats(&st[0], &st[1], &st[2]);
                st[3] = served_graphs_.load();
                rsp->add(st, 4);
                break;
            }
            default: rsp->type = kError; rsp->add_str("unknown graph request");
        }
    }

    void accept_loop() {
        while (!stopped_) {
            int c = ::accept(fd_, nullptr, nullptr);
            if (c < 0) { if (stopped_ || stop_flag_only_) break; continue; }
            int one = 1;
            setsockopt(c, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
            { std::unique_lock<std::mutex> l(conn_mu_); conn_fds_.insert(c); ++active_conns_; }
            std::thread([this, c] {
                Message req, rsp;
                bool stop_requested = false;
                while (hps::recv_msg(c, &req)) {
                    rsp = Message();
                    if (req.type == kStop) { rsp.type = kOk; try { hps::send_msg(c, rsp); } catch (...) {} stop_requested = true; break; }
                    try { handle(req, &rsp); }
                    catch (const std::exception& e) { rsp = Message(); rsp.type = kError; rsp.add_str(e.what()); }
                    try { hps::send_msg(c, rsp); } catch (...) { break; }
                }
                {
                    std::unique_lock<std::mutex> l(conn_mu_);
                    conn_fds_.erase(c);
                    ::close(c);
                    --active_conns_;
                    conn_cv_.notify_all();
                }
                if (stop_requested) stop_async();
            }).detach();
        }
    }

    int rank_, nparts_;
    int64_t f_len_, i_len_, n_local_ = 0;
    std::string meta_;
    int num_workers_;
    std::vector<int64_t> offset_, indptr_, nbr_;
    std::vector<float> f_;
    std::vector<int32_t> i_;
    std::vector<std::unique_ptr<Peer>> peers_;
    std::unique_ptr<NodeCache> cache_;
    std::vector<SamplerConf> confs_;
    std::vector<std::unique_ptr<Queue>> queues_;
    std::vector<std::thread> threads_;
    std::thread accept_th_;
    int fd_ = -1, port_ = 0;
    std::atomic<bool> stopped_{false}, stop_flag_only_{false};
    std::mutex conn_mu_;
    std::condition_variable conn_cv_;
    std::set<int> conn_fds_;
    int active_conns_ = 0;
    std::atomic<uint64_t> rr_{0};
    std::atomic<int64_t> worker_seq_{0}, served_graphs_{0};
    std::mutex ready_mu_, barrier_mu_;
    std::condition_variable ready_cv_, barrier_cv_;
    bool ready_ = false;
    int64_t barrier_cnt_ = 0, barrier_gen_ = 0;
};

// ----------------------------------------------------------------------------------- client
class Client {
  public:
    Client(const std::string& host, int port) {
        peer_.host = host; peer_.port = port;
        Message req; req.type = kMeta;
        Message rsp = peer_.call(req);
        meta_ = rsp.str(0);
        const int64_t* info = rsp.data<int64_t>(1);
        server_rank_ = (int)info[0]; nparts_ = (int)info[1]; num_workers_ = (int)info[2];
    }
    const std::string& meta() const { return meta_; }
    int num_workers() const { return num_workers_; }
    int num_servers() const { return nparts_; }

    int64_t pull_graph(const int32_t* types, size_t n) {
        std::vector<int32_t> t(types, types + n);
        return submit([this, t]() mutable {
            Message req; req.type = kPullGraph;
            if (!t.empty()) req.add(t.data(), t.size());
            return call(req);
        });
    }
    int64_t pull_node(const int64_t* ids, size_t n) {
        std::vector<int64_t> v(ids, ids + n);
        return submit([this, v]() mutable { Message req; req.type = kPullNode; req.add(v.data(), v.size()); return call(req); });
    }
    std::shared_ptr<Message> resolve(int64_t q) {
        std::future<std::shared_ptr<Message>> f;
        {
            std::unique_lock<std::mutex> l(mu_);
            auto it = pending_.find(q);
            if (it == pending_.end()) { auto d = done_.find(q); if (d != done_.end()) return d->second; throw std::runtime_error("unknown query id"); }
            f = std::move(it->second);
            pending_.erase(it);
        }
        auto m = f.get();
        std::unique_lock<std::mutex> l(mu_);
        done_[q] = m;
        return m;
    }
    void release(int64_t q) { std::unique_lock<std::mutex> l(mu_); done_.erase(q); }
    void barrier(int64_t count, const std::string& host0, int port0) {
        // all workers meet on server 0, whichever server they sample from
        Peer p; p.host = host0; p.port = port0;
        Message req; req.type = kBarrier; req.arg0 = count;
        p.call(req);
    }
    void stop_server() { Message req; req.type = kStop; try { peer_.call(req); } catch (...) {} }
    void server_stats(int64_t* out4) {
        Message req; req.type = kStats;
        Message rsp = peer_.call(req);
        memcpy(out4, rsp.arrays[0].data(), 4 * sizeof(int64_t));
    }

  private:
    // queries may be in flight concurrently: each gets a connection of its own from a small pool
    std::shared_ptr<Message> call(Message& req) {
        std::unique_ptr<Peer> p;
        { std::unique_lock<std::mutex> l(mu_); if (!pool_.empty()) { p = std::move(pool_.back()); pool_.pop_back(); } }
        if (!p) { p = std::make_unique<Peer>(); p->host = peer_.host; p->port = peer_.port; }
        auto rsp = std::make_shared<Message>(p->call(req));
        { std::unique_lock<std::mutex> l(mu_); pool_.push_back(std::move(p)); }
        return rsp;
    }
    int64_t submit(std::function<std::shared_ptr<Message>()> fn) {
        std::unique_lock<std::mutex> l(mu_);
        int64_t q = ++next_;
        pending_[q] = std::async(std::launch::async, std::move(fn));
        return q;
    }
    Peer peer_;
    std::string meta_;
    int server_rank_ = 0, nparts_ = 1, num_workers_ = 0;
    std::mutex mu_;
    int64_t next_ = 0;
    std::map<int64_t, std::future<std::shared_ptr<Message>>> pending_;
    std::map<int64_t, std::shared_ptr<Message>> done_;
    std::vector<std::unique_ptr<Peer>> pool_;
};

static thread_local std::string g_err;

}  // namespace hgm

using namespace hgm;

#define HGM_TRY(...)                                    \
    try { __VA_ARGS__; return 0; }                      \
    catch (const std::exception& e) { g_err = e.what(); return -1; }

HGM_API const char* hgm_last_error() { return g_err.c_str(); }

// ---- server ---------------------------------------------------------------------------------
HGM_API void* hgm_server_create(int rank, int nparts, const int64_t* offsets, int64_t f_len, int64_t i_len, const float* f,
                                const int32_t* ifeat, const int64_t* indptr, const int64_t* nbr, const char* meta,
                                int num_workers) {
    try { return new Server(rank, nparts, offsets, f_len, i_len, f, ifeat, indptr, nbr, meta ? meta : "", num_workers); }
    catch (const std::exception& e) { g_err = e.what(); return nullptr; }
}
// peers: "host:port,host:port,..." in server-rank order
HGM_API int hgm_server_set_peers(void* h, const char* peers) {
    HGM_TRY({
        std::vector<std::pair<std::string, int>> addr;
        std::string s(peers), item;
        size_t pos = 0;
        while (pos <= s.size()) {
            size_t c = s.find(',', pos);
            item = s.substr(pos, c == std::string::npos ? std::string::npos : c - pos);
            if (!item.empty()) { size_t k = item.rfind(':'); addr.push_back({item.substr(0, k), atoi(item.substr(k + 1).c_str())}); }
            if (c == std::string::npos) break;
            pos = c + 1;
        }
        static_cast<Server*>(h)->set_peers(addr);
    });
}
HGM_API int hgm_server_listen(void* h, int port) { HGM_TRY(static_cast<Server*>(h)->listen(port)); }
HGM_API int hgm_server_port(void* h) { return static_cast<Server*>(h)->port(); }
HGM_API int hgm_server_init_cache(void* h, double ratio, int policy) { HGM_TRY(static_cast<Server*>(h)->init_cache(ratio, policy)); }
HGM_API int hgm_server_add_sampler(void* h, int type, int batch_size, int depth, int width, int rw_head, int rw_length, int queue,
                                   uint64_t seed) {
    HGM_TRY({
        if (type < 0 || type >= kNumSampler) throw std::runtime_error("unknown sampler type");
        SamplerConf c;
        c.type = type; c.batch_size = std::max(1, batch_size); c.depth = depth; c.width = width; c.rw_head = std::max(1, rw_head);
        c.rw_length = rw_length; c.queue = std::max(1, queue); c.seed = seed;
        static_cast<Server*>(h)->add_sampler(c);
    });
}
HGM_API int hgm_server_ready(void* h) { HGM_TRY(static_cast<Server*>(h)->ready()); }
HGM_API int hgm_server_wait(void* h) { HGM_TRY(static_cast<Server*>(h)->wait()); }
HGM_API int hgm_server_stop(void* h) { HGM_TRY(static_cast<Server*>(h)->stop()); }
HGM_API void hgm_server_free(void* h) { delete static_cast<Server*>(h); }
HGM_API int hgm_server_cache_stats(void* h, int64_t* out3) { HGM_TRY(static_cast<Server*>(h)->cache_stats(out3, out3 + 1, out3 + 2)); }

// ---- client ---------------------------------------------------------------------------------
HGM_API void* hgm_client_create(const char* host, int port) {
    try { return new Client(host, port); }
    catch (const std::exception& e) { g_err = e.what(); return nullptr; }
}
HGM_API void hgm_client_free(void* h) { delete static_cast<Client*>(h); }
HGM_API int hgm_client_meta(void* h, char* buf, int cap) {
    const std::string& m = static_cast<Client*>(h)->meta();
    if (buf && cap > 0) { size_t n = std::min(m.size(), (size_t)cap - 1); memcpy(buf, m.data(), n); buf[n] = 0; }
    return (int)m.size();
}
HGM_API int hgm_client_num_workers(void* h) { return static_cast<Client*>(h)->num_workers(); }
HGM_API int hgm_client_num_servers(void* h) { return static_cast<Client*>(h)->num_servers(); }
HGM_API int64_t hgm_client_pull_graph(void* h, const int32_t* types, int n) { return static_cast<Client*>(h)->pull_graph(types, (size_t)n); }
HGM_API int64_t hgm_client_pull_node(void* h, const int64_t* ids, int64_t n) { return static_cast<Client*>(h)->pull_node(ids, (size_t)n); }
// wait for a graph query; dims = {type, n, f_len, i_len, extra_len, n_edges}
HGM_API int hgm_client_resolve_graph(void* h, int64_t q, int64_t* dims) {
    HGM_TRY({
        auto m = static_cast<Client*>(h)->resolve(q);
        const int64_t* hdr = m->data<int64_t>(0);
        for (int k = 0; k < 5; ++k) dims[k] = hdr[k];
        dims[5] = (int64_t)m->count<int64_t>(4) / 2;
    });
}
HGM_API int hgm_client_fetch_graph(void* h, int64_t q, float* f, int32_t* iv, int64_t* gid, int64_t* edges, int32_t* extra) {
    HGM_TRY({
        auto m = static_cast<Client*>(h)->resolve(q);
        memcpy(f, m->arrays[1].data(), m->arrays[1].size());
        memcpy(iv, m->arrays[2].data(), m->arrays[2].size());
        memcpy(gid, m->arrays[3].data(), m->arrays[3].size());
        memcpy(edges, m->arrays[4].data(), m->arrays[4].size());
        if (extra) memcpy(extra, m->arrays[5].data(), m->arrays[5].size());
        static_cast<Client*>(h)->release(q);
    });
}
// wait for a node query; dims = {n_f_values, n_i_values, n_ptr, n_nbr}
HGM_API int hgm_client_resolve_node(void* h, int64_t q, int64_t* dims) {
    HGM_TRY({
        auto m = static_cast<Client*>(h)->resolve(q);
        dims[0] = (int64_t)m->count<float>(0); dims[1] = (int64_t)m->count<int32_t>(1);
        dims[2] = (int64_t)m->count<int64_t>(2); dims[3] = (int64_t)m->count<int64_t>(3);
    });
}
HGM_API int hgm_client_fetch_node(void* h, int64_t q, float* f, int32_t* iv, int64_t* ptr, int64_t* nbr) {
    HGM_TRY({
        auto m = static_cast<Client*>(h)->resolve(q);
        memcpy(f, m->arrays[0].data(), m->arrays[0].size());
        memcpy(iv, m->arrays[1].data(), m->arrays[1].size());
        memcpy(ptr, m->arrays[2].data(), m->arrays[2].size());
        memcpy(nbr, m->arrays[3].data(), m->arrays[3].size());
        static_cast<Client*>(h)->release(q);
    });
}
HGM_API int hgm_client_barrier(void* h, int64_t count, const char* host0, int port0) { HGM_TRY(static_cast<Client*>(h)->barrier(count, host0, port0)); }
// out4 = {remote-node cache lookups, hits, cached nodes, sub-graphs served} of the server this client talks to
HGM_API int hgm_client_server_stats(void* h, int64_t* out4) { HGM_TRY(static_cast<Client*>(h)->server_stats(out4)); }
HGM_API int hgm_client_stop_server(void* h) { HGM_TRY(static_cast<Client*>(h)->stop_server()); }

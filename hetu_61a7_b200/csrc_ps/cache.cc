// HET embedding cache: bounded-staleness client cache in front of the parameter server.
//
// reference: src/hetu_cache/ (embedding.h Line<T>, cache.cc lookup/update/push-pull,
// lru_cache.cc / lfu_cache.cc / lfuopt_cache.cc, hetu_client.cc) - HET, VLDB'22.
// Semantics kept: every cached row carries (version, pending grad, #updates); a lookup syncs a
// row only when the server is more than `pull_bound` versions ahead (or the row is new), and
// re-applies the not-yet-pushed local gradient on top; an update accumulates locally and pushes
// rows whose pending updates exceed `push_bound` (plus evicted dirty rows).  Policies: LRU, LFU,
// LFUOpt (frequency buckets; rows reaching kUseCntMax are promoted to a never-evicted store).
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <list>
#include <map>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

#define HPS_API extern "C" __attribute__((visibility("default")))

namespace hps {
void worker_sync_embedding(uint32_t type, uint64_t key, const int64_t* rows, const int64_t* vers, size_t n, int64_t bound,
                           const int64_t* prow, const float* pdata, const int64_t* pupd, size_t pn,
                           std::vector<int64_t>* idx, std::vector<int64_t>* over, std::vector<float>* odata);
constexpr uint32_t kSyncEmbedding = 11, kPushEmbedding = 12, kPushSyncEmbedding = 13;

struct Line {
    int64_t key;
    std::vector<float> data, grad;
    int64_t version = -1, updates = 0;
    int64_t freq = 1;
    bool has_data = false;
    std::list<int64_t>::iterator pos;        // position in its policy list
};

enum Policy { LRU = 0, LFU = 1, LFUOPT = 2 };
constexpr int64_t kUseCntMax = 16;

struct Perf { int64_t lookups = 0, misses = 0, evicts = 0, pulled_rows = 0, pushed_rows = 0, requests = 0; };

class Cache {
  public:
    Cache(size_t limit, size_t length, size_t width, uint64_t node_id, int policy, int64_t pull_bound, int64_t push_bound)
        : limit_(limit), length_(length), width_(width), node_(node_id), policy_(policy), pull_bound_(pull_bound),
          push_bound_(push_bound) {}

    // out[i, :] = row keys[i]
    void lookup(const int64_t* keys, size_t n, float* out) {
        std::unique_lock<std::mutex> l(mu_);
        perf_.requests++;
        std::vector<int64_t> uniq(keys, keys + n);
        std::sort(uniq.begin(), uniq.end());
        uniq.erase(std::unique(uniq.begin(), uniq.end()), uniq.end());
        std::vector<int64_t> vers(uniq.size());
        for (size_t i = 0; i < uniq.size(); ++i) {
            perf_.lookups++;
            Line* ln = find(uniq[i]);
            if (!ln) { perf_.misses++; ln = insert(uniq[i]); }
            else touch(ln);
            vers[i] = (ln->has_data && !bypass_) ? ln->version : -1;
        }
        flush_evicted();
        std::vector<int64_t> idx, nver;
        std::vector<float> data;
        worker_sync_embedding(kSyncEmbedding, node_, uniq.data(), vers.data(), uniq.size(), pull_bound_, nullptr, nullptr,
                              nullptr, 0, &idx, &nver, &data);
        apply_sync(uniq, idx, nver, data);
        scatter(keys, n, out);
    }
    // grads[i, :] for keys[i] (already scaled by -lr by the caller, like the reference)
    void update(const int64_t* keys, size_t n, const float* grads) {
        std::unique_lock<std::mutex> l(mu_);
        perf_.requests++;
        std::vector<int64_t> prow, pupd;
        std::vector<float> pdata;
        accumulate(keys, n, grads, &prow, &pdata, &pupd);
        take_evicted(&prow, &pdata, &pupd);
        if (!prow.empty()) {
            worker_sync_embedding(kPushEmbedding, node_, nullptr, nullptr, 0, 0, prow.data(), pdata.data(), pupd.data(),
                                  prow.size(), nullptr, nullptr, nullptr);
            perf_.pushed_rows += (int64_t)prow.size();
        }
    }
    void push_pull(const int64_t* pull_keys, size_t n_pull, float* out, const int64_t* push_keys, size_t n_push,
                   const float* grads) {
        std::unique_lock<std::mutex> l(mu_);
        perf_.requests++;
        std::vector<int64_t> prow, pupd;
        std::vector<float> pdata;
        accumulate(push_keys, n_push, grads, &prow, &pdata, &pupd);
        std::vector<int64_t> uniq(pull_keys, pull_keys + n_pull);
        std::sort(uniq.begin(), uniq.end());
        uniq.erase(std::unique(uniq.begin(), uniq.end()), uniq.end());
        std::vector<int64_t> vers(uniq.size());
        for (size_t i = 0; i < uniq.size(); ++i) {
            perf_.lookups++;
            Line* ln = find(uniq[i]);
            if (!ln) { perf_.misses++; ln = insert(uniq[i]); }
            else touch(ln);
            vers[i] = (ln->has_data && !bypass_) ? ln->version : -1;
        }
        take_evicted(&prow, &pdata, &pupd);
        std::vector<int64_t> idx, nver;
        std::vector<float> data;
        worker_sync_embedding(kPushSyncEmbedding, node_, uniq.data(), vers.data(), uniq.size(), pull_bound_, prow.data(),
                              pdata.data(), pupd.data(), prow.size(), &idx, &nver, &data);
        perf_.pushed_rows += (int64_t)prow.size();
        apply_sync(uniq, idx, nver, data);
        scatter(pull_keys, n_pull, out);
    }
    // Push every not-yet-pushed local update (lines below the push bound and evicted dirty lines) to the servers,
    // so that a checkpoint taken right after sees them.  The reference's save does not do this (SURVEY 5.4:
    // "dirty lines are not flushed by save").  Returns the number of rows pushed.
    size_t flush() {
        std::unique_lock<std::mutex> l(mu_);
        std::vector<int64_t> prow, pupd;
        std::vector<float> pdata;
        for (auto& kv : lines_) {
            Line* ln = kv.second.get();
            if (ln->updates <= 0) continue;
            prow.push_back(ln->key);
            pdata.insert(pdata.end(), ln->grad.begin(), ln->grad.end());
            pupd.push_back(ln->updates);
            if (ln->version >= 0) ln->version += ln->updates;
            ln->updates = 0;
            std::fill(ln->grad.begin(), ln->grad.end(), 0.f);
        }
        take_evicted(&prow, &pdata, &pupd);
        if (!prow.empty()) {
            worker_sync_embedding(kPushEmbedding, node_, nullptr, nullptr, 0, 0, prow.data(), pdata.data(), pupd.data(),
                                  prow.size(), nullptr, nullptr, nullptr);
            perf_.pushed_rows += (int64_t)prow.size();
        }
        return prow.size();
    }
    // Forget every line, pending local gradients included (the server copy was just replaced, e.g. by a checkpoint
    // load: rows and un-pushed deltas held here describe the old parameters).
    void invalidate() {
        std::unique_lock<std::mutex> l(mu_);
        lines_.clear();
        order_.clear();
        evicted_.clear();
        permanent_ = 0;
    }
    size_t size() const { return lines_.size(); }
    Perf perf_;
    bool bypass_ = false;
    int count(int64_t key) { std::unique_lock<std::mutex> l(mu_); return lines_.count(key) ? 1 : 0; }
    void keys(std::vector<int64_t>* out) { std::unique_lock<std::mutex> l(mu_); for (auto& kv : lines_) out->push_back(kv.first); }
    bool peek(int64_t key, float* out) {
        std::unique_lock<std::mutex> l(mu_);
        auto it = lines_.find(key);
        if (it == lines_.end() || !it->second->has_data) return false;
        memcpy(out, it->second->data.data(), width_ * sizeof(float));
        return true;
    }
    void put(int64_t key, const float* row) {
        std::unique_lock<std::mutex> l(mu_);
        Line* ln = find(key);
        if (!ln) ln = insert(key);
        ln->data.assign(row, row + width_);
        ln->has_data = true;
        if (ln->version < 0) ln->version = 0;
        flush_evicted();
    }

  private:
    Line* find(int64_t key) { auto it = lines_.find(key); return it == lines_.end() ? nullptr : it->second.get(); }
    std::list<int64_t>& bucket(int64_t f) { return policy_ == LRU ? order_[0] : order_[f]; }
    void touch(Line* ln) {
        if (ln->freq < 0) return;                          // promoted to the permanent store (LFUOpt)
        bucket(ln->freq).erase(ln->pos);
        if (policy_ != LRU && order_[ln->freq].empty()) order_.erase(ln->freq);
        if (policy_ != LRU) ln->freq++;
        if (policy_ == LFUOPT && ln->freq >= kUseCntMax) { ln->freq = -1; permanent_++; return; }
        auto& b = bucket(ln->freq);
        b.push_front(ln->key);
        ln->pos = b.begin();
    }
    Line* insert(int64_t key) {
        while (lines_.size() - permanent_ >= limit_ && evict_one()) {}
        auto ln = std::make_unique<Line>();
        ln->key = key;
        ln->data.assign(width_, 0.f);
        ln->freq = 1;
        auto& b = bucket(1);
        b.push_front(key);
        ln->pos = b.begin();
        Line* raw = ln.get();
        lines_[key] = std::move(ln);
        return raw;
    }
    bool evict_one() {
        if (order_.empty()) return false;
        auto it = order_.begin();                          // LRU: the only list; LFU: lowest frequency
        while (it != order_.end() && it->second.empty()) it = order_.erase(it);
        if (it == order_.end()) return false;
        int64_t victim = it->second.back();
        it->second.pop_back();
        if (policy_ != LRU && it->second.empty()) order_.erase(it);
        auto lit = lines_.find(victim);
        if (lit != lines_.end()) {
            if (lit->second->updates > 0) evicted_.push_back(std::move(lit->second));   // dirty: push later
            lines_.erase(lit);
        }
        perf_.evicts++;
        return true;
    }
    void accumulate(const int64_t* keys, size_t n, const float* grads, std::vector<int64_t>* prow, std::vector<float>* pdata,
                    std::vector<int64_t>* pupd) {
        std::map<int64_t, std::vector<float>> sum;         // pre-sum duplicate ids (PSAgent.h:124-183)
        std::map<int64_t, int64_t> cnt;
        for (size_t i = 0; i < n; ++i) {
            auto& v = sum[keys[i]];
            if (v.empty()) v.assign(width_, 0.f);
            for (size_t j = 0; j < width_; ++j) v[j] += grads[i * width_ + j];
            cnt[keys[i]]++;
        }
        for (auto& kv : sum) {
            Line* ln = find(kv.first);
            if (!ln) ln = insert(kv.first);
            else touch(ln);
            if (ln->grad.empty()) ln->grad.assign(width_, 0.f);
            for (size_t j = 0; j < width_; ++j) { ln->grad[j] += kv.second[j]; if (ln->has_data) ln->data[j] += kv.second[j]; }
            ln->updates += 1;
            if (ln->updates > push_bound_ || !ln->has_data || bypass_) {
                prow->push_back(ln->key);
                pdata->insert(pdata->end(), ln->grad.begin(), ln->grad.end());
                pupd->push_back(ln->updates);
                if (ln->version >= 0) ln->version += ln->updates;
                ln->updates = 0;
                std::fill(ln->grad.begin(), ln->grad.end(), 0.f);
            }
        }
    }
    void take_evicted(std::vector<int64_t>* prow, std::vector<float>* pdata, std::vector<int64_t>* pupd) {
        for (auto& ln : evicted_) {
            prow->push_back(ln->key);
            pdata->insert(pdata->end(), ln->grad.begin(), ln->grad.end());
            pupd->push_back(ln->updates);
        }
        evicted_.clear();
    }
    void flush_evicted() {
        if (evicted_.empty()) return;
        std::vector<int64_t> prow, pupd;
        std::vector<float> pdata;
        take_evicted(&prow, &pdata, &pupd);
        worker_sync_embedding(kPushEmbedding, node_, nullptr, nullptr, 0, 0, prow.data(), pdata.data(), pupd.data(), prow.size(),
                              nullptr, nullptr, nullptr);
        perf_.pushed_rows += (int64_t)prow.size();
    }
    void apply_sync(const std::vector<int64_t>& uniq, const std::vector<int64_t>& idx, const std::vector<int64_t>& nver,
                    const std::vector<float>& data) {
        perf_.pulled_rows += (int64_t)idx.size();
        for (size_t i = 0; i < idx.size(); ++i) {
            Line* ln = find(uniq[(size_t)idx[i]]);
            if (!ln) continue;
            ln->data.assign(data.begin() + i * width_, data.begin() + (i + 1) * width_);
            if (!ln->grad.empty())                          // re-apply the not-yet-pushed local gradient
                for (size_t j = 0; j < width_; ++j) ln->data[j] += ln->grad[j];
            ln->version = nver[i];
            ln->has_data = true;
        }
    }
    void scatter(const int64_t* keys, size_t n, float* out) {
        for (size_t i = 0; i < n; ++i) {
            Line* ln = find(keys[i]);
            if (ln) memcpy(out + i * width_, ln->data.data(), width_ * sizeof(float));
            else memset(out + i * width_, 0, width_ * sizeof(float));
        }
    }

    size_t limit_, length_, width_;
    uint64_t node_;
    int policy_;
    int64_t pull_bound_, push_bound_;
    size_t permanent_ = 0;
    std::mutex mu_;
    std::unordered_map<int64_t, std::unique_ptr<Line>> lines_;
    std::map<int64_t, std::list<int64_t>> order_;          // freq -> keys (front = most recent)
    std::vector<std::unique_ptr<Line>> evicted_;
};

static thread_local std::string g_cerr;
}  // namespace hps

using namespace hps;
#define HC_TRY(body) try { body; return 0; } catch (const std::exception& e) { g_cerr = e.what(); return -1; }

HPS_API const char* hcache_last_error() { return g_cerr.c_str(); }
HPS_API void* hcache_create(uint64_t limit, uint64_t length, uint64_t width, uint64_t node_id, int policy, int64_t pull_bound,
                            int64_t push_bound) {
    return new Cache(limit, length, width, node_id, policy, pull_bound, push_bound);
}
HPS_API void hcache_destroy(void* h) { delete static_cast<Cache*>(h); }
HPS_API int hcache_lookup(void* h, const int64_t* keys, uint64_t n, float* out) { HC_TRY(static_cast<Cache*>(h)->lookup(keys, n, out)); }
HPS_API int hcache_update(void* h, const int64_t* keys, uint64_t n, const float* grads) { HC_TRY(static_cast<Cache*>(h)->update(keys, n, grads)); }
HPS_API int hcache_push_pull(void* h, const int64_t* pull_keys, uint64_t n_pull, float* out, const int64_t* push_keys,
                             uint64_t n_push, const float* grads) {
    HC_TRY(static_cast<Cache*>(h)->push_pull(pull_keys, n_pull, out, push_keys, n_push, grads));
}
HPS_API int64_t hcache_flush(void* h) {
    try { return (int64_t)static_cast<Cache*>(h)->flush(); }
    catch (const std::exception& e) { g_cerr = e.what(); return -1; }
}
HPS_API void hcache_invalidate(void* h) { static_cast<Cache*>(h)->invalidate(); }
HPS_API uint64_t hcache_size(void* h) { return static_cast<Cache*>(h)->size(); }
HPS_API void hcache_bypass(void* h, int on) { static_cast<Cache*>(h)->bypass_ = on != 0; }
HPS_API int hcache_count(void* h, int64_t key) { return static_cast<Cache*>(h)->count(key); }
HPS_API int hcache_peek(void* h, int64_t key, float* out) { return static_cast<Cache*>(h)->peek(key, out) ? 1 : 0; }
HPS_API int hcache_put(void* h, int64_t key, const float* row) { HC_TRY(static_cast<Cache*>(h)->put(key, row)); }
HPS_API uint64_t hcache_keys(void* h, int64_t* out, uint64_t cap) {
    std::vector<int64_t> k;
    static_cast<Cache*>(h)->keys(&k);
    for (size_t i = 0; i < k.size() && i < cap; ++i) out[i] = k[i];
    return k.size();
}
HPS_API void hcache_perf(void* h, int64_t* out6) {
    Perf& p = static_cast<Cache*>(h)->perf_;
    out6[0] = p.lookups; out6[1] = p.misses; out6[2] = p.evicts; out6[3] = p.pulled_rows; out6[4] = p.pushed_rows; out6[5] = p.requests;
}

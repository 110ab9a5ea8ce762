// Classical machine-learning core (libhetu_ml.so): linear models, factorization machine, gradient
// boosted trees / random forest, k-nearest-neighbours, Gaussian naive Bayes, matrix factorization
// and LDA over sparse (CSR / COO) inputs.  OpenMP on the host, plain C ABI for ctypes.
//
// reference: third_party/HetuML/hetuml/cc/ml (include/model/{linear,fm,gbdt,rf,knn,naive_bayes,mf,lda},
// include/objective/{loss,metric}.h, include/stats/quantile.h; 13.1 kLoC of header templates bound
// with pybind11 in src/python_binding/api.cc).  Same model families, hyper-parameters and
// prediction semantics; data stay in the caller's numpy buffers (no copies into wrapper classes),
// tree growth is histogram based with the sparse zero-bin residual and sibling subtraction, and
// the distributed variants plug in through one all-reduce callback instead of a second ps-lite.
#include <omp.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <functional>
#include <memory>
#include <numeric>
#include <queue>
#include <random>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

#define HML_API extern "C" __attribute__((visibility("default")))

namespace hml {

static thread_local std::string g_err;

struct Csr {
    const int64_t* indptr;
    const int32_t* indices;
    const float* vals;
    int64_t n;
};

inline float sigmoidf(float x) { return x >= 0 ? 1.f / (1.f + std::exp(-x)) : std::exp(x) / (1.f + std::exp(x)); }
inline float signf(float x) { return x > 0 ? 1.f : (x < 0 ? -1.f : 0.f); }

// ------------------------------------------------------------------------------ linear family
// loss 0: logistic on y in {-1,+1}; 1: hinge on y in {-1,+1}; 2: square.
// reference: include/model/linear/linear.h (mini-batch gradient descent over sparse rows,
// L1/L2 terms applied to the coordinates a batch touches), include/objective/loss.h:294-308.
inline float dloss(int loss, float margin, float y) {
    switch (loss) {
        case 0: return -y * sigmoidf(-y * margin);
        case 1: return y * margin < 1.f ? -y : 0.f;
        default: return margin - y;
    }
}

static void linear_fit(int loss, const Csr& X, const float* y, int64_t dim, float* w, int epochs, int batch,
                       float lr, float l1, float l2) {
    std::vector<float> grad((size_t)dim, 0.f);
    std::vector<int32_t> touched;
    std::vector<char> mark((size_t)dim, 0);
    std::vector<float> coef((size_t)batch);
    for (int ep = 0; ep < epochs; ++ep) {
        for (int64_t b0 = 0; b0 < X.n; b0 += batch) {
            int64_t b1 = std::min<int64_t>(X.n, b0 + batch);
#pragma omp parallel for schedule(static)
            for (int64_t i = b0; i < b1; ++i) {
                float m = 0.f;
                for (int64_t p = X.indptr[i]; p < X.indptr[i + 1]; ++p)
                    if (X.indices[p] < dim) m += w[X.indices[p]] * X.vals[p];
                coef[(size_t)(i - b0)] = dloss(loss, m, y[i]);
            }
            touched.clear();
            for (int64_t i = b0; i < b1; ++i) {
                float c = coef[(size_t)(i - b0)];
                for (int64_t p = X.indptr[i]; p < X.indptr[i + 1]; ++p) {
                    int32_t f = X.indices[p];
                    if (f >= dim) continue;
                    if (!mark[(size_t)f]) { mark[(size_t)f] = 1; touched.push_back(f); }
                    grad[(size_t)f] += c * X.vals[p];
                }
            }
            float inv = 1.f / (float)(b1 - b0);
            for (int32_t f : touched) {
                float g = grad[(size_t)f] * inv + l2 * w[f] + l1 * signf(w[f]);
                w[f] -= lr * g;
                grad[(size_t)f] = 0.f; mark[(size_t)f] = 0;
            }
        }
    }
}

static void linear_predict(const Csr& X, int64_t dim, const float* w, float* out) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < X.n; ++i) {
        float m = 0.f;
        for (int64_t p = X.indptr[i]; p < X.indptr[i + 1]; ++p)
            if (X.indices[p] < dim) m += w[X.indices[p]] * X.vals[p];
        out[i] = m;
    }
}

// ---------------------------------------------------------------------- factorization machine
// score = sum_i w_i x_i + 1/2 sum_f [ (sum_i v_if x_i)^2 - sum_i v_if^2 x_i^2 ]
// reference: include/model/fm/fm.h.
static float fm_score(const Csr& X, int64_t i, int64_t dim, int k, const float* w, const float* v, float* sums) {
    float s = 0.f;
    std::fill(sums, sums + k, 0.f);
    float sq = 0.f;
    for (int64_t p = X.indptr[i]; p < X.indptr[i + 1]; ++p) {
        int32_t f = X.indices[p];
        if (f >= dim) continue;
        float x = X.vals[p];
        s += w[f] * x;
        const float* vf = v + (size_t)f * k;
        for (int j = 0; j < k; ++j) { float t = vf[j] * x; sums[j] += t; sq += t * t; }
    }
    float inter = 0.f;
    for (int j = 0; j < k; ++j) inter += sums[j] * sums[j];
    return s + 0.5f * (inter - sq);
}

static void fm_fit(const Csr& X, const float* y, int64_t dim, int k, float* w, float* v, int epochs, int batch,
                   float lr, float l1, float l2) {
    std::vector<float> gw((size_t)dim, 0.f), gv((size_t)dim * k, 0.f);
    std::vector<char> mark((size_t)dim, 0);
    std::vector<int32_t> touched;
    std::vector<float> coef((size_t)batch), sums((size_t)batch * k);
    for (int ep = 0; ep < epochs; ++ep) {
        for (int64_t b0 = 0; b0 < X.n; b0 += batch) {
            int64_t b1 = std::min<int64_t>(X.n, b0 + batch);
#pragma omp parallel for schedule(static)
            for (int64_t i = b0; i < b1; ++i) {
                float m = fm_score(X, i, dim, k, w, v, &sums[(size_t)(i - b0) * k]);
                coef[(size_t)(i - b0)] = dloss(0, m, y[i]);
            }
            touched.clear();
            for (int64_t i = b0; i < b1; ++i) {
                float c = coef[(size_t)(i - b0)];
                const float* sm = &sums[(size_t)(i - b0) * k];
                for (int64_t p = X.indptr[i]; p < X.indptr[i + 1]; ++p) {
                    int32_t f = X.indices[p];
                    if (f >= dim) continue;
                    float x = X.vals[p];
                    if (!mark[(size_t)f]) { mark[(size_t)f] = 1; touched.push_back(f); }
                    gw[(size_t)f] += c * x;
                    float* g = &gv[(size_t)f * k];
                    const float* vf = v + (size_t)f * k;
                    for (int j = 0; j < k; ++j) g[j] += c * x * (sm[j] - vf[j] * x);
                }
            }
            float inv = 1.f / (float)(b1 - b0);
            for (int32_t f : touched) {
                w[f] -= lr * (gw[(size_t)f] * inv + l2 * w[f] + l1 * signf(w[f]));
                gw[(size_t)f] = 0.f; mark[(size_t)f] = 0;
                float* vf = v + (size_t)f * k;
                float* g = &gv[(size_t)f * k];
                for (int j = 0; j < k; ++j) { vf[j] -= lr * (g[j] * inv + l2 * vf[j] + l1 * signf(vf[j])); g[j] = 0.f; }
            }
        }
    }
}

// ------------------------------------------------------------------------------------- trees
// reference: include/model/gbdt/{gbdt.h,train/trainer.h,helper/*,hist/*}, include/model/tree/*,
// include/stats/quantile.h, include/model/rf/rf.h (random forest = the same trainer with
// instance/feature sampling and majority voting).
typedef void (*allreduce_fn)(double* buf, int64_t n);

struct TreeNode { int32_t feat = -1; float thr = 0.f; int32_t left = -1, right = -1; int32_t leaf = -1; };

struct Tree {
    std::vector<TreeNode> nodes;
    std::vector<float> leaf_vals;        // [num_leaves, out_dim]
    int out_dim = 1;
    const float* eval(const int32_t* idx, const float* val, int64_t nnz, int64_t /*dim*/) const {
        int cur = 0;
        while (nodes[(size_t)cur].leaf < 0) {
            const TreeNode& nd = nodes[(size_t)cur];
            float x = 0.f;
            const int32_t* it = std::lower_bound(idx, idx + nnz, nd.feat);
            if (it != idx + nnz && *it == nd.feat) x = val[it - idx];
            cur = x <= nd.thr ? nd.left : nd.right;
        }
        return &leaf_vals[(size_t)nodes[(size_t)cur].leaf * out_dim];
    }
};

struct GbdtParams {
    int max_depth = 4, max_node_num = 31, leafwise = 0, num_split = 10;
    double ins_sp_ratio = 1, feat_sp_ratio = 1;
    int is_regression = 0, num_label = 2, num_round = 10;
    double learning_rate = 0.1, min_split_gain = 0;
    int min_node_ins = 1024;
    double min_child_weight = 0, reg_alpha = 0, reg_lambda = 0, max_leaf_weight = 0;
    int multi_tree = 0, majority_voting = 0;
    uint64_t seed = 42;
};

struct Gbdt {
    GbdtParams p;
    int64_t dim = 0;
    int score_dim = 1;                   // 1 (binary / regression) or num_label
    std::vector<Tree> trees;
    std::vector<int> tree_class;         // class a single-output tree adds to (-1 = all / vector leaf)
    std::vector<double> round_metric;    // training loss after each round
};

struct Cuts {
    std::vector<int64_t> off;            // per feature offset into `vals`; bins of f = cuts+1
    std::vector<float> vals;
    std::vector<int64_t> bin_off;        // per feature offset into a histogram
    std::vector<int32_t> zero_bin;
    int64_t total_bins = 0;
    int bin_of(int32_t f, float x) const {
        const float* b = &vals[(size_t)off[(size_t)f]];
        const float* e = &vals[(size_t)off[(size_t)f + 1]];
        return (int)(std::lower_bound(b, e, x) - b);      // number of cuts < x ; "x <= cut_j" <=> bin <= j
    }
};

// Candidate splits: up to `num_split` quantiles of each feature's non-zero values plus 0.
// Distributed training passes `range` ([dim][lo, hi], agreed on by all workers) and equal-width
// cuts are used instead, so that every worker derives identical bins from its local shard.
static Cuts build_cuts(const Csr& X, int64_t dim, int num_split, const double* range) {
    std::vector<std::vector<float>> col((size_t)dim);
    if (!range)
        for (int64_t i = 0; i < X.n; ++i)
            for (int64_t p = X.indptr[i]; p < X.indptr[i + 1]; ++p)
                if (X.indices[p] < dim) col[(size_t)X.indices[p]].push_back(X.vals[p]);
    Cuts c;
    c.off.assign((size_t)dim + 1, 0);
    std::vector<std::vector<float>> cuts((size_t)dim);
    if (range) {
        for (int64_t f = 0; f < dim; ++f) {
            double lo = std::min(range[(size_t)f * 2], 0.0), hi = std::max(range[(size_t)f * 2 + 1], 0.0);
            if (hi <= lo) continue;
            auto& cf = cuts[(size_t)f];
            for (int j = 0; j < num_split; ++j) cf.push_back((float)(lo + (hi - lo) * j / num_split));
            cf.push_back(0.f);
            std::sort(cf.begin(), cf.end());
            cf.erase(std::unique(cf.begin(), cf.end()), cf.end());
        }
    } else {
#pragma omp parallel for schedule(dynamic, 64)
        for (int64_t f = 0; f < dim; ++f) {
            auto& v = col[(size_t)f];
            if (v.empty()) continue;
            std::sort(v.begin(), v.end());
            auto& cf = cuts[(size_t)f];
            size_t m = v.size();
            for (int j = 1; j <= num_split; ++j) {
                size_t q = std::min(m - 1, (size_t)((double)j / num_split * (double)m));
                if (q > 0) cf.push_back(v[q - 1]);
            }
            cf.push_back(0.f);
            std::sort(cf.begin(), cf.end());
            cf.erase(std::unique(cf.begin(), cf.end()), cf.end());
            if (!cf.empty() && cf.back() >= v.back()) cf.pop_back();     // a cut at the maximum separates nothing
        }
    }
    c.bin_off.assign((size_t)dim + 1, 0);
    c.zero_bin.assign((size_t)dim, 0);
    for (int64_t f = 0; f < dim; ++f) {
        c.off[(size_t)f + 1] = c.off[(size_t)f] + (int64_t)cuts[(size_t)f].size();
        c.bin_off[(size_t)f + 1] = c.bin_off[(size_t)f] + (int64_t)cuts[(size_t)f].size() + 1;
    }
    c.vals.resize((size_t)c.off[(size_t)dim]);
    for (int64_t f = 0; f < dim; ++f) {
        std::copy(cuts[(size_t)f].begin(), cuts[(size_t)f].end(), c.vals.begin() + c.off[(size_t)f]);
        c.zero_bin[(size_t)f] = c.bin_of((int32_t)f, 0.f);
    }
    c.total_bins = c.bin_off[(size_t)dim];
    return c;
}

struct Trainer {
    const Csr& X;
    int64_t dim;
    const GbdtParams& p;
    const Cuts& cuts;
    const std::vector<uint8_t>& bins;    // bin of every stored non-zero, aligned with X.vals
    allreduce_fn ar;
    int D;                               // gradient dimension of this tree
    const double* g;                     // [n, D]
    const double* h;
    std::vector<char> feat_on;

    double thr_l1(double G) const { return G > p.reg_alpha ? G - p.reg_alpha : (G < -p.reg_alpha ? G + p.reg_alpha : 0.0); }
    double score(const double* G, const double* H) const {
        double s = 0;
        for (int d = 0; d < D; ++d) { double t = thr_l1(G[d]); s += t * t / (H[d] + p.reg_lambda + 1e-12); }
        return s;
    }
    void leaf_value(const double* G, const double* H, float* out, double scale) const {
        for (int d = 0; d < D; ++d) {
            double w = -thr_l1(G[d]) / (H[d] + p.reg_lambda + 1e-12);
            if (p.max_leaf_weight > 0) w = std::max(-p.max_leaf_weight, std::min(p.max_leaf_weight, w));
            out[d] = (float)(w * scale);
        }
    }
    // histogram layout: [total_bins][D][2]; the two trailing slots hold the node totals [D][2] and count
    size_t hist_size() const { return (size_t)cuts.total_bins * D * 2 + (size_t)D * 2 + 1; }

    void build_hist(const std::vector<int64_t>& ins, std::vector<double>& hist) const {
        size_t hs = hist_size();
        hist.assign(hs, 0.0);
        int nt = std::max(1, std::min(omp_get_max_threads(), (int)(ins.size() / 2048 + 1)));
        std::vector<std::vector<double>> local((size_t)nt);
#pragma omp parallel num_threads(nt)
        {
            int t = omp_get_thread_num();
            std::vector<double>& H = t == 0 ? hist : local[(size_t)t];
            if (t != 0) H.assign(hs, 0.0);
            double* tot = &H[(size_t)cuts.total_bins * D * 2];
#pragma omp for schedule(static)
            for (int64_t k = 0; k < (int64_t)ins.size(); ++k) {
                int64_t i = ins[(size_t)k];
                const double* gi = g + (size_t)i * D;
                const double* hi = h + (size_t)i * D;
                for (int d = 0; d < D; ++d) { tot[d * 2] += gi[d]; tot[d * 2 + 1] += hi[d]; }
                tot[D * 2] += 1.0;
                for (int64_t q = X.indptr[i]; q < X.indptr[i + 1]; ++q) {
                    int32_t f = X.indices[q];
                    if (f >= dim || !feat_on[(size_t)f]) continue;
                    double* slot = &H[(size_t)(cuts.bin_off[(size_t)f] + bins[(size_t)q]) * D * 2];
                    for (int d = 0; d < D; ++d) { slot[d * 2] += gi[d]; slot[d * 2 + 1] += hi[d]; }
                }
            }
        }
        for (int t = 1; t < nt; ++t)
            if (local[(size_t)t].size() == hs)
                for (size_t j = 0; j < hs; ++j) hist[j] += local[(size_t)t][j];
        if (ar) ar(hist.data(), (int64_t)hs);
        // rows whose feature f is zero were never visited: their mass is the node total minus the visited mass
        const double* tot = &hist[(size_t)cuts.total_bins * D * 2];
#pragma omp parallel for schedule(static)
        for (int64_t f = 0; f < dim; ++f) {
            if (!feat_on[(size_t)f]) continue;
            int64_t b0 = cuts.bin_off[(size_t)f], b1 = cuts.bin_off[(size_t)f + 1];
            if (b1 - b0 < 2) continue;
            for (int d = 0; d < D; ++d) {
                double sg = 0, sh = 0;
                for (int64_t b = b0; b < b1; ++b) { sg += hist[(size_t)b * D * 2 + d * 2]; sh += hist[(size_t)b * D * 2 + d * 2 + 1]; }
                double* z = &hist[(size_t)(b0 + cuts.zero_bin[(size_t)f]) * D * 2];
                z[d * 2] += tot[d * 2] - sg;
                z[d * 2 + 1] += tot[d * 2 + 1] - sh;
            }
        }
    }

    struct Split { double gain = 0; int32_t feat = -1; int bin = 0; };
    Split best_split(const std::vector<double>& hist) const {
        const double* tot = &hist[(size_t)cuts.total_bins * D * 2];
        std::vector<double> G((size_t)D), H((size_t)D);
        for (int d = 0; d < D; ++d) { G[(size_t)d] = tot[d * 2]; H[(size_t)d] = tot[d * 2 + 1]; }
        double parent = score(G.data(), H.data());
        Split best;
#pragma omp parallel
        {
            Split mine;
            std::vector<double> GL((size_t)D), HL((size_t)D), GR((size_t)D), HR((size_t)D);
#pragma omp for schedule(dynamic, 64) nowait
            for (int64_t f = 0; f < dim; ++f) {
                if (!feat_on[(size_t)f]) continue;
                int64_t b0 = cuts.bin_off[(size_t)f], b1 = cuts.bin_off[(size_t)f + 1];
                if (b1 - b0 < 2) continue;
                std::fill(GL.begin(), GL.end(), 0.0); std::fill(HL.begin(), HL.end(), 0.0);
                for (int64_t b = b0; b + 1 < b1; ++b) {
                    double minl = 1e300, minr = 1e300;
                    for (int d = 0; d < D; ++d) {
                        GL[(size_t)d] += hist[(size_t)b * D * 2 + d * 2];
                        HL[(size_t)d] += hist[(size_t)b * D * 2 + d * 2 + 1];
                        GR[(size_t)d] = G[(size_t)d] - GL[(size_t)d];
                        HR[(size_t)d] = H[(size_t)d] - HL[(size_t)d];
                        minl = std::min(minl, HL[(size_t)d]); minr = std::min(minr, HR[(size_t)d]);
                    }
                    if (minl < p.min_child_weight || minr < p.min_child_weight) continue;
                    if (minl <= 0 || minr <= 0) continue;
                    double gain = 0.5 * (score(GL.data(), HL.data()) + score(GR.data(), HR.data()) - parent);
                    if (gain > mine.gain || (gain == mine.gain && mine.feat >= 0 && f < mine.feat)) { mine.gain = gain; mine.feat = (int32_t)f; mine.bin = (int)(b - b0); }
                }
            }
#pragma omp critical
            if (mine.feat >= 0 && (mine.gain > best.gain || (mine.gain == best.gain && (best.feat < 0 || mine.feat < best.feat)))) best = mine;
        }
        if (best.gain <= p.min_split_gain) best.feat = -1;
        return best;
    }

    struct Cand { int node; int depth; std::vector<int64_t> ins; std::vector<double> hist; Split sp; };

    Tree grow(const std::vector<int64_t>& root_ins, double leaf_scale) {
        Tree t;
        t.out_dim = D;
        t.nodes.emplace_back();
        auto cmp = [&](const std::shared_ptr<Cand>& a, const std::shared_ptr<Cand>& b) {
            if (p.leafwise) return a->sp.gain < b->sp.gain;
            return a->node > b->node;                     // level-wise = breadth-first by node id
        };
        std::priority_queue<std::shared_ptr<Cand>, std::vector<std::shared_ptr<Cand>>, decltype(cmp)> open(cmp);
        auto finish_leaf = [&](Cand& c) {
            const double* tot = &c.hist[(size_t)cuts.total_bins * D * 2];
            std::vector<double> G((size_t)D), H((size_t)D);
            for (int d = 0; d < D; ++d) { G[(size_t)d] = tot[d * 2]; H[(size_t)d] = tot[d * 2 + 1]; }
            t.nodes[(size_t)c.node].leaf = (int32_t)(t.leaf_vals.size() / (size_t)D);
            t.leaf_vals.resize(t.leaf_vals.size() + (size_t)D);
            leaf_value(G.data(), H.data(), &t.leaf_vals[t.leaf_vals.size() - (size_t)D], leaf_scale);
        };
        auto consider = [&](std::shared_ptr<Cand> c) {
            double cnt = c->hist[hist_size() - 1];        // global instance count of the node
            bool can = c->depth < p.max_depth && cnt >= (double)p.min_node_ins && cnt >= 2;
            if (can) c->sp = best_split(c->hist);
            if (!can || c->sp.feat < 0) finish_leaf(*c); else open.push(c);
        };
        auto root = std::make_shared<Cand>();
        root->node = 0; root->depth = 0; root->ins = root_ins;
        build_hist(root->ins, root->hist);
        consider(root);
        while (!open.empty()) {
            auto c = open.top(); open.pop();
            if ((int)t.nodes.size() + 2 > p.max_node_num) { finish_leaf(*c); continue; }
            int32_t f = c->sp.feat;
            float thr = cuts.vals[(size_t)(cuts.off[(size_t)f] + c->sp.bin)];
            auto L = std::make_shared<Cand>(), R = std::make_shared<Cand>();
            for (int64_t i : c->ins) {
                float x = 0.f;
                const int32_t* b = X.indices + X.indptr[i];
                const int32_t* e = X.indices + X.indptr[i + 1];
                const int32_t* it = std::lower_bound(b, e, f);
                if (it != e && *it == f) x = X.vals[it - X.indices];
                (x <= thr ? L : R)->ins.push_back(i);
            }
            int li = (int)t.nodes.size();
            t.nodes.emplace_back(); t.nodes.emplace_back();
            TreeNode& nd = t.nodes[(size_t)c->node];
            nd.feat = f; nd.thr = thr; nd.left = li; nd.right = li + 1;
            L->node = li; R->node = li + 1; L->depth = R->depth = c->depth + 1;
            // build the smaller child's histogram, derive the sibling by subtraction.  "Smaller" must
            // be agreed on by all workers, so in distributed mode the left child is always built.
            bool left_small = ar ? true : L->ins.size() <= R->ins.size();
            Cand& small = left_small ? *L : *R;
            Cand& big = left_small ? *R : *L;
            build_hist(small.ins, small.hist);
            big.hist.resize(c->hist.size());
            for (size_t j = 0; j < c->hist.size(); ++j) big.hist[j] = c->hist[j] - small.hist[j];
            c->hist.clear(); c->hist.shrink_to_fit(); c->ins.clear();
            consider(L);
            consider(R);
        }
        return t;
    }
};

static inline void softmax_inplace(double* s, int k) {
    double m = *std::max_element(s, s + k), z = 0;
    for (int j = 0; j < k; ++j) { s[j] = std::exp(s[j] - m); z += s[j]; }
    for (int j = 0; j < k; ++j) s[j] /= z;
}

static void gbdt_fit(Gbdt& M, const Csr& X, const float* y, int64_t dim, allreduce_fn ar, const double* range) {
    const GbdtParams& p = M.p;
    M.dim = dim;
    M.trees.clear(); M.tree_class.clear(); M.round_metric.clear();
    int K = (!p.is_regression && p.num_label > 2) ? p.num_label : 1;
    M.score_dim = K;
    Cuts cuts = build_cuts(X, dim, p.num_split, range);
    std::vector<uint8_t> bins((size_t)X.indptr[X.n]);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < X.n; ++i)
        for (int64_t q = X.indptr[i]; q < X.indptr[i + 1]; ++q)
            bins[(size_t)q] = X.indices[q] < dim ? (uint8_t)cuts.bin_of(X.indices[q], X.vals[q]) : 0;
    std::vector<double> score((size_t)X.n * K, 0.0), g((size_t)X.n * K), h((size_t)X.n * K);
    std::mt19937_64 rng(p.seed);
    std::uniform_real_distribution<double> U(0, 1);
    double lr = p.majority_voting ? 1.0 : p.learning_rate;
    for (int r = 0; r < p.num_round; ++r) {
        // gradients of the loss at the current scores (majority voting: always at zero scores)
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < X.n; ++i) {
            if (K == 1) {
                double s = p.majority_voting ? 0.0 : score[(size_t)i];
                if (p.is_regression) { g[(size_t)i] = s - y[i]; h[(size_t)i] = 1.0; }
                else { double pr = 1.0 / (1.0 + std::exp(-s)); g[(size_t)i] = pr - y[i]; h[(size_t)i] = std::max(pr * (1 - pr), 1e-16); }
            } else {
                double pr[256];
                for (int k = 0; k < K; ++k) pr[k] = p.majority_voting ? 0.0 : score[(size_t)i * K + k];
                softmax_inplace(pr, K);
                int yi = (int)y[i];
                for (int k = 0; k < K; ++k) {
                    g[(size_t)i * K + k] = pr[k] - (k == yi ? 1.0 : 0.0);
                    h[(size_t)i * K + k] = std::max(pr[k] * (1 - pr[k]), 1e-16);
                }
            }
        }
        std::vector<int64_t> ins;
        for (int64_t i = 0; i < X.n; ++i) if (p.ins_sp_ratio >= 1 || U(rng) < p.ins_sp_ratio) ins.push_back(i);
        std::vector<char> feat_on((size_t)dim, 1);
        if (p.feat_sp_ratio < 1) {
            std::mt19937_64 frng(p.seed * 7919 + (uint64_t)r);        // same mask on every worker
            for (int64_t f = 0; f < dim; ++f) feat_on[(size_t)f] = U(frng) < p.feat_sp_ratio;
        }
        int ntree = (K > 1 && p.multi_tree) ? K : 1;
        int D = (K > 1 && !p.multi_tree) ? K : 1;
        for (int tc = 0; tc < ntree; ++tc) {
            std::vector<double> gs, hs;
            const double* gp = g.data();
            const double* hp = h.data();
            if (K > 1 && p.multi_tree) {
                gs.resize((size_t)X.n); hs.resize((size_t)X.n);
                for (int64_t i = 0; i < X.n; ++i) { gs[(size_t)i] = g[(size_t)i * K + tc]; hs[(size_t)i] = h[(size_t)i * K + tc]; }
                gp = gs.data(); hp = hs.data();
            }
            Trainer T{X, dim, p, cuts, bins, ar, D, gp, hp, feat_on};
            Tree t = T.grow(ins, lr);
            // every instance (sampled or not) moves by the new tree
#pragma omp parallel for schedule(static)
            for (int64_t i = 0; i < X.n; ++i) {
                const float* lv = t.eval(X.indices + X.indptr[i], X.vals + X.indptr[i], X.indptr[i + 1] - X.indptr[i], dim);
                if (K > 1 && p.multi_tree) score[(size_t)i * K + tc] += lv[0];
                else for (int d = 0; d < D; ++d) score[(size_t)i * K + d] += lv[d];
            }
            M.trees.push_back(std::move(t));
            M.tree_class.push_back((K > 1 && p.multi_tree) ? tc : -1);
        }
        double loss = 0, cnt = (double)X.n;
        double avg = p.majority_voting ? 1.0 / (r + 1) : 1.0;
#pragma omp parallel for schedule(static) reduction(+ : loss)
        for (int64_t i = 0; i < X.n; ++i) {
            if (K == 1) {
                double s = score[(size_t)i] * avg;
                if (p.is_regression) loss += (s - y[i]) * (s - y[i]);
                else { double pr = std::min(std::max(1.0 / (1.0 + std::exp(-s)), 1e-15), 1 - 1e-15); loss += y[i] > 0.5 ? -std::log(pr) : -std::log(1 - pr); }
            } else {
                double pr[256];
                for (int k = 0; k < K; ++k) pr[k] = score[(size_t)i * K + k] * avg;
                softmax_inplace(pr, K);
                loss += -std::log(std::max(pr[(int)y[i]], 1e-15));
            }
        }
        if (ar) { double b[2] = {loss, cnt}; ar(b, 2); loss = b[0]; cnt = b[1]; }
        M.round_metric.push_back(loss / std::max(cnt, 1.0));
    }
}

static void gbdt_predict(const Gbdt& M, const Csr& X, float* out) {
    int K = M.score_dim;
    double avg = M.p.majority_voting && !M.trees.empty() ? 1.0 / ((double)M.trees.size() / ((K > 1 && M.p.multi_tree) ? K : 1)) : 1.0;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < X.n; ++i) {
        double s[256];
        std::fill(s, s + K, 0.0);
        for (size_t t = 0; t < M.trees.size(); ++t) {
            const Tree& tr = M.trees[t];
            const float* lv = tr.eval(X.indices + X.indptr[i], X.vals + X.indptr[i], X.indptr[i + 1] - X.indptr[i], M.dim);
            if (M.tree_class[t] >= 0) s[M.tree_class[t]] += lv[0];
            else for (int d = 0; d < tr.out_dim; ++d) s[d] += lv[d];
        }
        for (int k = 0; k < K; ++k) out[(size_t)i * K + k] = (float)(s[k] * avg);
    }
}

static void gbdt_save(const Gbdt& M, const char* path) {
    FILE* f = fopen(path, "w");
    if (!f) throw std::runtime_error(std::string("cannot open ") + path);
    const GbdtParams& p = M.p;
    fprintf(f, "hetuml-gbdt 1\n%lld %d %zu\n", (long long)M.dim, M.score_dim, M.trees.size());
    fprintf(f, "%d %d %d %d %.17g %.17g %d %d %d %.17g %.17g %d %.17g %.17g %.17g %.17g %d %d %llu\n", p.max_depth, p.max_node_num,
            p.leafwise, p.num_split, p.ins_sp_ratio, p.feat_sp_ratio, p.is_regression, p.num_label, p.num_round, p.learning_rate,
            p.min_split_gain, p.min_node_ins, p.min_child_weight, p.reg_alpha, p.reg_lambda, p.max_leaf_weight, p.multi_tree,
            p.majority_voting, (unsigned long long)p.seed);
    for (size_t t = 0; t < M.trees.size(); ++t) {
        const Tree& tr = M.trees[t];
        fprintf(f, "%d %d %zu %zu\n", M.tree_class[t], tr.out_dim, tr.nodes.size(), tr.leaf_vals.size());
        for (auto& n : tr.nodes) fprintf(f, "%d %.9g %d %d %d\n", n.feat, n.thr, n.left, n.right, n.leaf);
        for (float v : tr.leaf_vals) fprintf(f, "%.9g ", v);
        fprintf(f, "\n");
    }
    fclose(f);
}

static void gbdt_load(Gbdt& M, const char* path) {
    FILE* f = fopen(path, "r");
    if (!f) throw std::runtime_error(std::string("cannot open ") + path);
    char magic[32]; int ver; long long dim; size_t nt;
    bool ok = fscanf(f, "%31s %d %lld %d %zu", magic, &ver, &dim, &M.score_dim, &nt) == 5 && !strcmp(magic, "hetuml-gbdt");
    GbdtParams& p = M.p;
    unsigned long long seed = 0;
    ok = ok && fscanf(f, "%d %d %d %d %lf %lf %d %d %d %lf %lf %d %lf %lf %lf %lf %d %d %llu", &p.max_depth, &p.max_node_num, &p.leafwise,
                      &p.num_split, &p.ins_sp_ratio, &p.feat_sp_ratio, &p.is_regression, &p.num_label, &p.num_round, &p.learning_rate,
                      &p.min_split_gain, &p.min_node_ins, &p.min_child_weight, &p.reg_alpha, &p.reg_lambda, &p.max_leaf_weight,
                      &p.multi_tree, &p.majority_voting, &seed) == 19;
    p.seed = seed;
    M.dim = dim;
    M.trees.assign(ok ? nt : 0, Tree());
    M.tree_class.assign(ok ? nt : 0, -1);
    for (size_t t = 0; ok && t < nt; ++t) {
        Tree& tr = M.trees[t];
        size_t nn, nl;
        ok = fscanf(f, "%d %d %zu %zu", &M.tree_class[t], &tr.out_dim, &nn, &nl) == 4;
        if (!ok) break;
        tr.nodes.resize(nn); tr.leaf_vals.resize(nl);
        for (auto& n : tr.nodes) ok = ok && fscanf(f, "%d %f %d %d %d", &n.feat, &n.thr, &n.left, &n.right, &n.leaf) == 5;
        for (auto& v : tr.leaf_vals) ok = ok && fscanf(f, "%f", &v) == 1;
    }
    fclose(f);
    if (!ok) throw std::runtime_error(std::string("not a hetuml GBDT model file: ") + path);
}

// --------------------------------------------------------------------------------------- knn
// scores[q, label] = number of the k nearest labelled rows (squared Euclidean distance) carrying
// that label.  reference: include/model/knn/knn.h.
static void knn_predict(const Csr& L, const float* labels, const Csr& Q, int64_t dim, int k, int num_label, float* out) {
    std::vector<float> lnorm((size_t)L.n);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < L.n; ++i) {
        float s = 0;
        for (int64_t p = L.indptr[i]; p < L.indptr[i + 1]; ++p) s += L.vals[p] * L.vals[p];
        lnorm[(size_t)i] = s;
    }
#pragma omp parallel
    {
        std::vector<float> dense((size_t)dim, 0.f);
        std::vector<std::pair<float, int64_t>> dist((size_t)L.n);
#pragma omp for schedule(dynamic, 8)
        for (int64_t q = 0; q < Q.n; ++q) {
            float qn = 0;
            for (int64_t p = Q.indptr[q]; p < Q.indptr[q + 1]; ++p)
                if (Q.indices[p] < dim) { dense[(size_t)Q.indices[p]] = Q.vals[p]; qn += Q.vals[p] * Q.vals[p]; }
            for (int64_t i = 0; i < L.n; ++i) {
                float dot = 0;
                for (int64_t p = L.indptr[i]; p < L.indptr[i + 1]; ++p)
                    if (L.indices[p] < dim) dot += dense[(size_t)L.indices[p]] * L.vals[p];
                dist[(size_t)i] = {qn + lnorm[(size_t)i] - 2 * dot, i};
            }
            int kk = (int)std::min<int64_t>(k, L.n);
            std::partial_sort(dist.begin(), dist.begin() + kk, dist.end());
            float* o = out + (size_t)q * num_label;
            std::fill(o, o + num_label, 0.f);
            for (int j = 0; j < kk; ++j) {
                int lab = (int)labels[dist[(size_t)j].second];
                if (lab >= 0 && lab < num_label) o[lab] += 1.f;
            }
            for (int64_t p = Q.indptr[q]; p < Q.indptr[q + 1]; ++p)
                if (Q.indices[p] < dim) dense[(size_t)Q.indices[p]] = 0.f;
        }
    }
}

// ------------------------------------------------------------------------------- naive Bayes
// Gaussian likelihood per (feature, label) over the stored (non-zero) occurrences, class prior from
// label frequencies.  stats layout: count[dim*K], sum[dim*K], sumsq[dim*K], class_count[K] - additive,
// so distributed training is one all-reduce of `stats`.  reference: include/model/naive_bayes/naive_bayes.h.
static void nb_stats(const Csr& X, const float* y, int64_t dim, int K, double* stats) {
    double* cnt = stats; double* sum = stats + (size_t)dim * K; double* sq = stats + (size_t)dim * K * 2; double* cc = stats + (size_t)dim * K * 3;
    for (int64_t i = 0; i < X.n; ++i) {
        int k = (int)y[i];
        if (k < 0 || k >= K) continue;
        cc[k] += 1;
        for (int64_t p = X.indptr[i]; p < X.indptr[i + 1]; ++p) {
            int32_t f = X.indices[p];
            if (f >= dim) continue;
            double v = X.vals[p];
            cnt[(size_t)f * K + k] += 1; sum[(size_t)f * K + k] += v; sq[(size_t)f * K + k] += v * v;
        }
    }
}
// model layout: prior[K], mean[dim*K], var[dim*K]
static void nb_finalize(const double* stats, int64_t dim, int K, float* model) {
    const double* cnt = stats; const double* sum = stats + (size_t)dim * K; const double* sq = stats + (size_t)dim * K * 2; const double* cc = stats + (size_t)dim * K * 3;
    double total = 0;
    for (int k = 0; k < K; ++k) total += cc[k];
    for (int k = 0; k < K; ++k) model[k] = (float)(cc[k] / std::max(total, 1.0));
    float* mean = model + K; float* var = model + K + (size_t)dim * K;
    for (size_t j = 0; j < (size_t)dim * K; ++j) {
        double c = cnt[j];
        mean[j] = c > 0 ? (float)(sum[j] / c) : 0.f;
        var[j] = c > 0 ? (float)std::max(sq[j] / c - (sum[j] / c) * (sum[j] / c), 0.0) : 0.f;
    }
}
// out[n, K] = log prior + sum over stored features of log N(x; mean, var); features with (near) zero
// variance are skipped, likelihoods are floored at 1e-8 like the reference's EPSILON clamp.
static void nb_predict(const Csr& X, int64_t dim, int K, const float* model, float* out) {
    const float* mean = model + K; const float* var = model + K + (size_t)dim * K;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < X.n; ++i) {
        float* o = out + (size_t)i * K;
        for (int k = 0; k < K; ++k) o[k] = std::log(std::max(model[k], 1e-8f));
        for (int64_t p = X.indptr[i]; p < X.indptr[i + 1]; ++p) {
            int32_t f = X.indices[p];
            if (f >= dim) continue;
            for (int k = 0; k < K; ++k) {
                float v = var[(size_t)f * K + k];
                if (v < 1e-8f) continue;
                float d = X.vals[p] - mean[(size_t)f * K + k];
                float lp = -0.5f * std::log(6.2831853f * v) - d * d / (2 * v);
                o[k] += std::max(lp, std::log(1e-8f));
            }
        }
    }
}

// ----------------------------------------------------------------------- matrix factorization
// R[r, c] ~ P[r] . Q[c], lock-free parallel SGD over the shuffled non-zeros, factors kept
// non-negative (NMF).  reference: include/model/mf/* (LIBMF-style blocked SGD).
static void mf_fit(const int32_t* rows, const int32_t* cols, const float* vals, int64_t nnz, int k, float* P, float* Q,
                   int epochs, float lr, float l1, float l2, uint64_t seed, double* rmse_out) {
    std::vector<int64_t> order((size_t)nnz);
    std::iota(order.begin(), order.end(), 0);
    std::mt19937_64 rng(seed);
    for (int ep = 0; ep < epochs; ++ep) {
        std::shuffle(order.begin(), order.end(), rng);
        double se = 0;
#pragma omp parallel for schedule(static) reduction(+ : se)
        for (int64_t t = 0; t < nnz; ++t) {
            int64_t e = order[(size_t)t];
            float* p = P + (size_t)rows[e] * k;
            float* q = Q + (size_t)cols[e] * k;
            float pred = 0;
            for (int j = 0; j < k; ++j) pred += p[j] * q[j];
            float err = pred - vals[e];
            se += (double)err * err;
            for (int j = 0; j < k; ++j) {
                float pj = p[j], qj = q[j];
                p[j] = std::max(0.f, pj - lr * (err * qj + l2 * pj + l1));
                q[j] = std::max(0.f, qj - lr * (err * pj + l2 * qj + l1));
            }
        }
        if (rmse_out) rmse_out[ep] = std::sqrt(se / std::max<double>(1.0, (double)nnz));
    }
}

// --------------------------------------------------------------------------------------- LDA
// Collapsed Gibbs sampling.  nwk [V, K] word-topic counts, nk [K]; z [tokens] topic of every token.
// train = 1: fresh start (the counts do not contain these tokens yet), train = 2: continue from the
// assignments in z (the counts already contain them), train = 0: counts held fixed (inference on
// new documents).
// Returns the joint log likelihood of words given topics plus topics given documents.
// reference: include/model/lda/lda.h (sparse F+tree sampler; the sampler differs, the model and
// its hyper-parameters alpha / beta are the same).
static double lda_run(const int64_t* doc_ptr, const int32_t* words, int64_t ndocs, int V, int K, int iters, double alpha,
                      double beta, int32_t* z, int32_t* nwk, int32_t* nk, int train, uint64_t seed) {
    std::mt19937_64 rng(seed);
    std::uniform_real_distribution<double> U(0, 1);
    std::vector<int32_t> ndk((size_t)K);
    std::vector<double> pr((size_t)K);
    int64_t ntok = doc_ptr[ndocs];
    // initial assignment
    for (int64_t t = 0; t < ntok; ++t) {
        if (z[t] < 0 || z[t] >= K) z[t] = (int32_t)(rng() % (uint64_t)K);
        if (train == 1) { nwk[(size_t)words[t] * K + z[t]]++; nk[z[t]]++; }
    }
    double Vb = V * beta;
    for (int it = 0; it < iters; ++it) {
        for (int64_t d = 0; d < ndocs; ++d) {
            std::fill(ndk.begin(), ndk.end(), 0);
            for (int64_t t = doc_ptr[d]; t < doc_ptr[d + 1]; ++t) ndk[(size_t)z[t]]++;
            for (int64_t t = doc_ptr[d]; t < doc_ptr[d + 1]; ++t) {
                int w = words[t], old = z[t];
                ndk[(size_t)old]--;
                if (train) { nwk[(size_t)w * K + old]--; nk[old]--; }
                double tot = 0;
                const int32_t* nw = nwk + (size_t)w * K;
                for (int k = 0; k < K; ++k) { tot += (ndk[(size_t)k] + alpha) * (nw[k] + beta) / (nk[k] + Vb); pr[(size_t)k] = tot; }
                double u = U(rng) * tot;
                int nz = (int)(std::lower_bound(pr.begin(), pr.end(), u) - pr.begin());
                if (nz >= K) nz = K - 1;
                z[t] = nz;
                ndk[(size_t)nz]++;
                if (train) { nwk[(size_t)w * K + nz]++; nk[nz]++; }
            }
        }
    }
    double llh = 0;
    for (int64_t d = 0; d < ndocs; ++d) {
        std::fill(ndk.begin(), ndk.end(), 0);
        int64_t len = doc_ptr[d + 1] - doc_ptr[d];
        for (int64_t t = doc_ptr[d]; t < doc_ptr[d + 1]; ++t) ndk[(size_t)z[t]]++;
        for (int64_t t = doc_ptr[d]; t < doc_ptr[d + 1]; ++t) {
            int k = z[t];
            llh += std::log((nwk[(size_t)words[t] * K + k] + beta) / (nk[k] + Vb)) + std::log((ndk[(size_t)k] + alpha) / (len + K * alpha));
        }
    }
    return llh;
}

static void fm_predict(const Csr& X, int64_t dim, int k, const float* w, const float* v, float* out) {
#pragma omp parallel
    {
        std::vector<float> sums((size_t)k);
#pragma omp for schedule(static)
        for (int64_t i = 0; i < X.n; ++i) out[i] = fm_score(X, i, dim, k, w, v, sums.data());
    }
}
static void mf_predict(const int32_t* rows, const int32_t* cols, int64_t nnz, int k, const float* P, const float* Q, float* out) {
#pragma omp parallel for schedule(static)
    for (int64_t e = 0; e < nnz; ++e) {
        float s = 0;
        for (int j = 0; j < k; ++j) s += P[(size_t)rows[e] * k + j] * Q[(size_t)cols[e] * k + j];
        out[e] = s;
    }
}

}  // namespace hml

using namespace hml;

#define HML_TRY(body)                                   \
    try { body; return 0; }                             \
    catch (const std::exception& e) { g_err = e.what(); return -1; }

HML_API const char* hml_last_error() { return g_err.c_str(); }
HML_API int hml_num_threads() { return omp_get_max_threads(); }

HML_API int hml_linear_fit(int loss, const int64_t* indptr, const int32_t* indices, const float* vals, const float* y, int64_t n,
                           int64_t dim, float* w, int epochs, int batch, float lr, float l1, float l2) {
    HML_TRY(linear_fit(loss, Csr{indptr, indices, vals, n}, y, dim, w, epochs, std::max(batch, 1), lr, l1, l2));
}
HML_API int hml_linear_predict(const int64_t* indptr, const int32_t* indices, const float* vals, int64_t n, int64_t dim,
                               const float* w, float* out) {
    HML_TRY(linear_predict(Csr{indptr, indices, vals, n}, dim, w, out));
}
HML_API int hml_fm_fit(const int64_t* indptr, const int32_t* indices, const float* vals, const float* y, int64_t n, int64_t dim,
                       int k, float* w, float* v, int epochs, int batch, float lr, float l1, float l2) {
    HML_TRY(fm_fit(Csr{indptr, indices, vals, n}, y, dim, k, w, v, epochs, std::max(batch, 1), lr, l1, l2));
}
HML_API int hml_fm_predict(const int64_t* indptr, const int32_t* indices, const float* vals, int64_t n, int64_t dim, int k,
                           const float* w, const float* v, float* out) {
    HML_TRY(fm_predict(Csr{indptr, indices, vals, n}, dim, k, w, v, out));
}

// params (doubles, in this order): max_depth, max_node_num, leafwise, num_split, ins_sp_ratio, feat_sp_ratio,
// is_regression, num_label, num_round, learning_rate, min_split_gain, min_node_ins, min_child_weight, reg_alpha,
// reg_lambda, max_leaf_weight, multi_tree, majority_voting, seed
HML_API void* hml_gbdt_create(const double* a, int na) {
    auto* m = new Gbdt();
    if (a && na >= 19) {
        GbdtParams& p = m->p;
        p.max_depth = (int)a[0]; p.max_node_num = (int)a[1]; p.leafwise = (int)a[2]; p.num_split = (int)a[3];
        p.ins_sp_ratio = a[4]; p.feat_sp_ratio = a[5]; p.is_regression = (int)a[6]; p.num_label = (int)a[7];
        p.num_round = (int)a[8]; p.learning_rate = a[9]; p.min_split_gain = a[10]; p.min_node_ins = (int)a[11];
        p.min_child_weight = a[12]; p.reg_alpha = a[13]; p.reg_lambda = a[14]; p.max_leaf_weight = a[15];
        p.multi_tree = (int)a[16]; p.majority_voting = (int)a[17]; p.seed = (uint64_t)a[18];
    }
    return m;
}
HML_API void hml_gbdt_free(void* h) { delete static_cast<Gbdt*>(h); }
HML_API int hml_gbdt_fit(void* h, const int64_t* indptr, const int32_t* indices, const float* vals, const float* y, int64_t n,
                         int64_t dim, allreduce_fn ar, const double* range) {
    HML_TRY({
        Gbdt& M = *static_cast<Gbdt*>(h);
        if (M.p.num_label > 256) throw std::runtime_error("at most 256 labels");
        if (M.p.num_split > 254) throw std::runtime_error("at most 254 candidate splits per feature");
        if (ar && !range) throw std::runtime_error("distributed GBDT needs the agreed feature ranges");
        gbdt_fit(M, Csr{indptr, indices, vals, n}, y, dim, ar, range);
    });
}
HML_API int hml_gbdt_score_dim(void* h) { return static_cast<Gbdt*>(h)->score_dim; }
HML_API int hml_gbdt_num_trees(void* h) { return (int)static_cast<Gbdt*>(h)->trees.size(); }
HML_API int64_t hml_gbdt_max_dim(void* h) { return static_cast<Gbdt*>(h)->dim; }
HML_API int hml_gbdt_round_metric(void* h, double* out, int cap) {
    auto& v = static_cast<Gbdt*>(h)->round_metric;
    for (int i = 0; i < cap && i < (int)v.size(); ++i) out[i] = v[(size_t)i];
    return (int)v.size();
}
HML_API int hml_gbdt_get_params(void* h, double* a) {
    const GbdtParams& p = static_cast<Gbdt*>(h)->p;
    double v[19] = {(double)p.max_depth, (double)p.max_node_num, (double)p.leafwise, (double)p.num_split, p.ins_sp_ratio, p.feat_sp_ratio,
                    (double)p.is_regression, (double)p.num_label, (double)p.num_round, p.learning_rate, p.min_split_gain,
                    (double)p.min_node_ins, p.min_child_weight, p.reg_alpha, p.reg_lambda, p.max_leaf_weight, (double)p.multi_tree,
                    (double)p.majority_voting, (double)p.seed};
    memcpy(a, v, sizeof(v));
    return 19;
}
HML_API int hml_gbdt_predict(void* h, const int64_t* indptr, const int32_t* indices, const float* vals, int64_t n, float* out) {
    HML_TRY(gbdt_predict(*static_cast<Gbdt*>(h), Csr{indptr, indices, vals, n}, out));
}
HML_API int hml_gbdt_save(void* h, const char* path) { HML_TRY(gbdt_save(*static_cast<Gbdt*>(h), path)); }
HML_API int hml_gbdt_load(void* h, const char* path) { HML_TRY(gbdt_load(*static_cast<Gbdt*>(h), path)); }

// Candidate cuts and the bin of every stored entry, for trainers that run elsewhere (the device GBDT):
// off [dim + 1] offsets into cut_vals (capacity `cap`); returns the number of cuts, or -1 - needed when cap is short.
HML_API int64_t hml_gbdt_cuts(const int64_t* indptr, const int32_t* indices, const float* vals, int64_t n, int64_t dim,
                              int num_split, int64_t* off, float* cut_vals, int64_t cap) {
    try {
        Cuts c = build_cuts(Csr{indptr, indices, vals, n}, dim, num_split, nullptr);
        if ((int64_t)c.vals.size() > cap) return -1 - (int64_t)c.vals.size();
        std::copy(c.off.begin(), c.off.end(), off);
        std::copy(c.vals.begin(), c.vals.end(), cut_vals);
        return (int64_t)c.vals.size();
    } catch (const std::exception& e) { g_err = e.what(); return INT64_MIN; }
}
HML_API int hml_gbdt_bins(const int64_t* indptr, const int32_t* indices, const float* vals, int64_t n, int64_t dim,
                          const int64_t* off, const float* cut_vals, uint8_t* bins) {
    HML_TRY({
        for (int64_t i = 0; i < n; ++i)
            for (int64_t q = indptr[i]; q < indptr[i + 1]; ++q) {
                int32_t f = indices[q];
                if (f >= dim) { bins[q] = 0; continue; }
                const float* b = cut_vals + off[f];
                const float* e = cut_vals + off[f + 1];
                bins[q] = (uint8_t)(std::lower_bound(b, e, vals[q]) - b);
            }
    });
}

HML_API int hml_knn_predict(const int64_t* lptr, const int32_t* lidx, const float* lval, const float* labels, int64_t nl,
                            const int64_t* qptr, const int32_t* qidx, const float* qval, int64_t nq, int64_t dim, int k,
                            int num_label, float* out) {
    HML_TRY(knn_predict(Csr{lptr, lidx, lval, nl}, labels, Csr{qptr, qidx, qval, nq}, dim, k, num_label, out));
}
HML_API int hml_nb_stats(const int64_t* indptr, const int32_t* indices, const float* vals, const float* y, int64_t n, int64_t dim,
                         int K, double* stats) {
    HML_TRY(nb_stats(Csr{indptr, indices, vals, n}, y, dim, K, stats));
}
HML_API int hml_nb_finalize(const double* stats, int64_t dim, int K, float* model) { HML_TRY(nb_finalize(stats, dim, K, model)); }
HML_API int hml_nb_predict(const int64_t* indptr, const int32_t* indices, const float* vals, int64_t n, int64_t dim, int K,
                           const float* model, float* out) {
    HML_TRY(nb_predict(Csr{indptr, indices, vals, n}, dim, K, model, out));
}
HML_API int hml_mf_fit(const int32_t* rows, const int32_t* cols, const float* vals, int64_t nnz, int k, float* P, float* Q,
                       int epochs, float lr, float l1, float l2, uint64_t seed, double* rmse_out) {
    HML_TRY(mf_fit(rows, cols, vals, nnz, k, P, Q, epochs, lr, l1, l2, seed, rmse_out));
}
HML_API int hml_mf_predict(const int32_t* rows, const int32_t* cols, int64_t nnz, int k, const float* P, const float* Q, float* out) {
    HML_TRY(mf_predict(rows, cols, nnz, k, P, Q, out));
}
HML_API int hml_lda_run(const int64_t* doc_ptr, const int32_t* words, int64_t ndocs, int V, int K, int iters, double alpha,
                        double beta, int32_t* z, int32_t* nwk, int32_t* nk, int train, uint64_t seed, double* llh) {
    HML_TRY({ double v = lda_run(doc_ptr, words, ndocs, V, K, iters, alpha, beta, z, nwk, nk, train, seed); if (llh) *llh = v; });
}

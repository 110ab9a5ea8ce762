// Address / UB-sanitizer run of the CPU op library apart from the GEMM (cpu_gemm_asan.cc): convolution, pooling,
// batch norm, softmax / cross entropy, pad, concat, transpose, embedding lookup, optimizers, initialisers and the
// BFC allocator.  Buffers are exactly sized (any over-read or over-write at a window / panel edge is caught) and
// the convolution and pooling results are checked against direct loops.
//   g++ -std=c++17 -O1 -g -fsanitize=address,undefined -fopenmp -march=x86-64-v3 tests/native/cpu_ops_asan.cc \
//       hetu_61a7_b200/csrc_cpu/cpu_ops.cc hetu_61a7_b200/csrc_cpu/bfc_allocator.cc -o /tmp/ops_asan
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

extern "C" {
int hetu_cpu_conv2d(const float*, const float*, float*, int, int, int, int, int, int, int, int, int);
int hetu_cpu_conv2d_grad_data(const float*, const float*, float*, int, int, int, int, int, int, int, int, int);
int hetu_cpu_conv2d_grad_filter(const float*, const float*, float*, int, int, int, int, int, int, int, int, int);
int hetu_cpu_pool2d(const float*, float*, int64_t, int, int, int, int, int, int, int);
int hetu_cpu_pool2d_grad(const float*, const float*, const float*, float*, int64_t, int, int, int, int, int, int, int);
int hetu_cpu_batch_norm(const float*, const float*, const float*, float*, float*, float*, float*, float*, int, int, int, float, float,
                        int);
int hetu_cpu_batch_norm_grad(const float*, const float*, const float*, const float*, const float*, float*, float*, float*, int, int,
                             int);
int hetu_cpu_softmax(const float*, float*, int64_t, int64_t);
int hetu_cpu_softmax_cross_entropy(const float*, const float*, float*, int64_t, int64_t);
int hetu_cpu_softmax_cross_entropy_grad(const float*, const float*, const float*, float*, int64_t, int64_t);
int hetu_cpu_pad(const float*, float*, int64_t, int, int, int, int, int, int, float);
int hetu_cpu_pad_grad(const float*, float*, int64_t, int, int, int, int, int, int);
int hetu_cpu_concat(const float*, const float*, float*, int64_t, int64_t, int64_t);
int hetu_cpu_concat_grad(const float*, float*, int64_t, int64_t, int64_t, int64_t);
int hetu_cpu_transpose(const float*, float*, const int64_t*, const int*, int);
int hetu_cpu_broadcast_to(const float*, float*, int64_t, int64_t);
int hetu_cpu_reduce_sum_axis0(const float*, float*, int64_t, int64_t);
int hetu_cpu_embedding_lookup(const float*, const int64_t*, float*, int64_t, int64_t, int64_t);
int hetu_cpu_dropout(const float*, float*, int64_t, float, uint64_t);
int hetu_cpu_sgd_update(float*, const float*, int64_t, float, float);
int hetu_cpu_sgd_sparse_update(float*, const int64_t*, const float*, int64_t, int64_t, float);
int hetu_cpu_momentum_update(float*, const float*, float*, int64_t, float, float, int, float);
int hetu_cpu_adagrad_update(float*, const float*, float*, int64_t, float, float, float);
int hetu_cpu_adam_update(float*, const float*, float*, float*, int64_t, float, float, float, float, int64_t, float, int, float*);
int hetu_cpu_uniform_init(float*, int64_t, float, float, uint64_t);
int hetu_cpu_normal_init(float*, int64_t, float, float, uint64_t);
int hetu_cpu_truncated_normal_init(float*, int64_t, float, float, uint64_t);
void* hetu_bfc_create(uint64_t, uint64_t);
void hetu_bfc_destroy(void*);
int64_t hetu_bfc_alloc(void*, uint64_t);
int hetu_bfc_free(void*, uint64_t);
void hetu_bfc_stats(void*, uint64_t*);
}

static std::mt19937 rng(11);
static std::vector<float> rnd(size_t n) {
    std::uniform_real_distribution<float> U(-1.f, 1.f);
    std::vector<float> v(n);
    for (auto& x : v) x = U(rng);
    return v;
}
static bool close(float a, float b, float tol = 1e-3f) { return std::fabs(a - b) <= tol * (1.f + std::fabs(b)); }
#define CHECK(cond, ...)                                                  \
    do {                                                                  \
        if (!(cond)) { fprintf(stderr, "FAILED %s:%d: ", __FILE__, __LINE__); fprintf(stderr, __VA_ARGS__); fprintf(stderr, "\n"); return 1; } \
    } while (0)

// ---- convolution: forward against a direct loop; both gradients through the adjoint identities
//      <conv(x, w), gy> == <x, grad_data(w, gy)> == <w, grad_filter(x, gy)>
static int test_conv() {
    const int cfgs[][9] = {  // N C H W F kh kw pad stride
        {1, 1, 1, 1, 1, 1, 1, 0, 1}, {2, 3, 7, 5, 4, 3, 3, 1, 1}, {1, 2, 9, 9, 5, 3, 3, 0, 2},
        {2, 1, 8, 6, 3, 5, 5, 2, 1}, {1, 4, 6, 7, 2, 1, 1, 0, 1}, {1, 3, 11, 10, 33, 3, 2, 1, 3}};
    for (auto& c : cfgs) {
        const int N = c[0], C = c[1], H = c[2], W = c[3], F = c[4], kh = c[5], kw = c[6], pad = c[7], st = c[8];
        const int Ho = (H + 2 * pad - kh) / st + 1, Wo = (W + 2 * pad - kw) / st + 1;
        auto x = rnd((size_t)N * C * H * W), w = rnd((size_t)F * C * kh * kw), gy = rnd((size_t)N * F * Ho * Wo);
        std::vector<float> y((size_t)N * F * Ho * Wo), gx((size_t)N * C * H * W), gw((size_t)F * C * kh * kw);
        CHECK(hetu_cpu_conv2d(x.data(), w.data(), y.data(), N, C, H, W, F, kh, kw, pad, st) == 0, "conv2d rc");
        for (int n = 0; n < N; ++n)
            for (int f = 0; f < F; ++f)
                for (int ho = 0; ho < Ho; ++ho)
                    for (int wo = 0; wo < Wo; ++wo) {
                        double s = 0;
                        for (int ci = 0; ci < C; ++ci)
                            for (int i = 0; i < kh; ++i)
                                for (int j = 0; j < kw; ++j) {
                                    int hi = ho * st - pad + i, wi = wo * st - pad + j;
                                    if (hi < 0 || hi >= H || wi < 0 || wi >= W) continue;
                                    s += (double)x[(((size_t)n * C + ci) * H + hi) * W + wi] * w[(((size_t)f * C + ci) * kh + i) * kw + j];
                                }
                        float got = y[(((size_t)n * F + f) * Ho + ho) * Wo + wo];
                        CHECK(close(got, (float)s), "conv2d N=%d C=%d H=%d W=%d F=%d: %f vs %f", N, C, H, W, F, got, s);
                    }
        CHECK(hetu_cpu_conv2d_grad_data(w.data(), gy.data(), gx.data(), N, C, H, W, F, kh, kw, pad, st) == 0, "grad_data rc");
        CHECK(hetu_cpu_conv2d_grad_filter(x.data(), gy.data(), gw.data(), N, C, H, W, F, kh, kw, pad, st) == 0, "grad_filter rc");
        double a = 0, b = 0, d = 0;
        for (size_t k = 0; k < y.size(); ++k) a += (double)y[k] * gy[k];
        for (size_t k = 0; k < x.size(); ++k) b += (double)x[k] * gx[k];
        for (size_t k = 0; k < w.size(); ++k) d += (double)w[k] * gw[k];
        CHECK(close((float)b, (float)a, 2e-3f) && close((float)d, (float)a, 2e-3f), "conv adjoint %f %f %f", a, b, d);
    }
    return 0;
}

static int test_pool() {
    const int cfgs[][7] = {  // nc H W kh kw pad stride
        {1, 1, 1, 1, 1, 0, 1}, {3, 7, 5, 2, 2, 0, 2}, {2, 8, 8, 3, 3, 1, 2}, {5, 6, 9, 3, 2, 1, 1}, {2, 5, 5, 5, 5, 2, 3}};
    for (auto& c : cfgs)
        for (int mode = 0; mode < 2; ++mode) {
            const int nc = c[0], H = c[1], W = c[2], kh = c[3], kw = c[4], pad = c[5], st = c[6];
            const int Ho = (H + 2 * pad - kh) / st + 1, Wo = (W + 2 * pad - kw) / st + 1;
            auto x = rnd((size_t)nc * H * W), gy = rnd((size_t)nc * Ho * Wo);
            std::vector<float> y((size_t)nc * Ho * Wo), gx((size_t)nc * H * W);
            CHECK(hetu_cpu_pool2d(x.data(), y.data(), nc, H, W, kh, kw, pad, st, mode) == 0, "pool rc");
            CHECK(hetu_cpu_pool2d_grad(x.data(), y.data(), gy.data(), gx.data(), nc, H, W, kh, kw, pad, st, mode) == 0, "pool grad rc");
            for (float v : y) CHECK(std::isfinite(v), "pool mode %d produced a non-finite value", mode);
            // every output gradient lands somewhere exactly once (max) / is spread with weight 1/(kh*kw) per covered cell (avg)
            double sg = 0, sgx = 0;
            for (int p = 0; p < nc; ++p)
                for (int ho = 0; ho < Ho; ++ho)
                    for (int wo = 0; wo < Wo; ++wo) {
                        int covered = 0;
                        for (int i = 0; i < kh; ++i)
                            for (int j = 0; j < kw; ++j) {
                                int hi = ho * st - pad + i, wi = wo * st - pad + j;
                                covered += hi >= 0 && hi < H && wi >= 0 && wi < W;
                            }
                        sg += (double)gy[((size_t)p * Ho + ho) * Wo + wo] * (mode == 0 ? 1.0 : (double)covered / (kh * kw));
                    }
            for (float v : gx) sgx += v;
            CHECK(close((float)sgx, (float)sg, 2e-3f), "pool grad mass mode %d: %f vs %f", mode, sgx, sg);
        }
    return 0;
}

static int test_norm_softmax() {
    const int N = 3, C = 5, HW = 7;
    auto x = rnd((size_t)N * C * HW), gy = rnd((size_t)N * C * HW), scale = rnd(C), bias = rnd(C);
    std::vector<float> y(x.size()), gx(x.size()), rm(C, 0.f), rv(C, 1.f), sm(C), si(C), gs(C), gb(C);
    CHECK(hetu_cpu_batch_norm(x.data(), scale.data(), bias.data(), y.data(), rm.data(), rv.data(), sm.data(), si.data(), N, C, HW, 0.1f,
                              1e-5f, 1) == 0, "bn rc");
    for (int c = 0; c < C; ++c) {                         // normalised output has the affine's mean / spread per channel
        double s = 0;
        for (int n = 0; n < N; ++n)
            for (int i = 0; i < HW; ++i) s += y[((size_t)n * C + c) * HW + i];
        CHECK(close((float)(s / (N * HW)), bias[c], 1e-3f), "bn mean of channel %d", c);
    }
    CHECK(hetu_cpu_batch_norm_grad(x.data(), gy.data(), scale.data(), sm.data(), si.data(), gx.data(), gs.data(), gb.data(), N, C, HW) == 0,
          "bn grad rc");
    for (int c = 0; c < C; ++c) {                         // the input gradient of a batch-normalised channel sums to zero
        double s = 0;
        for (int n = 0; n < N; ++n)
            for (int i = 0; i < HW; ++i) s += gx[((size_t)n * C + c) * HW + i];
        CHECK(std::fabs(s) < 1e-3, "bn grad sum of channel %d = %g", c, s);
    }
    CHECK(hetu_cpu_batch_norm(x.data(), scale.data(), bias.data(), y.data(), rm.data(), rv.data(), nullptr, nullptr, N, C, HW, 0.1f, 1e-5f,
                              0) == 0, "bn inference rc");
    for (int64_t cols : {1, 3, 10, 17, 1000}) {
        const int64_t rows = 5;
        auto z = rnd((size_t)rows * cols);
        for (auto& v : z) v *= 30.f;                      // large logits: the max subtraction has to hold
        std::vector<float> p(z.size()), lab(z.size(), 0.f), loss(rows), g(rows, 1.f), gz(z.size());
        for (int64_t r = 0; r < rows; ++r) lab[r * cols + (r % cols)] = 1.f;
        CHECK(hetu_cpu_softmax(z.data(), p.data(), rows, cols) == 0, "softmax rc");
        CHECK(hetu_cpu_softmax_cross_entropy(z.data(), lab.data(), loss.data(), rows, cols) == 0, "sce rc");
        CHECK(hetu_cpu_softmax_cross_entropy_grad(z.data(), lab.data(), g.data(), gz.data(), rows, cols) == 0, "sce grad rc");
        for (int64_t r = 0; r < rows; ++r) {
            double s = 0;
            for (int64_t j = 0; j < cols; ++j) s += p[r * cols + j];
            CHECK(close((float)s, 1.f) && std::isfinite(loss[r]) && loss[r] >= -1e-5f, "softmax row %ld cols %ld", (long)r, (long)cols);
            CHECK(close(loss[r], -std::log(std::max(p[r * cols + (r % cols)], 1e-37f)), 1e-2f), "sce row %ld: %f", (long)r, loss[r]);
            for (int64_t j = 0; j < cols; ++j)
                CHECK(close(gz[r * cols + j], p[r * cols + j] - lab[r * cols + j]), "sce grad row %ld col %ld", (long)r, (long)j);
        }
    }
    return 0;
}

static int test_shape_ops() {
    {   // pad and its gradient are adjoint; the border carries the fill value
        const int nc = 3, H = 4, W = 5, pt = 1, pb = 2, pl = 0, pr = 3, Ho = H + pt + pb, Wo = W + pl + pr;
        auto x = rnd((size_t)nc * H * W), gy = rnd((size_t)nc * Ho * Wo);
        std::vector<float> y((size_t)nc * Ho * Wo), gx(x.size());
        CHECK(hetu_cpu_pad(x.data(), y.data(), nc, H, W, pt, pb, pl, pr, 7.f) == 0, "pad rc");
        CHECK(y[0] == 7.f && y[(size_t)pt * Wo + pl] == x[0] && y.back() == 7.f, "pad values");
        CHECK(hetu_cpu_pad_grad(gy.data(), gx.data(), nc, H, W, pt, pb, pl, pr) == 0, "pad grad rc");
        CHECK(gx[0] == gy[(size_t)pt * Wo + pl], "pad grad value");
    }
    {   // concat along an inner axis and the slice gradients
        const int64_t outer = 4, ai = 3, bi = 6;
        auto a = rnd(outer * ai), b = rnd(outer * bi);
        std::vector<float> y(outer * (ai + bi)), ga(a.size()), gb(b.size());
        CHECK(hetu_cpu_concat(a.data(), b.data(), y.data(), outer, ai, bi) == 0, "concat rc");
        CHECK(hetu_cpu_concat_grad(y.data(), ga.data(), outer, ai, ai + bi, 0) == 0, "concat grad a rc");
        CHECK(hetu_cpu_concat_grad(y.data(), gb.data(), outer, bi, ai + bi, ai) == 0, "concat grad b rc");
        CHECK(ga == a && gb == b, "concat round trip");
    }
    {   // 4-D transpose there and back
        const int64_t shp[4] = {2, 3, 5, 7}, shp2[4] = {7, 2, 5, 3};
        const int perm[4] = {3, 0, 2, 1}, inv[4] = {1, 3, 2, 0};
        auto x = rnd(2 * 3 * 5 * 7);
        std::vector<float> y(x.size()), z(x.size());
        CHECK(hetu_cpu_transpose(x.data(), y.data(), shp, perm, 4) == 0, "transpose rc");
        CHECK(y[1 * 2 * 5 * 3 + 1 * 5 * 3 + 2 * 3 + 1] == x[1 * 3 * 5 * 7 + 1 * 5 * 7 + 2 * 7 + 1], "transpose element");
        CHECK(hetu_cpu_transpose(y.data(), z.data(), shp2, inv, 4) == 0, "transpose back rc");
        CHECK(z == x, "transpose round trip");
    }
    {   // broadcast / reduce
        auto v = rnd(9);
        std::vector<float> big(9 * 11), red(9);
        CHECK(hetu_cpu_broadcast_to(v.data(), big.data(), 9, 9 * 11) == 0, "broadcast rc");
        CHECK(hetu_cpu_reduce_sum_axis0(big.data(), red.data(), 11, 9) == 0, "reduce rc");
        for (int j = 0; j < 9; ++j) CHECK(close(red[j], 11.f * v[j]), "broadcast/reduce col %d", j);
    }
    {   // embedding rows, out-of-range ids read as zeros; sparse update with repeated ids accumulates
        const int64_t rows = 6, width = 5;
        auto table = rnd(rows * width);
        const int64_t ids[] = {0, 5, -1, 6, 3, 3};
        std::vector<float> out(6 * width, 1.f);
        CHECK(hetu_cpu_embedding_lookup(table.data(), ids, out.data(), 6, width, rows) == 0, "lookup rc");
        CHECK(out[width] == table[5 * width] && out[2 * width] == 0.f && out[3 * width + 4] == 0.f, "lookup values");
        const int64_t uid[] = {3, 1, 3};
        std::vector<float> g(3 * width, 1.f), before = table;
        CHECK(hetu_cpu_sgd_sparse_update(table.data(), uid, g.data(), 3, width, 0.5f) == 0, "sparse update rc");
        CHECK(close(table[3 * width], before[3 * width] - 1.f) && close(table[1 * width], before[1 * width] - 0.5f) &&
              table[0] == before[0], "sparse update values");
    }
    return 0;
}

static int test_optimizers_init() {
    for (int64_t n : {1, 7, 64, 1001}) {
        auto p = rnd(n), g = rnd(n);
        std::vector<float> p0 = p, vel(n, 0.f), acc(n, 0.f), m(n, 0.f), v(n, 0.f), vmax(n, 0.f), y(n);
        CHECK(hetu_cpu_sgd_update(p.data(), g.data(), n, 0.1f, 0.f) == 0, "sgd rc");
        CHECK(close(p[n - 1], p0[n - 1] - 0.1f * g[n - 1]), "sgd value");
        CHECK(hetu_cpu_momentum_update(p.data(), g.data(), vel.data(), n, 0.1f, 0.9f, 0, 0.f) == 0, "momentum rc");
        CHECK(hetu_cpu_momentum_update(p.data(), g.data(), vel.data(), n, 0.1f, 0.9f, 1, 1e-4f) == 0, "nesterov rc");
        CHECK(hetu_cpu_adagrad_update(p.data(), g.data(), acc.data(), n, 0.1f, 1e-7f, 0.f) == 0, "adagrad rc");
        CHECK(hetu_cpu_adam_update(p.data(), g.data(), m.data(), v.data(), n, 1e-3f, 0.9f, 0.999f, 1e-8f, 1, 0.f, 0, nullptr) == 0, "adam rc");
        CHECK(hetu_cpu_adam_update(p.data(), g.data(), m.data(), v.data(), n, 1e-3f, 0.9f, 0.999f, 1e-8f, 2, 0.01f, 1, vmax.data()) == 0,
              "amsgrad rc");
        for (float x : p) CHECK(std::isfinite(x), "optimizer produced a non-finite parameter");
        CHECK(hetu_cpu_dropout(p.data(), y.data(), n, 0.5f, 42) == 0, "dropout rc");
        for (int64_t i = 0; i < n; ++i) CHECK(y[i] == 0.f || close(y[i], 2.f * p[i]), "dropout value");
        CHECK(hetu_cpu_uniform_init(p.data(), n, -0.5f, 0.5f, 1) == 0, "uniform rc");
        for (float x : p) CHECK(x >= -0.5f && x <= 0.5f, "uniform range");
        CHECK(hetu_cpu_normal_init(p.data(), n, 0.f, 1.f, 2) == 0, "normal rc");
        CHECK(hetu_cpu_truncated_normal_init(p.data(), n, 1.f, 0.5f, 3) == 0, "truncated normal rc");
        for (float x : p) CHECK(std::fabs(x - 1.f) <= 1.0f + 1e-5f, "truncated normal beyond two sigma: %f", x);
    }
    return 0;
}

// random alloc / free traffic against the best-fit allocator: offsets stay aligned, live blocks never overlap,
// everything coalesces back into one fragment
static int test_bfc() {
    const uint64_t cap = 1 << 20, align = 256;
    void* h = hetu_bfc_create(cap, align);
    struct Blk { uint64_t off, size; };
    std::vector<Blk> live;
    std::uniform_int_distribution<int> pick(1, 40000);
    for (int it = 0; it < 4000; ++it) {
        if (live.empty() || pick(rng) % 3 != 0) {
            uint64_t want = (uint64_t)pick(rng);
            int64_t off = hetu_bfc_alloc(h, want);
            if (off < 0) continue;                              // full: fine, frees will follow
            CHECK(off % (int64_t)align == 0 && (uint64_t)off + want <= cap, "bfc offset %ld for %lu bytes", (long)off, (unsigned long)want);
            for (auto& b : live)
                CHECK((uint64_t)off + want <= b.off || b.off + b.size <= (uint64_t)off, "bfc blocks overlap at %ld", (long)off);
            live.push_back({(uint64_t)off, want});
        } else {
            size_t k = (size_t)pick(rng) % live.size();
            CHECK(hetu_bfc_free(h, live[k].off) == 0, "bfc free rc");
            live[k] = live.back();
            live.pop_back();
        }
    }
    CHECK(hetu_bfc_free(h, cap + 123) != 0, "bfc accepted the free of an unknown offset");
    for (auto& b : live) CHECK(hetu_bfc_free(h, b.off) == 0, "bfc final free rc");
    uint64_t st[6];
    hetu_bfc_stats(h, st);
    CHECK(st[0] == 0 && st[2] == 0 && st[3] == 1 && st[4] == cap && st[1] > 0 && st[5] > 0, "bfc did not coalesce: in_use %lu frags %lu",
          (unsigned long)st[0], (unsigned long)st[3]);
    hetu_bfc_destroy(h);
    return 0;
}

int main() {
    if (test_conv() || test_pool() || test_norm_softmax() || test_shape_ops() || test_optimizers_init() || test_bfc()) return 1;
    printf("CPU OPS ASAN OK\n");
    return 0;
}

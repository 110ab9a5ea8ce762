// Address / undefined-behaviour sanitizer run of the classical-ML core on small random data.
//   g++ -std=c++17 -O1 -g -fsanitize=address,undefined -fopenmp tests/native/ml_asan.cc hetu_61a7_b200/csrc_ml/ml.cc -o /tmp/ml_asan && /tmp/ml_asan
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

extern "C" {
const char* hml_last_error();
int hml_linear_fit(int loss, const int64_t* indptr, const int32_t* indices, const float* vals, const float* y, int64_t n, int64_t dim,
                   float* w, int epochs, int batch, float lr, float l1, float l2);
int hml_linear_predict(const int64_t* indptr, const int32_t* indices, const float* vals, int64_t n, int64_t dim, const float* w, float* out);
int hml_fm_fit(const int64_t* indptr, const int32_t* indices, const float* vals, const float* y, int64_t n, int64_t dim, int k, float* w,
               float* v, int epochs, int batch, float lr, float l1, float l2);
void* hml_gbdt_create(const double* a, int na);
void hml_gbdt_free(void* h);
int hml_gbdt_fit(void* h, const int64_t* indptr, const int32_t* indices, const float* vals, const float* y, int64_t n, int64_t dim,
                 void (*ar)(double*, int64_t), const double* range);
int hml_gbdt_predict(void* h, const int64_t* indptr, const int32_t* indices, const float* vals, int64_t n, float* out);
int hml_gbdt_save(void* h, const char* path);
int hml_gbdt_load(void* h, const char* path);
int hml_knn_predict(const int64_t* lptr, const int32_t* lidx, const float* lval, const float* labels, int64_t nl, const int64_t* qptr,
                    const int32_t* qidx, const float* qval, int64_t nq, int64_t dim, int k, int num_label, float* out);
int hml_nb_stats(const int64_t* indptr, const int32_t* indices, const float* vals, const float* y, int64_t n, int64_t dim, int K, double* stats);
int hml_nb_finalize(const double* stats, int64_t dim, int K, float* model);
int hml_nb_predict(const int64_t* indptr, const int32_t* indices, const float* vals, int64_t n, int64_t dim, int K, const float* model, float* out);
int hml_mf_fit(const int32_t* rows, const int32_t* cols, const float* vals, int64_t nnz, int k, float* P, float* Q, int epochs, float lr,
               float l1, float l2, uint64_t seed, double* rmse_out);
int hml_lda_run(const int64_t* doc_ptr, const int32_t* words, int64_t ndocs, int V, int K, int iters, double alpha, double beta, int32_t* z,
                int32_t* nwk, int32_t* nk, int train, uint64_t seed, double* llh);
}
#define CHECK(x) do { if ((x) != 0) { fprintf(stderr, "FAILED %s: %s\n", #x, hml_last_error()); exit(1); } } while (0)

int main() {
    std::mt19937 rng(1);
    const int64_t n = 1500, dim = 23;
    std::vector<int64_t> ptr{0};
    std::vector<int32_t> idx;
    std::vector<float> val, y01, ypm, ymc;
    std::normal_distribution<float> N01(0.f, 1.f);
    for (int64_t i = 0; i < n; ++i) {
        float s = 0;
        for (int32_t f = 0; f < dim; ++f)
            if (rng() % 3 == 0) { float v = N01(rng); idx.push_back(f); val.push_back(v); if (f < 4) s += v; }
        ptr.push_back((int64_t)idx.size());
        y01.push_back(s > 0 ? 1.f : 0.f); ypm.push_back(s > 0 ? 1.f : -1.f); ymc.push_back((float)((s > 0) + (s > 1)));
    }
    std::vector<float> w(dim, 0.f), out(n * 3);
    for (int loss = 0; loss < 3; ++loss) CHECK(hml_linear_fit(loss, ptr.data(), idx.data(), val.data(), ypm.data(), n, dim, w.data(), 2, 97, 0.1f, 1e-4f, 1e-4f));
    CHECK(hml_linear_predict(ptr.data(), idx.data(), val.data(), n, dim, w.data(), out.data()));
    std::vector<float> v(dim * 5, 0.01f);
    CHECK(hml_fm_fit(ptr.data(), idx.data(), val.data(), ypm.data(), n, dim, 5, w.data(), v.data(), 2, 64, 0.05f, 0.f, 1e-4f));
    for (int variant = 0; variant < 4; ++variant) {
        //           depth nodes leafwise split ins  feat  reg  label round lr  gain min_ins mcw alpha lambda maxw multi major seed
        double a[19] = {4, 31, (double)(variant == 1), 12, 0.8, 0.9, (double)(variant == 2), variant == 3 ? 3.0 : 2.0, 4, 0.3, 0, 10, 0, 0.01, 1.0, 0, (double)(variant == 3), 0, 7};
        void* h = hml_gbdt_create(a, 19);
        const float* yy = variant == 3 ? ymc.data() : (variant == 2 ? ypm.data() : y01.data());
        CHECK(hml_gbdt_fit(h, ptr.data(), idx.data(), val.data(), yy, n, dim, nullptr, nullptr));
        CHECK(hml_gbdt_predict(h, ptr.data(), idx.data(), val.data(), n, out.data()));
        CHECK(hml_gbdt_save(h, "/tmp/ml_asan_gbdt.txt"));
        void* h2 = hml_gbdt_create(nullptr, 0);
        CHECK(hml_gbdt_load(h2, "/tmp/ml_asan_gbdt.txt"));
        CHECK(hml_gbdt_predict(h2, ptr.data(), idx.data(), val.data(), n, out.data()));
        hml_gbdt_free(h); hml_gbdt_free(h2);
    }
    std::vector<float> votes(100 * 3);
    CHECK(hml_knn_predict(ptr.data(), idx.data(), val.data(), ymc.data(), n, ptr.data(), idx.data(), val.data(), 100, dim, 7, 3, votes.data()));
    std::vector<double> stats(3 * dim * 3 + 3, 0.0);
    std::vector<float> model(3 + 2 * dim * 3);
    CHECK(hml_nb_stats(ptr.data(), idx.data(), val.data(), ymc.data(), n, dim, 3, stats.data()));
    CHECK(hml_nb_finalize(stats.data(), dim, 3, model.data()));
    CHECK(hml_nb_predict(ptr.data(), idx.data(), val.data(), n, dim, 3, model.data(), out.data()));
    std::vector<int32_t> r, c; std::vector<float> rv;
    for (int i = 0; i < 3000; ++i) { r.push_back((int32_t)(rng() % 40)); c.push_back((int32_t)(rng() % 30)); rv.push_back((float)(rng() % 5)); }
    std::vector<float> P(40 * 4, 0.3f), Q(30 * 4, 0.3f);
    std::vector<double> rm(3);
    CHECK(hml_mf_fit(r.data(), c.data(), rv.data(), 3000, 4, P.data(), Q.data(), 3, 0.01f, 0.f, 1e-3f, 3, rm.data()));
    std::vector<int64_t> dptr{0}; std::vector<int32_t> words;
    for (int d = 0; d < 50; ++d) { for (int t = 0; t < 30; ++t) words.push_back((int32_t)(rng() % 20)); dptr.push_back((int64_t)words.size()); }
    std::vector<int32_t> z(words.size(), -1), nwk(20 * 4, 0), nk(4, 0);
    double llh = 0;
    CHECK(hml_lda_run(dptr.data(), words.data(), 50, 20, 4, 5, 0.1, 0.01, z.data(), nwk.data(), nk.data(), 1, 9, &llh));
    CHECK(hml_lda_run(dptr.data(), words.data(), 50, 20, 4, 2, 0.1, 0.01, z.data(), nwk.data(), nk.data(), 2, 9, &llh));
    CHECK(hml_lda_run(dptr.data(), words.data(), 50, 20, 4, 2, 0.1, 0.01, z.data(), nwk.data(), nk.data(), 0, 9, &llh));
    if (!std::isfinite(llh)) { fprintf(stderr, "llh not finite\n"); return 1; }
    printf("ML ASAN OK\n");
    return 0;
}

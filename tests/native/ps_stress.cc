// In-process stress of the parameter server for the thread / address sanitizers: scheduler, two servers and one
// worker live in this process; several threads push, pull and use the async pool concurrently.
//   g++ -std=c++17 -O1 -g -fsanitize=thread  -fopenmp tests/native/ps_stress.cc hetu_61a7_b200/csrc_ps/*.cc -lpthread -o /tmp/ps_tsan && /tmp/ps_tsan
//   g++ -std=c++17 -O1 -g -fsanitize=address,undefined -fopenmp ... (same sources)
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <vector>

extern "C" {
const char* hps_last_error();
int hps_start_scheduler(int port, int nworkers, int nservers);
int hps_scheduler_port();
int hps_stop_scheduler();
int hps_start_server(const char* sched_host, int sched_port, const char* my_host, int nservers);
int hps_stop_servers();
int hps_worker_init(const char* sched_host, int sched_port);
int hps_worker_finalize();
int hps_init_tensor(uint64_t key, int type, uint64_t len, uint64_t width, int init_type, double a, double b, uint64_t seed);
int hps_pull(uint64_t key, float* out);
int hps_push(uint64_t key, const float* grad);
int hps_sparse_push(uint64_t key, const int64_t* rows, uint64_t n, const float* grads);
int hps_sparse_pull(uint64_t key, const int64_t* rows, uint64_t n, float* out);
int hps_push_async(uint64_t key, const float* grad);
int hps_pull_async(uint64_t key, float* out);
int hps_wait(uint64_t key);
int hps_set_priority(uint64_t key, int prio);
int hps_barrier(int group);
void* hcache_create(uint64_t limit, uint64_t length, uint64_t width, uint64_t node_id, int policy, int64_t pull_bound, int64_t push_bound);
void hcache_destroy(void* h);
int hcache_lookup(void* h, const int64_t* keys, uint64_t n, float* out);
int hcache_update(void* h, const int64_t* keys, uint64_t n, const float* grads);
int64_t hcache_flush(void* h);
const char* hcache_last_error();
}

#define CHECK(x) do { if ((x) != 0) { fprintf(stderr, "FAILED %s: %s\n", #x, hps_last_error()); exit(1); } } while (0)

int main() {
    CHECK(hps_start_scheduler(0, 1, 2));
    int port = hps_scheduler_port();
    CHECK(hps_start_server("127.0.0.1", port, "127.0.0.1", 2));
    CHECK(hps_start_server("127.0.0.1", port, "127.0.0.1", 2));
    CHECK(hps_worker_init("127.0.0.1", port));
    const int T = 4, N = 4096, ROUNDS = 30;
    for (int t = 0; t < T; ++t) CHECK(hps_init_tensor(100 + t, 0, N, 1, 0, 0.0, 0.0, 1));
    CHECK(hps_init_tensor(200, 1, 512, 8, 0, 0.0, 0.0, 1));
    std::vector<std::thread> th;
    for (int t = 0; t < T; ++t)
        th.emplace_back([t] {
            std::vector<float> g(N, 1.0f), out(N);
            std::vector<int64_t> rows = {1, 77, 300, 511};
            std::vector<float> sg(rows.size() * 8, 0.5f), so(rows.size() * 8);
            for (int r = 0; r < ROUNDS; ++r) {
                CHECK(hps_push(100 + t, g.data()));
                CHECK(hps_pull(100 + t, out.data()));
                if (std::fabs(out[0] - (r + 1) * 2.0f + 1.0f) > 1e-3f && std::fabs(out[0] - (r * 2.0f + 1.0f)) > 1e-3f) {
                    fprintf(stderr, "thread %d round %d: value %f\n", t, r, out[0]); exit(1);
                }
                CHECK(hps_push_async(100 + t, g.data()));      // same key: ordered behind nothing else of this thread
                CHECK(hps_wait(100 + t));
                CHECK(hps_sparse_push(200, rows.data(), rows.size(), sg.data()));
                CHECK(hps_sparse_pull(200, rows.data(), rows.size(), so.data()));
            }
        });
    for (auto& x : th) x.join();
    std::vector<float> out(N);
    for (int t = 0; t < T; ++t) {
        CHECK(hps_pull(100 + t, out.data()));
        if (std::fabs(out[N - 1] - 2.0f * ROUNDS) > 1e-3f) { fprintf(stderr, "tensor %d ends at %f\n", t, out[N - 1]); return 1; }
    }
    std::vector<int64_t> rows = {77};
    std::vector<float> so(8);
    CHECK(hps_sparse_pull(200, rows.data(), 1, so.data()));
    if (std::fabs(so[0] - 0.5f * T * ROUNDS) > 1e-2f) { fprintf(stderr, "sparse row ends at %f\n", so[0]); return 1; }
    // HET cache in front of a versioned table: three policies, concurrent lookups / updates, then a flush
    CHECK(hps_init_tensor(300, 2, 256, 4, 0, 0.0, 0.0, 1));
    std::vector<std::thread> ct;
    for (int pol = 0; pol < 3; ++pol)
        ct.emplace_back([pol] {
            void* c = hcache_create(40, 256, 4, 300, pol, 2, 2);
            std::vector<int64_t> keys(16);
            std::vector<float> rows(16 * 4), g(16 * 4, 0.25f);
            unsigned state = 17u * (pol + 1);
            for (int it = 0; it < 60; ++it) {
                for (auto& k : keys) { state = state * 1664525u + 1013904223u; k = (int64_t)((state >> 8) % 256); }
                if (hcache_lookup(c, keys.data(), keys.size(), rows.data()) != 0) { fprintf(stderr, "lookup: %s\n", hcache_last_error()); exit(1); }
                if (hcache_update(c, keys.data(), keys.size(), g.data()) != 0) { fprintf(stderr, "update: %s\n", hcache_last_error()); exit(1); }
            }
            if (hcache_flush(c) < 0) { fprintf(stderr, "flush: %s\n", hcache_last_error()); exit(1); }
            hcache_destroy(c);
        });
    for (auto& x : ct) x.join();
    {
        std::vector<int64_t> all(256);
        for (int i = 0; i < 256; ++i) all[i] = i;
        std::vector<float> tbl(256 * 4);
        CHECK(hps_sparse_pull(300, all.data(), all.size(), tbl.data()));
        double sum = 0;
        for (float v : tbl) sum += v;
        if (std::fabs(sum - 3.0 * 60 * 16 * 4 * 0.25) > 1e-2) { fprintf(stderr, "cache table sums to %f\n", sum); return 1; }
    }
    hps_worker_finalize();
    hps_stop_servers();
    hps_stop_scheduler();
    printf("PS STRESS OK\n");
    return 0;
}

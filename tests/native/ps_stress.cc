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
    hps_worker_finalize();
    hps_stop_servers();
    hps_stop_scheduler();
    printf("PS STRESS OK\n");
    return 0;
}

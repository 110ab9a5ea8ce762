// In-process stress of the graph servers for the sanitizers: two servers (each owns half of a ring-shaped graph),
// sampler threads for all four sampler kinds, two client threads pulling sub-graphs and node records.
//   g++ -std=c++17 -O1 -g -fsanitize=thread tests/native/graph_stress.cc hetu_61a7_b200/csrc_graph/graph.cc -lpthread -o /tmp/g_tsan && /tmp/g_tsan
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <thread>
#include <vector>

extern "C" {
const char* hgm_last_error();
void* hgm_server_create(int rank, int nparts, const int64_t* offsets, int64_t f_len, int64_t i_len, const float* f, const int32_t* ifeat,
                        const int64_t* indptr, const int64_t* nbr, const char* meta, int num_workers);
int hgm_server_set_peers(void* h, const char* peers);
int hgm_server_listen(void* h, int port);
int hgm_server_port(void* h);
int hgm_server_init_cache(void* h, double ratio, int policy);
int hgm_server_add_sampler(void* h, int type, int batch_size, int depth, int width, int rw_head, int rw_length, int queue, uint64_t seed);
int hgm_server_ready(void* h);
int hgm_server_stop(void* h);
void hgm_server_free(void* h);
void* hgm_client_create(const char* host, int port);
void hgm_client_free(void* h);
int64_t hgm_client_pull_graph(void* h, const int32_t* types, int n);
int64_t hgm_client_pull_node(void* h, const int64_t* ids, int64_t n);
int hgm_client_resolve_graph(void* h, int64_t q, int64_t* dims);
int hgm_client_fetch_graph(void* h, int64_t q, float* f, int32_t* iv, int64_t* gid, int64_t* edges, int32_t* extra);
int hgm_client_resolve_node(void* h, int64_t q, int64_t* dims);
int hgm_client_fetch_node(void* h, int64_t q, float* f, int32_t* iv, int64_t* ptr, int64_t* nbr);
int hgm_client_barrier(void* h, int64_t count, const char* host0, int port0);
}

#define CHECK(x) do { if ((x) != 0) { fprintf(stderr, "FAILED %s: %s\n", #x, hgm_last_error()); exit(1); } } while (0)

int main() {
    const int64_t N = 400, HALF = 200, F = 3, I = 2;
    int64_t offsets[3] = {0, HALF, N};
    void* srv[2];
    int ports[2];
    std::vector<std::vector<float>> feats(2);
    std::vector<std::vector<int32_t>> ifeats(2);
    std::vector<std::vector<int64_t>> ptrs(2), nbrs(2);
    // listen first on ephemeral ports, then tell both servers where the peers are
    for (int r = 0; r < 2; ++r) {
        for (int64_t l = 0; l < HALF; ++l) {
            int64_t g = r * HALF + l;
            for (int j = 0; j < F; ++j) feats[r].push_back((float)g + 0.1f * j);
            ifeats[r].push_back((int32_t)(g % 5)); ifeats[r].push_back((int32_t)g);
            ptrs[r].push_back((int64_t)nbrs[r].size());
            nbrs[r].push_back((g + 1) % N); nbrs[r].push_back((g + N - 1) % N); nbrs[r].push_back((g + 37) % N);
        }
        ptrs[r].push_back((int64_t)nbrs[r].size());
        srv[r] = hgm_server_create(r, 2, offsets, F, I, feats[r].data(), ifeats[r].data(), ptrs[r].data(), nbrs[r].data(), "{}", 2);
        if (!srv[r]) { fprintf(stderr, "create: %s\n", hgm_last_error()); return 1; }
        CHECK(hgm_server_listen(srv[r], 0));
        ports[r] = hgm_server_port(srv[r]);
    }
    std::string peers = "127.0.0.1:" + std::to_string(ports[0]) + ",127.0.0.1:" + std::to_string(ports[1]);
    for (int r = 0; r < 2; ++r) {
        CHECK(hgm_server_set_peers(srv[r], peers.c_str()));
        CHECK(hgm_server_init_cache(srv[r], 0.2, r));
        for (int type = 0; type < 4; ++type) CHECK(hgm_server_add_sampler(srv[r], type, 24, 2, 2, 12, 2, 3, 5 + type));
        CHECK(hgm_server_ready(srv[r]));
    }
    std::vector<std::thread> th;
    for (int w = 0; w < 2; ++w)
        th.emplace_back([&, w] {
            void* c = hgm_client_create("127.0.0.1", ports[w]);
            if (!c) { fprintf(stderr, "client: %s\n", hgm_last_error()); exit(1); }
            CHECK(hgm_client_barrier(c, 2, "127.0.0.1", ports[0]));
            for (int it = 0; it < 40; ++it) {
                int32_t type = it % 4;
                int64_t q = hgm_client_pull_graph(c, &type, 1), q2 = hgm_client_pull_graph(c, nullptr, 0);
                for (int64_t qq : {q, q2}) {
                    int64_t d[6];
                    CHECK(hgm_client_resolve_graph(c, qq, d));
                    std::vector<float> f((size_t)(d[1] * d[2]));
                    std::vector<int32_t> iv((size_t)(d[1] * d[3])), ex((size_t)(d[1] * d[4]));
                    std::vector<int64_t> gid((size_t)d[1]), e((size_t)(2 * d[5]));
                    CHECK(hgm_client_fetch_graph(c, qq, f.data(), iv.data(), gid.data(), e.data(), ex.data()));
                    for (int64_t k = 0; k < d[1]; ++k)
                        if (f[(size_t)(k * d[2])] != (float)gid[(size_t)k]) { fprintf(stderr, "feature / id mismatch\n"); exit(1); }
                }
                int64_t ids[3] = {it, HALF + it, N - 1 - it};
                int64_t qn = hgm_client_pull_node(c, ids, 3), dn[4];
                CHECK(hgm_client_resolve_node(c, qn, dn));
                std::vector<float> f((size_t)dn[0]);
                std::vector<int32_t> iv((size_t)dn[1]);
                std::vector<int64_t> p((size_t)dn[2]), nb((size_t)dn[3]);
                CHECK(hgm_client_fetch_node(c, qn, f.data(), iv.data(), p.data(), nb.data()));
                if (dn[3] != 9 || f[0] != (float)ids[0]) { fprintf(stderr, "node pack mismatch\n"); exit(1); }
            }
            CHECK(hgm_client_barrier(c, 2, "127.0.0.1", ports[0]));
            hgm_client_free(c);
        });
    for (auto& t : th) t.join();
    for (int r = 0; r < 2; ++r) { hgm_server_stop(srv[r]); hgm_server_free(srv[r]); }
    printf("GRAPH STRESS OK\n");
    return 0;
}

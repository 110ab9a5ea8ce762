// Address-sanitizer run of the CPU GEMM over ragged shapes and all transpose combinations, checked against a
// naive triple loop.  Buffers are exactly sized so that any over-read / over-write of a panel edge is caught.
//   g++ -std=c++17 -O1 -g -fsanitize=address,undefined -fopenmp -march=x86-64-v3 tests/native/cpu_gemm_asan.cc hetu_61a7_b200/csrc_cpu/cpu_ops.cc -o /tmp/gemm_asan
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

extern "C" int hetu_cpu_matmul(const float* A, const float* B, float* C, int M, int N, int K, int transA, int transB, float alpha, float beta);

int main() {
    std::mt19937 rng(3);
    std::uniform_real_distribution<float> U(-1.f, 1.f);
    const int dims[][3] = {{1, 1, 1}, {1, 7, 5}, {13, 33, 17}, {12, 32, 384}, {25, 65, 385}, {128, 10, 256}, {37, 96, 1000}, {64, 31, 700}, {5, 200, 3}};
    for (auto& d : dims)
        for (int tA = 0; tA < 2; ++tA)
            for (int tB = 0; tB < 2; ++tB)
                for (float beta : {0.f, 1.f, 0.5f}) {
                    const int M = d[0], N = d[1], K = d[2];
                    std::vector<float> A((size_t)M * K), B((size_t)K * N), C((size_t)M * N), R((size_t)M * N);
                    for (auto& v : A) v = U(rng);
                    for (auto& v : B) v = U(rng);
                    for (auto& v : C) v = U(rng);
                    for (int i = 0; i < M; ++i)
                        for (int j = 0; j < N; ++j) {
                            double s = 0;
                            for (int p = 0; p < K; ++p) s += (double)(tA ? A[(size_t)p * M + i] : A[(size_t)i * K + p]) * (tB ? B[(size_t)j * K + p] : B[(size_t)p * N + j]);
                            R[(size_t)i * N + j] = (float)(0.75 * s + beta * C[(size_t)i * N + j]);
                        }
                    if (hetu_cpu_matmul(A.data(), B.data(), C.data(), M, N, K, tA, tB, 0.75f, beta) != 0) return 1;
                    for (size_t k = 0; k < C.size(); ++k)
                        if (std::fabs(C[k] - R[k]) > 1e-3f * (1.f + std::fabs(R[k]))) {
                            fprintf(stderr, "mismatch M=%d N=%d K=%d tA=%d tB=%d beta=%g at %zu: %f vs %f\n", M, N, K, tA, tB, beta, k, C[k], R[k]);
                            return 1;
                        }
                }
    printf("CPU GEMM ASAN OK\n");
    return 0;
}

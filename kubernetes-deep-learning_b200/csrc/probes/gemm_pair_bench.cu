// Probe: the CTA-pair (cta_group::2) pointwise GEMM of probes/gemm_tc2.cuh against the product's single-CTA k_gemm_tc on the
// middle-flow shape (M = 92416, N = K = 768 incl. pitch padding), with per-role wait counters for the pair kernel
// (-DKDL_PAIR_PROF).  Stand-alone: no Python, no libkdl.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I. -Iprobes [-DKDL_PAIR_PROF] probes/gemm_pair_bench.cu -o probes/gemm_pair_bench
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#define KDL_OK 0
#define KDL_ECUDA (-2)
#include "../gemm_tc.cuh"
#include "gemm_tc2.cuh"

using namespace kdl;

int main(int argc, char **argv) {
    const int M = argc > 1 ? atoi(argv[1]) : 92416, N = argc > 2 ? atoi(argv[2]) : 768, K = argc > 3 ? atoi(argv[3]) : 768;
    char err[512] = "";
    tc::TmaEncoder enc;
    enc.init();
    if (!enc.ok() || tc::configure(err) || tc2::configure2(err)) { printf("setup failed %s\n", err); return 1; }
    bf16 *A, *W, *out;
    float *bias;
    cudaMalloc(&A, (size_t)(M + 256) * K * 2); cudaMalloc(&W, (size_t)N * K * 2); cudaMalloc(&out, (size_t)(M + 256) * N * 2); cudaMalloc(&bias, N * 4);
    cudaMemset(A, 0x3c, (size_t)(M + 256) * K * 2); cudaMemset(W, 0x3c, (size_t)N * K * 2); cudaMemset(bias, 0, N * 4);
    CUtensorMap tmB, tmB2;
    enc.encode_2d(&tmB, W, K, N, (uint64_t)K * 2, tc::BLOCK_K, tc::BLOCK_N, 128);
    enc.encode_2d(&tmB2, W, K, N, (uint64_t)K * 2, tc2::BK2, 128, 128);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const double flops = 2.0 * M * N * K;
    for (int which = 0; which < 2; ++which) {
        for (int i = 0; i < 3; ++i) {
            if (which == 0) tc::launch_gemm(enc, A, &tmB, bias, nullptr, out, M, N, K, 1, 0, err);
            else tc2::launch_gemm2(enc, A, &tmB2, bias, nullptr, out, M, N, K, 1, 0, err);
        }
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("kernel %d failed: %s %s\n", which, cudaGetErrorString(e), err); return 1; }
        cudaEventRecord(e0);
        const int iters = 20;
        for (int i = 0; i < iters; ++i) {
            if (which == 0) tc::launch_gemm(enc, A, &tmB, bias, nullptr, out, M, N, K, 1, 0, err);
            else tc2::launch_gemm2(enc, A, &tmB2, bias, nullptr, out, M, N, K, 1, 0, err);
        }
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        printf("%s: %.1f us  %.0f TFLOP/s  (M=%d N=%d K=%d)\n", which == 0 ? "k_gemm_tc  (1 CTA, 128x256)" : "k_gemm_tc2 (CTA pair, 256x256)", ms / iters * 1e3,
               flops / (ms / iters * 1e-3) / 1e12, M, N, K);
    }
#ifdef KDL_PAIR_PROF
    std::vector<unsigned long long> h(160 * 16);
    cudaMemcpyFromSymbol(h.data(), tc2::g_pair_prof, sizeof(unsigned long long) * 160 * 16);
    const char *names[] = {"prod wait empty", "prod issue", "mma wait tempty", "mma wait full", "mma issue", "epi wait tfull", "epi work", "total"};
    for (int r = 0; r < 2; ++r) {
        printf("rank %d CTAs:\n", r);
        for (int c = 0; c < 8; ++c) {
            double s = 0; int n = 0;
            for (int b = r; b < 148; b += 2) { s += (double)h[b * 16 + c]; ++n; }
            printf("  %-18s %10.0f clk\n", names[c], s / n);
        }
    }
#endif
    return 0;
}

// Probe: what does one tcgen05.mma (cta_group::1, kind::f16, M = 128, both operands in shared memory) cost on a B200 SM, alone and
// while other warps of the CTA move data through shared memory?  Decides whether the pointwise GEMM / the fused depthwise+pointwise
// kernel are bound by the tensor pipe, by the MMA's operand reads, or by the shared-memory bandwidth they share with everything else.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I.. probes/mma_rate.cu -o probes/mma_rate     (from csrc/)
//   probes/mma_rate
//
// One CTA per SM on every SM.  Thread 32 issues `iters` rounds of 4 k-steps x (256 / N) MMAs of N columns on fixed (zero) operand
// tiles -- A 128 x 64 bf16 and B 256 x 64 bf16, 128-byte swizzled K-major, the layout the real kernels use -- commits every round to an
// mbarrier and waits for the round before last (two rounds in flight, like a 2-deep operand ring).  Meanwhile `lw` warps stream
// 16-byte shared-memory loads and `sw` warps 16-byte stores over a separate 64 KB region until the MMA thread raises a flag.
// Output per configuration: cycles per MMA, the floor 128 * N / 256, and the bytes/clk the load / store warps moved.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#include "../gemm_tc.cuh"
#include "gemm_tc2.cuh"

using namespace kdl::tc;

struct Res {
    unsigned long long mma_clk, ld_bytes, st_bytes, elapsed;
};

__global__ void __launch_bounds__(1024, 1) k_probe(int N, int iters, int lw, int sw, Res *out) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t a_base = base, b_base = base + 16384, t_base = base + 16384 + 32768;   // traffic region: 64 KB
    const uint32_t bar = t_base + 65536, tmem_slot = bar + 16, flag = bar + 32;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (uint32_t i = threadIdx.x; i < (16384 + 32768 + 65536) / 4; i += blockDim.x)
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(base + 4 * i), "r"(0u) : "memory");
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(flag), "r"(0u) : "memory");
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, 512);
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
    const long long t_start = clock64();
    if (warp == 1) {
        if (lane == 0) {
            const uint64_t adesc = make_smem_desc<128>(a_base);
            const uint32_t idesc = make_idesc(128, (uint32_t)N);
            const int per = 256 / N;
            uint32_t phase = 0;
            const long long t0 = clock64();
            for (int it = 0; it < iters; ++it) {
                for (int k = 0; k < 4; ++k)
                    for (int s = 0; s < per; ++s)
                        umma_bf16(tmem_base + s * N, adesc + 2u * k, make_smem_desc<128>(b_base + s * N * 128) + 2u * k, idesc, 1);
                umma_commit(bar);
                if (it >= 1) { mbar_wait(bar, phase); phase ^= 1u; }   // wait for round it-1: two rounds in flight
            }
            mbar_wait(bar, phase);
            const long long t1 = clock64();
            out[blockIdx.x].mma_clk = (unsigned long long)(t1 - t0);
            asm volatile("st.shared.u32 [%0], %1;" ::"r"(flag), "r"(1u) : "memory");
        }
    } else if (warp >= 2 && warp < 2 + lw + sw) {
        const bool is_ld = warp < 2 + lw;
        unsigned long long n = 0;
        uint32_t acc = 0;
        uint32_t addr = t_base + (uint32_t)((warp - 2) * 512 + lane * 16) % 65536u;
        while (true) {
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                if (is_ld) {
                    uint4 v = ld_shared_v4(addr);
                    acc ^= v.x;
                } else {
                    st_shared_v4(addr, make_uint4(acc, 1, 2, 3));
                }
                addr += 16384;
                if (addr >= t_base + 65536) addr -= 65536;
            }
            n += 8 * 512;
            uint32_t f;
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(f) : "r"(flag));
            if (f) break;
        }
        if (lane == 0) {
            if (is_ld) atomicAdd(&out[blockIdx.x].ld_bytes, n + (acc == 0x12345 ? 1 : 0));
            else atomicAdd(&out[blockIdx.x].st_bytes, n);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0) out[blockIdx.x].elapsed = (unsigned long long)(clock64() - t_start);
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// Same measurement for the CTA-pair instruction (cta_group::2, M = 256 across two SMs): each CTA holds its 128 rows of A and HALF of
// the B tile (N/2 rows); the leader's thread issues, the commit is multicast to both CTAs.
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(1024, 1) k_probe_pair(int N, int iters, int lw, Res *out) {
    using namespace kdl::tc2;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t a_base = base, b_base = base + 16384, t_base = base + 16384 + 32768;
    const uint32_t bar = t_base + 65536, tmem_slot = bar + 16, flag = bar + 32;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    for (uint32_t i = threadIdx.x; i < (16384 + 32768 + 65536) / 4; i += blockDim.x)
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(base + 4 * i), "r"(0u) : "memory");
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(flag), "r"(0u) : "memory");
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc2(tmem_slot, 512);
    fence_proxy_async_smem();
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
    if (warp == 1) {
        if (lane == 0) {
            uint32_t phase = 0;
            const long long t0 = clock64();
            if (rank == 0) {
                const uint64_t adesc = make_smem_desc<128>(a_base), bdesc = make_smem_desc<128>(b_base);
                const uint32_t idesc = make_idesc(256, (uint32_t)N);
                for (int it = 0; it < iters; ++it) {
                    for (int k = 0; k < 4; ++k) umma_bf16_pair(tmem_base, adesc + 2u * k, bdesc + 2u * k, idesc, 1);
                    umma_commit_pair(bar);
                    if (it >= 1) { mbar_wait(bar, phase); phase ^= 1u; }
                }
                mbar_wait(bar, phase);
            } else {
                for (int it = 0; it < iters; ++it) { mbar_wait(bar, phase); phase ^= 1u; }   // the peer sees every multicast commit
            }
            out[blockIdx.x].mma_clk = (unsigned long long)(clock64() - t0);
            asm volatile("st.shared.u32 [%0], %1;" ::"r"(flag), "r"(1u) : "memory");
        }
    } else if (warp >= 2 && warp < 2 + lw) {
        unsigned long long n = 0;
        uint32_t acc = 0;
        uint32_t addr = t_base + (uint32_t)((warp - 2) * 512 + lane * 16) % 65536u;
        while (true) {
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                uint4 v = ld_shared_v4(addr);
                acc ^= v.x;
                addr += 16384;
                if (addr >= t_base + 65536) addr -= 65536;
            }
            n += 8 * 512;
            uint32_t f;
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(f) : "r"(flag));
            if (f) break;
        }
        if (lane == 0) atomicAdd(&out[blockIdx.x].ld_bytes, n + (acc == 0x12345 ? 1 : 0));
    }
    tc_fence_before();
    cluster_sync_all();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc2(tmem_base, 512);
    }
}

int main(int argc, char **argv) {
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int smem = 16384 + 32768 + 65536 + 2048;
    cudaFuncSetAttribute(k_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    Res *d;
    cudaMalloc(&d, sizeof(Res) * sms);
    Res *h = (Res *)malloc(sizeof(Res) * sms);
    const int iters = 2000;
    printf("%d SMs, %d rounds of 4 k-steps; cycles are SM clocks (clock64)\n", sms, iters);
    printf("%5s %4s %4s | %9s %7s | %9s %9s %9s\n", "N", "ldw", "stw", "clk/MMA", "floor", "ld B/clk", "st B/clk", "mma B/clk");
    for (int N : {256, 128, 192}) {
        if (256 % N && N != 192) continue;
        for (int cfg = 0; cfg < (N == 256 ? 6 : 1); ++cfg) {   // (smaller N with concurrent loads hung the MMA thread once: alone only)
            const int lw = (int[]){0, 4, 16, 0, 0, 4}[cfg], sw = (int[]){0, 0, 0, 1, 2, 1}[cfg];
            cudaMemset(d, 0, sizeof(Res) * sms);
            const int n_eff = N == 192 ? 192 : N;
            k_probe<<<sms, 32 * (2 + lw + sw), smem>>>(n_eff, iters, lw, sw, d);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("launch failed: %s\n", cudaGetErrorString(e)); return 1; }
            cudaMemcpy(h, d, sizeof(Res) * sms, cudaMemcpyDeviceToHost);
            double clk = 0, ldb = 0, stb = 0;
            const int per = 256 / n_eff;
            for (int i = 0; i < sms; ++i) {
                clk += (double)h[i].mma_clk;
                ldb += (double)h[i].ld_bytes / (double)h[i].mma_clk;
                stb += (double)h[i].st_bytes / (double)h[i].mma_clk;
            }
            clk /= sms; ldb /= sms; stb /= sms;
            const double per_mma = clk / ((double)iters * 4 * per);
            const double mma_bytes = (128.0 + n_eff) * 32.0 / per_mma;
            printf("%5d %4d %4d | %9.1f %7.1f | %9.1f %9.1f %9.1f\n", n_eff, lw, sw, per_mma, 128.0 * n_eff / 256.0, ldb, stb, mma_bytes);
        }
    }
    if (argc < 2) return 0;   // the pair section below hung once on a B200 (no trap fired); run it only on request: mma_rate pair
    printf("\ncta_group::2 (M = 256 over a CTA pair; per SM: 128 rows of A + N/2 rows of B)\n");
    cudaFuncSetAttribute(k_probe_pair, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    const bool quiet_only = argc >= 2 && argv[1][0] == 'q';   // "q": no competing ld.shared warps, more N values (the safe subset)
    for (int N : {256, 224, 192, 160, 128, 96, 64}) {
        if (!quiet_only && N != 256 && N != 128) continue;
        for (int lw : {0, 4, 16}) {
            if (quiet_only && lw != 0) continue;
            cudaMemset(d, 0, sizeof(Res) * sms);
            k_probe_pair<<<sms & ~1, 32 * (2 + lw), smem>>>(N, iters, lw, d);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("pair launch failed: %s\n", cudaGetErrorString(e)); return 1; }
            cudaMemcpy(h, d, sizeof(Res) * sms, cudaMemcpyDeviceToHost);
            double clk = 0, ldb = 0;
            for (int i = 0; i < (sms & ~1); ++i) { clk += (double)h[i].mma_clk; ldb += (double)h[i].ld_bytes / (double)h[i].mma_clk; }
            clk /= (sms & ~1); ldb /= (sms & ~1);
            const double per_mma = clk / ((double)iters * 4);
            printf("%5d %4d %4d | %9.1f %7.1f | %9.1f %9s %9.1f\n", N, lw, 0, per_mma, 256.0 * N / 512.0, ldb, "-", (128.0 + N / 2.0) * 32.0 / per_mma);
        }
    }
    return 0;
}

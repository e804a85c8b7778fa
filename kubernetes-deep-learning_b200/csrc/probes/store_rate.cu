// Probe: how fast can a B200 WRITE to HBM from a persistent one-CTA-per-SM kernel, by store flavour and stores in flight?
// Several kernels here are write-heavy (conv1 1:5, conv2 1:2, the K=64/128 pointwise layers 1:2) and all plateau near 2.1 TB/s of
// writes; a torch fill reaches 3.9 TB/s.  This separates the store path from everything else: no loads, no math.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I. probes/store_rate.cu -o probes/store_rate      (from csrc/)
//
// mode 0: st.global.v4 straight from registers, a warp writes 512 contiguous bytes per instruction
// mode 1: cp.async.bulk shared -> global, 2 KB per store (conv1_tc's epilogue), EB staging buffers per warp in flight
// mode 2: 2-D TMA tensor store, box 64 columns x 32 rows of a [M, 256]-column bf16 matrix (the GEMM epilogue's store), EB in flight
// Every warp owns a contiguous slice of the 1 GiB buffer.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define KDL_OK 0
#define KDL_ECUDA (-2)
#include "../gemm_tc.cuh"

using namespace kdl::tc;

__device__ __forceinline__ void bulk_store_1d(void *gdst, uint32_t ssrc, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(ssrc), "r"(bytes) : "memory");
}

template <int EB>
__global__ void __launch_bounds__(1024, 1) k_store(int mode, uint8_t *buf, size_t bytes, const __grid_constant__ CUtensorMap tm, int rows_total) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    const size_t gw = (size_t)blockIdx.x * nwarps + warp, total_warps = (size_t)gridDim.x * nwarps;
    const uint32_t ebuf = base + (uint32_t)warp * (uint32_t)(EB * 4096);
    for (uint32_t i = lane; i < EB * 1024; i += 32) asm volatile("st.shared.u32 [%0], %1;" ::"r"(ebuf + 4 * i), "r"(i) : "memory");
    fence_proxy_async_smem();
    __syncwarp();
    if (mode == 0) {
        const size_t per = bytes / total_warps / 512 * 512;
        uint4 *p = reinterpret_cast<uint4 *>(buf + gw * per) + lane;
        const uint4 v = make_uint4(lane, warp, 3, 4);
        for (size_t i = 0; i < per / 512; ++i) p[i * 32] = v;
    } else if (mode == 1) {
        const size_t per = bytes / total_warps / 2048 * 2048;
        uint8_t *p = buf + gw * per;
        if (lane == 0) {
            for (size_t i = 0; i < per / 2048; ++i) {
                bulk_wait_read<EB - 1>();
                bulk_store_1d(p + i * 2048, ebuf + (uint32_t)(i % EB) * 4096u, 2048);
                bulk_commit();
            }
            bulk_wait_all<0>();
        }
    } else {
        // [rows_total, 256] bf16; a warp owns row blocks of 32; per block four 64-column boxes
        const int blocks = rows_total / 32;
        if (lane == 0) {
            uint32_t n = 0;
            for (int rb = (int)gw; rb < blocks; rb += (int)total_warps) {
                for (int c = 0; c < 4; ++c, ++n) {
                    bulk_wait_read<EB - 1>();
                    tma_store_2d(&tm, ebuf + (n % EB) * 4096u, c * 64, rb * 32);
                    bulk_commit();
                }
            }
            bulk_wait_all<0>();
        }
    }
}

template <int EB> static void run(int mode, int warps, uint8_t *d, size_t bytes, const CUtensorMap &tm, int rows, int sms) {
    const int smem = warps * EB * 4096 + 2048;
    cudaFuncSetAttribute(k_store<EB>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k_store<EB><<<sms, warps * 32, smem>>>(mode, d, bytes, tm, rows);
    cudaEventRecord(e0);
    for (int i = 0; i < 5; ++i) k_store<EB><<<sms, warps * 32, smem>>>(mode, d, bytes, tm, rows);
    cudaEventRecord(e1);
    cudaError_t e = cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const char *names[] = {"st.global.v4 (registers)", "cp.async.bulk 2 KB (smem)", "TMA 2-D store 64x32 (smem)"};
    printf("%-28s warps/CTA %2d  in flight/warp %d : %7.0f GB/s %s\n", names[mode], warps, mode == 0 ? 0 : EB, bytes * 5.0 / (ms * 1e-3) / 1e9,
           e == cudaSuccess ? "" : cudaGetErrorString(e));
}

int main() {
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const size_t bytes = (size_t)1 << 30;
    uint8_t *d;
    cudaMalloc(&d, bytes);
    TmaEncoder enc;
    enc.init();
    const int rows = (int)(bytes / 512);
    CUtensorMap tm;
    enc.encode_2d(&tm, d, 256, rows, 512, 64, 32, 128);
    for (int warps : {4, 8, 16}) {
        run<2>(0, warps, d, bytes, tm, rows, sms);
        run<2>(1, warps, d, bytes, tm, rows, sms);
        run<4>(1, warps, d, bytes, tm, rows, sms);
        if (warps <= 4) run<8>(1, warps, d, bytes, tm, rows, sms);
        run<2>(2, warps, d, bytes, tm, rows, sms);
        run<4>(2, warps, d, bytes, tm, rows, sms);
        if (warps <= 4) run<8>(2, warps, d, bytes, tm, rows, sms);
    }
    return 0;
}

// Probe: what does a cudaEventRecord / kernel / cudaEventRecord bracket add to a kernel's duration?
// bench.py times every launch of a forward pass this way (kdl_profile_begin/end), and the brackets' sum came out ~8 % above the
// un-bracketed step.  Sequence per iteration, one stream: [a0] busy kernel (~100 us) [b0]  [a1] X [b1], where X is nothing, a
// 1 x 32 empty kernel, or a 148 x 512 empty kernel with 200 KB of dynamic shared memory.  Also: the busy kernel's bracket against
// the same kernel timed back to back without brackets.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 probes/event_overhead.cu -o probes/event_overhead      (from csrc/)
#include <algorithm>
#include <cstdio>
#include <vector>
#include <cuda_runtime.h>

__global__ void k_null() {}
__global__ void k_null_big() { extern __shared__ char s[]; if (threadIdx.x == 9999) s[0] = 0; }
__global__ void __launch_bounds__(512, 1) k_busy(float4 *p, size_t n) {
    extern __shared__ char s[];
    if (threadIdx.x == 9999) s[0] = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        float4 v = p[i];
        v.x += 1.f;
        p[i] = v;
    }
}

int main() {
    const size_t n = (size_t)16 << 20;   // 256 MB read + 256 MB written: ~90 us
    float4 *d;
    cudaMalloc(&d, n * sizeof(float4));
    cudaMemset(d, 0, n * sizeof(float4));
    cudaFuncSetAttribute(k_null_big, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaFuncSetAttribute(k_busy, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaStream_t st;
    cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
    const int IT = 200;
    std::vector<cudaEvent_t> ev(4 * IT);
    for (auto &e : ev) cudaEventCreate(&e);
    auto med = [](std::vector<float> v) { std::sort(v.begin(), v.end()); return v[v.size() / 2]; };
    const char *names[] = {"empty bracket (no kernel)", "1 x 32 empty kernel", "148 x 512 empty kernel, 200 KB smem"};
    for (int mode = 0; mode < 3; ++mode) {
        for (int i = 0; i < IT; ++i) {
            cudaEventRecord(ev[4 * i], st);
            k_busy<<<148, 512, 200 * 1024, st>>>(d, n);
            cudaEventRecord(ev[4 * i + 1], st);
            cudaEventRecord(ev[4 * i + 2], st);
            if (mode == 1) k_null<<<1, 32, 0, st>>>();
            if (mode == 2) k_null_big<<<148, 512, 200 * 1024, st>>>();
            cudaEventRecord(ev[4 * i + 3], st);
        }
        cudaStreamSynchronize(st);
        std::vector<float> busy, x, gap;
        for (int i = 10; i < IT; ++i) {
            float t;
            cudaEventElapsedTime(&t, ev[4 * i], ev[4 * i + 1]); busy.push_back(t * 1e3f);
            cudaEventElapsedTime(&t, ev[4 * i + 2], ev[4 * i + 3]); x.push_back(t * 1e3f);
            cudaEventElapsedTime(&t, ev[4 * i + 1], ev[4 * i + 2]); gap.push_back(t * 1e3f);
        }
        printf("%-40s bracket %6.2f us   (busy kernel's bracket %6.2f us, b->a gap %5.2f us)\n", names[mode], med(x), med(busy), med(gap));
    }
    // the busy kernel back to back, one pair of events around all of them
    for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(ev[0], st);
        for (int i = 0; i < IT; ++i) k_busy<<<148, 512, 200 * 1024, st>>>(d, n);
        cudaEventRecord(ev[1], st);
        cudaStreamSynchronize(st);
        float t;
        cudaEventElapsedTime(&t, ev[0], ev[1]);
        printf("busy kernel, %d launches back to back without brackets: %6.2f us per launch\n", IT, t * 1e3f / IT);
    }
    // every launch bracketed (as kdl_profile_begin does), nothing else in between
    for (int i = 0; i < IT; ++i) {
        cudaEventRecord(ev[2 * i], st);
        k_busy<<<148, 512, 200 * 1024, st>>>(d, n);
        cudaEventRecord(ev[2 * i + 1], st);
    }
    cudaStreamSynchronize(st);
    std::vector<float> b;
    for (int i = 10; i < IT; ++i) { float t; cudaEventElapsedTime(&t, ev[2 * i], ev[2 * i + 1]); b.push_back(t * 1e3f); }
    float tot;
    cudaEventElapsedTime(&tot, ev[20], ev[2 * IT - 1]);
    printf("busy kernel, every launch bracketed: median bracket %6.2f us, wall per launch %6.2f us\n", med(b), tot * 1e3f / (IT - 10));
    return 0;
}

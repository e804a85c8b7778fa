// Probe: CUDA-core arithmetic rates on B200 that bound the depthwise kernels (which sit at ~40 MAC/clk/SM whatever the warp count):
// scalar FFMA, packed FFMA2 (fma.rn.f32x2), HFMA2.BF16 (fma.rn.bf16x2), and the depthwise mix (FFMA2 + unpack shifts/masks).
// One 512-thread CTA per SM, 12 independent accumulator chains per thread, no memory traffic.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 probes/fma_rate.cu -o probes/fma_rate      (from csrc/)
#include <cstdio>
#include <cuda_runtime.h>
#include <cuda_bf16.h>

__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
    unsigned long long d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(*reinterpret_cast<unsigned long long *>(&a)), "l"(*reinterpret_cast<unsigned long long *>(&b)),
        "l"(*reinterpret_cast<unsigned long long *>(&c)));
    return *reinterpret_cast<float2 *>(&d);
}

constexpr int CH = 12;       // independent chains per thread
constexpr int INNER = 64;    // unrolled steps per loop trip

// mode 0: FFMA2   1: scalar FFMA (2 per step, same MACs)   2: bf16x2 HFMA2 (same MACs)   3: FFMA2 + one shift and one mask per 2.25 FFMA2
template <int MODE> __global__ void __launch_bounds__(512, 1) k_rate(float *out, int trips, float w0, unsigned seed) {
    float2 acc[CH];
    float2 w = make_float2(w0, w0 * 0.5f);
    unsigned u = seed + threadIdx.x;
    __nv_bfloat162 hacc[CH], hw = __floats2bfloat162_rn(w0, w0 * 0.5f);
#pragma unroll
    for (int i = 0; i < CH; ++i) { acc[i] = make_float2(1.f + i, 2.f + i); hacc[i] = __floats2bfloat162_rn(1.f + i, 2.f); }
    float2 x = make_float2(1.0001f, 0.9999f);
    __nv_bfloat162 hx = __floats2bfloat162_rn(1.0f, 0.99f);
    for (int t = 0; t < trips; ++t) {
#pragma unroll
        for (int s = 0; s < INNER; ++s) {
#pragma unroll
            for (int i = 0; i < CH; ++i) {
                if (MODE == 0 || MODE == 3) acc[i] = ffma2(x, w, acc[i]);
                if (MODE == 1) { acc[i].x = fmaf(x.x, w.x, acc[i].x); acc[i].y = fmaf(x.y, w.y, acc[i].y); }
                if (MODE == 2) hacc[i] = __hfma2(hx, hw, hacc[i]);
            }
            if (MODE == 3) {   // the unpack work that rides along in the depthwise kernel: 16 per 36 FFMA2 -> ~5.3 per 12
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    u = u * 1664525u + 1013904223u;   // keeps the compiler from folding; IMAD
                    x.x = __uint_as_float((u << 16) | 0x3f800000u & 0x3fffffffu);
                    x.y = __uint_as_float((u & 0xffff0000u) | 0x3f000000u);
                }
            }
        }
    }
    float r = 0.f;
#pragma unroll
    for (int i = 0; i < CH; ++i) r += acc[i].x + acc[i].y + __low2float(hacc[i]) + __high2float(hacc[i]);
    if (r == 123.456f) out[0] = r;
}

// mode 4: the depthwise scatter pattern itself, from registers: 4 input pixels x 4 channels (float2 lo/hi each) against 9 taps of
// per-channel weights into 12 accumulators = 36 FFMA2 per row with three DISTINCT 64-bit operands each (no operand reuse for free)
__device__ __forceinline__ void ffma2v(float2 &c, const float2 &a, const float2 &b) {   // volatile: ptxas keeps the order of these
    asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(*reinterpret_cast<unsigned long long *>(&c))
                 : "l"(*reinterpret_cast<const unsigned long long *>(&a)), "l"(*reinterpret_cast<const unsigned long long *>(&b)));
}
// ORDER 0: as the kernel writes it today (compiler's schedule)   1: grouped by INPUT operand (all uses of one pixel half back to back,
// so the operand-reuse cache can serve it)   2: grouped by WEIGHT operand
template <int ORDER> __global__ void __launch_bounds__(512, 1) k_scatter(float *out, int trips, const float *wsrc) {
    float2 wl[9], wh[9], in_lo[4], in_hi[4], a[3][2][2];
#pragma unroll
    for (int k = 0; k < 9; ++k) { wl[k] = make_float2(wsrc[k], wsrc[k + 9]); wh[k] = make_float2(wsrc[k + 18], wsrc[k + 27]); }
#pragma unroll
    for (int j = 0; j < 4; ++j) { in_lo[j] = make_float2(wsrc[40 + j], wsrc[44 + j]); in_hi[j] = make_float2(wsrc[48 + j], wsrc[52 + j]); }
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int j = 0; j < 2; ++j) { a[r][j][0] = make_float2(0.f, 0.f); a[r][j][1] = make_float2(0.f, 0.f); }
    for (int t = 0; t < trips; ++t) {
#pragma unroll
        for (int s = 0; s < 16; ++s) {
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int r = 0; r < 3; ++r)
#pragma unroll
                    for (int kx = 0; kx < 3; ++kx) {
                        if (ORDER == 3) {   // the same MACs as scalar FFMA
                            a[r][j][0].x = fmaf(in_lo[j + kx].x, wl[3 * r + kx].x, a[r][j][0].x);
                            a[r][j][0].y = fmaf(in_lo[j + kx].y, wl[3 * r + kx].y, a[r][j][0].y);
                            a[r][j][1].x = fmaf(in_hi[j + kx].x, wh[3 * r + kx].x, a[r][j][1].x);
                            a[r][j][1].y = fmaf(in_hi[j + kx].y, wh[3 * r + kx].y, a[r][j][1].y);
                        }
                        if (ORDER != 0) continue;
                        a[r][j][0] = ffma2(in_lo[j + kx], wl[3 * r + kx], a[r][j][0]);
                        a[r][j][1] = ffma2(in_hi[j + kx], wh[3 * r + kx], a[r][j][1]);
                    }
            if (ORDER == 1) {
#pragma unroll
                for (int p = 0; p < 4; ++p) {
#pragma unroll
                    for (int half = 0; half < 2; ++half)
#pragma unroll
                        for (int j = 0; j < 2; ++j) {
                            const int kx = p - j;
                            if (kx < 0 || kx > 2) continue;
#pragma unroll
                            for (int r = 0; r < 3; ++r) {
                                if (half == 0) ffma2v(a[r][j][0], in_lo[p], wl[3 * r + kx]);
                                else ffma2v(a[r][j][1], in_hi[p], wh[3 * r + kx]);
                            }
                        }
                }
            }
            if (ORDER == 2) {
#pragma unroll
                for (int r = 0; r < 3; ++r)
#pragma unroll
                    for (int kx = 0; kx < 3; ++kx)
#pragma unroll
                        for (int half = 0; half < 2; ++half)
#pragma unroll
                            for (int j = 0; j < 2; ++j) {
                                if (half == 0) ffma2v(a[r][j][0], in_lo[j + kx], wl[3 * r + kx]);
                                else ffma2v(a[r][j][1], in_hi[j + kx], wh[3 * r + kx]);
                            }
            }
            // rotate the inputs so that successive rows are not identical (4 cheap moves; no new arithmetic)
            const float2 t0 = in_lo[0], t1 = in_hi[0];
#pragma unroll
            for (int j = 0; j < 3; ++j) { in_lo[j] = in_lo[j + 1]; in_hi[j] = in_hi[j + 1]; }
            in_lo[3] = t0; in_hi[3] = t1;
        }
    }
    float r = 0.f;
#pragma unroll
    for (int q = 0; q < 3; ++q)
#pragma unroll
        for (int j = 0; j < 2; ++j) r += a[q][j][0].x + a[q][j][0].y + a[q][j][1].x + a[q][j][1].y;
    if (r == 123.456f) out[0] = r;
}

template <int MODE> static void run(const char *name, float *d, int sms) {
    const int trips = 2000;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int warps : {4, 8, 16}) {
        k_rate<MODE><<<sms, warps * 32>>>(d, 10, 1.0f, 1);
        cudaEventRecord(e0);
        k_rate<MODE><<<sms, warps * 32>>>(d, trips, 1.0f, 1);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        int clk_khz;
        cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
        const double macs = 2.0 * CH * INNER * (double)trips * warps * 32;       // per SM
        const double clks = ms * 1e-3 * clk_khz * 1e3;
        printf("%-42s %2d warps/SM: %7.1f MAC/clk/SM  (%.2f ms, nominal clock %d MHz)\n", name, warps, macs / clks, ms, clk_khz / 1000);
    }
}

int main() {
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    float *d;
    cudaMalloc(&d, 4);
    run<0>("FFMA2 (fma.rn.f32x2)", d, sms);
    run<1>("scalar FFMA", d, sms);
    run<2>("HFMA2.BF16 (fma.rn.bf16x2)", d, sms);
    run<3>("FFMA2 + unpack mix (shift/mask/IMAD)", d, sms);
    {
        float *w;
        cudaMalloc(&w, 64 * 4);
        cudaMemset(w, 0, 64 * 4);
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        for (int order = 0; order < 4; ++order)
        for (int warps : {4, 16}) {
            const int trips = 4000;
            auto go = [&](int tr) {
                if (order == 0) k_scatter<0><<<sms, warps * 32>>>(d, tr, w);
                if (order == 1) k_scatter<1><<<sms, warps * 32>>>(d, tr, w);
                if (order == 2) k_scatter<2><<<sms, warps * 32>>>(d, tr, w);
                if (order == 3) k_scatter<3><<<sms, warps * 32>>>(d, tr, w);
            };
            go(10);
            cudaEventRecord(e0);
            go(trips);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            int clk_khz;
            cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
            const double macs = 72.0 * 16 * (double)trips * warps * 32;
            const char *on[] = {"depthwise scatter pattern, compiler order", "scatter grouped by input operand", "scatter grouped by weight operand", "scatter as scalar FFMA, compiler order"};
            printf("%-42s %2d warps/SM: %7.1f MAC/clk/SM  (%.2f ms)\n", on[order], warps, macs / (ms * 1e-3 * clk_khz * 1e3), ms);
        }
    }
    return 0;
}

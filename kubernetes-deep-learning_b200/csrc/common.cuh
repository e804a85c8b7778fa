// Shared device helpers for the kdl kernels (sm_100a).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <utility>

namespace kdl {

typedef __nv_bfloat16 bf16;

// ---- 4-channel vector access, storage type T in {float, bf16}; math is always fp32 ----
template <typename T> struct V4;

template <> struct V4<float> {
    static __device__ __forceinline__ float4 ld(const float *p) { return *reinterpret_cast<const float4 *>(p); }
    static __device__ __forceinline__ void st(float *p, float4 v) { *reinterpret_cast<float4 *>(p) = v; }
};

template <> struct V4<bf16> {
    static __device__ __forceinline__ float4 ld(const bf16 *p) {
        uint2 r = *reinterpret_cast<const uint2 *>(p);
        float4 v;
        v.x = __uint_as_float(r.x << 16);
        v.y = __uint_as_float(r.x & 0xffff0000u);
        v.z = __uint_as_float(r.y << 16);
        v.w = __uint_as_float(r.y & 0xffff0000u);
        return v;
    }
    static __device__ __forceinline__ void st(bf16 *p, float4 v) {
        __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y);
        __nv_bfloat162 b = __floats2bfloat162_rn(v.z, v.w);
        uint2 r;
        r.x = *reinterpret_cast<uint32_t *>(&a);
        r.y = *reinterpret_cast<uint32_t *>(&b);
        *reinterpret_cast<uint2 *>(p) = r;
    }
};

// 8 consecutive channels: one 16-byte access for bf16, two for float
template <typename T> __device__ __forceinline__ void st8(T *p, float4 a, float4 b);
template <> __device__ __forceinline__ void st8<float>(float *p, float4 a, float4 b) {
    *reinterpret_cast<float4 *>(p) = a;
    *reinterpret_cast<float4 *>(p + 4) = b;
}
template <> __device__ __forceinline__ void st8<bf16>(bf16 *p, float4 a, float4 b) {
    __nv_bfloat162 v0 = __floats2bfloat162_rn(a.x, a.y), v1 = __floats2bfloat162_rn(a.z, a.w);
    __nv_bfloat162 v2 = __floats2bfloat162_rn(b.x, b.y), v3 = __floats2bfloat162_rn(b.z, b.w);
    uint4 r;
    r.x = *reinterpret_cast<uint32_t *>(&v0); r.y = *reinterpret_cast<uint32_t *>(&v1);
    r.z = *reinterpret_cast<uint32_t *>(&v2); r.w = *reinterpret_cast<uint32_t *>(&v3);
    *reinterpret_cast<uint4 *>(p) = r;
}
template <typename T> __device__ __forceinline__ void ld8(const T *p, float4 &a, float4 &b);
template <> __device__ __forceinline__ void ld8<float>(const float *p, float4 &a, float4 &b) {
    a = *reinterpret_cast<const float4 *>(p);
    b = *reinterpret_cast<const float4 *>(p + 4);
}
template <> __device__ __forceinline__ void ld8<bf16>(const bf16 *p, float4 &a, float4 &b) {
    const uint4 r = *reinterpret_cast<const uint4 *>(p);
    a = make_float4(__uint_as_float(r.x << 16), __uint_as_float(r.x & 0xffff0000u), __uint_as_float(r.y << 16),
                    __uint_as_float(r.y & 0xffff0000u));
    b = make_float4(__uint_as_float(r.z << 16), __uint_as_float(r.z & 0xffff0000u), __uint_as_float(r.w << 16),
                    __uint_as_float(r.w & 0xffff0000u));
}

template <typename T> __device__ __forceinline__ float ld1(const T *p);
template <> __device__ __forceinline__ float ld1<float>(const float *p) { return *p; }
template <> __device__ __forceinline__ float ld1<bf16>(const bf16 *p) { return __bfloat162float(*p); }
template <typename T> __device__ __forceinline__ void st1(T *p, float v);
template <> __device__ __forceinline__ void st1<float>(float *p, float v) { *p = v; }
template <> __device__ __forceinline__ void st1<bf16>(bf16 *p, float v) { *p = __float2bfloat16_rn(v); }

// Programmatic dependent launch: a kernel launched with the PDL attribute may start while its predecessor in the stream is still
// running; pdl_wait() blocks until the predecessor has completed and its writes are visible, pdl_launch_dependents() lets the
// successor be scheduled early.  Both are no-ops for a normal launch.  Every kernel of the forward pass runs its prologue (barrier
// init, TMEM allocation, descriptor prefetch), then pdl_wait(), then touches global memory -- so launch latency and prologues of
// the ~80 kernels overlap their predecessors' tails, which is what batch-1 latency is made of.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

inline bool &pdl_enabled() {
    static thread_local bool on = false;
    return on;
}
// cudaLaunchKernelEx wrapper: same as kernel<<<grid, block, smem, st>>>(args...) plus the PDL attribute when enabled
template <typename... Params, typename... Args>
inline cudaError_t launch_k(void (*kernel)(Params...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args &&...args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, Params(std::forward<Args>(args))...);
}

// packed fp32 FMA (Blackwell FFMA2): d = a * b + c on two lanes
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
#ifdef KDL_SCALAR_FMA
    return make_float2(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y));
#endif
    unsigned long long d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;"
        : "=l"(d)
        : "l"(*reinterpret_cast<unsigned long long *>(&a)), "l"(*reinterpret_cast<unsigned long long *>(&b)),
          "l"(*reinterpret_cast<unsigned long long *>(&c)));
    return *reinterpret_cast<float2 *>(&d);
}

__device__ __forceinline__ float4 relu4(float4 v) {
    return make_float4(fmaxf(v.x, 0.f), fmaxf(v.y, 0.f), fmaxf(v.z, 0.f), fmaxf(v.w, 0.f));
}
__device__ __forceinline__ float4 fma4(float4 a, float4 w, float4 acc) {
    acc.x = fmaf(a.x, w.x, acc.x);
    acc.y = fmaf(a.y, w.y, acc.y);
    acc.z = fmaf(a.z, w.z, acc.z);
    acc.w = fmaf(a.w, w.w, acc.w);
    return acc;
}
__device__ __forceinline__ float4 max4(float4 a, float4 b) {
    return make_float4(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z), fmaxf(a.w, b.w));
}
__device__ __forceinline__ float4 add4(float4 a, float4 b) {
    return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
}

// The gateway's scaling, bit-exact: np.float32 x /= 127.5 ; x -= 1.0  (two roundings, no FMA contraction).
__device__ __forceinline__ float scale_px(uint8_t v) {
    return __fsub_rn(__fdiv_rn((float)v, 127.5f), 1.0f);
}


#ifdef KDL_MID_PROF
// probe builds: effective SM clock while a kernel runs.  CTA 0's first thread records clock64 and %globaltimer when it starts and when
// it ends; slot 0 = pointwise GEMM, 1 = depthwise.  [slot][0..3] = cycles, nanoseconds, launches seen, reserved
static __device__ unsigned long long g_clk_meter[4][4];
struct ClkMeter {
    long long c0;
    unsigned long long t0;
    int slot;
    bool on;
    __device__ __forceinline__ ClkMeter(int s) : slot(s), on(blockIdx.x == 0 && threadIdx.x == 0) {
        if (on) {
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
            c0 = clock64();
        }
    }
    __device__ __forceinline__ void stop() {
        if (on) {
            const long long c1 = clock64();
            unsigned long long t1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            g_clk_meter[slot][0] += (unsigned long long)(c1 - c0);
            g_clk_meter[slot][1] += t1 - t0;
            g_clk_meter[slot][2] += 1;
        }
    }
};
#endif
}  // namespace kdl

// kdl_preprocess_u8_static: the gateway's resize + scale for callers that must not allocate.
//
// Who needs this: code that runs where device allocation is off the table -- a latency-critical request thread that may not
// take the allocator's locks or trigger an implicit device synchronisation, a plugin hosted inside another framework's memory
// budget, or a caller that never created a KdlModel (whose arena normally provides the index-table scratch).  The regular entry
// point kdl_preprocess_u8 keeps its per-image index tables in buffers the model allocated with cudaMalloc; this variant keeps
// them in one fixed, file-scope device array instead, so it calls no cudaMalloc / cudaFree / allocating library routine at all.
//
// Price of that: the array has a fixed, modest capacity (KDL_STATIC_MAX_BATCH images, target side <= KDL_STATIC_MAX_SIZE), it
// occupies KDL_STATIC_SCRATCH_BYTES on every GPU that loads this library until the process exits, and every call shares it --
// see the notes on the entry point.
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <mutex>

#include "../../include/kdl.h"
#include "common.cuh"

namespace {

constexpr int MAXB = KDL_STATIC_MAX_BATCH;
constexpr int MAXS = KDL_STATIC_MAX_SIZE;
constexpr size_t META_BYTES = 1024;                                   // MAXB * 24 B rounded up
constexpr size_t TAB_BYTES = (size_t)MAXB * MAXS * sizeof(int32_t);   // one axis
static_assert(META_BYTES + 2 * TAB_BYTES == KDL_STATIC_SCRATCH_BYTES, "header constant out of date");

struct Meta {
    long long offset, pitch;
    int h, w;
};
static_assert(sizeof(Meta) * MAXB <= META_BYTES, "meta block too small");
struct MetaPack {  // travels as a kernel argument: no host->device copy, no staging buffer
    Meta m[MAXB];
};

// The scratch itself.  In HOST code this name is only a stand-in for the symbol, never a device address: the address the
// kernels use is obtained once per device with cudaGetSymbolAddress (see device_scratch()).
static __device__ __align__(256) unsigned char g_static_scratch[KDL_STATIC_SCRATCH_BYTES];

// Same arithmetic as k_index_tables (kernels.cuh): Pillow's accumulated-double nearest-neighbour rule, one thread per
// (image, axis); thread 0..B-1 of axis 0 also drops the image's metadata into the scratch for the gather kernel.
__global__ void k_static_tables(MetaPack pack, int B, int S, unsigned char *scratch) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= 2 * B) return;
    int b = t >> 1, axis = t & 1;
    Meta *meta = reinterpret_cast<Meta *>(scratch);
    int32_t *xtab = reinterpret_cast<int32_t *>(scratch + META_BYTES);
    int32_t *ytab = reinterpret_cast<int32_t *>(scratch + META_BYTES + TAB_BYTES);
    if (axis == 0) meta[b] = pack.m[b];
    int n_in = axis ? pack.m[b].h : pack.m[b].w;
    int32_t *tab = (axis ? ytab : xtab) + (size_t)b * S;
    double a = __ddiv_rn((double)n_in, (double)S);
    double o = __dmul_rn(a, 0.5);
    for (int i = 0; i < S; ++i) {
        int s = o < 0.0 ? -1 : (int)o;
        s = max(0, min(s, n_in - 1));
        tab[i] = s;
        o = __dadd_rn(o, a);
    }
}

// Same gather + scale as k_resize<float>: 4 consecutive output floats per thread, one aligned 16-byte store.
__global__ void k_static_resize(const uint8_t *__restrict__ src, const unsigned char *__restrict__ scratch, int B, int S,
                                float *__restrict__ out) {
    const Meta *meta = reinterpret_cast<const Meta *>(scratch);
    const int32_t *xtab = reinterpret_cast<const int32_t *>(scratch + META_BYTES);
    const int32_t *ytab = reinterpret_cast<const int32_t *>(scratch + META_BYTES + TAB_BYTES);
    const long long per_img = (long long)S * S * 3;
    const long long total = per_img * B;
    long long g = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (g >= total) return;
    const int row_el = S * 3;
    int b = (int)(g / per_img);
    int r = (int)(g - (long long)b * per_img);
    int y = r / row_el;
    int j = r - y * row_el;
    int x = j / 3, c = j - x * 3;
    float v[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        if (g + e < total) {
            const Meta mt = meta[b];
            const uint8_t *p = src + mt.offset + (long long)ytab[(size_t)b * S + y] * mt.pitch + (long long)xtab[(size_t)b * S + x] * 3 + c;
            v[e] = kdl::scale_px(__ldg(p));
        } else {
            v[e] = 0.f;
        }
        if (++c == 3) { c = 0; if (++x == S) { x = 0; if (++y == S) { y = 0; ++b; } } }
    }
    if (g + 4 <= total) {
        *reinterpret_cast<float4 *>(out + g) = make_float4(v[0], v[1], v[2], v[3]);
    } else {
        for (int e = 0; g + e < total; ++e) out[g + e] = v[e];
    }
}

// Per-device set-up: resolve the symbol to this device's address, exactly once, with that device current.
constexpr int MAX_DEVICES = 64;
unsigned char *g_dev_ptr[MAX_DEVICES];
bool g_dev_ready[MAX_DEVICES];
std::mutex g_setup_mutex;

cudaError_t device_scratch(int dev, unsigned char **out) {
    std::lock_guard<std::mutex> lock(g_setup_mutex);
    if (!g_dev_ready[dev]) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);  // current device == dev (checked by the caller)
        if (e != cudaSuccess) return e;
        g_dev_ptr[dev] = static_cast<unsigned char *>(p);
        g_dev_ready[dev] = true;
    }
    *out = g_dev_ptr[dev];
    return cudaSuccess;
}

int fail(char *err, size_t err_len, int code, const char *fmt, ...) {
    if (err && err_len) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(err, err_len, fmt, ap);
        va_end(ap);
    }
    return code;
}

}  // namespace

extern "C" int kdl_preprocess_u8_static(const uint8_t *d_src, const int64_t *offset, const int32_t *h, const int32_t *w,
                                        const int64_t *pitch, int32_t B, int32_t S, float *d_out, void *stream, char *err,
                                        size_t err_len) {
    if (!d_src || !d_out || !offset || !h || !w || !pitch) return fail(err, err_len, KDL_EINVAL, "kdl_preprocess_u8_static: null argument");
    if (B < 1 || B > MAXB)
        return fail(err, err_len, KDL_EINVAL, "kdl_preprocess_u8_static: batch %d is outside [1, %d] -- the static scratch holds at most "
                    "KDL_STATIC_MAX_BATCH images per call; split the batch or use kdl_preprocess_u8", B, MAXB);
    if (S < 1 || S > MAXS)
        return fail(err, err_len, KDL_EINVAL, "kdl_preprocess_u8_static: target size %d is outside [1, %d] (KDL_STATIC_MAX_SIZE)", S, MAXS);
    if ((reinterpret_cast<uintptr_t>(d_out) & 15u) != 0)   // k_static_resize stores float4 vectors at d_out + 4k
        return fail(err, err_len, KDL_EINVAL, "kdl_preprocess_u8_static: d_out %p is not 16-byte aligned (kdl_preprocess_u8 takes any "
                    "alignment; this entry stores 16-byte vectors)", (void *)d_out);
    MetaPack pack;
    memset(&pack, 0, sizeof pack);
    for (int i = 0; i < B; ++i) {
        if (h[i] < 1 || w[i] < 1 || pitch[i] < (int64_t)w[i] * 3 || offset[i] < 0)
            return fail(err, err_len, KDL_EINVAL, "kdl_preprocess_u8_static: image %d has bad geometry h=%d w=%d pitch=%lld offset=%lld", i,
                        h[i], w[i], (long long)pitch[i], (long long)offset[i]);
        pack.m[i].offset = offset[i]; pack.m[i].pitch = pitch[i]; pack.m[i].h = h[i]; pack.m[i].w = w[i];
    }
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= MAX_DEVICES)
        return fail(err, err_len, KDL_ECUDA, "kdl_preprocess_u8_static: cudaGetDevice -> %s (device %d)", cudaGetErrorString(e), dev);
    unsigned char *scratch = nullptr;
    e = device_scratch(dev, &scratch);
    if (e != cudaSuccess)
        return fail(err, err_len, KDL_ECUDA, "kdl_preprocess_u8_static: cudaGetSymbolAddress on device %d -> %s", dev, cudaGetErrorString(e));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // Nothing in the scratch has to start at zero: every table entry and metadata slot a call reads is written by that call's
    // own first kernel, so leftovers of the previous call are harmless.
    k_static_tables<<<(2 * B + 63) / 64, 64, 0, st>>>(pack, B, S, scratch);
    const long long total = (long long)B * S * S * 3;
    k_static_resize<<<(unsigned)((total / 4 + 256) / 256), 256, 0, st>>>(d_src, scratch, B, S, d_out);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(err, err_len, KDL_ECUDA, "kdl_preprocess_u8_static: kernel launch -> %s", cudaGetErrorString(e));
    return KDL_OK;
}

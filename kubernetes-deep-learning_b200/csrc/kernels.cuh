// CUDA-core kernels of the Xception path (sm_100a): preprocessing, entry conv, depthwise, pooling, head,
// and the CUDA-core GEMM used by the fp32 path / as the kernel-vs-kernel reference of the tcgen05 GEMM.
// Storage type T in {float, bf16}; arithmetic is fp32.  Layout everywhere: NHWC, channels innermost.
#pragma once
#include "common.cuh"

namespace kdl {

// ---------------------------------------------------------------------------------------------------
// K1: gateway preprocessing (model_server.py:18,53 -> Pillow NEAREST resize + x/127.5-1)
// ---------------------------------------------------------------------------------------------------
struct ImgMeta {
    int64_t offset;  // byte offset of pixel (0,0) from the batch base pointer
    int64_t pitch;   // bytes per source row
    int32_t h, w;
};

// One thread per (image, axis): Pillow's serial accumulated-double rule, reproduced operation for operation
// (a = n_in/n_out; o = a*0.5; idx = (int)o; o += a).  A closed form is NOT bit-exact.
__global__ void k_index_tables(const ImgMeta *__restrict__ meta, int B, int S, int32_t *__restrict__ xtab,
                               int32_t *__restrict__ ytab) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= 2 * B) return;
    int b = t >> 1, axis = t & 1;
    int n_in = axis ? meta[b].h : meta[b].w;
    int32_t *tab = (axis ? ytab : xtab) + (size_t)b * S;
    double a = __ddiv_rn((double)n_in, (double)S);
    double o = __dmul_rn(a, 0.5);
    for (int i = 0; i < S; ++i) {
        int s = o < 0.0 ? -1 : (int)o;  // Pillow COORD()
        s = max(0, min(s, n_in - 1));
        tab[i] = s;
        o = __dadd_rn(o, a);  // accumulated, never i*a, never fused
    }
}

// Gather + (optional) scale.
template <typename TOut> struct PxOut;
template <> struct PxOut<float> {  // scaled: the gateway's x/127.5-1 through an exact 256-entry table
    static __device__ __forceinline__ float cvt1(uint8_t v, const float *lut) { return lut[v]; }
    static __device__ __forceinline__ void st16(float *p, const uint8_t *s, const float *lut) {
        *reinterpret_cast<float4 *>(p) = make_float4(lut[s[0]], lut[s[1]], lut[s[2]], lut[s[3]]);
    }
};
template <> struct PxOut<uint8_t> {  // raw bytes (scaling is applied inside block1_conv1's load)
    static __device__ __forceinline__ uint8_t cvt1(uint8_t v, const float *) { return v; }
    static __device__ __forceinline__ void st16(uint8_t *p, const uint8_t *s, const float *) {
        uint32_t w[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) w[j] = (uint32_t)s[4 * j] | ((uint32_t)s[4 * j + 1] << 8) | ((uint32_t)s[4 * j + 2] << 16) | ((uint32_t)s[4 * j + 3] << 24);
        *reinterpret_cast<uint4 *>(p) = make_uint4(w[0], w[1], w[2], w[3]);
    }
};

// One CTA per output row (b, y): the source pixels are gathered into shared memory (each thread fetches whole RGB pixels through
// the x-table; the y-table lookup and the image metadata are read once per row), then the row is written out with aligned
// vector stores -- head / body / tail, because a 897-element row starts at an arbitrary alignment.
template <typename TOut>
__global__ void __launch_bounds__(128)
k_resize(const uint8_t *__restrict__ src, const ImgMeta *__restrict__ meta, const int32_t *__restrict__ xtab,
         const int32_t *__restrict__ ytab, int B, int S, TOut *__restrict__ out) {
    extern __shared__ uint8_t rowbuf[];  // S*3 bytes (+ 256-entry float LUT for TOut = float)
    const int b = blockIdx.y, y = blockIdx.x;
    const int n = S * 3;
    float *lut = reinterpret_cast<float *>(rowbuf + ((n + 15) & ~15));
    if (sizeof(TOut) == 4)
        for (int i = threadIdx.x; i < 256; i += blockDim.x) lut[i] = scale_px((uint8_t)i);
    const ImgMeta mt = meta[b];
    const uint8_t *srow = src + mt.offset + (long long)ytab[(size_t)b * S + y] * mt.pitch;
    const int32_t *xt = xtab + (size_t)b * S;
    // Source bytes are fetched as ALIGNED 32-bit words and funnel-shifted into place (an aligned word that holds a wanted byte can
    // never leave the allocation or the page): a pixel costs one or two word loads instead of three byte loads, and when the
    // source row already has the target width (identity x-table) the row is streamed word by word, fully coalesced.
    if (mt.w == S) {
        const uintptr_t a0 = reinterpret_cast<uintptr_t>(srow);
        const uint32_t *wbase = reinterpret_cast<const uint32_t *>(a0 & ~uintptr_t(3));
        const int k0 = (int)(a0 & 3);
        const unsigned sh = (unsigned)k0 * 8u;
        uint32_t *rb = reinterpret_cast<uint32_t *>(rowbuf);
        const int nwords = (n + 3) >> 2;                        // the last word may carry up to 3 bytes past the row: rowbuf has room
        for (int i = threadIdx.x; i < nwords; i += blockDim.x) {
            const uint32_t lo = __ldg(wbase + i);
            const uint32_t hi = (k0 != 0 && 4 * i + 4 < n + k0) ? __ldg(wbase + i + 1) : 0u;   // only if it still holds row bytes
            rb[i] = __funnelshift_r(lo, hi, sh);
        }
    } else {
        for (int x = threadIdx.x; x < S; x += blockDim.x) {
            const uintptr_t a = reinterpret_cast<uintptr_t>(srow) + (uintptr_t)xt[x] * 3u;
            const uint32_t *wp = reinterpret_cast<const uint32_t *>(a & ~uintptr_t(3));
            const unsigned k = (unsigned)(a & 3);
            const uint32_t lo = __ldg(wp);
            const uint32_t hi = k > 1 ? __ldg(wp + 1) : 0u;     // bytes k..k+2 reach into the next word only for k = 2, 3
            const uint32_t px = __funnelshift_r(lo, hi, k * 8u);
            rowbuf[3 * x + 0] = (uint8_t)px;
            rowbuf[3 * x + 1] = (uint8_t)(px >> 8);
            rowbuf[3 * x + 2] = (uint8_t)(px >> 16);
        }
    }
    __syncthreads();
    TOut *orow = out + ((size_t)b * S + y) * n;
    constexpr int VEC = 16 / sizeof(TOut);  // elements per 16-byte store
    const int mis = (int)((reinterpret_cast<uintptr_t>(orow) / sizeof(TOut)) % VEC);
    const int head = mis ? min(VEC - mis, n) : 0;
    for (int i = threadIdx.x; i < head; i += blockDim.x) orow[i] = PxOut<TOut>::cvt1(rowbuf[i], lut);
    const int nvec = (n - head) / VEC;
    for (int v = threadIdx.x; v < nvec; v += blockDim.x) {
        const int i = head + v * VEC;
        PxOut<TOut>::st16(orow + i, rowbuf + i, lut);
    }
    for (int i = head + nvec * VEC + threadIdx.x; i < n; i += blockDim.x) orow[i] = PxOut<TOut>::cvt1(rowbuf[i], lut);
}

// ---------------------------------------------------------------------------------------------------
// K2: block1_conv1  3x3 s2 VALID, 3 -> 32, folded BN + ReLU.  Input is the serving_default float tensor or
// the resized uint8 image (scaled on load with the gateway's exact arithmetic).  Memory-bound, K=27.
// One thread = one output pixel x 8 output channels (4 threads per pixel -> 64 B contiguous per pixel).
// ---------------------------------------------------------------------------------------------------
template <typename TIn> struct InPx;
template <> struct InPx<float> {
    static __device__ __forceinline__ float get(const float *p, const float *) { return __ldg(p); }
};
template <> struct InPx<uint8_t> {  // the gateway's x/127.5-1 through a 256-entry table built with the exact arithmetic
    static __device__ __forceinline__ float get(const uint8_t *p, const float *lut) { return lut[__ldg(p)]; }
};

// One thread = two horizontally adjacent output pixels x all 32 output channels: every weight vector read from shared memory feeds
// two pixels (the kernel is bound by shared-memory weight reads + fp32 FMA issue, not by HBM), and the two 3x3 windows share a column.
template <typename TIn, typename T>
__global__ void __launch_bounds__(128)
k_conv1(const TIn *__restrict__ in, const float *__restrict__ w /*[27][32]*/, const float *__restrict__ bias,
        T *__restrict__ out, int B, int S, int Ho) {
    __shared__ __align__(16) float ws[27 * 32];
    __shared__ __align__(16) float bs[32];
    __shared__ float lut[256];
    for (int i = threadIdx.x; i < 27 * 32; i += blockDim.x) ws[i] = w[i];
    if (threadIdx.x < 32) bs[threadIdx.x] = bias[threadIdx.x];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) lut[i] = scale_px((uint8_t)i);
    __syncthreads();
    pdl_launch_dependents();
    pdl_wait();
    const int Wp = (Ho + 1) >> 1;  // pixel pairs per output row
    const long long npairs = (long long)B * Ho * Wp;
    // persistent: the 3.4 KB of weights and the LUT are staged once per CTA, not once per 256 output pixels
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < npairs; t += (long long)gridDim.x * blockDim.x) {
    const int b = (int)(t / ((long long)Ho * Wp));
    const int r = (int)(t - (long long)b * Ho * Wp);
    const int oy = r / Wp, ox = (r - oy * Wp) * 2;
    const bool two = ox + 1 < Ho;
    const TIn *base = in + (((long long)b * S + 2 * oy) * S + 2 * ox) * 3;
    // 3 rows x 5 input pixels x 3 channels; the second window starts 2 pixels (6 values) to the right
    float v[3][15];
#pragma unroll
    for (int ky = 0; ky < 3; ++ky)
#pragma unroll
        for (int kk = 0; kk < 15; ++kk) v[ky][kk] = (kk < 9 || two) ? InPx<TIn>::get(base + (long long)ky * S * 3 + kk, lut) : 0.f;
    float2 acc0[16], acc1[16];
#pragma unroll
    for (int c = 0; c < 16; ++c) acc0[c] = acc1[c] = *reinterpret_cast<const float2 *>(&bs[c * 2]);
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) {
#pragma unroll
        for (int kk = 0; kk < 9; ++kk) {
            const float4 *wp = reinterpret_cast<const float4 *>(&ws[(ky * 9 + kk) * 32]);
            const float2 a0 = make_float2(v[ky][kk], v[ky][kk]), a1 = make_float2(v[ky][kk + 6], v[ky][kk + 6]);
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const float4 w4 = wp[c];
                const float2 wlo = make_float2(w4.x, w4.y), whi = make_float2(w4.z, w4.w);
                acc0[2 * c] = ffma2(a0, wlo, acc0[2 * c]);
                acc0[2 * c + 1] = ffma2(a0, whi, acc0[2 * c + 1]);
                acc1[2 * c] = ffma2(a1, wlo, acc1[2 * c]);
                acc1[2 * c + 1] = ffma2(a1, whi, acc1[2 * c + 1]);
            }
        }
    }
    T *o = out + (((long long)b * Ho + oy) * Ho + ox) * 32;
#pragma unroll
    for (int c = 0; c < 4; ++c)
        st8<T>(o + c * 8, relu4(make_float4(acc0[4 * c].x, acc0[4 * c].y, acc0[4 * c + 1].x, acc0[4 * c + 1].y)),
               relu4(make_float4(acc0[4 * c + 2].x, acc0[4 * c + 2].y, acc0[4 * c + 3].x, acc0[4 * c + 3].y)));
    if (two) {
#pragma unroll
        for (int c = 0; c < 4; ++c)
            st8<T>(o + 32 + c * 8, relu4(make_float4(acc1[4 * c].x, acc1[4 * c].y, acc1[4 * c + 1].x, acc1[4 * c + 1].y)),
                   relu4(make_float4(acc1[4 * c + 2].x, acc1[4 * c + 2].y, acc1[4 * c + 3].x, acc1[4 * c + 3].y)));
    }
    }
}

// ---------------------------------------------------------------------------------------------------
// K4: depthwise 3x3 s1 SAME (zero pad 1), optional ReLU on load.  Memory-bound.
// In NHWC one image row is a contiguous run of W*C values, so the op is a 3-row stencil with taps at
// -C, 0, +C along the flattened row: every load and store is fully coalesced.  Each thread owns 4 channels
// of one column and slides down `rows` output rows with the 3x3 window and its 9x4 weights in registers.
// ---------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(128)
k_dw3x3(const T *__restrict__ in, const float *__restrict__ w /*[9][C]*/, T *__restrict__ out, int H, int W,
        int C, int relu_in, int rows) {
    const int CV = C >> 2;
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= W * CV) return;
    const int x = v / CV;
    const int c = (v - x * CV) << 2;
    const int y0 = blockIdx.y * rows;
    const int b = blockIdx.z;
    float4 wk[9];
#pragma unroll
    for (int t = 0; t < 9; ++t) wk[t] = __ldg(reinterpret_cast<const float4 *>(w + (size_t)t * C + c));
    const T *base = in + (size_t)b * H * W * C + (size_t)x * C + c;
    T *obase = out + (size_t)b * H * W * C + (size_t)x * C + c;
    const bool hasL = x > 0, hasR = x < W - 1;
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    float4 r0[3], r1[3], r2[3];
    auto loadrow = [&](int yy, float4 *r) {
        if (yy < 0 || yy >= H) {
            r[0] = z; r[1] = z; r[2] = z;
        } else {
            const T *p = base + (size_t)yy * W * C;
            r[1] = V4<T>::ld(p);
            r[0] = hasL ? V4<T>::ld(p - C) : z;
            r[2] = hasR ? V4<T>::ld(p + C) : z;
            if (relu_in) { r[0] = relu4(r[0]); r[1] = relu4(r[1]); r[2] = relu4(r[2]); }
        }
    };
    loadrow(y0 - 1, r0);
    loadrow(y0, r1);
    const int y1 = min(y0 + rows, H);
    for (int y = y0; y < y1; ++y) {
        loadrow(y + 1, r2);
        float4 acc = z;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            acc = fma4(r0[k], wk[k], acc);
            acc = fma4(r1[k], wk[3 + k], acc);
            acc = fma4(r2[k], wk[6 + k], acc);
        }
        V4<T>::st(obase + (size_t)y * W * C, acc);
#pragma unroll
        for (int k = 0; k < 3; ++k) { r0[k] = r1[k]; r1[k] = r2[k]; }
    }
}

// ---------------------------------------------------------------------------------------------------
// K6: MaxPooling2D(3, strides 2, 'same') with TF padding (pad_lo = 0 for even inputs, 1 for odd; -inf pad)
// fused with the residual add.  One thread = one output pixel x 4 channels.
// ---------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
k_maxpool_add(const T *__restrict__ in, const T *__restrict__ res, T *__restrict__ out, int B, int H, int W,
              int C, int ld /*pixel pitch in elements (>= C)*/, int Ho, int Wo, int pad_lo) {
    const int CV = C >> 3;  // 8 channels per thread
    long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    long long total = (long long)B * Ho * Wo * CV;
    if (t >= total) return;
    int cv = (int)(t % CV);
    long long p = t / CV;
    int ox = (int)(p % Wo);
    long long q = p / Wo;
    int oy = (int)(q % Ho);
    int b = (int)(q / Ho);
    const int c = cv << 3;
    const float NI = -INFINITY;
    float4 m0 = make_float4(NI, NI, NI, NI), m1 = m0;
    float4 v0[9], v1[9];
#pragma unroll
    for (int dy = 0; dy < 3; ++dy) {
#pragma unroll
        for (int dx = 0; dx < 3; ++dx) {
            int iy = 2 * oy - pad_lo + dy, ix = 2 * ox - pad_lo + dx;
            bool ok = iy >= 0 && iy < H && ix >= 0 && ix < W;
            const T *ptr = in + (((size_t)b * H + (ok ? iy : 0)) * W + (ok ? ix : 0)) * ld + c;
            v0[dy * 3 + dx] = m0;
            v1[dy * 3 + dx] = m0;
            if (ok) ld8<T>(ptr, v0[dy * 3 + dx], v1[dy * 3 + dx]);
        }
    }
    size_t o = (((size_t)b * Ho + oy) * Wo + ox) * ld + c;
    float4 r0, r1;
    ld8<T>(res + o, r0, r1);
#pragma unroll
    for (int k = 0; k < 9; ++k) { m0 = max4(m0, v0[k]); m1 = max4(m1, v1[k]); }
    st8<T>(out + o, add4(m0, r0), add4(m1, r1));
}

// Compact copy of the even-index pixels: the input of a Conv2D(1x1, strides 2, 'same') residual projection
// seen as a dense [B*Ho*Wo, C] matrix (so the projection is a plain GEMM for the tcgen05 kernel).
template <typename T>
__global__ void __launch_bounds__(256)
k_subsample2(const T *__restrict__ in, T *__restrict__ out, int B, int H, int W, int C, int ld, int Ho, int Wo) {
    const int CV = C >> 2;
    pdl_launch_dependents();
    pdl_wait();
    long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    long long total = (long long)B * Ho * Wo * CV;
    if (t >= total) return;
    int cv = (int)(t % CV);
    long long p = t / CV;
    int ox = (int)(p % Wo);
    long long q = p / Wo;
    int oy = (int)(q % Ho);
    int b = (int)(q / Ho);
    V4<T>::st(out + p * ld + cv * 4, V4<T>::ld(in + (((size_t)b * H + 2 * oy) * W + 2 * ox) * ld + cv * 4));
}

// ---------------------------------------------------------------------------------------------------
// CUDA-core GEMM with fused epilogue:  out[m,n] = act(sum_k A(m,k) W[k,n] + bias[n]) (+ res[m,n]).
// A is addressed through a row map so the same kernel serves
//   mode 0: dense [M,K] rows (pointwise 1x1 on a depthwise output),
//   mode 1: Conv2D(1x1, strides 2, 'same') -- row m = pixel (b, 2oy, 2ox) of an NHWC tensor,
//   mode 2: Conv2D(3x3, s1, VALID) with Cin = 32 as an implicit GEMM, K = 9*32 ordered (ky,kx,ci).
// fp32 path (precision contract <= 1e-4) and kernel-vs-kernel reference for the tcgen05 GEMM.
// ---------------------------------------------------------------------------------------------------
struct RowMap {
    int mode;
    int Hin, Win, Ho, Wo, C;  // modes 1, 2
};

__device__ __forceinline__ long long row_offset(const RowMap &rm, int m, int K) {
    if (rm.mode == 0) return (long long)m * K;
    int hw = rm.Ho * rm.Wo;
    int b = m / hw;
    int r = m - b * hw;
    int oy = r / rm.Wo, ox = r - oy * rm.Wo;
    int s = rm.mode == 1 ? 2 : 1;
    return (((long long)b * rm.Hin + s * oy) * rm.Win + s * ox) * rm.C;
}

constexpr int GS_BM = 128, GS_BN = 64, GS_BK = 16;

template <typename T>
__global__ void __launch_bounds__(256)
k_gemm_simple(const T *__restrict__ A, const T *__restrict__ Wkn, const float *__restrict__ bias,
              const T *res, T *out, int M, int N, int K, int relu, RowMap rm) {
    __shared__ __align__(16) float As[GS_BK][GS_BM + 4];
    __shared__ __align__(16) float Bs[GS_BK][GS_BN];
    const int tid = threadIdx.x;
    const int m0 = blockIdx.y * GS_BM, n0 = blockIdx.x * GS_BN;
    // A loader: thread -> (row = tid/2, 8 consecutive k)
    const int arow = tid >> 1, akoff = (tid & 1) * 8;
    const int am = m0 + arow;
    const long long aoff = am < M ? row_offset(rm, am, K) : 0;
    // B loader: thread -> (k = tid/16, 4 consecutive n)
    const int bk = tid >> 4, bn = (tid & 15) * 4;
    const int ty = tid >> 4, tx = tid & 15;  // 16 x 16 thread grid: 8 rows x 4 cols each
    float acc[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    for (int k0 = 0; k0 < K; k0 += GS_BK) {
        // implicit-GEMM tap offset (mode 2): k-block lies inside one (ky,kx) tap because 32 % 16 == 0
        long long koff = k0;
        if (rm.mode == 2) {
            int tap = k0 / rm.C, ci = k0 - tap * rm.C;
            int ky = tap / 3, kx = tap - ky * 3;
            koff = ((long long)ky * rm.Win + kx) * rm.C + ci;
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            int k = k0 + akoff + h * 4;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (am < M && k + 4 <= K) v = V4<T>::ld(A + aoff + koff + akoff + h * 4);
            As[akoff + h * 4 + 0][arow] = v.x;
            As[akoff + h * 4 + 1][arow] = v.y;
            As[akoff + h * 4 + 2][arow] = v.z;
            As[akoff + h * 4 + 3][arow] = v.w;
        }
        {
            int k = k0 + bk, n = n0 + bn;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (k < K && n + 4 <= N) v = V4<T>::ld(Wkn + (long long)k * N + n);
            *reinterpret_cast<float4 *>(&Bs[bk][bn]) = v;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < GS_BK; ++k) {
            float4 a0 = *reinterpret_cast<const float4 *>(&As[k][ty * 8]);
            float4 a1 = *reinterpret_cast<const float4 *>(&As[k][ty * 8 + 4]);
            float4 bv = *reinterpret_cast<const float4 *>(&Bs[k][tx * 4]);
            float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            float bb[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
        }
        __syncthreads();
    }
    const int n = n0 + tx * 4;
    if (n + 4 > N) return;
    float4 bv = *reinterpret_cast<const float4 *>(bias + n);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        int m = m0 + ty * 8 + i;
        if (m >= M) break;
        float4 v = make_float4(acc[i][0] + bv.x, acc[i][1] + bv.y, acc[i][2] + bv.z, acc[i][3] + bv.w);
        if (relu) v = relu4(v);
        size_t o = (size_t)m * N + n;
        if (res) v = add4(v, V4<T>::ld(res + o));
        V4<T>::st(out + o, v);
    }
}

// ---------------------------------------------------------------------------------------------------
// K7 + head: GlobalAveragePooling2D -> [Dense(N1)+bias, ReLU] -> Dense(N2)+bias, in two kernels so that a single image still
// spreads over 16 CTAs (batch-1 latency) and a large batch does not re-read the dense kernel once per image row-by-row:
//   k_gap_partial : CTA (slice s, image b): mean over H*W of channels [128 s, 128 s + 128) (warps split the pixels, shuffle-free
//                   smem reduction), then the slice's contribution to every first-dense output -> partial[b][s][*]
//   k_head_finish : CTA per image: fixed-order sum of the 16 partials (deterministic), +bias, ReLU, second dense (warp reductions)
// (N1 == 0 means the head is a single Dense(C -> N2): the partials are then the logits' partial sums.)
// ---------------------------------------------------------------------------------------------------
constexpr int HEAD_SLICE = 128;   // channels per slice
constexpr int HEAD_MAXN = 128;    // max width of the first dense layer

template <typename T>
__global__ void __launch_bounds__(256)
k_gap_partial(const T *__restrict__ in, int HW, int C, const float *__restrict__ Wd /*[C][Nd]*/, int Nd,
              float *__restrict__ feat_out /*nullable [B,C]*/, float *__restrict__ partial /*[B][C/128][HEAD_MAXN]*/) {
    __shared__ float4 red[8][32];
    __shared__ float feat[HEAD_SLICE];
    pdl_launch_dependents();
    pdl_wait();
    const int s = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int nslices = C / HEAD_SLICE;
    const T *base = in + (size_t)b * HW * C + (size_t)s * HEAD_SLICE + lane * 4;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int p = warp; p < HW; p += 8) acc = add4(acc, V4<T>::ld(base + (size_t)p * C));
    red[warp][lane] = acc;
    __syncthreads();
    if (warp == 0) {
        float4 t = red[0][lane];
#pragma unroll
        for (int w = 1; w < 8; ++w) t = add4(t, red[w][lane]);
        const float inv = 1.0f / (float)HW;
        feat[lane * 4 + 0] = t.x * inv; feat[lane * 4 + 1] = t.y * inv; feat[lane * 4 + 2] = t.z * inv; feat[lane * 4 + 3] = t.w * inv;
    }
    __syncthreads();
    if (feat_out && tid < HEAD_SLICE) feat_out[(size_t)b * C + (size_t)s * HEAD_SLICE + tid] = feat[tid];
    if (tid < Nd) {  // coalesced over the outputs: thread j walks the slice's 128 rows of the dense kernel
        const float *w = Wd + (size_t)s * HEAD_SLICE * Nd + tid;
        float a = 0.f;
#pragma unroll 8
        for (int k = 0; k < HEAD_SLICE; ++k) a = fmaf(feat[k], __ldg(w + (size_t)k * Nd), a);
        partial[((size_t)b * nslices + s) * HEAD_MAXN + tid] = a;
    }
}

__global__ void __launch_bounds__(128)
k_head_finish(const float *__restrict__ partial, int nslices, const float *__restrict__ b1, int N1, const float *__restrict__ W2,
              const float *__restrict__ b2, int N2, float *__restrict__ logits) {
    __shared__ float hid[HEAD_MAXN];
    pdl_launch_dependents();
    pdl_wait();
    const int b = blockIdx.x, tid = threadIdx.x;
    const int Nd = N1 > 0 ? N1 : N2;
    if (tid < Nd) {
        float s = 0.f;
        for (int i = 0; i < nslices; ++i) s += partial[((size_t)b * nslices + i) * HEAD_MAXN + tid];  // fixed order
        if (N1 > 0) hid[tid] = fmaxf(s + b1[tid], 0.f);
        else logits[(size_t)b * N2 + tid] = s + b2[tid];
    }
    if (N1 == 0) return;
    __syncthreads();
    const int warp = tid >> 5, lane = tid & 31;
    for (int n = warp; n < N2; n += 4) {
        float s = 0.f;
        for (int k = lane; k < N1; k += 32) s = fmaf(hid[k], __ldg(W2 + (size_t)k * N2 + n), s);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) logits[(size_t)b * N2 + n] = s + b2[n];
    }
}

// storage -> float32 copy (debug taps)
template <typename T>
__global__ void k_to_f32(const T *__restrict__ in, float *__restrict__ out, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = ld1<T>(in + i);
}

}  // namespace kdl

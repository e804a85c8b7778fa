// K6 (bf16 path): MaxPooling2D(3, strides 2, 'same') + residual add with TMA-staged input tiles.  Memory-bound.
//
// Same skeleton as dw_tma.cuh: the NHWC input is a 4-D tensor (C, W, H, B); one tile = 64 channels x (2*TWo+1) x (2*THo+1) input
// pixels fetched by one cp.async.bulk.tensor.4d into a 3-stage mbarrier ring by a producer warp; consumer threads own 4 channels of
// one output column and walk down a strip of output rows, reusing the column-max of the input row two windows share.  The max is
// taken on packed bf16x2 (exact), TF-SAME padding ((0,1) for even inputs, (1,1) for odd) is a per-tap validity mask with -inf --
// TMA's zero fill is never allowed to win a max.  The residual add is fp32; optionally the even-index pixels of the result are
// also written compactly as the next block's strided-projection input (replaces a separate subsample pass).
#pragma once
#include "dw_tma.cuh"

namespace kdl {
namespace poolt {

constexpr int STAGES = 3;
constexpr int CB = 64;
constexpr uint32_t TILE_BUDGET = 61440;

struct Params {
    int B, H, W, C, ld;      // input geometry / pixel pitch (elements)
    int Ho, Wo, pad_lo;
    int TWo, THo, BW, BH;    // output tile and staged input box
    int NS, RS;
    int ntx, nty, nch, total_tiles;
    int Hs, Ws;              // geometry of the optional even-pixel copy
    uint32_t tile_bytes, res_bytes, stage_bytes;
    int consumer_threads;
};

inline Params make_params(int B, int H, int W, int C, int ld, int Ho, int Wo, int pad_lo) {
    Params p;
    p.B = B; p.H = H; p.W = W; p.C = C; p.ld = ld; p.Ho = Ho; p.Wo = Wo; p.pad_lo = pad_lo;
    p.ntx = (Wo + 9) / 10;
    p.TWo = (Wo + p.ntx - 1) / p.ntx;
    p.BW = 2 * p.TWo + 1;
    int max_bh = (int)(TILE_BUDGET / (uint32_t)(p.BW * 128));
    int tho_max = (max_bh - 1) / 2;
    if (tho_max < 1) tho_max = 1;
    if (tho_max > Ho) tho_max = Ho;
    p.nty = (Ho + tho_max - 1) / tho_max;
    p.THo = (Ho + p.nty - 1) / p.nty;
    p.BH = 2 * p.THo + 1;
    int per_strip = 16 * p.TWo;
    int ns = 480 / per_strip;
    if (ns > 3) ns = 3;
    if (ns > p.THo) ns = p.THo;
    if (ns < 1) ns = 1;
    p.RS = (p.THo + ns - 1) / ns;
    p.NS = (p.THo + p.RS - 1) / p.RS;
    p.consumer_threads = ((per_strip * p.NS + 31) / 32) * 32;
    p.nch = (C + CB - 1) / CB;
    p.total_tiles = B * p.nty * p.ntx * p.nch;
    p.tile_bytes = (uint32_t)(p.BW * p.BH * 128);
    p.res_bytes = (uint32_t)(p.TWo * p.THo * 128);   // the residual tile rides in the same stage (its loads become asynchronous too)
    p.stage_bytes = p.tile_bytes + p.res_bytes;
    p.Hs = (Ho + 1) / 2; p.Ws = (Wo + 1) / 2;
    return p;
}
inline uint32_t smem_bytes(const Params &p) { return STAGES * p.stage_bytes + 256; }

__device__ __forceinline__ uint2 ld_shared_u2(uint32_t addr) {
    uint2 v;
    asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
    return v;
}
__device__ __forceinline__ uint2 max2(uint2 a, uint2 b) {
    asm("max.bf16x2 %0, %0, %1;" : "+r"(a.x) : "r"(b.x));
    asm("max.bf16x2 %0, %0, %1;" : "+r"(a.y) : "r"(b.y));
    return a;
}

__global__ void __launch_bounds__(512, 1)
k_pool_tma(const __grid_constant__ CUtensorMap tmIn, const __grid_constant__ CUtensorMap tmRes, bf16 *__restrict__ out,
           bf16 *__restrict__ sub, Params p) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (tc::smem_u32(smem_raw) + 127u) & ~127u;
    const uint32_t bar_base = base + STAGES * p.stage_bytes;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_cwarps = p.consumer_threads >> 5;
    if (threadIdx.x == 0) {
        tc::prefetch_tmap(&tmIn);
        tc::prefetch_tmap(&tmRes);
        for (int s = 0; s < STAGES; ++s) { tc::mbar_init(full_bar(s), 1); tc::mbar_init(empty_bar(s), n_cwarps); }
        tc::fence_barrier_init();
    }
    __syncthreads();
    pdl_launch_dependents();
    pdl_wait();
    if (warp == n_cwarps) {
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int t = blockIdx.x; t < p.total_tiles; t += gridDim.x) {
                const int ch = t % p.nch;
                int s = t / p.nch;
                const int tx = s % p.ntx; s /= p.ntx;
                const int ty = s % p.nty;
                const int b = s / p.nty;
                tc::mbar_wait(empty_bar(stage), phase ^ 1u);
                tc::mbar_expect_tx(full_bar(stage), p.stage_bytes);
                dwt::tma_load_4d(base + stage * p.stage_bytes, &tmIn, full_bar(stage), ch * CB, 2 * tx * p.TWo - p.pad_lo,
                                 2 * ty * p.THo - p.pad_lo, b);
                dwt::tma_load_4d(base + stage * p.stage_bytes + p.tile_bytes, &tmRes, full_bar(stage), ch * CB, tx * p.TWo, ty * p.THo, b);
                if (++stage == STAGES) { stage = 0; phase ^= 1u; }
            }
        }
        return;
    }
    if (warp > n_cwarps) return;

    const int cg = threadIdx.x & 15;
    const int rest = threadIdx.x >> 4;
    const int col = rest % p.TWo;
    const int strip = rest / p.TWo;
    const bool active = strip < p.NS;
    const uint2 NEG = make_uint2(0xFF80FF80u, 0xFF80FF80u);  // -inf, -inf (bf16x2)
    int stage = 0;
    uint32_t phase = 0;
    for (int t = blockIdx.x; t < p.total_tiles; t += gridDim.x) {
        const int ch = t % p.nch;
        int s = t / p.nch;
        const int tx = s % p.ntx; s /= p.ntx;
        const int ty = s % p.nty;
        const int b = s / p.nty;
        tc::mbar_wait(full_bar(stage), phase);
        const int c = ch * CB + cg * 4;
        const int ox = tx * p.TWo + col;
        const int oy_lo = ty * p.THo + strip * p.RS;
        int oy_hi = min(ty * p.THo + min((strip + 1) * p.RS, p.THo), p.Ho);
        if (active && c < p.C && ox < p.Wo && oy_lo < oy_hi) {
            const uint32_t tile = base + stage * p.stage_bytes;
            // residual tile: [THo][TWo][64ch]
            uint32_t raddr = tile + p.tile_bytes + ((uint32_t)((oy_lo - ty * p.THo) * p.TWo + col)) * 128u + (uint32_t)cg * 8u;
            const uint32_t rpitch = (uint32_t)p.TWo * 128u;
            // staged column of tap dx = 2*col + dx; global input column = 2*ox - pad_lo + dx
            bool vx[3];
#pragma unroll
            for (int dx = 0; dx < 3; ++dx) { int ix = 2 * ox - p.pad_lo + dx; vx[dx] = ix >= 0 && ix < p.W; }
            const uint32_t coladdr = tile + (uint32_t)(2 * col) * 128u + (uint32_t)cg * 8u;
            const uint32_t rowpitch = (uint32_t)p.BW * 128u;
            auto rowmax = [&](int srow, int iy) -> uint2 {  // max over the 3 columns of one staged row (-inf if the row is padding)
                if (iy < 0 || iy >= p.H) return NEG;
                const uint32_t a = coladdr + (uint32_t)srow * rowpitch;
                uint2 m = NEG;
#pragma unroll
                for (int dx = 0; dx < 3; ++dx)
                    if (vx[dx]) m = max2(m, ld_shared_u2(a + 128u * dx));
                return m;
            };
            int srow = 2 * (oy_lo - ty * p.THo);         // staged row of the window's first input row
            int iy = 2 * oy_lo - p.pad_lo;               // its global input row
            uint2 r0 = rowmax(srow, iy);
            size_t o = (((size_t)b * p.Ho + oy_lo) * p.Wo + ox) * p.ld + c;
            const size_t orow = (size_t)p.Wo * p.ld;
            for (int oy = oy_lo; oy < oy_hi; ++oy, srow += 2, iy += 2, o += orow, raddr += rpitch) {
                const uint2 r1 = rowmax(srow + 1, iy + 1);
                const uint2 r2 = rowmax(srow + 2, iy + 2);
                const uint2 m = max2(max2(r0, r1), r2);
                r0 = r2;  // shared with the next window
                const uint2 rr = ld_shared_u2(raddr);
                uint2 v;
                v.x = tc::pack_bf16(__uint_as_float(m.x << 16) + __uint_as_float(rr.x << 16),
                                    __uint_as_float(m.x & 0xffff0000u) + __uint_as_float(rr.x & 0xffff0000u));
                v.y = tc::pack_bf16(__uint_as_float(m.y << 16) + __uint_as_float(rr.y << 16),
                                    __uint_as_float(m.y & 0xffff0000u) + __uint_as_float(rr.y & 0xffff0000u));
                *reinterpret_cast<uint2 *>(out + o) = v;
                if (sub && !(oy & 1) && !(ox & 1))
                    *reinterpret_cast<uint2 *>(sub + (((size_t)b * p.Hs + (oy >> 1)) * p.Ws + (ox >> 1)) * p.ld + c) = v;
            }
        }
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(empty_bar(stage));
        if (++stage == STAGES) { stage = 0; phase ^= 1u; }
    }
}

inline int configure(char *err) {
    cudaError_t e = cudaFuncSetAttribute(k_pool_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) {
        snprintf(err, 512, "cudaFuncSetAttribute(k_pool_tma): %s", cudaGetErrorString(e));
        return KDL_ECUDA;
    }
    return KDL_OK;
}

inline int launch(const tc::TmaEncoder &enc, const bf16 *in, const bf16 *res, bf16 *out, bf16 *sub, int B, int H, int W, int C, int ld,
                  int Ho, int Wo, int pad_lo, cudaStream_t st, char *err) {
    Params p = make_params(B, H, W, C, ld, Ho, Wo, pad_lo);
    cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
    cuuint64_t strides[3] = {(cuuint64_t)ld * 2, (cuuint64_t)W * ld * 2, (cuuint64_t)H * W * ld * 2};
    cuuint32_t box[4] = {(cuuint32_t)CB, (cuuint32_t)p.BW, (cuuint32_t)p.BH, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUtensorMap tm;
    if (enc.fn(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<bf16 *>(in), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
        snprintf(err, 512, "cuTensorMapEncodeTiled(4d) failed for pool input [%d,%d,%d,%d] box [%d,%d]", B, H, W, C, p.BW, p.BH);
        return KDL_ECUDA;
    }
    cuuint64_t rdims[4] = {(cuuint64_t)C, (cuuint64_t)Wo, (cuuint64_t)Ho, (cuuint64_t)B};
    cuuint64_t rstrides[3] = {(cuuint64_t)ld * 2, (cuuint64_t)Wo * ld * 2, (cuuint64_t)Ho * Wo * ld * 2};
    cuuint32_t rbox[4] = {(cuuint32_t)CB, (cuuint32_t)p.TWo, (cuuint32_t)p.THo, 1};
    CUtensorMap tmr;
    if (enc.fn(&tmr, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<bf16 *>(res), rdims, rstrides, rbox, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
        snprintf(err, 512, "cuTensorMapEncodeTiled(4d) failed for pool residual [%d,%d,%d,%d]", B, Ho, Wo, C);
        return KDL_ECUDA;
    }
    int per_sm = (int)(220u * 1024u / smem_bytes(p));
    if (per_sm > 4) per_sm = 4;
    if (per_sm < 1) per_sm = 1;
    const int cap = tc::num_sms() * per_sm;
    const int grid = p.total_tiles < cap ? p.total_tiles : cap;
    launch_k(k_pool_tma, dim3(grid), dim3(p.consumer_threads + 32), smem_bytes(p), st, tm, tmr, out, sub, p);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        snprintf(err, 512, "k_pool_tma launch (B=%d H=%d C=%d threads=%d smem=%u): %s", B, H, C, p.consumer_threads + 32, smem_bytes(p),
                 cudaGetErrorString(e));
        return KDL_ECUDA;
    }
    return KDL_OK;
}

}  // namespace poolt
}  // namespace kdl

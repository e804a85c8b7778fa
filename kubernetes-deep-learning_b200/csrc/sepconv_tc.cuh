// Fused SeparableConv2D for the big early layers (K <= 256, N <= 256; HBM-bound): depthwise 3x3 on CUDA cores feeding the pointwise
// tcgen05 GEMM through shared memory, so the depthwise output never touches HBM (saves one full write + read of the widest tensors
// of the network).  One CTA = 14 warps:
//   warp 0      TMA producer : per 64-channel chunk, the input halo tile (10 x 18 pixels, OOB zero fill = SAME padding) + its 9x64
//                              fp32 depthwise weights into a 3-stage ring; the matching [N x 64] pointwise weight block into a 2-stage ring
//   warps 6..13 depthwise    : 4 channels x 2 columns x 4 rows per thread (the FFMA2 scatter of dw_tma.cuh); results are rounded to bf16
//                              and written straight into the K-major 128B-swizzled A tile the MMA reads (generic-proxy stores, then
//                              fence.proxy.async + mbarrier arrive)
//   warp 1      MMA issuer   : tcgen05.mma M128 x N x K16, accumulators double-buffered in TMEM
//   warps 2..5  epilogue     : tcgen05.ld -> +bias / ReLU -> swizzled smem -> 4-D TMA store of the 8 x 16 pixel tile (clipped at the edges)
// The GEMM row order inside a tile is r = ty*16 + tx, which is exactly the order a (64ch, 16, 2, 1) TMA store box expects.
#pragma once
#include "dw_tma.cuh"
#include "gemm_tc.cuh"

namespace kdl {
namespace sepc {

using namespace kdl::tc;

constexpr int TH = 8, TW = 16;
constexpr int XH = TH + 2, XW = TW + 2;
constexpr int CB = 64;
constexpr int NX = 3, NA = 2, NB = 2;
constexpr int BN = 256;
constexpr uint32_t X_BYTES = XH * XW * 128;
constexpr uint32_t DWW_BYTES = 9 * CB * 4;
constexpr uint32_t XSTAGE = X_BYTES + DWW_BYTES;
constexpr uint32_t A_BYTES = 128 * 128;
constexpr uint32_t B_BYTES = BN * 128;
#ifndef KDL_SEP_EPI_WARPS
#define KDL_SEP_EPI_WARPS 4
#endif
constexpr int EPI_WARPS = KDL_SEP_EPI_WARPS;          // 4, or 8 = two per TMEM lane quarter taking alternate 64-column chunks (needs NX = 2 to fit).
                                                      // Measured equal (2.14 vs 2.16 ms): the role is not short of warps, it shares the FMA/ALU
                                                      // pipes with the depthwise warps
constexpr uint32_t EPI_BYTES = EPI_WARPS * 2 * 4096;
constexpr uint32_t BIAS_BYTES = BN * 4;
constexpr uint32_t SMEM_BYTES = NA * A_BYTES + NB * B_BYTES + EPI_BYTES + NX * XSTAGE + BIAS_BYTES + 256 + 1024;
static_assert(SMEM_BYTES <= 232448, "exceeds 227 KB");
static_assert(XSTAGE % 128 == 0, "TMA destinations must stay 128-byte aligned");
#ifndef KDL_SEP_DW_WARPS
#define KDL_SEP_DW_WARPS 8
#endif
constexpr int DW_WARPS = KDL_SEP_DW_WARPS;             // 8 = two 4-row strips, 12 = three strips (3,3,2 rows): more warps to hide LDS/FMA latency
constexpr int DW_ROWS = DW_WARPS == 8 ? 4 : 3;
constexpr int FIRST_DW_WARP = 2 + EPI_WARPS;
constexpr int NUM_THREADS_SEP = (FIRST_DW_WARP + DW_WARPS) * 32;

struct Geom {
    int B, H, W, K, N;       // spatial size (stride 1, SAME), channels in / out
    int ntx, nty, num_tiles, kchunks;
    int relu_out;
    int prof_slot;           // probe builds (-DKDL_MID_PROF): which row block of g_sep_prof this launch writes
};

// -DKDL_MID_PROF (probe builds only): per-role wait / work cycle counters, [launch slot][CTA][counter]
#ifdef KDL_MID_PROF
__device__ unsigned long long g_sep_prof[4][160][16];
#define SP_T0() long long sp_t0 = clock64()
#define SP_ADD(var) do { long long sp_t1 = clock64(); var += sp_t1 - sp_t0; sp_t0 = sp_t1; } while (0)
#define SP_DECL(...) long long __VA_ARGS__
#define SP_STORE(idx, v) g_sep_prof[g.prof_slot & 3][blockIdx.x][idx] = (unsigned long long)(v)
#else
#define SP_T0()
#define SP_ADD(var)
#define SP_DECL(...)
#define SP_STORE(idx, v)
#endif

__device__ __forceinline__ void tma_store_4d(const CUtensorMap *tm, uint32_t src, int c0, int c1, int c2, int c3) {
    asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(tm), "r"(src), "r"(c0),
                 "r"(c1), "r"(c2), "r"(c3)
                 : "memory");
}

template <bool RELU_IN>
__global__ void __launch_bounds__(NUM_THREADS_SEP, 1)
k_sepconv_tc(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmOut,
             const float *__restrict__ dwpad, const float *__restrict__ bias, Geom g) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t a_base = smem_base;
    const uint32_t b_base = a_base + NA * A_BYTES;
    const uint32_t epi_base = b_base + NB * B_BYTES;
    const uint32_t x_base = epi_base + EPI_BYTES;
    const uint32_t bias_base = x_base + NX * XSTAGE;
    const uint32_t bar_base = bias_base + BIAS_BYTES;
    // barriers: x_full[NX] x_empty[NX] a_full[NA] a_empty[NA] b_full[NB] b_empty[NB] tfull[2] tempty[2]
    auto x_full = [&](int s) { return bar_base + 8u * s; };
    auto x_empty = [&](int s) { return bar_base + 8u * (NX + s); };
    auto a_full = [&](int s) { return bar_base + 8u * (2 * NX + s); };
    auto a_empty = [&](int s) { return bar_base + 8u * (2 * NX + NA + s); };
    auto b_full = [&](int s) { return bar_base + 8u * (2 * NX + 2 * NA + s); };
    auto b_empty = [&](int s) { return bar_base + 8u * (2 * NX + 2 * NA + NB + s); };
    auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * NX + 2 * NA + 2 * NB + a); };
    auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * NX + 2 * NA + 2 * NB + 2 + a); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * NX + 2 * NA + 2 * NB + 4);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmX);
        prefetch_tmap(&tmB);
        prefetch_tmap(&tmOut);
        for (int s = 0; s < NX; ++s) { mbar_init(x_full(s), 1); mbar_init(x_empty(s), DW_WARPS); }
        for (int s = 0; s < NA; ++s) { mbar_init(a_full(s), DW_WARPS); mbar_init(a_empty(s), 1); }
        for (int s = 0; s < NB; ++s) { mbar_init(b_full(s), 1); mbar_init(b_empty(s), 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(tfull_bar(a), 1); mbar_init(tempty_bar(a), EPI_WARPS); }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, 512);
    if (warp >= 2 && warp < FIRST_DW_WARP) {  // bias: one N tile for the whole kernel
        const int et = threadIdx.x - 64;
        for (int j = et; j < BN; j += 32 * EPI_WARPS) {
            float b = j < g.N ? __ldg(bias + j) : 0.f;
            asm volatile("st.shared.f32 [%0], %1;" ::"r"(bias_base + 4u * j), "f"(b) : "memory");
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
    pdl_launch_dependents();
    pdl_wait();

    if (warp == 0) {
        // ================= TMA producer =================
        if (lane == 0) {
            int xs = 0, bs = 0;
            uint32_t xph = 0, bph = 0;
            SP_DECL(w_xe = 0, w_be = 0, w_i = 0);
            SP_T0();
            for (int tile = blockIdx.x; tile < g.num_tiles; tile += gridDim.x) {
                const int tx = tile % g.ntx;
                const int r = tile / g.ntx;
                const int ty = r % g.nty, b = r / g.nty;
                for (int kc = 0; kc < g.kchunks; ++kc) {
                    SP_ADD(w_i);
                    mbar_wait(x_empty(xs), xph ^ 1u);
                    SP_ADD(w_xe);
                    mbar_expect_tx(x_full(xs), XSTAGE);
                    const uint32_t xd = x_base + xs * XSTAGE;
                    dwt::tma_load_4d(xd, &tmX, x_full(xs), kc * CB, tx * TW - 1, ty * TH - 1, b);
                    dwt::bulk_load_1d(xd + X_BYTES, dwpad + (size_t)kc * 9 * CB, DWW_BYTES, x_full(xs));
                    if (++xs == NX) { xs = 0; xph ^= 1u; }
                    SP_ADD(w_i);
                    mbar_wait(b_empty(bs), bph ^ 1u);
                    SP_ADD(w_be);
                    mbar_expect_tx(b_full(bs), B_BYTES);
                    tma_load_2d(b_base + bs * B_BYTES, &tmB, b_full(bs), kc * CB, 0);
                    if (++bs == NB) { bs = 0; bph ^= 1u; }
                }
            }
            SP_ADD(w_i);
            SP_STORE(0, w_xe); SP_STORE(1, w_be); SP_STORE(2, w_i);
        }
    } else if (warp == 1) {
        // ================= MMA issuer =================
        if (lane == 0) {
            int as = 0, bs = 0, it = 0;
            uint32_t aph = 0, bph = 0;
            const uint32_t idesc = make_idesc(BLOCK_M, (uint32_t)((g.N + 15) & ~15));
            SP_DECL(w_te = 0, w_af = 0, w_bf = 0, w_i = 0);
            SP_T0();
            for (int tile = blockIdx.x; tile < g.num_tiles; tile += gridDim.x, ++it) {
                const int acc = it & 1;
                SP_ADD(w_i);
                mbar_wait(tempty_bar(acc), (uint32_t)((it >> 1) & 1) ^ 1u);
                SP_ADD(w_te);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN;
                for (int kc = 0; kc < g.kchunks; ++kc) {
                    SP_ADD(w_i);
                    mbar_wait(a_full(as), aph);
                    SP_ADD(w_af);
                    mbar_wait(b_full(bs), bph);
                    SP_ADD(w_bf);
                    tc_fence_after();
                    const uint64_t adesc = make_smem_desc<128>(a_base + as * A_BYTES);
                    const uint64_t bdesc = make_smem_desc<128>(b_base + bs * B_BYTES);
                    const int k_left = g.K - kc * CB;
                    const int ksteps = k_left >= CB ? CB / UMMA_K : (k_left + UMMA_K - 1) / UMMA_K;
                    for (int k = 0; k < ksteps; ++k) umma_bf16(d_tmem, adesc + 2u * k, bdesc + 2u * k, idesc, (kc | k) != 0);
                    umma_commit(a_empty(as));
                    umma_commit(b_empty(bs));
                    if (++as == NA) { as = 0; aph ^= 1u; }
                    if (++bs == NB) { bs = 0; bph ^= 1u; }
                }
                umma_commit(tfull_bar(acc));
            }
            SP_ADD(w_i);
            SP_STORE(3, w_te); SP_STORE(4, w_af); SP_STORE(5, w_bf); SP_STORE(6, w_i);
        }
    } else if (warp < FIRST_DW_WARP) {
        // ================= epilogue (warps 2 .. FIRST_DW_WARP-1) =================
        // In the write-heavy launches (64 -> 128 @147, 128 -> 256 @74) this role's ~120 instructions per 64-column chunk take ~2 k clk:
        // its warps share the schedulers' FMA/ALU pipes with the depthwise warps, which have the higher warp ids and so the priority.
        const int q = warp & 3;
        const int half = (warp - 2) >> 2;
        const uint32_t ebuf = epi_base + (uint32_t)(warp - 2) * 8192u;
        uint32_t chunk_no = 0;
        int it = 0;
        SP_DECL(w_tf = 0, w_w = 0, w_bw = 0, w_ld = 0, w_st = 0);
        SP_T0();
        for (int tile = blockIdx.x; tile < g.num_tiles; tile += gridDim.x, ++it) {
            const int tx = tile % g.ntx;
            const int rr = tile / g.ntx;
            const int ty = rr % g.nty, b = rr / g.nty;
            const int acc = it & 1;
            const uint32_t t_row = tmem_base + acc * BN + ((uint32_t)(q * 32) << 16);
            SP_ADD(w_w);
            mbar_wait(tfull_bar(acc), (uint32_t)((it >> 1) & 1));
            SP_ADD(w_tf);
            tc_fence_after();
            // rows 32q..32q+31 of the tile = pixel rows (2q, 2q+1) x 16 columns; a warp whose rows lie below the image skips the
            // chunk entirely (no buffer write, no store) so that store groups and staging buffers stay in lock-step
            const bool do_store = ty * TH + 2 * q < g.H;
#pragma unroll
            for (int c = 0; c < BN / 64; ++c) {
                const int nc = c * 64;
                if (nc < g.N && do_store && (c % (EPI_WARPS / 4)) == half) {
                    uint32_t r[64];
                    tmem_ld32(t_row + c * 64, r);
                    tmem_ld32(t_row + c * 64 + 32, r + 32);
                    const uint32_t buf = ebuf + (uint32_t)(chunk_no++ & 1u) * 4096u;
                    SP_ADD(w_w);
                    if (lane == 0) bulk_wait_read<1>();
                    __syncwarp();
                    SP_ADD(w_bw);
                    tmem_ld_wait();
                    SP_ADD(w_ld);
#pragma unroll
                    for (int gidx = 0; gidx < 8; ++gidx) {
                        const uint4 o = epi8(r + gidx * 8, bias_base + (uint32_t)(c * 64 + gidx * 8) * 4u, g.relu_out, false, make_uint4(0, 0, 0, 0));
                        st_shared_v4(buf + (uint32_t)lane * 128u + (uint32_t)((gidx ^ (lane & 7)) << 4), o);
                    }
                    SP_ADD(w_w);
                    fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0) {
                        tma_store_4d(&tmOut, buf, nc, tx * TW, ty * TH + 2 * q, b);
                        bulk_commit();
                    }
                    SP_ADD(w_st);
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty_bar(acc));
        }
        if (lane == 0) bulk_wait_all<0>();
        SP_ADD(w_w);
        if (warp == 2 && lane == 0) { SP_STORE(7, w_tf); SP_STORE(8, w_w); SP_STORE(12, w_bw); SP_STORE(13, w_ld); SP_STORE(14, w_st); }
    } else {
        // ================= depthwise producers (warps FIRST_DW_WARP ..) =================
        const int dt = threadIdx.x - FIRST_DW_WARP * 32;      // 0..32*DW_WARPS-1
        const int cg = dt & 15;                   // 4-channel group of the 64-channel chunk
        const int cp = (dt >> 4) & 7;             // column pair
        const int y_lo = (dt >> 7) * DW_ROWS;     // strip of output rows
        const int y_hi = min(y_lo + DW_ROWS, TH);
        int xs = 0, as = 0;
        uint32_t xph = 0, aph = 0;
        SP_DECL(w_xf = 0, w_ae = 0, w_c = 0);
        SP_T0();
        for (int tile = blockIdx.x; tile < g.num_tiles; tile += gridDim.x) {
            for (int kc = 0; kc < g.kchunks; ++kc) {
                SP_ADD(w_c);
                mbar_wait(x_full(xs), xph);
                SP_ADD(w_xf);
                mbar_wait(a_empty(as), aph ^ 1u);
                SP_ADD(w_ae);
                const uint32_t xt = x_base + xs * XSTAGE;
                const uint32_t wb = xt + X_BYTES + (uint32_t)cg * 16u;
                float2 wl[9], wh[9];
#pragma unroll
                for (int k = 0; k < 9; ++k) {
                    const float4 w4 = ld_shared_f4(wb + (uint32_t)k * (CB * 4));
                    wl[k] = make_float2(w4.x, w4.y);
                    wh[k] = make_float2(w4.z, w4.w);
                }
                float2 aA[2][2], aB[2][2], aC[2][2];
                uint32_t rowaddr = xt + ((uint32_t)(y_lo * XW) + 2u * cp) * 128u + (uint32_t)cg * 8u;
                const uint32_t rowpitch = (uint32_t)XW * 128u;
                // A-tile address of (tile row y, column 2cp + j): row r = y*16 + 2cp + j, 16-byte chunk cg/2 XOR (r & 7), half cg & 1
                const uint32_t at = a_base + as * A_BYTES;
                int orow = y_lo;
                auto store_row = [&]() {
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        const uint32_t r = (uint32_t)(orow * TW + 2 * cp + j);
                        const uint32_t addr = at + r * 128u + ((((uint32_t)cg >> 1) ^ (r & 7u)) << 4) + ((uint32_t)cg & 1u) * 8u;
                        const uint32_t lo = pack_bf16(aA[j][0].x, aA[j][0].y), hi = pack_bf16(aA[j][1].x, aA[j][1].y);
                        asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(addr), "r"(lo), "r"(hi) : "memory");
                    }
                    ++orow;
                };
                auto rotate = [&]() {
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        aA[j][0] = aB[j][0]; aA[j][1] = aB[j][1];
                        aB[j][0] = aC[j][0]; aB[j][1] = aC[j][1];
                    }
                };
                dwt::scatter_row<RELU_IN, false, false, true>(rowaddr, wl, wh, aA, aB, aC);
                rotate();
                rowaddr += rowpitch;
                dwt::scatter_row<RELU_IN, false, true, true>(rowaddr, wl, wh, aA, aB, aC);
                rotate();
                rowaddr += rowpitch;
#pragma unroll
                for (int r = y_lo + 1; r < y_hi; ++r, rowaddr += rowpitch) {
                    dwt::scatter_row<RELU_IN, true, true, true>(rowaddr, wl, wh, aA, aB, aC);
                    store_row();
                    rotate();
                }
                dwt::scatter_row<RELU_IN, true, false, false>(rowaddr, wl, wh, aA, aB, aC);
                store_row();
                fence_proxy_async_smem();  // make the generic-proxy A-tile stores visible to the tensor core (async proxy)
                __syncwarp();
                if (lane == 0) {
                    mbar_arrive(a_full(as));
                    mbar_arrive(x_empty(xs));
                }
                if (++xs == NX) { xs = 0; xph ^= 1u; }
                if (++as == NA) { as = 0; aph ^= 1u; }
            }
        }
        SP_ADD(w_c);
        if (warp == FIRST_DW_WARP && lane == 0) { SP_STORE(9, w_xf); SP_STORE(10, w_ae); SP_STORE(11, w_c); }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// host ---------------------------------------------------------------------------------------------------
inline int configure(char *err) {
    cudaError_t e = cudaFuncSetAttribute(k_sepconv_tc<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_sepconv_tc<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES);
    if (e != cudaSuccess) {
        snprintf(err, 512, "cudaFuncSetAttribute(k_sepconv_tc): %s", cudaGetErrorString(e));
        return KDL_ECUDA;
    }
    return KDL_OK;
}

inline bool supported(int K, int N) { return K >= 64 && K <= 256 && (K % 64) == 0 && N <= 256 && (N % 8) == 0; }

// in [B,H,W,K] (pitch ld_in), dwpad [K/64][9][64] fp32, pointwise weights via tmB (box 64 x 256 over Wt[N][Kp]), out [B,H,W,N] (pitch ld_out)
inline int launch(const TmaEncoder &enc, const bf16 *in, const float *dwpad, const CUtensorMap *tmB, const float *bias, bf16 *out, int B,
                  int H, int W, int K, int N, int ld_in, int ld_out, int relu_in, int relu_out, cudaStream_t st, char *err, int prof_slot = 0) {
    Geom g;
    g.prof_slot = prof_slot;
    g.B = B; g.H = H; g.W = W; g.K = K; g.N = N;
    g.ntx = (W + TW - 1) / TW;
    g.nty = (H + TH - 1) / TH;
    g.num_tiles = B * g.ntx * g.nty;
    g.kchunks = (K + CB - 1) / CB;
    g.relu_out = relu_out;
    CUtensorMap tmX, tmOut;
    cuuint32_t estr[4] = {1, 1, 1, 1};
    {
        cuuint64_t dims[4] = {(cuuint64_t)K, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
        cuuint64_t strides[3] = {(cuuint64_t)ld_in * 2, (cuuint64_t)W * ld_in * 2, (cuuint64_t)H * W * ld_in * 2};
        cuuint32_t box[4] = {(cuuint32_t)CB, (cuuint32_t)XW, (cuuint32_t)XH, 1};
        if (enc.fn(&tmX, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<bf16 *>(in), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
            snprintf(err, 512, "cuTensorMapEncodeTiled failed for fused sepconv input [%d,%d,%d,%d]", B, H, W, K);
            return KDL_ECUDA;
        }
    }
    {
        cuuint64_t dims[4] = {(cuuint64_t)N, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
        cuuint64_t strides[3] = {(cuuint64_t)ld_out * 2, (cuuint64_t)W * ld_out * 2, (cuuint64_t)H * W * ld_out * 2};
        cuuint32_t box[4] = {64, (cuuint32_t)TW, 2, 1};
        if (enc.fn(&tmOut, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, out, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
            snprintf(err, 512, "cuTensorMapEncodeTiled failed for fused sepconv output [%d,%d,%d,%d]", B, H, W, N);
            return KDL_ECUDA;
        }
    }
    const int grid = g.num_tiles < num_sms() ? g.num_tiles : num_sms();
    if (relu_in) launch_k(k_sepconv_tc<true>, dim3(grid), dim3(NUM_THREADS_SEP), SMEM_BYTES, st, tmX, *tmB, tmOut, dwpad, bias, g);
    else launch_k(k_sepconv_tc<false>, dim3(grid), dim3(NUM_THREADS_SEP), SMEM_BYTES, st, tmX, *tmB, tmOut, dwpad, bias, g);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        snprintf(err, 512, "k_sepconv_tc launch (B=%d H=%d K=%d N=%d): %s", B, H, K, N, cudaGetErrorString(e));
        return KDL_ECUDA;
    }
    return KDL_OK;
}

}  // namespace sepc
}  // namespace kdl

// Fused SeparableConv2D for the middle flow (K = N = 728 at 19 x 19; any K, N <= 768 at W <= 20): depthwise 3x3 on CUDA cores
// feeding the pointwise tcgen05 GEMM through shared memory, so the depthwise output (142 MB per layer at batch 256, more than
// the L2) never touches HBM and the CUDA cores work while the tensor pipe does.
//
// Why it is shaped like this:
//   * GEMM rows are FLAT pixels (128 consecutive rows of the [B*H*W, C] activation matrix), not 2-D patches: 19 x 19 images do
//     not tile into 8 x 16 patches (47 % of the MMA rows would be padding), 92416 = 722 x 128 does.  A tile therefore covers
//     7-8 image rows, possibly the end of one image and the start of the next; the input halo is one 3-D TMA box over
//     (C, W, B*H) starting one row above and one column left (out-of-bounds columns / rows read as zero = SAME padding), and the
//     rows that belong to the neighbouring IMAGE are masked in the depthwise scatter instead of being zero in memory.
//   * One pass must produce all N outputs of a pixel (the depthwise result lives only in shared memory), but 128 x 768 fp32
//     accumulators are 768 TMEM columns and an SM has 512.  So a work item is (pixel tile, N half): 384 columns = two
//     M128 x N192 x K16 instructions per k-step, single-buffered in TMEM; the two halves of a tile go to neighbouring CTAs at
//     the same time (the second read of the input tile hits L2) and each recomputes the cheap depthwise part.
//   * 16 warps: warp 0 TMA producer (halo tile + 9x64 fp32 depthwise weights per 64-channel chunk, 2 stages; the matching
//     [384 x 64] pointwise weight block, 2 stages), warp 1 MMA issuer, warps 2-5 epilogue (tcgen05.ld -> +bias / ReLU /
//     +residual -> bf16 -> swizzled smem -> TMA store, the code path of gemm_tc.cuh), warps 6-15 depthwise (4 channels x 2
//     columns per thread, FFMA2 scatter of dw_tma.cuh, result rounded to bf16 straight into the K-major 128B-swizzled A tile).
// Arithmetic and its order are those of k_dw_tma followed by k_gemm_tc, so the results are identical to the unfused path.
#pragma once
#include "dw_tma.cuh"
#include "gemm_tc.cuh"

namespace kdl {
namespace sepm {

using namespace kdl::tc;

constexpr int BM = 128;                       // flat pixels per tile
constexpr int CB = 64;                        // channels per chunk (one 128-byte swizzle row of bf16)
constexpr int NHALF = 384;                    // output columns per work item
constexpr int NSUB = 192;                     // columns per MMA instruction (two per k-step)
constexpr int NX = 2, NA = 2, NB = 2;
constexpr int DW_WARPS = 10;
constexpr int DW_RS = 4;                      // output rows per depthwise strip (compile time: the row loop is fully unrolled)
constexpr int NUM_THREADS_MID = (6 + DW_WARPS) * 32;
constexpr uint32_t X_TILE_MAX = 27 * 1024;    // largest staged halo tile: rows x (W + 2) x 128 bytes
constexpr uint32_t DWW_BYTES = 9 * CB * 4;
constexpr uint32_t XSTAGE = X_TILE_MAX + DWW_BYTES;
constexpr uint32_t A_BYTES = BM * 128;
constexpr uint32_t B_BYTES = NHALF * 128;
constexpr uint32_t EPI_BYTES = 4 * 2 * 4096;
constexpr uint32_t BIAS_BYTES = 2 * NHALF * 4;   // both halves' bias, staged once
constexpr uint32_t SMEM_BYTES = NA * A_BYTES + NB * B_BYTES + EPI_BYTES + NX * XSTAGE + BIAS_BYTES + 256 + 1024;
static_assert(SMEM_BYTES <= 232448, "exceeds 227 KB");
static_assert(XSTAGE % 128 == 0, "TMA destinations must stay 128-byte aligned");

struct Geom {
    int M;            // B * H * W pixels = GEMM rows
    int H, W, K, N;
    int num_tiles;    // ceil(M / 128)
    int n_halves;     // ceil(N / 384)
    int items;        // num_tiles * n_halves
    int kchunks;      // ceil(K / 64)
    int BWp;          // staged tile width  = W + 2
    int BH;           // staged tile height = max rows a tile touches + 2
    int CP, NS, RS;   // column pairs, row strips per tile, rows per strip
    int relu_out;
    int ldo;          // row pitch (elements) of out / residual
    uint32_t x_bytes; // BH * BWp * 128
};

__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap *tm, uint32_t bar, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(dst),
        "l"(tm), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}

// One staged input row of a depthwise thread: 4 adjacent pixels x 4 channels, still packed bf16 (8 registers).  Loading the NEXT row's
// Raw before the current row is processed is what hides the shared-memory latency: only 10 depthwise warps run per SM here.
struct Raw {
    uint32_t v[8];
};
__device__ __forceinline__ Raw load_raw(uint32_t rowaddr) {
    Raw r;
#pragma unroll
    for (int j = 0; j < 4; ++j)
        asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(r.v[2 * j]), "=r"(r.v[2 * j + 1]) : "r"(rowaddr + 128u * j));
    return r;
}
template <bool RELU> __device__ __forceinline__ void unpack_raw(const Raw &r, dwt::Px (&in)[4]) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        uint32_t a = r.v[2 * j], b = r.v[2 * j + 1];
        if (RELU) {  // packed bf16x2 max against +0
            asm("max.bf16x2 %0, %0, %1;" : "+r"(a) : "r"(0u));
            asm("max.bf16x2 %0, %0, %1;" : "+r"(b) : "r"(0u));
        }
        in[j].lo = make_float2(__uint_as_float(a << 16), __uint_as_float(a & 0xffff0000u));
        in[j].hi = make_float2(__uint_as_float(b << 16), __uint_as_float(b & 0xffff0000u));
    }
}
// dw_tma.cuh's scatter_row on already unpacked pixels: bottom taps (ky = 2) complete the output row above (aA), middle taps go to
// the row itself (aB), top taps open the row below (aC; zero when the row below belongs to the next image).  Same operations in
// the same order per output as k_dw_tma.
template <bool DO_A, bool DO_B, bool DO_C>
__device__ __forceinline__ void scatter_px(const dwt::Px (&in)[4], const float2 (&wl)[9], const float2 (&wh)[9], float2 (&aA)[2][2],
                                           float2 (&aB)[2][2], float2 (&aC)[2][2]) {
#pragma unroll
    for (int j = 0; j < 2; ++j) {
        if (DO_A) {
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) {
                aA[j][0] = ffma2(in[j + kx].lo, wl[6 + kx], aA[j][0]);
                aA[j][1] = ffma2(in[j + kx].hi, wh[6 + kx], aA[j][1]);
            }
        }
        if (DO_B) {
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) {
                aB[j][0] = ffma2(in[j + kx].lo, wl[3 + kx], aB[j][0]);
                aB[j][1] = ffma2(in[j + kx].hi, wh[3 + kx], aB[j][1]);
            }
        }
        if (DO_C) {
            float2 cl = dwt::fmul2(in[j].lo, wl[0]), ch = dwt::fmul2(in[j].hi, wh[0]);
#pragma unroll
            for (int kx = 1; kx < 3; ++kx) {
                cl = ffma2(in[j + kx].lo, wl[kx], cl);
                ch = ffma2(in[j + kx].hi, wh[kx], ch);
            }
            aC[j][0] = cl;
            aC[j][1] = ch;
        } else {
            aC[j][0] = make_float2(0.f, 0.f);
            aC[j][1] = make_float2(0.f, 0.f);
        }
    }
}

// One input row of a strip, straight-line: which of the three tap rows it feeds is known at compile time (HAS_*), whether the image
// boundary cancels its upward (a) / downward (c) contribution is a warp-uniform run-time flag.
template <bool HAS_A, bool HAS_B, bool HAS_C>
__device__ __forceinline__ void scatter_rt(const dwt::Px (&in)[4], const float2 (&wl)[9], const float2 (&wh)[9], float2 (&aA)[2][2],
                                           float2 (&aB)[2][2], float2 (&aC)[2][2], bool a, bool c) {
    if (HAS_A && a) {
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) {
                aA[j][0] = ffma2(in[j + kx].lo, wl[6 + kx], aA[j][0]);
                aA[j][1] = ffma2(in[j + kx].hi, wh[6 + kx], aA[j][1]);
            }
    }
    if (HAS_B) {
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) {
                aB[j][0] = ffma2(in[j + kx].lo, wl[3 + kx], aB[j][0]);
                aB[j][1] = ffma2(in[j + kx].hi, wh[3 + kx], aB[j][1]);
            }
    }
    if (HAS_C) {
        if (c) {
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                float2 cl = dwt::fmul2(in[j].lo, wl[0]), ch = dwt::fmul2(in[j].hi, wh[0]);
#pragma unroll
                for (int kx = 1; kx < 3; ++kx) {
                    cl = ffma2(in[j + kx].lo, wl[kx], cl);
                    ch = ffma2(in[j + kx].hi, wh[kx], ch);
                }
                aC[j][0] = cl;
                aC[j][1] = ch;
            }
        } else {
#pragma unroll
            for (int j = 0; j < 2; ++j) aC[j][0] = aC[j][1] = make_float2(0.f, 0.f);
        }
    }
}

// -DKDL_MID_PROF (probe builds only): per-role wait / work cycle counters of the last launch, one row per CTA
#ifdef KDL_MID_PROF
__device__ unsigned long long g_mid_prof[160][32];
#define MP_T0() long long mp_t0 = clock64()
#define MP_ADD(var) do { long long mp_t1 = clock64(); var += mp_t1 - mp_t0; mp_t0 = mp_t1; } while (0)
#define MP_DECL(...) long long __VA_ARGS__
#define MP_STORE(idx, v) g_mid_prof[blockIdx.x][idx] = (unsigned long long)(v)
#else
#define MP_T0()
#define MP_ADD(var)
#define MP_DECL(...)
#define MP_STORE(idx, v)
#endif

template <bool RELU_IN>
__global__ void __launch_bounds__(NUM_THREADS_MID, 1)
k_sepconv_mid(const __grid_constant__ CUtensorMap tmX /*3-D (C, W, B*H), box 64 x (W+2) x BH*/,
              const __grid_constant__ CUtensorMap tmB /*Wt[N][Kp], box 64 x 192*/,
              const __grid_constant__ CUtensorMap tmOut /*[M, N], box 64 x 32*/,
              const float *__restrict__ dwpad, const float *__restrict__ bias, const bf16 *residual, Geom g) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t a_base = smem_base;
    const uint32_t b_base = a_base + NA * A_BYTES;
    const uint32_t epi_base = b_base + NB * B_BYTES;
    const uint32_t x_base = epi_base + EPI_BYTES;
    const uint32_t bias_base = x_base + NX * XSTAGE;
    const uint32_t bar_base = bias_base + BIAS_BYTES;
    // barriers: x_full[NX] x_empty[NX] a_full[NA] a_empty[NA] b_full[NB] b_empty[NB] tfull tempty
    auto x_full = [&](int s) { return bar_base + 8u * s; };
    auto x_empty = [&](int s) { return bar_base + 8u * (NX + s); };
    auto a_full = [&](int s) { return bar_base + 8u * (2 * NX + s); };
    auto a_empty = [&](int s) { return bar_base + 8u * (2 * NX + NA + s); };
    auto b_full = [&](int s) { return bar_base + 8u * (2 * NX + 2 * NA + s); };
    auto b_empty = [&](int s) { return bar_base + 8u * (2 * NX + 2 * NA + NB + s); };
    const uint32_t tfull_bar = bar_base + 8u * (2 * NX + 2 * NA + 2 * NB);
    const uint32_t tempty_bar = tfull_bar + 8u;
    const uint32_t tmem_slot = tempty_bar + 8u;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmX);
        prefetch_tmap(&tmB);
        prefetch_tmap(&tmOut);
        for (int s = 0; s < NX; ++s) { mbar_init(x_full(s), 1); mbar_init(x_empty(s), DW_WARPS); }
        for (int s = 0; s < NA; ++s) { mbar_init(a_full(s), DW_WARPS); mbar_init(a_empty(s), 1); }
        for (int s = 0; s < NB; ++s) { mbar_init(b_full(s), 1); mbar_init(b_empty(s), 1); }
        mbar_init(tfull_bar, 1);
        mbar_init(tempty_bar, 4);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
    pdl_launch_dependents();
    pdl_wait();
    if (warp >= 2 && warp < 6) {  // bias of every output column once (weights: not written by the previous kernel, but after pdl_wait anyway)
        const int et = threadIdx.x - 64;
        for (int j = et; j < 2 * NHALF; j += 128) {
            const float b = j < g.N ? __ldg(bias + j) : 0.f;
            asm volatile("st.shared.f32 [%0], %1;" ::"r"(bias_base + 4u * j), "f"(b) : "memory");
        }
        named_bar_sync(1, 128);
    }

    if (warp == 0) {
        // ================= TMA producer =================
        if (lane == 0) {
            int xs = 0, bs = 0;
            uint32_t xph = 0, bph = 0;
            MP_DECL(w_xe = 0, w_be = 0, w_iss = 0);
            MP_T0();
            for (int item = blockIdx.x; item < g.items; item += gridDim.x) {
                const int tile = item / g.n_halves, half = item - tile * g.n_halves;
                const int r_lo = (tile * BM) / g.W;        // first global image row (b*H + y) the tile touches
                const int n0 = half * NHALF;
                for (int kc = 0; kc < g.kchunks; ++kc) {
                    MP_ADD(w_iss);
                    mbar_wait(x_empty(xs), xph ^ 1u);
                    MP_ADD(w_xe);
                    mbar_expect_tx(x_full(xs), g.x_bytes + DWW_BYTES);
                    const uint32_t xd = x_base + xs * XSTAGE;
                    tma_load_3d(xd, &tmX, x_full(xs), kc * CB, -1, r_lo - 1);
                    dwt::bulk_load_1d(xd + g.x_bytes, dwpad + (size_t)kc * 9 * CB, DWW_BYTES, x_full(xs));
                    if (++xs == NX) { xs = 0; xph ^= 1u; }
                    MP_ADD(w_iss);
                    mbar_wait(b_empty(bs), bph ^ 1u);
                    MP_ADD(w_be);
                    mbar_expect_tx(b_full(bs), B_BYTES);
                    tma_load_2d(b_base + bs * B_BYTES, &tmB, b_full(bs), kc * CB, n0);
                    tma_load_2d(b_base + bs * B_BYTES + NSUB * 128, &tmB, b_full(bs), kc * CB, n0 + NSUB);
                    if (++bs == NB) { bs = 0; bph ^= 1u; }
                }
            }
            MP_ADD(w_iss);
            MP_STORE(0, w_xe); MP_STORE(1, w_be); MP_STORE(2, w_iss);
        }
    } else if (warp == 1) {
        // ================= MMA issuer =================
        if (lane == 0) {
            int as = 0, bs = 0, it = 0;
            uint32_t aph = 0, bph = 0;
            MP_DECL(w_te = 0, w_af = 0, w_bf = 0, w_iss = 0);
            MP_T0();
            for (int item = blockIdx.x; item < g.items; item += gridDim.x, ++it) {
                const int tile = item / g.n_halves, half = item - tile * g.n_halves;
                const int n_valid = min(NHALF, g.N - half * NHALF);
                const int n_first = min(NSUB, (n_valid + 15) & ~15);
                const int n_second = n_valid > NSUB ? ((n_valid - NSUB + 15) & ~15) : 0;
                const uint32_t idesc0 = make_idesc(BM, (uint32_t)n_first);
                const uint32_t idesc1 = make_idesc(BM, (uint32_t)(n_second ? n_second : 16));
                MP_ADD(w_iss);
                mbar_wait(tempty_bar, (uint32_t)(it & 1) ^ 1u);   // the epilogue has drained the (single) accumulator
                MP_ADD(w_te);
                tc_fence_after();
                for (int kc = 0; kc < g.kchunks; ++kc) {
                    MP_ADD(w_iss);
                    mbar_wait(a_full(as), aph);
                    MP_ADD(w_af);
                    mbar_wait(b_full(bs), bph);
                    MP_ADD(w_bf);
                    tc_fence_after();
                    const uint64_t adesc = make_smem_desc<128>(a_base + as * A_BYTES);
                    const uint64_t bdesc0 = make_smem_desc<128>(b_base + bs * B_BYTES);
                    const uint64_t bdesc1 = make_smem_desc<128>(b_base + bs * B_BYTES + NSUB * 128);
                    const int k_left = g.K - kc * CB;
                    const int ksteps = k_left >= CB ? CB / UMMA_K : (k_left + UMMA_K - 1) / UMMA_K;
                    for (int k = 0; k < ksteps; ++k) {
                        umma_bf16(tmem_base, adesc + 2u * k, bdesc0 + 2u * k, idesc0, (kc | k) != 0);
                        if (n_second) umma_bf16(tmem_base + NSUB, adesc + 2u * k, bdesc1 + 2u * k, idesc1, (kc | k) != 0);
                    }
                    umma_commit(a_empty(as));
                    umma_commit(b_empty(bs));
                    if (++as == NA) { as = 0; aph ^= 1u; }
                    if (++bs == NB) { bs = 0; bph ^= 1u; }
                }
                umma_commit(tfull_bar);
            }
            MP_ADD(w_iss);
            MP_STORE(3, w_te); MP_STORE(4, w_af); MP_STORE(5, w_bf); MP_STORE(6, w_iss);
        }
    } else if (warp < 6) {
        // ================= epilogue (warps 2..5): gemm_tc.cuh's standard path on a 128 x 384 tile =================
        const int q = warp & 3;
        const uint32_t ebuf = epi_base + (uint32_t)(warp - 2) * 8192u;
        const bool has_res = residual != nullptr;
        uint32_t chunk_no = 0;
        int it = 0;
        MP_DECL(w_tf = 0, w_epi = 0);
        MP_T0();
        for (int item = blockIdx.x; item < g.items; item += gridDim.x, ++it) {
            const int tile = item / g.n_halves, half = item - tile * g.n_halves;
            const int m0 = tile * BM, n0 = half * NHALF;
            const int row0 = m0 + q * 32;
            const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16);
            const int rl_row = row0 + (lane >> 3);
            const bf16 *res_lane = has_res ? residual + (size_t)rl_row * g.ldo + (lane & 7) * 8 : nullptr;
            uint4 rn[8];
            auto load_res = [&](int c) {
                const int n = n0 + c * 64 + (lane & 7) * 8;
#pragma unroll
                for (int gi = 0; gi < 8; ++gi)
                    rn[gi] = (rl_row + 4 * gi < g.M && n < g.N) ? *reinterpret_cast<const uint4 *>(res_lane + (size_t)(4 * gi) * g.ldo + n0 + c * 64)
                                                              : make_uint4(0, 0, 0, 0);
            };
            if (has_res) load_res(0);
            MP_ADD(w_epi);
            mbar_wait(tfull_bar, (uint32_t)(it & 1));
            MP_ADD(w_tf);
            tc_fence_after();
#pragma unroll
            for (int c = 0; c < NHALF / 64; ++c) {
                const int nc = n0 + c * 64;
                if (nc < g.N) {  // warp-uniform
                    uint32_t r[64];
                    tmem_ld32(t_row + c * 64, r);
                    tmem_ld32(t_row + c * 64 + 32, r + 32);
                    const uint32_t buf = ebuf + (uint32_t)(chunk_no++ & 1u) * 4096u;
                    if (lane == 0) bulk_wait_read<1>();   // the TMA store issued two chunks ago has finished reading this buffer
                    __syncwarp();
                    if (has_res) {   // transpose the coalesced residual chunk through the staging buffer (gemm_tc.cuh)
#pragma unroll
                        for (int gi = 0; gi < 8; ++gi) {
                            const uint32_t rr = (uint32_t)(4 * gi + (lane >> 3));
                            st_shared_v4(buf + rr * 128u + ((((uint32_t)lane & 7u) ^ (rr & 7u)) << 4), rn[gi]);
                        }
                        __syncwarp();
                        if (c + 1 < NHALF / 64 && nc + 64 < g.N) load_res(c + 1);
                    }
                    tmem_ld_wait();
#pragma unroll
                    for (int gi = 0; gi < 8; ++gi) {
                        // each lane reads back its own row's residual vector and overwrites it with the result (same address)
                        const uint32_t slot = buf + (uint32_t)lane * 128u + (uint32_t)((gi ^ (lane & 7)) << 4);
                        const uint4 rc = has_res ? ld_shared_v4(slot) : make_uint4(0, 0, 0, 0);
                        const uint4 o = epi8(r + gi * 8, bias_base + (uint32_t)(nc + gi * 8) * 4u, g.relu_out, has_res, rc);
                        st_shared_v4(slot, o);
                    }
                    fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0 && row0 < g.M) {
                        tma_store_2d(&tmOut, buf, nc, row0);
                        bulk_commit();
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty_bar);
        }
        if (lane == 0) bulk_wait_all<0>();
        MP_ADD(w_epi);
        if (warp == 2 && lane == 0) { MP_STORE(7, w_tf); MP_STORE(8, w_epi); }
    } else {
        // ================= depthwise producers (warps 6..15) =================
        const int dt = threadIdx.x - 6 * 32;      // 0 .. 32*DW_WARPS-1
        const int cg = dt & 15;                   // 4-channel group of the 64-channel chunk
        const int rest = dt >> 4;
        const int cp = rest % g.CP;               // column pair
        const int strip = rest / g.CP;
        const bool active = strip < g.NS && 2 * cp < g.W;
        const bool has_col1 = 2 * cp + 1 < g.W;
        int xs = 0, as = 0;
        uint32_t xph = 0, aph = 0;
        MP_DECL(w_xf = 0, w_ae = 0, w_cmp = 0);
        MP_T0();
        for (int item = blockIdx.x; item < g.items; item += gridDim.x) {
            const int tile = item / g.n_halves;
            const int g0 = tile * BM;
            const int r_lo = g0 / g.W;
            const int x_off = g0 - r_lo * g.W;                 // the tile's first pixel sits at column x_off of image row r_lo
            const int nr = (x_off + BM - 1) / g.W + 1;         // image rows the tile touches
            const int y_lo = strip * DW_RS;
            const int yimg_lo = (r_lo + y_lo) % g.H;           // position of the strip's first output row inside its image
            // input row i of the strip (i = 0 is the halo row above) has image position (yimg_lo - 1 + i) mod H: zero at i = ib
            const int ib0 = yimg_lo == 0 ? 1 : g.H - yimg_lo + 1;
            const int ib = ib0 <= DW_RS + 1 ? ib0 : -1;
            for (int kc = 0; kc < g.kchunks; ++kc) {
                MP_ADD(w_cmp);
                mbar_wait(x_full(xs), xph);
                MP_ADD(w_xf);
                mbar_wait(a_empty(as), aph ^ 1u);
                MP_ADD(w_ae);
                if (active && y_lo < nr) {
                    const uint32_t xt = x_base + xs * XSTAGE;
                    const uint32_t wb = xt + g.x_bytes + (uint32_t)cg * 16u;
                    float2 wl[9], wh[9];
#pragma unroll
                    for (int k = 0; k < 9; ++k) {
                        const float4 w4 = ld_shared_f4(wb + (uint32_t)k * (CB * 4));
                        wl[k] = make_float2(w4.x, w4.y);
                        wh[k] = make_float2(w4.z, w4.w);
                    }
                    float2 aA[2][2], aB[2][2], aC[2][2];
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        aA[j][0] = aA[j][1] = aB[j][0] = aB[j][1] = make_float2(0.f, 0.f);
                    }
                    // staged row s holds image row r_lo - 1 + s; staged column s holds image column s - 1
                    uint32_t rowaddr = xt + ((uint32_t)(y_lo * g.BWp) + 2u * cp) * 128u + (uint32_t)cg * 8u;
                    const uint32_t rowpitch = (uint32_t)g.BWp * 128u;
                    const uint32_t at = a_base + as * A_BYTES;
                    int p_row = y_lo * g.W + 2 * cp - x_off;   // A-tile row (flat pixel in the tile) of (output row, first column)
                    auto store_row = [&]() {                   // output row held in aA is complete
#pragma unroll
                        for (int j = 0; j < 2; ++j) {
                            const int p = p_row + j;
                            if ((j == 0 || has_col1) && p >= 0 && p < BM) {
                                const uint32_t r = (uint32_t)p;
                                const uint32_t addr = at + r * 128u + ((((uint32_t)cg >> 1) ^ (r & 7u)) << 4) + ((uint32_t)cg & 1u) * 8u;
                                const uint32_t lo = pack_bf16(aA[j][0].x, aA[j][0].y), hi = pack_bf16(aA[j][1].x, aA[j][1].y);
                                asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(addr), "r"(lo), "r"(hi) : "memory");
                            }
                        }
                        p_row += g.W;
                    };
                    auto rotate = [&]() {
#pragma unroll
                        for (int j = 0; j < 2; ++j) {
                            aA[j][0] = aB[j][0]; aA[j][1] = aB[j][1];
                            aB[j][0] = aC[j][0]; aB[j][1] = aC[j][1];
                        }
                    };
                    // A strip always computes DW_RS output rows from DW_RS + 2 staged rows, fully unrolled (rows past the tile's last
                    // pixel are computed from whatever is staged there and dropped by the p < 128 test in store_row).  ib = which of
                    // those input rows is the FIRST row of an image (or -1): that row must not feed the output row above it, and the row
                    // before it must not feed the output row below.  At most one boundary fits into DW_RS + 2 <= H rows.
                    dwt::Px in[4];
                    Raw cur = load_raw(rowaddr);
                    Raw nxt = load_raw(rowaddr + rowpitch);
                    rowaddr += 2u * rowpitch;
#pragma unroll
                    for (int i = 0; i < DW_RS + 2; ++i) {
                        unpack_raw<RELU_IN>(cur, in);
                        cur = nxt;
                        if (i + 2 < DW_RS + 2) { nxt = load_raw(rowaddr); rowaddr += rowpitch; }
                        const bool a = i != ib, c = i + 1 != ib;
                        if (i == 0) scatter_rt<false, false, true>(in, wl, wh, aA, aB, aC, a, c);
                        else if (i == 1) scatter_rt<false, true, true>(in, wl, wh, aA, aB, aC, a, c);
                        else if (i <= DW_RS) scatter_rt<true, true, true>(in, wl, wh, aA, aB, aC, a, c);
                        else scatter_rt<true, false, false>(in, wl, wh, aA, aB, aC, a, c);
                        if (i >= 2) store_row();
                        if (i <= DW_RS) rotate();
                    }
                }
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
        MP_ADD(w_cmp);
        if (lane == 0 && (warp == 6 || warp == 15)) {
            const int o = warp == 6 ? 9 : 12;
            MP_STORE(o, w_xf); MP_STORE(o + 1, w_ae); MP_STORE(o + 2, w_cmp);
        }
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
    cudaError_t e = cudaFuncSetAttribute(k_sepconv_mid<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_sepconv_mid<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES);
    if (e != cudaSuccess) {
        snprintf(err, 512, "cudaFuncSetAttribute(k_sepconv_mid): %s", cudaGetErrorString(e));
        return KDL_ECUDA;
    }
    return KDL_OK;
}

inline int tile_rows(int W) { return (W + BM - 2) / W + 1; }   // most image rows 128 consecutive pixels can touch

// K, N <= 768 (two N halves), rows narrow enough for one column-pair sweep (W <= 20) and a halo tile that fits its stage
inline bool supported(int H, int W, int K, int N) {
    if (K < 64 || K > 768 || (K % 8) || N < 8 || N > 2 * NHALF || (N % 8) || W < 3 || W > 20 || H < DW_RS + 2) return false;
    const int cp = (W + 1) / 2, ns = (DW_WARPS * 2) / cp;
    if (ns < 1 || ns * DW_RS < tile_rows(W)) return false;          // the strips of DW_RS rows must cover every row a tile touches
    const int rows = (ns * DW_RS > tile_rows(W) ? (tile_rows(W) + DW_RS - 1) / DW_RS * DW_RS : tile_rows(W)) + 2;
    return (uint32_t)(rows * (W + 2) * 128) <= X_TILE_MAX;
}

// in [B,H,W,K] (pitch ld_in), dwpad [K/64][9][64] fp32, pointwise weights via tmB192 (box 64 x 192 over Wt[N][Kp]),
// out / residual [B*H*W, N] (pitch ld_out); residual may alias out (the middle-flow skip tensor is updated in place) but `in` may not.
inline int launch(const TmaEncoder &enc, const bf16 *in, const float *dwpad, const CUtensorMap *tmB192, const float *bias,
                  const bf16 *residual, bf16 *out, int B, int H, int W, int K, int N, int ld_in, int ld_out, int relu_in, int relu_out,
                  cudaStream_t st, char *err) {
    Geom g;
    g.M = B * H * W; g.H = H; g.W = W; g.K = K; g.N = N;
    g.num_tiles = (g.M + BM - 1) / BM;
    g.n_halves = (N + NHALF - 1) / NHALF;
    g.items = g.num_tiles * g.n_halves;
    g.kchunks = (K + CB - 1) / CB;
    g.BWp = W + 2;
    g.CP = (W + 1) / 2;
    g.NS = (tile_rows(W) + DW_RS - 1) / DW_RS;     // strips of DW_RS output rows (supported() made sure the depthwise warps can hold them)
    g.RS = DW_RS;
    g.BH = g.NS * DW_RS + 2;                       // every strip reads DW_RS + 2 staged rows
    g.relu_out = relu_out;
    g.ldo = ld_out;
    g.x_bytes = (uint32_t)(g.BH * g.BWp * 128);
    CUtensorMap tmX, tmOut;
    {
        cuuint64_t dims[3] = {(cuuint64_t)K, (cuuint64_t)W, (cuuint64_t)B * H};
        cuuint64_t strides[2] = {(cuuint64_t)ld_in * 2, (cuuint64_t)W * ld_in * 2};
        cuuint32_t box[3] = {(cuuint32_t)CB, (cuuint32_t)g.BWp, (cuuint32_t)g.BH};
        cuuint32_t estr[3] = {1, 1, 1};
        if (enc.fn(&tmX, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<bf16 *>(in), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
            snprintf(err, 512, "cuTensorMapEncodeTiled failed for mid-flow sepconv input [%d,%d,%d,%d]", B, H, W, K);
            return KDL_ECUDA;
        }
    }
    if (!enc.encode_2d(&tmOut, out, (uint64_t)N, (uint64_t)g.M, (uint64_t)ld_out * 2, 64, 32, 128)) {
        snprintf(err, 512, "cuTensorMapEncodeTiled failed for mid-flow sepconv output [%d,%d]", g.M, N);
        return KDL_ECUDA;
    }
    const int grid = g.items < num_sms() ? g.items : num_sms();
    if (relu_in) launch_k(k_sepconv_mid<true>, dim3(grid), dim3(NUM_THREADS_MID), SMEM_BYTES, st, tmX, *tmB192, tmOut, dwpad, bias, residual, g);
    else launch_k(k_sepconv_mid<false>, dim3(grid), dim3(NUM_THREADS_MID), SMEM_BYTES, st, tmX, *tmB192, tmOut, dwpad, bias, residual, g);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        snprintf(err, 512, "k_sepconv_mid launch (B=%d H=%d K=%d N=%d): %s", B, H, K, N, cudaGetErrorString(e));
        return KDL_ECUDA;
    }
    return KDL_OK;
}

}  // namespace sepm
}  // namespace kdl

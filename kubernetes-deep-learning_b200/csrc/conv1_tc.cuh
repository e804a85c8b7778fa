// K2 on the tensor cores (bf16 path, uint8 input): block1_conv1, 3x3 stride 2 VALID, 3 -> 32, folded BatchNorm + ReLU, as an
// implicit GEMM  out[p, co] = relu( sum_k a[p, k] * w'[k, co] + b'[co] ),  p = flat output pixel, k = (ky, kx, c) in Keras HWIO order.
//
//   * The gateway's scaling x = v/127.5 - 1 is folded into the operands so that the A operand is the RAW uint8 value, which bf16
//     represents exactly: w' = bf16(w_folded / 127.5), b' = shift - sum_k w_folded (exact because the conv is
//     VALID: no padded position ever enters a window).  Only w' is rounded -- the same bf16 weight rounding every other layer has.
//   * A tile = 128 pixels x 32 (27 taps + 5 zeros) bf16, built by gather warps (three teams of four): one pixel per thread, three rows of nine
//     contiguous source bytes each, fetched as aligned 32-bit words + funnel shift, converted (I2F + pack) and written into the
//     K-major 128B-swizzled layout the MMA reads (rows are 128 bytes apart, only the first 64 bytes of a row are used).
//   * warp 1 issues two tcgen05.mma M128 x N32 x K16 per tile into a double-buffered 32-column accumulator; warps 2..5 drain it
//     (tcgen05.ld -> +b' -> ReLU -> bf16), stage their 32 rows x 64 bytes linearly and write them with one 2 KB bulk store (the
//     output rows of a tile are contiguous in memory).  The 32 x 64 weight tile is loaded once per CTA by TMA.
// The CUDA-core k_conv1 (kernels.cuh) stays for the fp32 path and for float inputs, which bf16 cannot hold exactly.
#pragma once
#include "gemm_tc.cuh"

namespace kdl {
namespace c1tc {

using namespace kdl::tc;

constexpr int BM = 128;
constexpr int NA = 6;                              // A-tile ring
constexpr int TEAMS = 3;                           // gather teams of 4 warps; team g builds the CTA's tiles g, g+3, g+6, ...
constexpr int GATHER_WARPS = 4 * TEAMS;
constexpr int EPI_SETS = 2;                        // epilogue warp sets (4 warps each); set h drains accumulator buffer h = every other tile.
                                                   // One set was the bottleneck: ~100 dependent instructions + three barrier round trips per tile
constexpr int FIRST_GATHER_WARP = 2 + 4 * EPI_SETS;
constexpr int NUM_THREADS_C1 = (FIRST_GATHER_WARP + GATHER_WARPS) * 32;
static_assert(NA % TEAMS == 0, "a ring slot always belongs to the same team");
constexpr uint32_t A_BYTES = BM * 128;             // 16 KB per stage
constexpr uint32_t W_BYTES = 32 * 128;             // [32 co][64 k] bf16, k >= 27 zero
constexpr int EB = 4;                              // output staging buffers per epilogue warp (bulk stores in flight: the kernel writes 5 bytes per byte read)
constexpr uint32_t EPI_BYTES = 4 * EPI_SETS * EB * 2048;   // per epilogue warp: EB x (32 rows x 64 B)
constexpr uint32_t SMEM_BYTES = NA * A_BYTES + W_BYTES + EPI_BYTES + 128 /*bias*/ + 256 /*barriers*/ + 1024 /*align*/;

__device__ __forceinline__ void bulk_store_1d(void *gdst, uint32_t ssrc, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(ssrc), "r"(bytes) : "memory");
}
__device__ __forceinline__ uint32_t pack_u8_bf16(uint32_t lo, uint32_t hi) {   // two integers 0..255 -> packed bf16x2 (exact)
    __nv_bfloat162 v = __floats2bfloat162_rn((float)lo, (float)hi);
    return *reinterpret_cast<uint32_t *>(&v);
}

__global__ void __launch_bounds__(NUM_THREADS_C1, 1)
k_conv1_tc(const uint8_t *__restrict__ in /*[B,S,S,3]*/, const __grid_constant__ CUtensorMap tmW, const float *__restrict__ bias,
           bf16 *__restrict__ out /*[B,Ho,Ho,32]*/, int B, int S, int Ho) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t a_base = smem_base;
    const uint32_t w_base = a_base + NA * A_BYTES;
    const uint32_t epi_base = w_base + W_BYTES;
    const uint32_t bias_base = epi_base + EPI_BYTES;
    const uint32_t bar_base = bias_base + 128;
    auto a_full = [&](int s) { return bar_base + 8u * s; };
    auto a_empty = [&](int s) { return bar_base + 8u * (NA + s); };
    auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * NA + a); };
    auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * NA + 2 + a); };
    const uint32_t w_bar = bar_base + 8u * (2 * NA + 4);
    const uint32_t tmem_slot = bar_base + 8u * (2 * NA + 5);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long M = (long long)B * Ho * Ho;
    const int num_tiles = (int)((M + BM - 1) / BM);

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmW);
        for (int s = 0; s < NA; ++s) { mbar_init(a_full(s), 4); mbar_init(a_empty(s), 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(tfull_bar(a), 1); mbar_init(tempty_bar(a), 4); }
        mbar_init(w_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, 64);
    if (warp == 2) asm volatile("st.shared.f32 [%0], %1;" ::"r"(bias_base + 4u * lane), "f"(__ldg(bias + lane)) : "memory");
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
    pdl_launch_dependents();
    pdl_wait();

    if (warp == 0) {
        if (lane == 0) {   // the weight tile, once
            mbar_expect_tx(w_bar, W_BYTES);
            tma_load_2d(w_base, &tmW, w_bar, 0, 0);
        }
    } else if (warp == 1) {
        // ================= MMA issuer =================
        if (lane == 0) {
            int as = 0, it = 0;
            uint32_t aph = 0;
            const uint32_t idesc = make_idesc(BM, 32);
            const uint64_t bdesc = make_smem_desc<128>(w_base);
            mbar_wait(w_bar, 0);
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
                const int acc = it & 1;
                mbar_wait(tempty_bar(acc), (uint32_t)((it >> 1) & 1) ^ 1u);
                mbar_wait(a_full(as), aph);
                tc_fence_after();
                const uint64_t adesc = make_smem_desc<128>(a_base + as * A_BYTES);
                umma_bf16(tmem_base + acc * 32, adesc, bdesc, idesc, 0);            // k = 0..15
                umma_bf16(tmem_base + acc * 32, adesc + 2u, bdesc + 2u, idesc, 1);  // k = 16..31 (27..31 are zero on both sides)
                umma_commit(a_empty(as));
                umma_commit(tfull_bar(acc));
                if (++as == NA) { as = 0; aph ^= 1u; }
            }
        }
    } else if (warp < FIRST_GATHER_WARP) {
        // ================= epilogue (warps 2..9; TMEM lane quarter = warp % 4; set = (warp - 2) / 4 takes the tiles of its accumulator) =================
        const int q = warp & 3;
        const int set = (warp - 2) >> 2;
        const uint32_t ebuf = epi_base + (uint32_t)(warp - 2) * (uint32_t)(EB * 2048);
        uint32_t n_store = 0;
        for (int it = set; blockIdx.x + it * (int)gridDim.x < num_tiles; it += EPI_SETS) {
            const int tile = blockIdx.x + it * (int)gridDim.x;
            const int acc = it & 1;
            mbar_wait(tfull_bar(acc), (uint32_t)((it >> 1) & 1));
            tc_fence_after();
            uint32_t r[32];
            tmem_ld32(tmem_base + acc * 32 + ((uint32_t)(q * 32) << 16), r);
            const uint32_t buf = ebuf + (n_store++ % (uint32_t)EB) * 2048u;
            if (lane == 0) bulk_wait_read<EB - 1>();   // the bulk store issued EB tiles ago has finished reading this buffer
            __syncwarp();
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty_bar(acc));   // accumulator is in registers: the MMA may reuse it
#pragma unroll
            for (int gidx = 0; gidx < 4; ++gidx) {
                const uint4 o = epi8(r + gidx * 8, bias_base + 32u * gidx, 1, false, make_uint4(0, 0, 0, 0));
                st_shared_v4(buf + (uint32_t)lane * 64u + 16u * gidx, o);
            }
            fence_proxy_async_smem();
            __syncwarp();
            const long long row0 = (long long)tile * BM + q * 32;
            if (lane == 0 && row0 < M) {
                const long long rows = M - row0 < 32 ? M - row0 : 32;
                bulk_store_1d(out + row0 * 32, buf, (uint32_t)rows * 64u);
                bulk_commit();
            }
        }
        if (lane == 0) bulk_wait_all<0>();
    } else {
        // ================= gather warps (10..21): three teams of 128 threads, one output pixel per thread =================
        // The source words of a team's NEXT tile are requested before the current tile is converted and stored, and the three teams
        // work on different tiles: about six tiles' loads are in flight per SM, which is what hides the ~1 us global-load latency
        // behind the ~0.3 us a tile may take at the HBM rate.
        const int team = (warp - FIRST_GATHER_WARP) >> 2;
        const int gt = threadIdx.x - FIRST_GATHER_WARP * 32 - team * 128;      // A-tile row
        const uint32_t sw = (uint32_t)gt & 7u;
        const long long hw = (long long)Ho * Ho;
        struct Words { uint32_t w[9]; unsigned sh[3]; bool ok; };
        auto fetch = [&](int tile) {
            Words f;
            const long long p = (long long)tile * BM + gt;
            f.ok = tile < num_tiles && p < M;
#pragma unroll
            for (int i = 0; i < 9; ++i) f.w[i] = 0u;
#pragma unroll
            for (int ky = 0; ky < 3; ++ky) f.sh[ky] = 0u;
            if (f.ok) {
                const int b = (int)(p / hw);
                const int rr = (int)(p - (long long)b * hw);
                const int oy = rr / Ho, ox = rr - oy * Ho;
                const uint8_t *base = in + (((long long)b * S + 2 * oy) * S + 2 * ox) * 3;
#pragma unroll
                for (int ky = 0; ky < 3; ++ky) {
                    const uintptr_t a = reinterpret_cast<uintptr_t>(base + (long long)ky * S * 3);
                    const uint32_t *wp = reinterpret_cast<const uint32_t *>(a & ~uintptr_t(3));
                    f.sh[ky] = (unsigned)(a & 3) * 8u;
                    f.w[3 * ky] = __ldg(wp); f.w[3 * ky + 1] = __ldg(wp + 1); f.w[3 * ky + 2] = __ldg(wp + 2);   // bytes k..k+8 lie in these words
                }
            }
            return f;
        };
        int it = team;
        Words cur = fetch(blockIdx.x + it * (int)gridDim.x);
        for (; blockIdx.x + it * (int)gridDim.x < num_tiles; it += TEAMS) {
            const Words nxt = fetch(blockIdx.x + (it + TEAMS) * (int)gridDim.x);
            uint32_t e[16];                        // 32 bf16: k = 2i, 2i+1
#pragma unroll
            for (int i = 0; i < 16; ++i) e[i] = 0u;
            if (cur.ok) {
                uint32_t v[27];
#pragma unroll
                for (int ky = 0; ky < 3; ++ky) {
                    const uint32_t w0 = cur.w[3 * ky], w1 = cur.w[3 * ky + 1], w2 = cur.w[3 * ky + 2];
                    const unsigned sh = cur.sh[ky];
                    const uint32_t b0 = __funnelshift_r(w0, w1, sh), b1 = __funnelshift_r(w1, w2, sh), b2 = w2 >> sh;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        v[ky * 9 + i] = (b0 >> (8 * i)) & 0xffu;
                        v[ky * 9 + 4 + i] = (b1 >> (8 * i)) & 0xffu;
                    }
                    v[ky * 9 + 8] = b2 & 0xffu;
                }
#pragma unroll
                for (int i = 0; i < 13; ++i) e[i] = pack_u8_bf16(v[2 * i], v[2 * i + 1]);
                e[13] = pack_u8_bf16(v[26], 0u);
            }
            const int as = it % NA;
            const uint32_t aph = (uint32_t)((it / NA) & 1);
            mbar_wait(a_empty(as), aph ^ 1u);
            const uint32_t row = a_base + as * A_BYTES + (uint32_t)gt * 128u;
#pragma unroll
            for (int c = 0; c < 4; ++c)   // 16-byte chunk c of the row sits at chunk position c ^ (row & 7) (128-byte swizzle)
                st_shared_v4(row + ((((uint32_t)c) ^ sw) << 4), make_uint4(e[4 * c], e[4 * c + 1], e[4 * c + 2], e[4 * c + 3]));
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(a_full(as));
            cur = nxt;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 64);
    }
}

inline int configure(char *err) {
    cudaError_t e = cudaFuncSetAttribute(k_conv1_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES);
    if (e != cudaSuccess) {
        snprintf(err, 512, "cudaFuncSetAttribute(k_conv1_tc): %s", cudaGetErrorString(e));
        return KDL_ECUDA;
    }
    return KDL_OK;
}

// tmW: box 64 x 32 (128-byte swizzle) over the [32][64] bf16 weight tile
inline int launch(const uint8_t *in, const CUtensorMap *tmW, const float *bias, bf16 *out, int B, int S, int Ho, cudaStream_t st, char *err) {
    const long long M = (long long)B * Ho * Ho;
    const int tiles = (int)((M + BM - 1) / BM);
    const int grid = tiles < num_sms() ? tiles : num_sms();
    launch_k(k_conv1_tc, dim3(grid), dim3(NUM_THREADS_C1), SMEM_BYTES, st, in, *tmW, bias, out, B, S, Ho);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        snprintf(err, 512, "k_conv1_tc launch (B=%d S=%d): %s", B, S, cudaGetErrorString(e));
        return KDL_ECUDA;
    }
    return KDL_OK;
}

}  // namespace c1tc
}  // namespace kdl

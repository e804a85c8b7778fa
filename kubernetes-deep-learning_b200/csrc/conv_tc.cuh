// K3: block1_conv2 (3x3, stride 1, VALID, 32 -> 64) as an implicit GEMM on tcgen05 with the input halo staged ONCE per tile.
//
// GEMM row m = input-grid position (b, y, x); output (b, y, x) needs input rows m + ky*Win + kx for the 9 taps.  Instead of
// fetching 9 shifted [128 x 32ch] boxes per tile (9x the L2 traffic in 64-byte requests), one tile stages the contiguous row
// range [m0, m0 + 128 + 2*Win + 2) of the [B*Hin*Win, 32] activation matrix (two TMA boxes, 64-byte swizzle) and the MMA reads
// tap (ky,kx) straight out of that buffer through a shared-memory descriptor whose start address is advanced by
// (ky*Win + kx) rows.  (The swizzle XOR is a function of the absolute shared-memory address, for the TMA write and for the UMMA
// read alike, so a row-shifted descriptor sees exactly the rows the TMA wrote.)  All 9 weight taps stay resident in shared memory.
#pragma once
#include "gemm_tc.cuh"

namespace kdl {
namespace convtc {

using namespace kdl::tc;

constexpr int CIN = 32, COUT = 64;
constexpr int ROWB = CIN * 2;                       // 64 bytes per activation row
constexpr int STAGES = 4;
constexpr uint32_t MAX_STAGE_BYTES = 2 * 256 * ROWB;  // two TMA boxes of <= 256 rows
constexpr uint32_t W_BYTES = 9 * COUT * ROWB;       // 36 KB
constexpr uint32_t BIAS_BYTES = COUT * 4;
constexpr int EPI_SETS = 2;                         // epilogue warp sets of 4; set h drains accumulator buffer h = every other tile
constexpr int THREADS_C2 = (2 + 4 * EPI_SETS) * 32;
constexpr uint32_t EPI_BYTES = 4 * EPI_SETS * 4096;   // per epilogue warp: 32 rows x 128 B, linear (bulk-store source)
constexpr uint32_t SMEM_BYTES = STAGES * MAX_STAGE_BYTES + W_BYTES + BIAS_BYTES + EPI_BYTES + 256 + 1024;
static_assert(SMEM_BYTES <= 232448, "exceeds 227 KB");

__global__ void __launch_bounds__(THREADS_C2, 1)
k_conv3x3_tc(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const float *__restrict__ bias,
             bf16 *__restrict__ out, bf16 *__restrict__ sub /*nullable: compact copy of the even pixels*/, int M, int relu, ConvGeom cg,
             int box_rows) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t stage_bytes = 2u * (uint32_t)box_rows * ROWB;  // multiple of 128
    const uint32_t wres_base = smem_base + STAGES * MAX_STAGE_BYTES;
    const uint32_t bias_base = wres_base + W_BYTES;
    const uint32_t epi_base = bias_base + BIAS_BYTES;
    const uint32_t bar_base = epi_base + EPI_BYTES;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
    auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + a); };
    auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + 2 + a); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES + 4);
    const uint32_t wres_bar = bar_base + 8u * (2 * STAGES + 5);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int num_tiles = (M + BLOCK_M - 1) / BLOCK_M;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmA);
        prefetch_tmap(&tmB);
        for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(tfull_bar(a), 1); mbar_init(tempty_bar(a), 4); }
        mbar_init(wres_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, 128);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
    pdl_launch_dependents();
    pdl_wait();

    if (warp == 0) {
        if (lane == 0) {
            mbar_expect_tx(wres_bar, W_BYTES);
            for (int t = 0; t < 9; ++t) tma_load_2d(wres_base + t * (COUT * ROWB), &tmB, wres_bar, t * CIN, 0);
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                const int m0 = tile * BLOCK_M;
                mbar_wait(empty_bar(stage), phase ^ 1u);
                mbar_expect_tx(full_bar(stage), stage_bytes);
                const uint32_t sa = smem_base + stage * MAX_STAGE_BYTES;
                tma_load_2d(sa, &tmA, full_bar(stage), 0, m0);
                tma_load_2d(sa + (uint32_t)box_rows * ROWB, &tmA, full_bar(stage), 0, m0 + box_rows);
                if (++stage == STAGES) { stage = 0; phase ^= 1u; }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            const uint32_t idesc = make_idesc(BLOCK_M, COUT);
            mbar_wait(wres_bar, 0);
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
                const int acc = it & 1;
                const uint32_t acc_phase = (it >> 1) & 1;
                mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
                mbar_wait(full_bar(stage), phase);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * COUT;
                const uint32_t sa = smem_base + stage * MAX_STAGE_BYTES;
#pragma unroll
                for (int t = 0; t < 9; ++t) {
                    const uint32_t a_addr = sa + (uint32_t)((t / 3) * cg.Win + (t % 3)) * ROWB;  // tap = row shift inside the halo
                    const uint64_t adesc = make_smem_desc<64>(a_addr);
                    const uint64_t bdesc = make_smem_desc<64>(wres_base + t * (COUT * ROWB));
                    umma_bf16(d_tmem, adesc, bdesc, idesc, t != 0);
                    umma_bf16(d_tmem, adesc + 2u, bdesc + 2u, idesc, 1);
                }
                umma_commit(empty_bar(stage));
                umma_commit(tfull_bar(acc));
                if (++stage == STAGES) { stage = 0; phase ^= 1u; }
            }
        }
    } else {
        const int q = warp & 3;
        const int set = (warp - 2) >> 2;
        const int et = threadIdx.x - 64;
        if (et < COUT) asm volatile("st.shared.f32 [%0], %1;" ::"r"(bias_base + 4u * et), "f"(__ldg(bias + et)) : "memory");
        named_bar_sync(1, 128 * EPI_SETS);
        const int hw = cg.Hin * cg.Win;
        const uint32_t ebuf = epi_base + (uint32_t)(warp - 2) * 4096u;
        for (int it = set; blockIdx.x + it * (int)gridDim.x < num_tiles; it += EPI_SETS) {
            const int tile = blockIdx.x + it * (int)gridDim.x;
            const int acc = it & 1;
            const uint32_t acc_phase = (it >> 1) & 1;
            const int row = tile * BLOCK_M + q * 32 + lane;
            const int b = row / hw;
            const int rem = row - b * hw;
            const int y = rem / cg.Win, x = rem - y * cg.Win;
            const uint32_t t_row = tmem_base + acc * COUT + ((uint32_t)(q * 32) << 16);
            mbar_wait(tfull_bar(acc), acc_phase);
            tc_fence_after();
            uint32_t r[64];
            tmem_ld32(t_row, r);
            tmem_ld32(t_row + 32, r + 32);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty_bar(acc));  // accumulator is in registers: release it before the stores
            // Output rows go through shared memory and leave as bulk stores: a lane owns one pixel (128 contiguous bytes), so direct
            // stores touched 32 cache lines per instruction with 16 bytes each (ncu round 1: L2 75 % busy).  The warp's 32 grid positions
            // are consecutive, and so are their outputs except for the two skipped columns at the end of an image row: at most two
            // contiguous runs [0, n1) and [s2, s2 + n2) of lanes, each one bulk store.
            const bool valid = row < M && y < cg.Hout && x < cg.Wout;
            const uint32_t buf = ebuf;
            if (lane == 0) bulk_wait_read<0>();   // this warp's previous store (two tiles ago: the other set took the tile between) has read the buffer
            __syncwarp();
            bf16 *os = (sub && valid && !(y & 1) && !(x & 1))
                           ? sub + ((size_t)(b * ((cg.Hout + 1) >> 1) + (y >> 1)) * ((cg.Wout + 1) >> 1) + (x >> 1)) * COUT
                           : nullptr;
#pragma unroll
            for (int g = 0; g < 8; ++g) {
                const uint4 v = epi8(r + g * 8, bias_base + 32u * g, relu, false, make_uint4(0, 0, 0, 0));
                st_shared_v4(buf + (uint32_t)lane * 128u + 16u * g, v);
                // the next block's Conv2D(1x1, strides 2) reads only the even pixels: emit them compactly here instead of a separate pass
                if (os) *reinterpret_cast<uint4 *>(os + g * 8) = v;
            }
            fence_proxy_async_smem();
            const unsigned vmask = __ballot_sync(0xffffffffu, valid);
            if (lane == 0 && vmask) {
                // runs of set bits in vmask (valid lanes); the output address of a run comes from its first lane's coordinates
                unsigned m = vmask;
                while (m) {
                    const int s0 = __ffs(m) - 1;
                    const unsigned from = m >> s0;
                    const int n = (~from) ? __ffs(~from) - 1 : 32 - s0;
                    const int rr0 = tile * BLOCK_M + q * 32 + s0;
                    const int b0 = rr0 / hw;
                    const int rem0 = rr0 - b0 * hw;
                    const int y0 = rem0 / cg.Win, x0 = rem0 - y0 * cg.Win;
                    bf16 *o = out + ((size_t)(b0 * cg.Hout + y0) * cg.Wout + x0) * COUT;
                    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(o), "r"(buf + (uint32_t)s0 * 128u),
                                 "r"((uint32_t)n * 128u)
                                 : "memory");
                    m = (s0 + n >= 32) ? 0u : (m & ~((1u << (s0 + n)) - 1u));
                }
                bulk_commit();
            }
        }
        if (lane == 0) bulk_wait_all<0>();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 128);
    }
}

inline int configure(char *err) {
    cudaError_t e = cudaFuncSetAttribute(k_conv3x3_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES);
    if (e != cudaSuccess) {
        snprintf(err, 512, "cudaFuncSetAttribute(k_conv3x3_tc): %s", cudaGetErrorString(e));
        return KDL_ECUDA;
    }
    return KDL_OK;
}

inline bool geometry_supported(int Win) { return 128 + 2 * Win + 2 <= 512; }

// in [B,Hin,Win,32] bf16; weights via tmB (box 32 x 64 over Wt[64][288(+pad)], 64B swizzle); out [B,Hin-2,Win-2,64]
inline int launch(const TmaEncoder &enc, const bf16 *in, const CUtensorMap *tmB, const float *bias, bf16 *out, bf16 *sub, int B, int Hin,
                  int Win, int relu, cudaStream_t st, char *err) {
    const int M = B * Hin * Win;
    const int halo_rows = BLOCK_M + 2 * Win + 2;
    const int box_rows = ((halo_rows + 1) / 2 + 7) & ~7;  // multiple of 8 rows: both boxes start on a 512-byte swizzle-pattern boundary
    CUtensorMap tmA;
    if (!enc.encode_2d(&tmA, in, CIN, (uint64_t)M, ROWB, CIN, (uint32_t)box_rows, 64)) {
        snprintf(err, 512, "cuTensorMapEncodeTiled failed for conv input [%d,%d] box_rows %d", M, CIN, box_rows);
        return KDL_ECUDA;
    }
    const int tiles = (M + BLOCK_M - 1) / BLOCK_M;
    const int grid = tiles < num_sms() ? tiles : num_sms();
    launch_k(k_conv3x3_tc, dim3(grid), dim3(THREADS_C2), SMEM_BYTES, st, tmA, *tmB, bias, out, sub, M, relu,
             ConvGeom{Hin, Win, Hin - 2, Win - 2}, box_rows);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        snprintf(err, 512, "k_conv3x3_tc launch (B=%d H=%d): %s", B, Hin, cudaGetErrorString(e));
        return KDL_ECUDA;
    }
    return KDL_OK;
}

}  // namespace convtc
}  // namespace kdl

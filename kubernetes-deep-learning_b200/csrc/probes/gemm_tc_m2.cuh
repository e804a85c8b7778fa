// K5 for the tensor-bound layers (K >= 512): 256 x 256 output tile per CTA.
//
// The single-CTA kernel of gemm_tc.cuh is limited by the L2 -> shared-memory operand feed (48 KB per 128x256x64 block, ~47 B/clk/SM
// arrive, tensor pipe ~56-66 % active).  Here one CTA works on TWO stacked 128-row M tiles that share every weight block: a stage holds
// A0 + A1 + B = 64 KB for twice the MMA work, i.e. 32 KB per 128x256x64 block (-33 % operand traffic) without any cross-CTA
// synchronisation.  The price: both 256-column accumulators fill all 512 TMEM columns, so the epilogue of a tile pair is not
// overlapped with the next pair's MMAs (the TMA ring keeps prefetching meanwhile); eight epilogue warps keep that phase short.
#pragma once
#include "gemm_tc.cuh"

namespace kdl {
namespace tcm2 {

using namespace kdl::tc;

constexpr int BN = 256, BK = 64, NST = 3;
constexpr int THREADS = 10 * 32;                     // TMA, MMA, 8 epilogue warps
constexpr uint32_t A_BYTES = 128 * 128;              // one 128-row half
constexpr uint32_t B_BYTES = BN * 128;
constexpr uint32_t STAGE_BYTES = 2 * A_BYTES + B_BYTES;   // 64 KB
constexpr uint32_t EPI_BYTES = 8 * 4096;             // one 32-row x 64-col staging buffer per epilogue warp
constexpr uint32_t BIAS_BYTES = BN * 4;
constexpr uint32_t SMEM_BYTES = NST * STAGE_BYTES + EPI_BYTES + BIAS_BYTES + 256 + 1024;
static_assert(SMEM_BYTES <= 232448, "exceeds 227 KB");

__global__ void __launch_bounds__(THREADS, 1)
k_gemm_tc_m256(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmOut,
               const float *__restrict__ bias, const bf16 *residual, int M, int N, int K, int relu, int ldo) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t epi_base = smem_base + NST * STAGE_BYTES;
    const uint32_t bias_base = epi_base + EPI_BYTES;
    const uint32_t bar_base = bias_base + BIAS_BYTES;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (NST + s); };
    const uint32_t tfull_bar = bar_base + 8u * (2 * NST);
    const uint32_t tempty_bar = bar_base + 8u * (2 * NST + 1);
    const uint32_t tmem_slot = bar_base + 8u * (2 * NST + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m_pairs = (M + 255) / 256;
    const int n_tiles = (N + BN - 1) / BN;
    const int num_tiles = m_pairs * n_tiles;
    const int k_blocks = (K + BK - 1) / BK;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmA);
        prefetch_tmap(&tmB);
        prefetch_tmap(&tmOut);
        for (int s = 0; s < NST; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        mbar_init(tfull_bar, 1);
        mbar_init(tempty_bar, 8);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

    if (warp == 0) {
        // ================= TMA producer =================
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                const int m0 = (tile / n_tiles) * 256, n0 = (tile % n_tiles) * BN;
                for (int kb = 0; kb < k_blocks; ++kb) {
                    mbar_wait(empty_bar(stage), phase ^ 1u);
                    mbar_expect_tx(full_bar(stage), STAGE_BYTES);
                    const uint32_t sa = smem_base + stage * STAGE_BYTES;
                    tma_load_2d(sa, &tmA, full_bar(stage), kb * BK, m0);
                    tma_load_2d(sa + A_BYTES, &tmA, full_bar(stage), kb * BK, m0 + 128);
                    tma_load_2d(sa + 2 * A_BYTES, &tmB, full_bar(stage), kb * BK, n0);
                    if (++stage == NST) { stage = 0; phase ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer =================
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
                const int n0 = (tile % n_tiles) * BN;
                const int n_valid = min(BN, N - n0);
                const uint32_t idesc = make_idesc(BLOCK_M, (uint32_t)((n_valid + 15) & ~15));
                mbar_wait(tempty_bar, (uint32_t)(it & 1) ^ 1u);  // all eight epilogue warps have drained both accumulators
                tc_fence_after();
                for (int kb = 0; kb < k_blocks; ++kb) {
                    mbar_wait(full_bar(stage), phase);
                    tc_fence_after();
                    const uint32_t sa = smem_base + stage * STAGE_BYTES;
                    const uint64_t a0 = make_smem_desc<128>(sa), a1 = make_smem_desc<128>(sa + A_BYTES);
                    const uint64_t bd = make_smem_desc<128>(sa + 2 * A_BYTES);
                    const int k_left = K - kb * BK;
                    const int ksteps = k_left >= BK ? BK / UMMA_K : (k_left + UMMA_K - 1) / UMMA_K;
                    for (int k = 0; k < ksteps; ++k) {
                        umma_bf16(tmem_base, a0 + 2u * k, bd + 2u * k, idesc, (kb | k) != 0);        // rows   0..127 -> columns   0..255
                        umma_bf16(tmem_base + BN, a1 + 2u * k, bd + 2u * k, idesc, (kb | k) != 0);   // rows 128..255 -> columns 256..511
                    }
                    umma_commit(empty_bar(stage));
                    if (++stage == NST) { stage = 0; phase ^= 1u; }
                }
                umma_commit(tfull_bar);
            }
        }
    } else {
        // ================= epilogue: warps 2..9; accumulator = (warp-2)/4, TMEM lane quarter = warp % 4 =================
        const int q = warp & 3;
        const int half = (warp - 2) >> 2;
        const int et = threadIdx.x - 64;  // 0..255
        const uint32_t buf = epi_base + (uint32_t)(warp - 2) * 4096u;
        const bool has_res = residual != nullptr;
        int staged_n0 = -1;
        int it = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
            const int m0 = (tile / n_tiles) * 256 + half * 128, n0 = (tile % n_tiles) * BN;
            if (n0 != staged_n0) {
                named_bar_sync(1, 256);
                float b = (n0 + et < N) ? __ldg(bias + n0 + et) : 0.f;
                asm volatile("st.shared.f32 [%0], %1;" ::"r"(bias_base + 4u * et), "f"(b) : "memory");
                named_bar_sync(1, 256);
                staged_n0 = n0;
            }
            const int row0 = m0 + q * 32;
            const int row = row0 + lane;
            const bool row_ok = row < M;
            const uint32_t t_row = tmem_base + half * BN + ((uint32_t)(q * 32) << 16);
            const bf16 *res_row = residual + (size_t)row * ldo;
            uint4 rn[8];
            auto load_res = [&](int c) {
#pragma unroll
                for (int g = 0; g < 8; ++g) {
                    const int n = n0 + c * 64 + g * 8;
                    rn[g] = (has_res && row_ok && n < N) ? *reinterpret_cast<const uint4 *>(res_row + n) : make_uint4(0, 0, 0, 0);
                }
            };
            if (has_res) load_res(0);
            mbar_wait(tfull_bar, (uint32_t)(it & 1));
            tc_fence_after();
#pragma unroll
            for (int c = 0; c < BN / 64; ++c) {
                const int nc = n0 + c * 64;
                if (nc < N) {
                    uint32_t r[64];
                    tmem_ld32(t_row + c * 64, r);
                    tmem_ld32(t_row + c * 64 + 32, r + 32);
                    uint4 rc[8];
#pragma unroll
                    for (int g = 0; g < 8; ++g) rc[g] = rn[g];
                    if (has_res && c + 1 < BN / 64 && nc + 64 < N) load_res(c + 1);
                    if (lane == 0) bulk_wait_read<0>();  // single staging buffer: the previous chunk's store must have read it
                    __syncwarp();
                    tmem_ld_wait();
#pragma unroll
                    for (int g = 0; g < 8; ++g) {
                        const uint4 o = epi8(r + g * 8, bias_base + (uint32_t)(c * 64 + g * 8) * 4u, relu, has_res, rc[g]);
                        st_shared_v4(buf + (uint32_t)lane * 128u + (uint32_t)((g ^ (lane & 7)) << 4), o);
                    }
                    fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0 && row0 < M) {
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
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

inline int configure(char *err) {
    cudaError_t e = cudaFuncSetAttribute(k_gemm_tc_m256, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES);
    if (e != cudaSuccess) {
        snprintf(err, 512, "cudaFuncSetAttribute(k_gemm_tc_m256): %s", cudaGetErrorString(e));
        return KDL_ECUDA;
    }
    return KDL_OK;
}

// same contract as tc::launch_gemm (weights via the box-64x256 map tmB)
inline int launch(const TmaEncoder &enc, const bf16 *A, const CUtensorMap *tmB, const float *bias, const bf16 *residual, bf16 *out, int M,
                  int N, int K, int relu, cudaStream_t st, char *err, int lda = 0, int ldo = 0) {
    if (lda == 0) lda = K;
    if (ldo == 0) ldo = N;
    CUtensorMap tmA, tmOut;
    if (!enc.encode_2d(&tmA, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda * 2, BK, 128, 128) ||
        !enc.encode_2d(&tmOut, out, (uint64_t)N, (uint64_t)M, (uint64_t)ldo * 2, 64, 32, 128)) {
        snprintf(err, 512, "cuTensorMapEncodeTiled failed for A [%d,%d] / out [%d,%d]", M, K, M, N);
        return KDL_ECUDA;
    }
    const int tiles = ((M + 255) / 256) * ((N + BN - 1) / BN);
    const int grid = tiles < num_sms() ? tiles : num_sms();
    k_gemm_tc_m256<<<grid, THREADS, SMEM_BYTES, st>>>(tmA, *tmB, tmOut, bias, residual, M, N, K, relu, ldo);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        snprintf(err, 512, "k_gemm_tc_m256 launch (M=%d N=%d K=%d): %s", M, N, K, cudaGetErrorString(e));
        return KDL_ECUDA;
    }
    return KDL_OK;
}

}  // namespace tcm2
}  // namespace kdl

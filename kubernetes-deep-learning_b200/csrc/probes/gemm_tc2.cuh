// K5, CTA-pair version: the pointwise GEMM with tcgen05.mma.cta_group::2.
//
// Two CTAs of a cluster (one TPC) compute one 256(M) x 256(N) tile: each CTA stages its own 128 rows of A and HALF of the weight
// tile (128 of the 256 N rows); the leader CTA's single MMA thread issues M=256 instructions that read A and B from both CTAs'
// shared memory and write each CTA's 128 accumulator rows into that CTA's TMEM.  Compared with the single-CTA kernel this cuts the
// L2->SM operand traffic per flop from 48 KB to 32 KB per (128 x 256 x 64) block and lets the same 192 KB of shared memory hold a
// 6-stage ring, which is what the K >= 728 layers are limited by (see profiles/).
//
// Synchronisation (s = stage, a = accumulator buffer):
//   full[s]   lives in the LEADER: both CTAs' TMA loads complete_tx on it (cp.async.bulk.tensor ... .cta_group::2)
//   empty[s]  one per CTA: the leader's tcgen05.commit multicasts the arrival to both
//   tfull[a]  one per CTA: same multicast commit; each CTA's epilogue drains its own TMEM half
//   tempty[a] lives in the LEADER, count 8: the four epilogue warps of BOTH CTAs arrive on it (remote arrive from the peer)
#pragma once
#include "gemm_tc.cuh"

namespace kdl {
namespace tc2 {

using namespace kdl::tc;

constexpr int BN2 = 256;
constexpr int BK2 = 64;
constexpr int STAGES2 = 6;
constexpr uint32_t A2_BYTES = 128 * 128;       // 128 rows x 64 bf16
constexpr uint32_t B2_BYTES = 128 * 128;       // 128 N-rows (half of the tile) x 64 bf16
constexpr uint32_t STAGE2_BYTES = A2_BYTES + B2_BYTES;
constexpr uint32_t EPI2_BYTES = 4 * 2 * 4096;
constexpr uint32_t BIAS2_BYTES = BN2 * 4;
constexpr uint32_t SMEM2_BYTES = STAGES2 * STAGE2_BYTES + EPI2_BYTES + BIAS2_BYTES + 256 + 1024;
static_assert(SMEM2_BYTES <= 232448, "exceeds 227 KB of shared memory per CTA");

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cta address of this CTA -> shared::cluster address of the same offset in CTA `rank`
__device__ __forceinline__ uint32_t map_to_cta(uint32_t saddr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap *tm, uint32_t leader_bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
            dst),
        "l"(tm), "r"(leader_bar), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tmem_alloc2(uint32_t smem_dst, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc2(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma_bf16_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrive (once all prior MMAs of this thread are done) on the barrier at this smem offset in BOTH CTAs of the pair
__device__ __forceinline__ void umma_commit_pair(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
                 "h"((uint16_t)3)
                 : "memory");
}

#ifdef KDL_PAIR_PROF
__device__ unsigned long long g_pair_prof[160][16];
#define PP_T0() long long pp_t0 = clock64()
#define PP_ADD(var) do { long long pp_t1 = clock64(); var += pp_t1 - pp_t0; pp_t0 = pp_t1; } while (0)
#define PP_DECL(...) long long __VA_ARGS__
#define PP_STORE(idx, v) g_pair_prof[blockIdx.x][idx] = (unsigned long long)(v)
#else
#define PP_T0()
#define PP_ADD(var)
#define PP_DECL(...)
#define PP_STORE(idx, v)
#endif

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS, 1)
k_gemm_tc2(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmOut,
           const float *__restrict__ bias, const bf16 *residual, int M, int N, int K, int relu, int ldo) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;  // same offset in both CTAs (same kernel, same layout)
    const uint32_t epi_base = smem_base + STAGES2 * STAGE2_BYTES;
    const uint32_t bias_base = epi_base + EPI2_BYTES;
    const uint32_t bar_base = bias_base + BIAS2_BYTES;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES2 + s); };
    auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * STAGES2 + a); };
    auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * STAGES2 + 2 + a); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES2 + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const bool leader = rank == 0;
    const int n_clusters = gridDim.x >> 1;
    const int cluster_id = blockIdx.x >> 1;
    const int m_tiles = (M + 255) / 256;
    const int n_tiles = (N + BN2 - 1) / BN2;
    const int num_tiles = m_tiles * n_tiles;
    const int k_blocks = (K + BK2 - 1) / BK2;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmA);
        prefetch_tmap(&tmB);
        prefetch_tmap(&tmOut);
        for (int s = 0; s < STAGES2; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(tfull_bar(a), 1); mbar_init(tempty_bar(a), 8); }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc2(tmem_slot, 512);
    tc_fence_before();
    cluster_sync_all();  // barriers of both CTAs initialised and TMEM allocated before anyone signals across the pair
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

    if (warp == 0) {
        // ================= TMA producer (both CTAs) =================
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            PP_DECL(w_e = 0, w_i = 0);
            PP_T0();
            const long long t_begin = clock64();
            for (int tile = cluster_id; tile < num_tiles; tile += n_clusters) {
                const int m0 = (tile / n_tiles) * 256 + (int)rank * 128;
                const int n0 = (tile % n_tiles) * BN2;
                const int n_inst = (min(BN2, N - n0) + 15) & ~15;           // instruction N; each CTA supplies n_inst/2 weight rows
                const int nb0 = n0 + (int)rank * (n_inst >> 1);
                for (int kb = 0; kb < k_blocks; ++kb) {
                    PP_ADD(w_i);
                    mbar_wait(empty_bar(stage), phase ^ 1u);
                    PP_ADD(w_e);
                    const uint32_t lfull = map_to_cta(full_bar(stage), 0);
                    if (leader) mbar_expect_tx(full_bar(stage), 2 * STAGE2_BYTES);
                    const uint32_t sa = smem_base + stage * STAGE2_BYTES;
                    tma_load_2d_pair(sa, &tmA, lfull, kb * BK2, m0);
                    tma_load_2d_pair(sa + A2_BYTES, &tmB, lfull, kb * BK2, nb0);
                    if (++stage == STAGES2) { stage = 0; phase ^= 1u; }
                }
            }
            PP_ADD(w_i);
            PP_STORE(0, w_e); PP_STORE(1, w_i); PP_STORE(7, clock64() - t_begin);
        }
    } else if (warp == 1) {
        // ================= MMA issuer (leader CTA only) =================
        if (leader && lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            PP_DECL(w_te = 0, w_f = 0, w_i = 0);
            PP_T0();
            for (int tile = cluster_id; tile < num_tiles; tile += n_clusters, ++it) {
                const int n0 = (tile % n_tiles) * BN2;
                const int acc = it & 1;
                const uint32_t acc_phase = (it >> 1) & 1;
                const int n_inst = (min(BN2, N - n0) + 15) & ~15;
                const uint32_t idesc = make_idesc(256, (uint32_t)n_inst);
                PP_ADD(w_i);
                mbar_wait(tempty_bar(acc), acc_phase ^ 1u);  // both CTAs' epilogues have drained this accumulator
                PP_ADD(w_te);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN2;
                for (int kb = 0; kb < k_blocks; ++kb) {
                    PP_ADD(w_i);
                    mbar_wait(full_bar(stage), phase);
                    PP_ADD(w_f);
                    tc_fence_after();
                    const uint32_t sa = smem_base + stage * STAGE2_BYTES;
                    const uint64_t adesc = make_smem_desc<128>(sa);
                    const uint64_t bdesc = make_smem_desc<128>(sa + A2_BYTES);
                    const int k_left = K - kb * BK2;
                    const int ksteps = k_left >= BK2 ? BK2 / UMMA_K : (k_left + UMMA_K - 1) / UMMA_K;
                    for (int k = 0; k < ksteps; ++k) umma_bf16_pair(d_tmem, adesc + 2u * k, bdesc + 2u * k, idesc, (kb | k) != 0);
                    umma_commit_pair(empty_bar(stage));
                    if (++stage == STAGES2) { stage = 0; phase ^= 1u; }
                }
                umma_commit_pair(tfull_bar(acc));
            }
            PP_ADD(w_i);
            PP_STORE(2, w_te); PP_STORE(3, w_f); PP_STORE(4, w_i);
        }
    } else {
        // ================= epilogue: warps 2..5 of both CTAs, each CTA drains its own 128 accumulator rows =================
        const int q = warp & 3;
        const int et = threadIdx.x - 64;
        const uint32_t ebuf = epi_base + (uint32_t)(warp - 2) * 8192u;
        const bool has_res = residual != nullptr;
        int staged_n0 = -1;
        int it = 0;
        PP_DECL(w_tf = 0, w_w = 0);
        PP_T0();
        for (int tile = cluster_id; tile < num_tiles; tile += n_clusters, ++it) {
            const int m0 = (tile / n_tiles) * 256 + (int)rank * 128;
            const int n0 = (tile % n_tiles) * BN2;
            const int acc = it & 1;
            const uint32_t acc_phase = (it >> 1) & 1;
            if (n0 != staged_n0) {
                named_bar_sync(1, 128);
                for (int j = et; j < BN2; j += 128) {
                    float b = (n0 + j < N) ? __ldg(bias + n0 + j) : 0.f;
                    asm volatile("st.shared.f32 [%0], %1;" ::"r"(bias_base + 4u * j), "f"(b) : "memory");
                }
                named_bar_sync(1, 128);
                staged_n0 = n0;
            }
            const int row0 = m0 + q * 32;
            const int row = row0 + lane;
            const uint32_t t_row = tmem_base + acc * BN2 + ((uint32_t)(q * 32) << 16);
            const bool row_ok = row < M;
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
            PP_ADD(w_w);
            mbar_wait(tfull_bar(acc), acc_phase);
            PP_ADD(w_tf);
            tc_fence_after();
#pragma unroll
            for (int c = 0; c < BN2 / 64; ++c) {
                const int nc = n0 + c * 64;
                if (nc < N) {
                    uint32_t r[64];
                    tmem_ld32(t_row + c * 64, r);
                    tmem_ld32(t_row + c * 64 + 32, r + 32);
                    uint4 rc[8];
#pragma unroll
                    for (int g = 0; g < 8; ++g) rc[g] = rn[g];
                    if (has_res && c + 1 < BN2 / 64 && nc + 64 < N) load_res(c + 1);
                    const uint32_t buf = ebuf + (uint32_t)(c & 1) * 4096u;
                    if (lane == 0) bulk_wait_read<1>();
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
            if (lane == 0) mbar_arrive_cluster(map_to_cta(tempty_bar(acc), 0));
        }
        if (lane == 0) bulk_wait_all<0>();
        PP_ADD(w_w);
        if (warp == 2 && lane == 0) { PP_STORE(5, w_tf); PP_STORE(6, w_w); }
    }
    tc_fence_before();
    cluster_sync_all();  // nobody leaves while the peer may still signal its barriers / read its shared memory
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc2(tmem_base, 512);
    }
}

inline int configure2(char *err) {
    cudaError_t e = cudaFuncSetAttribute(k_gemm_tc2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM2_BYTES);
    if (e != cudaSuccess) {
        snprintf(err, 512, "cudaFuncSetAttribute(k_gemm_tc2 smem): %s", cudaGetErrorString(e));
        return KDL_ECUDA;
    }
    return KDL_OK;
}

// weights via tmB2: box 64 x 128 over Wt[N,K] (each CTA of the pair loads its half of the N tile)
inline int launch_gemm2(const TmaEncoder &enc, const bf16 *A, const CUtensorMap *tmB2, const float *bias, const bf16 *residual, bf16 *out,
                        int M, int N, int K, int relu, cudaStream_t st, char *err, int lda = 0, int ldo = 0) {
    if (lda == 0) lda = K;
    if (ldo == 0) ldo = N;
    CUtensorMap tmA, tmOut;
    if (!enc.encode_2d(&tmA, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda * 2, BK2, 128, 128) ||
        !enc.encode_2d(&tmOut, out, (uint64_t)N, (uint64_t)M, (uint64_t)ldo * 2, 64, 32, 128)) {
        snprintf(err, 512, "cuTensorMapEncodeTiled failed for A [%d,%d] / out [%d,%d]", M, K, M, N);
        return KDL_ECUDA;
    }
    const int tiles = ((M + 255) / 256) * ((N + BN2 - 1) / BN2);
    const int max_clusters = num_sms() / 2;
    const int clusters = tiles < max_clusters ? tiles : max_clusters;
    k_gemm_tc2<<<2 * clusters, NUM_THREADS, SMEM2_BYTES, st>>>(tmA, *tmB2, tmOut, bias, residual, M, N, K, relu, ldo);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        snprintf(err, 512, "k_gemm_tc2 launch (M=%d N=%d K=%d): %s", M, N, K, cudaGetErrorString(e));
        return KDL_ECUDA;
    }
    return KDL_OK;
}

}  // namespace tc2
}  // namespace kdl

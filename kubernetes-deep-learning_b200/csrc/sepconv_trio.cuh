// Fused middle-flow SeparableConv2D on a thread-block cluster of THREE: sepconv_pair.cuh with the output columns split three ways, so
// that a CTA's accumulator is 256 columns and TMEM (512) holds two of them.
//
// What the pair kernel's counters said: with 384 columns per CTA the accumulator cannot be double-buffered, every
// tile's epilogue stops the tensor core, and the shallow A ring turns that stall into depthwise stalls as well.  Here
//   * the three CTAs of a cluster work on the same 128 flat pixels; CTA r owns output columns [256 r, 256 r + 256) (728 = 256 + 256
//     + 216): one M128 x N256 x K16 instruction per k-step, accumulators alternate between TMEM columns [0, 256) and [256, 512), the
//     epilogue of tile i runs under the MMAs of tile i + 1 exactly as in k_gemm_tc;
//   * the 64-channel depthwise chunks go round-robin over the three CTAs (sequence number % 3 = owner): each CTA loads and convolves
//     a third of the input channels, and a forwarder thread sends every finished 16 KB A tile to the same ring slot of both peers
//     (cp.async.bulk shared::cta -> shared::cluster, completing on the receiver's barrier);
//   * four A slots; a slot is reused when all three tensor cores have read it (tcgen05.commit multicast to the a_empty barrier
//     of the three CTAs, count 3).  Per (slot, producer) one "full" barrier, so every barrier sees a strictly periodic sequence.
//   * Everything that crosses CTAs is shared memory of THIS GPU's SMs; no TMA multicast is used.
// Arithmetic and its order are those of k_dw_tma followed by k_gemm_tc: bit-identical to the unfused path (tests/test_gpu_variants.py).
#pragma once
#include "sepconv_pair.cuh"

namespace kdl {
namespace sep3 {

using namespace kdl::tc;
using sepm::BM;
using sepm::CB;
using sepm::DW_RS;
using sepm::DW_WARPS;
using sepm::Geom;
using sepm::Raw;
#ifdef KDL_MID_PROF
using sepm::g_mid_prof;
#endif
using sepp::bulk_copy_to_peer;
using sepp::cluster_ctarank;
using sepp::cluster_sync_all;
using sepp::map_to_cta;
using sepp::mbar_expect_tx_remote;
using sepp::mbar_wait_cluster;

constexpr int CS = 3;                          // CTAs per cluster
constexpr int NC = 256;                        // output columns per CTA = one MMA instruction, two accumulators in TMEM
constexpr int NX = 2, NA = 4, NB = 2;
constexpr int NUM_THREADS_TRIO = (6 + DW_WARPS) * 32;
constexpr uint32_t DWW_BYTES = sepm::DWW_BYTES;
constexpr uint32_t XSTAGE = sepm::XSTAGE;
constexpr uint32_t A_BYTES = sepm::A_BYTES;
constexpr uint32_t B_BYTES = NC * 128;
constexpr uint32_t EPI_BYTES = 4 * 2 * 4096;   // two staging buffers per epilogue warp
constexpr uint32_t BIAS_BYTES = NC * 4;
constexpr uint32_t SMEM_BYTES = NA * A_BYTES + NB * B_BYTES + EPI_BYTES + NX * XSTAGE + BIAS_BYTES + 512 + 1024;
static_assert(SMEM_BYTES <= 232448, "exceeds 227 KB");

// arrive, once all MMAs issued so far by this thread are done, on the barrier at this offset in ALL THREE CTAs of the cluster
__device__ __forceinline__ void umma_commit_all3(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
                 "h"((uint16_t)7)
                 : "memory");
}

template <bool RELU_IN>
__global__ void __launch_bounds__(NUM_THREADS_TRIO, 1)
k_sepconv_trio(const __grid_constant__ CUtensorMap tmX /*3-D (C, W, B*H), box 64 x (W+2) x BH*/,
               const __grid_constant__ CUtensorMap tmB /*Wt[N][Kp], box 64 x 256*/,
               const __grid_constant__ CUtensorMap tmOut /*[M, N], box 64 x 32*/,
               const float *__restrict__ dwpad, const float *__restrict__ bias, const bf16 *residual, bf16 *out, Geom g) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;   // the same offset in every CTA (same kernel, same layout)
    const uint32_t a_base = smem_base;
    const uint32_t b_base = a_base + NA * A_BYTES;
    const uint32_t epi_base = b_base + NB * B_BYTES;
    const uint32_t x_base = epi_base + EPI_BYTES;
    const uint32_t bias_base = x_base + NX * XSTAGE;
    const uint32_t bar_base = bias_base + BIAS_BYTES;
    // barriers (8 bytes each)
    auto x_full = [&](int s) { return bar_base + 8u * s; };
    auto x_empty = [&](int s) { return bar_base + 8u * (NX + s); };
    auto a_fullp = [&](int s, int p) { return bar_base + 8u * (2 * NX + s * CS + p); };          // slot s filled by producer CTA p
    auto a_empty = [&](int s) { return bar_base + 8u * (2 * NX + NA * CS + s); };
    auto b_full = [&](int s) { return bar_base + 8u * (2 * NX + NA * CS + NA + s); };
    auto b_empty = [&](int s) { return bar_base + 8u * (2 * NX + NA * CS + NA + NB + s); };
    auto tfull = [&](int a) { return bar_base + 8u * (2 * NX + NA * CS + NA + 2 * NB + a); };
    auto tempty = [&](int a) { return bar_base + 8u * (2 * NX + NA * CS + NA + 2 * NB + 2 + a); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * NX + NA * CS + NA + 2 * NB + 4);
    static_assert(8 * (2 * NX + NA * CS + NA + 2 * NB + 4) + 8 <= 512, "barrier area");

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();          // 0..2: column third and chunk residue of this CTA
    const int cl_id = blockIdx.x / CS, n_cl = gridDim.x / CS;
    const int n0 = (int)rank * NC;                    // first output column of this CTA
    const bool has_cols = n0 < g.N;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmX);
        prefetch_tmap(&tmB);
        prefetch_tmap(&tmOut);
        for (int s = 0; s < NX; ++s) { mbar_init(x_full(s), 1); mbar_init(x_empty(s), DW_WARPS); }
        for (int s = 0; s < NA; ++s) {
            for (int p = 0; p < CS; ++p) mbar_init(a_fullp(s, p), p == (int)rank ? DW_WARPS : 1);   // own depthwise warps / one arrival + 16 KB of copy
            mbar_init(a_empty(s), CS);                                                               // one commit from each CTA's MMA thread
        }
        for (int s = 0; s < NB; ++s) { mbar_init(b_full(s), 1); mbar_init(b_empty(s), 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(tfull(a), 1); mbar_init(tempty(a), 4); }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, 512);
    tc_fence_before();
    cluster_sync_all();   // every CTA's barriers exist before anybody signals across the cluster
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
    pdl_launch_dependents();
    pdl_wait();
    if (warp >= 2 && warp < 6) {  // this CTA's bias slice, once
        const int et = threadIdx.x - 64;
        for (int j = et; j < NC; j += 128) {
            const float b = n0 + j < g.N ? __ldg(bias + n0 + j) : 0.f;
            asm volatile("st.shared.f32 [%0], %1;" ::"r"(bias_base + 4u * j), "f"(b) : "memory");
        }
        named_bar_sync(1, 128);
    }

    if (warp == 0) {
        if (lane == 0) {
            // ================= TMA producer, input side: the halo tile + depthwise weights of this CTA's chunks =================
            uint32_t xq = 0, seq = 0;
            for (int tile = cl_id; tile < g.num_tiles; tile += n_cl) {
                const int r_lo = (tile * BM) / g.W;        // first global image row (b*H + y) the tile touches
                for (int kc = 0; kc < g.kchunks; ++kc, ++seq) {
                    if (seq % CS != rank) continue;
                    const int xs = (int)(xq % NX);
                    mbar_wait(x_empty(xs), ((xq / NX) & 1u) ^ 1u);
                    mbar_expect_tx(x_full(xs), g.x_bytes + DWW_BYTES);
                    const uint32_t xd = x_base + xs * XSTAGE;
                    sepm::tma_load_3d(xd, &tmX, x_full(xs), kc * CB, -1, r_lo - 1);
                    dwt::bulk_load_1d(xd + g.x_bytes, dwpad + (size_t)kc * 9 * CB, DWW_BYTES, x_full(xs));
                    ++xq;
                }
            }
        } else if (lane == 1 && has_cols) {
            // ================= TMA producer, weight side: this CTA's [256 x 64] block of the pointwise weights for EVERY chunk =================
            uint32_t bq = 0;
            for (int tile = cl_id; tile < g.num_tiles; tile += n_cl) {
                for (int kc = 0; kc < g.kchunks; ++kc, ++bq) {
                    const int bs = (int)(bq % NB);
                    mbar_wait(b_empty(bs), ((bq / NB) & 1u) ^ 1u);
                    mbar_expect_tx(b_full(bs), B_BYTES);
                    tma_load_2d(b_base + bs * B_BYTES, &tmB, b_full(bs), kc * CB, n0);
                }
            }
        } else if (lane == 2) {
            // ================= forwarder: every chunk this CTA's depthwise warps finish goes to the same A slot of both peers =================
            uint32_t a_rem[CS - 1], f_rem[CS - 1];
#pragma unroll
            for (int i = 0; i < CS - 1; ++i) {
                const uint32_t q = (rank + 1u + (uint32_t)i) % CS;
                a_rem[i] = map_to_cta(a_base, q);
                f_rem[i] = map_to_cta(a_fullp(0, (int)rank), q);
            }
            uint32_t seq = 0;
            for (int tile = cl_id; tile < g.num_tiles; tile += n_cl) {
                for (int kc = 0; kc < g.kchunks; ++kc, ++seq) {
                    if (seq % CS != rank) continue;
                    const int as = (int)(seq % NA);
                    mbar_wait(a_fullp(as, (int)rank), ((seq / NA) / CS) & 1u);   // the depthwise warps fenced their stores for the async proxy before arriving
#pragma unroll
                    for (int i = 0; i < CS - 1; ++i) {
                        const uint32_t rb = f_rem[i] + 8u * (uint32_t)(as * CS);
                        mbar_expect_tx_remote(rb, A_BYTES);
                        bulk_copy_to_peer(a_rem[i] + as * A_BYTES, a_base + as * A_BYTES, A_BYTES, rb);
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer =================
        if (lane == 0) {
            const int n_valid = has_cols ? min(NC, g.N - n0) : 0;
            const uint32_t idesc = make_idesc(BM, (uint32_t)(n_valid ? ((n_valid + 15) & ~15) : 16));
            uint32_t seq = 0, bq = 0;
            int it = 0;
            MP_DECL(w_te = 0, w_af = 0, w_bf = 0, w_iss = 0);
            MP_T0();
            for (int tile = cl_id; tile < g.num_tiles; tile += n_cl, ++it) {
                const int acc = it & 1;
                if (has_cols) {
                    MP_ADD(w_iss);
                    mbar_wait(tempty(acc), (uint32_t)((it >> 1) & 1) ^ 1u);   // the epilogue has drained this accumulator
                    MP_ADD(w_te);
                    tc_fence_after();
                }
                const uint32_t d_tmem = tmem_base + (uint32_t)(acc * NC);
                for (int kc = 0; kc < g.kchunks; ++kc, ++seq) {
                    const int as = (int)(seq % NA);
                    const int owner = (int)(seq % CS);
                    MP_ADD(w_iss);
                    if (owner == (int)rank) mbar_wait(a_fullp(as, owner), ((seq / NA) / CS) & 1u);       // computed here
                    else mbar_wait_cluster(a_fullp(as, owner), ((seq / NA) / CS) & 1u);                // copied in by a peer
                    MP_ADD(w_af);
                    if (has_cols) {
                        const int bs = (int)(bq % NB);
                        mbar_wait(b_full(bs), (bq / NB) & 1u);
                        MP_ADD(w_bf);
                        tc_fence_after();
                        const uint64_t adesc = make_smem_desc<128>(a_base + as * A_BYTES);
                        const uint64_t bdesc = make_smem_desc<128>(b_base + bs * B_BYTES);
                        const int k_left = g.K - kc * CB;
                        const int ksteps = k_left >= CB ? CB / UMMA_K : (k_left + UMMA_K - 1) / UMMA_K;
                        for (int k = 0; k < ksteps; ++k) umma_bf16(d_tmem, adesc + 2u * k, bdesc + 2u * k, idesc, (kc | k) != 0);
                        umma_commit(b_empty(bs));
                        ++bq;
                    }
                    umma_commit_all3(a_empty(as));   // every CTA's depthwise warps learn that THIS tensor core is done with the slot
                }
                if (has_cols) umma_commit(tfull(acc));
            }
            MP_ADD(w_iss);
            MP_STORE(3, w_te); MP_STORE(4, w_af); MP_STORE(5, w_bf); MP_STORE(6, w_iss);
        }
    } else if (warp < 6) {
        // ================= epilogue (warps 2..5): gemm_tc.cuh's standard path on this CTA's 128 x 256 accumulators =================
        const int q = warp & 3;
        const uint32_t ebuf = epi_base + (uint32_t)(warp - 2) * 8192u;
        const bool has_res = residual != nullptr;
        uint32_t chunk_no = 0;
        int it = 0;
        MP_DECL(w_tf = 0, w_epi = 0);
        MP_T0();
        for (int tile = cl_id; has_cols && tile < g.num_tiles; tile += n_cl, ++it) {
            const int acc = it & 1;
            const int m0 = tile * BM;
            const int row0 = m0 + q * 32;
            const uint32_t t_row = tmem_base + (uint32_t)(acc * NC) + ((uint32_t)(q * 32) << 16);
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
            mbar_wait(tfull(acc), (uint32_t)((it >> 1) & 1));
            MP_ADD(w_tf);
            tc_fence_after();
#pragma unroll
            for (int c = 0; c < NC / 64; ++c) {
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
                        if (c + 1 < NC / 64 && nc + 64 < g.N) load_res(c + 1);
                    }
                    tmem_ld_wait();
#pragma unroll
                    for (int gi = 0; gi < 8; ++gi) {
                        const uint32_t slot = buf + (uint32_t)lane * 128u + (uint32_t)((gi ^ (lane & 7)) << 4);
                        const uint4 rc = has_res ? ld_shared_v4(slot) : make_uint4(0, 0, 0, 0);
                        const uint4 o = epi8(r + gi * 8, bias_base + (uint32_t)(c * 64 + gi * 8) * 4u, g.relu_out, has_res, rc);
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
            if (lane == 0) mbar_arrive(tempty(acc));
        }
        if (lane == 0) bulk_wait_all<0>();
        MP_ADD(w_epi);
        if (warp == 2 && lane == 0) { MP_STORE(7, w_tf); MP_STORE(8, w_epi); }
    } else {
        // ================= depthwise producers (warps 6..15): this CTA's chunks, into its own A ring =================
        const int dt = threadIdx.x - 6 * 32;      // 0 .. 32*DW_WARPS-1
        const int cg = dt & 15;                   // 4-channel group of the 64-channel chunk
        const int rest = dt >> 4;
        const int cp = rest % g.CP;               // column pair
        const int strip = rest / g.CP;
        const bool active = strip < g.NS && 2 * cp < g.W;
        const bool has_col1 = 2 * cp + 1 < g.W;
        uint32_t xq = 0, seq = 0;
        MP_DECL(w_xf = 0, w_ae = 0, w_cmp = 0);
        MP_T0();
        for (int tile = cl_id; tile < g.num_tiles; tile += n_cl) {
            const int g0 = tile * BM;
            const int r_lo = g0 / g.W;
            const int x_off = g0 - r_lo * g.W;                 // the tile's first pixel sits at column x_off of image row r_lo
            const int nr = (x_off + BM - 1) / g.W + 1;         // image rows the tile touches
            const int y_lo = strip * DW_RS;
            const int yimg_lo = (r_lo + y_lo) % g.H;           // position of the strip's first output row inside its image
            const int ib0 = yimg_lo == 0 ? 1 : g.H - yimg_lo + 1;
            const int ib = ib0 <= DW_RS + 1 ? ib0 : -1;        // which staged input row of the strip is the first row of an image
            for (int kc = 0; kc < g.kchunks; ++kc, ++seq) {
                if (seq % CS != rank) continue;
                const int xs = (int)(xq % NX), as = (int)(seq % NA);
                MP_ADD(w_cmp);
                mbar_wait(x_full(xs), (xq / NX) & 1u);
                MP_ADD(w_xf);
                mbar_wait_cluster(a_empty(as), ((seq / NA) & 1u) ^ 1u);   // all three tensor cores have finished with the slot's previous chunk
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
                                const uint32_t off = r * 128u + ((((uint32_t)cg >> 1) ^ (r & 7u)) << 4) + ((uint32_t)cg & 1u) * 8u;
                                const uint32_t lo = pack_bf16(aA[j][0].x, aA[j][0].y), hi = pack_bf16(aA[j][1].x, aA[j][1].y);
                                asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(at + off), "r"(lo), "r"(hi) : "memory");
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
                    dwt::Px in[4];
                    Raw cur = sepm::load_raw(rowaddr);
                    Raw nxt = sepm::load_raw(rowaddr + rowpitch);
                    rowaddr += 2u * rowpitch;
#pragma unroll
                    for (int i = 0; i < DW_RS + 2; ++i) {
                        sepm::unpack_raw<RELU_IN>(cur, in);
                        cur = nxt;
                        if (i + 2 < DW_RS + 2) { nxt = sepm::load_raw(rowaddr); rowaddr += rowpitch; }
                        const bool a = i != ib, c = i + 1 != ib;
                        if (i == 0) sepm::scatter_rt<false, false, true>(in, wl, wh, aA, aB, aC, a, c);
                        else if (i == 1) sepm::scatter_rt<false, true, true>(in, wl, wh, aA, aB, aC, a, c);
                        else if (i <= DW_RS) sepm::scatter_rt<true, true, true>(in, wl, wh, aA, aB, aC, a, c);
                        else sepm::scatter_rt<true, false, false>(in, wl, wh, aA, aB, aC, a, c);
                        if (i >= 2) store_row();
                        if (i <= DW_RS) rotate();
                    }
                }
                fence_proxy_async_smem();   // generic-proxy stores -> visible to the async proxy (this tensor core and the forwarder's bulk copies)
                __syncwarp();
                if (lane == 0) {
                    mbar_arrive(a_fullp(as, (int)rank));
                    mbar_arrive(x_empty(xs));
                }
                ++xq;
            }
        }
        MP_ADD(w_cmp);
        if (lane == 0 && (warp == 6 || warp == 15)) {
            const int o = warp == 6 ? 9 : 12;
            MP_STORE(o, w_xf); MP_STORE(o + 1, w_ae); MP_STORE(o + 2, w_cmp);
        }
    }
    tc_fence_before();
    cluster_sync_all();   // nobody leaves while a peer may still copy into its shared memory or signal its barriers
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// host ---------------------------------------------------------------------------------------------------
inline int configure(char *err) {
    cudaError_t e = cudaFuncSetAttribute(k_sepconv_trio<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_sepconv_trio<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES);
    if (e != cudaSuccess) {
        snprintf(err, 512, "cudaFuncSetAttribute(k_sepconv_trio): %s", cudaGetErrorString(e));
        return KDL_ECUDA;
    }
    return KDL_OK;
}

// the shapes of k_sepconv_mid whose output needs all three column blocks
inline bool supported(int H, int W, int K, int N) { return sepm::supported(H, W, K, N) && N > 2 * NC && N <= 3 * NC; }

template <bool RELU>
inline int launch_t(const CUtensorMap &tmX, const CUtensorMap &tmB, const CUtensorMap &tmOut, const float *dwpad, const float *bias,
                    const bf16 *residual, bf16 *out, const Geom &g, cudaStream_t st, char *err) {
    static int max_clusters = -1;   // how many 3-CTA clusters of this kernel the GPU holds at once (depends on the GPC layout)
    cudaLaunchConfig_t cfg = {};
    cfg.blockDim = dim3(NUM_THREADS_TRIO);
    cfg.dynamicSmemBytes = SMEM_BYTES;
    cfg.stream = st;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = CS;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    if (max_clusters < 0) {
        cfg.gridDim = dim3(CS * (num_sms() / CS));
        int n = 0;
        cudaError_t e = cudaOccupancyMaxActiveClusters(&n, k_sepconv_trio<RELU>, &cfg);
        if (e != cudaSuccess || n < 1) {
            snprintf(err, 512, "cudaOccupancyMaxActiveClusters(k_sepconv_trio): %s (%d)", cudaGetErrorString(e), n);
            cudaGetLastError();
            return KDL_ECUDA;
        }
        max_clusters = n < num_sms() / CS ? n : num_sms() / CS;
    }
    const int clusters = g.num_tiles < max_clusters ? g.num_tiles : max_clusters;
    cfg.gridDim = dim3(CS * clusters);
    cfg.numAttrs = pdl_enabled() ? 2 : 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, k_sepconv_trio<RELU>, tmX, tmB, tmOut, dwpad, bias, residual, out, g);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) {
        snprintf(err, 512, "k_sepconv_trio launch (%d clusters of %d): %s", clusters, CS, cudaGetErrorString(e));
        return KDL_ECUDA;
    }
    return KDL_OK;
}

// same contract as sepm::launch, except that tmB256 is the pointwise weights' 64 x 256 box (the plain GEMM's tensor map)
inline int launch(const TmaEncoder &enc, const bf16 *in, const float *dwpad, const CUtensorMap *tmB256, const float *bias,
                  const bf16 *residual, bf16 *out, int B, int H, int W, int K, int N, int ld_in, int ld_out, int relu_in, int relu_out,
                  cudaStream_t st, char *err) {
    Geom g;
    g.M = B * H * W; g.H = H; g.W = W; g.K = K; g.N = N;
    g.num_tiles = (g.M + BM - 1) / BM;
    g.n_halves = CS;
    g.items = g.num_tiles;
    g.kchunks = (K + CB - 1) / CB;
    g.BWp = W + 2;
    g.CP = (W + 1) / 2;
    g.NS = (sepm::tile_rows(W) + DW_RS - 1) / DW_RS;
    g.RS = DW_RS;
    g.BH = g.NS * DW_RS + 2;
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
            snprintf(err, 512, "cuTensorMapEncodeTiled failed for trio sepconv input [%d,%d,%d,%d]", B, H, W, K);
            return KDL_ECUDA;
        }
    }
    if (!enc.encode_2d(&tmOut, out, (uint64_t)N, (uint64_t)g.M, (uint64_t)ld_out * 2, 64, 32, 128)) {
        snprintf(err, 512, "cuTensorMapEncodeTiled failed for trio sepconv output [%d,%d]", g.M, N);
        return KDL_ECUDA;
    }
    return relu_in ? launch_t<true>(tmX, *tmB256, tmOut, dwpad, bias, residual, out, g, st, err)
                   : launch_t<false>(tmX, *tmB256, tmOut, dwpad, bias, residual, out, g, st, err);
}

}  // namespace sep3
}  // namespace kdl

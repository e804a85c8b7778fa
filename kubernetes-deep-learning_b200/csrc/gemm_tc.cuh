// K5 / K3: GEMM-class layers on the 5th-gen tensor cores (sm_100a), hand-written PTX.
//
//   pointwise / projection / (later) dense:  out[M,N] = act( A[M,K] * Wt[N,K]^T + bias[N] ) (+ residual[M,N])
//   block1_conv2 (3x3, s1, VALID, 32->64) :  the same kernel in implicit-GEMM mode -- K = 9 taps x 32 channels, the A tile
//                                            of tap (ky,kx) is the activation matrix shifted by ky*W+kx rows.
//   bf16 operands, fp32 accumulation in TMEM, bf16 output.
//
//   * operands staged by TMA (cp.async.bulk.tensor.2d, 128- or 64-byte swizzle) into a shared-memory ring,
//   * tcgen05.mma.cta_group::1.kind::f16 (M=128, N<=256, K=16 per instruction) issued by one elected thread,
//   * accumulators double-buffered in TMEM so the epilogue of tile i overlaps the MMAs of tile i+1,
//   * warp-specialised persistent CTAs (one per SM): warp 0 = TMA producer, warp 1 = MMA issuer + TMEM owner,
//     warps 2..5 = epilogue: tcgen05.ld -> +bias (folded BatchNorm shift, staged in smem once per tile) -> ReLU ->
//     +residual (prefetched one chunk ahead) -> bf16 -> swizzled smem -> TMA store (clipped at the tensor edge),
//   * ragged shapes (C = 728): K / N / M tails are zero-filled by TMA out-of-bounds handling and clipped on store.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdio.h>

#include "common.cuh"

#ifndef KDL_OK
#define KDL_OK 0
#define KDL_ECUDA (-2)
#endif

namespace kdl {
namespace tc {

constexpr int BLOCK_M = 128;
constexpr int UMMA_K = 16;
constexpr int NUM_THREADS = 192;
constexpr int BLOCK_N = 256;  // standard configuration (weight tensor-map box rows)
constexpr int BLOCK_K = 64;   // standard configuration: 64 bf16 = 128 bytes = one swizzle-128B row

// NST = operand ring depth, EB = output staging buffers per epilogue warp.  Tensor-bound layers want a deep operand ring
// (4 x 48 KB) and get by with 2 store buffers; the HBM-bound K<=256 layers are limited by TMA stores in flight, so they trade
// one ring stage for 4 store buffers per warp.
// EW = epilogue warps: 4 (one per TMEM lane quarter), or 8 = two per quarter taking alternate 64-column chunks.  The K <= 256 layers
// spend 16 MMAs per tile in the main loop and ~4 dependent epilogue chunks per warp: there the four warps are the bottleneck.
template <int SW, int BN, bool CONV, int NST = (CONV ? 12 : 4), int EB = 2, int EW = 4> struct Cfg {
    static constexpr int BLOCK_N = BN;
    static constexpr int BLOCK_K = SW / 2;  // elements per k-block = one swizzle row
    static constexpr int STAGES = NST;
    static constexpr int EPI_BUFS = EB;
    static constexpr uint32_t A_BYTES = BLOCK_M * SW;
    static constexpr uint32_t B_BYTES = CONV ? 0 : BN * SW;           // conv mode keeps all 9 weight taps resident instead
    static constexpr uint32_t W_RES_BYTES = CONV ? 9 * BN * SW : 0;   // 9 taps x [64 out-ch x 32 in-ch] bf16 = 36 KB
    static constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int EPI_WARPS = EW;
    static constexpr int THREADS = (2 + EW) * 32;
    static constexpr uint32_t EPI_BYTES = CONV ? 0 : EW * EB * 4096;  // per epilogue warp: EB x (32 rows x 128 B)
    static constexpr uint32_t BIAS_BYTES = BN * 4;
    static constexpr uint32_t BAR_BYTES = 256;
    static constexpr uint32_t SMEM_BYTES = STAGES * STAGE_BYTES + W_RES_BYTES + EPI_BYTES + BIAS_BYTES + BAR_BYTES + 1024 /*align slack*/;
    static constexpr uint32_t TMEM_COLS = 2 * BN <= 32 ? 32 : 2 * BN <= 64 ? 64 : 2 * BN <= 128 ? 128 : 2 * BN <= 256 ? 256 : 512;
    static_assert(SMEM_BYTES <= 232448, "exceeds 227 KB of shared memory per CTA");
    static_assert(2 * BN <= 512, "two accumulators must fit the 512 TMEM columns");
};
#ifndef KDL_GEMM_STD_EW
#define KDL_GEMM_STD_EW 4   /* 8 (one staging buffer per warp) measured slower on the K >= 728 layers: 3.49 vs 3.38 ms */
#endif
typedef Cfg<128, 256, false, 4, (KDL_GEMM_STD_EW == 8 ? 1 : 2), KDL_GEMM_STD_EW> CfgStd;
typedef Cfg<128, 256, false, 3, 2, 8> CfgHbm;
typedef Cfg<128, 64, false, 4, 2> CfgSmallM;   // 128 x 64 tiles: 4x more CTAs and half the operand bytes per CTA when M is small (latency)
typedef Cfg<64, 64, true> CfgConv;

// geometry of the implicit-GEMM conv mode (ignored otherwise)
struct ConvGeom {
    int Hin, Win, Hout, Wout;
};

// ---------------------------------------------------------------------------------------------------
// host: TMA descriptor encoder (driver entry point fetched at run time; no link-time libcuda dependency)
// ---------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

struct TmaEncoder {
    EncodeTiledFn fn = nullptr;
    void init() {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
        else
            (void)cudaGetLastError();
    }
    bool ok() const { return fn != nullptr; }
    // row-major bf16 matrix [rows, cols] with `pitch_bytes` per row; box = box_cols x box_rows; swizzle 128 or 64 bytes
    // (box_cols * 2 must equal the swizzle span); out-of-bounds elements read as zero / are not written.
    bool encode_2d(CUtensorMap *tm, const void *base, uint64_t cols, uint64_t rows, uint64_t pitch_bytes, uint32_t box_cols,
                   uint32_t box_rows, int swizzle = 128) const {
        cuuint64_t dims[2] = {cols, rows};
        cuuint64_t strides[1] = {pitch_bytes};
        cuuint32_t box[2] = {box_cols, box_rows};
        cuuint32_t estr[2] = {1, 1};
        CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        return r == CUDA_SUCCESS;
    }
};

inline bool shape_supported(int M, int N, int K) { return M >= 1 && N >= 8 && K >= 8 && (N % 8) == 0 && (K % 8) == 0; }

// ---------------------------------------------------------------------------------------------------
// device: PTX wrappers
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}
// Spins on the barrier; a wait longer than ~2 s of SM clocks means a protocol bug -> trap instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    if (mbar_try_wait(bar, parity)) return;
    long long t0 = clock64();
    while (!mbar_try_wait(bar, parity)) {
        if (clock64() - t0 > 4000000000LL) {
            printf("kdl gemm_tc: mbarrier wait timed out (block %d thread %d bar %u parity %u)\n", blockIdx.x, threadIdx.x, bar,
                   parity);
            __trap();
        }
    }
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

__device__ __forceinline__ void tma_prefetch_2d(const CUtensorMap *tm, int c0, int c1) {
    asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global.tile [%0, {%1, %2}];" ::"l"(tm), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *tm, uint32_t bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
        "l"(tm), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap *tm, uint32_t src, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(tm), "r"(src), "r"(c0),
                 "r"(c1)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N> __device__ __forceinline__ void bulk_wait_all() {
    asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void prefetch_tmap(const CUtensorMap *tm) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(tm) : "memory");
}

__device__ __forceinline__ void tmem_alloc(uint32_t smem_dst, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem] * B[smem]^T ; both operands K-major, described by 64-bit shared-memory descriptors
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrives on the mbarrier once all previously issued tcgen05.mma of this thread have completed
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t *r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// Shared-memory matrix descriptor: K-major operand tile written by TMA with SW-byte swizzle (rows of SW bytes).
//   bits [0,14)  start address >> 4          bits [16,30) leading byte offset >> 4 (unused for swizzled K-major: 1)
//   bits [32,46) stride byte offset >> 4 = 8 rows * SW bytes between 8-row groups
//   bits [46,48) descriptor version = 1 (sm_100)       bits [61,64) layout type: 2 = SWIZZLE_128B, 4 = SWIZZLE_64B
template <int SW> __device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)((8 * SW) >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)(SW == 128 ? 2 : 4) << 61;
    return d;
}
// Instruction descriptor for kind::f16: D = F32 (bits 4-5 = 1), A = B = BF16 (bits 7-9, 10-12 = 1), both K-major
// (bits 15, 16 = 0), N >> 3 at bits [17,23), M >> 4 at bits [24,29).
__device__ __forceinline__ uint32_t make_idesc(uint32_t M, uint32_t N) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((N >> 3) << 17) | ((M >> 4) << 24);
}

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t *>(&v);
}
__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint4 v) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ld_shared_v4(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ float4 ld_shared_f4(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}

// bias (+ReLU) (+residual) on 8 accumulator columns -> one 16-byte bf16 vector.  Instruction-lean on purpose (the epilogue of the
// HBM-bound K<=256 layers is issue-bound): packed fp32 adds (add.f32x2), ReLU as one packed bf16x2 max per pair AFTER rounding
// (max commutes with round-to-nearest), residual added in fp32 before the rounding.
__device__ __forceinline__ float2 fadd2(float2 a, float2 b) {
    unsigned long long d;
    asm("add.rn.f32x2 %0, %1, %2;"
        : "=l"(d)
        : "l"(*reinterpret_cast<unsigned long long *>(&a)), "l"(*reinterpret_cast<unsigned long long *>(&b)));
    return *reinterpret_cast<float2 *>(&d);
}
__device__ __forceinline__ uint4 epi8(const uint32_t *acc, uint32_t bias_addr, int relu, bool has_res, uint4 rr) {
    const float4 b0 = ld_shared_f4(bias_addr), b1 = ld_shared_f4(bias_addr + 16);
    float2 v0 = fadd2(make_float2(__uint_as_float(acc[0]), __uint_as_float(acc[1])), make_float2(b0.x, b0.y));
    float2 v1 = fadd2(make_float2(__uint_as_float(acc[2]), __uint_as_float(acc[3])), make_float2(b0.z, b0.w));
    float2 v2 = fadd2(make_float2(__uint_as_float(acc[4]), __uint_as_float(acc[5])), make_float2(b1.x, b1.y));
    float2 v3 = fadd2(make_float2(__uint_as_float(acc[6]), __uint_as_float(acc[7])), make_float2(b1.z, b1.w));
    if (has_res) {
        v0 = fadd2(v0, make_float2(__uint_as_float(rr.x << 16), __uint_as_float(rr.x & 0xffff0000u)));
        v1 = fadd2(v1, make_float2(__uint_as_float(rr.y << 16), __uint_as_float(rr.y & 0xffff0000u)));
        v2 = fadd2(v2, make_float2(__uint_as_float(rr.z << 16), __uint_as_float(rr.z & 0xffff0000u)));
        v3 = fadd2(v3, make_float2(__uint_as_float(rr.w << 16), __uint_as_float(rr.w & 0xffff0000u)));
    }
    uint4 o;
    o.x = pack_bf16(v0.x, v0.y); o.y = pack_bf16(v1.x, v1.y);
    o.z = pack_bf16(v2.x, v2.y); o.w = pack_bf16(v3.x, v3.y);
    if (relu) {
        asm("max.bf16x2 %0, %0, %1;" : "+r"(o.x) : "r"(0u));
        asm("max.bf16x2 %0, %0, %1;" : "+r"(o.y) : "r"(0u));
        asm("max.bf16x2 %0, %0, %1;" : "+r"(o.z) : "r"(0u));
        asm("max.bf16x2 %0, %0, %1;" : "+r"(o.w) : "r"(0u));
    }
    return o;
}

// ---------------------------------------------------------------------------------------------------
// the kernel
// ---------------------------------------------------------------------------------------------------
template <int SW, int BN, bool CONV, int NST, int EB, int EW>
__device__ __forceinline__ void gemm_tc_body(const CUtensorMap &tmA, const CUtensorMap &tmB, const CUtensorMap &tmOut, const CUtensorMap &tmRes,
                                             const float *__restrict__ bias, const bf16 *residual, bf16 *out, int M, int N, int K, int relu,
                                             ConvGeom cg, int ldo, int rev) {
    typedef Cfg<SW, BN, CONV, NST, EB, EW> C;
    constexpr int STAGES = C::STAGES;
    constexpr int BK = C::BLOCK_K;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;  // swizzled tiles need 1024-byte alignment
    const uint32_t wres_base = smem_base + STAGES * C::STAGE_BYTES;  // conv mode: resident weight taps
    const uint32_t epi_base = wres_base + C::W_RES_BYTES;
    const uint32_t bias_base = epi_base + C::EPI_BYTES;
    const uint32_t bar_base = bias_base + C::BIAS_BYTES;
    // barrier slots (8 B each): full[STAGES], empty[STAGES], tmem_full[2], tmem_empty[2]; then the TMEM base address
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
    auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + a); };
    auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + 2 + a); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES + 4);
    const uint32_t wres_bar = bar_base + 8u * (2 * STAGES + 5);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int m_tiles = (M + BLOCK_M - 1) / BLOCK_M;
    const int n_tiles = (N + BN - 1) / BN;
    const int num_tiles = m_tiles * n_tiles;
    const int k_blocks = (K + BK - 1) / BK;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmA);
        prefetch_tmap(&tmB);
        if (!CONV) prefetch_tmap(&tmOut);
        for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(tfull_bar(a), 1); mbar_init(tempty_bar(a), EW); }
        mbar_init(wres_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, C::TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
    pdl_launch_dependents();
    pdl_wait();  // everything above overlapped the previous kernel; global memory is touched only from here on
#ifdef KDL_MID_PROF
    ClkMeter clk_meter(CONV ? 2 : 0);
#endif

    if (warp == 0) {
        // ================= TMA producer =================
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            if (CONV) {  // all 9 weight taps once per CTA (persistent over its tiles)
                mbar_expect_tx(wres_bar, C::W_RES_BYTES);
                for (int t = 0; t < 9; ++t) tma_load_2d(wres_base + t * (BN * SW), &tmB, wres_bar, t * BK, 0);
            }
            for (int ti = blockIdx.x; ti < num_tiles; ti += gridDim.x) {
                const int tile = rev ? num_tiles - 1 - ti : ti;
                const int m0 = (tile / n_tiles) * BLOCK_M, n0 = (tile % n_tiles) * BN;
                if (!CONV && residual != nullptr) {
                    // pull this tile's residual block into L2 now: the epilogue reads it ~one tile (several us) from now with plain
                    // loads one 64-column chunk ahead, which hides an L2 hit but not an HBM miss (measured: +30 us per launch)
#pragma unroll
                    for (int c = 0; c < BN / 64; ++c)
                        if (n0 + c * 64 < N) tma_prefetch_2d(&tmRes, n0 + c * 64, m0);
                }
                for (int kb = 0; kb < k_blocks; ++kb) {
                    mbar_wait(empty_bar(stage), phase ^ 1u);
                    mbar_expect_tx(full_bar(stage), C::STAGE_BYTES);
                    const uint32_t sa = smem_base + stage * C::STAGE_BYTES;
                    if (CONV) {
                        // tap kb = (ky,kx): the activation matrix [B*Hin*Win, Cin] shifted by ky*Win + kx rows
                        tma_load_2d(sa, &tmA, full_bar(stage), 0, m0 + (kb / 3) * cg.Win + (kb % 3));
                    } else {
                        tma_load_2d(sa, &tmA, full_bar(stage), kb * BK, m0);
                    }
                    if (!CONV) tma_load_2d(sa + C::A_BYTES, &tmB, full_bar(stage), kb * BK, n0);
                    if (++stage == STAGES) { stage = 0; phase ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer =================
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            if (CONV) mbar_wait(wres_bar, 0);
            for (int ti = blockIdx.x; ti < num_tiles; ti += gridDim.x, ++it) {
                const int tile = rev ? num_tiles - 1 - ti : ti;
                const int n0 = (tile % n_tiles) * BN;
                const int acc = it & 1;
                const uint32_t acc_phase = (it >> 1) & 1;
                // shrink the instruction to the valid part of a ragged N tile (multiple of 16)
                const int n_valid = min(BN, N - n0);
                const uint32_t idesc = make_idesc(BLOCK_M, (uint32_t)((n_valid + 15) & ~15));
                mbar_wait(tempty_bar(acc), acc_phase ^ 1u);  // epilogue has drained this accumulator
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN;
                for (int kb = 0; kb < k_blocks; ++kb) {
                    mbar_wait(full_bar(stage), phase);
                    tc_fence_after();
                    const uint32_t sa = smem_base + stage * C::STAGE_BYTES;
                    const uint64_t adesc = make_smem_desc<SW>(sa);
                    const uint64_t bdesc = make_smem_desc<SW>(CONV ? wres_base + kb * (BN * SW) : sa + C::A_BYTES);
                    const int k_left = K - kb * BK;
                    const int ksteps = k_left >= BK ? BK / UMMA_K : (k_left + UMMA_K - 1) / UMMA_K;
                    for (int k = 0; k < ksteps; ++k) {
                        // advance 16 elements = 32 bytes inside the swizzle row: +2 in the (addr >> 4) field
                        umma_bf16(d_tmem, adesc + 2u * k, bdesc + 2u * k, idesc, (kb | k) != 0);
                    }
                    umma_commit(empty_bar(stage));  // frees the smem stage when these MMAs have read it
                    if (++stage == STAGES) { stage = 0; phase ^= 1u; }
                }
                umma_commit(tfull_bar(acc));  // accumulator complete -> epilogue
            }
        }
    } else {
        // ================= epilogue: warps 2..(1+EW); TMEM lane quarter = warp % 4; with EW = 8 warp pairs split the chunks =================
        const int q = warp & 3;
        const int half = (warp - 2) >> 2;            // 0 for EW = 4
        constexpr int CSTEP = EW / 4;                // chunk stride of a warp
        const int et = threadIdx.x - 64;  // 0..32*EW-1
        const uint32_t ebuf = epi_base + (uint32_t)(warp - 2) * (uint32_t)(C::EPI_BUFS * 4096);
        const bool has_res = residual != nullptr;
        int staged_n0 = -1, pref_n0 = -1;
        uint32_t chunk_no = 0;
        float bpre[(BN + 127) / 128];
        int it = 0;
        for (int ti = blockIdx.x; ti < num_tiles; ti += gridDim.x, ++it) {
            const int tile = rev ? num_tiles - 1 - ti : ti;
            const int m0 = (tile / n_tiles) * BLOCK_M, n0 = (tile % n_tiles) * BN;
            const int acc = it & 1;
            const uint32_t acc_phase = (it >> 1) & 1;
            // stage this tile's bias slice (all four epilogue warps share it); skipped while the N tile does not change, and the
            // values were prefetched into registers during the previous tile so no global-load latency sits between the barriers
            if (n0 != staged_n0) {
                named_bar_sync(1, 32 * EW);
                if (n0 != pref_n0) {
#pragma unroll
                    for (int j = 0; j < (BN + 127) / 128; ++j)
                        bpre[j] = (et + 128 * j < BN && n0 + et + 128 * j < N) ? __ldg(bias + n0 + et + 128 * j) : 0.f;
                }
#pragma unroll
                for (int j = 0; j < (BN + 127) / 128; ++j)
                    if (et + 128 * j < BN) asm volatile("st.shared.f32 [%0], %1;" ::"r"(bias_base + 4u * (et + 128 * j)), "f"(bpre[j]) : "memory");
                named_bar_sync(1, 32 * EW);
                staged_n0 = n0;
            }
            {
                const int nti = ti + gridDim.x;
                if (nti < num_tiles) {
                    const int next = rev ? num_tiles - 1 - nti : nti;
                    const int nn0 = (next % n_tiles) * BN;
                    if (nn0 != staged_n0) {
#pragma unroll
                        for (int j = 0; j < (BN + 127) / 128; ++j)
                            bpre[j] = (et + 128 * j < BN && nn0 + et + 128 * j < N) ? __ldg(bias + nn0 + et + 128 * j) : 0.f;
                        pref_n0 = nn0;
                    }
                }
            }
            const int row0 = m0 + q * 32;
            const int row = row0 + lane;
            const uint32_t t_row = tmem_base + acc * BN + ((uint32_t)(q * 32) << 16);
            if (CONV) {
                // ---- implicit-GEMM conv: row = input-grid position (b,y,x); valid outputs have y < Hout, x < Wout ----
                mbar_wait(tfull_bar(acc), acc_phase);
                tc_fence_after();
                uint32_t r[64];
                tmem_ld32(t_row, r);
                tmem_ld32(t_row + 32, r + 32);
                tmem_ld_wait();
                const int hw = cg.Hin * cg.Win;
                const int b = row / hw;
                const int rem = row - b * hw;
                const int y = rem / cg.Win, x = rem - y * cg.Win;
                if (row < M && y < cg.Hout && x < cg.Wout) {
                    bf16 *o = out + ((size_t)(b * cg.Hout + y) * cg.Wout + x) * N + n0;
#pragma unroll
                    for (int g = 0; g < 8; ++g)
                        *reinterpret_cast<uint4 *>(o + g * 8) = epi8(r + g * 8, bias_base + 32u * g, relu, false, make_uint4(0, 0, 0, 0));
                }
            } else {
                // ---- standard: 64-column chunks -> swizzled smem -> TMA store ----
                // Residual: a lane owns one output ROW, so loading its row directly would touch 32 cache lines per instruction.
                // Instead the warp loads the 32 x 128-byte residual chunk coalesced (instruction g covers rows 4g..4g+3, eight lanes
                // per row), one chunk ahead in registers, transposes it through the chunk's staging buffer (same swizzle as the
                // output) and each lane then reads back its own row.
                const int rl_row = row0 + (lane >> 3);                       // row of this lane in coalesced instruction 0
                const bf16 *res_lane = has_res ? residual + (size_t)rl_row * ldo + (lane & 7) * 8 : nullptr;
                uint4 rn[8];
                auto load_res = [&](int c) {
                    const int n = n0 + c * 64 + (lane & 7) * 8;
#pragma unroll
                    for (int g = 0; g < 8; ++g)
                        rn[g] = (rl_row + 4 * g < M && n < N) ? *reinterpret_cast<const uint4 *>(res_lane + (size_t)(4 * g) * ldo + n0 + c * 64)
                                                             : make_uint4(0, 0, 0, 0);
                };
                if (has_res) load_res(half);
                mbar_wait(tfull_bar(acc), acc_phase);
                tc_fence_after();
#ifndef KDL_ABLATE_EPI
#define KDL_ABLATE_EPI 0     // measurement builds only: 1 = no epilogue work at all, 2 = everything but the TMA stores
#endif
#pragma unroll
                for (int c = 0; c < BN / 64; ++c) {
                    const int nc = n0 + c * 64;
                    if (nc < N && KDL_ABLATE_EPI != 1 && (c % CSTEP) == half) {  // warp-uniform
                        uint32_t r[64];
                        tmem_ld32(t_row + c * 64, r);
                        tmem_ld32(t_row + c * 64 + 32, r + 32);
                        const uint32_t buf = ebuf + (uint32_t)(chunk_no++ % C::EPI_BUFS) * 4096u;  // ring over ALL chunks of all tiles
                        // the TMA store issued EPI_BUFS chunks ago from this buffer must have finished reading it
                        if (lane == 0) bulk_wait_read<C::EPI_BUFS - 1>();
                        __syncwarp();
                        uint4 rc[8];
                        if (has_res) {
#pragma unroll
                            for (int g = 0; g < 8; ++g) {
                                const uint32_t rr = (uint32_t)(4 * g + (lane >> 3));
                                st_shared_v4(buf + rr * 128u + ((((uint32_t)lane & 7u) ^ (rr & 7u)) << 4), rn[g]);
                            }
                            __syncwarp();
#pragma unroll
                            for (int g = 0; g < 8; ++g) rc[g] = ld_shared_v4(buf + (uint32_t)lane * 128u + (uint32_t)((g ^ (lane & 7)) << 4));
                            if (c + CSTEP < BN / 64 && nc + 64 * CSTEP < N) load_res(c + CSTEP);
                        } else {
#pragma unroll
                            for (int g = 0; g < 8; ++g) rc[g] = make_uint4(0, 0, 0, 0);
                        }
                        tmem_ld_wait();
#pragma unroll
                        for (int g = 0; g < 8; ++g) {
                            const uint4 o = epi8(r + g * 8, bias_base + (uint32_t)(c * 64 + g * 8) * 4u, relu, has_res, rc[g]);
                            st_shared_v4(buf + (uint32_t)lane * 128u + (uint32_t)((g ^ (lane & 7)) << 4), o);
                        }
                        fence_proxy_async_smem();
                        __syncwarp();
                        if (lane == 0 && row0 < M && KDL_ABLATE_EPI != 2) {
                            tma_store_2d(&tmOut, buf, nc, row0);
                            bulk_commit();
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty_bar(acc));
        }
        if (!CONV && lane == 0) bulk_wait_all<0>();
    }
    tc_fence_before();
    __syncthreads();
#ifdef KDL_MID_PROF
    clk_meter.stop();
#endif
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, C::TMEM_COLS);
    }
}

template <int SW, int BN, bool CONV, int NST = (CONV ? 12 : 4), int EB = 2, int EW = 4>
__global__ void __launch_bounds__((2 + EW) * 32, 1)
k_gemm_tc(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
          const __grid_constant__ CUtensorMap tmOut, const __grid_constant__ CUtensorMap tmRes /*residual, box 64 x 128; L2 prefetch only*/,
          const float *__restrict__ bias, const bf16 *residual, bf16 *out, int M, int N, int K, int relu, ConvGeom cg,
          int ldo /*row pitch of out / residual in elements*/, int rev = 0 /*1: walk the tiles from the last to the first*/) {
    gemm_tc_body<SW, BN, CONV, NST, EB, EW>(tmA, tmB, tmOut, tmRes, bias, residual, out, M, N, K, relu, cg, ldo, rev);
}
// "slim" build for the co-residency probe (engine.cu, probe builds): 3-stage ring (178 KB of shared memory) and at most 128 registers
// per thread, so that a depthwise CTA fits on the same SM
__global__ void __maxnreg__(128)
k_gemm_tc_slim(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmOut,
               const __grid_constant__ CUtensorMap tmRes, const float *__restrict__ bias, const bf16 *residual, bf16 *out, int M, int N, int K,
               int relu, int ldo, int rev) {
    gemm_tc_body<128, 256, false, 3, 2, 4>(tmA, tmB, tmOut, tmRes, bias, residual, out, M, N, K, relu, ConvGeom{0, 0, 0, 0}, ldo, rev);
}
typedef Cfg<128, 256, false, 3, 2, 4> CfgSlim;

// ---------------------------------------------------------------------------------------------------
// host launch
// ---------------------------------------------------------------------------------------------------
inline int configure(char *err) {
    cudaError_t e = cudaFuncSetAttribute(k_gemm_tc<128, 256, false, 4, CfgStd::EPI_BUFS, CfgStd::EPI_WARPS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)CfgStd::SMEM_BYTES);
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(k_gemm_tc<128, 256, false, 3, 2, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CfgHbm::SMEM_BYTES);
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(k_gemm_tc<128, 64, false, 4, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CfgSmallM::SMEM_BYTES);
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(k_gemm_tc<64, 64, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CfgConv::SMEM_BYTES);
    if (e != cudaSuccess) {
        snprintf(err, 512, "cudaFuncSetAttribute(k_gemm_tc smem): %s", cudaGetErrorString(e));
        return KDL_ECUDA;
    }
    return KDL_OK;
}

inline int num_sms() {
    static thread_local int n = 0;
    if (n == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (n <= 0) n = 148;
    }
    return n;
}

// standard GEMM: A [M,K] bf16 row-major, weights via tmB (box BLOCK_K x BLOCK_N over Wt[N,K]), out [M,N]
#ifndef KDL_HBM_CFG_MAX_K
#define KDL_HBM_CFG_MAX_K 256
#endif
// lda / ldo: row pitch (elements) of A and of out/residual; 0 = dense (K / N)
// tmB64 (optional): the same weights with a 64-row box; used when the 128x256 tiling would leave most SMs idle (small M)
inline int launch_gemm(const TmaEncoder &enc, const bf16 *A, const CUtensorMap *tmB, const float *bias, const bf16 *residual,
                       bf16 *out, int M, int N, int K, int relu, cudaStream_t st, char *err, int lda = 0, int ldo = 0,
                       const CUtensorMap *tmB64 = nullptr, int rev = 0) {
    if (lda == 0) lda = K;
    if (ldo == 0) ldo = N;
    CUtensorMap tmA, tmOut, tmRes;
    if (!enc.encode_2d(&tmA, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda * 2, BLOCK_K, BLOCK_M, 128) ||
        !enc.encode_2d(&tmOut, out, (uint64_t)N, (uint64_t)M, (uint64_t)ldo * 2, 64, 32, 128)) {
        snprintf(err, 512, "cuTensorMapEncodeTiled failed for A [%d,%d] / out [%d,%d]", M, K, M, N);
        return KDL_ECUDA;
    }
    if (residual == nullptr) tmRes = tmOut;   // never dereferenced
    else if (!enc.encode_2d(&tmRes, residual, (uint64_t)N, (uint64_t)M, (uint64_t)ldo * 2, 64, BLOCK_M, 128)) {
        snprintf(err, 512, "cuTensorMapEncodeTiled failed for the residual [%d,%d]", M, N);
        return KDL_ECUDA;
    }
    const int tiles = ((M + BLOCK_M - 1) / BLOCK_M) * ((N + BLOCK_N - 1) / BLOCK_N);
    if (tmB64 && tiles * 2 < num_sms()) {
        const int tiles64 = ((M + BLOCK_M - 1) / BLOCK_M) * ((N + 63) / 64);
        const int grid64 = tiles64 < num_sms() ? tiles64 : num_sms();
        launch_k(k_gemm_tc<128, 64, false, 4, 2>, dim3(grid64), dim3(NUM_THREADS), CfgSmallM::SMEM_BYTES, st, tmA, *tmB64, tmOut, tmRes, bias, residual,
                 out, M, N, K, relu, ConvGeom{0, 0, 0, 0}, ldo, rev);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) {
            snprintf(err, 512, "k_gemm_tc<small-M> launch (M=%d N=%d K=%d): %s", M, N, K, cudaGetErrorString(e));
            return KDL_ECUDA;
        }
        return KDL_OK;
    }
    const int grid = tiles < num_sms() ? tiles : num_sms();
    if (K <= KDL_HBM_CFG_MAX_K)  // HBM-bound: more stores in flight matter more than ring depth
        launch_k(k_gemm_tc<128, 256, false, 3, 2, 8>, dim3(grid), dim3(CfgHbm::THREADS), CfgHbm::SMEM_BYTES, st, tmA, *tmB, tmOut, tmRes, bias, residual, out,
                 M, N, K, relu, ConvGeom{0, 0, 0, 0}, ldo, rev);
    else
        launch_k(k_gemm_tc<128, 256, false, 4, CfgStd::EPI_BUFS, CfgStd::EPI_WARPS>, dim3(grid), dim3(CfgStd::THREADS), CfgStd::SMEM_BYTES, st, tmA, *tmB, tmOut, tmRes, bias, residual, out, M, N,
                 K, relu, ConvGeom{0, 0, 0, 0}, ldo, rev);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        snprintf(err, 512, "k_gemm_tc launch (M=%d N=%d K=%d): %s", M, N, K, cudaGetErrorString(e));
        return KDL_ECUDA;
    }
    return KDL_OK;
}

// block1_conv2 as implicit GEMM: in [B,Hin,Win,32] bf16, weights via tmB (box 32 x 64 over Wt[64, 288], 64B swizzle),
// out [B,Hout,Wout,64]
inline int launch_conv3x3(const TmaEncoder &enc, const bf16 *in, const CUtensorMap *tmB, const float *bias, bf16 *out, int B, int Hin,
                          int Win, int relu, cudaStream_t st, char *err) {
    const int Cin = 32, Cout = 64;
    const int M = B * Hin * Win;  // one GEMM row per input-grid position; the 2 rightmost / bottom positions are discarded
    CUtensorMap tmA;
    if (!enc.encode_2d(&tmA, in, Cin, (uint64_t)M, Cin * 2, 32, BLOCK_M, 64)) {
        snprintf(err, 512, "cuTensorMapEncodeTiled failed for conv input [%d,%d]", M, Cin);
        return KDL_ECUDA;
    }
    const int tiles = (M + BLOCK_M - 1) / BLOCK_M;
    const int grid = tiles < num_sms() ? tiles : num_sms();
    k_gemm_tc<64, 64, true><<<grid, NUM_THREADS, CfgConv::SMEM_BYTES, st>>>(tmA, *tmB, tmA /*unused*/, tmA /*unused*/, bias, nullptr, out, M, Cout,
                                                                            9 * Cin, relu, ConvGeom{Hin, Win, Hin - 2, Win - 2}, Cout);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        snprintf(err, 512, "k_gemm_tc<conv> launch (B=%d H=%d): %s", B, Hin, cudaGetErrorString(e));
        return KDL_ECUDA;
    }
    return KDL_OK;
}

}  // namespace tc
}  // namespace kdl

// K5: pointwise-convolution GEMM on the 5th-gen tensor cores (sm_100a), hand-written PTX:
//   out[M,N] = act( A[M,K] * Wt[N,K]^T + bias[N] ) (+ residual[M,N]),  bf16 in / fp32 accumulate / bf16 out.
//
//   * operands staged by TMA (cp.async.bulk.tensor.2d, 128-byte swizzle) into a 4-stage shared-memory ring,
//   * tcgen05.mma.cta_group::1.kind::f16 (M=128, N<=256, K=16 per instruction) issued by one elected thread,
//   * fp32 accumulators live in TMEM (2 x 256 columns, double buffered so the epilogue of tile i overlaps
//     the MMAs of tile i+1), read back with tcgen05.ld.32x32b,
//   * warp-specialised persistent CTAs (one per SM): warp 0 = TMA producer, warp 1 = MMA issuer + TMEM owner,
//     warps 2..5 = epilogue (bias = folded BatchNorm shift, optional ReLU, optional residual add, bf16 store),
//   * ragged shapes (C = 728 is neither a multiple of 16 nor of any legal N tile): the K tail and the N / M
//     tails are zero-filled by TMA out-of-bounds handling and clipped in the epilogue; no padded copies.
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
constexpr int BLOCK_N = 256;
constexpr int BLOCK_K = 64;  // 64 bf16 = 128 bytes = one swizzle-128B row
constexpr int STAGES = 4;
constexpr int UMMA_K = 16;
constexpr int NUM_THREADS = 192;
constexpr uint32_t A_BYTES = BLOCK_M * BLOCK_K * 2;  // 16 KB
constexpr uint32_t B_BYTES = BLOCK_N * BLOCK_K * 2;  // 32 KB
constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
constexpr uint32_t SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
constexpr uint32_t TMEM_COLS = 512;

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
    // row-major bf16 matrix [rows, cols] with `pitch_bytes` per row; box = box_cols x box_rows, 128B swizzle,
    // out-of-bounds elements read as zero.
    bool encode_2d(CUtensorMap *tm, const void *base, uint64_t cols, uint64_t rows, uint64_t pitch_bytes, uint32_t box_cols,
                   uint32_t box_rows) const {
        cuuint64_t dims[2] = {cols, rows};
        cuuint64_t strides[1] = {pitch_bytes};
        cuuint32_t box[2] = {box_cols, box_rows};
        cuuint32_t estr[2] = {1, 1};
        CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
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

__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *tm, uint32_t bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
        "l"(tm), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
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

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
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

// Shared-memory matrix descriptor: K-major operand tile written by TMA with 128-byte swizzle.
//   bits [0,14)  start address >> 4          bits [16,30) leading byte offset >> 4 (unused for swizzled K-major: 1)
//   bits [32,46) stride byte offset >> 4 = 1024 B between 8-row groups      bits [46,48) descriptor version = 1 (sm_100)
//   bits [61,64) layout type: 2 = SWIZZLE_128B
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
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

// ---------------------------------------------------------------------------------------------------
// the kernel
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(NUM_THREADS, 1)
k_gemm_tc(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const float *__restrict__ bias,
          const bf16 *residual, bf16 *out, int M, int N, int K, int relu) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;  // swizzle-128B tiles need 1024-byte alignment
    const uint32_t bar_base = smem_base + STAGES * STAGE_BYTES;
    // barrier slots (8 B each): full[STAGES], empty[STAGES], tmem_full[2], tmem_empty[2]; then the TMEM base address
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
    auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + a); };
    auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + 2 + a); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int m_tiles = (M + BLOCK_M - 1) / BLOCK_M;
    const int n_tiles = (N + BLOCK_N - 1) / BLOCK_N;
    const int num_tiles = m_tiles * n_tiles;
    const int k_blocks = (K + BLOCK_K - 1) / BLOCK_K;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmA);
        prefetch_tmap(&tmB);
        for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(tfull_bar(a), 1); mbar_init(tempty_bar(a), 4); }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
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
                const int m0 = (tile / n_tiles) * BLOCK_M, n0 = (tile % n_tiles) * BLOCK_N;
                for (int kb = 0; kb < k_blocks; ++kb) {
                    mbar_wait(empty_bar(stage), phase ^ 1u);
                    mbar_expect_tx(full_bar(stage), STAGE_BYTES);
                    const uint32_t sa = smem_base + stage * STAGE_BYTES;
                    tma_load_2d(sa, &tmA, full_bar(stage), kb * BLOCK_K, m0);
                    tma_load_2d(sa + A_BYTES, &tmB, full_bar(stage), kb * BLOCK_K, n0);
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
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
                const int n0 = (tile % n_tiles) * BLOCK_N;
                const int acc = it & 1;
                const uint32_t acc_phase = (it >> 1) & 1;
                // shrink the instruction to the valid part of a ragged N tile (multiple of 16)
                int n_valid = min(BLOCK_N, N - n0);
                const uint32_t idesc = make_idesc(BLOCK_M, (uint32_t)((n_valid + 15) & ~15));
                mbar_wait(tempty_bar(acc), acc_phase ^ 1u);  // epilogue has drained this accumulator
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BLOCK_N;
                for (int kb = 0; kb < k_blocks; ++kb) {
                    mbar_wait(full_bar(stage), phase);
                    tc_fence_after();
                    const uint32_t sa = smem_base + stage * STAGE_BYTES;
                    const uint64_t adesc = make_smem_desc(sa);
                    const uint64_t bdesc = make_smem_desc(sa + A_BYTES);
                    const int k_left = K - kb * BLOCK_K;
                    const int ksteps = k_left >= BLOCK_K ? BLOCK_K / UMMA_K : (k_left + UMMA_K - 1) / UMMA_K;
                    for (int k = 0; k < ksteps; ++k) {
                        // advance 16 elements = 32 bytes inside the swizzle atom: +2 in the (addr >> 4) field
                        umma_bf16(d_tmem, adesc + 2u * k, bdesc + 2u * k, idesc, (kb | k) != 0);
                    }
                    umma_commit(empty_bar(stage));  // frees the smem stage when these MMAs have read it
                    if (++stage == STAGES) { stage = 0; phase ^= 1u; }
                }
                umma_commit(tfull_bar(acc));  // accumulator complete -> epilogue
            }
        }
    } else {
        // ================= epilogue: warps 2..5; TMEM lane quarter = warp % 4 =================
        const int q = warp & 3;
        int it = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
            const int m0 = (tile / n_tiles) * BLOCK_M, n0 = (tile % n_tiles) * BLOCK_N;
            const int acc = it & 1;
            const uint32_t acc_phase = (it >> 1) & 1;
            mbar_wait(tfull_bar(acc), acc_phase);
            tc_fence_after();
            const int row = m0 + q * 32 + lane;
            const bool row_ok = row < M;
            const uint32_t t_row = tmem_base + acc * BLOCK_N + ((uint32_t)(q * 32) << 16);
            const size_t row_off = (size_t)row * N;
#pragma unroll 1
            for (int c = 0; c < BLOCK_N / 32; ++c) {
                const int nc = n0 + c * 32;
                if (nc >= N) break;  // warp-uniform
                uint32_t r[32];
                tmem_ld32(t_row + c * 32, r);
                tmem_ld_wait();
                if (row_ok) {
#pragma unroll
                    for (int g = 0; g < 4; ++g) {  // 8 columns = one 16-byte bf16 vector
                        const int n = nc + g * 8;
                        if (n < N) {
                            const float4 b0 = __ldg(reinterpret_cast<const float4 *>(bias + n));
                            const float4 b1 = __ldg(reinterpret_cast<const float4 *>(bias + n + 4));
                            float v[8];
                            v[0] = __uint_as_float(r[g * 8 + 0]) + b0.x; v[1] = __uint_as_float(r[g * 8 + 1]) + b0.y;
                            v[2] = __uint_as_float(r[g * 8 + 2]) + b0.z; v[3] = __uint_as_float(r[g * 8 + 3]) + b0.w;
                            v[4] = __uint_as_float(r[g * 8 + 4]) + b1.x; v[5] = __uint_as_float(r[g * 8 + 5]) + b1.y;
                            v[6] = __uint_as_float(r[g * 8 + 6]) + b1.z; v[7] = __uint_as_float(r[g * 8 + 7]) + b1.w;
                            if (relu) {
#pragma unroll
                                for (int j = 0; j < 8; ++j) v[j] = fmaxf(v[j], 0.f);
                            }
                            if (residual) {
                                const uint4 rr = *reinterpret_cast<const uint4 *>(residual + row_off + n);
                                v[0] += __uint_as_float(rr.x << 16); v[1] += __uint_as_float(rr.x & 0xffff0000u);
                                v[2] += __uint_as_float(rr.y << 16); v[3] += __uint_as_float(rr.y & 0xffff0000u);
                                v[4] += __uint_as_float(rr.z << 16); v[5] += __uint_as_float(rr.z & 0xffff0000u);
                                v[6] += __uint_as_float(rr.w << 16); v[7] += __uint_as_float(rr.w & 0xffff0000u);
                            }
                            uint4 o;
                            o.x = pack_bf16(v[0], v[1]); o.y = pack_bf16(v[2], v[3]);
                            o.z = pack_bf16(v[4], v[5]); o.w = pack_bf16(v[6], v[7]);
                            *reinterpret_cast<uint4 *>(out + row_off + n) = o;
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty_bar(acc));
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

// ---------------------------------------------------------------------------------------------------
// host launch
// ---------------------------------------------------------------------------------------------------
inline int configure(char *err) {
    cudaError_t e = cudaFuncSetAttribute(k_gemm_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES);
    if (e != cudaSuccess) {
        snprintf(err, 512, "cudaFuncSetAttribute(k_gemm_tc, %u B smem): %s", SMEM_BYTES, cudaGetErrorString(e));
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

inline int launch_gemm(const TmaEncoder &enc, const bf16 *A, const bf16 *Wt, const CUtensorMap *tmB, const float *bias,
                       const bf16 *residual, bf16 *out, int M, int N, int K, int relu, cudaStream_t st, char *err) {
    (void)Wt;
    CUtensorMap tmA;
    if (!enc.encode_2d(&tmA, A, (uint64_t)K, (uint64_t)M, (uint64_t)K * 2, BLOCK_K, BLOCK_M)) {
        snprintf(err, 512, "cuTensorMapEncodeTiled failed for A [%d,%d]", M, K);
        return KDL_ECUDA;
    }
    const int tiles = ((M + BLOCK_M - 1) / BLOCK_M) * ((N + BLOCK_N - 1) / BLOCK_N);
    const int grid = tiles < num_sms() ? tiles : num_sms();
    k_gemm_tc<<<grid, NUM_THREADS, SMEM_BYTES, st>>>(tmA, *tmB, bias, residual, out, M, N, K, relu);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        snprintf(err, 512, "k_gemm_tc launch (M=%d N=%d K=%d): %s", M, N, K, cudaGetErrorString(e));
        return KDL_ECUDA;
    }
    return KDL_OK;
}

}  // namespace tc
}  // namespace kdl

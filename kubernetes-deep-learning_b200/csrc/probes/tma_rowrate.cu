// Probe: how fast does one SM's TMA unit move tiled boxes, as a function of the box geometry / swizzle mode?
// Persistent CTAs (one per SM); one thread issues cp.async.bulk.tensor loads into a 4-deep ring and a second thread retires
// them as soon as they land (no compute), so the time per box is the TMA pipeline's own cost.  Prints bytes/s and cycles per
// box row for: GEMM-style 2-D boxes, the depthwise kernels' 4-D halo boxes, with and without swizzle, 128- and 256-byte rows.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I.. -I../../../include probes/tma_rowrate.cu -o probes/tma_rowrate
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../gemm_tc.cuh"
#include "../dw_tma.cuh"

using namespace kdl;

constexpr int MAX_RING = 4;

struct Job {
    int ring;        // stages in flight (as many as fit in 200 KB, at most 4)
    int mode;        // 0: 2-D boxes over [rows, cols]; 1: 4-D boxes over (C, W, H, B)
    int boxes_per_cta;
    uint32_t box_bytes;
    // 2-D: row blocks of box_rows, column chunks
    int n_rowblk, n_colchunk, box_rows, box_cols;
    // 4-D
    int nch, B, x0, y0, chunk_elems;
    const void *raw;   // mode 2: contiguous blocks of box_bytes
};

__global__ void __launch_bounds__(64, 1) k_probe(const __grid_constant__ CUtensorMap tm, Job j, unsigned long long *cycles) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (tc::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t stage_bytes = (j.box_bytes + 1023u) & ~1023u;
    const int RING = j.ring;
    const uint32_t bar = base + RING * stage_bytes;
    if (threadIdx.x == 0) {
        for (int s = 0; s < RING; ++s) { tc::mbar_init(bar + 8 * s, 1); tc::mbar_init(bar + 8 * (MAX_RING + s), 1); }
        tc::fence_barrier_init();
    }
    __syncthreads();
    const long long t0 = clock64();
    if (threadIdx.x == 0) {
        int stage = 0; uint32_t ph = 0;
        for (int i = 0; i < j.boxes_per_cta; ++i) {
            const long long t = (long long)blockIdx.x + (long long)i * gridDim.x;
            tc::mbar_wait(bar + 8 * (MAX_RING + stage), ph ^ 1u);
            tc::mbar_expect_tx(bar + 8 * stage, j.box_bytes);
            if (j.mode == 0) {
                const int cc = (int)(t % j.n_colchunk), rb = (int)((t / j.n_colchunk) % j.n_rowblk);
                tc::tma_load_2d(base + stage * stage_bytes, &tm, bar + 8 * stage, cc * j.box_cols, rb * j.box_rows);
            } else if (j.mode == 2) {
                const unsigned long long nblk = (unsigned long long)j.n_rowblk;
                dwt::bulk_load_1d(base + stage * stage_bytes, (const char *)j.raw + (size_t)((unsigned long long)t % nblk) * j.box_bytes, j.box_bytes, bar + 8 * stage);
            } else {
                const int ch = (int)(t % j.nch), b = (int)((t / j.nch) % j.B);
                dwt::tma_load_4d(base + stage * stage_bytes, &tm, bar + 8 * stage, ch * j.chunk_elems, j.x0, j.y0, b);
            }
            if (++stage == RING) { stage = 0; ph ^= 1u; }
        }
    } else if (threadIdx.x == 32) {
        int stage = 0; uint32_t ph = 0;
        for (int i = 0; i < j.boxes_per_cta; ++i) {
            tc::mbar_wait(bar + 8 * stage, ph);
            tc::mbar_arrive(bar + 8 * (MAX_RING + stage));
            if (++stage == RING) { stage = 0; ph ^= 1u; }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) cycles[blockIdx.x] = (unsigned long long)(clock64() - t0);
}


// GEMM-stage pattern: each ring stage receives an A box (128 rows of the big activation matrix) and a B box (256 rows of the small,
// L2-resident weight matrix), either as two 2-D boxes of one 64-element K atom (what k_gemm_tc does) or as two 3-D boxes covering
// KA consecutive K atoms (fewer, larger TMA operations for the same bytes).
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap *tm, uint32_t bar, int c0, int c1, int c2) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(dst),
                 "l"(tm), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
                 : "memory");
}
__global__ void __launch_bounds__(64, 1) k_probe_gemm(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                                                      int ring, int KA, int kblocks, int m_blocks, int tiles_per_cta, unsigned long long *cycles) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (tc::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t a_bytes = 128u * 128u * KA, b_bytes = 256u * 128u * KA, stage_bytes = a_bytes + b_bytes;
    const uint32_t bar = base + ring * stage_bytes;
    if (threadIdx.x == 0) {
        for (int s = 0; s < ring; ++s) { tc::mbar_init(bar + 8 * s, 1); tc::mbar_init(bar + 8 * (8 + s), 1); }
        tc::fence_barrier_init();
    }
    __syncthreads();
    const long long t0 = clock64();
    const int steps = kblocks / KA;
    if (threadIdx.x == 0) {
        int stage = 0; uint32_t ph = 0;
        for (int i = 0; i < tiles_per_cta; ++i) {
            const int tile = blockIdx.x + i * gridDim.x;
            const int mb = (tile / 3) % m_blocks, nb = tile % 3;
            for (int k = 0; k < steps; ++k) {
                tc::mbar_wait(bar + 8 * (8 + stage), ph ^ 1u);
                tc::mbar_expect_tx(bar + 8 * stage, stage_bytes);
                const uint32_t sa = base + stage * stage_bytes;
                if (KA == 1) {
                    tc::tma_load_2d(sa, &tmA, bar + 8 * stage, k * 64, mb * 128);
                    tc::tma_load_2d(sa + a_bytes, &tmB, bar + 8 * stage, k * 64, nb * 256);
                } else {
                    tma_load_3d(sa, &tmA, bar + 8 * stage, 0, mb * 128, k * KA);
                    tma_load_3d(sa + a_bytes, &tmB, bar + 8 * stage, 0, nb * 256, k * KA);
                }
                if (++stage == ring) { stage = 0; ph ^= 1u; }
            }
        }
    } else if (threadIdx.x == 32) {
        int stage = 0; uint32_t ph = 0;
        for (int i = 0; i < tiles_per_cta * steps; ++i) {
            tc::mbar_wait(bar + 8 * stage, ph);
            tc::mbar_arrive(bar + 8 * (8 + stage));
            if (++stage == ring) { stage = 0; ph ^= 1u; }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) cycles[blockIdx.x] = (unsigned long long)(clock64() - t0);
}

static tc::TmaEncoder enc;

static CUtensorMap map3d_k(void *p, uint64_t rows, uint64_t pitch_bytes, int katoms_total, int box_rows, int KA) {
    // (64 elements, rows, K atoms): atom stride 128 bytes
    CUtensorMap tm;
    cuuint64_t dims[3] = {64, rows, (cuuint64_t)katoms_total};
    cuuint64_t strides[2] = {pitch_bytes, 128};
    cuuint32_t box[3] = {64, (cuuint32_t)box_rows, (cuuint32_t)KA};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc.fn(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, p, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode 3d failed (%d)\n", (int)r); exit(1); }
    return tm;
}


static void run(const char *name, const CUtensorMap &tm, Job j, int rows_per_box, double valid_frac) {
    unsigned long long *d_cyc;
    cudaMalloc(&d_cyc, 148 * 8);
    const uint32_t sb = (j.box_bytes + 1023u) & ~1023u;
    j.ring = (int)(200u * 1024u / sb) < MAX_RING ? (int)(200u * 1024u / sb) : MAX_RING;
    const uint32_t smem = j.ring * sb + 1024 + 128;
    cudaFuncSetAttribute(k_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(e0);
        k_probe<<<148, 64, smem>>>(tm, j, d_cyc);
        cudaEventRecord(e1);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("%-58s FAILED: %s\n", name, cudaGetErrorString(e)); exit(1); }
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (rep && ms < best) best = ms;
    }
    std::vector<unsigned long long> cyc(148);
    cudaMemcpy(cyc.data(), d_cyc, 148 * 8, cudaMemcpyDeviceToHost);
    double avg = 0; for (auto c : cyc) avg += (double)c; avg /= 148;
    const double boxes = 148.0 * j.boxes_per_cta;
    const double bytes = boxes * j.box_bytes * valid_frac;
    printf("%-58s %8.3f ms  %7.1f GB/s  %6.2f us/box  %5.2f clk/row  (box %u B, %d rows, ring %d)\n", name, best, bytes / best * 1e-6,
           best * 1e3 / j.boxes_per_cta, avg / j.boxes_per_cta / rows_per_box, j.box_bytes, rows_per_box, j.ring);
    cudaFree(d_cyc);
}

static CUtensorMap map4d(void *p, int C, int W, int H, int B, int ld, int bc, int bw, int bh, CUtensorMapSwizzle sw,
                         CUtensorMapL2promotion l2 = CU_TENSOR_MAP_L2_PROMOTION_L2_256B) {
    CUtensorMap tm;
    cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
    cuuint64_t strides[3] = {(cuuint64_t)ld * 2, (cuuint64_t)W * ld * 2, (cuuint64_t)H * W * ld * 2};
    cuuint32_t box[4] = {(cuuint32_t)bc, (cuuint32_t)bw, (cuuint32_t)bh, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc.fn(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, p, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw, l2,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode 4d failed (%d)\n", (int)r); exit(1); }
    return tm;
}
static CUtensorMap map2d(void *p, uint64_t cols, uint64_t rows, uint64_t pitch, int bc, int br, CUtensorMapSwizzle sw) {
    CUtensorMap tm;
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {pitch};
    cuuint32_t box[2] = {(cuuint32_t)bc, (cuuint32_t)br};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc.fn(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, p, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode 2d failed (%d)\n", (int)r); exit(1); }
    return tm;
}

int main(int argc, char **argv) {
    enc.init();
    if (!enc.ok()) { printf("no cuTensorMapEncodeTiled\n"); return 1; }
    // B = 2048 images = 1.14 GB: every box below is read once per launch, so the rates are HBM-cold (pass a smaller B to see L2 hits)
    const int B = argc > 1 ? atoi(argv[1]) : 2048, H = 19, W = 19, LD = 768;
    const size_t n = (size_t)B * H * W * LD;
    void *p;
    cudaMalloc(&p, n * 2);
    cudaMemset(p, 0, n * 2);
    const int reps = argc > 2 ? atoi(argv[2]) : 1;   // passes over the tensor (use > 1 with a small B for L2-resident rates)
    const int per = 12 * B / 148 * reps;   // boxes per CTA: one pass over the tensor for the 64-channel boxes
    {   // GEMM A operand: [M = B*H*W, K = 768], box 64 x 128
        Job j{}; j.mode = 0; j.boxes_per_cta = per; j.box_rows = 128; j.box_cols = 64; j.box_bytes = 128 * 128;
        j.n_colchunk = 12; j.n_rowblk = B * H * W / 128; j.boxes_per_cta = 12 * (B * H * W / 128) / 148 * reps;
        run("2-D  64ch x 128 rows  SWIZZLE_128B (GEMM A operand)", map2d(p, 768, (uint64_t)B * H * W, LD * 2, 64, 128, CU_TENSOR_MAP_SWIZZLE_128B), j, 128, 1.0);
        run("2-D  64ch x 128 rows  SWIZZLE_NONE", map2d(p, 768, (uint64_t)B * H * W, LD * 2, 64, 128, CU_TENSOR_MAP_SWIZZLE_NONE), j, 128, 1.0);
        Job k = j; k.box_rows = 256; k.box_bytes = 256 * 128; k.n_rowblk = B * H * W / 256; k.boxes_per_cta = j.boxes_per_cta / 2;
        run("2-D  64ch x 256 rows  SWIZZLE_128B", map2d(p, 768, (uint64_t)B * H * W, LD * 2, 64, 256, CU_TENSOR_MAP_SWIZZLE_128B), k, 256, 1.0);
        Job m = j; m.box_cols = 128; m.box_bytes = 128 * 256; m.n_colchunk = 6; m.boxes_per_cta = j.boxes_per_cta / 2;
        run("2-D 128ch x 128 rows  SWIZZLE_NONE (256-byte rows)", map2d(p, 768, (uint64_t)B * H * W, LD * 2, 128, 128, CU_TENSOR_MAP_SWIZZLE_NONE), m, 128, 1.0);
        Job q = j; q.box_cols = 256; q.box_bytes = 128 * 512; q.n_colchunk = 3; q.boxes_per_cta = j.boxes_per_cta / 4;
        run("2-D 256ch x 128 rows  SWIZZLE_NONE (512-byte rows)", map2d(p, 768, (uint64_t)B * H * W, LD * 2, 256, 128, CU_TENSOR_MAP_SWIZZLE_NONE), q, 128, 1.0);
    }
    for (uint32_t bytes : {16384u, 32768u, 49152u}) {   // the same bytes as contiguous blocks: one cp.async.bulk per block, no tensor map
        Job j{}; j.mode = 2; j.box_bytes = bytes; j.raw = p; j.n_rowblk = (int)(n * 2 / bytes); j.boxes_per_cta = (int)(n * 2 / bytes / 148) * reps;
        char nm[96]; snprintf(nm, sizeof nm, "1-D bulk copy of %u-byte contiguous blocks", bytes);
        run(nm, map2d(p, 768, (uint64_t)B * H * W, LD * 2, 64, 128, CU_TENSOR_MAP_SWIZZLE_128B), j, (int)(bytes / 128), 1.0);
    }
    {   // depthwise halo box on the 19x19 map: 64ch x 22 x 21, origin (-1,-1)
        Job j{}; j.mode = 1; j.boxes_per_cta = per; j.nch = 12; j.B = B; j.chunk_elems = 64; j.x0 = -1; j.y0 = -1; j.box_bytes = 22 * 21 * 128;
        const double valid = 361.0 / 462.0;
        run("4-D  64ch x 22 x 21  SWIZZLE_NONE  origin -1 (dw 19x19 now)", map4d(p, 728, W, H, B, LD, 64, 22, 21, CU_TENSOR_MAP_SWIZZLE_NONE), j, 462, valid);
        run("4-D  64ch x 22 x 21  SWIZZLE_128B  origin -1", map4d(p, 728, W, H, B, LD, 64, 22, 21, CU_TENSOR_MAP_SWIZZLE_128B), j, 462, valid);
        run("4-D  64ch x 22 x 21  SWIZZLE_NONE  no L2 promotion", map4d(p, 728, W, H, B, LD, 64, 22, 21, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE), j, 462, valid);
        run("4-D  64ch x 22 x 21  SWIZZLE_NONE  L2 promotion 128B", map4d(p, 728, W, H, B, LD, 64, 22, 21, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B), j, 462, valid);
        Job k = j; k.x0 = 0; k.y0 = 0; k.box_bytes = 19 * 19 * 128;
        run("4-D  64ch x 19 x 19  SWIZZLE_NONE  origin 0 (no OOB)", map4d(p, 728, W, H, B, LD, 64, 19, 19, CU_TENSOR_MAP_SWIZZLE_NONE), k, 361, 1.0);
        run("4-D  64ch x 19 x 19  SWIZZLE_128B  origin 0 (no OOB)", map4d(p, 728, W, H, B, LD, 64, 19, 19, CU_TENSOR_MAP_SWIZZLE_128B), k, 361, 1.0);
        Job m = j; m.nch = 6; m.chunk_elems = 128; m.box_bytes = 22 * 21 * 256; m.boxes_per_cta = per / 2;
        run("4-D 128ch x 22 x 21  SWIZZLE_NONE  origin -1 (256-byte rows)", map4d(p, 728, W, H, B, LD, 128, 22, 21, CU_TENSOR_MAP_SWIZZLE_NONE), m, 462, valid);
        Job r = j; r.nch = 3; r.chunk_elems = 256; r.box_bytes = 22 * 7 * 512; r.boxes_per_cta = per / 4;
        run("4-D 256ch x 22 x 7   SWIZZLE_NONE  origin -1 (512-byte rows)", map4d(p, 728, W, H, B, LD, 256, 22, 7, CU_TENSOR_MAP_SWIZZLE_NONE), r, 154, (19.0 * 6) / 154.0);
    }
    {   // the fused kernel's box on a big map: view the same memory as [B=64, 37, 37, 768]
        Job j{}; j.mode = 1; j.boxes_per_cta = per; j.nch = 12; j.B = B / 4; j.chunk_elems = 64; j.x0 = 7; j.y0 = 7; j.box_bytes = 18 * 10 * 128;
        run("4-D  64ch x 18 x 10  SWIZZLE_NONE  interior (fused sepconv)", map4d(p, 728, 37, 37, B / 4, LD, 64, 18, 10, CU_TENSOR_MAP_SWIZZLE_NONE), j, 180, 1.0);
        run("4-D  64ch x 18 x 10  SWIZZLE_128B  interior", map4d(p, 728, 37, 37, B / 4, LD, 64, 18, 10, CU_TENSOR_MAP_SWIZZLE_128B), j, 180, 1.0);
    }
    {   // GEMM stage pattern, M = B*361 rows, K = 768 (12 atoms), N = 768
        const int M = B * H * W, m_blocks = M / 128;
        void *w; cudaMalloc(&w, 768 * 768 * 2); cudaMemset(w, 0, 768 * 768 * 2);
        unsigned long long *d_cyc; cudaMalloc(&d_cyc, 148 * 8);
        const int tiles_per_cta = m_blocks * 3 / 148;
        for (int cfg = 0; cfg < 5; ++cfg) {
            const int KA = cfg == 0 ? 1 : cfg <= 2 ? 2 : cfg == 3 ? 4 : 3, ring = cfg == 0 ? 4 : cfg == 1 ? 2 : cfg == 2 ? 2 : 1, dummy = 0; (void)dummy;
            if (KA == 4 || KA == 3) { if (48 * 1024 * KA * ring > 220 * 1024) { } }
            const uint32_t smem = ring * 48u * 1024u * KA + 1024 + 256;
            if (smem > 227u * 1024u) { printf("cfg %d skipped (smem)\n", cfg); continue; }
            CUtensorMap tmA = KA == 1 ? map2d(p, 768, (uint64_t)M, LD * 2, 64, 128, CU_TENSOR_MAP_SWIZZLE_128B) : map3d_k(p, (uint64_t)M, LD * 2, 12, 128, KA);
            CUtensorMap tmB = KA == 1 ? map2d(w, 768, 768, 768 * 2, 64, 256, CU_TENSOR_MAP_SWIZZLE_128B) : map3d_k(w, 768, 768 * 2, 12, 256, KA);
            cudaFuncSetAttribute(k_probe_gemm, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
            float best = 1e30f;
            for (int rep = 0; rep < 3; ++rep) {
                cudaEventRecord(e0);
                k_probe_gemm<<<148, 64, smem>>>(tmA, tmB, ring, KA, 12, m_blocks, tiles_per_cta, d_cyc);
                cudaEventRecord(e1);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("gemm probe cfg %d FAILED: %s\n", cfg, cudaGetErrorString(e)); exit(1); }
                float ms; cudaEventElapsedTime(&ms, e0, e1);
                if (rep && ms < best) best = ms;
            }
            const double bytes = 148.0 * tiles_per_cta * 12 * 48.0 * 1024;
            const double flops = 148.0 * tiles_per_cta * 2.0 * 128 * 256 * 768;
            printf("GEMM operand feed: %d K-atoms per TMA box, ring %d (%3u KB in flight): %7.3f ms  %7.1f GB/s = %5.1f B/clk/SM @1.85GHz  -> feeds %6.0f TFLOP/s\n",
                   KA, ring, ring * 48u * KA, best, bytes / best * 1e-6, bytes / best * 1e-6 / 148 / 1.85, flops / best * 1e-9);
        }
        cudaFree(w); cudaFree(d_cyc);
    }
    cudaFree(p);
    return 0;
}

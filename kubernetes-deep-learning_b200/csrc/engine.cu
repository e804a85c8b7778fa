// kdl engine: the C ABI of include/kdl.h on top of the kernels in kernels.cuh / gemm_tc.cuh.
// One KdlModel = one GPU replica of the clothing-model graph (convert.py:4-6) with its weights folded and
// resident, an activation arena sized for max_batch, and the kernel launch sequence of the forward pass that
// TF-Serving would run for model_server.py:55.
#include <cuda.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include "../../include/kdl.h"
#include "gemm_tc.cuh"
#include "conv_tc.cuh"
#include "conv1_tc.cuh"
#include "dw_tma.cuh"
#include "pool_tma.cuh"
#include "sepconv_tc.cuh"
#include "sepconv_mid.cuh"
#include "sepconv_pair.cuh"
#include "sepconv_trio.cuh"
#include "kernels.cuh"

using namespace kdl;

static thread_local char g_err[512] = "";

static void set_err(char *dst, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(dst, 512, fmt, ap);
    va_end(ap);
    if (dst != g_err) { strncpy(g_err, dst, 511); g_err[511] = 0; }
}

#define CK(call)                                                                                         \
    do {                                                                                                 \
        cudaError_t e_ = (call);                                                                         \
        if (e_ != cudaSuccess) {                                                                         \
            set_err(err, "%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_));           \
            return KDL_ECUDA;                                                                            \
        }                                                                                                \
    } while (0)

static const float BN_EPS = 1e-3f;  // Keras BatchNormalization(epsilon=1e-3) used by applications.Xception

static inline int conv_valid(int n, int k, int s) { return (n - k) / s + 1; }
static inline int same_out(int n, int s) { return (n + s - 1) / s; }

// ---------------------------------------------------------------------------------------------------
// canonical weight order (mirror of kubernetes-deep-learning_b200/arch.py weight_spec)
// ---------------------------------------------------------------------------------------------------
struct SepSpec { int block, idx, cin, cout; };

static std::vector<SepSpec> sep_specs() {
    std::vector<SepSpec> v;
    v.push_back({2, 1, 64, 128}); v.push_back({2, 2, 128, 128});
    v.push_back({3, 1, 128, 256}); v.push_back({3, 2, 256, 256});
    v.push_back({4, 1, 256, 728}); v.push_back({4, 2, 728, 728});
    for (int b = 5; b <= 12; ++b) for (int j = 1; j <= 3; ++j) v.push_back({b, j, 728, 728});
    v.push_back({13, 1, 728, 728}); v.push_back({13, 2, 728, 1024});
    v.push_back({14, 1, 1024, 1536}); v.push_back({14, 2, 1536, 2048});
    return v;
}
struct ResSpec { int block, cin, cout; };
static const ResSpec RES_SPECS[4] = {{2, 64, 128}, {3, 128, 256}, {4, 256, 728}, {13, 728, 1024}};

extern "C" size_t kdl_weight_count(int32_t head_inner) {
    size_t n = 27 * 32 + 4 * 32 + 288 * 64 + 4 * 64;
    for (auto &r : RES_SPECS) n += (size_t)r.cin * r.cout + 4 * r.cout;
    for (auto &s : sep_specs()) n += 9 * (size_t)s.cin + (size_t)s.cin * s.cout + 4 * s.cout;
    if (head_inner > 0) n += 2048 * (size_t)head_inner + head_inner + (size_t)head_inner * 10 + 10;
    else n += 2048 * 10 + 10;
    return n;
}

// ---------------------------------------------------------------------------------------------------
// model
// ---------------------------------------------------------------------------------------------------
struct GemmW {
    int K = 0, N = 0;
    void *Wkn = nullptr;   // storage T, [K][N]  (CUDA-core GEMM)
    bf16 *Wnk = nullptr;   // bf16, [N][K]       (tcgen05 GEMM, K-major B operand)
    float *bias = nullptr; // folded BN shift
    CUtensorMap tmB;       // TMA descriptor of Wnk (box 64 x 256)
    CUtensorMap tmB64;     // TMA descriptor of Wnk (box 64 x 64: small-M tiling)
    CUtensorMap tmB192;    // TMA descriptor of Wnk (box 64 x 192: the fused middle-flow kernel's two N sub-tiles)
};

struct Tap {
    std::string name;
    int relu_applied;
    int H, C;          // spatial side (1 for vectors) and channels
    void *buf = nullptr;  // storage T (or float for vectors), sized max_batch
    bool is_float;
};

struct KdlModel {
    KdlConfig cfg;
    char err[512];
    bool weights_loaded = false;
    KdlTuning tune;                  // kernel-selection switches (kdl_set_tuning; defaults in default_tuning())
    int launches = 0;
    int last_B = 0;
    int sp[7];  // spatial chain S,H1..H6
    size_t esz; // storage element size

    // weights
    float *w_conv1 = nullptr, *b_conv1 = nullptr;
    bf16 *w_conv1_tc = nullptr;       // [32 co][64 k] bf16: w_folded / 127.5, k >= 27 zero (tensor-core conv1 on raw uint8 pixels)
    float *b_conv1_tc = nullptr;      // shift - sum_k w_folded
    CUtensorMap tm_conv1;
    GemmW conv2, res[4];
    std::vector<float *> dw;      // [34] [9][C]
    std::vector<float *> dwpad;   // [34] [ceil(C/64)][9][64] zero padded (TMA depthwise kernel)
    std::vector<GemmW> pw;        // [34]
    float *W1 = nullptr, *b1 = nullptr, *W2 = nullptr, *b2 = nullptr;
    std::vector<void *> owned;    // every device allocation (freed in destroy)

    // arena
    void *X = nullptr, *T = nullptr, *U = nullptr, *R = nullptr, *S = nullptr;
    uint8_t *u8img = nullptr;     // resized uint8 batch (fused path)
    uint8_t *d_src = nullptr;     // staging for *_host calls
    size_t d_src_cap = 0;
    float *d_in_f32 = nullptr;    // staging for kdl_forward_host
    float *d_logits = nullptr;
    float *head_partial = nullptr;  // [max_batch][16][HEAD_MAXN] partial sums of the first dense layer
    ImgMeta *d_meta = nullptr;
    int32_t *d_xtab = nullptr, *d_ytab = nullptr;
    static const int NSLOT = 4;
    ImgMeta *h_meta[NSLOT] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t slot_ev[NSLOT];
    int slot = 0;
    cudaStream_t own_stream = nullptr;
    // pipelined host path (kdl_submit_u8_host / kdl_wait): H2D of call i+1 on copy_stream overlaps the kernels of call i
    cudaStream_t copy_stream = nullptr;
    uint8_t *pipe_src[2] = {nullptr, nullptr};
    size_t pipe_cap[2] = {0, 0};
    float *pipe_logits[2] = {nullptr, nullptr};
    cudaEvent_t pipe_copied[2], pipe_done[2];
    bool pipe_init = false;
    int pipe_head = 0, pipe_inflight = 0;

    // TMA descriptors of the activation buffers viewed as [rows, K] matrices are built per call site
    std::vector<Tap> taps;
    tc::TmaEncoder tma;

    // CUDA graphs of the forward pass, keyed by (batch, input pointer)
    struct GraphEntry { int B; const float *in_f32; const uint8_t *in_u8; cudaGraphExec_t exec; int launches; uint64_t last_use; };
    std::vector<GraphEntry> graphs;
    uint64_t graph_clock = 0;
    int graph_replays = 0, graph_captures = 0, graph_evictions = 0;   // kdl_graph_stats

    // per-launch CUDA-event profiling (kdl_profile_begin / kdl_profile_end)
    bool profiling = false;
    struct ProfRec { int cls; cudaEvent_t a, b; double flops, bytes; char label[40]; };
    std::vector<ProfRec> prof_recs;
    std::vector<cudaEvent_t> ev_pool;
    size_t ev_used = 0;
    double cls_flops[16], cls_bytes[16];
    int cls_launches[16];
};

enum ProfClass { PC_RESIZE = 0, PC_CONV1, PC_CONV2, PC_DW, PC_PW_TC_HBM, PC_PW_TC_TENSOR, PC_GEMM_SIMPLE, PC_SUBSAMPLE, PC_POOL, PC_HEAD, PC_SEPFUSED, PC_SEPMID, PC_COUNT };
static const char *PROF_NAMES[PC_COUNT] = {"k_resize", "k_conv1_tc|k_conv1", "conv2_implicit_gemm", "k_dw3x3", "k_gemm_tc[K<=256,hbm-bound]",
                                           "k_gemm_tc[K>=728,tensor-bound]", "k_gemm_simple", "k_subsample2", "k_maxpool_add", "k_gap_partial+k_head_finish",
                                           "k_sepconv_tc[fused dw+pw,hbm-bound]", "k_sepconv_mid[fused dw+pw,tensor-bound]"};

static cudaEvent_t prof_event(KdlModel *m) {
    if (m->ev_used == m->ev_pool.size()) {
        cudaEvent_t e;
        cudaEventCreate(&e);
        m->ev_pool.push_back(e);
    }
    return m->ev_pool[m->ev_used++];
}
static void prof_open(KdlModel *m, cudaStream_t st, int cls, double flops, double bytes, const char *fmt = nullptr, int a0 = 0,
                      int a1 = 0, int a2 = 0) {
    if (!m->profiling) return;
    KdlModel::ProfRec r{cls, prof_event(m), prof_event(m), flops, bytes, {0}};
    if (fmt) snprintf(r.label, sizeof r.label, fmt, a0, a1, a2);
    cudaEventRecord(r.a, st);
    m->prof_recs.push_back(r);
    m->cls_flops[cls] += flops; m->cls_bytes[cls] += bytes; m->cls_launches[cls]++;
}
static void prof_close(KdlModel *m, cudaStream_t st) {
    if (!m->profiling) return;
    cudaEventRecord(m->prof_recs.back().b, st);
}

template <typename T> struct Runner;

// Kernel-selection defaults (KdlTuning in kdl.h).  Programmatic dependent launch pays in the launch-latency-bound regime
// (-9 % at batch 1, -5 % at 8, -1 % at 64; at 256 +0.4 % on the device but -1.5 % end to end), hence the batch threshold.
static KdlTuning default_tuning() {
    KdlTuning t;
    memset(&t, 0, sizeof t);
    t.pdl_max_batch = 64;
    t.serpentine = 1;
    t.conv2_shifted_taps = 0;
    t.fused_sepconv = 1;
    t.conv1_tc = 1;
    t.dw_tma_store = 0;
    t.fused_midflow = 0;   // built, bit-identical, measured: 246 us per 728->728 layer at batch 256 against 165 us for k_dw_tma + k_gemm_tc
    return t;
}

static int dev_alloc(KdlModel *m, void **p, size_t bytes) {
    char *err = m->err;
    CK(cudaMalloc(p, bytes ? bytes : 16));
    m->owned.push_back(*p);
    return KDL_OK;
}

#define LAUNCH_CHECK()                                                    \
    do {                                                                  \
        m->launches++;                                                    \
        cudaError_t e_ = cudaGetLastError();                              \
        if (e_ != cudaSuccess) {                                          \
            set_err(m->err, "%s:%d launch failed: %s", __FILE__, __LINE__, cudaGetErrorString(e_)); \
            return KDL_ECUDA;                                             \
        }                                                                 \
    } while (0)

static int find_tap(KdlModel *m, const char *name) {
    for (size_t i = 0; i < m->taps.size(); ++i)
        if (m->taps[i].name == name) return (int)i;
    return -1;
}

#define RUN(x)                          \
    do {                                \
        int rc_ = (x);                  \
        if (rc_ != KDL_OK) return rc_;  \
    } while (0)

template <typename T> struct Runner {
    KdlModel *m;
    cudaStream_t st;
    int B;
    // Serpentine traversal: most activations are larger than the 126 MB L2, so a consumer that walks a tensor in the producer's
    // order finds the start already evicted and thrashes (ncu: the pointwise GEMM re-read ALL of its input from DRAM).  A kernel
    // that can walk its tiles backwards therefore runs opposite to the kernel before it: it starts on the freshest data.
    int prev_rev = 0;
    int flip() { prev_rev = m->tune.serpentine ? 1 - prev_rev : 0; return prev_rev; }
    void no_flip() { prev_rev = 0; }   // a kernel without a reverse mode ran forwards

    int alloc_tap(const char *name, int H, int C, int relu_applied, bool is_float, int *idx) {
        char *err = m->err;
        int i = find_tap(m, name);
        if (i < 0) {
            Tap t;
            t.name = name; t.relu_applied = relu_applied; t.H = H; t.C = C; t.is_float = is_float;
            size_t bytes = (size_t)m->cfg.max_batch * H * H * C * (is_float ? 4 : sizeof(T));
            CK(cudaMalloc(&t.buf, bytes));
            m->owned.push_back(t.buf);
            m->taps.push_back(t);
            i = (int)m->taps.size() - 1;
        }
        *idx = i;
        return KDL_OK;
    }

    int tap(const char *name, const void *src, int H, int C, int relu_applied, bool is_float = false) {
        if (!m->cfg.keep_taps) return KDL_OK;
        char *err = m->err;
        int i;
        int rc = alloc_tap(name, H, C, relu_applied, is_float, &i);
        if (rc) return rc;
        const size_t esz = is_float ? 4 : sizeof(T);
        const size_t pitch = (is_float ? (size_t)C : (size_t)ld(C)) * esz;   // taps are stored dense
        CK(cudaMemcpy2DAsync(m->taps[i].buf, (size_t)C * esz, src, pitch, (size_t)C * esz, (size_t)B * H * H,
                             cudaMemcpyDeviceToDevice, st));
        return KDL_OK;
    }

    int dwconv(const T *in, const float *w, const float *wpad, T *out, int H, int C, int relu_in, int nb = -1) {
        const int B = nb < 0 ? this->B : nb;
        int rows = H >= 64 ? 8 : (H >= 16 ? 5 : H);
        int vec = H * (C / 4);
        dim3 grid((vec + 127) / 128, (H + rows - 1) / rows, B);
        prof_open(m, st, PC_DW, 18.0 * B * H * H * C, 2.0 * B * H * H * C * sizeof(T) + 36.0 * C, "dw H=%d C=%d relu=%d", H, C, relu_in);
        if (use_tc()) {
            int rc = dwt::launch(m->tma, (const bf16 *)in, wpad, (bf16 *)out, B, H, H, C, ld(C), relu_in, st, m->err, flip(), m->tune.dw_tma_store);
            prof_close(m, st);
            if (rc != KDL_OK) return rc;
            m->launches++;
            return KDL_OK;
        }
        k_dw3x3<T><<<grid, 128, 0, st>>>(in, w, out, H, H, C, relu_in, rows);
        prof_close(m, st);
        LAUNCH_CHECK();
        return KDL_OK;
    }

    bool use_tc() const {
        return sizeof(T) == 2 && m->cfg.gemm_impl == KDL_GEMM_AUTO;
    }
    // Pixel pitch (elements) of a C-channel activation tensor.  The tensor-core path pads it to a multiple of 64 so that every
    // 64-channel TMA row starts 128-byte aligned: with C = 728 (1456-byte rows) each row straddles two L2 lines, which doubled
    // the L2 request count and cost the K = 728 GEMMs ~35 % (measured: 716 -> 1116 TFLOP/s at pitch 768).  Only the pitch
    // changes -- the pad lanes are never read (TMA boxes stop at C) nor written.
    int ld(int C) const { return (use_tc() && C > 64) ? ((C + 63) / 64) * 64 : C; }

    // out[M,N] = act(A*W + bias) (+res).  A rows through RowMap (mode 0 = dense).
    int gemm(const T *A, const GemmW &g, const T *resid, T *out, int M, int relu, RowMap rm) {
        const double flops = 2.0 * M * g.N * g.K;
        const double bytes = ((double)M * g.K + (double)M * g.N * (resid ? 2 : 1) + (double)g.K * g.N) * sizeof(T) + 4.0 * g.N;
        if (use_tc() && rm.mode == 0 && tc::shape_supported(M, g.N, g.K)) {
            prof_open(m, st, g.K >= 512 ? PC_PW_TC_TENSOR : PC_PW_TC_HBM, flops, bytes, "gemm M=%d N=%d K=%d", M, g.N, g.K);
            int rc = tc::launch_gemm(m->tma, (const bf16 *)A, &g.tmB, g.bias, (const bf16 *)resid, (bf16 *)out,
                                     M, g.N, g.K, relu, st, m->err, ld(g.K), ld(g.N), &g.tmB64, flip());
            prof_close(m, st);
            if (rc != KDL_OK) return rc;
            m->launches++;
            return KDL_OK;
        }
        dim3 grid((g.N + GS_BN - 1) / GS_BN, (M + GS_BM - 1) / GS_BM);
        prof_open(m, st, rm.mode == 2 ? PC_CONV2 : PC_GEMM_SIMPLE, flops, bytes);
        k_gemm_simple<T><<<grid, 256, 0, st>>>(A, (const T *)g.Wkn, g.bias, resid, out, M, g.N, g.K, relu, rm);
        prof_close(m, st);
        LAUNCH_CHECK();
        return KDL_OK;
    }

    // depthwise -> pointwise of one SeparableConv2D as two kernels (T holds the depthwise output)
    int sepconv(const T *in, int si, T *out, int H, int relu_in, int relu_out) {
        const GemmW &g = m->pw[si];
        const RowMap dense{0, 0, 0, 0, 0, 0};
        RUN(dwconv(in, m->dw[si], m->dwpad[si], (T *)m->T, H, g.K, relu_in));
        return gemm((T *)m->T, g, nullptr, out, B * H * H, relu_out, dense);
    }

    bool can_fuse(int si) const { return use_tc() && m->tune.fused_sepconv && sepc::supported(m->pw[si].K, m->pw[si].N); }
    // fused depthwise+pointwise (sepconv_tc.cuh); in and out must be different buffers (tiles read a halo)
    int sepconv_fused(const T *in, int si, T *out, int H, int relu_in, int relu_out) {
        const GemmW &g = m->pw[si];
        const double px = (double)B * H * H;
        prof_open(m, st, PC_SEPFUSED, 2.0 * px * (9.0 * g.K + (double)g.K * g.N), (px * g.K + px * g.N + (double)g.K * g.N) * 2.0 + 36.0 * g.K,
                  "sepfused H=%d K=%d N=%d", H, g.K, g.N);
        no_flip();
        int rc = sepc::launch(m->tma, (const bf16 *)in, m->dwpad[si], &g.tmB, g.bias, (bf16 *)out, B, H, H, g.K, g.N, ld(g.K), ld(g.N),
                              relu_in, relu_out, st, m->err, si);
        prof_close(m, st);
        if (rc != KDL_OK) return rc;
        m->launches++;
        return KDL_OK;
    }

    // middle-flow shape (K, N <= 768 at W <= 20): fused depthwise + pointwise with flat-pixel tiles (sepconv_mid.cuh); resid may alias out
    bool can_fuse_mid(int si, int H) const {
        return use_tc() && m->tune.fused_midflow && sepm::supported(H, H, m->pw[si].K, m->pw[si].N);
    }
    // fused_midflow == 2: the CTA-pair kernel (sepconv_pair.cuh) where it applies, k_sepconv_mid elsewhere
    bool use_pair(int si, int H) const { return m->tune.fused_midflow == 2 && sepp::supported(H, H, m->pw[si].K, m->pw[si].N); }
    // fused_midflow == 3: clusters of three CTAs, 256 output columns each (sepconv_trio.cuh)
    bool use_trio(int si, int H) const { return m->tune.fused_midflow == 3 && sep3::supported(H, H, m->pw[si].K, m->pw[si].N); }
    int sepconv_mid(const T *in, int si, const T *resid, T *out, int H, int relu_in, int relu_out) {
        const GemmW &g = m->pw[si];
        const double px = (double)B * H * H;
        prof_open(m, st, PC_SEPMID, 2.0 * px * (9.0 * g.K + (double)g.K * g.N),
                  (px * g.K + px * g.N * (resid ? 2 : 1) + (double)g.K * g.N) * 2.0 + 36.0 * g.K, "sepmid H=%d K=%d N=%d", H, g.K, g.N);
        no_flip();
        int rc = use_trio(si, H)
                     ? sep3::launch(m->tma, (const bf16 *)in, m->dwpad[si], &g.tmB, g.bias, (const bf16 *)resid, (bf16 *)out, B, H, H, g.K, g.N,
                                    ld(g.K), ld(g.N), relu_in, relu_out, st, m->err)
                 : use_pair(si, H)
                     ? sepp::launch(m->tma, (const bf16 *)in, m->dwpad[si], &g.tmB192, g.bias, (const bf16 *)resid, (bf16 *)out, B, H, H, g.K,
                                    g.N, ld(g.K), ld(g.N), relu_in, relu_out, st, m->err)
                     : sepm::launch(m->tma, (const bf16 *)in, m->dwpad[si], &g.tmB192, g.bias, (const bf16 *)resid, (bf16 *)out, B, H, H, g.K,
                                    g.N, ld(g.K), ld(g.N), relu_in, relu_out, st, m->err);
        prof_close(m, st);
        if (rc != KDL_OK) return rc;
        m->launches++;
        return KDL_OK;
    }

    // Conv2D(1x1, strides 2, 'same') + BN : residual projection.  in: [B,H,H,Cin] -> out [B,Ho,Ho,Cout]
    int projection(const T *in, const GemmW &g, T *out, int H, int Ho, bool have_sub = false) {
        int M = B * Ho * Ho;
        if (use_tc() && have_sub) return gemm((const T *)m->S, g, nullptr, out, M, 0, RowMap{0, 0, 0, 0, 0, 0});
        if (use_tc() && tc::shape_supported(M, g.N, g.K)) {
            long long total = (long long)M * (g.K / 4);
            prof_open(m, st, PC_SUBSAMPLE, 0.0, 2.0 * M * g.K * sizeof(T));
            launch_k(k_subsample2<T>, dim3((unsigned)((total + 255) / 256)), dim3(256), 0, st, in, (T *)m->S, B, H, H, g.K, ld(g.K), Ho, Ho);
            prof_close(m, st);
            LAUNCH_CHECK();
            return gemm((const T *)m->S, g, nullptr, out, M, 0, RowMap{0, 0, 0, 0, 0, 0});
        }
        return gemm(in, g, nullptr, out, M, 0, RowMap{1, H, H, Ho, Ho, g.K});
    }

    // sub != nullptr: also write the even-index pixels of the result compactly (input of the next strided projection)
    int pool_add(const T *in, const T *resid, T *out, T *sub, int H, int C, int Ho) {
        int pad_total = (Ho - 1) * 2 + 3 - H;
        if (pad_total < 0) pad_total = 0;
        int pad_lo = pad_total / 2;  // TF SAME: extra padding goes to the end
        prof_open(m, st, PC_POOL, 0.0, ((double)B * H * H * C + 2.0 * B * Ho * Ho * C) * sizeof(T), "pool H=%d C=%d", H, C);
        if (use_tc()) {
            no_flip();
            int rc = poolt::launch(m->tma, (const bf16 *)in, (const bf16 *)resid, (bf16 *)out, (bf16 *)sub, B, H, H, C, ld(C), Ho, Ho, pad_lo,
                                   st, m->err);
            prof_close(m, st);
            if (rc != KDL_OK) return rc;
            m->launches++;
            return KDL_OK;
        }
        long long total = (long long)B * Ho * Ho * (C / 8);
        k_maxpool_add<T><<<(unsigned)((total + 255) / 256), 256, 0, st>>>(in, resid, out, B, H, H, C, ld(C), Ho, Ho, pad_lo);
        prof_close(m, st);
        LAUNCH_CHECK();
        return KDL_OK;
    }


    // in_f32: serving_default tensor, or in_u8: resized uint8 batch (exactly one non-null)
    int forward(const float *in_f32, const uint8_t *in_u8, float *d_logits) {
        const int *sp = m->sp;
        T *X = (T *)m->X, *Tb = (T *)m->T, *U = (T *)m->U, *R = (T *)m->R;
        const RowMap dense{0, 0, 0, 0, 0, 0};
        char name[64];
        // ---- block1 ----
        {
            long long threads = (long long)B * sp[1] * ((sp[1] + 1) / 2);  // one thread per pair of output pixels
            unsigned grid = (unsigned)((threads + 127) / 128);
            const unsigned persistent = (unsigned)tc::num_sms() * 8;    // k_conv1 loops over pixel pairs (4 CTAs of 128 threads per SM, 2 waves)
            if (grid > persistent) grid = persistent;
            prof_open(m, st, PC_CONV1, 2.0 * 27 * 32 * B * sp[1] * sp[1],
                      (double)B * sp[0] * sp[0] * 3 * (in_f32 ? 4 : 1) + (double)B * sp[1] * sp[1] * 32 * sizeof(T));
            if (in_f32) launch_k(k_conv1<float, T>, dim3(grid), dim3(128), 0, st, in_f32, m->w_conv1, m->b_conv1, U, B, sp[0], sp[1]);
            else if (use_tc() && m->tune.conv1_tc && m->w_conv1_tc) {   // raw uint8 pixels are exact in bf16: tensor-core implicit GEMM
                int rc = c1tc::launch(in_u8, &m->tm_conv1, m->b_conv1_tc, (bf16 *)U, B, sp[0], sp[1], st, m->err);
                if (rc != KDL_OK) return rc;
            } else launch_k(k_conv1<uint8_t, T>, dim3(grid), dim3(128), 0, st, in_u8, m->w_conv1, m->b_conv1, U, B, sp[0], sp[1]);
            prof_close(m, st);
            LAUNCH_CHECK();
            RUN(tap("block1_conv1_act", U, sp[1], 32, 1));
            // block1_conv2 as an implicit GEMM (K = 9 taps x 32 channels)
            if (use_tc()) {
                prof_open(m, st, PC_CONV2, 2.0 * B * sp[2] * sp[2] * 288.0 * 64.0,
                          ((double)B * sp[1] * sp[1] * 32 + (double)B * sp[2] * sp[2] * 64 + 288.0 * 64) * 2.0);
                int rc = (convtc::geometry_supported(sp[1]) && !m->tune.conv2_shifted_taps)
                             ? convtc::launch(m->tma, (const bf16 *)U, &m->conv2.tmB, m->conv2.bias, (bf16 *)X, (bf16 *)m->S, B, sp[1], sp[1], 1, st, m->err)
                             : tc::launch_conv3x3(m->tma, (const bf16 *)U, &m->conv2.tmB, m->conv2.bias, (bf16 *)X, B, sp[1], sp[1], 1, st, m->err);
                prof_close(m, st);
                if (rc != KDL_OK) return rc;
                m->launches++;
            } else {
                RUN(gemm(U, m->conv2, nullptr, X, B * sp[2] * sp[2], 1, RowMap{2, sp[1], sp[1], sp[2], sp[2], 32}));
            }
            RUN(tap("block1_conv2_act", X, sp[2], 64, 1));
        }
        // ---- entry/exit blocks with projection + pool: 2,3,4 then middle, then 13 ----
        int si = 0;  // sepconv index
        // S already holds the even pixels of X when the producer of X wrote them (conv2's epilogue, the pool kernels of blocks 2-3)
        bool have_sub = use_tc() && convtc::geometry_supported(sp[1]) && !m->tune.conv2_shifted_taps;
        auto entry_block = [&](int blk, int ri, int H, int Ho) -> int {
            const GemmW &rg = m->res[ri];
            RUN(projection(X, rg, R, H, Ho, have_sub));
            snprintf(name, sizeof name, "block%d_res_bn", blk);
            RUN(tap(name, R, Ho, rg.N, 0));
            // sepconv1: ReLU on load except block2 (input is already post-ReLU there)
            if (can_fuse(si)) RUN(sepconv_fused(X, si, U, H, blk != 2, 1));
            else if (can_fuse_mid(si, H)) RUN(sepconv_mid(X, si, nullptr, U, H, blk != 2, 1));   // block13_sepconv1 (728 -> 728 at 19 x 19)
            else RUN(sepconv(X, si, U, H, blk != 2, 1));  // next sepconv's ReLU folded into this GEMM's epilogue
            snprintf(name, sizeof name, "block%d_sepconv1_bn", blk);
            RUN(tap(name, U, H, m->pw[si].N, 1));
            ++si;
            T *V = U;  // where sepconv2's output lives
            if (can_fuse(si)) {
                V = Tb;  // the fused kernel cannot run in place, and T is free when nothing is written between dw and pw
                RUN(sepconv_fused(U, si, V, H, 0, 0));
            } else {
                RUN(sepconv(U, si, U, H, 0, 0));
            }
            snprintf(name, sizeof name, "block%d_sepconv2_bn", blk);
            RUN(tap(name, V, H, m->pw[si].N, 0));
            int C = m->pw[si].N;
            ++si;
            // blocks 2 and 3 feed another strided projection: let the pool kernel emit its input (tensor-core path only)
            const bool emit_sub = use_tc() && (blk == 2 || blk == 3);
            RUN(pool_add(V, R, X, emit_sub ? (T *)m->S : nullptr, H, C, Ho));
            have_sub = emit_sub;
            snprintf(name, sizeof name, "block%d_out", blk);
            RUN(tap(name, X, Ho, C, 0));
            return KDL_OK;
        };
        RUN(entry_block(2, 0, sp[2], sp[3]));
        RUN(entry_block(3, 1, sp[3], sp[4]));
        RUN(entry_block(4, 2, sp[4], sp[5]));
        // ---- middle flow ----
        {
            const int H = sp[5], M = B * H * H;
            const bool fuse_mid = can_fuse_mid(si, H);
            for (int blk = 5; blk <= 12; ++blk) {
                if (fuse_mid) {   // X -> U -> T -> X (+X): the fused kernel reads a halo, so it never runs in place on its input
                    RUN(sepconv_mid(X, si, nullptr, U, H, 1, 1));
                    snprintf(name, sizeof name, "block%d_sepconv1_bn", blk);
                    RUN(tap(name, U, H, 728, 1));
                    ++si;
                    RUN(sepconv_mid(U, si, nullptr, Tb, H, 0, 1));
                    snprintf(name, sizeof name, "block%d_sepconv2_bn", blk);
                    RUN(tap(name, Tb, H, 728, 1));
                    ++si;
                    RUN(sepconv_mid(Tb, si, X, X, H, 0, 0));      // + residual, in place on the skip tensor
                    ++si;
                    snprintf(name, sizeof name, "block%d_out", blk);
                    RUN(tap(name, X, H, 728, 0));
                    continue;
                }
                RUN(dwconv(X, m->dw[si], m->dwpad[si], Tb, H, 728, 1));
                RUN(gemm(Tb, m->pw[si], nullptr, U, M, 1, dense));
                snprintf(name, sizeof name, "block%d_sepconv1_bn", blk);
                RUN(tap(name, U, H, 728, 1));
                ++si;
                RUN(dwconv(U, m->dw[si], m->dwpad[si], Tb, H, 728, 0));
                RUN(gemm(Tb, m->pw[si], nullptr, U, M, 1, dense));
                snprintf(name, sizeof name, "block%d_sepconv2_bn", blk);
                RUN(tap(name, U, H, 728, 1));
                ++si;
                RUN(dwconv(U, m->dw[si], m->dwpad[si], Tb, H, 728, 0));
                RUN(gemm(Tb, m->pw[si], X, X, M, 0, dense));  // + residual, in place on the skip tensor
                ++si;
                snprintf(name, sizeof name, "block%d_out", blk);
                RUN(tap(name, X, H, 728, 0));
            }
        }
        RUN(entry_block(13, 3, sp[5], sp[6]));
        // ---- block14: ReLU after BN ----
        {
            const int H = sp[6], M = B * H * H;
            RUN(dwconv(X, m->dw[si], m->dwpad[si], Tb, H, 1024, 0));
            RUN(gemm(Tb, m->pw[si], nullptr, U, M, 1, dense));
            RUN(tap("block14_sepconv1_act", U, H, 1536, 1));
            ++si;
            RUN(dwconv(U, m->dw[si], m->dwpad[si], Tb, H, 1536, 0));
            RUN(gemm(Tb, m->pw[si], nullptr, U, M, 1, dense));
            RUN(tap("block14_sepconv2_act", U, H, 2048, 1));
            ++si;
        }
        // ---- head ----
        {
            const int H = sp[6];
            float *feat = nullptr;
            if (m->cfg.keep_taps) {
                int i;
                RUN(alloc_tap("avg_pool", 1, 2048, 0, true, &i));
                feat = (float *)m->taps[i].buf;
            }
            const int hi = m->cfg.head_inner;
            prof_open(m, st, PC_HEAD, 2.0 * B * (2048.0 * (hi > 0 ? hi : 10)), (double)B * H * H * 2048 * sizeof(T));
            launch_k(k_gap_partial<T>, dim3(2048 / HEAD_SLICE, B), dim3(256), 0, st, U, H * H, 2048, hi > 0 ? m->W1 : m->W2, hi > 0 ? hi : 10, feat,
                     m->head_partial);
            m->launches++;
            launch_k(k_head_finish, dim3(B), dim3(128), 0, st, m->head_partial, 2048 / HEAD_SLICE, m->b1, hi, m->W2, m->b2, 10, d_logits);
            prof_close(m, st);
            LAUNCH_CHECK();
        }
        return KDL_OK;
    }
};

// ---------------------------------------------------------------------------------------------------
// weight loading: fold BN, pack, upload
// ---------------------------------------------------------------------------------------------------
static inline uint16_t f2bf(float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);
    u += 0x7fffu + ((u >> 16) & 1u);  // round to nearest even
    return (uint16_t)(u >> 16);
}

struct BnFold { std::vector<float> scale, shift; };

static BnFold fold_bn(const float *&p, int c) {
    const float *g = p, *be = p + c, *mu = p + 2 * c, *var = p + 3 * c;
    p += 4 * (size_t)c;
    BnFold f;
    f.scale.resize(c); f.shift.resize(c);
    for (int i = 0; i < c; ++i) {
        double s = (double)g[i] / sqrt((double)var[i] + (double)BN_EPS);
        f.scale[i] = (float)s;
        f.shift[i] = (float)((double)be[i] - (double)mu[i] * s);
    }
    return f;
}

static int upload(KdlModel *m, void **dst, const void *src, size_t bytes) {
    char *err = m->err;
    int rc = dev_alloc(m, dst, bytes);
    if (rc) return rc;
    CK(cudaMemcpy(*dst, src, bytes, cudaMemcpyHostToDevice));
    return KDL_OK;
}

// kernel [K][N] fp32 (Keras (kh,kw,cin,cout) flattened) + BN -> GemmW
static int make_gemm(KdlModel *m, GemmW &g, const float *kern, int K, int N, const BnFold *bn, const float *bias,
                     bool conv3x3 = false) {
    g.K = K; g.N = N;
    std::vector<float> w((size_t)K * N);
    for (int k = 0; k < K; ++k)
        for (int n = 0; n < N; ++n) w[(size_t)k * N + n] = kern[(size_t)k * N + n] * (bn ? bn->scale[n] : 1.f);
    std::vector<float> b(N);
    for (int n = 0; n < N; ++n) b[n] = bn ? bn->shift[n] : (bias ? bias[n] : 0.f);
    int rc = upload(m, (void **)&g.bias, b.data(), N * sizeof(float));
    if (rc) return rc;
    if (m->cfg.precision == KDL_PRECISION_F32) {
        return upload(m, &g.Wkn, w.data(), w.size() * sizeof(float));
    }
    const int Kp = ((K + 63) / 64) * 64;  // K-major rows padded to a 128-byte multiple (aligned TMA rows); pad is never read
    std::vector<uint16_t> wkn((size_t)K * N), wnk((size_t)Kp * N, 0);
    for (int k = 0; k < K; ++k)
        for (int n = 0; n < N; ++n) {
            uint16_t v = f2bf(w[(size_t)k * N + n]);
            wkn[(size_t)k * N + n] = v;
            wnk[(size_t)n * Kp + k] = v;
        }
    rc = upload(m, &g.Wkn, wkn.data(), wkn.size() * 2);
    if (rc) return rc;
    rc = upload(m, (void **)&g.Wnk, wnk.data(), wnk.size() * 2);
    if (rc) return rc;
    if (m->tma.ok()) {
        bool ok = conv3x3 ? m->tma.encode_2d(&g.tmB, g.Wnk, (uint64_t)K, (uint64_t)N, (uint64_t)Kp * 2, 32, 64, 64)
                          : m->tma.encode_2d(&g.tmB, g.Wnk, (uint64_t)K, (uint64_t)N, (uint64_t)Kp * 2, tc::BLOCK_K, tc::BLOCK_N, 128);
        if (ok && !conv3x3) ok = m->tma.encode_2d(&g.tmB64, g.Wnk, (uint64_t)K, (uint64_t)N, (uint64_t)Kp * 2, tc::BLOCK_K, 64, 128);
        if (ok && !conv3x3) ok = m->tma.encode_2d(&g.tmB192, g.Wnk, (uint64_t)K, (uint64_t)N, (uint64_t)Kp * 2, tc::BLOCK_K, sepm::NSUB, 128);
        if (!ok) {
            set_err(m->err, "cuTensorMapEncodeTiled failed for weight [%d,%d]", N, K);
            return KDL_ECUDA;
        }
    }
    return KDL_OK;
}

extern "C" int kdl_load_weights(KdlModel *m, const float *blob, size_t n_floats) {
    if (!m || !blob) { set_err(g_err, "kdl_load_weights: null argument"); return KDL_EINVAL; }
    char *err = m->err;
    if (m->weights_loaded) { set_err(err, "kdl_load_weights: weights already loaded"); return KDL_ESTATE; }
    size_t want = kdl_weight_count(m->cfg.head_inner);
    if (n_floats != want) {
        set_err(err, "kdl_load_weights: blob has %zu floats, head_inner=%d needs %zu", n_floats, m->cfg.head_inner, want);
        return KDL_EINVAL;
    }
    CK(cudaSetDevice(m->cfg.device));
    const float *p = blob;
    int rc;
    // block1_conv1 (3,3,3,32) + bn : CUDA-core kernel, fp32 weights [27][32]
    {
        const float *k = p; p += 27 * 32;
        BnFold f = fold_bn(p, 32);
        std::vector<float> w(27 * 32);
        for (int i = 0; i < 27; ++i) for (int c = 0; c < 32; ++c) w[i * 32 + c] = k[i * 32 + c] * f.scale[c];
        if ((rc = upload(m, (void **)&m->w_conv1, w.data(), w.size() * 4))) return rc;
        if ((rc = upload(m, (void **)&m->b_conv1, f.shift.data(), 32 * 4))) return rc;
        if (m->cfg.precision == KDL_PRECISION_BF16 && m->tma.ok()) {
            // x = v/127.5 - 1 folded into the operands (conv1_tc.cuh): the MMA multiplies raw pixel values, exact in bf16
            std::vector<uint16_t> wt(32 * 64, 0);
            std::vector<float> bt(32);
            for (int c = 0; c < 32; ++c) {
                double sum = 0.0;
                for (int i = 0; i < 27; ++i) {
                    const double wf = (double)k[i * 32 + c] * (double)f.scale[c];
                    sum += wf;
                    wt[c * 64 + i] = f2bf((float)(wf / 127.5));
                }
                bt[c] = (float)((double)f.shift[c] - sum);
            }
            if ((rc = upload(m, (void **)&m->w_conv1_tc, wt.data(), wt.size() * 2))) return rc;
            if ((rc = upload(m, (void **)&m->b_conv1_tc, bt.data(), 32 * 4))) return rc;
            if (!m->tma.encode_2d(&m->tm_conv1, m->w_conv1_tc, 64, 32, 128, 64, 32, 128)) {
                set_err(m->err, "cuTensorMapEncodeTiled failed for the conv1 weight tile");
                return KDL_ECUDA;
            }
        }
    }
    {
        const float *k = p; p += 288 * 64;
        BnFold f = fold_bn(p, 64);
        if ((rc = make_gemm(m, m->conv2, k, 288, 64, &f, nullptr, true))) return rc;
    }
    auto seps = sep_specs();
    m->dw.assign(seps.size(), nullptr);
    m->dwpad.assign(seps.size(), nullptr);
    m->pw.assign(seps.size(), GemmW());
    size_t si = 0;
    auto load_sep = [&](const SepSpec &s) -> int {
        const float *dwk = p; p += 9 * (size_t)s.cin;
        const float *pwk = p; p += (size_t)s.cin * s.cout;
        BnFold f = fold_bn(p, s.cout);
        int r = upload(m, (void **)&m->dw[si], dwk, 9 * (size_t)s.cin * 4);
        if (r) return r;
        {
            int nch = (s.cin + 63) / 64;
            std::vector<float> wp((size_t)nch * 9 * 64, 0.f);
            for (int t = 0; t < 9; ++t)
                for (int c = 0; c < s.cin; ++c) wp[((size_t)(c / 64) * 9 + t) * 64 + (c % 64)] = dwk[(size_t)t * s.cin + c];
            r = upload(m, (void **)&m->dwpad[si], wp.data(), wp.size() * 4);
            if (r) return r;
        }
        r = make_gemm(m, m->pw[si], pwk, s.cin, s.cout, &f, nullptr);
        ++si;
        return r;
    };
    auto load_res = [&](int ri) -> int {
        const ResSpec &r = RES_SPECS[ri];
        const float *k = p; p += (size_t)r.cin * r.cout;
        BnFold f = fold_bn(p, r.cout);
        return make_gemm(m, m->res[ri], k, r.cin, r.cout, &f, nullptr);
    };
    for (int ri = 0; ri < 3; ++ri) {
        if ((rc = load_res(ri))) return rc;
        if ((rc = load_sep(seps[si]))) return rc;
        if ((rc = load_sep(seps[si]))) return rc;
    }
    for (int i = 0; i < 24; ++i)
        if ((rc = load_sep(seps[si]))) return rc;
    if ((rc = load_res(3))) return rc;
    for (int i = 0; i < 4; ++i)
        if ((rc = load_sep(seps[si]))) return rc;
    int hi = m->cfg.head_inner;
    if (hi > 0) {
        if ((rc = upload(m, (void **)&m->W1, p, 2048 * (size_t)hi * 4))) return rc; p += 2048 * (size_t)hi;
        if ((rc = upload(m, (void **)&m->b1, p, hi * 4))) return rc; p += hi;
        if ((rc = upload(m, (void **)&m->W2, p, (size_t)hi * 10 * 4))) return rc; p += (size_t)hi * 10;
    } else {
        if ((rc = upload(m, (void **)&m->W2, p, 2048 * 10 * 4))) return rc; p += 2048 * 10;
    }
    if ((rc = upload(m, (void **)&m->b2, p, 10 * 4))) return rc; p += 10;
    if ((size_t)(p - blob) != want) { set_err(err, "internal: weight cursor mismatch"); return KDL_EINVAL; }
    m->weights_loaded = true;
    return KDL_OK;
}

// ---------------------------------------------------------------------------------------------------
// lifecycle
// ---------------------------------------------------------------------------------------------------
extern "C" int kdl_create(const KdlConfig *cfg, KdlModel **out) {
    char *err = g_err;
    if (!cfg || !out) { set_err(g_err, "kdl_create: null argument"); return KDL_EINVAL; }
    if (cfg->max_batch < 1 || cfg->input_size < 71 || cfg->input_size > 1024 || cfg->head_inner < 0 ||
        cfg->head_inner > 128 || (cfg->precision != KDL_PRECISION_F32 && cfg->precision != KDL_PRECISION_BF16)) {
        set_err(g_err, "kdl_create: bad config (max_batch=%d input_size=%d head_inner=%d precision=%d)",
                cfg->max_batch, cfg->input_size, cfg->head_inner, cfg->precision);
        return KDL_EINVAL;
    }
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    if (cfg->device < 0 || cfg->device >= ndev) {
        set_err(g_err, "kdl_create: device %d not present (%d CUDA devices)", cfg->device, ndev);
        return KDL_EINVAL;
    }
    CK(cudaSetDevice(cfg->device));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, cfg->device));
    if (prop.major != 10) {
        set_err(g_err, "kdl_create: device %d is sm_%d%d; this library is built for sm_100a (B200) only",
                cfg->device, prop.major, prop.minor);
        return KDL_ECUDA;
    }
    KdlModel *m = new (std::nothrow) KdlModel();
    if (!m) { set_err(g_err, "kdl_create: out of host memory"); return KDL_ENOMEM; }
    for (int i = 0; i < KdlModel::NSLOT; ++i) m->slot_ev[i] = nullptr;
    // from here on a failure must release what was built: CKM = CK + kdl_destroy
#define CKM(call)                                                                                        \
    do {                                                                                                 \
        cudaError_t e_ = (call);                                                                         \
        if (e_ != cudaSuccess) {                                                                         \
            set_err(g_err, "%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_));         \
            kdl_destroy(m);                                                                              \
            return KDL_ECUDA;                                                                            \
        }                                                                                                \
    } while (0)
    m->cfg = *cfg;
    m->err[0] = 0;
    err = m->err;
    int S = cfg->input_size;
    m->sp[0] = S;
    m->sp[1] = conv_valid(S, 3, 2);
    m->sp[2] = conv_valid(m->sp[1], 3, 1);
    for (int i = 3; i < 7; ++i) m->sp[i] = same_out(m->sp[i - 1], 2);
    m->esz = cfg->precision == KDL_PRECISION_F32 ? 4 : 2;
    m->tma.init();
    if (cfg->precision == KDL_PRECISION_BF16 && cfg->gemm_impl == KDL_GEMM_AUTO && !m->tma.ok()) {
        set_err(g_err, "kdl_create: cuTensorMapEncodeTiled entry point not available");
        delete m;
        return KDL_ECUDA;
    }
    const int *sp = m->sp;
    const size_t MB = cfg->max_batch;
    auto sq = [](int h) { return (size_t)h * h; };
    size_t maxel = 0;
    size_t cands[] = {sq(sp[1]) * 32, sq(sp[2]) * 128, sq(sp[3]) * 256, sq(sp[4]) * 768, sq(sp[5]) * 1024, sq(sp[6]) * 2048};
    for (size_t c : cands) maxel = c > maxel ? c : maxel;
    size_t rel = 0;
    size_t rc_[] = {sq(sp[3]) * 128, sq(sp[4]) * 256, sq(sp[5]) * 768, sq(sp[6]) * 1024};
    for (size_t c : rc_) rel = c > rel ? c : rel;
    size_t sel = 0;
    size_t sc_[] = {sq(sp[3]) * 64, sq(sp[4]) * 128, sq(sp[5]) * 256, sq(sp[6]) * 768};
    for (size_t c : sc_) sel = c > sel ? c : sel;
    int rc = KDL_OK;
    // +tile slack: TMA boxes of the last M tile may read past M rows (results are discarded)
    size_t slack = 256 * 2048;
    if ((rc = dev_alloc(m, &m->X, (MB * maxel + slack) * m->esz)) || (rc = dev_alloc(m, &m->T, (MB * maxel + slack) * m->esz)) ||
        (rc = dev_alloc(m, &m->U, (MB * maxel + slack) * m->esz)) || (rc = dev_alloc(m, &m->R, (MB * rel + slack) * m->esz)) ||
        (rc = dev_alloc(m, &m->S, (MB * sel + slack) * m->esz)) ||
        (rc = dev_alloc(m, (void **)&m->u8img, MB * sq(S) * 3 + 16)) ||
        (rc = dev_alloc(m, (void **)&m->d_logits, MB * 10 * 4)) ||
        (rc = dev_alloc(m, (void **)&m->head_partial, MB * (2048 / HEAD_SLICE) * HEAD_MAXN * 4)) ||
        (rc = dev_alloc(m, (void **)&m->d_meta, MB * sizeof(ImgMeta) * KdlModel::NSLOT)) ||
        (rc = dev_alloc(m, (void **)&m->d_xtab, MB * S * 4)) || (rc = dev_alloc(m, (void **)&m->d_ytab, MB * S * 4))) {
        set_err(g_err, "%s", m->err);
        kdl_destroy(m);
        return rc;
    }
    for (int i = 0; i < KdlModel::NSLOT; ++i) {
        CKM(cudaMallocHost((void **)&m->h_meta[i], MB * sizeof(ImgMeta)));
        CKM(cudaEventCreateWithFlags(&m->slot_ev[i], cudaEventDisableTiming));
    }
    CKM(cudaStreamCreateWithFlags(&m->own_stream, cudaStreamNonBlocking));
    m->tune = default_tuning();
    if ((rc = tc::configure(m->err)) != KDL_OK || (rc = convtc::configure(m->err)) != KDL_OK || (rc = dwt::configure(m->err)) != KDL_OK || (rc = poolt::configure(m->err)) != KDL_OK || (rc = sepc::configure(m->err)) != KDL_OK || (rc = sepm::configure(m->err)) != KDL_OK || (rc = sepp::configure(m->err)) != KDL_OK || (rc = sep3::configure(m->err)) != KDL_OK || (rc = c1tc::configure(m->err)) != KDL_OK) { set_err(g_err, "%s", m->err); kdl_destroy(m); return rc; }
#undef CKM
    *out = m;
    return KDL_OK;
}

extern "C" int kdl_destroy(KdlModel *m) {
    if (!m) return KDL_OK;
    cudaSetDevice(m->cfg.device);
    cudaDeviceSynchronize();
    for (void *p : m->owned) cudaFree(p);
    if (m->d_src) cudaFree(m->d_src);
    if (m->d_in_f32) cudaFree(m->d_in_f32);
    for (int i = 0; i < KdlModel::NSLOT; ++i) {
        if (m->h_meta[i]) cudaFreeHost(m->h_meta[i]);
        if (m->slot_ev[i]) cudaEventDestroy(m->slot_ev[i]);
    }
    if (m->own_stream) cudaStreamDestroy(m->own_stream);
    for (cudaEvent_t e : m->ev_pool) cudaEventDestroy(e);
    for (auto &g : m->graphs) if (g.exec) cudaGraphExecDestroy(g.exec);
    if (m->pipe_init) {
        for (int i = 0; i < 2; ++i) {
            if (m->pipe_src[i]) cudaFree(m->pipe_src[i]);
            if (m->pipe_logits[i]) cudaFree(m->pipe_logits[i]);
            cudaEventDestroy(m->pipe_copied[i]);
            cudaEventDestroy(m->pipe_done[i]);
        }
        cudaStreamDestroy(m->copy_stream);
    }
    delete m;
    return KDL_OK;
}

extern "C" int kdl_get_tuning(const KdlModel *m, KdlTuning *out) {
    if (!m || !out) { set_err(g_err, "kdl_get_tuning: null argument"); return KDL_EINVAL; }
    *out = m->tune;
    return KDL_OK;
}

// Takes effect from the next forward pass; cached CUDA graphs were captured under the old selection and are dropped.
static void drop_graphs(KdlModel *m);
extern "C" int kdl_set_tuning(KdlModel *m, const KdlTuning *t) {
    if (!m || !t) { set_err(g_err, "kdl_set_tuning: null argument"); return KDL_EINVAL; }
    if (t->pdl_max_batch < 0) { set_err(m->err, "kdl_set_tuning: pdl_max_batch must be >= 0"); return KDL_EINVAL; }
    cudaSetDevice(m->cfg.device);
    drop_graphs(m);
    m->tune = *t;
    return KDL_OK;
}

extern "C" const char *kdl_last_error(const KdlModel *m) { return m ? m->err : g_err; }
extern "C" int kdl_last_launch_count(const KdlModel *m) { return m ? m->launches : 0; }
extern "C" const char *kdl_version(void) {
    return "kdl 0.3 sm_100a: k_resize k_conv1 k_conv1_tc k_conv3x3_tc k_dw_tma k_sepconv_tc k_sepconv_mid k_sepconv_pair k_sepconv_trio k_pool_tma k_gemm_tc(tcgen05/TMEM/TMA) k_gap_partial k_head_finish";
}

// ---------------------------------------------------------------------------------------------------
// calls
// ---------------------------------------------------------------------------------------------------
static int check_call(KdlModel *m, int B, bool need_weights) {
    if (!m) { set_err(g_err, "null model"); return KDL_EINVAL; }
    if (B < 1 || B > m->cfg.max_batch) {
        set_err(m->err, "batch %d outside [1, max_batch=%d]", B, m->cfg.max_batch);
        return KDL_EINVAL;
    }
    if (need_weights && !m->weights_loaded) { set_err(m->err, "kdl_load_weights has not been called"); return KDL_ESTATE; }
    cudaError_t e = cudaSetDevice(m->cfg.device);
    if (e != cudaSuccess) { set_err(m->err, "cudaSetDevice: %s", cudaGetErrorString(e)); return KDL_ECUDA; }
    return KDL_OK;
}

// Everything a *_host entry point must know before it queues a copy: non-null arrays, sane geometry, and every image inside
// the `total_bytes` the caller says it owns (a read past the staging buffer would be an illegal address, which is sticky and
// poisons the CUDA context of the serving process).
static int check_host_images(KdlModel *m, const char *fn, const uint8_t *h_src, size_t total_bytes, const int64_t *offset,
                             const int32_t *h, const int32_t *w, const int64_t *pitch, int B, const float *h_out) {
    char *err = m->err;
    if (!h_src || !h_out || total_bytes == 0) { set_err(err, "%s: null buffer", fn); return KDL_EINVAL; }
    if (!offset || !h || !w || !pitch) { set_err(err, "%s: null metadata array", fn); return KDL_EINVAL; }
    for (int i = 0; i < B; ++i) {
        if (h[i] < 1 || w[i] < 1 || pitch[i] < (int64_t)w[i] * 3 || offset[i] < 0) {
            set_err(err, "%s: image %d: bad geometry h=%d w=%d pitch=%lld offset=%lld", fn, i, h[i], w[i], (long long)pitch[i],
                    (long long)offset[i]);
            return KDL_EINVAL;
        }
        const unsigned long long end = (unsigned long long)offset[i] + (unsigned long long)(h[i] - 1) * (unsigned long long)pitch[i] +
                                       (unsigned long long)w[i] * 3ull;
        if (end > (unsigned long long)total_bytes) {
            set_err(err, "%s: image %d extends past total_bytes (%llu > %zu)", fn, i, end, total_bytes);
            return KDL_EINVAL;
        }
    }
    return KDL_OK;
}

// uploads per-image metadata and builds the index tables
static int stage_meta(KdlModel *m, const int64_t *offset, const int32_t *h, const int32_t *w, const int64_t *pitch,
                      int B, cudaStream_t st, ImgMeta **d_meta_out) {
    char *err = m->err;
    if (!offset || !h || !w || !pitch) { set_err(err, "null metadata array"); return KDL_EINVAL; }
    for (int i = 0; i < B; ++i) {
        if (h[i] < 1 || w[i] < 1 || pitch[i] < (int64_t)w[i] * 3 || offset[i] < 0) {
            set_err(err, "image %d: bad geometry h=%d w=%d pitch=%lld offset=%lld", i, h[i], w[i], (long long)pitch[i],
                    (long long)offset[i]);
            return KDL_EINVAL;
        }
    }
    int s = m->slot;
    m->slot = (m->slot + 1) % KdlModel::NSLOT;
    CK(cudaEventSynchronize(m->slot_ev[s]));  // previous user of this staging slot has been copied
    for (int i = 0; i < B; ++i) m->h_meta[s][i] = ImgMeta{offset[i], pitch[i], h[i], w[i]};
    ImgMeta *dm = m->d_meta + (size_t)s * m->cfg.max_batch;
    CK(cudaMemcpyAsync(dm, m->h_meta[s], B * sizeof(ImgMeta), cudaMemcpyHostToDevice, st));
    CK(cudaEventRecord(m->slot_ev[s], st));
    k_index_tables<<<(2 * B + 63) / 64, 64, 0, st>>>(dm, B, m->cfg.input_size, m->d_xtab, m->d_ytab);
    LAUNCH_CHECK();
    *d_meta_out = dm;
    return KDL_OK;
}

extern "C" int kdl_preprocess_u8(KdlModel *m, const uint8_t *d_src, const int64_t *offset, const int32_t *h,
                                 const int32_t *w, const int64_t *pitch, int32_t B, float *d_out, void *stream) {
    int rc = check_call(m, B, false);
    if (rc) return rc;
    if (!d_src || !d_out) { set_err(m->err, "kdl_preprocess_u8: null buffer"); return KDL_EINVAL; }
    cudaStream_t st = (cudaStream_t)stream;
    m->launches = 0;
    ImgMeta *dm;
    if ((rc = stage_meta(m, offset, h, w, pitch, B, st, &dm))) return rc;
    int S = m->cfg.input_size;
    long long total = (long long)B * S * S * 3;
    prof_open(m, st, PC_RESIZE, 0.0, 5.0 * total);
    k_resize<float><<<dim3(S, B), 128, ((S * 3 + 15) & ~15) + 1024, st>>>(d_src, dm, m->d_xtab, m->d_ytab, B, S, d_out);
    prof_close(m, st);
    LAUNCH_CHECK();
    return KDL_OK;
}

static int run_forward_eager(KdlModel *m, const float *in_f32, const uint8_t *in_u8, int B, float *d_logits, cudaStream_t st) {
    struct PdlScope {  // programmatic dependent launch for the kernels of this forward pass (tensor-core path only)
        PdlScope(bool on) { pdl_enabled() = on; }
        ~PdlScope() { pdl_enabled() = false; }
    } pdl_scope(B <= m->tune.pdl_max_batch && m->cfg.precision == KDL_PRECISION_BF16 && m->cfg.gemm_impl == KDL_GEMM_AUTO && !m->profiling);
    if (m->cfg.precision == KDL_PRECISION_F32) {
        Runner<float> r{m, st, B};
        return r.forward(in_f32, in_u8, d_logits);
    }
    Runner<bf16> r{m, st, B};
    return r.forward(in_f32, in_u8, d_logits);
}

// The forward pass is a fixed sequence of ~80 launches whose arguments (arena pointers, TMA descriptors passed by value) depend
// only on (batch, input pointer).  With cfg.use_graphs the second call with a given key captures that sequence into a CUDA graph
// (on the model's own stream) and later calls replay it on the caller's stream: what matters for batch-1 latency, where the
// ~30 us of device work would otherwise sit behind ~400 us of launch overhead.  The graph writes to the model's logits buffer.
// The cache is LRU with room for every batch size a dynamic batcher can produce (max_batch + 8 keys, at least 16, at most 264):
// on overflow ONE least-recently-used entry is evicted, after a device synchronise so that no replay of it is still in flight.
static size_t graph_capacity(const KdlModel *m) {
    size_t c = (size_t)m->cfg.max_batch + 8;
    return c < 16 ? 16 : (c > 264 ? 264 : c);
}
static void drop_graphs(KdlModel *m) {
    if (m->graphs.empty()) return;
    cudaDeviceSynchronize();
    for (auto &g : m->graphs) if (g.exec) cudaGraphExecDestroy(g.exec);
    m->graphs.clear();
}
static int run_forward(KdlModel *m, const float *in_f32, const uint8_t *in_u8, int B, float *d_logits, cudaStream_t st) {
    m->last_B = B;
    char *err = m->err;
    if (!m->cfg.use_graphs || m->cfg.keep_taps || m->profiling) return run_forward_eager(m, in_f32, in_u8, B, d_logits, st);
    KdlModel::GraphEntry *ge = nullptr;
    for (auto &g : m->graphs)
        if (g.B == B && g.in_f32 == in_f32 && g.in_u8 == in_u8) { ge = &g; break; }
    if (!ge) {  // first sighting: run eagerly (also finishes any lazy initialisation), remember the key
        if (m->graphs.size() >= graph_capacity(m)) {
            size_t lru = 0;
            for (size_t i = 1; i < m->graphs.size(); ++i)
                if (m->graphs[i].last_use < m->graphs[lru].last_use) lru = i;
            if (m->graphs[lru].exec) {
                CK(cudaDeviceSynchronize());
                cudaGraphExecDestroy(m->graphs[lru].exec);
            }
            m->graphs.erase(m->graphs.begin() + (long)lru);
            m->graph_evictions++;
        }
        KdlModel::GraphEntry g;
        g.B = B; g.in_f32 = in_f32; g.in_u8 = in_u8; g.exec = nullptr; g.launches = 0; g.last_use = ++m->graph_clock;
        m->graphs.push_back(g);
        return run_forward_eager(m, in_f32, in_u8, B, d_logits, st);
    }
    ge->last_use = ++m->graph_clock;
    if (!ge->exec) {
        const int before = m->launches;
        cudaGraph_t graph = nullptr;
        CK(cudaStreamBeginCapture(m->own_stream, cudaStreamCaptureModeThreadLocal));
        int rc = run_forward_eager(m, in_f32, in_u8, B, m->d_logits, m->own_stream);
        cudaError_t e = cudaStreamEndCapture(m->own_stream, &graph);
        if (rc != KDL_OK) { if (graph) cudaGraphDestroy(graph); return rc; }
        if (e != cudaSuccess) { set_err(err, "cudaStreamEndCapture -> %s", cudaGetErrorString(e)); return KDL_ECUDA; }
        e = cudaGraphInstantiate(&ge->exec, graph, 0);
        cudaGraphDestroy(graph);
        if (e != cudaSuccess) { ge->exec = nullptr; set_err(err, "cudaGraphInstantiate -> %s", cudaGetErrorString(e)); return KDL_ECUDA; }
        ge->launches = m->launches - before;
        m->launches = before;
        m->graph_captures++;
    } else {
        m->graph_replays++;
    }
    CK(cudaGraphLaunch(ge->exec, st));
    m->launches += ge->launches;
    if (d_logits != m->d_logits)
        CK(cudaMemcpyAsync(d_logits, m->d_logits, (size_t)B * 10 * sizeof(float), cudaMemcpyDeviceToDevice, st));
    return KDL_OK;
}

extern "C" int kdl_graph_stats(const KdlModel *m, int32_t *cached, int32_t *captures, int32_t *replays, int32_t *evictions) {
    if (!m) { set_err(g_err, "kdl_graph_stats: null model"); return KDL_EINVAL; }
    if (cached) *cached = (int32_t)m->graphs.size();
    if (captures) *captures = m->graph_captures;
    if (replays) *replays = m->graph_replays;
    if (evictions) *evictions = m->graph_evictions;
    return KDL_OK;
}

extern "C" int kdl_forward(KdlModel *m, const float *d_input_8, int32_t B, float *d_dense_7, void *stream) {
    int rc = check_call(m, B, true);
    if (rc) return rc;
    if (!d_input_8 || !d_dense_7) { set_err(m->err, "kdl_forward: null buffer"); return KDL_EINVAL; }
    m->launches = 0;
    return run_forward(m, d_input_8, nullptr, B, d_dense_7, (cudaStream_t)stream);
}

extern "C" int kdl_predict_u8(KdlModel *m, const uint8_t *d_src, const int64_t *offset, const int32_t *h,
                              const int32_t *w, const int64_t *pitch, int32_t B, float *d_dense_7, void *stream) {
    int rc = check_call(m, B, true);
    if (rc) return rc;
    if (!d_src || !d_dense_7) { set_err(m->err, "kdl_predict_u8: null buffer"); return KDL_EINVAL; }
    cudaStream_t st = (cudaStream_t)stream;
    m->launches = 0;
    ImgMeta *dm;
    if ((rc = stage_meta(m, offset, h, w, pitch, B, st, &dm))) return rc;
    int S = m->cfg.input_size;
    long long total = (long long)B * S * S * 3;
    // resized uint8 stays 1 byte/px; the x/127.5-1 scaling is applied (bit-exactly) inside block1_conv1's load
    prof_open(m, st, PC_RESIZE, 0.0, 2.0 * total);
    k_resize<uint8_t><<<dim3(S, B), 128, ((S * 3 + 15) & ~15) + 1024, st>>>(d_src, dm, m->d_xtab, m->d_ytab, B, S, m->u8img);
    prof_close(m, st);
    LAUNCH_CHECK();
    return run_forward(m, nullptr, m->u8img, B, d_dense_7, st);
}

extern "C" int kdl_predict_u8_host(KdlModel *m, const uint8_t *h_src, size_t total_bytes, const int64_t *offset,
                                   const int32_t *h, const int32_t *w, const int64_t *pitch, int32_t B,
                                   float *h_dense_7) {
    int rc = check_call(m, B, true);
    if (rc) return rc;
    char *err = m->err;
    if ((rc = check_host_images(m, "kdl_predict_u8_host", h_src, total_bytes, offset, h, w, pitch, B, h_dense_7))) return rc;
    if (total_bytes > m->d_src_cap) {
        if (m->d_src) CK(cudaFree(m->d_src));
        m->d_src = nullptr; m->d_src_cap = 0;
        size_t cap = total_bytes + total_bytes / 4;
        CK(cudaMalloc((void **)&m->d_src, cap));
        m->d_src_cap = cap;
    }
    cudaStream_t st = m->own_stream;
    CK(cudaMemcpyAsync(m->d_src, h_src, total_bytes, cudaMemcpyHostToDevice, st));
    rc = kdl_predict_u8(m, m->d_src, offset, h, w, pitch, B, m->d_logits, st);
    if (rc) return rc;
    CK(cudaMemcpyAsync(h_dense_7, m->d_logits, (size_t)B * 10 * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    return KDL_OK;
}

extern "C" int kdl_wait(KdlModel *m) {
    if (!m) { set_err(g_err, "kdl_wait: null model"); return KDL_EINVAL; }
    char *err = m->err;
    if (m->pipe_inflight == 0) return KDL_OK;
    CK(cudaSetDevice(m->cfg.device));
    const int slot = (m->pipe_head - m->pipe_inflight + 4) & 1;  // oldest in-flight submit
    CK(cudaEventSynchronize(m->pipe_done[slot]));
    m->pipe_inflight--;
    return KDL_OK;
}

extern "C" int kdl_submit_u8_host(KdlModel *m, const uint8_t *h_src, size_t total_bytes, const int64_t *offset, const int32_t *h,
                                  const int32_t *w, const int64_t *pitch, int32_t B, float *h_dense_7) {
    int rc = check_call(m, B, true);
    if (rc) return rc;
    char *err = m->err;
    if ((rc = check_host_images(m, "kdl_submit_u8_host", h_src, total_bytes, offset, h, w, pitch, B, h_dense_7))) return rc;
    if (!m->pipe_init) {
        CK(cudaStreamCreateWithFlags(&m->copy_stream, cudaStreamNonBlocking));
        for (int i = 0; i < 2; ++i) {
            CK(cudaEventCreateWithFlags(&m->pipe_copied[i], cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&m->pipe_done[i], cudaEventDisableTiming));
            CK(cudaMalloc((void **)&m->pipe_logits[i], (size_t)m->cfg.max_batch * 10 * sizeof(float)));
        }
        m->pipe_init = true;
    }
    if (m->pipe_inflight == 2) { if ((rc = kdl_wait(m))) return rc; }  // both staging buffers busy: retire the oldest
    const int slot = m->pipe_head & 1;
    if (total_bytes > m->pipe_cap[slot]) {
        CK(cudaStreamSynchronize(m->own_stream));
        if (m->pipe_src[slot]) CK(cudaFree(m->pipe_src[slot]));
        m->pipe_src[slot] = nullptr; m->pipe_cap[slot] = 0;
        size_t cap = total_bytes + total_bytes / 4;
        CK(cudaMalloc((void **)&m->pipe_src[slot], cap));
        m->pipe_cap[slot] = cap;
    }
    // the staging buffer of this slot was last read by the submit two calls ago, which has been retired above
    CK(cudaMemcpyAsync(m->pipe_src[slot], h_src, total_bytes, cudaMemcpyHostToDevice, m->copy_stream));
    CK(cudaEventRecord(m->pipe_copied[slot], m->copy_stream));
    CK(cudaStreamWaitEvent(m->own_stream, m->pipe_copied[slot], 0));
    rc = kdl_predict_u8(m, m->pipe_src[slot], offset, h, w, pitch, B, m->pipe_logits[slot], m->own_stream);
    if (rc) return rc;
    CK(cudaMemcpyAsync(h_dense_7, m->pipe_logits[slot], (size_t)B * 10 * sizeof(float), cudaMemcpyDeviceToHost, m->own_stream));
    CK(cudaEventRecord(m->pipe_done[slot], m->own_stream));
    m->pipe_head++;
    m->pipe_inflight++;
    return KDL_OK;
}

extern "C" int kdl_forward_host(KdlModel *m, const float *h_input_8, int32_t B, float *h_dense_7) {
    int rc = check_call(m, B, true);
    if (rc) return rc;
    char *err = m->err;
    if (!h_input_8 || !h_dense_7) { set_err(err, "kdl_forward_host: null buffer"); return KDL_EINVAL; }
    size_t S = m->cfg.input_size;
    if (!m->d_in_f32) CK(cudaMalloc((void **)&m->d_in_f32, (size_t)m->cfg.max_batch * S * S * 3 * 4));
    cudaStream_t st = m->own_stream;
    CK(cudaMemcpyAsync(m->d_in_f32, h_input_8, (size_t)B * S * S * 3 * 4, cudaMemcpyHostToDevice, st));
    rc = kdl_forward(m, m->d_in_f32, B, m->d_logits, st);
    if (rc) return rc;
    CK(cudaMemcpyAsync(h_dense_7, m->d_logits, (size_t)B * 10 * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    return KDL_OK;
}

extern "C" int kdl_num_taps(const KdlModel *m) { return m ? (int)m->taps.size() : 0; }
extern "C" const char *kdl_tap_name(const KdlModel *m, int32_t i) {
    if (!m || i < 0 || i >= (int)m->taps.size()) return nullptr;
    return m->taps[i].name.c_str();
}

extern "C" int kdl_layer_output(KdlModel *m, const char *tapname, float *d_out, size_t capacity, int32_t shape4[4],
                                int32_t *relu_applied, void *stream) {
    if (!m || !tapname) { set_err(g_err, "kdl_layer_output: null argument"); return KDL_EINVAL; }
    char *err = m->err;
    if (!m->cfg.keep_taps) { set_err(err, "kdl_layer_output: model was created with keep_taps=0"); return KDL_ESTATE; }
    int i = find_tap(m, tapname);
    if (i < 0) { set_err(err, "kdl_layer_output: unknown tap '%s' (run a forward first)", tapname); return KDL_EINVAL; }
    const Tap &t = m->taps[i];
    size_t n = (size_t)m->last_B * t.H * t.H * t.C;
    if (shape4) { shape4[0] = m->last_B; shape4[1] = t.H; shape4[2] = t.H; shape4[3] = t.C; }
    if (relu_applied) *relu_applied = t.relu_applied;
    if (!d_out) return KDL_OK;
    if (capacity < n) { set_err(err, "kdl_layer_output: capacity %zu < %zu", capacity, n); return KDL_EINVAL; }
    CK(cudaSetDevice(m->cfg.device));
    cudaStream_t st = (cudaStream_t)stream;
    if (t.is_float) {
        CK(cudaMemcpyAsync(d_out, t.buf, n * 4, cudaMemcpyDeviceToDevice, st));
    } else if (m->cfg.precision == KDL_PRECISION_F32) {
        k_to_f32<float><<<(unsigned)((n + 255) / 256), 256, 0, st>>>((const float *)t.buf, d_out, n);
    } else {
        k_to_f32<bf16><<<(unsigned)((n + 255) / 256), 256, 0, st>>>((const bf16 *)t.buf, d_out, n);
    }
    CK(cudaGetLastError());
    return KDL_OK;
}

extern "C" int kdl_gemm_bf16(const void *d_A, const void *d_Wt, const float *d_bias, const void *d_residual, void *d_out,
                             int32_t M, int32_t N, int32_t K, int32_t relu, void *stream) {
    static thread_local tc::TmaEncoder enc;
    if (!d_A || !d_Wt || !d_bias || !d_out || M < 1) { set_err(g_err, "kdl_gemm_bf16: null argument"); return KDL_EINVAL; }
    if (!tc::shape_supported(M, N, K)) { set_err(g_err, "kdl_gemm_bf16: unsupported shape M=%d N=%d K=%d (N,K must be multiples of 8)", M, N, K); return KDL_EINVAL; }
    if (!enc.ok()) enc.init();
    if (!enc.ok()) { set_err(g_err, "kdl_gemm_bf16: cuTensorMapEncodeTiled unavailable"); return KDL_ECUDA; }
    int rc = tc::configure(g_err);
    if (rc) return rc;
    CUtensorMap tmB;
    if (!enc.encode_2d(&tmB, d_Wt, (uint64_t)K, (uint64_t)N, (uint64_t)K * 2, tc::BLOCK_K, tc::BLOCK_N, 128)) {
        set_err(g_err, "kdl_gemm_bf16: tensor map encode failed");
        return KDL_ECUDA;
    }
    return tc::launch_gemm(enc, (const bf16 *)d_A, &tmB, d_bias, (const bf16 *)d_residual, (bf16 *)d_out, M, N, K, relu & 1,
                           (cudaStream_t)stream, g_err);
}

// ---------------------------------------------------------------------------------------------------
// per-launch profiling with CUDA events on the launching stream
// ---------------------------------------------------------------------------------------------------
extern "C" int kdl_profile_begin(KdlModel *m) {
    if (!m) { set_err(g_err, "kdl_profile_begin: null model"); return KDL_EINVAL; }
    m->profiling = true;
    m->prof_recs.clear();
    m->ev_used = 0;
    for (int i = 0; i < 16; ++i) { m->cls_flops[i] = 0; m->cls_bytes[i] = 0; m->cls_launches[i] = 0; }
    return KDL_OK;
}

extern "C" int kdl_profile_end(KdlModel *m, KdlProfileRow *rows, int32_t capacity, int32_t *n_rows) {
    if (!m || !rows || !n_rows) { set_err(g_err, "kdl_profile_end: null argument"); return KDL_EINVAL; }
    char *err = m->err;
    m->profiling = false;
    CK(cudaSetDevice(m->cfg.device));
    CK(cudaDeviceSynchronize());
    double ms[16] = {0};
    for (auto &r : m->prof_recs) {
        float t = 0.f;
        CK(cudaEventElapsedTime(&t, r.a, r.b));
        ms[r.cls] += t;
    }
    if (const char *path = getenv("KDL_PROFILE_DUMP")) {  // per-launch rows for offline analysis
        if (FILE *f = fopen(path, "a")) {
            for (auto &r : m->prof_recs) {
                float t = 0.f;
                cudaEventElapsedTime(&t, r.a, r.b);
                fprintf(f, "%s;%s;%.6f;%.0f;%.0f\n", PROF_NAMES[r.cls], r.label, t, r.flops, r.bytes);
            }
            fclose(f);
        }
    }
    int n = 0;
    for (int c = 0; c < PC_COUNT && n < capacity; ++c) {
        if (!m->cls_launches[c]) continue;
        KdlProfileRow &o = rows[n++];
        memset(&o, 0, sizeof o);
        strncpy(o.name, PROF_NAMES[c], sizeof(o.name) - 1);
        o.launches = m->cls_launches[c];
        o.ms = ms[c];
        o.flops = m->cls_flops[c];
        o.bytes = m->cls_bytes[c];
    }
    *n_rows = n;
    m->prof_recs.clear();
    m->ev_used = 0;
    return KDL_OK;
}

#ifdef KDL_MID_PROF
// probe builds only (make prof): cycle counters of the last k_sepconv_mid launch, [ctas][32]
extern "C" int kdl_mid_prof_read(unsigned long long *out, int ctas) {
    return cudaMemcpyFromSymbol(out, kdl::sepm::g_mid_prof, sizeof(unsigned long long) * 32 * (size_t)ctas) == cudaSuccess ? 0 : -2;
}
// [4 launches][160 CTAs][16 counters] of the k_sepconv_tc launches of the last forward (slot = sepconv index 0..3)
extern "C" int kdl_sep_prof_read(unsigned long long *out) {
    return cudaMemcpyFromSymbol(out, kdl::sepc::g_sep_prof, sizeof(unsigned long long) * 4 * 160 * 16) == cudaSuccess ? 0 : -2;
}
#endif

#ifdef KDL_MID_PROF
// Probe (probe builds only): can a pointwise GEMM and a depthwise kernel of DIFFERENT half-batches share the SMs?  Runs, on the
// middle-flow shape (B images of 19 x 19 x 728, pitch 768), `iters` rounds of { GEMM(A) ; depthwise(B) } three ways -- the product
// kernels back to back on one stream, the slim kernels (3-stage GEMM at <= 128 registers; depthwise with two strips, ~22 KB tiles,
// 2 stages, <= 96 registers: 222 KB of shared memory and ~58 k registers together) back to back, and the slim kernels on two streams
// so that their CTAs can be co-resident -- and returns the three times per round in microseconds.
extern "C" int kdl_corun_probe(int B, int iters, float *out_us /*[3]*/) {
    using namespace kdl;
    char err[512];
    tc::TmaEncoder enc;
    enc.init();
    if (!enc.ok() || tc::configure(err) || dwt::configure(err)) return -2;
    const int H = 19, C = 728, LD = 768, M = B * H * H;
    bf16 *X[2], *T[2], *U[2], *Wt;
    float *bias, *wpad;
    for (int i = 0; i < 2; ++i) {
        cudaMalloc(&X[i], (size_t)(M + 256) * LD * 2); cudaMalloc(&T[i], (size_t)(M + 256) * LD * 2); cudaMalloc(&U[i], (size_t)(M + 256) * LD * 2);
        cudaMemset(X[i], 0x3c, (size_t)(M + 256) * LD * 2); cudaMemset(T[i], 0x3c, (size_t)(M + 256) * LD * 2);
    }
    cudaMalloc(&Wt, (size_t)768 * 768 * 2); cudaMemset(Wt, 0x3c, (size_t)768 * 768 * 2);
    cudaMalloc(&bias, 768 * 4); cudaMemset(bias, 0, 768 * 4);
    cudaMalloc(&wpad, 12 * 9 * 64 * 4); cudaMemset(wpad, 0, 12 * 9 * 64 * 4);
    CUtensorMap tmB;
    enc.encode_2d(&tmB, Wt, C, C, 768 * 2, tc::BLOCK_K, tc::BLOCK_N, 128);
    cudaFuncSetAttribute(tc::k_gemm_tc_slim, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tc::CfgSlim::SMEM_BYTES);
    cudaFuncSetAttribute(dwt::k_dw_tma_slim<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    cudaStream_t s1, s2;
    cudaStreamCreateWithFlags(&s1, cudaStreamNonBlocking); cudaStreamCreateWithFlags(&s2, cudaStreamNonBlocking);
    cudaEvent_t e0, e1, j2;
    cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventCreateWithFlags(&j2, cudaEventDisableTiming);

    auto gemm_slim = [&](int h, cudaStream_t st) {
        CUtensorMap tmA, tmOut;
        enc.encode_2d(&tmA, T[h], C, M, LD * 2, tc::BLOCK_K, tc::BLOCK_M, 128);
        enc.encode_2d(&tmOut, U[h], C, M, LD * 2, 64, 32, 128);
        const int tiles = ((M + 127) / 128) * 3;
        const int grid = tiles < tc::num_sms() ? tiles : tc::num_sms();
        tc::k_gemm_tc_slim<<<grid, 192, tc::CfgSlim::SMEM_BYTES, st>>>(tmA, tmB, tmOut, tmOut, bias, nullptr, U[h], M, C, C, 1, LD, 0);
    };
    auto dw_slim = [&](int h, cudaStream_t st) {
        dwt::Params p = dwt::make_params(B, H, H, C, LD, 0, 22528, 2);
        p.stages = 2;
        CUtensorMap tm;
        dwt::encode_input_map(enc, &tm, X[h], p);
        const int grid = p.total_tiles < tc::num_sms() ? p.total_tiles : tc::num_sms();
        dwt::k_dw_tma_slim<false><<<grid, p.consumer_threads + 32, dwt::smem_bytes(p), st>>>(tm, wpad, T[h], p);
    };
    auto timed = [&](int mode) -> float {
        for (int w = 0; w < 2 + iters; ++w) {
            if (w == 2) { cudaDeviceSynchronize(); cudaEventRecord(e0, s1); }
            if (mode == 0) {
                tc::launch_gemm(enc, T[0], &tmB, bias, nullptr, U[0], M, C, C, 1, s1, err, LD, LD);
                dwt::launch(enc, X[1], wpad, T[1], B, H, H, C, LD, 0, s1, err);
            } else if (mode == 1) {
                gemm_slim(0, s1);
                dw_slim(1, s1);
            } else {
                gemm_slim(0, s1);
                dw_slim(1, s2);
            }
        }
        if (mode == 2) { cudaEventRecord(j2, s2); cudaStreamWaitEvent(s1, j2, 0); }
        cudaEventRecord(e1, s1);
        cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        return ms * 1e3f / iters;
    };
    for (int mode = 0; mode < 3; ++mode) out_us[mode] = timed(mode);
    cudaError_t e = cudaDeviceSynchronize();
    for (int i = 0; i < 2; ++i) { cudaFree(X[i]); cudaFree(T[i]); cudaFree(U[i]); }
    cudaFree(Wt); cudaFree(bias); cudaFree(wpad);
    cudaStreamDestroy(s1); cudaStreamDestroy(s2);
    return e == cudaSuccess ? 0 : -2;
}

// Probe: does a chunk-major activation layout ([C/64][B][H][W][64], every 19x19x64 tile one contiguous 46 KB run) move the depthwise
// kernel closer to the HBM roofline than the pixel-major pitch-768 layout (128-byte pieces 1536 bytes apart)?  Same kernel, same
// bytes: (a) B images of 728 channels at pitch 768, (b) 12*B "images" of 64 channels at pitch 64.  out_us[0..1].
__global__ void k_probe_fill(unsigned short *p, size_t n, unsigned seed, float scale) {   // pseudo-random bf16 in (-scale, scale)
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        unsigned h = (unsigned)i * 2654435761u + seed;
        h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
        const float v = ((float)(h & 0xffff) / 32768.f - 1.f) * scale;
        p[i] = (unsigned short)(__float_as_uint(v) >> 16);
    }
}
__global__ void k_probe_fill_f32(float *p, size_t n, unsigned seed, float scale) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        unsigned h = (unsigned)i * 2654435761u + seed;
        h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
        p[i] = ((float)(h & 0xffff) / 32768.f - 1.f) * scale;
    }
}
static int g_probe_random = 0;
extern "C" void kdl_layout_probe_data(int random) { g_probe_random = random; }
extern "C" void kdl_dw_ablate(int mode) { kdl::dwt::g_dw_ablate = mode; }
// effective SM clock while the GEMM (slot 0) / depthwise (slot 1) / conv2 (slot 2) kernels ran since the last reset: [slot][cycles, ns, launches, -]
extern "C" int kdl_clk_meter_read(unsigned long long *out /*[16]*/, int reset) {
    cudaDeviceSynchronize();
    if (cudaMemcpyFromSymbol(out, kdl::g_clk_meter, sizeof(unsigned long long) * 16) != cudaSuccess) return -2;
    if (reset) {
        unsigned long long z[16] = {0};
        if (cudaMemcpyToSymbol(kdl::g_clk_meter, z, sizeof z) != cudaSuccess) return -2;
    }
    return 0;
}
extern "C" int kdl_layout_probe(int B, int iters, float *out_us /*[6]*/) {
    using namespace kdl;
    char err[512];
    tc::TmaEncoder enc;
    enc.init();
    if (!enc.ok() || dwt::configure(err) || tc::configure(err)) return -2;
    const int H = 19, LD = 768, C = 728, M = B * H * H;
    const size_t elems = (size_t)B * H * H * LD;
    bf16 *X[2], *T[2], *Wt;
    float *wpad, *bias;
    for (int i = 0; i < 2; ++i) {
        cudaMalloc(&X[i], elems * 2 + 65536); cudaMalloc(&T[i], elems * 2 + 65536);
        cudaMemset(X[i], 0x3c, elems * 2); cudaMemset(T[i], 0, elems * 2);
    }
    cudaMalloc(&wpad, 12 * 9 * 64 * 4); cudaMemset(wpad, 0, 12 * 9 * 64 * 4);
    cudaMalloc(&Wt, (size_t)768 * 768 * 2); cudaMemset(Wt, 0, (size_t)768 * 768 * 2);
    cudaMalloc(&bias, 768 * 4); cudaMemset(bias, 0, 768 * 4);
    if (g_probe_random) {   // realistic operands: data-dependent power is part of what a kernel costs inside the forward pass
        for (int i = 0; i < 2; ++i) {
            k_probe_fill<<<592, 256>>>((unsigned short *)X[i], elems, 17u + i, 1.f);
            k_probe_fill<<<592, 256>>>((unsigned short *)T[i], elems, 91u + i, 1.f);
        }
        k_probe_fill<<<592, 256>>>((unsigned short *)Wt, (size_t)768 * 768, 5u, 0.05f);
        k_probe_fill_f32<<<64, 256>>>(wpad, 12 * 9 * 64, 3u, 0.3f);
        k_probe_fill_f32<<<4, 256>>>(bias, 768, 7u, 0.1f);
        cudaDeviceSynchronize();
    }
    CUtensorMap tmB;
    enc.encode_2d(&tmB, Wt, C, C, 768 * 2, tc::BLOCK_K, tc::BLOCK_N, 128);
    cudaStream_t s1;
    cudaStreamCreateWithFlags(&s1, cudaStreamNonBlocking);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    int rc = 0;
    // 0: dw, pixel-major  1: dw, chunk-major  2: GEMM alone (two buffer pairs)  3: dw -> GEMM chain as in the middle flow, serpentine
    // 4: the same chain, every kernel forwards  5: GEMM alone, same buffers every time (A partly L2-resident)
    for (int mode = 0; mode < 6; ++mode) {
        int rev = 0;
        for (int w = 0; w < 2 + iters; ++w) {
            if (w == 2) { cudaDeviceSynchronize(); cudaEventRecord(e0, s1); }
            const int h = w & 1;
            if (mode == 0) rc |= dwt::launch(enc, X[h], wpad, T[h], B, H, H, C, LD, 0, s1, err);
            else if (mode == 1) rc |= dwt::launch(enc, X[h], wpad, T[h], B * 12, H, H, 64, 64, 0, s1, err);
            else if (mode == 2) rc |= tc::launch_gemm(enc, T[h], &tmB, bias, nullptr, X[h], M, C, C, 1, s1, err, LD, LD);
            else if (mode == 5) rc |= tc::launch_gemm(enc, T[0], &tmB, bias, nullptr, X[0], M, C, C, 1, s1, err, LD, LD);
            else {
                const bool serp = mode == 3;
                if (serp) rev ^= 1;
                rc |= dwt::launch(enc, X[h], wpad, T[h], B, H, H, C, LD, 0, s1, err, rev);
                if (serp) rev ^= 1;
                rc |= tc::launch_gemm(enc, T[h], &tmB, bias, nullptr, X[h ^ 1], M, C, C, 1, s1, err, LD, LD, nullptr, rev);
            }
        }
        cudaEventRecord(e1, s1);
        cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        out_us[mode] = ms * 1e3f / iters;
    }
    cudaError_t e = cudaDeviceSynchronize();
    for (int i = 0; i < 2; ++i) { cudaFree(X[i]); cudaFree(T[i]); }
    cudaFree(wpad); cudaFree(Wt); cudaFree(bias);
    cudaStreamDestroy(s1);
    return (e == cudaSuccess && rc == 0) ? 0 : -2;
}
#endif

/*
 * kdl.h -- C ABI of the B200-native replacement for the one hot path of
 * alexeygrigorev/kubernetes-deep-learning: gateway preprocessing + Xception-299 forward.
 *
 * Everything is `extern "C"`, plain pointers and sizes.  No exceptions or aborts cross this
 * boundary: every call returns 0 on success or a negative KDL_E* code, and the message is
 * available from kdl_last_error().  The reference binds none of this today (it talks gRPC to a
 * TF-Serving container); INTEGRATION.md shows the ctypes stub a maintainer would drop into
 * model_server.py in place of the grpc stub.
 *
 * Threading: a KdlModel is bound to one CUDA device and is not re-entrant -- one model per GPU per
 * host thread.  Calls taking a `stream` are asynchronous with respect to the host and ordered on
 * that stream; the *_host calls synchronise before returning.
 * Ownership: the model owns weights, the activation arena, TMA descriptors; the caller owns all
 * input / output buffers.
 */
#ifndef KDL_H_
#define KDL_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KDL_OK 0
#define KDL_EINVAL (-1)   /* bad argument (null pointer, batch > max_batch, unknown tap ...) */
#define KDL_ECUDA (-2)    /* a CUDA runtime / driver call failed; see kdl_last_error */
#define KDL_ESTATE (-3)   /* call order violated (e.g. forward before kdl_load_weights) */
#define KDL_ENOMEM (-4)   /* host or device allocation failed */

#define KDL_PRECISION_F32 0   /* fp32 storage + fp32 math on CUDA cores: the <=1e-4 "fp32 path" */
#define KDL_PRECISION_BF16 1  /* bf16 storage, fp32 accumulate, tcgen05 GEMMs: the production path */

#define KDL_GEMM_AUTO 0       /* tcgen05/TMEM/TMA GEMM for bf16, CUDA-core GEMM for f32 */
#define KDL_GEMM_SIMPLE 1     /* force the CUDA-core GEMM (debug / kernel-vs-kernel parity) */

typedef struct KdlModel KdlModel;

/* Replaces the deployment constants of tf-serving.dockerfile:2-5 (which model, which version) and the
 * module-level set-up of model_server.py:13-18 (channel, stub, preprocessor target size). */
typedef struct KdlConfig {
    int32_t device;       /* CUDA device ordinal */
    int32_t max_batch;    /* largest batch any call will pass; sizes the activation arena */
    int32_t input_size;   /* network input H=W: 299 (model_server.py:18) or e.g. 150 */
    int32_t head_inner;   /* width of the hidden Dense(+ReLU) of the clothing head; 0 = Dense(2048->10) only */
    int32_t precision;    /* KDL_PRECISION_* */
    int32_t gemm_impl;    /* KDL_GEMM_* */
    int32_t keep_taps;    /* !=0: keep a copy of every tapped layer output for kdl_layer_output (debug; slow) */
    int32_t use_graphs;   /* !=0: replay the forward pass as a CUDA graph from the 3rd call with the same (batch, input buffer) on */
} KdlConfig;

/* Kernel-selection switches of one model (no reference counterpart: TF-Serving has no such knobs; these exist so that A/B
 * measurements and the bit-identity tests of tests/test_gpu_variants.py do not depend on hidden environment variables).
 * kdl_create installs the defaults; kdl_set_tuning takes effect from the next forward pass and drops the cached CUDA graphs.
 * Except where noted, variants compute the same arithmetic in the same order: results are bit-identical whatever is selected. */
typedef struct KdlTuning {
    int32_t pdl_max_batch;       /* programmatic dependent launch for forwards of at most this many images (default 64; 0 = never) */
    int32_t serpentine;          /* !=0 (default): kernels walk their tiles in alternating directions so a consumer starts on the
                                    rows its producer wrote last (activations exceed the 126 MB L2 at large batch) */
    int32_t conv2_shifted_taps;  /* !=0: block1_conv2 as 9 shifted TMA boxes per tile instead of the staged halo (default 0) */
    int32_t fused_sepconv;       /* !=0 (default): blocks 2-3 run the fused depthwise+pointwise kernel */
    int32_t fused_midflow;       /* !=0: 728 -> 728 layers at 19 x 19 (middle flow, block13_sepconv1) run a fused depthwise+pointwise
                                    kernel with flat-pixel tiles: 1 = one CTA per (tile, N half), the depthwise tile computed twice
                                    (sepconv_mid.cuh); 2 = a CTA pair (cluster of 2) per tile, the depthwise chunks split between
                                    the two CTAs and exchanged through distributed shared memory (sepconv_pair.cuh); 3 = a cluster
                                    of three, 256 output columns and two TMEM accumulators per CTA (sepconv_trio.cuh).  Default 0:
                                    all three are bit-identical to but measured slower than k_dw_tma + k_gemm_tc on a B200
                                    (DESIGN.md §4 has the cycle breakdowns) */
    int32_t conv1_tc;            /* !=0 (default): block1_conv1 of the uint8 path runs as a tcgen05 implicit GEMM on the raw pixel
                                    values (exact in bf16; x/127.5-1 folded into bf16 weights + fp32 bias).  0: the CUDA-core kernel
                                    with fp32 weights (always used for float inputs and the fp32 path).  The ONE switch that
                                    changes bits: conv1's weights are rounded to bf16 like every other layer's */
    int32_t dw_tma_store;        /* !=0: the depthwise kernel stages each output tile in shared memory and writes it with one 4-D TMA
                                    store (2 input stages) instead of 8-byte stores from every thread (3 input stages) */
    int32_t reserved[5];         /* must be 0 */
} KdlTuning;
int kdl_get_tuning(const KdlModel *m, KdlTuning *out);
int kdl_set_tuning(KdlModel *m, const KdlTuning *t);

/* CUDA-graph cache counters (cfg.use_graphs): entries cached, captures, replays and LRU evictions so far.  The cache holds
 * max_batch + 8 keys (16..264), so a dynamic batcher cycling through every batch size keeps replaying. */
int kdl_graph_stats(const KdlModel *m, int32_t *cached, int32_t *captures, int32_t *replays, int32_t *evictions);

/* Model lifecycle.  Replaces TF-Serving loading /models/clothing-model/1 at container start
 * (tf-serving.dockerfile:4-5). */
int kdl_create(const KdlConfig *cfg, KdlModel **out);
int kdl_destroy(KdlModel *m);

/* Last error text of this model (or of the library when m == NULL).  Owned by the library. */
const char *kdl_last_error(const KdlModel *m);

/* Number of float32 values kdl_load_weights expects for a given head (canonical Keras order, see
 * kubernetes-deep-learning_b200/arch.py weight_spec). */
size_t kdl_weight_count(int32_t head_inner);

/* Replaces keras.models.load_model + tf.saved_model.save (convert.py:4-6) and TF-Serving's variable
 * restore: takes the Keras-layout fp32 arrays (HOST pointer), folds BatchNorm (eps 1e-3) into the
 * preceding conv, packs kernel-native layouts and uploads.  The caller keeps ownership of `blob`. */
int kdl_load_weights(KdlModel *m, const float *blob, size_t n_floats);

/* Replaces preprocessor.from_url(url) minus download/decode (model_server.py:53, keras-image-helper 0.0.1 ->
 * Pillow Image.resize(NEAREST) + float32 x/127.5-1): B decoded RGB uint8 HWC images of arbitrary sizes,
 * already on the device at d_src + offset[i] with row pitch pitch[i] bytes, -> float32 (B,S,S,3) in d_out
 * (S = cfg.input_size), bit-exact.  offset/h/w/pitch are HOST arrays of length B. */
int kdl_preprocess_u8(KdlModel *m, const uint8_t *d_src, const int64_t *offset, const int32_t *h,
                      const int32_t *w, const int64_t *pitch, int32_t B, float *d_out, void *stream);

/* Replaces stub.Predict on signature serving_default (model_server.py:38-43,55; guide.md:220-231):
 * input_8 float32 (B,S,S,3) in [-1,1] -> dense_7 float32 (B,10) raw logits.  Device pointers. */
int kdl_forward(KdlModel *m, const float *d_input_8, int32_t B, float *d_dense_7, void *stream);

/* apply_model without the RPC hop (model_server.py:52-56): fused preprocess + forward, device-resident
 * source images (same addressing as kdl_preprocess_u8), logits to device memory. */
int kdl_predict_u8(KdlModel *m, const uint8_t *d_src, const int64_t *offset, const int32_t *h,
                   const int32_t *w, const int64_t *pitch, int32_t B, float *d_dense_7, void *stream);

/* Same, end to end from HOST memory: copies `total_bytes` of packed source pixels (pinned memory
 * recommended) to the device, runs, copies (B,10) logits back to h_dense_7 and synchronises. */
int kdl_predict_u8_host(KdlModel *m, const uint8_t *h_src, size_t total_bytes, const int64_t *offset,
                        const int32_t *h, const int32_t *w, const int64_t *pitch, int32_t B,
                        float *h_dense_7);

/* Pipelined form of kdl_predict_u8_host for throughput: returns as soon as the work is queued; up to two submits may be in
 * flight, so the host->device copy of call i+1 (own copy stream, double-buffered staging) overlaps the kernels of call i.
 * h_src and h_dense_7 should be pinned and must stay valid until the matching kdl_wait, which retires the OLDEST submit.
 * A third submit waits for the oldest one itself.  No reference counterpart (the reference gateway is strictly one request
 * at a time, gateway.dockerfile:17). */
int kdl_submit_u8_host(KdlModel *m, const uint8_t *h_src, size_t total_bytes, const int64_t *offset,
                       const int32_t *h, const int32_t *w, const int64_t *pitch, int32_t B, float *h_dense_7);
int kdl_wait(KdlModel *m);

/* Same as kdl_forward but from/to HOST memory (the PredictRequest tensor_content / PredictResponse
 * float_val payloads of model_server.py:35-36,46-47). */
int kdl_forward_host(KdlModel *m, const float *h_input_8, int32_t B, float *h_dense_7);

/* Debug / layer-wise parity (needs cfg.keep_taps): copies the named intermediate tensor of the LAST
 * forward as float32 NHWC into d_out (capacity in floats).  shape4 = {B,H,W,C} (H=W=1 for vectors);
 * *relu_applied tells whether the stored tensor already has the next layer's ReLU folded in. */
int kdl_num_taps(const KdlModel *m);
const char *kdl_tap_name(const KdlModel *m, int32_t i);
int kdl_layer_output(KdlModel *m, const char *tap, float *d_out, size_t capacity, int32_t shape4[4],
                     int32_t *relu_applied, void *stream);

/* Per-launch timing with CUDA events on the launching stream (bench.py's roofline numbers).  Between
 * kdl_profile_begin and kdl_profile_end every kernel launched by this model is bracketed by two events;
 * kdl_profile_end synchronises and returns one row per kernel class: launches, summed device ms, and the
 * summed ALGORITHMIC flops / bytes of those launches (DESIGN.md states the per-launch formulas). */
typedef struct KdlProfileRow {
    char name[48];
    int32_t launches;
    float ms;
    double flops;
    double bytes;
} KdlProfileRow;
int kdl_profile_begin(KdlModel *m);
int kdl_profile_end(KdlModel *m, KdlProfileRow *rows, int32_t capacity, int32_t *n_rows);

/* Number of kernels the last forward/predict call launched (the bench's gpu_launches claim). */
int kdl_last_launch_count(const KdlModel *m);

/* Stand-alone pointwise GEMM entry (kernel-level parity and roofline measurement):
 * out[M,N] = act( A[M,K] * W[K,N] + bias[N] ) (+ residual[M,N]); bf16 device buffers, W given K-major
 * as Wt[N,K]; residual may be NULL.  `relu` bit 0 = apply ReLU (other bits ignored).  Runs the same k_gemm_tc the model uses. */
int kdl_gemm_bf16(const void *d_A, const void *d_Wt, const float *d_bias, const void *d_residual,
                  void *d_out, int32_t M, int32_t N, int32_t K, int32_t relu, void *stream);

/* No counterpart in the reference.  kdl_preprocess_u8 for callers that cannot allocate device memory: a request thread that
 * must stay clear of the allocator (its locks and implicit device synchronisation), a plugin living inside someone else's
 * memory budget, or code that has no KdlModel (whose arena normally holds the index tables).  Same result as
 * kdl_preprocess_u8, bit for bit; the per-image index tables live in one fixed device array built into the library instead
 * of in cudaMalloc'ed buffers, and the call performs no cudaMalloc / cudaFree / allocating library call and no host->device
 * copy (the image metadata travels as a kernel argument).
 *   - Limits: B <= KDL_STATIC_MAX_BATCH images per call, target side S <= KDL_STATIC_MAX_SIZE; larger values are refused
 *     with KDL_EINVAL and a message.  The array takes KDL_STATIC_SCRATCH_BYTES on every GPU that loads the library, until exit.
 *   - Works on the CURRENT CUDA device (d_src / d_out must belong to it).  The first call on each device resolves the array's
 *     device address (cudaGetSymbolAddress); make that first call outside any stream capture.
 *   - All calls on a device share the array and it keeps the previous call's contents.  Nothing needs clearing (each call
 *     rewrites every entry it reads), but two calls whose kernels overlap on the GPU (different streams, or threads racing on
 *     one stream) corrupt each other: serialise them on one stream or with events.
 *   - d_out must be 16-byte aligned (the kernel writes float4 vectors; kdl_preprocess_u8 accepts any alignment, this entry
 *     refuses a misaligned pointer with KDL_EINVAL rather than fault on the device).
 *   - offset/h/w/pitch are HOST arrays of length B, as for kdl_preprocess_u8.  Errors: text is copied into err (may be NULL). */
#define KDL_STATIC_MAX_BATCH 32
#define KDL_STATIC_MAX_SIZE 512
#define KDL_STATIC_SCRATCH_BYTES (1024 + 2 * KDL_STATIC_MAX_BATCH * KDL_STATIC_MAX_SIZE * 4)
int kdl_preprocess_u8_static(const uint8_t *d_src, const int64_t *offset, const int32_t *h, const int32_t *w,
                             const int64_t *pitch, int32_t B, int32_t S, float *d_out, void *stream, char *err,
                             size_t err_len);

/* Library / build info string (arch, kernels compiled in). */
const char *kdl_version(void);

#ifdef __cplusplus
}
#endif
#endif /* KDL_H_ */

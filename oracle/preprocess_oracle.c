/*
 * ORACLE — TEST INFRASTRUCTURE ONLY. Not part of the product path.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this.
 *
 * CPU restatement of the gateway preprocessing the reference performs at
 * model_server.py:18 (create_preprocessor('xception', target_size=(299, 299))) and
 * model_server.py:53 (preprocessor.from_url(url)).  The arithmetic itself lives in two
 * un-vendored dependencies pinned by the reference:
 *   - Pillow==8.1.0 (Pipfile.lock:314-349): Image.resize(size, NEAREST) -> libImaging/Geometry.c
 *     ImagingScaleAffine: a = n_in / n_out (double); xo = a*0.5; for each output x: idx = (int)xo; xo += a.
 *     Rows and columns use the same rule independently.
 *   - keras-image-helper==0.0.1 (Pipfile.lock:214-220): np.array(img, dtype='float32'), then the
 *     "tf"-mode Xception scaling  x /= 127.5 ; x -= 1.0  (two separate fp32 roundings, in place).
 *
 * Parity pin: checked in tests/test_oracle_preprocess.py against Pillow itself (importable in this image)
 * on >1400 (n_in, n_out) pairs and against the media/image2.png known-answer hashes.
 */
#include <stdint.h>
#include <stddef.h>

/* Pillow's COORD(): negative -> -1, otherwise C truncation. */
static int coord(double v) { return v < 0.0 ? -1 : (int)v; }

/* Index table for one axis: tab[i] = source index feeding output i. Returns 0 on success. */
int kdl_oracle_index_table(int n_in, int n_out, int32_t *tab)
{
    if (n_in <= 0 || n_out <= 0 || !tab) return -1;
    double a = (double)n_in / (double)n_out;
    double o = a * 0.5;                 /* a[2] (= box origin 0) + a[0]*0.5 */
    for (int i = 0; i < n_out; ++i) {
        int s = coord(o);
        if (s < 0) s = 0;               /* cannot happen for n_in>0; kept for safety */
        if (s >= n_in) s = n_in - 1;    /* cannot happen either (o < n_in) */
        tab[i] = s;
        o += a;                         /* accumulated, NOT i*a: this is what pins bit-exactness */
    }
    return 0;
}

/* uint8 HWC (h, w, 3) with row pitch (bytes) -> uint8 HWC (out_h, out_w, 3), nearest neighbour. */
int kdl_oracle_resize_u8(const uint8_t *src, int h, int w, int64_t pitch,
                         uint8_t *dst, int out_h, int out_w,
                         int32_t *xtab_scratch, int32_t *ytab_scratch)
{
    if (!src || !dst) return -1;
    if (kdl_oracle_index_table(w, out_w, xtab_scratch)) return -1;
    if (kdl_oracle_index_table(h, out_h, ytab_scratch)) return -1;
    for (int y = 0; y < out_h; ++y) {
        const uint8_t *row = src + (size_t)ytab_scratch[y] * (size_t)pitch;
        uint8_t *o = dst + (size_t)y * out_w * 3;
        for (int x = 0; x < out_w; ++x) {
            const uint8_t *p = row + (size_t)xtab_scratch[x] * 3;
            o[3 * x + 0] = p[0]; o[3 * x + 1] = p[1]; o[3 * x + 2] = p[2];
        }
    }
    return 0;
}

/* float32 cast then x /= 127.5f ; x -= 1.0f  (numpy in-place float32 ops: two roundings). */
int kdl_oracle_scale_f32(const uint8_t *u8, size_t n, float *out)
{
    if (!u8 || !out) return -1;
    for (size_t i = 0; i < n; ++i) {
        volatile float v = (float)u8[i];
        v = v / 127.5f;                 /* volatile: forbid contraction / reassociation */
        v = v - 1.0f;
        out[i] = v;
    }
    return 0;
}

/* Full gateway preprocessing for one image: resize to (out_h,out_w) then scale -> float32 HWC. */
int kdl_oracle_preprocess(const uint8_t *src, int h, int w, int64_t pitch,
                          float *out, int out_h, int out_w,
                          uint8_t *u8_scratch, int32_t *xtab_scratch, int32_t *ytab_scratch)
{
    int rc = kdl_oracle_resize_u8(src, h, w, pitch, u8_scratch, out_h, out_w, xtab_scratch, ytab_scratch);
    if (rc) return rc;
    return kdl_oracle_scale_f32(u8_scratch, (size_t)out_h * out_w * 3, out);
}

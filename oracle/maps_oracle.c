/*
 * oracle/maps_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement (plain C, single thread) of the TF1 graph ops that
 * tf-pose-estimation runs after the network:
 *     channel split      tf_pose/estimator.py:321-322
 *     tf.image.resize_area   tf_pose/estimator.py:326-329
 *     Smoother 25x25 depthwise conv (SAME)  tf_pose/estimator.py:332-333,
 *                                           tf_pose/tensblur/smoother.py:62-95
 *     3x3 max-pool + equal + where          tf_pose/estimator.py:335-337
 *
 * PARITY UNPINNED for this file: TensorFlow is an un-vendored, unpinned
 * dependency of the reference (README.md:30 "1.4.1+"), it is not installed
 * here and the reference ships no golden vectors for these ops.  The op
 * semantics below restate TF's published behaviour (ResizeAreaOp,
 * DepthwiseConv2dNative SAME/stride 1, MaxPool SAME) at the reference's own
 * call sites.  The only part that IS pinned is the 25x25 kernel: the literal
 * reference gauss_kernel() is executed by tests/golden/make_golden.py and
 * its output is committed (tests/golden/gauss_kernel_25_3.npy).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library.
 */
#include <math.h>
/* Runtime-dispatched AVX2 clone of the hot loops: the inner channel loop vectorises across the 19
 * channels, so each channel keeps its own tap-by-tap float32 accumulation order and (with
 * -ffp-contract=off) the results are bit-identical to the scalar build. */
#if defined(__x86_64__) && defined(__GNUC__)
#define TARGET_CLONES __attribute__((target_clones("avx2", "default")))
#else
#define TARGET_CLONES
#endif
#include <stdlib.h>
#include <string.h>

/* Phi(x), standard normal CDF (scipy.stats.norm.cdf used at smoother.py:65). */
static double std_normal_cdf(double x) { return 0.5 * erfc(-x / sqrt(2.0)); }

/*
 * smoother.py:62-71 gauss_kernel(kernlen, nsig):
 *   interval = (2*nsig+1)/kernlen
 *   x        = linspace(-nsig-interval/2, nsig+interval/2, kernlen+1)
 *   kern1d   = diff(cdf(x))
 *   raw      = sqrt(outer(kern1d, kern1d));  kernel = raw / raw.sum() -> float32
 * Outputs: k2d32[kernlen*kernlen] (the float32 weights TF convolves with) and
 * s1d[kernlen] (float64 separable factor sqrt(kern1d)/sum(sqrt(kern1d))).
 */
void oracle_gauss_kernel(int kernlen, double nsig, float *k2d32, double *s1d)
{
    double interval = (2.0 * nsig + 1.0) / (double)kernlen;
    double lo = -nsig - interval / 2.0, hi = nsig + interval / 2.0;
    int n1 = kernlen + 1;
    double *cdf = (double *)malloc(sizeof(double) * (size_t)n1);
    double *k1 = (double *)malloc(sizeof(double) * (size_t)kernlen);
    double step = (hi - lo) / (double)kernlen;
    for (int i = 0; i < n1; ++i) {
        /* numpy.linspace: start + i*step, last sample forced to stop */
        double x = (i == kernlen) ? hi : lo + (double)i * step;
        cdf[i] = std_normal_cdf(x);
    }
    for (int i = 0; i < kernlen; ++i) k1[i] = cdf[i + 1] - cdf[i];
    double total = 0.0;
    for (int r = 0; r < kernlen; ++r)
        for (int q = 0; q < kernlen; ++q) total += sqrt(k1[r] * k1[q]);
    for (int r = 0; r < kernlen; ++r)
        for (int q = 0; q < kernlen; ++q)
            k2d32[r * kernlen + q] = (float)(sqrt(k1[r] * k1[q]) / total);
    if (s1d) {
        double ssum = 0.0;
        for (int i = 0; i < kernlen; ++i) ssum += sqrt(k1[i]);
        for (int i = 0; i < kernlen; ++i) s1d[i] = sqrt(k1[i]) / ssum;
    }
    free(cdf);
    free(k1);
}

/* estimator.py:321-322: heat = X[...,:19], paf = X[...,19:] (HWC, C=57). */
void oracle_split(const float *x, int h, int w, int c_total, int c0, int c1, float *out)
{
    int c = c1 - c0;
    for (long p = 0; p < (long)h * w; ++p)
        memcpy(out + p * c, x + p * c_total + c0, sizeof(float) * (size_t)c);
}

static int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* one-axis area weight of source cell i for the output span [a,b), scale s */
static float area_wt(int i, float a, float b, float s)
{
    if ((float)i < a) return ((float)(i + 1) > b) ? s : ((float)(i + 1) - a);
    return ((float)(i + 1) > b) ? (b - (float)i) : 1.0f;
}

/*
 * tf.image.resize_area(images,[OH,OW],align_corners=False) as called at
 * estimator.py:326-329.  All arithmetic float32; inner (x) sum first, then the
 * y sum, then the 1/(hs*ws) scale -- SURVEY.md A.1.
 */
void oracle_resize_area(const float *in, int ih, int iw, int c, float *out, int oh, int ow)
{
    float hs = (float)ih / (float)oh, ws = (float)iw / (float)ow;
    float scale = 1.0f / (hs * ws);
    for (int y = 0; y < oh; ++y) {
        float ya = (float)y * hs, yb = (float)(y + 1) * hs;
        int y0 = (int)floorf(ya), y1 = (int)ceilf(yb);
        for (int x = 0; x < ow; ++x) {
            float xa = (float)x * ws, xb = (float)(x + 1) * ws;
            int x0 = (int)floorf(xa), x1 = (int)ceilf(xb);
            for (int k = 0; k < c; ++k) {
                float sum = 0.0f;
                for (int iy = y0; iy < y1; ++iy) {
                    float wy = area_wt(iy, ya, yb, hs);
                    const float *row = in + (long)clampi(iy, 0, ih - 1) * iw * c;
                    float sx = 0.0f;
                    for (int ix = x0; ix < x1; ++ix) {
                        float wx = area_wt(ix, xa, xb, ws);
                        sx += row[(long)clampi(ix, 0, iw - 1) * c + k] * wx;
                    }
                    sum += sx * wy;
                }
                out[((long)y * ow + x) * c + k] = sum * scale;
            }
        }
    }
}

/*
 * smoother.py:91 tf.nn.depthwise_conv2d(x, K, [1,1,1,1], 'SAME'):
 * g[Y,X,c] = sum_{r,q} K[r,q] * up[Y+r-pad, X+q-pad, c], zero outside, float32
 * accumulation in r-major / q-minor tap order (TF's own order is build
 * dependent; hence the 1e-5 tolerance on g, never bit-exactness).
 */
TARGET_CLONES void oracle_blur_f32(const float *up, int H, int W, int C, const float *k2d, int ks, float *g)
{
    int pad = ks / 2;
    float *acc = (float *)malloc(sizeof(float) * (size_t)C);
    for (int Y = 0; Y < H; ++Y)
        for (int X = 0; X < W; ++X) {
            for (int c = 0; c < C; ++c) acc[c] = 0.0f;
            for (int r = 0; r < ks; ++r) {
                int yy = Y + r - pad;
                if (yy < 0 || yy >= H) continue;
                for (int q = 0; q < ks; ++q) {
                    int xx = X + q - pad;
                    if (xx < 0 || xx >= W) continue;
                    float kv = k2d[r * ks + q];
                    const float *src = up + ((long)yy * W + xx) * C;
                    for (int c = 0; c < C; ++c) acc[c] += kv * src[c];
                }
            }
            memcpy(g + ((long)Y * W + X) * C, acc, sizeof(float) * (size_t)C);
        }
    free(acc);
}

/* Same conv with the float32 weights widened to double and double accumulation:
 * the "truth" the float32 results (CPU and GPU) are measured against. */
void oracle_blur_f64(const float *up, int H, int W, int C, const float *k2d, int ks, double *g)
{
    int pad = ks / 2;
    double *acc = (double *)malloc(sizeof(double) * (size_t)C);
    for (int Y = 0; Y < H; ++Y)
        for (int X = 0; X < W; ++X) {
            for (int c = 0; c < C; ++c) acc[c] = 0.0;
            for (int r = 0; r < ks; ++r) {
                int yy = Y + r - pad;
                if (yy < 0 || yy >= H) continue;
                for (int q = 0; q < ks; ++q) {
                    int xx = X + q - pad;
                    if (xx < 0 || xx >= W) continue;
                    double kv = (double)k2d[r * ks + q];
                    const float *src = up + ((long)yy * W + xx) * C;
                    for (int c = 0; c < C; ++c) acc[c] += kv * (double)src[c];
                }
            }
            memcpy(g + ((long)Y * W + X) * C, acc, sizeof(double) * (size_t)C);
        }
    free(acc);
}

/*
 * estimator.py:335-337: m = max-pool 3x3 SAME (window clipped at the border),
 * peaks = where(g == m, g, 0).  Plateau pixels are all kept.
 * margin (optional): g - (max over the in-range 8-neighbourhood), used by the
 * tests to find near-ties (|margin| tiny relative to g).
 */
void oracle_nms(const float *g, int H, int W, int C, float *peaks, float *margin)
{
    for (int Y = 0; Y < H; ++Y)
        for (int X = 0; X < W; ++X)
            for (int c = 0; c < C; ++c) {
                float v = g[((long)Y * W + X) * C + c];
                float m = v, nb = -INFINITY;
                for (int dy = -1; dy <= 1; ++dy)
                    for (int dx = -1; dx <= 1; ++dx) {
                        int yy = Y + dy, xx = X + dx;
                        if (yy < 0 || yy >= H || xx < 0 || xx >= W) continue;
                        float u = g[((long)yy * W + xx) * C + c];
                        if (u > m) m = u;
                        if ((dy || dx) && u > nb) nb = u;
                    }
                long o = ((long)Y * W + X) * C + c;
                peaks[o] = (v == m) ? v : 0.0f;
                if (margin) margin[o] = v - nb;
            }
}

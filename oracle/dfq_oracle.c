/*
 * dfq_oracle.c -- TEST INFRASTRUCTURE ONLY.
 *
 * Scalar CPU restatement (plain C, fp32 arithmetic, no FMA contraction) of the
 * reference's Data-Free-Quantization hot path.  It is the checker used by tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline leg; nothing in the product
 * package (data_free_quantization_b200/) may import, link or call it.
 *
 * Parity pin: the reference ships no golden vectors, so this file
 * is pinned against the reference's own Python run in the build container:
 * oracle/gen_golden.py imports /root/reference, runs every step on seeded weights,
 * asserts this restatement equals it (bit-exact for quantise / CLE / clip / BN-merge,
 * 1e-6 rel for the GEMV-like reductions) and commits the digests under tests/golden/.
 *
 * Build: gcc -O2 -fPIC -shared -ffp-contract=off -fno-fast-math  (see oracle/Makefile)
 *
 * Each function cites the reference lines it follows (paths relative to the
 * reference repository root).
 */
#include <math.h>
#include <stdint.h>
#include <stddef.h>
#include <string.h>
#include <stdlib.h>

#define EXPORT __attribute__((visibility("default")))

/* ---------------------------------------------------------------- min / max */
/* float(param.min()), float(param.max())  -- utils/layer_transform.py:298 */
EXPORT void dfqo_minmax_f32(const float *x, int64_t n, float *mn, float *mx)
{
    float lo = INFINITY, hi = -INFINITY;
    int nan = 0;
    for (int64_t i = 0; i < n; ++i) {
        float v = x[i];
        if (v != v) nan = 1;                /* torch.min / torch.max propagate NaN */
        if (v < lo) lo = v;
        if (v > hi) hi = v;
    }
    *mn = nan ? NAN : lo;
    *mx = nan ? NAN : hi;
}

/* ------------------------------------------------------------- quant params */
/*
 * utils/quantize.py:51-68.  range_f32 == 0: min/max are Python floats (double
 * arithmetic for the scale); range_f32 == 1: the `None` path where min/max are 0-dim
 * fp32 tensors and the scale is formed in fp32 (:32,:37,:66).
 * Outputs the three fp32 constants of the elementwise pass.
 */
EXPORT void dfqo_quant_params(double min_value, double max_value, int bits, int symmetric,
                              int range_f32, float *neg_min, float *scale, float *qmin,
                              float *qmax, float *add_back)
{
    double qlo, qhi, sc, mn = min_value;
    if (symmetric) {
        qlo = -pow(2.0, bits - 1);
        qhi = pow(2.0, bits - 1) - 1.0;
        double a = fabs(max_value), b = fabs(min_value);
        if (a < b) a = b;
        sc = range_f32 ? (double)((float)a / (float)qhi) : a / qhi;
        mn = 0.0;
    } else {
        qlo = 0.0;
        qhi = pow(2.0, bits) - 1.0;
        if (range_f32)
            sc = (double)(((float)max_value - (float)min_value) / (float)(qhi - qlo));
        else
            sc = (max_value - min_value) / (qhi - qlo);
    }
    if (1e-8 > sc) sc = 1e-8;             /* scale = max(scale, 1e-8)  :68 -- Python max(): a NaN scale stays NaN */
    *neg_min = (float)(-mn);
    *scale = (float)sc;
    *qmin = (float)qlo;
    *qmax = (float)qhi;
    *add_back = (float)mn;
}

/* --------------------------------------------------------------- fake quant */
/*
 * utils/quantize.py:72-76: add_(-min).div_(scale).clamp_(qmin,qmax).round_()
 * .mul_(scale).add_(min), six separate fp32 roundings, round half to even.
 * codes (optional) receives the integer grid index q - qmin.
 */
EXPORT void dfqo_fake_quant_f32(const float *x, float *y, int32_t *codes, int64_t n,
                                float neg_min, float scale, float qmin, float qmax,
                                float add_back)
{
    for (int64_t i = 0; i < n; ++i) {
        volatile float t = x[i] + neg_min;
        t = t / scale;
        /* torch clamp: min(max(x, lo), hi); NaN propagates */
        float c = t;
        if (c < qmin) c = qmin;
        if (c > qmax) c = qmax;
        float q = rintf(c);
        if (codes) codes[i] = (int32_t)(q - qmin);
        volatile float d = q * scale;
        y[i] = d + add_back;
    }
}

/* quantize_targ_layer for one tensor -- utils/layer_transform.py:297-303 */
EXPORT void dfqo_quantize_tensor(const float *x, float *y, int32_t *codes, int64_t n, int bits,
                                 float *out_min, float *out_scale)
{
    float mn, mx, nm, sc, ql, qh, ab;
    dfqo_minmax_f32(x, n, &mn, &mx);
    dfqo_quant_params((double)mn, (double)mx, bits, 0, 0, &nm, &sc, &ql, &qh, &ab);
    dfqo_fake_quant_f32(x, y, codes, n, nm, sc, ql, qh, ab);
    if (out_min) *out_min = mn;
    if (out_scale) *out_scale = sc;
}

/* per-channel variant (new capability; no reference implementation -- parity unpinned,
 * SURVEY.md section 8c(vi)): the same formula applied to each row of W.view(O,-1). */
EXPORT void dfqo_quantize_per_channel(const float *x, float *y, int32_t *codes, int64_t rows,
                                      int64_t cols, int bits)
{
    for (int64_t r = 0; r < rows; ++r)
        dfqo_quantize_tensor(x + r * cols, y + r * cols, codes ? codes + r * cols : NULL, cols,
                             bits, NULL, NULL);
}

/* ----------------------------------------------------------------- BN merge */
/*
 * utils/layer_transform.py:255-270.  Per out-channel o:
 *   f = gamma / sqrt(var + eps)            (fp32, eps cast to fp32)
 *   W[o,:] *= f
 *   b[o]   = b*f + (beta - (gamma*mean)/sqrt(var+eps))
 * fake_weight = |gamma|, fake_bias = beta (:273-274) are produced by the caller.
 */
EXPORT void dfqo_merge_bn(float *W, float *b, int64_t O, int64_t K, const float *gamma,
                          const float *beta, const float *mean, const float *var, float eps)
{
    for (int64_t o = 0; o < O; ++o) {
        volatile float ve = var[o] + eps;
        volatile float sd = sqrtf(ve);
        volatile float f = gamma[o] / sd;
        for (int64_t k = 0; k < K; ++k) W[o * K + k] = W[o * K + k] * f;
        volatile float bf = b[o] * f;
        volatile float gm = gamma[o] * mean[o];
        volatile float gms = gm / sd;
        volatile float sh = beta[o] - gms;
        b[o] = bf + sh;
    }
}

/* ------------------------------------------------- cross-layer equalisation */
/*
 * One relation, Cross_layer_equal.py:11-59 (unsigned ranges :38-39 or signed :34-35).
 * W1: [C1, K1] rows; W2: [O2, I2, KK] with C1 = G*I2, O2 = G*Og.
 * Channel c = g*I2 + it owns row c of W1 and the "column" W2[g*Og:(g+1)*Og, it, :].
 * s = (1/(r1+eps)) * sqrt(r1*r2+eps) in fp32, clamped with Python's min/max semantics
 * (:43: NaN -> smax because `s < smax` is False), then five in-place multiplies
 * (:46-57) and W2 *= 1/s with 1/s rounded to fp32 first (fp32 when s is a tensor,
 * double->fp32 when the clamp replaced it by a Python float).
 */
EXPORT void dfqo_cle_pair(float *W1, int64_t C1, int64_t K1, float *W2, int64_t O2, int64_t I2,
                          int64_t KK, float *b1, float *bn_w, float *bn_b, float *S,
                          double smin, double smax, int is_signed, float eps)
{
    int64_t G = 1;
    if (C1 != I2) G = C1 / I2;
    int64_t Og = O2 / G;
    for (int64_t g = 0; g < G; ++g) {
        for (int64_t it = 0; it < I2; ++it) {
            int64_t c = g * I2 + it;
            float *row = W1 + c * K1;
            float lo1 = INFINITY, hi1 = -INFINITY, lo2 = INFINITY, hi2 = -INFINITY;
            float a1 = 0.f, a2 = 0.f;
            for (int64_t k = 0; k < K1; ++k) {
                float v = row[k];
                if (v < lo1) lo1 = v;
                if (v > hi1) hi1 = v;
                if (fabsf(v) > a1) a1 = fabsf(v);
            }
            for (int64_t o = 0; o < Og; ++o) {
                float *col = W2 + ((g * Og + o) * I2 + it) * KK;
                for (int64_t k = 0; k < KK; ++k) {
                    float v = col[k];
                    if (v < lo2) lo2 = v;
                    if (v > hi2) hi2 = v;
                    if (fabsf(v) > a2) a2 = fabsf(v);
                }
            }
            volatile float r1 = is_signed ? a1 : hi1 - lo1;
            volatile float r2 = is_signed ? a2 : hi2 - lo2;
            volatile float den = r1 + eps;
            volatile float inv = 1.0f / den;
            volatile float prod = r1 * r2;
            volatile float rad = prod + eps;
            volatile float rt = sqrtf(rad);
            volatile float st = inv * rt;
            float s, sinv;
            /* s = max(smin, min(smax, s)) with Python builtin semantics */
            if ((double)st < smax) {            /* min() keeps the tensor */
                if ((double)st > smin) {        /* max() keeps the tensor */
                    s = st;
                    sinv = 1.0f / s;
                } else {
                    s = (float)smin;
                    sinv = (float)(1.0 / smin);
                }
            } else {                            /* includes NaN */
                if (smax > smin) {
                    s = (float)smax;
                    sinv = (float)(1.0 / smax);
                } else {
                    s = (float)smin;
                    sinv = (float)(1.0 / smin);
                }
            }
            S[c] = s;
            for (int64_t k = 0; k < K1; ++k) row[k] = row[k] * s;
            if (b1) b1[c] = b1[c] * s;
            if (bn_w) bn_w[c] = bn_w[c] * s;
            if (bn_b) bn_b[c] = bn_b[c] * s;
            for (int64_t o = 0; o < Og; ++o) {
                float *col = W2 + ((g * Og + o) * I2 + it) * KK;
                for (int64_t k = 0; k < KK; ++k) col[k] = col[k] * sinv;
            }
        }
    }
}

/* mean(|W - W_old|) of one layer as fp32 -- Cross_layer_equal.py:107.
 * Accumulated in double: the reference's fp32 tree sum differs from the exact mean by
 * ~1e-7 relative, far inside the convergence margin. */
EXPORT float dfqo_mean_abs_diff(const float *a, const float *b, int64_t n)
{
    double acc = 0.0;
    for (int64_t i = 0; i < n; ++i) {
        volatile float d = a[i] - b[i];
        acc += (double)fabsf(d);
    }
    return (float)(acc / (double)n);
}

/* ----------------------------------------------------- bias absorption + clip */
/*
 * bias_absorption.py:42-80 for one relation.
 *   c  = clamp(beta - N*gamma, min=0)                        (:48-49, fp32)
 *   wc[g*Og+o] = sum_it ( sum_k W2[g*Og+o, it, k] ) * c[g*I2+it]   (:52-67)
 *   b1 -= c ; fake_bias -= c ; b2 += wc                      (:78-80)
 * The reference reduces with ATen sum + BLAS matmul (order unspecified); this
 * restatement accumulates in double and rounds once, so parity is <=1e-5 relative,
 * not bit-exact.  c_out (optional) returns c.
 */
EXPORT void dfqo_absorb_pair(const float *W2, int64_t O2, int64_t I2, int64_t KK, int64_t C1,
                             float *b1, float *b2, const float *bn_w, float *bn_b, float N,
                             float *c_out)
{
    int64_t G = C1 / I2;
    int64_t Og = O2 / G;
    float *c = (float *)malloc(sizeof(float) * (size_t)C1);
    for (int64_t i = 0; i < C1; ++i) {
        volatile float ng = N * bn_w[i];
        volatile float v = bn_b[i] - ng;
        c[i] = v < 0.f ? 0.f : v;          /* clamp_(0): NaN stays NaN */
        if (v != v) c[i] = v;
    }
    for (int64_t g = 0; g < G; ++g)
        for (int64_t o = 0; o < Og; ++o) {
            double acc = 0.0;
            for (int64_t it = 0; it < I2; ++it) {
                const float *p = W2 + ((g * Og + o) * I2 + it) * KK;
                double rs = 0.0;
                for (int64_t k = 0; k < KK; ++k) rs += (double)p[k];
                acc += (double)(float)rs * (double)c[g * I2 + it];
            }
            b2[g * Og + o] = b2[g * Og + o] + (float)acc;
        }
    for (int64_t i = 0; i < C1; ++i) {
        b1[i] = b1[i] + (-c[i]);
        bn_b[i] = bn_b[i] + (-c[i]);
        if (c_out) c_out[i] = c[i];
    }
    free(c);
}

/* clip_weight.py:29 -- weight.clamp_(lo, hi) */
EXPORT void dfqo_clip(float *w, int64_t n, float lo, float hi)
{
    for (int64_t i = 0; i < n; ++i) {
        float v = w[i];
        if (v < lo) v = lo;
        if (v > hi) v = hi;
        w[i] = v;
    }
}

/* ------------------------------------------------------------ bias correction */
/*
 * eps = Q(W) - W (bias_correction.py:125-130), summed over the kernel window to
 * [O, I2] (:231), then reduced against a per-input-channel vector v:
 *   out[o] = sum_i epsI[o,i] * v[i]            (paper form, eps . E[x])
 * Weight range = the tensor's own min/max as in _quantize_error (:128).
 * Double accumulation, one rounding: <=1e-5 relative parity.
 * eps_out (optional, [O*I2]) returns the window-summed error for the ref-body mode.
 */
EXPORT void dfqo_bias_correct(const float *W, int64_t O, int64_t I2, int64_t KK, int bits,
                              int is_signed, const float *v, float *out, float *eps_out)
{
    float mn, mx, nm, sc, ql, qh, ab;
    int64_t n = O * I2 * KK;
    dfqo_minmax_f32(W, n, &mn, &mx);
    dfqo_quant_params((double)mn, (double)mx, bits, is_signed, 0, &nm, &sc, &ql, &qh, &ab);
    for (int64_t o = 0; o < O; ++o) {
        double acc = 0.0;
        for (int64_t i = 0; i < I2; ++i) {
            const float *p = W + (o * I2 + i) * KK;
            float es = 0.f;                      /* torch.sum over <=9 taps, sequential fp32 */
            for (int64_t k = 0; k < KK; ++k) {
                float q;
                dfqo_fake_quant_f32(p + k, &q, NULL, 1, nm, sc, ql, qh, ab);
                volatile float e = q - p[k];
                volatile float t = es + e;
                es = t;
            }
            if (eps_out) eps_out[o * I2 + i] = es;
            if (v) acc += (double)es * (double)v[i];
        }
        if (out) out[o] = (float)acc;
    }
}

/* ------------------------------------------------ QuantMeasure statistics */
/* utils/quantize.py:110-111: mean over samples of the per-sample max / min. */
EXPORT void dfqo_sample_minmax_mean(const float *x, int64_t N, int64_t chw, float *mn, float *mx)
{
    double slo = 0.0, shi = 0.0;
    for (int64_t n = 0; n < N; ++n) {
        float lo, hi;
        dfqo_minmax_f32(x + n * chw, chw, &lo, &hi);
        slo += lo;
        shi += hi;
    }
    *mn = (float)(slo / (double)N);
    *mx = (float)(shi / (double)N);
}

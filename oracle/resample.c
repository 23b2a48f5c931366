/* oracle/resample.c -- TEST INFRASTRUCTURE ONLY.  CPU restatement of the reference's rational resampler as the
 * STREAMING state machine it is (history, skipped front samples, per-input-sample subfilter lists, flush), so that
 * the CUDA path's closed form  y[m] = sum_j h[(fd + m M) % L + j L] x[(fd + m M) / L - j]  is checked against the
 * original bookkeeping:
 *   lib/src/phy/resample/resampler.cpp:55-460        constructor, get_N_samples_after_resampling, resample,
 *                                                    resample_final_samples, reset, resample_LXMX_generic_impl
 *   lib/src/phy/filter/kaiser.cpp:37-118, rectangular.cpp:28-46, bessel.cpp:28-62   Kaiser low pass design
 *   lib/include/dectnrp/phy/resample/resampler_param.hpp:61-86                      filter definitions per os
 * The manually unrolled variants (resampler.cpp:462-564) compute the same dot products and are not restated.
 * PARITY UNPINNED: the reference cannot be built here and ships no vectors for this path; VOLK's SIMD summation order
 * inside one dot product is not reproduced (plain left-to-right float sum).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct { float re, im; } cf_t;

/* bessel.cpp:28-62: 8-term series in double, returned as float */
static float bessel_i0(float zf) {
    const double z = zf;
    double ret = 0.0;
    for (uint32_t k = 0; k <= 8; ++k) {
        double fac = 1.0;
        for (uint32_t i = 1; i <= k; ++i) fac *= (double)i;
        ret += pow(0.25 * z * z, (double)k) / (fac * fac);
    }
    return (float)ret;
}

static float sincf_(float n) { return n == 0.0f ? 1.0f : (float)(sin(M_PI * n) / (M_PI * n)); }

/* kaiser.cpp:37-118 (SampleRate = 1, N forced odd); returns number of taps written (<= cap) */
int orc_kaiser(float f_pass, float f_stop, float ripple_dB, float att_dB, float* out, int cap) {
    const float d0 = powf(10.0f, -att_dB / 20.0f), d1 = powf(10.0f, ripple_dB / 20.0f) - 1.0f;
    const float delta = d0 < d1 ? d0 : d1;
    const float A = -20.0f * log10f(delta);
    float beta = 0.0f;
    if (A > 50.0f) beta = 0.1102f * (A - 8.7f);
    else if (21.0f <= A && A <= 50.0f) beta = (float)(0.5842f * powf(A - 21.0f, 0.4f) + 0.07886 * (A - 21.0f));
    const float b = f_stop - f_pass;
    const float order = (float)((A - 7.95f) / (2.285f * 2.0f * M_PI * b));
    uint32_t N = (uint32_t)ceilf(order + 1.0f);
    if (N % 2 == 0) N += 1;
    if ((int)N > cap) return -1;
    const float fc = f_pass + b / 2.0f, Nf = (float)N;
    float norm = 0.0f;
    for (uint32_t n = 0; n < N; ++n) {
        const float nf = (float)n;
        const float arg = beta * sqrtf(1.0f - powf(2.0f * nf / (Nf - 1.0f) - 1.0f, 2.0f));
        const float w = bessel_i0(arg) / bessel_i0(beta);
        const float r = 2.0f * fc * sincf_(2.0f * fc * (nf - (Nf - 1.0f) / 2.0f));
        out[n] = w * r;
        norm += out[n];
    }
    for (uint32_t n = 0; n < N; ++n) out[n] = out[n] / norm;
    return (int)N;
}

typedef struct {
    uint32_t L, M, filter_length, subfilter_length, n_skip_front, n_skip_current, history_length, input_sample_cnt;
    float* lowpass;            /* [subfilter_length * L] scaled by L, zero padded */
    float** subfilters;        /* [L][subfilter_length], mirrored */
    uint32_t* idx_cnt;         /* [rows] number of subfilters used at input_sample_cnt = row */
    uint32_t* idx;             /* [rows][L] */
    uint32_t rows;
    cf_t* history;             /* [2 * history_length] */
} orc_resampler_t;

void orc_resampler_reset(orc_resampler_t* r) {
    memset(r->history, 0, sizeof(cf_t) * r->history_length);
    r->n_skip_current = r->n_skip_front;
    r->input_sample_cnt = 0;
}

orc_resampler_t* orc_resampler_new(int L, int M, float f_pass_norm, float f_stop_norm, float ripple_dB, float att_dB) {
    orc_resampler_t* r = (orc_resampler_t*)calloc(1, sizeof(*r));
    r->L = L; r->M = M;
    const float lm = (float)(L > M ? L : M);
    float* tmp = (float*)malloc(sizeof(float) * 65536);
    const int fl = orc_kaiser(f_pass_norm / lm, f_stop_norm / lm, ripple_dB, att_dB, tmp, 65536);
    if (fl < 0) { free(tmp); free(r); return NULL; }
    r->filter_length = fl;
    const uint32_t delay = (fl - 1) / 2;
    r->subfilter_length = (fl + L - 1) / L;
    r->lowpass = (float*)calloc((size_t)r->subfilter_length * L, sizeof(float));
    for (int i = 0; i < fl; ++i) r->lowpass[i] = tmp[i] * (float)L;
    free(tmp);
    r->subfilters = (float**)malloc(sizeof(float*) * L);
    for (int l = 0; l < L; ++l) {
        r->subfilters[l] = (float*)malloc(sizeof(float) * r->subfilter_length);
        for (uint32_t k = 0; k < r->subfilter_length; ++k)
            r->subfilters[l][r->subfilter_length - 1 - k] = r->lowpass[l + k * L];
    }
    r->n_skip_front = delay / L;
    uint32_t sub = delay % L;
    r->idx_cnt = (uint32_t*)calloc(2 * M + 2, sizeof(uint32_t));
    r->idx = (uint32_t*)calloc((size_t)(2 * M + 2) * L, sizeof(uint32_t));
    r->rows = 0;
    for (uint32_t period = 0; period < (uint32_t)M; ++period) {
        while (sub < (uint32_t)L) { r->idx[r->rows * L + r->idx_cnt[r->rows]++] = sub; sub += M; }
        r->rows++;
        if (sub >= 2 * (uint32_t)L) { r->rows++; ++period; }   /* an input sample without output sample */
        sub %= L;
    }
    r->history_length = r->subfilter_length - 1;
    r->history = (cf_t*)calloc(2 * r->history_length + 1, sizeof(cf_t));
    orc_resampler_reset(r);
    return r;
}

void orc_resampler_free(orc_resampler_t* r) {
    if (!r) return;
    for (uint32_t l = 0; l < r->L; ++l) free(r->subfilters[l]);
    free(r->subfilters); free(r->lowpass); free(r->idx_cnt); free(r->idx); free(r->history); free(r);
}

void orc_resampler_info(const orc_resampler_t* r, uint32_t* out5) {
    out5[0] = r->filter_length; out5[1] = r->subfilter_length; out5[2] = r->n_skip_front; out5[3] = r->history_length;
    out5[4] = r->rows;
}
void orc_resampler_lowpass(const orc_resampler_t* r, float* out) {
    memcpy(out, r->lowpass, sizeof(float) * r->subfilter_length * r->L);
}

/* resampler.cpp:310-328 */
uint32_t orc_resampler_n_after(const orc_resampler_t* r, uint32_t n_in) {
    if (r->L == 1 && r->M == 1) return n_in;
    const uint32_t eff = r->history_length + n_in - r->n_skip_front;
    const uint32_t mult = eff / r->M, res = eff - mult * r->M;
    uint32_t n = mult * r->L;
    for (uint32_t i = 0; i < res; ++i) n += r->idx_cnt[i];
    return n;
}

/* resampler.cpp:437-460 */
static uint32_t generic_impl(orc_resampler_t* r, const cf_t* in, cf_t* out, uint32_t n_new) {
    uint32_t cnt_in = 0, cnt_out = 0;
    for (uint32_t i = 0; i < n_new; ++i) {
        for (uint32_t j = 0; j < r->idx_cnt[r->input_sample_cnt]; ++j) {
            const float* f = r->subfilters[r->idx[r->input_sample_cnt * r->L + j]];
            float re = 0.0f, im = 0.0f;
            for (uint32_t k = 0; k < r->subfilter_length; ++k) { re += in[cnt_in + k].re * f[k]; im += in[cnt_in + k].im * f[k]; }
            out[cnt_out].re = re; out[cnt_out].im = im; ++cnt_out;
        }
        ++cnt_in;
        r->input_sample_cnt = (r->input_sample_cnt + 1) % r->M;
    }
    return cnt_out;
}

/* resampler.cpp:392-435, one antenna; n_new >= history_length is the caller's duty (asserted in the reference) */
int orc_resampler_resample(orc_resampler_t* r, const cf_t* in, cf_t* out, uint32_t n_new) {
    if (r->L == 1 && r->M == 1) { memcpy(out, in, sizeof(cf_t) * n_new); return (int)n_new; }
    if (n_new < r->history_length) return -1;
    const uint32_t H = r->history_length;
    memcpy(r->history + H, in, sizeof(cf_t) * H);
    uint32_t n = generic_impl(r, r->history + r->n_skip_current, out, H - r->n_skip_current);
    n += generic_impl(r, in, out + n, n_new - H);
    memcpy(r->history, in + n_new - H, sizeof(cf_t) * H);
    r->n_skip_current = 0;
    return (int)n;
}

/* resampler.cpp:337-358 */
int orc_resampler_final(orc_resampler_t* r, cf_t* out) {
    if (r->L == 1 && r->M == 1) return 0;
    memset(r->history + r->history_length, 0, sizeof(cf_t) * r->history_length);
    return (int)generic_impl(r, r->history, out, r->history_length);
}

/* oracle/sync.c -- TEST INFRASTRUCTURE ONLY.  CPU restatement of the reference's packet synchronisation for one
 * IQ window, written as the same streaming algorithm (moving sums that pop and push, streaks, per-sample peak
 * search), so that the CUDA path's batched prefix-sum formulation is checked against the original bookkeeping:
 *   lib/src/phy/rx/sync/sync_chunk.cpp:120-215        search(): detection -> coarse peak -> fine peak
 *   lib/src/phy/rx/sync/autocorrelator_detection.cpp:152-283, movsum.hpp, movsum_uw.cpp
 *   lib/src/phy/rx/sync/autocorrelator_peak.cpp:145-384
 *   lib/src/phy/rx/sync/crosscorrelator.cpp:89-249, stf_template.cpp:81-204 (templates are passed in)
 *   lib/include/dectnrp/phy/rx/sync/sync_param.hpp (constants as shipped, cover sequence active)
 * PARITY UNPINNED: the reference cannot be built here and ships no vectors for this path.
 * Not restated: the periodic "resum for numerical stability" (float drift control, no effect on the algorithm),
 * resampling: see orc_sync_search2 (detection/coarse on the M/L-resampled window, xcorr at the hardware rate), beta / integer CFO search (disabled or unimplemented in the reference).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct { float re, im; } cf_t;

typedef struct {
    int32_t detected;                 /* 0: nothing found in the window */
    int32_t detection_ant_idx, detection_time_local, detection_time_with_jump_back_local;
    float detection_rms, detection_metric;
    int32_t coarse_peak_time_local;
    float coarse_peak_array[8];
    float rms_array[8];
    float cfo_fractional_rad, cfo_integer_rad;
    int32_t N_eff_TX, fine_peak_time_local;   /* fine peak relative to the window start */
    int32_t n_detections;
    int32_t coarse_peak_time_hw;              /* coarse peak at the hardware rate (sync_chunk.cpp:204-205) */
} orc_sync_report_t;

/* ---- movsum_t<float> / movsum_uw_t (movsum.hpp:33-104, movsum_uw.cpp:28-97) */
typedef struct { float* reg; int len, ptr; float sum; } msf_t;
typedef struct { cf_t* reg; int len, ptr, n_rep, n_uw; const float* uw; cf_t sum; } msu_t;

static int front_idx(int len, int ptr, long n_backward) {
    const int nbw = (int)(n_backward % len);
    return nbw > ptr ? len - (nbw - ptr) : ptr - nbw;
}
static void msf_init(msf_t* m, int len) { m->reg = (float*)calloc(len, sizeof(float)); m->len = len; m->ptr = 0; m->sum = 0; }
static void msf_pop_push(msf_t* m, float v) {
    m->sum -= m->reg[m->ptr]; m->sum += v; m->reg[m->ptr] = v; m->ptr = (m->ptr + 1) % m->len;
}
static float msf_sum_front(const msf_t* m, int n) {
    float r = 0; for (int i = 0; i < n; ++i) r += m->reg[front_idx(m->len, m->ptr, i)]; return r;
}
static float msf_sum_back(const msf_t* m, int n) {
    float r = 0; for (int i = 0; i < n; ++i) r += m->reg[(m->ptr + 1 + i) % m->len]; return r;
}
static void msf_resum(msf_t* m) { float s = 0; for (int i = 0; i < m->len; ++i) s += m->reg[i]; m->sum = s; }

static void msu_init(msu_t* m, const float* uw, int n_uw, int n_rep) {
    m->len = n_uw * n_rep; m->reg = (cf_t*)calloc(m->len, sizeof(cf_t)); m->ptr = 0; m->n_rep = n_rep; m->n_uw = n_uw;
    m->uw = uw; m->sum.re = m->sum.im = 0;
}
static void msu_pop_push(msu_t* m, cf_t v) {
    /* prefactor_pop = uw.front(), prefactor_push = uw.back() = 1 */
    m->sum.re -= m->uw[0] * m->reg[m->ptr].re; m->sum.im -= m->uw[0] * m->reg[m->ptr].im;
    m->reg[m->ptr] = v;
    m->sum.re += v.re; m->sum.im += v.im;
    for (int i = 1; i < m->n_uw; ++i) {
        const long n_backward = (long)m->len - (long)i * m->n_rep;
        const cf_t x = m->reg[front_idx(m->len, m->ptr, n_backward)];
        if (m->uw[i - 1] == -1.0f && m->uw[i] == 1.0f) { m->sum.re -= 2.0f * x.re; m->sum.im -= 2.0f * x.im; }   /* pos_neg list */
        if (m->uw[i - 1] == 1.0f && m->uw[i] == -1.0f) { m->sum.re += 2.0f * x.re; m->sum.im += 2.0f * x.im; }   /* neg_pos list */
    }
    m->ptr = (m->ptr + 1) % m->len;
}
static void msu_resum(msu_t* m) {
    cf_t s = {0, 0};
    for (int i = 0; i < m->n_uw; ++i) {
        const long n_backward = (long)m->len - (long)i * m->n_rep;
        int B = front_idx(m->len, m->ptr, n_backward);
        cf_t p = {0, 0};
        for (int k = 0; k < m->n_rep; ++k) { const cf_t x = m->reg[(B + k) % m->len]; p.re += x.re; p.im += x.im; }
        s.re += m->uw[i] * p.re; s.im += m->uw[i] * p.im;
    }
    m->sum = s;
}
static cf_t conj_dot(const cf_t* a, const cf_t* b, int n);

static const float COVER[9] = {1, -1, 1, 1, -1, -1, -1, -1, -1};  /* stf.hpp:146-148 */

/* One window.  iq [n_ant][stride] at the DECT rate (x os); len = usable samples; search_length = number of samples
 * the detection may consume (A + B of sync_chunk.hpp:108-124).  The cross-correlation runs on the hardware-rate
 * samples iq_hw [n_ant][stride_hw] (NULL: = iq, L = M = 1) against templates [n_tpl][tpl_len] that went through the
 * TX resampler (tpl_len = stf_len * L / M), hardware rate = DECT rate * L / M. */
int orc_sync_search3(const cf_t* iq, long stride, int len, int n_ant, int u, int b, int os, const cf_t* templates,
                     int n_tpl, int search_length, double samp_rate, const cf_t* iq_hw, long stride_hw, int len_hw,
                     int L, int M, int tpl_len, int ignore_before_min, orc_sync_report_t* rep);
int orc_sync_search2(const cf_t* iq, long stride, int len, int n_ant, int u, int b, int os, const cf_t* templates,
                     int n_tpl, int search_length, double samp_rate, const cf_t* iq_hw, long stride_hw, int len_hw,
                     int L, int M, int tpl_len, orc_sync_report_t* rep) {
    return orc_sync_search3(iq, stride, len, n_ant, u, b, os, templates, n_tpl, search_length, samp_rate, iq_hw, stride_hw,
                            len_hw, L, M, tpl_len, 0, rep);
}
/* ignore_before_min: the caller resumes a longer walk inside this window (autocorrelator_detection_t::skip_after_peak
 * moved ignore_before_index, autocorrelator_detection.cpp:137-146): no detection may end before that sample */
int orc_sync_search3(const cf_t* iq, long stride, int len, int n_ant, int u, int b, int os, const cf_t* templates,
                     int n_tpl, int search_length, double samp_rate, const cf_t* iq_hw, long stride_hw, int len_hw,
                     int L, int M, int tpl_len, int ignore_before_min, orc_sync_report_t* rep) {
    if (!iq_hw) { iq_hw = iq; stride_hw = stride; len_hw = len; L = M = 1; }
    const int bos = b * os;
    const int n_pat = u == 1 ? 7 : 9;
    const int stf_len = (u == 1 ? 112 : 144) * bos, pat = stf_len / n_pat;
    const int step = pat / 4;
    float uw[8];
    for (int i = 0; i + 1 < n_pat; ++i) uw[i] = COVER[i] * COVER[i + 1];
    const float rms_min = 0.005f * (float)sqrt(samp_rate / 30.72e6), rms_max = 2.0f;
    const float prefactor = (float)n_pat / (float)(n_pat - 1);
    const int jump_back = pat, search_start = stf_len + jump_back, skip_after_peak = 2 * stf_len;
    memset(rep, 0, sizeof(*rep));

    msu_t mc[8]; msf_t mp[8];
    for (int a = 0; a < n_ant; ++a) { msu_init(&mc[a], uw, n_pat - 1, 4); msf_init(&mp[a], n_pat * 4); }
    int ret = 0;
    /* set_power_of_first_stf_pattern */
    int cnt_r = 0;
    for (int a = 0; a < n_ant; ++a)
        for (int i = 0; i < 4; ++i) {
            const cf_t* A = iq + a * stride + i * step;
            msf_pop_push(&mp[a], conj_dot(A, A, step).re);
        }
    cnt_r = pat;
    int ignore_before = search_start > ignore_before_min ? search_start : ignore_before_min;
    while (cnt_r < search_length && !rep->detected) {
        /* ---- autocorrelator_detection_t::search_by_correlation, one step */
        if (cnt_r + step > len) break;
        for (int a = 0; a < n_ant; ++a) {
            const cf_t* A = iq + a * stride + cnt_r - pat;
            const cf_t* B = iq + a * stride + cnt_r;
            msu_pop_push(&mc[a], conj_dot(A, B, step));
            msf_pop_push(&mp[a], conj_dot(B, B, step).re);
        }
        cnt_r += step;
        if (ignore_before > cnt_r) continue;
        int found = 0;
        for (int a = 0; a < n_ant && !found; ++a) {
            const float power = mp[a].sum;
            const float rms = sqrtf(power / (float)stf_len);
            if (rms < rms_min || rms_max < rms) continue;
            const float rms_back = sqrtf(msf_sum_back(&mp[a], 2)), rms_front = sqrtf(msf_sum_front(&mp[a], 2));
            if (rms_back * 0.5 >= rms_front) continue;
            const float metric = powf(prefactor * hypotf(mc[a].sum.re, mc[a].sum.im) / power, 2.0f);
            if (metric < 0.18f || 1.50f < metric) continue;      /* streak of length 1 ...            */
            if (!(0.18f < metric)) continue;                     /* ... streak_t::check: value < metric */
            rep->detection_ant_idx = a; rep->detection_rms = rms; rep->detection_metric = metric;
            rep->detection_time_local = cnt_r; rep->detection_time_with_jump_back_local = cnt_r - jump_back;
            found = 1;
        }
        if (!found) continue;
        rep->n_detections++;

        /* ---- autocorrelator_peak_t: per-sample search over one STF after (detection - 1 pattern) */
        const int r0 = rep->detection_time_with_jump_back_local, D = stf_len;
        if (r0 + D > len) break;
        float peak_metric[8]; int peak_index[8];
        const int sm_right = 1 * bos, sm_len = 1 * bos + 1 + sm_right;
        for (int a = 0; a < n_ant; ++a) {
            msu_t c2; msf_t p2, sm;
            msu_init(&c2, uw, n_pat - 1, pat); msf_init(&p2, stf_len); msf_init(&sm, sm_len);
            const cf_t* x = iq + a * stride;
            /* set_initial_movsums(r0 - stf_len) */
            const int A0 = r0 - stf_len;
            for (int i = 0; i < c2.len; ++i) { const cf_t v = conj_dot(x + A0 + i, x + A0 + pat + i, 1); c2.reg[i] = v; }
            c2.ptr = 0; msu_resum(&c2);
            for (int i = 0; i < stf_len; ++i) p2.reg[i] = conj_dot(x + A0 + i, x + A0 + i, 1).re;
            p2.ptr = 0; msf_resum(&p2);
            peak_metric[a] = 0.0f; peak_index[a] = 0;
            for (int n = r0; n < r0 + D; ++n) {
                msu_pop_push(&c2, conj_dot(x + n - pat, x + n, 1));
                msf_pop_push(&p2, conj_dot(x + n, x + n, 1).re);
                const float metric = powf(prefactor * hypotf(c2.sum.re, c2.sum.im) / p2.sum, 2.0f);
                msf_pop_push(&sm, metric);
                const float mean = sm.sum / (float)sm.len;
                if (mean > peak_metric[a]) { peak_metric[a] = mean; peak_index[a] = n - sm_right; }
            }
            free(c2.reg); free(p2.reg); free(sm.reg);
        }
        /* post_processing_validity */
        float widx = 0.0f, msum = 0.0f; int valid = 0;
        for (int a = 0; a < 8; ++a) rep->coarse_peak_array[a] = 0.0f;
        for (int a = 0; a < n_ant; ++a) {
            if (rep->detection_metric + (-0.25f) >= peak_metric[a]) continue;
            if ((long)rep->detection_time_local + (long)(-0.3 * (double)stf_len) >= (long)peak_index[a]) continue;
            rep->coarse_peak_array[a] = peak_metric[a];
            widx += peak_metric[a] * (float)peak_index[a]; msum += peak_metric[a]; ++valid;
        }
        if (!valid) { continue; }   /* sync_chunk.cpp:200-205: report dropped, detection goes on */
        rep->coarse_peak_time_local = (int)roundf(widx / msum) - (stf_len - 1);
        /* post_processing_at_coarse_peak: rms and fractional CFO over the STF at the coarse peak */
        float wcfo = 0.0f, ms2 = 0.0f;
        for (int a = 0; a < n_ant; ++a) {
            if (rep->coarse_peak_array[a] <= 0.0f) continue;
            const cf_t* x = iq + a * stride + rep->coarse_peak_time_local;
            float power = 0; for (int i = 0; i < stf_len; ++i) power += x[i].re * x[i].re + x[i].im * x[i].im;
            rep->rms_array[a] = sqrtf(power / (float)stf_len);
            cf_t C = {0, 0};
            for (int i = 0; i + 1 < n_pat; ++i) { const cf_t v = conj_dot(x + i * pat, x + (i + 1) * pat, pat); C.re += uw[i] * v.re; C.im += uw[i] * v.im; }
            wcfo += peak_metric[a] * atan2f(C.im, C.re) / (float)pat; ms2 += peak_metric[a];
        }
        rep->cfo_fractional_rad = wcfo / ms2;
        rep->cfo_integer_rad = 0.0f;   /* coarse_peak_f_domain.cpp:94-201: not implemented in the reference */

        /* ---- crosscorrelator_t on the strongest antenna (ALL_ANTENNAS switch is off, sync_param.hpp:286) */
        /* sync_chunk.cpp:204-205 -> rx_pacer.cpp:311-318: local (resampled) time -> hardware-rate time */
        rep->coarse_peak_time_hw = (int)round((double)rep->coarse_peak_time_local * (double)L / (double)M);
        const int sl = 16 * b * os * L / M, sr = sl, n_lag = sl + 1 + sr;   /* crosscorrelator.cpp:52-56 */
        const int w0 = rep->coarse_peak_time_hw - sl;
        if (w0 < 0 || w0 + n_lag + tpl_len - 1 > len_hw) { ret = -2; break; }
        int amax = 0; for (int a = 1; a < n_ant; ++a) if (rep->coarse_peak_array[a] > rep->coarse_peak_array[amax]) amax = a;
        cf_t* mix = (cf_t*)malloc(sizeof(cf_t) * (n_lag + tpl_len));
        /* crosscorrelator.cpp:97-105: CFO per hardware sample */
        const float inc = (rep->cfo_fractional_rad + rep->cfo_integer_rad) * (float)M / (float)L;
        for (int i = 0; i < n_lag + tpl_len - 1; ++i) {
            const cf_t v = iq_hw[amax * stride_hw + w0 + i];
            const double ph = (double)inc * i; const float c = (float)cos(ph), s = (float)sin(ph);
            mix[i].re = v.re * c - v.im * s; mix[i].im = v.re * s + v.im * c;
        }
        float best_sum = 0.0f; int best_tpl = 0, best_idx = 0;
        for (int tpl = 0; tpl < n_tpl; ++tpl) {
            float mmax = -1.0f; int imax = 0;
            for (int j = 0; j < n_lag; ++j) {
                const cf_t v = conj_dot(mix + j, templates + (long)tpl * tpl_len, tpl_len);
                const float m2 = v.re * v.re + v.im * v.im;
                if (m2 > mmax) { mmax = m2; imax = j; }
            }
            const float metric = sqrtf(mmax);
            if (best_sum < metric) { best_sum = metric; best_tpl = tpl; best_idx = imax; }
        }
        free(mix);
        rep->N_eff_TX = 1 << best_tpl;
        rep->fine_peak_time_local = w0 + best_idx;   /* = round(metric * idx / metric) for one antenna */
        rep->detected = 1;
        (void)skip_after_peak; (void)ignore_before;
    }
    for (int a = 0; a < n_ant; ++a) { free(mc[a].reg); free(mp[a].reg); }
    return ret;
}

int orc_sync_search(const cf_t* iq, long stride, int len, int n_ant, int u, int b, int os, const cf_t* templates,
                    int n_tpl, int search_length, double samp_rate, orc_sync_report_t* rep) {
    const int stf_len = (u == 1 ? 112 : 144) * b * os;
    return orc_sync_search2(iq, stride, len, n_ant, u, b, os, templates, n_tpl, search_length, samp_rate, NULL, 0, 0,
                            1, 1, stf_len, rep);
}

static cf_t conj_dot(const cf_t* a, const cf_t* b, int n) {  /* sum a * conj(b) */
    float re = 0, im = 0;
    for (int i = 0; i < n; ++i) { re += a[i].re * b[i].re + a[i].im * b[i].im; im += a[i].im * b[i].re - a[i].re * b[i].im; }
    cf_t r = {re, im}; return r;
}

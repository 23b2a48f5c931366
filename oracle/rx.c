/* oracle/rx.c -- TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  CPU restatement of the reference's
 * synchronised receiver for one packet, written as the same sequential state machine the reference runs
 * (absolute OFDM symbol index, processing stage index, relative symbol index, chestim buffers that are
 * overwritten by every interpolation), so that the CUDA path's closed-form indexing is checked against
 * the original bookkeeping and not against itself.
 *
 *   lib/src/phy/rx/rx_synced/rx_synced.cpp:186-323   demoddecod_rx_pcc
 *   lib/src/phy/rx/rx_synced/rx_synced.cpp:325-436   demoddecod_rx_pdc
 *   lib/src/phy/rx/rx_synced/rx_synced.cpp:503-1390  run_stf ... run_pxx_mode_transmit_diversity
 *   lib/src/phy/rx/rx_synced/channel_estimation/channel_lut.cpp:169-634 (Wiener LUT)
 *   lib/src/phy/rx/rx_synced/offsets/estimator_sto.cpp, snr/estimator_snr.cpp
 *
 * PARITY UNPINNED: the reference cannot be built here (srsRAN, VOLK, FFTW, Eigen are absent) and its
 * tests hold no golden vectors for this path.  The soft demapper restates srsRAN 23.x
 * lib/src/phy/modem/demod_soft.c (demod_*_lte_s, scalar branch) from its published algorithm.
 * The FFT is a plain mixed radix 2/3 DFT evaluated in double precision.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct { float re, im; } cf_t;
typedef struct { double re, im; } cd_t;

static inline cf_t cf(float a, float b) { cf_t r = {a, b}; return r; }
static inline cf_t cf_add(cf_t a, cf_t b) { return cf(a.re + b.re, a.im + b.im); }
static inline cf_t cf_sub(cf_t a, cf_t b) { return cf(a.re - b.re, a.im - b.im); }
static inline cf_t cf_mul(cf_t a, cf_t b) { return cf(a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re); }
static inline cf_t cf_mulc(cf_t a, cf_t b) { return cf(a.re * b.re + a.im * b.im, a.im * b.re - a.re * b.im); }
static inline cf_t cf_scale(cf_t a, float s) { return cf(a.re * s, a.im * s); }
static inline cf_t cf_div(cf_t a, cf_t b) {
    const float d = b.re * b.re + b.im * b.im;
    return cf_scale(cf_mulc(a, b), 1.0f / d);
}

/* ---- forward DFT, N = 2^a 3^b, decimation in time, double precision ---------------------------- */
static void dft_rec(const cd_t* in, int stride, cd_t* out, int N, const cd_t* tw, int twstride) {
    if (N == 1) { out[0] = in[0]; return; }
    const int r = (N % 2 == 0) ? 2 : 3;
    const int M = N / r;
    for (int q = 0; q < r; ++q) dft_rec(in + q * stride, stride * r, out + q * M, M, tw, twstride * r);
    /* combine: X[k + p*M] = sum_q W_N^{q (k + p M)} Xq[k] */
    cd_t tmp[3];
    for (int k = 0; k < M; ++k) {
        for (int p = 0; p < r; ++p) {
            double sr = 0, si = 0;
            for (int q = 0; q < r; ++q) {
                const long idx = ((long)q * (k + p * M)) % N;
                const cd_t w = tw[idx * twstride];
                const cd_t x = out[q * M + k];
                sr += w.re * x.re - w.im * x.im;
                si += w.re * x.im + w.im * x.re;
            }
            tmp[p].re = sr; tmp[p].im = si;
        }
        for (int p = 0; p < r; ++p) out[p * M + k] = tmp[p];
    }
}

typedef struct { int N; cd_t* tw; cd_t* a; cd_t* b; } fft_t;

static fft_t* fft_new(int N) {
    fft_t* f = (fft_t*)calloc(1, sizeof(fft_t));
    f->N = N;
    f->tw = (cd_t*)malloc(sizeof(cd_t) * N);
    f->a = (cd_t*)malloc(sizeof(cd_t) * N);
    f->b = (cd_t*)malloc(sizeof(cd_t) * N);
    for (int k = 0; k < N; ++k) {
        f->tw[k].re = cos(-2.0 * M_PI * k / N);
        f->tw[k].im = sin(-2.0 * M_PI * k / N);
    }
    return f;
}
static void fft_free(fft_t* f) { free(f->tw); free(f->a); free(f->b); free(f); }
static void fft_run(fft_t* f, const cf_t* in, cf_t* out) {
    for (int i = 0; i < f->N; ++i) { f->a[i].re = in[i].re; f->a[i].im = in[i].im; }
    dft_rec(f->a, 1, f->b, f->N, f->tw, 1);
    for (int i = 0; i < f->N; ++i) out[i] = cf((float)f->b[i].re, (float)f->b[i].im);
}

/* ---- Wiener LUT (channel_lut.cpp:232-634, channel_statistics.cpp:29-35, bessel.cpp:28-62,
 * rectangular.cpp:29, wiener.hpp:113-139) -------------------------------------------------------- */
typedef struct {
    int n_interp, n_t, Nk, n_vec;
    uint32_t* idx_pilot;   /* [n_t][4][Nk] */
    uint32_t* idx_weight;  /* [n_t][4][Nk] */
    float* weights;        /* [n_vec][n_interp] */
} orc_wiener_t;

static float flt_sinc(float n) { return n == 0.0f ? 1.0f : (float)(sin(M_PI * n) / (M_PI * n)); }
static float flt_bessel_j0(float z) {
    /* series sum_{k=0}^{8} (-z^2/4)^k / (k!)^2 */
    double ret = 0.0, fac = 1.0;
    for (int k = 0; k <= 8; ++k) {
        if (k > 0) fac *= (double)k;
        ret += pow(-0.25 * (double)z * (double)z, (double)k) / (fac * fac);
    }
    return (float)ret;
}
static float r_f_uni(float tau_rms_sec, float delta_f) { return flt_sinc((float)M_PI * tau_rms_sec * delta_f); }
static float r_t_jakes(float nu_max_hz, float delta_t) { return flt_bessel_j0(2.0f * (float)M_PI * nu_max_hz * delta_t); }

/* least squares solution of the (full rank, symmetric positive definite) system by Cholesky */
static int chol_solve(const double* A, const double* rhs, double* x, int n) {
    double L[14 * 14];
    for (int i = 0; i < n; ++i)
        for (int j = 0; j <= i; ++j) {
            double s = A[i * n + j];
            for (int k = 0; k < j; ++k) s -= L[i * n + k] * L[j * n + k];
            if (i == j) {
                if (s <= 0.0) return -1;
                L[i * n + i] = sqrt(s);
            } else
                L[i * n + j] = s / L[j * n + j];
        }
    double y[14];
    for (int i = 0; i < n; ++i) {
        double s = rhs[i];
        for (int k = 0; k < i; ++k) s -= L[i * n + k] * y[k];
        y[i] = s / L[i * n + i];
    }
    for (int i = n - 1; i >= 0; --i) {
        double s = y[i];
        for (int k = i + 1; k < n; ++k) s -= L[k * n + i] * x[k];
        x[i] = s / L[i * n + i];
    }
    return 0;
}

static int drs_k(int b, int i, int off) {  /* subcarrier index 0..N_OCC of DRS pilot i with offset off */
    const int q = 4 * i + off;
    return q < 28 * b ? q : q + 1;
}

void orc_wiener_free(orc_wiener_t* L) {
    if (!L) return;
    free(L->idx_pilot); free(L->idx_weight); free(L->weights); free(L);
}

orc_wiener_t* orc_wiener_build(int b, int u, int N_step_virtual, double nu_max_hz, double tau_rms_sec,
                               double snr_db, int n_interp) {
    const int N_OCC = 56 * b, Nk = N_OCC + 1, NP = 14 * b, n_t = N_step_virtual + 1;
    const double delta_u_f = 27000.0 * u;
    const double T_u_symb = (64.0 + 8.0) / 64.0 / delta_u_f;
    const double sigma = 1.0 / pow(10.0, snr_db / 10.0);
    orc_wiener_t* L = (orc_wiener_t*)calloc(1, sizeof(*L));
    L->n_interp = n_interp; L->n_t = n_t; L->Nk = Nk;
    L->idx_pilot = (uint32_t*)calloc((size_t)n_t * 4 * Nk, 4);
    L->idx_weight = (uint32_t*)calloc((size_t)n_t * 4 * Nk, 4);
    int cap = 1024;
    L->weights = (float*)malloc(sizeof(float) * cap * n_interp);
    const int n_coord = N_step_virtual > 0 ? 2 * NP : NP;
    double* cf_ = (double*)malloc(sizeof(double) * n_coord);
    double* ct_ = (double*)malloc(sizeof(double) * n_coord);
    double Rpp[14 * 14], rdp[14], w[14];
    for (int ts = 0; ts < 4; ++ts) {
        /* s0: left symbol offsets ts, right symbol offsets (ts + 2) % 4 (drs.cpp advance rule) */
        for (int p = 0; p < NP; ++p) {
            const double kl = drs_k(b, p, ts), kr = drs_k(b, p, (ts + 2) % 4);
            if (N_step_virtual == 0) { cf_[p] = kl; ct_[p] = 1.0; }
            else if (ts <= 1) { cf_[2 * p] = kl; ct_[2 * p] = 1.0; cf_[2 * p + 1] = kr; ct_[2 * p + 1] = 1.0 + N_step_virtual; }
            else { cf_[2 * p] = kr; ct_[2 * p] = 1.0 + N_step_virtual; cf_[2 * p + 1] = kl; ct_[2 * p + 1] = 1.0; }
        }
        for (int t = 1; t <= n_t; ++t) {
            uint32_t prev = 0;
            for (int f = 0; f < Nk; ++f) {
                /* s1 */
                double min_sum = 1e9;
                uint32_t opt = 0;
                for (uint32_t i = prev; i + n_interp <= (uint32_t)n_coord; ++i) {
                    double s = 0;
                    for (uint32_t j = i; j < i + n_interp; ++j)
                        s += sqrt(pow((double)f - cf_[j], 2.0) + pow((double)t - ct_[j], 2.0));
                    if (s < min_sum) { min_sum = s; opt = i; }
                    if (s > min_sum * 1.1) break;
                }
                /* s2 (Rpp is rebuilt every time here; the reference caches it, same values) */
                for (int r = 0; r < n_interp; ++r) {
                    for (int c = 0; c < n_interp; ++c) {
                        const double df = cf_[opt + c] - cf_[opt + r], dt = ct_[opt + c] - ct_[opt + r];
                        const double v = (double)r_f_uni((float)tau_rms_sec, (float)(df * delta_u_f)) *
                                         (double)r_t_jakes((float)nu_max_hz, (float)(dt * T_u_symb));
                        Rpp[r * n_interp + c] = v + (r == c ? sigma : 0.0);
                    }
                    const double df = cf_[opt + r] - (double)f, dt = ct_[opt + r] - (double)t;
                    rdp[r] = (double)r_f_uni((float)tau_rms_sec, (float)(df * delta_u_f)) *
                             (double)r_t_jakes((float)nu_max_hz, (float)(dt * T_u_symb));
                }
                if (chol_solve(Rpp, rdp, w, n_interp) != 0) { orc_wiener_free(L); free(cf_); free(ct_); return NULL; }
                double norm = 0;
                for (int r = 0; r < n_interp; ++r) norm += w[r];
                float wf[14];
                for (int r = 0; r < n_interp; ++r) wf[r] = (float)(w[r] / norm);
                /* s3 */
                int known = -1;
                for (int v = 0; v < L->n_vec && known < 0; ++v) {
                    float md = 0.f, sub_at_max = 0.f;
                    for (int r = 0; r < n_interp; ++r) {
                        const float sub = wf[r] - L->weights[v * n_interp + r];
                        if (fabsf(sub) > md) { md = fabsf(sub); sub_at_max = sub; }
                    }
                    if (fabs((double)sub_at_max) < 1e-4) known = v;
                }
                if (known < 0) {
                    if (L->n_vec == cap) { cap *= 2; L->weights = (float*)realloc(L->weights, sizeof(float) * cap * n_interp); }
                    memcpy(L->weights + (size_t)L->n_vec * n_interp, wf, sizeof(float) * n_interp);
                    known = L->n_vec++;
                }
                const size_t o = ((size_t)(t - 1) * 4 + ts) * Nk + f;
                L->idx_pilot[o] = opt;
                L->idx_weight[o] = (uint32_t)known;
                prev = opt;
            }
        }
    }
    free(cf_); free(ct_);
    return L;
}

/* ---- receiver configuration (tables come from oracle/spec.py) and state ------------------------- */
typedef struct {
    int32_t b, u, N_eff_TX, N_RX, N_os, cp_os, stf_cp_os, n_pat;
    int32_t N_DF, N_bps, chestim_lr, stride, n_pdc, literal_neff8;
    int32_t N_SS;              /* spatial streams of the PDC (1 = MRC / transmit diversity like the reference) */
    const int8_t* stf_sign;   /* [Nk] sign s of the STF cell value s*(1+j)/sqrt(2), 0 = no STF cell */
    const uint8_t* cell_type;  /* [(N_DF+1)][Nk] 0 none, 1 DRS, 2 PCC, 3 PDC */
    const int8_t* cell_ts;     /* DRS cells: transmit stream */
    const int8_t* cell_sign;   /* DRS cells: +-1 */
    const int32_t* pcc_l; const int32_t* pcc_k;   /* [98] in collection order */
    const int32_t* pdc_l; const int32_t* pdc_k;   /* [n_pdc] in collection order */
    orc_wiener_t* lut_l[3];
    orc_wiener_t* lut_lr[3];
} orc_rx_cfg_t;

typedef struct {
    float cfo_left, sto_inc, sto_fractional, snr_db_pcc, snr_db_pdc;
    float rms[8];
    int32_t n_cls;
    uint8_t cls[512];          /* LUT class picked after each OFDM symbol with DRS cells */
} orc_rx_report_t;

typedef struct {
    const orc_rx_cfg_t* c;
    int Nk, NP, N_OCC, N_step, ps_len, stf_len;
    const cf_t* iq; long iq_stride; long rd;     /* read cursor relative to the window start */
    double mix_phase; float mix_inc;
    fft_t* fft;
    cf_t* mixer_stage;   /* [N_RX][stf_len] */
    cf_t* fft_stage;     /* [N_os] */
    cf_t* stage;         /* [ps_len][N_RX][Nk] processing stage */
    cf_t** now;          /* [N_RX] ofdm_symbol_now */
    cf_t* zf;            /* [N_RX][8][NP]  chestim_drs_zf */
    cf_t* zfi;           /* [N_RX][8][2 NP] chestim_drs_zf_interlaced */
    cf_t* chestim;       /* [N_RX][8][Nk] */
    int ps_idx, ofdm_symb_ps_idx, ofdm_symb_idx, chestim_mode_lr, TS_idx_first, TS_idx_last;
    float SpN_sum, N_sum; uint32_t SpN_cnt, N_cnt;
    int lut_cls;
    float sto_inc;
    int PCC_idx, PDC_subc_idx, PDC_bits_idx;
    int16_t* llr_pcc; int16_t* llr_pdc;
    float* sm_rho;       /* [n_pdc][N_SS] spatial multiplexing: loss of every (cell, stream) against the matched filter bound */
    cf_t* Y_out;         /* optional [N_RX][N_DF][Nk] */
    cf_t* H_out;         /* optional [N_RX][N_eff_TX][N_DF][Nk]: chestim at the time symbol l was collected */
    orc_rx_report_t* rep;
} rxs_t;

static int N_step_of(int N_eff_TX) { return N_eff_TX <= 2 ? 5 : 10; }
static int ps_len_of(int N_eff_TX) { return N_eff_TX <= 2 ? 6 : (N_eff_TX == 4 ? 11 : 12); }

/* rx_synced.cpp:712-740 + mixer: y[n] = x[n] * exp(j (phase + inc n)), phase continuous */
static void run_mix(rxs_t* s, int n_cp) {
    const int len = n_cp + s->c->N_os;
    for (int a = 0; a < s->c->N_RX; ++a) {
        const cf_t* x = s->iq + (long)a * s->iq_stride + s->rd;
        cf_t* y = s->mixer_stage + (long)a * s->stf_len;
        for (int n = 0; n < len; ++n) {
            const double ph = s->mix_phase + (double)s->mix_inc * n;
            y[n] = cf_mul(x[n], cf((float)cos(ph), (float)sin(ph)));
        }
    }
    s->mix_phase += (double)s->mix_inc * len;
    s->mix_phase = fmod(s->mix_phase, 2.0 * M_PI);
    s->rd += len;
    for (int a = 0; a < s->c->N_RX; ++a)
        s->now[a] = s->stage + ((long)s->ofdm_symb_ps_idx * s->c->N_RX + a) * s->Nk;
}

/* estimator_sto_t::apply_full_phase_rotation (estimator_sto.cpp:103-114, 163-171) */
static void sto_rotate(rxs_t* s) {
    const float start = -s->sto_inc * (float)((s->Nk - 1) / 2);
    for (int a = 0; a < s->c->N_RX; ++a)
        for (int k = 0; k < s->Nk; ++k) {
            const double ph = (double)start + (double)s->sto_inc * k;
            s->now[a][k] = cf_mul(s->now[a][k], cf((float)cos(ph), (float)sin(ph)));
        }
}

/* rx_synced.cpp:742-771 */
static void run_cp_fft_scale(rxs_t* s, int n_cp) {
    const int N_os = s->c->N_os, h = s->N_OCC / 2;
    const float scale = sqrtf((float)s->N_OCC) / (float)N_os;
    for (int a = 0; a < s->c->N_RX; ++a) {
        fft_run(s->fft, s->mixer_stage + (long)a * s->stf_len + n_cp, s->fft_stage);
        for (int k = 0; k <= h; ++k) s->now[a][h + k] = cf_scale(s->fft_stage[k], scale);
        for (int k = 0; k < h; ++k) s->now[a][k] = cf_scale(s->fft_stage[N_os - h + k], scale);
    }
    sto_rotate(s);
}

/* estimator_snr_t::process_stf_or_drs_packed (estimator_snr.cpp:105-145) */
static void snr_packed(rxs_t* s, const cf_t* z, int n) {
    float p = 0.f, q = 0.f;
    for (int i = 0; i < n; ++i) p += z[i].re * z[i].re + z[i].im * z[i].im;
    for (int i = 0; i + 1 < n; ++i) {
        const cf_t d = cf_sub(z[i], z[i + 1]);
        q += d.re * d.re + d.im * d.im;
    }
    s->SpN_sum += p; s->SpN_cnt += n;
    s->N_sum += q / 2.0f; s->N_cnt += n - 1;
}
static float snr_db_now(const rxs_t* s) {  /* estimator_snr.cpp:88-97 */
    if (s->SpN_sum <= 0.0f || s->N_sum <= 0.0f) return 0.0f;
    const float S_avg = (s->SpN_sum - s->N_sum) / (float)s->SpN_cnt;
    const float N_avg = s->N_sum / (float)s->N_cnt;
    return 10.0f * log10f(S_avg / N_avg);
}

/* rx_synced.cpp:863-891 */
static void run_drs_channel_lut_pick(rxs_t* s) {
    static const float snr_ref[3] = {-5.0f, 15.0f, 35.0f};
    const float snr_now = snr_db_now(s);
    int idx = 0;
    float diff = fabsf(snr_now - snr_ref[0]);
    for (int i = 1; i < 3; ++i) {
        const float d = fabsf(snr_now - snr_ref[i]);
        if (d <= diff) { diff = d; idx = i; }
    }
    s->lut_cls = idx;
    if (s->rep->n_cls < 512) s->rep->cls[s->rep->n_cls++] = (uint8_t)idx;
}

/* rx_synced.cpp:661-709 */
static void run_stf_chestim_zf(rxs_t* s) {
    const int n_stf = s->N_OCC / 4;
    for (int a = 0; a < s->c->N_RX; ++a) {
        cf_t* z = s->zf + ((long)a * 8 + 0) * s->NP;
        int r = 0, w = 0;
        for (int i = 0; i < n_stf; ++i) {
            const float sg = (float)s->c->stf_sign[r] * (float)M_SQRT1_2;
            z[w++] = cf_div(s->now[a][r], cf(sg, sg));
            r += (i == n_stf / 2 - 1) ? 8 : 4;
        }
    }
}

/* rx_synced.cpp:503-619 with the compile-time switches of rx_synced_param.hpp as shipped */
static void run_stf(rxs_t* s) {
    const orc_rx_cfg_t* c = s->c;
    run_mix(s, c->stf_cp_os);
    const int pat_len = s->stf_len / c->n_pat;
    /* run_stf_rms_estimation: whole STF, values provided by sync are kept */
    for (int a = 0; a < c->N_RX; ++a) {
        if (s->rep->rms[a] > 0.0f) continue;
        const cf_t* y = s->mixer_stage + (long)a * s->stf_len;
        float p = 0.f;
        for (int n = 0; n < s->stf_len; ++n) p += y[n].re * y[n].re + y[n].im * y[n].im;
        s->rep->rms[a] = sqrtf(p / (float)s->stf_len);
    }
    /* cover sequence (stf.cpp apply_cover_sequence) */
    static const float cover[9] = {1, -1, 1, 1, -1, -1, -1, -1, -1};
    for (int a = 0; a < c->N_RX; ++a) {
        cf_t* y = s->mixer_stage + (long)a * s->stf_len;
        for (int n = 0; n < s->stf_len; ++n) y[n] = cf_scale(y[n], cover[n / pat_len]);
    }
    /* fractional CFO adjustment from the repeating patterns */
    float sr = 0.f, si = 0.f;
    for (int a = 0; a < c->N_RX; ++a) {
        const cf_t* y = s->mixer_stage + (long)a * s->stf_len;
        for (int j = 0; j + 1 < c->n_pat; ++j)
            for (int n = 0; n < pat_len; ++n) {
                const cf_t v = cf_mulc(y[j * pat_len + n], y[(j + 1) * pat_len + n]);
                sr += v.re; si += v.im;
            }
    }
    s->rep->cfo_left = atan2f(si, sr) / (float)pat_len;
    s->mix_inc += s->rep->cfo_left;
    run_cp_fft_scale(s, c->stf_cp_os);
    run_stf_chestim_zf(s);
    /* estimator_sto_t::process_stf (estimator_sto.cpp:47-60, 125-147) */
    const int n_stf = s->N_OCC / 4;
    float inc = 0.f;
    for (int a = 0; a < c->N_RX; ++a) {
        const cf_t* z = s->zf + ((long)a * 8 + 0) * s->NP;
        float br = 0.f, bi = 0.f;
        for (int i = 0; i + 1 < n_stf; ++i) {
            cf_t v = cf_mulc(z[i], z[i + 1]);
            if (i == n_stf / 2 - 1) {
                const float arg = atan2f(v.im, v.re);
                v = cf_mul(v, cf(cosf(-arg / 2.0f), sinf(-arg / 2.0f)));
            }
            br += v.re; bi += v.im;
        }
        inc += atan2f(bi, br) / 4.0f;
    }
    s->sto_inc = inc / (float)c->N_RX;
    s->rep->sto_inc = s->sto_inc;
    sto_rotate(s);
    s->rep->sto_fractional = s->sto_inc / 2.0f / (float)M_PI * (float)c->N_os;
    run_stf_chestim_zf(s);
    /* estimator_snr_t::process_stf (estimator_snr.cpp:48-66) */
    for (int a = 0; a < c->N_RX; ++a) snr_packed(s, s->zf + ((long)a * 8 + 0) * s->NP, n_stf);
    const float S = s->SpN_sum - s->N_sum;
    s->SpN_sum = S / 4.0f + s->N_sum;
}

/* rx_synced.cpp:773-861 */
static void run_drs_chestim_zf(rxs_t* s) {
    const orc_rx_cfg_t* c = s->c;
    const int l = s->ofdm_symb_idx;
    const uint8_t* type = c->cell_type + (long)l * s->Nk;
    const int8_t* tsv = c->cell_ts + (long)l * s->Nk;
    const int8_t* sgn = c->cell_sign + (long)l * s->Nk;
    int first = 99, last = -1;
    for (int k = 0; k < s->Nk; ++k)
        if (type[k] == 1) { if (tsv[k] < first) first = tsv[k]; if (tsv[k] > last) last = tsv[k]; }
    s->TS_idx_first = first; s->TS_idx_last = last;
    /* channel_antenna_t::get_write_offsets (channel_antenna.hpp:45-71) */
    static const int t0011[8] = {0, 0, 1, 1, 0, 0, 1, 1}, t1100[8] = {1, 1, 0, 0, 1, 1, 0, 0};
    const int* woff = (s->ps_idx % 2 == 0) ? (s->ofdm_symb_ps_idx <= 1 ? t0011 : t1100)
                                           : (s->ofdm_symb_ps_idx <= 1 ? t1100 : t0011);
    for (int a = 0; a < c->N_RX; ++a)
        for (int ts = first; ts <= last; ++ts) {
            cf_t* z = s->zf + ((long)a * 8 + ts) * s->NP;
            cf_t* zi = s->zfi + ((long)a * 8 + ts) * 2 * s->NP;
            int w = 0, wi = woff[ts];
            for (int k = 0; k < s->Nk; ++k) {
                if (type[k] != 1 || tsv[k] != ts) continue;
                const cf_t A = cf_div(s->now[a][k], cf((float)sgn[k], 0.0f));
                z[w++] = A;
                zi[wi] = A;
                wi += 2;
            }
        }
    /* estimator_snr_t::process_drs (estimator_snr.cpp:68-86) */
    for (int a = 0; a < c->N_RX; ++a)
        for (int ts = first; ts <= (last < 7 ? last : 7); ++ts) snr_packed(s, s->zf + ((long)a * 8 + ts) * s->NP, s->NP);
    run_drs_channel_lut_pick(s);
}

/* rx_synced.cpp:893-960, channel_lut.cpp:71-150 */
static void run_drs_ch_interpolation(rxs_t* s) {
    const orc_rx_cfg_t* c = s->c;
    const orc_wiener_t* L = s->chestim_mode_lr ? c->lut_lr[s->lut_cls] : c->lut_l[s->lut_cls];
    int t = s->ofdm_symb_ps_idx;
    if (t >= L->n_t) t = L->n_t - 1;   /* N_eff_TX = 8, mode l: the reference reads past its one-symbol LUT */
    const int first = s->chestim_mode_lr ? 0 : s->TS_idx_first;
    for (int a = 0; a < c->N_RX; ++a)
        for (int ts = first; ts <= s->TS_idx_last; ++ts) {
            int row = ts % 4;
            if (s->ps_idx % 2 == 1) row = (row + 2) % 4;   /* swap_upper_lower_half */
            const uint32_t* ip = L->idx_pilot + ((long)t * 4 + row) * s->Nk;
            const uint32_t* iw = L->idx_weight + ((long)t * 4 + row) * s->Nk;
            const cf_t* z = s->chestim_mode_lr ? s->zfi + ((long)a * 8 + ts) * 2 * s->NP
                                               : s->zf + ((long)a * 8 + ts) * s->NP;
            cf_t* h = s->chestim + ((long)a * 8 + ts) * s->Nk;
            for (int k = 0; k < s->Nk; ++k) {
                const float* w = L->weights + (long)iw[k] * L->n_interp;
                float re = 0.f, im = 0.f;
                for (int j = 0; j < L->n_interp; ++j) { re += w[j] * z[ip[k] + j].re; im += w[j] * z[ip[k] + j].im; }
                h[k] = cf(re, im);
            }
        }
}

/* ---- soft demapper, srsRAN demod_soft.c demod_{bpsk,qpsk,16qam,64qam,256qam}_lte_s restated ------- */
static int16_t to_short(float v) {  /* C cast semantics (toward zero); out of range is UB in C, saturate */
    v = truncf(v);
    if (v > 32767.f) v = 32767.f;
    if (v < -32768.f) v = -32768.f;
    return (int16_t)v;
}
static int16_t neg_short(int16_t v) { return v == -32768 ? 32767 : (int16_t)-v; }
static int abs_i(int v) { return v < 0 ? -v : v; }

static void demap(cf_t z, int N_bps, int16_t* out) {
    const float SCALE_QPSK = 100.f, SCALE_16 = 400.f, SCALE_64 = 700.f, SCALE_256 = 1000.f;
    /* 1024-QAM (MCS 10, 11) has no srsRAN counterpart (tx.cpp:53-60 stops at 256-QAM): own extension of the same
     * recursion, scale 1300 continuing 100 sqrt(2) / 400 / 700 / 1000, thresholds 16, 8, 4, 2 over sqrt(682) */
    const float SCALE_1024 = 1300.f;
    switch (N_bps) {
        case 10: {
            const int16_t yre = to_short(SCALE_1024 * z.re), yim = to_short(SCALE_1024 * z.im);
            out[0] = neg_short(yre); out[1] = neg_short(yim);
            out[2] = to_short((float)abs_i(yre) - 16.f * SCALE_1024 / sqrtf(682.f));
            out[3] = to_short((float)abs_i(yim) - 16.f * SCALE_1024 / sqrtf(682.f));
            for (int i = 4, t = 8; i < 10; i += 2, t /= 2) {
                out[i] = to_short((float)abs_i(out[i - 2]) - (float)t * SCALE_1024 / sqrtf(682.f));
                out[i + 1] = to_short((float)abs_i(out[i - 1]) - (float)t * SCALE_1024 / sqrtf(682.f));
            }
        } break;
        case 1: out[0] = to_short(-(SCALE_QPSK * (z.re + z.im) * (float)M_SQRT1_2)); break;
        case 2:
            out[0] = to_short(z.re * (-SCALE_QPSK * sqrtf(2.f)));
            out[1] = to_short(z.im * (-SCALE_QPSK * sqrtf(2.f)));
            break;
        case 4: {
            const int16_t yre = to_short(SCALE_16 * z.re), yim = to_short(SCALE_16 * z.im);
            out[0] = neg_short(yre); out[1] = neg_short(yim);
            out[2] = to_short((float)abs_i(yre) - 2.f * SCALE_16 / sqrtf(10.f));
            out[3] = to_short((float)abs_i(yim) - 2.f * SCALE_16 / sqrtf(10.f));
        } break;
        case 6: {
            const int16_t yre = to_short(SCALE_64 * z.re), yim = to_short(SCALE_64 * z.im);
            out[0] = neg_short(yre); out[1] = neg_short(yim);
            out[2] = to_short((float)abs_i(yre) - 4.f * SCALE_64 / sqrtf(42.f));
            out[3] = to_short((float)abs_i(yim) - 4.f * SCALE_64 / sqrtf(42.f));
            out[4] = to_short((float)abs_i(out[2]) - 2.f * SCALE_64 / sqrtf(42.f));
            out[5] = to_short((float)abs_i(out[3]) - 2.f * SCALE_64 / sqrtf(42.f));
        } break;
        default: {
            const int16_t yre = to_short(SCALE_256 * z.re), yim = to_short(SCALE_256 * z.im);
            out[0] = neg_short(yre); out[1] = neg_short(yim);
            out[2] = to_short((float)abs_i(yre) - 8.f * SCALE_256 / sqrtf(170.f));
            out[3] = to_short((float)abs_i(yim) - 8.f * SCALE_256 / sqrtf(170.f));
            out[4] = to_short((float)abs_i(out[2]) - 4.f * SCALE_256 / sqrtf(170.f));
            out[5] = to_short((float)abs_i(out[3]) - 4.f * SCALE_256 / sqrtf(170.f));
            out[6] = to_short((float)abs_i(out[4]) - 2.f * SCALE_256 / sqrtf(170.f));
            out[7] = to_short((float)abs_i(out[5]) - 2.f * SCALE_256 / sqrtf(170.f));
        } break;
    }
}

/* transmit_diversity_precoding.cpp:45-72 index_mat_N_TS_x */
static void txdiv_pair(int N_TS, int i_mod, int* A, int* B) {
    static const int p4[6][2] = {{0, 1}, {2, 3}, {0, 2}, {1, 3}, {0, 3}, {1, 2}};
    static const int p8[12][2] = {{0, 1}, {2, 3}, {4, 5}, {6, 7}, {0, 4}, {1, 5}, {2, 6}, {3, 7}, {0, 2}, {1, 3}, {4, 6}, {5, 7}};
    if (N_TS == 2) { *A = 0; *B = 1; }
    else if (N_TS == 4) { *A = p4[i_mod][0]; *B = p4[i_mod][1]; }
    else { *A = p8[i_mod][0]; *B = p8[i_mod][1]; }
}

/* combine the n cells k[0..n) of the current OFDM symbol: MRC (rx_synced.cpp:962-996, 1204-1273) or
 * Alamouti (rx_synced.cpp:1343-1390); z[n] receives numerator / denominator */
static void combine_cells(rxs_t* s, const int32_t* k, int n, int PXX_idx, cf_t* z) {
    const orc_rx_cfg_t* c = s->c;
    if (c->N_eff_TX == 1) {
        for (int i = 0; i < n; ++i) {
            cf_t num = cf(0, 0), den = cf(0, 0);
            for (int a = 0; a < c->N_RX; ++a) {
                const cf_t y = s->now[a][k[i]], h = s->chestim[((long)a * 8 + 0) * s->Nk + k[i]];
                num = cf_add(num, cf_mulc(y, h));
                den = cf_add(den, cf_mulc(h, h));
            }
            z[i] = cf_div(num, den);
        }
        return;
    }
    const int modulo = c->N_eff_TX == 2 ? 1 : (c->N_eff_TX == 4 ? 6 : 12);
    for (int cnt = 0; cnt < n; cnt += 2) {
        cf_t n0 = cf(0, 0), n1 = cf(0, 0), den = cf(0, 0);
        int A, B;
        txdiv_pair(c->N_eff_TX, ((PXX_idx + cnt) / 2) % modulo, &A, &B);
        for (int a = 0; a < c->N_RX; ++a) {
            const cf_t r0 = s->now[a][k[cnt]], r1 = s->now[a][k[cnt + 1]];
            const cf_t* hA = s->chestim + ((long)a * 8 + A) * s->Nk;
            const cf_t* hB = s->chestim + ((long)a * 8 + B) * s->Nk;
            const cf_t h0 = cf_scale(cf_add(hA[k[cnt]], hA[k[cnt + 1]]), 0.5f);
            const cf_t h1 = cf_scale(cf_add(hB[k[cnt]], hB[k[cnt + 1]]), 0.5f);
            const cf_t r0c = cf(r0.re, -r0.im), r1c = cf(r1.re, -r1.im);
            n0 = cf_add(n0, cf_add(cf_mulc(r0, h0), cf_mul(h1, r1c)));
            n1 = cf_add(n1, cf_sub(cf_mulc(r1, h0), cf_mul(h1, r0c)));
            den = cf_add(den, cf_add(cf_mulc(h0, h0), cf_mulc(h1, h1)));
        }
        z[cnt] = cf_div(n0, den);
        z[cnt + 1] = cf_div(n1, den);
    }
}

/* Spatial multiplexing (transmission modes 2, 4, 6, 8, 9, 11).  The reference has only a stub here
 * (rx_synced.cpp:1331-1333 run_pdc_mode_AxA_MIMO, "todo"), so this is the repo's OWN specification (DESIGN.md
 * "Spatial multiplexing receive"), restated sequentially: per cell zero-forcing through the Cholesky factor of
 * A = H^H H, every stream's LLRs weighted by rho_s = 1 / ([A^-1]_ss A_ss) in (0, 1], the loss against the matched
 * filter bound; rho = 1 for orthogonal columns, where the result equals the reference's MRC arithmetic.
 * Stream s of cell q owns the LLRs (q N_SS + s) N_bps ... (tx.cpp:1054-1060 layer order). */
static void detect_sm(rxs_t* s, const int32_t* k, int n, int PDC_idx) {
    const orc_rx_cfg_t* c = s->c;
    const int N = c->N_SS, R = c->N_RX;
    for (int i = 0; i < n; ++i) {
        cf_t A[8][8], L[8][8], M[8][8], b[8], z[8], x[8];
        float g[8];
        for (int p = 0; p < N; ++p) {
            for (int q = 0; q <= p; ++q) {
                cf_t acc = cf(0, 0);
                for (int a = 0; a < R; ++a)   /* conj(H[a][p]) H[a][q] */
                    acc = cf_add(acc, cf_mulc(s->chestim[((long)a * 8 + q) * s->Nk + k[i]], s->chestim[((long)a * 8 + p) * s->Nk + k[i]]));
                A[p][q] = acc;
            }
            cf_t acc = cf(0, 0);
            for (int a = 0; a < R; ++a) acc = cf_add(acc, cf_mulc(s->now[a][k[i]], s->chestim[((long)a * 8 + p) * s->Nk + k[i]]));
            b[p] = acc;
            g[p] = A[p][p].re;
        }
        for (int j = 0; j < N; ++j) {
            float d = A[j][j].re;
            for (int q = 0; q < j; ++q) d -= L[j][q].re * L[j][q].re + L[j][q].im * L[j][q].im;
            const float floor_ = 1.0e-6f * g[j] + 1.0e-30f;
            if (!(d > floor_)) d = floor_;
            const float inv = 1.0f / sqrtf(d);
            L[j][j] = cf(inv, 0);   /* the diagonal is kept as its reciprocal */
            for (int p = j + 1; p < N; ++p) {
                cf_t acc = A[p][j];
                for (int q = 0; q < j; ++q) acc = cf_sub(acc, cf_mulc(L[p][q], L[j][q]));
                L[p][j] = cf_scale(acc, inv);
            }
        }
        for (int j = 0; j < N; ++j) {   /* M = L^-1, column by column */
            M[j][j] = L[j][j];
            for (int p = j + 1; p < N; ++p) {
                cf_t acc = cf(0, 0);
                for (int q = j; q < p; ++q) acc = cf_add(acc, cf_mul(L[p][q], M[q][j]));
                M[p][j] = cf_scale(acc, -L[p][p].re);
            }
        }
        for (int p = 0; p < N; ++p) {
            cf_t acc = cf(0, 0);
            for (int q = 0; q <= p; ++q) acc = cf_add(acc, cf_mul(M[p][q], b[q]));
            z[p] = acc;
        }
        for (int p = 0; p < N; ++p) {
            cf_t acc = cf(0, 0);
            float v = 0.f;
            for (int q = p; q < N; ++q) {   /* conj(M[q][p]) z[q] */
                acc = cf_add(acc, cf_mulc(z[q], M[q][p]));
                v += M[q][p].re * M[q][p].re + M[q][p].im * M[q][p].im;
            }
            x[p] = acc;
            float rho = 1.0f / (v * g[p] + 1.0e-30f);
            if (!(rho < 1.0f)) rho = 1.0f;
            /* unweighted LLRs now, the weights once the packet's reference loss is known (finish_sm) */
            demap(x[p], c->N_bps, s->llr_pdc + ((long)(PDC_idx + i) * N + p) * c->N_bps);
            s->sm_rho[(long)(PDC_idx + i) * N + p] = rho;
        }
    }
}

/* second half of the spatial multiplexing specification: the packet's reference loss is the mean of rho over every
 * 8th cell (all streams); each LLR is weighted by min(rho / rho_ref, 4) and truncated.  A flat channel (one rho for
 * the whole packet) therefore keeps the demapper's own scale, which is what the decoder's +-1023 input range expects. */
static void finish_sm(rxs_t* s) {
    const orc_rx_cfg_t* c = s->c;
    const int N = c->N_SS;
    float sum = 0.f;
    int cnt = 0;
    for (int q = 0; q < c->n_pdc; q += 8)
        for (int p = 0; p < N; ++p) { sum += s->sm_rho[(long)q * N + p]; ++cnt; }
    const float ref = sum / (float)cnt;
    const float inv = 1.0f / (ref > 1.0e-12f ? ref : 1.0e-12f);
    for (long i = 0; i < (long)c->n_pdc * N; ++i) {
        float w = s->sm_rho[i] * inv;
        if (!(w < 4.0f)) w = 4.0f;
        int16_t* out = s->llr_pdc + i * c->N_bps;
        for (int e = 0; e < c->N_bps; ++e) out[e] = to_short(w * (float)out[e]);
    }
}

static void snapshot_H(rxs_t* s, int l) {
    if (!s->H_out) return;
    const orc_rx_cfg_t* c = s->c;
    for (int a = 0; a < c->N_RX; ++a)
        for (int ts = 0; ts < c->N_eff_TX; ++ts)
            memcpy(s->H_out + ((((long)a * c->N_eff_TX + ts) * c->N_DF) + (l - 1)) * s->Nk,
                   s->chestim + ((long)a * 8 + ts) * s->Nk, sizeof(cf_t) * s->Nk);
}

/* rx_synced.cpp:962-1008; l = absolute symbol index of the cells */
static void run_pcc_collection_and_demapping(rxs_t* s, int l) {
    int32_t k[98];
    cf_t z[98];
    int n = 0;
    while (s->PCC_idx + n < 98 && s->c->pcc_l[s->PCC_idx + n] == l) { k[n] = s->c->pcc_k[s->PCC_idx + n]; ++n; }
    if (n == 0) return;
    combine_cells(s, k, n, s->PCC_idx, z);
    for (int i = 0; i < n; ++i) demap(z[i], 2, s->llr_pcc + 2 * (s->PCC_idx + i));
    s->PCC_idx += n;
}

/* rx_synced.cpp:1154-1337 (the decoder call is outside this function: the oracle hands the complete
 * LLR vector to oracle/fec_chain.c afterwards) */
static void run_pdc(rxs_t* s, int l) {
    const orc_rx_cfg_t* c = s->c;
    int n = 0;
    while (s->PDC_subc_idx + n < c->n_pdc && c->pdc_l[s->PDC_subc_idx + n] == l) ++n;
    if (n == 0) return;
    if (c->N_SS > 1) {
        detect_sm(s, c->pdc_k + s->PDC_subc_idx, n, s->PDC_subc_idx);
    } else {
        cf_t* z = (cf_t*)malloc(sizeof(cf_t) * n);
        combine_cells(s, c->pdc_k + s->PDC_subc_idx, n, s->PDC_subc_idx, z);
        for (int i = 0; i < n; ++i) demap(z[i], c->N_bps, s->llr_pdc + (long)(s->PDC_subc_idx + i) * c->N_bps);
        free(z);
    }
    snapshot_H(s, l);
    s->PDC_subc_idx += n;
    s->PDC_bits_idx += n * c->N_bps * (c->N_SS > 1 ? c->N_SS : 1);
}

static int is_drs_symbol(const rxs_t* s, int l) {
    if (l > s->c->N_DF) return 0;
    const uint8_t* type = s->c->cell_type + (long)l * s->Nk;
    for (int k = 0; k < s->Nk; ++k) if (type[k] == 1) return 1;
    return 0;
}
static void keep_Y(rxs_t* s) {
    if (!s->Y_out) return;
    for (int a = 0; a < s->c->N_RX; ++a)
        memcpy(s->Y_out + (((long)a * s->c->N_DF) + (s->ofdm_symb_idx - 1)) * s->Nk, s->now[a], sizeof(cf_t) * s->Nk);
}

/* rx_synced.cpp:1027-1105 */
static void run_pdc_ps_in_chestim_mode_lr_t(rxs_t* s) {
    const orc_rx_cfg_t* c = s->c;
    const int nof_full_ps = (c->N_DF - (s->ps_len - s->N_step)) / s->N_step;
    if (c->N_DF < s->ps_len - s->N_step || nof_full_ps == 0) return;
    s->chestim_mode_lr = 1;
    while (s->ps_idx < nof_full_ps) {
        const int first_idx = 1 + s->ps_idx * s->N_step;
        const int final_idx = s->ps_len + s->ps_idx * s->N_step;
        if (s->ps_idx > 0) s->ofdm_symb_ps_idx = 1;
        /* DEVIATION (documented in DESIGN.md "N_eff_TX = 8"): with the 12 symbol stage of N_eff_TX = 8
         * the previous stage already consumed first_idx AND first_idx + 1, so the next symbol has the
         * relative index 2.  The reference keeps 1 (rx_synced.cpp:1063-1065), which shifts every
         * later symbol by one row and demaps garbage from stage 1 on; its 8 stream receive path was
         * never exercised (no golden data, asserts in run_drs_chestim_zf).  literal_neff8 = 1
         * reproduces the literal behaviour. */
        if (s->ps_idx > 0 && c->N_eff_TX == 8 && !c->literal_neff8) s->ofdm_symb_ps_idx = 2;
        for (; s->ofdm_symb_idx <= final_idx; ++s->ofdm_symb_idx) {
            run_mix(s, c->cp_os);
            run_cp_fft_scale(s, c->cp_os);
            keep_Y(s);
            if (is_drs_symbol(s, s->ofdm_symb_idx)) run_drs_chestim_zf(s);
            ++s->ofdm_symb_ps_idx;
        }
        const int relative_start = s->ps_idx == 0 ? 0 : 1;
        for (s->ofdm_symb_ps_idx = relative_start; s->ofdm_symb_ps_idx <= s->N_step; ++s->ofdm_symb_ps_idx) {
            for (int a = 0; a < c->N_RX; ++a)
                s->now[a] = s->stage + ((long)s->ofdm_symb_ps_idx * c->N_RX + a) * s->Nk;
            if ((s->ofdm_symb_ps_idx - relative_start) % c->stride == 0) run_drs_ch_interpolation(s);
            run_pdc(s, first_idx + s->ofdm_symb_ps_idx);
        }
        ++s->ps_idx;
        s->ofdm_symb_ps_idx = 0;
    }
}

/* rx_synced.cpp:1107-1152 */
static void run_pdc_ps_in_chestim_mode_lr_f(rxs_t* s) {
    const orc_rx_cfg_t* c = s->c;
    s->chestim_mode_lr = 0;
    const int ps_t_length = s->N_step;
    if (s->ps_idx == 0) {
        s->ofdm_symb_ps_idx = 0;
        for (int i = 1; i < s->ofdm_symb_idx; ++i) {
            run_pdc(s, i);   /* ofdm_symbol_now is NOT moved here, exactly like the reference */
            ++s->ofdm_symb_ps_idx;
        }
    }
    for (; s->ofdm_symb_idx <= c->N_DF; ++s->ofdm_symb_idx) {
        run_mix(s, c->cp_os);
        run_cp_fft_scale(s, c->cp_os);
        keep_Y(s);
        if (is_drs_symbol(s, s->ofdm_symb_idx)) {
            run_drs_chestim_zf(s);
            run_drs_ch_interpolation(s);
        }
        run_pdc(s, s->ofdm_symb_idx);
        ++s->ofdm_symb_ps_idx;
        if (s->ofdm_symb_ps_idx == ps_t_length) {
            ++s->ps_idx;
            s->ofdm_symb_ps_idx = 0;
        }
    }
}

/* One packet: demoddecod_rx_pcc (rx_synced.cpp:186-323) followed by demoddecod_rx_pdc
 * (rx_synced.cpp:325-436) without the FEC calls.  iq: [N_RX][iq_stride], the STF starts at sample t0.
 * Returns 0, or -1 if the cell counters do not add up. */
int orc_rx_packet(const orc_rx_cfg_t* c, const cf_t* iq, long iq_stride, int t0, float cfo_fractional_rad,
                  float cfo_integer_rad, const float* rms_in, int16_t* llr_pcc, int16_t* llr_pdc, cf_t* Y_out,
                  cf_t* H_out, orc_rx_report_t* rep) {
    rxs_t S, *s = &S;
    memset(s, 0, sizeof(S));
    memset(rep, 0, sizeof(*rep));
    s->c = c; s->rep = rep;
    s->N_OCC = 56 * c->b; s->Nk = s->N_OCC + 1; s->NP = 14 * c->b;
    s->N_step = N_step_of(c->N_eff_TX); s->ps_len = ps_len_of(c->N_eff_TX);
    s->stf_len = c->stf_cp_os + c->N_os;
    s->iq = iq; s->iq_stride = iq_stride; s->rd = t0;
    s->mix_phase = 0.0; s->mix_inc = cfo_fractional_rad + cfo_integer_rad;
    s->fft = fft_new(c->N_os);
    s->mixer_stage = (cf_t*)calloc((size_t)c->N_RX * s->stf_len, sizeof(cf_t));
    s->fft_stage = (cf_t*)calloc(c->N_os, sizeof(cf_t));
    s->stage = (cf_t*)calloc((size_t)s->ps_len * c->N_RX * s->Nk, sizeof(cf_t));
    s->now = (cf_t**)calloc(c->N_RX, sizeof(cf_t*));
    s->zf = (cf_t*)calloc((size_t)c->N_RX * 8 * s->NP, sizeof(cf_t));
    s->zfi = (cf_t*)calloc((size_t)c->N_RX * 8 * 2 * s->NP, sizeof(cf_t));
    s->chestim = (cf_t*)calloc((size_t)c->N_RX * 8 * s->Nk, sizeof(cf_t));
    s->llr_pcc = llr_pcc; s->llr_pdc = llr_pdc; s->Y_out = Y_out; s->H_out = H_out;
    if (c->N_SS > 1) s->sm_rho = (float*)calloc((size_t)c->n_pdc * c->N_SS, sizeof(float));
    for (int a = 0; a < 8; ++a) rep->rms[a] = rms_in ? rms_in[a] : 0.0f;

    /* ---- PCC */
    s->chestim_mode_lr = 0;
    s->ps_idx = 0;
    s->ofdm_symb_ps_idx = 0;
    run_stf(s);
    int pcc_symbol_idx_max = 0;
    for (int i = 0; i < 98; ++i) if (c->pcc_l[i] > pcc_symbol_idx_max) pcc_symbol_idx_max = c->pcc_l[i];
    for (s->ofdm_symb_idx = 1; s->ofdm_symb_idx <= pcc_symbol_idx_max; ++s->ofdm_symb_idx) {
        run_mix(s, c->cp_os);
        run_cp_fft_scale(s, c->cp_os);
        keep_Y(s);
        if (is_drs_symbol(s, s->ofdm_symb_idx)) {
            run_drs_chestim_zf(s);
            run_drs_ch_interpolation(s);
        }
        run_pcc_collection_and_demapping(s, s->ofdm_symb_idx);
        ++s->ofdm_symb_ps_idx;
    }
    rep->snr_db_pcc = snr_db_now(s);

    /* ---- PDC */
    if (c->chestim_lr) run_pdc_ps_in_chestim_mode_lr_t(s);
    run_pdc_ps_in_chestim_mode_lr_f(s);
    rep->snr_db_pdc = snr_db_now(s);

    const int ok = s->PCC_idx == 98 && s->PDC_subc_idx == c->n_pdc;
    if (ok && c->N_SS > 1) finish_sm(s);
    free(s->sm_rho);
    fft_free(s->fft);
    free(s->mixer_stage); free(s->fft_stage); free(s->stage); free(s->now); free(s->zf); free(s->zfi); free(s->chestim);
    return ok ? 0 : -1;
}

/* direct access for unit tests */
void orc_demap(const cf_t* z, int n, int N_bps, int16_t* out) {
    for (int i = 0; i < n; ++i) demap(z[i], N_bps, out + (long)i * N_bps);
}
void orc_fft(const cf_t* in, cf_t* out, int N) {
    fft_t* f = fft_new(N);
    fft_run(f, in, out);
    fft_free(f);
}
int orc_wiener_dims(const orc_wiener_t* L, int* n_interp, int* n_t, int* Nk, int* n_vec) {
    *n_interp = L->n_interp; *n_t = L->n_t; *Nk = L->Nk; *n_vec = L->n_vec;
    return 0;
}
const uint32_t* orc_wiener_idx_pilot(const orc_wiener_t* L) { return L->idx_pilot; }
const uint32_t* orc_wiener_idx_weight(const orc_wiener_t* L) { return L->idx_weight; }
const float* orc_wiener_weights(const orc_wiener_t* L) { return L->weights; }

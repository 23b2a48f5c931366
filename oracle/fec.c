/* ORACLE (test infrastructure, NOT product code): plain-C restatement of the FEC chain of the
 * reference's PCC/PDC path.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library.
 *
 * Reference call sites restated (paths relative to /root/reference/lib/src):
 *   phy/fec/pcc_enc.cpp:145-213  (PLCF -> CRC16^mask -> turbo -> RM(196) -> scramble)
 *   phy/fec/pcc_enc.cpp:215-364  (blind PLCF decode, <=5 iterations, 4 CRC masks)
 *   phy/fec/pdc_enc.cpp:127-229  (TB -> CBs -> CRC24A/B -> turbo -> RM(rv) -> scramble at bit offset)
 *   phy/fec/pdc_enc.cpp:291-492  (descramble -> de-RM accumulate -> turbo 2..10 its -> CRC logic)
 *   sections_part3/fix/scrambling.cpp:35-61, sections_part3/fix/cbsegm.cpp:65-123
 * The arithmetic inside srsRAN_4G release_23_11 (crc.c, turbocoder.c, rm_turbo.c, sequence.c,
 * turbodecoder*.c) is NOT in the reference tree; it is restated from TS 36.212 5.1.1/5.1.3/5.1.4 and
 * TS 36.211 7.2.  PARITY UNPINNED against srsRAN itself (no golden vectors exist): the
 * encoder side is pinned by the 3GPP definitions (tests/test_oracle_fec.py holds the KATs), the int16
 * max-log-MAP decoder is this project's own windowed formulation (DESIGN.md "turbo decoder contract").
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define MAXK 6144
#define NEG_METRIC (-8192)
#define CLAMP_CH 1023
#define CLAMP_EXT 2047

/* ------------------------------------------------------------------------------------------- CRC */
uint32_t orc_crc(const uint8_t* bits, uint32_t n, uint32_t poly, uint32_t order) {
    /* bits: one bit per byte; zero init; returns remainder of bits * x^order */
    uint32_t r = 0, top = 1u << order;
    for (uint32_t i = 0; i < n + order; ++i) {
        r = (r << 1) | (i < n ? bits[i] : 0);
        if (r & top) r ^= poly;
    }
    return r & (top - 1);
}

void orc_unpack(const uint8_t* bytes, uint8_t* bits, uint32_t nbits) {
    for (uint32_t i = 0; i < nbits; ++i) bits[i] = (bytes[i >> 3] >> (7 - (i & 7))) & 1;
}
void orc_pack(const uint8_t* bits, uint8_t* bytes, uint32_t nbits) {
    memset(bytes, 0, (nbits + 7) / 8);
    for (uint32_t i = 0; i < nbits; ++i) bytes[i >> 3] |= (uint8_t)(bits[i] << (7 - (i & 7)));
}

/* --------------------------------------------------------------------------------- turbo encoder */
/* TS 36.212 5.1.3.2: g0 = 1+D^2+D^3, g1 = 1+D+D^3; d: [3][K+4] one bit per byte */
static void rsc(const uint8_t* in, uint32_t K, uint8_t* par, uint8_t xt[3], uint8_t zt[3]) {
    uint8_t s0 = 0, s1 = 0, s2 = 0;
    for (uint32_t k = 0; k < K; ++k) {
        uint8_t a = in[k] ^ s1 ^ s2;
        par[k] = a ^ s0 ^ s2;
        s2 = s1; s1 = s0; s0 = a;
    }
    for (int t = 0; t < 3; ++t) {
        uint8_t u = s1 ^ s2;
        xt[t] = u;
        zt[t] = s0 ^ s2;
        s2 = s1; s1 = s0; s0 = 0;
    }
}

void orc_turbo_encode(const uint8_t* c, uint32_t K, uint32_t f1, uint32_t f2, uint8_t* d) {
    uint8_t* d0 = d; uint8_t* d1 = d + (K + 4); uint8_t* d2 = d + 2 * (K + 4);
    static __thread uint8_t ci[MAXK];
    for (uint32_t i = 0; i < K; ++i) ci[i] = c[((uint64_t)f1 * i + (uint64_t)f2 * i * i) % K];
    uint8_t x[3], z[3], xp[3], zp[3];
    memcpy(d0, c, K);
    rsc(c, K, d1, x, z);
    rsc(ci, K, d2, xp, zp);
    d0[K] = x[0];      d1[K] = z[0];      d2[K] = x[1];
    d0[K + 1] = z[1];  d1[K + 1] = x[2];  d2[K + 1] = z[2];
    d0[K + 2] = xp[0]; d1[K + 2] = zp[0]; d2[K + 2] = xp[1];
    d0[K + 3] = zp[1]; d1[K + 3] = xp[2]; d2[K + 3] = zp[2];
}

/* --------------------------------------------------------------------------------- rate matching */
/* TS 36.212 5.1.4.1, textbook form: builds w (NULL = -1 as index) for one K; wmap[k] = d address */
static const uint8_t PERM[32] = {0, 16, 8, 24, 4, 20, 12, 28, 2, 18, 10, 26, 6, 22, 14, 30,
                                 1, 17, 9, 25, 5, 21, 13, 29, 3, 19, 11, 27, 7, 23, 15, 31};

static uint32_t build_w(uint32_t K, int32_t* w, uint32_t* Rout) {
    const uint32_t D = K + 4, R = (D + 31) / 32, Kp = 32 * R, ND = Kp - D;
    int32_t* y = (int32_t*)malloc(sizeof(int32_t) * Kp);
    for (int s = 0; s < 3; ++s) {
        for (uint32_t i = 0; i < Kp; ++i) y[i] = i < ND ? -1 : (int32_t)(s * D + (i - ND));
        for (uint32_t k = 0; k < Kp; ++k) {
            int32_t v;
            if (s < 2) {
                v = y[(k % R) * 32 + PERM[k / R]];
            } else {
                v = y[(PERM[k / R] + 32 * (k % R) + 1) % Kp];
            }
            if (s == 0) w[k] = v;
            else if (s == 1) w[Kp + 2 * k] = v;
            else w[Kp + 2 * k + 1] = v;
        }
    }
    free(y);
    *Rout = R;
    return 3 * Kp;
}

/* Limited soft buffer (TS 36.212 5.1.4.1.2; ETSI TS 103 636-3 6.1.5.3; the reference: ToDo at pdc_enc.cpp:122-123, always
 * N_cb = K_w): orc_N_cb > 0 ends the circular buffer of the NEXT orc_rm_tx / orc_rm_rx calls at min(orc_N_cb, K_w).  The
 * PDC chain (fec_chain.c) sets it per transport block from orc_N_IR (= N_soft / M_DL_HARQ, 0 = no limitation) and clears
 * it again; the PCC never limits. */
uint32_t orc_N_IR = 0;
__thread uint32_t orc_N_cb = 0;

/* e[0..E) <- bits selected from d (one bit per byte) with redundancy version rv */
void orc_rm_tx(const uint8_t* d, uint32_t K, uint32_t E, uint32_t rv, uint8_t* e) {
    static __thread int32_t w[3 * (MAXK + 36)];
    uint32_t R, Kw = build_w(K, w, &R);
    if (orc_N_cb && orc_N_cb < Kw) Kw = orc_N_cb;
    uint32_t k0 = R * (2 * ((Kw + 8 * R - 1) / (8 * R)) * rv + 2);
    uint32_t k = 0, j = 0;
    while (k < E) {
        int32_t a = w[(k0 + j) % Kw];
        if (a >= 0) e[k++] = d[a];
        ++j;
    }
}

/* soft[a] += e[k] (plain int16 wrap-around add) */
void orc_rm_rx(const int16_t* e, uint32_t K, uint32_t E, uint32_t rv, int16_t* soft) {
    static __thread int32_t w[3 * (MAXK + 36)];
    uint32_t R, Kw = build_w(K, w, &R);
    if (orc_N_cb && orc_N_cb < Kw) Kw = orc_N_cb;
    uint32_t k0 = R * (2 * ((Kw + 8 * R - 1) / (8 * R)) * rv + 2);
    uint32_t k = 0, j = 0;
    while (k < E) {
        int32_t a = w[(k0 + j) % Kw];
        if (a >= 0) {
            soft[a] = (int16_t)(uint16_t)((uint16_t)soft[a] + (uint16_t)e[k]);
            ++k;
        }
        ++j;
    }
}

/* --------------------------------------------------------------------------------- turbo decoder */
/* Windowed int16 max-log-MAP with next-iteration initialisation; the CONTRACT shared with the CUDA
 * kernel.  state S = s0<<2 | s1<<1 | s2. */
typedef struct {
    uint32_t K, L, P, f1, f2;
    int16_t sys[MAXK], p1[MAXK], p2[MAXK];
    int16_t xt1[3], zt1[3], xt2[3], zt2[3];
    int16_t ext[MAXK];               /* extrinsic in natural order (a-priori of the next decoder) */
    int16_t a_end[2][MAXK / 4][8];   /* [decoder][window] alpha at window end (previous iteration) */
    int16_t b_start[2][MAXK / 4][8]; /* [decoder][window] beta at window start (previous iteration) */
    int32_t range_violations;
} tdec_t;

static inline int next_state(int S, int u, int* z) {
    int s0 = (S >> 2) & 1, s1 = (S >> 1) & 1, s2 = S & 1;
    int a = u ^ s1 ^ s2;
    *z = a ^ s0 ^ s2;
    return (a << 2) | (s0 << 1) | s1;
}

static inline int32_t chk(tdec_t* t, int32_t v) {
    if (v < -32768 || v > 32767) t->range_violations++;
    return v;
}

static inline int16_t clamp16(int32_t v, int32_t lim) { return (int16_t)(v > lim ? lim : (v < -lim ? -lim : v)); }

uint32_t orc_window_override = 0; /* test hook: 0 = contract rule, else forced window length */
uint32_t orc_tdec_window(uint32_t K) {
    if (orc_window_override && (K / 2) % orc_window_override == 0) return orc_window_override;
    if (orc_window_override >= K) return K;
    /* divisor of K/2 closest to 32 among divisors >= 16 (ties: larger); K/2 itself if K/2 < 16 */
    uint32_t h = K / 2, best = h;
    int32_t bestd = 1 << 30;
    if (h < 16) return h;
    for (uint32_t d = 16; d <= h; ++d) {
        if (h % d) continue;
        int32_t dd = (int32_t)d - 32;
        if (dd < 0) dd = -dd;
        if (dd < bestd || (dd == bestd && d > best)) { bestd = dd; best = d; }
    }
    return best;
}

void orc_tdec_new_cb(tdec_t* t, const int16_t* soft, uint32_t K, uint32_t f1, uint32_t f2) {
    const int16_t* d0 = soft; const int16_t* d1 = soft + K + 4; const int16_t* d2 = soft + 2 * (K + 4);
    t->K = K; t->f1 = f1; t->f2 = f2;
    t->L = orc_tdec_window(K);
    t->P = K / t->L;
    for (uint32_t i = 0; i < K; ++i) {
        t->sys[i] = clamp16(d0[i], CLAMP_CH);
        t->p1[i] = clamp16(d1[i], CLAMP_CH);
        t->p2[i] = clamp16(d2[i], CLAMP_CH);
        t->ext[i] = 0;
    }
    t->xt1[0] = clamp16(d0[K], CLAMP_CH);     t->zt1[0] = clamp16(d1[K], CLAMP_CH);
    t->xt1[1] = clamp16(d2[K], CLAMP_CH);     t->zt1[1] = clamp16(d0[K + 1], CLAMP_CH);
    t->xt1[2] = clamp16(d1[K + 1], CLAMP_CH); t->zt1[2] = clamp16(d2[K + 1], CLAMP_CH);
    t->xt2[0] = clamp16(d0[K + 2], CLAMP_CH); t->zt2[0] = clamp16(d1[K + 2], CLAMP_CH);
    t->xt2[1] = clamp16(d2[K + 2], CLAMP_CH); t->zt2[1] = clamp16(d0[K + 3], CLAMP_CH);
    t->xt2[2] = clamp16(d1[K + 3], CLAMP_CH); t->zt2[2] = clamp16(d2[K + 3], CLAMP_CH);
    memset(t->a_end, 0, sizeof(t->a_end));
    memset(t->b_start, 0, sizeof(t->b_start));
    t->range_violations = 0;
}

/* Normalisation schedule of the product kernel (csrc/fec_rx.cu k_turbo_decode): in the training passes and the
 * plain backward pass the metrics are normalised to state 0 only every fourth step (and on every step of the
 * L % 4 remainder).  Normalisation shifts all eight metrics alike, so the schedule changes no decision and no
 * stored value; it is restated here so that chk() proves the 16-bit range of the un-normalised values too. */
static inline int norm_train(uint32_t i, uint32_t L) { return (i & 3) == 3 || i >= (L & ~3u); }
static inline int norm_back(uint32_t k, uint32_t L) { return (k & 3) == 0 || k >= (L & ~3u); }

/* one constituent decoder over all windows; x = sys + a-priori (decoder order), y = parity;
 * writes llr (decoder order) */
static void map_half(tdec_t* t, int dec, const int16_t* x, const int16_t* y, const int16_t* xt,
                     const int16_t* zt, int16_t* llr) {
    const uint32_t K = t->K, L = t->L, P = t->P;
    static __thread int16_t beta[MAXK + 1][8];
    int16_t a_end_new[MAXK / 4][8], b_start_new[MAXK / 4][8];
    /* beta at the trellis end from the three termination steps */
    int32_t b_tail[8];
    {
        int32_t b[8], nb[8];
        for (int S = 0; S < 8; ++S) b[S] = S == 0 ? 0 : NEG_METRIC;
        for (int tt = 2; tt >= 0; --tt) {
            for (int S = 0; S < 8; ++S) {
                int s1 = (S >> 1) & 1, s2 = S & 1, u = s1 ^ s2, z;
                int Sn = next_state(S, u, &z);
                nb[S] = chk(t, b[Sn] + u * xt[tt] + z * zt[tt]);
            }
            for (int S = 0; S < 8; ++S) b[S] = chk(t, nb[S] - nb[0]);
        }
        for (int S = 0; S < 8; ++S) b_tail[S] = b[S];
    }
    for (uint32_t w = 0; w < P; ++w) {
        int32_t b[8], a[8], nb[8], na[8];
        /* ---- beta initialisation: exact at the trellis end, otherwise a training pass over window
         * w+1 that itself starts from the boundary stored by window w+2 in the previous iteration
         * (zeros in the first one) or from the exact tail metrics */
        if (w == P - 1) {
            for (int S = 0; S < 8; ++S) b[S] = b_tail[S];
        } else {
            for (int S = 0; S < 8; ++S) b[S] = (w + 2 == P) ? b_tail[S] : t->b_start[dec][w + 2][S];
            for (int32_t k = (int32_t)L - 1; k >= 0; --k) {
                int32_t xx = x[(w + 1) * L + k], yy = y[(w + 1) * L + k];
                for (int S = 0; S < 8; ++S) {
                    int z0, z1;
                    int S0 = next_state(S, 0, &z0), S1 = next_state(S, 1, &z1);
                    int32_t m0 = chk(t, b[S0] + z0 * yy);
                    int32_t m1 = chk(t, b[S1] + xx + z1 * yy);
                    nb[S] = m0 > m1 ? m0 : m1;
                }
                const uint32_t i = L - 1 - (uint32_t)k;
                const int32_t ref = norm_train(i, L) ? nb[0] : 0;
                for (int S = 0; S < 8; ++S) b[S] = chk(t, nb[S] - ref);
            }
        }
        for (int S = 0; S < 8; ++S) beta[L][S] = (int16_t)b[S];
        /* ---- backward recursion */
        for (int32_t k = (int32_t)L - 1; k >= 0; --k) {
            int32_t xx = x[w * L + k], yy = y[w * L + k];
            for (int S = 0; S < 8; ++S) {
                int z0, z1;
                int S0 = next_state(S, 0, &z0), S1 = next_state(S, 1, &z1);
                int32_t m0 = chk(t, b[S0] + z0 * yy);
                int32_t m1 = chk(t, b[S1] + xx + z1 * yy);
                nb[S] = m0 > m1 ? m0 : m1;
            }
            /* beta[k] holds the normalised metrics (the product recomputes them, normalised, from normalised
             * checkpoints); the running b[] follows the product's schedule for the range proof */
            const int32_t ref = norm_back((uint32_t)k, L) ? nb[0] : 0;
            for (int S = 0; S < 8; ++S) { b[S] = chk(t, nb[S] - ref); beta[k][S] = (int16_t)chk(t, nb[S] - nb[0]); }
        }
        for (int S = 0; S < 8; ++S) b_start_new[w][S] = (int16_t)b[S];
        /* ---- alpha initialisation: exact at the trellis start, otherwise a training pass over
         * window w-1 starting from the boundary stored by window w-2 in the previous iteration */
        if (w == 0) {
            for (int S = 0; S < 8; ++S) a[S] = S == 0 ? 0 : NEG_METRIC;
        } else {
            for (int S = 0; S < 8; ++S)
                a[S] = (w == 1) ? (S == 0 ? 0 : NEG_METRIC) : t->a_end[dec][w - 2][S];
            for (uint32_t k = 0; k < L; ++k) {
                int32_t xx = x[(w - 1) * L + k], yy = y[(w - 1) * L + k];
                for (int S = 0; S < 8; ++S) na[S] = -(1 << 30);
                for (int S = 0; S < 8; ++S)
                    for (int u = 0; u < 2; ++u) {
                        int z;
                        int Sn = next_state(S, u, &z);
                        int32_t g = chk(t, a[S] + u * xx + z * yy);
                        if (g > na[Sn]) na[Sn] = g;
                    }
                const int32_t ref = norm_train(k, L) ? na[0] : 0;
                for (int S = 0; S < 8; ++S) a[S] = chk(t, na[S] - ref);
            }
        }
        /* ---- forward recursion + LLR */
        for (uint32_t k = 0; k < L; ++k) {
            int32_t xx = x[w * L + k], yy = y[w * L + k];
            int32_t m1max = -(1 << 30), m0max = -(1 << 30);
            for (int S = 0; S < 8; ++S) na[S] = -(1 << 30);
            for (int S = 0; S < 8; ++S) {
                for (int u = 0; u < 2; ++u) {
                    int z;
                    int Sn = next_state(S, u, &z);
                    int32_t g = chk(t, a[S] + u * xx + z * yy);
                    if (g > na[Sn]) na[Sn] = g;
                    int32_t tot = chk(t, g + beta[k + 1][Sn]);
                    if (u) { if (tot > m1max) m1max = tot; }
                    else   { if (tot > m0max) m0max = tot; }
                }
            }
            llr[w * L + k] = (int16_t)chk(t, m1max - m0max);
            for (int S = 0; S < 8; ++S) a[S] = chk(t, na[S] - na[0]);
        }
        for (int S = 0; S < 8; ++S) a_end_new[w][S] = (int16_t)a[S];
    }
    (void)K;
    memcpy(t->a_end[dec], a_end_new, sizeof(int16_t) * 8 * P);
    memcpy(t->b_start[dec], b_start_new, sizeof(int16_t) * 8 * P);
}

/* one full iteration; hard decisions (one bit per byte, natural order) into bits[K] */
void orc_tdec_iteration(tdec_t* t, uint8_t* bits) {
    const uint32_t K = t->K;
    static __thread int16_t x[MAXK], llr[MAXK];
    static __thread uint32_t pi[MAXK];
    for (uint32_t i = 0; i < K; ++i) pi[i] = (uint32_t)(((uint64_t)t->f1 * i + (uint64_t)t->f2 * i * i) % K);
    /* decoder 1, natural order */
    for (uint32_t i = 0; i < K; ++i) x[i] = (int16_t)(t->sys[i] + t->ext[i]);
    map_half(t, 0, x, t->p1, t->xt1, t->zt1, llr);
    for (uint32_t i = 0; i < K; ++i) t->ext[i] = clamp16((int32_t)llr[i] - x[i], CLAMP_EXT);
    /* decoder 2, interleaved order */
    for (uint32_t j = 0; j < K; ++j) x[j] = (int16_t)(t->sys[pi[j]] + t->ext[pi[j]]);
    map_half(t, 1, x, t->p2, t->xt2, t->zt2, llr);
    for (uint32_t j = 0; j < K; ++j) {
        t->ext[pi[j]] = clamp16((int32_t)llr[j] - x[j], CLAMP_EXT);
        bits[pi[j]] = llr[j] > 0;
    }
}

tdec_t* orc_tdec_alloc(void) { return (tdec_t*)calloc(1, sizeof(tdec_t)); }
void orc_tdec_free(tdec_t* t) { free(t); }
int32_t orc_tdec_range_violations(const tdec_t* t) { return t->range_violations; }
const int16_t* orc_tdec_ext(const tdec_t* t) { return t->ext; }

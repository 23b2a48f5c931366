/* oracle/channel.c -- TEST INFRASTRUCTURE ONLY.  CPU restatement of the simulation channel for one packet:
 * superposition of all TX antennas on every RX antenna (reference
 * lib/src/simulation/wireless/channel_awgn.cpp:39-64, channel_flat.cpp:47-88), integer delay into the RX
 * window, white noise at n0_dBFS = -10 log10(net_bandwidth_norm) - snr_dB over the whole window
 * (lib/src/simulation/vspace.cpp:487-497, lib/src/simulation/hardware/noise.cpp:31-41; srsRAN
 * ch_awgn.c: std_dev = 10^(n0/20), complex sample = std_dev/sqrt(2) * (N(0,1) + j N(0,1))).
 * The random numbers are NOT srsRAN's: both this file and the CUDA kernel draw from Philox4x32-10 with
 * counter = (sample pair | tx antenna, rx antenna, packet, stream) so that realisations can be compared.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

typedef struct { float re, im; } cf_t;

static void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
    for (int r = 0; r < 10; ++r) {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c[0], p1 = (uint64_t)0xCD9E8D57u * c[2];
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0, n1 = (uint32_t)p1;
        const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1, n3 = (uint32_t)p0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
}

static cf_t cn01(uint32_t a, uint32_t b) {  /* Box-Muller, E|z|^2 = 1 */
    const float u1 = ((float)(a >> 8) + 0.5f) * (1.0f / 16777216.0f);
    const float u2 = ((float)(b >> 8) + 0.5f) * (1.0f / 16777216.0f);
    const float r = sqrtf(-logf(u1));
    cf_t z = {r * (float)cos(2.0 * M_PI * (double)u2), r * (float)sin(2.0 * M_PI * (double)u2)};
    return z;
}

/* tx [N_TX][tx_stride], rx [N_RX][rx_stride] */
void orc_channel(const cf_t* tx, long tx_stride, int N_TX, int n_samples, cf_t* rx, long rx_stride, int N_RX,
                 int t0, int mode, float snr_dB, float net_bandwidth_norm, uint64_t seed, uint32_t pkt) {
    const float n0_dBFS = -10.0f * log10f(net_bandwidth_norm) - snr_dB;
    const float sigma = powf(10.0f, n0_dBFS / 20.0f);
    const uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
    for (int a = 0; a < N_RX; ++a) {
        cf_t g[8];
        for (int t = 0; t < N_TX; ++t) {
            g[t].re = 1.f; g[t].im = 0.f;
            if (mode == 1) {
                uint32_t c[4] = {(uint32_t)t, (uint32_t)a, pkt, 1u};
                philox4x32_10(c, k0, k1);
                g[t] = cn01(c[0], c[1]);
            }
        }
        for (long pair = 0; 2 * pair < rx_stride; ++pair) {
            uint32_t c[4] = {(uint32_t)pair, (uint32_t)a, pkt, 0u};
            philox4x32_10(c, k0, k1);
            cf_t y[2] = {cn01(c[0], c[1]), cn01(c[2], c[3])};
            for (int i = 0; i < 2; ++i) {
                const long n = 2 * pair + i;
                if (n >= rx_stride) break;
                y[i].re *= sigma; y[i].im *= sigma;
                if (n >= t0 && n - t0 < n_samples)
                    for (int t = 0; t < N_TX; ++t) {
                        const cf_t x = tx[(long)t * tx_stride + (n - t0)];
                        y[i].re += g[t].re * x.re - g[t].im * x.im;
                        y[i].im += g[t].re * x.im + g[t].im * x.re;
                    }
                rx[(long)a * rx_stride + n] = y[i];
            }
        }
    }
}

/* ---- doubly selective channel: tapped delay line with a sum of 40 sinusoids per tap
 * (lib/src/simulation/wireless/channel_doubly.cpp:59-87, link.cpp:72-146 set_pdp, :165-215 randomize, :217-282
 * pass_through_link; profiles link.hpp:96-109).  The reference draws angles / phases from a time seeded generator
 * and rotates chunk by chunk (spp_size) with a recursive VOLK rotator restarted from time_64 % |period| at every
 * chunk; here every sample's phase is that closed form, phase = 2 pi / P * ((time_64 + n) % |P|) + phase0, and the
 * draws come from Philox (counter = (sinusoid | tap << 8 | tx << 16, rx antenna, packet, 2)).  Pathloss = 1. */
static const double TDL_DELAY_NS[3][9] = {{0, 30, 70, 90, 110, 190, 410, NAN, NAN},
                                          {0, 30, 150, 310, 370, 710, 1090, 1730, 2510},
                                          {0, 50, 120, 200, 230, 500, 1600, 2300, 5000}};
static const double TDL_POWER_DB[3][9] = {{+0.0f, -1.0f, -2.0f, -3.0f, -8.0f, -17.2f, -20.8f, NAN, NAN},
                                          {+0.0f, -1.5f, -1.4f, -3.6f, -0.6f, -9.1f, -7.0f, -12.0f, -16.9f},
                                          {-1.0f, -1.0f, -1.0f, +0.0f, +0.0f, +0.0f, -3.0f, -5.0f, -7.0f}};

/* link_t::get_tau_rms_ns (link.cpp:290-322) of a generic profile; the reference quotes 43.13 / 356.65 / 990.93 ns
 * (link.hpp:84-88) */
double orc_tdl_tau_rms_ns(int pdp_idx) {
    int nfin = 0;
    for (int i = 0; i < 9; ++i) { if (isfinite(TDL_DELAY_NS[pdp_idx][i])) ++nfin; else break; }
    double pl[9], sum = 0, mean = 0, var = 0;
    for (int i = 0; i < nfin; ++i) { pl[i] = pow(10.0, TDL_POWER_DB[pdp_idx][i] / 10.0); sum += pl[i]; }
    for (int i = 0; i < nfin; ++i) mean += TDL_DELAY_NS[pdp_idx][i] * (pl[i] / sum);
    for (int i = 0; i < nfin; ++i) { const double A = TDL_DELAY_NS[pdp_idx][i] - mean; var += A * A * (pl[i] / sum); }
    return sqrt(var);
}

/* link_t::set_pdp: returns the number of merged taps; delays in samples, powers normalised to sum 1 */
int orc_tdl_pdp(double samp_rate, int pdp_idx, float tau_rms_ns, uint32_t* delay_smpl, double* power_lin) {
    int nfin = 0;
    for (int i = 0; i < 9; ++i) { if (isfinite(TDL_DELAY_NS[pdp_idx][i])) ++nfin; else break; }
    double pl[9], sum = 0, mean = 0, var = 0;
    for (int i = 0; i < nfin; ++i) { pl[i] = pow(10.0, TDL_POWER_DB[pdp_idx][i] / 10.0); sum += pl[i]; }
    for (int i = 0; i < nfin; ++i) mean += TDL_DELAY_NS[pdp_idx][i] * (pl[i] / sum);
    for (int i = 0; i < nfin; ++i) { const double A = TDL_DELAY_NS[pdp_idx][i] - mean; var += A * A * (pl[i] / sum); }
    const double scale = (double)tau_rms_ns / sqrt(var), Ts = 1.0 / samp_rate;
    int n = 0;
    for (int i = 0; i < nfin; ++i) {
        const uint32_t A = (uint32_t)floor(TDL_DELAY_NS[pdp_idx][i] * 1.0e-9 * scale / Ts);
        int idx = -1;
        for (int j = 0; j < n; ++j) if (delay_smpl[j] == A) { idx = j; break; }
        if (idx < 0) { delay_smpl[n] = A; power_lin[n] = pl[i]; ++n; } else power_lin[idx] += pl[i];
    }
    double s2 = 0;
    for (int i = 0; i < n; ++i) s2 += power_lin[i];
    for (int i = 0; i < n; ++i) power_lin[i] /= s2;
    return n;
}

static float m1p1(uint32_t r) { return ((float)(r >> 8) + 0.5f) * (1.0f / 8388608.0f) - 1.0f; }

void orc_channel_doubly(const cf_t* tx, long tx_stride, int N_TX, int n_samples, cf_t* rx, long rx_stride, int N_RX,
                        int t0, float snr_dB, float net_bandwidth_norm, uint64_t seed, uint32_t pkt,
                        double samp_rate, int pdp_idx, float tau_rms_ns, float fD_Hz, int64_t time_64) {
    const float n0_dBFS = -10.0f * log10f(net_bandwidth_norm) - snr_dB;
    const float sigma = powf(10.0f, n0_dBFS / 20.0f);
    const uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
    uint32_t delay[9]; double power[9];
    const int n_taps = orc_tdl_pdp(samp_rate, pdp_idx, tau_rms_ns, delay, power);
    for (int a = 0; a < N_RX; ++a) {
        /* noise, same stream as orc_channel */
        for (long pair = 0; 2 * pair < rx_stride; ++pair) {
            uint32_t c[4] = {(uint32_t)pair, (uint32_t)a, pkt, 0u};
            philox4x32_10(c, k0, k1);
            const cf_t y[2] = {cn01(c[0], c[1]), cn01(c[2], c[3])};
            for (int i = 0; i < 2 && 2 * pair + i < rx_stride; ++i) {
                rx[(long)a * rx_stride + 2 * pair + i].re = y[i].re * sigma;
                rx[(long)a * rx_stride + 2 * pair + i].im = y[i].im * sigma;
            }
        }
        for (int t = 0; t < N_TX; ++t)
            for (int i = 0; i < n_taps; ++i) {
                int64_t P[40]; float ph0[40];
                for (int j = 0; j < 40; ++j) {
                    uint32_t c[4] = {(uint32_t)j | ((uint32_t)i << 8) | ((uint32_t)t << 16), (uint32_t)a, pkt, 2u};
                    philox4x32_10(c, k0, k1);
                    const double ang = (double)m1p1(c[0]) * 2.0 * M_PI;
                    const double fd = (double)fD_Hz * cos(ang);
                    P[j] = (-0.01f < fd && fd < 0.01f) ? INT64_MAX : (int64_t)(samp_rate / fd);
                    ph0[j] = m1p1(c[1]) * 2.0f * (float)M_PI;
                }
                float scale = 1.0f / sqrtf(40.0f);
                scale *= sqrt(power[i]);
                for (long n = t0 + delay[i]; n < rx_stride && n - t0 - (long)delay[i] < n_samples; ++n) {
                    const cf_t x = tx[(long)t * tx_stride + (n - t0 - delay[i])];
                    float sr = 0.f, si = 0.f;
                    for (int j = 0; j < 40; ++j) {
                        const double s2s = 2.0f * (float)M_PI / (double)P[j];
                        const double off = (double)((time_64 + n) % llabs(P[j]));
                        const float ph = (float)(s2s * off) + ph0[j];
                        sr += cosf(ph); si += sinf(ph);
                    }
                    sr *= scale; si *= scale;
                    rx[(long)a * rx_stride + n].re += sr * x.re - si * x.im;
                    rx[(long)a * rx_stride + n].im += sr * x.im + si * x.re;
                }
            }
    }
}

/* ---- hardware effects of the simulated radio and large-scale fading (row a38 of SURVEY.md section 8) ------------------------
 * hw_simulator_t::clip_and_quantize (hw_simulator.cpp:710-731): symmetric clipping of I and Q at clip_limit
 * (simulation/hardware/clip.cpp:27-37), then rounding to multiples of 2 clip_limit / 2^n_bits
 * (simulation/hardware/quantize.cpp:27-45, roundf = half away from zero); n_bits = 0: clip only; clip_limit <= 0: nothing. */
void orc_hw_clip_quantize(cf_t* x, long n, float clip_limit, uint32_t n_bits) {
    if (!(clip_limit > 0.0f)) return;
    const float w = 2.0f * clip_limit / powf(2.0f, (float)n_bits);
    for (long i = 0; i < n; ++i) {
        float re = x[i].re < -clip_limit ? -clip_limit : (x[i].re > clip_limit ? clip_limit : x[i].re);
        float im = x[i].im < -clip_limit ? -clip_limit : (x[i].im > clip_limit ? clip_limit : x[i].im);
        if (n_bits > 0) { re = roundf(re / w) * w; im = roundf(im / w) * w; }
        x[i].re = re; x[i].im = im;
    }
}
/* channel_t::get_large_scale_via_pathloss (channel.cpp:53-73) with pathloss::fspl (pathloss.cpp:28-38): magnitude gain of
 * one TX antenna into one RX antenna; same_node: the node's own TX leaking into its RX */
float orc_large_scale_via_pathloss(float tx_power_ant_0dBFS, float rx_power_ant_0dBFS, float distance_m, float freq_Hz,
                                   int same_node, float tx_into_rx_leakage_dB) {
    float fading_dB;
    if (!same_node) {
        if (distance_m <= 0.0f || freq_Hz <= 0.0f) fading_dB = 0.0f;
        else {
            const double v = 20.0 * log10((double)distance_m) + 20.0f * log10((double)freq_Hz) - 147.55;
            fading_dB = v < 0.0 ? 0.0f : (float)v;
        }
    } else fading_dB = tx_into_rx_leakage_dB;
    return powf(10.0f, (tx_power_ant_0dBFS - fading_dB - rx_power_ant_0dBFS) / 20.0f);
}

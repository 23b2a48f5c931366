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

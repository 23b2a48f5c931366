// phy.cuh -- device-visible description of one packet geometry (flat tables built by plan_phy.cu) and
// the launch interface of the OFDM / channel / receiver kernels.
#pragma once
#include <cuda_runtime.h>
#include <cstddef>
#include <cstdint>
#include "../../include/dnrp_c_api.h"

namespace dnrp {

struct phy_dims_t {
    uint32_t u, b, N_eff_TX, N_TS, N_TX, N_SS, N_RX, N_bps, txdiv;
    uint32_t N_OCC, Nk, NP;                 // occupied subcarriers, N_OCC + 1, pilots per DRS symbol and TS
    uint32_t N_os, cp_os, stf_cp_os;        // FFT size and CP lengths at the oversampled rate
    uint32_t stf_len_os, sym_len_os;        // samples of the STF symbol / of a data field symbol
    uint32_t n_pat, pat_len_os;             // STF patterns and their length
    uint32_t N_DF, N_step, n_drs;           // data field symbols, DRS distance, DRS occurrences per TS
    uint32_t N_PDC_subc, G;
    uint32_t n_no_gi_os, n_packet_os;       // samples without / with full guard interval
    uint32_t n_cb_idx;                      // codebook entries for (N_TS, N_TX)
    uint32_t txdiv_modulo;
    // device tables
    const uint32_t* cell;      // [(N_DF+1)][Nk] cell words (spec_re.hpp)
    const uint8_t* cell_ab;    // [(N_DF+1)][Nk] transmit diversity: streams A | B << 4 of the cell's pair (Table 6.3.3.2), else 0
    const float2* qam_lut;     // constellation points, LTE Gray mapping: entry qam_lut_base(N_bps) + v (v = the N_bps bits, MSB first)
    const uint32_t* pcc_lk;    // [98]
    const uint32_t* pdc_lk;    // [N_PDC_subc]
    // channel-estimation plan per demapper unit (cell, or cell pair with transmit diversity) and stream slot (A, B):
    // everything rx_synced's processing-stage bookkeeping derives from (symbol, stream), packed by ce_plan_word()
    const uint32_t* pcc_plan;  // [units][2]
    const uint32_t* pdc_plan;  // [units][2]; spatial multiplexing (N_SS > 1): [cells][N_SS], one word per stream
    const int8_t* stf_sign;    // [Nk]
    const int8_t* W;           // [n_cb_idx][N_TX][N_TS] codes 0, +-1, +-2 (= +-j)
    const float* W_scale;      // [n_cb_idx]
    const float* W_scale_opt;  // [n_cb_idx] "optimal DAC" variant
    const float2* twiddle;     // [N_os]
    // the same values per pass in the order the lanes read them (fft.cuh fft_pass_tables): radix-4 plan / radix-16 plan;
    // N_os = 1024, radix 16: [15][16] W_256^(k r), then [3][256] W_1024^(j r)
    const float2* twiddle_p4;
    const float2* twiddle_p16;
    // receiver
    uint32_t chestim_lr, chestim_stride;    // phy.json chestim_mode_lr_default / _t_stride_default
    uint32_t n_full_ps;                     // full processing stages in lr mode (rx_synced.cpp:1031-1032)
    uint32_t N_step_inv, chestim_stride_inv;  // ceil(2^32 / x): i / x == __umulhi(i, inv) for the symbol indices used
    uint32_t pcc_l_max;                     // last OFDM symbol carrying PCC cells
    uint32_t last_plan[8];                  // plan word of (stream ts, last data field symbol): the estimate left at packet end
    struct wiener_dev_t {
        const uint16_t* idx_pilot;          // [n_t][4][Nk]
        const uint16_t* idx_weight;         // [n_t][4][Nk]
        const float* weights;               // [n_vec][n_interp]
        uint32_t n_interp, n_t;
    } wiener_l[3], wiener_lr[3];            // per SNR class {-5, 15, 35} dB (rx_synced_param.hpp:216-232)
    // transmit diversity: the two Wiener windows of a cell pair merged into ONE window per (unit, stream slot) on the host
    // (chan_est_rx_pair): first pilot and tap count, and the weights 0.5 (w_k0 + w_k1) stored tap-major so that the
    // threads of a warp (adjacent units) read one contiguous line per tap
    struct pair_lut_t {
        const uint16_t* info;               // [2][units]: first pilot m | taps << 10; 0xffff: windows too far apart
        const float* w;                     // [n_interp + 1][2][units]
        uint32_t units;
    } pcc_pair[3], pdc_pair[3];             // per SNR class; info == nullptr: not built (single stream / spatial multiplexing PDC)
};

// first entry of the N_bps-bit constellation in phy_dims_t::qam_lut (1, 2, 4, 6, 8, 10 bits: 2 + 4 + 16 + 64 + 256 + 1024 entries)
__host__ __device__ __forceinline__ uint32_t qam_lut_base(uint32_t N_bps) {
    return N_bps == 1 ? 0u : (N_bps == 2 ? 2u : (N_bps == 4 ? 6u : (N_bps == 6 ? 22u : (N_bps == 8 ? 86u : 342u))));
}
// plan word: LUT slice base (o = base * Nk + k) | class slot j | pilot row | interlace flip | stream | mode lr
__host__ __device__ __forceinline__ uint32_t ce_plan_word(uint32_t base, uint32_t j, uint32_t row, uint32_t flip, uint32_t ts,
                                                          uint32_t lr) {
    return base | j << 6 | row << 12 | flip << 17 | ts << 18 | lr << 21;
}

// per-packet receiver state, produced by the STF stage and consumed by every later stage
// (the mutable part of sync_report_t + estimator_sto/snr state, reference rx_synced.cpp:503-619)
struct rx_state_t {
    int32_t t0;              // first STF sample inside the packet's RX window (fine peak)
    float inc0;              // mixer phase increment for the STF (cfo_fractional + cfo_integer), rad/sample
    float inc1;              // increment after the STF based residual correction
    float cfo_left;          // residual estimated on the STF (added to sync_report.cfo_fractional_rad)
    float sto_inc;           // estimator_sto phase_increment_rad (per subcarrier)
    float sto_fractional;    // in samples
    float rms[8];
    float SpN_sum, N_sum;    // estimator_snr accumulators after the STF (boost removed)
    uint32_t SpN_cnt, N_cnt;
    float snr_db_pcc, snr_db_pdc;
    float sm_rho_ref;        // spatial multiplexing: the packet's reference loss (phy_rx.cu k_rx_sm_ref)
};

struct rx_params_t {
    phy_dims_t d;
    const float2* iq;        // [packet][N_RX][iq_stride]
    size_t iq_stride;
    uint32_t t0_zero;        // 1: iq holds the packet resampled from its fine peak (rx_state_t::t0 is a hardware-rate index)
    rx_state_t* state;       // [packet]
    float2* Y;               // [packet][N_RX][N_DF][Nk] equalised-domain OFDM symbols 1..N_DF
    float2* hzf;             // [packet][N_RX][N_eff_TX][n_drs][NP] zero-forced DRS pilots
    float2* hstf;            // [packet][N_RX][NP] zero-forced STF cells (after STO derotation)
    float2* snr_part;        // [packet][N_DF][N_RX] (S+N sum, N sum) of the DRS cells in that symbol
    uint8_t* cls;            // [packet][cls_stride] Wiener class valid after the j-th OFDM symbol with DRS cells
    uint32_t cls_stride;
    int16_t* llr_pcc;        // [packet][196]
    int16_t* llr;            // [packet][llr_stride]
    size_t llr_stride;
    // residual CFO / STO tracking on the DRS (k_rx_track; nullptr = off): per packet the phase advance per N_step symbols,
    // .x common to all subcarriers (residual CFO), .y per subcarrier (drift of the symbol timing)
    float2* track;           // [packet]
    uint32_t track_cfo, track_sto;
};

// mimo_report_t (lib/include/dectnrp/phy/rx/rx_synced/mimo/mimo_report.hpp:27-59)
struct mimo_report_dev_t {
    uint32_t N_RX, N_TS_other, tm_3_7_beamforming_idx, tm_3_7_beamforming_reciprocal_idx;
};
// single-stream codebooks W(1, N) for the codebook search of estimator_mimo_t: side 0 = N = N_eff_TX (matrix for the
// other side), side 1 = N = N_RX (reciprocal); n = 0 where no beamforming exists (N = 1 or 8)
struct mimo_params_t {
    phy_dims_t d;
    const float2* hzf;               // [packet][N_RX][N_eff_TX][n_drs][NP]
    const int8_t* W[2];              // [n][N] codes 0, +-1, +-2
    const float* W_scale[2];         // [n]
    uint32_t n[2], first[2];         // number of matrices, first matrix without zero entries
    mimo_report_dev_t* out;          // [packet]
};
cudaError_t launch_rx_mimo(const mimo_params_t& p, uint32_t n_packets, cudaStream_t st);

cudaError_t launch_rx_stf(const rx_params_t& p, uint32_t n_packets, cudaStream_t st);
cudaError_t launch_rx_fft(const rx_params_t& p, uint32_t n_packets, cudaStream_t st);
cudaError_t launch_rx_snr_class(const rx_params_t& p, uint32_t n_packets, cudaStream_t st);
cudaError_t launch_rx_track(const rx_params_t& p, uint32_t n_packets, cudaStream_t st);
cudaError_t launch_rx_demap(const rx_params_t& p, uint32_t n_packets, bool pcc, cudaStream_t st);
// out [packet][N_RX][N_eff_TX][Nk]: the channel estimate the reference holds in channel_antennas[a]->chestim[ts] when the
// packet is done (rx_synced.cpp:444-480 exports exactly that)
cudaError_t launch_rx_chestim_tap(const rx_params_t& p, float2* out, uint32_t n_packets, cudaStream_t st);

// per-packet transmit parameters: tx_descriptor_t::codebook_index + tx_meta_t
// (reference lib/include/dectnrp/phy/tx/tx_descriptor.hpp:43-78, tx_meta.hpp:32-58)
struct tx_meta_dev_t {
    uint32_t codebook_index;
    uint32_t optimal_scaling_DAC;
    float DAC_scale;
    float iq_phase_rad;
    float iq_phase_increment;  // rad per (oversampled, post-resampling) sample
    uint32_t GI_percentage;
};

struct txofdm_params_t {
    phy_dims_t d;
    const uint8_t* pcc_bits;  size_t pcc_stride;   // [packet] packed 196 bits
    const uint8_t* pdc_bits;  size_t pdc_stride;   // [packet] packed G bits
    const tx_meta_dev_t* meta;                     // [packet]
    float2* iq;                                    // [packet][N_TX][iq_stride]
    size_t iq_stride;
    uint32_t no_mix;                               // 1: the mixer runs after the resampler instead
};

cudaError_t launch_tx_ofdm(const txofdm_params_t& p, uint32_t n_packets, cudaStream_t st);

// rational resampler (phy_resample.cu): y[m] = sum_j taps[j*L + l] x[i0 - j], p = delay + m*M, l = p % L, i0 = p / L,
// x = 0 outside [0, n_in) -- the closed form of resampler_t's streaming state machine incl. its final flush
struct resample_params_t {
    uint32_t L, M, SL, delay;
    const float* taps;                       // [SL][L]
    const float2* in;  size_t in_stride;     // [packet][n_ant][in_stride]
    const int32_t* start; uint32_t start_stride_words;  // optional per-packet first input sample (else 0)
    uint32_t n_in;                           // input samples per stream (clipped to in_stride - start)
    float2* out;  size_t out_stride;         // [packet][n_ant][out_stride]
    uint32_t n_out;                          // filtered output samples
    uint32_t n_fill;                         // outputs [n_out, n_fill) are zeroed (guard interval / padding)
    uint32_t n_ant;
    const tx_meta_dev_t* meta;               // optional: mixer after resampling (tx.cpp:913-934)
};
cudaError_t launch_resample(const resample_params_t& p, uint32_t n_packets, cudaStream_t st);
// int16 I/Q pairs -> float2 * scale; n_samples (complex) must be a multiple of 4
cudaError_t launch_sc16_to_cf32(const int16_t* in, float2* out, size_t n_samples, float scale, cudaStream_t st);
cudaError_t launch_sc16_rows_to_cf32(const int16_t* in, size_t in_stride_samples, float2* out, size_t out_stride_samples,
                                     uint32_t n_samples, uint32_t n_rows, float scale, cudaStream_t st);
cudaError_t launch_cf32_to_sc16(const float2* in, int16_t* out, size_t n_samples, float scale, cudaStream_t st);
cudaError_t launch_sc8_to_cf32(const int8_t* in, float2* out, size_t n_samples, float scale, cudaStream_t st);

// packet synchronisation (phy_sync.cu): detection, coarse peak, fine peak of the first packet in each window
struct sync_result_dev_t {
    uint32_t detected;
    uint32_t detection_time_local, detection_ant_idx;
    float detection_metric, detection_rms;
    uint32_t coarse_peak_time_local;
    float coarse_peak_array[8];
    float rms_array[8];
    float cfo_fractional_rad, cfo_integer_rad;
    uint32_t N_eff_TX, fine_peak_time_local;
    uint32_t n_detections, strongest_ant;
    uint32_t coarse_peak_time_hw;
};
struct sync_params_t {
    phy_dims_t d;
    const float2* iq;  size_t iq_stride;  // [packet][N_RX][iq_stride] windows at the DECT rate (x oversampling)
    uint32_t len;                          // usable samples per window in iq
    const float2* iq_x;  size_t x_stride;  // [packet][N_RX][x_stride] the same windows at the hardware rate (= iq if L = M)
    uint32_t x_len;
    uint32_t rs_L, rs_M;                   // hardware rate = DECT rate * L / M
    uint32_t xc_S, xc_sl;                  // template length and one-sided search range at the hardware rate
    uint32_t n_no_gi_hw;                   // packet length the receiver needs behind the fine peak (hardware rate)
    uint32_t n_ant;                        // antennas used (RX_SYNC_PARAM_AUTOCORRELATOR_ANTENNA_LIMIT = 8)
    uint32_t step, n_steps;                // detection step (pattern / 4) and number of steps in the window
    uint32_t n_steps_use;                  // steps the detection can reach (<= n_steps: those that start inside search_length);
                                           // only these are summed and scanned, 0 = all
    uint32_t search_length;                // samples the detection may consume
    float rms_min;
    float2* c_steps;                       // [packet][n_ant][n_steps]
    float* p_steps;                        // [packet][n_ant][n_steps]
    const float2* templates;               // [n_tpl][xc_S]
    uint32_t n_tpl;
    // fast correlation (optional): FFT size >= lags + template length - 1, conj(FFT(template)) / N_fft, twiddles
    uint32_t xc_fft_n;                     // 0: direct correlation
    const float2* tpl_spec;                // [n_tpl][xc_fft_n]
    const float2* xc_twiddle;              // [xc_fft_n]
    const float2* xc_twiddle_p16;          // per-pass tables of that size, radix-16 plan (fft.cuh fft_pass_tables)
    float2* xc;                            // [packet][n_tpl] (max abs(xcorr), lag)
    sync_result_dev_t* result;             // [packet]
    rx_state_t* rx_state;                  // [packet] optional: receives t0 / CFO / rms for the synced receiver
    // streaming front door (dnrp_stream_sync): "packets" are overlapping chunks of one continuous stream
    // [N_RX][iq_stride]; chunk c starts iq_pkt_stride samples after chunk c - 1 (0: dense windows as above)
    size_t iq_pkt_stride, x_pkt_stride;
    const uint32_t* resume;                // [packet] optional: first sample the detection may end on (skip_after_peak);
                                           // 0xffffffff = chunk finished
    const uint32_t* limit;                 // [packet] optional: end of the chunk's detection range (else search_length)
    const uint32_t* chunk_map;             // optional: block i of the search kernels works on chunk chunk_map[i] (the
                                           // chunks still active in a continuation round)
    // frequency-domain stage between coarse and fine peak (coarse_peak_f_domain_t, k_sync_fdomain), optional:
    float fd_beta_thr;                     // linear threshold of the beta search (db2mag of the dB value), 0 = off: b = context beta
    uint32_t fd_cfo_range;                 // integer CFO search range in subcarriers (multiples of 4 are tested), 0 = off
    const int8_t* fd_seq;                  // [6][224] STF base sequences y_b for b = 1, 2, 4, 8, 12, 16 (no cyclic shift)
    uint32_t* fd_b;                        // [packet] estimated beta
    int32_t* fd_cfo;                       // [packet] integer CFO in subcarriers
};
__device__ __forceinline__ uint32_t sync_chunk(const sync_params_t& p, uint32_t i) { return p.chunk_map ? p.chunk_map[i] : i; }
// first sample of antenna a of window / chunk pkt
__device__ __forceinline__ const float2* sync_iq(const sync_params_t& p, uint32_t pkt, uint32_t a) {
    return p.iq + (p.iq_pkt_stride ? (size_t)pkt * p.iq_pkt_stride + (size_t)a * p.iq_stride
                                   : ((size_t)pkt * p.d.N_RX + a) * p.iq_stride);
}
__device__ __forceinline__ const float2* sync_iq_x(const sync_params_t& p, uint32_t pkt, uint32_t a) {
    return p.iq_x + (p.x_pkt_stride ? (size_t)pkt * p.x_pkt_stride + (size_t)a * p.x_stride
                                    : ((size_t)pkt * p.d.N_RX + a) * p.x_stride);
}
cudaError_t launch_sync(const sync_params_t& p, uint32_t n_packets, cudaStream_t st);
// the two halves of launch_sync: per-step sums of every window (independent of resume / limit), and detection -> coarse
// peak -> fine peak over n_active windows (all of them, or the ones listed in chunk_map)
cudaError_t launch_sync_steps(const sync_params_t& p, uint32_t n_packets, cudaStream_t st);
cudaError_t launch_sync_search(const sync_params_t& p, uint32_t n_active, cudaStream_t st);
// dynamic shared memory k_sync_peak needs for n_steps detection steps (bounded by the device's opt-in limit)
size_t sync_peak_smem(const phy_dims_t& d, uint32_t n_steps);

// simulation channel (phy_channel.cu)
struct chan_meta_dev_t {
    uint32_t t0;        // first sample of the packet inside the RX window
    uint32_t mode;      // 0 = awgn (unit links), 1 = flat (one CN(0,1) coefficient per link and packet),
                        // 2 = doubly selective (tapped delay line, 40 sinusoids per tap and link)
    float sigma;        // standard deviation of the complex noise sample, sqrt(10^(n0_dBFS / 10))
    float large_scale;  // magnitude gain of every link (channel_t::get_large_scale_via_pathloss), 1 = none
    float clip;         // simulated radio (hw_simulator_t::clip_and_quantize): symmetric clipping amplitude, 0 = off
    float dac_lsb, adc_lsb;  // quantisation steps 2 clip / 2^bits of the TX / RX samples, 0 = clip only
    uint32_t seed_lo, seed_hi;
    // mode 2: link_t::set_pdp result (host), Doppler, absolute time of the window's first sample
    uint32_t n_taps;
    uint32_t tap_delay[9];
    float tap_scale[9];  // sqrt(tap power) / sqrt(40)
    float fD_Hz;
    double samp_rate;
    long long time_64;
    // knot distance (samples, a power of two >= 8) for the interpolated evaluation of the tap coefficients, 0 = every
    // sample exactly (k_channel_tdl; chosen by the host from the largest Doppler increment per sample)
    uint32_t interp_h;
};
// one sinusoid of one tap of one link (link.cpp:165-215): period in samples (sign = sense of rotation), start phase
struct tdl_sin_t {
    long long period;
    float phase0;
    float s2s;          // (float)(2 pi / period)
};
struct channel_params_t {
    uint32_t N_TX, N_RX;
    const float2* tx;  size_t tx_stride;   // [packet][N_TX][tx_stride]
    float2* rx;        size_t rx_stride;   // [packet][N_RX][rx_stride]
    uint32_t n_samples;                    // TX samples superimposed (rest of the window is noise only)
    const chan_meta_dev_t* meta;           // [packet]
    tdl_sin_t* sin_tab;                    // [packet][N_RX][N_TX][9][40], mode 2 only
    uint32_t tdl_min_h;                    // smallest non-zero chan_meta_dev_t::interp_h of the batch (sizes the knot buffer), 0 = none
};
constexpr uint32_t TDL_HALO = 512;         // largest tap delay (+ nothing else) the interpolated path stages in front of a tile
cudaError_t launch_channel(const channel_params_t& p, uint32_t n_packets, cudaStream_t st);
cudaError_t launch_channel_tdl(const channel_params_t& p, uint32_t n_packets, cudaStream_t st);

}  // namespace dnrp

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
    const uint32_t* pcc_lk;    // [98]
    const uint32_t* pdc_lk;    // [N_PDC_subc]
    const int8_t* stf_sign;    // [Nk]
    const int8_t* W;           // [n_cb_idx][N_TX][N_TS] codes 0, +-1, +-2 (= +-j)
    const float* W_scale;      // [n_cb_idx]
    const float* W_scale_opt;  // [n_cb_idx] "optimal DAC" variant
    const float2* twiddle;     // [N_os]
};

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
};

cudaError_t launch_tx_ofdm(const txofdm_params_t& p, uint32_t n_packets, cudaStream_t st);

}  // namespace dnrp

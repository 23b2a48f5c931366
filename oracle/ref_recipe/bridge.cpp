// bridge.cpp -- extern "C" test bridge into the reference's OWN PHY classes (compiled unmodified from /root/reference by
// the Makefile in this directory, with the third-party stand-ins of shim/).  TEST INFRASTRUCTURE ONLY: loaded by tests/,
// __graft_entry__.smoke() and bench.py's reference arm through ctypes (oracle/ref.py); the product never links it.
//
// What it drives, each through the reference's public entry point:
//   sp3::get_packet_sizes, pcc_t / pdc_t::get_k_i_linear, stf_t::get_stf, W_t::get_W          (ref_get_packet_sizes ...)
//   resampler_t::resample / resample_final_samples                                            (ref_resample)
//   tx_t::generate_tx_packet(tx_descriptor_t, radio::buffer_tx_t)                              (ref_tx_*)
//   sync_chunk_t::search over a radio::buffer_rx_t ring                                        (ref_sync_stream)
//   rx_synced_t::demoddecod_rx_pcc / demoddecod_rx_pdc, fec_t::decode_plcf_test                (ref_rx_*)
// Compiled with -fno-access-control: the IQ buffers of radio::buffer_tx_t / buffer_rx_t and a few intermediate stages of
// rx_synced_t are private members (the reference's radio layer reaches them as a friend).
#include <omp.h>

#include <algorithm>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "dectnrp/common/adt/miscellaneous.hpp"
#include "dectnrp/constants.hpp"
#include "dectnrp/phy/harq/process_pool.hpp"
#include "dectnrp/phy/interfaces/maclow_phy.hpp"
#include "dectnrp/phy/resample/resampler.hpp"
#include "dectnrp/phy/rx/rx_synced/rx_synced.hpp"
#include "dectnrp/phy/rx/sync/sync_chunk.hpp"
#include "dectnrp/phy/tx/tx.hpp"
#include "dectnrp/phy/worker_pool_config.hpp"
#include "dectnrp/radio/buffer_rx.hpp"
#include "dectnrp/radio/buffer_tx.hpp"
#include "dectnrp/sections_part3/beamforming_and_antenna_port_mapping.hpp"
#include "dectnrp/sections_part3/derivative/packet_sizes.hpp"
#include "dectnrp/sections_part3/fix/cbsegm.hpp"
#include "dectnrp/sections_part3/pcc.hpp"
#include "dectnrp/sections_part3/pdc.hpp"
#include "dectnrp/sections_part3/stf.hpp"

using namespace dectnrp;

extern "C" void shim_rm_warm(uint32_t cb_idx);

namespace {

sp3::packet_sizes_def_t to_psdef(const uint32_t* p) {
    sp3::packet_sizes_def_t q;
    q.u = p[0]; q.b = p[1]; q.PacketLengthType = p[2]; q.PacketLength = p[3];
    q.tm_mode_index = p[4]; q.mcs_index = p[5]; q.Z = p[6];
    return q;
}

void warm_rm_tables() {   // rate-matching tables of the shim are built on first use: do it before any thread runs
    static std::once_flag once;
    std::call_once(once, [] { for (uint32_t i = 0; i < 188; ++i) shim_rm_warm(i); });
}

// worker_pool_config_t as phy_config / worker_pool_t would fill it for a pool whose maximum packet is `ps`
phy::worker_pool_config_t make_config(const sp3::packet_sizes_t& ps, uint32_t os_min, uint32_t L, uint32_t M,
                                      bool chestim_lr, uint32_t chestim_stride) {
    phy::worker_pool_config_t c{};
    c.id = 0;
    c.radio_device_class_string = "bridge";
    c.radio_device_class.radio_device_class_string = "bridge";
    c.radio_device_class.u_min = ps.psdef.u;
    c.radio_device_class.b_min = ps.psdef.b;
    c.radio_device_class.N_TX_min = ps.tm_mode.N_TX;
    c.radio_device_class.mcs_index_min = ps.psdef.mcs_index;
    c.radio_device_class.M_DL_HARQ_min = 8;
    c.radio_device_class.M_connection_DL_HARQ_min = 8;
    c.radio_device_class.N_soft_min = 0;
    c.radio_device_class.Z_min = ps.psdef.Z;
    c.radio_device_class.PacketLength_min = ps.psdef.PacketLength;
    c.maximum_packet_sizes = ps;
    c.os_min = os_min;
    c.enforce_dectnrp_samp_rate_by_resampling = L != M;
    c.nof_jobs = 64;
    c.rx_ant_streams_length_slots = 24;
    c.rx_chunk_length_u8subslot = 32;
    c.rx_chunk_unit_length_u8subslot = 2;
    c.rx_job_regular_period = 0;
    c.threads_core_prio_config_sync_vec.resize(1);
    c.threads_core_prio_config_tx_rx_vec.resize(1);
    c.chestim_mode_lr_default = chestim_lr;
    c.chestim_mode_lr_t_stride_default = chestim_stride;
    c.json_export_length = 0;
    const uint32_t dect_rate = ps.numerology.B_u_b_DFT * os_min;
    c.resampler_param = phy::resampler_param_t(
        phy::resampler_t::get_samp_rate_converted_with_temporary_overflow(dect_rate, L, M), L, M);
    return c;
}

struct tx_handle_t {
    sp3::packet_sizes_t ps_max;
    std::unique_ptr<phy::harq::process_pool_t> hpp;
    std::unique_ptr<phy::tx_t> tx;
    std::unique_ptr<radio::buffer_tx_t> buf;
    std::mutex m;
    std::condition_variable cv;
    uint32_t cnt = 0;
    phy::harq::process_tx_t* hp = nullptr;
    uint32_t hp_id = 0;
};

struct rx_handle_t {
    sp3::packet_sizes_t ps_max;
    phy::worker_pool_config_t cfg;
    uint32_t N_RX = 1, ring_len = 0, unit = 0;
    std::unique_ptr<radio::buffer_rx_t> ring;
    std::unique_ptr<phy::rx_synced_t> rx;
    std::unique_ptr<phy::harq::process_pool_t> hpp;
    phy::sync_report_t sync;
    phy::harq::process_rx_t* hp = nullptr;
    uint32_t hp_id = 0;
};

uint32_t round_up(uint32_t v, uint32_t m) { return (v + m - 1) / m * m; }

}  // namespace

extern "C" {

int ref_bridge_version(void) { return 2; }
int ref_max_threads(void) { return omp_get_max_threads(); }

// ---- sections_part3 ------------------------------------------------------------------------------------------------------
// out: the fields of dnrp_packet_sizes_t behind psdef, in its order (31 words)
int ref_get_packet_sizes(const uint32_t* psdef, uint32_t* out) {
    const auto q = sp3::get_packet_sizes(to_psdef(psdef));
    if (!q.has_value()) return -1;
    const sp3::packet_sizes_t& p = q.value();
    srsran_cbsegm_t seg{};
    if (sp3::fix::srsran_cbsegm_FIX(&seg, p.N_TB_bits, p.psdef.Z) != SRSRAN_SUCCESS) return -2;
    uint32_t i = 0;
    out[i++] = p.tm_mode.N_eff_TX; out[i++] = p.tm_mode.N_SS; out[i++] = p.tm_mode.N_TS; out[i++] = p.tm_mode.N_TX;
    out[i++] = p.tm_mode.cl ? 1 : 0; out[i++] = p.mcs.N_bps;
    out[i++] = p.numerology.N_b_DFT; out[i++] = p.numerology.N_b_CP; out[i++] = p.numerology.N_b_OCC;
    out[i++] = p.N_PACKET_symb; out[i++] = p.N_DF_symb; out[i++] = p.N_PDC_subc; out[i++] = p.N_DRS_subc;
    out[i++] = p.G; out[i++] = p.N_PDC_bits; out[i++] = p.N_TB_bits; out[i++] = p.N_TB_byte; out[i++] = p.C;
    out[i++] = seg.K1; out[i++] = seg.K2; out[i++] = seg.C1; out[i++] = seg.C2;
    out[i++] = p.N_samples_STF; out[i++] = p.N_samples_STF_CP_only; out[i++] = p.N_samples_DF; out[i++] = p.N_samples_GI;
    out[i++] = p.N_samples_packet_no_GI; out[i++] = p.N_samples_packet;
    return (int)i;
}

// kind 0: pcc_t::get_k_i_linear, 1: pdc_t::get_k_i_linear; returns the number of entries (written up to cap)
int ref_k_i_linear(int kind, uint32_t b, uint32_t N_TS, uint32_t* out, uint32_t cap) {
    const std::vector<uint32_t> v = kind == 0 ? sp3::pcc_t::get_k_i_linear(b, N_TS) : sp3::pdc_t::get_k_i_linear(b, N_TS);
    for (size_t i = 0; i < v.size() && i < cap; ++i) out[i] = v[i];
    return (int)v.size();
}

// stf_t::get_stf(b, N_eff_TX): complex values at the occupied cells (re, im interleaved); returns the count
int ref_stf(uint32_t b, uint32_t N_eff_TX, float* out, uint32_t cap) {
    sp3::stf_t stf(b, N_eff_TX, 1.0f);
    const std::vector<cf_t>& v = stf.get_stf(b, N_eff_TX);
    for (size_t i = 0; i < v.size() && 2 * i + 1 < cap; ++i) { out[2 * i] = __real__(v[i]); out[2 * i + 1] = __imag__(v[i]); }
    return (int)v.size();
}

// W_t::get_W(N_TS, N_TX)[idx]: N_TX * N_TS complex weights; returns the number of matrices of that table
int ref_W(uint32_t N_TS, uint32_t N_TX, uint32_t idx, float* out, uint32_t cap) {
    sp3::W_t W;
    const auto& t = W.get_W(N_TS, N_TX);
    if (idx < t.size())
        for (size_t i = 0; i < t[idx].size() && 2 * i + 1 < cap; ++i) { out[2 * i] = __real__(t[idx][i]); out[2 * i + 1] = __imag__(t[idx][i]); }
    return (int)t.size();
}

// ---- resampler_t: one antenna, input fed in `unit`-sized calls, then flushed; user 0 TX, 1 SYNC, 2 RX_SYNCED --------------
int ref_resample(uint32_t L, uint32_t M, uint32_t os_min, uint32_t user, const float* in, uint32_t n_in, uint32_t unit, int flush,
                 float* out, uint32_t cap) {
    phy::resampler_t rs(1, L, M, phy::resampler_param_t::f_pass_norm[user][os_min], phy::resampler_param_t::f_stop_norm[user][os_min],
                        phy::resampler_param_t::PASSBAND_RIPPLE_DONT_CARE, phy::resampler_param_t::f_stop_att_dB[user][os_min]);
    rs.reset();
    const uint32_t n_out_max = rs.get_N_samples_after_resampling(n_in) + L + 8;
    std::vector<cf_t> o(n_out_max + 64);
    uint32_t w = 0, r = 0;
    if (unit == 0) unit = n_in;
    while (r < n_in) {
        const uint32_t n = std::min(unit, n_in - r);
        std::vector<const cf_t*> iv{reinterpret_cast<const cf_t*>(in) + r};
        std::vector<cf_t*> ov{o.data() + w};
        w += rs.resample(iv, ov, n);
        r += n;
    }
    if (flush) {
        std::vector<cf_t*> ov{o.data() + w};
        w += rs.resample_final_samples(ov);
    }
    if (w > cap) return -(int)w;
    std::memcpy(out, o.data(), sizeof(cf_t) * w);
    return (int)w;
}
uint32_t ref_resample_n_after(uint32_t L, uint32_t M, uint32_t os_min, uint32_t user, uint32_t n_in) {
    phy::resampler_t rs(1, L, M, phy::resampler_param_t::f_pass_norm[user][os_min], phy::resampler_param_t::f_stop_norm[user][os_min],
                        phy::resampler_param_t::PASSBAND_RIPPLE_DONT_CARE, phy::resampler_param_t::f_stop_att_dB[user][os_min]);
    return rs.get_N_samples_after_resampling(n_in);
}

// ---- transmitter --------------------------------------------------------------------------------------------------------
void* ref_tx_create(const uint32_t* psdef_max, uint32_t os_min, uint32_t L, uint32_t M, uint32_t buffer_samples) {
    warm_rm_tables();
    const auto q = sp3::get_packet_sizes(to_psdef(psdef_max));
    if (!q.has_value()) return nullptr;
    auto* h = new tx_handle_t();
    h->ps_max = q.value();
    h->hpp = std::make_unique<phy::harq::process_pool_t>(h->ps_max, 1, 1);
    const uint32_t dect_rate = h->ps_max.numerology.B_u_b_DFT * os_min;
    const phy::resampler_param_t rp(phy::resampler_t::get_samp_rate_converted_with_temporary_overflow(dect_rate, L, M), L, M);
    h->tx = std::make_unique<phy::tx_t>(h->ps_max, os_min, rp);
    h->buf = std::make_unique<radio::buffer_tx_t>(0, h->ps_max.tm_mode.N_TX, buffer_samples, h->m, h->cv, h->cnt);
    return h;
}
void ref_tx_destroy(void* hv) { delete static_cast<tx_handle_t*>(hv); }
void ref_tx_add_network_id(void* hv, uint32_t network_id) { static_cast<tx_handle_t*>(hv)->tx->add_new_network_id(network_id); }

// tx_t::generate_tx_packet for one packet.  new_tb != 0: a fresh HARQ process (rv must then be 0: the reference encodes
// only at rv 0 and reuses the stored code block for later redundancy versions); new_tb == 0: the running process with the
// given rv.  meta = {optimal_scaling_DAC, DAC_scale, iq_phase_rad, iq_phase_increment_s2s_post_resampling_rad,
// GI_percentage}.  Outputs (any may be NULL): iq [N_TX][iq_stride] cf32, n_samples, pcc_d [25] and pdc_d [ceil(G/8)] (the
// coded, scrambled bits as the HARQ buffers hold them).
int ref_tx_packet(void* hv, const uint32_t* psdef, uint32_t plcf_type, uint32_t network_id, const uint8_t* plcf, const uint8_t* tb,
                  uint32_t codebook_index, const float* meta, uint32_t rv, int new_tb, float* iq, uint32_t iq_stride,
                  uint32_t* n_samples, uint8_t* pcc_d, uint8_t* pdc_d) {
    auto* h = static_cast<tx_handle_t*>(hv);
    const auto q = sp3::get_packet_sizes(to_psdef(psdef));
    if (!q.has_value()) return -1;
    const sp3::packet_sizes_t& ps = q.value();
    if (new_tb) {
        if (h->hp != nullptr) {   // terminate the previous process
            auto* old = h->hpp->get_process_tx_running(h->hp_id, phy::harq::finalize_tx_t::reset_and_terminate);
            if (old) old->finalize();
            h->hp = nullptr;
        }
        h->hp = h->hpp->get_process_tx(plcf_type, network_id, ps.psdef, phy::harq::finalize_tx_t::keep_rv_and_keep_running);
        if (!h->hp) return -2;
        h->hp_id = h->hp->get_id();
        std::memcpy(h->hp->get_a_plcf(), plcf, 10);
        std::memcpy(h->hp->get_a_tb(), tb, ps.N_TB_byte);
    } else {
        if (!h->hp) return -3;
        h->hp = h->hpp->get_process_tx_running(h->hp_id, phy::harq::finalize_tx_t::keep_rv_and_keep_running);
        if (!h->hp) return -4;
    }
    h->hp->rv = rv;
    const phy::tx_meta_t tx_meta = {.optimal_scaling_DAC = meta[0] != 0.0f, .DAC_scale = meta[1], .iq_phase_rad = meta[2],
                                    .iq_phase_increment_s2s_post_resampling_rad = meta[3], .GI_percentage = (uint32_t)meta[4]};
    const radio::buffer_tx_meta_t btm = {.tx_order_id = 0, .tx_order_id_expect_next = -1, .tx_time_64 = 0, .busy_wait_us = 0};
    const phy::tx_descriptor_t desc(*h->hp, codebook_index, tx_meta, btm);
    radio::buffer_tx_t& b = *h->buf;
    if (!b.try_lock_outer()) return -5;
    for (uint32_t a = 0; a < b.nof_antennas; ++a) std::memset(b.ant_streams[a], 0, sizeof(cf_t) * b.ant_streams_length_samples);
    h->tx->generate_tx_packet(desc, b);
    const uint32_t n = b.tx_length_samples.load();
    if (n_samples) *n_samples = n;
    if (iq) {
        if (n > iq_stride) return -6;
        for (uint32_t a = 0; a < ps.tm_mode.N_TX; ++a) {
            std::memset(iq + (size_t)2 * a * iq_stride, 0, sizeof(float) * 2 * iq_stride);
            std::memcpy(iq + (size_t)2 * a * iq_stride, b.ant_streams[a], sizeof(cf_t) * n);
        }
    }
    if (pcc_d) std::memcpy(pcc_d, h->hp->get_hb_plcf()->get_d(), 25);
    if (pdc_d) std::memcpy(pdc_d, h->hp->get_hb_tb()->get_d(), (ps.G + 7) / 8);
    b.set_transmitted_or_abort();   // what the radio thread does after transmission
    h->hp->finalize();   // keep_rv_and_keep_running: unlocks the inner lock only
    return 0;
}

// ---- receiver -----------------------------------------------------------------------------------------------------------
// One rx_synced_t (+ its ring buffer of n_window samples per antenna at the hardware rate) for a worker pool whose maximum
// packet is psdef_max; N_RX = N_TX of that maximum, as in the reference (rx_synced.cpp: N_RX(maximum_packet_sizes.tm_mode.N_TX)).
void* ref_rx_create(const uint32_t* psdef_max, uint32_t os_min, uint32_t L, uint32_t M, uint32_t n_window, uint32_t chestim_lr,
                    uint32_t chestim_stride) {
    warm_rm_tables();
    const auto q = sp3::get_packet_sizes(to_psdef(psdef_max));
    if (!q.has_value()) return nullptr;
    auto* h = new rx_handle_t();
    h->ps_max = q.value();
    h->cfg = make_config(h->ps_max, os_min, L, M, chestim_lr != 0, chestim_stride);
    h->N_RX = h->ps_max.tm_mode.N_TX;
    const uint32_t u8 = h->cfg.resampler_param.hw_samp_rate / constants::u8_subslots_per_sec;
    h->unit = u8 * h->cfg.rx_chunk_unit_length_u8subslot;
    // ring: the window, a unit of slack for the resampler's look-ahead, rounded to units; >= 8 units
    h->ring_len = std::max(round_up(n_window + 2 * h->unit, h->unit), 8 * h->unit);
    h->ring = std::make_unique<radio::buffer_rx_t>(0, h->N_RX, h->ring_len, h->cfg.resampler_param.hw_samp_rate, h->unit, 0, 100);
    h->ring->rx_time_passed_64.store((int64_t)h->ring_len);
    h->rx = std::make_unique<phy::rx_synced_t>(*h->ring, h->cfg, h->unit);
    h->hpp = std::make_unique<phy::harq::process_pool_t>(h->ps_max, 1, 1);
    return h;
}
void ref_rx_destroy(void* hv) { delete static_cast<rx_handle_t*>(hv); }
void ref_rx_add_network_id(void* hv, uint32_t network_id) { static_cast<rx_handle_t*>(hv)->rx->add_new_network_id(network_id); }
uint32_t ref_rx_n_rx(void* hv) { return static_cast<rx_handle_t*>(hv)->N_RX; }

// demoddecod_rx_pcc on the window iq [N_RX][iq_stride] (n_window valid samples, hardware rate); sync_in = {fine_peak_time,
// u, b, N_eff_TX, cfo_fractional_rad, cfo_integer_rad, rms[0..7]} (the reference asserts that synchronisation provided the RMS
// of at least one antenna; antennas with rms 0 are estimated by the STF stage).  Outputs (may be NULL): plcf_ok [2], plcf [2][10] (both PLCF types,
// from fec_t::decode_plcf_test on the receiver's own LLRs), cl_bf [2] (bit 0 closed loop, bit 1 beamforming), llr_pcc [196],
// rep = {snr_dB, cfo_fractional_rad (after the STF stage), sto_fractional, rms[0..7]}, has_plcf [2]: what plcf_decoder_t
// accepted after its radio-device-class check.
int ref_rx_pcc(void* hv, const float* iq, uint32_t iq_stride, uint32_t n_window, const double* sync_in, uint8_t* plcf_ok, uint8_t* plcf,
               uint8_t* cl_bf, int16_t* llr_pcc, float* rep, uint8_t* has_plcf) {
    auto* h = static_cast<rx_handle_t*>(hv);
    if (n_window > h->ring_len) return -1;
    for (uint32_t a = 0; a < h->N_RX; ++a) {
        std::memset(h->ring->ant_streams[a], 0, sizeof(cf_t) * (h->ring_len + h->unit));
        std::memcpy(h->ring->ant_streams[a], iq + (size_t)2 * a * iq_stride, sizeof(cf_t) * n_window);
    }
    h->rx->reset_for_next_pcc();
    phy::sync_report_t s(h->N_RX);
    s.u = (uint32_t)sync_in[1]; s.b = (uint32_t)sync_in[2]; s.N_eff_TX = (uint32_t)sync_in[3];
    s.fine_peak_time_64 = (int64_t)sync_in[0];
    s.coarse_peak_time_64 = s.fine_peak_time_64;
    s.cfo_fractional_rad = (float)sync_in[4];
    s.cfo_integer_rad = (float)sync_in[5];
    for (uint32_t a = 0; a < h->N_RX && a < 8; ++a) s.rms_array.at(a) = (float)sync_in[6 + a];
    h->sync = s;
    const phy::pcc_report_t pr = h->rx->demoddecod_rx_pcc(h->sync);
    if (has_plcf) {
        has_plcf[0] = pr.plcf_decoder.get_plcf_base(1) != nullptr;
        has_plcf[1] = pr.plcf_decoder.get_plcf_base(2) != nullptr;
    }
    if (rep) {
        rep[0] = pr.snr_dB; rep[1] = h->sync.cfo_fractional_rad; rep[2] = h->sync.sto_fractional;
        for (uint32_t a = 0; a < 8; ++a) rep[3 + a] = a < h->N_RX ? h->sync.rms_array.at(a) : 0.0f;
    }
    auto* hb = h->rx->hb_rx_plcf.get();
    if (llr_pcc) std::memcpy(llr_pcc, hb->get_d(), sizeof(int16_t) * constants::pcc_bits);
    if (plcf_ok || plcf || cl_bf)
        for (uint32_t ty = 1; ty <= 2; ++ty) {
            sp3::fec_cfg_t cfg{};
            cfg.PLCF_type = ty;
            const bool ok = h->rx->fec->decode_plcf_test(cfg, *hb, ty);
            if (plcf_ok) plcf_ok[ty - 1] = ok;
            if (plcf) std::memcpy(plcf + 10 * (ty - 1), hb->get_a(), ty == 1 ? 5 : 10);
            if (cl_bf) cl_bf[ty - 1] = (uint8_t)((cfg.closed_loop ? 1 : 0) | (cfg.beamforming ? 2 : 0));
        }
    return 0;
}

// demoddecod_rx_pdc for the packet of the preceding ref_rx_pcc.  new_tb != 0: fresh HARQ process (soft buffer reset),
// otherwise the running one (soft combining).  Outputs (may be NULL): tb [N_TB_byte], llr_pdc [G] (descrambled in place by
// the decoder, as the reference leaves them), rep = {crc_status, snr_dB, tm_3_7 idx, tm_3_7 reciprocal idx}.
int ref_rx_pdc(void* hv, const uint32_t* psdef, uint32_t plcf_type, uint32_t network_id, uint32_t rv, int new_tb, uint8_t* tb,
               int16_t* llr_pdc, float* rep) {
    auto* h = static_cast<rx_handle_t*>(hv);
    const auto q = sp3::get_packet_sizes(to_psdef(psdef));
    if (!q.has_value()) return -1;
    const sp3::packet_sizes_t& ps = q.value();
    if (new_tb) {
        if (h->hp != nullptr) {
            auto* old = h->hpp->get_process_rx_running(h->hp_id, 0, phy::harq::finalize_rx_t::reset_and_terminate);
            if (old) old->finalize(false);
            h->hp = nullptr;
        }
        h->hp = h->hpp->get_process_rx(plcf_type, network_id, ps.psdef, rv, phy::harq::finalize_rx_t::keep_running);
        if (!h->hp) return -2;
        h->hp_id = h->hp->get_id();
    } else {
        if (!h->hp) return -3;
        h->hp = h->hpp->get_process_rx_running(h->hp_id, rv, phy::harq::finalize_rx_t::keep_running);
        if (!h->hp) return -4;
    }
    const phy::maclow_phy_t mp(h->hp);
    const phy::pdc_report_t pr = h->rx->demoddecod_rx_pdc(mp);
    if (tb) std::memcpy(tb, h->hp->get_hb_tb()->get_a(), ps.N_TB_byte);
    if (llr_pdc) std::memcpy(llr_pdc, h->hp->get_hb_tb()->get_d(), sizeof(int16_t) * ps.G);
    if (rep) {
        rep[0] = pr.crc_status ? 1.0f : 0.0f; rep[1] = pr.snr_dB;
        rep[2] = (float)pr.mimo_report.tm_3_7_beamforming_idx; rep[3] = (float)pr.mimo_report.tm_3_7_beamforming_reciprocal_idx;
    }
    h->hp->finalize(pr.crc_status);
    return pr.crc_status ? 1 : 0;
}

// ---- packet synchronisation: sync_chunk_t::search over ONE stream -----------------------------------------------------------
// iq [N_RX][iq_stride] with n_samples valid samples at the hardware rate (zeros behind).  The whole stream is one chunk
// whose detection range is chunk_length hardware samples (0: all of it) plus the reference's overlap B; search() is called
// until the chunk is used up.  out [max_out][12] doubles per packet: fine_peak_time_64, coarse_peak_time_64, N_eff_TX,
// detection_time_local, detection_ant_idx, detection_metric, detection_rms, coarse_peak_time_local, cfo_fractional_rad,
// fine_peak_time_local, detection_time_with_jump_back_local, u; coarse [max_out][8], rms [max_out][8] (may be NULL).
int ref_sync_stream(const uint32_t* psdef_max, uint32_t os_min, uint32_t L, uint32_t M, uint32_t N_RX, const float* iq, uint32_t iq_stride,
                    uint32_t n_samples, uint32_t chunk_length, uint32_t max_out, double* out, float* coarse, float* rms) {
    const auto q = sp3::get_packet_sizes(to_psdef(psdef_max));
    if (!q.has_value()) return -1;
    phy::worker_pool_config_t cfg = make_config(q.value(), os_min, L, M, true, 2);
    const uint32_t u8 = cfg.resampler_param.hw_samp_rate / constants::u8_subslots_per_sec;
    const uint32_t unit = u8 * cfg.rx_chunk_unit_length_u8subslot;
    if (chunk_length == 0) chunk_length = n_samples;
    chunk_length = round_up(chunk_length, unit * L);
    // samples the chunk reads behind its end: overlap + one pattern + peak search + cross-correlation, generously 8 STFs
    const uint32_t stf_hw = sp3::stf_t::get_N_samples_stf(q->psdef.u) * q->psdef.b * os_min * L / M + L;
    const uint32_t ring_len = std::max(round_up(std::max(chunk_length, n_samples) + 8 * stf_hw + 4 * unit, unit), 8 * unit);
    radio::buffer_rx_t ring(0, N_RX, ring_len, cfg.resampler_param.hw_samp_rate, unit, 0, 100);
    for (uint32_t a = 0; a < N_RX; ++a) std::memcpy(ring.ant_streams[a], iq + (size_t)2 * a * iq_stride, sizeof(cf_t) * n_samples);
    ring.rx_time_passed_64.store((int64_t)ring_len);
    phy::sync_chunk_t chunk(ring, cfg, chunk_length, 4 * ring_len, 0, unit, [](int64_t) {});
    chunk.wait_for_first_chunk_nto(0);
    int n = 0;
    while ((uint32_t)n < max_out) {
        const auto r = chunk.search();
        if (!r.has_value()) break;
        double* o = out + 12 * (size_t)n;
        o[0] = (double)r->fine_peak_time_64; o[1] = (double)r->coarse_peak_time_64; o[2] = r->N_eff_TX;
        o[3] = r->detection_time_local; o[4] = r->detection_ant_idx; o[5] = r->detection_metric; o[6] = r->detection_rms;
        o[7] = r->coarse_peak_time_local; o[8] = r->cfo_fractional_rad; o[9] = r->fine_peak_time_local;
        o[10] = r->detection_time_with_jump_back_local; o[11] = r->u;
        for (uint32_t a = 0; a < 8; ++a) {
            if (coarse) coarse[8 * (size_t)n + a] = a < r->coarse_peak_array.get_nof_antennas() ? r->coarse_peak_array.at(a) : 0.0f;
            if (rms) rms[8 * (size_t)n + a] = a < r->rms_array.get_nof_antennas() ? r->rms_array.at(a) : 0.0f;
        }
        ++n;
        if (chunk.is_chunk_completely_processed()) break;
    }
    return n;
}

// ---- a pool of workers for the bench's reference arm ----------------------------------------------------------------------
// The reference scales by one worker thread = one tx_t / rx_synced_t pair (worker_tx_rx.cpp:43-67) fed by sync workers
// (worker_sync.cpp).  Here: n_threads workers, each with its own ring buffer, sync_chunk_t and rx_synced_t; a window is
// handled start to finish by one worker: sync_chunk_t::search (first packet of the window) -> demoddecod_rx_pcc ->
// demoddecod_rx_pdc.  No real-time pacing (the ring is complete before the search starts).
struct worker_t {
    std::unique_ptr<radio::buffer_rx_t> ring;
    std::unique_ptr<phy::sync_chunk_t> chunk;
    std::unique_ptr<phy::rx_synced_t> rx;
    std::unique_ptr<phy::harq::process_pool_t> hpp;
    double ms[3] = {0, 0, 0};
};
struct pool_t {
    sp3::packet_sizes_t ps_max;
    phy::worker_pool_config_t cfg;
    uint32_t N_RX = 1, ring_len = 0, unit = 0, chunk_length = 0;
    std::vector<std::unique_ptr<worker_t>> w;
};

void* ref_pool_create(const uint32_t* psdef_max, uint32_t os_min, uint32_t L, uint32_t M, uint32_t n_window, uint32_t chestim_lr,
                      uint32_t chestim_stride, uint32_t n_threads, uint32_t network_id) {
    warm_rm_tables();
    const auto q = sp3::get_packet_sizes(to_psdef(psdef_max));
    if (!q.has_value()) return nullptr;
    auto* p = new pool_t();
    p->ps_max = q.value();
    p->cfg = make_config(p->ps_max, os_min, L, M, chestim_lr != 0, chestim_stride);
    p->N_RX = p->ps_max.tm_mode.N_TX;
    const uint32_t u8 = p->cfg.resampler_param.hw_samp_rate / constants::u8_subslots_per_sec;
    p->unit = u8 * p->cfg.rx_chunk_unit_length_u8subslot;
    p->chunk_length = round_up(n_window, p->unit * L);
    const uint32_t stf_hw = sp3::stf_t::get_N_samples_stf(q->psdef.u) * q->psdef.b * os_min * L / M + L;
    p->ring_len = std::max(round_up(p->chunk_length + 8 * stf_hw + 4 * p->unit, p->unit), 8 * p->unit);
    // every worker builds its own Wiener LUTs (channel_lut.cpp, tens of seconds at b = 16 -- the reference's start-up
    // cost per worker thread); the constructors touch instance state only, so the workers are built concurrently
    const uint32_t nw = std::max(n_threads, 1u);
    p->w.resize(nw);
#pragma omp parallel for schedule(static, 1) num_threads((int)nw)
    for (uint32_t t = 0; t < nw; ++t) {
        auto w = std::make_unique<worker_t>();
        w->ring = std::make_unique<radio::buffer_rx_t>(t, p->N_RX, p->ring_len, p->cfg.resampler_param.hw_samp_rate, p->unit, 0, 100);
        w->ring->rx_time_passed_64.store((int64_t)p->ring_len);
        w->chunk = std::make_unique<phy::sync_chunk_t>(*w->ring, p->cfg, p->chunk_length, 4 * p->ring_len, 0, p->unit, [](int64_t) {});
        w->rx = std::make_unique<phy::rx_synced_t>(*w->ring, p->cfg, p->unit);
        w->rx->add_new_network_id(network_id);
        w->hpp = std::make_unique<phy::harq::process_pool_t>(p->ps_max, 1, 1);
        p->w[t] = std::move(w);
    }
    return p;
}
void ref_pool_destroy(void* pv) { delete static_cast<pool_t*>(pv); }
uint32_t ref_pool_threads(void* pv) { return (uint32_t)static_cast<pool_t*>(pv)->w.size(); }

// n windows iq [n][N_RX][iq_stride] (n_window valid samples each).  Outputs per window (any may be NULL): crc_ok, tb
// [n][N_TB_byte], fine_peak (-1: no packet found), plcf_ok (type 1).  stage_ms [3]: sync / PCC / PDC milliseconds summed
// over all workers.  Returns the number of windows with a passing transport block CRC.
int ref_pool_rx_windows(void* pv, uint32_t n, const float* iq, uint32_t iq_stride, uint32_t n_window, const uint32_t* psdef,
                        uint32_t plcf_type, uint32_t network_id, uint8_t* crc_ok, uint8_t* tb, int32_t* fine_peak, uint8_t* plcf_ok,
                        double* stage_ms) {
    auto* p = static_cast<pool_t*>(pv);
    const auto q = sp3::get_packet_sizes(to_psdef(psdef));
    if (!q.has_value() || n_window > p->chunk_length) return -1;
    const sp3::packet_sizes_t ps = q.value();
    for (auto& w : p->w) w->ms[0] = w->ms[1] = w->ms[2] = 0;
    int n_ok = 0;
#pragma omp parallel for schedule(dynamic, 1) num_threads((int)p->w.size()) reduction(+ : n_ok)
    for (int64_t i = 0; i < (int64_t)n; ++i) {
        worker_t& w = *p->w[(size_t)omp_get_thread_num()];
        for (uint32_t a = 0; a < p->N_RX; ++a) {
            std::memcpy(w.ring->ant_streams[a], iq + ((size_t)i * p->N_RX + a) * 2 * iq_stride, sizeof(cf_t) * n_window);
            std::memset(w.ring->ant_streams[a] + n_window, 0, sizeof(cf_t) * (p->ring_len + p->unit - n_window));
        }
        if (crc_ok) crc_ok[i] = 0;
        if (fine_peak) fine_peak[i] = -1;
        if (plcf_ok) plcf_ok[i] = 0;
        double t0 = omp_get_wtime();
        w.chunk->wait_for_first_chunk_nto(0);
        auto r = w.chunk->search();
        double t1 = omp_get_wtime();
        w.ms[0] += (t1 - t0) * 1e3;
        if (!r.has_value()) continue;
        phy::sync_report_t s = r.value();
        if (fine_peak) fine_peak[i] = (int32_t)s.fine_peak_time_64;
        if (s.fine_peak_time_64 < 0 || s.fine_peak_time_64 + (int64_t)(ps.N_samples_packet_no_GI * p->cfg.os_min) > (int64_t)p->ring_len) continue;
        w.rx->reset_for_next_pcc();
        const phy::pcc_report_t pr = w.rx->demoddecod_rx_pcc(s);
        double t2 = omp_get_wtime();
        w.ms[1] += (t2 - t1) * 1e3;
        const bool has = pr.plcf_decoder.get_plcf_base(plcf_type) != nullptr || w.rx->hb_rx_plcf->get_softbuffer_d_type_1()->cb_crc[0];
        if (plcf_ok) plcf_ok[i] = w.rx->hb_rx_plcf->get_softbuffer_d_type_1()->cb_crc[0] ? 1 : 0;
        (void)has;
        // the MAC's decision between PCC and PDC: the bench knows the packet geometry (as its firmware would from the PLCF)
        auto* hp = w.hpp->get_process_rx(plcf_type, network_id, ps.psdef, 0, phy::harq::finalize_rx_t::reset_and_terminate);
        if (!hp) continue;
        const phy::maclow_phy_t mp(hp);
        const phy::pdc_report_t dr = w.rx->demoddecod_rx_pdc(mp);
        if (dr.crc_status) {
            ++n_ok;
            if (crc_ok) crc_ok[i] = 1;
            if (tb) std::memcpy(tb + (size_t)i * ps.N_TB_byte, hp->get_hb_tb()->get_a(), ps.N_TB_byte);
        }
        hp->finalize(dr.crc_status);
        w.ms[2] += (omp_get_wtime() - t2) * 1e3;
    }
    if (stage_ms) {
        stage_ms[0] = stage_ms[1] = stage_ms[2] = 0;
        for (auto& w : p->w)
            for (int k = 0; k < 3; ++k) stage_ms[k] += w->ms[k];
    }
    return n_ok;
}

}  // extern "C"

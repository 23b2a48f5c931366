// multi_plan_firmware.cpp -- test program for include/dnrp_host.hpp: ONE process_windows call over packets of several
// geometries (two packet lengths x two MCS), two network ids and both PLCF types, the way the reference's firmware
// learns every packet's sizes from its PLCF (tpoint.cpp:119-141, worksub_psdef) and hands them to the PHY through
// hp_rx (rx_synced.cpp:325-350).  With two or more GPUs it also runs one batch_phy_t per GPU inside this one process.
// Prints one JSON line; exit code 0 = all checks passed, 3 = no usable GPU (the library has no CPU fallback).
#include <cuda_runtime_api.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <random>

#include "dnrp_host.hpp"

using namespace dectnrp_b200;

static const uint32_t NETS[2] = {123456789u, 0x11223344u};

struct firmware_t final : public upper::tpoint_t {
    firmware_t(const sp3::packet_sizes_t& ps_max, uint32_t n_proc) : pool(ps_max, n_proc, n_proc), max_psdef(ps_max.psdef) {}

    phy::maclow_phy_t work_pcc(const phy::phy_maclow_t& in) override {
        ++n_work_pcc;
        // this firmware sends PLCF type 1 for even tags and type 2 for odd ones; both carry length, MCS and short network id
        for (uint32_t type = 1; type <= 2; ++type) {
            const uint8_t* plcf = in.pcc_report.get_plcf_bytes(type);
            if (!plcf) continue;
            const uint32_t tag = type == 1 ? (((uint32_t)plcf[2] << 8) | plcf[3]) : (((uint32_t)plcf[3] << 8) | plcf[4]);
            auto it = sent.find(tag);
            if (it == sent.end() || it->second.plcf_type != type) continue;
            // the geometry as announced by the PLCF: packet length and MCS fields (byte 0 / byte 1 for type 1 like tfw_loopback)
            sp3::packet_sizes_def_t psdef = max_psdef;
            psdef.PacketLength = (uint32_t)(plcf[0] & 0x0f) + 1;
            psdef.mcs_index = type == 1 ? (uint32_t)(plcf[4] & 0x07) : (uint32_t)(plcf[5] & 0x0f);
            if (std::memcmp(&psdef, &it->second.psdef, sizeof(psdef)) != 0) { ++n_plcf_mismatch; continue; }
            const uint32_t net = (plcf[1] == (NETS[0] & 0xff)) ? NETS[0] : NETS[1];
            if (net != it->second.network_id) { ++n_plcf_mismatch; continue; }
            auto* hp = pool.get_process_rx(type, net, psdef, 0, phy::harq::finalize_rx_t::reset_and_terminate);
            if (!hp) return phy::maclow_phy_t();
            return phy::maclow_phy_t(hp, phy::maclow_phy_handle_t(phy::handle_pcc2pdc_t::th10, tag));
        }
        return phy::maclow_phy_t();
    }
    phy::machigh_phy_t work_pdc(const phy::phy_machigh_t& in) override {
        ++n_work_pdc;
        const sent_t& s = sent[in.maclow_phy.get_handle_lrdid()];
        if (in.pdc_report.N_TB_byte != s.tb.size() || std::memcmp(in.pdc_report.a_tb, s.tb.data(), s.tb.size()) != 0) ++n_tb_mismatch;
        return phy::machigh_phy_t();
    }
    phy::machigh_phy_t work_pdc_error(const phy::phy_machigh_t&) override {
        ++n_work_pdc_error;
        return phy::machigh_phy_t();
    }
    struct sent_t { sp3::packet_sizes_def_t psdef; uint32_t network_id, plcf_type; std::vector<uint8_t> tb; };
    phy::harq::process_pool_t pool;
    sp3::packet_sizes_def_t max_psdef;
    std::map<uint32_t, sent_t> sent;
    uint32_t n_work_pcc{0}, n_work_pdc{0}, n_work_pdc_error{0}, n_plcf_mismatch{0}, n_tb_mismatch{0};
};

// one loopback round on `device`: n packets cycling through 4 geometries x 2 networks x 2 PLCF types
static int run(int device, uint32_t n, firmware_t*& fw_out, std::unique_ptr<phy::batch_phy_t>& bp_out) {
    dnrp_config_t cfg{};
    cfg.psdef = {1, 2, 1, 2, 0, 4, 2048};   // the maximum geometry: mu 1, beta 2, two slots, 16-QAM 3/4
    cfg.N_RX = 1; cfg.oversampling = 1; cfg.resample_L = 1; cfg.resample_M = 1;
    cfg.max_batch = n; cfg.network_id = NETS[0]; cfg.plcf_type = 1;
    cfg.chestim_mode_lr = 1; cfg.chestim_mode_lr_t_stride = 2;
    bp_out = std::make_unique<phy::batch_phy_t>(cfg, device, n);
    phy::batch_phy_t& bp = *bp_out;
    const sp3::packet_sizes_def_t geo[4] = {{1, 2, 1, 2, 0, 4, 2048}, {1, 2, 1, 1, 0, 4, 2048}, {1, 2, 1, 2, 0, 2, 2048}, {1, 2, 1, 1, 0, 2, 2048}};
    for (const auto& g : geo) bp.add_plan(g);
    bp.add_network_id(NETS[1]);
    const sp3::packet_sizes_t ps_max = bp.get_packet_sizes();
    fw_out = new firmware_t(ps_max, n);
    firmware_t& fw = *fw_out;
    std::mt19937 rng(5 + device);
    std::vector<phy::tx_descriptor_t> descs;
    const phy::tx_meta_t tx_meta{false, 1.0f, 0.0f, 0.0f, 5};
    int fails = 0;
    for (uint32_t i = 0; i < n; ++i) {
        firmware_t::sent_t s{};
        s.psdef = geo[i % 4]; s.network_id = NETS[(i / 4) % 2]; s.plcf_type = 1 + (i % 2);
        dnrp_packet_sizes_t ps{};
        if (dnrp_get_packet_sizes(&s.psdef, &ps) != DNRP_OK) return 100;
        auto* hp = fw.pool.get_process_tx(s.plcf_type, s.network_id, s.psdef, phy::harq::finalize_tx_t::reset_and_terminate);
        if (!hp) return 101;
        const uint32_t tag = 0x100u + 2 * i + (s.plcf_type - 1);   // even tags type 1, odd tags type 2
        uint8_t* a = hp->get_a_plcf();
        std::memset(a, 0, 10);
        a[0] = (uint8_t)((s.plcf_type == 1 ? 0x00 : 0x20) | 0x10 | ((s.psdef.PacketLength - 1) & 0x0f));   // format, length type, length
        a[1] = (uint8_t)(s.network_id & 0xff);
        if (s.plcf_type == 1) { a[2] = (uint8_t)(tag >> 8); a[3] = (uint8_t)tag; a[4] = (uint8_t)(0x50 | s.psdef.mcs_index); }
        else { a[3] = (uint8_t)(tag >> 8); a[4] = (uint8_t)tag; a[5] = (uint8_t)(0xa0 | s.psdef.mcs_index); }
        s.tb.resize(ps.N_TB_byte);
        for (auto& v : s.tb) v = (uint8_t)rng();
        std::memcpy(hp->get_a_tb(), s.tb.data(), ps.N_TB_byte);
        fw.sent[tag] = s;
        descs.emplace_back(*hp, 0, tx_meta, radio::buffer_tx_meta_t{});
    }
    const uint32_t tx_stride = bp.get_tx_stride(), rx_stride = bp.get_rx_stride();
    std::vector<cf_t> iq((size_t)n * tx_stride);
    std::vector<uint32_t> ns;
    bp.generate_tx_packets(descs, iq.data(), &ns);
    // windows of the maximum geometry: packet at a random offset, noise about 25 dB below
    std::vector<cf_t> win((size_t)n * rx_stride);
    std::normal_distribution<float> g(0.f, 0.04f);
    std::vector<int64_t> start(n);
    for (uint32_t i = 0; i < n; ++i) {
        cf_t* w = &win[(size_t)i * rx_stride];
        for (uint32_t q = 0; q < rx_stride; ++q) { w[q].re = g(rng); w[q].im = g(rng); }
        const uint32_t t0 = 120 + (uint32_t)(rng() % 200);
        for (uint32_t q = 0; q < ns[i] && t0 + q < rx_stride; ++q) { w[t0 + q].re += iq[(size_t)i * tx_stride + q].re; w[t0 + q].im += iq[(size_t)i * tx_stride + q].im; }
        start[i] = 1000000 * (int64_t)(i + 1);
    }
    bp.process_windows(n, win.data(), start.data(), fw);
    if (fw.n_work_pcc != n || fw.n_work_pdc != n || fw.n_work_pdc_error || fw.n_plcf_mismatch || fw.n_tb_mismatch) ++fails;
    // the same windows in the radio's wire format (int16 I/Q at -20 dBFS rms): half the bytes over the host link, the same
    // callbacks with the same transport blocks
    std::vector<int16_t> w16((size_t)n * rx_stride * 2);
    double pw = 0.0;
    for (size_t q = 0; q < win.size(); ++q) pw += (double)win[q].re * win[q].re + (double)win[q].im * win[q].im;
    const float gain = (float)(0.1 * 32768.0 / std::sqrt(pw / (double)win.size()));
    auto q16 = [&](float v) { const float r = std::nearbyint(v * gain); return (int16_t)(r > 32767.f ? 32767.f : (r < -32768.f ? -32768.f : r)); };
    for (size_t q = 0; q < win.size(); ++q) { w16[2 * q] = q16(win[q].re); w16[2 * q + 1] = q16(win[q].im); }
    bp.set_tracking(true, true);   // residual CFO / STO stage on in every plan's context: no impairment, no change
    bp.process_windows_sc16(n, w16.data(), 1.0f / gain, start.data(), fw);
    if (fw.n_work_pcc != 2 * n || fw.n_work_pdc != 2 * n || fw.n_work_pdc_error || fw.n_plcf_mismatch || fw.n_tb_mismatch) ++fails;
    return fails;
}

int main(int argc, char** argv) {
    const uint32_t n = argc > 1 ? (uint32_t)std::atoi(argv[1]) : 32;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        std::fprintf(stderr, "multi_plan_firmware: no CUDA device (this library has no CPU fallback)\n");
        return 3;
    }
    int fails = 0;
    const int used = ndev >= 2 ? 2 : 1;
    firmware_t* fw[2] = {nullptr, nullptr};
    std::unique_ptr<phy::batch_phy_t> bp[2];
    try {
        // contexts on both GPUs alive at the same time inside one process (per-device kernel attributes)
        for (int d = 0; d < used; ++d) fails += run(d, n, fw[d], bp[d]);
    } catch (const std::exception& e) {
        std::fprintf(stderr, "multi_plan_firmware: %s\n", e.what());
        return 2;
    }
    // the counters of all PHYs of this process in one NCCL all-reduce (one rank per GPU; a single GPU reduces over itself)
    try {
        std::vector<phy::batch_phy_t*> all;
        uint64_t crc_sum = 0, win_sum = 0;
        for (int d = 0; d < used; ++d) {
            all.push_back(bp[d].get());
            crc_sum += bp[d]->get_stats().pdc_crc_ok;
            win_sum += bp[d]->get_stats().windows;
        }
        const phy::batch_phy_t::stats_t tot = phy::batch_phy_t::reduce_stats(all);
        if (tot.pdc_crc_ok != crc_sum || tot.windows != win_sum || tot.pdc_crc_ok == 0) {
            std::fprintf(stderr, "multi_plan_firmware: reduced counters %llu / %llu differ from the host sums %llu / %llu\n",
                         (unsigned long long)tot.pdc_crc_ok, (unsigned long long)tot.windows, (unsigned long long)crc_sum,
                         (unsigned long long)win_sum);
            ++fails;
        }
    } catch (const std::exception& e) {
        std::fprintf(stderr, "multi_plan_firmware: counter reduction: %s\n", e.what());
        return 2;
    }
    std::printf("{\"n\": %u, \"gpus\": %d, \"work_pcc\": %u, \"work_pdc\": %u, \"work_pdc_error\": %u, \"plcf_mismatch\": %u, "
                "\"tb_mismatch\": %u, \"fails\": %d}\n", n, used, fw[0]->n_work_pcc, fw[0]->n_work_pdc, fw[0]->n_work_pdc_error,
                fw[0]->n_plcf_mismatch, fw[0]->n_tb_mismatch, fails);
    return fails ? 1 : 0;
}

// multi_beta_firmware.cpp -- test program for include/dnrp_host.hpp (multi_beta_phy_t): ONE receiver configured for
// beta = 4 receives a mix of packets of beta 4, 2 and 1 that all arrive at its sample rate.  The reference's
// synchronisation "is triggered by packets for b=8,4,2,1" of a radio device class with b = 8 and can "find the best
// fitting b in frequency domain" (sync_param.hpp:238-244, compiled out there); here the beta search of the packet
// synchronisation names every window's beta, the window goes to the PHY of that beta, and the firmware learns the beta
// from sync_report_t::b.  Prints one JSON line; exit code 0 = all checks passed, 3 = no usable GPU.
#include <cuda_runtime_api.h>

#include <cstdio>
#include <cstdlib>
#include <map>
#include <random>
#include <string>

#include "dnrp_host.hpp"

using namespace dectnrp_b200;

static const uint32_t NET = 123456789u;

struct firmware_t final : public upper::tpoint_t {
    firmware_t(const sp3::packet_sizes_t& ps_max, uint32_t n_proc) : pool(ps_max, n_proc, n_proc) {}

    phy::maclow_phy_t work_pcc(const phy::phy_maclow_t& in) override {
        ++n_work_pcc;
        const uint8_t* plcf = in.pcc_report.get_plcf_bytes(1);
        if (!plcf) return phy::maclow_phy_t();
        const uint32_t tag = ((uint32_t)plcf[2] << 8) | plcf[3];
        auto it = sent.find(tag);
        if (it == sent.end()) return phy::maclow_phy_t();
        // the geometry: mu and beta from the synchronisation, MCS from the PLCF
        sp3::packet_sizes_def_t psdef = it->second.psdef;
        psdef.u = in.sync_report.u;
        psdef.b = in.sync_report.b;
        psdef.mcs_index = (uint32_t)(plcf[4] & 0x07);
        if (std::memcmp(&psdef, &it->second.psdef, sizeof(psdef)) != 0) { ++n_mismatch; return phy::maclow_phy_t(); }
        auto* hp = pool.get_process_rx(1, NET, psdef, 0, phy::harq::finalize_rx_t::reset_and_terminate);
        if (!hp) return phy::maclow_phy_t();
        return phy::maclow_phy_t(hp, phy::maclow_phy_handle_t(phy::handle_pcc2pdc_t::th10, tag));
    }
    phy::machigh_phy_t work_pdc(const phy::phy_machigh_t& in) override {
        ++n_work_pdc;
        const sent_t& s = sent[in.maclow_phy.get_handle_lrdid()];
        if (in.pdc_report.N_TB_byte != s.tb.size() || std::memcmp(in.pdc_report.a_tb, s.tb.data(), s.tb.size()) != 0) ++n_mismatch;
        return phy::machigh_phy_t();
    }
    phy::machigh_phy_t work_pdc_error(const phy::phy_machigh_t&) override {
        ++n_work_pdc_error;
        return phy::machigh_phy_t();
    }
    struct sent_t { sp3::packet_sizes_def_t psdef; std::vector<uint8_t> tb; };
    phy::harq::process_pool_t pool;
    std::map<uint32_t, sent_t> sent;
    uint32_t n_work_pcc{0}, n_work_pdc{0}, n_work_pdc_error{0}, n_mismatch{0};
};

int main(int argc, char** argv) {
    const uint32_t n = argc > 1 ? (uint32_t)std::atoi(argv[1]) : 24;
    // "beta": a mu 1 / beta 4 receiver and packets of beta 4, 2, 1;  "mu": a mu 2 / beta 2 receiver and packets of
    // (mu, beta) = (2, 2), (2, 1), (1, 4), (1, 2) -- mu detection is future work in the reference (README.md:198)
    const bool with_mu = argc > 2 && std::string(argv[2]) == "mu";
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        std::fprintf(stderr, "multi_beta_firmware: no CUDA device (this library has no CPU fallback)\n");
        return 3;
    }
    int fails = 0;
    try {
        struct num_t { uint32_t u, b; };
        const std::vector<num_t> nums = with_mu ? std::vector<num_t>{{2, 2}, {2, 1}, {1, 4}, {1, 2}}
                                                : std::vector<num_t>{{1, 4}, {1, 2}, {1, 1}};
        const uint32_t NN = (uint32_t)nums.size(), rate = nums[0].u * nums[0].b;
        dnrp_config_t cfg{};
        // the receiver: windows for two slots (the other numerologies' search margins are longer than its own), 16-QAM 1/2
        cfg.psdef = {nums[0].u, nums[0].b, 1, 2, 0, 3, 2048};
        cfg.N_RX = 1; cfg.oversampling = 1; cfg.resample_L = 1; cfg.resample_M = 1;
        cfg.max_batch = n; cfg.network_id = NET; cfg.plcf_type = 1;
        cfg.chestim_mode_lr = 1; cfg.chestim_mode_lr_t_stride = 2;
        phy::multi_beta_phy_t rx(cfg, 0, n);
        for (uint32_t k = 1; k < NN; ++k) {
            const sp3::packet_sizes_def_t d{nums[k].u, nums[k].b, 1, 1, 0, 3, 2048};
            if (nums[k].u == nums[0].u) rx.add_beta(d); else rx.add_mu(d);
        }
        const uint32_t rx_stride = rx.get(nums[0].u, nums[0].b).get_rx_stride();
        firmware_t fw(rx.get(nums[0].u, nums[0].b).get_packet_sizes(), n);
        std::mt19937 rng(17);
        std::vector<cf_t> win((size_t)n * rx_stride);
        std::normal_distribution<float> g(0.f, 0.03f);
        for (auto& v : win) { v.re = g(rng); v.im = g(rng); }
        std::vector<int64_t> start(n);
        std::vector<uint32_t> beta_sent(n), mu_sent(n);
        // transmitters: one PHY per numerology at the common rate (mu x beta x oversampling constant)
        for (uint32_t k = 0; k < NN; ++k) {
            const uint32_t b = nums[k].b;
            dnrp_config_t tc = cfg;
            tc.psdef = {nums[k].u, b, 1, 1, 0, 3, 2048};   // packets of one slot
            tc.oversampling = rate / (nums[k].u * b);
            phy::batch_phy_t txp(tc, 0, n);
            std::vector<phy::tx_descriptor_t> descs;
            std::vector<uint32_t> where;
            const phy::tx_meta_t tx_meta{false, 1.0f, 0.0f, 0.0f, 5};
            for (uint32_t i = k; i < n; i += NN) {         // windows i = k, k + NN, ... carry this numerology
                firmware_t::sent_t s{};
                s.psdef = tc.psdef;
                s.psdef.mcs_index = 1 + (i % 3);
                dnrp_packet_sizes_t ps{};
                if (dnrp_get_packet_sizes(&s.psdef, &ps) != DNRP_OK) return 100;
                txp.add_plan(s.psdef);
                auto* hp = fw.pool.get_process_tx(1, NET, s.psdef, phy::harq::finalize_tx_t::reset_and_terminate);
                if (!hp) return 101;
                const uint32_t tag = 0x200u + i;
                uint8_t* a = hp->get_a_plcf();
                std::memset(a, 0, 10);
                a[0] = 0x10; a[1] = (uint8_t)(NET & 0xff); a[2] = (uint8_t)(tag >> 8); a[3] = (uint8_t)tag;
                a[4] = (uint8_t)(0x50 | s.psdef.mcs_index);
                s.tb.resize(ps.N_TB_byte);
                for (auto& v : s.tb) v = (uint8_t)rng();
                std::memcpy(hp->get_a_tb(), s.tb.data(), ps.N_TB_byte);
                fw.sent[tag] = s;
                descs.emplace_back(*hp, 0, tx_meta, radio::buffer_tx_meta_t{});
                where.push_back(i);
                beta_sent[i] = b;
                mu_sent[i] = nums[k].u;
            }
            const uint32_t tx_stride = txp.get_tx_stride();
            std::vector<cf_t> iq((size_t)descs.size() * tx_stride);
            std::vector<uint32_t> ns;
            txp.generate_tx_packets(descs, iq.data(), &ns);
            for (size_t z = 0; z < where.size(); ++z) {
                cf_t* w = &win[(size_t)where[z] * rx_stride];
                const uint32_t t0 = 160 + (uint32_t)(rng() % 200);
                for (uint32_t q = 0; q < ns[z] && t0 + q < rx_stride; ++q) { w[t0 + q].re += iq[z * tx_stride + q].re; w[t0 + q].im += iq[z * tx_stride + q].im; }
                start[where[z]] = 1000000 * (int64_t)(where[z] + 1);
            }
        }
        // add the other geometries (three MCS each) to the receiving PHYs
        for (uint32_t k = 0; k < NN; ++k)
            for (uint32_t mcs = 1; mcs <= 3; ++mcs) rx.get(nums[k].u, nums[k].b).add_plan({nums[k].u, nums[k].b, 1, 1, 0, mcs, 2048});
        rx.process_windows(n, win.data(), start.data(), fw);
        uint32_t beta_wrong = 0, mu_wrong = 0;
        for (uint32_t i = 0; i < n; ++i) beta_wrong += rx.get_last_beta()[i] != beta_sent[i];
        for (uint32_t i = 0; i < n; ++i) mu_wrong += rx.get_last_mu()[i] != mu_sent[i];
        if (beta_wrong || mu_wrong || fw.n_work_pdc != n || fw.n_work_pdc_error || fw.n_mismatch) ++fails;
        std::printf("{\"n\": %u, \"beta_wrong\": %u, \"mu_wrong\": %u, \"work_pcc\": %u, \"work_pdc\": %u, \"work_pdc_error\": %u, \"mismatch\": %u, "
                    "\"fails\": %d}\n", n, beta_wrong, mu_wrong, fw.n_work_pcc, fw.n_work_pdc, fw.n_work_pdc_error, fw.n_mismatch, fails);
    } catch (const std::exception& e) {
        std::fprintf(stderr, "multi_beta_firmware: %s\n", e.what());
        return 2;
    }
    return fails ? 1 : 0;
}

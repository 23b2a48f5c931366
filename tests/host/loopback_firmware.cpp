// loopback_firmware.cpp -- test program for include/dnrp_host.hpp: a firmware in the style of the reference's
// tfw_loopback (lib/src/upper/tpoint_firmware/loopback/tfw_loopback.cpp) written against the mirrored tpoint_t
// callbacks.  Scenario 1: n packets at a good SNR -> every work_pcc sees its PLCF, every work_pdc its transport block.
// Scenario 2 (HARQ): first transmission at a poor SNR fails (work_pdc_error, process keeps running), the
// retransmission with the next redundancy version is combined in the same soft buffer slot and decodes.
// Prints one JSON line; exit code 0 = all checks passed, 3 = no usable GPU (the library has no CPU fallback).
#include <cstdio>
#include <cstdlib>
#include <map>
#include <random>

#include "dnrp_host.hpp"

using namespace dectnrp_b200;

struct firmware_t final : public upper::tpoint_t {
    firmware_t(const sp3::packet_sizes_t& ps, uint32_t n_proc) : packet_sizes(ps), pool(ps, n_proc, n_proc) {}

    phy::maclow_phy_t work_pcc(const phy::phy_maclow_t& in) override {
        ++n_work_pcc;
        order.push_back(in.sync_report.fine_peak_time_64);
        const uint8_t* plcf = in.pcc_report.get_plcf_bytes(1);
        if (!plcf) return phy::maclow_phy_t();
        // field decoding + radio device class check like rx_synced.cpp:1011-1026, then tfw_loopback-style reading
        if (in.pcc_report.decode_plcf(plcf_decoder) != 1) { ++n_plcf_mismatch; return phy::maclow_phy_t(); }
        const auto* p10 = static_cast<const sp4::plcf_10_t*>(plcf_decoder.get_plcf_base(1));
        const uint32_t key = p10->TransmitterIdentity;             // the firmware's packet tag
        auto it = sent.find(key);
        if (it == sent.end()) return phy::maclow_phy_t();
        if (std::memcmp(plcf, it->second.plcf, 5) != 0 || p10->ShortNetworkID != (network_id & 0xff) ||
            p10->get_PacketLength() != packet_sizes.psdef.PacketLength || p10->DFMCS != packet_sizes.psdef.mcs_index ||
            p10->get_TransmitPower_dBm() != 10 || p10->get_N_SS() != 1)
            ++n_plcf_mismatch;
        phy::harq::process_rx_t* hp;
        auto run = running_rx.find(key);
        if (run != running_rx.end()) {
            hp = pool.get_process_rx_running(run->second, it->second.rv, frx);
        } else {
            hp = pool.get_process_rx(1, network_id, packet_sizes.psdef, it->second.rv, frx);
            if (hp) running_rx[key] = hp->get_id();
        }
        if (!hp) return phy::maclow_phy_t();
        return phy::maclow_phy_t(hp, phy::maclow_phy_handle_t(phy::handle_pcc2pdc_t::th10, key));
    }
    phy::machigh_phy_t work_pdc(const phy::phy_machigh_t& in) override {
        ++n_work_pdc;
        const uint32_t key = in.maclow_phy.get_handle_lrdid();
        if (std::memcmp(in.pdc_report.a_tb, sent[key].tb.data(), packet_sizes.N_TB_byte) != 0) ++n_tb_mismatch;
        // MAC PDU framing: scenario 1 sends DATA MAC PDUs with one user plane data SDU (tag 1 << 8 | i + 1)
        if ((key >> 8) == 1) {
            if (!in.pdc_report.decode_mac_pdu(mac_pdu_decoder)) ++n_mac_mismatch;
            const auto& v = mac_pdu_decoder.get_mmie_decoded_vec();
            const auto* mch = mac_pdu_decoder.get_mac_common_header();
            if (v.size() != 1 || !v[0].is_user_plane_data() || mch == nullptr || mch->Sequence_number != (key & 0xff) ||
                v[0].N_bytes != packet_sizes.N_TB_byte - 10 || std::memcmp(v[0].data_ptr, sent[key].tb.data() + 6, v[0].N_bytes) != 0)
                ++n_mac_mismatch;
        }
        snr_sum += in.pdc_report.snr_dB;
        running_rx.erase(key);
        return phy::machigh_phy_t();
    }
    phy::machigh_phy_t work_pdc_error(const phy::phy_machigh_t& in) override {
        ++n_work_pdc_error;
        (void)in;
        return phy::machigh_phy_t();
    }

    struct sent_t { uint8_t plcf[10]; std::vector<uint8_t> tb; uint32_t rv; };
    /// PLCF type 1 of a packet like tfw_loopback.cpp:253-289; tag = TransmitterIdentity (bytes 2, 3)
    void make_plcf(uint8_t* a, uint32_t tag) const {
        sp4::plcf_10_t p;
        p.HeaderFormat = 0;
        p.PacketLengthType = packet_sizes.psdef.PacketLengthType;
        p.set_PacketLength_m1((int32_t)packet_sizes.psdef.PacketLength);
        p.ShortNetworkID = network_id & 0xff;
        p.TransmitterIdentity = tag;
        p.set_TransmitPower(10);
        p.DFMCS = packet_sizes.psdef.mcs_index;
        std::memset(a, 0, 10);
        p.pack(a);
    }
    static uint32_t key_of(const uint8_t* a) { return ((uint32_t)a[2] << 8) | a[3]; }
    sp4::plcf_decoder_t plcf_decoder{16, 11, 8};
    sp4::mac_pdu_decoder_t mac_pdu_decoder;
    uint32_t n_mac_mismatch{0};
    sp3::packet_sizes_t packet_sizes;
    phy::harq::process_pool_t pool;
    phy::harq::finalize_rx_t frx{phy::harq::finalize_rx_t::keep_running_or_reset_and_terminate_if_crc_correct};
    uint32_t network_id{123456789};
    std::map<uint32_t, sent_t> sent;
    std::map<uint32_t, uint32_t> running_rx;
    std::vector<int64_t> order;
    uint32_t n_work_pcc{0}, n_work_pdc{0}, n_work_pdc_error{0}, n_plcf_mismatch{0}, n_tb_mismatch{0};
    double snr_sum{0};
};

int main(int argc, char** argv) {
    const uint32_t n = argc > 1 ? (uint32_t)std::atoi(argv[1]) : 32;
    dnrp_config_t cfg{};
    cfg.psdef = {1, 4, 1, 1, 0, 4, 2048};   // mu 1, beta 4, one slot, SISO, 16-QAM 3/4
    cfg.N_RX = 2; cfg.oversampling = 1; cfg.resample_L = 10; cfg.resample_M = 9;
    cfg.max_batch = n; cfg.network_id = 123456789; cfg.plcf_type = 1;
    cfg.chestim_mode_lr = 1; cfg.chestim_mode_lr_t_stride = 2;
    std::unique_ptr<phy::batch_phy_t> bp;
    try {
        bp = std::make_unique<phy::batch_phy_t>(cfg, 0, n);
    } catch (const std::exception& e) {
        std::fprintf(stderr, "loopback_firmware: %s\n", e.what());
        return 3;
    }
    // optional second argument: prefix of the per-packet JSON export (the reference's worker_pool_* files), 8 entries per file
    if (argc > 2) bp->enable_json_export(8, argv[2]);
    const sp3::packet_sizes_t ps = bp->get_packet_sizes();
    firmware_t fw(ps, n);
    std::mt19937 rng(7);
    int fails = 0;
    const phy::tx_meta_t tx_meta{false, 1.0f, 0.0f, 0.0f, 5};

    auto transmit = [&](std::vector<phy::harq::process_tx_t*>& procs, float snr_dB, uint64_t seed) {
        std::vector<phy::tx_descriptor_t> descs;
        for (auto* hp : procs) descs.emplace_back(*hp, 0, tx_meta, radio::buffer_tx_meta_t{});
        for (auto* hp : procs) fw.sent[firmware_t::key_of(hp->get_a_plcf())].rv = hp->get_rv();
        bp->generate_tx_packets(descs, nullptr);
        std::vector<dnrp_channel_desc_t> ch(procs.size());
        std::vector<int64_t> start(procs.size());
        for (size_t i = 0; i < ch.size(); ++i) {
            ch[i] = dnrp_channel_desc_t{};
            ch[i].snr_dB = snr_dB; ch[i].net_bandwidth_norm = 0.9f; ch[i].t0 = 150 + (uint32_t)(rng() % 300);
            ch[i].mode = 0; ch[i].seed = seed;
            start[i] = 1000000 * (int64_t)(i + 1);
        }
        bp->channel(ch);
        auto back = bp->process_windows((uint32_t)procs.size(), nullptr, start.data(), fw);
        return back.size();
    };

    // ---- scenario 1: n fresh packets, good SNR
    std::vector<phy::harq::process_tx_t*> procs;
    for (uint32_t i = 0; i < n; ++i) {
        auto* hp = fw.pool.get_process_tx(1, fw.network_id, ps.psdef, phy::harq::finalize_tx_t::reset_and_terminate);
        if (!hp) { ++fails; break; }
        firmware_t::sent_t s{};
        fw.make_plcf(s.plcf, (1u << 8) | (i + 1));
        s.tb.resize(ps.N_TB_byte);
        {   // DATA MAC PDU: header type (1) + common header (2) + multiplexing header (3) + SDU + padding IE (4)
            std::vector<uint8_t> payload(ps.N_TB_byte - 10);
            for (auto& v : payload) v = (uint8_t)rng();
            sp4::mac_header_type_t mht;
            mht.Version = sp4::mac_header_type_t::version_ec::v00;
            mht.MAC_security = sp4::mac_header_type_t::mac_security_ec::macsec_not_used;
            mht.MAC_header_type = sp4::mac_header_type_t::mac_header_type_ec::DATA_MAC_PDU;
            sp4::mac_common_header_t mch;
            mch.kind = mht.MAC_header_type; mch.Sequence_number = i + 1;
            sp4::mac_pdu_writer_t w(s.tb.data(), ps.N_TB_byte);
            if (!w.begin(mht, mch) || !w.add_sdu(sp4::mac_multiplexing_header_t::User_Plane_Data_Flow_1, payload.data(),
                                                 (uint32_t)payload.size()))
                ++fails;
            w.finish();
        }
        std::memcpy(hp->get_a_plcf(), s.plcf, 10);
        std::memcpy(hp->get_a_tb(), s.tb.data(), ps.N_TB_byte);
        fw.sent[firmware_t::key_of(s.plcf)] = s;
        procs.push_back(hp);
    }
    transmit(procs, 28.0f, 11);
    const uint32_t pcc1 = fw.n_work_pcc, pdc1 = fw.n_work_pdc, err1 = fw.n_work_pdc_error;
    if (pcc1 != n || pdc1 != n || err1 != 0 || fw.n_plcf_mismatch || fw.n_tb_mismatch || fw.n_mac_mismatch) ++fails;
    for (size_t i = 1; i < fw.order.size(); ++i) if (fw.order[i] <= fw.order[i - 1]) ++fails;   // work_pcc FIFO
    for (auto* hp : procs) if (hp->is_running()) ++fails;                                        // reset_and_terminate

    // ---- scenario 2: HARQ, poor SNR first, retransmission with the next rv
    procs.clear();
    const uint32_t m = n < 8 ? n : 8;
    for (uint32_t i = 0; i < m; ++i) {
        auto* hp = fw.pool.get_process_tx(1, fw.network_id, ps.psdef, phy::harq::finalize_tx_t::increase_rv_and_keep_running);
        firmware_t::sent_t s{};
        fw.make_plcf(s.plcf, (2u << 8) | (i + 1));
        s.tb.resize(ps.N_TB_byte);
        for (auto& v : s.tb) v = (uint8_t)rng();
        std::memcpy(hp->get_a_plcf(), s.plcf, 10);
        std::memcpy(hp->get_a_tb(), s.tb.data(), ps.N_TB_byte);
        fw.sent[firmware_t::key_of(s.plcf)] = s;
        procs.push_back(hp);
    }
    uint32_t rounds = 0;
    const uint32_t pdc_before = fw.n_work_pdc;
    while (fw.n_work_pdc - pdc_before < m && rounds < 4) {
        std::vector<phy::harq::process_tx_t*> todo;
        for (auto* hp : procs)
            if (fw.running_rx.count(firmware_t::key_of(hp->get_a_plcf())) || rounds == 0) todo.push_back(hp);
        if (todo.empty()) break;
        transmit(todo, 5.0f, 100 + rounds);   // 16-QAM 3/4 on two antennas does not decode at 5 dB in one shot
        ++rounds;
    }
    const uint32_t err2 = fw.n_work_pdc_error - err1, pdc2 = fw.n_work_pdc - pdc_before;
    if (pdc2 != m || err2 == 0 || rounds < 2 || fw.n_tb_mismatch) ++fails;

    // ---- scenario 3: streaming front door -- the packets sit at random gaps in ONE continuous two-antenna stream
    // (context without resampling); process_stream finds them, hands them to the synced receiver and the callbacks
    uint32_t stream_found = 0;
    {
        dnrp_config_t cfg3 = cfg;
        cfg3.resample_L = 1; cfg3.resample_M = 1;
        phy::batch_phy_t bp3(cfg3, 0, n);
        const uint32_t k = n < 12 ? n : 12, S = ps.N_samples_packet_no_GI, stride3 = bp3.get_tx_stride();
        std::vector<phy::harq::process_tx_t*> p3;
        for (uint32_t i = 0; i < k; ++i) {
            auto* hp = fw.pool.get_process_tx(1, fw.network_id, ps.psdef, phy::harq::finalize_tx_t::reset_and_terminate);
            if (!hp) { ++fails; break; }
            firmware_t::sent_t s{};
            fw.make_plcf(s.plcf, (3u << 8) | (i + 1));
            s.tb.resize(ps.N_TB_byte);
            for (auto& v : s.tb) v = (uint8_t)rng();
            std::memcpy(hp->get_a_plcf(), s.plcf, 10);
            std::memcpy(hp->get_a_tb(), s.tb.data(), ps.N_TB_byte);
            s.rv = hp->get_rv();
            fw.sent[firmware_t::key_of(s.plcf)] = s;
            p3.push_back(hp);
        }
        std::vector<phy::tx_descriptor_t> descs;
        for (auto* hp : p3) descs.emplace_back(*hp, 0, tx_meta, radio::buffer_tx_meta_t{});
        std::vector<cf_t> iq((size_t)p3.size() * stride3);
        bp3.generate_tx_packets(descs, iq.data());
        std::vector<size_t> at(p3.size());
        size_t t = 3 * ps.N_samples_STF;
        for (size_t i = 0; i < p3.size(); ++i) { at[i] = t; t += S + ps.N_samples_GI + (i % 3 ? rng() % (2 * S) : 0); }
        const size_t ns = t + 4 * ps.N_samples_STF;
        std::vector<cf_t> stream(2 * ns);
        std::normal_distribution<float> g(0.f, 0.05f);   // about 23 dB below the packets
        for (auto& v : stream) { v.re = g(rng); v.im = g(rng); }
        for (size_t i = 0; i < p3.size(); ++i)
            for (uint32_t a = 0; a < 2; ++a)
                for (uint32_t q = 0; q < S; ++q) {
                    const cf_t x = iq[i * stride3 + q];
                    cf_t& y = stream[a * ns + at[i] + q];
                    y.re += a ? -x.im : x.re; y.im += a ? x.re : x.im;   // antenna 1: rotated by 90 degrees
                }
        const uint32_t pcc0 = fw.n_work_pcc, pdc0 = fw.n_work_pdc;
        bp3.process_stream(ns, stream.data(), 5000000, fw, 2 * S);
        stream_found = fw.n_work_pdc - pdc0;
        if (fw.n_work_pcc - pcc0 != p3.size() || stream_found != p3.size() || fw.n_tb_mismatch || fw.n_plcf_mismatch) ++fails;
        for (size_t i = 0; i < p3.size(); ++i)   // sync_report.fine_peak_time_64 = global time of the packet start
            if (fw.order[fw.order.size() - p3.size() + i] != (int64_t)(5000000 + at[i])) { ++fails; break; }
    }

    // ---- scenario 4: the same for a radio that delivers its samples piece by piece, at 10/9 of the DECT rate (the main
    // context): open_stream / push_stream keep the overlap, the resampler history and unfinished packets between pushes
    uint32_t push_found = 0;
    {
        const uint32_t k = n < 12 ? n : 12, stride4 = bp->get_tx_stride();
        std::vector<phy::harq::process_tx_t*> p4;
        for (uint32_t i = 0; i < k; ++i) {
            auto* hp = fw.pool.get_process_tx(1, fw.network_id, ps.psdef, phy::harq::finalize_tx_t::reset_and_terminate);
            if (!hp) { ++fails; break; }
            firmware_t::sent_t s{};
            fw.make_plcf(s.plcf, (4u << 8) | (i + 1));
            s.tb.resize(ps.N_TB_byte);
            for (auto& v : s.tb) v = (uint8_t)rng();
            std::memcpy(hp->get_a_plcf(), s.plcf, 10);
            std::memcpy(hp->get_a_tb(), s.tb.data(), ps.N_TB_byte);
            s.rv = hp->get_rv();
            fw.sent[firmware_t::key_of(s.plcf)] = s;
            p4.push_back(hp);
        }
        std::vector<phy::tx_descriptor_t> descs;
        for (auto* hp : p4) descs.emplace_back(*hp, 0, tx_meta, radio::buffer_tx_meta_t{});
        std::vector<cf_t> iq((size_t)p4.size() * stride4);
        std::vector<uint32_t> ns4;
        bp->generate_tx_packets(descs, iq.data(), &ns4);
        const uint32_t S = ns4.empty() ? 0 : ns4[0], stf_hw = ps.N_samples_STF * 10 / 9 + 1;
        std::vector<size_t> at(p4.size());
        size_t t = 3 * stf_hw;
        for (size_t i = 0; i < p4.size(); ++i) { at[i] = t; t += S + stf_hw / 2 + (i % 3 ? rng() % (2 * S) : 0); }
        const size_t total = t + 6 * stf_hw;
        std::vector<cf_t> stream(2 * total);
        std::normal_distribution<float> g(0.f, 0.05f);
        for (auto& v : stream) { v.re = g(rng); v.im = g(rng); }
        for (size_t i = 0; i < p4.size(); ++i)
            for (uint32_t a = 0; a < 2; ++a)
                for (uint32_t q = 0; q < S; ++q) {
                    const cf_t x = iq[i * stride4 + q];
                    cf_t& y = stream[a * total + at[i] + q];
                    y.re += a ? -x.im : x.re; y.im += a ? x.re : x.im;
                }
        const uint32_t pcc0 = fw.n_work_pcc, pdc0 = fw.n_work_pdc;
        const uint32_t max_push = 3 * S;
        bp->open_stream(max_push, 7000000);
        std::vector<cf_t> piece;
        for (size_t pos = 0; pos < total;) {   // pushes of uneven length, cuts wherever they fall
            const size_t len = std::min<size_t>(total - pos, 1 + rng() % max_push);
            piece.resize(2 * len);
            for (uint32_t a = 0; a < 2; ++a) std::memcpy(&piece[a * len], &stream[a * total + pos], len * sizeof(cf_t));
            bp->push_stream((uint32_t)len, piece.data(), fw);
            pos += len;
        }
        bp->close_stream();
        push_found = fw.n_work_pdc - pdc0;
        if (fw.n_work_pcc - pcc0 != p4.size() || push_found != p4.size() || fw.n_tb_mismatch || fw.n_plcf_mismatch) ++fails;
        for (size_t i = 0; i < p4.size() && fw.order.size() >= p4.size(); ++i)
            if (fw.order[fw.order.size() - p4.size() + i] != (int64_t)(7000000 + at[i])) { ++fails; break; }
        // the same stream in the radio's wire format (int16 I/Q, push_stream_sc16): the same packets at the same times
        const uint32_t pdc1 = fw.n_work_pdc;
        const float gain16 = 0.1f * 32768.0f / 0.5f;     // the packets have an rms of about 0.5 here
        bp->open_stream(max_push, 9000000);
        std::vector<int16_t> piece16;
        for (size_t pos = 0; pos < total;) {
            const size_t len = std::min<size_t>(total - pos, 1 + rng() % max_push);
            piece16.resize(4 * len);
            for (uint32_t a = 0; a < 2; ++a)
                for (size_t q = 0; q < len; ++q) {
                    const cf_t v = stream[a * total + pos + q];
                    auto q16 = [&](float x) { const float r = std::nearbyint(x * gain16); return (int16_t)(r > 32767.f ? 32767.f : (r < -32768.f ? -32768.f : r)); };
                    piece16[2 * (a * len + q)] = q16(v.re);
                    piece16[2 * (a * len + q) + 1] = q16(v.im);
                }
            bp->push_stream_sc16((uint32_t)len, piece16.data(), 1.0f / gain16, fw);
            pos += len;
        }
        bp->close_stream();
        if (fw.n_work_pdc - pdc1 != p4.size() || fw.n_tb_mismatch || fw.n_plcf_mismatch) ++fails;
        for (size_t i = 0; i < p4.size() && fw.order.size() >= p4.size(); ++i)
            if (fw.order[fw.order.size() - p4.size() + i] != (int64_t)(9000000 + at[i])) { ++fails; break; }
    }

    const size_t json_files = bp->flush_json_export().size();
    const auto& st = bp->get_stats();
    std::printf("{\"n\": %u, \"work_pcc\": %u, \"work_pdc\": %u, \"work_pdc_error\": %u, \"plcf_mismatch\": %u, "
                "\"tb_mismatch\": %u, \"harq_rounds\": %u, \"harq_first_round_errors\": %u, \"harq_decoded\": %u, "
                "\"windows\": %llu, \"detections\": %llu, \"mean_snr_dB\": %.2f, \"stream_decoded\": %u, \"push_decoded\": %u, \"json_files\": %zu, \"fails\": %d}\n",
                n, fw.n_work_pcc, fw.n_work_pdc, fw.n_work_pdc_error, fw.n_plcf_mismatch, fw.n_tb_mismatch, rounds, err2,
                pdc2, (unsigned long long)st.windows, (unsigned long long)st.detections,
                fw.n_work_pdc ? fw.snr_sum / fw.n_work_pdc : 0.0, stream_found, push_found, json_files, fails);
    return fails ? 1 : 0;
}

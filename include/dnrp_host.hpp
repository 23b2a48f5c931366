// dnrp_host.hpp -- C++ host side above the C ABI (dnrp_c_api.h): the reference's PHY-to-MAC surface for the PCC/PDC
// path with the same type names, members and call order, so that firmware written against the reference's
// tpoint_t::work_pcc / work_pdc callbacks runs on batches processed by libdnrp_cuda.so.  Header only, C++17.
//
// Mirrored reference interfaces (paths under /root/reference/lib/include/dectnrp):
//   phy/tx/tx_meta.hpp:32-58, phy/tx/tx_descriptor.hpp:43-78, radio/buffer_tx_meta.hpp:32-50
//   phy/harq/{process,process_tx,process_rx,process_pool,finalize}.hpp
//   phy/rx/sync/sync_report.hpp:31-98, phy/rx/rx_synced/{pcc_report,pdc_report}.hpp, mimo/mimo_report.hpp
//   phy/interfaces/{phy_maclow,maclow_phy,maclow_phy_handle,phy_machigh,machigh_phy}.hpp
//   upper/tpoint.hpp:128-170 (work_pcc, work_pdc, work_pdc_error)
//   phy/pool/worker_tx_rx.cpp:117-240 (order of calls: work_pcc FIFO by job order, PDC only if continue_with_pdc,
//   hp_rx->finalize(crc_status) after the callback, TX descriptors generated afterwards)
//   sections_part4/physical_header_field/*.hpp (plcf_10/20/21_t, feedback info, plcf_decoder_t): dnrp_plcf.hpp;
//   pcc_report_t::decode_plcf() runs rx_synced.cpp:1011-1026 on the bytes the device delivered
//   sections_part4/mac_pdu/*.hpp (MAC header type, common headers, multiplexing header, mac_pdu_decoder_t framing):
//   dnrp_mac_pdu.hpp; pdc_report_t::decode_mac_pdu() demultiplexes the SDUs of a transport block
// Not mirrored (MAC layer): the parsers of the individual MAC messages / IEs (sections_part4/mac_messages_and_ie).
// There is no CPU fallback: construction throws if libdnrp_cuda.so cannot create its context.
#pragma once
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <optional>
#include <stdexcept>
#include <string>
#include <vector>

extern "C" {
#include "dnrp_c_api.h"
}
#include "dnrp_json_export.hpp"
#include "dnrp_plcf.hpp"
#include "dnrp_mac_pdu.hpp"

namespace dectnrp_b200 {

using cf_t = dnrp_cf32_t;

namespace sp3 {
using packet_sizes_def_t = dnrp_psdef_t;
using packet_sizes_t = dnrp_packet_sizes_t;
}  // namespace sp3

namespace common {
/// common::ant_t = std::array<float, 8> + number of antennas
struct ant_t {
    std::array<float, 8> ary{};
    uint32_t nof_antennas{0};
    float at(uint32_t i) const { return ary.at(i); }
    float& at(uint32_t i) { return ary.at(i); }
    uint32_t get_nof_antennas() const { return nof_antennas; }
    bool has_any_larger(float v) const {
        for (uint32_t i = 0; i < nof_antennas; ++i) if (ary[i] > v) return true;
        return false;
    }
    uint32_t get_index_of_max() const {
        uint32_t m = 0;
        for (uint32_t i = 1; i < nof_antennas; ++i) if (ary[i] > ary[m]) m = i;
        return m;
    }
};
}  // namespace common

namespace radio {
struct buffer_tx_meta_t {
    int64_t tx_order_id{0};
    int64_t tx_order_id_expect_next{0};
    int64_t tx_time_64{0};
    uint32_t busy_wait_us{0};
};
}  // namespace radio

namespace phy {

namespace harq {

enum class finalize_tx_t { reset_and_terminate, increase_rv_and_keep_running, keep_rv_and_keep_running };
enum class finalize_rx_t { reset_and_terminate, keep_running, keep_running_or_reset_and_terminate_if_crc_correct };

class process_pool_t;

class process_t {
    public:
        explicit process_t(uint32_t id_) : id(id_) {}
        virtual ~process_t() = default;
        uint32_t get_id() const { return id; }
        uint32_t get_PLCF_type() const { return PLCF_type; }
        uint32_t get_network_id() const { return network_id; }
        const sp3::packet_sizes_def_t& get_psdef() const { return psdef; }
        uint32_t get_rv() const { return rv; }
        bool is_running() const { return running; }

    protected:
        friend class process_pool_t;
        const uint32_t id;
        uint32_t PLCF_type{0}, network_id{0}, rv{0};
        sp3::packet_sizes_def_t psdef{};
        bool running{false};
        void reset() { PLCF_type = 0; network_id = 0; rv = 0; running = false; }
};

class process_tx_t final : public process_t {
    public:
        process_tx_t(uint32_t id_, uint32_t N_TB_byte_max) : process_t(id_), a_tb(N_TB_byte_max, 0) {}
        uint8_t* get_a_plcf() { return a_plcf.data(); }
        uint8_t* get_a_tb() { return a_tb.data(); }
        /// process_tx.cpp:43-62
        void finalize() {
            switch (finalize_tx) {
                case finalize_tx_t::reset_and_terminate: reset_and_terminate_func(); break;
                case finalize_tx_t::increase_rv_and_keep_running: set_next_rv(); break;
                case finalize_tx_t::keep_rv_and_keep_running: break;
            }
        }

    private:
        friend class process_pool_t;
        std::array<uint8_t, 10> a_plcf{};
        std::vector<uint8_t> a_tb;
        uint32_t rv_unwrapped{0};
        finalize_tx_t finalize_tx{finalize_tx_t::reset_and_terminate};
        /// process_tx.cpp:64-88: 0 -> 2 -> 3 -> 1 -> 0
        void set_next_rv() {
            static const uint32_t next[4] = {2, 0, 3, 1};
            rv = next[rv & 3];
            ++rv_unwrapped;
        }
        void reset_and_terminate_func() {
            rv_unwrapped = 0;
            finalize_tx = finalize_tx_t::reset_and_terminate;
            reset();
        }
};

class process_rx_t final : public process_t {
    public:
        process_rx_t(uint32_t id_, uint32_t N_TB_byte_max) : process_t(id_), a_tb(N_TB_byte_max, 0) {}
        /// decoded transport block (reference: get_hb_tb()->get_a())
        uint8_t* get_a_tb() { return a_tb.data(); }
        const uint8_t* get_a_tb() const { return a_tb.data(); }
        /// true until the first PDC of this process was decoded: the soft buffer slot has to be cleared
        bool is_first_transmission() const { return nof_rx == 0; }
        /// process_rx.cpp:45-66
        void finalize(bool crc_status) {
            ++nof_rx;
            switch (finalize_rx) {
                case finalize_rx_t::reset_and_terminate: reset_and_terminate_func(); break;
                case finalize_rx_t::keep_running: break;
                case finalize_rx_t::keep_running_or_reset_and_terminate_if_crc_correct:
                    if (crc_status) reset_and_terminate_func();
                    break;
            }
        }

    private:
        friend class process_pool_t;
        std::vector<uint8_t> a_tb;
        uint32_t nof_rx{0};
        finalize_rx_t finalize_rx{finalize_rx_t::reset_and_terminate};
        void reset_and_terminate_func() {
            nof_rx = 0;
            finalize_rx = finalize_rx_t::reset_and_terminate;
            reset();
        }
};

/// process_pool.hpp:36-70.  A receive process id doubles as the soft buffer slot on the device.
class process_pool_t {
    public:
        process_pool_t(const sp3::packet_sizes_t& maximum_packet_sizes, uint32_t nof_process_tx, uint32_t nof_process_rx) {
            for (uint32_t i = 0; i < nof_process_tx; ++i)
                hp_tx_vec.push_back(std::make_unique<process_tx_t>(i, maximum_packet_sizes.N_TB_byte));
            for (uint32_t i = 0; i < nof_process_rx; ++i)
                hp_rx_vec.push_back(std::make_unique<process_rx_t>(i, maximum_packet_sizes.N_TB_byte));
        }
        /// nullptr if every process is running (the reference returns nullptr as well, process_pool.cpp:60-70)
        process_tx_t* get_process_tx(uint32_t PLCF_type, uint32_t network_id, const sp3::packet_sizes_def_t psdef,
                                     finalize_tx_t ftx) const {
            for (auto& p : hp_tx_vec)
                if (!p->running) {
                    p->running = true; p->PLCF_type = PLCF_type; p->network_id = network_id; p->psdef = psdef;
                    p->rv = 0; p->finalize_tx = ftx;
                    return p.get();
                }
            return nullptr;
        }
        process_rx_t* get_process_rx(uint32_t PLCF_type, uint32_t network_id, const sp3::packet_sizes_def_t psdef,
                                     uint32_t rv, finalize_rx_t frx) const {
            for (auto& p : hp_rx_vec)
                if (!p->running) {
                    p->running = true; p->PLCF_type = PLCF_type; p->network_id = network_id; p->psdef = psdef;
                    p->rv = rv; p->finalize_rx = frx; p->nof_rx = 0;
                    return p.get();
                }
            return nullptr;
        }
        process_tx_t* get_process_tx_running(uint32_t id, finalize_tx_t ftx) const {
            if (id >= hp_tx_vec.size() || !hp_tx_vec[id]->running) return nullptr;
            hp_tx_vec[id]->finalize_tx = ftx;
            return hp_tx_vec[id].get();
        }
        process_rx_t* get_process_rx_running(uint32_t id, uint32_t rv, finalize_rx_t frx) const {
            if (id >= hp_rx_vec.size() || !hp_rx_vec[id]->running) return nullptr;
            hp_rx_vec[id]->rv = rv; hp_rx_vec[id]->finalize_rx = frx;
            return hp_rx_vec[id].get();
        }

    private:
        std::vector<std::unique_ptr<process_tx_t>> hp_tx_vec;
        std::vector<std::unique_ptr<process_rx_t>> hp_rx_vec;
};

}  // namespace harq

/// tx_meta.hpp:32-58
struct tx_meta_t {
    bool optimal_scaling_DAC{false};
    float DAC_scale{1.0f};
    float iq_phase_rad{0.0f};
    float iq_phase_increment_s2s_post_resampling_rad{0.0f};
    uint32_t GI_percentage{5};
};

/// tx_descriptor.hpp:43-78
class tx_descriptor_t {
    public:
        tx_descriptor_t(harq::process_tx_t& hp_tx_, uint32_t codebook_index_, tx_meta_t tx_meta_,
                        radio::buffer_tx_meta_t buffer_tx_meta_, int32_t hw_id_ = -1)
            : hp_tx(hp_tx_), codebook_index(codebook_index_), tx_meta(tx_meta_), buffer_tx_meta(buffer_tx_meta_),
              hw_id(hw_id_) {}
        harq::process_tx_t& hp_tx;
        uint32_t codebook_index;
        tx_meta_t tx_meta;
        radio::buffer_tx_meta_t buffer_tx_meta;
        int32_t hw_id;
};

/// sync_report.hpp:31-98
struct sync_report_t {
    explicit sync_report_t(uint32_t nof_antennas = 0) {
        coarse_peak_array.nof_antennas = nof_antennas;
        rms_array.nof_antennas = nof_antennas;
    }
    uint32_t detection_ant_idx{0};
    float detection_rms{0.f}, detection_metric{0.f};
    uint32_t detection_time_local{0};
    uint32_t u{0}, b{0};
    common::ant_t coarse_peak_array{};
    uint32_t coarse_peak_time_local{0};
    int64_t coarse_peak_time_64{0};
    common::ant_t rms_array{};
    float cfo_fractional_rad{0.f}, cfo_integer_rad{0.f};
    uint32_t N_eff_TX{0};
    uint32_t fine_peak_time_local{0};
    int64_t fine_peak_time_64{0};
    float sto_fractional{0.f};
    int64_t fine_peak_time_corrected_by_sto_fractional_64{0};
};

/// mimo/mimo_report.hpp:27-59
struct mimo_report_t {
    uint32_t N_RX{0}, N_TS_other{0};
    uint32_t tm_3_7_beamforming_idx{0}, tm_3_7_beamforming_reciprocal_idx{0};
    uint32_t RI{0}, CQI{0}, PMI{0};
};

/// pcc_report.hpp:29-43 -- plcf_decoder_t replaced by the raw PLCF bytes per type (index 0 = type 1, 1 = type 2)
struct pcc_report_t {
    bool plcf_ok[2]{false, false};
    std::array<uint8_t, 10> plcf[2]{};
    uint32_t crc_mask[2]{0, 0};      // 0 none, 0x5555 closed loop, 0xAAAA beamforming, 0xFFFF both
    float snr_dB{0.f};
    bool has_any_plcf() const { return plcf_ok[0] || plcf_ok[1]; }
    /// nullptr if that type did not pass its CRC (plcf_decoder_t::get_plcf_base semantics)
    const uint8_t* get_plcf_bytes(uint32_t PLCF_type) const {
        return (PLCF_type == 1 || PLCF_type == 2) && plcf_ok[PLCF_type - 1] ? plcf[PLCF_type - 1].data() : nullptr;
    }
    /// rx_synced_t::run_pcc_decoding_and_candidate_search (rx_synced.cpp:1011-1026): every type whose CRC passed is
    /// unpacked and checked against the radio device class; returns plcf_decoder.has_any_plcf()
    uint32_t decode_plcf(sp4::plcf_decoder_t& plcf_decoder) const {
        plcf_decoder.set_configuration();
        for (uint32_t type = 1; type <= 2; ++type)
            if (plcf_ok[type - 1]) plcf_decoder.decode_and_rdc_check(type, plcf[type - 1].data());
        return plcf_decoder.has_any_plcf();
    }
};

/// pdc_report.hpp:28-62 -- mac_pdu_decoder_t replaced by the decoded transport block bytes (decode_mac_pdu feeds one)
struct pdc_report_t {
    bool crc_status{false};
    const uint8_t* a_tb{nullptr};
    uint32_t N_TB_byte{0};
    float snr_dB{0.f};
    mimo_report_t mimo_report{};
    /// what rx_synced_t does code block by code block (rx_synced.cpp:394-400): MAC header type, common header and the
    /// demultiplexed SDUs of the transport block; true if the decoder reached a valid final state
    bool decode_mac_pdu(sp4::mac_pdu_decoder_t& mac_pdu_decoder, uint32_t mu = 1) const {
        if (!crc_status || a_tb == nullptr) return false;
        mac_pdu_decoder.set_configuration(a_tb, N_TB_byte, mu);
        mac_pdu_decoder.decode(N_TB_byte);
        return mac_pdu_decoder.has_reached_valid_final_state();
    }
};

/// maclow_phy_handle.hpp:28-48
enum class handle_pcc2pdc_t : uint32_t { th10 = 0, th20, th21, CARDINALITY };
struct maclow_phy_handle_t {
    maclow_phy_handle_t() = default;
    maclow_phy_handle_t(handle_pcc2pdc_t h, uint32_t lrdid_) : handle_pcc2pdc(h), lrdid(lrdid_) {}
    handle_pcc2pdc_t handle_pcc2pdc{handle_pcc2pdc_t::CARDINALITY};
    uint32_t lrdid{0};
};

/// phy_maclow.hpp:28-36
struct phy_maclow_t {
    const sync_report_t& sync_report;
    const pcc_report_t& pcc_report;
};

/// maclow_phy.hpp:29-85
struct maclow_phy_t {
    maclow_phy_t() = default;
    explicit maclow_phy_t(harq::process_rx_t* hp_rx_) : continue_with_pdc(true), hp_rx(hp_rx_) {}
    maclow_phy_t(harq::process_rx_t* hp_rx_, maclow_phy_handle_t h) : continue_with_pdc(true), hp_rx(hp_rx_), maclow_phy_handle(h) {}
    bool continue_with_pdc{false};
    harq::process_rx_t* hp_rx{nullptr};
    maclow_phy_handle_t maclow_phy_handle{};
    handle_pcc2pdc_t get_handle_pcc2pdc() const { return maclow_phy_handle.handle_pcc2pdc; }
    uint32_t get_handle_lrdid() const { return maclow_phy_handle.lrdid; }
};

/// phy_machigh.hpp:29-41
struct phy_machigh_t {
    const phy_maclow_t& phy_maclow;
    const maclow_phy_t& maclow_phy;
    const pdc_report_t& pdc_report;
};

/// machigh_phy.hpp:33-61 (irregular reports and channel scans are not part of this path)
struct machigh_phy_t {
    typedef std::vector<tx_descriptor_t> tx_descriptor_vec_t;
    tx_descriptor_vec_t tx_descriptor_vec{};
};

}  // namespace phy

namespace upper {
/// upper/tpoint.hpp:128-170: the three callbacks of the PCC/PDC path
class tpoint_t {
    public:
        virtual ~tpoint_t() = default;
        [[nodiscard]] virtual phy::maclow_phy_t work_pcc(const phy::phy_maclow_t& phy_maclow) = 0;
        [[nodiscard]] virtual phy::machigh_phy_t work_pdc(const phy::phy_machigh_t& phy_machigh) = 0;
        [[nodiscard]] virtual phy::machigh_phy_t work_pdc_error(const phy::phy_machigh_t& phy_machigh) = 0;
};
}  // namespace upper

namespace phy {

/// What worker_pool_t + worker_tx_rx_t + tx_t + rx_synced_t + sync_chunk_t are to the reference for this path, for a
/// BATCH of packets of one geometry on one GPU.
class batch_phy_t {
    public:
        struct stats_t {
            uint64_t windows{0}, detections{0}, pcc_ok{0}, pdc_requested{0}, pdc_crc_ok{0}, tx_packets{0};
        };

        /// cfg.max_batch = windows per process_windows call; nof_process_rx_ = receive HARQ processes of the firmware's
        /// process_pool_t (their ids address the first soft buffer slots of the context)
        batch_phy_t(const dnrp_config_t& cfg, int device, uint32_t nof_process_rx_) : config(cfg), max_windows(cfg.max_batch),
                                                                                     nof_process_rx(nof_process_rx_) {
            config.max_batch = max_windows + nof_process_rx;
            device_index = device;
            if (dnrp_ctx_create(&config, device, &ctx) != DNRP_OK)
                throw std::runtime_error(std::string("dnrp_ctx_create: ") + dnrp_last_error_string());
            dnrp_ctx_get_packet_sizes(ctx, &packet_sizes);
            tx_stride = dnrp_tx_stride(ctx);
            rx_stride = dnrp_rx_stride(ctx);
            plans.push_back(plan_t{config.psdef, ctx, packet_sizes, tx_stride, rx_stride});
            network_ids.push_back(config.network_id);
        }
        ~batch_phy_t() {
            for (auto& pl : plans) dnrp_ctx_destroy(pl.ctx);
        }

        /// A further packet geometry this PHY accepts (the reference allocates for the radio device class' maximum and
        /// takes every packet's own sizes from its HARQ process: tpoint.cpp:119-141, rx_synced.cpp:325-350).  The
        /// constructor's psdef is the MAXIMUM: sync and PCC run under it, its IQ window holds the longest packet.  A plan
        /// shares mu, beta, the number of effective TX antennas (the PCC layout) and the antenna / rate configuration.
        void add_plan(const sp3::packet_sizes_def_t& psdef) {
            if (find_plan(psdef) >= 0) return;
            dnrp_config_t c = config;
            c.psdef = psdef;
            plan_t pl{};
            pl.psdef = psdef;
            if (dnrp_ctx_create(&c, device_index, &pl.ctx) != DNRP_OK)
                throw std::runtime_error(std::string("dnrp_ctx_create (add_plan): ") + dnrp_last_error_string());
            dnrp_ctx_get_packet_sizes(pl.ctx, &pl.ps);
            pl.tx_stride = dnrp_tx_stride(pl.ctx);
            pl.rx_stride = dnrp_rx_stride(pl.ctx);
            if (pl.ps.psdef.u != packet_sizes.psdef.u || pl.ps.psdef.b != packet_sizes.psdef.b ||
                pl.ps.N_eff_TX != packet_sizes.N_eff_TX || pl.rx_stride > rx_stride) {
                dnrp_ctx_destroy(pl.ctx);
                throw std::invalid_argument("add_plan: geometry must share mu, beta and N_eff_TX with the maximum geometry and "
                                            "must not be longer than it");
            }
            for (uint32_t id : network_ids) check(dnrp_add_network_id(pl.ctx, id), "dnrp_add_network_id");
            if (track_cfo || track_sto) check(dnrp_rx_set_tracking(pl.ctx, track_cfo, track_sto), "dnrp_rx_set_tracking");
            plans.push_back(pl);
        }

        /// worker_pool_t::add_network_id (worker_pool.cpp:109-114): scrambling sequences of a further network
        void add_network_id(uint32_t network_id) {
            for (uint32_t id : network_ids) if (id == network_id) return;
            network_ids.push_back(network_id);
            for (auto& pl : plans) check(dnrp_add_network_id(pl.ctx, network_id), "dnrp_add_network_id");
        }
        batch_phy_t(const batch_phy_t&) = delete;
        batch_phy_t& operator=(const batch_phy_t&) = delete;

        const sp3::packet_sizes_t& get_packet_sizes() const { return packet_sizes; }
        uint32_t get_tx_stride() const { return tx_stride; }
        uint32_t get_rx_stride() const { return rx_stride; }
        uint32_t get_N_RX() const { return config.N_RX ? config.N_RX : 1; }
        const stats_t& get_stats() const { return stats; }

        /// samples of a window the packet search may consume (sync_chunk_t's search range); 0 = the whole window.  The
        /// device only sums the detection steps inside it: with windows of one packet plus slack,
        /// window - packet + STF + 2 patterns covers every packet that still fits
        void set_search_length(uint32_t n_samples) { search_length = n_samples; }

        /// coarse_peak_f_domain_t of the packet search (sync_param.hpp:238-252, compiled out in the reference): beta search
        /// with the given threshold (< 0 dB switches it on, the reference suggests -10) and integer CFO search over
        /// +- cfo_search_range_subc subcarriers (>= 4 switches it on, the reference suggests 4) for every following search
        void set_sync_fdomain(float beta_threshold_dB, uint32_t cfo_search_range_subc) {
            check(dnrp_sync_set_fdomain(ctx, beta_threshold_dB, cfo_search_range_subc), "dnrp_sync_set_fdomain");
        }
        /// residual CFO / STO tracking on the DRS (rx_synced_param.hpp:156-181, compiled out in the reference) for every
        /// following reception
        void set_tracking(bool residual_cfo, bool residual_sto) {
            track_cfo = residual_cfo; track_sto = residual_sto;
            for (auto& pl : plans) check(dnrp_rx_set_tracking(pl.ctx, residual_cfo, residual_sto), "dnrp_rx_set_tracking");
        }
        /// beta estimates and integer CFOs (subcarriers) of the windows of the last search (set_sync_fdomain switched on)
        void read_sync_fdomain(uint32_t n, std::vector<uint32_t>& beta, std::vector<int32_t>& cfo_subc) {
            beta.assign(n, 0);
            cfo_subc.assign(n, 0);
            check(dnrp_sync_fdomain_read(ctx, n, beta.data(), cfo_subc.data()), "dnrp_sync_fdomain_read");
        }

        /// Several batch_phy_t of ONE process, one per GPU: the sum of their counters in one NCCL all-reduce over the
        /// GPUs (dnrp_comm_init_all / dnrp_counters_allreduce_all; the reference adds into the members of one
        /// worker_pool_t / result_t of one process, worker_tx_rx.cpp:96-198, result.cpp:78-95).  The communicator is
        /// created by the first call.  Every PHY's own stats stay untouched; the totals are returned.
        static stats_t reduce_stats(const std::vector<batch_phy_t*>& phys) {
            std::vector<dnrp_ctx_t*> ctxs;
            for (batch_phy_t* p : phys) ctxs.push_back(p->ctx);
            bool have = true;
            for (batch_phy_t* p : phys) have = have && p->comm_ready;
            if (!have) {
                for (batch_phy_t* p : phys) if (p->comm_ready) { dnrp_comm_destroy(p->ctx); p->comm_ready = false; }
                check(dnrp_comm_init_all(ctxs.data(), (int)ctxs.size()), "dnrp_comm_init_all");
                for (batch_phy_t* p : phys) p->comm_ready = true;
            }
            constexpr uint32_t NC = 6;
            std::vector<int64_t> c(NC * phys.size());
            for (size_t i = 0; i < phys.size(); ++i) {
                const stats_t& s = phys[i]->stats;
                const uint64_t v[NC] = {s.windows, s.detections, s.pcc_ok, s.pdc_requested, s.pdc_crc_ok, s.tx_packets};
                for (uint32_t j = 0; j < NC; ++j) c[i * NC + j] = (int64_t)v[j];
            }
            check(dnrp_counters_allreduce_all(ctxs.data(), (int)ctxs.size(), c.data(), NC, 0), "dnrp_counters_allreduce_all");
            stats_t t;
            t.windows = (uint64_t)c[0]; t.detections = (uint64_t)c[1]; t.pcc_ok = (uint64_t)c[2];
            t.pdc_requested = (uint64_t)c[3]; t.pdc_crc_ok = (uint64_t)c[4]; t.tx_packets = (uint64_t)c[5];
            return t;
        }
        dnrp_ctx_t* get_ctx() { return ctx; }

        /// tx_t::generate_tx_packet for every descriptor (worker_tx_rx.cpp:300-335): IQ [n][N_TX][tx_stride] into iq_out
        /// (nullptr: stays on the device for channel()), then hp_tx.finalize().  All descriptors must share the context's
        /// geometry (the reference asserts psdef against maximum_packet_sizes, tx.cpp:430-482).
        void generate_tx_packets(std::vector<tx_descriptor_t>& descs, cf_t* iq_out, std::vector<uint32_t>* n_samples = nullptr) {
            const uint32_t n = (uint32_t)descs.size();
            if (n == 0) return;
            std::vector<uint32_t> ns(n);
            // one device batch per geometry; network id and PLCF type ride per packet (tx.cpp:176-186)
            std::vector<std::vector<uint32_t>> bucket(plans.size());
            for (uint32_t i = 0; i < n; ++i) {
                const int pl = find_plan(descs[i].hp_tx.get_psdef());
                if (pl < 0) throw std::invalid_argument("tx_descriptor_t: packet geometry not registered (batch_phy_t::add_plan)");
                bucket[(size_t)pl].push_back(i);
            }
            if (bucket[0].size() != n && iq_out == nullptr)
                throw std::invalid_argument("packets of several geometries need a host IQ buffer (device hand-over is per context)");
            for (size_t pl = 0; pl < plans.size(); ++pl) {
                const std::vector<uint32_t>& idx = bucket[pl];
                const uint32_t m = (uint32_t)idx.size();
                if (m == 0) continue;
                const plan_t& P = plans[pl];
                std::vector<dnrp_tx_desc_t> d(m);
                std::vector<uint32_t> ids(m), nsm(m);
                std::vector<uint8_t> tys(m), plcf(10 * (size_t)m), tb((size_t)m * P.ps.N_TB_byte);
                for (uint32_t q = 0; q < m; ++q) {
                    tx_descriptor_t& D = descs[idx[q]];
                    const tx_meta_t& mt = D.tx_meta;
                    d[q] = {D.codebook_index, mt.optimal_scaling_DAC ? 1u : 0u, mt.DAC_scale, mt.iq_phase_rad,
                            mt.iq_phase_increment_s2s_post_resampling_rad, mt.GI_percentage, D.hp_tx.get_rv()};
                    ids[q] = D.hp_tx.get_network_id();
                    tys[q] = (uint8_t)D.hp_tx.get_PLCF_type();
                    std::memcpy(&plcf[10 * (size_t)q], D.hp_tx.get_a_plcf(), 10);
                    std::memcpy(&tb[(size_t)q * P.ps.N_TB_byte], D.hp_tx.get_a_tb(), P.ps.N_TB_byte);
                }
                const bool direct = pl == 0 && m == n;      // one geometry: IQ straight into the caller's buffer
                std::vector<cf_t> tmp;
                if (!direct && iq_out) tmp.resize((size_t)m * P.ps.N_TX * P.tx_stride);
                check(dnrp_tx_batch_ex(P.ctx, m, d.data(), ids.data(), tys.data(), plcf.data(), tb.data(),
                                       direct ? iq_out : (iq_out ? tmp.data() : nullptr), nsm.data()), "dnrp_tx_batch_ex");
                for (uint32_t q = 0; q < m; ++q) {
                    ns[idx[q]] = nsm[q];
                    if (!direct && iq_out)     // packet slots of the caller's buffer have the maximum geometry's stride
                        for (uint32_t a = 0; a < P.ps.N_TX; ++a) {
                            cf_t* dst = iq_out + ((size_t)idx[q] * packet_sizes.N_TX + a) * tx_stride;
                            std::memset(dst, 0, sizeof(cf_t) * tx_stride);
                            std::memcpy(dst, &tmp[((size_t)q * P.ps.N_TX + a) * P.tx_stride], sizeof(cf_t) * std::min(P.tx_stride, tx_stride));
                        }
                }
            }
            for (uint32_t i = 0; i < n; ++i) descs[i].hp_tx.finalize();
            if (n_samples) *n_samples = ns;
            stats.tx_packets += n;
        }

        /// simulation only: the packets of the last generate_tx_packets call through the channel into the RX windows
        void channel(const std::vector<dnrp_channel_desc_t>& descs, cf_t* iq_rx_out = nullptr) {
            check(dnrp_channel_batch(ctx, (uint32_t)descs.size(), descs.data(), iq_rx_out), "dnrp_channel_batch");
        }

        /// One scheduling quantum of the receive path for n IQ windows [n][N_RX][rx_stride] at the hardware rate
        /// (nullptr: the windows left on the device by channel()); window_start_64[i] = global time of window i's first
        /// sample.  sync_chunk_t::search -> rx_synced_t::demoddecod_rx_pcc -> work_pcc (FIFO) -> demoddecod_rx_pdc ->
        /// work_pdc / work_pdc_error -> hp_rx->finalize.  Returns the TX descriptors the firmware handed back.
        machigh_phy_t::tx_descriptor_vec_t process_windows(uint32_t n, const cf_t* windows, const int64_t* window_start_64,
                                                           upper::tpoint_t& tpoint) {
            machigh_phy_t::tx_descriptor_vec_t tx_out;
            if (n == 0) return tx_out;
            if (n > max_windows) throw std::invalid_argument("more windows than the context was created for");
            std::vector<dnrp_sync_result_t> sres(n);
            check(dnrp_sync_batch(ctx, n, windows, search_length, sres.data()), "dnrp_sync_batch");
            std::vector<dnrp_pcc_report_t> pcc(n);
            check(dnrp_rx_pcc_batch(ctx, n, nullptr, nullptr, pcc.data()), "dnrp_rx_pcc_batch");
            stats.windows += n;
            last_claimed.assign(n, 0);
            for (uint32_t i = 0; i < n; ++i) last_claimed[i] = sres[i].detected && (pcc[i].plcf_ok[0] || pcc[i].plcf_ok[1]);
            deliver(n, sres, pcc, window_start_64, tpoint, tx_out);
            return tx_out;
        }
        /// process_windows for windows in the radio's wire format: [n][N_RX][rx_stride][2] int16 (I, Q), sample = (I, Q) x
        /// scale (1 / 32768 = UHD's sc16 -> fc32); half the bytes over the host link, converted on the device
        machigh_phy_t::tx_descriptor_vec_t process_windows_sc16(uint32_t n, const int16_t* windows, float scale,
                                                                const int64_t* window_start_64, upper::tpoint_t& tpoint) {
            machigh_phy_t::tx_descriptor_vec_t tx_out;
            if (n == 0) return tx_out;
            if (n > max_windows) throw std::invalid_argument("more windows than the context was created for");
            std::vector<dnrp_sync_result_t> sres(n);
            check(dnrp_sync_batch_sc16(ctx, n, windows, scale, search_length, sres.data()), "dnrp_sync_batch_sc16");
            std::vector<dnrp_pcc_report_t> pcc(n);
            check(dnrp_rx_pcc_batch(ctx, n, nullptr, nullptr, pcc.data()), "dnrp_rx_pcc_batch");
            stats.windows += n;
            last_claimed.assign(n, 0);
            for (uint32_t i = 0; i < n; ++i) last_claimed[i] = sres[i].detected && (pcc[i].plcf_ok[0] || pcc[i].plcf_ok[1]);
            deliver(n, sres, pcc, window_start_64, tpoint, tx_out);
            return tx_out;
        }
        /// per window of the last process_windows call: a packet was found AND one of its PLCF CRCs passed
        const std::vector<uint8_t>& get_last_claimed() const { return last_claimed; }

        /// The streaming front door: one continuous stream [N_RX][n_samples] at the hardware rate (first sample at
        /// global time stream_start_64) instead of pre-cut windows -- worker_sync_t / sync_chunk_t / baton_t
        /// (worker_sync.cpp:112-221): every unique packet found is handed to the synced receiver and the firmware in
        /// time order, max_windows packets per device batch.
        machigh_phy_t::tx_descriptor_vec_t process_stream(uint64_t n_samples, const cf_t* stream, int64_t stream_start_64,
                                                          upper::tpoint_t& tpoint, uint32_t chunk_hop = 0) {
            machigh_phy_t::tx_descriptor_vec_t tx_out;
            const uint32_t max_out = (uint32_t)(n_samples / std::max<uint32_t>(packet_sizes.N_samples_STF, 1u)) + 1;
            std::vector<dnrp_stream_packet_t> found(max_out);
            uint32_t n_found = 0;
            check(dnrp_stream_sync(ctx, n_samples, stream, chunk_hop, max_out, found.data(), &n_found), "dnrp_stream_sync");
            deliver_found(found.data(), n_found, stream_start_64, tpoint, tx_out);
            return tx_out;
        }

        /// The same for a radio that delivers its samples piece by piece (radio/buffer_rx.hpp read chunk after chunk by
        /// worker_sync_t, resampled with carried history by rx_pacer_t): open_stream sizes the ring on the device and
        /// starts the stream clock at stream_start_64; every push_stream appends hardware-rate samples [N_RX][n_samples]
        /// and hands every packet that is complete by then to the synced receiver and the firmware, in time order.
        /// Works with and without resampling.
        void open_stream(uint32_t max_push_samples, int64_t stream_start_64, uint32_t chunk_hop = 0) {
            check(dnrp_stream_open(ctx, max_push_samples, chunk_hop), "dnrp_stream_open");
            stream_clock_64 = stream_start_64;
        }
        machigh_phy_t::tx_descriptor_vec_t push_stream(uint32_t n_samples, const cf_t* samples, upper::tpoint_t& tpoint) {
            machigh_phy_t::tx_descriptor_vec_t tx_out;
            std::vector<dnrp_stream_packet_t> found(max_windows);
            const cf_t* src = samples;
            for (uint32_t n = n_samples;; n = 0, src = nullptr) {   // further rounds drain packets beyond max_windows
                uint32_t n_found = 0;
                check(dnrp_stream_push(ctx, n, src, max_windows, found.data(), &n_found), "dnrp_stream_push");
                deliver_found(found.data(), n_found, stream_clock_64, tpoint, tx_out);
                if (n_found < max_windows) break;
            }
            return tx_out;
        }
        /// push_stream for samples in the radio's wire format: [N_RX][n_samples][2] int16 (I, Q), sample = (I, Q) x scale
        machigh_phy_t::tx_descriptor_vec_t push_stream_sc16(uint32_t n_samples, const int16_t* samples, float scale,
                                                            upper::tpoint_t& tpoint) {
            machigh_phy_t::tx_descriptor_vec_t tx_out;
            std::vector<dnrp_stream_packet_t> found(max_windows);
            const int16_t* src = samples;
            for (uint32_t n = n_samples;; n = 0, src = nullptr) {
                uint32_t n_found = 0;
                check(dnrp_stream_push_sc16(ctx, n, src, scale, max_windows, found.data(), &n_found), "dnrp_stream_push_sc16");
                deliver_found(found.data(), n_found, stream_clock_64, tpoint, tx_out);
                if (n_found < max_windows) break;
            }
            return tx_out;
        }
        void close_stream() { check(dnrp_stream_close(ctx), "dnrp_stream_close"); }

        /// PHY_JSON_SWITCH_IMPLEMENT_ANY_JSON_FUNCTIONALITY of the reference (worker_tx_rx.cpp:356-414): from now on every
        /// packet whose PDC is decoded adds one entry in the reference's layout; a file "<prefix_file><0000000000>" every
        /// json_length entries (json_export_t, lib/src/common/json/json_export.cpp)
        void enable_json_export(uint32_t json_length, const std::string& prefix_file) {
            json_export = std::make_unique<common::json_export_t>(json_length, prefix_file);
        }
        /// writes the entries that do not fill a file yet; returns the files written so far
        std::vector<std::string> flush_json_export() {
            if (!json_export) return {};
            json_export->flush();
            return json_export->get_files();
        }

    private:
        int64_t stream_clock_64{0};

        /// packets found in the device-resident stream -> receiver windows -> PCC -> firmware, max_windows per batch
        void deliver_found(const dnrp_stream_packet_t* found, uint32_t n_found, int64_t stream_start_64, upper::tpoint_t& tpoint,
                           machigh_phy_t::tx_descriptor_vec_t& tx_out) {
            // room in front of the packet inside its window (the window has two hardware-rate STFs of slack)
            const uint32_t t0 = packet_sizes.N_samples_STF / 2;
            for (uint32_t first = 0; first < n_found; first += max_windows) {
                const uint32_t n = std::min(max_windows, n_found - first);
                std::vector<int64_t> start(n), window_start(n);
                std::vector<dnrp_sync_result_t> sres(n);
                std::vector<dnrp_sync_report_t> sync_in(n);
                for (uint32_t i = 0; i < n; ++i) {
                    const dnrp_stream_packet_t& f = found[first + i];
                    start[i] = f.fine_peak_time_64;
                    window_start[i] = stream_start_64 + f.fine_peak_time_64 - (int64_t)t0;
                    sres[i] = f.sync;                               // chunk-local record -> window-local times
                    sres[i].fine_peak_time_local = t0;
                    sres[i].coarse_peak_time_hw = (uint32_t)((int64_t)t0 + f.coarse_peak_time_64 - f.fine_peak_time_64);
                    sres[i].coarse_peak_time_local = sres[i].coarse_peak_time_hw;
                    sync_in[i] = dnrp_sync_report_t{};
                    sync_in[i].fine_peak_time_local = t0;
                    sync_in[i].cfo_fractional_rad = f.sync.cfo_fractional_rad;
                    sync_in[i].cfo_integer_rad = f.sync.cfo_integer_rad;
                    for (uint32_t a = 0; a < 8; ++a) sync_in[i].rms[a] = f.sync.rms_array[a];
                }
                check(dnrp_stream_extract(ctx, n, start.data(), t0), "dnrp_stream_extract");
                std::vector<dnrp_pcc_report_t> pcc(n);
                check(dnrp_rx_pcc_batch(ctx, n, nullptr, sync_in.data(), pcc.data()), "dnrp_rx_pcc_batch");
                stats.windows += n;
                deliver(n, sres, pcc, window_start.data(), tpoint, tx_out);
            }
        }
        /// sync results + PCC reports of n windows -> work_pcc (FIFO) -> demoddecod_rx_pdc -> work_pdc / work_pdc_error
        void deliver(uint32_t n, const std::vector<dnrp_sync_result_t>& sres, const std::vector<dnrp_pcc_report_t>& pcc,
                     const int64_t* window_start_64, upper::tpoint_t& tpoint, machigh_phy_t::tx_descriptor_vec_t& tx_out) {

            std::vector<sync_report_t> sync_reports;
            std::vector<pcc_report_t> pcc_reports(n);
            std::vector<maclow_phy_t> maclow(n);
            std::vector<uint8_t> keep(n, 0);
            sync_reports.reserve(n);
            const uint32_t N_RX = get_N_RX();
            for (uint32_t i = 0; i < n; ++i) {               // FIFO: job order = window order (worker_tx_rx.cpp:165-170)
                sync_reports.emplace_back(N_RX);
                sync_report_t& s = sync_reports.back();
                const dnrp_sync_result_t& r = sres[i];
                if (!r.detected) continue;
                ++stats.detections;
                const int64_t t0 = window_start_64 ? window_start_64[i] : 0;
                s.detection_ant_idx = r.detection_ant_idx; s.detection_rms = r.detection_rms;
                s.detection_metric = r.detection_metric; s.detection_time_local = r.detection_time_local;
                s.u = config.psdef.u; s.b = config.psdef.b; s.N_eff_TX = r.N_eff_TX;
                for (uint32_t a = 0; a < N_RX; ++a) { s.coarse_peak_array.at(a) = r.coarse_peak_array[a]; s.rms_array.at(a) = pcc[i].rms[a]; }
                s.coarse_peak_time_local = r.coarse_peak_time_local;
                s.coarse_peak_time_64 = t0 + r.coarse_peak_time_hw;
                s.cfo_fractional_rad = r.cfo_fractional_rad + pcc[i].cfo_residual_rad;   // rx_synced.cpp:552
                s.cfo_integer_rad = r.cfo_integer_rad;
                s.fine_peak_time_local = r.fine_peak_time_local;
                s.fine_peak_time_64 = t0 + r.fine_peak_time_local;
                s.sto_fractional = pcc[i].sto_fractional;                                   // rx_synced.cpp:581
                s.fine_peak_time_corrected_by_sto_fractional_64 = s.fine_peak_time_64 + (int64_t)std::lround(s.sto_fractional);
                pcc_report_t& p = pcc_reports[i];
                for (uint32_t t = 0; t < 2; ++t) {
                    p.plcf_ok[t] = pcc[i].plcf_ok[t] != 0;
                    std::memcpy(p.plcf[t].data(), pcc[i].plcf[t], 10);
                    p.crc_mask[t] = pcc[i].crc_mask[t];
                }
                p.snr_dB = pcc[i].snr_dB;
                if (!p.has_any_plcf()) continue;                // worker_tx_rx.cpp:129: no callback without a valid PLCF
                ++stats.pcc_ok;
                maclow[i] = tpoint.work_pcc(phy_maclow_t{s, p});
                if (!maclow[i].continue_with_pdc || maclow[i].hp_rx == nullptr) continue;
                keep[i] = 1;
                ++stats.pdc_requested;
            }
            // one PDC batch per geometry: the maximum geometry's context holds every window already (sync + PCC ran
            // there); packets that announced another registered geometry travel to its context on the device, where
            // the front end runs again under that geometry (rx_synced.cpp:325-350 takes the sizes from hp_rx)
            std::vector<std::vector<uint32_t>> bucket(plans.size());
            for (uint32_t i = 0; i < n; ++i) {
                if (!keep[i]) continue;
                const int pl = find_plan(maclow[i].hp_rx->get_psdef());
                if (pl < 0) throw std::runtime_error("work_pcc chose a packet geometry that is not registered (batch_phy_t::add_plan)");
                bucket[(size_t)pl].push_back(i);
            }
            for (size_t pl = 0; pl < plans.size(); ++pl) {
                const std::vector<uint32_t>& idx = bucket[pl];
                if (idx.empty()) continue;
                const plan_t& P = plans[pl];
                // windows of this batch: plan 0 decodes all n in place (the unwanted ones on scratch slots), another
                // plan its own m
                const uint32_t m = pl == 0 ? n : (uint32_t)idx.size();
                std::vector<uint32_t> slot(m), ids(m, config.network_id), win(m);
                std::vector<uint8_t> rv(m, 0), reset(m, 1), tys(m, (uint8_t)config.plcf_type), mine(m, 0);
                std::vector<uint8_t> used(nof_process_rx, 0);
                for (uint32_t q = 0; q < m; ++q) { slot[q] = nof_process_rx + q; win[q] = pl == 0 ? q : idx[q]; }
                for (uint32_t z = 0; z < (uint32_t)idx.size(); ++z) {
                    const uint32_t i = idx[z], q = pl == 0 ? i : z;
                    harq::process_rx_t* hp = maclow[i].hp_rx;
                    const uint32_t id = hp->get_id();
                    // a receive process id IS its soft buffer slot [0, nof_process_rx) in the context of its geometry
                    if (id >= nof_process_rx || used[id])
                        throw std::runtime_error("HARQ process assigned twice in one batch or id beyond nof_process_rx");
                    used[id] = 1; slot[q] = id; rv[q] = (uint8_t)hp->get_rv();
                    reset[q] = hp->is_first_transmission() ? 1 : 0;
                    ids[q] = hp->get_network_id(); tys[q] = (uint8_t)hp->get_PLCF_type(); mine[q] = 1;
                }
                if (pl != 0) {
                    check(dnrp_rx_import_windows(P.ctx, ctx, m, win.data()), "dnrp_rx_import_windows");
                    std::vector<dnrp_sync_report_t> sync_in(m);
                    std::vector<dnrp_pcc_report_t> pcc2(m);
                    for (uint32_t q = 0; q < m; ++q) {
                        const dnrp_sync_result_t& r = sres[win[q]];
                        sync_in[q] = dnrp_sync_report_t{};
                        sync_in[q].fine_peak_time_local = r.fine_peak_time_local;
                        sync_in[q].cfo_fractional_rad = r.cfo_fractional_rad;
                        sync_in[q].cfo_integer_rad = r.cfo_integer_rad;
                        for (uint32_t a = 0; a < 8; ++a) sync_in[q].rms[a] = r.rms_array[a];
                    }
                    check(dnrp_rx_pcc_batch(P.ctx, m, nullptr, sync_in.data(), pcc2.data()), "dnrp_rx_pcc_batch");
                }
                std::vector<uint8_t> tb((size_t)m * P.ps.N_TB_byte);
                std::vector<dnrp_pdc_report_t> pdc(m);
                check(dnrp_rx_pdc_batch_ex(P.ctx, m, slot.data(), rv.data(), reset.data(), ids.data(), tys.data(), tb.data(),
                                           pdc.data(), nullptr), "dnrp_rx_pdc_batch_ex");
                std::vector<float> chestim;
                const uint32_t Nk = P.ps.N_b_OCC + 1, N_eff = P.ps.N_eff_TX;
                if (json_export) {   // worker_tx_rx.cpp:356-414: one JSON entry per packet whose PDC was decoded
                    chestim.resize(2 * (size_t)m * N_RX * N_eff * Nk);
                    check(dnrp_tap_read(P.ctx, DNRP_TAP_CHESTIM, m, chestim.data(), chestim.size() * sizeof(float)), "dnrp_tap_read");
                }
                for (uint32_t q = 0; q < m; ++q) {
                    if (!mine[q]) continue;
                    const uint32_t i = win[q];
                    harq::process_rx_t* hp = maclow[i].hp_rx;
                    if (json_export) {
                        common::json_packet_t j;
                        const sync_report_t& sr = sync_reports[i];
                        j.worker_id = (uint32_t)device_index;
                        const double dect_os = (double)config.psdef.u * config.psdef.b * 1.728e6 * config.oversampling;
                        j.samp_rate = (uint32_t)std::lround(dect_os * config.resample_L / config.resample_M);
                        j.N_TX_min = N_RX;
                        j.enforce_dectnrp_samp_rate_by_resampling = config.resample_L != config.resample_M;
                        j.L = config.resample_L; j.M = config.resample_M;
                        j.dect_samp_rate_max_oversampled = (uint32_t)std::lround(dect_os);
                        j.detection_ant_idx = sr.detection_ant_idx; j.detection_rms = sr.detection_rms;
                        j.detection_metric = sr.detection_metric;
                        j.u = sr.u; j.b = sr.b; j.N_eff_TX = sr.N_eff_TX;
                        for (uint32_t a = 0; a < N_RX && a < 8; ++a) {
                            j.coarse_peak_array[a] = sr.coarse_peak_array.at(a); j.rms_array[a] = sr.rms_array.at(a);
                        }
                        j.cfo_f = sr.cfo_fractional_rad; j.cfo_i = sr.cfo_integer_rad; j.sto_fractional = sr.sto_fractional;
                        j.coarse_peak_time = sr.coarse_peak_time_64; j.fine_peak_time = sr.fine_peak_time_64;
                        j.fine_peak_time_corrected_by_sto_fractional = sr.fine_peak_time_corrected_by_sto_fractional_64;
                        j.snr = pdc[q].snr_dB; j.mcs = P.psdef.mcs_index;
                        j.N_b_DFT = P.ps.N_b_DFT; j.N_b_OCC_plus_DC = Nk; j.N_RX = N_RX;
                        // the export loops over sync_report.N_eff_TX streams; the estimate holds the geometry's
                        j.N_eff_TX = N_eff;
                        j.chestim = &chestim[2 * (size_t)q * N_RX * N_eff * Nk];
                        j.PLCF_type = hp->get_PLCF_type();
                        const uint8_t* pb = pcc_reports[i].get_plcf_bytes(j.PLCF_type);
                        j.HeaderFormat = pb ? (uint32_t)(pb[0] >> 5) : 0u;
                        json_export->append(j);
                    }
                    std::memcpy(hp->get_a_tb(), &tb[(size_t)q * P.ps.N_TB_byte], P.ps.N_TB_byte);
                    pdc_report_t rep;
                    rep.crc_status = pdc[q].crc_ok != 0;
                    rep.a_tb = hp->get_a_tb(); rep.N_TB_byte = P.ps.N_TB_byte;
                    rep.snr_dB = pdc[q].snr_dB;
                    rep.mimo_report.N_RX = pdc[q].mimo_N_RX; rep.mimo_report.N_TS_other = pdc[q].mimo_N_TS_other;
                    rep.mimo_report.tm_3_7_beamforming_idx = pdc[q].tm_3_7_beamforming_idx;
                    rep.mimo_report.tm_3_7_beamforming_reciprocal_idx = pdc[q].tm_3_7_beamforming_reciprocal_idx;
                    const phy_maclow_t pm{sync_reports[i], pcc_reports[i]};
                    const phy_machigh_t in{pm, maclow[i], rep};
                    machigh_phy_t out = rep.crc_status ? tpoint.work_pdc(in) : tpoint.work_pdc_error(in);
                    if (rep.crc_status) ++stats.pdc_crc_ok;
                    hp->finalize(rep.crc_status);                   // worker_tx_rx.cpp:228
                    for (auto& d : out.tx_descriptor_vec) tx_out.push_back(d);
                }
            }
        }

        std::unique_ptr<common::json_export_t> json_export;

        struct plan_t {
            sp3::packet_sizes_def_t psdef;
            dnrp_ctx_t* ctx;
            sp3::packet_sizes_t ps;
            uint32_t tx_stride, rx_stride;
        };
        std::vector<plan_t> plans;           // [0] = the maximum geometry (sync + PCC run there)
        std::vector<uint32_t> network_ids;
        int device_index{0};
        int find_plan(const sp3::packet_sizes_def_t& psdef) const {
            for (size_t i = 0; i < plans.size(); ++i)
                if (std::memcmp(&plans[i].psdef, &psdef, sizeof(psdef)) == 0) return (int)i;
            return -1;
        }
        dnrp_config_t config;
        uint32_t max_windows, nof_process_rx;
        dnrp_ctx_t* ctx{nullptr};
        sp3::packet_sizes_t packet_sizes{};
        uint32_t tx_stride{0}, rx_stride{0};
        stats_t stats;
        bool comm_ready{false};
        uint32_t search_length{0};
        std::vector<uint8_t> last_claimed;
        bool track_cfo{false}, track_sto{false};
        static void check(int rc, const char* what) {
            if (rc != DNRP_OK) throw std::runtime_error(std::string(what) + ": " + dnrp_last_error_string());
        }
};

/// Several beta behind one receiver.  The reference's synchronisation, configured for the radio device class' largest
/// beta, "is triggered by packets for b=8,4,2,1" and can "find the best fitting b in frequency domain" (sync_param.hpp:
/// 238-244; the option is compiled out there).  Here: one batch_phy_t per beta, all at the same sample rate (beta x
/// oversampling constant, so windows, STF length and detection are common).  The largest beta's PHY searches every
/// window with the beta search on and receives its own packets; the windows whose estimate names a smaller beta are
/// handed, gathered, to that beta's PHY, which searches and receives them under its own geometry.  The firmware learns a
/// packet's beta from sync_report_t::b.  work_pcc order: FIFO inside every beta, the largest beta first.
class multi_beta_phy_t {
    public:
        multi_beta_phy_t(const dnrp_config_t& cfg_largest_beta, int device_, uint32_t nof_process_rx_,
                         float beta_threshold_dB = -10.0f)
            : device(device_), nof_process_rx(nof_process_rx_), base(cfg_largest_beta) {
            phys.emplace_back(std::make_unique<batch_phy_t>(cfg_largest_beta, device, nof_process_rx));
            phys[0]->set_sync_fdomain(beta_threshold_dB, 0);
        }
        /// a smaller beta (its largest packet geometry, as for batch_phy_t); oversampling follows from the common rate
        batch_phy_t& add_beta(const sp3::packet_sizes_def_t& psdef_largest_of_beta) {
            const uint32_t b = psdef_largest_of_beta.b, rate = base.psdef.b * base.oversampling;
            if (psdef_largest_of_beta.u != base.psdef.u || b == 0 || b >= base.psdef.b || rate % b != 0)
                throw std::invalid_argument("add_beta: same mu, a smaller beta that divides beta x oversampling of the receiver");
            dnrp_config_t c = base;
            c.psdef = psdef_largest_of_beta;
            c.oversampling = rate / b;
            phys.emplace_back(std::make_unique<batch_phy_t>(c, device, nof_process_rx));
            if (phys.back()->get_rx_stride() > phys[0]->get_rx_stride()) {
                phys.pop_back();
                throw std::invalid_argument("add_beta: the packet must fit the windows of the largest beta");
            }
            return *phys.back();
        }
        /// Another subcarrier scaling factor mu ("mu detection" is future work in the reference: README.md:198, the
        /// synchronisation assumes the radio device class' mu, sync_chunk.cpp:168-172).  A packet of another mu has another
        /// STF pattern length at the receiver's sample rate, so the receiver's detector does not fire on it (the lag-L
        /// autocorrelation of a signal on every 4th subcarrier of a 2L / 4L grid sums to zero) and the beta search cannot
        /// name it: the windows no PHY has claimed so far (no packet found, or no PLCF CRC passed) are searched again by
        /// the PHY of that mu.  psdef = the largest geometry of that mu; mu x beta x oversampling = the receiver's.
        batch_phy_t& add_mu(const sp3::packet_sizes_def_t& psdef_largest_of_mu) {
            const uint32_t ub = psdef_largest_of_mu.u * psdef_largest_of_mu.b, rate = base.psdef.u * base.psdef.b * base.oversampling;
            if (psdef_largest_of_mu.u == base.psdef.u || ub == 0 || rate % ub != 0)
                throw std::invalid_argument("add_mu: another mu with mu x beta dividing mu x beta x oversampling of the receiver");
            dnrp_config_t c = base;
            c.psdef = psdef_largest_of_mu;
            c.oversampling = rate / ub;
            phys.emplace_back(std::make_unique<batch_phy_t>(c, device, nof_process_rx));
            if (phys.back()->get_rx_stride() > phys[0]->get_rx_stride()) {
                phys.pop_back();
                throw std::invalid_argument("add_mu: the packet must fit the windows of the receiver");
            }
            return *phys.back();
        }
        batch_phy_t& get(uint32_t b) { return get(base.psdef.u, b); }
        batch_phy_t& get(uint32_t u, uint32_t b) {
            for (auto& p : phys) if (p->get_packet_sizes().psdef.u == u && p->get_packet_sizes().psdef.b == b) return *p;
            throw std::invalid_argument("no PHY for this mu / beta");
        }
        /// n host windows [n][N_RX][rx_stride of the receiver]; returns all TX descriptors handed back
        machigh_phy_t::tx_descriptor_vec_t process_windows(uint32_t n, const cf_t* windows, const int64_t* window_start_64,
                                                           upper::tpoint_t& tpoint) {
            if (!windows) throw std::invalid_argument("multi_beta_phy_t needs the windows in host memory");
            machigh_phy_t::tx_descriptor_vec_t tx = phys[0]->process_windows(n, windows, window_start_64, tpoint);
            std::vector<uint32_t> beta;
            std::vector<int32_t> cfo;
            phys[0]->read_sync_fdomain(n, beta, cfo);
            last_beta = beta;
            last_mu.assign(n, 0);
            std::vector<uint8_t> claimed = phys[0]->get_last_claimed();
            for (uint32_t i = 0; i < n; ++i) if (claimed[i]) last_mu[i] = base.psdef.u;
            const size_t wlen = (size_t)phys[0]->get_N_RX() * phys[0]->get_rx_stride();
            for (size_t k = 1; k < phys.size(); ++k) {
                batch_phy_t& P = *phys[k];
                const uint32_t u = P.get_packet_sizes().psdef.u, b = P.get_packet_sizes().psdef.b;
                const bool same_mu = u == base.psdef.u;
                const size_t plen = (size_t)P.get_N_RX() * P.get_rx_stride();
                std::vector<cf_t> sub;
                std::vector<int64_t> start;
                std::vector<uint32_t> from;
                for (uint32_t i = 0; i < n; ++i) {
                    if (same_mu ? beta[i] != b : claimed[i] != 0) continue;
                    // antenna by antenna: this PHY's window may be shorter than the receiver's
                    for (uint32_t a = 0; a < P.get_N_RX(); ++a) {
                        const cf_t* src = windows + (size_t)i * wlen + (size_t)a * phys[0]->get_rx_stride();
                        sub.insert(sub.end(), src, src + P.get_rx_stride());
                    }
                    start.push_back(window_start_64 ? window_start_64[i] : 0);
                    from.push_back(i);
                }
                const uint32_t m = (uint32_t)(sub.size() / std::max<size_t>(plen, 1));
                if (m == 0) continue;
                machigh_phy_t::tx_descriptor_vec_t t = P.process_windows(m, sub.data(), start.data(), tpoint);
                for (const auto& d : t) tx.push_back(d);
                const std::vector<uint8_t>& c = P.get_last_claimed();
                for (uint32_t z = 0; z < m; ++z)
                    if (c[z]) { claimed[from[z]] = 1; last_mu[from[z]] = u; if (!same_mu) last_beta[from[z]] = b; }
            }
            return tx;
        }
        const std::vector<uint32_t>& get_last_beta() const { return last_beta; }
        /// mu of the PHY that claimed the window (0: none did)
        const std::vector<uint32_t>& get_last_mu() const { return last_mu; }

    private:
        int device;
        uint32_t nof_process_rx;
        dnrp_config_t base;
        std::vector<std::unique_ptr<batch_phy_t>> phys;
        std::vector<uint32_t> last_beta, last_mu;
};

}  // namespace phy
}  // namespace dectnrp_b200

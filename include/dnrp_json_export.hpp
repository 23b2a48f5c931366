// dnrp_json_export.hpp -- per-packet JSON export of the PHY, in the reference's file format, so that its MATLAB / Python
// viewers (json/ in the reference tree) read what this library decoded.  Mirrors, without third-party dependencies:
//   common::json_export_t                   lib/src/common/json/json_export.cpp:30-100 (entries "packet_<n>", a file every
//                                           json_length entries, "<prefix><10 digit number>")
//   worker_tx_rx_t::collect_and_write_json  lib/src/phy/pool/worker_tx_rx.cpp:356-414 (RADIO / PHY / MAC sections)
//   rx_synced_t::get_json                   lib/src/phy/rx/rx_synced/rx_synced.cpp:444-480 (snr, mcs, ch_<rx>_<ts> as
//                                           integers scaled by 1e5, adt/cast.hpp:30-60)
//   plcf_decoder_t::get_json                lib/src/sections_part4/physical_header_field/plcf_decoder.cpp:154-164
// Key order and nesting are the reference's (nlohmann::ordered_json keeps insertion order); numbers are printed with
// enough digits to round-trip a float / double.
#pragma once
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <fstream>
#include <limits>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace dectnrp_b200 {
namespace common {

/// adt/cast.hpp:44-66: float -> int32 with the scale 1e5, saturating
inline int32_t float_to_int_1e5(float v) {
    const float s = v * 1.0e5f;
    if (s <= (float)std::numeric_limits<int32_t>::min()) return std::numeric_limits<int32_t>::min();
    if (s >= (float)std::numeric_limits<int32_t>::max()) return std::numeric_limits<int32_t>::max();
    return (int32_t)s;
}

/// what worker_tx_rx_t::collect_and_write_json gathers for one packet
struct json_packet_t {
    uint32_t worker_id{0};
    // RADIO
    uint32_t samp_rate{0}, N_TX_min{1};
    float tx_power_ant_0dBFS{0.f};
    float rx_power_ant_0dBFS[8]{};
    // PHY.worker_pool_config
    bool enforce_dectnrp_samp_rate_by_resampling{false};
    uint32_t L{1}, M{1}, dect_samp_rate_max_oversampled{0};
    // PHY.sync_report
    uint32_t detection_ant_idx{0};
    float detection_rms{0.f}, detection_metric{0.f};
    uint32_t u{1}, b{1}, N_eff_TX{1};
    float coarse_peak_array[8]{}, rms_array[8]{};
    float cfo_f{0.f}, cfo_i{0.f}, sto_fractional{0.f};
    int64_t coarse_peak_time{0}, fine_peak_time{0}, fine_peak_time_corrected_by_sto_fractional{0};
    // PHY.rx_synced
    float snr{0.f};
    uint32_t mcs{0}, N_b_DFT{64}, N_b_OCC_plus_DC{57};
    uint32_t N_RX{1};
    const float* chestim{nullptr};   ///< interleaved re, im: [N_RX][N_eff_TX][N_b_OCC_plus_DC] (DNRP_TAP_CHESTIM)
    // MAC.plcf
    uint32_t PLCF_type{1}, HeaderFormat{0};
};

class json_export_t {
    public:
        /// json_length entries per file, files "<prefix_file><0000000000>", entries "<prefix_entry><n>"
        json_export_t(uint32_t json_length_, std::string prefix_file_, std::string prefix_entry_ = "packet_")
            : json_length(json_length_ ? json_length_ : 1), prefix_file(std::move(prefix_file_)), prefix_entry(std::move(prefix_entry_)) {}
        ~json_export_t() { flush(); }

        void append(const json_packet_t& p) {
            std::ostringstream o;
            o.precision(17);
            const char* I2 = "        ";
            o << "    \"" << prefix_entry << postfix_entry << "\": {\n";
            o << I2 << "\"worker_id\": " << p.worker_id << ",\n";
            o << I2 << "\"elapsed_since_epoch\": "
              << std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::system_clock::now().time_since_epoch()).count() << ",\n";
            o << I2 << "\"32fc_int32_scale\": 100000.0,\n";
            o << I2 << "\"RADIO\": {\n";
            o << I2 << "    \"samp_rate\": " << p.samp_rate << ",\n";
            o << I2 << "    \"N_TX_min\": " << p.N_TX_min << ",\n";
            o << I2 << "    \"tx_power_ant_0dBFS\": " << num(p.tx_power_ant_0dBFS) << ",\n";
            o << I2 << "    \"rx_power_ant_0dBFS\": " << arr(p.rx_power_ant_0dBFS, 8) << "\n";
            o << I2 << "},\n";
            o << I2 << "\"PHY\": {\n";
            o << I2 << "    \"worker_pool_config\": {\n";
            o << I2 << "        \"enforce_dectnrp_samp_rate_by_resampling\": " << (p.enforce_dectnrp_samp_rate_by_resampling ? "true" : "false") << ",\n";
            o << I2 << "        \"L\": " << p.L << ",\n";
            o << I2 << "        \"M\": " << p.M << ",\n";
            o << I2 << "        \"dect_samp_rate_max_oversampled\": " << p.dect_samp_rate_max_oversampled << "\n";
            o << I2 << "    },\n";
            o << I2 << "    \"sync_report\": {\n";
            const char* I4 = "                ";
            o << I4 << "\"detection_ant_idx\": " << p.detection_ant_idx << ",\n";
            o << I4 << "\"detection_rms\": " << num(p.detection_rms) << ",\n";
            o << I4 << "\"detection_metric\": " << num(p.detection_metric) << ",\n";
            o << I4 << "\"u\": " << p.u << ",\n";
            o << I4 << "\"coarse_peak_array\": " << arr(p.coarse_peak_array, 8) << ",\n";
            o << I4 << "\"rms_array\": " << arr(p.rms_array, 8) << ",\n";
            o << I4 << "\"cfo_f\": " << num(p.cfo_f) << ",\n";
            o << I4 << "\"b\": " << p.b << ",\n";
            o << I4 << "\"cfo_i\": " << num(p.cfo_i) << ",\n";
            o << I4 << "\"coarse_peak_time\": " << p.coarse_peak_time << ",\n";
            o << I4 << "\"N_eff_TX\": " << p.N_eff_TX << ",\n";
            o << I4 << "\"fine_peak_time\": " << p.fine_peak_time << ",\n";
            o << I4 << "\"sto_fractional\": " << num(p.sto_fractional) << ",\n";
            o << I4 << "\"fine_peak_time_corrected_by_sto_fractional\": " << p.fine_peak_time_corrected_by_sto_fractional << "\n";
            o << I2 << "    },\n";
            o << I2 << "    \"rx_synced\": {\n";
            o << I4 << "\"snr\": " << num(p.snr) << ",\n";
            o << I4 << "\"mcs\": " << p.mcs << ",\n";
            o << I4 << "\"N_b_DFT\": " << p.N_b_DFT << ",\n";
            o << I4 << "\"N_b_OCC_plus_DC\": " << p.N_b_OCC_plus_DC << ",\n";
            o << I4 << "\"stride\": 1";
            const uint32_t n_rx = p.N_RX < 8 ? p.N_RX : 8, n_ts = p.N_eff_TX < 8 ? p.N_eff_TX : 8;   // limits::dectnrp_max_nof_antennas
            for (uint32_t a = 0; a < n_rx && p.chestim; ++a)
                for (uint32_t t = 0; t < n_ts; ++t) {
                    const float* h = p.chestim + 2 * ((size_t)a * p.N_eff_TX + t) * p.N_b_OCC_plus_DC;
                    o << ",\n" << I4 << "\"ch_" << a << "_" << t << "\": {\n";
                    for (int part = 0; part < 2; ++part) {
                        o << I4 << "    \"" << (part ? "im" : "re") << "\": [";
                        for (uint32_t k = 0; k < p.N_b_OCC_plus_DC; ++k) o << (k ? ", " : "") << float_to_int_1e5(h[2 * k + part]);
                        o << "]" << (part ? "\n" : ",\n");
                    }
                    o << I4 << "}";
                }
            o << "\n" << I2 << "    }\n";
            o << I2 << "},\n";
            o << I2 << "\"MAC\": {\n";
            o << I2 << "    \"plcf\": {\n";
            o << I4 << "\"PLCF_type\": " << p.PLCF_type << ",\n";
            o << I4 << "\"HeaderFormat\": " << p.HeaderFormat << "\n";
            o << I2 << "    }\n";
            o << I2 << "}\n";
            o << "    }";
            entries.push_back(o.str());
            ++postfix_entry;
            if (postfix_entry % json_length == 0) flush();
        }

        /// writes the entries collected so far (the reference only writes full files; the rest would be lost at exit)
        void flush() {
            if (entries.empty()) return;
            char num10[16];
            std::snprintf(num10, sizeof(num10), "%010u", postfix_file++);
            const std::string filename = prefix_file + num10;
            std::ofstream f(filename);
            if (!f) throw std::runtime_error("json_export_t: cannot write " + filename);
            f << "{\n";
            for (size_t i = 0; i < entries.size(); ++i) f << entries[i] << (i + 1 < entries.size() ? ",\n" : "\n");
            f << "}\n";
            entries.clear();
            files.push_back(filename);
        }
        const std::vector<std::string>& get_files() const { return files; }

    private:
        static std::string num(double v) {
            if (!std::isfinite(v)) return "null";   // nlohmann prints null for NaN / inf
            char b[40];
            std::snprintf(b, sizeof(b), "%.17g", v);
            std::string s(b);
            if (s.find_first_of(".eEn") == std::string::npos) s += ".0";   // floats stay floats for the viewers
            return s;
        }
        static std::string arr(const float* v, uint32_t n) {
            std::string s = "[";
            for (uint32_t i = 0; i < n; ++i) s += (i ? ", " : "") + num(v[i]);
            return s + "]";
        }
        uint32_t json_length;
        std::string prefix_file, prefix_entry;
        uint32_t postfix_entry{0}, postfix_file{0};
        std::vector<std::string> entries, files;
};

}  // namespace common
}  // namespace dectnrp_b200

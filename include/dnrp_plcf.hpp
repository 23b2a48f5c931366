// dnrp_plcf.hpp -- Physical Layer Control Field codec at the PHY-to-MAC boundary (header-only C++17, host side).
//
// Mirrors, with the reference's type and field names, what firmware reads out of phy::pcc_report_t:
//   sp4::plcf_base_t / plcf_10_t / plcf_20_t / plcf_21_t   lib/include/dectnrp/sections_part4/physical_header_field/
//                                                          plcf_{base,10,20,21}.hpp, src/.../plcf_{base,10,20,21}.cpp
//   sp4::feedback_info_f1_t ... f6_t, feedback_info_pool_t  feedback_info.hpp, feedback_info.cpp:28-427
//   sp4::plcf_decoder_t (decode_and_rdc_check, has_any_plcf, get_plcf_base)   plcf_decoder.cpp:44-150
// (ETSI TS 103 636-4 clause 6.2, Tables 6.2.1-1, 6.2.1-2, 6.2.1-2a, 6.2.1-3a, 6.2.2-1...6.2.2-3).
//
// Design: every header is a list of bit-field descriptors {member, first bit, width} over the MSB-first bit string;
// one generic routine packs / unpacks any of them (the reference hand-writes the shifts per class).  Field semantics,
// validity rules and the radio-device-class check are the reference's, so firmware written against
// plcf_decoder_t::get_plcf_base() reads the same values.  The CUDA receiver delivers the PLCF bytes whose CRC16
// (with its closed-loop / beamforming mask) passed; this codec is pure host code and never touches the device.
#pragma once
#include <array>
#include <cassert>
#include <cstddef>
#include <cstdint>
#include <limits>

namespace dectnrp_b200::sp4 {

inline constexpr uint32_t UNDEFINED_NUMERIC_32 = std::numeric_limits<uint32_t>::max();

namespace detail {
/// value of `width` bits starting at MSB-first bit position `first` of the byte string a
inline uint32_t get_bits(const uint8_t* a, uint32_t first, uint32_t width) {
    uint32_t v = 0;
    for (uint32_t b = first; b < first + width; ++b) v = (v << 1) | ((a[b >> 3] >> (7 - (b & 7))) & 1u);
    return v;
}
inline void put_bits(uint8_t* a, uint32_t first, uint32_t width, uint32_t v) {
    for (uint32_t i = 0; i < width; ++i) {
        const uint32_t b = first + i, bit = (v >> (width - 1 - i)) & 1u;
        a[b >> 3] = (uint8_t)((a[b >> 3] & ~(0x80u >> (b & 7))) | (bit ? (0x80u >> (b & 7)) : 0u));
    }
}
template <class T>
struct field_t {
    uint32_t T::*member;
    uint32_t first, width;
};
template <class T, size_t N>
void pack_fields(const T& t, const std::array<field_t<T>, N>& f, uint8_t* a) {
    for (const auto& e : f) put_bits(a, e.first, e.width, t.*(e.member));
}
template <class T, size_t N>
void unpack_fields(T& t, const std::array<field_t<T>, N>& f, const uint8_t* a) {
    for (const auto& e : f) t.*(e.member) = get_bits(a, e.first, e.width);
}
/// every member fits its field
template <class T, size_t N>
bool fields_fit(const T& t, const std::array<field_t<T>, N>& f) {
    for (const auto& e : f)
        if (e.width < 32 && (t.*(e.member) >> e.width) != 0) return false;
    return true;
}
}  // namespace detail

// ---------------------------------------------------------------------------------------------------------------
// feedback info (12 bits behind the 4-bit FeedbackFormat of a type 2 header), feedback_info.cpp
struct feedback_info_t {
    static constexpr uint32_t No_feedback = 0;
    enum class transmission_feedback_t : uint32_t { NACK = 0, ACK = 1, not_defined = UNDEFINED_NUMERIC_32 };
    enum class mimo_feedback_t : uint32_t { single_layer = 0, dual_layer = 1, four_layer = 2, reserved = 3,
                                            not_defined = UNDEFINED_NUMERIC_32 };
    static constexpr uint32_t MCS_out_of_range = UNDEFINED_NUMERIC_32;
    /// Table 6.2.2-4: buffer status code -> lower bound of the buffer size in bytes (feedback_info.hpp:58-77)
    static constexpr std::array<uint32_t, 16> buffer_status_2_buffer_size_lower = {
        0, 0, 16, 32, 64, 128, 256, 512, 1024, 2048, 4096, 8192, 16384, 32768, 65536, 131072};
    /// smallest code whose range (lower, upper] holds the size; 0 for an empty buffer, 15 above 131072
    static uint32_t buffer_size_2_buffer_status(uint32_t buffer_size) {
        if (buffer_size == 0) return 0;
        uint32_t code = 1;
        for (uint32_t upper = 16; code < 15 && buffer_size > upper; upper <<= 1) ++code;
        return code;
    }
    static uint32_t mcs_2_cqi(uint32_t mcs) { return mcs + 1; }
    static uint32_t cqi_2_mcs(uint32_t cqi) { return (cqi == 0 || cqi > 12) ? MCS_out_of_range : cqi - 1; }
};

/// The six formats share one record: the members a format does not carry keep their zero() state.
/// Bit positions are relative to the first feedback info bit (bit 68 of the header).
struct feedback_info_fields_t : feedback_info_t {
    uint32_t HARQ_Process_number = 0;                       // f1 f5 f6; f3: process 0
    transmission_feedback_t Transmission_feedback = transmission_feedback_t::not_defined;  // f1 f5; f3: process 0
    uint32_t HARQ_Process_number_1 = 0;                     // f3
    transmission_feedback_t Transmission_feedback_1 = transmission_feedback_t::not_defined;  // f3
    uint32_t HARQ_feedback_bitmap = 0;                      // f4
    uint32_t Codebook_index = 0;                            // f2 (3 bits) f5 (6 bits)
    mimo_feedback_t MIMO_feedback = mimo_feedback_t::not_defined;  // f2 (1 bit) f5 (2 bits)
    uint32_t Reserved = 0;                                  // f6
    uint32_t Buffer_Size = 0;                               // f1 f2 f6, in bytes (coded as buffer status)
    uint32_t MCS = MCS_out_of_range;                        // f1 f2 f3 f4 f6 (coded as CQI = MCS + 1)

    void zero() { *this = feedback_info_fields_t{}; }
};

class feedback_info_pool_t {
    public:
        /// one record per format like the reference's pool members feedback_info_f1 ... f6
        feedback_info_fields_t feedback_info_f1, feedback_info_f2, feedback_info_f3, feedback_info_f4, feedback_info_f5,
            feedback_info_f6;

        feedback_info_fields_t& by_format(uint32_t f) {
            feedback_info_fields_t* const t[6] = {&feedback_info_f1, &feedback_info_f2, &feedback_info_f3,
                                                  &feedback_info_f4, &feedback_info_f5, &feedback_info_f6};
            return *t[f - 1];
        }
        const feedback_info_fields_t& by_format(uint32_t f) const {
            return const_cast<feedback_info_pool_t*>(this)->by_format(f);
        }

        static bool is_valid(uint32_t format, const feedback_info_fields_t& v) {
            using tf = feedback_info_t::transmission_feedback_t;
            using mf = feedback_info_t::mimo_feedback_t;
            const bool mcs_ok = v.MCS <= 11;
            switch (format) {
                case 1: return v.HARQ_Process_number <= 7 && v.Transmission_feedback != tf::not_defined && mcs_ok;
                case 2: return v.Codebook_index <= 7 && (v.MIMO_feedback == mf::single_layer || v.MIMO_feedback == mf::dual_layer) && mcs_ok;
                case 3: return v.HARQ_Process_number <= 7 && v.Transmission_feedback != tf::not_defined &&
                               v.HARQ_Process_number_1 <= 7 && v.Transmission_feedback_1 != tf::not_defined && mcs_ok;
                case 4: return v.HARQ_feedback_bitmap <= 0xff && mcs_ok;
                case 5: return v.HARQ_Process_number <= 7 && v.Transmission_feedback != tf::not_defined &&
                               v.MIMO_feedback != mf::not_defined && v.Codebook_index <= 63;
                case 6: return v.HARQ_Process_number <= 7 && v.Reserved == 0 && mcs_ok;
                default: return false;
            }
        }

        /// a_ptr = byte 8 of the header: its upper nibble (FeedbackFormat) is kept (feedback_info.cpp:379-411)
        void pack(uint32_t format, uint8_t* a_ptr) const {
            using d = feedback_info_t;
            detail::put_bits(a_ptr, 4, 12, 0);
            if (format == d::No_feedback) return;
            const feedback_info_fields_t& v = by_format(format);
            const uint32_t tf0 = (uint32_t)v.Transmission_feedback & 1u, tf1 = (uint32_t)v.Transmission_feedback_1 & 1u;
            const uint32_t bs = d::buffer_size_2_buffer_status(v.Buffer_Size), cqi = d::mcs_2_cqi(v.MCS);
            uint32_t w = 0;  // the 12 bits
            switch (format) {
                case 1: w = v.HARQ_Process_number << 9 | tf0 << 8 | bs << 4 | cqi; break;
                case 2: w = v.Codebook_index << 9 | ((uint32_t)v.MIMO_feedback & 1u) << 8 | bs << 4 | cqi; break;
                case 3: w = v.HARQ_Process_number << 9 | tf0 << 8 | v.HARQ_Process_number_1 << 5 | tf1 << 4 | cqi; break;
                case 4: w = v.HARQ_feedback_bitmap << 4 | cqi; break;
                case 5: w = v.HARQ_Process_number << 9 | tf0 << 8 | ((uint32_t)v.MIMO_feedback & 3u) << 6 | v.Codebook_index; break;
                case 6: w = v.HARQ_Process_number << 9 | v.Reserved << 8 | bs << 4 | cqi; break;
                default: break;
            }
            detail::put_bits(a_ptr, 4, 12, w & 0xfffu);
        }

        /// false for an unknown format, non-zero bits behind "no feedback" or an invalid field (feedback_info.cpp:413-425)
        bool unpack(uint32_t format, const uint8_t* a_ptr) {
            using d = feedback_info_t;
            const uint32_t w = detail::get_bits(a_ptr, 4, 12);
            if (format == d::No_feedback) return w == 0;
            if (format > 6) return false;
            feedback_info_fields_t& v = by_format(format);
            v.zero();
            const uint32_t top3 = w >> 9, b8 = (w >> 8) & 1u, bs = (w >> 4) & 0xfu, cqi = w & 0xfu;
            const uint32_t size = d::buffer_status_2_buffer_size_lower[bs];
            switch (format) {
                case 1:
                    v.HARQ_Process_number = top3; v.Transmission_feedback = (d::transmission_feedback_t)b8;
                    v.Buffer_Size = size; v.MCS = d::cqi_2_mcs(cqi);
                    break;
                case 2:
                    v.Codebook_index = top3; v.MIMO_feedback = (d::mimo_feedback_t)b8;
                    v.Buffer_Size = size; v.MCS = d::cqi_2_mcs(cqi);
                    break;
                case 3:
                    v.HARQ_Process_number = top3; v.Transmission_feedback = (d::transmission_feedback_t)b8;
                    v.HARQ_Process_number_1 = (w >> 5) & 7u;
                    v.Transmission_feedback_1 = (d::transmission_feedback_t)((w >> 4) & 1u);
                    v.MCS = d::cqi_2_mcs(cqi);
                    break;
                case 4:
                    v.HARQ_feedback_bitmap = w >> 4; v.MCS = d::cqi_2_mcs(cqi);
                    break;
                case 5:
                    v.HARQ_Process_number = top3; v.Transmission_feedback = (d::transmission_feedback_t)b8;
                    v.MIMO_feedback = (d::mimo_feedback_t)((w >> 6) & 3u); v.Codebook_index = w & 63u;
                    break;
                case 6:
                    v.HARQ_Process_number = top3; v.Reserved = b8; v.Buffer_Size = size; v.MCS = d::cqi_2_mcs(cqi);
                    break;
            }
            return is_valid(format, v);
        }
};

// ---------------------------------------------------------------------------------------------------------------
// headers
class plcf_base_t {
    public:
        virtual ~plcf_base_t() = default;
        virtual void zero() = 0;
        virtual bool is_valid() const = 0;
        virtual uint32_t get_packed_size() const = 0;  // bytes: 5 (type 1) or 10 (type 2)
        virtual void pack(uint8_t* plcf_front) const = 0;
        virtual bool unpack(const uint8_t* plcf_front) = 0;

        uint32_t HeaderFormat = 0;
        uint32_t PacketLengthType = 0;  // 0: subslots, 1: slots
        uint32_t PacketLength_m1 = 0;
        uint32_t ShortNetworkID = 0;
        uint32_t TransmitterIdentity = 0;
        uint32_t TransmitPower = 0;
        uint32_t DFMCS = 0;

        void set_PacketLength_m1(int32_t PacketLength) { PacketLength_m1 = (uint32_t)(PacketLength - 1); }
        uint32_t get_PacketLength() const { return PacketLength_m1 + 1; }

        /// Table 6.2.1-3a; smallest code whose power is >= the request, the last code beyond the table (plcf_base.cpp:70-76)
        static constexpr std::array<int32_t, 16> tx_power_table = {-40, -30, -20, -16, -12, -8, -4, 0,
                                                                   4,   7,   10,  13,  16,  19, 21, 23};
        void set_TransmitPower(int32_t TransmitPower_dBm) {
            uint32_t i = 0;
            while (i + 1 < tx_power_table.size() && tx_power_table[i] < TransmitPower_dBm) ++i;
            TransmitPower = i;
        }
        int32_t get_TransmitPower_dBm() const { return tx_power_table.at(TransmitPower); }

        virtual uint32_t get_Type() const = 0;
        virtual uint32_t get_HeaderFormat() const = 0;
        virtual uint32_t get_N_SS() const = 0;
        virtual uint32_t get_DFRedundancyVersion() const = 0;

    protected:
        /// identity.cpp:55-87: 0 is reserved, 8 / 16 bit ranges
        static bool id_ok(uint32_t v, uint32_t bits) { return v != 0 && (v >> bits) == 0; }
        bool common_valid(uint32_t header_format, uint32_t mcs_max) const {
            return HeaderFormat == header_format && PacketLengthType <= 1 && PacketLength_m1 <= 15 &&
                   id_ok(ShortNetworkID, 8) && id_ok(TransmitterIdentity, 16) && TransmitPower <= 15 && DFMCS <= mcs_max;
        }
        static constexpr std::array<uint32_t, 9> N_SS_coded_lut = {0, 0, 1, 0, 2, 0, 0, 0, 3};
        static constexpr std::array<uint32_t, 4> N_SS_coded_lut_rev = {1, 2, 4, 8};
};

/// type 1, header format 0 (40 bits, Table 6.2.1-1), plcf_10.cpp
class plcf_10_t final : public plcf_base_t {
    public:
        uint32_t Reserved = 0;

        void zero() override { *this = plcf_10_t{}; }
        bool is_valid() const override { return common_valid(0, 7) && Reserved == 0; }  // DFMCS has three bits
        uint32_t get_packed_size() const override { return 5; }
        void pack(uint8_t* a) const override { assert(is_valid()); detail::pack_fields(*this, fields(), a); }
        bool unpack(const uint8_t* a) override { detail::unpack_fields(*this, fields(), a); return is_valid(); }
        uint32_t get_Type() const override { return 1; }
        uint32_t get_HeaderFormat() const override { return 0; }
        uint32_t get_N_SS() const override { return 1; }
        uint32_t get_DFRedundancyVersion() const override { return 0; }

    private:
        using F = detail::field_t<plcf_10_t>;
        static const std::array<F, 8>& fields() {
            static const std::array<F, 8> f = {{{&plcf_10_t::HeaderFormat, 0, 3}, {&plcf_10_t::PacketLengthType, 3, 1},
                                                {&plcf_10_t::PacketLength_m1, 4, 4}, {&plcf_10_t::ShortNetworkID, 8, 8},
                                                {&plcf_10_t::TransmitterIdentity, 16, 16}, {&plcf_10_t::TransmitPower, 32, 4},
                                                {&plcf_10_t::Reserved, 36, 1}, {&plcf_10_t::DFMCS, 37, 3}}};
            return f;
        }
};

/// type 2, header format 0 (80 bits, Table 6.2.1-2): HARQ fields + feedback, plcf_20.cpp
class plcf_20_t final : public plcf_base_t {
    public:
        uint32_t ReceiverIdentity = 0;
        uint32_t NumberOfSpatialStreams = 0;
        uint32_t DFRedundancyVersion = 0;
        uint32_t DFNewDataIndication = 0;
        uint32_t DFHARQProcessNumber = 0;
        uint32_t FeedbackFormat = 0;
        feedback_info_pool_t feedback_info_pool;

        void zero() override { *this = plcf_20_t{}; }
        bool is_valid() const override {  // DFMCS up to 11 = 1024-QAM R=5/6
            return common_valid(0, 11) && id_ok(ReceiverIdentity, 16) && detail::fields_fit(*this, fields());
        }
        uint32_t get_packed_size() const override { return 10; }
        void pack(uint8_t* a) const override {
            assert(is_valid());
            detail::pack_fields(*this, fields(), a);
            feedback_info_pool.pack(FeedbackFormat, a + 8);
        }
        bool unpack(const uint8_t* a) override {
            detail::unpack_fields(*this, fields(), a);
            if (!feedback_info_pool.unpack(FeedbackFormat, a + 8)) return false;
            return is_valid();
        }
        void set_NumberOfSpatialStreams(int32_t N_SS) { NumberOfSpatialStreams = N_SS_coded_lut.at((size_t)N_SS); }
        uint32_t get_Type() const override { return 2; }
        uint32_t get_HeaderFormat() const override { return 0; }
        uint32_t get_N_SS() const override { return N_SS_coded_lut_rev.at(NumberOfSpatialStreams); }
        uint32_t get_DFRedundancyVersion() const override { return DFRedundancyVersion; }

    private:
        using F = detail::field_t<plcf_20_t>;
        static const std::array<F, 13>& fields() {
            static const std::array<F, 13> f = {
                {{&plcf_20_t::HeaderFormat, 0, 3}, {&plcf_20_t::PacketLengthType, 3, 1}, {&plcf_20_t::PacketLength_m1, 4, 4},
                 {&plcf_20_t::ShortNetworkID, 8, 8}, {&plcf_20_t::TransmitterIdentity, 16, 16},
                 {&plcf_20_t::TransmitPower, 32, 4}, {&plcf_20_t::DFMCS, 36, 4}, {&plcf_20_t::ReceiverIdentity, 40, 16},
                 {&plcf_20_t::NumberOfSpatialStreams, 56, 2}, {&plcf_20_t::DFRedundancyVersion, 58, 2},
                 {&plcf_20_t::DFNewDataIndication, 60, 1}, {&plcf_20_t::DFHARQProcessNumber, 61, 3},
                 {&plcf_20_t::FeedbackFormat, 64, 4}}};
            return f;
        }
};

/// type 2, header format 1 (80 bits, Table 6.2.1-2a): no HARQ fields, plcf_21.cpp
class plcf_21_t final : public plcf_base_t {
    public:
        uint32_t ReceiverIdentity = 0;
        uint32_t NumberOfSpatialStreams = 0;
        uint32_t Reserved = 0;
        uint32_t FeedbackFormat = 0;
        feedback_info_pool_t feedback_info_pool;

        void zero() override { *this = plcf_21_t{}; HeaderFormat = 0; }
        bool is_valid() const override {
            return common_valid(1, 11) && id_ok(ReceiverIdentity, 16) && Reserved == 0 && detail::fields_fit(*this, fields());
        }
        uint32_t get_packed_size() const override { return 10; }
        void pack(uint8_t* a) const override {
            assert(is_valid());
            detail::pack_fields(*this, fields(), a);
            feedback_info_pool.pack(FeedbackFormat, a + 8);
        }
        bool unpack(const uint8_t* a) override {
            detail::unpack_fields(*this, fields(), a);
            if (!feedback_info_pool.unpack(FeedbackFormat, a + 8)) return false;
            return is_valid();
        }
        void set_NumberOfSpatialStreams(int32_t N_SS) { NumberOfSpatialStreams = N_SS_coded_lut.at((size_t)N_SS); }
        uint32_t get_Type() const override { return 2; }
        uint32_t get_HeaderFormat() const override { return 1; }
        uint32_t get_N_SS() const override { return N_SS_coded_lut_rev.at(NumberOfSpatialStreams); }
        uint32_t get_DFRedundancyVersion() const override { return 0; }

    private:
        using F = detail::field_t<plcf_21_t>;
        static const std::array<F, 11>& fields() {
            static const std::array<F, 11> f = {
                {{&plcf_21_t::HeaderFormat, 0, 3}, {&plcf_21_t::PacketLengthType, 3, 1}, {&plcf_21_t::PacketLength_m1, 4, 4},
                 {&plcf_21_t::ShortNetworkID, 8, 8}, {&plcf_21_t::TransmitterIdentity, 16, 16},
                 {&plcf_21_t::TransmitPower, 32, 4}, {&plcf_21_t::DFMCS, 36, 4}, {&plcf_21_t::ReceiverIdentity, 40, 16},
                 {&plcf_21_t::NumberOfSpatialStreams, 56, 2}, {&plcf_21_t::Reserved, 58, 6},
                 {&plcf_21_t::FeedbackFormat, 64, 4}}};
            return f;
        }
};

// ---------------------------------------------------------------------------------------------------------------
/// plcf_decoder.cpp:44-150: fed once per packet and PLCF type with the bytes whose CRC passed; keeps the header of the
/// decoded format and rejects fields beyond the radio device class
class plcf_decoder_t {
    public:
        plcf_decoder_t(uint32_t PacketLength_max_, uint32_t mcs_index_max_, uint32_t N_SS_max_)
            : PacketLength_max(PacketLength_max_), mcs_index_max(mcs_index_max_), N_SS_max(N_SS_max_) {}

        /// before every packet: both types undefined
        void set_configuration() { HeaderFormat_type1 = HeaderFormat_type2 = UNDEFINED_NUMERIC_32; }

        void decode_and_rdc_check(uint32_t PLCF_type, const uint8_t* plcf_front) {
            const uint32_t candidate = detail::get_bits(plcf_front, 0, 3);
            plcf_base_t* h = nullptr;
            if (PLCF_type == 1 && candidate == 0) h = &plcf_10;
            if (PLCF_type == 2 && candidate == 0) h = &plcf_20;
            if (PLCF_type == 2 && candidate == 1) h = &plcf_21;
            if (h == nullptr || !h->unpack(plcf_front)) return;   // format not implemented / field invalid
            if (h->get_PacketLength() > PacketLength_max || h->DFMCS > mcs_index_max || h->get_N_SS() > N_SS_max) return;
            (PLCF_type == 1 ? HeaderFormat_type1 : HeaderFormat_type2) = candidate;
        }

        /// 0: none, 1: type 1, 2: type 2, 3: both
        uint32_t has_any_plcf() const {
            return (HeaderFormat_type1 != UNDEFINED_NUMERIC_32 ? 1u : 0u) + (HeaderFormat_type2 != UNDEFINED_NUMERIC_32 ? 2u : 0u);
        }

        /// header of the decoded format (cast by get_HeaderFormat()), nullptr if that type was not decoded
        const plcf_base_t* get_plcf_base(uint32_t PLCF_type) const {
            if (PLCF_type == 1 && HeaderFormat_type1 == 0) return &plcf_10;
            if (PLCF_type == 2 && HeaderFormat_type2 == 0) return &plcf_20;
            if (PLCF_type == 2 && HeaderFormat_type2 == 1) return &plcf_21;
            return nullptr;
        }

    private:
        const uint32_t PacketLength_max, mcs_index_max, N_SS_max;
        uint32_t HeaderFormat_type1 = UNDEFINED_NUMERIC_32, HeaderFormat_type2 = UNDEFINED_NUMERIC_32;
        plcf_10_t plcf_10;
        plcf_20_t plcf_20;
        plcf_21_t plcf_21;
};

}  // namespace dectnrp_b200::sp4

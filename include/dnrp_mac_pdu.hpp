// dnrp_mac_pdu.hpp -- MAC PDU framing at the PHY-to-MAC boundary (header-only C++17, host side): what firmware reads out
// of phy::pdc_report_t through the reference's mac_pdu_decoder_t, down to the demultiplexed MAC SDUs.
//
// Mirrors (names and byte layouts):
//   sp4::mac_header_type_t                      sections_part4/mac_pdu/mac_header_type.{hpp,cpp}          (clause 6.3.2)
//   sp4::data_mac_pdu_header_t, beacon_header_t, unicast_header_t, rd_broadcasting_header_t, mch_empty_t
//                                               sections_part4/mac_pdu/mac_common_header.{hpp,cpp}        (clause 6.3.3)
//   sp4::mac_multiplexing_header_t              sections_part4/mac_pdu/mac_multiplexing_header.{hpp,cpp}  (clause 6.3.4)
//   sp4::mac_pdu_decoder_t (set_configuration / decode / has_reached_valid_final_state)
//                                               sections_part4/mac_pdu/mac_pdu_decoder.cpp:58-318
// Scope: framing and demultiplexing.  SDUs whose size is carried by the multiplexing header (8 / 16 bit length field,
// one-bit length) are delivered as (IE type, pointer, length) -- user plane data and higher layer signalling, the
// "flowing" IEs of the reference, are exactly that.  MAC messages without a length field size themselves by their
// content (mmie_packing_peeking_t); their parsers belong to the MAC layer and are not part of this path: the decoder
// stops in front of the first one and reports where (unparsed_offset), like the reference stops at an IE it has no
// pool entry for.  A padding IE ends the PDU (clause 6.4.3.8).
#pragma once
#include <cstdint>
#include <cstring>
#include <vector>

#include "dnrp_plcf.hpp"

namespace dectnrp_b200::sp4 {

struct mac_header_type_t {
    enum class version_ec : uint32_t { v00 = 0, not_defined = UNDEFINED_NUMERIC_32 };
    enum class mac_security_ec : uint32_t { macsec_not_used = 0, macsec_used_no_ie = 1, macsec_used_with_ie = 2, reserved = 3,
                                            not_defined = UNDEFINED_NUMERIC_32 };
    enum class mac_header_type_ec : uint32_t { DATA_MAC_PDU = 0, Beacon = 1, Unicast = 2, RD_Broadcasting = 3, mch_empty = 4,
                                               Escape = 15, not_defined = UNDEFINED_NUMERIC_32 };
    version_ec Version = version_ec::not_defined;
    mac_security_ec MAC_security = mac_security_ec::not_defined;
    mac_header_type_ec MAC_header_type = mac_header_type_ec::not_defined;

    void zero() { *this = mac_header_type_t{}; }
    bool is_valid() const {
        return Version != version_ec::not_defined && MAC_security != mac_security_ec::not_defined &&
               MAC_header_type != mac_header_type_ec::not_defined;
    }
    static constexpr uint32_t get_packed_size() { return 1; }
    void pack(uint8_t* a) const { a[0] = (uint8_t)((uint32_t)Version << 6 | (uint32_t)MAC_security << 4 | (uint32_t)MAC_header_type); }
    bool unpack(const uint8_t* a) {   // version 0 only, header types 0..4 (mac_header_type.hpp:64-78)
        const uint32_t v = a[0] >> 6, s = (a[0] >> 4) & 3, t = a[0] & 15;
        Version = v == 0 ? version_ec::v00 : version_ec::not_defined;
        MAC_security = (mac_security_ec)s;
        MAC_header_type = t <= 4 ? (mac_header_type_ec)t : mac_header_type_ec::not_defined;
        return is_valid();
    }
};

/// the four common headers share one record; members a header does not carry stay zero
struct mac_common_header_t {
    mac_header_type_t::mac_header_type_ec kind = mac_header_type_t::mac_header_type_ec::not_defined;
    uint32_t Reserved = 0, Reset = 0, Sequence_number = 0;       // DATA, Unicast, RD broadcasting
    uint32_t Network_ID_3_lsb = 0;                                 // Beacon
    uint32_t Receiver_Address = 0, Transmitter_Address = 0;        // long radio device ids

    static uint32_t packed_size_of(mac_header_type_t::mac_header_type_ec k) {
        using e = mac_header_type_t::mac_header_type_ec;
        return k == e::DATA_MAC_PDU ? 2 : k == e::Beacon ? 7 : k == e::Unicast ? 10 : k == e::RD_Broadcasting ? 6 : 0;
    }
    uint32_t get_packed_size() const { return packed_size_of(kind); }
    void set_Network_ID_3_lsb(uint32_t Network_ID) { Network_ID_3_lsb = Network_ID & 0xffffffu; }
    bool is_valid() const {
        using e = mac_header_type_t::mac_header_type_ec;
        const bool seq_ok = Reserved == 0 && Reset <= 1 && Sequence_number <= 0xfff;
        switch (kind) {
            case e::DATA_MAC_PDU: return seq_ok;
            case e::Beacon: return Network_ID_3_lsb <= 0xffffffu && Transmitter_Address != 0;
            case e::Unicast: return seq_ok && Transmitter_Address != 0 && Receiver_Address != 0;
            case e::RD_Broadcasting: return seq_ok && Transmitter_Address != 0;
            case e::mch_empty: return true;
            default: return false;
        }
    }
    void pack(uint8_t* a) const {
        using e = mac_header_type_t::mac_header_type_ec;
        auto be = [](uint8_t* p, uint32_t v, int n) { for (int i = 0; i < n; ++i) p[i] = (uint8_t)(v >> (8 * (n - 1 - i))); };
        if (kind == e::Beacon) { be(a, Network_ID_3_lsb, 3); be(a + 3, Transmitter_Address, 4); return; }
        if (kind == e::mch_empty) return;
        a[0] = (uint8_t)(Reserved << 5 | Reset << 4 | Sequence_number >> 8);
        a[1] = (uint8_t)(Sequence_number & 0xff);
        if (kind == e::Unicast) { be(a + 2, Receiver_Address, 4); be(a + 6, Transmitter_Address, 4); }
        if (kind == e::RD_Broadcasting) be(a + 2, Transmitter_Address, 4);
    }
    bool unpack(const uint8_t* a) {
        using e = mac_header_type_t::mac_header_type_ec;
        auto be = [](const uint8_t* p, int n) { uint32_t v = 0; for (int i = 0; i < n; ++i) v = v << 8 | p[i]; return v; };
        if (kind == e::Beacon) { Network_ID_3_lsb = be(a, 3); Transmitter_Address = be(a + 3, 4); return is_valid(); }
        if (kind == e::mch_empty) return true;
        Reserved = a[0] >> 5; Reset = (a[0] >> 4) & 1; Sequence_number = (uint32_t)(a[0] & 15) << 8 | a[1];
        if (kind == e::Unicast) { Receiver_Address = be(a + 2, 4); Transmitter_Address = be(a + 6, 4); }
        if (kind == e::RD_Broadcasting) Transmitter_Address = be(a + 2, 4);
        return is_valid();
    }
};

struct mac_multiplexing_header_t {
    enum class mac_ext_t : uint32_t { No_Length_Field = 0, Length_Field_8_Bit = 1, Length_Field_16_Bit = 2, Length_Field_1_Bit = 3,
                                      not_defined = UNDEFINED_NUMERIC_32 };
    /// IE type codes for MAC_Ext 00 / 01 / 10 (Table 6.3.4-2), mac_multiplexing_header.hpp:43-82
    enum ie_type_mac_ext_00_01_10_t : uint32_t {
        Padding_IE = 0, Higher_Layer_Signalling_Flow_1 = 1, Higher_Layer_Signalling_Flow_2 = 2, User_Plane_Data_Flow_1 = 3,
        User_Plane_Data_Flow_2 = 4, User_Plane_Data_Flow_3 = 5, User_Plane_Data_Flow_4 = 6, Network_Beacon_Message = 8,
        Cluster_Beacon_Message = 9, Association_Request_Message = 10, Association_Response_Message = 11,
        Association_Release_Message = 12, Reconfiguration_Request_Message = 13, Reconfiguration_Response_Message = 14,
        Additional_MAC_Messages = 15, Security_Info_IE = 16, Route_Info_IE = 17, Resource_Allocation_IE = 18,
        Random_Access_Resource_IE = 19, RD_Capability_IE = 20, Neighbouring_IE = 21, Broadcast_Indication_IE = 22,
        Group_Assignment_IE = 23, Load_Info_IE = 24, Measurement_Report_IE = 25, Power_Target_IE = 29, Time_Announce_IE = 30,
        Escape = 62, IE_Type_Extension = 63};
    static constexpr uint32_t packed_size_min_to_peek = 1;

    mac_ext_t mac_ext = mac_ext_t::not_defined;
    uint32_t ie_type = UNDEFINED_NUMERIC_32;   // 6 bits (MAC_Ext 00 / 01 / 10) or 5 bits (MAC_Ext 11)
    uint32_t length = UNDEFINED_NUMERIC_32;    // bytes of the SDU; undefined without a length field

    void zero() { *this = mac_multiplexing_header_t{}; }
    uint32_t get_packed_size() const {
        return mac_ext == mac_ext_t::Length_Field_8_Bit ? 2 : mac_ext == mac_ext_t::Length_Field_16_Bit ? 3 : 1;
    }
    bool is_valid() const {
        switch (mac_ext) {
            case mac_ext_t::Length_Field_1_Bit: return length <= 1 && ie_type <= 31;
            case mac_ext_t::Length_Field_8_Bit: return length <= 0xff && ie_type <= 63;
            case mac_ext_t::Length_Field_16_Bit: return length <= 0xffff && ie_type <= 63;
            case mac_ext_t::No_Length_Field: return ie_type <= 63;
            default: return false;
        }
    }
    void pack(uint8_t* a) const {
        a[0] = (uint8_t)((uint32_t)mac_ext << 6);
        switch (mac_ext) {
            case mac_ext_t::Length_Field_1_Bit: a[0] |= (uint8_t)(length << 5 | ie_type); break;
            case mac_ext_t::Length_Field_8_Bit: a[0] |= (uint8_t)ie_type; a[1] = (uint8_t)length; break;
            case mac_ext_t::Length_Field_16_Bit: a[0] |= (uint8_t)ie_type; a[1] = (uint8_t)(length >> 8); a[2] = (uint8_t)length; break;
            default: a[0] |= (uint8_t)ie_type; break;
        }
    }
    /// first byte: MAC extension and IE type (mac_multiplexing_header.cpp: unpack_mac_ext_ie_type)
    void unpack_mac_ext_ie_type(const uint8_t* a) {
        mac_ext = (mac_ext_t)(a[0] >> 6);
        ie_type = mac_ext == mac_ext_t::Length_Field_1_Bit ? (a[0] & 31u) : (a[0] & 63u);
    }
    /// needs get_packed_size() bytes (unpack_length)
    void unpack_length(const uint8_t* a) {
        switch (mac_ext) {
            case mac_ext_t::Length_Field_8_Bit: length = a[1]; break;
            case mac_ext_t::Length_Field_16_Bit: length = (uint32_t)a[1] << 8 | a[2]; break;
            case mac_ext_t::Length_Field_1_Bit: length = (a[0] >> 5) & 1; break;
            default: length = UNDEFINED_NUMERIC_32; break;
        }
    }
    bool is_padding() const { return ie_type == 0; }
};

/// one demultiplexed SDU: its multiplexing header and where its bytes are inside the transport block
struct mmie_ref_t {
    mac_multiplexing_header_t mac_multiplexing_header;
    const uint8_t* data_ptr = nullptr;
    uint32_t N_bytes = 0;
    bool is_user_plane_data() const {
        using h = mac_multiplexing_header_t;
        return mac_multiplexing_header.mac_ext != h::mac_ext_t::Length_Field_1_Bit &&
               mac_multiplexing_header.ie_type >= h::User_Plane_Data_Flow_1 && mac_multiplexing_header.ie_type <= h::User_Plane_Data_Flow_4;
    }
    bool is_higher_layer_signalling() const {
        using h = mac_multiplexing_header_t;
        return mac_multiplexing_header.mac_ext != h::mac_ext_t::Length_Field_1_Bit &&
               (mac_multiplexing_header.ie_type == h::Higher_Layer_Signalling_Flow_1 ||
                mac_multiplexing_header.ie_type == h::Higher_Layer_Signalling_Flow_2);
    }
};

class mac_pdu_decoder_t {
    public:
        /// before every transport block (mac_pdu_decoder.cpp:58-83); mu kept for parity with the reference's signature
        void set_configuration(const uint8_t* a_, uint32_t a_cnt_w_tb_, uint32_t mu_ = 1) {
            a = a_; a_cnt_w_tb = a_cnt_w_tb_; mu = mu_;
            a_cnt_r = 0; done = aborted = false; unparsed_offset = UNDEFINED_NUMERIC_32;
            mht.zero(); mch = mac_common_header_t{}; has_mht = has_mch = false;
            mmie_decoded_vec.clear();
        }
        /// a_cnt_w = bytes of the transport block available so far (the reference decodes code block by code block;
        /// the batched receiver delivers whole transport blocks: a_cnt_w = N_TB_byte)
        void decode(uint32_t a_cnt_w) {
            if (done || aborted) return;
            auto abort = [&]() { aborted = true; a_cnt_r = a_cnt_w; };
            auto have = [&](uint32_t n) { return a_cnt_r + n <= a_cnt_w_tb && a_cnt_r + n <= a_cnt_w; };
            if (!has_mht) {
                if (!have(1)) { if (a_cnt_r + 1 > a_cnt_w_tb) abort(); return; }
                if (!mht.unpack(a + a_cnt_r)) { abort(); return; }
                has_mht = true; a_cnt_r += 1;
                mch.kind = mht.MAC_header_type;
            }
            if (!has_mch) {
                const uint32_t n = mch.get_packed_size();
                if (!have(n)) { if (a_cnt_r + n > a_cnt_w_tb) abort(); return; }
                if (!mch.unpack(a + a_cnt_r)) { abort(); return; }
                has_mch = true; a_cnt_r += n;
            }
            while (a_cnt_r < a_cnt_w_tb) {
                if (!have(1)) return;
                mac_multiplexing_header_t mmh;
                mmh.unpack_mac_ext_ie_type(a + a_cnt_r);
                if (mmh.ie_type == 0) { abort(); return; }                       // padding: the rest is padding (6.4.3.8)
                if (mmh.mac_ext == mac_multiplexing_header_t::mac_ext_t::No_Length_Field) {
                    unparsed_offset = a_cnt_r;                                    // self-sizing MAC message: MAC layer's job
                    abort();
                    return;
                }
                const uint32_t nh = mmh.get_packed_size();
                if (!have(nh)) { if (a_cnt_r + nh > a_cnt_w_tb) abort(); return; }
                mmh.unpack_length(a + a_cnt_r);
                if (!have(nh + mmh.length)) { if (a_cnt_r + nh + mmh.length > a_cnt_w_tb) abort(); return; }
                mmie_ref_t r;
                r.mac_multiplexing_header = mmh;
                r.data_ptr = a + a_cnt_r + nh;
                r.N_bytes = mmh.length;
                mmie_decoded_vec.push_back(r);
                a_cnt_r += nh + mmh.length;
            }
            done = true;
        }
        bool has_reached_valid_final_state() const { return a_cnt_r == a_cnt_w_tb && (done || aborted); }
        const mac_header_type_t* get_mac_header_type() const { return has_mht ? &mht : nullptr; }
        const mac_common_header_t* get_mac_common_header() const { return has_mch ? &mch : nullptr; }
        const std::vector<mmie_ref_t>& get_mmie_decoded_vec() const { return mmie_decoded_vec; }
        /// offset of the first MAC message this decoder left to the MAC layer, UNDEFINED_NUMERIC_32 if none
        uint32_t get_unparsed_offset() const { return unparsed_offset; }

    private:
        const uint8_t* a = nullptr;
        uint32_t a_cnt_w_tb = 0, a_cnt_r = 0, mu = 1, unparsed_offset = UNDEFINED_NUMERIC_32;
        bool done = false, aborted = false, has_mht = false, has_mch = false;
        mac_header_type_t mht;
        mac_common_header_t mch;
        std::vector<mmie_ref_t> mmie_decoded_vec;
};

/// builds a MAC PDU in a transport block: header type + common header, SDUs with the shortest length field that fits,
/// padding to the end (the writer side of tfw_p2p-style firmware; mirrors the pack() calls of the reference's headers)
class mac_pdu_writer_t {
    public:
        mac_pdu_writer_t(uint8_t* a_, uint32_t N_TB_byte_) : a(a_), N(N_TB_byte_) {}
        bool begin(const mac_header_type_t& mht, const mac_common_header_t& mch) {
            if (!mht.is_valid() || mch.kind != mht.MAC_header_type || !mch.is_valid() || 1 + mch.get_packed_size() > N) return false;
            mht.pack(a);
            mch.pack(a + 1);
            w = 1 + mch.get_packed_size();
            return true;
        }
        /// SDU with an 8 or 16 bit length field (user plane data, higher layer signalling ...)
        bool add_sdu(uint32_t ie_type, const uint8_t* data, uint32_t n) {
            mac_multiplexing_header_t h;
            h.mac_ext = n <= 0xff ? mac_multiplexing_header_t::mac_ext_t::Length_Field_8_Bit
                                  : mac_multiplexing_header_t::mac_ext_t::Length_Field_16_Bit;
            h.ie_type = ie_type; h.length = n;
            if (!h.is_valid() || ie_type == 0 || w + h.get_packed_size() + n > N) return false;
            h.pack(a + w);
            std::memcpy(a + w + h.get_packed_size(), data, n);
            w += h.get_packed_size() + n;
            return true;
        }
        /// padding IE over the rest (padding_ie_t: one byte with MAC_Ext 11 / length 0, two with length 1, else 8 bit length)
        void finish() {
            const uint32_t rest = N - w;
            if (rest == 0) return;
            std::memset(a + w, 0, rest);
            mac_multiplexing_header_t h;
            h.ie_type = 0;
            if (rest == 1) { h.mac_ext = mac_multiplexing_header_t::mac_ext_t::Length_Field_1_Bit; h.length = 0; }
            else if (rest == 2) { h.mac_ext = mac_multiplexing_header_t::mac_ext_t::Length_Field_1_Bit; h.length = 1; }
            else { h.mac_ext = mac_multiplexing_header_t::mac_ext_t::Length_Field_8_Bit; h.length = rest - 2 > 0xff ? 0xff : rest - 2; }
            h.pack(a + w);
            w = N;
        }
        uint32_t bytes_written() const { return w; }

    private:
        uint8_t* a;
        uint32_t N, w = 0;
};

}  // namespace dectnrp_b200::sp4

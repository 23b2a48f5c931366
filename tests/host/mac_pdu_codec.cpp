// mac_pdu_codec.cpp -- line-driven test driver for include/dnrp_mac_pdu.hpp (tests/test_mac_pdu_cpu.py compares with
// oracle/mac_pdu.py).
//   B <N_TB> <type> <sec> <reset> <seq> <rx> <tx> <nid> <k> {<ie> <hexdata>}*k    -> hex of the built transport block
//   D <hex>                                                                       -> decoded summary
#include <cstdio>
#include <iostream>
#include <sstream>
#include <string>
#include <vector>
#include "dnrp_mac_pdu.hpp"

using namespace dectnrp_b200::sp4;

static std::vector<uint8_t> unhex(const std::string& h) {
    std::vector<uint8_t> v;
    for (size_t i = 0; i + 1 < h.size(); i += 2) v.push_back((uint8_t)std::stoul(h.substr(i, 2), nullptr, 16));
    return v;
}

int main() {
    std::string line;
    while (std::getline(std::cin, line)) {
        std::istringstream in(line);
        std::string cmd;
        in >> cmd;
        if (cmd == "B") {
            uint32_t N, type, sec, k;
            mac_common_header_t mch;
            in >> N >> type >> sec >> mch.Reset >> mch.Sequence_number >> mch.Receiver_Address >> mch.Transmitter_Address >>
                mch.Network_ID_3_lsb >> k;
            mac_header_type_t mht;
            mht.Version = mac_header_type_t::version_ec::v00;
            mht.MAC_security = (mac_header_type_t::mac_security_ec)sec;
            mht.MAC_header_type = (mac_header_type_t::mac_header_type_ec)type;
            mch.kind = mht.MAC_header_type;
            std::vector<uint8_t> a(N, 0xEE);
            mac_pdu_writer_t w(a.data(), N);
            bool ok = w.begin(mht, mch);
            for (uint32_t i = 0; ok && i < k; ++i) {
                uint32_t ie;
                std::string h;
                in >> ie >> h;
                const std::vector<uint8_t> d = unhex(h == "-" ? "" : h);
                ok = w.add_sdu(ie, d.data(), (uint32_t)d.size());
            }
            if (!ok) { std::printf("INVALID\n"); continue; }
            w.finish();
            for (uint8_t b : a) std::printf("%02x", b);
            std::printf("\n");
        } else if (cmd == "D") {
            std::string h;
            in >> h;
            const std::vector<uint8_t> a = unhex(h);
            mac_pdu_decoder_t dec;
            dec.set_configuration(a.data(), (uint32_t)a.size());
            dec.decode((uint32_t)a.size());
            const mac_header_type_t* mht = dec.get_mac_header_type();
            const mac_common_header_t* mch = dec.get_mac_common_header();
            std::printf("final=%d", dec.has_reached_valid_final_state() ? 1 : 0);
            if (mht) std::printf(" type %u sec %u", (uint32_t)mht->MAC_header_type, (uint32_t)mht->MAC_security);
            if (mch) std::printf(" reset %u seq %u rx %u tx %u nid %u", mch->Reset, mch->Sequence_number, mch->Receiver_Address,
                                 mch->Transmitter_Address, mch->Network_ID_3_lsb);
            for (const auto& m : dec.get_mmie_decoded_vec())
                std::printf(" [%u %u %u @%ld]", (uint32_t)m.mac_multiplexing_header.mac_ext, m.mac_multiplexing_header.ie_type, m.N_bytes,
                            (long)(m.data_ptr - a.data()));
            if (dec.get_unparsed_offset() != UNDEFINED_NUMERIC_32) std::printf(" unparsed %u", dec.get_unparsed_offset());
            std::printf("\n");
        }
    }
    return 0;
}

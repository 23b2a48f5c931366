// plcf_codec.cpp -- line-driven test driver for include/dnrp_plcf.hpp (tests/test_plcf_cpu.py compares its output with
// oracle/plcf.py, the literal restatement of the reference's shifts and masks).
//   P10 <HeaderFormat PacketLengthType PacketLength_m1 ShortNetworkID TransmitterIdentity TransmitPower Reserved DFMCS>
//   P20 <PLT PLm1 NID TID TXP MCS RID NSS RV NDI HARQ FF> <10 feedback numbers>
//   P21 <PLT PLm1 NID TID TXP MCS RID NSS Reserved FF> <10 feedback numbers>
//       feedback numbers: HARQ TF HARQ1 TF1 bitmap Codebook MIMO Reserved Buffer_Size MCS  (-1 = not defined / out of range)
//   U <type> <hex bytes> <PacketLength_max mcs_index_max N_SS_max>
//   S <dBm>   B <buffer size>
#include <cstdio>
#include <iostream>
#include <sstream>
#include <string>
#include <vector>
#include "dnrp_plcf.hpp"

using namespace dectnrp_b200::sp4;

static void read_fb(std::istringstream& in, uint32_t ff, feedback_info_pool_t& pool) {
    long long v[10];
    for (auto& x : v) in >> x;
    if (ff < 1 || ff > 6) return;
    feedback_info_fields_t& f = pool.by_format(ff);
    f.HARQ_Process_number = (uint32_t)v[0];
    f.Transmission_feedback = (feedback_info_t::transmission_feedback_t)(uint32_t)v[1];
    f.HARQ_Process_number_1 = (uint32_t)v[2];
    f.Transmission_feedback_1 = (feedback_info_t::transmission_feedback_t)(uint32_t)v[3];
    f.HARQ_feedback_bitmap = (uint32_t)v[4];
    f.Codebook_index = (uint32_t)v[5];
    f.MIMO_feedback = (feedback_info_t::mimo_feedback_t)(uint32_t)v[6];
    f.Reserved = (uint32_t)v[7];
    f.Buffer_Size = (uint32_t)v[8];
    f.MCS = (uint32_t)v[9];
}

static void print_fb(uint32_t ff, const feedback_info_pool_t& pool) {
    if (ff < 1 || ff > 6) { std::printf(" fb none"); return; }
    const feedback_info_fields_t& f = pool.by_format(ff);
    std::printf(" fb %d %d %d %d %d %d %d %d %d %d", (int)f.HARQ_Process_number, (int)f.Transmission_feedback,
                (int)f.HARQ_Process_number_1, (int)f.Transmission_feedback_1, (int)f.HARQ_feedback_bitmap,
                (int)f.Codebook_index, (int)f.MIMO_feedback, (int)f.Reserved, (int)f.Buffer_Size, (int)f.MCS);
}

static void hex(const uint8_t* a, size_t n) {
    for (size_t i = 0; i < n; ++i) std::printf("%02x", a[i]);
    std::printf("\n");
}

int main() {
    std::string line;
    while (std::getline(std::cin, line)) {
        std::istringstream in(line);
        std::string cmd;
        in >> cmd;
        if (cmd == "P10") {
            plcf_10_t p;
            in >> p.HeaderFormat >> p.PacketLengthType >> p.PacketLength_m1 >> p.ShortNetworkID >> p.TransmitterIdentity >>
                p.TransmitPower >> p.Reserved >> p.DFMCS;
            if (!p.is_valid()) { std::printf("INVALID\n"); continue; }
            uint8_t a[5] = {0};
            p.pack(a);
            hex(a, 5);
        } else if (cmd == "P20") {
            plcf_20_t p;
            in >> p.PacketLengthType >> p.PacketLength_m1 >> p.ShortNetworkID >> p.TransmitterIdentity >> p.TransmitPower >>
                p.DFMCS >> p.ReceiverIdentity >> p.NumberOfSpatialStreams >> p.DFRedundancyVersion >> p.DFNewDataIndication >>
                p.DFHARQProcessNumber >> p.FeedbackFormat;
            read_fb(in, p.FeedbackFormat, p.feedback_info_pool);
            const bool fb_ok = p.FeedbackFormat == 0 || (p.FeedbackFormat <= 6 &&
                feedback_info_pool_t::is_valid(p.FeedbackFormat, p.feedback_info_pool.by_format(p.FeedbackFormat)));
            if (!p.is_valid() || !fb_ok) { std::printf("INVALID\n"); continue; }
            uint8_t a[10] = {0};
            p.pack(a);
            hex(a, 10);
        } else if (cmd == "P21") {
            plcf_21_t p;
            p.HeaderFormat = 1;
            in >> p.PacketLengthType >> p.PacketLength_m1 >> p.ShortNetworkID >> p.TransmitterIdentity >> p.TransmitPower >>
                p.DFMCS >> p.ReceiverIdentity >> p.NumberOfSpatialStreams >> p.Reserved >> p.FeedbackFormat;
            read_fb(in, p.FeedbackFormat, p.feedback_info_pool);
            const bool fb_ok = p.FeedbackFormat == 0 || (p.FeedbackFormat <= 6 &&
                feedback_info_pool_t::is_valid(p.FeedbackFormat, p.feedback_info_pool.by_format(p.FeedbackFormat)));
            if (!p.is_valid() || !fb_ok) { std::printf("INVALID\n"); continue; }
            uint8_t a[10] = {0};
            p.pack(a);
            hex(a, 10);
        } else if (cmd == "U") {
            uint32_t type, plmax, mcsmax, nssmax;
            std::string h;
            in >> type >> h >> plmax >> mcsmax >> nssmax;
            std::vector<uint8_t> a(10, 0);
            for (size_t i = 0; i + 1 < h.size() && i / 2 < 10; i += 2) a[i / 2] = (uint8_t)std::stoul(h.substr(i, 2), nullptr, 16);
            plcf_decoder_t dec(plmax, mcsmax, nssmax);
            dec.set_configuration();
            dec.decode_and_rdc_check(type, a.data());
            const plcf_base_t* b = dec.get_plcf_base(type);
            std::printf("has=%u", dec.has_any_plcf());
            if (b != nullptr) {
                std::printf(" hf %u plt %u pl %u nid %u tid %u txp %u dBm %d mcs %u nss %u rv %u", b->get_HeaderFormat(),
                            b->PacketLengthType, b->get_PacketLength(), b->ShortNetworkID, b->TransmitterIdentity,
                            b->TransmitPower, b->get_TransmitPower_dBm(), b->DFMCS, b->get_N_SS(), b->get_DFRedundancyVersion());
                if (type == 2 && b->get_HeaderFormat() == 0) {
                    const auto* p = static_cast<const plcf_20_t*>(b);
                    std::printf(" rid %u ndi %u harq %u ff %u", p->ReceiverIdentity, p->DFNewDataIndication, p->DFHARQProcessNumber,
                                p->FeedbackFormat);
                    print_fb(p->FeedbackFormat, p->feedback_info_pool);
                } else if (type == 2) {
                    const auto* p = static_cast<const plcf_21_t*>(b);
                    std::printf(" rid %u ff %u", p->ReceiverIdentity, p->FeedbackFormat);
                    print_fb(p->FeedbackFormat, p->feedback_info_pool);
                }
            }
            std::printf("\n");
        } else if (cmd == "S") {
            int dbm;
            in >> dbm;
            plcf_10_t p;
            p.set_TransmitPower(dbm);
            std::printf("%u\n", p.TransmitPower);
        } else if (cmd == "B") {
            uint32_t sz;
            in >> sz;
            std::printf("%u\n", feedback_info_t::buffer_size_2_buffer_status(sz));
        }
    }
    return 0;
}

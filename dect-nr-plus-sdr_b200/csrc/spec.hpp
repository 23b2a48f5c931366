// spec.hpp -- host-side (C++17) restatement of ETSI TS 103 636-3 packet dimensions and of the LTE
// (TS 36.212 / 36.211) coding tables the DECT NR+ physical layer is built on.  Everything here runs
// once per packet *configuration* (not per packet) and produces FLAT look-up tables that are uploaded
// to HBM and indexed by the CUDA kernels.  No per-symbol state machines (the reference walks
// pcc_t/drs_t/pdc_t cursors symbol by symbol, we classify every resource element in closed form).
//
// Reference behaviour followed (files under /root/reference/lib/src/sections_part3 unless noted):
//   numerologies.cpp:30-70, mcs.cpp:27-107, tm_mode.cpp:27-131, transport_block_size.cpp:27-79,
//   transmission_packet_structure.cpp:28-96, derivative/packet_sizes.cpp:97-227,
//   fix/cbsegm.cpp:34-155, stf.cpp:34-287, drs.cpp:31-258, pcc.cpp:110-246, pdc.cpp:155-310,
//   transmit_diversity_precoding.cpp:34-95, beamforming_and_antenna_port_mapping.cpp:30-321.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

namespace dnrp {
namespace spec {

static constexpr uint32_t PCC_CELLS = 98;
static constexpr uint32_t PCC_BITS = 196;
static constexpr uint32_t SUBCARRIER_SPACING_MIN = 27000;
static constexpr uint32_t SAMP_RATE_MIN = 1728000;

// ------------------------------------------------------------------------------------------------
// packet definition and derived sizes
// ------------------------------------------------------------------------------------------------
struct psdef_t {
    uint32_t u, b, PacketLengthType, PacketLength, tm_mode_index, mcs_index, Z;
};

struct tm_mode_t {
    uint32_t index, N_eff_TX, N_SS, cl, N_TS, N_TX;
};

struct mcs_t {
    uint32_t index, N_bps, R_num, R_den;
};

struct cbsegm_t {
    uint32_t tbs, Z, C, K1, K2, K1_idx, K2_idx, C1, C2, F;
};

struct packet_sizes_t {
    psdef_t psdef;
    tm_mode_t tm;
    mcs_t mcs;
    uint32_t N_b_DFT, N_b_CP, N_b_OCC, N_guards_top, N_guards_bottom;
    uint32_t N_PACKET_symb, N_DF_symb, N_PDC_subc, N_DRS_subc;
    uint32_t G, N_PDC_bits, N_TB_bits, N_TB_byte, C;
    uint32_t N_samples_OFDM_symbol, N_samples_STF, N_samples_STF_CP_only, N_samples_DF,
        N_samples_GI, N_samples_packet_no_GI, N_samples_packet;
    cbsegm_t seg;
};

inline bool is_pow2(uint32_t x) { return x && !(x & (x - 1)); }

inline bool get_tm_mode(uint32_t index, tm_mode_t& q) {
    // Table 7.2-1: {N_eff_TX, N_SS, closed loop, N_TS, N_TX}
    static const uint32_t t[12][5] = {{1, 1, 0, 1, 1}, {2, 1, 0, 2, 2}, {2, 2, 0, 2, 2}, {1, 1, 1, 1, 2},
                                      {2, 2, 1, 2, 2}, {4, 1, 0, 4, 4}, {4, 4, 0, 4, 4}, {1, 1, 1, 1, 4},
                                      {2, 2, 1, 2, 4}, {4, 4, 1, 4, 4}, {8, 1, 0, 8, 8}, {8, 8, 0, 8, 8}};
    if (index > 11) return false;
    q = {index, t[index][0], t[index][1], t[index][2], t[index][3], t[index][4]};
    return true;
}

inline bool get_mcs(uint32_t index, mcs_t& q) {
    static const uint32_t t[12][3] = {{1, 1, 2}, {2, 1, 2}, {2, 3, 4}, {4, 1, 2}, {4, 3, 4}, {6, 2, 3},
                                      {6, 3, 4}, {6, 5, 6}, {8, 3, 4}, {8, 5, 6}, {10, 3, 4}, {10, 5, 6}};
    if (index > 11) return false;
    q = {index, t[index][0], t[index][1], t[index][2]};
    return true;
}

// TS 36.212 Table 5.1.3-3, block sizes K_i
inline uint32_t tc_cb_size(uint32_t idx) {
    if (idx < 60) return 40 + 8 * idx;
    if (idx < 92) return 512 + 16 * (idx - 59);
    if (idx < 124) return 1024 + 32 * (idx - 91);
    return 2048 + 64 * (idx - 123);
}
static constexpr uint32_t NOF_TC_CB_SIZES = 188;

inline int tc_cb_index(uint32_t len) {
    for (uint32_t j = 0; j < NOF_TC_CB_SIZES; ++j)
        if (tc_cb_size(j) >= len) return (int)j;
    return -1;
}

// TS 36.212 Table 5.1.3-3, QPP interleaver parameters (f1, f2) for the 188 block sizes.
inline void qpp_params(uint32_t idx, uint32_t& f1, uint32_t& f2) {
    static const uint16_t t[188][2] = {
        {3, 10},    {7, 12},    {19, 42},   {7, 16},    {7, 18},    {11, 20},   {5, 22},    {11, 24},
        {7, 26},    {41, 84},   {103, 90},  {15, 32},   {9, 34},    {17, 108},  {9, 38},    {21, 120},
        {101, 84},  {21, 44},   {57, 46},   {23, 48},   {13, 50},   {27, 52},   {11, 36},   {27, 56},
        {85, 58},   {29, 60},   {33, 62},   {15, 32},   {17, 198},  {33, 68},   {103, 210}, {19, 36},
        {19, 74},   {37, 76},   {19, 78},   {21, 120},  {21, 82},   {115, 84},  {193, 86},  {21, 44},
        {133, 90},  {81, 46},   {45, 94},   {23, 48},   {243, 98},  {151, 40},  {155, 102}, {25, 52},
        {51, 106},  {47, 72},   {91, 110},  {29, 168},  {29, 114},  {247, 58},  {29, 118},  {89, 180},
        {91, 122},  {157, 62},  {55, 84},   {31, 64},   {17, 66},   {35, 68},   {227, 420}, {65, 96},
        {19, 74},   {37, 76},   {41, 234},  {39, 80},   {185, 82},  {43, 252},  {21, 86},   {155, 44},
        {79, 120},  {139, 92},  {23, 94},   {217, 48},  {25, 98},   {17, 80},   {127, 102}, {25, 52},
        {239, 106}, {17, 48},   {137, 110}, {215, 112}, {29, 114},  {15, 58},   {147, 118}, {29, 60},
        {59, 122},  {65, 124},  {55, 84},   {31, 64},   {17, 66},   {171, 204}, {67, 140},  {35, 72},
        {19, 74},   {39, 76},   {19, 78},   {199, 240}, {21, 82},   {211, 252}, {21, 86},   {43, 88},
        {149, 60},  {45, 92},   {49, 846},  {71, 48},   {13, 28},   {17, 80},   {25, 102},  {183, 104},
        {55, 954},  {127, 96},  {27, 110},  {29, 112},  {29, 114},  {57, 116},  {45, 354},  {31, 120},
        {59, 610},  {185, 124}, {113, 420}, {31, 64},   {17, 66},   {171, 136}, {209, 420}, {253, 216},
        {367, 444}, {265, 456}, {181, 468}, {39, 80},   {27, 164},  {127, 504}, {143, 172}, {43, 88},
        {29, 300},  {45, 92},   {157, 188}, {47, 96},   {13, 28},   {111, 240}, {443, 204}, {51, 104},
        {51, 212},  {451, 192}, {257, 220}, {57, 336},  {313, 228}, {271, 232}, {179, 236}, {331, 120},
        {363, 244}, {375, 248}, {127, 168}, {31, 64},   {33, 130},  {43, 264},  {33, 134},  {477, 408},
        {35, 138},  {233, 280}, {357, 142}, {337, 480}, {37, 146},  {71, 444},  {71, 120},  {37, 152},
        {39, 462},  {127, 234}, {39, 158},  {39, 80},   {31, 96},   {113, 902}, {41, 166},  {251, 336},
        {43, 170},  {21, 86},   {43, 174},  {45, 176},  {45, 178},  {161, 120}, {89, 182},  {323, 184},
        {47, 186},  {23, 94},   {47, 190},  {263, 480}};
    f1 = t[idx][0];
    f2 = t[idx][1];
}

// code block segmentation with configurable Z (reference: fix/cbsegm.cpp:65-123)
inline bool cbsegm(uint32_t tbs, uint32_t Z, cbsegm_t& s) {
    std::memset(&s, 0, sizeof(s));
    s.Z = Z;
    if (tbs == 0) return true;
    const uint32_t B = tbs + 24;
    s.tbs = tbs;
    uint32_t Bp;
    if (B <= Z) {
        s.C = 1;
        Bp = B;
    } else {
        s.C = (B + (Z - 24) - 1) / (Z - 24);
        Bp = B + 24 * s.C;
    }
    const int idx1 = tc_cb_index((Bp - 1) / s.C + 1);
    if (idx1 < 0) return false;
    s.K1 = tc_cb_size(idx1);
    s.K1_idx = idx1;
    if (s.C == 1) {
        s.C1 = 1;
    } else {
        if (idx1 == 0) return false;
        s.K2 = tc_cb_size(idx1 - 1);
        s.K2_idx = idx1 - 1;
        s.C2 = (s.C * s.K1 - Bp) / (s.K1 - s.K2);
        s.C1 = s.C - s.C2;
    }
    s.F = s.C1 * s.K1 + s.C2 * s.K2 - Bp;
    return true;
}

inline uint32_t get_N_DF_symb(uint32_t u, uint32_t N_PACKET_symb) {
    return N_PACKET_symb - (u == 1 ? 2 : (u == 8 ? 4 : 3));
}

inline uint32_t get_N_step(uint32_t N_eff_TX) { return N_eff_TX <= 2 ? 5 : 10; }

inline uint32_t get_nof_drs_symbols_per_TS(uint32_t N_PACKET_symb, uint32_t N_eff_TX) {
    const uint32_t N_step = get_N_step(N_eff_TX);
    uint32_t n = N_PACKET_symb / N_step;
    if (N_step == 10 && N_PACKET_symb % 10 != 0) ++n;
    return n;
}

inline uint32_t get_N_TB_bits(uint32_t N_PDC_bits, uint32_t Z) {
    const uint32_t L = 24;
    const uint32_t M = N_PDC_bits <= 512 ? 8 : (N_PDC_bits <= 1024 ? 16 : (N_PDC_bits <= 2048 ? 32 : 64));
    const uint32_t N_M = (N_PDC_bits / M) * M;
    if (N_M <= L) return 0;
    if (N_M <= Z) return N_M - L;
    const uint32_t C = (N_M - L + Z - 1) / Z;
    return N_M - (C + 1) * L;
}

// returns false where the reference returns std::nullopt (packet_sizes.cpp:97-227)
inline bool get_packet_sizes(const psdef_t& d, packet_sizes_t& q) {
    std::memset(&q, 0, sizeof(q));
    if (!(is_pow2(d.u) && d.u <= 8)) return false;
    if (!((is_pow2(d.b) && d.b <= 16) || d.b == 12)) return false;
    if (d.PacketLengthType > 1 || d.PacketLength == 0 || d.PacketLength > 16) return false;
    if (!(d.Z == 2048 || d.Z == 6144)) return false;
    if (!get_tm_mode(d.tm_mode_index, q.tm) || !get_mcs(d.mcs_index, q.mcs)) return false;
    q.psdef = d;
    q.N_b_DFT = 64 * d.b;
    q.N_b_CP = 8 * d.b;
    q.N_b_OCC = 56 * d.b;
    q.N_guards_top = (q.N_b_DFT - q.N_b_OCC) / 2 - 1;
    q.N_guards_bottom = q.N_guards_top + 1;
    const uint32_t N_SLOT_symb = d.u * 10, N_SLOT_subslot = d.u * 2;
    q.N_PACKET_symb =
        d.PacketLengthType == 0 ? d.PacketLength * N_SLOT_symb / N_SLOT_subslot : d.PacketLength * N_SLOT_symb;
    const uint32_t N_eff_TX = q.tm.N_eff_TX;
    if (N_eff_TX == 4 && q.N_PACKET_symb < 15) return false;
    if (d.u == 8 && N_eff_TX == 8 && (q.N_PACKET_symb < 20 || q.N_PACKET_symb % 10 != 0)) return false;
    q.N_DF_symb = get_N_DF_symb(d.u, q.N_PACKET_symb);
    q.N_DRS_subc = N_eff_TX * q.N_b_OCC / 4 * get_nof_drs_symbols_per_TS(q.N_PACKET_symb, N_eff_TX);
    if (q.N_DF_symb * q.N_b_OCC <= q.N_DRS_subc + PCC_CELLS) return false;
    q.N_PDC_subc = q.N_DF_symb * q.N_b_OCC - q.N_DRS_subc - PCC_CELLS;
    q.G = q.tm.N_SS * q.N_PDC_subc * q.mcs.N_bps;
    q.N_PDC_bits = (q.G * q.mcs.R_num) / q.mcs.R_den;
    q.N_TB_bits = get_N_TB_bits(q.N_PDC_bits, d.Z);
    if (q.N_TB_bits == 0) return false;
    q.N_TB_byte = (q.N_TB_bits + 7) / 8;
    if (!cbsegm(q.N_TB_bits, d.Z, q.seg) || q.seg.F > 0) return false;
    q.C = q.seg.C;
    q.N_samples_OFDM_symbol = 72 * d.b;
    q.N_samples_STF = d.u == 1 ? q.N_samples_OFDM_symbol * 14 / 9 : q.N_samples_OFDM_symbol * 2;
    q.N_samples_STF_CP_only = q.N_samples_STF - 64 * d.b;
    q.N_samples_DF = q.N_samples_OFDM_symbol * q.N_DF_symb;
    q.N_samples_GI = d.u == 1 ? q.N_samples_OFDM_symbol * 4 / 9
                              : (d.u == 8 ? q.N_samples_OFDM_symbol * 2 : q.N_samples_OFDM_symbol);
    q.N_samples_packet_no_GI = q.N_samples_STF + q.N_samples_DF;
    q.N_samples_packet = q.N_samples_packet_no_GI + q.N_samples_GI;
    return true;
}

}  // namespace spec
}  // namespace dnrp

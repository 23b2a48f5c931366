// spec_fec.hpp -- host-side tables for the FEC chain: LTE turbo rate matching (TS 36.212 5.1.4.1),
// Gold-31 scrambling sequence (TS 36.211 7.2), CRC polynomials (TS 36.212 5.1.1) and the transmit
// codebooks.  The reference delegates these to srsRAN_4G release_23_11 (rm_turbo.c, sequence.c, crc.c;
// call sites reference lib/src/phy/fec/pdc_enc.cpp:191-217, pcc_enc.cpp:187-212,
// lib/src/sections_part3/scrambling_pdc.cpp:43-56); the arithmetic below is the published 3GPP one.
#pragma once
#include "spec.hpp"

namespace dnrp {
namespace spec {

static constexpr uint32_t CRC24A_POLY = 0x1864CFB;
static constexpr uint32_t CRC24B_POLY = 0x1800063;
static constexpr uint32_t CRC16_POLY = 0x11021;
static constexpr uint32_t PCC_SCRAMBLING_INIT = 0x44454354;  // reference pcc_enc.cpp:46

// bitwise CRC over `nbits` MSB-first bits of `data`, zero initial state
inline uint32_t crc_bits(const uint8_t* data, uint32_t nbits, uint32_t poly, uint32_t order) {
    uint32_t r = 0;
    const uint32_t top = 1u << order;
    for (uint32_t i = 0; i < nbits; ++i) {
        const uint32_t bit = (data[i >> 3] >> (7 - (i & 7))) & 1;
        r = (r << 1) | bit;
        if (r & top) r ^= poly;
    }
    for (uint32_t i = 0; i < order; ++i) {
        r <<= 1;
        if (r & top) r ^= poly;
    }
    return r & (top - 1);
}

// Rate matching for one code block size K.  d-streams are addressed as n = s*(K+4) + i with
// s = 0 systematic, 1 parity 1, 2 parity 2 (tail bits in i = K..K+3 per TS 36.212 5.1.3.2.2).
// `src[p]` lists the d-stream address of the p-th NON-NULL entry of the circular buffer w (in buffer
// order), `start[rv]` is the position in that list of the first non-NULL entry at or after k0(rv).
// Transmitted bit e of a code block is d[src[(start[rv] + e) % (3K+12)]].
// Limited soft buffer (TS 36.212 5.1.4.1.2, which ETSI TS 103 636-3 6.1.5.3 follows; the reference leaves it as a ToDo,
// pdc_enc.cpp:122-123, README.md:180): the circular buffer ends at N_cb = min(floor(N_IR / C), K_w) instead of K_w.  The
// first Nc entries of src[] are then the usable ones, k0 and the wrap-around follow N_cb.
struct rm_table_t {
    uint32_t K, N;  // N = 3K+12
    uint32_t Nc;    // non-NULL entries inside the first N_cb positions of the circular buffer (= N without limitation)
    std::vector<uint16_t> src;
    std::vector<uint16_t> pos;  // inverse of src: d-stream address -> position in the non-NULL list
    uint32_t start[4];
};

inline void build_rm_table(uint32_t K, rm_table_t& t, uint32_t N_cb_limit = 0) {
    static const uint8_t P[32] = {0, 16, 8, 24, 4, 20, 12, 28, 2, 18, 10, 26, 6, 22, 14, 30,
                                  1, 17, 9, 25, 5, 21, 13, 29, 3, 19, 11, 27, 7, 23, 15, 31};
    const uint32_t D = K + 4, R = (D + 31) / 32, Kp = 32 * R, ND = Kp - D, Kw = 3 * Kp;
    t.K = K;
    t.N = 3 * D;
    std::vector<int32_t> w(Kw, -1);
    for (uint32_t k = 0; k < Kp; ++k) {
        // v0/v1: column-permuted block interleaver read column by column; y has ND leading NULLs
        const uint32_t col = P[k / R], row = k % R;
        const int32_t y = (int32_t)(row * 32 + col) - (int32_t)ND;
        if (y >= 0) {
            w[k] = y;                  // stream 0
            w[Kp + 2 * k] = D + y;     // stream 1
        }
        const uint32_t pi = (P[k / R] + 32 * (k % R) + 1) % Kp;
        const int32_t y2 = (int32_t)pi - (int32_t)ND;
        if (y2 >= 0) w[Kp + 2 * k + 1] = 2 * D + y2;  // stream 2
    }
    t.src.clear();
    std::vector<uint32_t> nn_before(Kw + 1, 0);  // number of non-NULL entries before w index k
    for (uint32_t k = 0; k < Kw; ++k) {
        nn_before[k] = (uint32_t)t.src.size();
        if (w[k] >= 0) t.src.push_back((uint16_t)w[k]);
    }
    nn_before[Kw] = (uint32_t)t.src.size();
    t.pos.assign(t.N, 0);
    for (uint32_t p = 0; p < t.N; ++p) t.pos[t.src[p]] = (uint16_t)p;
    // N_cb = K_w unless the soft buffer is limited (the reference: always K_w, README.md:180)
    const uint32_t N_cb = (N_cb_limit && N_cb_limit < Kw) ? N_cb_limit : Kw;
    t.Nc = nn_before[N_cb];
    for (uint32_t rv = 0; rv < 4; ++rv) {
        const uint32_t k0 = R * (2 * ((N_cb + 8 * R - 1) / (8 * R)) * rv + 2);
        t.start[rv] = nn_before[k0 % N_cb] % t.Nc;
    }
}

// Gold-31 sequence c(n), n = 0..len-1, packed MSB-first into bytes (TS 36.211 7.2, Nc = 1600)
inline std::vector<uint8_t> gold_sequence_packed(uint32_t c_init, uint32_t len) {
    std::vector<uint8_t> out((len + 7) / 8, 0);
    uint32_t x1 = 1, x2 = c_init & 0x7fffffff;
    auto step = [&]() {
        const uint32_t n1 = ((x1 >> 3) ^ x1) & 1;
        const uint32_t n2 = ((x2 >> 3) ^ (x2 >> 2) ^ (x2 >> 1) ^ x2) & 1;
        x1 = (x1 >> 1) | (n1 << 30);
        x2 = (x2 >> 1) | (n2 << 30);
    };
    for (uint32_t i = 0; i < 1600; ++i) step();
    for (uint32_t n = 0; n < len; ++n) {
        const uint32_t c = (x1 ^ x2) & 1;
        out[n >> 3] |= (uint8_t)(c << (7 - (n & 7)));
        step();
    }
    return out;
}

// ---- codebooks W (Tables 6.3.4-1..6), entries coded 0, 1, -1, 2 (= j), -2 (= -j), row-major
// [N_TX][N_TS]; reference beamforming_and_antenna_port_mapping.cpp:251-281
inline const int8_t* codebook(uint32_t N_TS, uint32_t N_TX, uint32_t idx, uint32_t* n_idx) {
    static const int8_t j = 2;
    static const int8_t W0[1][1] = {{1}};
    static const int8_t W1[6][2] = {{1, 0}, {0, 1}, {1, 1}, {1, -1}, {1, j}, {1, -j}};
    static const int8_t W2[28][4] = {
        {1, 0, 0, 0},   {0, 1, 0, 0},  {0, 0, 1, 0},   {0, 0, 0, 1},   {1, 0, 1, 0},   {1, 0, -1, 0},
        {1, 0, j, 0},   {1, 0, -j, 0}, {0, 1, 0, 1},   {0, 1, 0, -1},  {0, 1, 0, j},   {0, 1, 0, -j},
        {1, 1, 1, 1},   {1, 1, j, j},  {1, 1, -1, -1}, {1, 1, -j, -j}, {1, j, 1, j},   {1, j, j, -1},
        {1, j, -1, -j}, {1, j, -j, 1}, {1, -1, 1, -1}, {1, -1, j, -j}, {1, -1, -1, 1}, {1, -1, -j, j},
        {1, -j, 1, -j}, {1, -j, j, 1}, {1, -j, -1, j}, {1, -j, -j, -1}};
    static const int8_t W3[3][4] = {{1, 0, 0, 1}, {1, 1, 1, -1}, {1, 1, j, -j}};
    static const int8_t W4[22][8] = {
        {1, 0, 0, 1, 0, 0, 0, 0},     {1, 0, 0, 0, 0, 1, 0, 0},     {1, 0, 0, 0, 0, 0, 0, 1},
        {0, 0, 1, 0, 0, 1, 0, 0},     {0, 0, 1, 0, 0, 0, 0, 1},     {0, 0, 0, 0, 1, 0, 0, 1},
        {1, 0, 0, 1, 1, 0, 0, -j},    {1, 0, 0, 1, 1, 0, 0, j},     {1, 0, 0, 1, -j, 0, 0, 1},
        {1, 0, 0, 1, -j, 0, 0, -1},   {1, 0, 0, 1, -1, 0, 0, -j},   {1, 0, 0, 1, -1, 0, 0, j},
        {1, 0, 0, 1, j, 0, 0, 1},     {1, 0, 0, 1, j, 0, 0, -1},    {1, 1, 1, 1, 1, -1, 1, -1},
        {1, 1, 1, 1, j, -j, j, -j},   {1, 1, j, j, 1, -1, j, -j},   {1, 1, j, j, j, -j, -1, 1},
        {1, 1, -1, -1, 1, -1, -1, 1}, {1, 1, -1, -1, j, -j, -j, j}, {1, 1, -j, -j, 1, -1, -j, j},
        {1, 1, -j, -j, j, -j, 1, -1}};
    static const int8_t W5[5][16] = {{1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1},
                                     {1, 1, 0, 0, 0, 0, 1, 1, 1, -1, 0, 0, 0, 0, 1, -1},
                                     {1, 1, 0, 0, 0, 0, 1, 1, j, -j, 0, 0, 0, 0, j, -j},
                                     {1, 1, 1, 1, 1, -1, 1, -1, 1, 1, -1, -1, 1, -1, -1, 1},
                                     {1, 1, 1, 1, 1, -1, 1, -1, j, j, -j, -j, j, -j, -j, j}};
    static int8_t W6[1][64];
    static bool w6 = false;
    if (!w6) {
        std::memset(W6, 0, sizeof(W6));
        for (int i = 0; i < 8; ++i) W6[0][i * 8 + i] = 1;
        w6 = true;
    }
#define DNRP_CB(W, n)            \
    do {                         \
        if (n_idx) *n_idx = n;   \
        return idx < n ? W[idx] : nullptr; \
    } while (0)
    if (N_TS == 1 && N_TX == 1) DNRP_CB(W0, 1);
    if (N_TS == 1 && N_TX == 2) DNRP_CB(W1, 6);
    if (N_TS == 1 && N_TX == 4) DNRP_CB(W2, 28);
    if (N_TS == 2 && N_TX == 2) DNRP_CB(W3, 3);
    if (N_TS == 2 && N_TX == 4) DNRP_CB(W4, 22);
    if (N_TS == 4 && N_TX == 4) DNRP_CB(W5, 5);
    if (N_TS == 8 && N_TX == 8) DNRP_CB(W6, 1);
#undef DNRP_CB
    if (n_idx) *n_idx = 0;
    return nullptr;
}

// 1/sqrt(number of non-zero entries) (reference beamforming_and_antenna_port_mapping.cpp:307-319)
inline float codebook_scale(const int8_t* W, uint32_t n) {
    uint32_t nz = 0;
    for (uint32_t i = 0; i < n; ++i) nz += W[i] != 0;
    return 1.0f / std::sqrt((float)nz);
}

// "optimal DAC" scaling (reference beamforming_and_antenna_port_mapping.cpp:155-190)
inline float codebook_scale_optimal_dac(uint32_t N_TS, uint32_t N_TX, uint32_t idx) {
    const float r2 = 1.0f / std::sqrt(2.0f);
    if (N_TS == 2 && N_TX == 2) return idx == 0 ? 1.0f : r2;
    if (N_TS == 2 && N_TX == 4) return idx < 14 ? 1.0f : r2;
    if (N_TS == 4 && N_TX == 4) return idx == 0 ? 1.0f : (idx <= 2 ? r2 : 0.5f);
    return 1.0f;
}

}  // namespace spec
}  // namespace dnrp

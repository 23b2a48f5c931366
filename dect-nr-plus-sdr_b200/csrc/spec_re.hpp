// spec_re.hpp -- resource-element classification for one packet geometry (b, N_eff_TX, u,
// N_PACKET_symb) as flat tables for the GPU.  The reference keeps per-symbol cursors
// (pcc_t/drs_t/pdc_t::is_symbol_index, reference lib/src/sections_part3/{pcc,drs,pdc}.cpp); here every
// cell (l, k) of the data field is classified once, in closed form, and the kernels index the result.
//
// Subcarrier index k runs 0..N_b_OCC (N_b_OCC+1 entries): k = N_b_OCC/2 is DC, k < N_b_OCC/2 is the
// lower half of the spectrum (reference convention, physical_resources.cpp:25-35, drs.cpp:222-227).
#pragma once
#include "spec.hpp"

namespace dnrp {
namespace spec {

enum cell_type_t : uint32_t { CELL_NONE = 0, CELL_DRS = 1, CELL_PCC = 2, CELL_PDC = 3 };

// cell word: [31:30] type, [29:27] DRS transmit stream, [26] DRS sign (1 = -1), [25:0] running index
// (PCC symbol index 0..97, PDC symbol index 0..N_PDC_subc-1, DRS pilot index 0..N_b_OCC/4-1)
inline uint32_t cell_word(uint32_t type, uint32_t ts, uint32_t neg, uint32_t idx) {
    return (type << 30) | (ts << 27) | (neg << 26) | idx;
}

struct re_map_t {
    uint32_t b, N_eff_TX, u, N_PACKET_symb, N_DF_symb, N_b_OCC;
    // [(N_DF_symb+1) * (N_b_OCC+1)], row 0 (STF symbol) is all CELL_NONE
    std::vector<uint32_t> cell;
    // (l << 16) | k for every PCC / PDC symbol in mapping order
    std::vector<uint32_t> pcc_lk, pdc_lk;
    // DRS: for transmit stream t and DRS occurrence n: symbol index and pilot subcarriers
    uint32_t n_drs_symb;                 // occurrences per transmit stream
    std::vector<uint32_t> drs_l;         // [N_eff_TX][n_drs_symb]
    std::vector<uint16_t> drs_k;         // [N_eff_TX][n_drs_symb][N_b_OCC/4]
    std::vector<int8_t> drs_sign;        // [N_eff_TX][N_b_OCC/4] (+1/-1), independent of n
};

// base sequence of the DRS (TS 103 636-3 clause 5.2.3; same numbers as reference drs.hpp:124-132)
inline int drs_base(uint32_t i) {
    static const int8_t y[56] = {1,  1,  1,  1,  -1, 1,  1,  -1, -1, 1, 1,  1,  1,  -1, 1,  -1, 1,  1,  -1,
                                 1,  -1, 1,  -1, 1,  1,  1,  1,  1,  -1, 1, -1, -1, 1,  1,  -1, -1, -1, -1,
                                 1,  -1, -1, -1, -1, -1, 1,  1,  1,  -1, 1, 1,  -1, -1, 1,  -1, -1, -1};
    return y[i % 56];
}

// occupied-list position q (0..N_b_OCC-1, DC excluded) -> subcarrier index k (0..N_b_OCC, DC included)
inline uint32_t occ2k(uint32_t q, uint32_t N_b_OCC) { return q < N_b_OCC / 2 ? q : q + 1; }

inline bool build_re_map(uint32_t b, uint32_t N_eff_TX, uint32_t u, uint32_t N_PACKET_symb, re_map_t& m) {
    if (N_eff_TX == 4 && N_PACKET_symb < 15) return false;
    if (u == 8 && N_eff_TX == 8 && (N_PACKET_symb < 20 || N_PACKET_symb % 10 != 0)) return false;
    m.b = b;
    m.N_eff_TX = N_eff_TX;
    m.u = u;
    m.N_PACKET_symb = N_PACKET_symb;
    m.N_DF_symb = get_N_DF_symb(u, N_PACKET_symb);
    m.N_b_OCC = 56 * b;
    const uint32_t Nk = m.N_b_OCC + 1, NP = m.N_b_OCC / 4;
    const uint32_t N_step = get_N_step(N_eff_TX);
    // PCC allocation may look at symbols past a very short data field, classify a few more rows
    const uint32_t rows = std::max(m.N_DF_symb + 1, 21u);
    std::vector<uint32_t> cell(rows * Nk, 0);

    // ---- DRS (Figure 4.5-2 / 4.5-3): TS t, occurrence n sits in symbol 1 + t/4 + n*N_step on every
    // 4th occupied subcarrier with offset (t + 2*(n%2)) % 4
    m.n_drs_symb = get_nof_drs_symbols_per_TS(N_PACKET_symb, N_eff_TX);
    const uint32_t n_mark = std::max(m.n_drs_symb, rows / N_step + 1);
    m.drs_l.assign(N_eff_TX * m.n_drs_symb, 0);
    m.drs_k.assign(N_eff_TX * m.n_drs_symb * NP, 0);
    m.drs_sign.assign(N_eff_TX * NP, 0);
    for (uint32_t t = 0; t < N_eff_TX; ++t) {
        for (uint32_t i = 0; i < NP; ++i) m.drs_sign[t * NP + i] = (t < 4 ? 1 : -1) * drs_base(4 * i + t % 4);
        for (uint32_t n = 0; n < n_mark; ++n) {
            const uint32_t l = 1 + t / 4 + n * N_step;
            for (uint32_t i = 0; i < NP; ++i) {
                const uint32_t k = occ2k(4 * i + (t + 2 * (n % 2)) % 4, m.N_b_OCC);
                if (n < m.n_drs_symb) {
                    m.drs_l[t * m.n_drs_symb + n] = l;
                    m.drs_k[(t * m.n_drs_symb + n) * NP + i] = (uint16_t)k;
                }
                if (l < rows) cell[l * Nk + k] = cell_word(CELL_DRS, t, m.drs_sign[t * NP + i] < 0, i);
            }
        }
    }

    // ---- PCC, clause 5.2.4 (98 cells): fill symbols from l = 1; the last, partially used symbol is
    // read out of a 7-row block interleaver; the final set is used in ascending (l, k) order
    // (reference pcc.cpp:160-246, the sort at :238).
    std::vector<uint32_t> pcc;
    {
        uint32_t l = 1, remaining = PCC_CELLS;
        while (remaining) {
            std::vector<uint32_t> freek;
            for (uint32_t k = 0; k < Nk; ++k)
                if (k != m.N_b_OCC / 2 && cell[l * Nk + k] == 0) freek.push_back(k);
            const uint32_t U = (uint32_t)freek.size();
            if (U < remaining) {
                for (uint32_t k : freek) pcc.push_back((l << 16) | k);
                remaining -= U;
                ++l;
                if (l >= rows) return false;
            } else {
                if (U % 7) return false;
                const uint32_t Ccol = U / 7;
                for (uint32_t c = 0; c < Ccol && remaining; ++c)
                    for (uint32_t r = 0; r < 7 && remaining; ++r) {
                        pcc.push_back((l << 16) | freek[r * Ccol + c]);
                        --remaining;
                    }
            }
        }
        std::sort(pcc.begin(), pcc.end());
    }
    for (uint32_t i = 0; i < pcc.size(); ++i) {
        const uint32_t l = pcc[i] >> 16, k = pcc[i] & 0xffff;
        cell[l * Nk + k] = cell_word(CELL_PCC, 0, 0, i);
    }
    if ((pcc.back() >> 16) > m.N_DF_symb) return false;  // PCC does not fit into the data field
    m.pcc_lk = pcc;

    // ---- PDC, clause 5.2.5: every remaining occupied cell of the data field in ascending order
    m.pdc_lk.clear();
    for (uint32_t l = 1; l <= m.N_DF_symb; ++l)
        for (uint32_t k = 0; k < Nk; ++k) {
            if (k == m.N_b_OCC / 2 || cell[l * Nk + k] != 0) continue;
            cell[l * Nk + k] = cell_word(CELL_PDC, 0, 0, (uint32_t)m.pdc_lk.size());
            m.pdc_lk.push_back((l << 16) | k);
        }
    m.cell.assign(cell.begin(), cell.begin() + (m.N_DF_symb + 1) * Nk);
    return true;
}

// ---- STF (clause 5.2.2; reference stf.cpp:159-287): values (+-1)*exp(j*pi/4) on every 4th occupied
// subcarrier, returned as sign per k (0 where empty); the caller multiplies by exp(j*pi/4)*scale.
inline std::vector<int8_t> build_stf_sign(uint32_t b, uint32_t N_eff_TX) {
    static const int8_t y1[14] = {1, -1, 1, 1, -1, 1, 1, -1, 1, 1, 1, -1, -1, -1};
    static const int8_t y2[28] = {-1, 1,  -1, 1, 1,  -1, 1,  1, -1, 1, 1,  1,  -1, 1,
                                  -1, -1, -1, 1, -1, -1, -1, 1, 1,  1, -1, -1, -1, -1};
    static const int8_t y4[56] = {-1, -1, -1, 1,  -1, 1,  -1, -1, 1,  1,  1,  1,  -1, 1,  -1, -1, -1, 1,  -1,
                                  1,  1,  -1, -1, -1, -1, -1, 1,  -1, 1,  1,  1,  -1, 1,  -1, 1,  1,  -1, -1,
                                  -1, -1, 1,  -1, -1, -1, -1, 1,  -1, 1,  1,  -1, -1, -1, -1, -1, 1,  -1};
    const uint32_t N_b_OCC = 56 * b, NS = N_b_OCC / 4;
    std::vector<int> y;
    auto mirror = [](const std::vector<int>& in) {  // y_r(k) = (-1)^k y(N-1-k)
        std::vector<int> out(in.size());
        for (size_t i = 0; i < in.size(); ++i) out[i] = in[in.size() - 1 - i] * (i % 2 ? -1 : 1);
        return out;
    };
    if (b == 1) y.assign(y1, y1 + 14);
    else if (b == 2) y.assign(y2, y2 + 28);
    else {
        y.assign(y4, y4 + 56);
        if (b >= 8) {
            auto r = mirror(y);
            y.insert(y.end(), r.begin(), r.end());
        }
        if (b >= 12) {
            auto r = mirror(y);
            y.insert(y.end(), r.begin(), r.end());
        }
        if (b == 12) y = std::vector<int>(y.begin() + 28, y.begin() + 28 + 168);
    }
    uint32_t lg = 0;
    for (uint32_t n = N_eff_TX; n > 1; n >>= 1) ++lg;
    std::vector<int8_t> out(N_b_OCC + 1, 0);
    for (uint32_t i = 0; i < NS; ++i) {
        // occupied-list position of STF cell i: lower half every 4th from 0, upper half from +3
        const uint32_t q = i < NS / 2 ? 4 * i : N_b_OCC / 2 + 3 + 4 * (i - NS / 2);
        out[occ2k(q, N_b_OCC)] = (int8_t)y[(i + 2 * lg) % NS];
    }
    return out;
}

// cover sequence of the STF patterns (active in the reference snapshot, stf.hpp:146-151)
inline float stf_cover(uint32_t pattern_idx) {
    static const float c[9] = {1, -1, 1, 1, -1, -1, -1, -1, -1};
    return c[pattern_idx];
}
inline uint32_t stf_n_pattern(uint32_t u) { return u == 1 ? 7 : 9; }

// ---- transmit diversity pairing (clause 6.3.3.2; reference transmit_diversity_precoding.cpp:45-72)
inline uint32_t txdiv_modulo(uint32_t N_TS) { return N_TS == 2 ? 1 : (N_TS == 4 ? 6 : 12); }
inline void txdiv_pair(uint32_t N_TS, uint32_t i_mod, uint32_t& A, uint32_t& B) {
    static const uint8_t p4[6][2] = {{0, 1}, {2, 3}, {0, 2}, {1, 3}, {0, 3}, {1, 2}};
    static const uint8_t p8[12][2] = {{0, 1}, {2, 3}, {4, 5}, {6, 7}, {0, 4}, {1, 5},
                                      {2, 6}, {3, 7}, {0, 2}, {1, 3}, {4, 6}, {5, 7}};
    if (N_TS == 2) { A = 0; B = 1; }
    else if (N_TS == 4) { A = p4[i_mod][0]; B = p4[i_mod][1]; }
    else { A = p8[i_mod][0]; B = p8[i_mod][1]; }
}

}  // namespace spec
}  // namespace dnrp

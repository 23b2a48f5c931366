// plan_fec.cu -- host-side construction of the FEC plan: code block descriptors for the encoder and the
// decoder kernels, rate-matching tables, CRC combination tables.  Segmentation and the per-CB bit
// budgets follow reference lib/src/phy/fec/pdc_enc.cpp:148-175 (encoder) and :322-337 (decoder,
// including its '>' comparison inherited from srsRAN's sch.c), lib/src/phy/fec/pcc_enc.cpp:187-208.
#include <algorithm>
#include <map>
#include "ctx.hpp"

namespace dnrp {

uint32_t tdec_window(uint32_t K) {
    // divisor of K/2 closest to 32 among divisors >= 16 (ties: larger); K/2 itself if K/2 < 16
    const uint32_t h = K / 2;
    if (h < 16) return h;
    uint32_t best = h;
    int bestd = 1 << 30;
    for (uint32_t d = 16; d <= h; ++d) {
        if (h % d) continue;
        int dd = (int)d - 32;
        if (dd < 0) dd = -dd;
        if (dd < bestd || (dd == bestd && d > best)) { bestd = dd; best = d; }
    }
    return best;
}

static uint32_t gf2_mulmod_host(uint32_t a, uint32_t b, uint32_t poly, uint32_t order) {
    uint32_t r = 0;
    const uint32_t top = 1u << order;
    for (uint32_t i = 0; i < order; ++i) {
        r <<= 1;
        if (r & top) r ^= poly;
        if ((b >> (order - 1 - i)) & 1) r ^= a;
    }
    return r;
}

uint32_t gf2_xpow_host(uint32_t n, uint32_t poly, uint32_t order) {
    uint32_t result = 1, base = 2;
    while (n) {
        if (n & 1) result = gf2_mulmod_host(result, base, poly, order);
        base = gf2_mulmod_host(base, base, poly, order);
        n >>= 1;
    }
    return result;
}

int build_fec_plan(const spec::packet_sizes_t& ps, fec_plan_t& f, uint32_t N_IR) {
    const spec::cbsegm_t& s = ps.seg;
    // K -> offset of src[] (transmitter: circular buffer position -> source bit); the receiver's tables follow at +N
    // (four, one per redundancy version): pos[] in the order of the DECODER's soft-buffer layout (soft_layout_index below), so that k_derm, which walks
    // table and soft buffer in lockstep, writes the streams exactly as k_turbo_decode stages them
    std::map<uint32_t, uint32_t> rm_off;
    std::map<uint32_t, spec::rm_table_t> rm_tab;
    // limit = N_cb of this K (0: none).  One plan uses a K with one limit only (the transport block's, or none: PLCF)
    // key: K for the transport block's tables, K + PLCF_KEY for the PLCF's (never limited, even where a tiny transport
    // block shares its K)
    constexpr uint32_t PLCF_KEY = 100000u;
    auto add_rm = [&](uint32_t K, uint32_t limit = 0, uint32_t key_off = 0) {
        if (rm_off.count(K + key_off)) return;
        spec::rm_table_t t;
        spec::build_rm_table(K, t, limit);
        rm_off[K + key_off] = (uint32_t)f.rm_arena.size();
        f.rm_arena.insert(f.rm_arena.end(), t.src.begin(), t.src.end());
        // soft buffer of a code block (3K + 12 int16): three streams d0 d1 d2 of K/2 words, word a = k * T + t =
        // (element t*L + k | element K/2 + t*L + k << 16) -- the window-transposed pair packing of the decoder's shared
        // arrays --, then the 12 termination values d0[K..K+3] d1[K..K+3] d2[K..K+3]
        const uint32_t L = tdec_window(K), T = K / L / 2, KH = K / 2, D = K + 4;
        std::vector<uint16_t> posT(3 * D);
        for (uint32_t st = 0; st < 3; ++st) {
            for (uint32_t a = 0; a < KH; ++a) {
                const uint32_t k = a / T, tt = a % T, i = tt * L + k;
                posT[st * K + 2 * a] = t.pos[st * D + i];
                posT[st * K + 2 * a + 1] = t.pos[st * D + i + KH];
            }
            for (uint32_t j = 0; j < 4; ++j) posT[3 * K + 4 * st + j] = t.pos[st * D + K + j];
        }
        // one copy per redundancy version with the start offset folded in: entry = index of the received LLR (mod N,
        // first pass over the circular buffer) that lands on this soft-buffer position
        // A soft-buffer position outside a limited circular buffer (posT >= Nc) is never sent: entry 0xffff, which the
        // gather of k_derm skips
        const uint32_t N = 3 * D, Nc = t.Nc;
        for (int rv = 0; rv < 4; ++rv)
            for (uint32_t n = 0; n < N; ++n)
                f.rm_arena.push_back(posT[n] < Nc ? (uint16_t)((posT[n] + Nc - t.start[rv]) % Nc) : (uint16_t)0xffffu);
        // and the inverse per redundancy version: received LLR e (first pass over the circular buffer) -> soft-buffer
        // position; k_derm scatters the n_e <= N LLRs of a code block through it into a zeroed shared image of the soft
        // buffer (fewer random accesses than gathering all N positions at the usual code rates)
        for (int rv = 0; rv < 4; ++rv) {
            std::vector<uint16_t> inv(N, 0);
            for (uint32_t n = 0; n < N; ++n)
                if (posT[n] < Nc) inv[(posT[n] + Nc - t.start[rv]) % Nc] = (uint16_t)n;
            f.rm_arena.insert(f.rm_arena.end(), inv.begin(), inv.end());
        }
        rm_tab[K + key_off] = t;
    };
    auto add_xp = [&](uint32_t K, uint32_t L, uint32_t poly, uint32_t order) -> uint32_t {
        const uint32_t off = (uint32_t)f.xp_arena.size();
        for (uint32_t w = 0; w < K / L; ++w) f.xp_arena.push_back(gf2_xpow_host(K - (w + 1) * L, poly, order));
        return off;
    };

    // TX CRC tables: thread t of TX_FEC_THREADS owns bits [min(n, t c), min(n, t c + c)), c = ceil(n / threads); its partial
    // remainder is multiplied by x^(n - end + order) mod g (fec_tx.cu: block_crc)
    // unit = 1: chunks of bits (code block CRC over unpacked bits); unit = 8: chunks of whole bytes (message CRC)
    auto add_xp_tx = [&](uint32_t nbits, uint32_t poly, uint32_t order, uint32_t unit) -> uint32_t {
        const uint32_t off = (uint32_t)f.xp_arena.size();
        const uint32_t n = nbits / unit, chunk = (n + TX_FEC_THREADS - 1) / TX_FEC_THREADS;
        for (uint32_t t = 0; t < TX_FEC_THREADS; ++t) {
            const uint32_t b0 = std::min(n, t * chunk), b1 = std::min(n, b0 + chunk);
            f.xp_arena.push_back(gf2_xpow_host((n - b1) * unit + order, poly, order));
        }
        return off;
    };

    f.Kmax = 0;
    f.tdec_smem = 0;
    f.tdec_threads = 32;
    auto track = [&](uint32_t K, uint32_t L) {
        f.Kmax = std::max(f.Kmax, K);
        f.tdec_smem = std::max(f.tdec_smem, tdec_smem_bytes(K, L));
        f.tdec_threads = std::max(f.tdec_threads, ((K / L / 2 + 31) / 32) * 32);
    };

    // ---- PDC
    const uint32_t Qm = ps.mcs.N_bps, Gp = ps.G / Qm, gamma = Gp % s.C;
    uint32_t rp_tx = 0, wp_tx = 0, out_byte = 0, soft_off = 0, saved_off = 0;
    std::map<uint32_t, uint32_t> xp_off;
    for (uint32_t cb = 0; cb < s.C; ++cb) {
        const uint32_t K = cb < s.C2 ? s.K2 : s.K1;
        const uint32_t Kidx = cb < s.C2 ? s.K2_idx : s.K1_idx;
        const uint32_t rlen = s.C > 1 ? K - 24 : K;
        const bool last = cb + 1 == s.C;
        uint32_t f1, f2;
        spec::qpp_params(Kidx, f1, f2);
        add_rm(K, N_IR ? std::max(N_IR / s.C, 1u) : 0u);   // N_cb = min(floor(N_IR / C), K_w)
        const uint32_t L = tdec_window(K);
        track(K, L);
        const uint32_t poly = s.C > 1 ? spec::CRC24B_POLY : spec::CRC24A_POLY;
        if (!xp_off.count(K)) xp_off[K] = add_xp(K, L, poly, 24);

        // encoder (pdc_enc.cpp:168-175)
        txcb_desc_t t{};
        t.K = K; t.f1 = f1; t.f2 = f2;
        t.rp = rp_tx;
        t.n_data = last ? rlen - 24 : rlen;
        if (last) { t.tb_crc_bits = 24; t.tb_crc_poly = spec::CRC24A_POLY; t.tb_crc_len = s.tbs; }
        t.cb_crc = s.C > 1;
        t.wp = wp_tx;
        t.n_e = (cb <= s.C - gamma - 1) ? Qm * (Gp / s.C) : Qm * ((Gp + s.C - 1) / s.C);
        t.n_pad = 0;
        t.rm_src_off = rm_off[K];
        for (int rv = 0; rv < 4; ++rv) t.rm_start[rv] = rm_tab[K].start[rv];
        t.Nc = rm_tab[K].Nc;
        t.xp_tb_off = t.tb_crc_bits ? add_xp_tx(t.tb_crc_len, t.tb_crc_poly, t.tb_crc_bits, 8) : 0;
        t.xp_cb_off = t.cb_crc ? add_xp_tx(t.n_data + t.tb_crc_bits, spec::CRC24B_POLY, 24, 1) : 0;
        f.tx_cbs.push_back(t);
        rp_tx += t.n_data;
        wp_tx += t.n_e;

        // decoder (pdc_enc.cpp:322-337)
        cb_desc_t r{};
        r.K = K; r.L = L; r.P = K / L; r.f1 = f1; r.f2 = f2;
        r.Linv = (uint32_t)(((1ull << 32) + L - 1) / L);
        const uint32_t n_e = Qm * (Gp / s.C);
        r.rp = cb * n_e;
        r.n_e = n_e;
        if (cb > s.C - gamma) {
            r.n_e = n_e + Qm;
            r.rp = (s.C - gamma) * n_e + (cb - (s.C - gamma)) * r.n_e;
        }
        r.out_byte = out_byte;
        r.rlen_byte = rlen / 8;
        r.soft_off = soft_off;
        r.rm_off = rm_off[K] + 3 * (K + 4);  // pos[]
        r.rm_inv_off = r.rm_off + 4 * 3 * (K + 4);
        for (int rv = 0; rv < 4; ++rv) r.rm_start[rv] = rm_tab[K].start[rv];
        r.Nc = rm_tab[K].Nc;
        r.crc_poly = poly;
        r.crc_order = 24;
        r.xp_off = xp_off[K];
        r.saved_off = saved_off;
        f.rx_cbs.push_back(r);
        out_byte += rlen / 8;
        soft_off += (3 * (K + 4) + 7) & ~7u;  // code blocks start 16-byte aligned (16-byte loads in the decoder)
        saved_off += rlen / 8;
    }
    if (wp_tx != ps.G || rp_tx != s.tbs || s.tbs % 8 != 0) return DNRP_ERR_INVALID;  // (whole bytes: fec_tx.cu CRC chunks)
    f.soft_stride = (soft_off + 7) & ~7u;
    f.saved_stride = (saved_off + 15) & ~15u;
    f.tb_stride = (out_byte + 15) & ~15u;          // N_TB_byte + 3 (TB CRC) rounded up
    f.dbits_stride = ((ps.G + 7) / 8 + 3) & ~3u;

    // TB-level CRC24A over tbs + 24 bits, chunked by bytes
    {
        const uint32_t nbytes = s.tbs / 8 + 3, ch = f.tb_chunk_bytes;
        const uint32_t nchunk = (nbytes + ch - 1) / ch;
        for (uint32_t i = 0; i < nchunk; ++i) {
            const uint32_t b1 = std::min(nbytes, (i + 1) * ch);
            f.xp_tb.push_back(gf2_xpow_host(8 * (nbytes - b1), spec::CRC24A_POLY, 24));
        }
    }

    // ---- PCC: PLCF type 1 (40 bits) / type 2 (80 bits) + CRC16, K = 56 / 96, E = 196, rv 0
    for (uint32_t ty = 0; ty < 2; ++ty) {
        const uint32_t nb = ty == 0 ? 40 : 80, K = nb + 16;
        const int Kidx = spec::tc_cb_index(K);
        uint32_t f1, f2;
        spec::qpp_params(Kidx, f1, f2);
        add_rm(K, 0, PLCF_KEY);
        const uint32_t L = tdec_window(K);
        track(K, L);
        txcb_desc_t t{};
        t.K = K; t.f1 = f1; t.f2 = f2;
        t.rp = 0; t.n_data = nb;
        t.tb_crc_bits = 16; t.tb_crc_poly = spec::CRC16_POLY; t.tb_crc_len = nb;
        t.cb_crc = 0;
        t.wp = 0; t.n_e = spec::PCC_BITS; t.n_pad = 4;  // the reference scrambles 200 bits (pcc_enc.cpp:212)
        t.rm_src_off = rm_off[K + PLCF_KEY];
        for (int rv = 0; rv < 4; ++rv) t.rm_start[rv] = rm_tab[K + PLCF_KEY].start[rv];
        t.Nc = rm_tab[K + PLCF_KEY].Nc;
        t.xp_tb_off = add_xp_tx(t.tb_crc_len, t.tb_crc_poly, t.tb_crc_bits, 8);
        t.xp_cb_off = 0;
        f.tx_pcc[ty] = t;
        cb_desc_t r{};
        r.K = K; r.L = L; r.P = K / L; r.f1 = f1; r.f2 = f2;
        r.Linv = (uint32_t)(((1ull << 32) + L - 1) / L);
        r.rp = 0; r.n_e = spec::PCC_BITS;
        r.out_byte = ty * 16;      // plcf_out [packet][2][16]
        r.rlen_byte = nb / 8;
        r.soft_off = ty * 320;     // soft_pcc [packet][2][320]
        r.rm_off = rm_off[K + PLCF_KEY] + 3 * (K + 4);
        r.rm_inv_off = r.rm_off + 4 * 3 * (K + 4);
        for (int rv = 0; rv < 4; ++rv) r.rm_start[rv] = rm_tab[K + PLCF_KEY].start[rv];
        r.Nc = rm_tab[K + PLCF_KEY].Nc;
        r.crc_poly = spec::CRC16_POLY;
        r.crc_order = 16;
        r.xp_off = add_xp(K, L, spec::CRC16_POLY, 16);
        r.saved_off = 0;
        f.rx_pcc[ty] = r;
    }
    return DNRP_OK;
}

}  // namespace dnrp

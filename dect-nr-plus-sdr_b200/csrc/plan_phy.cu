// plan_phy.cu -- host-side construction of the OFDM / receiver plan of one packet geometry: flat
// resource-element tables, STF signs, codebooks, FFT twiddles, Wiener look-up tables; uploaded once.
#include <algorithm>
#include <cmath>
#include <cstring>
#include "ctx.hpp"
#include "fft.cuh"
#include "spec_wiener.hpp"

namespace dnrp {

int build_phy_plan(dnrp_ctx* c) {
    const spec::packet_sizes_t& ps = c->ps;
    const uint32_t os = c->cfg.oversampling;
    phy_dims_t& d = c->dims;
    std::memset(&d, 0, sizeof(d));
    d.u = ps.psdef.u; d.b = ps.psdef.b;
    d.N_eff_TX = ps.tm.N_eff_TX; d.N_TS = ps.tm.N_TS; d.N_TX = ps.tm.N_TX; d.N_SS = ps.tm.N_SS;
    d.N_RX = c->cfg.N_RX;
    d.N_bps = ps.mcs.N_bps;
    d.txdiv = (ps.tm.index == 1 || ps.tm.index == 5 || ps.tm.index == 10) ? 1 : 0;
    d.N_OCC = ps.N_b_OCC; d.Nk = ps.N_b_OCC + 1; d.NP = ps.N_b_OCC / 4;
    d.N_os = ps.N_b_DFT * os; d.cp_os = ps.N_b_CP * os; d.stf_cp_os = ps.N_samples_STF_CP_only * os;
    d.stf_len_os = ps.N_samples_STF * os; d.sym_len_os = ps.N_samples_OFDM_symbol * os;
    d.n_pat = spec::stf_n_pattern(d.u); d.pat_len_os = 16 * d.b * os;
    d.N_DF = ps.N_DF_symb; d.N_step = spec::get_N_step(d.N_eff_TX);
    d.N_PDC_subc = ps.N_PDC_subc; d.G = ps.G;
    d.n_no_gi_os = ps.N_samples_packet_no_GI * os; d.n_packet_os = ps.N_samples_packet * os;
    d.txdiv_modulo = spec::txdiv_modulo(d.N_TS);
    {
        // 64 ... 4096 in one pass structure, 6144 ... 16384 four-step in place (fft.cuh; tx_rx.hpp:66-69 lists the sizes
        // the reference plans); radices 2 and 3 only
        uint32_t r = d.N_os;
        while (r % 2 == 0) r /= 2;
        while (r % 3 == 0) r /= 3;
        if (r != 1 || d.N_os > 16384) {
            set_last_error("FFT size " + std::to_string(d.N_os) + " (64 * beta * oversampling) must be 2^a 3^b and at most 16384");
            return DNRP_ERR_INVALID;
        }
    }

    spec::re_map_t& m = c->re_map;
    if (!spec::build_re_map(d.b, d.N_eff_TX, d.u, ps.N_PACKET_symb, m)) {
        set_last_error("resource element map failed");
        return DNRP_ERR_INVALID;
    }
    if (m.pdc_lk.size() != ps.N_PDC_subc || m.pcc_lk.size() != spec::PCC_CELLS) {
        set_last_error("resource element counts inconsistent");
        return DNRP_ERR_INVALID;
    }
    d.n_drs = m.n_drs_symb;

    // codebooks
    uint32_t n_idx = 0;
    spec::codebook(d.N_TS, d.N_TX, 0, &n_idx);
    if (n_idx == 0) {
        set_last_error("no codebook for (N_TS, N_TX)");
        return DNRP_ERR_INVALID;
    }
    d.n_cb_idx = n_idx;
    std::vector<int8_t> W;
    std::vector<float> Ws, Wso;
    for (uint32_t i = 0; i < n_idx; ++i) {
        const int8_t* w = spec::codebook(d.N_TS, d.N_TX, i, nullptr);
        W.insert(W.end(), w, w + d.N_TX * d.N_TS);
        Ws.push_back(spec::codebook_scale(w, d.N_TX * d.N_TS));
        Wso.push_back(spec::codebook_scale_optimal_dac(d.N_TS, d.N_TX, i));
    }
    std::vector<float2> tw(d.N_os);
    for (uint32_t k = 0; k < d.N_os; ++k) {
        const double a = -2.0 * M_PI * (double)k / (double)d.N_os;
        tw[k] = make_float2((float)std::cos(a), (float)std::sin(a));
    }
    bool ok = true;
#define UP(x) ok = ok && ((x) == cudaSuccess)
    UP(c->d_cell.upload(m.cell));
    {
        // transmitter look-ups: stream pair of every PCC / PDC cell under transmit diversity (the pair index is
        // (symbol index / 2) mod txdiv_modulo, transmit_diversity_precoding.cpp:45-72) and the constellations
        static const uint8_t a4[6] = {0, 2, 0, 1, 0, 1}, b4[6] = {1, 3, 2, 3, 3, 2};
        static const uint8_t a8[12] = {0, 2, 4, 6, 0, 1, 2, 3, 0, 1, 4, 5}, b8[12] = {1, 3, 5, 7, 4, 5, 6, 7, 2, 3, 6, 7};
        std::vector<uint8_t> ab(m.cell.size(), 0);
        for (size_t i = 0; i < m.cell.size(); ++i) {
            const uint32_t cw = m.cell[i], type = cw >> 30, idx = cw & 0x03ffffffu;
            if ((type != 2 && type != 3) || d.N_TS == 1) continue;
            const uint32_t im = (idx >> 1) % d.txdiv_modulo;
            const uint32_t A = d.N_TS == 2 ? 0u : (d.N_TS == 4 ? a4[im] : a8[im]);
            const uint32_t Bv = d.N_TS == 2 ? 1u : (d.N_TS == 4 ? b4[im] : b8[im]);
            ab[i] = (uint8_t)(A | Bv << 4);
        }
        UP(c->d_cell_ab.upload(ab));
        // the arithmetic of the former device function, in float: (1-2b0)(2^(h-1) - (1-2b2)(2^(h-2) - ...)) * norm
        std::vector<float2> lut(342 + 1024);
        for (uint32_t nb : {1u, 2u, 4u, 6u, 8u, 10u}) {
            const uint32_t base = qam_lut_base(nb), half = nb >> 1;
            for (uint32_t v = 0; v < (1u << nb); ++v) {
                if (nb == 1) {
                    const float a = v ? -0.70710678118654752f : 0.70710678118654752f;
                    lut[base + v] = make_float2(a, a);
                    continue;
                }
                float re = 0.f, im = 0.f;
                for (int i = (int)half - 1; i >= 1; --i) {
                    const float bi = ((v >> (nb - 1 - 2 * i)) & 1u) ? -1.f : 1.f;
                    const float bq = ((v >> (nb - 2 - 2 * i)) & 1u) ? -1.f : 1.f;
                    const float cc = (float)(1u << (half - 1 - i));
                    re = bi * (cc - re);
                    im = bq * (cc - im);
                }
                const float c0 = (float)(1u << (half - 1));
                re = (((v >> (nb - 1)) & 1u) ? -1.f : 1.f) * (c0 - re);
                im = (((v >> (nb - 2)) & 1u) ? -1.f : 1.f) * (c0 - im);
                const float norm = half == 1 ? 0.70710678118654752f
                                 : (half == 2 ? 0.31622776601683794f
                                 : (half == 3 ? 0.15430334996209191f : (half == 4 ? 0.07669649888473704f : 0.03829197905337418f)));
                lut[base + v] = make_float2(re * norm, im * norm);
            }
        }
        UP(c->d_qam_lut.upload(lut));
    }
    UP(c->d_pcc_lk.upload(m.pcc_lk));
    UP(c->d_pdc_lk.upload(m.pdc_lk));
    UP(c->d_stf_sign.upload(spec::build_stf_sign(d.b, d.N_eff_TX)));
    UP(c->d_W.upload(W));
    UP(c->d_W_scale.upload(Ws));
    UP(c->d_W_scale_opt.upload(Wso));
    UP(c->d_twiddle.upload(tw));
    // the same twiddles per pass, lane-contiguous (a warp's load then touches one or two 128-byte lines instead of up to
    // sixteen: those scattered loads were half of the FFT kernels' load/store wavefronts)
    if (d.N_os <= 4096) {
        UP(c->d_twiddle_p4.upload(fft_pass_tables(tw, false)));
        UP(c->d_twiddle_p16.upload(fft_pass_tables(tw, true)));
    }
    // ---- resampling to the hardware rate (tx.cpp:519-560 sample counts; resampler.cpp design)
    {
        const uint32_t L = c->cfg.resample_L, M = c->cfg.resample_M;
        c->resampling = L != M;
        if (!spec::design_resampler(L, M, os, c->rs_tx) || !spec::design_resampler(M, L, os, c->rs_rx)) {
            set_last_error("no resampler filter definition for this oversampling (resampler_param.hpp:71-86)");
            return DNRP_ERR_INVALID;
        }
        if (c->resampling) {
            if (L < M) {
                set_last_error("the hardware rate must not be below the DECT rate (resample_L >= resample_M)");
                return DNRP_ERR_INVALID;
            }
            if (d.n_packet_os % M != 0) {
                set_last_error("oversampled packet length with GI not a multiple of M (tx.cpp:537)");
                return DNRP_ERR_INVALID;
            }
            c->n_no_gi_rs = (uint32_t)(((uint64_t)d.n_no_gi_os * L + M - 1) / M);
            c->n_packet_rs = d.n_packet_os / M * L;
            c->stf_len_rs = (uint32_t)((uint64_t)d.stf_len_os * L / M);
            if (c->n_no_gi_rs > c->rs_tx.n_after(d.n_no_gi_os) || c->n_no_gi_rs >= c->n_packet_rs) {
                set_last_error("resampler creates fewer samples than the packet needs (tx.cpp:530-545)");
                return DNRP_ERR_INVALID;
            }
            c->tx_os_stride = (d.n_no_gi_os + 3) & ~3u;
            UP(c->d_iq_tx_os.alloc((size_t)c->B * d.N_TX * c->tx_os_stride));
            UP(c->d_rs_tx_taps.upload(c->rs_tx.taps));
            UP(c->d_rs_rx_taps.upload(c->rs_rx.taps));
        } else {
            c->n_no_gi_rs = d.n_no_gi_os; c->n_packet_rs = d.n_packet_os; c->stf_len_rs = d.stf_len_os;
        }
    }
    c->tx_stride = (c->n_packet_rs + 3) & ~3u;
    UP(c->d_iq_tx.alloc((size_t)c->B * d.N_TX * c->tx_stride));
    UP(c->d_tx_meta.alloc(c->B));

    // ---- receiver
    d.chestim_lr = c->cfg.chestim_mode_lr ? 1 : 0;
    d.chestim_stride = c->cfg.chestim_mode_lr_t_stride ? c->cfg.chestim_mode_lr_t_stride : 1;
    d.N_step_inv = (uint32_t)(((1ull << 32) + d.N_step - 1) / d.N_step);
    d.chestim_stride_inv = d.chestim_stride > 1 ? (uint32_t)(((1ull << 32) + d.chestim_stride - 1) / d.chestim_stride) : 0;
    {
        // rx_synced.cpp:1027-1032: stage length 6 / 11 / 12 symbols for N_step 5 / 10 / 10 (N_eff_TX 8)
        const uint32_t ps_len = d.N_eff_TX <= 2 ? 6 : (d.N_eff_TX == 4 ? 11 : 12);
        d.n_full_ps = d.N_DF >= ps_len - d.N_step ? (d.N_DF - (ps_len - d.N_step)) / d.N_step : 0;
        if (d.n_full_ps + 1 > d.n_drs) d.n_full_ps = d.n_drs > 0 ? d.n_drs - 1 : 0;
        d.pcc_l_max = 0;
        for (uint32_t v : m.pcc_lk) d.pcc_l_max = std::max(d.pcc_l_max, v >> 16);
    }
    std::vector<uint32_t> pcc_plan_h, pdc_plan_h;
    {
        // which interpolation serves (stream ts, symbol l): closed form of rx_synced.cpp:1027-1152,
        // evaluated once per demapper unit here instead of per cell and packet on the device
        auto plan_word = [&](uint32_t ts, uint32_t l, bool pcc_stage) -> uint32_t {
            if (pcc_stage || !d.chestim_lr || d.n_full_ps == 0) {
                uint32_t n = (l - 1) / d.N_step;
                if (n >= d.n_drs) n = d.n_drs - 1;
                const uint32_t j = d.N_eff_TX == 8 ? 2 * n + (ts >= 4 ? 1 : 0) : n;
                const uint32_t pc = ((ts & 3) + 2 * (n & 1)) & 3;
                return ce_plan_word(pc, j, n, 0, ts, 0);
            }
            uint32_t ps_ = 0, r = 0;
            if (l > 1) {
                ps_ = (l - 2) / d.N_step;
                if (ps_ > d.n_full_ps - 1) ps_ = d.n_full_ps - 1;
                r = l - 1 - ps_ * d.N_step;
            }
            const uint32_t rs = ps_ == 0 ? 0 : 1, re = r < d.N_step ? r : d.N_step;
            const uint32_t rstar = rs + ((re - rs) / d.chestim_stride) * d.chestim_stride;
            const uint32_t j = d.N_eff_TX == 8 ? 2 * (ps_ + 1) + 1 : ps_ + 1;
            const uint32_t pc = ((ts & 3) + 2 * (ps_ & 1)) & 3;
            return ce_plan_word(rstar * 4 + pc, j, ps_, pc >= 2 ? 1 : 0, ts, 1);
        };
        auto pair_streams = [&](uint32_t im, uint32_t& A, uint32_t& B) {   // Table 6.3.3.2 (phy_rx.cu: txdiv_pair_rx)
            static const uint8_t a4[6] = {0, 2, 0, 1, 0, 1}, b4[6] = {1, 3, 2, 3, 3, 2};
            static const uint8_t a8[12] = {0, 2, 4, 6, 0, 1, 2, 3, 0, 1, 4, 5}, b8[12] = {1, 3, 5, 7, 4, 5, 6, 7, 2, 3, 6, 7};
            if (d.N_eff_TX == 2) { A = 0; B = 1; }
            else if (d.N_eff_TX == 4) { A = a4[im]; B = b4[im]; }
            else { A = a8[im]; B = b8[im]; }
        };
        auto build = [&](const std::vector<uint32_t>& lk, bool pcc_stage) {
            const bool txdiv = d.N_eff_TX != 1;
            const uint32_t units = (uint32_t)lk.size() / (txdiv ? 2 : 1);
            std::vector<uint32_t> t(2 * (size_t)units, 0u);
            for (uint32_t u = 0; u < units; ++u) {
                const uint32_t l = lk[txdiv ? 2 * u : u] >> 16;
                uint32_t A = 0, B = 0;
                if (txdiv) pair_streams(u % d.txdiv_modulo, A, B);
                t[2 * u] = plan_word(A, l, pcc_stage);
                t[2 * u + 1] = plan_word(B, l, pcc_stage);
            }
            return t;
        };
        // spatial multiplexing: every cell carries all N_SS streams, one plan word per (cell, stream)
        auto build_sm = [&](const std::vector<uint32_t>& lk) {
            std::vector<uint32_t> t(lk.size() * (size_t)d.N_SS, 0u);
            for (size_t q = 0; q < lk.size(); ++q)
                for (uint32_t s_ = 0; s_ < d.N_SS; ++s_) t[q * d.N_SS + s_] = plan_word(s_, lk[q] >> 16, false);
            return t;
        };
        for (uint32_t ts = 0; ts < 8; ++ts) d.last_plan[ts] = ts < d.N_eff_TX ? plan_word(ts, d.N_DF, false) : 0u;
        pcc_plan_h = build(m.pcc_lk, true);
        pdc_plan_h = d.N_SS > 1 ? build_sm(m.pdc_lk) : build(m.pdc_lk, false);
        UP(c->d_pcc_plan.upload(pcc_plan_h));
        UP(c->d_pdc_plan.upload(pdc_plan_h));
    }
    static const double nu_max[3] = {100.0, 100.0, 500.0}, tau_rms[3] = {0.1e-6, 0.1e-6, 1.0e-6},
                        snr_db[3] = {-5.0, 15.0, 35.0};
    static const uint32_t n_lr[3] = {14, 8, 3}, n_l[3] = {7, 4, 2};
    for (uint32_t i = 0; i < 3 && ok; ++i) {
        spec::wiener_lut_t Ll, Llr;
        if (!spec::build_wiener_lut(d.b, d.u, 0, nu_max[i], tau_rms[i], snr_db[i], n_l[i], Ll) ||
            !spec::build_wiener_lut(d.b, d.u, d.N_step, nu_max[i], tau_rms[i], snr_db[i], n_lr[i], Llr)) {
            set_last_error("Wiener LUT construction failed");
            return DNRP_ERR_INVALID;
        }
        UP(c->d_wiener_l[i].idx_pilot.upload(Ll.idx_pilot));
        UP(c->d_wiener_l[i].idx_weight.upload(Ll.idx_weight));
        UP(c->d_wiener_l[i].weights.upload(Ll.weights));
        UP(c->d_wiener_lr[i].idx_pilot.upload(Llr.idx_pilot));
        UP(c->d_wiener_lr[i].idx_weight.upload(Llr.idx_weight));
        UP(c->d_wiener_lr[i].weights.upload(Llr.weights));
        d.wiener_l[i] = {c->d_wiener_l[i].idx_pilot.p, c->d_wiener_l[i].idx_weight.p, c->d_wiener_l[i].weights.p,
                         Ll.n_interp, Ll.n_t};
        d.wiener_lr[i] = {c->d_wiener_lr[i].idx_pilot.p, c->d_wiener_lr[i].idx_weight.p,
                          c->d_wiener_lr[i].weights.p, Llr.n_interp, Llr.n_t};
        // transmit diversity: merged window per (unit, stream slot), the arithmetic of chan_est_rx_pair done once here
        auto build_pair = [&](const std::vector<uint32_t>& lk, const std::vector<uint32_t>& plan, dnrp_ctx::pair_bufs_t& buf,
                              phy_dims_t::pair_lut_t& out) -> int {
            const uint32_t units = (uint32_t)lk.size() / 2;
            out = {nullptr, nullptr, units};
            if (units == 0) return DNRP_OK;
            const bool lr = ((plan[0] >> 21) & 1u) != 0u;     // one mode per list (plan_word above)
            const spec::wiener_lut_t& L = lr ? Llr : Ll;
            const uint32_t NT = L.n_interp + 1;
            std::vector<uint16_t> info(2 * (size_t)units, 0xffffu);
            std::vector<float> w((size_t)NT * 2 * units, 0.f);
            for (uint32_t u = 0; u < units; ++u)
                for (uint32_t s_ = 0; s_ < 2; ++s_) {
                    const uint32_t pw = plan[2 * u + s_], base = pw & 63u;
                    if ((((pw >> 21) & 1u) != 0u) != lr) return DNRP_ERR_INVALID;
                    const uint32_t k0 = lk[2 * u] & 0xffffu, k1 = lk[2 * u + 1] & 0xffffu;
                    const uint32_t o0 = base * d.Nk + k0, o1 = base * d.Nk + k1;
                    const uint32_t m0 = L.idx_pilot[o0], m1 = L.idx_pilot[o1];
                    const uint32_t mmin = std::min(m0, m1), d0 = m0 - mmin, d1 = m1 - mmin;
                    if (d0 + d1 > 1u || mmin > 1023u) continue;
                    const uint32_t nt = L.n_interp + d0 + d1;
                    const float* w0 = L.weights.data() + (size_t)L.idx_weight[o0] * L.n_interp;
                    const float* w1 = L.weights.data() + (size_t)L.idx_weight[o1] * L.n_interp;
                    for (uint32_t t = 0; t < nt; ++t) {
                        const float a0 = (t >= d0 && t - d0 < L.n_interp) ? w0[t - d0] : 0.f;
                        const float a1 = (t >= d1 && t - d1 < L.n_interp) ? w1[t - d1] : 0.f;
                        w[((size_t)t * 2 + s_) * units + u] = 0.5f * (a0 + a1);
                    }
                    info[(size_t)s_ * units + u] = (uint16_t)(mmin | nt << 10);
                }
            UP(buf.info.upload(info));
            UP(buf.w.upload(w));
            out = {buf.info.p, buf.w.p, units};
            return DNRP_OK;
        };
        d.pcc_pair[i] = {nullptr, nullptr, 0};
        d.pdc_pair[i] = {nullptr, nullptr, 0};
        if (d.N_eff_TX != 1) {
            if (build_pair(m.pcc_lk, pcc_plan_h, c->d_pcc_pair[i], d.pcc_pair[i]) != DNRP_OK ||
                (d.N_SS == 1 && build_pair(m.pdc_lk, pdc_plan_h, c->d_pdc_pair[i], d.pdc_pair[i]) != DNRP_OK)) {
                set_last_error("merged Wiener window tables: construction failed");
                return DNRP_ERR_INVALID;
            }
        }
    }
    // RX window: the packet plus room for two STFs (the detection ignores the first STF + 1 pattern of a window
    // and the fine search looks one pattern to either side of the coarse peak)
    c->rx_stride = (c->n_packet_rs + 2 * c->stf_len_rs + 3) & ~3u;   // hardware rate
    if (c->resampling) {
        // detection / coarse peak run on the window resampled to the DECT rate (no flush: only outputs whose newest
        // input sample lies inside the window), the synchronised receiver on the packet resampled from its fine peak
        const spec::resampler_design_t& r = c->rs_rx;
        c->sync_len = (uint32_t)(((uint64_t)c->rx_stride * r.L - 1 - r.delay) / r.M + 1);
        c->sync_stride = (c->sync_len + 3) & ~3u;
        c->rs_stride = (d.n_no_gi_os + 3) & ~3u;
        UP(c->d_iq_sync.alloc((size_t)c->B * d.N_RX * c->sync_stride));
        UP(c->d_iq_rs.alloc((size_t)c->B * d.N_RX * c->rs_stride));
    } else {
        c->sync_len = c->sync_stride = c->rx_stride;
    }
    // packet synchronisation: time-domain STF of a single stream for every N_eff_TX the receiver has to
    // tell apart (stf_template.cpp:81-204 for L = M = 1): scale 1/sqrt(N_OCC/4), unnormalised inverse DFT,
    // long CP, cover sequence.  Only every 4th subcarrier is occupied, so a direct sparse IDFT is cheap.
    {
        std::vector<float2> tpl;
        c->sync_n_tpl = 0;
        const double scale = 1.0 / std::sqrt((double)(d.N_OCC / 4)), pi = 3.14159265358979323846;
        for (uint32_t ntx = 1; ntx <= d.N_eff_TX; ntx <<= 1, ++c->sync_n_tpl) {
            const std::vector<int8_t> sgn = spec::build_stf_sign(d.b, ntx);
            std::vector<double> xr(d.N_os, 0.0), xi(d.N_os, 0.0);
            for (uint32_t k = 0; k < d.Nk; ++k) {
                if (!sgn[k]) continue;
                const int bin = (int)k - (int)(d.N_OCC / 2);
                const double yr = sgn[k] * scale * std::sqrt(0.5), yi = yr;  // +-exp(j pi/4) * scale
                for (uint32_t n = 0; n < d.N_os; ++n) {
                    const long long m = ((long long)bin * n) % (long long)d.N_os;
                    const double ph = 2.0 * pi * (double)m / (double)d.N_os;
                    const double cs = std::cos(ph), sn = std::sin(ph);
                    xr[n] += yr * cs - yi * sn;
                    xi[n] += yr * sn + yi * cs;
                }
            }
            std::vector<float2> t(d.stf_len_os);
            for (uint32_t n = 0; n < d.stf_len_os; ++n) {
                const uint32_t src = (n + d.N_os - d.stf_cp_os % d.N_os) % d.N_os;
                const float cv = spec::stf_cover(n / d.pat_len_os);
                t[n] = make_float2((float)xr[src] * cv, (float)xi[src] * cv);
            }
            if (!c->resampling) {
                tpl.insert(tpl.end(), t.begin(), t.end());
            } else {
                // through the TX resampler, truncated to stf_len * L / M (stf_template.cpp:169-194)
                const spec::resampler_design_t& r = c->rs_tx;
                for (uint32_t mo = 0; mo < c->stf_len_rs; ++mo) {
                    const uint64_t pp = (uint64_t)r.delay + (uint64_t)mo * r.M;
                    const uint32_t l = (uint32_t)(pp % r.L);
                    const int64_t i0 = (int64_t)(pp / r.L);
                    float re = 0.f, im = 0.f;
                    for (uint32_t j = 0; j < r.subfilter_length; ++j) {
                        const int64_t i = i0 - j;
                        if (i < 0 || i >= (int64_t)d.stf_len_os) continue;
                        re += r.taps[j * r.L + l] * t[i].x;
                        im += r.taps[j * r.L + l] * t[i].y;
                    }
                    tpl.push_back(make_float2(re, im));
                }
            }
        }
        UP(c->d_sync_tpl.upload(tpl));
        {
            // fast cross-correlation: smallest FFT size 2^a or 3 * 2^a (the sizes fft.cuh handles in one block) that
            // covers lags + template length - 1; spectra by a direct DFT in double (init time only)
            const uint32_t S = c->stf_len_rs;
            const uint32_t sl = (uint32_t)((uint64_t)d.pat_len_os * c->cfg.resample_L / c->cfg.resample_M);
            const uint32_t need = 2 * sl + 1 + S - 1;
            uint32_t N = 0;
            for (uint32_t cand = 64; cand <= 4096 && !N; cand *= 2) {
                if (cand >= need) N = cand;
                else if (cand / 2 * 3 >= need && cand / 2 * 3 <= 4096 && cand / 2 * 3 < cand * 2) N = cand / 2 * 3;
            }
            c->sync_fft_n = N;
            if (N) {
                std::vector<float2> twx(N), spec((size_t)c->sync_n_tpl * N);
                std::vector<double> cs(N), sn(N);
                for (uint32_t k = 0; k < N; ++k) {
                    const double a = -2.0 * pi * (double)k / (double)N;
                    cs[k] = std::cos(a); sn[k] = std::sin(a);
                    twx[k] = make_float2((float)cs[k], (float)sn[k]);
                }
                for (uint32_t t = 0; t < c->sync_n_tpl; ++t)
                    for (uint32_t k = 0; k < N; ++k) {
                        double re = 0.0, im = 0.0;
                        for (uint32_t i = 0; i < S; ++i) {
                            const uint32_t m = (uint32_t)(((uint64_t)k * i) % N);
                            const float2 v = tpl[(size_t)t * S + i];
                            re += v.x * cs[m] - v.y * sn[m];
                            im += v.x * sn[m] + v.y * cs[m];
                        }
                        spec[(size_t)t * N + k] = make_float2((float)(re / N), (float)(-im / N));   // conj / N
                    }
                UP(c->d_sync_tw.upload(twx));
                UP(c->d_sync_tw_p16.upload(fft_pass_tables(twx, true)));
                UP(c->d_sync_tpl_spec.upload(spec));
            }
        }
        c->sync_step = d.pat_len_os / 4;
        c->sync_n_steps = c->sync_len / c->sync_step;
        const size_t nb = (size_t)c->B * d.N_RX * c->sync_n_steps;
        UP(c->d_sync_c.alloc(nb));
        UP(c->d_sync_p.alloc(nb));
        UP(c->d_sync_xc.alloc((size_t)c->B * c->sync_n_tpl));
        UP(c->d_sync_result.alloc(c->B));
    }
    c->cls_stride = (d.N_eff_TX == 8 ? 2 : 1) * d.n_drs + 2;
    const size_t B = c->B;
    UP(c->d_iq_rx.alloc(B * d.N_RX * c->rx_stride));
    UP(c->d_Y.alloc(B * d.N_RX * d.N_DF * d.Nk));
    UP(c->d_hzf.alloc(B * d.N_RX * d.N_eff_TX * d.n_drs * d.NP));
    UP(c->d_hstf.alloc(B * d.N_RX * d.NP));
    UP(c->d_snr_part.alloc(B * d.N_DF * d.N_RX));
    UP(c->d_rx_state.alloc(B));
    UP(c->d_cls.alloc(B * c->cls_stride));
    UP(c->d_chan_meta.alloc(B));
    // MIMO report: W(1, N) for N = N_eff_TX and N = N_RX; the search starts at the first matrix without zero entries
    // (beamforming_and_antenna_port_mapping.hpp:110-119: index 2 for N = 2, 12 for N = 4)
    for (uint32_t side = 0; side < 2; ++side) {
        const uint32_t N = side == 0 ? d.N_eff_TX : d.N_RX;
        c->mimo_n[side] = 0;
        if (N != 2 && N != 4) continue;
        uint32_t cnt = 0;
        spec::codebook(1, N, 0, &cnt);
        std::vector<int8_t> Wm;
        std::vector<float> sc;
        for (uint32_t i = 0; i < cnt; ++i) {
            const int8_t* w = spec::codebook(1, N, i, nullptr);
            Wm.insert(Wm.end(), w, w + N);
            sc.push_back(spec::codebook_scale(w, N));
        }
        c->mimo_n[side] = cnt;
        c->mimo_first[side] = N == 2 ? 2 : 12;
        UP(c->d_mimo_W[side].upload(Wm));
        UP(c->d_mimo_scale[side].upload(sc));
    }
    UP(c->d_mimo.alloc(B));
    c->h_rx_state.resize(B);
#undef UP
    if (!ok) {
        set_last_error("device allocation failed (phy plan)");
        return DNRP_ERR_CUDA;
    }
    d.cell_ab = c->d_cell_ab.p; d.qam_lut = c->d_qam_lut.p;
    d.cell = c->d_cell.p; d.pcc_lk = c->d_pcc_lk.p; d.pdc_lk = c->d_pdc_lk.p;
    d.pcc_plan = c->d_pcc_plan.p; d.pdc_plan = c->d_pdc_plan.p;
    d.stf_sign = c->d_stf_sign.p; d.W = c->d_W.p; d.W_scale = c->d_W_scale.p; d.W_scale_opt = c->d_W_scale_opt.p;
    d.twiddle = c->d_twiddle.p;
    d.twiddle_p4 = c->d_twiddle_p4.p;     // nullptr for the four-step sizes above 4096
    d.twiddle_p16 = c->d_twiddle_p16.p;
    return DNRP_OK;
}

}  // namespace dnrp

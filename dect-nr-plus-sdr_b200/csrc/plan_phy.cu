// plan_phy.cu -- host-side construction of the OFDM / receiver plan of one packet geometry: flat
// resource-element tables, STF signs, codebooks, FFT twiddles, Wiener look-up tables; uploaded once.
#include <cmath>
#include "ctx.hpp"

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
    if (d.N_os > 4096) {
        set_last_error("FFT sizes above 4096 are not implemented yet");
        return DNRP_ERR_INVALID;
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
    UP(c->d_pcc_lk.upload(m.pcc_lk));
    UP(c->d_pdc_lk.upload(m.pdc_lk));
    UP(c->d_stf_sign.upload(spec::build_stf_sign(d.b, d.N_eff_TX)));
    UP(c->d_W.upload(W));
    UP(c->d_W_scale.upload(Ws));
    UP(c->d_W_scale_opt.upload(Wso));
    UP(c->d_twiddle.upload(tw));
    c->tx_stride = (d.n_packet_os + 3) & ~3u;
    UP(c->d_iq_tx.alloc((size_t)c->B * d.N_TX * c->tx_stride));
    UP(c->d_tx_meta.alloc(c->B));
#undef UP
    if (!ok) {
        set_last_error("device allocation failed (phy plan)");
        return DNRP_ERR_CUDA;
    }
    d.cell = c->d_cell.p; d.pcc_lk = c->d_pcc_lk.p; d.pdc_lk = c->d_pdc_lk.p;
    d.stf_sign = c->d_stf_sign.p; d.W = c->d_W.p; d.W_scale = c->d_W_scale.p; d.W_scale_opt = c->d_W_scale_opt.p;
    d.twiddle = c->d_twiddle.p;
    return DNRP_OK;
}

}  // namespace dnrp

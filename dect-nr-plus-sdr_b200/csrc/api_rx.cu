// api_rx.cu -- receiver side of the C ABI: batched rx_synced_t::demoddecod_rx_pcc / _pdc
// (reference lib/src/phy/rx/rx_synced/rx_synced.cpp:186-436) on the context's device buffers.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include "ctx.hpp"
#include "spec_channel.hpp"

using namespace dnrp;

#define DNRP_REQUIRE(cond, msg)             \
    do {                                    \
        if (!(cond)) {                      \
            dnrp::set_last_error(msg);      \
            return DNRP_ERR_INVALID;        \
        }                                   \
    } while (0)

namespace dnrp {

rx_params_t make_rx_params(dnrp_ctx* c) {
    rx_params_t p{};
    const phy_dims_t& d = c->dims;
    const size_t off = c->off;
    p.d = d;
    if (c->resampling) {
        p.iq = c->d_iq_rs.p + off * d.N_RX * c->rs_stride; p.iq_stride = c->rs_stride; p.t0_zero = 1;
    } else {
        p.iq = c->d_iq_rx.p + off * d.N_RX * c->rx_stride; p.iq_stride = c->rx_stride; p.t0_zero = 0;
    }
    p.state = c->d_rx_state.p + off;
    p.Y = c->d_Y.p + off * d.N_RX * d.N_DF * d.Nk;
    p.hzf = c->d_hzf.p + off * d.N_RX * d.N_eff_TX * d.n_drs * d.NP;
    p.hstf = c->d_hstf.p + off * d.N_RX * d.NP;
    p.snr_part = c->d_snr_part.p + off * d.N_DF * d.N_RX;
    p.cls = c->d_cls.p + off * c->cls_stride; p.cls_stride = c->cls_stride;
    p.llr_pcc = c->d_llr_pcc.p + off * 196;
    p.llr = c->d_llr.p + off * c->ps.G; p.llr_stride = c->ps.G;
    p.track = (c->track_cfo || c->track_sto) ? c->d_track.p + off : nullptr;
    p.track_cfo = c->track_cfo; p.track_sto = c->track_sto;
    return p;
}

// The reference interleaves FFTs, estimation and decoding symbol by symbol for latency; for a batch
// the whole front end of all packets runs first (every later stage only reads its results).
int run_rx_front(dnrp_ctx* c, uint32_t n) {
    const int rc = run_rx_resample(c, n);
    if (rc) return rc;
    const rx_params_t p = make_rx_params(c);
    DNRP_LAUNCH(c, DNRP_KERNEL_RX_STF, launch_rx_stf(p, n, c->stream));
    DNRP_LAUNCH(c, DNRP_KERNEL_RX_FFT, launch_rx_fft(p, n, c->stream));
    if (p.track) DNRP_LAUNCH(c, -1, launch_rx_track(p, n, c->stream));
    DNRP_LAUNCH(c, -1, launch_rx_snr_class(p, n, c->stream));
    return DNRP_OK;
}

int run_rx_pcc(dnrp_ctx* c, uint32_t n) {
    const rx_params_t p = make_rx_params(c);
    DNRP_LAUNCH(c, -1, launch_rx_demap(p, n, true, c->stream));
    return run_fec_decode_plcf(c, n);
}

int run_rx_pdc(dnrp_ctx* c, uint32_t n, bool have_slot, bool have_rv, bool have_reset) {
    if (c->dims.N_SS > 1 && c->dims.N_RX < c->dims.N_SS) {
        set_last_error("spatial multiplexing with " + std::to_string(c->dims.N_SS) + " streams needs at least as many receive antennas (N_RX = " +
                       std::to_string(c->dims.N_RX) + ")");
        return DNRP_ERR_INVALID;
    }
    const rx_params_t p = make_rx_params(c);
    DNRP_LAUNCH(c, DNRP_KERNEL_RX_DEMAP_PDC, launch_rx_demap(p, n, false, c->stream));
    mimo_params_t mp{};
    mp.d = c->dims; mp.hzf = p.hzf; mp.out = c->d_mimo.p + c->off;
    for (int s = 0; s < 2; ++s) {
        mp.W[s] = c->d_mimo_W[s].p; mp.W_scale[s] = c->d_mimo_scale[s].p;
        mp.n[s] = c->mimo_n[s]; mp.first[s] = c->mimo_first[s];
    }
    DNRP_LAUNCH(c, -1, launch_rx_mimo(mp, n, c->stream));
    return run_fec_decode_tb(c, n, have_slot, have_rv, have_reset);
}

int run_channel(dnrp_ctx* c, uint32_t n) {
    const phy_dims_t& d = c->dims;
    channel_params_t p{};
    p.N_TX = d.N_TX; p.N_RX = d.N_RX;
    p.tx = c->d_iq_tx.p; p.tx_stride = c->tx_stride;
    p.rx = c->d_iq_rx.p; p.rx_stride = c->rx_stride;
    p.n_samples = c->n_packet_rs;
    p.meta = c->d_chan_meta.p;
    p.sin_tab = c->d_chan_sin.p;
    p.tdl_min_h = c->chan_min_h;
    DNRP_LAUNCH(c, -1, launch_channel(p, n, c->stream));
    if (c->chan_any_tdl) {
        ++c->launches;  // setup + main kernel
        DNRP_LAUNCH(c, -1, launch_channel_tdl(p, n, c->stream));
    }
    return DNRP_OK;
}

int run_sync(dnrp_ctx* c, uint32_t n, uint32_t search_length, bool hand_over) {
    const phy_dims_t& d = c->dims;
    {
        // the coarse peak stage keeps prefix sums over two STFs in shared memory; at oversampling factors that make the
        // STF longer than about 8000 samples this no longer fits: refuse with a reason instead of a launch failure
        int lim = 0;
        DNRP_CUDA(cudaDeviceGetAttribute(&lim, cudaDevAttrMaxSharedMemoryPerBlockOptin, c->device));
        if (sync_peak_smem(d, c->sync_n_steps) > (size_t)lim || (!c->sync_fft_n && 2 * ((uint64_t)d.pat_len_os * c->cfg.resample_L / c->cfg.resample_M) + 1 > 4096)) {
            set_last_error("packet synchronisation is not available for an STF of " + std::to_string(d.stf_len_os) +
                           " samples (oversampling too high for the shared-memory search); hand sync reports to dnrp_rx_pcc_batch");
            return DNRP_ERR_INVALID;
        }
    }
    sync_params_t p{};
    p.d = d;
    const size_t off = c->off;
    const float2* iq_hw = c->d_iq_rx.p + off * d.N_RX * c->rx_stride;
    p.iq_x = iq_hw; p.x_stride = c->rx_stride; p.x_len = c->rx_stride;
    p.rs_L = c->cfg.resample_L; p.rs_M = c->cfg.resample_M;
    p.xc_S = c->stf_len_rs;
    p.xc_sl = (uint32_t)((uint64_t)d.pat_len_os * p.rs_L / p.rs_M);  // crosscorrelator.cpp:52-56
    p.n_no_gi_hw = c->n_no_gi_rs;
    if (c->resampling) {
        // rx_pacer_t::resample_single_unit on the whole window: hardware rate -> DECT rate, zero history at the start
        resample_params_t r{};
        r.L = c->rs_rx.L; r.M = c->rs_rx.M; r.SL = c->rs_rx.subfilter_length; r.delay = c->rs_rx.delay;
        r.taps = c->d_rs_rx_taps.p;
        r.in = iq_hw; r.in_stride = c->rx_stride; r.n_in = c->rx_stride;
        r.out = c->d_iq_sync.p + off * d.N_RX * c->sync_stride; r.out_stride = c->sync_stride;
        r.n_out = c->sync_len; r.n_fill = c->sync_stride;
        r.n_ant = d.N_RX;
        DNRP_LAUNCH(c, DNRP_KERNEL_RESAMPLE, launch_resample(r, n, c->stream));
        p.iq = r.out; p.iq_stride = c->sync_stride; p.len = c->sync_len;
    } else {
        p.iq = iq_hw; p.iq_stride = c->rx_stride; p.len = c->rx_stride;
    }
    p.n_ant = d.N_RX;  // RX_SYNC_PARAM_AUTOCORRELATOR_ANTENNA_LIMIT = 8 >= N_RX
    p.step = c->sync_step; p.n_steps = c->sync_n_steps;
    p.search_length = search_length ? search_length : p.len - d.stf_len_os - d.pat_len_os;
    // the detection accepts step j only if it starts inside search_length (j step < search_length, k_sync_peak): the steps
    // behind are neither summed nor scanned -- a caller who knows where a packet can still fit into the window
    // (search_length) does not pay for reading the rest of it
    p.n_steps_use = std::min(p.n_steps, (p.search_length + p.step - 1) / p.step + 1);
    if (c->fd_beta_thr > 0.f || c->fd_cfo_range >= 4) {
        p.fd_beta_thr = c->fd_beta_thr; p.fd_cfo_range = c->fd_cfo_range;
        p.fd_seq = c->d_fd_seq.p; p.fd_b = c->d_fd_b.p + off; p.fd_cfo = c->d_fd_cfo.p + off;
    }
    // autocorrelator_detection.cpp:60-62: minimum rms scaled with the sample rate relative to 30.72 MS/s
    p.rms_min = 0.005f * (float)std::sqrt((double)d.u * d.b * c->cfg.oversampling * 1.728e6 / 30.72e6);
    p.c_steps = c->d_sync_c.p + off * d.N_RX * c->sync_n_steps; p.p_steps = c->d_sync_p.p + off * d.N_RX * c->sync_n_steps;
    p.templates = c->d_sync_tpl.p; p.n_tpl = c->sync_n_tpl;
    p.xc = c->d_sync_xc.p + off * c->sync_n_tpl;
    p.xc_fft_n = c->sync_fft_n; p.tpl_spec = c->d_sync_tpl_spec.p; p.xc_twiddle = c->d_sync_tw.p;
    p.xc_twiddle_p16 = c->d_sync_tw_p16.p;
    p.result = c->d_sync_result.p + off;
    p.rx_state = hand_over ? c->d_rx_state.p + off : nullptr;
    c->launches += 3;  // launch_sync issues four kernels
    DNRP_LAUNCH(c, DNRP_KERNEL_SYNC, launch_sync(p, n, c->stream));
    return DNRP_OK;
}

// rx_synced.cpp:214-220 + rx_pacer.cpp: the resampler restarts with empty history at the fine peak
int run_rx_resample(dnrp_ctx* c, uint32_t n) {
    if (!c->resampling) return DNRP_OK;
    const phy_dims_t& d = c->dims;
    resample_params_t r{};
    r.L = c->rs_rx.L; r.M = c->rs_rx.M; r.SL = c->rs_rx.subfilter_length; r.delay = c->rs_rx.delay;
    r.taps = c->d_rs_rx_taps.p;
    const size_t off = c->off;
    r.in = c->d_iq_rx.p + off * d.N_RX * c->rx_stride; r.in_stride = c->rx_stride; r.n_in = c->rx_stride;
    r.start = &(c->d_rx_state.p + off)->t0; r.start_stride_words = sizeof(rx_state_t) / 4;
    r.out = c->d_iq_rs.p + off * d.N_RX * c->rs_stride; r.out_stride = c->rs_stride;
    r.n_out = d.n_no_gi_os; r.n_fill = c->rs_stride;
    r.n_ant = d.N_RX;
    DNRP_LAUNCH(c, DNRP_KERNEL_RESAMPLE, launch_resample(r, n, c->stream));
    return DNRP_OK;
}

// device -> host: results of the PCC stage for packets [0, n)
int fetch_pcc_reports(dnrp_ctx* c, uint32_t n, dnrp_pcc_report_t* out) {
    std::vector<uint8_t> ok(2 * n), plcf(32 * n);
    std::vector<uint32_t> mask(2 * n);
    DNRP_CUDA(cudaMemcpyAsync(c->h_rx_state.data(), c->d_rx_state.p, n * sizeof(rx_state_t),
                              cudaMemcpyDeviceToHost, c->stream));
    DNRP_CUDA(cudaMemcpyAsync(ok.data(), c->d_pcc_crc.p, 2 * n, cudaMemcpyDeviceToHost, c->stream));
    DNRP_CUDA(cudaMemcpyAsync(plcf.data(), c->d_plcf_out.p, 32 * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
    DNRP_CUDA(cudaMemcpyAsync(mask.data(), c->d_pcc_mask.p, 8 * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
    DNRP_CUDA(cudaStreamSynchronize(c->stream));
    for (uint32_t i = 0; i < n; ++i) {
        const rx_state_t& s = c->h_rx_state[i];
        dnrp_pcc_report_t& r = out[i];
        std::memset(&r, 0, sizeof(r));
        for (uint32_t t = 0; t < 2; ++t) {
            r.plcf_ok[t] = ok[2 * i + t];
            std::memcpy(r.plcf[t], &plcf[32 * i + 16 * t], 10);
            r.crc_mask[t] = mask[2 * i + t];
        }
        r.snr_dB = s.snr_db_pcc;
        r.cfo_residual_rad = s.cfo_left;
        r.sto_fractional = s.sto_fractional;
        for (uint32_t a = 0; a < 8; ++a) r.rms[a] = s.rms[a];
    }
    return DNRP_OK;
}

// device -> host: transport blocks and reports of the PDC stage for packets [0, n)
int fetch_pdc_reports(dnrp_ctx* c, uint32_t n, uint8_t* tb_out, dnrp_pdc_report_t* out, uint8_t* n_iter,
                      bool tb_already_copied) {
    const fec_plan_t& f = c->fec;
    std::vector<uint8_t> ok(n);
    std::vector<mimo_report_dev_t> mimo(n);
    DNRP_CUDA(cudaMemcpyAsync(mimo.data(), c->d_mimo.p, n * sizeof(mimo_report_dev_t), cudaMemcpyDeviceToHost, c->stream));
    if (!tb_already_copied)
        DNRP_CUDA(cudaMemcpy2DAsync(tb_out, c->ps.N_TB_byte, c->d_tb_out.p, f.tb_stride, c->ps.N_TB_byte, n,
                                    cudaMemcpyDeviceToHost, c->stream));
    DNRP_CUDA(cudaMemcpyAsync(ok.data(), c->d_tb_ok.p, n, cudaMemcpyDeviceToHost, c->stream));
    if (n_iter) DNRP_CUDA(cudaMemcpyAsync(n_iter, c->d_n_iter.p, (size_t)n * c->ps.C, cudaMemcpyDeviceToHost, c->stream));
    DNRP_CUDA(cudaStreamSynchronize(c->stream));
    for (uint32_t i = 0; i < n; ++i) {
        out[i].crc_ok = ok[i];
        out[i].snr_dB = c->h_rx_state[i].snr_db_pdc;  // read back by the PCC stage (whole packet is known)
        out[i].mimo_N_RX = mimo[i].N_RX; out[i].mimo_N_TS_other = mimo[i].N_TS_other;
        out[i].tm_3_7_beamforming_idx = mimo[i].tm_3_7_beamforming_idx;
        out[i].tm_3_7_beamforming_reciprocal_idx = mimo[i].tm_3_7_beamforming_reciprocal_idx;
    }
    return DNRP_OK;
}

}  // namespace dnrp

extern "C" {

// coarse_peak_f_domain_t between coarse and fine peak (k_sync_fdomain): the reference switches both parts at compile time
// (sync_param.hpp:244, 252, both commented out); here per context at run time, off by default
int dnrp_sync_set_fdomain(dnrp_ctx_t* c, float beta_threshold_dB, uint32_t cfo_search_range_subc) {
    DNRP_REQUIRE(c, "null argument");
    DNRP_REQUIRE(cfo_search_range_subc <= 32, "integer CFO search range: at most 32 subcarriers");
    const bool on = beta_threshold_dB < 0.f || cfo_search_range_subc >= 4;
    DNRP_REQUIRE(!on || (c->dims.N_os <= 4096 && !c->resampling),
                 "the frequency-domain stage needs N_b_DFT_os <= 4096 and a context without resampling");
    DNRP_CUDA(cudaSetDevice(c->device));
    if (on && c->d_fd_b.n == 0) {
        std::vector<int8_t> seq(6 * 224, 0);
        const uint32_t tab[6] = {1, 2, 4, 8, 12, 16};
        for (uint32_t bi = 0; bi < 6; ++bi) {
            const std::vector<int8_t> sg = spec::build_stf_sign(tab[bi], 1);   // by subcarrier; cell i of the comb:
            const uint32_t NS = 14 * tab[bi], N_OCC = 56 * tab[bi];
            for (uint32_t i = 0; i < NS; ++i) {
                const uint32_t q = i < NS / 2 ? 4 * i : N_OCC / 2 + 3 + 4 * (i - NS / 2);
                seq[bi * 224 + i] = sg[spec::occ2k(q, N_OCC)];
            }
        }
        DNRP_CUDA(c->d_fd_seq.upload(seq));
        DNRP_CUDA(c->d_fd_b.alloc(c->B));
        DNRP_CUDA(c->d_fd_cfo.alloc(c->B));
    }
    // the reference converts its dB threshold with db2mag (coarse_peak_f_domain.cpp:140-141)
    c->fd_beta_thr = beta_threshold_dB < 0.f ? powf(10.0f, beta_threshold_dB / 20.0f) : 0.f;
    c->fd_cfo_range = cfo_search_range_subc >= 4 ? cfo_search_range_subc : 0;
    return DNRP_OK;
}

// residual CFO / STO tracking on the DRS: RX_SYNCED_PARAM_CFO_RESIDUAL_BASED_ON_DRS / RX_SYNCED_PARAM_STO_RESIDUAL_BASED_ON_DRS
// of the reference (rx_synced_param.hpp:156-181, both commented out, "not functional yet"); here per context at run time
int dnrp_rx_set_tracking(dnrp_ctx_t* c, int residual_cfo, int residual_sto) {
    DNRP_REQUIRE(c, "null argument");
    DNRP_CUDA(cudaSetDevice(c->device));
    if ((residual_cfo || residual_sto) && c->d_track.n == 0) DNRP_CUDA(c->d_track.alloc(c->B));
    c->track_cfo = residual_cfo != 0;
    c->track_sto = residual_sto != 0;
    return DNRP_OK;
}

int dnrp_rx_tracking_read(dnrp_ctx_t* c, uint32_t n, float* out) {
    DNRP_REQUIRE(c && out && n > 0 && n <= c->B, "invalid argument");
    if (c->d_track.n == 0 || !(c->track_cfo || c->track_sto)) {
        set_last_error("residual CFO / STO tracking is off (dnrp_rx_set_tracking)");
        return DNRP_ERR_STATE;
    }
    DNRP_CUDA(cudaSetDevice(c->device));
    DNRP_CUDA(cudaMemcpyAsync(out, c->d_track.p, n * sizeof(float2), cudaMemcpyDeviceToHost, c->stream));
    DNRP_CUDA(cudaStreamSynchronize(c->stream));
    return DNRP_OK;
}

int dnrp_sync_fdomain_read(dnrp_ctx_t* c, uint32_t n, uint32_t* beta_out, int32_t* cfo_subc_out) {
    DNRP_REQUIRE(c && n > 0 && n <= c->B, "invalid argument");
    if (c->d_fd_b.n == 0) {
        set_last_error("the frequency-domain stage was never switched on (dnrp_sync_set_fdomain)");
        return DNRP_ERR_STATE;
    }
    DNRP_CUDA(cudaSetDevice(c->device));
    if (beta_out) DNRP_CUDA(cudaMemcpyAsync(beta_out, c->d_fd_b.p, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
    if (cfo_subc_out) DNRP_CUDA(cudaMemcpyAsync(cfo_subc_out, c->d_fd_cfo.p, n * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
    DNRP_CUDA(cudaStreamSynchronize(c->stream));
    return DNRP_OK;
}

int dnrp_sync_batch(dnrp_ctx_t* c, uint32_t n, const dnrp_cf32_t* iq_in, uint32_t search_length,
                    dnrp_sync_result_t* out) {
    static_assert(sizeof(dnrp_sync_result_t) == sizeof(sync_result_dev_t), "sync result layouts differ");
    DNRP_REQUIRE(c, "null argument");
    DNRP_REQUIRE(n > 0 && n <= c->B, "batch size out of range");
    DNRP_REQUIRE(search_length <= c->sync_len, "search_length exceeds the (resampled) IQ window");
    DNRP_CUDA(cudaSetDevice(c->device));
    if (iq_in)
        DNRP_CUDA(cudaMemcpyAsync(c->d_iq_rx.p, iq_in, (size_t)n * c->dims.N_RX * c->rx_stride * sizeof(float2),
                                  cudaMemcpyHostToDevice, c->stream));
    c->last_search_length = search_length;
    const int rc = run_sync(c, n, search_length, true);
    if (rc) return rc;
    if (out)
        DNRP_CUDA(cudaMemcpyAsync(out, c->d_sync_result.p, n * sizeof(dnrp_sync_result_t), cudaMemcpyDeviceToHost,
                                  c->stream));
    DNRP_CUDA(cudaStreamSynchronize(c->stream));
    return DNRP_OK;
}

uint32_t dnrp_rx_stride(const dnrp_ctx_t* c) { return c ? c->rx_stride : 0; }

// dnrp_sync_batch for windows in the radio's wire format (int16 I/Q): the windows cross the host link at half the size and
// are converted on the device, where dnrp_rx_pcc_batch / dnrp_rx_pdc_batch continue on them (iq_in = NULL)
int dnrp_sync_batch_sc16(dnrp_ctx_t* c, uint32_t n, const int16_t* iq_in, float scale, uint32_t search_length,
                         dnrp_sync_result_t* out) {
    DNRP_REQUIRE(c && iq_in, "null argument");
    DNRP_REQUIRE(n > 0 && n <= c->B, "batch size out of range");
    DNRP_REQUIRE(scale > 0.f, "scale must be > 0");
    DNRP_REQUIRE(search_length <= c->sync_len, "search_length exceeds the (resampled) IQ window");
    DNRP_REQUIRE(c->rx_stride % 4 == 0, "sc16 ingest needs a window stride that is a multiple of 4");
    DNRP_CUDA(cudaSetDevice(c->device));
    const size_t win = (size_t)c->dims.N_RX * c->rx_stride;
    if (!c->d_iq_sc16.p) DNRP_CUDA(c->d_iq_sc16.alloc((size_t)c->B * win * 2));
    DNRP_CUDA(cudaMemcpyAsync(c->d_iq_sc16.p, iq_in, (size_t)n * win * 2 * sizeof(int16_t), cudaMemcpyHostToDevice, c->stream));
    DNRP_LAUNCH(c, -1, launch_sc16_to_cf32(c->d_iq_sc16.p, c->d_iq_rx.p, (size_t)n * win, scale, c->stream));
    c->last_search_length = search_length;
    const int rc = run_sync(c, n, search_length, true);
    if (rc) return rc;
    if (out)
        DNRP_CUDA(cudaMemcpyAsync(out, c->d_sync_result.p, n * sizeof(dnrp_sync_result_t), cudaMemcpyDeviceToHost,
                                  c->stream));
    DNRP_CUDA(cudaStreamSynchronize(c->stream));
    return DNRP_OK;
}

int dnrp_rx_pcc_batch(dnrp_ctx_t* c, uint32_t n, const dnrp_cf32_t* iq_in, const dnrp_sync_report_t* sync,
                      dnrp_pcc_report_t* out) {
    DNRP_REQUIRE(c && out, "null argument");
    DNRP_REQUIRE(n > 0 && n <= c->B, "batch size out of range");
    DNRP_CUDA(cudaSetDevice(c->device));
    const phy_dims_t& d = c->dims;
    if (sync) {
        for (uint32_t i = 0; i < n; ++i) {
            DNRP_REQUIRE(sync[i].fine_peak_time_local + c->n_no_gi_rs <= c->rx_stride, "packet exceeds the IQ window");
            rx_state_t& s = c->h_rx_state[i];
            std::memset(&s, 0, sizeof(s));
            s.t0 = (int32_t)sync[i].fine_peak_time_local;
            s.inc0 = sync[i].cfo_fractional_rad + sync[i].cfo_integer_rad;  // rx_synced.cpp:240
            for (uint32_t a = 0; a < 8; ++a) s.rms[a] = sync[i].rms[a];
        }
        DNRP_CUDA(cudaMemcpyAsync(c->d_rx_state.p, c->h_rx_state.data(), n * sizeof(rx_state_t),
                                  cudaMemcpyHostToDevice, c->stream));
    }  // else: d_rx_state was written by the preceding dnrp_sync_batch (k_sync_final)
    if (iq_in)
        DNRP_CUDA(cudaMemcpyAsync(c->d_iq_rx.p, iq_in, (size_t)n * d.N_RX * c->rx_stride * sizeof(float2),
                                  cudaMemcpyHostToDevice, c->stream));
    int rc = run_rx_front(c, n);
    if (rc) return rc;
    if ((rc = run_rx_pcc(c, n)) != DNRP_OK) return rc;
    return fetch_pcc_reports(c, n, out);
}

int dnrp_rx_pdc_batch(dnrp_ctx_t* c, uint32_t n, const uint32_t* harq_slot, const uint8_t* rv,
                      const uint8_t* harq_reset, uint8_t* tb_out, dnrp_pdc_report_t* out, uint8_t* n_iter) {
    DNRP_REQUIRE(c && tb_out && out, "null argument");
    DNRP_REQUIRE(n > 0 && n <= c->B, "batch size out of range");
    DNRP_CUDA(cudaSetDevice(c->device));
    const fec_plan_t& f = c->fec;
    int rc = validate_harq_slots(c, n, harq_slot);
    if (rc) return rc;
    if (harq_slot) DNRP_CUDA(cudaMemcpyAsync(c->d_slot.p, harq_slot, 4 * n, cudaMemcpyHostToDevice, c->stream));
    if (rv) DNRP_CUDA(cudaMemcpyAsync(c->d_rv.p, rv, n, cudaMemcpyHostToDevice, c->stream));
    if (harq_reset) DNRP_CUDA(cudaMemcpyAsync(c->d_reset.p, harq_reset, n, cudaMemcpyHostToDevice, c->stream));
    rc = run_rx_pdc(c, n, harq_slot != nullptr, rv != nullptr, harq_reset != nullptr);
    if (rc) return rc;
    return fetch_pdc_reports(c, n, tb_out, out, n_iter);
}

// iq_in: float2 windows, or (sc16_scale != 0) int16 (sc_bytes = 2) / int8 (sc_bytes = 1) I/Q pairs that are converted on the
// device chunk by chunk
static int rx_batch_impl(dnrp_ctx_t* c, uint32_t n, const void* iq_in, float sc16_scale, uint32_t sc_bytes, uint32_t search_length,
                         const uint32_t* harq_slot, const uint8_t* rv, const uint8_t* harq_reset,
                         dnrp_sync_result_t* sync_out, dnrp_pcc_report_t* pcc_out, uint8_t* tb_out,
                         dnrp_pdc_report_t* pdc_out, uint8_t* n_iter) {
    DNRP_REQUIRE(c && iq_in && tb_out && pdc_out, "null argument");
    DNRP_REQUIRE(n > 0 && n <= c->B, "batch size out of range");
    DNRP_REQUIRE(search_length <= c->sync_len, "search_length exceeds the (resampled) IQ window");
    c->last_search_length = search_length;
    DNRP_CUDA(cudaSetDevice(c->device));
    const phy_dims_t& d = c->dims;
    const bool sc16 = sc16_scale != 0.f;
    DNRP_REQUIRE(!(sc16 && sc_bytes == 1) || c->rx_stride % 8 == 0, "sc8 ingest needs a window stride that is a multiple of 8");
    if (sc16 && !c->d_iq_sc16.p) {
        DNRP_REQUIRE(c->rx_stride % 4 == 0, "sc16 / sc8 ingest needs a window stride that is a multiple of 4");
        DNRP_CUDA(c->d_iq_sc16.alloc((size_t)c->B * d.N_RX * c->rx_stride * 2));
    }
    {
        const int rcs = validate_harq_slots(c, n, harq_slot);
        if (rcs) return rcs;
    }
    // from here on every early return leaves the context as it found it: window offset back to 0, the three streams
    // drained (a later call must not see half a batch in flight or work on offset buffers)
    struct batch_guard_t {
        dnrp_ctx* c;
        bool done = false;
        ~batch_guard_t() {
            c->off = 0;
            if (done) return;
            cudaStreamSynchronize(c->stream_copy);
            cudaStreamSynchronize(c->stream);
            cudaStreamSynchronize(c->stream_d2h);
        }
    } guard{c};
    if (harq_slot) DNRP_CUDA(cudaMemcpyAsync(c->d_slot.p, harq_slot, 4 * n, cudaMemcpyHostToDevice, c->stream));
    if (rv) DNRP_CUDA(cudaMemcpyAsync(c->d_rv.p, rv, n, cudaMemcpyHostToDevice, c->stream));
    if (harq_reset) DNRP_CUDA(cudaMemcpyAsync(c->d_reset.p, harq_reset, n, cudaMemcpyHostToDevice, c->stream));
    // chunks: the H2D copy of chunk k+1 runs on the copy stream while chunk k is processed on the compute stream and
    // the records of chunk k-1 travel back on the third stream.  Smaller chunks at both ends, larger ones in the middle:
    // the first kernels start after a short copy, the last copy is followed by a short computation (whichever of
    // the two streams is the bottleneck), and the bulk runs in launches large enough to fill the GPU.
    std::vector<uint32_t> bounds{0};
    {
        // Measured on C2 / 10 000 windows (scratch/e2e_weights.py): one chunk costs 0.19 ms + 1.51 us per packet.
        // cf32 windows: the copy (2.17 us per packet over PCIe) is the bottleneck -> many gently tapered chunks, the
        // last copy is followed by a short computation.  sc16 windows: the kernels are the bottleneck -> few chunks
        // growing by <= 1.4x, so that the first kernels start early and no chunk waits for its copy.  (r01k: with the faster
        // kernels copy and compute of the sc16 path are balanced too -- seven chunks tapered at both ends measured best:
        // 15.7 -> 14.6 ms per 10 000 C2 windows.)
        static const uint32_t wc[] = {2, 2, 3, 3, 4, 4, 4, 4, 4, 4, 4, 4, 3, 3, 2, 2}, ws[] = {2, 3, 4, 4, 4, 3, 2},
                              w6[] = {2, 3, 4, 4, 3, 2}, w3[] = {1, 1, 1}, w1[] = {1};
        // (r02w: a call that is bound by its copy -- long windows of several antennas, the north-star workload moves
        // 1.7 GB per 4000 windows -- ends with the computation of its last chunk: ten chunks with a long taper took
        // 32.72 -> 31.86 ms there; with short windows (C2, 0.24 GB) the 0.19 ms per chunk weigh more: 6.28 -> 6.76 ms,
        // so the taper is for calls of 768 MiB and more only.  profiles/tools/e2e_chunks.sh)
        // 8192 windows and more in a packed format: fifteen chunks with the same taper, 72.49 -> 70.63 ms per 9000 north-star
        // windows (profiles/tools/e2e_chunks_large.sh); complex64 windows already go in sixteen chunks
        static const uint32_t w10[] = {1, 2, 3, 4, 4, 4, 3, 2, 1, 1}, w15[] = {1, 2, 3, 4, 4, 4, 4, 4, 4, 4, 4, 3, 2, 1, 1};
        const uint64_t call_bytes = (uint64_t)n * c->dims.N_RX * c->rx_stride * (sc16 ? 2u * sc_bytes : 8u);
        const bool long_copy = n >= 2048 && n < 8192 && call_bytes >= (768ull << 20);
        const bool long_copy15 = n >= 8192 && sc16 && call_bytes >= (768ull << 20);
        const uint32_t* w = long_copy ? w10 : long_copy15 ? w15 : n >= 8192 ? (sc16 ? ws : wc) : (n >= 2048 ? w6 : (n >= 256 ? w3 : w1));
        uint32_t nw = long_copy ? 10 : long_copy15 ? 15 : n >= 8192 ? (sc16 ? 7 : 16) : (n >= 2048 ? 6 : (n >= 256 ? 3 : 1));
        const char* env_chunks = std::getenv("DNRP_RX_CHUNKS");   // tuning knobs: that many equal chunks, or weights
        const char* env_w = std::getenv("DNRP_RX_WEIGHTS");       // "3,4,5,..."
        std::vector<uint32_t> eq;
        if (env_w && *env_w) {
            for (const char* q = env_w; *q;) {
                eq.push_back((uint32_t)std::max(1L, std::strtol(q, const_cast<char**>(&q), 10)));
                while (*q == ',' || *q == ' ') ++q;
            }
            w = eq.data(); nw = (uint32_t)eq.size();
        } else if (env_chunks && std::atoi(env_chunks) > 0) {
            eq.assign(std::min<uint32_t>((uint32_t)std::atoi(env_chunks), n), 1);
            w = eq.data(); nw = (uint32_t)eq.size();
        }
        uint64_t tot = 0, acc = 0;
        for (uint32_t k = 0; k < nw; ++k) tot += w[k];
        for (uint32_t k = 0; k < nw; ++k) {
            acc += w[k];
            const uint32_t e = (uint32_t)((uint64_t)n * acc / tot);
            if (e > bounds.back()) bounds.push_back(e);
        }
    }
    const uint32_t n_chunks = (uint32_t)bounds.size() - 1;
    while (c->chunk_events.size() < n_chunks) {
        cudaEvent_t e, g;
        DNRP_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        DNRP_CUDA(cudaEventCreateWithFlags(&g, cudaEventDisableTiming));
        c->chunk_events.push_back(e);
        c->chunk_done.push_back(g);
    }
    // pinned staging of the per-packet records (pageable destinations would serialise the copies)
    const size_t B = c->B, C_cb = c->ps.C;
    size_t so = 0;
    auto carve = [&](size_t bytes) { const size_t o = so; so += (bytes + 63) & ~(size_t)63; return o; };
    const size_t o_state = carve(B * sizeof(rx_state_t)), o_ok = carve(2 * B), o_plcf = carve(32 * B), o_mask = carve(8 * B),
                 o_tbok = carve(B), o_mimo = carve(B * sizeof(mimo_report_dev_t)), o_nit = carve(B * C_cb),
                 o_sync = carve(B * sizeof(dnrp_sync_result_t));
    if (c->h_stage_bytes < so) {
        if (c->h_stage) cudaFreeHost(c->h_stage);
        c->h_stage = nullptr; c->h_stage_bytes = 0;
        DNRP_CUDA(cudaMallocHost(&c->h_stage, so));
        c->h_stage_bytes = so;
    }
    uint8_t* hs = static_cast<uint8_t*>(c->h_stage);
    const size_t win = (size_t)d.N_RX * c->rx_stride;
    for (uint32_t k = 0; k < n_chunks; ++k) {
        const size_t o = bounds[k], m = bounds[k + 1] - bounds[k];
        if (sc16)   // staged at byte offset o * win * 2 * sc_bytes of the int16 staging buffer
            DNRP_CUDA(cudaMemcpyAsync(reinterpret_cast<uint8_t*>(c->d_iq_sc16.p) + o * win * 2 * sc_bytes,
                                      reinterpret_cast<const uint8_t*>(iq_in) + o * win * 2 * sc_bytes,
                                      m * win * 2 * sc_bytes, cudaMemcpyHostToDevice, c->stream_copy));
        else
            DNRP_CUDA(cudaMemcpyAsync(c->d_iq_rx.p + o * win, reinterpret_cast<const float2*>(iq_in) + o * win,
                                      m * win * sizeof(float2), cudaMemcpyHostToDevice, c->stream_copy));
        DNRP_CUDA(cudaEventRecord(c->chunk_events[k], c->stream_copy));
    }
    int rc = DNRP_OK;
    for (uint32_t k = 0; k < n_chunks && rc == DNRP_OK; ++k) {
        const size_t o = bounds[k];
        const uint32_t m = bounds[k + 1] - bounds[k];
        DNRP_CUDA(cudaStreamWaitEvent(c->stream, c->chunk_events[k], 0));
        c->off = (uint32_t)o;
        if (sc16 && sc_bytes == 2)
            DNRP_LAUNCH(c, -1, launch_sc16_to_cf32(c->d_iq_sc16.p + o * win * 2, c->d_iq_rx.p + o * win, m * win, sc16_scale,
                                                   c->stream));
        if (sc16 && sc_bytes == 1)
            DNRP_LAUNCH(c, -1, launch_sc8_to_cf32(reinterpret_cast<const int8_t*>(c->d_iq_sc16.p) + o * win * 2,
                                                  c->d_iq_rx.p + o * win, m * win, sc16_scale, c->stream));
        rc = run_sync(c, m, search_length, true);
        if (rc == DNRP_OK) rc = run_rx_front(c, m);
        if (rc == DNRP_OK) rc = run_rx_pcc(c, m);
        if (rc == DNRP_OK) rc = run_rx_pdc(c, m, harq_slot != nullptr, rv != nullptr, harq_reset != nullptr);
        if (rc == DNRP_OK) {
            // this chunk's transport blocks and records travel back while the next chunk is processed
            DNRP_CUDA(cudaEventRecord(c->chunk_done[k], c->stream));
            cudaStream_t sd = c->stream_d2h;
            DNRP_CUDA(cudaStreamWaitEvent(sd, c->chunk_done[k], 0));
            DNRP_CUDA(cudaMemcpy2DAsync(tb_out + o * c->ps.N_TB_byte, c->ps.N_TB_byte,
                                        c->d_tb_out.p + o * c->fec.tb_stride, c->fec.tb_stride, c->ps.N_TB_byte, m,
                                        cudaMemcpyDeviceToHost, sd));
            auto back = [&](size_t off, const void* dev, size_t unit) {
                return cudaMemcpyAsync(hs + off + o * unit, static_cast<const uint8_t*>(dev) + o * unit, m * unit,
                                       cudaMemcpyDeviceToHost, sd);
            };
            DNRP_CUDA(back(o_state, c->d_rx_state.p, sizeof(rx_state_t)));
            DNRP_CUDA(back(o_ok, c->d_pcc_crc.p, 2));
            DNRP_CUDA(back(o_plcf, c->d_plcf_out.p, 32));
            DNRP_CUDA(back(o_mask, c->d_pcc_mask.p, 8));
            DNRP_CUDA(back(o_tbok, c->d_tb_ok.p, 1));
            DNRP_CUDA(back(o_mimo, c->d_mimo.p, sizeof(mimo_report_dev_t)));
            if (n_iter) DNRP_CUDA(back(o_nit, c->d_n_iter.p, C_cb));
            if (sync_out) DNRP_CUDA(back(o_sync, c->d_sync_result.p, sizeof(dnrp_sync_result_t)));
        }
    }
    c->off = 0;
    if (rc) return rc;
    DNRP_CUDA(cudaStreamSynchronize(c->stream_d2h));
    DNRP_CUDA(cudaStreamSynchronize(c->stream));
    guard.done = true;
    // records -> ABI structs (same fields as fetch_pcc_reports / fetch_pdc_reports)
    const rx_state_t* st = reinterpret_cast<const rx_state_t*>(hs + o_state);
    const mimo_report_dev_t* mimo = reinterpret_cast<const mimo_report_dev_t*>(hs + o_mimo);
    const uint32_t* mask = reinterpret_cast<const uint32_t*>(hs + o_mask);
    std::memcpy(c->h_rx_state.data(), st, n * sizeof(rx_state_t));
    if (sync_out) std::memcpy(sync_out, hs + o_sync, n * sizeof(dnrp_sync_result_t));
    if (n_iter) std::memcpy(n_iter, hs + o_nit, n * C_cb);
    for (uint32_t i = 0; i < n; ++i) {
        if (pcc_out) {
            dnrp_pcc_report_t& r = pcc_out[i];
            std::memset(&r, 0, sizeof(r));
            for (uint32_t t = 0; t < 2; ++t) {
                r.plcf_ok[t] = hs[o_ok + 2 * i + t];
                std::memcpy(r.plcf[t], hs + o_plcf + 32 * (size_t)i + 16 * t, 10);
                r.crc_mask[t] = mask[2 * i + t];
            }
            r.snr_dB = st[i].snr_db_pcc;
            r.cfo_residual_rad = st[i].cfo_left;
            r.sto_fractional = st[i].sto_fractional;
            for (uint32_t a = 0; a < 8; ++a) r.rms[a] = st[i].rms[a];
        }
        dnrp_pdc_report_t& q = pdc_out[i];
        q.crc_ok = hs[o_tbok + i];
        q.snr_dB = st[i].snr_db_pdc;
        q.mimo_N_RX = mimo[i].N_RX; q.mimo_N_TS_other = mimo[i].N_TS_other;
        q.tm_3_7_beamforming_idx = mimo[i].tm_3_7_beamforming_idx;
        q.tm_3_7_beamforming_reciprocal_idx = mimo[i].tm_3_7_beamforming_reciprocal_idx;
    }
    return DNRP_OK;
}

int dnrp_rx_batch(dnrp_ctx_t* c, uint32_t n, const dnrp_cf32_t* iq_in, uint32_t search_length, const uint32_t* harq_slot,
                  const uint8_t* rv, const uint8_t* harq_reset, dnrp_sync_result_t* sync_out, dnrp_pcc_report_t* pcc_out,
                  uint8_t* tb_out, dnrp_pdc_report_t* pdc_out, uint8_t* n_iter) {
    return rx_batch_impl(c, n, iq_in, 0.f, 0, search_length, harq_slot, rv, harq_reset, sync_out, pcc_out, tb_out, pdc_out,
                         n_iter);
}

int dnrp_rx_batch_sc16(dnrp_ctx_t* c, uint32_t n, const int16_t* iq_in, float scale, uint32_t search_length,
                       const uint32_t* harq_slot, const uint8_t* rv, const uint8_t* harq_reset,
                       dnrp_sync_result_t* sync_out, dnrp_pcc_report_t* pcc_out, uint8_t* tb_out,
                       dnrp_pdc_report_t* pdc_out, uint8_t* n_iter) {
    DNRP_REQUIRE(scale > 0.f, "scale must be > 0");
    return rx_batch_impl(c, n, iq_in, scale, 2, search_length, harq_slot, rv, harq_reset, sync_out, pcc_out, tb_out,
                         pdc_out, n_iter);
}

int dnrp_rx_batch_sc8(dnrp_ctx_t* c, uint32_t n, const int8_t* iq_in, float scale, uint32_t search_length,
                      const uint32_t* harq_slot, const uint8_t* rv, const uint8_t* harq_reset,
                      dnrp_sync_result_t* sync_out, dnrp_pcc_report_t* pcc_out, uint8_t* tb_out,
                      dnrp_pdc_report_t* pdc_out, uint8_t* n_iter) {
    DNRP_REQUIRE(scale > 0.f, "scale must be > 0");
    return rx_batch_impl(c, n, iq_in, scale, 1, search_length, harq_slot, rv, harq_reset, sync_out, pcc_out, tb_out,
                         pdc_out, n_iter);
}

static bool tap_desc(const dnrp_ctx_t* c, int tap, uint32_t n, const void*& src, size_t& bytes) {
    const phy_dims_t& d = c->dims;
    switch (tap) {
        case DNRP_TAP_Y: src = c->d_Y.p; bytes = (size_t)n * d.N_RX * d.N_DF * d.Nk * sizeof(float2); return true;
        case DNRP_TAP_DRS_ZF:
            src = c->d_hzf.p; bytes = (size_t)n * d.N_RX * d.N_eff_TX * d.n_drs * d.NP * sizeof(float2); return true;
        case DNRP_TAP_STF_ZF: src = c->d_hstf.p; bytes = (size_t)n * d.N_RX * d.NP * sizeof(float2); return true;
        case DNRP_TAP_LLR_PCC: src = c->d_llr_pcc.p; bytes = (size_t)n * 196 * 2; return true;
        case DNRP_TAP_LLR_PDC: src = c->d_llr.p; bytes = (size_t)n * c->ps.G * 2; return true;
        case DNRP_TAP_WIENER_CLASS: src = c->d_cls.p; bytes = (size_t)n * c->cls_stride; return true;
        case DNRP_TAP_TB_OK: src = c->d_tb_ok.p; bytes = n; return true;
        default: return false;
    }
}

size_t dnrp_tap_bytes(const dnrp_ctx_t* c, int tap, uint32_t n) {
    const void* src;
    size_t bytes = 0;
    if (c && tap == DNRP_TAP_CHESTIM) return (size_t)n * c->dims.N_RX * c->dims.N_eff_TX * c->dims.Nk * sizeof(float2);
    if (!c || !tap_desc(c, tap, n, src, bytes)) return 0;
    return bytes;
}

int dnrp_tap_read(dnrp_ctx_t* c, int tap, uint32_t n, void* dst, size_t dst_bytes) {
    DNRP_REQUIRE(c && dst, "null argument");
    DNRP_REQUIRE(n > 0 && n <= c->B, "batch size out of range");
    const void* src;
    size_t bytes = 0;
    DNRP_CUDA(cudaSetDevice(c->device));
    if (tap == DNRP_TAP_CHESTIM) {
        // computed on demand from the zero-forced pilots and Wiener classes the last receiver call left on the device
        bytes = dnrp_tap_bytes(c, tap, n);
        DNRP_REQUIRE(dst_bytes >= bytes, "destination too small");
        const size_t need = (size_t)c->B * c->dims.N_RX * c->dims.N_eff_TX * c->dims.Nk;
        if (c->d_chestim_tap.n < need) DNRP_CUDA(c->d_chestim_tap.alloc(need));
        const rx_params_t p = make_rx_params(c);
        DNRP_LAUNCH(c, -1, launch_rx_chestim_tap(p, c->d_chestim_tap.p, n, c->stream));
        DNRP_CUDA(cudaMemcpyAsync(dst, c->d_chestim_tap.p, bytes, cudaMemcpyDeviceToHost, c->stream));
        DNRP_CUDA(cudaStreamSynchronize(c->stream));
        return DNRP_OK;
    }
    DNRP_REQUIRE(tap_desc(c, tap, n, src, bytes), "unknown tap");
    DNRP_REQUIRE(dst_bytes >= bytes, "destination too small");
    DNRP_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, c->stream));
    DNRP_CUDA(cudaStreamSynchronize(c->stream));
    return DNRP_OK;
}

float dnrp_large_scale_via_pathloss(float tx_power_ant_0dBFS, float rx_power_ant_0dBFS, float distance_m, float freq_Hz,
                                    int same_node, float tx_into_rx_leakage_dB) {
    float fading_dB = tx_into_rx_leakage_dB;   // the node's own transmitter leaking into its receiver
    if (!same_node) {
        // pathloss::fspl (pathloss.cpp:28-38): free space, never below 0 dB, 0 for a null request
        fading_dB = 0.0f;
        if (distance_m > 0.0f && freq_Hz > 0.0f) {
            const double v = 20.0 * std::log10((double)distance_m) + 20.0f * std::log10((double)freq_Hz) - 147.55;
            fading_dB = v < 0.0 ? 0.0f : (float)v;
        }
    }
    return powf(10.0f, (tx_power_ant_0dBFS - fading_dB - rx_power_ant_0dBFS) / 20.0f);   // adt::db2mag
}

// Knot distance of k_channel_tdl's interpolated tap coefficients.  A coefficient is a sum of 40 unit phasors whose phase
// advances by at most w = 2 pi fD / fs per sample; linear interpolation between knots h samples apart is off by at most
// (h w)^2 / 8 per phasor, i.e. relative to the coefficient's rms (sqrt(40)) by sqrt(40) (h w)^2 / 8.  h = the largest power
// of two <= 256 that keeps this below 2e-6 (two orders of magnitude inside the 2e-4 the IQ is compared at); below 8 the
// kernel evaluates every sample exactly.  DNRP_TDL_EXACT=1 forces that path (A/B runs, tests).
static uint32_t tdl_interp_h(float fD_Hz, double samp_rate) {
    static const bool exact = [] { const char* e = getenv("DNRP_TDL_EXACT"); return e && e[0] == '1'; }();
    if (exact) return 0;
    const double w = 2.0 * 3.14159265358979323846 * (double)fD_Hz / samp_rate;
    const double hmax = w > 0.0 ? std::sqrt(8.0 * 2e-6 / std::sqrt(40.0)) / w : 256.0;
    uint32_t h = 256;
    while (h > 1 && (double)h > hmax) h >>= 1;
    return h >= 8 ? h : 0;
}

int dnrp_channel_batch(dnrp_ctx_t* c, uint32_t n, const dnrp_channel_desc_t* descs, dnrp_cf32_t* iq_rx_out) {
    DNRP_REQUIRE(c && descs, "null argument");
    DNRP_REQUIRE(n > 0 && n <= c->B, "batch size out of range");
    DNRP_CUDA(cudaSetDevice(c->device));
    const phy_dims_t& d = c->dims;
    std::vector<chan_meta_dev_t> meta(n);
    bool any_tdl = false;
    uint32_t min_h = 0;
    const double samp_rate_hw = (double)d.u * d.b * c->cfg.oversampling * 1.728e6 * c->cfg.resample_L / c->cfg.resample_M;
    for (uint32_t i = 0; i < n; ++i) {
        DNRP_REQUIRE(descs[i].t0 + c->n_packet_rs <= c->rx_stride, "offset exceeds the IQ window");
        DNRP_REQUIRE(descs[i].mode <= 2, "unknown channel mode");
        DNRP_REQUIRE(descs[i].net_bandwidth_norm > 0.f, "net_bandwidth_norm must be > 0");
        const float n0_dBFS = -10.0f * log10f(descs[i].net_bandwidth_norm) - descs[i].snr_dB;  // noise.cpp:37-38
        chan_meta_dev_t& m = meta[i];
        std::memset(&m, 0, sizeof(m));
        m.t0 = descs[i].t0; m.mode = descs[i].mode; m.sigma = powf(10.0f, n0_dBFS / 20.0f);
        m.seed_lo = (uint32_t)descs[i].seed; m.seed_hi = (uint32_t)(descs[i].seed >> 32);
        DNRP_REQUIRE(descs[i].large_scale >= 0.f && descs[i].clip_limit >= 0.f && descs[i].dac_bits <= 24 && descs[i].adc_bits <= 24,
                     "large_scale / clip_limit must not be negative, at most 24 converter bits");
        m.large_scale = descs[i].large_scale > 0.f ? descs[i].large_scale : 1.0f;
        m.clip = descs[i].clip_limit;
        // hw_simulator.cpp:720-725: step 2 clip / 2^bits, only together with clipping
        m.dac_lsb = m.clip > 0.f && descs[i].dac_bits ? 2.0f * m.clip / powf(2.0f, (float)descs[i].dac_bits) : 0.f;
        m.adc_lsb = m.clip > 0.f && descs[i].adc_bits ? 2.0f * m.clip / powf(2.0f, (float)descs[i].adc_bits) : 0.f;
        if (descs[i].mode == 2) {
            // link.cpp:44-46 asserts these bounds
            DNRP_REQUIRE(descs[i].pdp_idx < spec::TDL_PROFILES, "pdp_idx out of bound");
            DNRP_REQUIRE(descs[i].tau_rms_ns >= 0.f && descs[i].tau_rms_ns <= 2000.f, "tau_rms_ns out of bound");
            DNRP_REQUIRE(descs[i].fD_Hz >= 0.f && descs[i].fD_Hz <= 2000.f, "fD_Hz out of bound");
            DNRP_REQUIRE(descs[i].time_64 >= 0, "time_64 must not be negative");
            double pw[spec::TDL_MAX_TAPS];
            m.samp_rate = samp_rate_hw;
            m.n_taps = spec::tdl_taps(samp_rate_hw, descs[i].pdp_idx, descs[i].tau_rms_ns, m.tap_delay, pw);
            for (uint32_t k = 0; k < m.n_taps; ++k) {
                float sc = 1.0f / sqrtf(40.0f);   // link.cpp:268-274
                sc *= sqrt(pw[k]);
                m.tap_scale[k] = sc;
            }
            m.fD_Hz = descs[i].fD_Hz; m.time_64 = descs[i].time_64;
            m.interp_h = tdl_interp_h(m.fD_Hz, samp_rate_hw);
            for (uint32_t k = 0; k < m.n_taps; ++k)
                if (m.tap_delay[k] > TDL_HALO) m.interp_h = 0;   // the staged tile has room for delays up to TDL_HALO samples
            if (m.interp_h && (min_h == 0 || m.interp_h < min_h)) min_h = m.interp_h;
            any_tdl = true;
        }
    }
    c->chan_any_tdl = any_tdl;
    c->chan_min_h = min_h;
    if (any_tdl && c->d_chan_sin.n == 0)
        DNRP_CUDA(c->d_chan_sin.alloc((size_t)c->B * d.N_RX * d.N_TX * 9 * 40));
    DNRP_CUDA(cudaMemcpyAsync(c->d_chan_meta.p, meta.data(), n * sizeof(chan_meta_dev_t), cudaMemcpyHostToDevice,
                              c->stream));
    DNRP_CUDA(cudaStreamSynchronize(c->stream));  // meta is a local pageable vector
    int rc = run_channel(c, n);
    if (rc) return rc;
    if (iq_rx_out)
        DNRP_CUDA(cudaMemcpyAsync(iq_rx_out, c->d_iq_rx.p, (size_t)n * d.N_RX * c->rx_stride * sizeof(float2),
                                  cudaMemcpyDeviceToHost, c->stream));
    DNRP_CUDA(cudaStreamSynchronize(c->stream));
    return DNRP_OK;
}

int dnrp_kernel_time_enable(dnrp_ctx_t* c, int enable) {
    DNRP_REQUIRE(c, "null argument");
    DNRP_CUDA(cudaSetDevice(c->device));
    DNRP_CUDA(cudaStreamSynchronize(c->stream));
    c->ktime_on = enable != 0;
    for (auto& t : c->ktimer) { t.used = 0; t.acc_ms = 0.f; t.count = 0; }
    return DNRP_OK;
}

int dnrp_kernel_time_read(dnrp_ctx_t* c, int kernel, float* ms_total, uint32_t* launches) {
    DNRP_REQUIRE(c && ms_total && launches, "null argument");
    DNRP_REQUIRE(kernel >= 0 && kernel < DNRP_KERNEL_COUNT, "unknown kernel");
    DNRP_CUDA(cudaSetDevice(c->device));
    DNRP_CUDA(cudaStreamSynchronize(c->stream));
    auto& t = c->ktimer[kernel];
    for (size_t i = 0; i + 1 < t.used; i += 2) {
        float ms = 0.f;
        DNRP_CUDA(cudaEventElapsedTime(&ms, t.ev[i], t.ev[i + 1]));
        t.acc_ms += ms;
        ++t.count;
    }
    t.used = 0;
    *ms_total = t.acc_ms;
    *launches = t.count;
    return DNRP_OK;
}

uint64_t dnrp_launch_count(const dnrp_ctx_t* c) { return c ? c->launches : 0; }

int dnrp_loopback_tx_to_rx(dnrp_ctx_t* c, uint32_t n, uint32_t t0) {
    DNRP_REQUIRE(c, "null argument");
    DNRP_REQUIRE(n > 0 && n <= c->B, "batch size out of range");
    const phy_dims_t& d = c->dims;
    DNRP_REQUIRE(t0 + c->n_packet_rs <= c->rx_stride, "offset exceeds the IQ window");
    DNRP_CUDA(cudaSetDevice(c->device));
    DNRP_CUDA(cudaMemsetAsync(c->d_iq_rx.p, 0, (size_t)n * d.N_RX * c->rx_stride * sizeof(float2), c->stream));
    for (uint32_t a = 0; a < d.N_RX; ++a) {
        // one strided copy per RX antenna: rows = packets
        const float2* src = c->d_iq_tx.p + (size_t)(a % d.N_TX) * c->tx_stride;
        float2* dst = c->d_iq_rx.p + (size_t)a * c->rx_stride + t0;
        DNRP_CUDA(cudaMemcpy2DAsync(dst, (size_t)d.N_RX * c->rx_stride * sizeof(float2), src,
                                    (size_t)d.N_TX * c->tx_stride * sizeof(float2),
                                    (size_t)c->n_packet_rs * sizeof(float2), n, cudaMemcpyDeviceToDevice, c->stream));
    }
    return DNRP_OK;
}

}  // extern "C"

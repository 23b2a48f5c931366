// ctx.hpp -- the per-device, per-geometry context behind the C ABI: host-side plan (tables derived
// from spec*.hpp) + device workspaces.  Host C++ only; kernels are reached through the launch_*
// functions of the .cuh headers.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <string>
#include <vector>
#include "../../include/dnrp_c_api.h"
#include "fec_rx.cuh"
#include "fec_tx.cuh"
#include "phy.cuh"
#include "spec.hpp"
#include "spec_fec.hpp"
#include "spec_re.hpp"
#include "spec_resample.hpp"

namespace dnrp {

void set_last_error(const std::string& s);

#define DNRP_CUDA(expr)                                                                       \
    do {                                                                                      \
        cudaError_t e__ = (expr);                                                             \
        if (e__ != cudaSuccess) {                                                             \
            ::dnrp::set_last_error(std::string(#expr) + ": " + cudaGetErrorString(e__));      \
            return DNRP_ERR_CUDA;                                                             \
        }                                                                                     \
    } while (0)

// RAII device allocation
template <typename T>
struct dbuf {
    T* p = nullptr;
    size_t n = 0;
    dbuf() = default;
    dbuf(const dbuf&) = delete;
    dbuf& operator=(const dbuf&) = delete;
    ~dbuf() { if (p) cudaFree(p); }
    cudaError_t alloc(size_t count) {
        if (p) cudaFree(p);
        p = nullptr;
        n = count;
        if (count == 0) return cudaSuccess;
        cudaError_t e = cudaMalloc((void**)&p, count * sizeof(T));
        if (e == cudaSuccess) e = cudaMemset(p, 0, count * sizeof(T));
        // the zero fill runs on the default stream, the contexts' streams are non-blocking: without this barrier a
        // lazily allocated buffer could be zeroed AFTER the first asynchronous copy / kernel wrote to it
        if (e == cudaSuccess) e = cudaDeviceSynchronize();
        return e;
    }
    cudaError_t upload(const std::vector<T>& v) {
        cudaError_t e = alloc(v.size());
        if (e != cudaSuccess || v.empty()) return e;
        return cudaMemcpy(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice);
    }
};

uint32_t tdec_window(uint32_t K);
uint32_t gf2_xpow_host(uint32_t n, uint32_t poly, uint32_t order);

struct fec_plan_t {
    // rate matching / CRC-combination tables shared by all code block sizes of the plan
    std::vector<uint16_t> rm_arena;   // per K: src[N], then pos[N] in the decoder soft-buffer order (plan_fec.cu)
    std::vector<uint32_t> xp_arena;
    std::vector<cb_desc_t> rx_cbs;    // PDC decode descriptors [C]
    std::vector<txcb_desc_t> tx_cbs;  // PDC encode descriptors [C]
    cb_desc_t rx_pcc[2];              // PLCF type 1 / type 2
    txcb_desc_t tx_pcc[2];
    std::vector<uint32_t> xp_tb;      // TB CRC chunk multipliers
    uint32_t tb_chunk_bytes = 64;
    uint32_t soft_stride = 0, saved_stride = 0, tb_stride = 0, dbits_stride = 0;
    uint32_t Kmax = 0, tdec_threads = 0;
    size_t tdec_smem = 0;
};

// N_IR: soft bits one transport block may use (N_soft / M_DL_HARQ, TS 36.212 5.1.4.1.2); 0 = no limitation
int build_fec_plan(const spec::packet_sizes_t& ps, fec_plan_t& f, uint32_t N_IR = 0);

}  // namespace dnrp

// one kernel launch: counted, and timed with events when dnrp_kernel_time_enable is on (k < 0: untimed)
#define DNRP_LAUNCH(c, k, expr)          \
    do {                                 \
        ::dnrp::kscope ks__((c), (k));   \
        DNRP_CUDA(expr);                 \
    } while (0)

struct dnrp_ctx {
    dnrp_config_t cfg;
    dnrp::spec::packet_sizes_t ps;
    int device = 0;
    cudaStream_t stream = nullptr;
    uint32_t B = 0;  // max batch
    uint32_t off = 0;  // first packet the run_* stage functions work on (pipelined host calls process chunk by chunk)
    cudaStream_t stream_copy = nullptr;      // H2D copies of the pipelined receive call
    cudaStream_t stream_d2h = nullptr;       // transport blocks of finished chunks back to the host
    std::vector<cudaEvent_t> chunk_events, chunk_done;
    void* h_stage = nullptr;                 // pinned staging of the per-packet records of dnrp_rx_batch (first use)
    size_t h_stage_bytes = 0;

    dnrp::fec_plan_t fec;
    // device tables
    dnrp::dbuf<uint16_t> d_rm_arena;
    dnrp::dbuf<uint32_t> d_xp_arena, d_xp_tb;
    dnrp::dbuf<dnrp::cb_desc_t> d_rx_cbs, d_rx_pcc;
    dnrp::dbuf<dnrp::txcb_desc_t> d_tx_cbs, d_tx_pcc;
    dnrp::dbuf<uint8_t> d_gold_pdc, d_gold_pcc;
    // PDC scrambling sequences: entry 0 = the context's default (cfg.network_id, cfg.plcf_type), further entries added by
    // dnrp_add_network_id (both PLCF types of the id: c_init = id & 0xff and id >> 8, scrambling_pdc.cpp:43-56)
    std::vector<uint32_t> gold_cinit;
    uint32_t gold_stride = 0;                      // bytes per sequence in d_gold_pdc
    dnrp::dbuf<uint32_t> d_gold_sel;               // [B] per-packet sequence index of the *_ex calls
    dnrp::dbuf<uint8_t> d_pcc_sel;                 // [B] per-packet PLCF type - 1 of dnrp_tx_batch_ex
    bool use_gold_sel = false, use_pcc_sel = false;   // set for the duration of a *_ex call
    // device workspaces, [B] leading dimension
    dnrp::dbuf<uint8_t> d_tb_in, d_plcf_in, d_rv, d_reset, d_pdc_bits, d_pcc_bits;
    dnrp::dbuf<uint32_t> d_plcf_mask, d_slot;
    dnrp::dbuf<int16_t> d_llr, d_llr_pcc, d_soft, d_soft_pcc;
    dnrp::dbuf<uint8_t> d_cb_crc, d_saved, d_tb_out, d_n_iter, d_tb_ok;
    dnrp::dbuf<uint8_t> d_pcc_crc, d_plcf_out, d_pcc_n_iter;
    dnrp::dbuf<uint32_t> d_pcc_mask;

    // OFDM / receiver plan
    dnrp::phy_dims_t dims;
    dnrp::spec::re_map_t re_map;
    dnrp::dbuf<uint32_t> d_cell, d_pcc_lk, d_pdc_lk, d_pcc_plan, d_pdc_plan;
    dnrp::dbuf<uint8_t> d_cell_ab;
    dnrp::dbuf<float2> d_qam_lut;
    dnrp::dbuf<int8_t> d_stf_sign, d_W;
    dnrp::dbuf<float> d_W_scale, d_W_scale_opt;
    dnrp::dbuf<float2> d_twiddle, d_twiddle_p4, d_twiddle_p16;
    uint32_t tx_stride = 0;  // samples per antenna in d_iq_tx
    dnrp::dbuf<float2> d_iq_tx;
    dnrp::dbuf<dnrp::tx_meta_dev_t> d_tx_meta;

    // receiver: Wiener LUTs (mode l / lr, three SNR classes) and workspaces
    struct wiener_bufs_t {
        dnrp::dbuf<uint16_t> idx_pilot, idx_weight;
        dnrp::dbuf<float> weights;
    } d_wiener_l[3], d_wiener_lr[3];
    struct pair_bufs_t {
        dnrp::dbuf<uint16_t> info;
        dnrp::dbuf<float> w;
    } d_pcc_pair[3], d_pdc_pair[3];
    uint32_t rx_stride = 0;  // samples per antenna in d_iq_rx
    uint32_t cls_stride = 0;
    dnrp::dbuf<float2> d_iq_rx, d_Y, d_hzf, d_hstf, d_snr_part;
    dnrp::dbuf<float2> d_track;            // residual CFO / STO estimates (dnrp_rx_set_tracking), allocated on first use
    bool track_cfo{false}, track_sto{false};
    dnrp::dbuf<dnrp::rx_state_t> d_rx_state;
    dnrp::dbuf<uint8_t> d_cls;
    dnrp::dbuf<float2> d_chestim_tap;           // DNRP_TAP_CHESTIM, allocated by the first read
    std::vector<dnrp::rx_state_t> h_rx_state;
    dnrp::dbuf<dnrp::chan_meta_dev_t> d_chan_meta;
    // MIMO report: single-stream codebooks W(1, N_eff_TX) and W(1, N_RX)
    dnrp::dbuf<int8_t> d_mimo_W[2];
    dnrp::dbuf<float> d_mimo_scale[2];
    uint32_t mimo_n[2] = {0, 0}, mimo_first[2] = {0, 0};
    dnrp::dbuf<dnrp::mimo_report_dev_t> d_mimo;
    dnrp::dbuf<dnrp::tdl_sin_t> d_chan_sin;   // allocated by the first doubly selective channel call
    bool chan_any_tdl = false;
    uint32_t chan_min_h = 0;                  // smallest non-zero interp_h of the resident channel descriptors

    // resampling between the DECT rate (x oversampling) and the hardware rate: hw = dect * L / M (TX direction);
    // the receive direction uses M / L.  All IQ crossing the C ABI is at the hardware rate.
    bool resampling = false;
    dnrp::spec::resampler_design_t rs_tx, rs_rx;
    dnrp::dbuf<float> d_rs_tx_taps, d_rs_rx_taps;
    uint32_t n_no_gi_rs = 0, n_packet_rs = 0;   // packet without / with full guard interval at the hardware rate
    uint32_t stf_len_rs = 0;                    // STF template length at the hardware rate
    uint32_t tx_os_stride = 0, sync_stride = 0, sync_len = 0, rs_stride = 0;
    dnrp::dbuf<float2> d_iq_tx_os;              // [B][N_TX][tx_os_stride] TX IQ before resampling
    dnrp::dbuf<float2> d_iq_sync;               // [B][N_RX][sync_stride] window resampled for detection / coarse peak
    dnrp::dbuf<int16_t> d_iq_tx_sc16;           // [B][N_TX][tx_stride][2] staging of dnrp_tx_batch_sc16 (allocated on first use)
    dnrp::dbuf<int16_t> d_stream_sc16;          // [N_RX][max_push][2] staging of dnrp_stream_push_sc16 (allocated on first use)
    dnrp::dbuf<int16_t> d_iq_sc16;              // [B][N_RX][rx_stride][2] staging of dnrp_rx_batch_sc16 (allocated on first use)
    dnrp::dbuf<float2> d_iq_rs;                 // [B][N_RX][rs_stride] packet resampled from the fine peak

    // packet synchronisation: STF templates for N_eff_TX = 1, 2, 4, 8 (up to the context's), step sums, results
    uint32_t sync_n_tpl = 0, sync_step = 0, sync_n_steps = 0;
    float fd_beta_thr = 0.f;                       // dnrp_sync_set_fdomain: linear threshold of the beta search, 0 = off
    uint32_t fd_cfo_range = 0;                     // integer CFO search range in subcarriers, 0 = off
    dnrp::dbuf<int8_t> d_fd_seq;                   // [6][224] STF base sequences
    dnrp::dbuf<uint32_t> d_fd_b;                   // [B]
    dnrp::dbuf<int32_t> d_fd_cfo;                  // [B]
    uint32_t last_search_length = 0;               // of the last dnrp_sync_batch / dnrp_rx_batch call (dnrp_time_stage re-runs with it)
    dnrp::dbuf<float2> d_sync_tpl, d_sync_c, d_sync_xc;
    uint32_t sync_fft_n = 0;                       // FFT size of the fast cross-correlation (0: direct)
    dnrp::dbuf<float2> d_sync_tw_p16;
    dnrp::dbuf<float2> d_sync_tpl_spec, d_sync_tw; // conj(FFT(template)) / N, twiddles of that size
    dnrp::dbuf<float> d_sync_p;
    dnrp::dbuf<dnrp::sync_result_dev_t> d_sync_result;

    // streaming front door (api_stream.cu): the stream on the device, per-chunk work buffers (grown on demand)
    dnrp::dbuf<float2> d_stream;                // [N_RX][stream_alloc] zero padded behind the last sample
    size_t stream_alloc = 0, stream_len = 0;    // samples per antenna allocated / valid
    uint32_t stream_chunks_alloc = 0, stream_steps_alloc = 0;
    dnrp::dbuf<float2> d_stream_c, d_stream_xc;
    dnrp::dbuf<float> d_stream_p;
    dnrp::dbuf<dnrp::sync_result_dev_t> d_stream_result;
    dnrp::dbuf<uint32_t> d_stream_resume, d_stream_limit, d_stream_map;
    // what dnrp_stream_extract reads: the one-shot stream above or the ring of the stateful front door below
    struct stream_view_t {
        const float2* p = nullptr;   // [N_RX][stride] hardware-rate samples
        size_t stride = 0, len = 0;  // antenna stride, valid samples per antenna
        int64_t base = 0;            // absolute index of element 0
    } stream_view;
    // stateful streaming front door (dnrp_stream_open / _push): hardware-rate ring, the same stream at the DECT rate
    // (x oversampling) when L != M, search cursor, packets found but not yet complete
    struct stream_state_t {
        bool open = false;
        uint32_t hop = 0, lead = 0, tail = 0, Q = 0, len = 0;   // DECT-rate samples; Q = grid of chunk starts
        uint32_t max_push = 0, max_chunks = 0;
        size_t cap_hw = 0, cap_rs = 0;
        int cur = 0;
        dnrp::dbuf<float2> hw[2], rs[2];
        int64_t hw_base = 0, rs_base = 0;    // absolute index of element 0 of the current buffers
        size_t hw_len = 0, rs_len = 0;       // samples held per antenna
        int64_t search_from = 0;             // absolute DECT-rate index of the next detection position
        int64_t carry_resume = 0;            // skip_after_peak of the last packet, when it reaches behind search_from
        int64_t last_fine = INT64_MIN / 2;   // baton_t::is_sync_time_unique across pushes
        bool first = true;
        std::vector<dnrp_stream_packet_t> pending;
    } ss;

    // measurement: launch counter and optional per-kernel event timers (dnrp_kernel_time_*)
    uint64_t launches = 0;
    bool ktime_on = false;
    struct ktimer_t {
        std::vector<cudaEvent_t> ev;  // start/stop pairs
        size_t used = 0;
        float acc_ms = 0.f;
        uint32_t count = 0;
    } ktimer[DNRP_KERNEL_COUNT];
};

namespace dnrp {
// brackets one kernel launch with events when timing is enabled; always counts the launch
struct kscope {
    dnrp_ctx* c;
    int k;
    cudaEvent_t stop = nullptr;
    kscope(dnrp_ctx* c_, int k_) : c(c_), k(k_) {
        ++c->launches;
        if (!c->ktime_on || k < 0) return;
        auto& t = c->ktimer[k];
        if (t.used + 2 > t.ev.size()) {
            cudaEvent_t a, b;
            if (cudaEventCreate(&a) != cudaSuccess || cudaEventCreate(&b) != cudaSuccess) return;
            t.ev.push_back(a);
            t.ev.push_back(b);
        }
        cudaEventRecord(t.ev[t.used], c->stream);
        stop = t.ev[t.used + 1];
        t.used += 2;
    }
    ~kscope() {
        if (stop) cudaEventRecord(stop, c->stream);
    }
};
}  // namespace dnrp

namespace dnrp {
int build_phy_plan(dnrp_ctx* c);
int upload_gold_table(dnrp_ctx* c);
int gold_index(dnrp_ctx* c, uint32_t network_id, uint32_t plcf_type);   // -1: not registered
// every slot < max_batch and no slot twice in one batch (two packets on one slot would race on its soft buffer)
int validate_harq_slots(dnrp_ctx* c, uint32_t n, const uint32_t* harq_slot);
int run_fec_encode(dnrp_ctx* c, uint32_t n, bool have_mask, bool have_rv);
int run_fec_decode_tb(dnrp_ctx* c, uint32_t n, bool have_slot, bool have_rv, bool have_reset);
int run_fec_decode_plcf(dnrp_ctx* c, uint32_t n);
int run_tx_ofdm(dnrp_ctx* c, uint32_t n);
dnrp::rx_params_t make_rx_params(dnrp_ctx* c);
int run_rx_front(dnrp_ctx* c, uint32_t n);             // STF stage, FFTs, DRS zero forcing, SNR classes
int run_rx_pcc(dnrp_ctx* c, uint32_t n);               // PCC demap + blind PLCF decode
int run_rx_pdc(dnrp_ctx* c, uint32_t n, bool have_slot, bool have_rv, bool have_reset);
int run_channel(dnrp_ctx* c, uint32_t n);
int run_sync(dnrp_ctx* c, uint32_t n, uint32_t search_length, bool hand_over);
int fetch_pcc_reports(dnrp_ctx* c, uint32_t n, dnrp_pcc_report_t* out);
int fetch_pdc_reports(dnrp_ctx* c, uint32_t n, uint8_t* tb_out, dnrp_pdc_report_t* out, uint8_t* n_iter,
                      bool tb_already_copied = false);
int run_rx_resample(dnrp_ctx* c, uint32_t n);          // hardware rate -> DECT rate from each packet's fine peak
}  // namespace dnrp

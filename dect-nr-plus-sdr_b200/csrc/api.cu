// api.cu -- the extern "C" boundary of libdnrp_cuda.so (declared in include/dnrp_c_api.h).
#include <cstring>
#include <mutex>
#include "ctx.hpp"

namespace dnrp {
static thread_local std::string g_last_error;
void set_last_error(const std::string& s) { g_last_error = s; }
}  // namespace dnrp

using namespace dnrp;

#define DNRP_REQUIRE(cond, msg)             \
    do {                                    \
        if (!(cond)) {                      \
            dnrp::set_last_error(msg);      \
            return DNRP_ERR_INVALID;        \
        }                                   \
    } while (0)

static void fill_packet_sizes(const spec::packet_sizes_t& q, dnrp_packet_sizes_t* o) {
    std::memset(o, 0, sizeof(*o));
    std::memcpy(&o->psdef, &q.psdef, sizeof(o->psdef));
    o->N_eff_TX = q.tm.N_eff_TX; o->N_SS = q.tm.N_SS; o->N_TS = q.tm.N_TS; o->N_TX = q.tm.N_TX;
    o->closed_loop = q.tm.cl; o->N_bps = q.mcs.N_bps;
    o->N_b_DFT = q.N_b_DFT; o->N_b_CP = q.N_b_CP; o->N_b_OCC = q.N_b_OCC;
    o->N_PACKET_symb = q.N_PACKET_symb; o->N_DF_symb = q.N_DF_symb;
    o->N_PDC_subc = q.N_PDC_subc; o->N_DRS_subc = q.N_DRS_subc;
    o->G = q.G; o->N_PDC_bits = q.N_PDC_bits; o->N_TB_bits = q.N_TB_bits; o->N_TB_byte = q.N_TB_byte;
    o->C = q.C; o->K_plus = q.seg.K1; o->K_minus = q.seg.K2; o->C_plus = q.seg.C1; o->C_minus = q.seg.C2;
    o->N_samples_STF = q.N_samples_STF; o->N_samples_STF_CP_only = q.N_samples_STF_CP_only;
    o->N_samples_DF = q.N_samples_DF; o->N_samples_GI = q.N_samples_GI;
    o->N_samples_packet_no_GI = q.N_samples_packet_no_GI; o->N_samples_packet = q.N_samples_packet;
}

extern "C" {

const char* dnrp_last_error_string(void) { return g_last_error.c_str(); }

int dnrp_get_packet_sizes(const dnrp_psdef_t* psdef, dnrp_packet_sizes_t* out) {
    DNRP_REQUIRE(psdef && out, "null argument");
    spec::psdef_t d;
    std::memcpy(&d, psdef, sizeof(d));
    spec::packet_sizes_t q;
    if (!spec::get_packet_sizes(d, q)) {
        set_last_error("packet configuration not possible");
        return DNRP_ERR_INVALID;
    }
    fill_packet_sizes(q, out);
    return DNRP_OK;
}

int dnrp_resampler_design(uint32_t L, uint32_t M, uint32_t oversampling, float* taps_out, uint32_t taps_cap,
                          uint32_t* info4) {
    DNRP_REQUIRE(info4 && L > 0 && M > 0, "invalid argument");
    spec::resampler_design_t r;
    DNRP_REQUIRE(spec::design_resampler(L, M, oversampling, r), "no filter definition for this oversampling");
    info4[0] = r.filter_length; info4[1] = r.subfilter_length; info4[2] = r.delay; info4[3] = (uint32_t)r.taps.size();
    if (taps_out) {
        DNRP_REQUIRE(taps_cap >= r.taps.size(), "taps_out too small");
        std::memcpy(taps_out, r.taps.data(), r.taps.size() * sizeof(float));
    }
    return DNRP_OK;
}

uint32_t dnrp_resampler_n_after(uint32_t L, uint32_t M, uint32_t oversampling, uint32_t n_in) {
    spec::resampler_design_t r;
    if (!spec::design_resampler(L, M, oversampling, r)) return 0;
    return r.n_after(n_in);
}

int dnrp_ctx_create(const dnrp_config_t* cfg, int device, dnrp_ctx_t** out) {
    DNRP_REQUIRE(cfg && out, "null argument");
    DNRP_REQUIRE(cfg->max_batch > 0, "max_batch must be > 0");
    DNRP_REQUIRE(cfg->max_batch <= 65535, "max_batch must be <= 65535 (packets ride in a CUDA grid dimension); process larger "
                                          "workloads as successive batches");
    DNRP_REQUIRE(cfg->plcf_type == 1 || cfg->plcf_type == 2, "plcf_type must be 1 or 2");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
        set_last_error("no CUDA device (this library has no CPU fallback)");
        return DNRP_ERR_NO_DEVICE;
    }
    DNRP_CUDA(cudaSetDevice(device));
    dnrp_ctx* c = new dnrp_ctx();
    c->cfg = *cfg;
    if (c->cfg.tdec_min_iter == 0) c->cfg.tdec_min_iter = 2;
    if (c->cfg.tdec_max_iter == 0) c->cfg.tdec_max_iter = 10;
    if (c->cfg.oversampling == 0) c->cfg.oversampling = 1;
    if (c->cfg.resample_L == 0 || c->cfg.resample_M == 0) c->cfg.resample_L = c->cfg.resample_M = 1;
    if (c->cfg.N_RX == 0) c->cfg.N_RX = 1;
    if (c->cfg.chestim_mode_lr_t_stride == 0) c->cfg.chestim_mode_lr_t_stride = 2;
    c->device = device;
    c->B = cfg->max_batch;
    spec::psdef_t d;
    std::memcpy(&d, &cfg->psdef, sizeof(d));
    int rc = DNRP_OK;
    auto fail = [&](int code, const char* msg) {
        if (msg) set_last_error(msg);
        delete c;
        return code;
    };
    if (!spec::get_packet_sizes(d, c->ps)) return fail(DNRP_ERR_INVALID, "packet configuration not possible");
    uint32_t N_IR = 0;
    if (c->cfg.N_soft) {
        // TS 36.212 5.1.4.1.2 with K_C = K_MIMO = 1; every code block must keep at least its systematic part
        N_IR = c->cfg.N_soft / (c->cfg.M_DL_HARQ ? c->cfg.M_DL_HARQ : 1u);
        const uint32_t Kp = 32 * ((c->ps.seg.K1 + 4 + 31) / 32);
        if (N_IR / c->ps.seg.C < Kp)
            return fail(DNRP_ERR_INVALID, "N_soft / M_DL_HARQ / C is smaller than the systematic part of a code block");
    }
    if ((rc = build_fec_plan(c->ps, c->fec, N_IR)) != DNRP_OK) return fail(rc, "FEC plan failed");
    if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&c->stream_copy, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&c->stream_d2h, cudaStreamNonBlocking) != cudaSuccess)
        return fail(DNRP_ERR_CUDA, "cudaStreamCreate failed");

    const fec_plan_t& f = c->fec;
    const size_t B = c->B;
    const uint32_t c_init = cfg->plcf_type == 1 ? (cfg->network_id & 0xff) : (cfg->network_id >> 8);
    bool ok = true;
#define UP(x) ok = ok && ((x) == cudaSuccess)
    UP(c->d_rm_arena.upload(f.rm_arena));
    UP(c->d_xp_arena.upload(f.xp_arena));
    UP(c->d_xp_tb.upload(f.xp_tb));
    UP(c->d_rx_cbs.upload(f.rx_cbs));
    UP(c->d_tx_cbs.upload(f.tx_cbs));
    UP(c->d_rx_pcc.upload(std::vector<cb_desc_t>(f.rx_pcc, f.rx_pcc + 2)));
    UP(c->d_tx_pcc.upload(std::vector<txcb_desc_t>(f.tx_pcc, f.tx_pcc + 2)));
    c->gold_cinit.assign(1, c_init);   // entry 0: the context's default; entry 1: the other PLCF type of the same id
    {
        const uint32_t other = cfg->plcf_type == 1 ? (cfg->network_id >> 8) : (cfg->network_id & 0xff);
        if (other != c_init) c->gold_cinit.push_back(other);
    }
    ok = ok && upload_gold_table(c) == DNRP_OK;
    UP(c->d_gold_sel.alloc(B));
    UP(c->d_pcc_sel.alloc(B));
    UP(c->d_gold_pcc.upload(spec::gold_sequence_packed(spec::PCC_SCRAMBLING_INIT, 256)));
    UP(c->d_tb_in.alloc(B * f.tb_stride));
    UP(c->d_plcf_in.alloc(B * 16));
    UP(c->d_rv.alloc(B));
    UP(c->d_reset.alloc(B));
    UP(c->d_plcf_mask.alloc(B));
    UP(c->d_slot.alloc(B));
    UP(c->d_pdc_bits.alloc(B * f.dbits_stride));
    UP(c->d_pcc_bits.alloc(B * 28));
    UP(c->d_llr.alloc(B * (size_t)c->ps.G));
    UP(c->d_llr_pcc.alloc(B * 196));
    UP(c->d_soft.alloc(B * (size_t)f.soft_stride));
    UP(c->d_soft_pcc.alloc(B * 640));
    UP(c->d_cb_crc.alloc(B * c->ps.C));
    UP(c->d_saved.alloc(B * (size_t)f.saved_stride));
    UP(c->d_tb_out.alloc(B * (size_t)f.tb_stride));
    UP(c->d_n_iter.alloc(B * c->ps.C));
    UP(c->d_tb_ok.alloc(B));
    UP(c->d_pcc_crc.alloc(B * 2));
    UP(c->d_plcf_out.alloc(B * 32));
    UP(c->d_pcc_n_iter.alloc(B * 2));
    UP(c->d_pcc_mask.alloc(B * 2));
#undef UP
    if (!ok) return fail(DNRP_ERR_CUDA, "device allocation failed");
    if ((rc = build_phy_plan(c)) != DNRP_OK) return fail(rc, nullptr);
    *out = c;
    return DNRP_OK;
}

void dnrp_ctx_destroy(dnrp_ctx_t* ctx) {
    if (!ctx) return;
    dnrp_comm_destroy(ctx);   // communicator and staging buffer of the counter reduction, if any (api_comm.cu)
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    if (ctx->stream_copy) cudaStreamDestroy(ctx->stream_copy);
    if (ctx->stream_d2h) cudaStreamDestroy(ctx->stream_d2h);
    if (ctx->h_stage) cudaFreeHost(ctx->h_stage);
    for (cudaEvent_t e : ctx->chunk_events) cudaEventDestroy(e);
    for (cudaEvent_t e : ctx->chunk_done) cudaEventDestroy(e);
    for (auto& t : ctx->ktimer)
        for (cudaEvent_t e : t.ev) cudaEventDestroy(e);
    delete ctx;
}

int dnrp_ctx_get_packet_sizes(const dnrp_ctx_t* ctx, dnrp_packet_sizes_t* out) {
    DNRP_REQUIRE(ctx && out, "null argument");
    fill_packet_sizes(ctx->ps, out);
    return DNRP_OK;
}

int dnrp_ctx_synchronize(dnrp_ctx_t* ctx) {
    DNRP_REQUIRE(ctx, "null argument");
    DNRP_CUDA(cudaStreamSynchronize(ctx->stream));
    return DNRP_OK;
}

}  // extern "C"

// ---- device-side FEC stages on the context's own buffers (used by the full chain as well)
namespace dnrp {

int upload_gold_table(dnrp_ctx* c) {
    const uint32_t bits = c->ps.G + 64;
    c->gold_stride = ((bits + 7) / 8 + 15u) & ~15u;
    std::vector<uint8_t> t((size_t)c->gold_stride * c->gold_cinit.size(), 0);
    for (size_t i = 0; i < c->gold_cinit.size(); ++i) {
        const std::vector<uint8_t> g = spec::gold_sequence_packed(c->gold_cinit[i], bits);
        std::memcpy(t.data() + i * c->gold_stride, g.data(), std::min<size_t>(g.size(), c->gold_stride));
    }
    DNRP_CUDA(cudaStreamSynchronize(c->stream));   // nothing in flight may still read the old table
    DNRP_CUDA(c->d_gold_pdc.upload(t));
    return DNRP_OK;
}

int gold_index(dnrp_ctx* c, uint32_t network_id, uint32_t plcf_type) {
    const uint32_t c_init = plcf_type == 1 ? (network_id & 0xff) : (network_id >> 8);
    for (size_t i = 0; i < c->gold_cinit.size(); ++i)
        if (c->gold_cinit[i] == c_init) return (int)i;
    return -1;
}

int run_fec_encode(dnrp_ctx* c, uint32_t n, bool have_mask, bool have_rv) {
    const fec_plan_t& f = c->fec;
    DNRP_CUDA(cudaMemsetAsync(c->d_pdc_bits.p, 0, (size_t)n * f.dbits_stride, c->stream));
    DNRP_CUDA(cudaMemsetAsync(c->d_pcc_bits.p, 0, (size_t)n * 28, c->stream));
    txfec_params_t p{};
    p.cbs = c->d_tx_cbs.p; p.C = c->ps.C;
    p.in = c->d_tb_in.p; p.in_stride = f.tb_stride;
    p.rv = have_rv ? c->d_rv.p : nullptr;
    p.plcf_mask = nullptr;
    p.gold = c->d_gold_pdc.p; p.gold_bit_off = 0;
    p.gold_sel = c->use_gold_sel ? c->d_gold_sel.p : nullptr; p.gold_stride = c->gold_stride;
    p.out = c->d_pdc_bits.p; p.out_stride = f.dbits_stride;
    p.rm_arena = c->d_rm_arena.p; p.xp_arena = c->d_xp_arena.p;
    DNRP_LAUNCH(c, -1, launch_tx_fec(p, n, f.Kmax, c->stream));
    // PLCF: one launch, the descriptor (type 1 or 2) is the plan's plcf_type or, in dnrp_tx_batch_ex, the packet's
    txfec_params_t q{};
    q.cbs = c->d_tx_pcc.p + (c->use_pcc_sel ? 0 : (c->cfg.plcf_type - 1)); q.C = 1;
    q.cb_sel = c->use_pcc_sel ? c->d_pcc_sel.p : nullptr;
    q.in = c->d_plcf_in.p; q.in_stride = 16;
    q.rv = nullptr;
    q.plcf_mask = have_mask ? c->d_plcf_mask.p : nullptr;
    q.gold = c->d_gold_pcc.p; q.gold_bit_off = 0;
    q.out = c->d_pcc_bits.p; q.out_stride = 28;
    q.rm_arena = c->d_rm_arena.p; q.xp_arena = c->d_xp_arena.p;
    DNRP_LAUNCH(c, -1, launch_tx_fec(q, n, 96, c->stream));
    return DNRP_OK;
}

int run_fec_decode_tb(dnrp_ctx* c, uint32_t n, bool have_slot, bool have_rv, bool have_reset) {
    const fec_plan_t& f = c->fec;
    tdec_params_t p{};
    p.cbs = c->d_rx_cbs.p; p.C = c->ps.C;
    // packets [off, off + n): per-packet buffers are offset; HARQ state is addressed by slot (= packet without a slot
    // array, then offset alike)
    const size_t off = c->off, soff = have_slot ? 0 : off;
    p.llr = c->d_llr.p + off * c->ps.G; p.llr_stride = c->ps.G;
    p.gold = c->d_gold_pdc.p; p.gold_bit_off = 0;
    p.gold_sel = c->use_gold_sel ? c->d_gold_sel.p + off : nullptr; p.gold_stride = c->gold_stride;
    p.soft = c->d_soft.p + soff * f.soft_stride; p.soft_stride = f.soft_stride;
    p.cb_crc = c->d_cb_crc.p + soff * c->ps.C;
    p.saved = c->d_saved.p + soff * f.saved_stride; p.saved_stride = f.saved_stride;
    p.slot = have_slot ? c->d_slot.p + off : nullptr;
    p.rv = have_rv ? c->d_rv.p + off : nullptr;
    p.reset = have_reset ? c->d_reset.p + off : nullptr;
    p.tb_out = c->d_tb_out.p + off * f.tb_stride; p.tb_stride = f.tb_stride;
    p.n_iter = c->d_n_iter.p + off * c->ps.C;
    p.rm_arena = c->d_rm_arena.p; p.xp_arena = c->d_xp_arena.p;
    p.min_it = c->cfg.tdec_min_iter; p.max_it = c->cfg.tdec_max_iter;
    p.crc_mode = 0; p.plcf_mask = nullptr;
    uint32_t max_n_e = 0;
    for (const cb_desc_t& d : f.rx_cbs) max_n_e = std::max(max_n_e, d.n_e);
    {   // room for the whole circular buffer (zero padded staging) if every block's fits into 48 KB
        uint32_t max_N = 0;
        for (const cb_desc_t& d : f.rx_cbs) max_N = std::max(max_N, 3 * (d.K + 4));
        if (max_N <= 24576u) max_n_e = std::max(max_n_e, max_N);
    }
    DNRP_LAUNCH(c, DNRP_KERNEL_DERM_PDC, launch_derm(p, n, max_n_e, c->stream));
    const tdec_geom_t geom = tdec_geometry(f.rx_cbs.data(), (uint32_t)f.rx_cbs.size());
    DNRP_LAUNCH(c, DNRP_KERNEL_TURBO_DECODE_PDC, launch_turbo_decode(p, n, geom, c->stream));
    tbfin_params_t t{};
    t.C = c->ps.C; t.tbs_bits = c->ps.N_TB_bits; t.chunk_bytes = f.tb_chunk_bytes;
    t.slot = p.slot; t.cb_crc = p.cb_crc;
    t.tb_out = p.tb_out; t.tb_stride = f.tb_stride;
    t.xp_tb = c->d_xp_tb.p; t.tb_ok = c->d_tb_ok.p + off;
    DNRP_LAUNCH(c, -1, launch_tb_finalize(t, n, c->stream));
    return DNRP_OK;
}

int run_fec_decode_plcf(dnrp_ctx* c, uint32_t n) {
    const fec_plan_t& f = c->fec;
    tdec_params_t p{};
    p.cbs = c->d_rx_pcc.p; p.C = 2;
    const size_t off = c->off;
    p.llr = c->d_llr_pcc.p + off * 196; p.llr_stride = 196;
    p.gold = c->d_gold_pcc.p; p.gold_bit_off = 0;
    p.soft = c->d_soft_pcc.p + off * 640; p.soft_stride = 640;
    p.cb_crc = c->d_pcc_crc.p + off * 2;
    p.saved = nullptr; p.saved_stride = 0;
    p.slot = nullptr; p.rv = nullptr; p.reset = nullptr;
    p.tb_out = c->d_plcf_out.p + off * 32; p.tb_stride = 32;
    p.n_iter = c->d_pcc_n_iter.p + off * 2;
    p.rm_arena = c->d_rm_arena.p; p.xp_arena = c->d_xp_arena.p;
    p.min_it = 1; p.max_it = 5;  // pcc_enc.cpp:39
    p.crc_mode = 1; p.plcf_mask = c->d_pcc_mask.p + off * 2;
    DNRP_LAUNCH(c, -1, launch_derm(p, n, 304, c->stream));  // >= 196 LLRs and >= 3 (96 + 4) buffer positions
    DNRP_LAUNCH(c, -1, launch_turbo_decode(p, n, tdec_geometry(f.rx_pcc, 2), c->stream));
    return DNRP_OK;
}

}  // namespace dnrp

extern "C" {

int dnrp_fec_encode(dnrp_ctx_t* c, uint32_t n, const uint8_t* plcf, const uint32_t* cl_bf_mask,
                    const uint8_t* tb, const uint8_t* rv, uint8_t* pcc_bits, uint8_t* pdc_bits) {
    DNRP_REQUIRE(c && plcf && tb && pcc_bits && pdc_bits, "null argument");
    DNRP_REQUIRE(n > 0 && n <= c->B, "batch size out of range");
    DNRP_CUDA(cudaSetDevice(c->device));
    const fec_plan_t& f = c->fec;
    DNRP_CUDA(cudaMemcpy2DAsync(c->d_tb_in.p, f.tb_stride, tb, c->ps.N_TB_byte, c->ps.N_TB_byte, n,
                                cudaMemcpyHostToDevice, c->stream));
    DNRP_CUDA(cudaMemcpy2DAsync(c->d_plcf_in.p, 16, plcf, 10, 10, n, cudaMemcpyHostToDevice, c->stream));
    if (cl_bf_mask) DNRP_CUDA(cudaMemcpyAsync(c->d_plcf_mask.p, cl_bf_mask, 4 * n, cudaMemcpyHostToDevice, c->stream));
    if (rv) DNRP_CUDA(cudaMemcpyAsync(c->d_rv.p, rv, n, cudaMemcpyHostToDevice, c->stream));
    int rc = run_fec_encode(c, n, cl_bf_mask != nullptr, rv != nullptr);
    if (rc) return rc;
    DNRP_CUDA(cudaMemcpyAsync(pcc_bits, c->d_pcc_bits.p, (size_t)n * 28, cudaMemcpyDeviceToHost, c->stream));
    DNRP_CUDA(cudaMemcpyAsync(pdc_bits, c->d_pdc_bits.p, (size_t)n * f.dbits_stride, cudaMemcpyDeviceToHost, c->stream));
    DNRP_CUDA(cudaStreamSynchronize(c->stream));
    return DNRP_OK;
}

}  // extern "C"

int dnrp::validate_harq_slots(dnrp_ctx* c, uint32_t n, const uint32_t* harq_slot) {
    if (!harq_slot) return DNRP_OK;   // packet i uses slot i
    std::vector<uint8_t> seen(c->B, 0);
    for (uint32_t i = 0; i < n; ++i) {
        if (harq_slot[i] >= c->B) {
            set_last_error("harq slot out of range");
            return DNRP_ERR_INVALID;
        }
        if (seen[harq_slot[i]]) {
            set_last_error("harq slot " + std::to_string(harq_slot[i]) + " appears twice in one batch (its soft buffer would be "
                           "combined by two packets at once)");
            return DNRP_ERR_INVALID;
        }
        seen[harq_slot[i]] = 1;
    }
    return DNRP_OK;
}

extern "C" {

int dnrp_fec_decode_tb(dnrp_ctx_t* c, uint32_t n, const int16_t* llr, const uint32_t* harq_slot,
                       const uint8_t* rv, const uint8_t* harq_reset, uint8_t* tb_out, uint8_t* crc_ok,
                       uint8_t* n_iter) {
    DNRP_REQUIRE(c && llr && tb_out && crc_ok, "null argument");
    DNRP_REQUIRE(n > 0 && n <= c->B, "batch size out of range");
    DNRP_CUDA(cudaSetDevice(c->device));
    const fec_plan_t& f = c->fec;
    {
        const int rcs = validate_harq_slots(c, n, harq_slot);
        if (rcs) return rcs;
    }
    DNRP_CUDA(cudaMemcpyAsync(c->d_llr.p, llr, (size_t)n * c->ps.G * 2, cudaMemcpyHostToDevice, c->stream));
    if (harq_slot) DNRP_CUDA(cudaMemcpyAsync(c->d_slot.p, harq_slot, 4 * n, cudaMemcpyHostToDevice, c->stream));
    if (rv) DNRP_CUDA(cudaMemcpyAsync(c->d_rv.p, rv, n, cudaMemcpyHostToDevice, c->stream));
    if (harq_reset) DNRP_CUDA(cudaMemcpyAsync(c->d_reset.p, harq_reset, n, cudaMemcpyHostToDevice, c->stream));
    int rc = run_fec_decode_tb(c, n, harq_slot != nullptr, rv != nullptr, harq_reset != nullptr);
    if (rc) return rc;
    DNRP_CUDA(cudaMemcpy2DAsync(tb_out, c->ps.N_TB_byte, c->d_tb_out.p, f.tb_stride, c->ps.N_TB_byte, n,
                                cudaMemcpyDeviceToHost, c->stream));
    DNRP_CUDA(cudaMemcpyAsync(crc_ok, c->d_tb_ok.p, n, cudaMemcpyDeviceToHost, c->stream));
    if (n_iter) DNRP_CUDA(cudaMemcpyAsync(n_iter, c->d_n_iter.p, (size_t)n * c->ps.C, cudaMemcpyDeviceToHost, c->stream));
    DNRP_CUDA(cudaStreamSynchronize(c->stream));
    return DNRP_OK;
}

int dnrp_fec_decode_plcf(dnrp_ctx_t* c, uint32_t n, const int16_t* llr196, uint8_t* ok, uint8_t* plcf,
                         uint32_t* mask, uint8_t* n_iter) {
    DNRP_REQUIRE(c && llr196 && ok && plcf && mask, "null argument");
    DNRP_REQUIRE(n > 0 && n <= c->B, "batch size out of range");
    DNRP_CUDA(cudaSetDevice(c->device));
    DNRP_CUDA(cudaMemcpyAsync(c->d_llr_pcc.p, llr196, (size_t)n * 196 * 2, cudaMemcpyHostToDevice, c->stream));
    int rc = run_fec_decode_plcf(c, n);
    if (rc) return rc;
    DNRP_CUDA(cudaMemcpyAsync(ok, c->d_pcc_crc.p, (size_t)n * 2, cudaMemcpyDeviceToHost, c->stream));
    DNRP_CUDA(cudaMemcpy2DAsync(plcf, 10, c->d_plcf_out.p, 16, 10, (size_t)n * 2, cudaMemcpyDeviceToHost, c->stream));
    DNRP_CUDA(cudaMemcpyAsync(mask, c->d_pcc_mask.p, (size_t)n * 8, cudaMemcpyDeviceToHost, c->stream));
    if (n_iter) DNRP_CUDA(cudaMemcpyAsync(n_iter, c->d_pcc_n_iter.p, (size_t)n * 2, cudaMemcpyDeviceToHost, c->stream));
    DNRP_CUDA(cudaStreamSynchronize(c->stream));
    return DNRP_OK;
}

}  // extern "C"

extern "C" int dnrp_time_stage(dnrp_ctx_t* c, int stage, uint32_t n, uint32_t reps, float* ms_per_rep) {
    DNRP_REQUIRE(c && ms_per_rep && reps > 0, "invalid argument");
    DNRP_REQUIRE(n > 0 && n <= c->B, "batch size out of range");
    DNRP_CUDA(cudaSetDevice(c->device));
    cudaEvent_t e0, e1;
    DNRP_CUDA(cudaEventCreate(&e0));
    DNRP_CUDA(cudaEventCreate(&e1));
    int rc = DNRP_OK;
    DNRP_CUDA(cudaEventRecord(e0, c->stream));
    for (uint32_t r = 0; r < reps && rc == DNRP_OK; ++r) {
        switch (stage) {
            case DNRP_STAGE_FEC_DECODE_TB: rc = run_fec_decode_tb(c, n, false, false, false); break;
            case DNRP_STAGE_FEC_ENCODE: rc = run_fec_encode(c, n, false, false); break;
            case DNRP_STAGE_FEC_DECODE_PLCF: rc = run_fec_decode_plcf(c, n); break;
            case DNRP_STAGE_RX_CHAIN:
                rc = run_rx_front(c, n);
                if (rc == DNRP_OK) rc = run_rx_pcc(c, n);
                if (rc == DNRP_OK) rc = run_rx_pdc(c, n, false, false, false);
                break;
            case DNRP_STAGE_TX_CHAIN:
                rc = run_fec_encode(c, n, true, true);
                if (rc == DNRP_OK) rc = run_tx_ofdm(c, n);
                break;
            case DNRP_STAGE_CHANNEL: rc = run_channel(c, n); break;
            case DNRP_STAGE_SYNC: rc = run_sync(c, n, c->last_search_length, true); break;
            case DNRP_STAGE_SYNC_RX_CHAIN:
                rc = run_sync(c, n, c->last_search_length, true);
                if (rc == DNRP_OK) rc = run_rx_front(c, n);
                if (rc == DNRP_OK) rc = run_rx_pcc(c, n);
                if (rc == DNRP_OK) rc = run_rx_pdc(c, n, false, false, false);
                break;
            default: set_last_error("unknown stage"); rc = DNRP_ERR_INVALID;
        }
    }
    if (rc != DNRP_OK) return rc;
    DNRP_CUDA(cudaEventRecord(e1, c->stream));
    DNRP_CUDA(cudaEventSynchronize(e1));
    float ms = 0;
    DNRP_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *ms_per_rep = ms / reps;
    return DNRP_OK;
}

// ---- transmitter ---------------------------------------------------------------------------------
namespace dnrp {
int run_tx_ofdm(dnrp_ctx* c, uint32_t n) {
    const phy_dims_t& d = c->dims;
    DNRP_CUDA(cudaMemsetAsync(c->d_iq_tx.p, 0, (size_t)n * d.N_TX * c->tx_stride * sizeof(float2), c->stream));
    txofdm_params_t p{};
    p.d = d;
    p.pcc_bits = c->d_pcc_bits.p; p.pcc_stride = 28;
    p.pdc_bits = c->d_pdc_bits.p; p.pdc_stride = c->fec.dbits_stride;
    p.meta = c->d_tx_meta.p;
    if (!c->resampling) {
        p.iq = c->d_iq_tx.p; p.iq_stride = c->tx_stride;
        DNRP_LAUNCH(c, -1, launch_tx_ofdm(p, n, c->stream));
        return DNRP_OK;
    }
    // tx.cpp:913-934 + 679-704: IFFT output -> polyphase resampler -> mixer at the hardware rate; the guard interval
    // stays zero (memset above)
    p.iq = c->d_iq_tx_os.p; p.iq_stride = c->tx_os_stride; p.no_mix = 1;
    DNRP_LAUNCH(c, -1, launch_tx_ofdm(p, n, c->stream));
    resample_params_t r{};
    r.L = c->rs_tx.L; r.M = c->rs_tx.M; r.SL = c->rs_tx.subfilter_length; r.delay = c->rs_tx.delay;
    r.taps = c->d_rs_tx_taps.p;
    r.in = c->d_iq_tx_os.p; r.in_stride = c->tx_os_stride; r.n_in = d.n_no_gi_os;
    r.out = c->d_iq_tx.p; r.out_stride = c->tx_stride; r.n_out = c->n_no_gi_rs; r.n_fill = 0;
    r.n_ant = d.N_TX;
    r.meta = c->d_tx_meta.p;
    DNRP_LAUNCH(c, DNRP_KERNEL_RESAMPLE, launch_resample(r, n, c->stream));
    return DNRP_OK;
}
}  // namespace dnrp

extern "C" {

uint32_t dnrp_tx_stride(const dnrp_ctx_t* c) { return c ? c->tx_stride : 0; }

static int tx_batch_impl(dnrp_ctx_t* c, uint32_t n, const dnrp_tx_desc_t* descs, const uint8_t* plcf, const uint8_t* tb,
                         void* iq_out, float sc16_scale, uint32_t* n_samples_out);

int dnrp_tx_batch(dnrp_ctx_t* c, uint32_t n, const dnrp_tx_desc_t* descs, const uint8_t* plcf, const uint8_t* tb,
                  dnrp_cf32_t* iq_out, uint32_t* n_samples_out) {
    return tx_batch_impl(c, n, descs, plcf, tb, iq_out, 0.f, n_samples_out);
}

int dnrp_tx_batch_sc16(dnrp_ctx_t* c, uint32_t n, const dnrp_tx_desc_t* descs, const uint8_t* plcf, const uint8_t* tb,
                       int16_t* iq_out, float scale, uint32_t* n_samples_out) {
    DNRP_REQUIRE(iq_out && scale > 0.f, "dnrp_tx_batch_sc16 needs an output buffer and a scale > 0");
    return tx_batch_impl(c, n, descs, plcf, tb, iq_out, scale, n_samples_out);
}

static int tx_batch_impl(dnrp_ctx_t* c, uint32_t n, const dnrp_tx_desc_t* descs, const uint8_t* plcf, const uint8_t* tb,
                         void* iq_out, float sc16_scale, uint32_t* n_samples_out) {
    DNRP_REQUIRE(c && descs && plcf && tb, "null argument");
    DNRP_REQUIRE(n > 0 && n <= c->B, "batch size out of range");
    DNRP_CUDA(cudaSetDevice(c->device));
    const fec_plan_t& f = c->fec;
    const phy_dims_t& d = c->dims;
    std::vector<tx_meta_dev_t> meta(n);
    std::vector<uint32_t> mask(n);
    std::vector<uint8_t> rv(n);
    for (uint32_t i = 0; i < n; ++i) {
        // the reference asserts these (tx.cpp:560-579); here they are argument errors
        DNRP_REQUIRE(descs[i].codebook_index < d.n_cb_idx, "codebook index out of bound");
        DNRP_REQUIRE(descs[i].GI_percentage <= 100, "GI percentage too large");
        DNRP_REQUIRE(descs[i].rv < 4, "rv out of range");
        meta[i] = {descs[i].codebook_index, descs[i].optimal_scaling_DAC, descs[i].DAC_scale, descs[i].iq_phase_rad,
                   descs[i].iq_phase_increment_s2s_post_resampling_rad, descs[i].GI_percentage};
        const bool cl = c->ps.tm.cl != 0, bf = descs[i].codebook_index > 0;  // tx.cpp:179-181, pcc_enc.cpp:168-178
        mask[i] = (!cl && !bf) ? 0x0000u : (cl && !bf) ? 0x5555u : (!cl && bf) ? 0xAAAAu : 0xFFFFu;
        rv[i] = (uint8_t)descs[i].rv;
        if (n_samples_out)
            n_samples_out[i] = c->n_no_gi_rs + (c->n_packet_rs - c->n_no_gi_rs) * descs[i].GI_percentage / 100;  // tx.cpp:556
    }
    DNRP_CUDA(cudaMemcpy2DAsync(c->d_tb_in.p, f.tb_stride, tb, c->ps.N_TB_byte, c->ps.N_TB_byte, n,
                                cudaMemcpyHostToDevice, c->stream));
    DNRP_CUDA(cudaMemcpy2DAsync(c->d_plcf_in.p, 16, plcf, 10, 10, n, cudaMemcpyHostToDevice, c->stream));
    DNRP_CUDA(cudaMemcpyAsync(c->d_plcf_mask.p, mask.data(), 4 * n, cudaMemcpyHostToDevice, c->stream));
    DNRP_CUDA(cudaMemcpyAsync(c->d_rv.p, rv.data(), n, cudaMemcpyHostToDevice, c->stream));
    DNRP_CUDA(cudaMemcpyAsync(c->d_tx_meta.p, meta.data(), n * sizeof(tx_meta_dev_t), cudaMemcpyHostToDevice, c->stream));
    DNRP_CUDA(cudaStreamSynchronize(c->stream));  // the staging vectors above are pageable and local
    int rc = run_fec_encode(c, n, true, true);
    if (rc) return rc;
    if ((rc = run_tx_ofdm(c, n)) != DNRP_OK) return rc;
    const size_t n_iq = (size_t)n * d.N_TX * c->tx_stride;
    if (iq_out && sc16_scale != 0.f) {
        DNRP_REQUIRE(c->tx_stride % 4 == 0, "sc16 output needs a transmit stride that is a multiple of 4");
        if (!c->d_iq_tx_sc16.p) DNRP_CUDA(c->d_iq_tx_sc16.alloc((size_t)c->B * d.N_TX * c->tx_stride * 2));
        DNRP_LAUNCH(c, -1, launch_cf32_to_sc16(c->d_iq_tx.p, c->d_iq_tx_sc16.p, n_iq, sc16_scale, c->stream));
        DNRP_CUDA(cudaMemcpyAsync(iq_out, c->d_iq_tx_sc16.p, n_iq * 2 * sizeof(int16_t), cudaMemcpyDeviceToHost, c->stream));
    } else if (iq_out)
        DNRP_CUDA(cudaMemcpyAsync(iq_out, c->d_iq_tx.p, n_iq * sizeof(float2), cudaMemcpyDeviceToHost, c->stream));
    DNRP_CUDA(cudaStreamSynchronize(c->stream));
    return DNRP_OK;
}

}  // extern "C"

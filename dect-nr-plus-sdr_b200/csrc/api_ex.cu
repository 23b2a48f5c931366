// api_ex.cu -- the parts of the C ABI that make one context serve several networks and PLCF types per batch:
// dnrp_add_network_id, dnrp_tx_batch_ex, dnrp_rx_pdc_batch_ex, dnrp_harq_reset.  The reference takes network id, PLCF
// type and redundancy version per packet from the HARQ process (rx_synced.cpp:325-350, tx.cpp:176-186) and registers
// network ids up front (worker_pool.cpp:109-114, tx_rx.cpp add_new_network_id).
#include <cstring>
#include "ctx.hpp"

using namespace dnrp;

#define DNRP_REQUIRE(cond, msg)             \
    do {                                    \
        if (!(cond)) {                      \
            dnrp::set_last_error(msg);      \
            return DNRP_ERR_INVALID;        \
        }                                   \
    } while (0)

namespace {
// per-packet (network id, PLCF type) -> index of the scrambling sequence in the context's table
int fill_gold_sel(dnrp_ctx* c, uint32_t n, const uint32_t* network_id, const uint8_t* plcf_type, std::vector<uint32_t>& sel) {
    sel.resize(n);
    for (uint32_t i = 0; i < n; ++i) {
        const uint32_t id = network_id ? network_id[i] : c->cfg.network_id;
        const uint32_t ty = plcf_type ? plcf_type[i] : c->cfg.plcf_type;
        DNRP_REQUIRE(ty == 1 || ty == 2, "plcf_type must be 1 or 2");
        const int g = gold_index(c, id, ty);
        if (g < 0) {
            set_last_error("network id " + std::to_string(id) + " is not registered with this context (dnrp_add_network_id)");
            return DNRP_ERR_STATE;
        }
        sel[i] = (uint32_t)g;
    }
    return DNRP_OK;
}
struct flag_guard {   // the *_ex calls switch the per-packet selection on for their own duration only
    bool& f;
    explicit flag_guard(bool& f_) : f(f_) { f = true; }
    ~flag_guard() { f = false; }
};
}  // namespace

extern "C" {

int dnrp_add_network_id(dnrp_ctx_t* c, uint32_t network_id) {
    DNRP_REQUIRE(c, "null argument");
    DNRP_CUDA(cudaSetDevice(c->device));
    bool added = false;
    for (uint32_t ty = 1; ty <= 2; ++ty)
        if (gold_index(c, network_id, ty) < 0) {
            c->gold_cinit.push_back(ty == 1 ? (network_id & 0xff) : (network_id >> 8));
            added = true;
        }
    return added ? upload_gold_table(c) : DNRP_OK;
}

int dnrp_harq_reset(dnrp_ctx_t* c, uint32_t harq_slot) {
    DNRP_REQUIRE(c, "null argument");
    DNRP_REQUIRE(harq_slot < c->B, "harq slot out of range");
    DNRP_CUDA(cudaSetDevice(c->device));
    const fec_plan_t& f = c->fec;
    // what process_rx_t::reset_and_terminate_func does (process_rx.cpp:56-66): soft values, code block CRC flags
    DNRP_CUDA(cudaMemsetAsync(c->d_soft.p + (size_t)harq_slot * f.soft_stride, 0, f.soft_stride * sizeof(int16_t), c->stream));
    DNRP_CUDA(cudaMemsetAsync(c->d_cb_crc.p + (size_t)harq_slot * c->ps.C, 0, c->ps.C, c->stream));
    if (c->d_saved.p) DNRP_CUDA(cudaMemsetAsync(c->d_saved.p + (size_t)harq_slot * f.saved_stride, 0, f.saved_stride, c->stream));
    DNRP_CUDA(cudaStreamSynchronize(c->stream));
    return DNRP_OK;
}

int dnrp_tx_batch_ex(dnrp_ctx_t* c, uint32_t n, const dnrp_tx_desc_t* descs, const uint32_t* network_id, const uint8_t* plcf_type,
                     const uint8_t* plcf, const uint8_t* tb, dnrp_cf32_t* iq_out, uint32_t* n_samples_out) {
    DNRP_REQUIRE(c && descs && plcf && tb, "null argument");
    DNRP_REQUIRE(n > 0 && n <= c->B, "batch size out of range");
    DNRP_CUDA(cudaSetDevice(c->device));
    std::vector<uint32_t> sel;
    int rc = fill_gold_sel(c, n, network_id, plcf_type, sel);
    if (rc) return rc;
    std::vector<uint8_t> pcc_sel(n);
    for (uint32_t i = 0; i < n; ++i) pcc_sel[i] = (uint8_t)((plcf_type ? plcf_type[i] : c->cfg.plcf_type) - 1);
    DNRP_CUDA(cudaMemcpyAsync(c->d_gold_sel.p, sel.data(), 4 * (size_t)n, cudaMemcpyHostToDevice, c->stream));
    DNRP_CUDA(cudaMemcpyAsync(c->d_pcc_sel.p, pcc_sel.data(), n, cudaMemcpyHostToDevice, c->stream));
    DNRP_CUDA(cudaStreamSynchronize(c->stream));
    flag_guard g1(c->use_gold_sel), g2(c->use_pcc_sel);
    return dnrp_tx_batch(c, n, descs, plcf, tb, iq_out, n_samples_out);
}

int dnrp_rx_pdc_batch_ex(dnrp_ctx_t* c, uint32_t n, const uint32_t* harq_slot, const uint8_t* rv, const uint8_t* harq_reset,
                         const uint32_t* network_id, const uint8_t* plcf_type, uint8_t* tb_out, dnrp_pdc_report_t* out,
                         uint8_t* n_iter) {
    DNRP_REQUIRE(c && tb_out && out, "null argument");
    DNRP_REQUIRE(n > 0 && n <= c->B, "batch size out of range");
    DNRP_CUDA(cudaSetDevice(c->device));
    std::vector<uint32_t> sel;
    int rc = fill_gold_sel(c, n, network_id, plcf_type, sel);
    if (rc) return rc;
    DNRP_CUDA(cudaMemcpyAsync(c->d_gold_sel.p, sel.data(), 4 * (size_t)n, cudaMemcpyHostToDevice, c->stream));
    DNRP_CUDA(cudaStreamSynchronize(c->stream));
    flag_guard g(c->use_gold_sel);
    return dnrp_rx_pdc_batch(c, n, harq_slot, rv, harq_reset, tb_out, out, n_iter);
}

}  // extern "C"

// ---- several packet geometries behind one receiver: windows travel between contexts on the device ---------------------
namespace {
__global__ void k_copy_windows(float2* __restrict__ dst, uint32_t dst_stride, const float2* __restrict__ src, uint32_t src_stride,
                               const uint32_t* __restrict__ src_idx, uint32_t n_ant) {
    const uint32_t i = blockIdx.z, a = blockIdx.y;
    const float2* s = src + ((size_t)src_idx[i] * n_ant + a) * src_stride;
    float2* d = dst + ((size_t)i * n_ant + a) * dst_stride;
    for (uint32_t n = blockIdx.x * blockDim.x + threadIdx.x; n < dst_stride; n += gridDim.x * blockDim.x)
        d[n] = n < src_stride ? s[n] : make_float2(0.f, 0.f);
}
}  // namespace

extern "C" int dnrp_rx_import_windows(dnrp_ctx_t* dst, dnrp_ctx_t* src, uint32_t n, const uint32_t* src_idx) {
    DNRP_REQUIRE(dst && src && src_idx, "null argument");
    DNRP_REQUIRE(dst->device == src->device, "both contexts must live on the same device");
    DNRP_REQUIRE(dst->dims.N_RX == src->dims.N_RX, "both contexts must have the same number of receive antennas");
    DNRP_REQUIRE(dst->cfg.oversampling == src->cfg.oversampling && dst->cfg.resample_L == src->cfg.resample_L &&
                     dst->cfg.resample_M == src->cfg.resample_M && dst->cfg.psdef.u == src->cfg.psdef.u &&
                     dst->cfg.psdef.b == src->cfg.psdef.b,
                 "both contexts must share mu, beta, oversampling and resampling (one sample rate)");
    DNRP_REQUIRE(n > 0 && n <= dst->B, "batch size out of range");
    for (uint32_t i = 0; i < n; ++i) DNRP_REQUIRE(src_idx[i] < src->B, "source window index out of range");
    DNRP_CUDA(cudaSetDevice(dst->device));
    DNRP_CUDA(cudaStreamSynchronize(src->stream));   // the source windows are complete
    DNRP_CUDA(cudaMemcpyAsync(dst->d_slot.p, src_idx, 4 * (size_t)n, cudaMemcpyHostToDevice, dst->stream));
    const dim3 grid((dst->rx_stride + 1023) / 1024, dst->dims.N_RX, n);
    k_copy_windows<<<grid, 256, 0, dst->stream>>>(dst->d_iq_rx.p, dst->rx_stride, src->d_iq_rx.p, src->rx_stride,
                                                  dst->d_slot.p, dst->dims.N_RX);
    DNRP_CUDA(cudaGetLastError());
    DNRP_CUDA(cudaStreamSynchronize(dst->stream));   // src_idx is caller memory
    ++dst->launches;
    return DNRP_OK;
}

// api_stream.cu -- streaming front door: packet search over one continuous IQ stream and hand-over of the found
// packets to the batched receiver.  Batched restatement of the reference's chunked, multi-threaded search:
//   worker_sync_t::work          lib/src/phy/pool/worker_sync.cpp:112-221  (search a chunk, queue every packet found)
//   sync_chunk_t::search         lib/src/phy/rx/sync/sync_chunk.cpp:146-283 (detection -> coarse peak -> skip_after_peak
//                                -> fine peak, continue in the same chunk)
//   baton_t::is_sync_time_unique lib/src/phy/pool/baton.cpp:157-169        (double detections in chunk overlaps)
//   rx_pacer_t                   lib/src/phy/rx/rx_pacer.cpp:106-143,265-327 (resampling of the stream with carried
//                                history, conversion of resampled to global times)
// Every chunk is one "window" of the batched sync kernels (phy_sync.cu), addressed inside the stream through
// sync_params_t::iq_pkt_stride; the per-chunk continuation is the resume array read by k_sync_peak.  Two entry points
// share the search: dnrp_stream_sync (the whole stream in one call, L = M) and dnrp_stream_open / dnrp_stream_push
// (a ring on the device that keeps the overlap and unfinished packets between calls, any L / M).
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <vector>
#include "ctx.hpp"

using namespace dnrp;

#define DNRP_REQUIRE(cond, msg)             \
    do {                                    \
        if (!(cond)) {                      \
            dnrp::set_last_error(msg);      \
            return DNRP_ERR_INVALID;        \
        }                                   \
    } while (0)

namespace dnrp {

// window i of the receiver <- stream[start_i - t0, start_i - t0 + stride) per antenna, zeros outside the stream
__global__ void __launch_bounds__(256) k_stream_extract(const float2* __restrict__ stream, size_t stream_stride,
                                                        long long stream_base, long long stream_len,
                                                        const long long* __restrict__ start, uint32_t t0,
                                                        float2* __restrict__ win, uint32_t stride, uint32_t N_RX) {
    const uint32_t pkt = blockIdx.z, a = blockIdx.y;
    const long long s0 = start[pkt] - (long long)t0 - stream_base;
    const float2* src = stream + (size_t)a * stream_stride;
    float2* dst = win + ((size_t)pkt * N_RX + a) * stride;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < stride; i += gridDim.x * blockDim.x) {
        const long long s = s0 + i;
        dst[i] = (s >= 0 && s < stream_len) ? src[s] : make_float2(0.f, 0.f);
    }
}

namespace {

struct found_t {
    uint32_t chunk, round;
    sync_result_dev_t r;
};

// work buffers for n_chunks chunks of n_steps detection steps (grown on demand, never on a warm path)
int stream_work_alloc(dnrp_ctx* c, uint32_t n_chunks, uint32_t n_steps) {
    if (n_chunks <= c->stream_chunks_alloc && n_steps <= c->stream_steps_alloc) return DNRP_OK;
    const phy_dims_t& d = c->dims;
    const uint32_t nc = std::max(n_chunks, c->stream_chunks_alloc), ns = std::max(n_steps, c->stream_steps_alloc);
    DNRP_CUDA(c->d_stream_c.alloc((size_t)nc * d.N_RX * ns));
    DNRP_CUDA(c->d_stream_p.alloc((size_t)nc * d.N_RX * ns));
    DNRP_CUDA(c->d_stream_xc.alloc((size_t)nc * c->sync_n_tpl));
    DNRP_CUDA(c->d_stream_result.alloc(nc));
    DNRP_CUDA(c->d_stream_resume.alloc(nc));
    DNRP_CUDA(c->d_stream_limit.alloc(nc));
    DNRP_CUDA(c->d_stream_map.alloc(nc));
    c->stream_chunks_alloc = nc; c->stream_steps_alloc = ns;
    return DNRP_OK;
}

// the part of sync_params_t that does not depend on where the chunks lie
sync_params_t stream_sync_params(dnrp_ctx* c, uint32_t len, uint32_t n_steps) {
    const phy_dims_t& d = c->dims;
    sync_params_t p{};
    p.d = d;
    p.len = len;
    p.rs_L = c->cfg.resample_L; p.rs_M = c->cfg.resample_M;
    p.xc_S = c->stf_len_rs;
    p.xc_sl = (uint32_t)((uint64_t)d.pat_len_os * p.rs_L / p.rs_M);   // crosscorrelator.cpp:52-56
    p.n_no_gi_hw = c->n_no_gi_rs;
    p.n_ant = d.N_RX;
    p.step = c->sync_step; p.n_steps = n_steps;
    p.rms_min = 0.005f * (float)std::sqrt((double)d.u * d.b * c->cfg.oversampling * 1.728e6 / 30.72e6);
    p.c_steps = c->d_stream_c.p; p.p_steps = c->d_stream_p.p;
    p.templates = c->d_sync_tpl.p; p.n_tpl = c->sync_n_tpl;
    p.xc = c->d_stream_xc.p;
    p.xc_fft_n = c->sync_fft_n; p.tpl_spec = c->d_sync_tpl_spec.p; p.xc_twiddle = c->d_sync_tw.p;
    p.xc_twiddle_p16 = c->d_sync_tw_p16.p;
    p.result = c->d_stream_result.p;
    p.rx_state = nullptr;
    p.resume = c->d_stream_resume.p;
    p.limit = c->d_stream_limit.p;
    return p;
}

// Search n_chunks chunks: the per-step sums once, then rounds of detection -> coarse peak -> fine peak over the chunks
// that are still active (a chunk that found a packet continues behind it, skip_after_peak = 2 STFs, until its detection
// range [resume, limit) is used up).  over[k] = how far behind limit[k] the skip of the chunk's last packet reaches.
int stream_search_chunks(dnrp_ctx* c, sync_params_t p, uint32_t n_chunks, std::vector<uint32_t>& resume,
                         const std::vector<uint32_t>& limit, std::vector<found_t>& found, std::vector<uint32_t>& over) {
    const phy_dims_t& d = c->dims;
    const uint32_t stf = d.stf_len_os, step = p.step, skip = 2 * stf;   // ..._DETECTION_SKIP_AFTER_PEAK_IN_STFS_DP
    const bool dbg = std::getenv("DNRP_STREAM_DEBUG") != nullptr;
    DNRP_CUDA(cudaMemcpyAsync(c->d_stream_limit.p, limit.data(), n_chunks * sizeof(uint32_t), cudaMemcpyHostToDevice, c->stream));
    DNRP_LAUNCH(c, DNRP_KERNEL_SYNC, launch_sync_steps(p, n_chunks, c->stream));
    over.assign(n_chunks, 0u);
    std::vector<uint32_t> active(n_chunks);
    std::iota(active.begin(), active.end(), 0u);
    std::vector<sync_result_dev_t> res(n_chunks);
    for (uint32_t round = 0; round < 4096 && !active.empty(); ++round) {
        const uint32_t na = (uint32_t)active.size();
        DNRP_CUDA(cudaMemcpyAsync(c->d_stream_resume.p, resume.data(), n_chunks * sizeof(uint32_t), cudaMemcpyHostToDevice, c->stream));
        DNRP_CUDA(cudaMemcpyAsync(c->d_stream_map.p, active.data(), na * sizeof(uint32_t), cudaMemcpyHostToDevice, c->stream));
        sync_params_t q = p;
        q.chunk_map = c->d_stream_map.p;
        c->launches += 2;
        DNRP_LAUNCH(c, DNRP_KERNEL_SYNC, launch_sync_search(q, na, c->stream));
        DNRP_CUDA(cudaMemcpyAsync(res.data(), c->d_stream_result.p, n_chunks * sizeof(sync_result_dev_t), cudaMemcpyDeviceToHost,
                                  c->stream));
        DNRP_CUDA(cudaStreamSynchronize(c->stream));
        std::vector<uint32_t> next_active;
        for (uint32_t k : active) {
            const sync_result_dev_t& r = res[k];
            if (dbg) std::fprintf(stderr, "round %u chunk %u resume %u limit %u: det %u t_det %u coarse %u fine %u ndet %u\n", round, k,
                                  resume[k], limit[k], r.detected, r.detection_time_local, r.coarse_peak_time_local,
                                  r.fine_peak_time_local, r.n_detections);
            uint32_t next;
            if (!r.detected) {
                // detection range used up (0), eight invalid peaks in a row (go on behind the last one), or a coarse
                // peak whose fine search range left the chunk (skip it like a packet)
                next = r.coarse_peak_time_local ? r.coarse_peak_time_local + skip : r.detection_time_local;
                if (next != 0 && next <= resume[k]) next = resume[k] + step;   // always advance
                if (next == 0) next = 0xffffffffu;
            } else {
                found.push_back({k, round, r});
                next = r.coarse_peak_time_local + skip;                         // skip_after_peak
            }
            if (next != 0xffffffffu && next >= limit[k]) {
                // finished; a packet at the very end of the range reaches behind it with its skip: kept for the caller
                if (r.detected) over[k] = next - limit[k];
                next = 0xffffffffu;
            }
            resume[k] = next;
            if (next != 0xffffffffu) next_active.push_back(k);
        }
        active.swap(next_active);
    }
    return DNRP_OK;
}

}  // namespace
}  // namespace dnrp


namespace {

// largest dynamic shared memory a block may opt into on the context's device
int smem_optin_limit(int device, size_t& out) {
    int v = 0;
    DNRP_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
    out = (size_t)v;
    return DNRP_OK;
}

// time order, unique packets only (baton_t::is_sync_time_unique: later than the predecessor by more than one STF pattern)
void order_and_dedupe(std::vector<dnrp_stream_packet_t>& v, int64_t& last, int64_t pat_hw) {
    std::stable_sort(v.begin(), v.end(), [](const dnrp_stream_packet_t& a, const dnrp_stream_packet_t& b) {
        return a.fine_peak_time_64 < b.fine_peak_time_64;
    });
    std::vector<dnrp_stream_packet_t> u;
    for (const auto& f : v) {
        if (f.fine_peak_time_64 - last <= pat_hw) continue;
        last = f.fine_peak_time_64;
        u.push_back(f);
    }
    v.swap(u);
}

}  // namespace

extern "C" {

int dnrp_stream_sync(dnrp_ctx_t* c, uint64_t n_samples, const dnrp_cf32_t* iq, uint32_t chunk_hop, uint32_t max_out,
                     dnrp_stream_packet_t* out, uint32_t* n_out) {
    DNRP_REQUIRE(c && iq && out && n_out, "null argument");
    DNRP_REQUIRE(!c->resampling, "dnrp_stream_sync needs a context without resampling (L = M); use dnrp_stream_open / dnrp_stream_push");
    DNRP_CUDA(cudaSetDevice(c->device));
    const phy_dims_t& d = c->dims;
    const uint32_t stf = d.stf_len_os, pat = d.pat_len_os, step = c->sync_step;
    // history in front of the first detection position: the detector's STF + one pattern, plus one pattern so that the
    // cross-correlation range (+- one pattern around the coarse peak) of a packet detected there stays inside the chunk
    const uint32_t lead = stf + 2 * pat;
    uint32_t hop = chunk_hop ? chunk_hop : 4 * c->n_packet_rs;
    hop = std::max(hop, 2 * stf);
    hop = (hop + step - 1) / step * step;                  // chunks start on the detection step grid
    // behind the last detection position: coarse peak search (1 STF from detection - 1 pattern), cross-correlation
    // range (+- one pattern) and the STF template
    const uint32_t tail = 2 * stf + 3 * pat;
    const uint32_t len = (lead + hop + tail + step - 1) / step * step;
    DNRP_REQUIRE(n_samples > lead + stf, "stream shorter than one STF plus its history");
    const uint64_t span = n_samples - lead;
    const uint64_t n_chunks64 = (span + hop - 1) / hop;
    DNRP_REQUIRE(n_chunks64 <= 65535, "more than 65535 chunks: raise chunk_hop");
    const uint32_t n_chunks = (uint32_t)n_chunks64;
    const uint32_t n_steps = len / step;
    size_t smem_max = 0;
    int rc = smem_optin_limit(c->device, smem_max);
    if (rc) return rc;
    if (sync_peak_smem(d, n_steps) > smem_max) {
        set_last_error("chunk_hop " + std::to_string(hop) + " needs " + std::to_string(sync_peak_smem(d, n_steps)) +
                       " B of shared memory per block for the detection prefix sums, the device allows " + std::to_string(smem_max));
        return DNRP_ERR_INVALID;
    }
    *n_out = 0;

    // ---- stream to the device, zero padded so that every chunk has the full length
    const size_t need = (size_t)(n_chunks - 1) * hop + len;
    if (need > c->stream_alloc) {
        DNRP_CUDA(c->d_stream.alloc((size_t)d.N_RX * need));   // alloc() zero fills and synchronises
        c->stream_alloc = need;
    } else {
        DNRP_CUDA(cudaMemsetAsync(c->d_stream.p, 0, (size_t)d.N_RX * c->stream_alloc * sizeof(float2), c->stream));
    }
    c->stream_len = (size_t)n_samples;
    c->stream_view = {c->d_stream.p, c->stream_alloc, (size_t)n_samples, 0};
    DNRP_CUDA(cudaMemcpy2DAsync(c->d_stream.p, c->stream_alloc * sizeof(float2), iq, (size_t)n_samples * sizeof(float2),
                                (size_t)n_samples * sizeof(float2), d.N_RX, cudaMemcpyHostToDevice, c->stream));
    rc = stream_work_alloc(c, n_chunks, n_steps);
    if (rc) return rc;

    sync_params_t p = stream_sync_params(c, len, n_steps);
    p.iq = p.iq_x = c->d_stream.p;
    p.iq_stride = p.x_stride = c->stream_alloc;            // antenna stride
    p.iq_pkt_stride = p.x_pkt_stride = hop;                // chunk stride
    p.x_len = len;
    p.search_length = lead + hop;                          // detection positions [lead, lead + hop) of every chunk

    std::vector<uint32_t> resume(n_chunks, lead), limit(n_chunks, lead + hop), over;
    resume[0] = stf + pat;   // nothing in front of the first chunk: the detector's own ignore_before_index
    std::vector<found_t> hits;
    rc = stream_search_chunks(c, p, n_chunks, resume, limit, hits, over);
    if (rc) return rc;
    std::vector<dnrp_stream_packet_t> found;
    for (const found_t& h : hits) {
        const int64_t base = (int64_t)h.chunk * hop;
        if ((uint64_t)(base + h.r.fine_peak_time_local) + stf > n_samples) continue;
        dnrp_stream_packet_t f{};
        f.fine_peak_time_64 = base + h.r.fine_peak_time_local;
        f.coarse_peak_time_64 = base + h.r.coarse_peak_time_local;
        static_assert(sizeof(f.sync) == sizeof(h.r), "sync result layouts differ");
        std::memcpy(&f.sync, &h.r, sizeof(h.r));
        f.chunk = h.chunk; f.round = h.round;
        found.push_back(f);
    }
    int64_t last = INT64_MIN / 2;
    order_and_dedupe(found, last, (int64_t)pat);
    const uint32_t n = (uint32_t)found.size();
    for (uint32_t i = 0; i < n && i < max_out; ++i) out[i] = found[i];
    *n_out = n;
    if (n > max_out) {
        dnrp::set_last_error("dnrp_stream_sync: more packets found than max_out");
        return DNRP_ERR_INVALID;
    }
    return DNRP_OK;
}

int dnrp_stream_extract(dnrp_ctx_t* c, uint32_t n, const int64_t* start_64, uint32_t t0) {
    DNRP_REQUIRE(c && start_64, "null argument");
    DNRP_REQUIRE(n > 0 && n <= c->B, "batch size out of range");
    DNRP_REQUIRE(c->stream_view.p && c->stream_view.len, "no stream on the device (dnrp_stream_sync / dnrp_stream_push first)");
    DNRP_REQUIRE(t0 + c->n_no_gi_rs <= c->rx_stride, "t0 leaves no room for the packet inside the window");
    DNRP_CUDA(cudaSetDevice(c->device));
    static_assert(sizeof(long long) == sizeof(int64_t), "64-bit times");
    dbuf<long long> d_start;
    DNRP_CUDA(d_start.alloc(n));
    DNRP_CUDA(cudaMemcpyAsync(d_start.p, start_64, n * sizeof(int64_t), cudaMemcpyHostToDevice, c->stream));
    const phy_dims_t& d = c->dims;
    const uint32_t bx = std::min<uint32_t>((c->rx_stride + 255) / 256, 16);
    ++c->launches;
    k_stream_extract<<<dim3(bx, d.N_RX, n), 256, 0, c->stream>>>(c->stream_view.p, c->stream_view.stride, (long long)c->stream_view.base,
                                                                  (long long)c->stream_view.len, d_start.p, t0,
                                                                  c->d_iq_rx.p + c->off * d.N_RX * c->rx_stride, c->rx_stride, d.N_RX);
    DNRP_CUDA(cudaGetLastError());
    DNRP_CUDA(cudaStreamSynchronize(c->stream));
    return DNRP_OK;
}

}  // extern "C"


// ---- stateful front door: a ring of the stream on the device ------------------------------------------------------------
// Absolute sample indices count from the first sample ever pushed: h at the hardware rate, r at the DECT rate (x
// oversampling); r corresponds to h = r L / M (rx_pacer.cpp:311-318).  Chunk windows start on a grid Q = lcm(step, 2 M)
// so that every window starts on the detection step grid AND on an integer, even hardware-rate sample.
extern "C" {

int dnrp_stream_open(dnrp_ctx_t* c, uint32_t max_push_samples, uint32_t chunk_hop) {
    DNRP_REQUIRE(c, "null argument");
    DNRP_REQUIRE(max_push_samples > 0, "max_push_samples must be positive");
    DNRP_CUDA(cudaSetDevice(c->device));
    const phy_dims_t& d = c->dims;
    auto& ss = c->ss;
    const uint32_t L = c->cfg.resample_L, M = c->cfg.resample_M;
    const uint32_t stf = d.stf_len_os, pat = d.pat_len_os, step = c->sync_step;
    ss.Q = c->resampling ? (uint32_t)std::lcm<uint64_t>(step, 2ull * M) : step;
    ss.lead = (stf + 2 * pat + ss.Q - 1) / ss.Q * ss.Q;
    uint32_t hop = chunk_hop ? chunk_hop : 4 * d.n_packet_os;
    hop = std::max(hop, 2 * stf);
    ss.hop = (hop + ss.Q - 1) / ss.Q * ss.Q;
    ss.tail = 2 * stf + 3 * pat;
    ss.len = (ss.lead + ss.hop + ss.tail + step - 1) / step * step;
    size_t smem_max = 0;
    int rc = smem_optin_limit(c->device, smem_max);
    if (rc) return rc;
    if (sync_peak_smem(d, ss.len / step) > smem_max) {
        set_last_error("chunk_hop " + std::to_string(ss.hop) + " needs " + std::to_string(sync_peak_smem(d, ss.len / step)) +
                       " B of shared memory per block for the detection prefix sums, the device allows " + std::to_string(smem_max));
        return DNRP_ERR_INVALID;
    }
    ss.max_push = max_push_samples;
    const uint64_t push_rs = ((uint64_t)max_push_samples * M + L - 1) / L + 2;
    // between two pushes the ring holds the history of the next chunk, what could not be searched yet (< tail + Q) and
    // the samples of packets that are not complete yet; then one push arrives; one chunk length of zeros behind the data
    ss.cap_rs = ((size_t)ss.lead + ss.tail + ss.Q + push_rs + ss.len + 64 + 3) & ~(size_t)3;
    const uint64_t keep_hw = ((uint64_t)(ss.lead + ss.tail + ss.Q) * L) / M + c->n_packet_rs + 4ull * c->stf_len_rs +
                             (uint64_t)c->rs_rx.subfilter_length + 8;
    ss.cap_hw = ((size_t)keep_hw + max_push_samples + ((size_t)ss.len * L) / M + 64 + 3) & ~(size_t)3;
    ss.max_chunks = (uint32_t)((push_rs + ss.tail + ss.Q) / ss.hop + 2);
    for (int i = 0; i < 2; ++i) {
        DNRP_CUDA(ss.hw[i].alloc((size_t)d.N_RX * ss.cap_hw));
        if (c->resampling) DNRP_CUDA(ss.rs[i].alloc((size_t)d.N_RX * ss.cap_rs));
    }
    rc = stream_work_alloc(c, ss.max_chunks, ss.len / step);
    if (rc) return rc;
    ss.cur = 0;
    ss.hw_base = ss.rs_base = 0;
    ss.hw_len = ss.rs_len = 0;
    ss.search_from = ss.lead;   // the first chunk window starts at sample 0
    ss.carry_resume = 0;
    ss.last_fine = INT64_MIN / 2;
    ss.first = true;
    ss.pending.clear();
    ss.open = true;
    c->stream_view = {ss.hw[0].p, ss.cap_hw, 0, 0};
    return DNRP_OK;
}

int dnrp_stream_close(dnrp_ctx_t* c) {
    DNRP_REQUIRE(c, "null argument");
    DNRP_CUDA(cudaSetDevice(c->device));
    DNRP_CUDA(cudaStreamSynchronize(c->stream));
    auto& ss = c->ss;
    for (int i = 0; i < 2; ++i) {
        DNRP_CUDA(ss.hw[i].alloc(0));
        DNRP_CUDA(ss.rs[i].alloc(0));
    }
    ss.pending.clear();
    ss.open = false;
    c->stream_view = {};
    return DNRP_OK;
}

// sc16_scale != 0: iq holds int16 I/Q pairs [N_RX][n_samples][2] (the radio's wire format), converted on the device
static int stream_push_impl(dnrp_ctx_t* c, uint32_t n_samples, const void* iq, float sc16_scale, uint32_t max_out,
                            dnrp_stream_packet_t* out, uint32_t* n_out);

int dnrp_stream_push(dnrp_ctx_t* c, uint32_t n_samples, const dnrp_cf32_t* iq, uint32_t max_out, dnrp_stream_packet_t* out,
                     uint32_t* n_out) {
    return stream_push_impl(c, n_samples, iq, 0.f, max_out, out, n_out);
}

int dnrp_stream_push_sc16(dnrp_ctx_t* c, uint32_t n_samples, const int16_t* iq, float scale, uint32_t max_out,
                          dnrp_stream_packet_t* out, uint32_t* n_out) {
    DNRP_REQUIRE(scale > 0.f, "scale must be > 0");
    return stream_push_impl(c, n_samples, iq, scale, max_out, out, n_out);
}

static int stream_push_impl(dnrp_ctx_t* c, uint32_t n_samples, const void* iq, float sc16_scale, uint32_t max_out,
                            dnrp_stream_packet_t* out, uint32_t* n_out) {
    DNRP_REQUIRE(c && out && n_out, "null argument");
    DNRP_REQUIRE(n_samples == 0 || iq, "null argument");
    auto& ss = c->ss;
    DNRP_REQUIRE(ss.open, "dnrp_stream_open first");
    DNRP_REQUIRE(n_samples <= ss.max_push, "push larger than max_push_samples of dnrp_stream_open");
    DNRP_CUDA(cudaSetDevice(c->device));
    const phy_dims_t& d = c->dims;
    const uint32_t L = c->cfg.resample_L, M = c->cfg.resample_M;   // hardware rate = DECT rate * L / M
    const uint32_t stf = d.stf_len_os, pat = d.pat_len_os, step = c->sync_step;
    const bool rsm = c->resampling;
    const spec::resampler_design_t& r = c->rs_rx;
    *n_out = 0;

    // ---- 1. drop what nobody needs any more: everything in front of the next chunk's history, of the packets still
    // waiting for their last samples, and of the resampler's history
    const int64_t win_r = ss.search_from - ss.lead;                    // first sample of the next chunk window (on Q)
    int64_t keep_h = win_r * L / M;                                     // exact: Q is a multiple of 2 M
    for (const auto& f : ss.pending) keep_h = std::min<int64_t>(keep_h, f.fine_peak_time_64 - 2 * (int64_t)c->stf_len_rs);
    if (rsm) {
        const int64_t next_out = ss.rs_base + (int64_t)ss.rs_len;      // first DECT-rate sample not produced yet
        keep_h = std::min<int64_t>(keep_h, ((int64_t)r.delay + next_out * r.M) / r.L - (int64_t)r.subfilter_length - 2);
    }
    keep_h = std::max<int64_t>(keep_h & ~(int64_t)1, ss.hw_base);
    const int64_t keep_r = rsm ? std::max<int64_t>(std::min<int64_t>(win_r, ss.rs_base + (int64_t)ss.rs_len), ss.rs_base) : 0;
    if (keep_h > ss.hw_base || (rsm && keep_r > ss.rs_base)) {
        // the kept parts move to the front of the other buffer pair (no overlapping copy)
        const int nxt = 1 - ss.cur;
        const size_t drop = (size_t)(keep_h - ss.hw_base), left = ss.hw_len - std::min(ss.hw_len, drop);
        if (left)
            DNRP_CUDA(cudaMemcpy2DAsync(ss.hw[nxt].p, ss.cap_hw * sizeof(float2), ss.hw[ss.cur].p + drop, ss.cap_hw * sizeof(float2),
                                        left * sizeof(float2), d.N_RX, cudaMemcpyDeviceToDevice, c->stream));
        ss.hw_base = keep_h; ss.hw_len = left;
        if (rsm) {
            const size_t drop_r = (size_t)(keep_r - ss.rs_base), left_r = ss.rs_len - drop_r;
            if (left_r)
                DNRP_CUDA(cudaMemcpy2DAsync(ss.rs[nxt].p, ss.cap_rs * sizeof(float2), ss.rs[ss.cur].p + drop_r,
                                            ss.cap_rs * sizeof(float2), left_r * sizeof(float2), d.N_RX, cudaMemcpyDeviceToDevice,
                                            c->stream));
            ss.rs_base = keep_r; ss.rs_len = left_r;
        }
        ss.cur = nxt;
    }
    float2* hw = ss.hw[ss.cur].p;
    float2* rs = rsm ? ss.rs[ss.cur].p : hw;
    DNRP_REQUIRE(ss.hw_len + n_samples + ((size_t)ss.len * L) / M + 8 <= ss.cap_hw, "stream ring overflow (hardware rate)");

    // ---- 2. append the new samples, zeros behind them (the last chunk window reads past the data)
    if (n_samples && sc16_scale != 0.f) {
        if (c->d_stream_sc16.n < (size_t)d.N_RX * ss.max_push * 2) DNRP_CUDA(c->d_stream_sc16.alloc((size_t)d.N_RX * ss.max_push * 2));
        DNRP_CUDA(cudaMemcpyAsync(c->d_stream_sc16.p, iq, (size_t)d.N_RX * n_samples * 2 * sizeof(int16_t), cudaMemcpyHostToDevice, c->stream));
        DNRP_LAUNCH(c, -1, launch_sc16_rows_to_cf32(c->d_stream_sc16.p, n_samples, hw + ss.hw_len, ss.cap_hw, n_samples, d.N_RX,
                                                    sc16_scale, c->stream));
    } else if (n_samples)
        DNRP_CUDA(cudaMemcpy2DAsync(hw + ss.hw_len, ss.cap_hw * sizeof(float2), iq, (size_t)n_samples * sizeof(float2),
                                    (size_t)n_samples * sizeof(float2), d.N_RX, cudaMemcpyHostToDevice, c->stream));
    ss.hw_len += n_samples;
    {
        const size_t z = std::min(ss.cap_hw - ss.hw_len, ((size_t)ss.len * L) / M + 8);
        DNRP_CUDA(cudaMemset2DAsync(hw + ss.hw_len, ss.cap_hw * sizeof(float2), 0, z * sizeof(float2), d.N_RX, c->stream));
    }
    const int64_t hw_total = ss.hw_base + (int64_t)ss.hw_len;

    // ---- 3. the stream at the DECT rate: every output whose newest input sample has arrived (rx_pacer_t, no flush)
    int64_t rs_total = hw_total;
    if (rsm) {
        const int64_t m_a = ss.rs_base + (int64_t)ss.rs_len;
        const int64_t t_in = hw_total * r.L - 1 - (int64_t)r.delay;   // outputs m with delay + m M <= t_in
        const int64_t m_hi = t_in >= 0 ? t_in / r.M + 1 : 0;
        if (m_hi > m_a) {
            DNRP_REQUIRE(ss.rs_len + (size_t)(m_hi - m_a) + ss.len + 8 <= ss.cap_rs, "stream ring overflow (DECT rate)");
            const int64_t delay = (int64_t)r.delay + m_a * r.M - ss.hw_base * r.L;   // phase of output m_a against element 0
            DNRP_REQUIRE(delay >= 0 && delay < (1ll << 30), "resampler history lost");
            resample_params_t q{};
            q.L = r.L; q.M = r.M; q.SL = r.subfilter_length; q.delay = (uint32_t)delay;
            q.taps = c->d_rs_rx_taps.p;
            // the antennas of the ring are k_resample's "packets": [antenna][cap] with one stream each
            q.in = hw; q.in_stride = ss.cap_hw; q.n_in = (uint32_t)ss.hw_len;
            q.out = rs + ss.rs_len; q.out_stride = ss.cap_rs;
            q.n_out = (uint32_t)(m_hi - m_a); q.n_fill = q.n_out;
            q.n_ant = 1;
            DNRP_LAUNCH(c, DNRP_KERNEL_RESAMPLE, launch_resample(q, d.N_RX, c->stream));
            ss.rs_len += (size_t)(m_hi - m_a);
        }
        const size_t z = std::min(ss.cap_rs - ss.rs_len, (size_t)ss.len + 8);
        DNRP_CUDA(cudaMemset2DAsync(rs + ss.rs_len, ss.cap_rs * sizeof(float2), 0, z * sizeof(float2), d.N_RX, c->stream));
        rs_total = ss.rs_base + (int64_t)ss.rs_len;
    } else {
        ss.rs_base = ss.hw_base; ss.rs_len = ss.hw_len;
    }
    c->stream_view = {hw, ss.cap_hw, ss.hw_len, ss.hw_base};

    // ---- 4. search the detection positions [search_from, search_from + span): whole chunks of hop, the last one shorter
    std::vector<dnrp_stream_packet_t> fresh;
    const int64_t span64 = (rs_total - (int64_t)ss.tail - ss.search_from) / (int64_t)ss.Q * (int64_t)ss.Q;
    if (span64 > 0) {
        const uint32_t span = (uint32_t)span64;
        const uint32_t n_chunks = (span + ss.hop - 1) / ss.hop;
        DNRP_REQUIRE(n_chunks <= ss.max_chunks, "more chunks than dnrp_stream_open sized the work buffers for");
        sync_params_t p = stream_sync_params(c, ss.len, ss.len / step);
        const int64_t win_h = win_r * L / M;
        p.iq = rs + (size_t)(win_r - ss.rs_base);
        p.iq_stride = rsm ? ss.cap_rs : ss.cap_hw;
        p.iq_pkt_stride = ss.hop;
        p.iq_x = hw + (size_t)(win_h - ss.hw_base);
        p.x_stride = ss.cap_hw;
        p.x_pkt_stride = (size_t)ss.hop * L / M;
        p.x_len = (uint32_t)(((uint64_t)ss.len * L) / M);
        p.search_length = ss.lead + ss.hop;
        std::vector<uint32_t> resume(n_chunks, ss.lead), limit(n_chunks), over;
        for (uint32_t k = 0; k < n_chunks; ++k) limit[k] = ss.lead + std::min(ss.hop, span - k * ss.hop);
        if (ss.first) resume[0] = stf + pat;   // nothing in front of the very first chunk: the detector's own ignore_before_index
        if (ss.carry_resume > ss.search_from) resume[0] = std::max<uint32_t>(resume[0], (uint32_t)(ss.carry_resume - win_r));
        if (resume[0] >= limit[0]) resume[0] = 0xffffffffu;
        ss.first = false;
        std::vector<found_t> hits;
        const int rc = stream_search_chunks(c, p, n_chunks, resume, limit, hits, over);
        if (rc) return rc;
        for (const found_t& h : hits) {
            const int64_t base_h = win_h + (int64_t)h.chunk * (int64_t)p.x_pkt_stride;
            dnrp_stream_packet_t f{};
            f.fine_peak_time_64 = base_h + h.r.fine_peak_time_local;
            f.coarse_peak_time_64 = base_h + h.r.coarse_peak_time_hw;
            std::memcpy(&f.sync, &h.r, sizeof(h.r));
            f.chunk = h.chunk; f.round = h.round;
            fresh.push_back(f);
        }
        ss.search_from += span;
        if (over[n_chunks - 1]) ss.carry_resume = ss.search_from + over[n_chunks - 1];
    }

    // ---- 5. unique packets in time order; a packet is handed out once its last sample is in the ring
    order_and_dedupe(fresh, ss.last_fine, (int64_t)((uint64_t)pat * L / M));
    ss.pending.insert(ss.pending.end(), fresh.begin(), fresh.end());
    uint32_t n = 0;
    std::vector<dnrp_stream_packet_t> wait;
    for (const auto& f : ss.pending) {
        // complete ones beyond max_out stay queued: a call with n_samples = 0 drains them
        if (f.fine_peak_time_64 + (int64_t)c->n_packet_rs <= hw_total && n < max_out) out[n++] = f;
        else wait.push_back(f);
    }
    ss.pending.swap(wait);
    *n_out = n;
    return DNRP_OK;
}

}  // extern "C"

// api_stream.cu -- streaming front door: packet search over one continuous IQ stream and hand-over of the found
// packets to the batched receiver.  Batched restatement of the reference's chunked, multi-threaded search:
//   worker_sync_t::work          lib/src/phy/pool/worker_sync.cpp:112-221  (search a chunk, queue every packet found)
//   sync_chunk_t::search         lib/src/phy/rx/sync/sync_chunk.cpp:146-283 (detection -> coarse peak -> skip_after_peak
//                                -> fine peak, continue in the same chunk)
//   baton_t::is_sync_time_unique lib/src/phy/pool/baton.cpp:157-169        (double detections in chunk overlaps)
// Every chunk is one "window" of the batched sync kernels (phy_sync.cu), addressed inside the stream through
// sync_params_t::iq_pkt_stride; the per-chunk continuation is the resume array read by k_sync_peak.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
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
                                                        long long stream_len, const long long* __restrict__ start,
                                                        uint32_t t0, float2* __restrict__ win, uint32_t stride,
                                                        uint32_t N_RX) {
    const uint32_t pkt = blockIdx.z, a = blockIdx.y;
    const long long s0 = start[pkt] - (long long)t0;
    const float2* src = stream + (size_t)a * stream_stride;
    float2* dst = win + ((size_t)pkt * N_RX + a) * stride;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < stride; i += gridDim.x * blockDim.x) {
        const long long s = s0 + i;
        dst[i] = (s >= 0 && s < stream_len) ? src[s] : make_float2(0.f, 0.f);
    }
}

}  // namespace dnrp

extern "C" {

int dnrp_stream_sync(dnrp_ctx_t* c, uint64_t n_samples, const dnrp_cf32_t* iq, uint32_t chunk_hop, uint32_t max_out,
                     dnrp_stream_packet_t* out, uint32_t* n_out) {
    DNRP_REQUIRE(c && iq && out && n_out, "null argument");
    DNRP_REQUIRE(!c->resampling, "the streaming front door needs a context without resampling (L = M)");
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
    const uint32_t n_chunks = (uint32_t)((span + hop - 1) / hop);
    const uint32_t n_steps = len / step;
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
    DNRP_CUDA(cudaMemcpy2DAsync(c->d_stream.p, c->stream_alloc * sizeof(float2), iq, (size_t)n_samples * sizeof(float2),
                                (size_t)n_samples * sizeof(float2), d.N_RX, cudaMemcpyHostToDevice, c->stream));
    if (n_chunks > c->stream_chunks_alloc || n_steps > c->stream_steps_alloc) {
        const uint32_t nc = std::max(n_chunks, c->stream_chunks_alloc), ns = std::max(n_steps, c->stream_steps_alloc);
        DNRP_CUDA(c->d_stream_c.alloc((size_t)nc * d.N_RX * ns));
        DNRP_CUDA(c->d_stream_p.alloc((size_t)nc * d.N_RX * ns));
        DNRP_CUDA(c->d_stream_xc.alloc((size_t)nc * c->sync_n_tpl));
        DNRP_CUDA(c->d_stream_result.alloc(nc));
        DNRP_CUDA(c->d_stream_resume.alloc(nc));
        c->stream_chunks_alloc = nc; c->stream_steps_alloc = ns;
    }

    sync_params_t p{};
    p.d = d;
    p.iq = p.iq_x = c->d_stream.p;
    p.iq_stride = p.x_stride = c->stream_alloc;            // antenna stride
    p.iq_pkt_stride = p.x_pkt_stride = hop;                // chunk stride
    p.len = p.x_len = len;
    p.rs_L = p.rs_M = 1;
    p.xc_S = c->stf_len_rs;
    p.xc_sl = pat;
    p.n_no_gi_hw = c->n_no_gi_rs;
    p.n_ant = d.N_RX;
    p.step = step; p.n_steps = n_steps;
    p.search_length = lead + hop;                          // detection positions [lead, lead + hop) of every chunk
    p.rms_min = 0.005f * (float)std::sqrt((double)d.u * d.b * c->cfg.oversampling * 1.728e6 / 30.72e6);
    p.c_steps = c->d_stream_c.p; p.p_steps = c->d_stream_p.p;
    p.templates = c->d_sync_tpl.p; p.n_tpl = c->sync_n_tpl;
    p.xc = c->d_stream_xc.p;
    p.xc_fft_n = c->sync_fft_n; p.tpl_spec = c->d_sync_tpl_spec.p; p.xc_twiddle = c->d_sync_tw.p;
    p.result = c->d_stream_result.p;
    p.rx_state = nullptr;
    p.resume = c->d_stream_resume.p;

    const bool dbg = std::getenv("DNRP_STREAM_DEBUG") != nullptr;
    std::vector<uint32_t> resume(n_chunks, lead);
    resume[0] = stf + pat;   // nothing in front of the first chunk: the detector's own ignore_before_index
    std::vector<sync_result_dev_t> res(n_chunks);
    std::vector<dnrp_stream_packet_t> found;
    const uint32_t skip = 2 * stf;                         // RX_SYNC_PARAM_AUTOCORRELATOR_DETECTION_SKIP_AFTER_PEAK_IN_STFS_DP
    for (uint32_t round = 0; round < 256; ++round) {
        DNRP_CUDA(cudaMemcpyAsync(c->d_stream_resume.p, resume.data(), n_chunks * sizeof(uint32_t), cudaMemcpyHostToDevice,
                                  c->stream));
        c->launches += 3;
        DNRP_LAUNCH(c, DNRP_KERNEL_SYNC, launch_sync(p, n_chunks, c->stream));
        DNRP_CUDA(cudaMemcpyAsync(res.data(), c->d_stream_result.p, n_chunks * sizeof(sync_result_dev_t),
                                  cudaMemcpyDeviceToHost, c->stream));
        DNRP_CUDA(cudaStreamSynchronize(c->stream));
        uint32_t active = 0;
        for (uint32_t k = 0; k < n_chunks; ++k) {
            if (resume[k] == 0xffffffffu) continue;
            const sync_result_dev_t& r = res[k];
            if (dbg) std::fprintf(stderr, "round %u chunk %u resume %u: det %u t_det %u coarse %u fine %u ndet %u\n", round, k,
                                  resume[k], r.detected, r.detection_time_local, r.coarse_peak_time_local, r.fine_peak_time_local,
                                  r.n_detections);
            if (!r.detected) {
                // detection range used up (0), eight invalid peaks in a row (go on behind the last one), or a coarse
                // peak whose fine search range left the chunk (skip it like a packet)
                uint32_t next = r.coarse_peak_time_local ? r.coarse_peak_time_local + skip : r.detection_time_local;
                if (next != 0 && next <= resume[k]) next = resume[k] + step;   // always advance
                if (next == 0 || next >= lead + hop) resume[k] = 0xffffffffu;
                else { resume[k] = next; ++active; }
                continue;
            }
            const int64_t base = (int64_t)k * hop;
            if ((uint64_t)(base + r.fine_peak_time_local) + stf <= n_samples) {
                dnrp_stream_packet_t f{};
                f.fine_peak_time_64 = base + r.fine_peak_time_local;
                f.coarse_peak_time_64 = base + r.coarse_peak_time_local;
                static_assert(sizeof(f.sync) == sizeof(r), "sync result layouts differ");
                std::memcpy(&f.sync, &r, sizeof(r));
                f.chunk = k; f.round = round;
                found.push_back(f);
            }
            const uint32_t next = r.coarse_peak_time_local + skip;     // skip_after_peak
            if (next >= lead + hop) resume[k] = 0xffffffffu;
            else { resume[k] = next; ++active; }
        }
        if (!active) break;
    }
    // ---- time order, unique packets only (baton_t::is_sync_time_unique: later than the predecessor by more than one
    // STF pattern)
    std::stable_sort(found.begin(), found.end(), [](const dnrp_stream_packet_t& a, const dnrp_stream_packet_t& b) {
        return a.fine_peak_time_64 < b.fine_peak_time_64;
    });
    int64_t last = INT64_MIN / 2;
    uint32_t n = 0;
    for (const auto& f : found) {
        if (f.fine_peak_time_64 - last <= (int64_t)pat) continue;
        last = f.fine_peak_time_64;
        if (n < max_out) out[n] = f;
        ++n;
    }
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
    DNRP_REQUIRE(c->d_stream.p && c->stream_len, "no stream on the device (dnrp_stream_sync first)");
    DNRP_REQUIRE(!c->resampling, "the streaming front door needs a context without resampling (L = M)");
    DNRP_REQUIRE(t0 + c->n_no_gi_rs <= c->rx_stride, "t0 leaves no room for the packet inside the window");
    DNRP_CUDA(cudaSetDevice(c->device));
    static_assert(sizeof(long long) == sizeof(int64_t), "64-bit times");
    dbuf<long long> d_start;
    DNRP_CUDA(d_start.alloc(n));
    DNRP_CUDA(cudaMemcpyAsync(d_start.p, start_64, n * sizeof(int64_t), cudaMemcpyHostToDevice, c->stream));
    const phy_dims_t& d = c->dims;
    const uint32_t bx = std::min<uint32_t>((c->rx_stride + 255) / 256, 16);
    ++c->launches;
    k_stream_extract<<<dim3(bx, d.N_RX, n), 256, 0, c->stream>>>(c->d_stream.p, c->stream_alloc, (long long)c->stream_len,
                                                                  d_start.p, t0, c->d_iq_rx.p + c->off * d.N_RX * c->rx_stride,
                                                                  c->rx_stride, d.N_RX);
    DNRP_CUDA(cudaGetLastError());
    DNRP_CUDA(cudaStreamSynchronize(c->stream));
    return DNRP_OK;
}

}  // extern "C"

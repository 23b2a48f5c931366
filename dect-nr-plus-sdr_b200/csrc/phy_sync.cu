// phy_sync.cu -- packet synchronisation on the GPU for a batch of IQ windows (one search per window):
//   k_sync_steps   per-step autocorrelation and power sums          (autocorrelator_detection.cpp:152-196)
//   k_sync_peak    detection gates, per-sample coarse peak search, validity, fractional CFO, rms
//                  (autocorrelator_detection.cpp:197-250, autocorrelator_peak.cpp:145-384)
//   k_sync_xcorr   CFO pre-corrected cross-correlation with every N_eff_TX template (crosscorrelator.cpp:89-249)
//   k_sync_final   N_eff_TX decision and fine peak, hand-over to the synchronised receiver
// The reference streams samples through moving sums that pop and push; here every candidate position is
// evaluated directly from prefix sums (same sums, different association order).  Constants:
// lib/include/dectnrp/phy/rx/sync/sync_param.hpp as shipped (cover sequence active).
#include "fft.cuh"
#include "phy.cuh"

namespace dnrp {

namespace {
__device__ __constant__ float c_cover[9] = {1.f, -1.f, 1.f, 1.f, -1.f, -1.f, -1.f, -1.f, -1.f};

__device__ __forceinline__ float warp_sum(float v) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// exclusive offsets of the per-thread partial sums v (one per thread, blockDim.x <= 1024): warp shuffle scan, warp
// totals through s_part[32 * slot ...]; returns the sum of the partials of all lower threads.  All threads call.
template <typename F>
__device__ __forceinline__ F block_exclusive(F v, F* s_part) {
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    F inc = v;
    for (int o = 1; o < 32; o <<= 1) {
        const F u = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= (uint32_t)o) inc += u;
    }
    if (lane == 31) s_part[w] = inc;
    __syncthreads();
    F off = inc - v;
    for (uint32_t i = 0; i < nw; ++i) if (i < w) off += s_part[i];
    __syncthreads();
    return off;
}

// inclusive block-wide prefix sum of a double array in shared memory, in place (n arbitrary)
__device__ void block_scan(double* a, uint32_t n, double* s_part) {
    const uint32_t tid = threadIdx.x, nthr = blockDim.x;
    const uint32_t per = (n + nthr - 1) / nthr, lo = min(tid * per, n), hi = min(lo + per, n);
    double s = 0.0;
    for (uint32_t i = lo; i < hi; ++i) { s += a[i]; a[i] = s; }
    const double off = block_exclusive(s, s_part);
    for (uint32_t i = lo; i < hi; ++i) a[i] += off;
    __syncthreads();
}
}  // namespace

// block-wide sum of a float2 (result in every thread); s_red: 8 entries
__device__ float2 block_sum2f(float2 v, float2* s_red) {
    v.x = warp_sum(v.x); v.y = warp_sum(v.y);
    const uint32_t w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    __syncthreads();
    if ((threadIdx.x & 31) == 0) s_red[w] = v;
    __syncthreads();
    float2 t = make_float2(0.f, 0.f);
    for (uint32_t i = 0; i < nw; ++i) { t.x += s_red[i].x; t.y += s_red[i].y; }
    return t;
}

// block-wide sum of three floats (result in every thread)
__device__ float3 block_sum3(float3 v, float3* s_red) {
    v.x = warp_sum(v.x); v.y = warp_sum(v.y); v.z = warp_sum(v.z);
    const uint32_t w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    __syncthreads();
    if ((threadIdx.x & 31) == 0) s_red[w] = v;
    __syncthreads();
    float3 t = make_float3(0.f, 0.f, 0.f);
    for (uint32_t i = 0; i < nw; ++i) { t.x += s_red[i].x; t.y += s_red[i].y; t.z += s_red[i].z; }
    return t;
}

// ---- one block per packet: detection -> coarse peak -> validity -> fractional CFO and rms
__global__ void __launch_bounds__(256) k_sync_peak(const sync_params_t p, uint32_t len) {
    extern __shared__ double smem_d[];
    __shared__ double s_part[384];  // also used as float [3][256]
    __shared__ float3 s_red[8];
    __shared__ uint32_t s_min;
    __shared__ unsigned long long s_best;
    const phy_dims_t& d = p.d;
    const uint32_t pkt = sync_chunk(p, blockIdx.x), tid = threadIdx.x, nthr = blockDim.x;
    const uint32_t n_pat = d.n_pat, pat = d.pat_len_os, stf_len = d.stf_len_os, step = p.step;
    const uint32_t W = 4 * n_pat, Wc = 4 * (n_pat - 1);
    const float prefactor = (float)n_pat / (float)(n_pat - 1);
    const uint32_t bos = pat / 16;  // b * os
    sync_result_dev_t& res = p.result[pkt];
    float uw[8];
    for (uint32_t i = 0; i + 1 < n_pat; ++i) uw[i] = c_cover[i] * c_cover[i + 1];

    const uint32_t search_length = p.limit ? p.limit[pkt] : p.search_length;
    uint32_t j_from = (stf_len + pat + step - 1) / step - 1;  // first step whose end reaches ignore_before_index
    if (p.resume) {  // continuation inside a chunk: autocorrelator_detection_t::skip_after_peak moved ignore_before_index
        const uint32_t r = p.resume[pkt];
        if (r == 0xffffffffu) {
            if (tid == 0) { res.detected = 0; res.n_detections = 0; }
            return;
        }
        const uint32_t jr = r ? (r + step - 1) / step - 1 : 0;
        if (jr > j_from) j_from = jr;
    }
    uint32_t n_det_total = 0;
    uint32_t attempt = 0;
    for (; attempt < 8; ++attempt) {
        // ---- detection: first (step, antenna) passing the gates
        uint32_t best_j = 0xffffffffu, best_a = 0;
        float best_metric = 0.f, best_rms = 0.f;
        for (uint32_t a = 0; a < p.n_ant; ++a) {
            double* Pp = smem_d;                    // [n_steps + 1]
            double* Pr = Pp + p.n_steps + 1;
            double* Pi = Pr + p.n_steps + 1;
            const size_t o = ((size_t)pkt * p.n_ant + a) * p.n_steps;
            const uint32_t n_use = p.n_steps_use ? p.n_steps_use : p.n_steps;   // steps that start inside search_length
            for (uint32_t j = tid; j < n_use; j += nthr) {
                Pp[j + 1] = p.p_steps[o + j];
                Pr[j + 1] = p.c_steps[o + j].x;
                Pi[j + 1] = p.c_steps[o + j].y;
            }
            if (tid == 0) { Pp[0] = Pr[0] = Pi[0] = 0.0; s_min = 0xffffffffu; }
            __syncthreads();
            block_scan(Pp + 1, n_use, s_part);
            block_scan(Pr + 1, n_use, s_part);
            block_scan(Pi + 1, n_use, s_part);
            for (uint32_t j = j_from + tid; j < n_use; j += nthr) {
                const uint32_t n = (j + 1) * step;
                if (n - step >= search_length || n > len || j + 1 < W) continue;
                const float power = (float)(Pp[j + 1] - Pp[j + 1 - W]);
                const float rms = sqrtf(power / (float)stf_len);
                if (rms < p.rms_min || 2.0f < rms) continue;
                // movsum_t::get_sum_back(2) / get_sum_front(2) as implemented by the reference (movsum.hpp:55-70):
                // "front" = oldest + newest step, "back" = the two steps after the oldest
                const float front = (float)(Pp[j + 2 - W] - Pp[j + 1 - W]) + (float)(Pp[j + 1] - Pp[j]);
                const float back = (float)(Pp[j + 4 - W] - Pp[j + 2 - W]);
                if (sqrtf(back) * 0.5 >= sqrtf(front)) continue;
                float cr = 0.f, ci = 0.f;
                for (uint32_t i = 0; i + 1 < n_pat; ++i) {
                    const uint32_t lo = j + 1 - Wc + 4 * i;
                    cr += uw[i] * (float)(Pr[lo + 4] - Pr[lo]);
                    ci += uw[i] * (float)(Pi[lo + 4] - Pi[lo]);
                }
                const float t = prefactor * hypotf(cr, ci) / power;
                const float metric = t * t;
                if (metric < 0.18f || 1.50f < metric || !(0.18f < metric)) continue;
                atomicMin(&s_min, j);
            }
            __syncthreads();
            const uint32_t ja = s_min;
            if (ja < best_j) {
                best_j = ja; best_a = a;
                const float power = (float)(Pp[ja + 1] - Pp[ja + 1 - W]);
                float cr = 0.f, ci = 0.f;
                for (uint32_t i = 0; i + 1 < n_pat; ++i) {
                    const uint32_t lo = ja + 1 - Wc + 4 * i;
                    cr += uw[i] * (float)(Pr[lo + 4] - Pr[lo]);
                    ci += uw[i] * (float)(Pi[lo + 4] - Pi[lo]);
                }
                const float t = prefactor * hypotf(cr, ci) / power;
                best_metric = t * t;
                best_rms = sqrtf(power / (float)stf_len);
            }
            __syncthreads();
        }
        if (best_j == 0xffffffffu) break;  // nothing (more) in this window
        ++n_det_total;
        const uint32_t n_det = (best_j + 1) * step, r0 = n_det - pat;
        if (r0 + stf_len > len) break;

        // ---- coarse peak: per-sample metric over one STF after (detection - 1 pattern), smoothed, arg max
        const uint32_t A0 = r0 - stf_len, S2 = 2 * stf_len;
        float* Qr = reinterpret_cast<float*>(smem_d);   // prefix of q[m] = r[m - pat] conj(r[m]), [S2 + 1]
        float* Qi = Qr + S2 + 1;
        float* PW = Qi + S2 + 1;                        // prefix of abs(r[m])^2
        float* M = PW + S2 + 1;                         // metric [stf_len]
        float peak_metric[8];
        uint32_t peak_index[8];
        const uint32_t sm_right = bos, sm_len = 2 * bos + 1;
        for (uint32_t a = 0; a < p.n_ant; ++a) {
            const float2* x = sync_iq(p, pkt, a) + A0;
            // serial prefix per thread chunk + chunk offsets (three arrays at once)
            const uint32_t per = (S2 + nthr - 1) / nthr, lo = min(tid * per, S2), hi = min(lo + per, S2);
            float sr = 0.f, si = 0.f, sp = 0.f;
            auto acc1 = [&](uint32_t i, float2 b, float2 e) {   // one sample, in index order (the sums are order sensitive)
                sp += b.x * b.x + b.y * b.y;
                if (i >= pat) {
                    sr += e.x * b.x + e.y * b.y;
                    si += e.y * b.x - e.x * b.y;
                }
                Qr[i + 1] = sr; Qi[i + 1] = si; PW[i + 1] = sp;
            };
            // a thread walks `per` consecutive samples, so a warp's load touches 32 different lines whatever its width:
            // two samples per 16-byte load halve the number of such loads (same values, same order of additions)
            const bool vec = ((per | lo | hi | pat) & 1u) == 0u && ((reinterpret_cast<uintptr_t>(x) & 15u) == 0u);
            if (vec) {
                const float4* x4 = reinterpret_cast<const float4*>(x);
                for (uint32_t i = lo; i < hi; i += 2) {
                    const float4 b = x4[i >> 1];
                    const float4 e = i >= pat ? x4[(i - pat) >> 1] : make_float4(0.f, 0.f, 0.f, 0.f);
                    acc1(i, make_float2(b.x, b.y), make_float2(e.x, e.y));
                    acc1(i + 1, make_float2(b.z, b.w), make_float2(e.z, e.w));
                }
            } else {
                for (uint32_t i = lo; i < hi; ++i) acc1(i, x[i], i >= pat ? x[i - pat] : make_float2(0.f, 0.f));
            }
            float* part = reinterpret_cast<float*>(s_part);
            if (tid == 0) { Qr[0] = Qi[0] = PW[0] = 0.f; }
            const float o_r = block_exclusive(sr, part), o_i = block_exclusive(si, part + 32), o_p = block_exclusive(sp, part + 64);
            for (uint32_t i = lo; i < hi; ++i) { Qr[i + 1] += o_r; Qi[i + 1] += o_i; PW[i + 1] += o_p; }
            __syncthreads();
            for (uint32_t k = tid; k < stf_len; k += nthr) {
                const uint32_t e = stf_len + k + 1;  // prefix index just after sample n = r0 + k
                const float power = PW[e] - PW[e - stf_len];
                float cr = 0.f, ci = 0.f;
                for (uint32_t i = 0; i + 1 < n_pat; ++i) {
                    const uint32_t hi2 = e - (n_pat - 2 - i) * pat, lo2 = hi2 - pat;
                    cr += uw[i] * (Qr[hi2] - Qr[lo2]);
                    ci += uw[i] * (Qi[hi2] - Qi[lo2]);
                }
                const float t = prefactor * hypotf(cr, ci) / power;
                M[k] = t * t;
            }
            if (tid == 0) s_best = 0ull;
            __syncthreads();
            unsigned long long mine = 0ull;
            auto consider = [&](uint32_t k, float s) {
                const float mean = s / (float)sm_len;
                // larger mean wins, first position on ties: key = (float bits, inverted position)
                const unsigned long long key = ((unsigned long long)__float_as_uint(mean) << 32) | (0xffffffffu - k);
                if (mean > 0.f && key > mine) mine = key;
            };
            constexpr uint32_t SMK = 7, SML = 33;   // beta x oversampling = 16 with 256 threads: 7 positions per thread
            if (sm_len == SML && stf_len == SMK * nthr) {
                // a thread takes SMK CONSECUTIVE positions: their windows overlap, the SML + SMK - 1 metrics are read once
                // into registers; every window is still summed from its oldest element upwards (same rounding)
                const uint32_t kf = tid * SMK;
                if (kf + 1 >= SML) {
                    float m[SML + SMK - 1];
#pragma unroll
                    for (uint32_t q = 0; q < SML + SMK - 1; ++q) m[q] = M[kf + 1 - SML + q];
#pragma unroll
                    for (uint32_t i = 0; i < SMK; ++i) {
                        float s = 0.f;
#pragma unroll
                        for (uint32_t q = 0; q < SML; ++q) s += m[i + q];
                        consider(kf + i, s);
                    }
                } else {
                    for (uint32_t k = kf; k < kf + SMK; ++k) {
                        float s = 0.f;
                        const uint32_t k0 = k + 1 >= sm_len ? k + 1 - sm_len : 0;
                        for (uint32_t q = k0; q <= k; ++q) s += M[q];
                        consider(k, s);
                    }
                }
            } else
            for (uint32_t k = tid; k < stf_len; k += nthr) {
                float s = 0.f;
                const uint32_t k0 = k + 1 >= sm_len ? k + 1 - sm_len : 0;
                for (uint32_t q = k0; q <= k; ++q) s += M[q];
                consider(k, s);
            }
            atomicMax(&s_best, mine);
            __syncthreads();
            const unsigned long long bk = s_best;
            peak_metric[a] = __uint_as_float((uint32_t)(bk >> 32));
            peak_index[a] = bk ? r0 + (0xffffffffu - (uint32_t)bk) - sm_right : 0u;
            __syncthreads();
        }
        // ---- validity (autocorrelator_peak.cpp:286-326)
        float widx = 0.f, msum = 0.f, cpa[8];
        uint32_t valid = 0;
        for (uint32_t a = 0; a < 8; ++a) cpa[a] = 0.f;
        for (uint32_t a = 0; a < p.n_ant; ++a) {
            if (best_metric + (-0.25f) >= peak_metric[a]) continue;
            if ((long long)n_det + (long long)(-0.3 * (double)stf_len) >= (long long)peak_index[a]) continue;
            cpa[a] = peak_metric[a];
            widx += peak_metric[a] * (float)peak_index[a];
            msum += peak_metric[a];
            ++valid;
        }
        if (!valid) { j_from = best_j + 1; continue; }
        const uint32_t coarse = (uint32_t)roundf(widx / msum) - (stf_len - 1);
        // ---- rms and fractional CFO over the STF at the coarse peak (autocorrelator_peak.cpp:328-360)
        float wcfo = 0.f, ms2 = 0.f, rms_arr[8];
        uint32_t strongest = 0;
        for (uint32_t a = 0; a < 8; ++a) rms_arr[a] = 0.f;
        for (uint32_t a = 0; a < p.n_ant; ++a) {
            if (cpa[a] <= 0.f) continue;
            if (cpa[a] > cpa[strongest]) strongest = a;
            const float2* x = sync_iq(p, pkt, a) + coarse;
            float3 acc = make_float3(0.f, 0.f, 0.f);
            for (uint32_t i = tid; i < stf_len; i += nthr) {
                const float2 b = x[i];
                acc.z += b.x * b.x + b.y * b.y;
                if (i + pat < stf_len) {
                    const float2 e = x[i + pat];
                    const float w = uw[i / pat];
                    acc.x += w * (b.x * e.x + b.y * e.y);
                    acc.y += w * (b.y * e.x - b.x * e.y);
                }
            }
            acc = block_sum3(acc, s_red);
            rms_arr[a] = sqrtf(acc.z / (float)stf_len);
            wcfo += peak_metric[a] * atan2f(acc.y, acc.x) / (float)pat;
            ms2 += peak_metric[a];
        }
        if (tid == 0) {
            res.detected = 1;
            res.detection_time_local = n_det; res.detection_ant_idx = best_a;
            res.detection_metric = best_metric; res.detection_rms = best_rms;
            res.coarse_peak_time_local = coarse;
            // rx_pacer_t::convert_length_resampled_to_global (rx_pacer.cpp:311-318)
            res.coarse_peak_time_hw = (uint32_t)round((double)coarse * (double)p.rs_L / (double)p.rs_M);
            for (uint32_t a = 0; a < 8; ++a) { res.coarse_peak_array[a] = cpa[a]; res.rms_array[a] = rms_arr[a]; }
            res.cfo_fractional_rad = wcfo / ms2;
            res.cfo_integer_rad = 0.f;  // coarse_peak_f_domain.cpp:94-201: not implemented in the reference
            res.N_eff_TX = 0; res.fine_peak_time_local = 0;
            res.n_detections = n_det_total; res.strongest_ant = strongest;
        }
        return;
    }
    if (tid == 0) {
        res.detected = 0;
        res.n_detections = n_det_total;
        // streaming search: all attempts ended in invalid peaks, the chunk is not used up -- where to go on
        res.detection_time_local = (p.resume && attempt == 8) ? (j_from + 1) * step : 0;
        res.N_eff_TX = 0; res.fine_peak_time_local = 0; res.coarse_peak_time_local = 0; res.coarse_peak_time_hw = 0;
        res.cfo_fractional_rad = 0.f; res.cfo_integer_rad = 0.f; res.strongest_ant = 0;
        for (uint32_t a = 0; a < 8; ++a) { res.coarse_peak_array[a] = 0.f; res.rms_array[a] = 0.f; }
    }
}

// ---- frequency-domain stage between coarse and fine peak (reference coarse_peak_f_domain_t::process,
// coarse_peak_f_domain.cpp:70-201; both parts are compiled out there, sync_param.hpp:238-252).  One block per window: the
// N_os samples behind the STF's cyclic prefix at the coarse peak, fractional CFO removed, cover sequence undone, are
// transformed per antenna (the reference leaves the CFO correction as a to-do and does not undo the cover sequence; the
// power comparison below does not depend on either).
//  * beta search (estimate_beta, :94-190): power per subcarrier summed over the antennas; starting from b = 1 the next
//    larger b of {1, 2, 4, 8, 12, 16} is accepted while the mean power of the two side bands it adds exceeds
//    threshold x the mean power of the band found so far.  The side bands are taken symmetrically (the reference
//    advances its right-hand index by the NEW band width instead of the old one, :150-151: a slip in disabled code).
//  * integer CFO (estimate_integer_cfo is "#error not implemented", :193-200; own specification): the
//    fractional estimate of the pattern autocorrelation is unambiguous within +-2 subcarriers, what is left is a multiple
//    of 4.  For every shift 4 s inside the search range and every cyclic shift of the STF sequence (N_eff_TX = 1, 2, 4,
//    8 <= the context's) the differential correlation D = sum_a sum_i X_a[c_i + 4 s] conj(X_a[c_(i+1) + 4 s]) y_i y_(i+1)
//    over neighbouring STF cells (c_i = multiples of 4 without 0) is formed; timing offset and a smooth channel turn all
//    products alike, so |D| peaks at the true shift.  fd_cfo = 4 s (where the signal sits, in subcarriers),
//    cfo_integer_rad = -2 pi 4 s / N_os per sample (a correction increment like cfo_fractional_rad).
constexpr uint32_t FD_THREADS = 256;

__device__ __forceinline__ uint32_t fd_b_idx(uint32_t b) { return b == 1 ? 0u : (b == 2 ? 1u : (b == 4 ? 2u : (b == 8 ? 3u : (b == 12 ? 4u : 5u)))); }

__global__ void __launch_bounds__(FD_THREADS) k_sync_fdomain(const sync_params_t p, uint32_t len) {
    extern __shared__ __align__(16) float2 smem_fd[];
    __shared__ float2 s_red2[8];
    __shared__ float2 s_hyp[4 * 17];     // [cyclic shift][s + S], S <= 8
    __shared__ uint32_t s_b;
    const phy_dims_t& d = p.d;
    const uint32_t pkt = sync_chunk(p, blockIdx.x), tid = threadIdx.x, N = d.N_os;
    sync_result_dev_t& res = p.result[pkt];
    const uint32_t start = res.coarse_peak_time_local + d.stf_cp_os;
    if (!res.detected || start + N > len) {
        if (tid == 0) { p.fd_b[pkt] = d.b; p.fd_cfo[pkt] = 0; }
        return;
    }
    float2* A = smem_fd;
    float2* Bf = smem_fd + fft_buf_len(N);
    float* P = reinterpret_cast<float*>(smem_fd + 2 * fft_buf_len(N));   // [N] power per bin over the antennas
    const uint32_t os = N / (64u * d.b);       // oversampling of the context
    const uint32_t S = p.fd_cfo_range / 4 > 8 ? 8u : p.fd_cfo_range / 4;
    uint32_t n_lg = 0;
    for (uint32_t n = d.N_eff_TX; n >= 1; n >>= 1) ++n_lg;     // cyclic shifts 0, 2, 4, 6 for N_eff_TX = 1, 2, 4, 8
    for (uint32_t i = tid; i < N; i += FD_THREADS) P[i] = 0.f;
    for (uint32_t i = tid; i < 4 * 17; i += FD_THREADS) s_hyp[i] = make_float2(0.f, 0.f);
    uint32_t b_est = d.b;
    for (uint32_t a = 0; a < p.n_ant; ++a) {
        const float2* x = sync_iq(p, pkt, a) + start;
        __syncthreads();
        for (uint32_t n = tid; n < N; n += FD_THREADS) {
            double sd, cd;
            sincos((double)res.cfo_fractional_rad * (double)n, &sd, &cd);   // the correction increment, as in the receiver
            // patterns 3 .. 6 of the STF (the cyclic prefix holds 0 .. 2): cover sequence 1 -1 -1 -1 undone
            const float cv = n < N / 4 ? 1.f : -1.f;
            A[n] = cmul(x[n], make_float2((float)cd * cv, (float)sd * cv));
        }
        const float2* X = block_fft<false, false>(A, Bf, N, d.twiddle, d.twiddle_p4);
        for (uint32_t i = tid; i < N; i += FD_THREADS) P[i] += X[i].x * X[i].x + X[i].y * X[i].y;
        __syncthreads();
        if (a + 1 == p.n_ant && tid == 0) {
            // ---- beta search on the summed power (bin of subcarrier c: (c + N) % N)
            uint32_t be = d.b;
            if (p.fd_beta_thr > 0.f && d.b > 1) {
                auto band = [&](int lo, int hi) {   // power of the subcarriers lo .. hi - 1 and -hi .. -lo - 1 (the reference's two arrays)
                    float s = 0.f;
                    for (int c = lo; c < hi; ++c) s += P[(uint32_t)(c + (int)N) % N] + P[(uint32_t)(-c - 1 + (int)N) % N];
                    return s;
                };
                const uint32_t tab[6] = {1, 2, 4, 8, 12, 16};
                be = 1;
                float central = band(0, 32);
                for (uint32_t bi = 1; bi < 6 && tab[bi] <= d.b; ++bi) {
                    const uint32_t bt = tab[bi];
                    const float side = band(32 * (int)be, 32 * (int)bt);
                    const float cn = central / (float)(be * 64u), sn = side / (float)((bt - be) * 64u);
                    if (sn / cn > p.fd_beta_thr) { be = bt; central += side; }
                    else break;
                }
            }
            s_b = be;
        }
        if (a + 1 == p.n_ant) { __syncthreads(); b_est = s_b; }
        if (S == 0) continue;
        // ---- differential correlation of this antenna for every hypothesis; the sequence and the comb belong to beta:
        // the context's unless the beta search is on (then the estimate is only known after the last antenna, and the
        // antennas are revisited below)
        if (p.fd_beta_thr > 0.f && d.b > 1) continue;
        const uint32_t NS = 14 * b_est;
        const int8_t* y = p.fd_seq + fd_b_idx(b_est) * 224;
        for (uint32_t lg = 0; lg < n_lg; ++lg)
            for (int s = -(int)S; s <= (int)S; ++s) {
                float2 acc = make_float2(0.f, 0.f);
                for (uint32_t i = tid; i + 1 < NS; i += FD_THREADS) {
                    if (i + 1 == NS / 2) continue;   // the pair across DC is 8 subcarriers apart
                    const int c0 = (i < NS / 2 ? -(int)(2 * NS) + 4 * (int)i : 4 * (int)(i - NS / 2) + 4) + 4 * s;
                    const float sg = (float)(y[(i + 2 * lg) % NS] * y[(i + 1 + 2 * lg) % NS]);
                    const float2 v = cmulc(X[(uint32_t)(c0 + (int)N) % N], X[(uint32_t)(c0 + 4 + (int)N) % N]);
                    acc.x += sg * v.x; acc.y += sg * v.y;
                }
                acc = block_sum2f(acc, s_red2);
                if (tid == 0) { float2& h = s_hyp[lg * 17 + (uint32_t)(s + (int)S)]; h.x += acc.x; h.y += acc.y; }
            }
    }
    (void)os;
    if (S != 0 && p.fd_beta_thr > 0.f && d.b > 1) {
        // beta search on: second sweep over the antennas with the comb and sequence of the estimated beta
        const uint32_t NS = 14 * b_est;
        const int8_t* y = p.fd_seq + fd_b_idx(b_est) * 224;
        for (uint32_t a = 0; a < p.n_ant; ++a) {
            const float2* x = sync_iq(p, pkt, a) + start;
            __syncthreads();
            for (uint32_t n = tid; n < N; n += FD_THREADS) {
                double sd, cd;
                sincos((double)res.cfo_fractional_rad * (double)n, &sd, &cd);
                const float cv = n < N / 4 ? 1.f : -1.f;
                A[n] = cmul(x[n], make_float2((float)cd * cv, (float)sd * cv));
            }
            const float2* X = block_fft<false, false>(A, Bf, N, d.twiddle, d.twiddle_p4);
            for (uint32_t lg = 0; lg < n_lg; ++lg)
                for (int s = -(int)S; s <= (int)S; ++s) {
                    float2 acc = make_float2(0.f, 0.f);
                    for (uint32_t i = tid; i + 1 < NS; i += FD_THREADS) {
                        if (i + 1 == NS / 2) continue;
                        const int c0 = (i < NS / 2 ? -(int)(2 * NS) + 4 * (int)i : 4 * (int)(i - NS / 2) + 4) + 4 * s;
                        const float sg = (float)(y[(i + 2 * lg) % NS] * y[(i + 1 + 2 * lg) % NS]);
                        const float2 v = cmulc(X[(uint32_t)(c0 + (int)N) % N], X[(uint32_t)(c0 + 4 + (int)N) % N]);
                        acc.x += sg * v.x; acc.y += sg * v.y;
                    }
                    acc = block_sum2f(acc, s_red2);
                    if (tid == 0) { float2& h = s_hyp[lg * 17 + (uint32_t)(s + (int)S)]; h.x += acc.x; h.y += acc.y; }
                }
        }
    }
    __syncthreads();
    if (tid == 0) {
        int s_best = 0;
        if (S != 0) {
            float best = -1.f;
            for (uint32_t lg = 0; lg < n_lg; ++lg)
                for (int s = -(int)S; s <= (int)S; ++s) {
                    const float2 h = s_hyp[lg * 17 + (uint32_t)(s + (int)S)];
                    const float m = h.x * h.x + h.y * h.y;
                    if (m > best) { best = m; s_best = s; }
                }
        }
        p.fd_b[pkt] = b_est;
        p.fd_cfo[pkt] = 4 * s_best;
        // the signal sits 4 s subcarriers up: the correction increment (same sense as cfo_fractional_rad) is negative
        res.cfo_integer_rad = -6.283185307179586f * (float)(4 * s_best) / (float)N;
    }
}

// ---- eight lanes per detection step, four steps per warp: c_j = sum r[m - pat] conj(r[m]), p_j = sum abs(r[m])^2 over
// the step.  All 16-byte loads of a lane are independent (issued back to back), the reduction runs over 8 lanes.
__global__ void __launch_bounds__(256) k_sync_steps(const sync_params_t p) {
    const uint32_t lane = threadIdx.x & 31, sub = lane & 7;
    const uint32_t j = (blockIdx.x * 8 + (threadIdx.x >> 5)) * 4 + (lane >> 3);
    const uint32_t a = blockIdx.y, pkt = blockIdx.z;
    const bool valid = j < (p.n_steps_use ? p.n_steps_use : p.n_steps);
    const float2* x = sync_iq(p, pkt, a);
    const uint32_t pat = p.d.pat_len_os, m0 = valid ? j * p.step : 0;
    float cr = 0.f, ci = 0.f, pw = 0.f;
    // two samples per 16-byte load (steps, pattern length and window strides are even; windows are 16-byte aligned)
    const float4* xb = reinterpret_cast<const float4*>(x + m0);
    const float4* xe = reinterpret_cast<const float4*>(x + m0 - (m0 >= pat ? pat : 0));
    const bool corr = m0 >= pat;
    const uint32_t n4 = valid ? p.step / 2 : 0;
#pragma unroll 4
    for (uint32_t i = sub; i < n4; i += 8) {
        const float4 b = xb[i];
        const float4 e = corr ? xe[i] : make_float4(0.f, 0.f, 0.f, 0.f);
        pw += b.x * b.x + b.y * b.y + b.z * b.z + b.w * b.w;
        cr += e.x * b.x + e.y * b.y + e.z * b.z + e.w * b.w;
        ci += e.y * b.x - e.x * b.y + e.w * b.z - e.z * b.w;
    }
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) {
        cr += __shfl_xor_sync(0xffffffffu, cr, o);
        ci += __shfl_xor_sync(0xffffffffu, ci, o);
        pw += __shfl_xor_sync(0xffffffffu, pw, o);
    }
    if (sub == 0 && valid) {
        const size_t o = ((size_t)pkt * p.n_ant + a) * p.n_steps + j;
        p.c_steps[o] = make_float2(cr, ci);
        p.p_steps[o] = pw;
    }
}

// ---- fine peak: cross-correlation of the CFO pre-corrected window on the strongest antenna with one STF template.
// grid (template, packet).  A thread owns XC_LAGS consecutive lags and a slice of the template; the window
// slides through registers, so eight complex MACs cost two shared loads.  The window is stored with one pad
// element per eight (threads of a warp start eight elements apart).
constexpr uint32_t XC_LAGS = 8, XC_THREADS = 512;
__device__ __forceinline__ uint32_t xc_pos(uint32_t e) { return e + (e >> 3); }

__global__ void __launch_bounds__(XC_THREADS) k_sync_xcorr(const sync_params_t p, uint32_t len, uint32_t IC) {
    extern __shared__ float2 smem_x[];
    __shared__ unsigned long long s_best;
    const phy_dims_t& d = p.d;
    const uint32_t tpl = blockIdx.x, pkt = sync_chunk(p, blockIdx.y), tid = threadIdx.x;
    const uint32_t S = p.xc_S, sl = p.xc_sl, n_lag = 2 * sl + 1;  // 16 b os L/M samples to either side
    const uint32_t LT = (n_lag + XC_LAGS - 1) / XC_LAGS, n_lag_pad = LT * XC_LAGS;
    const sync_result_dev_t& res = p.result[pkt];
    float2* out = p.xc + (size_t)pkt * p.n_tpl + tpl;
    const long long w0 = (long long)res.coarse_peak_time_hw - sl;
    if (!res.detected || w0 < 0 || w0 + n_lag + S - 1 > len) {
        if (tid == 0) *out = make_float2(-1.f, 0.f);
        return;
    }
    const uint32_t S_pad = (S + 7) & ~7u;              // the MAC loop is unrolled by 8: template zero padded
    float2* mix = smem_x;                              // [xc_pos(n_lag_pad + S_pad)]
    float2* tp = mix + xc_pos(n_lag_pad + S_pad) + 1;  // [S_pad]
    float2* part = tp + S_pad;                         // [IC][n_lag_pad]
    const float2* x = sync_iq_x(p, pkt, res.strongest_ant) + w0;
    // the CFO was measured at the DECT rate; per hardware sample it is smaller by M / L (crosscorrelator.cpp:97-105)
    const double inc = (double)((res.cfo_fractional_rad + res.cfo_integer_rad) * (float)p.rs_M / (float)p.rs_L);
    const uint32_t n_mix = n_lag + S - 1;
    for (uint32_t i = tid; i < n_lag_pad + S_pad; i += XC_THREADS) {
        float2 r = make_float2(0.f, 0.f);
        if (i < n_mix) {
            const float2 v = x[i];
            double sd, cd;
            sincos(inc * (double)i, &sd, &cd);
            const float cs = (float)cd, sn = (float)sd;
            r = make_float2(v.x * cs - v.y * sn, v.x * sn + v.y * cs);
        }
        mix[xc_pos(i)] = r;
    }
    const float2* T = p.templates + (size_t)tpl * S;
    for (uint32_t i = tid; i < S_pad; i += XC_THREADS) tp[i] = i < S ? T[i] : make_float2(0.f, 0.f);
    if (tid == 0) s_best = 0ull;
    __syncthreads();
    if (tid < LT * IC) {
        const uint32_t lt = tid % LT, ic = tid / LT, j0 = lt * XC_LAGS;
        const uint32_t chunk = ((S_pad + IC - 1) / IC + 7) & ~7u;
        const uint32_t i0 = min(ic * chunk, S_pad), i1 = min(i0 + chunk, S_pad);
        float ar[XC_LAGS], ai[XC_LAGS];
        float2 w[XC_LAGS];
#pragma unroll
        for (uint32_t k = 0; k < XC_LAGS; ++k) { ar[k] = ai[k] = 0.f; }
#pragma unroll
        for (uint32_t k = 0; k + 1 < XC_LAGS; ++k) w[k] = mix[xc_pos(j0 + i0 + k)];
        for (uint32_t i = i0; i < i1; i += 8) {
#pragma unroll
            for (uint32_t u = 0; u < 8; ++u) {
                w[(u + 7) & 7] = mix[xc_pos(j0 + i + u + 7)];
                const float2 t = tp[i + u];
#pragma unroll
                for (uint32_t k = 0; k < XC_LAGS; ++k) {
                    const float2 m = w[(u + k) & 7];
                    ar[k] = fmaf(m.x, t.x, fmaf(m.y, t.y, ar[k]));
                    ai[k] = fmaf(m.y, t.x, fmaf(-m.x, t.y, ai[k]));
                }
            }
        }
#pragma unroll
        for (uint32_t k = 0; k < XC_LAGS; ++k) part[ic * n_lag_pad + j0 + k] = make_float2(ar[k], ai[k]);
    }
    __syncthreads();
    unsigned long long mine = 0ull;
    for (uint32_t j = tid; j < n_lag; j += XC_THREADS) {
        float re = 0.f, im = 0.f;
        for (uint32_t ic = 0; ic < IC; ++ic) { re += part[ic * n_lag_pad + j].x; im += part[ic * n_lag_pad + j].y; }
        const float m2 = re * re + im * im;
        const unsigned long long key = ((unsigned long long)__float_as_uint(m2) << 32) | (0xffffffffu - j);
        if (key > mine) mine = key;
    }
    atomicMax(&s_best, mine);
    __syncthreads();
    if (tid == 0) {
        const unsigned long long bk = s_best;
        *out = make_float2(sqrtf(__uint_as_float((uint32_t)(bk >> 32))), (float)(0xffffffffu - (uint32_t)bk));
    }
}

// ---- the same cross-correlation through the frequency domain: corr[j] = sum_i x[j + i] conj(t[i]) =
// IFFT(FFT(x) . conj(FFT(t)))[j] for the lags j < n_lag when the FFT size covers lags + template length - 1 (no wrap
// around).  One block per packet: one forward FFT of the CFO pre-rotated window, one inverse FFT per template.
// 5 N log2 N flops per FFT instead of 8 n_lag S per template (14 x fewer for beta = 16).  The reference notes that
// "frequency domain methods ... are not applicable" to its SIMD path (sync_param.hpp:283-285); on this hardware the
// shared-memory FFT is the cheaper way to the same numbers (fp32 rounding differs, the arg max does not).
constexpr uint32_t XF_THREADS = 256, XF_MAX_PER_THREAD = 16;

__global__ void __launch_bounds__(XF_THREADS, 3) k_sync_xcorr_fft(const sync_params_t p, uint32_t len) {
    extern __shared__ __align__(16) float2 smem_f[];
    __shared__ unsigned long long s_best;
    const phy_dims_t& d = p.d;
    const uint32_t pkt = sync_chunk(p, blockIdx.x), tid = threadIdx.x, N = p.xc_fft_n;
    const uint32_t S = p.xc_S, sl = p.xc_sl, n_lag = 2 * sl + 1, n_mix = n_lag + S - 1;
    const sync_result_dev_t& res = p.result[pkt];
    const long long w0 = (long long)res.coarse_peak_time_hw - sl;
    if (!res.detected || w0 < 0 || w0 + n_mix > len) {
        for (uint32_t t = tid; t < p.n_tpl; t += XF_THREADS) p.xc[(size_t)pkt * p.n_tpl + t] = make_float2(-1.f, 0.f);
        return;
    }
    float2* A = smem_f;
    float2* B = smem_f + fft_buf_len(N);
    const float2* x = sync_iq_x(p, pkt, res.strongest_ant) + w0;
    const double inc = (double)((res.cfo_fractional_rad + res.cfo_integer_rad) * (float)p.rs_M / (float)p.rs_L);
    // CFO pre-rotation: one double-precision phasor per thread, then a float rotation by XF_THREADS samples per
    // iteration (at most 16 of them)
    double sd, cd;
    sincos(inc * (double)tid, &sd, &cd);
    float2 w = make_float2((float)cd, (float)sd);
    sincos(inc * (double)XF_THREADS, &sd, &cd);
    const float2 ws = make_float2((float)cd, (float)sd);
    for (uint32_t i = tid; i < N; i += XF_THREADS) {
        float2 r = make_float2(0.f, 0.f);
        if (i < n_mix) r = cmul(x[i], w);
        w = cmul(w, ws);
        A[i] = r;
    }
    const float2* X = block_fft<false, true>(A, B, N, p.xc_twiddle, p.xc_twiddle_p16);
    float2 Xr[XF_MAX_PER_THREAD];
#pragma unroll
    for (uint32_t r = 0; r < XF_MAX_PER_THREAD; ++r) {
        const uint32_t k = tid + r * XF_THREADS;
        Xr[r] = k < N ? X[k] : make_float2(0.f, 0.f);
    }
    for (uint32_t t = 0; t < p.n_tpl; ++t) {
        const float2* T = p.tpl_spec + (size_t)t * N;
        __syncthreads();
        if (tid == 0) s_best = 0ull;
#pragma unroll
        for (uint32_t r = 0; r < XF_MAX_PER_THREAD; ++r) {
            const uint32_t k = tid + r * XF_THREADS;
            if (k < N) A[k] = cmul(Xr[r], T[k]);
        }
        const float2* c = block_fft<true, true>(A, B, N, p.xc_twiddle, p.xc_twiddle_p16);
        unsigned long long mine = 0ull;
        for (uint32_t j = tid; j < n_lag; j += XF_THREADS) {
            const float2 v = c[j];
            const float m2 = v.x * v.x + v.y * v.y;
            const unsigned long long key = ((unsigned long long)__float_as_uint(m2) << 32) | (0xffffffffu - j);
            if (key > mine) mine = key;
        }
        atomicMax(&s_best, mine);
        __syncthreads();
        if (tid == 0) {
            const unsigned long long bk = s_best;
            p.xc[(size_t)pkt * p.n_tpl + t] =
                make_float2(sqrtf(__uint_as_float((uint32_t)(bk >> 32))), (float)(0xffffffffu - (uint32_t)bk));
        }
    }
}

// ---- N_eff_TX decision, fine peak, hand-over to the synchronised receiver (crosscorrelator.cpp:205-249)
__global__ void k_sync_final(const sync_params_t p, uint32_t n, uint32_t len) {
    const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n) return;
    const uint32_t pkt = sync_chunk(p, idx);
    sync_result_dev_t& res = p.result[pkt];
    const uint32_t sl = p.xc_sl;
    if (res.detected) {
        float best = 0.f;
        uint32_t best_tpl = 0, best_idx = 0;
        bool ok = true;
        for (uint32_t t = 0; t < p.n_tpl; ++t) {
            const float2 v = p.xc[(size_t)pkt * p.n_tpl + t];
            if (v.x < 0.f) ok = false;  // window left the buffer
            if (best < v.x) { best = v.x; best_tpl = t; best_idx = (uint32_t)v.y; }
        }
        if (ok) {
            res.N_eff_TX = 1u << best_tpl;
            res.fine_peak_time_local = res.coarse_peak_time_hw - sl + best_idx;
        } else {
            res.detected = 0;
        }
    }
    if (p.rx_state) {
        rx_state_t s = {};
        if (res.detected && (size_t)res.fine_peak_time_local + p.n_no_gi_hw <= p.x_stride) {
            s.t0 = (int32_t)res.fine_peak_time_local;
            s.inc0 = res.cfo_fractional_rad + res.cfo_integer_rad;
            for (uint32_t a = 0; a < 8; ++a) s.rms[a] = res.rms_array[a];
        } else {
            // nothing found, or the packet does not fit behind its fine peak: the receiver runs on the window start and
            // fails its CRCs; the record says "not detected" so that no caller mistakes offset 0 for a packet
            res.detected = 0;
            s.t0 = 0;
            for (uint32_t a = 0; a < 8; ++a) s.rms[a] = 1.f;
        }
        p.rx_state[pkt] = s;
    }
}

size_t sync_peak_smem(const phy_dims_t& d, uint32_t n_steps) {
    const size_t sm_det = 3 * (size_t)(n_steps + 1) * sizeof(double);
    const size_t sm_peak = (3 * (size_t)(2 * d.stf_len_os + 1) + d.stf_len_os) * sizeof(float);
    return sm_det > sm_peak ? sm_det : sm_peak;
}

cudaError_t launch_sync_steps(const sync_params_t& p, uint32_t n, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    const uint32_t n_use = p.n_steps_use ? p.n_steps_use : p.n_steps;
    k_sync_steps<<<dim3((n_use + 31) / 32, p.n_ant, n), 256, 0, st>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_sync(const sync_params_t& p, uint32_t n, cudaStream_t st) {
    cudaError_t e = launch_sync_steps(p, n, st);
    return e != cudaSuccess ? e : launch_sync_search(p, n, st);
}

cudaError_t launch_sync_search(const sync_params_t& p, uint32_t n, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    const phy_dims_t& d = p.d;
    const uint32_t len = p.len;
    const size_t sm2 = sync_peak_smem(d, p.n_steps);
    cudaError_t e = cudaFuncSetAttribute(k_sync_peak, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm2);
    if (e != cudaSuccess) return e;
    k_sync_peak<<<n, 256, sm2, st>>>(p, len);
    if ((p.fd_beta_thr > 0.f || p.fd_cfo_range >= 4) && p.fd_b && d.N_os <= 4096) {
        const size_t smfd = 2 * (size_t)fft_buf_len(d.N_os) * sizeof(float2) + (size_t)d.N_os * sizeof(float);
        e = cudaFuncSetAttribute(k_sync_fdomain, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smfd);
        if (e != cudaSuccess) return e;
        k_sync_fdomain<<<n, FD_THREADS, smfd, st>>>(p, len);
    }
    if (p.xc_fft_n) {
        const size_t smf = 2 * (size_t)fft_buf_len(p.xc_fft_n) * sizeof(float2);
        e = cudaFuncSetAttribute(k_sync_xcorr_fft, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smf);
        if (e != cudaSuccess) return e;
        k_sync_xcorr_fft<<<n, XF_THREADS, smf, st>>>(p, p.x_len);
        k_sync_final<<<(n + 127) / 128, 128, 0, st>>>(p, n, len);
        return cudaGetLastError();
    }
    const uint32_t n_lag = 2 * p.xc_sl + 1, LT = (n_lag + XC_LAGS - 1) / XC_LAGS, n_lag_pad = LT * XC_LAGS;
    uint32_t IC = XC_THREADS / LT;
    const uint32_t ic_max = p.xc_S / 32 ? p.xc_S / 32 : 1;
    if (IC > ic_max) IC = ic_max;
    if (IC == 0) return cudaErrorInvalidConfiguration;  // more than XC_THREADS * XC_LAGS lags
    const uint32_t S_pad = (p.xc_S + 7) & ~7u;
    const size_t sm3 = ((size_t)(n_lag_pad + S_pad) * 9 / 8 + 2 + S_pad + (size_t)IC * n_lag_pad) * sizeof(float2);
    e = cudaFuncSetAttribute(k_sync_xcorr, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm3);
    if (e != cudaSuccess) return e;
    k_sync_xcorr<<<dim3(p.n_tpl, n), XC_THREADS, sm3, st>>>(p, p.x_len, IC);
    k_sync_final<<<(n + 127) / 128, 128, 0, st>>>(p, n, len);
    return cudaGetLastError();
}

}  // namespace dnrp

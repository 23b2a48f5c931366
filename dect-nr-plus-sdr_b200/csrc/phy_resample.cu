// phy_resample.cu -- rational resampler L/M with a polyphase Kaiser low pass for a batch of antenna streams
// (reference lib/src/phy/resample/resampler.cpp; callers tx.cpp:913-934 + 679-704, rx_pacer.cpp:106-143,
// rx_synced.cpp:711-746, stf_template.cpp:169-186).
// The reference streams symbol by symbol through a history buffer; its output is the plain convolution of the
// zero-extended stream with the up-sampled filter, decimated by M and advanced by the filter delay, so every output
// sample is independent:  y[m] = sum_j taps[j*L + l] x[i0 - j],  p = delay + m*M,  l = p % L,  i0 = p / L.
// One block = RS_TILE consecutive outputs of one stream: the input span (about RS_TILE*M/L + SL samples) is staged in
// shared memory with coalesced loads and zero extension, taps are read as [j][l] rows (at most L distinct banks per
// warp), outputs are written coalesced.  HBM traffic = input once + output once.
#include <algorithm>
#include "phy.cuh"

namespace dnrp {

constexpr uint32_t RS_TILE = 1024, RS_THREADS = 256;

__global__ void __launch_bounds__(RS_THREADS) k_resample(const resample_params_t p, uint32_t span_max) {
    extern __shared__ float2 s_x[];
    float* s_t = reinterpret_cast<float*>(s_x + span_max);
    const uint32_t tile0 = blockIdx.x * RS_TILE, a = blockIdx.y, pkt = blockIdx.z, tid = threadIdx.x;
    const int32_t start = p.start ? p.start[(size_t)pkt * p.start_stride_words] : 0;
    const long long avail = (long long)p.in_stride - start;
    const uint32_t n_in = avail <= 0 ? 0u : (avail < (long long)p.n_in ? (uint32_t)avail : p.n_in);
    const float2* x = p.in + ((size_t)pkt * p.n_ant + a) * p.in_stride + start;
    float2* y = p.out + ((size_t)pkt * p.n_ant + a) * p.out_stride;
    const uint32_t m_end = min(tile0 + RS_TILE, p.n_out);
    if (tile0 < p.n_out) {
        const uint32_t p_lo = p.delay + tile0 * p.M, p_hi = p.delay + (m_end - 1) * p.M;
        const int32_t i_lo = (int32_t)(p_lo / p.L) - (int32_t)(p.SL - 1);
        const uint32_t span = (uint32_t)((int32_t)(p_hi / p.L) - i_lo + 1);
        for (uint32_t i = tid; i < span; i += RS_THREADS) {
            const int32_t idx = i_lo + (int32_t)i;
            s_x[i] = (idx >= 0 && (uint32_t)idx < n_in) ? x[idx] : make_float2(0.f, 0.f);
        }
        for (uint32_t i = tid; i < p.SL * p.L; i += RS_THREADS) s_t[i] = p.taps[i];
        __syncthreads();
        float ph0 = 0.f, inc = 0.f;
        bool mix = false;
        if (p.meta) {
            ph0 = p.meta[pkt].iq_phase_rad; inc = p.meta[pkt].iq_phase_increment;
            mix = ph0 != 0.f || inc != 0.f;
        }
        for (uint32_t m = tile0 + tid; m < m_end; m += RS_THREADS) {
            const uint32_t pp = p.delay + m * p.M, l = pp % p.L;
            const float2* v = s_x + ((int32_t)(pp / p.L) - i_lo);
            const float* t = s_t + l;
            float re = 0.f, im = 0.f;
            for (uint32_t j = 0; j < p.SL; ++j) {
                const float c = t[j * p.L];
                const float2 s = v[-(int32_t)j];
                re = fmaf(c, s.x, re);
                im = fmaf(c, s.y, im);
            }
            if (mix) {
                double ph = (double)ph0 + (double)inc * (double)m;
                ph -= 6.283185307179586476925 * rint(ph * 0.15915494309189533577);
                float sn, cs;
                sincosf((float)ph, &sn, &cs);
                const float r2 = re * cs - im * sn;
                im = re * sn + im * cs;
                re = r2;
            }
            y[m] = make_float2(re, im);
        }
    }
    const uint32_t z0 = max(tile0, p.n_out), z1 = min(tile0 + RS_TILE, p.n_fill);
    for (uint32_t m = z0 + tid; m < z1; m += RS_THREADS) y[m] = make_float2(0.f, 0.f);
}

// ---- register-blocked form for the common ratios (10/9 and 9/10 at the DECT rate): the outputs m, m + L, m + 2L, m + 3L
// share their polyphase branch (p = delay + m M, (m + L) M = m M + L M: same p mod L) and their input windows overlap
// (shift M < SL), so a thread that computes all four keeps the SL taps in registers and loads every input sample once:
// 3 M + SL sample loads + SL tap loads per four outputs instead of 4 SL of each.  Every output is accumulated in the
// same order as in k_resample (tap 0 first): bit-identical results.
template <int L, int M, int SL>
__global__ void __launch_bounds__(RS_THREADS) k_resample_blk(const resample_params_t p, uint32_t span_max) {
    constexpr int R = 4, G = RS_THREADS / L, TILE = G * L * R;
    extern __shared__ float2 s_x[];
    float* s_t = reinterpret_cast<float*>(s_x + span_max);
    const uint32_t tile0 = blockIdx.x * TILE, a = blockIdx.y, pkt = blockIdx.z, tid = threadIdx.x;
    const int32_t start = p.start ? p.start[(size_t)pkt * p.start_stride_words] : 0;
    const long long avail = (long long)p.in_stride - start;
    const uint32_t n_in = avail <= 0 ? 0u : (avail < (long long)p.n_in ? (uint32_t)avail : p.n_in);
    const float2* x = p.in + ((size_t)pkt * p.n_ant + a) * p.in_stride + start;
    float2* y = p.out + ((size_t)pkt * p.n_ant + a) * p.out_stride;
    const uint32_t m_end = min(tile0 + (uint32_t)TILE, p.n_out);
    if (tile0 < p.n_out) {
        const uint32_t p_lo = p.delay + tile0 * M, p_hi = p.delay + (m_end - 1) * M;
        const int32_t i_lo = (int32_t)(p_lo / L) - (int32_t)(SL - 1);
        const uint32_t span = (uint32_t)((int32_t)(p_hi / L) - i_lo + 1);
        for (uint32_t i = tid; i < span; i += RS_THREADS) {
            const int32_t idx = i_lo + (int32_t)i;
            s_x[i] = (idx >= 0 && (uint32_t)idx < n_in) ? x[idx] : make_float2(0.f, 0.f);
        }
        for (uint32_t i = tid; i < SL * L; i += RS_THREADS) s_t[i] = p.taps[i];
        __syncthreads();
        float ph0 = 0.f, inc = 0.f;
        bool mix = false;
        if (p.meta) {
            ph0 = p.meta[pkt].iq_phase_rad; inc = p.meta[pkt].iq_phase_increment;
            mix = ph0 != 0.f || inc != 0.f;
        }
        const uint32_t m0 = tile0 + (tid / L) * (L * R) + tid % L;
        if (tid < G * L && m0 < m_end) {
            const uint32_t pp = p.delay + m0 * M, l = pp % L;
            const float2* v = s_x + ((int32_t)(pp / L) - i_lo);   // sample of tap 0 of output m0; output j: v[j M]
            float c[SL];
#pragma unroll
            for (int k = 0; k < SL; ++k) c[k] = s_t[l + k * L];
            float2 acc[R];
#pragma unroll
            for (int j = 0; j < R; ++j) acc[j] = make_float2(0.f, 0.f);
            // input offset e relative to v, newest first: output j uses it with tap k = j M - e
#pragma unroll
            for (int e = (R - 1) * M; e >= -(SL - 1); --e) {
                // outputs beyond the tile's last one read samples that were not staged: their sums are never stored,
                // the reads stay inside the shared array (span_max is sized for a full tile)
                const float2 sv = v[e];
#pragma unroll
                for (int j = 0; j < R; ++j) {
                    const int k = j * M - e;
                    if (k >= 0 && k < SL) {
                        acc[j].x = fmaf(c[k], sv.x, acc[j].x);
                        acc[j].y = fmaf(c[k], sv.y, acc[j].y);
                    }
                }
            }
#pragma unroll
            for (int j = 0; j < R; ++j) {
                const uint32_t m = m0 + j * L;
                if (m >= m_end) break;
                float re = acc[j].x, im = acc[j].y;
                if (mix) {
                    double ph = (double)ph0 + (double)inc * (double)m;
                    ph -= 6.283185307179586476925 * rint(ph * 0.15915494309189533577);
                    float sn, cs;
                    sincosf((float)ph, &sn, &cs);
                    const float r2 = re * cs - im * sn;
                    im = re * sn + im * cs;
                    re = r2;
                }
                y[m] = make_float2(re, im);
            }
        }
    }
    const uint32_t z0 = max(tile0, p.n_out), z1 = min(tile0 + (uint32_t)TILE, p.n_fill);
    for (uint32_t m = z0 + tid; m < z1; m += RS_THREADS) y[m] = make_float2(0.f, 0.f);
}

template <int L, int M, int SL>
static cudaError_t launch_resample_blk(const resample_params_t& p, uint32_t n_packets, uint32_t n_total, cudaStream_t st) {
    constexpr uint32_t TILE = (RS_THREADS / L) * L * 4;
    const uint32_t span_max = (TILE * M) / L + SL + 2 + 4 * M;   // + the reach of the unused partial outputs
    const size_t smem = (size_t)span_max * sizeof(float2) + (size_t)SL * L * sizeof(float);
    cudaError_t e = cudaFuncSetAttribute(k_resample_blk<L, M, SL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    k_resample_blk<L, M, SL><<<dim3((n_total + TILE - 1) / TILE, p.n_ant, n_packets), RS_THREADS, smem, st>>>(p, span_max);
    return cudaGetLastError();
}

cudaError_t launch_resample(const resample_params_t& p, uint32_t n_packets, cudaStream_t st) {
    if (n_packets == 0) return cudaSuccess;
    const uint32_t n_total = p.n_out > p.n_fill ? p.n_out : p.n_fill;
    if (n_total == 0) return cudaSuccess;
    if ((uint64_t)p.delay + (uint64_t)p.n_out * p.M >= (1ull << 31)) return cudaErrorInvalidValue;
    // the ratios of the reference's radio configurations (hardware = 10/9 DECT rate, resampler_param.hpp) at oversampling 1
    if (p.L == 9 && p.M == 10 && p.SL == 25) return launch_resample_blk<9, 10, 25>(p, n_packets, n_total, st);
    if (p.L == 10 && p.M == 9 && p.SL == 23) return launch_resample_blk<10, 9, 23>(p, n_packets, n_total, st);
    const uint32_t span_max = (RS_TILE * p.M) / p.L + p.SL + 2;
    const size_t smem = (size_t)span_max * sizeof(float2) + (size_t)p.SL * p.L * sizeof(float);
    cudaError_t e = cudaFuncSetAttribute(k_resample, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    k_resample<<<dim3((n_total + RS_TILE - 1) / RS_TILE, p.n_ant, n_packets), RS_THREADS, smem, st>>>(p, span_max);
    return cudaGetLastError();
}

// ---- sc16 -> cf32 (the radio's wire format: int16 I/Q pairs; the reference lets UHD convert on the host CPU,
// hw_usrp.cpp "fc32" cpu format): 4 complex samples per thread, 16 B in, 32 B out
__global__ void __launch_bounds__(256) k_sc16_to_cf32(const int4* __restrict__ in, float4* __restrict__ out, size_t n4,
                                                       float scale) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        const int4 v = __ldg(in + i);
        const int w[4] = {v.x, v.y, v.z, v.w};
        float f[8];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            f[2 * k] = (float)(int16_t)(w[k] & 0xffff) * scale;
            f[2 * k + 1] = (float)(int16_t)(w[k] >> 16) * scale;
        }
        out[2 * i] = make_float4(f[0], f[1], f[2], f[3]);
        out[2 * i + 1] = make_float4(f[4], f[5], f[6], f[7]);
    }
}

// ---- sc8 -> cf32 (UHD's 8-bit wire format, otw_format "sc8": half the bytes of sc16 on the radio link and on the host
// link; the reference's radio layer asks for sc16, hw_usrp.cpp:424,438): 8 complex samples per thread, 16 B in, 64 B out
__global__ void __launch_bounds__(256) k_sc8_to_cf32(const int4* __restrict__ in, float4* __restrict__ out, size_t n8,
                                                      float scale) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (size_t)gridDim.x * blockDim.x) {
        const int4 v = __ldg(in + i);
        const int w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int k = 0; k < 4; ++k)
            out[4 * i + k] = make_float4((float)(int8_t)(w[k] & 0xff) * scale, (float)(int8_t)((w[k] >> 8) & 0xff) * scale,
                                         (float)(int8_t)((w[k] >> 16) & 0xff) * scale, (float)(int8_t)(w[k] >> 24) * scale);
    }
}

cudaError_t launch_sc8_to_cf32(const int8_t* in, float2* out, size_t n_samples, float scale, cudaStream_t st) {
    if (n_samples == 0) return cudaSuccess;
    if (n_samples % 8) return cudaErrorInvalidValue;
    const size_t n8 = n_samples / 8;
    const unsigned grid = (unsigned)std::min<size_t>((n8 + 255) / 256, 148 * 16);
    k_sc8_to_cf32<<<grid, 256, 0, st>>>(reinterpret_cast<const int4*>(in), reinterpret_cast<float4*>(out), n8, scale);
    return cudaGetLastError();
}

// ---- cf32 -> sc16 (the transmit direction of the wire format: UHD converts fc32 -> sc16 on the host CPU, hw_usrp.cpp:422-424):
// round to nearest, saturate; 4 complex samples per thread, 32 B in, 16 B out
__global__ void __launch_bounds__(256) k_cf32_to_sc16(const float4* __restrict__ in, int4* __restrict__ out, size_t n4, float scale) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        const float4 a = __ldg(in + 2 * i), b = __ldg(in + 2 * i + 1);
        auto q = [scale](float v) { return (uint32_t)(uint16_t)(int16_t)max(-32768, min(32767, __float2int_rn(v * scale))); };
        out[i] = make_int4((int)(q(a.x) | (q(a.y) << 16)), (int)(q(a.z) | (q(a.w) << 16)), (int)(q(b.x) | (q(b.y) << 16)),
                           (int)(q(b.z) | (q(b.w) << 16)));
    }
}

cudaError_t launch_cf32_to_sc16(const float2* in, int16_t* out, size_t n_samples, float scale, cudaStream_t st) {
    if (n_samples == 0) return cudaSuccess;
    if (n_samples % 4) return cudaErrorInvalidValue;
    const size_t n4 = n_samples / 4;
    const unsigned grid = (unsigned)std::min<size_t>((n4 + 255) / 256, 148 * 16);
    k_cf32_to_sc16<<<grid, 256, 0, st>>>(reinterpret_cast<const float4*>(in), reinterpret_cast<int4*>(out), n4, scale);
    return cudaGetLastError();
}

// rows of int16 I/Q pairs -> rows of float2 at any offset (the stream ring is appended at arbitrary sample positions):
// one sample per thread, 4 B in, 8 B out
__global__ void __launch_bounds__(256) k_sc16_rows_to_cf32(const uint32_t* __restrict__ in, size_t in_stride, float2* __restrict__ out,
                                                            size_t out_stride, uint32_t n, float scale) {
    const uint32_t a = blockIdx.y;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t w = __ldg(in + a * in_stride + i);
        out[a * out_stride + i] = make_float2((float)(int16_t)(w & 0xffff) * scale, (float)(int16_t)(w >> 16) * scale);
    }
}

cudaError_t launch_sc16_rows_to_cf32(const int16_t* in, size_t in_stride_samples, float2* out, size_t out_stride_samples,
                                     uint32_t n_samples, uint32_t n_rows, float scale, cudaStream_t st) {
    if (n_samples == 0 || n_rows == 0) return cudaSuccess;
    const unsigned gx = (unsigned)std::min<size_t>(((size_t)n_samples + 255) / 256, 148 * 8);
    k_sc16_rows_to_cf32<<<dim3(gx, n_rows), 256, 0, st>>>(reinterpret_cast<const uint32_t*>(in), in_stride_samples, out,
                                                          out_stride_samples, n_samples, scale);
    return cudaGetLastError();
}

cudaError_t launch_sc16_to_cf32(const int16_t* in, float2* out, size_t n_samples, float scale, cudaStream_t st) {
    if (n_samples == 0) return cudaSuccess;
    if (n_samples % 4) return cudaErrorInvalidValue;
    const size_t n4 = n_samples / 4;
    const unsigned grid = (unsigned)std::min<size_t>((n4 + 255) / 256, 148 * 16);
    k_sc16_to_cf32<<<grid, 256, 0, st>>>(reinterpret_cast<const int4*>(in), reinterpret_cast<float4*>(out), n4, scale);
    return cudaGetLastError();
}

}  // namespace dnrp

// phy_channel.cu -- simulation channel on the GPU: superposition of all TX antennas onto every RX antenna
// with per-link small scale coefficients (AWGN: 1, flat: CN(0,1) per link), integer delay into the RX
// window, and complex white Gaussian noise at an absolute level n0 (dBFS).
// Reference: lib/src/simulation/wireless/channel_awgn.cpp:39-64, channel_flat.cpp:47-88,
// lib/src/simulation/vspace.cpp:487-497, lib/src/simulation/hardware/noise.cpp:31-41.
// Random numbers: Philox4x32-10, counter = (sample pair, rx antenna, packet, stream), key = seed; the
// oracle (oracle/channel.c) draws from the same generator so that realisations agree.
#include "phy.cuh"
#include "fft.cuh"

namespace dnrp {

__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += 0x9E3779B9u;
        k.y += 0xBB67AE85u;
    }
    return c;
}

// two uniform 32-bit words -> one complex Gaussian sample with E|z|^2 = 1 (Box-Muller)
__device__ __forceinline__ float2 cn01(uint32_t a, uint32_t b) {
    const float u1 = ((float)(a >> 8) + 0.5f) * (1.0f / 16777216.0f);   // (0, 1)
    const float u2 = ((float)(b >> 8) + 0.5f) * (1.0f / 16777216.0f);
    const float r = sqrtf(-logf(u1));                                     // sqrt(-2 ln u / 2)
    float s, c;
    sincospif(2.0f * u2, &s, &c);
    return make_float2(r * c, r * s);
}

// grid (ceil(rx_stride / 512), N_RX, packets), 256 threads, two samples per thread
__global__ void __launch_bounds__(256) k_channel(const channel_params_t p) {
    const uint32_t a = blockIdx.y, pkt = blockIdx.z;
    const chan_meta_dev_t m = p.meta[pkt];
    __shared__ float2 g[8];
    if (threadIdx.x < p.N_TX) {
        float2 v = make_float2(1.f, 0.f);
        if (m.mode == 1) {
            const uint4 r = philox4x32_10(make_uint4(threadIdx.x, a, pkt, 1u), make_uint2(m.seed_lo, m.seed_hi));
            v = cn01(r.x, r.y);
        }
        g[threadIdx.x] = v;
    }
    __syncthreads();
    const uint32_t pair = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t n0 = 2 * pair;
    if (n0 >= p.rx_stride) return;
    const uint4 r = philox4x32_10(make_uint4(pair, a, pkt, 0u), make_uint2(m.seed_lo, m.seed_hi));
    float2 y[2];
    y[0] = cn01(r.x, r.y);
    y[1] = cn01(r.z, r.w);
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        y[i].x *= m.sigma;
        y[i].y *= m.sigma;
        const uint32_t n = n0 + i;
        if (n >= m.t0 && n - m.t0 < p.n_samples) {
            for (uint32_t t = 0; t < p.N_TX; ++t) {
                const float2 x = p.tx[((size_t)pkt * p.N_TX + t) * p.tx_stride + (n - m.t0)];
                y[i] = cadd(y[i], cmul(g[t], x));
            }
        }
    }
    float2* out = p.rx + ((size_t)pkt * p.N_RX + a) * p.rx_stride + n0;
    if (n0 + 1 < p.rx_stride) *reinterpret_cast<float4*>(out) = make_float4(y[0].x, y[0].y, y[1].x, y[1].y);
    else out[0] = y[0];
}

cudaError_t launch_channel(const channel_params_t& p, uint32_t n_packets, cudaStream_t st) {
    const uint32_t pairs = (p.rx_stride + 1) / 2;
    k_channel<<<dim3((pairs + 255) / 256, p.N_RX, n_packets), 256, 0, st>>>(p);
    return cudaGetLastError();
}

}  // namespace dnrp

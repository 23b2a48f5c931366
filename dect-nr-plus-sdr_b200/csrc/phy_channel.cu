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
// hw_simulator_t::clip_and_quantize (hw_simulator.cpp:710-731, simulation/hardware/clip.cpp, quantize.cpp)
__device__ __forceinline__ float2 hw_clip_quantize(float2 v, float clip, float lsb) {
    if (clip > 0.f) {
        v.x = fminf(fmaxf(v.x, -clip), clip);
        v.y = fminf(fmaxf(v.y, -clip), clip);
        if (lsb > 0.f) {
            v.x = roundf(__fdiv_rn(v.x, lsb)) * lsb;
            v.y = roundf(__fdiv_rn(v.y, lsb)) * lsb;
        }
    }
    return v;
}

__global__ void __launch_bounds__(256) k_channel(const channel_params_t p) {
    const uint32_t a = blockIdx.y, pkt = blockIdx.z;
    const chan_meta_dev_t m = p.meta[pkt];
    if (m.mode == 2) return;   // k_channel_tdl
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
                const float2 x = hw_clip_quantize(p.tx[((size_t)pkt * p.N_TX + t) * p.tx_stride + (n - m.t0)], m.clip, m.dac_lsb);
                const float2 gx = cmul(g[t], x);
                y[i] = cadd(y[i], make_float2(gx.x * m.large_scale, gx.y * m.large_scale));
            }
        }
        y[i] = hw_clip_quantize(y[i], m.clip, m.adc_lsb);
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

// ---- doubly selective channel (channel_doubly.cpp:59-87, link.cpp:165-282): per link a tapped delay line whose tap
// coefficients are sums of 40 unit sinusoids with Doppler fD cos(angle).  The reference rotates chunk by chunk with a
// recursive rotator restarted from time_64 % |period|; here a block evaluates that closed form once per sinusoid at
// its first sample and every thread advances it with fixed rotations (by its lane offset, then by the block width).
constexpr uint32_t TDL_R = 8, TDL_THREADS = 256, TDL_SIN = 40, TDL_TAPS = 9;

__device__ __forceinline__ float rand_m1p1(uint32_t r) {
    return ((float)(r >> 8) + 0.5f) * (1.0f / 8388608.0f) - 1.0f;
}

// grid (packet), one thread per (rx, tx, tap, sinusoid): draw angle of arrival and start phase
__global__ void __launch_bounds__(256) k_channel_tdl_setup(const channel_params_t p) {
    const uint32_t pkt = blockIdx.x;
    const chan_meta_dev_t& m = p.meta[pkt];
    if (m.mode != 2) return;
    const uint32_t per_ant = p.N_TX * TDL_TAPS * TDL_SIN, total = p.N_RX * per_ant;
    for (uint32_t e = threadIdx.x; e < total; e += blockDim.x) {
        const uint32_t a = e / per_ant, rem = e % per_ant;
        const uint32_t t = rem / (TDL_TAPS * TDL_SIN), i = (rem / TDL_SIN) % TDL_TAPS, j = rem % TDL_SIN;
        tdl_sin_t s;
        s.period = 0x7fffffffffffffffLL; s.phase0 = 0.f; s.s2s = 0.f;
        if (i < m.n_taps) {
            const uint4 r = philox4x32_10(make_uint4(j | (i << 8) | (t << 16), a, pkt, 2u),
                                          make_uint2(m.seed_lo, m.seed_hi));
            const double ang = (double)rand_m1p1(r.x) * 2.0 * 3.14159265358979323846;
            const double fd = (double)m.fD_Hz * cos(ang);
            if (!(-0.01f < fd && fd < 0.01f)) s.period = (long long)(m.samp_rate / fd);   // else: dead band
            s.phase0 = rand_m1p1(r.y) * 2.0f * 3.14159265358979323846f;
            s.s2s = (float)((double)(2.0f * 3.14159265358979323846f) / (double)s.period);
        }
        p.sin_tab[(size_t)pkt * total + e] = s;
    }
}

// grid (ceil(rx_stride / (TDL_THREADS * TDL_R)), N_RX, packet); thread = samples base + tid + 256 r
__global__ void __launch_bounds__(TDL_THREADS) k_channel_tdl(const channel_params_t p) {
    extern __shared__ float4 s_rot[];   // per (tx, tap, sinusoid): phasor at the block's first sample, rotation by 256
    const uint32_t a = blockIdx.y, pkt = blockIdx.z, tid = threadIdx.x;
    const chan_meta_dev_t& m = p.meta[pkt];
    if (m.mode != 2) return;
    float* s_s2s = reinterpret_cast<float*>(s_rot + p.N_TX * TDL_TAPS * TDL_SIN);
    const uint32_t base = blockIdx.x * TDL_THREADS * TDL_R;
    const uint32_t per_ant = p.N_TX * TDL_TAPS * TDL_SIN;
    const tdl_sin_t* tab = p.sin_tab + ((size_t)pkt * p.N_RX + a) * per_ant;
    for (uint32_t e = tid; e < per_ant; e += TDL_THREADS) {
        const tdl_sin_t s = tab[e];
        const long long ap = s.period < 0 ? -s.period : s.period;
        const double s2s = (double)(2.0f * 3.14159265358979323846f) / (double)s.period;
        const float ph = (float)(s2s * (double)((m.time_64 + (long long)base) % ap)) + s.phase0;
        float sn, cs, sn2, cs2;
        sincosf(ph, &sn, &cs);
        sincosf(s.s2s * (float)TDL_THREADS, &sn2, &cs2);
        s_rot[e] = make_float4(cs, sn, cs2, sn2);
        s_s2s[e] = s.s2s;
    }
    __syncthreads();
    float2 y[TDL_R];
#pragma unroll
    for (uint32_t r = 0; r < TDL_R; ++r) y[r] = make_float2(0.f, 0.f);
    const uint32_t h = m.interp_h;
    if (h) {
        // ---- slowly varying coefficients (chan_meta_dev_t::interp_h): every tap coefficient of the tile is evaluated
        // exactly at knots h samples apart (one thread per knot, the 40 sinusoids in a fixed order) and interpolated
        // linearly in between -- 40 / h sinusoid evaluations per sample and tap instead of 40
        constexpr uint32_t TILE = TDL_THREADS * TDL_R;
        const uint32_t nk = TILE / h + 1, lg = 31u - (uint32_t)__clz(h);
        const float hinv = 1.0f / (float)h;
        // knots of one TX antenna: two partial sums (sinusoids 0..19 / 20..39, one thread each) [2][n_taps][nk_max], added
        // into the first half afterwards -- twice the threads busy, half the serial chain, fixed order of additions.
        // Behind them the antenna's samples of the tile, clipped / quantised once, zero outside the packet, with
        // TDL_HALO samples in front for the tap delays: the taps then read shared memory without range tests.
        const uint32_t nk_max = TILE / p.tdl_min_h + 1;
        float2* knot = reinterpret_cast<float2*>(s_s2s + per_ant);
        float2* xs = knot + 2 * TDL_TAPS * nk_max;          // [TDL_HALO + TILE]
        const uint32_t nkt = m.n_taps * nk;
        uint32_t kk[TDL_R];
        float fr[TDL_R];
#pragma unroll
        for (uint32_t r = 0; r < TDL_R; ++r) {
            const uint32_t pos = tid + r * TDL_THREADS;
            kk[r] = pos >> lg;
            fr[r] = (float)(pos & (h - 1u)) * hinv;
        }
        const int k_first = (int)base - (int)m.t0 - (int)TDL_HALO;   // TX sample behind xs[0]
        for (uint32_t t = 0; t < p.N_TX; ++t) {
            const float2* x = p.tx + ((size_t)pkt * p.N_TX + t) * p.tx_stride;
            __syncthreads();   // the previous antenna's knots and samples are consumed
            for (uint32_t j = tid; j < TDL_HALO + TILE; j += TDL_THREADS) {
                const int k = k_first + (int)j;
                xs[j] = (k >= 0 && k < (int)p.n_samples) ? hw_clip_quantize(x[k], m.clip, m.dac_lsb) : make_float2(0.f, 0.f);
            }
            for (uint32_t e2 = tid; e2 < 2 * nkt; e2 += TDL_THREADS) {
                const uint32_t hf = e2 >= nkt ? 1u : 0u, e = e2 - hf * nkt;
                const uint32_t i = e / nk, k = e - i * nk;
                const uint32_t e0 = (t * TDL_TAPS + i) * TDL_SIN + hf * (TDL_SIN / 2);
                const float dn = (float)(k * h);
                float2 c = make_float2(0.f, 0.f);
                for (uint32_t j = 0; j < TDL_SIN / 2; ++j) {
                    const float4 rt = s_rot[e0 + j];
                    float sn, cs;
                    // the knot path is only taken for w h <= 1.6e-3 rad, i.e. arguments below 0.4 rad over the tile:
                    // the fast intrinsic is exact to 1e-7 there
                    __sincosf(s_s2s[e0 + j] * dn, &sn, &cs);
                    c = cadd(c, cmul(make_float2(rt.x, rt.y), make_float2(cs, sn)));
                }
                const float sc = m.tap_scale[i];
                knot[e2] = make_float2(c.x * sc, c.y * sc);
            }
            __syncthreads();
            for (uint32_t e = tid; e < nkt; e += TDL_THREADS) knot[e] = cadd(knot[e], knot[nkt + e]);
            __syncthreads();
            for (uint32_t i = 0; i < m.n_taps; ++i) {
                const float2* kn = knot + i * nk;
                const float2* xi = xs + TDL_HALO - m.tap_delay[i] + tid;
#pragma unroll
                for (uint32_t r = 0; r < TDL_R; ++r) {
                    const float2 c0 = kn[kk[r]], c1 = kn[kk[r] + 1];
                    const float2 c = make_float2(c0.x + fr[r] * (c1.x - c0.x), c0.y + fr[r] * (c1.y - c0.y));
                    y[r] = cadd(y[r], cmul(c, xi[r * TDL_THREADS]));
                }
            }
        }
    } else
    for (uint32_t t = 0; t < p.N_TX; ++t) {
        const float2* x = p.tx + ((size_t)pkt * p.N_TX + t) * p.tx_stride;
        for (uint32_t i = 0; i < m.n_taps; ++i) {
            float2 acc[TDL_R];
#pragma unroll
            for (uint32_t r = 0; r < TDL_R; ++r) acc[r] = make_float2(0.f, 0.f);
            const uint32_t e0 = (t * TDL_TAPS + i) * TDL_SIN;
            for (uint32_t j = 0; j < TDL_SIN; ++j) {
                const float4 rt = s_rot[e0 + j];
                float sn, cs;
                sincosf(s_s2s[e0 + j] * (float)tid, &sn, &cs);
                float2 ph = cmul(make_float2(rt.x, rt.y), make_float2(cs, sn));
                const float2 step = make_float2(rt.z, rt.w);
#pragma unroll
                for (uint32_t r = 0; r < TDL_R; ++r) {
                    acc[r] = cadd(acc[r], ph);
                    ph = cmul(ph, step);
                }
            }
            const float sc = m.tap_scale[i];
            const long long off = (long long)m.t0 + m.tap_delay[i];
#pragma unroll
            for (uint32_t r = 0; r < TDL_R; ++r) {
                const long long k = (long long)(base + tid + r * TDL_THREADS) - off;
                if (k >= 0 && k < (long long)p.n_samples)
                    y[r] = cadd(y[r], cmul(make_float2(acc[r].x * sc, acc[r].y * sc), hw_clip_quantize(x[k], m.clip, m.dac_lsb)));
            }
        }
    }
    float2* out = p.rx + ((size_t)pkt * p.N_RX + a) * p.rx_stride;
#pragma unroll
    for (uint32_t r = 0; r < TDL_R; ++r) {
        const uint32_t n = base + tid + r * TDL_THREADS;
        if (n >= p.rx_stride) continue;
        const uint4 q = philox4x32_10(make_uint4(n >> 1, a, pkt, 0u), make_uint2(m.seed_lo, m.seed_hi));
        const float2 w = (n & 1u) ? cn01(q.z, q.w) : cn01(q.x, q.y);
        out[n] = hw_clip_quantize(make_float2(y[r].x * m.large_scale + w.x * m.sigma, y[r].y * m.large_scale + w.y * m.sigma),
                                  m.clip, m.adc_lsb);
    }
}

cudaError_t launch_channel_tdl(const channel_params_t& p, uint32_t n_packets, cudaStream_t st) {
    k_channel_tdl_setup<<<n_packets, 256, 0, st>>>(p);
    // phasors and increments per (tx, tap, sinusoid); interpolated path (some packet has a knot distance): the knots of
    // one TX antenna [2 partial sums][9 taps][2048 / h_min + 1] and the staged samples of the tile [TDL_HALO + 2048]
    size_t smem = (size_t)p.N_TX * TDL_TAPS * TDL_SIN * (sizeof(float4) + sizeof(float));
    if (p.tdl_min_h)
        smem += (2 * (size_t)TDL_TAPS * (TDL_THREADS * TDL_R / p.tdl_min_h + 1) + TDL_HALO + TDL_THREADS * TDL_R) * sizeof(float2);
    cudaError_t e = cudaFuncSetAttribute(k_channel_tdl, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    const uint32_t per_block = TDL_THREADS * TDL_R;
    k_channel_tdl<<<dim3((p.rx_stride + per_block - 1) / per_block, p.N_RX, n_packets), TDL_THREADS, smem, st>>>(p);
    return cudaGetLastError();
}

}  // namespace dnrp

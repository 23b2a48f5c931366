// phy_tx.cu -- transmitter OFDM stage, one thread block per (OFDM symbol, TX antenna, packet):
// bits -> QAM symbols -> transmit-diversity precoding -> resource mapping (STF / PCC / DRS / PDC) ->
// codebook beamforming -> scaling -> IFFT -> cyclic prefix (+ STF cover sequence, + mixer), all in one
// pass through shared memory.  Replaces the per-symbol loop of reference
// lib/src/phy/tx/tx.cpp:216-279 (run_stf, run_pcc, run_drs, run_pdc, run_beamforming,
// run_ifft_cp_scale, stf_t::apply_cover_sequence, mixer) for L = M = 1.
#include "launch_util.hpp"
#include <cstdlib>
#include "fft.cuh"
#include "phy.cuh"

namespace dnrp {

// LTE Gray mapping, TS 36.211 7.1 (srsRAN lte_tables.c; reference fix/mod.cpp:137-174), bits MSB first
__device__ __forceinline__ float2 qam_symbol(const uint8_t* __restrict__ bits, uint32_t off, uint32_t N_bps,
                                             const float2* __restrict__ lut) {
    // N_bps <= 10 bits span at most three bytes: one 24-bit window instead of a byte load per bit (a later byte is
    // only touched when the symbol reaches into it); the constellation point comes from the host-built table
    // (plan_phy.cu: the recursion (1-2b0)(2^(h-1) - (1-2b2)(2^(h-2) - ...)) / sqrt(2 / 10 / 42 / 170 / 682); 1024-QAM, MCS
    // 10 and 11, is outside the reference, tx.cpp:53-60 -- own extension)
    const uint32_t b = off >> 3, end = (off & 7) + N_bps;
    const uint32_t win = ((uint32_t)bits[b] << 16) | (end > 8 ? (uint32_t)bits[b + 1] << 8 : 0u) | (end > 16 ? (uint32_t)bits[b + 2] : 0u);
    const uint32_t v = (win >> (24 - end)) & ((1u << N_bps) - 1u);
    return __ldg(lut + v);
}

__device__ __forceinline__ float2 apply_w(int8_t w, float2 v) {
    switch (w) {
        case 1: return v;
        case -1: return make_float2(-v.x, -v.y);
        case 2: return make_float2(-v.y, v.x);    // * j
        case -2: return make_float2(v.y, -v.x);   // * -j
        default: return make_float2(0.f, 0.f);
    }
}

// value of subcarrier k (0 = lowest occupied) of OFDM symbol l on one TX antenna before scaling: STF (l = 0), DRS, PCC
// and PDC cells through the antenna's row W of the beamforming matrix (tx.cpp:602-677, 729-860, 936-1116)
__device__ __forceinline__ float2 tx_cell(const phy_dims_t& d, uint32_t l, uint32_t k, const int8_t* W, const uint8_t* pcc_bits,
                                          const uint8_t* pdc_bits) {
    float2 acc = make_float2(0.f, 0.f);
    if (l == 0) {
        const int s = d.stf_sign[k];
        if (s != 0) {  // only transmit stream 0 carries the STF (run_beamforming(1))
            const float a = 0.70710678118654752f * (float)s;
            acc = apply_w(W[0], make_float2(a, a));
        }
        return acc;
    }
    const uint32_t cw = d.cell[l * d.Nk + k];
    const uint32_t type = cw >> 30, idx = cw & 0x03ffffffu;
    if (type == 1) {  // DRS of transmit stream ts
        const uint32_t ts = (cw >> 27) & 7u;
        acc = apply_w(W[ts], make_float2((cw >> 26) & 1u ? -1.f : 1.f, 0.f));
    } else if (type == 2 || type == 3) {
        const uint8_t* bits = type == 2 ? pcc_bits : pdc_bits;
        const uint32_t nbps = type == 2 ? 2u : d.N_bps;
        const float2* lut = d.qam_lut + qam_lut_base(nbps);
        if (d.N_TS == 1) {
            acc = apply_w(W[0], qam_symbol(bits, idx * nbps, nbps, lut));
        } else if (type == 2 || d.txdiv) {
            // Alamouti pair (q even, q odd): stream A carries s[q], stream B the flipped partner; the pair's streams
            // (Table 6.3.3.2) come from the per-cell table
            const uint32_t ab = d.cell_ab[l * d.Nk + k], A = ab & 15u, B = ab >> 4;
            if (W[A] != 0) acc = apply_w(W[A], qam_symbol(bits, idx * nbps, nbps, lut));
            if (W[B] != 0) {
                const float2 o = qam_symbol(bits, (idx ^ 1u) * nbps, nbps, lut);
                const float2 f = (idx & 1u) ? make_float2(o.x, -o.y) : make_float2(-o.x, o.y);
                acc = cadd(acc, apply_w(W[B], f));
            }
        } else {
            // spatial multiplexing: symbol idx*N_SS + s on stream s (tx.cpp:1054-1060)
            for (uint32_t s = 0; s < d.N_SS; ++s)
                if (W[s] != 0) acc = cadd(acc, apply_w(W[s], qam_symbol(bits, (idx * d.N_SS + s) * nbps, nbps, lut)));
        }
    }
    return acc;
}

template <bool BIG>
__global__ void __launch_bounds__(256) k_tx_ofdm(const txofdm_params_t p) {
    extern __shared__ __align__(16) float2 sm[];
    const phy_dims_t& d = p.d;
    const uint32_t l = blockIdx.x, ant = blockIdx.y, pkt = blockIdx.z;
    const uint32_t tid = threadIdx.x, nthr = blockDim.x;
    // N_os <= 4096: two ping-pong buffers of fft_buf_len(N_os); longer: one array of N_os + two row scratch buffers (fft.cuh)
    constexpr bool big = BIG;
    const uint32_t Mb = d.N_os / 16;
    float2* X = sm;
    float2* Y = sm + (big ? d.N_os : fft_buf_len(d.N_os));
    const tx_meta_dev_t meta = p.meta[pkt];
    const uint8_t* pcc_bits = p.pcc_bits + (size_t)pkt * p.pcc_stride;
    const uint8_t* pdc_bits = p.pdc_bits + (size_t)pkt * p.pdc_stride;
    const int8_t* W = d.W + (size_t)meta.codebook_index * d.N_TX * d.N_TS + ant * d.N_TS;
    const float wscale = meta.DAC_scale * (meta.optimal_scaling_DAC ? d.W_scale_opt[meta.codebook_index]
                                                                    : d.W_scale[meta.codebook_index]);
    const float scale = (l == 0 ? rsqrtf((float)(d.N_OCC / 4)) : rsqrtf((float)d.N_OCC)) * wscale;

    for (uint32_t i = tid; i < d.N_os; i += nthr) X[i] = make_float2(0.f, 0.f);
    __syncthreads();

    // ---- frequency domain of this antenna
    for (uint32_t k = tid; k < d.Nk; k += nthr) {
        const float2 acc = tx_cell(d, l, k, W, pcc_bits, pdc_bits);
        // FFT bin: DC and the upper half from bin 0, the lower half at the top (tx_rx.cpp:197-240)
        const uint32_t bin = k >= d.N_OCC / 2 ? k - d.N_OCC / 2 : d.N_os - d.N_OCC / 2 + k;
        X[bin] = make_float2(acc.x * scale, acc.y * scale);
    }

    const float2* x;
    if (big) {
        big_fft<true>(X, Y, Y + fft_buf_len(Mb), d.N_os, d.twiddle);
        x = X;   // time sample n at big_fft_at(n, Mb)
    } else {
        x = block_fft<true>(X, Y, d.N_os, d.twiddle, d.twiddle_p4);
    }

    // ---- cyclic prefix, STF cover sequence, mixer
    const uint32_t cp = l == 0 ? d.stf_cp_os : d.cp_os;
    const uint32_t len = cp + d.N_os;
    const uint32_t start = l == 0 ? 0 : d.stf_len_os + (l - 1) * d.sym_len_os;
    float2* out = p.iq + ((size_t)pkt * d.N_TX + ant) * p.iq_stride + start;
    const bool mix = !p.no_mix && (meta.iq_phase_rad != 0.f || meta.iq_phase_increment != 0.f);
    for (uint32_t n = tid; n < len; n += nthr) {
        uint32_t src = (n + d.N_os - (cp % d.N_os)) % d.N_os;
        float2 v = x[big ? big_fft_at(src, Mb) : src];
        if (l == 0) {
            const float c[9] = {1.f, -1.f, 1.f, 1.f, -1.f, -1.f, -1.f, -1.f, -1.f};  // stf.hpp:146-151
            const float cs = c[n / d.pat_len_os];
            v.x *= cs; v.y *= cs;
        }
        if (mix) {
            // closed-form phase (the reference's VOLK rotator is recursive; tolerance item in DESIGN.md)
            double ph = (double)meta.iq_phase_rad + (double)meta.iq_phase_increment * (double)(start + n);
            ph -= 6.283185307179586476925 * rint(ph * 0.15915494309189533577);
            float s, c;
            sincosf((float)ph, &s, &c);
            v = cmul(v, make_float2(c, s));
        }
        out[n] = v;
    }
}

// ---- the same for N_os = 1024 (beta = 16 at the DECT rate) with every size fixed at compile time, the transmit-side twin
// of k_rx_fft1024 (phy_rx.cu): 64 threads per (symbol, antenna), four groups per block; every thread forms its 16 bins
// t + 64 m straight in registers (no zeroed spectrum in shared memory), inverse passes 16 x 16 x 4 with the data in
// registers between the passes and one padded buffer per group, named group barriers, twiddles from the per-pass tables.
__device__ __forceinline__ void tx_group_bar(uint32_t grp) {
    switch (grp) {
        case 0: asm volatile("bar.sync 1, 64;" ::: "memory"); break;
        case 1: asm volatile("bar.sync 2, 64;" ::: "memory"); break;
        case 2: asm volatile("bar.sync 3, 64;" ::: "memory"); break;
        default: asm volatile("bar.sync 4, 64;" ::: "memory"); break;
    }
}

__global__ void __launch_bounds__(256, 4) k_tx_ofdm1024(const txofdm_params_t p) {
    constexpr uint32_t N = 1024, GT = 64, G = 4, BL = N + N / 16;
    extern __shared__ __align__(16) float2 sm[];   // [G][BL]
    const phy_dims_t& d = p.d;
    const uint32_t grp = threadIdx.x >> 6, tid = threadIdx.x & 63;
    const uint32_t item = blockIdx.x * G + grp, pkt = blockIdx.y;
    if (item >= (d.N_DF + 1) * d.N_TX) return;   // whole groups leave; only group barriers below
    const uint32_t l = item / d.N_TX, ant = item - l * d.N_TX;
    float2* A = sm + grp * BL;
    const float2* __restrict__ tw16 = d.twiddle_p16;         // [15][16] W_256^(k r)
    const float2* __restrict__ tw4 = d.twiddle_p16 + 240;    // [3][256] W_1024^(j r)
    const tx_meta_dev_t meta = p.meta[pkt];
    const uint8_t* pcc_bits = p.pcc_bits + (size_t)pkt * p.pcc_stride;
    const uint8_t* pdc_bits = p.pdc_bits + (size_t)pkt * p.pdc_stride;
    const int8_t* W = d.W + (size_t)meta.codebook_index * d.N_TX * d.N_TS + ant * d.N_TS;
    const float wscale = meta.DAC_scale * (meta.optimal_scaling_DAC ? d.W_scale_opt[meta.codebook_index]
                                                                    : d.W_scale[meta.codebook_index]);
    const float scale = (l == 0 ? rsqrtf((float)(d.N_OCC / 4)) : rsqrtf((float)d.N_OCC)) * wscale;

    // bins t + 64 m: DC and the upper half from bin 0, the lower half at the top (tx_rx.cpp:197-240), zero in between
    const uint32_t half = d.N_OCC / 2;
    // one copy of the cell code (a rolled loop; sixteen inlined copies thrash the instruction cache: measured 24 % slower
    // than the generic kernel): every thread parks its own bins in the group's buffer and reads them back itself
#pragma unroll 1
    for (uint32_t m = 0; m < 16; ++m) {
        const uint32_t bin = tid + 64 * m;
        float2 acc = make_float2(0.f, 0.f);
        if (bin <= half) acc = tx_cell(d, l, bin + half, W, pcc_bits, pdc_bits);
        else if (bin >= N - half) acc = tx_cell(d, l, bin - (N - half), W, pcc_bits, pdc_bits);
        A[bin] = make_float2(acc.x * scale, acc.y * scale);
    }
    float2 v[16];
#pragma unroll
    for (int m = 0; m < 16; ++m) v[m] = A[tid + 64 * m];
    tx_group_bar(grp);   // every thread has its bins: the buffer is free for the first exchange
    // pass 1 (Ns = 1)
    dft16<true>(v);
#pragma unroll
    for (int r = 0; r < 16; ++r) A[17 * tid + r] = v[r];
    tx_group_bar(grp);
    // pass 2 (Ns = 16)
    {
        const uint32_t k = tid & 15u, q = tid >> 4;
        const float2* xj = A + tid + q;
#pragma unroll
        for (int r = 0; r < 16; ++r) v[r] = xj[68 * r];
        tx_group_bar(grp);
#pragma unroll
        for (int r = 1; r < 16; ++r) v[r] = cmul(v[r], tw<true>(tw16, (r - 1) * 16 + k));
        dft16<true>(v);
        float2* yj = A + 272 * q + k;
#pragma unroll
        for (int r = 0; r < 16; ++r) yj[17 * r] = v[r];
    }
    tx_group_bar(grp);
    // pass 3 (Ns = 256, radix 4)
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const uint32_t j = tid + 64 * i;
        const float2* xj = A + j + (j >> 4);
#pragma unroll
        for (int r = 0; r < 4; ++r) v[4 * i + r] = xj[272 * r];
    }
    tx_group_bar(grp);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const uint32_t j = tid + 64 * i;
#pragma unroll
        for (int r = 1; r < 4; ++r) v[4 * i + r] = cmul(v[4 * i + r], tw<true>(tw4, (r - 1) * 256 + j));
        dft4<true>(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
#pragma unroll
        for (int r = 0; r < 4; ++r) A[j + 256 * r] = v[4 * i + r];
    }
    tx_group_bar(grp);

    // ---- cyclic prefix, STF cover sequence, mixer
    const uint32_t cp = l == 0 ? d.stf_cp_os : d.cp_os;
    const uint32_t len = cp + N;
    const uint32_t start = l == 0 ? 0 : d.stf_len_os + (l - 1) * d.sym_len_os;
    float2* out = p.iq + ((size_t)pkt * d.N_TX + ant) * p.iq_stride + start;
    const bool mix = !p.no_mix && (meta.iq_phase_rad != 0.f || meta.iq_phase_increment != 0.f);
    const uint32_t shift = N - (cp & (N - 1));
    for (uint32_t n = tid; n < len; n += GT) {
        float2 x = A[(n + shift) & (N - 1)];
        if (l == 0 && ((0x1F2u >> (n / d.pat_len_os)) & 1u)) {   // cover sequence 1 -1 1 1 -1 -1 -1 -1 -1 (stf.hpp:146-151)
            x.x = -x.x; x.y = -x.y;
        }
        if (mix) {
            double ph = (double)meta.iq_phase_rad + (double)meta.iq_phase_increment * (double)(start + n);
            ph -= 6.283185307179586476925 * rint(ph * 0.15915494309189533577);
            float s, c;
            sincosf((float)ph, &s, &c);
            x = cmul(x, make_float2(c, s));
        }
        out[n] = x;
    }
}

cudaError_t launch_tx_ofdm(const txofdm_params_t& p, uint32_t n_packets, cudaStream_t st) {
    const size_t smem = fft_is_big(p.d.N_os) ? sizeof(float2) * (p.d.N_os + 2 * (size_t)fft_buf_len(p.d.N_os / 16))
                                             : 2 * sizeof(float2) * fft_buf_len(p.d.N_os);
    const bool big = fft_is_big(p.d.N_os);
    if (smem > 48 * 1024) {
        static smem_optin_t optin[2];   // per device and variant
        cudaError_t e = optin[big].ensure(smem, [big](size_t b) {
            return big ? cudaFuncSetAttribute(k_tx_ofdm<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)b)
                       : cudaFuncSetAttribute(k_tx_ofdm<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)b);
        });
        if (e != cudaSuccess) return e;
    }
    // DNRP_FFT_GENERIC=1: the run-time sized kernel also at N_os = 1024 (A/B measurements)
    static const bool generic = [] { const char* e = getenv("DNRP_FFT_GENERIC"); return e && e[0] == '1'; }();
    if (p.d.N_os == 1024 && !generic) {
        const uint32_t items = (p.d.N_DF + 1) * p.d.N_TX;
        k_tx_ofdm1024<<<dim3((items + 3) / 4, n_packets), 256, sizeof(float2) * 4 * (1024 + 64), st>>>(p);
        return cudaGetLastError();
    }
    const dim3 grid(p.d.N_DF + 1, p.d.N_TX, n_packets);
    if (big) k_tx_ofdm<true><<<grid, 256, smem, st>>>(p);
    else k_tx_ofdm<false><<<grid, 256, smem, st>>>(p);
    return cudaGetLastError();
}

}  // namespace dnrp

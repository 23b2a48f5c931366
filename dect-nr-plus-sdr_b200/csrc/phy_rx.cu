// phy_rx.cu -- synchronised receiver front end on the GPU (reference lib/src/phy/rx/rx_synced/
// rx_synced.cpp): STF stage (CFO derotation, cover sequence, residual CFO, FFT, zero-forcing at the STF
// cells, fractional STO, SNR), data field FFTs with DRS zero-forcing, Wiener class selection, and the
// fused channel interpolation + MRC/Alamouti combining + soft demapping kernel.
#include <cstdlib>
#include "launch_util.hpp"
#include "fft.cuh"
#include "phy.cuh"

namespace dnrp {

__device__ __forceinline__ float2 rot(double ph) {  // exp(j*ph), argument reduced in double
    ph -= 6.283185307179586476925 * rint(ph * 0.15915494309189533577);
    float s, c;
    sincosf((float)ph, &s, &c);
    return make_float2(c, s);
}

// block-wide sum of a float2 (all threads call; result valid in every thread)
__device__ float2 block_sum2(float2 v, float2* s_red) {
    for (int o = 16; o > 0; o >>= 1) {
        v.x += __shfl_xor_sync(0xffffffffu, v.x, o);
        v.y += __shfl_xor_sync(0xffffffffu, v.y, o);
    }
    const uint32_t w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    __syncthreads();
    if ((threadIdx.x & 31) == 0) s_red[w] = v;
    __syncthreads();
    float2 t = make_float2(0.f, 0.f);
    for (uint32_t i = 0; i < nw; ++i) { t.x += s_red[i].x; t.y += s_red[i].y; }
    return t;
}

// subcarrier index of STF cell i (every 4th occupied subcarrier, centre gap of 8; stf.cpp:219-231)
__device__ __forceinline__ uint32_t stf_cell_k(uint32_t i, uint32_t N_OCC) {
    const uint32_t NS = N_OCC / 4;
    return i < NS / 2 ? 4 * i : N_OCC / 2 + 4 + 4 * (i - NS / 2);
}

// FFT bin of subcarrier k (0 .. N_OCC, DC at N_OCC/2), inverse of the TX mapping
__device__ __forceinline__ uint32_t bin_of_k(uint32_t k, const phy_dims_t& d) {
    return k >= d.N_OCC / 2 ? k - d.N_OCC / 2 : d.N_os - d.N_OCC / 2 + k;
}

// ---- STF stage, one block per packet (rx_synced.cpp:503-709, estimator_sto.cpp:47-173,
// estimator_snr.cpp:48-145)
template <bool BIG>
__global__ void __launch_bounds__(256) k_rx_stf(const rx_params_t p) {
    extern __shared__ __align__(16) float2 sm[];
    const phy_dims_t& d = p.d;
    const uint32_t pkt = blockIdx.x, tid = threadIdx.x, nthr = blockDim.x;
    rx_state_t& st = p.state[pkt];
    const uint32_t NS = d.N_OCC / 4;
    // N_os <= 4096: the derotated STF is staged once (S) and read three times; longer transforms (fft.cuh, four-step in
    // place) have no room for it: the samples are derotated from global memory again where they are needed
    constexpr bool big = BIG;
    const uint32_t Mb = d.N_os / 16;
    float2* S = sm;                      // [stf_len_os] derotated STF with the cover sequence undone (not when big)
    float2* X = big ? sm : S + d.stf_len_os;   // [N_os]
    float2* Yb = X + (big ? d.N_os : fft_buf_len(d.N_os));   // ping-pong partner of X, or the two row scratch buffers
    float2* H = Yb + (big ? 2 * fft_buf_len(Mb) : fft_buf_len(d.N_os));   // [N_RX][NS] zero-forced STF cells
    __shared__ float2 s_red[8];
    __shared__ float s_ang[8];

    const int32_t t0 = p.t0_zero ? 0 : st.t0;
    const float inc0 = st.inc0;
    const float c[9] = {1.f, -1.f, 1.f, 1.f, -1.f, -1.f, -1.f, -1.f, -1.f};
    const float fscale = sqrtf((float)d.N_OCC) / (float)d.N_os;  // rx_synced.cpp:761
    float2 corr = make_float2(0.f, 0.f);

    for (uint32_t a = 0; a < d.N_RX; ++a) {
        const float2* x = p.iq + ((size_t)pkt * d.N_RX + a) * p.iq_stride + t0;
        float2 pw = make_float2(0.f, 0.f);
        float2 w = rot((double)inc0 * (double)tid);
        const float2 ws = rot((double)inc0 * (double)nthr);
        const uint32_t npair = (d.n_pat - 1) * d.pat_len_os;
        float2 cr = make_float2(0.f, 0.f);
        if (big) {
            const float2 wp = rot((double)inc0 * (double)d.pat_len_os);   // one pattern further
            for (uint32_t n = tid; n < d.stf_len_os; n += nthr) {
                const float2 v = cmul(x[n], w);
                pw.x += v.x * v.x + v.y * v.y;
                if (n < npair) {
                    const float cc = c[n / d.pat_len_os] * c[n / d.pat_len_os + 1];
                    const float2 v2 = cmul(x[n + d.pat_len_os], cmul(w, wp));
                    const float2 pr = cmulc(v, v2);
                    cr = cadd(cr, make_float2(pr.x * cc, pr.y * cc));
                }
                w = cmul(w, ws);
            }
        } else {
            for (uint32_t n = tid; n < d.stf_len_os; n += nthr) {
                float2 v = cmul(x[n], w);
                w = cmul(w, ws);
                pw.x += v.x * v.x + v.y * v.y;
                const float cs = c[n / d.pat_len_os];
                S[n] = make_float2(v.x * cs, v.y * cs);
            }
        }
        pw = block_sum2(pw, s_red);
        if (tid == 0 && !(st.rms[a] > 0.f)) st.rms[a] = sqrtf(pw.x / (float)d.stf_len_os);
        // residual CFO: sum over pattern pairs of <pattern j, conj(pattern j+1)> (rx_synced.cpp:536-550)
        if (!big)
            for (uint32_t n = tid; n < npair; n += nthr) cr = cadd(cr, cmulc(S[n], S[n + d.pat_len_os]));
        cr = block_sum2(cr, s_red);
        corr = cadd(corr, cr);
        // FFT of the last N_os samples, un-mirror, scale, zero-forcing at the STF cells
        const float2* F;
        if (big) {
            float2 w2 = rot((double)inc0 * (double)(d.stf_cp_os + tid));
            for (uint32_t n = tid; n < d.N_os; n += nthr) {
                const float cs = c[(d.stf_cp_os + n) / d.pat_len_os];
                const float2 v = cmul(x[d.stf_cp_os + n], w2);
                w2 = cmul(w2, ws);
                X[n] = make_float2(v.x * cs, v.y * cs);
            }
            big_fft<false>(X, Yb, Yb + fft_buf_len(Mb), d.N_os, d.twiddle);
            F = X;   // bin k at big_fft_at(k, Mb)
        } else {
            for (uint32_t n = tid; n < d.N_os; n += nthr) X[n] = S[d.stf_cp_os + n];
            F = block_fft<false>(X, Yb, d.N_os, d.twiddle, d.twiddle_p4);
        }
        for (uint32_t i = tid; i < NS; i += nthr) {
            const uint32_t k = stf_cell_k(i, d.N_OCC);
            const uint32_t bin = bin_of_k(k, d);
            const float2 y = F[big ? big_fft_at(bin, Mb) : bin];
            const float sg = 0.70710678118654752f * (float)d.stf_sign[k];
            // y / (sg + j*sg) = y * conj(ref) / |ref|^2, |ref|^2 = 1
            H[a * NS + i] = cmulc(make_float2(y.x * fscale, y.y * fscale), make_float2(sg, sg));
        }
        __syncthreads();
    }

    // ---- fractional STO from the phase gradient over the STF cells, angles averaged over antennas
    for (uint32_t a = 0; a < d.N_RX; ++a) {
        float2 acc = make_float2(0.f, 0.f);
        for (uint32_t i = tid; i + 1 < NS; i += nthr) {
            float2 pr = cmulc(H[a * NS + i], H[a * NS + i + 1]);
            if (i == NS / 2 - 1) {  // centre pair spans 8 subcarriers: rotate back by half its angle
                const float ang = atan2f(pr.y, pr.x);
                float s_, c_;
                sincosf(-ang * 0.5f, &s_, &c_);
                pr = cmul(pr, make_float2(c_, s_));
            }
            acc = cadd(acc, pr);
        }
        acc = block_sum2(acc, s_red);
        if (tid == 0) s_ang[a] = atan2f(acc.y, acc.x) / 4.0f;
    }
    __syncthreads();
    float sto_inc = 0.f;
    for (uint32_t a = 0; a < d.N_RX; ++a) sto_inc += s_ang[a];
    sto_inc /= (float)d.N_RX;

    // ---- derotate the STF cells (zero rotation at DC), then SNR on the derotated cells
    float2 sn = make_float2(0.f, 0.f);  // (S+N sum, N sum)
    for (uint32_t a = 0; a < d.N_RX; ++a) {
        for (uint32_t i = tid; i < NS; i += nthr) {
            const int k = (int)stf_cell_k(i, d.N_OCC) - (int)(d.N_OCC / 2);
            H[a * NS + i] = cmul(H[a * NS + i], rot((double)sto_inc * (double)k));
        }
        __syncthreads();
        for (uint32_t i = tid; i < NS; i += nthr) {
            const float2 h = H[a * NS + i];
            sn.x += h.x * h.x + h.y * h.y;
            if (i + 1 < NS) {
                const float2 dlt = csub(h, H[a * NS + i + 1]);
                sn.y += 0.5f * (dlt.x * dlt.x + dlt.y * dlt.y);
            }
            p.hstf[((size_t)pkt * d.N_RX + a) * NS + i] = h;
        }
    }
    sn = block_sum2(sn, s_red);
    if (tid == 0) {
        const float cfo_left = atan2f(corr.y, corr.x) / (float)d.pat_len_os;
        st.cfo_left = cfo_left;
        st.inc1 = inc0 + cfo_left;
        st.sto_inc = sto_inc;
        st.sto_fractional = sto_inc / 6.283185307179586f * (float)d.N_os;
        // STF cells are boosted by 6 dB: keep the noise, divide the signal part by 4
        const float S_ = sn.x - sn.y;
        st.SpN_sum = S_ * 0.25f + sn.y;
        st.N_sum = sn.y;
        st.SpN_cnt = d.N_RX * NS;
        st.N_cnt = d.N_RX * (NS - 1);
    }
}

// sign of the DRS cell at (symbol l, subcarrier k): the transmitter's own resource element table
// (spec_re.hpp cell words, bit 26 = negative pilot; clause 5.2.3, reference drs.cpp:60-101)
__device__ __forceinline__ float drs_sign_dev(const phy_dims_t& d, uint32_t l, uint32_t k) {
    return ((d.cell[(size_t)l * d.Nk + k] >> 26) & 1u) ? -1.f : 1.f;
}

// ---- data field: a GROUP of N_os / 16 threads per (symbol l = 1..N_DF, RX antenna), several groups per block, packet in
// gridDim.y: CFO derotation, CP removal, FFT, un-mirroring, amplitude scaling, STO derotation (rx_synced.cpp:711-771);
// DRS symbols also deliver the zero-forced pilots and their SNR sums (rx_synced.cpp:773-861, estimator_snr.cpp:76-86).
// Every thread loads its 16 samples (stride N_os / 16: coalesced across the group) straight into registers, derotates
// them there and runs the first radix-16 pass without touching shared memory (fft.cuh, group_fft16).
// BIG (N_os > 4096): one group of N_os / 16 (up to 1024) threads per block, four-step in-place transform (fft.cuh).
template <bool BIG>
__global__ void __launch_bounds__(BIG ? 1024 : 256) k_rx_fft(const rx_params_t p, uint32_t GT, uint32_t groups) {
    extern __shared__ __align__(16) float2 sm[];
    const phy_dims_t& d = p.d;
    const uint32_t grp = threadIdx.x / GT, tid = threadIdx.x - grp * GT, nthr = GT;
    const uint32_t item = blockIdx.x * groups + grp;          // (symbol, antenna) of this group
    const uint32_t pkt = blockIdx.y;
    const bool active = grp < groups && item < d.N_DF * d.N_RX;
    const uint32_t l = (active ? item / d.N_RX : 0) + 1, a = active ? item % d.N_RX : 0;
    const rx_state_t& st = p.state[pkt];
    // per group: two ping-pong buffers of fft_buf_len(N_os); BIG: the array of N_os, two row scratch buffers and the
    // output symbol (Nk)
    const uint32_t Mb = d.N_os / 16;
    const uint32_t blen = BIG ? (d.N_os + 2 * fft_buf_len(Mb) + d.Nk + 1) / 2 : fft_buf_len(d.N_os);
    float2* A = sm + (size_t)grp * 2 * blen;
    float2* B = BIG ? A + d.N_os : A + blen;
    float* s_part = reinterpret_cast<float*>(sm + (size_t)groups * 2 * blen);   // [blockDim][2] SNR partial sums

    const uint32_t n0 = d.stf_len_os + (l - 1) * d.sym_len_os + d.cp_os;  // relative to the first STF sample
    float2 v[16];
    if (active) {
        const float2* x = p.iq + ((size_t)pkt * d.N_RX + a) * p.iq_stride + (p.t0_zero ? 0 : st.t0) + n0;
        // mixer phase is continuous: inc0 over the STF, inc1 afterwards (rx_synced.cpp:239-245, 552-555); one exact
        // phasor per thread, then 15 float rotations by N_os / 16 samples
        const double ph0 = (double)st.inc0 * (double)d.stf_len_os + (double)st.inc1 * (double)(n0 - d.stf_len_os);
        float2 w = rot(ph0 + (double)st.inc1 * (double)tid);
        const float2 ws = rot((double)st.inc1 * (double)GT);
#pragma unroll
        for (int m = 0; m < 16; ++m) {
            v[m] = cmul(__ldcs(x + tid + m * GT), w);
            w = cmul(w, ws);
        }
    }
    const float2* F;
    if constexpr (BIG) {
        if (active) big_fft_col<false>(v, A, tid, Mb, d.twiddle);   // thread tid holds column n2 = tid
        big_fft_rows<false>(A, B, B + fft_buf_len(Mb), d.N_os, d.twiddle, tid, nthr, active);
        F = A;   // bin k at big_fft_at(k, Mb)
    } else {
        F = group_fft16<false>(v, A, B, d.N_os, d.twiddle, tid, nthr, active, d.twiddle_p16);
    }

    float2 sn = make_float2(0.f, 0.f);
    bool any = false;
    if (active) {
        const float fscale = sqrtf((float)d.N_OCC) / (float)d.N_os;
        float2* Yout = p.Y + (((size_t)pkt * d.N_RX + a) * d.N_DF + (l - 1)) * d.Nk;
        // keep the output symbol in shared for the pilot stage
        float2* Ys = BIG ? B + 2 * fft_buf_len(Mb) : (float2*)(F == A ? B : A);
        float2 wk = rot((double)st.sto_inc * (double)((int)tid - (int)(d.N_OCC / 2)));
        const float2 wks = rot((double)st.sto_inc * (double)nthr);
        wk = make_float2(wk.x * fscale, wk.y * fscale);
        for (uint32_t k = tid; k < d.Nk; k += nthr) {
            const uint32_t bin = bin_of_k(k, d);
            const float2 y = cmul(F[BIG ? big_fft_at(bin, Mb) : bin], wk);
            wk = cmul(wk, wks);
            Yout[k] = y;
            Ys[k] = y;
        }
    }
    __syncthreads();
    if (active) {
        const float2* Ys = BIG ? B + 2 * fft_buf_len(Mb) : (F == A ? B : A);
        // ---- DRS of this symbol (if any): stream ts at occurrence n
        for (uint32_t ts = 0; ts < d.N_eff_TX; ++ts) {
            const uint32_t lt = l - 1 - ts / 4;  // symbols since the first DRS symbol of this stream
            if (l < 1 + ts / 4 || lt % d.N_step != 0) continue;
            const uint32_t n = lt / d.N_step;
            if (n >= d.n_drs) continue;
            any = true;
            const uint32_t pc = ((ts & 3) + 2 * (n & 1)) & 3;
            float2* hz = p.hzf + ((((size_t)pkt * d.N_RX + a) * d.N_eff_TX + ts) * d.n_drs + n) * d.NP;
            for (uint32_t i = tid; i < d.NP; i += nthr) {
                const uint32_t q = 4 * i + pc;
                const uint32_t k = q < d.N_OCC / 2 ? q : q + 1;
                const float sg = drs_sign_dev(d, l, k);
                const float2 h = make_float2(Ys[k].x * sg, Ys[k].y * sg);  // division by +-1
                hz[i] = h;
                sn.x += h.x * h.x + h.y * h.y;
                if (i + 1 < d.NP) {
                    const uint32_t q2 = q + 4, k2 = q2 < d.N_OCC / 2 ? q2 : q2 + 1;
                    const float sg2 = drs_sign_dev(d, l, k2);
                    const float2 dlt = csub(h, make_float2(Ys[k2].x * sg2, Ys[k2].y * sg2));
                    sn.y += 0.5f * (dlt.x * dlt.x + dlt.y * dlt.y);
                }
            }
        }
    }
    // group-wide sum in a fixed order (deterministic): partials to shared memory, thread 0 of the group adds them up
    s_part[2 * threadIdx.x] = sn.x;
    s_part[2 * threadIdx.x + 1] = sn.y;
    __syncthreads();
    if (active && tid == 0) {
        float sx = 0.f, sy = 0.f;
        for (uint32_t i = 0; i < GT; ++i) { sx += s_part[2 * (grp * GT + i)]; sy += s_part[2 * (grp * GT + i) + 1]; }
        p.snr_part[((size_t)pkt * d.N_DF + (l - 1)) * d.N_RX + a] = any ? make_float2(sx, sy) : make_float2(0.f, 0.f);
    }
}

// ---- the same for N_os = 1024 (beta = 16 at the DECT rate: BASELINE configs C2, C3, C4) with every size fixed at
// compile time: 64 threads per (symbol, antenna), four groups per block, passes 16 x 16 x 4 with the data staying in
// registers between the passes -- ONE padded buffer per group instead of a ping-pong pair (read all, group barrier,
// write back), so four blocks instead of three are resident per SM; groups synchronise among themselves through named
// barriers (bar.sync id, 64) and never wait for each other.  The DRS stage reads the symbol back from global memory
// (L1 / L2 hits; only three symbols in twenty carry DRS cells) instead of keeping a copy in shared memory.  Same
// butterflies, twiddles and operation order as group_fft16 (fft.cuh): bit-identical Y; the SNR partial sums are
// reduced in a different (fixed) order.
__device__ __forceinline__ void group_bar(uint32_t grp) {   // literal ids: the kernel reserves five barriers, not sixteen
    switch (grp) {
        case 0: asm volatile("bar.sync 1, 64;" ::: "memory"); break;
        case 1: asm volatile("bar.sync 2, 64;" ::: "memory"); break;
        case 2: asm volatile("bar.sync 3, 64;" ::: "memory"); break;
        default: asm volatile("bar.sync 4, 64;" ::: "memory"); break;
    }
}

__global__ void __launch_bounds__(256, 4) k_rx_fft1024(const rx_params_t p) {
    constexpr uint32_t N = 1024, GT = 64, G = 4, BL = N + N / 16;
    extern __shared__ __align__(16) float2 sm[];   // [G][BL], then [G][2] warp partial sums
    const phy_dims_t& d = p.d;
    const uint32_t grp = threadIdx.x >> 6, tid = threadIdx.x & 63;
    const uint32_t item = blockIdx.x * G + grp, pkt = blockIdx.y;
    if (item >= d.N_DF * d.N_RX) return;   // whole groups leave; only group barriers below
    const uint32_t l = item / d.N_RX + 1, a = item - (l - 1) * d.N_RX;
    const rx_state_t& st = p.state[pkt];
    float2* A = sm + grp * BL;
    const float2* __restrict__ tw16 = d.twiddle_p16;         // [15][16] W_256^(k r)
    const float2* __restrict__ tw4 = d.twiddle_p16 + 240;    // [3][256] W_1024^(j r)

    const uint32_t n0 = d.stf_len_os + (l - 1) * d.sym_len_os + d.cp_os;
    float2 v[16];
    {
        const float2* x = p.iq + ((size_t)pkt * d.N_RX + a) * p.iq_stride + (p.t0_zero ? 0 : st.t0) + n0;
        const double ph0 = (double)st.inc0 * (double)d.stf_len_os + (double)st.inc1 * (double)(n0 - d.stf_len_os);
        float2 w = rot(ph0 + (double)st.inc1 * (double)tid);
        const float2 ws = rot((double)st.inc1 * (double)GT);
#pragma unroll
        for (int m = 0; m < 16; ++m) {
            v[m] = cmul(__ldcs(x + tid + m * GT), w);
            w = cmul(w, ws);
        }
    }
    // pass 1 (Ns = 1): X[r] of butterfly tid at element 16 tid + r (padded 17 tid + r)
    dft16<false>(v);
#pragma unroll
    for (int r = 0; r < 16; ++r) A[17 * tid + r] = v[r];
    group_bar(grp);
    // pass 2 (Ns = 16): inputs at tid + 64 r, twiddle W_256^(k r), outputs at 256 (tid / 16) + k + 16 r
    {
        const uint32_t k = tid & 15u, q = tid >> 4;
        const float2* xj = A + tid + q;
#pragma unroll
        for (int r = 0; r < 16; ++r) v[r] = xj[68 * r];
        group_bar(grp);
#pragma unroll
        for (int r = 1; r < 16; ++r) v[r] = cmul(v[r], __ldg(tw16 + (r - 1) * 16 + k));
        dft16<false>(v);
        float2* yj = A + 272 * q + k;
#pragma unroll
        for (int r = 0; r < 16; ++r) yj[17 * r] = v[r];
    }
    group_bar(grp);
    // pass 3 (Ns = 256, radix 4): butterflies j = tid + 64 i, inputs at j + 256 r (padded), outputs at j + 256 r (plain)
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const uint32_t j = tid + 64 * i;
        const float2* xj = A + j + (j >> 4);
#pragma unroll
        for (int r = 0; r < 4; ++r) v[4 * i + r] = xj[272 * r];
    }
    group_bar(grp);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const uint32_t j = tid + 64 * i;
#pragma unroll
        for (int r = 1; r < 4; ++r) v[4 * i + r] = cmul(v[4 * i + r], __ldg(tw4 + (r - 1) * 256 + j));
        dft4<false>(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
#pragma unroll
        for (int r = 0; r < 4; ++r) A[j + 256 * r] = v[4 * i + r];
    }
    group_bar(grp);

    // un-mirror, scale, STO derotation -> Y
    float2* Yout = p.Y + (((size_t)pkt * d.N_RX + a) * d.N_DF + (l - 1)) * d.Nk;
    {
        const float fscale = sqrtf((float)d.N_OCC) / (float)N;
        float2 wk = rot((double)st.sto_inc * (double)((int)tid - (int)(d.N_OCC / 2)));
        const float2 wks = rot((double)st.sto_inc * (double)GT);
        wk = make_float2(wk.x * fscale, wk.y * fscale);
        const uint32_t half = d.N_OCC / 2;
        for (uint32_t k = tid; k < d.Nk; k += GT) {
            const uint32_t bin = k >= half ? k - half : N - half + k;
            Yout[k] = cmul(A[bin], wk);
            wk = cmul(wk, wks);
        }
    }
    // ---- DRS of this symbol (if any): streams 0..3 at symbols 1 + n N_step, streams 4..7 one symbol later
    const uint32_t q0 = (l - 1) / d.N_step, r0 = (l - 1) - q0 * d.N_step;
    const bool drs_lo = r0 == 0 && q0 < d.n_drs;
    const bool drs_hi = d.N_eff_TX > 4 && r0 == 1 && q0 < d.n_drs;   // (l - 2) % N_step == 0, N_step >= 5
    float2* part = sm + G * BL + 2 * grp;
    if (!(drs_lo || drs_hi)) {
        if (tid == 0) p.snr_part[((size_t)pkt * d.N_DF + (l - 1)) * d.N_RX + a] = make_float2(0.f, 0.f);
        return;
    }
    group_bar(grp);   // the group's stores to Yout are visible to the group
    float2 sn = make_float2(0.f, 0.f);
    const uint32_t ts0 = drs_lo ? 0u : 4u, ts1 = drs_lo ? min(d.N_eff_TX, 4u) : d.N_eff_TX;
    const uint32_t n = q0;
    const volatile float2* Yv = Yout;
    for (uint32_t ts = ts0; ts < ts1; ++ts) {
        const uint32_t pc = ((ts & 3) + 2 * (n & 1)) & 3;
        float2* hz = p.hzf + ((((size_t)pkt * d.N_RX + a) * d.N_eff_TX + ts) * d.n_drs + n) * d.NP;
        for (uint32_t i = tid; i < d.NP; i += GT) {
            const uint32_t q = 4 * i + pc;
            const uint32_t k = q < d.N_OCC / 2 ? q : q + 1;
            const float sg = drs_sign_dev(d, l, k);
            const float yx = Yv[k].x, yy = Yv[k].y;
            const float2 h = make_float2(yx * sg, yy * sg);  // division by +-1
            hz[i] = h;
            sn.x += h.x * h.x + h.y * h.y;
            if (i + 1 < d.NP) {
                const uint32_t q2 = q + 4, k2 = q2 < d.N_OCC / 2 ? q2 : q2 + 1;
                const float sg2 = drs_sign_dev(d, l, k2);
                const float2 dlt = csub(h, make_float2(Yv[k2].x * sg2, Yv[k2].y * sg2));
                sn.y += 0.5f * (dlt.x * dlt.x + dlt.y * dlt.y);
            }
        }
    }
    // group-wide sum in a fixed order: butterfly inside each warp, then warp 0 + warp 1
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sn.x += __shfl_xor_sync(0xffffffffu, sn.x, o);
        sn.y += __shfl_xor_sync(0xffffffffu, sn.y, o);
    }
    if ((tid & 31u) == 0) part[tid >> 5] = sn;
    group_bar(grp);
    if (tid == 0)
        p.snr_part[((size_t)pkt * d.N_DF + (l - 1)) * d.N_RX + a] = make_float2(part[0].x + part[1].x, part[0].y + part[1].y);
}

// ---- Wiener class per DRS symbol, one thread per packet: replays estimator_snr_t in symbol order and
// picks the nearest of {-5, 15, 35} dB, later index winning ties (rx_synced.cpp:863-891)
__global__ void k_rx_snr_class(const rx_params_t p, uint32_t n_packets) {
    const uint32_t pkt = blockIdx.x * blockDim.x + threadIdx.x;
    if (pkt >= n_packets) return;
    const phy_dims_t& d = p.d;
    rx_state_t& st = p.state[pkt];
    float SpN = st.SpN_sum, Ns = st.N_sum;
    uint32_t c1 = st.SpN_cnt, c2 = st.N_cnt;
    uint32_t j = 0;
    const uint32_t ts_per_sym = d.N_eff_TX < 4 ? d.N_eff_TX : 4;
    float snr_db = 0.f;
    st.snr_db_pcc = 0.f;
    for (uint32_t l = 1; l <= d.N_DF; ++l) {
        bool is_drs = (l - 1) % d.N_step == 0 && (l - 1) / d.N_step < d.n_drs;
        if (d.N_eff_TX == 8 && l >= 2 && (l - 2) % d.N_step == 0 && (l - 2) / d.N_step < d.n_drs) is_drs = true;
        if (is_drs) {
            for (uint32_t a = 0; a < d.N_RX; ++a) {
                const float2 v = p.snr_part[((size_t)pkt * d.N_DF + (l - 1)) * d.N_RX + a];
                SpN += v.x;
                Ns += v.y;
            }
            c1 += d.N_RX * ts_per_sym * d.NP;
            c2 += d.N_RX * ts_per_sym * (d.NP - 1);
            if (SpN <= 0.f || Ns <= 0.f) snr_db = 0.f;
            else snr_db = 10.0f * log10f(((SpN - Ns) / (float)c1) / (Ns / (float)c2));
            const float ref[3] = {-5.f, 15.f, 35.f};
            uint32_t idx = 0;
            float best = fabsf(snr_db - ref[0]);
            for (uint32_t i = 1; i < 3; ++i) {
                const float s = fabsf(snr_db - ref[i]);
                if (s <= best) { best = s; idx = i; }
            }
            p.cls[(size_t)pkt * p.cls_stride + j] = (uint8_t)idx;
            ++j;
        }
        if (l == d.pcc_l_max) st.snr_db_pcc = snr_db;
    }
    st.snr_db_pdc = snr_db;
}

// ---- residual CFO / STO tracking on the DRS (own specification, DESIGN.md "Residual CFO / STO on the DRS"; the
// reference plans it: rx_synced_param.hpp:156-181, estimator_cfo.cpp:31-70, rx_synced.cpp:839-851, "not functional yet").
// Batched form: one estimate per packet over ALL its DRS symbols instead of a feedback loop from symbol to symbol.
//   pilot i of occurrence n + 1 lies midway between two pilots of occurrence n (interlaced by two subcarriers):
//               P_i = h_{n+1}[i] conj((h_n[i'] + h_n[i' + 1]) / 2)   has the phase dphi + (k_i - k_DC) dpsi
//   STO drift:  dpsi = arg(sum P_{i+1} conj(P_i)) / 4   (the pair across DC, 5 subcarriers apart, left out)
//   CFO:        dphi = arg(sum P_i e^{-j (k_i - k_DC) dpsi})          (sums over antennas, streams, occurrences, pilots)
// both are phase advances per N_step OFDM symbols.  Cell (l, k) is taken back to the phase reference of the first DRS
// symbol by e^{-j f (dphi + (k - k_DC) dpsi)}, f = (l - 1) / N_step: the pilots here in place, Y where it is loaded.
__device__ __forceinline__ float2 track_phasor(const phy_dims_t& d, float2 tr, uint32_t l, uint32_t k) {
    const float f = (float)(l - 1) / (float)d.N_step;
    float turns = -f * (tr.x + (float)((int)k - (int)(d.N_OCC / 2)) * tr.y) * 0.15915494309189535f;
    turns -= rintf(turns);
    float sn, cs;
    __sincosf(6.283185307179586f * turns, &sn, &cs);
    return make_float2(cs, sn);
}

__global__ void __launch_bounds__(128) k_rx_track(const rx_params_t p) {
    __shared__ float2 s_red[4];
    const phy_dims_t& d = p.d;
    const uint32_t pkt = blockIdx.x, tid = threadIdx.x;
    const uint32_t stride_a = d.N_eff_TX * d.n_drs * d.NP;
    float2* hz_pkt = p.hzf + (size_t)pkt * d.N_RX * stride_a;
    float2 tr = make_float2(0.f, 0.f);
    if (d.n_drs >= 2) {
        // P(n, ts, a, i) = h_{n+1}[i] conj(midpoint of its two neighbours in occurrence n): the channel cancels, the phase
        // that is left is dphi + (k - k_DC) dpsi
        auto P = [&](const float2* h0, const float2* h1, uint32_t i, uint32_t lo) {
            const float2 ref = make_float2(0.5f * (h0[i - lo].x + h0[i - lo + 1].x), 0.5f * (h0[i - lo].y + h0[i - lo + 1].y));
            return cmulc(h1[i], ref);
        };
        for (int pass = p.track_sto ? 0 : 1; pass < (p.track_cfo ? 2 : 1); ++pass) {
            float2 Z = make_float2(0.f, 0.f);
            for (uint32_t n = 0; n + 1 < d.n_drs; ++n)
                for (uint32_t ts = 0; ts < d.N_eff_TX; ++ts) {
                    const uint32_t pc = ((ts & 3) + 2 * (n & 1)) & 3, pc1 = (pc + 2) & 3;   // occurrences n, n + 1
                    const uint32_t lo = pc < 2 ? 0u : 1u;          // pilot i of n + 1 between i - lo and i - lo + 1 of n
                    for (uint32_t a = 0; a < d.N_RX; ++a) {
                        const float2* h0 = hz_pkt + a * stride_a + (ts * d.n_drs + n) * d.NP;
                        const float2* h1 = h0 + d.NP;
                        if (pass == 0) {
                            // slope of the phase of P over the subcarriers: neighbours 4 apart (5 across DC: left out)
                            for (uint32_t i = tid + lo; i + 2 - lo < d.NP; i += blockDim.x) {
                                const uint32_t q = 4 * i + pc1;
                                if (q < d.N_OCC / 2 && q + 4 >= d.N_OCC / 2) continue;
                                Z = cadd(Z, cmulc(P(h0, h1, i + 1, lo), P(h0, h1, i, lo)));
                            }
                        } else {
                            for (uint32_t i = tid + lo; i + 1 - lo < d.NP; i += blockDim.x) {
                                float2 z = P(h0, h1, i, lo);
                                if (p.track_sto) {
                                    const uint32_t q = 4 * i + pc1, k = q < d.N_OCC / 2 ? q : q + 1;
                                    z = cmul(z, track_phasor(d, make_float2(0.f, tr.y), 1 + d.N_step, k));
                                }
                                Z = cadd(Z, z);
                            }
                        }
                    }
                }
            Z = block_sum2(Z, s_red);
            const float ang = (Z.x != 0.f || Z.y != 0.f) ? atan2f(Z.y, Z.x) : 0.f;
            if (pass == 0) tr.y = 0.25f * ang; else tr.x = ang;
        }
    }
    if (tid == 0) p.track[pkt] = tr;
    if (tr.x == 0.f && tr.y == 0.f) return;
    // the pilots back to the phase reference of the first DRS symbol (streams 4..7 sit one symbol later)
    for (uint32_t a = 0; a < d.N_RX; ++a)
        for (uint32_t ts = 0; ts < d.N_eff_TX; ++ts)
            for (uint32_t n = 0; n < d.n_drs; ++n) {
                const uint32_t pc = ((ts & 3) + 2 * (n & 1)) & 3, l = 1 + n * d.N_step + (ts >= 4 ? 1u : 0u);
                float2* h = hz_pkt + a * stride_a + (ts * d.n_drs + n) * d.NP;
                for (uint32_t i = tid; i < d.NP; i += blockDim.x) {
                    const uint32_t q = 4 * i + pc, k = q < d.N_OCC / 2 ? q : q + 1;
                    h[i] = cmul(h[i], track_phasor(d, tr, l, k));
                }
            }
}

cudaError_t launch_rx_track(const rx_params_t& p, uint32_t n_packets, cudaStream_t st) {
    k_rx_track<<<n_packets, 128, 0, st>>>(p);
    return cudaGetLastError();
}

// ---- channel estimate of transmit stream ts at (symbol l, subcarrier k) seen by RX antenna a: the value
// rx_synced_t::run_drs_ch_interpolation (rx_synced.cpp:893-960) holds in chestim[ts][k] at the moment
// the reference collects the cells of symbol l.  Closed form of the processing stage bookkeeping of
// rx_synced.cpp:1027-1152, see DESIGN.md "Receiver: which interpolation serves which symbol".
// All NRX receive antennas at once: h[a] = estimate seen by antenna a.  Symbol plan, LUT indices and the weights are
// the same for every antenna, so they are computed / loaded once and every tap is applied to the NRX pilot sets.
template <int NRX>
__device__ __forceinline__ void chan_est_rx(const rx_params_t& p, uint32_t pkt, uint32_t plan, uint32_t k, float2 (&h)[NRX]) {
    const phy_dims_t& d = p.d;
    // plan word (plan_phy.cu, ce_plan_word): the symbol / stream bookkeeping was done on the host
    const uint32_t base = plan & 63u, j = (plan >> 6) & 63u, row = (plan >> 12) & 31u, flip = (plan >> 17) & 1u,
                   ts = (plan >> 18) & 7u, lr = (plan >> 21) & 1u;
    const phy_dims_t::wiener_dev_t& W = lr ? d.wiener_lr[p.cls[(size_t)pkt * p.cls_stride + j]]
                                           : d.wiener_l[p.cls[(size_t)pkt * p.cls_stride + j]];
    // all offsets inside one packet's pilots fit 32 bits: one 64-bit base per packet, 32-bit element offsets per tap
    const uint32_t stride_a = d.N_eff_TX * d.n_drs * d.NP;          // pilots of the next antenna
    const float2* hz_pkt = p.hzf + (size_t)pkt * d.N_RX * stride_a;
    const uint32_t off_row = (ts * d.n_drs + row) * d.NP;
    const uint32_t o = base * d.Nk + k;
    const float* w = W.weights + (uint32_t)W.idx_weight[o] * W.n_interp;
    const uint32_t m0 = W.idx_pilot[o];
    const uint32_t n_interp = W.n_interp;
#pragma unroll
    for (int a = 0; a < NRX; ++a) h[a] = make_float2(0.f, 0.f);
    if (!lr) {
        // mode "l": non-interlaced pilots of DRS occurrence row
        const float2* hz = hz_pkt + off_row + m0;
        for (uint32_t i = 0; i < n_interp; ++i) {
            const float wi = w[i];
#pragma unroll
            for (int a = 0; a < NRX; ++a) {
                const float2 v = hz[a * stride_a + i];
                h[a].x += wi * v.x;
                h[a].y += wi * v.y;
            }
        }
    } else {
        // mode "lr": pilots interlaced over DRS occurrences row (left) and row + 1 (right): tap m sits on side
        // (m & 1) ^ flip at index m >> 1.  Two running offsets, one per parity of m.
        uint32_t off_e = off_row + (flip ? d.NP : 0u) + ((m0 + 1) >> 1);      // next even m
        uint32_t off_o = off_row + (flip ? 0u : d.NP) + (m0 >> 1);            // next odd m
        bool odd = (m0 & 1u) != 0u;
        for (uint32_t i = 0; i < n_interp; ++i) {
            const uint32_t off = odd ? off_o : off_e;
            if (odd) ++off_o; else ++off_e;
            odd = !odd;
            const float wi = w[i];
#pragma unroll
            for (int a = 0; a < NRX; ++a) {
                const float2 v = hz_pkt[off + a * stride_a];
                h[a].x += wi * v.x;
                h[a].y += wi * v.y;
            }
        }
    }
}

// The transmit diversity combiner uses, per stream, the MEAN of the estimates at the two subcarriers of a cell pair
// (rx_synced.cpp:1360-1369).  Both Wiener windows cover the same pilots or are shifted by one pilot against each other
// (the subcarriers are at most two apart, pilots four), so the mean is ONE dot product over the union window with the
// weights 0.5 (w_k0 + w_k1): n_interp + 1 taps instead of 2 n_interp, i.e. about half the pilot loads and MACs.
// Differs from averaging two separately rounded estimates by fp32 rounding only.
// pair_luts: the host-merged windows of this list (phy_dims_t::pair_lut_t, per SNR class), slot = stream slot (A / B) of
// unit u: first pilot, tap count and weights come straight from the tables (one coalesced load per tap).
template <int NRX>
__device__ __forceinline__ void chan_est_rx_pair(const rx_params_t& p, uint32_t pkt, uint32_t plan, uint32_t k0, uint32_t k1,
                                                 float2 (&h)[NRX], const phy_dims_t::pair_lut_t* pair_luts, uint32_t u,
                                                 uint32_t slot) {
    const phy_dims_t& d = p.d;
    const uint32_t base = plan & 63u, j = (plan >> 6) & 63u, row = (plan >> 12) & 31u, flip = (plan >> 17) & 1u,
                   ts = (plan >> 18) & 7u, lr = (plan >> 21) & 1u;
    {
        const phy_dims_t::pair_lut_t& PL = pair_luts[p.cls[(size_t)pkt * p.cls_stride + j]];
        const uint32_t info = PL.info ? PL.info[slot * PL.units + u] : 0xffffu;
        if (info != 0xffffu) {
            const uint32_t mmin = info & 1023u, nt = info >> 10;
            const float* wp = PL.w + slot * PL.units + u;
            const uint32_t wstride = 2 * PL.units;
            const uint32_t stride_a = d.N_eff_TX * d.n_drs * d.NP;
            const float2* hz_pkt = p.hzf + (size_t)pkt * d.N_RX * stride_a;
            const uint32_t off_row = (ts * d.n_drs + row) * d.NP;
#pragma unroll
            for (int a = 0; a < NRX; ++a) h[a] = make_float2(0.f, 0.f);
            if (lr) {
                // tap m sits on side (m & 1) ^ flip at index m >> 1: the taps alternate between two rows of pilots
                uint32_t off_a = off_row + (flip ? d.NP : 0u) + ((mmin + 1) >> 1);      // even m
                uint32_t off_b = off_row + (flip ? 0u : d.NP) + (mmin >> 1);            // odd m
                if (mmin & 1u) { const uint32_t t_ = off_a; off_a = off_b; off_b = t_; }   // off_a: taps 0, 2, ...
                for (uint32_t t = 0; t < nt; ++t) {
                    const uint32_t off = (t & 1u) ? off_b + (t >> 1) : off_a + (t >> 1);
                    const float wi = wp[t * wstride];
#pragma unroll
                    for (int a = 0; a < NRX; ++a) {
                        const float2 v = hz_pkt[off + a * stride_a];
                        h[a].x += wi * v.x;
                        h[a].y += wi * v.y;
                    }
                }
            } else {
                const float2* hz = hz_pkt + off_row + mmin;
                for (uint32_t t = 0; t < nt; ++t) {
                    const float wi = wp[t * wstride];
#pragma unroll
                    for (int a = 0; a < NRX; ++a) {
                        const float2 v = hz[t + a * stride_a];
                        h[a].x += wi * v.x;
                        h[a].y += wi * v.y;
                    }
                }
            }
            return;
        }
    }
    const phy_dims_t::wiener_dev_t& W = lr ? d.wiener_lr[p.cls[(size_t)pkt * p.cls_stride + j]]
                                           : d.wiener_l[p.cls[(size_t)pkt * p.cls_stride + j]];
    const uint32_t o0 = base * d.Nk + k0, o1 = base * d.Nk + k1;
    const uint32_t m0 = W.idx_pilot[o0], m1 = W.idx_pilot[o1];
    const uint32_t mmin = min(m0, m1), d0 = m0 - mmin, d1 = m1 - mmin;
    if (d0 + d1 > 1u) {   // windows further apart (not with the DECT pilot grid): the plain way
        float2 t[NRX];
        chan_est_rx<NRX>(p, pkt, plan, k0, h);
        chan_est_rx<NRX>(p, pkt, plan, k1, t);
#pragma unroll
        for (int a = 0; a < NRX; ++a) h[a] = make_float2(0.5f * (h[a].x + t[a].x), 0.5f * (h[a].y + t[a].y));
        return;
    }
    const uint32_t n_interp = W.n_interp, nt = n_interp + d0 + d1;
    const float* w0 = W.weights + (uint32_t)W.idx_weight[o0] * n_interp - d0;   // tap t of the union uses w0[t - d0]
    const float* w1 = W.weights + (uint32_t)W.idx_weight[o1] * n_interp - d1;
    const uint32_t stride_a = d.N_eff_TX * d.n_drs * d.NP;
    const float2* hz_pkt = p.hzf + (size_t)pkt * d.N_RX * stride_a;
    const uint32_t off_row = (ts * d.n_drs + row) * d.NP;
#pragma unroll
    for (int a = 0; a < NRX; ++a) h[a] = make_float2(0.f, 0.f);
    // lr: tap m sits on side (m & 1) ^ flip at index m >> 1 (two running offsets); l: plain index
    uint32_t off_e = off_row + (lr ? (flip ? d.NP : 0u) + ((mmin + 1) >> 1) : 0u);
    uint32_t off_o = off_row + (lr ? (flip ? 0u : d.NP) + (mmin >> 1) : 0u);
    uint32_t off_l = off_row + mmin;
    bool odd = (mmin & 1u) != 0u;
    for (uint32_t t = 0; t < nt; ++t) {
        uint32_t off;
        if (lr) {
            off = odd ? off_o : off_e;
            if (odd) ++off_o; else ++off_e;
            odd = !odd;
        } else {
            off = off_l++;
        }
        const float a0 = (t >= d0 && t - d0 < n_interp) ? w0[t] : 0.f;
        const float a1 = (t >= d1 && t - d1 < n_interp) ? w1[t] : 0.f;
        const float wi = 0.5f * (a0 + a1);
#pragma unroll
        for (int a = 0; a < NRX; ++a) {
            const float2 v = hz_pkt[off + a * stride_a];
            h[a].x += wi * v.x;
            h[a].y += wi * v.y;
        }
    }
}

// float -> int16 like a C cast (toward zero), saturating
__device__ __forceinline__ float to_short(float v) {
    v = truncf(v);
    return fminf(fmaxf(v, -32768.f), 32767.f);
}

// soft demapper to int16 LLRs, LLR > 0 <=> bit 1 (srsRAN demod_soft_demodulate_s restated, DESIGN.md
// "Demapper"; called by rx_synced.cpp:997, 1287, 1297, 1322).  The N_bps LLRs of the cell come back packed two per
// word, first LLR in the low half (BPSK: one LLR in the low half of w[0]); no dynamically indexed storage.
__device__ __forceinline__ uint32_t pack_llr(float lo, float hi) {
    return ((uint32_t)(uint16_t)(int16_t)lo) | ((uint32_t)(uint16_t)(int16_t)hi << 16);
}
__device__ __forceinline__ void demap_cell(float2 z, uint32_t N_bps, uint32_t (&w)[5]) {
    w[1] = w[2] = w[3] = w[4] = 0u;
    if (N_bps == 1) {
        w[0] = pack_llr(to_short(-(100.f * (z.x + z.y) * 0.70710678118654752f)), 0.f);
        return;
    }
    if (N_bps == 2) {
        w[0] = pack_llr(to_short(z.x * (-100.f * 1.41421356237309505f)), to_short(z.y * (-100.f * 1.41421356237309505f)));
        return;
    }
    // 1024-QAM (MCS 10, 11) is outside srsRAN and the reference (tx.cpp:53-60): the same recursion one level deeper,
    // scale 1300, first threshold 16 / sqrt(682) -- own extension, DESIGN.md "1024-QAM"
    const float S = N_bps == 4 ? 400.f : (N_bps == 6 ? 700.f : (N_bps == 8 ? 1000.f : 1300.f));
    const float nrm = N_bps == 4 ? 0.31622776601683794f
                                 : (N_bps == 6 ? 0.15430334996209191f : (N_bps == 8 ? 0.07669649888473704f : 0.03829197905337418f));
    const float yre = to_short(S * z.x), yim = to_short(S * z.y);
    // -(-32768) saturates
    w[0] = pack_llr(yre == -32768.f ? 32767.f : -yre, yim == -32768.f ? 32767.f : -yim);
    float th = (float)(1u << (N_bps / 2 - 1)) * S * nrm;
    float a0 = to_short(fabsf(yre) - th), a1 = to_short(fabsf(yim) - th);
    w[1] = pack_llr(a0, a1);
#pragma unroll
    for (uint32_t i = 2; i < 5; ++i) {
        if (N_bps >= 2 * i + 2) {
            th *= 0.5f;
            a0 = to_short(fabsf(a0) - th);
            a1 = to_short(fabsf(a1) - th);
            w[i] = pack_llr(a0, a1);
        }
    }
}

// ---- fused channel interpolation + MRC / Alamouti combining + demapping, one thread per cell (single
// stream) or per cell pair (transmit diversity).  grid (ceil(units / 128), packets).  A block's LLRs are one
// contiguous span of the d-stream: they are collected in shared memory and stored as coalesced 32-bit words
// (N_bps even; BPSK stores its int16 values directly).
template <bool TXDIV, int NRX>
__global__ void __launch_bounds__(128, TXDIV ? (NRX > 2 ? 6 : 8) : 10) k_rx_demap(const rx_params_t p, uint32_t pcc) {
    __shared__ uint32_t s_out[128 * 10];
    const phy_dims_t& d = p.d;
    const uint32_t pkt = blockIdx.y;
    const uint32_t n_cells = pcc ? 98u : d.N_PDC_subc;
    const uint32_t N_bps = pcc ? 2u : d.N_bps;
    const uint32_t* lk = pcc ? d.pcc_lk : d.pdc_lk;
    const uint32_t* plan = pcc ? d.pcc_plan : d.pdc_plan;
    int16_t* out = pcc ? p.llr_pcc + (size_t)pkt * 196 : p.llr + (size_t)pkt * p.llr_stride;
    const uint32_t u = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t cells_per_unit = TXDIV ? 2u : 1u;
    const uint32_t n_units = n_cells / cells_per_unit;
    const uint32_t n_llr = cells_per_unit * N_bps;  // LLRs per thread
    uint32_t w0[5] = {0u, 0u, 0u, 0u, 0u}, w1[5] = {0u, 0u, 0u, 0u, 0u};
    if (u < n_units) {
        if (!TXDIV) {
            const uint32_t lkv = lk[u];
            const uint32_t l = lkv >> 16, k = lkv & 0xffffu;
            float2 num = make_float2(0.f, 0.f);
            float den = 0.f;
            float2 h[NRX];
            chan_est_rx<NRX>(p, pkt, plan[2 * u], k, h);
            const bool trk = p.track != nullptr;
            const float2 c0 = trk ? track_phasor(d, p.track[pkt], l, k) : make_float2(1.f, 0.f);
#pragma unroll
            for (int a = 0; a < NRX; ++a) {
                // streaming loads / stores (evict first): Y and the LLRs pass through once, the pilots stay in L1
                float2 y = __ldcs(&p.Y[(((size_t)pkt * NRX + a) * d.N_DF + (l - 1)) * d.Nk + k]);
                if (trk) y = cmul(y, c0);
                num = cadd(num, cmulc(y, h[a]));
                den += h[a].x * h[a].x + h[a].y * h[a].y;
            }
            demap_cell(make_float2(num.x / den, num.y / den), N_bps, w0);
        } else {
            const uint32_t q = 2 * u;
            const uint32_t l = lk[q] >> 16, k0 = lk[q] & 0xffffu, k1 = lk[q + 1] & 0xffffu;
            const uint32_t planA = plan[2 * u], planB = plan[2 * u + 1];   // streams A, B of this pair (Table 6.3.3.2)
            float2 n0 = make_float2(0.f, 0.f), n1 = n0;
            float den = 0.f;
            // per antenna: h0 = mean of stream A's estimates at the two cells, h1 = the same for stream B
            float2 h0[NRX], h1[NRX];
            const phy_dims_t::pair_lut_t* pl = pcc ? d.pcc_pair : d.pdc_pair;
            chan_est_rx_pair<NRX>(p, pkt, planA, k0, k1, h0, pl, u, 0);
            chan_est_rx_pair<NRX>(p, pkt, planB, k0, k1, h1, pl, u, 1);
            const bool trk = p.track != nullptr;
            float2 c0 = make_float2(1.f, 0.f), c1 = c0;
            if (trk) {
                const float2 tr = p.track[pkt];
                c0 = track_phasor(d, tr, l, k0);
                c1 = track_phasor(d, tr, l, k1);
            }
#pragma unroll
            for (int a = 0; a < NRX; ++a) {
                const float2* Y = p.Y + (((size_t)pkt * NRX + a) * d.N_DF + (l - 1)) * d.Nk;
                float2 r0 = __ldcs(&Y[k0]), r1 = __ldcs(&Y[k1]);
                if (trk) { r0 = cmul(r0, c0); r1 = cmul(r1, c1); }
                // conj(h0) r0 + h1 conj(r1)   and   -h1 conj(r0) + conj(h0) r1   (rx_synced.cpp:1372-1373)
                n0 = cadd(n0, cadd(cmulc(r0, h0[a]), cmulc(h1[a], r1)));
                n1 = cadd(n1, csub(cmulc(r1, h0[a]), cmulc(h1[a], r0)));
                den += h0[a].x * h0[a].x + h0[a].y * h0[a].y + h1[a].x * h1[a].x + h1[a].y * h1[a].y;
            }
            demap_cell(make_float2(n0.x / den, n0.y / den), N_bps, w0);
            demap_cell(make_float2(n1.x / den, n1.y / den), N_bps, w1);
        }
    }
    if (n_llr & 1) {  // BPSK, single stream: one int16 per cell
        if (u < n_units) out[u] = (int16_t)(w0[0] & 0xffffu);
        return;
    }
    const uint32_t W = n_llr / 2;  // words per thread: 1 .. 10
    if (u < n_units) {
        uint32_t* so = s_out + threadIdx.x * W;
        if (TXDIV && N_bps == 1) {
            so[0] = (w0[0] & 0xffffu) | (w1[0] << 16);
        } else {
            const uint32_t wc = N_bps / 2;  // words per cell
#pragma unroll
            for (uint32_t i = 0; i < 5; ++i) {
                if (i < wc) so[i] = w0[i];
                if (TXDIV && i < wc) so[wc + i] = w1[i];
            }
        }
    }
    __syncthreads();
    const uint32_t u0 = blockIdx.x * blockDim.x;
    const uint32_t n_here = min(blockDim.x, n_units > u0 ? n_units - u0 : 0u);
    uint32_t* out32 = reinterpret_cast<uint32_t*>(out) + (size_t)u0 * W;  // G even: 4-byte aligned
    for (uint32_t i = threadIdx.x; i < n_here * W; i += blockDim.x) __stcs(&out32[i], s_out[i]);
}

// ---- spatial multiplexing (transmission modes 2, 4, 6, 8, 9, 11).  The reference stops at a stub
// (rx_synced.cpp:1331-1333 run_pdc_mode_AxA_MIMO), so the specification is this repo's own, DESIGN.md "Spatial
// multiplexing receive", restated in oracle/rx.c detect_sm / finish_sm and oracle/rx_mimo.py:
//   per cell: A = H^H H, b = H^H y, Cholesky A = L L^H (pivot floor 1e-6 A_jj), M = L^-1, x = M^H M b (zero forcing),
//             rho_s = min(1, 1 / ([A^-1]_ss A_ss)) = loss of stream s against its matched filter bound;
//   per packet: rho_ref = mean of rho over the cells 0, 8, 16, ... (all streams);
//   LLRs of (cell q, stream s) at (q N_SS + s) N_bps = trunc(min(rho_s / rho_ref, 4) * demap(x_s)).
// Everything of one cell lives in the registers of one thread (8 x 8 spills to local memory, a rare mode).
template <int NSS, int NRX>
__device__ __forceinline__ void sm_solve(const rx_params_t& p, uint32_t pkt, uint32_t q, float2 (&x)[NSS], float (&rho)[NSS]) {
    const phy_dims_t& d = p.d;
    const uint32_t lkv = d.pdc_lk[q], l = lkv >> 16, k = lkv & 0xffffu;
    float2 H[NSS][NRX];   // H[s][a]: stream s seen by antenna a
#pragma unroll
    for (int s = 0; s < NSS; ++s) chan_est_rx<NRX>(p, pkt, d.pdc_plan[q * NSS + s], k, H[s]);
    float2 y[NRX];
#pragma unroll
    for (int a = 0; a < NRX; ++a) y[a] = __ldcs(&p.Y[(((size_t)pkt * NRX + a) * d.N_DF + (l - 1)) * d.Nk + k]);
    if (p.track) {
        const float2 c0 = track_phasor(d, p.track[pkt], l, k);
#pragma unroll
        for (int a = 0; a < NRX; ++a) y[a] = cmul(y[a], c0);
    }
    float2 L[NSS][NSS], b[NSS];   // lower triangle; after the factorisation the diagonal holds 1 / L_jj in .x
    float g[NSS];
#pragma unroll
    for (int i = 0; i < NSS; ++i) {
#pragma unroll
        for (int j = 0; j <= i; ++j) {
            float2 acc = make_float2(0.f, 0.f);
#pragma unroll
            for (int a = 0; a < NRX; ++a) acc = cadd(acc, cmulc(H[j][a], H[i][a]));   // conj(H[a][i]) H[a][j]
            L[i][j] = acc;
        }
        float2 acc = make_float2(0.f, 0.f);
#pragma unroll
        for (int a = 0; a < NRX; ++a) acc = cadd(acc, cmulc(y[a], H[i][a]));
        b[i] = acc;
        g[i] = L[i][i].x;
    }
#pragma unroll
    for (int j = 0; j < NSS; ++j) {
        float dj = L[j][j].x;
#pragma unroll
        for (int c = 0; c < j; ++c) dj -= L[j][c].x * L[j][c].x + L[j][c].y * L[j][c].y;
        const float fl = 1.0e-6f * g[j] + 1.0e-30f;
        if (!(dj > fl)) dj = fl;
        const float inv = 1.0f / sqrtf(dj);
        L[j][j] = make_float2(inv, 0.f);
#pragma unroll
        for (int i = j + 1; i < NSS; ++i) {
            float2 acc = L[i][j];
#pragma unroll
            for (int c = 0; c < j; ++c) acc = csub(acc, cmulc(L[i][c], L[j][c]));
            L[i][j] = make_float2(acc.x * inv, acc.y * inv);
        }
    }
    // M = L^-1 column by column, kept in the upper triangle of the same array: M(i, j) -> L[j][i] for i > j
    float2 z[NSS];
#pragma unroll
    for (int j = 0; j < NSS; ++j) {
#pragma unroll
        for (int i = j + 1; i < NSS; ++i) {
            float2 acc = make_float2(0.f, 0.f);
#pragma unroll
            for (int c = j; c < i; ++c) acc = cadd(acc, cmul(L[i][c], c == j ? L[j][j] : L[j][c]));
            const float s = -L[i][i].x;
            L[j][i] = make_float2(acc.x * s, acc.y * s);
        }
    }
#pragma unroll
    for (int i = 0; i < NSS; ++i) {
        float2 acc = make_float2(0.f, 0.f);
#pragma unroll
        for (int c = 0; c <= i; ++c) acc = cadd(acc, cmul(c == i ? L[i][i] : L[c][i], b[c]));
        z[i] = acc;
    }
#pragma unroll
    for (int i = 0; i < NSS; ++i) {
        float2 acc = make_float2(0.f, 0.f);
        float v = 0.f;
#pragma unroll
        for (int c = i; c < NSS; ++c) {
            const float2 m = c == i ? L[i][i] : L[i][c];   // M(c, i)
            acc = cadd(acc, cmulc(z[c], m));
            v += m.x * m.x + m.y * m.y;
        }
        x[i] = acc;
        const float r = 1.0f / (v * g[i] + 1.0e-30f);
        rho[i] = r < 1.0f ? r : 1.0f;
    }
}

// reference loss of every packet: one block per packet over the cells 0, 8, 16, ...; fixed summation tree
template <int NSS, int NRX>
__global__ void __launch_bounds__(128) k_rx_sm_ref(const rx_params_t p) {
    __shared__ float2 s_red[4];
    const uint32_t pkt = blockIdx.x;
    const uint32_t n_sub = (p.d.N_PDC_subc + 7) / 8;
    float sum = 0.f;
    for (uint32_t i = threadIdx.x; i < n_sub; i += blockDim.x) {
        float2 x[NSS];
        float rho[NSS];
        sm_solve<NSS, NRX>(p, pkt, 8 * i, x, rho);
#pragma unroll
        for (int s = 0; s < NSS; ++s) sum += rho[s];
    }
    const float2 t = block_sum2(make_float2(sum, 0.f), s_red);
    if (threadIdx.x == 0) p.state[pkt].sm_rho_ref = t.x / (float)(n_sub * NSS);
}

template <int NSS, int NRX>
__global__ void __launch_bounds__(128) k_rx_demap_sm(const rx_params_t p) {
    extern __shared__ uint32_t s_sm[];   // [128][NSS * N_bps / 2] packed LLR pairs
    const phy_dims_t& d = p.d;
    const uint32_t pkt = blockIdx.y, N_bps = d.N_bps;
    const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t W = NSS * N_bps / 2;   // words per cell (N_SS is even)
    if (q < d.N_PDC_subc) {
        float2 x[NSS];
        float rho[NSS];
        sm_solve<NSS, NRX>(p, pkt, q, x, rho);
        const float ref = p.state[pkt].sm_rho_ref;
        const float inv = 1.0f / (ref > 1.0e-12f ? ref : 1.0e-12f);
        int16_t* so = reinterpret_cast<int16_t*>(s_sm + threadIdx.x * W);
#pragma unroll
        for (int s = 0; s < NSS; ++s) {
            uint32_t w[5];
            demap_cell(x[s], N_bps, w);
            float wt = rho[s] * inv;
            wt = wt < 4.0f ? wt : 4.0f;
#pragma unroll
            for (uint32_t e = 0; e < 10; ++e)
                if (e < N_bps) {
                    const int16_t raw = (int16_t)((w[e >> 1] >> (16 * (e & 1))) & 0xffffu);
                    so[s * N_bps + e] = (int16_t)to_short(wt * (float)raw);
                }
        }
    }
    __syncthreads();
    const uint32_t q0 = blockIdx.x * blockDim.x;
    const uint32_t n_here = min(blockDim.x, d.N_PDC_subc > q0 ? d.N_PDC_subc - q0 : 0u);
    uint32_t* out32 = reinterpret_cast<uint32_t*>(p.llr + (size_t)pkt * p.llr_stride) + (size_t)q0 * W;
    for (uint32_t i = threadIdx.x; i < n_here * W; i += blockDim.x) __stcs(&out32[i], s_sm[i]);
}

template <int NSS, int NRX>
static cudaError_t launch_sm(const rx_params_t& p, uint32_t n_packets, cudaStream_t st) {
    k_rx_sm_ref<NSS, NRX><<<n_packets, 128, 0, st>>>(p);
    const size_t smem = 128 * (size_t)(NSS * p.d.N_bps / 2) * sizeof(uint32_t);   // at most 20 KiB
    k_rx_demap_sm<NSS, NRX><<<dim3((p.d.N_PDC_subc + 127) / 128, n_packets), 128, smem, st>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_rx_demap(const rx_params_t& p, uint32_t n_packets, bool pcc, cudaStream_t st) {
    if (!pcc && p.d.N_SS > 1) {
        // more streams than receive antennas cannot be separated; N_RX = N_TX of the largest packet in the reference
        // (rx_synced.cpp:74), so this only triggers on a hand-made configuration
        if (p.d.N_RX < p.d.N_SS) return cudaErrorInvalidValue;
        switch (p.d.N_SS * 16 + p.d.N_RX) {
            case 2 * 16 + 2: return launch_sm<2, 2>(p, n_packets, st);
            case 2 * 16 + 4: return launch_sm<2, 4>(p, n_packets, st);
            case 2 * 16 + 8: return launch_sm<2, 8>(p, n_packets, st);
            case 4 * 16 + 4: return launch_sm<4, 4>(p, n_packets, st);
            case 4 * 16 + 8: return launch_sm<4, 8>(p, n_packets, st);
            case 8 * 16 + 8: return launch_sm<8, 8>(p, n_packets, st);
            default: return cudaErrorInvalidValue;
        }
    }
    const uint32_t n_cells = pcc ? 98u : p.d.N_PDC_subc;
    const uint32_t units = p.d.N_eff_TX == 1 ? n_cells : n_cells / 2;
    const dim3 grid((units + 127) / 128, n_packets);
    const uint32_t pf = pcc ? 1u : 0u;
    const bool txdiv = p.d.N_eff_TX != 1;
#define DNRP_DEMAP(NRX)                                                          \
    case NRX:                                                                    \
        if (txdiv) k_rx_demap<true, NRX><<<grid, 128, 0, st>>>(p, pf);           \
        else k_rx_demap<false, NRX><<<grid, 128, 0, st>>>(p, pf);                \
        break;
    switch (p.d.N_RX) {
        DNRP_DEMAP(1) DNRP_DEMAP(2) DNRP_DEMAP(4) DNRP_DEMAP(8)
        default: return cudaErrorInvalidValue;   // N_RX is 1, 2, 4 or 8 (radio device classes)
    }
#undef DNRP_DEMAP
    return cudaGetLastError();
}

// ---- MIMO report: codebook search for single-stream beamforming (modes 3 and 7) on the last zero-forced DRS cells
// of the packet (estimator_mimo.cpp:80-272, rx_synced.cpp:416-431).  Four wideband cells (rx_synced_param.hpp:271),
// metric "highest minimum RX power" (:286), matrices without zero entries only (:277).  One thread per packet.
__device__ uint32_t mimo_search(const float2 (*h)[8][4], uint32_t n_tx, uint32_t n_rx, bool transpose, const int8_t* W,
                                const float* scale, uint32_t n, uint32_t first) {
    float best = -1.0e6f;
    uint32_t ret = 0;
    for (uint32_t wm = first; wm < n; ++wm) {
        float inner = 1.0e6f;
        for (uint32_t rx = 0; rx < n_rx; ++rx) {
            float2 sum = make_float2(0.f, 0.f);
            for (uint32_t tx = 0; tx < n_tx; ++tx) {
                const int8_t w = W[wm * n_tx + tx];
                float2 part = make_float2(0.f, 0.f);
                for (uint32_t c = 0; c < 4; ++c) {
                    const float2 v = transpose ? h[tx][rx][c] : h[rx][tx][c];
                    float2 q = make_float2(0.f, 0.f);
                    if (w == 1) q = v;
                    else if (w == -1) q = make_float2(-v.x, -v.y);
                    else if (w == 2) q = make_float2(-v.y, v.x);
                    else if (w == -2) q = make_float2(v.y, -v.x);
                    part.x += q.x; part.y += q.y;
                }
                sum.x += part.x; sum.y += part.y;
            }
            const float mag = hypotf(sum.x, sum.y);
            if (mag < inner) inner = mag;
        }
        inner *= scale[wm];
        if (best < inner) { best = inner; ret = wm; }
    }
    return ret;
}

__global__ void k_rx_mimo(const mimo_params_t p, uint32_t n) {
    const uint32_t pkt = blockIdx.x * blockDim.x + threadIdx.x;
    if (pkt >= n) return;
    const phy_dims_t& d = p.d;
    const uint32_t step = d.NP / 4, off = step / 2;   // estimator_mimo.cpp:117-122
    float2 h[8][8][4];
    for (uint32_t rx = 0; rx < d.N_RX; ++rx)
        for (uint32_t ts = 0; ts < d.N_eff_TX; ++ts) {
            const float2* src = p.hzf + ((((size_t)pkt * d.N_RX + rx) * d.N_eff_TX + ts) * d.n_drs + (d.n_drs - 1)) * d.NP;
            for (uint32_t c = 0; c < 4; ++c) h[rx][ts][c] = src[off + c * step];
        }
    mimo_report_dev_t r;
    r.N_RX = d.N_RX; r.N_TS_other = d.N_eff_TX;
    r.tm_3_7_beamforming_idx = (d.N_eff_TX == 1 || p.n[0] == 0) ? 0u
        : mimo_search(h, d.N_eff_TX, d.N_RX, false, p.W[0], p.W_scale[0], p.n[0], p.first[0]);
    r.tm_3_7_beamforming_reciprocal_idx = (d.N_RX == 1 || p.n[1] == 0) ? 0u
        : mimo_search(h, d.N_RX, d.N_eff_TX, true, p.W[1], p.W_scale[1], p.n[1], p.first[1]);
    p.out[pkt] = r;
}

// ---- scope / JSON tap: the interpolated channel of every (RX antenna, stream) over all subcarriers as it stands after
// the last data field symbol (rx_synced_t::get_json, rx_synced.cpp:444-480: "last known channel measurement")
template <int NRX>
__global__ void __launch_bounds__(128) k_rx_chestim_tap(const rx_params_t p, float2* __restrict__ out) {
    const phy_dims_t& d = p.d;
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x, ts = blockIdx.y, pkt = blockIdx.z;
    if (k >= d.Nk) return;
    float2 h[NRX];
    chan_est_rx<NRX>(p, pkt, d.last_plan[ts], k, h);
    if (p.track) {   // the estimate as it stands at the last symbol: the tracked rotation put back
        const float2 c0 = track_phasor(d, p.track[pkt], d.N_DF, k);
#pragma unroll
        for (int a = 0; a < NRX; ++a) h[a] = cmulc(h[a], c0);
    }
#pragma unroll
    for (int a = 0; a < NRX; ++a) out[(((size_t)pkt * NRX + a) * d.N_eff_TX + ts) * d.Nk + k] = h[a];
}

cudaError_t launch_rx_chestim_tap(const rx_params_t& p, float2* out, uint32_t n_packets, cudaStream_t st) {
    const dim3 grid((p.d.Nk + 127) / 128, p.d.N_eff_TX, n_packets);
    switch (p.d.N_RX) {
        case 1: k_rx_chestim_tap<1><<<grid, 128, 0, st>>>(p, out); break;
        case 2: k_rx_chestim_tap<2><<<grid, 128, 0, st>>>(p, out); break;
        case 4: k_rx_chestim_tap<4><<<grid, 128, 0, st>>>(p, out); break;
        case 8: k_rx_chestim_tap<8><<<grid, 128, 0, st>>>(p, out); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

cudaError_t launch_rx_mimo(const mimo_params_t& p, uint32_t n_packets, cudaStream_t st) {
    k_rx_mimo<<<(n_packets + 63) / 64, 64, 0, st>>>(p, n_packets);
    return cudaGetLastError();
}

cudaError_t launch_rx_snr_class(const rx_params_t& p, uint32_t n_packets, cudaStream_t st) {
    k_rx_snr_class<<<(n_packets + 127) / 128, 128, 0, st>>>(p, n_packets);
    return cudaGetLastError();
}

cudaError_t launch_rx_fft(const rx_params_t& p, uint32_t n_packets, cudaStream_t st) {
    // one group of N_os / 16 threads per symbol; as many groups per block as fit into 256 threads (at most 8)
    const uint32_t GT = p.d.N_os / 16;
    uint32_t groups = GT >= 256 ? 1 : 256 / GT;
    if (groups > 8) groups = 8;
    const uint32_t items = p.d.N_DF * p.d.N_RX;
    if (groups > items) groups = items;
    const uint32_t threads = ((groups * GT + 31) / 32) * 32;
    if (fft_is_big(p.d.N_os)) {
        const uint32_t Mb = p.d.N_os / 16;
        const uint32_t blen = (p.d.N_os + 2 * fft_buf_len(Mb) + p.d.Nk + 1) / 2;
        const size_t smem = sizeof(float2) * 2 * blen + sizeof(float) * 2 * threads;
        static smem_optin_t optin;   // per device
        cudaError_t e = optin.ensure(smem, [](size_t b) {
            return cudaFuncSetAttribute(k_rx_fft<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)b);
        });
        if (e != cudaSuccess) return e;
        k_rx_fft<true><<<dim3(items, n_packets), threads, smem, st>>>(p, GT, 1);
        return cudaGetLastError();
    }
    // DNRP_FFT_GENERIC=1: the run-time sized kernel also at N_os = 1024 (A/B measurements; Y is bit-identical)
    static const bool generic = [] { const char* e = getenv("DNRP_FFT_GENERIC"); return e && e[0] == '1'; }();
    if (p.d.N_os == 1024 && !generic) {
        const size_t smem1k = sizeof(float2) * (4 * (1024 + 64) + 2 * 4);
        k_rx_fft1024<<<dim3((items + 3) / 4, n_packets), 256, smem1k, st>>>(p);
        return cudaGetLastError();
    }
    const size_t smem = sizeof(float2) * 2 * fft_buf_len(p.d.N_os) * groups + sizeof(float) * 2 * threads;
    if (smem > 48 * 1024) {
        static smem_optin_t optin;   // per device
        cudaError_t e = optin.ensure(smem, [](size_t b) {
            return cudaFuncSetAttribute(k_rx_fft<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)b);
        });
        if (e != cudaSuccess) return e;
    }
    k_rx_fft<false><<<dim3((items + groups - 1) / groups, n_packets), threads, smem, st>>>(p, GT, groups);
    return cudaGetLastError();
}

cudaError_t launch_rx_stf(const rx_params_t& p, uint32_t n_packets, cudaStream_t st) {
    const size_t smem = fft_is_big(p.d.N_os)
        ? sizeof(float2) * ((size_t)p.d.N_os + 2 * fft_buf_len(p.d.N_os / 16) + (size_t)p.d.N_RX * (p.d.N_OCC / 4))
        : sizeof(float2) * ((size_t)p.d.stf_len_os + 2 * fft_buf_len(p.d.N_os) + (size_t)p.d.N_RX * (p.d.N_OCC / 4));
    if (smem > 48 * 1024) {
        static smem_optin_t optin[2];   // per device and variant
        const bool big = fft_is_big(p.d.N_os);
        cudaError_t e = optin[big].ensure(smem, [big](size_t b) {
            return big ? cudaFuncSetAttribute(k_rx_stf<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)b)
                       : cudaFuncSetAttribute(k_rx_stf<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)b);
        });
        if (e != cudaSuccess) return e;
    }
    if (fft_is_big(p.d.N_os)) k_rx_stf<true><<<n_packets, 256, smem, st>>>(p);
    else k_rx_stf<false><<<n_packets, 256, smem, st>>>(p);
    return cudaGetLastError();
}

}  // namespace dnrp

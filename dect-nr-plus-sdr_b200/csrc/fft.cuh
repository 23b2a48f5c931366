// fft.cuh -- block-wide shared-memory Stockham FFT for the OFDM sizes N = 2^a * 3^b (64 ... 4096 in one
// block).  Unnormalised in both directions, natural order in and out, like the FFTW3f plans behind the
// reference's dft::single_symbol_{tx,rx}_ofdm_zero_copy (lib/src/phy/dft/ofdm.cpp:62-92).
// Twiddles W_N^k = exp(-2*pi*i*k/N) come from a table computed in double precision on the host.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace dnrp {

__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
    return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
__device__ __forceinline__ float2 cmulc(float2 a, float2 b) {  // a * conj(b)
    return make_float2(a.x * b.x + a.y * b.y, a.y * b.x - a.x * b.y);
}
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }

// twiddle W_N^(idx) for the forward transform, conjugated for the inverse
template <bool INV>
__device__ __forceinline__ float2 tw(const float2* __restrict__ table, uint32_t idx) {
    float2 w = __ldg(table + idx);
    if (INV) w.y = -w.y;
    return w;
}

// Buffers handed to block_fft hold fft_buf_len(N) elements: intermediate passes store with one pad element per 16
// (index i at i + i/16), which spreads the stride-16 writes of the radix-16 passes over all banks.  The first pass
// reads and the last pass writes plain indices, so callers never see the padding.
__host__ __device__ __forceinline__ uint32_t fft_buf_len(uint32_t N) { return N + (N >> 4); }
__device__ __forceinline__ uint32_t fft_pad(uint32_t i) { return i + (i >> 4); }

// radix-4 butterfly, forward (INV: inverse) DFT of four values in place
template <bool INV>
__device__ __forceinline__ void dft4(float2& v0, float2& v1, float2& v2, float2& v3) {
    const float2 a = cadd(v0, v2), b = csub(v0, v2);
    const float2 c = cadd(v1, v3), d = csub(v1, v3);
    const float2 id = INV ? make_float2(-d.y, d.x) : make_float2(d.y, -d.x);  // forward: -i*d, inverse: +i*d
    v0 = cadd(a, c);
    v1 = cadd(b, id);
    v2 = csub(a, c);
    v3 = csub(b, id);
}

// 16-point DFT in registers: n = c + 4m, k = q + 4p: X[q + 4p] = sum_c W4^(cp) W16^(cq) sum_m x[c + 4m] W4^(mq)
template <bool INV>
__device__ __forceinline__ void dft16(float2 (&v)[16]) {
#pragma unroll
    for (int c = 0; c < 4; ++c) dft4<INV>(v[c], v[c + 4], v[c + 8], v[c + 12]);  // v[c + 4q] = inner sum for (c, q)
    // W16^(cq), c, q = 1..3 (forward: exp(-2 pi i cq / 16))
    constexpr float C1 = 0.92387953251128674f, S1 = 0.38268343236508977f, R2 = 0.70710678118654752f;
    const float sg = INV ? 1.f : -1.f;
    auto rotm = [&](float2& x, float cr, float ci) { x = make_float2(x.x * cr - x.y * (sg * ci), x.x * (sg * ci) + x.y * cr); };
    rotm(v[1 + 4], C1, S1);                       // cq = 1
    rotm(v[2 + 4], R2, R2);                       // 2
    rotm(v[3 + 4], S1, C1);                       // 3
    rotm(v[1 + 8], R2, R2);                       // 2
    rotm(v[2 + 8], 0.f, 1.f);                     // 4
    rotm(v[3 + 8], -R2, R2);                      // 6
    rotm(v[1 + 12], S1, C1);                      // 3
    rotm(v[2 + 12], -R2, R2);                     // 6
    rotm(v[3 + 12], -C1, -S1);                    // 9
#pragma unroll
    for (int q = 0; q < 4; ++q) dft4<INV>(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);  // over c -> p
    // now v[4q + p] = X[q + 4p]: transpose to natural order
    float2 o[16];
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int pp = 0; pp < 4; ++pp) o[q + 4 * pp] = v[4 * q + pp];
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = o[i];
}

// One Stockham pass of radix R (Bainville's formulation): butterfly j of N/R takes x[j + r N/R], multiplies by
// W_{Ns R}^(k r), k = j mod Ns, and stores X[r] at (j / Ns) Ns R + k + r Ns.  Ns = product of the radices of the
// previous passes.  POW2: Ns is a power of two: index arithmetic by mask and shift instead of integer division.
// k * r * tstep < Ns * R * N / (Ns * R) = N, so the twiddle index needs no reduction; the first pass (Ns == 1) has unit
// twiddles.  pin / pout: the source / destination buffer is padded (fft_pad); N/R and, from the second pass on, Ns are
// multiples of 16, so the padded position of element r is base + r * padded stride.
// ptab (optional): this pass's twiddles in the order the lanes read them, W_{Ns R}^(k r) at (r - 1) Ns + k
// (fft_pass_tables below) -- consecutive butterflies read consecutive entries, one or two 128-byte lines per warp load,
// where table[k r tstep] scatters a warp over up to sixteen lines.  Same values, same results.
template <bool INV, int R, bool POW2, bool PAD>
__device__ __forceinline__ void fft_pass(const float2* __restrict__ x, float2* __restrict__ y, uint32_t N,
                                         uint32_t Ns, uint32_t log2Ns, const float2* __restrict__ table, bool pin_, bool pout_,
                                         uint32_t tid = threadIdx.x, uint32_t nthr = blockDim.x, uint32_t tmul = 1,
                                         const float2* __restrict__ ptab = nullptr) {
    const bool pin = PAD && pin_, pout = PAD && pout_;   // PAD = false: plain buffers, no index selects in the code
    const uint32_t M = N / R;
    const uint32_t tstep = POW2 ? M >> log2Ns : N / (Ns * R);  // W_{Ns*R}^k = W_N^{k*tstep}
    const uint32_t xs = pin ? fft_pad(M) : M;                  // stride between the R inputs
    const uint32_t ys = (pout && Ns >= 16) ? fft_pad(Ns) : Ns;  // stride between the R outputs
    for (uint32_t j = tid; j < M; j += nthr) {
        const uint32_t k = POW2 ? (j & (Ns - 1)) : j % Ns;
        float2 v[R];
        const float2* xj = x + (pin ? fft_pad(j) : j);
#pragma unroll
        for (int r = 0; r < R; ++r) v[r] = xj[r * xs];
        if (Ns != 1) {
            if (ptab) {
#pragma unroll
                for (int r = 1; r < R; ++r) v[r] = cmul(v[r], tw<INV>(ptab, (r - 1) * Ns + k));
            } else {
                const uint32_t kt = k * tstep * tmul;   // tmul: the table belongs to a transform tmul times longer
#pragma unroll
                for (int r = 1; r < R; ++r) v[r] = cmul(v[r], tw<INV>(table, kt * r));
            }
        }
        if constexpr (R == 2) {
            const float2 a = v[0], b = v[1];
            v[0] = cadd(a, b);
            v[1] = csub(a, b);
        } else if constexpr (R == 4) {
            dft4<INV>(v[0], v[1], v[2], v[3]);
        } else if constexpr (R == 16) {
            dft16<INV>(v);
        } else if constexpr (R == 3) {
            const float s = INV ? 0.8660254037844386f : -0.8660254037844386f;  // sin(-2pi/3) forward
            const float2 t1 = cadd(v[1], v[2]);
            const float2 t2 = make_float2(v[0].x - 0.5f * t1.x, v[0].y - 0.5f * t1.y);
            const float2 d = csub(v[1], v[2]);
            const float2 t3 = make_float2(-s * d.y, s * d.x);  // i*s*d
            v[0] = cadd(v[0], t1);
            v[1] = cadd(t2, t3);
            v[2] = csub(t2, t3);
        }
        const uint32_t j0 = POW2 ? (((j >> log2Ns) << log2Ns) * R + k) : ((j / Ns) * Ns * R + k);
        if (Ns == 1) {  // first pass: R consecutive outputs (padded: R j + r + j for R = 16)
            float2* yj = y + (pout ? fft_pad(R * j) : R * j);
            if constexpr (!PAD && R == 4) {  // two 16-byte stores (buffers are 16-byte aligned)
                float4* y4 = reinterpret_cast<float4*>(yj);
                y4[0] = make_float4(v[0].x, v[0].y, v[1].x, v[1].y);
                y4[1] = make_float4(v[2].x, v[2].y, v[3].x, v[3].y);
            } else {
#pragma unroll
                for (int r = 0; r < R; ++r) yj[r] = v[r];
            }
        } else {
            float2* yj = y + (pout ? fft_pad(j0) : j0);
#pragma unroll
            for (int r = 0; r < R; ++r) yj[r * ys] = v[r];
        }
    }
}

// In-block FFT of a[0..N); a and b hold fft_buf_len(N) elements each, b is scratch.  Returns the buffer holding the
// result (plain indices).  All threads of the block must call it; contains __syncthreads().  N = 2^a 3^b, a >= 4 (the
// OFDM sizes are multiples of 64): radix-16 passes in registers first, then radix 4, 2 and 3 -- Ns stays a power of
// two through the last power-of-two pass and the first radix-3 pass, so no pass divides by a runtime value.
// RADIX16 = false: radix-4 passes only and no padding (kernels whose block has four times more threads than N / 16
// butterflies: measured slower with radix 16 -- k_rx_fft 0.84 -> 1.13 ms at N = 1024 with 256 threads, 32 -> 70
// registers); true: the long transforms of the fine synchronisation (N = 3072 at beta = 16: 0.60 -> 0.45 ms).
// ptab: per-pass twiddle tables of this size and plan (fft_pass_tables(N, RADIX16)), or nullptr.
template <bool INV, bool RADIX16 = false>
__device__ float2* block_fft(float2* a, float2* b, uint32_t N, const float2* __restrict__ table,
                             const float2* __restrict__ ptab = nullptr) {
    uint32_t Ns = 1, lg = 0;
    uint32_t rem = N;
    bool first = true;
    __syncthreads();
#define DNRP_FFT_PASS(R, P2, LGINC)                                                          \
    {                                                                                        \
        const bool last = rem == R;                                                          \
        fft_pass<INV, R, P2, RADIX16>(a, b, N, Ns, lg, table, !first, !last, threadIdx.x, blockDim.x, 1, ptab); \
        if (ptab && Ns != 1) ptab += (R - 1) * Ns;                                           \
        Ns *= R; lg += LGINC; rem /= R; first = false;                                       \
        float2* t = a; a = b; b = t;                                                         \
        __syncthreads();                                                                     \
    }
    if constexpr (RADIX16) {
        while (rem % 16 == 0) DNRP_FFT_PASS(16, true, 4)
    }
    while (rem % 4 == 0) DNRP_FFT_PASS(4, true, 2)
    if (rem % 2 == 0) DNRP_FFT_PASS(2, true, 1)
    bool pow2 = true;
    while (rem % 3 == 0) {
        if (pow2) DNRP_FFT_PASS(3, true, 0)
        else DNRP_FFT_PASS(3, false, 0)
        pow2 = false;
    }
#undef DNRP_FFT_PASS
    return a;
}

// FFT of one symbol by a GROUP of N / 16 threads (several groups per block, all moving through the same barriers):
// thread t of the group enters with v[m] = x[t + (N / 16) m], m = 0..15, in registers -- the first radix-16 pass
// (Ns = 1, unit twiddles) needs no shared memory on its input side -- and the remaining passes alternate between the
// group's two padded buffers a and b (fft_buf_len(N) elements each).  N = 1024: passes 16, 16, 4 = two exchanges
// through shared memory instead of the four of the radix-4 plan.  Returns the buffer holding the result (plain
// indices).  Every thread of the block must call it (active = false: only takes part in the barriers).
template <bool INV>
__device__ float2* group_fft16(float2 (&v)[16], float2* a, float2* b, uint32_t N, const float2* __restrict__ table,
                               uint32_t tid, uint32_t nthr, bool active, const float2* __restrict__ ptab = nullptr) {
    uint32_t Ns = 16, lg = 4, rem = N / 16;
    if (active) {
        dft16<INV>(v);
        const bool last = rem == 1;
        float2* yj = a + (last ? 16 * tid : fft_pad(16 * tid));
#pragma unroll
        for (int r = 0; r < 16; ++r) yj[r] = v[r];
    }
    __syncthreads();
#define DNRP_GFFT_PASS(R, P2, LGINC)                                                              \
    {                                                                                             \
        const bool last = rem == R;                                                               \
        if (active) fft_pass<INV, R, P2, true>(a, b, N, Ns, lg, table, true, !last, tid, nthr, 1, ptab);   \
        if (ptab) ptab += (R - 1) * Ns;                                                           \
        Ns *= R; lg += LGINC; rem /= R;                                                           \
        float2* t = a; a = b; b = t;                                                              \
        __syncthreads();                                                                          \
    }
    while (rem % 16 == 0) DNRP_GFFT_PASS(16, true, 4)
    while (rem % 4 == 0) DNRP_GFFT_PASS(4, true, 2)
    if (rem % 2 == 0) DNRP_GFFT_PASS(2, true, 1)
    bool pow2 = true;
    while (rem % 3 == 0) {
        if (pow2) DNRP_GFFT_PASS(3, true, 0)
        else DNRP_GFFT_PASS(3, false, 0)
        pow2 = false;
    }
#undef DNRP_GFFT_PASS
    return a;
}

// ---- transforms longer than 4096 (6144 ... 16384: a packet received at several times its own bandwidth, tx_rx.hpp:66-69):
// two full ping-pong buffers no longer fit the shared memory of a block, so N = 16 M is split four-step style and done
// IN PLACE in one array a[N]:
//   step A  for every column n2 < M: 16-point DFT over x[n2 + M n1], times W_N^(n2 k1), stored at a[k1 M + n2]
//           (the same 16 positions the column was read from);
//   step B  for every row k1 < 16: M-point FFT of a[k1 M .. k1 M + M) through two small padded scratch buffers
//           (s1, s2: fft_buf_len(M) each), back into the row.
// Result: X[k1 + 16 k2] = a[k1 M + k2], i.e. bin k sits at big_fft_at(k, M).  Shared memory N + 2 fft_buf_len(M).
__host__ __device__ __forceinline__ bool fft_is_big(uint32_t N) { return N > 4096; }
__device__ __forceinline__ uint32_t big_fft_at(uint32_t k, uint32_t M) { return (k & 15u) * M + (k >> 4); }

// step A for one column held in registers: v[n1] = x[n2 + M n1] -> a[k1 M + n2]
template <bool INV>
__device__ __forceinline__ void big_fft_col(float2 (&v)[16], float2* a, uint32_t n2, uint32_t M, const float2* __restrict__ table) {
    dft16<INV>(v);
    a[n2] = v[0];
#pragma unroll
    for (int k1 = 1; k1 < 16; ++k1) a[k1 * M + n2] = cmul(v[k1], tw<INV>(table, n2 * k1));   // n2 k1 < N
}

// step B: all nthr threads (ids tid) of the block / group take part; contains __syncthreads()
template <bool INV>
__device__ void big_fft_rows(float2* a, float2* s1, float2* s2, uint32_t N, const float2* __restrict__ table, uint32_t tid,
                             uint32_t nthr, bool active) {
    const uint32_t M = N / 16;
    __syncthreads();
    for (uint32_t k1 = 0; k1 < 16; ++k1) {
        float2* row = a + k1 * M;
        const float2* src = row;
        uint32_t Ns = 1, lg = 0, rem = M, pass = 0;
#define DNRP_BFFT_PASS(R, P2, LGINC)                                                                              \
        {                                                                                                         \
            const bool last = rem == R;                                                                           \
            float2* dst = last ? row : ((pass & 1) ? s2 : s1);                                                    \
            if (active) fft_pass<INV, R, P2, true>(src, dst, M, Ns, lg, table, pass != 0, !last, tid, nthr, 16);  \
            Ns *= R; lg += LGINC; rem /= R; ++pass; src = dst;                                                    \
            __syncthreads();                                                                                      \
        }
        while (rem % 16 == 0 && rem > 16) DNRP_BFFT_PASS(16, true, 4)   // never as the only pass: row -> row would alias
        while (rem % 4 == 0) DNRP_BFFT_PASS(4, true, 2)
        if (rem % 2 == 0) DNRP_BFFT_PASS(2, true, 1)
        bool pow2 = true;
        while (rem % 3 == 0) {
            if (pow2) DNRP_BFFT_PASS(3, true, 0)
            else DNRP_BFFT_PASS(3, false, 0)
            pow2 = false;
        }
#undef DNRP_BFFT_PASS
    }
}

// whole transform of a[0..N) in shared memory (block-wide; every thread calls)
template <bool INV>
__device__ void big_fft(float2* a, float2* s1, float2* s2, uint32_t N, const float2* __restrict__ table) {
    const uint32_t M = N / 16;
    __syncthreads();
    for (uint32_t n2 = threadIdx.x; n2 < M; n2 += blockDim.x) {
        float2 v[16];
#pragma unroll
        for (int n1 = 0; n1 < 16; ++n1) v[n1] = a[n2 + n1 * M];
        big_fft_col<INV>(v, a, n2, M, table);
    }
    big_fft_rows<INV>(a, s1, s2, N, table, threadIdx.x, blockDim.x, true);
}

}  // namespace dnrp

#include <vector>
namespace dnrp {
// Host: the per-pass twiddle tables of one transform size, concatenated in pass order.  tw = W_N^k (k < N); radix16 = the
// plan of block_fft<.., true> / group_fft16 (16s, then 4s, 2, 3s), else that of block_fft<.., false> (4s, 2, 3s).  A pass
// of radix R entered with Ns > 1 contributes W_{Ns R}^(k r) = tw[k r N / (Ns R)] at (r - 1) Ns + k; the first pass
// (Ns = 1) has unit twiddles and no entries.  N = 1024, radix16: [15][16], then [3][256] (k_rx_fft1024 reads these).
inline std::vector<float2> fft_pass_tables(const std::vector<float2>& tw, bool radix16) {
    const uint32_t N = (uint32_t)tw.size();
    std::vector<float2> out;
    uint32_t Ns = 1, rem = N;
    auto pass = [&](uint32_t R) {
        if (Ns != 1) {
            const uint32_t tstep = N / (Ns * R);
            for (uint32_t r = 1; r < R; ++r)
                for (uint32_t k = 0; k < Ns; ++k) out.push_back(tw[(size_t)k * tstep * r]);
        }
        Ns *= R; rem /= R;
    };
    if (radix16) while (rem % 16 == 0) pass(16);
    while (rem % 4 == 0) pass(4);
    if (rem % 2 == 0) pass(2);
    while (rem % 3 == 0) pass(3);
    return out;
}
}  // namespace dnrp

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

// One Stockham pass of radix R (Bainville's formulation): element j of N/R butterflies,
// Ns = product of the radices of the previous passes.  POW2: Ns is a power of two (every pass of a
// power-of-two transform): index arithmetic by mask and shift instead of integer division.
// k * r * tstep < Ns * R * N / (Ns * R) = N, so the twiddle index needs no reduction; the first pass
// (Ns == 1) has unit twiddles.
template <bool INV, int R, bool POW2>
__device__ __forceinline__ void fft_pass(const float2* __restrict__ x, float2* __restrict__ y, uint32_t N,
                                         uint32_t Ns, uint32_t log2Ns, const float2* __restrict__ table) {
    const uint32_t M = N / R;
    const uint32_t tstep = POW2 ? M >> log2Ns : N / (Ns * R);  // W_{Ns*R}^k = W_N^{k*tstep}
    for (uint32_t j = threadIdx.x; j < M; j += blockDim.x) {
        const uint32_t k = POW2 ? (j & (Ns - 1)) : j % Ns;
        float2 v[R];
#pragma unroll
        for (int r = 0; r < R; ++r) v[r] = x[j + r * M];
        if (Ns != 1) {
            const uint32_t kt = k * tstep;
#pragma unroll
            for (int r = 1; r < R; ++r) v[r] = cmul(v[r], tw<INV>(table, kt * r));
        }
        if (R == 2) {
            const float2 a = v[0], b = v[1];
            v[0] = cadd(a, b);
            v[1] = csub(a, b);
        } else if (R == 4) {
            const float2 a = cadd(v[0], v[2]), b = csub(v[0], v[2]);
            const float2 c = cadd(v[1], v[3]), d = csub(v[1], v[3]);
            // forward: -i*d, inverse: +i*d
            const float2 id = INV ? make_float2(-d.y, d.x) : make_float2(d.y, -d.x);
            v[0] = cadd(a, c);
            v[1] = cadd(b, id);
            v[2] = csub(a, c);
            v[3] = csub(b, id);
        } else if (R == 3) {
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
        if (R == 4 && Ns == 1) {  // first pass: four consecutive outputs, two 16-byte stores (y is 16-byte aligned)
            float4* y4 = reinterpret_cast<float4*>(y + 4 * j);
            y4[0] = make_float4(v[0].x, v[0].y, v[1].x, v[1].y);
            y4[1] = make_float4(v[2].x, v[2].y, v[3].x, v[3].y);
        } else {
#pragma unroll
            for (int r = 0; r < R; ++r) y[j0 + r * Ns] = v[r];
        }
    }
}

// In-block FFT of a[0..N); b is scratch of the same size.  Returns the buffer holding the result.
// All threads of the block must call it; contains __syncthreads().  Radix-4 passes first, then one radix-2,
// then the radix-3 passes: Ns stays a power of two through the last power-of-two pass and the first radix-3 pass
// (the OFDM sizes have at most one factor 3), so no pass divides by a runtime value.
template <bool INV>
__device__ float2* block_fft(float2* a, float2* b, uint32_t N, const float2* __restrict__ table) {
    uint32_t Ns = 1, lg = 0;
    uint32_t rem = N;
    __syncthreads();
    while (rem % 4 == 0) {
        fft_pass<INV, 4, true>(a, b, N, Ns, lg, table);
        Ns *= 4; lg += 2; rem /= 4;
        float2* t = a; a = b; b = t;
        __syncthreads();
    }
    if (rem % 2 == 0) {
        fft_pass<INV, 2, true>(a, b, N, Ns, lg, table);
        Ns *= 2; lg += 1; rem /= 2;
        float2* t = a; a = b; b = t;
        __syncthreads();
    }
    bool pow2 = true;
    while (rem % 3 == 0) {
        if (pow2) fft_pass<INV, 3, true>(a, b, N, Ns, lg, table);
        else fft_pass<INV, 3, false>(a, b, N, Ns, 0, table);
        Ns *= 3; rem /= 3; pow2 = false;
        float2* t = a; a = b; b = t;
        __syncthreads();
    }
    return a;
}

}  // namespace dnrp

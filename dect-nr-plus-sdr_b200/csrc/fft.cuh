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
// Ns = product of the radices of the previous passes.
template <bool INV, int R>
__device__ __forceinline__ void fft_pass(const float2* __restrict__ x, float2* __restrict__ y, uint32_t N,
                                         uint32_t Ns, const float2* __restrict__ table) {
    const uint32_t M = N / R;
    for (uint32_t j = threadIdx.x; j < M; j += blockDim.x) {
        const uint32_t k = j % Ns;
        const uint32_t tstep = N / (Ns * R);  // W_{Ns*R}^k = W_N^{k*tstep}
        float2 v[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            v[r] = x[j + r * M];
            if (r > 0) v[r] = cmul(v[r], tw<INV>(table, (k * r * tstep) % N));
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
        const uint32_t j0 = (j / Ns) * Ns * R + k;
#pragma unroll
        for (int r = 0; r < R; ++r) y[j0 + r * Ns] = v[r];
    }
}

// In-block FFT of a[0..N); b is scratch of the same size.  Returns the buffer holding the result.
// All threads of the block must call it; contains __syncthreads().
template <bool INV>
__device__ float2* block_fft(float2* a, float2* b, uint32_t N, const float2* __restrict__ table) {
    uint32_t Ns = 1;
    uint32_t rem = N;
    __syncthreads();
    while (rem % 3 == 0) {
        fft_pass<INV, 3>(a, b, N, Ns, table);
        Ns *= 3; rem /= 3;
        float2* t = a; a = b; b = t;
        __syncthreads();
    }
    while (rem % 4 == 0) {
        fft_pass<INV, 4>(a, b, N, Ns, table);
        Ns *= 4; rem /= 4;
        float2* t = a; a = b; b = t;
        __syncthreads();
    }
    if (rem == 2) {
        fft_pass<INV, 2>(a, b, N, Ns, table);
        float2* t = a; a = b; b = t;
        __syncthreads();
    }
    return a;
}

}  // namespace dnrp

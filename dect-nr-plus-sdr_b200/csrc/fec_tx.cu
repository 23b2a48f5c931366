// fec_tx.cu -- TX side of the FEC chain on the GPU, one thread block per (code block, packet):
// CRC24A (TB) / CRC24B (CB) / CRC16^mask (PLCF) -> LTE turbo encoder -> rate matching with redundancy
// version -> scrambling at the running bit offset -> packed bits.  Reference:
// lib/src/phy/fec/pdc_enc.cpp:127-229 and pcc_enc.cpp:145-213 (relative to /root/reference); the
// arithmetic they delegate to srsRAN (turbocoder.c, rm_turbo.c, crc.c) is TS 36.212 5.1.1-5.1.4.
#include "fec_tx.cuh"

namespace dnrp {

__device__ __forceinline__ uint32_t gf2_mulmod_tx(uint32_t a, uint32_t b, uint32_t poly, uint32_t order) {
    uint32_t r = 0;
    const uint32_t top = 1u << order;
    for (uint32_t i = 0; i < order; ++i) {
        r <<= 1;
        if (r & top) r ^= poly;
        if ((b >> (order - 1 - i)) & 1) r ^= a;
    }
    return r;
}

// x^n mod g
__device__ uint32_t gf2_xpow(uint32_t n, uint32_t poly, uint32_t order) {
    uint32_t result = 1, base = 2;
    while (n) {
        if (n & 1) result = gf2_mulmod_tx(result, base, poly, order);
        base = gf2_mulmod_tx(base, base, poly, order);
        n >>= 1;
    }
    return result;
}

// CRC (zero init) of nbits message bits fetched through getbit(i), computed by the whole block:
// remainder of M(x) * x^order.  s_acc must be a shared word.
template <typename F>
__device__ uint32_t block_crc(F getbit, uint32_t nbits, uint32_t poly, uint32_t order, uint32_t* s_acc) {
    const uint32_t tid = threadIdx.x, nthr = blockDim.x;
    const uint32_t chunk = (nbits + nthr - 1) / nthr;
    const uint32_t b0 = min(nbits, tid * chunk), b1 = min(nbits, b0 + chunk);
    const uint32_t top = 1u << order;
    if (tid == 0) *s_acc = 0;
    __syncthreads();
    uint32_t r = 0;
    for (uint32_t i = b0; i < b1; ++i) {
        r = (r << 1) | getbit(i);
        if (r & top) r ^= poly;
    }
    if (r) r = gf2_mulmod_tx(r, gf2_xpow(nbits - b1 + order, poly, order), poly, order);
    for (int o = 16; o > 0; o >>= 1) r ^= __shfl_xor_sync(0xffffffffu, r, o);
    if ((tid & 31) == 0 && r) atomicXor(s_acc, r);
    __syncthreads();
    const uint32_t out = *s_acc;
    __syncthreads();
    return out;
}

// recursive systematic convolutional encoder over bits in[perm(i)], chunk-parallel:
// the register evolution is GF(2)-linear, so chunk end states combine through the zero-input map.
template <typename F>
__device__ void block_rsc(F getbit, uint32_t K, uint8_t* par, uint8_t* xt, uint8_t* zt, uint8_t* s_state) {
    const uint32_t tid = threadIdx.x, nthr = blockDim.x;
    const uint32_t chunk = (K + nthr - 1) / nthr;
    const uint32_t k0 = min(K, tid * chunk), k1 = min(K, k0 + chunk);
    // pass 1: end state of the chunk started from the all-zero register
    uint32_t s0 = 0, s1 = 0, s2 = 0;
    for (uint32_t k = k0; k < k1; ++k) {
        const uint32_t a = getbit(k) ^ s1 ^ s2;
        s2 = s1; s1 = s0; s0 = a;
    }
    s_state[tid] = (uint8_t)((s0 << 2) | (s1 << 1) | s2);
    __syncthreads();
    if (tid == 0) {
        // zero-input evolution of the three basis states over one full chunk
        uint8_t zi[8];
        for (uint32_t b = 0; b < 8; ++b) {
            uint32_t q0 = (b >> 2) & 1, q1 = (b >> 1) & 1, q2 = b & 1;
            for (uint32_t k = 0; k < chunk; ++k) {
                const uint32_t a = q1 ^ q2;
                q2 = q1; q1 = q0; q0 = a;
            }
            zi[b] = (uint8_t)((q0 << 2) | (q1 << 1) | q2);
        }
        uint8_t st = 0;  // state at the start of chunk t
        for (uint32_t t = 0; t < nthr; ++t) {
            const uint8_t e = s_state[t];
            s_state[t] = st;
            const uint32_t kk0 = min(K, t * chunk), kk1 = min(K, kk0 + chunk);
            if (kk1 - kk0 == chunk) st = zi[st] ^ e;
            else if (kk1 > kk0) {  // short last chunk: evolve explicitly
                uint32_t q0 = (st >> 2) & 1, q1 = (st >> 1) & 1, q2 = st & 1;
                for (uint32_t k = kk0; k < kk1; ++k) { const uint32_t a = q1 ^ q2; q2 = q1; q1 = q0; q0 = a; }
                st = (uint8_t)(((q0 << 2) | (q1 << 1) | q2) ^ e);
            }
        }
        s_state[nthr] = st;  // final register content -> termination
    }
    __syncthreads();
    // pass 2: true start state, emit parity
    uint32_t st = s_state[tid];
    s0 = (st >> 2) & 1; s1 = (st >> 1) & 1; s2 = st & 1;
    for (uint32_t k = k0; k < k1; ++k) {
        const uint32_t a = getbit(k) ^ s1 ^ s2;
        par[k] = (uint8_t)(a ^ s0 ^ s2);
        s2 = s1; s1 = s0; s0 = a;
    }
    if (tid == 0) {
        st = s_state[nthr];
        s0 = (st >> 2) & 1; s1 = (st >> 1) & 1; s2 = st & 1;
        for (int t = 0; t < 3; ++t) {
            xt[t] = (uint8_t)(s1 ^ s2);
            zt[t] = (uint8_t)(s0 ^ s2);
            s2 = s1; s1 = s0; s0 = 0;
        }
    }
    __syncthreads();
}

__global__ void __launch_bounds__(256) k_tx_fec(const txfec_params_t p) {
    extern __shared__ uint8_t sm8[];
    const txcb_desc_t cb = p.cbs[blockIdx.x];
    const uint32_t pkt = blockIdx.y, tid = threadIdx.x, nthr = blockDim.x;
    const uint32_t K = cb.K, D = K + 4, N = 3 * D;
    uint8_t* c = sm8;            // [K] input bits of the code block
    uint8_t* d = c + K;          // [3][K+4]
    uint8_t* s_state = d + N;    // [nthr+1]
    __shared__ uint32_t s_acc;
    __shared__ uint8_t s_tail[12];

    const uint8_t* in = p.in + (size_t)pkt * p.in_stride;
    const uint32_t rv = p.rv ? p.rv[pkt] : 0;
    auto inbit = [&](uint32_t i) -> uint32_t { return (in[i >> 3] >> (7 - (i & 7))) & 1u; };

    // ---- payload bits of this code block
    for (uint32_t i = tid; i < cb.n_data; i += nthr) c[i] = (uint8_t)inbit(cb.rp + i);
    uint32_t fill = cb.n_data;
    if (cb.tb_crc_bits) {
        // whole-message CRC (CRC24A over the TB in the last CB; CRC16 ^ mask for the PLCF)
        uint32_t crc = block_crc(inbit, cb.tb_crc_len, cb.tb_crc_poly, cb.tb_crc_bits, &s_acc);
        if (p.plcf_mask) crc ^= p.plcf_mask[pkt];
        if (tid < cb.tb_crc_bits) c[fill + tid] = (uint8_t)((crc >> (cb.tb_crc_bits - 1 - tid)) & 1);
        fill += cb.tb_crc_bits;
    }
    __syncthreads();
    if (cb.cb_crc) {
        const uint32_t crc = block_crc([&](uint32_t i) -> uint32_t { return c[i]; }, fill, 0x1800063u, 24, &s_acc);
        if (tid < 24) c[fill + tid] = (uint8_t)((crc >> (23 - tid)) & 1);
        fill += 24;
    }
    __syncthreads();

    // ---- turbo encoder: d0 systematic, d1 parity of c, d2 parity of c[pi(i)]
    uint8_t* d0 = d; uint8_t* d1 = d + D; uint8_t* d2 = d + 2 * D;
    for (uint32_t i = tid; i < K; i += nthr) d0[i] = c[i];
    block_rsc([&](uint32_t i) -> uint32_t { return c[i]; }, K, d1, s_tail, s_tail + 3, s_state);
    block_rsc([&](uint32_t i) -> uint32_t {
        return c[(uint32_t)(((uint64_t)cb.f1 * i + (uint64_t)cb.f2 * i * i) % K)]; },
        K, d2, s_tail + 6, s_tail + 9, s_state);
    if (tid == 0) {
        const uint8_t* x = s_tail; const uint8_t* z = s_tail + 3;
        const uint8_t* xp = s_tail + 6; const uint8_t* zp = s_tail + 9;
        d0[K] = x[0];      d1[K] = z[0];      d2[K] = x[1];
        d0[K + 1] = z[1];  d1[K + 1] = x[2];  d2[K + 1] = z[2];
        d0[K + 2] = xp[0]; d1[K + 2] = zp[0]; d2[K + 2] = xp[1];
        d0[K + 3] = zp[1]; d1[K + 3] = xp[2]; d2[K + 3] = zp[2];
    }
    __syncthreads();

    // ---- rate matching + scrambling + bit packing at bit offset wp of the packet's d buffer
    uint8_t* out = p.out + (size_t)pkt * p.out_stride;
    const uint16_t* src = p.rm_arena + cb.rm_src_off;
    const uint32_t start = cb.rm_start[rv];
    const uint32_t wp = cb.wp, we = cb.wp + cb.n_e + cb.n_pad;  // n_pad: scrambled zero bits (PLCF: 196 -> 200)
    const uint32_t B0 = wp >> 3, B1 = (we + 7) >> 3;
    for (uint32_t B = B0 + tid; B < B1; B += nthr) {
        uint32_t v = 0;
        for (uint32_t q = 0; q < 8; ++q) {
            const uint32_t g = B * 8 + q;
            uint32_t bit = 0;
            if (g >= wp && g < we) {
                const uint32_t e = g - wp;
                if (e < cb.n_e) {
                    uint32_t pos = start + e % N;
                    if (pos >= N) pos -= N;
                    bit = d[src[pos]];
                }
                const uint32_t gg = g + p.gold_bit_off;
                bit ^= (p.gold[gg >> 3] >> (7 - (gg & 7))) & 1u;
            }
            v = (v << 1) | bit;
        }
        if (B * 8 >= wp && B * 8 + 8 <= we) out[B] = (uint8_t)v;
        else atomicOr((unsigned int*)(out + (B & ~3u)), v << (8 * (B & 3u)));
    }
}

cudaError_t launch_tx_fec(const txfec_params_t& p, uint32_t n_packets, uint32_t Kmax, cudaStream_t st) {
    const uint32_t threads = 256;
    const size_t smem = (size_t)Kmax + 3 * (Kmax + 4) + threads + 16;
    k_tx_fec<<<dim3(p.C, n_packets), threads, smem, st>>>(p);
    return cudaGetLastError();
}

}  // namespace dnrp

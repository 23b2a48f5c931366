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

// CRC (zero init) of nbits message bits fetched through getbit(i), computed by the whole block:
// remainder of M(x) * x^order.  s_acc must be a shared word.
// The same over packed message bytes (MSB first), chunks of whole bytes: one load per eight bits.  xp as below with
// byte chunks (plan_fec.cu: add_xp_tx with byte_chunks).
__device__ uint32_t block_crc_bytes(const uint8_t* __restrict__ msg, uint32_t nbytes, uint32_t poly, uint32_t order,
                                    const uint32_t* __restrict__ xp, uint32_t* s_acc) {
    const uint32_t tid = threadIdx.x, nthr = blockDim.x;
    const uint32_t chunk = (nbytes + nthr - 1) / nthr;
    const uint32_t b0 = min(nbytes, tid * chunk), b1 = min(nbytes, b0 + chunk);
    const uint32_t top = 1u << order;
    if (tid == 0) *s_acc = 0;
    __syncthreads();
    uint32_t r = 0;
    for (uint32_t i = b0; i < b1; ++i) {
        const uint32_t v = msg[i];
#pragma unroll
        for (int q = 7; q >= 0; --q) {
            r = (r << 1) | ((v >> q) & 1u);
            if (r & top) r ^= poly;
        }
    }
    if (r) r = gf2_mulmod_tx(r, xp[tid], poly, order);
    for (int o = 16; o > 0; o >>= 1) r ^= __shfl_xor_sync(0xffffffffu, r, o);
    if ((tid & 31) == 0 && r) atomicXor(s_acc, r);
    __syncthreads();
    const uint32_t out = *s_acc;
    __syncthreads();
    return out;
}

// xp[t] = x^(nbits - end of thread t's chunk + order) mod g (host table, plan_fec.cu: add_xp_tx).
template <typename F>
__device__ uint32_t block_crc(F getbit, uint32_t nbits, uint32_t poly, uint32_t order, const uint32_t* __restrict__ xp,
                              uint32_t* s_acc) {
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
    if (r) r = gf2_mulmod_tx(r, xp[tid], poly, order);
    for (int o = 16; o > 0; o >>= 1) r ^= __shfl_xor_sync(0xffffffffu, r, o);
    if ((tid & 31) == 0 && r) atomicXor(s_acc, r);
    __syncthreads();
    const uint32_t out = *s_acc;
    __syncthreads();
    return out;
}

// recursive systematic convolutional encoder over bits in[perm(i)], chunk-parallel:
// the register evolution is GF(2)-linear, so chunk end states combine through the zero-input map.
// make_reader(k0) returns a callable that delivers the input bits k0, k0 + 1, ... one per call (the interleaved stream
// walks the QPP permutation incrementally instead of reducing a 64-bit polynomial per bit).
template <typename MR>
__device__ void block_rsc(MR make_reader, uint32_t K, uint8_t* par, uint8_t* xt, uint8_t* zt, uint8_t* s_state) {
    const uint32_t tid = threadIdx.x, nthr = blockDim.x;
    const uint32_t chunk = (K + nthr - 1) / nthr;
    const uint32_t k0 = min(K, tid * chunk), k1 = min(K, k0 + chunk);
    // pass 1: end state of the chunk started from the all-zero register
    uint32_t s0 = 0, s1 = 0, s2 = 0;
    {
        auto next = make_reader(k0);
        for (uint32_t k = k0; k < k1; ++k) {
            const uint32_t a = next() ^ s1 ^ s2;
            s2 = s1; s1 = s0; s0 = a;
        }
    }
    // The register evolution over a chunk is the affine map s -> Z(s) ^ e (Z = zero-input response over the chunk's
    // length, e = end state from the all-zero register).  A map is eight 3-bit images packed in 24 bits; maps compose
    // associatively, so the start state of every chunk comes out of a block-wide prefix scan of the maps.
    const uint32_t e_own = (s0 << 2) | (s1 << 1) | s2;
    if (tid < 8) {  // zero-input response of state tid over one full chunk
        uint32_t q0 = (tid >> 2) & 1, q1 = (tid >> 1) & 1, q2 = tid & 1;
        for (uint32_t k = 0; k < chunk; ++k) {
            const uint32_t a = q1 ^ q2;
            q2 = q1; q1 = q0; q0 = a;
        }
        s_state[nthr + 1 + tid] = (uint8_t)((q0 << 2) | (q1 << 1) | q2);
    }
    __syncthreads();
    constexpr uint32_t IDENT = 0xFAC688u;  // s -> s: sum of s << 3 s
    auto compose = [](uint32_t first, uint32_t then) -> uint32_t {  // s -> then(first(s))
        uint32_t r = 0;
#pragma unroll
        for (int sidx = 0; sidx < 8; ++sidx) r |= ((then >> (3 * ((first >> (3 * sidx)) & 7u))) & 7u) << (3 * sidx);
        return r;
    };
    uint32_t m = IDENT;  // map of this thread's chunk
    if (k1 > k0) {
        m = 0;
        if (k1 - k0 == chunk) {
#pragma unroll
            for (int sidx = 0; sidx < 8; ++sidx) m |= ((uint32_t)s_state[nthr + 1 + sidx] ^ e_own) << (3 * sidx);
        } else {  // short last chunk: evolve the eight states explicitly
            for (uint32_t sidx = 0; sidx < 8; ++sidx) {
                uint32_t q0 = (sidx >> 2) & 1, q1 = (sidx >> 1) & 1, q2 = sidx & 1;
                for (uint32_t k = k0; k < k1; ++k) { const uint32_t a = q1 ^ q2; q2 = q1; q1 = q0; q0 = a; }
                m |= (((q0 << 2) | (q1 << 1) | q2) ^ e_own) << (3 * sidx);
            }
        }
    }
    // inclusive scan inside the warp (lower lanes act first), warp totals through shared memory
    const uint32_t lane = tid & 31, warp = tid >> 5, nwarp = (nthr + 31) >> 5;
    uint32_t inc = m;
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t lower = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= (uint32_t)o) inc = compose(lower, inc);
    }
    uint32_t* s_wtot = reinterpret_cast<uint32_t*>(s_state + ((nthr + 16 + 3) & ~3u));  // [nwarp] behind the byte states
    if (lane == 31) s_wtot[warp] = inc;
    __syncthreads();
    uint32_t before = IDENT;  // all chunks of the lower warps
    for (uint32_t w = 0; w < nwarp; ++w)
        if (w < warp) before = compose(before, s_wtot[w]);
    uint32_t excl = __shfl_up_sync(0xffffffffu, inc, 1);
    if (lane == 0) excl = IDENT;
    excl = compose(before, excl);
    s_state[tid] = (uint8_t)(excl & 7u);                       // start state of this chunk = map of all earlier chunks at 0
    if (tid == nthr - 1) s_state[nthr] = (uint8_t)(compose(before, inc) & 7u);  // final register content -> termination
    __syncthreads();
    // pass 2: true start state, emit parity
    uint32_t st = s_state[tid];
    s0 = (st >> 2) & 1; s1 = (st >> 1) & 1; s2 = st & 1;
    {
        auto next = make_reader(k0);
        for (uint32_t k = k0; k < k1; ++k) {
            const uint32_t a = next() ^ s1 ^ s2;
            par[k] = (uint8_t)(a ^ s0 ^ s2);
            s2 = s1; s1 = s0; s0 = a;
        }
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
    const uint32_t pkt = blockIdx.y, tid = threadIdx.x, nthr = blockDim.x;
    const txcb_desc_t cb = p.cbs[blockIdx.x + (p.cb_sel ? p.cb_sel[pkt] : 0u)];
    const uint8_t* gold = p.gold + (p.gold_sel ? (size_t)p.gold_sel[pkt] * p.gold_stride : 0);
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
    if (((cb.rp | cb.n_data) & 7u) == 0) {  // byte aligned (every TB / PLCF segmentation is): one load per eight bits
        const uint8_t* inb = in + (cb.rp >> 3);
        for (uint32_t B = tid; B < cb.n_data / 8; B += nthr) {
            const uint32_t v = inb[B];
#pragma unroll
            for (uint32_t q = 0; q < 8; ++q) c[8 * B + q] = (uint8_t)((v >> (7 - q)) & 1u);
        }
    } else {
        for (uint32_t i = tid; i < cb.n_data; i += nthr) c[i] = (uint8_t)inbit(cb.rp + i);
    }
    uint32_t fill = cb.n_data;
    if (cb.tb_crc_bits) {
        // whole-message CRC (CRC24A over the TB in the last CB; CRC16 ^ mask for the PLCF)
        // (tb_crc_len is a whole number of bytes for TBs and PLCFs; the table is built for byte chunks)
        uint32_t crc = block_crc_bytes(in, cb.tb_crc_len / 8, cb.tb_crc_poly, cb.tb_crc_bits, p.xp_arena + cb.xp_tb_off, &s_acc);
        if (p.plcf_mask) crc ^= p.plcf_mask[pkt];
        if (tid < cb.tb_crc_bits) c[fill + tid] = (uint8_t)((crc >> (cb.tb_crc_bits - 1 - tid)) & 1);
        fill += cb.tb_crc_bits;
    }
    __syncthreads();
    if (cb.cb_crc) {
        const uint32_t crc = block_crc([&](uint32_t i) -> uint32_t { return c[i]; }, fill, 0x1800063u, 24,
                                       p.xp_arena + cb.xp_cb_off, &s_acc);
        if (tid < 24) c[fill + tid] = (uint8_t)((crc >> (23 - tid)) & 1);
        fill += 24;
    }
    __syncthreads();

    // ---- turbo encoder: d0 systematic, d1 parity of c, d2 parity of c[pi(i)]
    uint8_t* d0 = d; uint8_t* d1 = d + D; uint8_t* d2 = d + 2 * D;
    for (uint32_t i = tid; i < K; i += nthr) d0[i] = c[i];
    block_rsc([&](uint32_t k0) { return [&c, k = k0]() mutable -> uint32_t { return c[k++]; }; }, K, d1, s_tail, s_tail + 3,
              s_state);
    {
        // QPP walk of this thread's chunk: pi(i + 1) = pi(i) + g(i), g(i + 1) = g(i) + 2 f2 (mod K); the two 64-bit
        // reductions are done once per thread
        const uint32_t chunk = (K + nthr - 1) / nthr, i0 = min(K, tid * chunk);
        const uint32_t pi0 = (uint32_t)(((uint64_t)cb.f1 * i0 + (uint64_t)cb.f2 * i0 * i0) % K);
        const uint32_t g0 = (uint32_t)((cb.f1 + (uint64_t)cb.f2 * (2 * (uint64_t)i0 + 1)) % K);
        const uint32_t g2 = (2 * cb.f2) % K;
        block_rsc([&](uint32_t) {
            return [&c, K, g2, pi = pi0, g = g0]() mutable -> uint32_t {
                const uint32_t v = c[pi];
                pi += g; if (pi >= K) pi -= K;
                g += g2; if (g >= K) g -= K;
                return v;
            }; }, K, d2, s_tail + 6, s_tail + 9, s_state);
    }
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
                    const uint32_t Nc = cb.Nc;
                    uint32_t pos = start + (e < Nc ? e : e % Nc);   // repetition (e >= Nc): below rate 1/3 or a limited buffer
                    if (pos >= Nc) pos -= Nc;
                    bit = d[src[pos]];
                }
                const uint32_t gg = g + p.gold_bit_off;
                bit ^= (gold[gg >> 3] >> (7 - (gg & 7))) & 1u;
            }
            v = (v << 1) | bit;
        }
        if (B * 8 >= wp && B * 8 + 8 <= we) out[B] = (uint8_t)v;
        else atomicOr((unsigned int*)(out + (B & ~3u)), v << (8 * (B & 3u)));
    }
}

cudaError_t launch_tx_fec(const txfec_params_t& p, uint32_t n_packets, uint32_t Kmax, cudaStream_t st) {
    const uint32_t threads = TX_FEC_THREADS;
    // s_state: threads + 1 states + 8 zero-input entries, then (4-byte aligned) one packed map per warp
    const size_t smem = (size_t)Kmax + 3 * (Kmax + 4) + threads + 16 + 4 + 4 * ((threads + 31) / 32) + 4;
    k_tx_fec<<<dim3(p.C, n_packets), threads, smem, st>>>(p);
    return cudaGetLastError();
}

}  // namespace dnrp

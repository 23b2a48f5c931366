// fec_rx.cu -- RX side of the FEC chain on the GPU: descramble + de-rate-match + HARQ combine + turbo
// decode + CRC in ONE kernel per batch (grid = code blocks x packets), then a small per-packet kernel
// for the transport-block level CRC logic.  Reference: lib/src/phy/fec/pdc_enc.cpp:291-492,
// lib/src/phy/fec/pcc_enc.cpp:215-364 (file:line relative to /root/reference).
#include <cstdlib>
#include <type_traits>
#include "fec_rx.cuh"
#include "launch_util.hpp"
#include "turbo_dec.cuh"

namespace dnrp {
using namespace tdec;
// k_derm: LLR loads in flight per thread in the scatter loop (the loop waits on HBM latency: ncu long_scoreboard 17.6)
constexpr int DERM_UNROLL = 4;   // 8: 64 instead of 40 registers, fewer resident blocks: 1.74 -> 1.79 ms (measured)

__device__ __forceinline__ int clampi(int v, int lim) { return max(-lim, min(lim, v)); }

// a*b mod g over GF(2), operands < 2^order
__device__ __forceinline__ uint32_t gf2_mulmod(uint32_t a, uint32_t b, uint32_t poly, uint32_t order) {
    uint32_t r = 0;
    const uint32_t top = 1u << order;
    for (uint32_t i = 0; i < order; ++i) {
        r <<= 1;
        if (r & top) r ^= poly;
        if ((b >> (order - 1 - i)) & 1) r ^= a;
    }
    return r;
}

// ---- descrambling + de-rate-matching + HARQ soft combining, one block per (code block, packet):
// soft[n] (+)= sum of the descrambled LLRs that rate matching mapped onto circular buffer position n
// (srsran_scrambling_s_offset + srsran_rm_turbo_rx_lut_, reference pdc_enc.cpp:339-363); plain int16
// wrap-around adds.  A streaming kernel of its own so that it runs at full occupancy: inside the decoder
// (three warps per block) its dependent loads were the largest stall source (profiles/).
// The code block's LLRs are first descrambled into shared memory with coalesced loads; the table driven
// gather then reads shared memory.  cap = int16 capacity of the dynamic shared buffer.
__global__ void __launch_bounds__(256) k_derm(const tdec_params_t p, uint32_t cap) {
    extern __shared__ __align__(16) uint32_t smem[];
    const cb_desc_t cb = p.cbs[blockIdx.x];
    const uint32_t pkt = blockIdx.y;
    const uint32_t slot = p.slot ? p.slot[pkt] : pkt;
    const uint32_t tid = threadIdx.x, nthr = blockDim.x;
    const uint32_t N = 3 * (cb.K + 4);
    const bool reset = p.reset ? p.reset[pkt] != 0 : true;
    const uint32_t rv = p.rv ? p.rv[pkt] : 0;
    const uint8_t* gold = p.gold + (p.gold_sel ? (size_t)p.gold_sel[pkt] * p.gold_stride : 0);
    // code block already decoded in an earlier transmission: nothing to combine (pdc_enc.cpp:404-406)
    if (!reset && p.cb_crc[(size_t)slot * p.C + blockIdx.x]) return;

    int16_t* soft = p.soft + (size_t)slot * p.soft_stride + cb.soft_off;
    const int16_t* llr = p.llr + (size_t)pkt * p.llr_stride + cb.rp;
    const uint16_t* rm_pos = p.rm_arena + cb.rm_off + (size_t)rv * N;  // table of this redundancy version
    int16_t* DSC = (int16_t*)smem;
    const bool staged = cb.n_e <= cap;
    // rate >= 1/3 (n_e <= N) and room for N entries: zero pad the staged LLRs up to N, every circular-buffer
    // position then reads exactly one entry (punctured positions read a zero) -- no loop, no bounds test
    const bool single = staged && cb.n_e <= cb.Nc && N <= cap;
    auto descr = [&](uint32_t e, int v) -> int {   // srsran_scrambling_s_offset: sign flip where the Gold bit is 1
        const uint32_t g = cb.rp + e + p.gold_bit_off;
        return ((gold[g >> 3] >> (7 - (g & 7))) & 1) ? -v : v;
    };
    if (single && (reinterpret_cast<uintptr_t>(llr) & 3) == 0) {
        // ---- usual case (code rate >= 1/3, LLR pairs aligned): build the soft-buffer image in shared memory by
        // SCATTERING the n_e descrambled LLRs through the inverse table into a zeroed image (n_e random 2-byte stores
        // instead of N random 2-byte loads: N / n_e = 2.5 at R = 5/6), then stream the image out with 8-byte accesses
        uint4* Z = reinterpret_cast<uint4*>(DSC);
        for (uint32_t i = tid; i < (N + 7) / 8; i += nthr) Z[i] = make_uint4(0u, 0u, 0u, 0u);   // cap >= N rounded up to 8
        __syncthreads();
        const uint32_t* llr2 = reinterpret_cast<const uint32_t*>(llr);
        const uint16_t* inv = p.rm_arena + cb.rm_inv_off + (size_t)rv * N;
        const uint32_t g0 = cb.rp + p.gold_bit_off;
#pragma unroll DERM_UNROLL
        for (uint32_t e2 = tid; e2 < cb.n_e / 2; e2 += nthr) {
            const uint32_t w = llr2[e2];
            const uint32_t g = g0 + 2 * e2;
            const uint32_t gw = ((uint32_t)gold[g >> 3] << 8) | gold[(g >> 3) + 1];
            const uint32_t b0 = (gw >> (15 - (g & 7))) & 1u, b1 = (gw >> (14 - (g & 7))) & 1u;
            const uint32_t m = (b0 ? 0xffffu : 0u) | (b1 ? 0xffff0000u : 0u);
            const uint32_t v = __vsub2(w ^ m, m);
            DSC[inv[2 * e2]] = (int16_t)(v & 0xffffu);
            DSC[inv[2 * e2 + 1]] = (int16_t)(v >> 16);
        }
        if ((cb.n_e & 1) && tid == 0) DSC[inv[cb.n_e - 1]] = (int16_t)descr(cb.n_e - 1, llr[cb.n_e - 1]);
        __syncthreads();
        const uint2* img = reinterpret_cast<const uint2*>(DSC);
        uint2* soft4 = reinterpret_cast<uint2*>(soft);
        for (uint32_t q = tid; q < N / 4; q += nthr) {
            uint2 v = img[q];
            if (!reset) {
                const uint2 old = soft4[q];
                v = make_uint2(__vadd2(v.x, old.x), __vadd2(v.y, old.y));   // int16 wrap-around adds, as below
            }
            soft4[q] = v;
        }
        return;
    }
    if (staged) {
        if ((reinterpret_cast<uintptr_t>(llr) & 3) == 0) {
            // two LLRs per 4-byte load
            const uint32_t* llr2 = reinterpret_cast<const uint32_t*>(llr);
            uint32_t* DSC2 = reinterpret_cast<uint32_t*>(DSC);
            // the two Gold bits of a pair come from one 16-bit window of the sequence (independent of the LLR load)
            const uint32_t g0 = cb.rp + p.gold_bit_off;
#pragma unroll 4
            for (uint32_t e2 = tid; e2 < cb.n_e / 2; e2 += nthr) {
                const uint32_t w = llr2[e2];
                const uint32_t g = g0 + 2 * e2;
                const uint32_t gw = ((uint32_t)gold[g >> 3] << 8) | gold[(g >> 3) + 1];
                const uint32_t b0 = (gw >> (15 - (g & 7))) & 1u, b1 = (gw >> (14 - (g & 7))) & 1u;
                // sign flip of each half where its bit is set: (v ^ m) - m with m = 0xffff per flipped half
                const uint32_t m = (b0 ? 0xffffu : 0u) | (b1 ? 0xffff0000u : 0u);
                DSC2[e2] = __vsub2(w ^ m, m);
            }
            if ((cb.n_e & 1) && tid == 0) DSC[cb.n_e - 1] = (int16_t)descr(cb.n_e - 1, llr[cb.n_e - 1]);
        } else {
            for (uint32_t e = tid; e < cb.n_e; e += nthr) DSC[e] = (int16_t)descr(e, llr[e]);
        }
        if (single) {  // zero padding: a few int16 up to the next 16-byte boundary, then 16-byte stores
            const uint32_t z8 = min((cb.n_e + 7u) & ~7u, N);
            for (uint32_t e = cb.n_e + tid; e < z8; e += nthr) DSC[e] = 0;
            uint4* Z = reinterpret_cast<uint4*>(DSC);
            for (uint32_t e8 = z8 / 8 + tid; e8 < N / 8; e8 += nthr) Z[e8] = make_uint4(0u, 0u, 0u, 0u);
            for (uint32_t e = (N & ~7u) + tid; e < N; e += nthr) if (e >= z8) DSC[e] = 0;
        }
        __syncthreads();
    }
    // four consecutive circular-buffer positions per thread: 8-byte table loads and soft-buffer stores (N is a
    // multiple of 4 for every turbo block size, table and soft-buffer offsets are multiples of 4 elements)
    const uint2* rm4 = reinterpret_cast<const uint2*>(rm_pos);
    uint2* soft4 = reinterpret_cast<uint2*>(soft);
#pragma unroll 2
    for (uint32_t q = tid; q < N / 4; q += nthr) {
        const uint2 rp = rm4[q];
        uint2 old = make_uint2(0u, 0u);
        if (!reset) old = soft4[q];
        const uint32_t pos[4] = {rp.x & 0xffffu, rp.x >> 16, rp.y & 0xffffu, rp.y >> 16};
        const int prev[4] = {(int)(int16_t)(old.x & 0xffff), (int)(int16_t)(old.x >> 16), (int)(int16_t)(old.y & 0xffff),
                             (int)(int16_t)(old.y >> 16)};
        int acc[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const uint32_t v = pos[u];
            acc[u] = prev[u];
            if (v == 0xffffu) continue;   // outside a limited circular buffer: never sent
            if (single) {
                acc[u] += DSC[v];
            } else if (staged) {
                for (uint32_t ee = v; ee < cb.n_e; ee += cb.Nc) acc[u] += DSC[ee];
            } else {
                for (uint32_t ee = v; ee < cb.n_e; ee += cb.Nc) acc[u] += descr(ee, llr[ee]);
            }
        }
        soft4[q] = make_uint2(((uint32_t)acc[0] & 0xffffu) | ((uint32_t)acc[1] << 16),
                              ((uint32_t)acc[2] & 0xffffu) | ((uint32_t)acc[3] << 16));
    }
}

cudaError_t launch_derm(const tdec_params_t& p, uint32_t n_packets, uint32_t max_n_e, cudaStream_t st) {
    // shared buffer for the largest code block's LLRs, at most 48 KB (no opt-in); larger ones gather from global
    // (callers pass max(n_e, 3 (K + 4)) where that fits, so that the kernel's "single" path applies)
    const uint32_t cap = max_n_e < 24576u ? ((max_n_e + 7u) & ~7u) : 24576u;
    k_derm<<<dim3(p.C, n_packets), 256, (size_t)cap * 2, st>>>(p, cap);
    return cudaGetLastError();
}

// One thread GROUP of tg threads decodes one item = (code block, packet); a block holds
// blockDim.x / tg groups.  Large code blocks: one group per block (tg = blockDim.x, a multiple of 32).
// Small ones (PLCF, K < ~2000: fewer than 32 window pairs): tg is a power of two <= 32 and many items
// share a block, all groups moving through the same barriers in lockstep.
// BULK (one item per block only): the streams of the soft buffer are staged by the copy engine (cp.async.bulk + mbarrier,
// turbo_dec.cuh) and clamped in place, instead of a load / clamp / store loop per thread; the copy of the next half
// iteration's parity stream is issued as soon as the shared arrays are free, so it overlaps the boundary / CRC work.
// CKALIAS (one item per block only): the first two checkpoints of the backward pass are kept in the running decoder's
// rows of AE / BS (dead between the boundary fetch and the store of the new boundaries) at the price of one more barrier
// per half iteration: 14 rows of T words less, which lets K = 6080 ... 6144 keep four resident blocks per SM.  Sizes that
// fit four blocks anyway run without it (the barrier and seven more registers cost 6 % there; launch_turbo_decode).
template <bool GROUPED, bool BULK = false, bool CKALIAS = false>
__global__ void __launch_bounds__(GROUPED ? 128 : 96, GROUPED ? 1 : 4) k_turbo_decode(const tdec_params_t p, uint32_t tg, uint32_t smem_item_words,
                                                      uint32_t n_items) {
    static_assert(!(GROUPED && BULK), "bulk staging is for one item per block");
    extern __shared__ __align__(16) uint32_t smem_all[];
    __shared__ uint32_t s_crc_all[GROUPED ? 128 : 1];
    __shared__ __align__(8) uint64_t s_bar;   // BULK: completion barrier of the staging copies
    __shared__ int s_tail_all[(GROUPED ? 128 : 1) * 16];
    const uint32_t grp = GROUPED ? threadIdx.x / tg : 0;
    const uint32_t tid = GROUPED ? threadIdx.x % tg : threadIdx.x, nthr = GROUPED ? tg : blockDim.x;
    const uint32_t item = GROUPED ? blockIdx.x * (blockDim.x / tg) + grp : blockIdx.x;
    const bool valid = item < n_items;
    const uint32_t cbi = valid ? item % p.C : 0, pkt = valid ? item / p.C : 0;
    const cb_desc_t cb = p.cbs[cbi];
    uint32_t* smem = smem_all + (size_t)grp * smem_item_words;
    uint32_t& s_crc = s_crc_all[grp];
    int* s_tail = s_tail_all + grp * 16;
    const uint32_t slot = p.slot ? p.slot[pkt] : pkt;
    const uint32_t K = cb.K, L = cb.L, T = cb.P / 2, KH = K / 2, D = K + 4;
    const uint32_t nck = (L + S_CK - 1) / S_CK;
    const bool reset = p.reset ? p.reset[pkt] != 0 : true;

    uint8_t* cb_crc = p.cb_crc + (size_t)slot * p.C + cbi;
    uint8_t* saved = p.saved ? p.saved + (size_t)slot * p.saved_stride + cb.saved_off : nullptr;
    uint8_t* tb_out = p.tb_out + (size_t)pkt * p.tb_stride + cb.out_byte;
    uint8_t* n_iter_out = p.n_iter + (size_t)pkt * p.C + cbi;

    // ---- code block already decoded in an earlier transmission: copy the saved bytes
    // (reference pdc_enc.cpp:404-406); the group then idles through the barriers below
    bool done = !valid;
    if (valid && !reset && *cb_crc) {
        for (uint32_t i = tid; i < cb.rlen_byte; i += nthr) tb_out[i] = saved[i];
        if (tid == 0) *n_iter_out = 0;
        done = true;
    }
    const bool skipped = done;
    if (!GROUPED && done) return;  // single item per block: nothing left to do
    if (BULK && threadIdx.x == 0) mbar_init(&s_bar, 1);   // visible to the block after the barrier below

    // the systematic stream stays in the (L2 resident) soft buffer and is re-read once per half
    // iteration: one K/2-word array less per block = one more resident block per SM
    uint32_t* EXT = smem;
    uint32_t* XIN = EXT + KH;
    uint32_t* PAR = XIN + KH;
    const uint32_t nck_own = CKALIAS ? (nck > 2 ? nck - 2 : 0) : nck;   // CKALIAS: checkpoints 0 and 1 live in AE / BS (ck_rows)
    uint32_t* CK = PAR + KH;                 // [nck_own][7][T]
    uint32_t* AE = CK + nck_own * 7 * T;     // [2][7][T]
    uint32_t* BS = AE + 2 * 7 * T;           // [2][7][T]
    // The hard decisions need no array of their own: the second decoder leaves 2 e + (LLR > 0) in EXT (the extrinsic
    // values are clamped to +-2047, so the doubled value fits an int16); the CRC and output stages read bit 0 of each
    // half, the next iteration's staging halves the words back (ext_strip).  K = 5952: 59.5 -> 56.6 KB per block, i.e.
    // four instead of three resident blocks per SM.
    // checkpoint c of the running decoder dec: c = 0 in its rows of AE, c = 1 in its rows of BS, the others in CK
    auto ck_rows = [&](uint32_t dec, uint32_t c) -> uint32_t* {
        if (!CKALIAS) return CK + c * 7 * T;
        return c == 0 ? AE + dec * 7 * T : (c == 1 ? BS + dec * 7 * T : CK + (c - 2) * 7 * T);
    };
    auto ext_strip = [](uint32_t w) { return ((w >> 1) & 0x7fff7fffu) | (w & 0x80008000u); };   // per half: >> 1, signed
    auto strip4 = [&](uint4 v) { return make_uint4(ext_strip(v.x), ext_strip(v.y), ext_strip(v.z), ext_strip(v.w)); };

    // soft buffer of this code block, written by k_derm in the layout of plan_fec.cu: three streams of K/2 packed
    // words in shared-array order, then the 12 termination values
    const int16_t* soft = p.soft + (size_t)slot * p.soft_stride + cb.soft_off;
    // the whole soft buffer of the code block towards L2 now (k_derm's writes went through to HBM): the staging loops
    // of the first iteration then wait for one HBM latency instead of one per loop iteration
    if (!done) {
        const char* sb = reinterpret_cast<const char*>(soft);
        for (uint32_t o = tid * 128u; o < (3 * K + 12) * 2u; o += nthr * 128u)
            asm volatile("prefetch.global.L2 [%0];" ::"l"(sb + o));
    }
    for (uint32_t i = tid; i < KH; i += nthr) EXT[i] = 0;
    for (uint32_t i = tid; i < 4 * 7 * T; i += nthr) AE[i] = 0;  // AE and BS are contiguous
    // ---- termination metrics of the two constituent decoders (three tail steps back from state 0), thread dec of
    // the group computes those of decoder dec (a one-thread group both)
    for (uint32_t dec = tid; dec < 2 && !done; dec += nthr) {
        // termination values d_s[K + j] at soft[3 K + 4 s + j] (plan_fec.cu)
        const int16_t* d0 = soft + 3 * K; const int16_t* d1 = d0 + 4; const int16_t* d2 = d0 + 8;
        int xt[3], zt[3];
        if (dec == 0) {
            xt[0] = d0[0]; zt[0] = d1[0]; xt[1] = d2[0]; zt[1] = d0[1]; xt[2] = d1[1]; zt[2] = d2[1];
        } else {
            xt[0] = d0[2]; zt[0] = d1[2]; xt[1] = d2[2]; zt[1] = d0[3]; xt[2] = d1[3]; zt[2] = d2[3];
        }
        int b[8], nb[8];
#pragma unroll
        for (int s = 0; s < 8; ++s) b[s] = s == 0 ? 0 : -8192;
#pragma unroll
        for (int tt = 2; tt >= 0; --tt) {
            const int x = clampi(xt[tt], 1023), z = clampi(zt[tt], 1023);
            // allowed transition: u = s1^s2 (a = 0), parity z = s0^s2, next = (0, s0, s1)
            nb[0] = b[0];          nb[1] = b[0] + x + z;  nb[2] = b[1] + x;      nb[3] = b[1] + z;
            nb[4] = b[2] + z;      nb[5] = b[2] + x;      nb[6] = b[3] + x + z;  nb[7] = b[3];
#pragma unroll
            for (int s = 0; s < 8; ++s) b[s] = nb[s] - nb[0];
        }
#pragma unroll
        for (int s = 0; s < 8; ++s) s_tail[dec * 8 + s] = b[s];
    }
    __syncthreads();

    // stage the parity stream of a constituent decoder (and, for the first decoder, the systematic stream plus the
    // a-priori values in EXT) into the shared arrays, clamped to the channel range.  k_derm wrote the streams in the
    // layout of the shared arrays (plan_fec.cu: word a = k * T + t), so this is a straight 16-byte copy.
    auto clamp4 = [](uint4 v) {
        return make_uint4(vclamp(v.x, 1023), vclamp(v.y, 1023), vclamp(v.z, 1023), vclamp(v.w, 1023));
    };
    auto stage = [&](uint32_t stream, bool with_x) {
        const uint4* p4 = reinterpret_cast<const uint4*>(soft + (size_t)stream * K);
        const uint4* x4 = reinterpret_cast<const uint4*>(soft);
        uint4* PAR4 = reinterpret_cast<uint4*>(PAR);
        uint4* XIN4 = reinterpret_cast<uint4*>(XIN);
        uint4* EXT4 = reinterpret_cast<uint4*>(EXT);
#pragma unroll 1  // measured: any unrolling here is slower (instruction-cache bound kernel)
        for (uint32_t i4 = tid; i4 < KH / 4; i4 += nthr) {
            PAR4[i4] = clamp4(p4[i4]);
            if (with_x) {
                const uint4 x = clamp4(x4[i4]), e = strip4(EXT4[i4]);
                EXT4[i4] = e;
                XIN4[i4] = make_uint4(vadd(x.x, e.x), vadd(x.y, e.y), vadd(x.z, e.z), vadd(x.w, e.w));
            }
        }
    };

    // BULK: thread 0 issues the copies of a half iteration (parity stream, with_x: also the systematic stream into XIN);
    // all threads wait for them and clamp / add the a-priori values in place (each thread its own words)
    uint32_t bar_phase = 0;
    auto stage_issue = [&](uint32_t stream, bool with_x) {
        if (threadIdx.x == 0) {
            fence_proxy_async();   // the block's earlier reads / writes of PAR and XIN (ordered by the barrier before)
            mbar_expect_tx(&s_bar, (with_x ? 2u : 1u) * K * 2u);
            bulk_g2s(PAR, soft + (size_t)stream * K, K * 2u, &s_bar);
            if (with_x) bulk_g2s(XIN, soft, K * 2u, &s_bar);
        }
    };
    auto stage_finish = [&](bool with_x) {
        mbar_wait(&s_bar, bar_phase);
        bar_phase ^= 1u;
        uint4* PAR4 = reinterpret_cast<uint4*>(PAR);
        uint4* XIN4 = reinterpret_cast<uint4*>(XIN);
        uint4* EXT4 = reinterpret_cast<uint4*>(EXT);
#pragma unroll 1
        for (uint32_t i4 = tid; i4 < KH / 4; i4 += nthr) {
            PAR4[i4] = clamp4(PAR4[i4]);
            if (with_x) {
                const uint4 x = clamp4(XIN4[i4]), e = strip4(EXT4[i4]);
                EXT4[i4] = e;
                XIN4[i4] = make_uint4(vadd(x.x, e.x), vadd(x.y, e.y), vadd(x.z, e.z), vadd(x.w, e.w));
            }
        }
    };
    if (BULK) stage_issue(1, true);

    const uint32_t t = tid;
    const uint32_t wl = t, wh = t + T, P = cb.P;
    // QPP position pi(j0) of this thread's first step j0 = t * L and its first difference g = pi(j0 + 1) - pi(j0)
    // (second difference g2 = 2 f2): the 64-bit reductions are done once per code block, not per half iteration
    const uint32_t qpp_j0 = t * L;
    const uint32_t qpp_pi0 = (uint32_t)(((uint64_t)cb.f1 * qpp_j0 + (uint64_t)cb.f2 * qpp_j0 * qpp_j0) % K);
    const uint32_t qpp_g0 = (uint32_t)((cb.f1 + (uint64_t)cb.f2 * (2 * qpp_j0 + 1)) % K);
    const uint32_t qpp_g2 = (2 * cb.f2) % K;
    uint32_t n_iter = 0;
    bool crc_ok = false;
    uint32_t plcf_mask = 0;

    for (uint32_t it = 0; it < p.max_it; ++it) {
        if (GROUPED) {
            if (!__syncthreads_or(!done)) break;  // every group of the block has finished
        } else if (done) {
            break;
        }
        const bool work = GROUPED ? !done : true;  // uniform within a group
        const bool active = work && tid < T;
#pragma unroll 1  // one copy of the half-iteration body: the kernel is instruction-cache bound otherwise
        for (uint32_t dec = 0; dec < 2; ++dec) {
            // ---- parity of this constituent decoder -> shared, window-transposed
            // (first decoder: with its input x = systematic + a-priori in natural order)
            if (BULK) stage_finish(dec == 0);
            else if (work) stage(1 + dec, dec == 0);
            // ---- input of the second decoder, interleaved order
            if (dec != 0 && active) {
                uint32_t pi = qpp_pi0, g = qpp_g0;
                const uint32_t g2 = qpp_g2;
#pragma unroll 4
                for (uint32_t k = 0; k < L; ++k) {
                    const bool up = pi >= KH;
                    const uint32_t pw = up ? pi - KH : pi;
                    const uint32_t q = __umulhi(pw, cb.Linv);
                    const uint32_t a = (pw - q * L) * T + q;
                    uint32_t w = EXT[a];  // systematic + extrinsic, left there by the first decoder's LLR pass
                    if (up) w = swap16(w);
                    XIN[k * T + t] = w;
                    pi += g; if (pi >= K) pi -= K;
                    g += g2; if (g >= K) g -= K;
                }
            }
            __syncthreads();

            uint32_t ae_new[8], bs_new[8];
            uint32_t b[8], a[8];
            int tail[8];
            if (active) {
                // ---- termination metrics of this decoder, computed once per code block (s_tail)
#pragma unroll
                for (int s = 0; s < 8; ++s) tail[s] = s_tail[dec * 8 + s];
                const uint32_t* BSd = BS + dec * 7 * T;
                const uint32_t* AEd = AE + dec * 7 * T;
                // boundary metric of window v as int16: word v (low half) for v < T, word v - T (high half)
                // otherwise; row s - 1 of the [7][T] word array
                auto hidx = [&](uint32_t v) -> uint32_t { return v < T ? 2 * v : 2 * (v - T) + 1; };
                const bool b_lo_end = wl + 2 >= P, b_hi_end = wh + 2 >= P, a_lo_beg = wl <= 1, a_hi_beg = wh <= 1;
                const int16_t* pb_lo = (const int16_t*)BSd + hidx(b_lo_end ? wl : wl + 2);
                const int16_t* pb_hi = (const int16_t*)BSd + hidx(b_hi_end ? wh : wh + 2);
                const int16_t* pa_lo = (const int16_t*)AEd + hidx(a_lo_beg ? wl : wl - 2);
                const int16_t* pa_hi = (const int16_t*)AEd + hidx(a_hi_beg ? wh : wh - 2);

                // ---- training passes: beta over window w+1 and alpha over window w-1 are independent
                // recursions, run in one loop (two dependency chains in flight) with the next step's
                // shared-memory words prefetched one iteration ahead
                b[0] = pack(b_lo_end ? tail[0] : 0, b_hi_end ? tail[0] : 0);
                a[0] = 0;
#pragma unroll
                for (int s = 1; s < 8; ++s) {
                    const uint32_t row = (s - 1) * 2 * T;
                    const int lo = b_lo_end ? tail[s] : (int)pb_lo[row];
                    const int hi = b_hi_end ? tail[s] : (int)pb_hi[row];
                    b[s] = pack(lo, hi);
                    const int alo = a_lo_beg ? -8192 : (int)pa_lo[row];
                    const int ahi = a_hi_beg ? -8192 : (int)pa_hi[row];
                    a[s] = pack(alo, ahi);
                }
            }
            // CKALIAS: every boundary metric of this decoder is in registers now; until the new ones are stored at the end
            // of the half iteration its rows of AE and BS hold the first two checkpoints of the backward pass (ck_rows)
            if (CKALIAS) __syncthreads();
            if (active) {
                {
                    const bool wrapn = (t + 1 == T), wrapp = (t == 0);
                    const uint32_t* Xn = XIN + (wrapn ? 0 : t + 1);
                    const uint32_t* Yn = PAR + (wrapn ? 0 : t + 1);
                    const uint32_t* Xp = XIN + (wrapp ? T - 1 : t - 1);
                    const uint32_t* Yp = PAR + (wrapp ? T - 1 : t - 1);
                    const uint32_t shn = wrapn ? 16 : 0, shp = wrapp ? 16 : 0;
                    // running pointers; the prefetch of the last iteration reads one row outside XIN / PAR,
                    // which is still inside this item's shared arrays (EXT | XIN | PAR | CK) and unused
                    const uint32_t* pxb = Xn + (size_t)(L - 1) * T;
                    const uint32_t* pyb = Yn + (size_t)(L - 1) * T;
                    uint32_t xb = *pxb, yb = *pyb, xa = *Xp, ya = *Yp;
                    // three steps without normalisation, the fourth with (turbo_dec.cuh: overflow bound); a remainder
                    // of L % 4 steps is normalised every step, so the pass always ends normalised
                    auto train_step = [&](auto norm_tag) {
                        constexpr bool NORM = decltype(norm_tag)::value;
                        pxb -= T; pyb -= T; Xp += T; Yp += T;
                        const uint32_t xb_n = *pxb, yb_n = *pyb, xa_n = *Xp, ya_n = *Yp;
                        beta_step<NORM>(b, xb >> shn, yb >> shn);
                        alpha_step<NORM>(a, xa << shp, ya << shp);
                        xb = xb_n; yb = yb_n; xa = xa_n; ya = ya_n;
                    };
                    uint32_t i = 0;
#pragma unroll 1
                    for (; i + 4 <= L; i += 4) {
                        train_step(std::false_type{}); train_step(std::false_type{}); train_step(std::false_type{});
                        train_step(std::true_type{});
                    }
#pragma unroll 1
                    for (; i < L; ++i) train_step(std::true_type{});
                    if (wh == P - 1) {
#pragma unroll
                        for (int s = 0; s < 8; ++s) b[s] = (b[s] & 0xffffu) | ((uint32_t)tail[s] << 16);
                    }
                    if (wl == 0) {
#pragma unroll
                        for (int s = 0; s < 8; ++s) a[s] = (a[s] & 0xffff0000u) | (s == 0 ? 0u : 0xE000u);
                    }
                }
                // ---- backward pass over the own windows, checkpoint every S_CK steps
                {
                    const uint32_t* X = XIN + t + (size_t)(L - 1) * T;
                    const uint32_t* Y = PAR + t + (size_t)(L - 1) * T;
                    uint32_t x = *X, y = *Y;
                    // normalised after the steps k with k % 4 == 0: every checkpoint ((k + 1) % S_CK == 0) and the final
                    // metrics are normalised values, at most three steps in a row are not
                    auto back_step = [&](int k, auto norm_tag) {
                        constexpr bool NORM = decltype(norm_tag)::value;
                        X -= T; Y -= T;  // last prefetch: one row below, inside EXT / XIN, unused
                        const uint32_t x_n = *X, y_n = *Y;
                        if (((k + 1) % S_CK) == 0 || k + 1 == (int)L) {
                            uint32_t* ckc = ck_rows(dec, k / S_CK);
#pragma unroll
                            for (int s = 1; s < 8; ++s) ckc[(s - 1) * T + t] = b[s];
                        }
                        beta_step<NORM>(b, x, y);
                        x = x_n; y = y_n;
                    };
                    int k = (int)L - 1;
#pragma unroll 1
                    for (; (k & 3) != 3 && k >= 0; --k) back_step(k, std::true_type{});
#pragma unroll 1
                    for (; k >= 3; k -= 4) {
                        back_step(k, std::false_type{}); back_step(k - 1, std::false_type{});
                        back_step(k - 2, std::false_type{}); back_step(k - 3, std::true_type{});
                    }
                }
#pragma unroll
                for (int s = 0; s < 8; ++s) bs_new[s] = b[s];
                // QPP position of the first step of this thread (decoder 2 scatters to natural order)
                uint32_t pi = qpp_pi0, g = qpp_g0;
                const uint32_t g2 = qpp_g2;
                // one chunk of S_CK steps; FULL = compile-time knowledge that the chunk is complete (always the case when
                // L is a multiple of S_CK, e.g. L = 32): no per-step bounds predicates in the unrolled body
                auto chunk = [&](uint32_t c, auto full_tag) {
                    constexpr bool FULL = decltype(full_tag)::value;
                    const int k0 = c * S_CK;
                    const int n = FULL ? (int)S_CK : min((int)S_CK, (int)L - k0);
                    uint32_t bb[S_CK][8], xs[S_CK], ys[S_CK];
                    const uint32_t* ckc = ck_rows(dec, c);
#pragma unroll
                    for (int i = 0; i < S_CK; ++i) {
                        bb[i][0] = 0;
                        const int ki = k0 + (i < n ? i : 0);  // the chunk's inputs, loaded once
                        xs[i] = XIN[ki * T + t];
                        ys[i] = PAR[ki * T + t];
                    }
                    // beta at time k0+n from the checkpoint, then recompute down to k0+1
#pragma unroll
                    for (int i = S_CK - 1; i >= 0; --i) {
                        if (i == n - 1) {
#pragma unroll
                            for (int s = 1; s < 8; ++s) bb[i][s] = ckc[(s - 1) * T + t];
                        } else if (i < n - 1) {
#pragma unroll
                            for (int s = 0; s < 8; ++s) bb[i][s] = bb[i + 1][s];
                            constexpr int S1 = S_CK - 1;
                            beta_step(bb[i], xs[i < S1 ? i + 1 : S1], ys[i < S1 ? i + 1 : S1]);
                        }
                    }
#pragma unroll
                    for (int i = 0; i < S_CK; ++i) {
                        if (i < n) {
                            const uint32_t k = k0 + i;
                            const uint32_t x = xs[i], y = ys[i];
                            const uint32_t l = llr_step(a, bb[i], x, y);
                            const uint32_t e = vclamp(vsub(l, x), 2047);
                            if (dec == 0) {
                                // x = systematic + a-priori (EXT): store extrinsic + systematic, the second decoder's
                                // input in natural order -- its gather then reads shared memory only
                                EXT[k * T + t] = vadd(e, vsub(x, EXT[k * T + t]));
                            } else {
                                const bool up = pi >= KH;
                                const uint32_t pw = up ? pi - KH : pi;
                                const uint32_t q = __umulhi(pw, cb.Linv);
                                const uint32_t ad = (pw - q * L) * T + q;
                                // extrinsic doubled, hard decision (LLR > 0) in bit 0 of each half
                                const uint32_t w = vadd(e, e) | (__vcmpgts2(l, 0u) & 0x00010001u);
                                EXT[ad] = up ? swap16(w) : w;
                                pi += g; if (pi >= K) pi -= K;
                                g += g2; if (g >= K) g -= K;
                            }
                            alpha_step(a, x, y);
                        }
                    }
                };
                const uint32_t n_full = L / S_CK;
#pragma unroll 1
                for (uint32_t c = 0; c < n_full; ++c) chunk(c, std::true_type{});
                if (n_full < nck) chunk(n_full, std::false_type{});
#pragma unroll
                for (int s = 0; s < 8; ++s) ae_new[s] = a[s];
            }
            if (tid == 0) s_crc = 0;
            __syncthreads();
            if (BULK && dec == 0) stage_issue(2, false);   // PAR is free: the second decoder's parity stream
            if (active) {
#pragma unroll
                for (int s = 1; s < 8; ++s) {
                    AE[(dec * 7 + s - 1) * T + t] = ae_new[s];
                    BS[(dec * 7 + s - 1) * T + t] = bs_new[s];
                }
            }
        }
        // ---- CRC of the hard decisions (natural order), remainder of the whole block mod g
        if (work) ++n_iter;
        {
            uint32_t r = 0;
            // below the minimum iteration count the remainder cannot stop the decoder: not computed (crc_mode 0)
            if (active && (p.crc_mode != 0 || n_iter >= p.min_it)) {
                uint32_t r_lo = 0, r_hi = 0;
                const uint32_t top = 1u << cb.crc_order;
                for (uint32_t k = 0; k < L; ++k) {
                    const uint32_t hb = EXT[k * T + t];
                    r_lo = (r_lo << 1) | (hb & 1);
                    if (r_lo & top) r_lo ^= cb.crc_poly;
                    r_hi = (r_hi << 1) | ((hb >> 16) & 1);
                    if (r_hi & top) r_hi ^= cb.crc_poly;
                }
                const uint32_t* xp = p.xp_arena + cb.xp_off;
                r = gf2_mulmod(r_lo, xp[wl], cb.crc_poly, cb.crc_order) ^
                    gf2_mulmod(r_hi, xp[wh], cb.crc_poly, cb.crc_order);
            }
            // xor over the group: whole warps, or an aligned power-of-two slice of one warp
            const uint32_t width = GROUPED ? (tg >= 32 ? 32u : tg) : 32u;
            for (int o = (int)width / 2; o > 0; o >>= 1) r ^= __shfl_xor_sync(0xffffffffu, r, o);
            if ((tid & (width - 1)) == 0 && r) atomicXor(&s_crc, r);
        }
        __syncthreads();
        if (work) {
            const uint32_t rem = s_crc;
            if (p.crc_mode == 0) {
                crc_ok = (rem == 0) && n_iter >= p.min_it;
            } else {  // PLCF: remainder equals the CL/BF mask (pcc_enc.cpp:326-346)
                crc_ok = (rem == 0x0000 || rem == 0x5555 || rem == 0xAAAA || rem == 0xFFFF);
                plcf_mask = rem;
            }
            done = crc_ok;
        }
        if (BULK && !done && it + 1 < p.max_it) stage_issue(1, true);   // next iteration (nothing may be left in flight at exit)
        __syncthreads();
    }
    if (skipped) return;  // no barrier below

    // ---- outputs: packed bytes (MSB first), flags
    if (L == 32) {
        // a window is one 32-bit word of the code block: every thread packs its own two windows (conflict-free column
        // reads, no index arithmetic) and stores their four bytes each
        if (tid < T) {
            uint32_t w_lo = 0, w_hi = 0;
#pragma unroll 4
            for (uint32_t k = 0; k < 32; ++k) {
                const uint32_t hb = EXT[k * T + t];
                w_lo = (w_lo << 1) | (hb & 1u);
                w_hi = (w_hi << 1) | ((hb >> 16) & 1u);
            }
#pragma unroll
            for (uint32_t j = 0; j < 8; ++j) {
                const uint32_t bi = (j < 4 ? 0u : KH / 8) + 4 * t + (j & 3u);
                const uint8_t v = (uint8_t)((j < 4 ? w_lo : w_hi) >> (24 - 8 * (j & 3u)));
                if (bi < cb.rlen_byte) {
                    tb_out[bi] = v;
                    if (saved && crc_ok) saved[bi] = v;
                }
            }
        }
    } else
    for (uint32_t bi = tid; bi < K / 8; bi += nthr) {
        uint32_t v = 0;
#pragma unroll
        for (uint32_t q = 0; q < 8; ++q) {
            const uint32_t pos = bi * 8 + q;
            const uint32_t pw = pos >= KH ? pos - KH : pos;
            const uint32_t qq = __umulhi(pw, cb.Linv);
            const uint32_t hb = EXT[(pw - qq * L) * T + qq];
            v = (v << 1) | (pos >= KH ? ((hb >> 16) & 1u) : (hb & 1u));
        }
        if (bi < cb.rlen_byte) {
            tb_out[bi] = (uint8_t)v;
            if (saved && crc_ok) saved[bi] = (uint8_t)v;
        }
    }
    if (tid == 0) {
        *n_iter_out = (uint8_t)n_iter;
        *cb_crc = crc_ok ? 1 : 0;
        if (p.plcf_mask) p.plcf_mask[(size_t)pkt * p.C + cbi] = crc_ok ? plcf_mask : 0xffffffffu;
    }
}

// ---- transport-block level logic, one block per packet (reference pdc_enc.cpp:415-489):
// all CB CRCs ok -> (C > 1: check CRC24A over the whole TB, failure resets the CB flags)
__global__ void __launch_bounds__(256) k_tb_finalize(const tbfin_params_t p) {
    const uint32_t pkt = blockIdx.x, tid = threadIdx.x;
    const uint32_t slot = p.slot ? p.slot[pkt] : pkt;
    uint8_t* cb_crc = p.cb_crc + (size_t)slot * p.C;
    __shared__ uint32_t s_all, s_crc;
    if (tid == 0) { s_all = 1; s_crc = 0; }
    __syncthreads();
    for (uint32_t c = tid; c < p.C; c += blockDim.x)
        if (!cb_crc[c]) s_all = 0;
    __syncthreads();
    bool ok = s_all != 0;
    if (ok && p.C > 1) {
        // CRC24A over tbs + 24 bits, chunked: chunk i covers bytes [i*CH, (i+1)*CH)
        const uint8_t* tb = p.tb_out + (size_t)pkt * p.tb_stride;
        const uint32_t nbytes = p.tbs_bits / 8 + 3;
        const uint32_t nchunk = (nbytes + p.chunk_bytes - 1) / p.chunk_bytes;
        uint32_t r = 0;
        for (uint32_t i = tid; i < nchunk; i += blockDim.x) {
            uint32_t c = 0;
            const uint32_t b0 = i * p.chunk_bytes, b1 = min(nbytes, b0 + p.chunk_bytes);
            for (uint32_t b = b0; b < b1; ++b) {
                const uint32_t v = tb[b];
#pragma unroll
                for (int q = 7; q >= 0; --q) {
                    c = (c << 1) | ((v >> q) & 1);
                    if (c & 0x1000000u) c ^= 0x1864CFBu;
                }
            }
            // multiply by x^(8 * bytes after this chunk)
            r ^= gf2_mulmod(c, p.xp_tb[i], 0x1864CFBu, 24);
        }
        for (int o = 16; o > 0; o >>= 1) r ^= __shfl_xor_sync(0xffffffffu, r, o);
        if ((tid & 31) == 0) atomicXor(&s_crc, r);
        __syncthreads();
        ok = s_crc == 0;
        if (!ok)
            for (uint32_t c = tid; c < p.C; c += blockDim.x) cb_crc[c] = 0;
    }
    if (tid == 0) p.tb_ok[pkt] = ok ? 1 : 0;
}

size_t tdec_smem_bytes(uint32_t K, uint32_t L, bool ck_alias) {
    const uint32_t T = K / L / 2, nck = (L + tdec::S_CK - 1) / tdec::S_CK;
    return (size_t)4 * (3 * (K / 2) + (ck_alias ? (nck > 2 ? nck - 2 : 0) : nck) * 7 * T + 4 * 7 * T) + 16;
}

tdec_geom_t tdec_geometry(const cb_desc_t* cbs, uint32_t n) {
    tdec_geom_t g{32, 32, 0, false};
    uint32_t Tmax = 1;
    size_t aliased = 0;
    for (uint32_t i = 0; i < n; ++i) {
        Tmax = std::max(Tmax, cbs[i].P / 2);
        g.smem_item = std::max(g.smem_item, tdec_smem_bytes(cbs[i].K, cbs[i].L, false));
        aliased = std::max(aliased, tdec_smem_bytes(cbs[i].K, cbs[i].L, true));
    }
    g.smem_item = (g.smem_item + 15) & ~(size_t)15;
    aliased = (aliased + 15) & ~(size_t)15;
    // four blocks per SM: 228 KB of shared memory, 1 KB of it reserved per block, 80 B static per block
    const size_t four = 233472 / 4 - 1024 - 128;
    if (Tmax >= 32 && g.smem_item > four && aliased <= four) { g.smem_item = aliased; g.ck_alias = true; }
    if (Tmax >= 32) {  // one item per block
        g.threads = g.tg = ((Tmax + 31) / 32) * 32;
    } else {           // many small items per block: power-of-two groups inside 128 threads
        g.tg = 1;
        while (g.tg < Tmax) g.tg *= 2;
        g.threads = 128;
        while (g.threads > g.tg && (size_t)(g.threads / g.tg) * g.smem_item > 96 * 1024) g.threads /= 2;
    }
    return g;
}

cudaError_t launch_turbo_decode(const tdec_params_t& p, uint32_t n_packets, const tdec_geom_t& g, cudaStream_t st) {
    const uint32_t n_items = n_packets * p.C;
    uint32_t threads = g.threads;
    // grouped launch: prefer at least two blocks per SM over full 128-thread blocks
    while (threads > 32 && threads > g.tg && n_items / (threads / g.tg) < 2 * 148) threads /= 2;
    const uint32_t per_block = threads / g.tg;
    const size_t smem = per_block * g.smem_item;
    static smem_optin_t optin;   // per device (a second context may live on another GPU of this process)
    cudaError_t e0 = optin.ensure(smem, [](size_t b) {
        cudaError_t e = cudaFuncSetAttribute(k_turbo_decode<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)b);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(k_turbo_decode<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)b);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(k_turbo_decode<false, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)b);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(k_turbo_decode<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)b);
        return e;
    });
    if (e0 != cudaSuccess) return e0;
    const dim3 grid((n_items + per_block - 1) / per_block);
    // DNRP_TDEC_BULK=0: the per-thread staging loop instead of the copy engine (for A/B measurements; same results)
    static const bool bulk = [] { const char* e = getenv("DNRP_TDEC_BULK"); return !(e && e[0] == '0'); }();
    if (per_block == 1 && g.ck_alias)
        k_turbo_decode<false, true, true><<<grid, threads, smem, st>>>(p, g.tg, (uint32_t)(g.smem_item / 4), n_items);
    else if (per_block == 1 && bulk)
        k_turbo_decode<false, true><<<grid, threads, smem, st>>>(p, g.tg, (uint32_t)(g.smem_item / 4), n_items);
    else if (per_block == 1)
        k_turbo_decode<false><<<grid, threads, smem, st>>>(p, g.tg, (uint32_t)(g.smem_item / 4), n_items);
    else
        k_turbo_decode<true><<<grid, threads, smem, st>>>(p, g.tg, (uint32_t)(g.smem_item / 4), n_items);
    return cudaGetLastError();
}

cudaError_t launch_tb_finalize(const tbfin_params_t& p, uint32_t n_packets, cudaStream_t st) {
    k_tb_finalize<<<n_packets, 256, 0, st>>>(p);
    return cudaGetLastError();
}

}  // namespace dnrp

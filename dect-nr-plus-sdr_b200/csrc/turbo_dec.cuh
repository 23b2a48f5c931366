// turbo_dec.cuh -- batched int16 max-log-MAP turbo decoder for sm_100a, one thread block per code
// block, fused with descrambling, de-rate-matching and HARQ soft combining (reference call chain:
// lib/src/phy/fec/pdc_enc.cpp:291-411 and pcc_enc.cpp:291-351; the arithmetic the reference delegates
// to srsRAN's rm_turbo.c / turbodecoder*.c).
//
// Decoder contract (identical in oracle/fec.c, DESIGN.md "turbo decoder contract"):
//  * trellis split into P = K/L windows; every window is decoded independently per half-iteration;
//  * boundary metrics come from a training pass over the neighbouring window, which itself starts from
//    the boundary that the window two positions away stored in the previous iteration (zero at first),
//    except at the trellis ends where the exact start state / termination metrics are used;
//  * metrics are int16, normalised to state 0 after every step; channel LLRs are clamped to +-1023,
//    extrinsics to +-2047, so no int16 overflow is possible.
//
// Mapping to the SM: thread t decodes windows t and t+P/2 at once, packed in the two halves of one
// 32-bit register (VIADD.16x2 / VIADDMNMX.S16x2 / VIMNMX3.S16x2), because the QPP interleaver maps the
// pair (j, j+K/2) onto the pair (pi(j), pi(j)+K/2): one shared-memory word serves both halves.
// Shared arrays are stored "window-transposed" (word of trellis step k of thread t at k*T + t) so
// every per-step access of a warp is one conflict-free row.  Backward metrics are only checkpointed
// every 8 steps and recomputed into registers (shared memory, not issue slots, limits residency).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace dnrp {

struct cb_desc_t {
    uint32_t K, L, P;          // code block size, window length, number of windows
    uint32_t f1, f2;           // QPP parameters
    uint32_t rp, n_e;          // first LLR of this CB inside the packet's d-stream, LLR count
    uint32_t out_byte, rlen_byte;  // where the decoded bytes go inside the TB, how many (CB CRC dropped)
    uint32_t soft_off;         // offset (int16 elements) of this CB inside the packet's soft buffer
    uint32_t rm_off;           // offset of this K's rate-matching table (pos[]) in the table arena
    uint32_t rm_inv_off;       // offset of the inverse tables (per redundancy version: LLR index -> soft-buffer position)
    uint32_t rm_start[4];      // start position per redundancy version
    uint32_t crc_poly, crc_order;  // check polynomial of this CB (24B, 24A when C == 1, 16 for PLCF)
    uint32_t xp_off;           // offset of the per-window x^n mod g table in the table arena
    uint32_t saved_off;        // byte offset inside the packet's saved-bytes area
    uint32_t Linv;             // ceil(2^32 / L): i / L == __umulhi(i, Linv) for every i < 2^32 / L
    uint32_t Nc;               // circular-buffer entries in use (3K + 12 unless the soft buffer is limited): wrap-around length
};

namespace tdec {

static constexpr int S_CK = 8;             // checkpoint distance
static constexpr uint32_t NEG2 = 0xE000E000u;  // (-8192, -8192)

__device__ __forceinline__ uint32_t vadd(uint32_t a, uint32_t b) { return __vadd2(a, b); }
__device__ __forceinline__ uint32_t vsub(uint32_t a, uint32_t b) { return __vsub2(a, b); }
__device__ __forceinline__ uint32_t vneg(uint32_t a) { return __vsub2(0u, a); }
__device__ __forceinline__ uint32_t vmax(uint32_t a, uint32_t b) { return __vmaxs2(a, b); }
__device__ __forceinline__ uint32_t vmin(uint32_t a, uint32_t b) { return __vmins2(a, b); }
__device__ __forceinline__ uint32_t vaddmax(uint32_t a, uint32_t b, uint32_t c) { return __viaddmax_s16x2(a, b, c); }
__device__ __forceinline__ uint32_t vmax3(uint32_t a, uint32_t b, uint32_t c) { return __vimax3_s16x2(a, b, c); }
__device__ __forceinline__ uint32_t vclamp(uint32_t a, int lim) {
    const uint32_t hi = ((uint32_t)lim & 0xffff) * 0x00010001u, lo = ((uint32_t)(-lim) & 0xffff) * 0x00010001u;
    return vmax(vmin(a, hi), lo);
}
__device__ __forceinline__ uint32_t swap16(uint32_t a) { return __byte_perm(a, 0, 0x1032); }
__device__ __forceinline__ uint32_t pack(int lo, int hi) { return ((uint32_t)lo & 0xffff) | ((uint32_t)hi << 16); }

// ---- bulk staging of one stream of the soft buffer into shared memory by the copy engine (cp.async.bulk, SASS
// UBLKCP): one elected thread arms the mbarrier with the byte count and issues the copy, every thread of the block
// waits on the barrier's phase.  dst / src / bytes are multiples of 16.
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                     : "=r"(ok) : "r"(smem_addr(bar)), "r"(parity) : "memory");
    } while (!ok);
}
// generic-proxy accesses to shared memory issued so far are ordered before the async-proxy writes that follow
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_addr(dst)), "l"(src), "r"(bytes), "r"(smem_addr(bar)) : "memory");
}

// one backward step: b[] <- beta at the previous trellis time, normalised to state 0 (NORM) or left as computed.
// Normalisation only shifts all eight metrics alike, so skipping it changes no decision and no stored (normalised)
// value as long as nothing overflows: normalised metrics lie within +-12.3k (three steps connect any two states,
// |gamma| <= 4093), each un-normalised step moves them by at most 4093, so three un-normalised steps followed by a
// normalised one stay below 12.3k + 4 * 4.1k = 28.7k < 2^15.  Used in the training and plain backward passes
// (shorter dependency chains); every step that feeds an LLR is normalised.
template <bool NORM = true>
__device__ __forceinline__ void beta_step(uint32_t (&b)[8], uint32_t x, uint32_t y) {
    const uint32_t xy = vadd(x, y);
    uint32_t n[8];
    n[0] = vaddmax(b[4], xy, b[0]);              // u0->S0 (0) | u1->S4 (x+y)
    n[1] = vaddmax(b[0], xy, b[4]);              // u0->S4 (0) | u1->S0 (x+y)
    n[2] = vaddmax(b[5], y, vadd(b[1], x));      // u0->S5 (y) | u1->S1 (x)
    n[3] = vaddmax(b[1], y, vadd(b[5], x));      // u0->S1 (y) | u1->S5 (x)
    n[4] = vaddmax(b[2], y, vadd(b[6], x));      // u0->S2 (y) | u1->S6 (x)
    n[5] = vaddmax(b[6], y, vadd(b[2], x));      // u0->S6 (y) | u1->S2 (x)
    n[6] = vaddmax(b[3], xy, b[7]);              // u0->S7 (0) | u1->S3 (x+y)
    n[7] = vaddmax(b[7], xy, b[3]);              // u0->S3 (0) | u1->S7 (x+y)
    if (NORM) {
        b[0] = 0;
#pragma unroll
        for (int s = 1; s < 8; ++s) b[s] = vsub(n[s], n[0]);
    } else {
#pragma unroll
        for (int s = 0; s < 8; ++s) b[s] = n[s];
    }
}

// one forward step: a[] <- alpha at the next trellis time, normalised to state 0
template <bool NORM = true>
__device__ __forceinline__ void alpha_step(uint32_t (&a)[8], uint32_t x, uint32_t y) {
    const uint32_t xy = vadd(x, y);
    uint32_t n[8];
    n[0] = vaddmax(a[1], xy, a[0]);
    n[4] = vaddmax(a[0], xy, a[1]);
    n[5] = vaddmax(a[2], y, vadd(a[3], x));
    n[1] = vaddmax(a[3], y, vadd(a[2], x));
    n[2] = vaddmax(a[4], y, vadd(a[5], x));
    n[6] = vaddmax(a[5], y, vadd(a[4], x));
    n[7] = vaddmax(a[7], xy, a[6]);
    n[3] = vaddmax(a[6], xy, a[7]);
    if (NORM) {
        a[0] = 0;
#pragma unroll
        for (int s = 1; s < 8; ++s) a[s] = vsub(n[s], n[0]);
    } else {
#pragma unroll
        for (int s = 0; s < 8; ++s) a[s] = n[s];
    }
}

// a-posteriori LLR of the step with alpha a[] (before), beta bn[] (after).  Both are normalised where an LLR is formed
// (a[0] == bn[0] == 0), which removes three of the sixteen additions; the fourth term of every group is folded into
// a fused add-max.
__device__ __forceinline__ uint32_t llr_step(const uint32_t (&a)[8], const uint32_t (&bn)[8], uint32_t x, uint32_t y) {
    const uint32_t xy = vadd(x, y);
    const uint32_t m_xy = vaddmax(a[7], bn[7], vmax3(bn[4], a[1], vadd(a[6], bn[3])));
    const uint32_t m_x = vaddmax(a[5], bn[2], vmax3(vadd(a[2], bn[1]), vadd(a[3], bn[5]), vadd(a[4], bn[6])));
    const uint32_t m_0 = vaddmax(a[7], bn[3], vmax3(0u, vadd(a[1], bn[4]), vadd(a[6], bn[7])));
    const uint32_t m_y = vaddmax(a[5], bn[6], vmax3(vadd(a[2], bn[5]), vadd(a[3], bn[1]), vadd(a[4], bn[2])));
    const uint32_t l1 = vaddmax(m_xy, xy, vadd(m_x, x));
    const uint32_t l0 = vaddmax(m_y, y, m_0);
    return vsub(l1, l0);
}

}  // namespace tdec
}  // namespace dnrp

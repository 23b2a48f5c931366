/* ORACLE (test infrastructure, NOT product code): PCC / PDC coding chains on top of fec.c.
 * Control flow follows reference lib/src/phy/fec/pdc_enc.cpp:127-229 (encode), :291-492 (decode),
 * pcc_enc.cpp:145-213 (encode), :215-364 (blind decode).  See fec.c header for the parity status. */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define MAXK 6144
typedef struct tdec_s tdec_t;
uint32_t orc_crc(const uint8_t* bits, uint32_t n, uint32_t poly, uint32_t order);
void orc_unpack(const uint8_t* bytes, uint8_t* bits, uint32_t nbits);
void orc_pack(const uint8_t* bits, uint8_t* bytes, uint32_t nbits);
void orc_turbo_encode(const uint8_t* c, uint32_t K, uint32_t f1, uint32_t f2, uint8_t* d);
void orc_rm_tx(const uint8_t* d, uint32_t K, uint32_t E, uint32_t rv, uint8_t* e);
void orc_rm_rx(const int16_t* e, uint32_t K, uint32_t E, uint32_t rv, int16_t* soft);
extern uint32_t orc_N_IR;
extern __thread uint32_t orc_N_cb;
tdec_t* orc_tdec_alloc(void);
void orc_tdec_free(tdec_t* t);
void orc_tdec_new_cb(tdec_t* t, const int16_t* soft, uint32_t K, uint32_t f1, uint32_t f2);
void orc_tdec_iteration(tdec_t* t, uint8_t* bits);

#define CRC24A 0x1864CFB
#define CRC24B 0x1800063
#define CRC16 0x11021

typedef struct {
    uint32_t tbs, C, K1, K2, C1, C2;
    uint32_t f1_K1, f2_K1, f1_K2, f2_K2; /* QPP parameters of K1 (=K+) and K2 (=K-) */
} orc_seg_t;

static void put_crc(uint8_t* bits, uint32_t crc, uint32_t order) {
    for (uint32_t i = 0; i < order; ++i) bits[i] = (crc >> (order - 1 - i)) & 1;
}

/* ---------------------------------------------------------------- PDC encode, pdc_enc.cpp:127-229 */
/* tb: N_TB_byte packed; gold: G scrambling bits (one per byte); d_out: ceil(G/8) packed bytes */
void orc_pdc_encode(const uint8_t* tb, const orc_seg_t* s, uint32_t Qm, uint32_t rv, uint32_t G,
                    const uint8_t* gold, uint8_t* d_out) {
    uint8_t* tbbits = (uint8_t*)malloc(s->tbs + 24);
    uint8_t* dbits = (uint8_t*)calloc(G + 8, 1);
    static __thread uint8_t c[MAXK], d[3 * (MAXK + 4)], e[3 * (MAXK + 4) * 8];
    orc_unpack(tb, tbbits, s->tbs);
    put_crc(tbbits + s->tbs, orc_crc(tbbits, s->tbs, CRC24A, 24), 24);
    const uint32_t Gp = G / Qm, gamma = Gp % s->C;
    uint32_t rp = 0, wp = 0;
    for (uint32_t cb = 0; cb < s->C; ++cb) {
        const uint32_t K = cb < s->C2 ? s->K2 : s->K1;
        const uint32_t f1 = cb < s->C2 ? s->f1_K2 : s->f1_K1, f2 = cb < s->C2 ? s->f2_K2 : s->f2_K1;
        const uint32_t rlen = s->C > 1 ? K - 24 : K;
        uint32_t n_e;
        if (cb <= s->C - gamma - 1) n_e = Qm * (Gp / s->C);
        else n_e = Qm * ((Gp + s->C - 1) / s->C);
        /* the last CB carries the TB CRC inside its rlen bits (tbbits already has it appended) */
        memcpy(c, tbbits + rp, rlen);
        if (s->C > 1) put_crc(c + rlen, orc_crc(c, rlen, CRC24B, 24), 24);
        orc_turbo_encode(c, K, f1, f2, d);
        uint8_t* ee = n_e <= sizeof(e) ? e : (uint8_t*)malloc(n_e);
        orc_N_cb = orc_N_IR ? orc_N_IR / s->C : 0;
        orc_rm_tx(d, K, n_e, rv, ee);
        orc_N_cb = 0;
        for (uint32_t i = 0; i < n_e; ++i) dbits[wp + i] = ee[i] ^ gold[wp + i];
        if (ee != e) free(ee);
        rp += rlen;
        wp += n_e;
    }
    orc_pack(dbits, d_out, G);
    free(tbbits);
    free(dbits);
}

/* ---------------------------------------------------------------- PDC decode, pdc_enc.cpp:291-492 */
/* e: G int16 LLRs (descrambled IN PLACE like the reference); soft: C x 3*(K1+4) int16 (HARQ state);
 * cb_crc: C flags (HARQ state); saved: C x K1/8 bytes (HARQ state); data: >= ceil((tbs+24)/8)+K1/8
 * bytes; n_iter[C] receives the iterations spent per CB (0 = skipped).  returns 1 if TB CRC ok. */
int orc_pdc_decode(int16_t* e, const orc_seg_t* s, uint32_t Qm, uint32_t rv, uint32_t G,
                   const uint8_t* gold, int16_t* soft, uint8_t* cb_crc, uint8_t* saved, uint8_t* data,
                   uint32_t* n_iter, uint32_t min_it, uint32_t max_it) {
    tdec_t* t = orc_tdec_alloc();
    static __thread uint8_t bits[MAXK];
    const uint32_t stride = 3 * (s->K1 + 4), sstride = s->K1 / 8;
    uint32_t wp = 0;
    for (uint32_t cb = 0; cb < s->C; ++cb) {
        const uint32_t K = cb < s->C2 ? s->K2 : s->K1;
        const uint32_t f1 = cb < s->C2 ? s->f1_K2 : s->f1_K1, f2 = cb < s->C2 ? s->f2_K2 : s->f2_K1;
        const uint32_t rlen = s->C == 1 ? K : K - 24;
        const uint32_t Gp = G / Qm, gamma = Gp % s->C;
        const uint32_t n_e = Qm * (Gp / s->C);
        uint32_t rp = cb * n_e, n_e2 = n_e;
        if (cb > s->C - gamma) { /* sic: '>' as in the reference (srsRAN sch.c heritage) */
            n_e2 = n_e + Qm;
            rp = (s->C - gamma) * n_e + (cb - (s->C - gamma)) * n_e2;
        }
        for (uint32_t i = 0; i < n_e2; ++i)
            if (gold[rp + i]) e[rp + i] = (int16_t)-e[rp + i];
        n_iter[cb] = 0;
        if (!cb_crc[cb]) {
            orc_N_cb = orc_N_IR ? orc_N_IR / s->C : 0;
            orc_rm_rx(e + rp, K, n_e2, rv, soft + cb * stride);
            orc_N_cb = 0;
            orc_tdec_new_cb(t, soft + cb * stride, K, f1, f2);
            uint32_t noi = 0;
            int stop = 0;
            do {
                orc_tdec_iteration(t, bits);
                ++noi;
                uint32_t len_crc = s->C > 1 ? K : s->tbs + 24;
                uint32_t rem = orc_crc(bits, len_crc - 24, s->C > 1 ? CRC24B : CRC24A, 24);
                uint32_t got = 0;
                for (uint32_t i = 0; i < 24; ++i) got = (got << 1) | bits[len_crc - 24 + i];
                if (rem == got && noi >= min_it) {
                    cb_crc[cb] = 1;
                    stop = 1;
                }
            } while (noi < max_it && !stop);
            n_iter[cb] = noi;
            orc_pack(bits, data + wp / 8, K);
        } else {
            memcpy(data + wp / 8, saved + cb * sstride, rlen / 8);
        }
        wp += rlen;
    }
    orc_tdec_free(t);
    int tb_ok = 1;
    for (uint32_t i = 0; i < s->C; ++i) tb_ok = tb_ok && cb_crc[i];
    if (!tb_ok) {
        uint32_t off = 0;
        for (uint32_t cb = 0; cb < s->C; ++cb) {
            const uint32_t K = cb < s->C2 ? s->K2 : s->K1;
            const uint32_t rlen = s->C == 1 ? K : K - 24;
            if (cb_crc[cb]) memcpy(saved + cb * sstride, data + off / 8, rlen / 8);
            off += rlen;
        }
        return 0;
    }
    if (s->C == 1) return 1;
    uint8_t* tbbits = (uint8_t*)malloc(s->tbs + 24);
    orc_unpack(data, tbbits, s->tbs + 24);
    uint32_t rem = orc_crc(tbbits, s->tbs, CRC24A, 24), got = 0;
    for (uint32_t i = 0; i < 24; ++i) got = (got << 1) | tbbits[s->tbs + i];
    free(tbbits);
    if (rem == got) return 1;
    memset(cb_crc, 0, s->C);
    return 0;
}

/* ---------------------------------------------------------------- PCC encode, pcc_enc.cpp:145-213 */
/* plcf: 5 or 10 bytes; gold200: 200 scrambling bits (c_init 0x44454354); d_out: 25 bytes */
void orc_pcc_encode(const uint8_t* plcf, uint32_t PLCF_type, uint32_t cl, uint32_t bf, uint32_t f1,
                    uint32_t f2, const uint8_t* gold200, uint8_t* d_out) {
    const uint32_t nb = PLCF_type == 1 ? 40 : 80, K = nb + 16;
    uint8_t c[96], d[3 * 100], e[200];
    orc_unpack(plcf, c, nb);
    uint32_t crc = orc_crc(c, nb, CRC16, 16);
    const uint32_t mask = (!cl && !bf) ? 0x0000 : (cl && !bf) ? 0x5555 : (!cl && bf) ? 0xAAAA : 0xFFFF;
    put_crc(c + nb, crc ^ mask, 16);
    orc_turbo_encode(c, K, f1, f2, d);
    memset(e, 0, sizeof(e));
    orc_rm_tx(d, K, 196, 0, e);
    for (uint32_t i = 0; i < 200; ++i) e[i] ^= gold200[i]; /* reference scrambles 200 bits */
    orc_pack(e, d_out, 200);
}

/* ---------------------------------------------------------------- PCC decode, pcc_enc.cpp:215-364 */
/* d_rx: 196 int16 LLRs (not modified); returns 1 and fills plcf_out/cl/bf when a masked CRC matches */
int orc_pcc_decode(const int16_t* d_rx, uint32_t PLCF_type, uint32_t f1, uint32_t f2,
                   const uint8_t* gold200, uint8_t* plcf_out, uint32_t* cl, uint32_t* bf,
                   uint32_t* n_iter, uint32_t max_it) {
    const uint32_t nb = PLCF_type == 1 ? 40 : 80, K = nb + 16;
    int16_t e[196], soft[3 * 100];
    uint8_t bits[96];
    for (uint32_t i = 0; i < 196; ++i) e[i] = gold200[i] ? (int16_t)-d_rx[i] : d_rx[i];
    memset(soft, 0, sizeof(soft));
    orc_rm_rx(e, K, 196, 0, soft);
    tdec_t* t = orc_tdec_alloc();
    orc_tdec_new_cb(t, soft, K, f1, f2);
    uint32_t noi = 0;
    int found = 0;
    do {
        orc_tdec_iteration(t, bits);
        ++noi;
        uint32_t rx = 0;
        for (uint32_t i = 0; i < 16; ++i) rx = (rx << 1) | bits[nb + i];
        const uint32_t calc = orc_crc(bits, nb, CRC16, 16);
        static const uint32_t masks[4] = {0x0000, 0x5555, 0xAAAA, 0xFFFF};
        for (int m = 0; m < 4 && !found; ++m)
            if ((rx ^ masks[m]) == calc) {
                *cl = (m == 1 || m == 3);
                *bf = (m == 2 || m == 3);
                found = 1;
            }
    } while (noi < max_it && !found);
    *n_iter = noi;
    orc_tdec_free(t);
    if (found) orc_pack(bits, plcf_out, nb);
    return found;
}

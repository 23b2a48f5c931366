/* shim_srsran.c -- plain-C bodies for the srsRAN_4G (release_23_11) functions the reference's PHY calls.
 *
 * TEST INFRASTRUCTURE ONLY (part of the oracle/_ref build, see README.md in this directory).  srsRAN is not installed on
 * this image and cannot be fetched, so these bodies restate the PUBLISHED algorithms srsRAN implements:
 *   CRC (TS 36.212 5.1.1), turbo encoder + QPP interleaver (5.1.3.2), rate matching (5.1.4.1), Gold sequence
 *   (TS 36.211 7.2), constellations (TS 36.211 7.1), max-log soft demapper, max-log-MAP turbo decoder, DFT.
 * Where srsRAN has implementation freedom (demapper int16 scale constants, the decoder's fixed-point details, the
 * Gaussian generator) the choice made here is marked UNPINNED: it follows srsRAN as far as its documentation and this
 * author's knowledge go, it is not verified against srsRAN's object code.  Everything the reference itself computes
 * around these calls is the reference's own, unmodified source.
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "srsran/srsran.h"

#define RE(z) __real__(z)
#define IM(z) __imag__(z)

/* ---- vector helpers ---------------------------------------------------------------------------------------------- */
void* srsran_vec_malloc(uint32_t size) {
    /* 64 bytes of slack behind every buffer: the reference lets the decoder write whole code blocks (payload + CRCs) into
     * HARQ buffers sized for the payload alone (pdc_enc.cpp:378 writes K / 8 bytes at data[wp / 8], the buffer holds
     * N_TB_byte), which srsRAN's 64-byte aligned allocations absorb */
    void* p = NULL;
    if (posix_memalign(&p, 64, (size_t)size + 64)) return NULL;
    return p;
}
cf_t* srsran_vec_cf_malloc(uint32_t n) { return (cf_t*)srsran_vec_malloc(n * sizeof(cf_t)); }
float* srsran_vec_f_malloc(uint32_t n) { return (float*)srsran_vec_malloc(n * sizeof(float)); }
uint8_t* srsran_vec_u8_malloc(uint32_t n) { return (uint8_t*)srsran_vec_malloc(n); }
int16_t* srsran_vec_i16_malloc(uint32_t n) { return (int16_t*)srsran_vec_malloc(n * sizeof(int16_t)); }
void srsran_vec_cf_zero(cf_t* p, uint32_t n) { memset(p, 0, n * sizeof(cf_t)); }
void srsran_vec_f_zero(float* p, uint32_t n) { memset(p, 0, n * sizeof(float)); }
void srsran_vec_u8_zero(uint8_t* p, uint32_t n) { memset(p, 0, n); }
void srsran_vec_cf_copy(cf_t* d, const cf_t* s, uint32_t n) { memmove(d, s, n * sizeof(cf_t)); }
void srsran_vec_f_copy(float* d, const float* s, uint32_t n) { memmove(d, s, n * sizeof(float)); }
void srsran_vec_u8_copy(uint8_t* d, const uint8_t* s, uint32_t n) { memmove(d, s, n); }
void srsran_vec_sum_ccc(const cf_t* x, const cf_t* y, cf_t* z, uint32_t n) { for (uint32_t i = 0; i < n; ++i) z[i] = x[i] + y[i]; }
void srsran_vec_sub_ccc(const cf_t* x, const cf_t* y, cf_t* z, uint32_t n) { for (uint32_t i = 0; i < n; ++i) z[i] = x[i] - y[i]; }
void srsran_vec_sub_fff(const float* x, const float* y, float* z, uint32_t n) { for (uint32_t i = 0; i < n; ++i) z[i] = x[i] - y[i]; }
void srsran_vec_prod_fff(const float* x, const float* y, float* z, uint32_t n) { for (uint32_t i = 0; i < n; ++i) z[i] = x[i] * y[i]; }
void srsran_vec_sc_prod_cfc(const cf_t* x, float h, cf_t* z, uint32_t n) { for (uint32_t i = 0; i < n; ++i) z[i] = x[i] * h; }
void srsran_vec_sc_prod_ccc(const cf_t* x, cf_t h, cf_t* z, uint32_t n) { for (uint32_t i = 0; i < n; ++i) z[i] = x[i] * h; }
void srsran_vec_xor_bbb(const uint8_t* x, const uint8_t* y, uint8_t* z, uint32_t n) { for (uint32_t i = 0; i < n; ++i) z[i] = x[i] ^ y[i]; }
void srsran_vec_sc_prod_cfc_simd(const cf_t* x, float h, cf_t* z, int n) { srsran_vec_sc_prod_cfc(x, h, z, (uint32_t)n); }
void srsran_vec_sc_prod_ccc_simd(const cf_t* x, cf_t h, cf_t* z, int n) { srsran_vec_sc_prod_ccc(x, h, z, (uint32_t)n); }
void srsran_vec_add_fff_simd(const float* x, const float* y, float* z, int n) { for (int i = 0; i < n; ++i) z[i] = x[i] + y[i]; }
uint32_t srsran_vec_max_abs_fi(const float* x, uint32_t n) {
    uint32_t p = 0;
    float m = -1.f;
    for (uint32_t i = 0; i < n; ++i)
        if (fabsf(x[i]) > m) { m = fabsf(x[i]); p = i; }
    return p;
}
uint32_t srsran_vec_max_abs_ci(const cf_t* x, uint32_t n) {
    uint32_t p = 0;
    float m = -1.f;
    for (uint32_t i = 0; i < n; ++i) {
        const float a = RE(x[i]) * RE(x[i]) + IM(x[i]) * IM(x[i]);
        if (a > m) { m = a; p = i; }
    }
    return p;
}

/* ---- bits -------------------------------------------------------------------------------------------------------- */
void srsran_bit_unpack_vector(const uint8_t* packed, uint8_t* unpacked, int nof_bits) {
    for (int i = 0; i < nof_bits; ++i) unpacked[i] = (packed[i >> 3] >> (7 - (i & 7))) & 1;
}
void srsran_bit_pack_vector(uint8_t* unpacked, uint8_t* packed, int nof_bits) {
    /* whole bytes, then the remaining bits into the top of the last byte (its low bits are cleared) */
    const int nbytes = nof_bits / 8;
    for (int b = 0; b < nbytes; ++b) {
        uint8_t v = 0;
        for (int q = 0; q < 8; ++q) v = (uint8_t)((v << 1) | (unpacked[8 * b + q] & 1));
        packed[b] = v;
    }
    if (nof_bits % 8) {
        uint8_t v = 0;
        for (int q = 0; q < nof_bits % 8; ++q) v = (uint8_t)((v << 1) | (unpacked[8 * nbytes + q] & 1));
        packed[nbytes] = (uint8_t)(v << (8 - nof_bits % 8));
    }
}

/* ---- Gold sequence, TS 36.211 7.2 -------------------------------------------------------------------------------- */
int srsran_sequence_LTE_pr(srsran_sequence_t* q, uint32_t len, uint32_t seed) {
    const uint32_t Nc = 1600;
    if (q->c && len > q->max_len) srsran_sequence_free(q);
    if (!q->c) {
        q->c = (uint8_t*)calloc(len + 8, 1);
        q->c_bytes = (uint8_t*)calloc(len / 8 + 8, 1);
        q->c_float = (float*)calloc(len + 8, sizeof(float));
        q->c_short = (short*)calloc(len + 8, sizeof(short));
        q->c_char = (int8_t*)calloc(len + 8, 1);
        if (!q->c || !q->c_bytes || !q->c_float || !q->c_short || !q->c_char) return SRSRAN_ERROR;
        q->max_len = len;
    }
    q->cur_len = len;
    uint8_t* x1 = (uint8_t*)calloc(Nc + len + 31, 1);
    uint8_t* x2 = (uint8_t*)calloc(Nc + len + 31, 1);
    if (!x1 || !x2) return SRSRAN_ERROR;
    x1[0] = 1;
    for (int i = 0; i < 31; ++i) x2[i] = (seed >> i) & 1;
    for (uint32_t n = 0; n < Nc + len; ++n) {
        x1[n + 31] = x1[n + 3] ^ x1[n];
        x2[n + 31] = x2[n + 3] ^ x2[n + 2] ^ x2[n + 1] ^ x2[n];
    }
    memset(q->c_bytes, 0, len / 8 + 8);
    for (uint32_t n = 0; n < len; ++n) {
        const uint8_t c = x1[n + Nc] ^ x2[n + Nc];
        q->c[n] = c;
        q->c_float[n] = c ? -1.f : 1.f;
        q->c_short[n] = c ? -1 : 1;
        q->c_char[n] = c ? -1 : 1;
        q->c_bytes[n >> 3] |= (uint8_t)(c << (7 - (n & 7)));
    }
    free(x1);
    free(x2);
    return SRSRAN_SUCCESS;
}
void srsran_sequence_free(srsran_sequence_t* q) {
    free(q->c); free(q->c_bytes); free(q->c_float); free(q->c_short); free(q->c_char);
    memset(q, 0, sizeof(*q));
}
void srsran_scrambling_s_offset(srsran_sequence_t* s, short* data, int offset, int len) {
    for (int i = 0; i < len; ++i)
        if (s->c[offset + i]) data[i] = (short)-data[i];
}
void srsran_scrambling_sb_offset(srsran_sequence_t* s, int8_t* data, int offset, int len) {
    for (int i = 0; i < len; ++i)
        if (s->c[offset + i]) data[i] = (int8_t)-data[i];
}
void srsran_scrambling_bytes(srsran_sequence_t* s, uint8_t* data, int len) {
    for (int i = 0; i < len / 8; ++i) data[i] ^= s->c_bytes[i];
    if (len % 8) data[len / 8] ^= (uint8_t)(s->c_bytes[len / 8] & (0xff << (8 - len % 8)));
}

/* ---- constellations, TS 36.211 7.1 -------------------------------------------------------------------------------- */
static cf_t qam_point(uint32_t idx, uint32_t nbits) {
    /* bit b_j = bit (nbits - 1 - j) of idx; even j -> I, odd j -> Q; e.g. 64-QAM: I = (1 - 2 b0)(4 - (1 - 2 b2)(2 - (1 - 2 b4))) */
    const int half = (int)nbits / 2;
    float amp[2];
    for (int c = 0; c < 2; ++c) {
        float acc = 1.f;
        for (int l = half - 1; l >= 1; --l) {
            const float s = ((idx >> (nbits - 1 - (uint32_t)(2 * l + c))) & 1) ? -1.f : 1.f;
            acc = (float)(1 << (half - l)) - s * acc;
        }
        amp[c] = (((idx >> (nbits - 1 - (uint32_t)c)) & 1) ? -1.f : 1.f) * acc;
    }
    static const float norm[5] = {0.f, 1.41421356237f, 3.16227766017f, 6.48074069841f, 13.0384048104f};
    cf_t z;
    RE(z) = amp[0] / norm[half];
    IM(z) = amp[1] / norm[half];
    return z;
}
int srsran_modem_table_lte(srsran_modem_table_t* q, srsran_mod_t mod) {
    memset(q, 0, sizeof(*q));
    static const uint32_t nb[] = {1, 2, 4, 6, 8, 10};
    if (mod > SRSRAN_MOD_256QAM) return SRSRAN_ERROR;
    q->nbits_x_symbol = nb[mod];
    q->nsymbols = 1u << q->nbits_x_symbol;
    q->symbol_table = srsran_vec_cf_malloc(q->nsymbols);
    if (mod == SRSRAN_MOD_BPSK) {
        RE(q->symbol_table[0]) = (float)M_SQRT1_2; IM(q->symbol_table[0]) = (float)M_SQRT1_2;
        RE(q->symbol_table[1]) = -(float)M_SQRT1_2; IM(q->symbol_table[1]) = -(float)M_SQRT1_2;
    } else {
        for (uint32_t i = 0; i < q->nsymbols; ++i) q->symbol_table[i] = qam_point(i, q->nbits_x_symbol);
    }
    return SRSRAN_SUCCESS;
}
void srsran_modem_table_bytes(srsran_modem_table_t* q) {
    if (q->byte_tables_init) return;
    if (q->nbits_x_symbol == 1) {
        q->symbol_table_bpsk = (bpsk_packed_t*)srsran_vec_malloc(256 * sizeof(bpsk_packed_t));
        for (uint32_t b = 0; b < 256; ++b)
            for (int i = 0; i < 8; ++i) q->symbol_table_bpsk[b].symbol[i] = q->symbol_table[(b >> (7 - i)) & 1];
    } else if (q->nbits_x_symbol == 2) {
        q->symbol_table_qpsk = (qpsk_packed_t*)srsran_vec_malloc(256 * sizeof(qpsk_packed_t));
        for (uint32_t b = 0; b < 256; ++b)
            for (int i = 0; i < 4; ++i) q->symbol_table_qpsk[b].symbol[i] = q->symbol_table[(b >> (6 - 2 * i)) & 3];
    } else if (q->nbits_x_symbol == 4) {
        q->symbol_table_16qam = (qam16_packed_t*)srsran_vec_malloc(256 * sizeof(qam16_packed_t));
        for (uint32_t b = 0; b < 256; ++b)
            for (int i = 0; i < 2; ++i) q->symbol_table_16qam[b].symbol[i] = q->symbol_table[(b >> (4 - 4 * i)) & 15];
    }
    q->byte_tables_init = true;
}
void srsran_modem_table_free(srsran_modem_table_t* q) {
    free(q->symbol_table); free(q->symbol_table_bpsk); free(q->symbol_table_qpsk); free(q->symbol_table_16qam);
    memset(q, 0, sizeof(*q));
}
int srsran_mod_modulate(const srsran_modem_table_t* q, uint8_t* bits, cf_t* symbols, uint32_t nbits) {
    uint32_t j = 0;
    for (uint32_t i = 0; i + q->nbits_x_symbol <= nbits; i += q->nbits_x_symbol) {
        uint32_t idx = 0;
        for (uint32_t b = 0; b < q->nbits_x_symbol; ++b) idx = (idx << 1) | (bits[i + b] & 1);
        symbols[j++] = q->symbol_table[idx];
    }
    return (int)j;
}

/* ---- soft demapper: UNPINNED scale constants (100 sqrt2 / 400 / 700 / 1000), float -> int16 by C cast semantics ---- */
static int16_t to_short(float v) {
    v = truncf(v);
    if (v > 32767.f) v = 32767.f;
    if (v < -32768.f) v = -32768.f;
    return (int16_t)v;
}
static int16_t neg_short(int16_t v) { return v == -32768 ? 32767 : (int16_t)-v; }
static void demap_qam(const cf_t* sym, short* llr, int n, int nbits, float S, float root) {
    const int half = nbits / 2;
    for (int i = 0; i < n; ++i) {
        short* o = llr + (size_t)i * nbits;
        const int16_t yre = to_short(S * RE(sym[i])), yim = to_short(S * IM(sym[i]));
        o[0] = neg_short(yre);
        o[1] = neg_short(yim);
        int16_t pre = yre, pim = yim;
        for (int l = 1; l < half; ++l) {
            const float thr = (float)(1 << (half - l)) * S / root;
            pre = to_short((float)abs((int)pre) - thr);
            pim = to_short((float)abs((int)pim) - thr);
            o[2 * l] = pre;
            o[2 * l + 1] = pim;
        }
    }
}
int srsran_demod_soft_demodulate_s(srsran_mod_t mod, const cf_t* sym, short* llr, int n) {
    switch (mod) {
        case SRSRAN_MOD_BPSK:
            for (int i = 0; i < n; ++i) llr[i] = to_short(-(100.f * (RE(sym[i]) + IM(sym[i])) * (float)M_SQRT1_2));
            break;
        case SRSRAN_MOD_QPSK:
            for (int i = 0; i < n; ++i) {
                llr[2 * i] = to_short(RE(sym[i]) * (-100.f * sqrtf(2.f)));
                llr[2 * i + 1] = to_short(IM(sym[i]) * (-100.f * sqrtf(2.f)));
            }
            break;
        case SRSRAN_MOD_16QAM: demap_qam(sym, llr, n, 4, 400.f, sqrtf(10.f)); break;
        case SRSRAN_MOD_64QAM: demap_qam(sym, llr, n, 6, 700.f, sqrtf(42.f)); break;
        case SRSRAN_MOD_256QAM: demap_qam(sym, llr, n, 8, 1000.f, sqrtf(170.f)); break;
        default: return SRSRAN_ERROR;
    }
    return SRSRAN_SUCCESS;
}
int srsran_demod_soft_demodulate_b(srsran_mod_t mod, const cf_t* sym, int8_t* llr, int n) {
    /* 8-bit variant (PHY_LLR_BIT_WIDTH 8 is not the shipped configuration): int16 result scaled down by 8, saturated */
    static const uint32_t nb[] = {1, 2, 4, 6, 8};
    if (mod > SRSRAN_MOD_256QAM) return SRSRAN_ERROR;
    short* tmp = (short*)malloc(sizeof(short) * (size_t)n * nb[mod]);
    srsran_demod_soft_demodulate_s(mod, sym, tmp, n);
    for (size_t i = 0; i < (size_t)n * nb[mod]; ++i) {
        int v = tmp[i] / 8;
        llr[i] = (int8_t)(v > 127 ? 127 : v < -128 ? -128 : v);
    }
    free(tmp);
    return SRSRAN_SUCCESS;
}

/* ---- CRC ------------------------------------------------------------------------------------------------------------ */
int srsran_crc_init(srsran_crc_t* h, uint32_t poly, int order) {
    memset(h, 0, sizeof(*h));
    h->polynom = (int)poly;
    h->order = order;
    h->crchighbit = 1ull << (order - 1);
    h->crcmask = (h->crchighbit << 1) - 1;
    for (uint32_t b = 0; b < 256; ++b) {
        uint64_t r = (uint64_t)b << (order - 8);
        for (int i = 0; i < 8; ++i) r = (r & h->crchighbit) ? ((r << 1) ^ poly) : (r << 1);
        h->table[b] = r & h->crcmask;
    }
    return SRSRAN_SUCCESS;
}
int srsran_crc_set_init(srsran_crc_t* h, uint64_t v) { h->crcinit = v; return SRSRAN_SUCCESS; }
void srsran_crc_checksum_put_byte(srsran_crc_t* h, uint8_t byte) {
    const uint64_t idx = ((h->crcinit >> (h->order - 8)) ^ byte) & 0xff;
    h->crcinit = ((h->crcinit << 8) ^ h->table[idx]) & h->crcmask;
}
uint64_t srsran_crc_checksum_get(srsran_crc_t* h) { return h->crcinit & h->crcmask; }
uint32_t srsran_crc_checksum_byte(srsran_crc_t* h, const uint8_t* data, int len) {
    h->crcinit = 0;
    for (int i = 0; i < len / 8; ++i) srsran_crc_checksum_put_byte(h, data[i]);
    return (uint32_t)srsran_crc_checksum_get(h);
}
uint32_t srsran_crc_attach_byte(srsran_crc_t* h, uint8_t* data, int len) {
    const uint32_t c = srsran_crc_checksum_byte(h, data, len);
    for (int i = 0; i < h->order / 8; ++i) data[len / 8 + i] = (uint8_t)(c >> (h->order - 8 - 8 * i));
    return c;
}
bool srsran_crc_match_byte(srsran_crc_t* h, uint8_t* data, int len) {
    const uint32_t c = srsran_crc_checksum_byte(h, data, len);
    uint32_t r = 0;
    for (int i = 0; i < h->order / 8; ++i) r = (r << 8) | data[len / 8 + i];
    return c == r;
}

/* ---- soft buffers --------------------------------------------------------------------------------------------------- */
int srsran_softbuffer_rx_init_guru(srsran_softbuffer_rx_t* q, uint32_t max_cb, uint32_t max_cb_size) {
    memset(q, 0, sizeof(*q));
    q->max_cb = max_cb;
    q->max_cb_size = max_cb_size;
    q->buffer_f = (int16_t**)calloc(max_cb, sizeof(int16_t*));
    q->data = (uint8_t**)calloc(max_cb, sizeof(uint8_t*));
    q->cb_crc = (bool*)calloc(max_cb, sizeof(bool));
    if (!q->buffer_f || !q->data || !q->cb_crc) return SRSRAN_ERROR;
    for (uint32_t i = 0; i < max_cb; ++i) {
        q->buffer_f[i] = srsran_vec_i16_malloc(max_cb_size);
        q->data[i] = srsran_vec_u8_malloc(6144 / 8);
        if (!q->buffer_f[i] || !q->data[i]) return SRSRAN_ERROR;
    }
    srsran_softbuffer_rx_reset(q);
    return SRSRAN_SUCCESS;
}
void srsran_softbuffer_rx_free(srsran_softbuffer_rx_t* q) {
    if (q->buffer_f) for (uint32_t i = 0; i < q->max_cb; ++i) free(q->buffer_f[i]);
    if (q->data) for (uint32_t i = 0; i < q->max_cb; ++i) free(q->data[i]);
    free(q->buffer_f); free(q->data); free(q->cb_crc);
    memset(q, 0, sizeof(*q));
}
void srsran_softbuffer_rx_reset_cb(srsran_softbuffer_rx_t* q, uint32_t nof_cb) {
    if (nof_cb > q->max_cb) nof_cb = q->max_cb;
    for (uint32_t i = 0; i < nof_cb; ++i) {
        memset(q->buffer_f[i], 0, sizeof(int16_t) * q->max_cb_size);
        memset(q->data[i], 0, 6144 / 8);
    }
    srsran_softbuffer_rx_reset_cb_crc(q, nof_cb);
    q->tb_crc = false;
}
void srsran_softbuffer_rx_reset(srsran_softbuffer_rx_t* q) { srsran_softbuffer_rx_reset_cb(q, q->max_cb); }
void srsran_softbuffer_rx_reset_cb_crc(srsran_softbuffer_rx_t* q, uint32_t nof_cb) {
    if (nof_cb > q->max_cb) nof_cb = q->max_cb;
    for (uint32_t i = 0; i < nof_cb; ++i) q->cb_crc[i] = false;
}
int srsran_softbuffer_tx_init_guru(srsran_softbuffer_tx_t* q, uint32_t max_cb, uint32_t max_cb_size) {
    memset(q, 0, sizeof(*q));
    q->max_cb = max_cb;
    q->max_cb_size = max_cb_size;
    q->buffer_b = (uint8_t**)calloc(max_cb, sizeof(uint8_t*));
    if (!q->buffer_b) return SRSRAN_ERROR;
    for (uint32_t i = 0; i < max_cb; ++i) {
        q->buffer_b[i] = srsran_vec_u8_malloc(max_cb_size);
        if (!q->buffer_b[i]) return SRSRAN_ERROR;
    }
    srsran_softbuffer_tx_reset(q);
    return SRSRAN_SUCCESS;
}
void srsran_softbuffer_tx_free(srsran_softbuffer_tx_t* q) {
    if (q->buffer_b) for (uint32_t i = 0; i < q->max_cb; ++i) free(q->buffer_b[i]);
    free(q->buffer_b);
    memset(q, 0, sizeof(*q));
}
void srsran_softbuffer_tx_reset_cb(srsran_softbuffer_tx_t* q, uint32_t nof_cb) {
    if (nof_cb > q->max_cb) nof_cb = q->max_cb;
    for (uint32_t i = 0; i < nof_cb; ++i) memset(q->buffer_b[i], 0, q->max_cb_size);
}
void srsran_softbuffer_tx_reset(srsran_softbuffer_tx_t* q) { srsran_softbuffer_tx_reset_cb(q, q->max_cb); }

/* ---- code block sizes and QPP interleaver parameters, TS 36.212 Table 5.1.3-3 ------------------------------------------ */
static const uint16_t QPP_F1[188] = {
    3, 7, 19, 7, 7, 11, 5, 11, 7, 41, 103, 15, 9, 17, 9, 21, 101, 21, 57, 23, 13, 27, 11, 27, 85, 29, 33, 15, 17, 33, 103, 19,
    19, 37, 19, 21, 21, 115, 193, 21, 133, 81, 45, 23, 243, 151, 155, 25, 51, 47, 91, 29, 29, 247, 29, 89, 91, 157, 55, 31, 17, 35, 227, 65,
    19, 37, 41, 39, 185, 43, 21, 155, 79, 139, 23, 217, 25, 17, 127, 25, 239, 17, 137, 215, 29, 15, 147, 29, 59, 65, 55, 31, 17, 171, 67, 35,
    19, 39, 19, 199, 21, 211, 21, 43, 149, 45, 49, 71, 13, 17, 25, 183, 55, 127, 27, 29, 29, 57, 45, 31, 59, 185, 113, 31, 17, 171, 209, 253,
    367, 265, 181, 39, 27, 127, 143, 43, 29, 45, 157, 47, 13, 111, 443, 51, 51, 451, 257, 57, 313, 271, 179, 331, 363, 375, 127, 31, 33, 43, 33, 477,
    35, 233, 357, 337, 37, 71, 71, 37, 39, 127, 39, 39, 31, 113, 41, 251, 43, 21, 43, 45, 45, 161, 89, 323, 47, 23, 47, 263};
static const uint16_t QPP_F2[188] = {
    10, 12, 42, 16, 18, 20, 22, 24, 26, 84, 90, 32, 34, 108, 38, 120, 84, 44, 46, 48, 50, 52, 36, 56, 58, 60, 62, 32, 198, 68, 210, 36,
    74, 76, 78, 120, 82, 84, 86, 44, 90, 46, 94, 48, 98, 40, 102, 52, 106, 72, 110, 168, 114, 58, 118, 180, 122, 62, 84, 64, 66, 68, 420, 96,
    74, 76, 234, 80, 82, 252, 86, 44, 120, 92, 94, 48, 98, 80, 102, 52, 106, 48, 110, 112, 114, 58, 118, 60, 122, 124, 84, 64, 66, 204, 140, 72,
    74, 76, 78, 240, 82, 252, 86, 88, 60, 92, 846, 48, 28, 80, 102, 104, 954, 96, 110, 112, 114, 116, 354, 120, 610, 124, 420, 64, 66, 136, 420, 216,
    444, 456, 468, 80, 164, 504, 172, 88, 300, 92, 188, 96, 28, 240, 204, 104, 212, 192, 220, 336, 228, 232, 236, 120, 244, 248, 168, 64, 130, 264, 134, 408,
    138, 280, 142, 480, 146, 444, 120, 152, 462, 234, 158, 80, 96, 902, 166, 336, 170, 86, 174, 176, 178, 120, 182, 184, 186, 94, 190, 480};
static uint32_t cb_size(uint32_t idx) {
    if (idx < 60) return 40 + 8 * idx;
    if (idx < 92) return 512 + 16 * (idx - 59);
    if (idx < 124) return 1024 + 32 * (idx - 91);
    return 2048 + 64 * (idx - 123);
}

/* ---- turbo encoder, TS 36.212 5.1.3.2 ------------------------------------------------------------------------------------- */
int srsran_tcod_init(srsran_tcod_t* h, uint32_t max_long_cb) {
    h->max_long_cb = max_long_cb;
    h->temp = srsran_vec_u8_malloc(max_long_cb + 8);
    return h->temp ? SRSRAN_SUCCESS : SRSRAN_ERROR;
}
void srsran_tcod_free(srsran_tcod_t* h) { free(h->temp); h->temp = NULL; }

static inline void put_bit(uint8_t* p, uint32_t i, uint32_t b) {
    if (b) p[i >> 3] |= (uint8_t)(0x80 >> (i & 7));
    else p[i >> 3] &= (uint8_t)~(0x80 >> (i & 7));
}
static inline uint32_t get_bit(const uint8_t* p, uint32_t i) { return (p[i >> 3] >> (7 - (i & 7))) & 1; }

int srsran_tcod_encode_lut(srsran_tcod_t* h, srsran_crc_t* crc_tb, srsran_crc_t* crc_cb, uint8_t* input, uint8_t* parity,
                           uint32_t cblen_idx, bool last_cb) {
    if (cblen_idx >= 188) return SRSRAN_ERROR;
    const uint32_t K = cb_size(cblen_idx), nbytes = K / 8;
    if (K > h->max_long_cb) return SRSRAN_ERROR;
    /* payload bytes of this block (before any CRC): K/8 minus the CRCs that are appended here */
    uint32_t payload = nbytes - (crc_cb ? 3 : 0) - (last_cb ? 3 : 0);
    if (crc_tb) {
        for (uint32_t i = 0; i < payload; ++i) srsran_crc_checksum_put_byte(crc_tb, input[i]);
        if (last_cb) {
            const uint32_t c = (uint32_t)srsran_crc_checksum_get(crc_tb);
            for (int i = 0; i < crc_tb->order / 8; ++i) input[payload + i] = (uint8_t)(c >> (crc_tb->order - 8 - 8 * i));
        }
    }
    if (crc_cb) {
        const uint32_t n = nbytes - 3;
        const uint32_t c = srsran_crc_checksum_byte(crc_cb, input, (int)(8 * n));
        input[n] = (uint8_t)(c >> 16); input[n + 1] = (uint8_t)(c >> 8); input[n + 2] = (uint8_t)c;
    }
    /* two RSC encoders g0 = 1 + D^2 + D^3 (feedback), g1 = 1 + D + D^3 */
    const uint32_t f1 = QPP_F1[cblen_idx], f2 = QPP_F2[cblen_idx];
    memset(parity, 0, (2 * K + 12 + 7) / 8);
    uint32_t r1 = 0, r2 = 0;  /* registers: bit0 = D, bit1 = D^2, bit2 = D^3 */
    uint32_t w = 0;
    for (uint32_t k = 0; k < K; ++k) {
        const uint32_t c = get_bit(input, k);
        const uint32_t pk = (uint32_t)(((uint64_t)f1 * k + (uint64_t)f2 * k * k) % K);
        const uint32_t ci = get_bit(input, pk);
        uint32_t fb = c ^ ((r1 >> 1) & 1) ^ ((r1 >> 2) & 1);
        uint32_t z = fb ^ (r1 & 1) ^ ((r1 >> 2) & 1);
        r1 = ((r1 << 1) | fb) & 7;
        put_bit(parity, w++, z);
        fb = ci ^ ((r2 >> 1) & 1) ^ ((r2 >> 2) & 1);
        z = fb ^ (r2 & 1) ^ ((r2 >> 2) & 1);
        r2 = ((r2 << 1) | fb) & 7;
        put_bit(parity, w++, z);
    }
    /* trellis termination: x_K z_K x_K+1 z_K+1 x_K+2 z_K+2 x'_K z'_K x'_K+1 z'_K+1 x'_K+2 z'_K+2 */
    for (int e = 0; e < 2; ++e) {
        uint32_t* r = e ? &r2 : &r1;
        for (int t = 0; t < 3; ++t) {
            const uint32_t x = ((*r >> 1) & 1) ^ ((*r >> 2) & 1);   /* input that makes the feedback zero */
            const uint32_t z = (*r & 1) ^ ((*r >> 2) & 1);
            *r = (*r << 1) & 7;
            put_bit(parity, w++, x);
            put_bit(parity, w++, z);
        }
    }
    return SRSRAN_SUCCESS;
}

/* ---- rate matching, TS 36.212 5.1.4.1 ---------------------------------------------------------------------------------------- */
static const uint8_t RM_PERM[32] = {0, 16, 8, 24, 4, 20, 12, 28, 2, 18, 10, 26, 6, 22, 14, 30,
                                    1, 17, 9, 25, 5, 21, 13, 29, 3, 19, 11, 27, 7, 23, 15, 31};
/* circular buffer position -> index into the d-streams (3 * k + s), or -1 for a <NULL> */
static int32_t* rm_w2d(uint32_t K, uint32_t* Kw_out) {
    const uint32_t D = K + 4, R = (D + 31) / 32, Kp = 32 * R, ND = Kp - D, Kw = 3 * Kp;
    int32_t* t = (int32_t*)malloc(sizeof(int32_t) * Kw);
    for (uint32_t j = 0; j < Kp; ++j) {
        const uint32_t col = j / R, row = j % R;
        /* streams 0 and 1: y read column by column after the permutation */
        const int32_t y = (int32_t)(RM_PERM[col] + 32 * row) - (int32_t)ND;
        t[j] = y >= 0 ? 3 * y : -1;
        t[Kp + 2 * j] = y >= 0 ? 3 * y + 1 : -1;
        /* stream 2: pi(j) = (P[floor(j / R)] + 32 (j mod R) + 1) mod K_pi */
        const int32_t y2 = (int32_t)((RM_PERM[col] + 32 * row + 1) % Kp) - (int32_t)ND;
        t[Kp + 2 * j + 1] = y2 >= 0 ? 3 * y2 + 2 : -1;
    }
    *Kw_out = Kw;
    return t;
}
static int32_t* g_w2d[188];
static uint32_t g_Kw[188];
static const int32_t* w2d(uint32_t cb_idx) {
    if (!g_w2d[cb_idx]) {   /* tables built on first use (single-threaded set-up; the bridge warms them before threads) */
        g_w2d[cb_idx] = rm_w2d(cb_size(cb_idx), &g_Kw[cb_idx]);
    }
    return g_w2d[cb_idx];
}
void srsran_rm_turbo_gentables(void) {}
void srsran_rm_turbo_free_tables(void) {}
void shim_rm_warm(uint32_t cb_idx) { (void)w2d(cb_idx); }
static uint32_t rm_k0(uint32_t K, uint32_t rv) {
    const uint32_t R = (K + 4 + 31) / 32, Kw = 3 * 32 * R;
    return R * (2 * ((Kw + 8 * R - 1) / (8 * R)) * rv + 2);
}
/* d-stream bit (3 k + s) of the encoder output held as (systematic, parity) */
static inline uint32_t d_bit(const uint8_t* sys, const uint8_t* par, uint32_t K, uint32_t di) {
    const uint32_t k = di / 3, s = di % 3;
    if (k < K) return s == 0 ? get_bit(sys, k) : get_bit(par, 2 * k + (s - 1));
    /* termination, 36.212: d0 = x_K z_K+1 x'_K z'_K+1; d1 = z_K x_K+2 z'_K x'_K+2; d2 = x_K+1 z_K+2 x'_K+1 z'_K+2;
     * tail order in parity[2K ..]: x_K z_K x_K+1 z_K+1 x_K+2 z_K+2 x'_K z'_K x'_K+1 z'_K+1 x'_K+2 z'_K+2 */
    static const uint8_t tail_of[3][4] = {{0, 3, 6, 9}, {1, 4, 7, 10}, {2, 5, 8, 11}};
    return get_bit(par, 2 * K + tail_of[s][k - K]);
}
int srsran_rm_turbo_tx_lut(uint8_t* w_buff, uint8_t* systematic, uint8_t* parity, uint8_t* output, uint32_t cb_idx,
                           uint32_t out_len, uint32_t w_offset, uint32_t rv_idx) {
    if (rv_idx > 3 || cb_idx >= 188) return SRSRAN_ERROR;
    const uint32_t K = cb_size(cb_idx);
    const int32_t* t = w2d(cb_idx);
    const uint32_t Kw = g_Kw[cb_idx];
    if (rv_idx == 0)   /* the circular buffer is built once per code block and reused by later redundancy versions */
        for (uint32_t j = 0; j < Kw; ++j) w_buff[j] = t[j] < 0 ? 2 : (uint8_t)d_bit(systematic, parity, K, (uint32_t)t[j]);
    uint32_t j = rm_k0(K, rv_idx) % Kw, k = 0;
    while (k < out_len) {
        if (w_buff[j] != 2) { put_bit(output, w_offset + k, w_buff[j]); ++k; }
        if (++j == Kw) j = 0;
    }
    return SRSRAN_SUCCESS;
}
int srsran_rm_turbo_rx_lut_(int16_t* input, int16_t* output, uint32_t in_len, uint32_t cb_idx, uint32_t rv_idx, bool tdec) {
    (void)tdec;
    if (rv_idx > 3 || cb_idx >= 188) return SRSRAN_ERROR;
    const uint32_t K = cb_size(cb_idx);
    const int32_t* t = w2d(cb_idx);
    const uint32_t Kw = g_Kw[cb_idx];
    uint32_t j = rm_k0(K, rv_idx) % Kw, k = 0;
    while (k < in_len) {
        if (t[j] >= 0) { output[t[j]] = (int16_t)(output[t[j]] + input[k]); ++k; }   /* int16 wrap-around add */
        if (++j == Kw) j = 0;
    }
    return SRSRAN_SUCCESS;
}
int srsran_rm_turbo_rx_lut(int16_t* input, int16_t* output, uint32_t in_len, uint32_t cb_idx, uint32_t rv_idx) {
    return srsran_rm_turbo_rx_lut_(input, output, in_len, cb_idx, rv_idx, false);
}
int srsran_rm_turbo_rx_lut_8bit(int8_t* input, int8_t* output, uint32_t in_len, uint32_t cb_idx, uint32_t rv_idx) {
    if (rv_idx > 3 || cb_idx >= 188) return SRSRAN_ERROR;
    const uint32_t K = cb_size(cb_idx);
    const int32_t* t = w2d(cb_idx);
    const uint32_t Kw = g_Kw[cb_idx];
    uint32_t j = rm_k0(K, rv_idx) % Kw, k = 0;
    while (k < in_len) {
        if (t[j] >= 0) { output[t[j]] = (int8_t)(output[t[j]] + input[k]); ++k; }
        if (++j == Kw) j = 0;
    }
    return SRSRAN_SUCCESS;
}

/* ---- turbo decoder: max-log-MAP over the full trellis, int16 values with 32-bit intermediate sums ---------------------------
 * UNPINNED against srsRAN's srsran_tdec_iteration (fixed-point details differ between its generic / SSE / AVX variants
 * and are not reproducible from documentation).  Data flow as in srsRAN: decoder 1 on systematic + a-priori, its
 * a-posteriori minus the a-priori goes interleaved to decoder 2 as "systematic" input, decoder 2's a-posteriori comes back
 * de-interleaved; a-priori of the next iteration = that minus decoder 1's contribution; hard decision from decoder 2.
 * Metrics are normalised to state 0 every step and kept in int32, soft outputs saturate to int16. */
int srsran_tdec_init(srsran_tdec_t* h, uint32_t max_long_cb) {
    memset(h, 0, sizeof(*h));
    h->max_long_cb = max_long_cb;
    const uint32_t n = max_long_cb + 8;
    h->syst = srsran_vec_i16_malloc(n); h->par0 = srsran_vec_i16_malloc(n); h->par1 = srsran_vec_i16_malloc(n);
    h->apri = srsran_vec_i16_malloc(n); h->ext1 = srsran_vec_i16_malloc(n); h->llr2 = srsran_vec_i16_malloc(n);
    h->xin = srsran_vec_i16_malloc(n);
    h->alpha = srsran_vec_i16_malloc(2 * 8 * (n + 1));   /* int32 [K + 1][8] */
    h->pi = (uint16_t*)srsran_vec_malloc(sizeof(uint16_t) * n);
    return (h->syst && h->par0 && h->par1 && h->apri && h->ext1 && h->llr2 && h->xin && h->alpha && h->pi) ? SRSRAN_SUCCESS : SRSRAN_ERROR;
}
void srsran_tdec_free(srsran_tdec_t* h) {
    free(h->syst); free(h->par0); free(h->par1); free(h->apri); free(h->ext1); free(h->llr2); free(h->xin); free(h->alpha); free(h->pi);
    memset(h, 0, sizeof(*h));
}
void srsran_tdec_force_not_sb(srsran_tdec_t* h) { h->force_not_sb = true; }
int srsran_tdec_new_cb(srsran_tdec_t* h, uint32_t long_cb) {
    uint32_t idx = 0;
    while (idx < 188 && cb_size(idx) != long_cb) ++idx;
    if (idx == 188 || long_cb > h->max_long_cb + 8) return SRSRAN_ERROR;
    if (h->long_cb != long_cb) {
        const uint32_t f1 = QPP_F1[idx], f2 = QPP_F2[idx];
        for (uint32_t k = 0; k < long_cb; ++k) h->pi[k] = (uint16_t)(((uint64_t)f1 * k + (uint64_t)f2 * k * k) % long_cb);
    }
    h->long_cb = long_cb;
    h->n_iter = 0;
    return SRSRAN_SUCCESS;
}
static inline int16_t sat16(int v) { return (int16_t)(v > 32767 ? 32767 : v < -32768 ? -32768 : v); }
/* one constituent decoder: x[k] (systematic + a-priori), y[k] parity, tail (x, z) x 3; writes the a-posteriori LLR.
 * State s = s0 | s1 << 1 | s2 << 2 (register bits D, D^2, D^3); from s with input u: fb = u ^ s1 ^ s2, parity z = fb ^ s0 ^ s2,
 * next = (fb, s0, s1).  Seen from the next state ns = (fb, s0, s1) and the dropped bit s2 = c: s = ns1 | ns2 << 1 | c << 2,
 * u = ns0 ^ ns2 ^ c, z = ns0 ^ ns1 ^ c.  LLR > 0 <=> bit 1; branch metric (u ? x : 0) + (z ? y : 0); metrics in int32,
 * normalised to state 0 every step (state 0 is reachable at every step in both directions). */
#define TD_BIG (1 << 22)
#ifdef __AVX2__
#include <immintrin.h>
/* The same arithmetic with the eight state metrics of a trellis step in one 256-bit register (8 x int32): what a production
 * build gets from srsRAN's SIMD decoders, here only to make the CPU baseline less of a straw man.  Bit-identical to the
 * scalar form below (int32 add / max, same normalisation and clamps; tests/test_ref_cpu.py pins both against the oracle). */
static inline int32_t hmax8(__m256i v) {
    __m128i m = _mm_max_epi32(_mm256_castsi256_si128(v), _mm256_extracti128_si256(v, 1));
    m = _mm_max_epi32(m, _mm_shuffle_epi32(m, 0x4e));
    m = _mm_max_epi32(m, _mm_shuffle_epi32(m, 0xb1));
    return _mm_cvtsi128_si32(m);
}
static void map_dec(const int16_t* x, const int16_t* y, const int16_t* tail, uint32_t K, int32_t* alpha, int16_t* out) {
    int32_t ps0[8], ps1[8], pg0[8], pg1[8], nu0[8], nu1[8], gu0[8], gu1[8];
    for (int ns = 0; ns < 8; ++ns) {   /* predecessors of every next state (forward) */
        const int n0 = ns & 1, n1 = (ns >> 1) & 1, n2 = (ns >> 2) & 1;
        ps0[ns] = n1 | (n2 << 1);            pg0[ns] = ((n0 ^ n2) << 1) | (n0 ^ n1);
        ps1[ns] = n1 | (n2 << 1) | 4;        pg1[ns] = ((n0 ^ n2 ^ 1) << 1) | (n0 ^ n1 ^ 1);
    }
    for (int st = 0; st < 8; ++st) {   /* successors of every state for u = 0, 1 (backward) */
        const int s0 = st & 1, s1 = (st >> 1) & 1, s2 = (st >> 2) & 1;
        for (int u = 0; u < 2; ++u) {
            const int fb = u ^ s1 ^ s2, z = fb ^ s0 ^ s2, ns = fb | (s0 << 1) | (s1 << 2);
            if (u) { nu1[st] = ns; gu1[st] = 2 | z; } else { nu0[st] = ns; gu0[st] = z; }
        }
    }
    const __m256i iPS0 = _mm256_loadu_si256((const __m256i*)ps0), iPS1 = _mm256_loadu_si256((const __m256i*)ps1);
    const __m256i iPG0 = _mm256_loadu_si256((const __m256i*)pg0), iPG1 = _mm256_loadu_si256((const __m256i*)pg1);
    const __m256i iNU0 = _mm256_loadu_si256((const __m256i*)nu0), iNU1 = _mm256_loadu_si256((const __m256i*)nu1);
    const __m256i iGU0 = _mm256_loadu_si256((const __m256i*)gu0), iGU1 = _mm256_loadu_si256((const __m256i*)gu1);
    const __m256i vmin = _mm256_set1_epi32(-TD_BIG), zero = _mm256_setzero_si256();
    __m256i a = _mm256_setr_epi32(0, -TD_BIG, -TD_BIG, -TD_BIG, -TD_BIG, -TD_BIG, -TD_BIG, -TD_BIG);
    _mm256_storeu_si256((__m256i*)alpha, a);
    for (uint32_t k = 0; k < K; ++k) {
        const int32_t xv = x[k], yv = y[k];
        const __m256i g = _mm256_setr_epi32(0, yv, xv, xv + yv, 0, yv, xv, xv + yv);
        const __m256i m0 = _mm256_add_epi32(_mm256_permutevar8x32_epi32(a, iPS0), _mm256_permutevar8x32_epi32(g, iPG0));
        const __m256i m1 = _mm256_add_epi32(_mm256_permutevar8x32_epi32(a, iPS1), _mm256_permutevar8x32_epi32(g, iPG1));
        const __m256i t = _mm256_max_epi32(m0, m1);
        a = _mm256_max_epi32(_mm256_sub_epi32(t, _mm256_permutevar8x32_epi32(t, zero)), vmin);
        _mm256_storeu_si256((__m256i*)(alpha + 8 * (k + 1)), a);
    }
    int32_t b8[8], bn8[8];
    for (int st = 0; st < 8; ++st) b8[st] = st == 0 ? 0 : -TD_BIG;
    for (int t = 2; t >= 0; --t) {   /* termination, three steps: scalar */
        for (int st = 0; st < 8; ++st) {
            const int u = ((st >> 1) & 1) ^ ((st >> 2) & 1), z = (st & 1) ^ ((st >> 2) & 1), ns = (st << 1) & 7;
            const int32_t v = b8[ns] + (u ? tail[2 * t] : 0) + (z ? tail[2 * t + 1] : 0);
            bn8[st] = v < -TD_BIG ? -TD_BIG : v;
        }
        memcpy(b8, bn8, sizeof(b8));
    }
    __m256i b = _mm256_loadu_si256((const __m256i*)b8);
    for (int k = (int)K - 1; k >= 0; --k) {
        const __m256i ak = _mm256_loadu_si256((const __m256i*)(alpha + 8 * k));
        const int32_t xv = x[k], yv = y[k];
        const __m256i g = _mm256_setr_epi32(0, yv, xv, xv + yv, 0, yv, xv, xv + yv);
        const __m256i bm0 = _mm256_add_epi32(_mm256_permutevar8x32_epi32(b, iNU0), _mm256_permutevar8x32_epi32(g, iGU0));
        const __m256i bm1 = _mm256_add_epi32(_mm256_permutevar8x32_epi32(b, iNU1), _mm256_permutevar8x32_epi32(g, iGU1));
        const int32_t l0 = hmax8(_mm256_add_epi32(ak, bm0)), l1 = hmax8(_mm256_add_epi32(ak, bm1));
        out[k] = sat16(l1 - l0);
        const __m256i bn = _mm256_max_epi32(bm0, bm1);
        b = _mm256_max_epi32(_mm256_sub_epi32(bn, _mm256_permutevar8x32_epi32(bn, zero)), vmin);
    }
}
#define map_dec_scalar map_dec_unused
#else
#define map_dec_scalar map_dec
#endif
__attribute__((unused)) static void map_dec_scalar(const int16_t* x, const int16_t* y, const int16_t* tail, uint32_t K, int32_t* alpha, int16_t* out) {
    /* predecessor pairs of every next state: [ns][c] -> (s, u, z) */
    int ps[8][2], pg[8][2];
    for (int ns = 0; ns < 8; ++ns)
        for (int c = 0; c < 2; ++c) {
            const int n0 = ns & 1, n1 = (ns >> 1) & 1, n2 = (ns >> 2) & 1;
            ps[ns][c] = n1 | (n2 << 1) | (c << 2);
            pg[ns][c] = ((n0 ^ n2 ^ c) << 1) | (n0 ^ n1 ^ c);   /* index into g[] = {0, y, x, x + y} */
        }
    int32_t* a = alpha;
    a[0] = 0;
    for (int s = 1; s < 8; ++s) a[s] = -TD_BIG;
    for (uint32_t k = 0; k < K; ++k) {
        const int32_t g[4] = {0, y[k], x[k], (int32_t)x[k] + y[k]};
        int32_t* an = a + 8;
        int32_t t[8];
        for (int ns = 0; ns < 8; ++ns) {
            const int32_t m0 = a[ps[ns][0]] + g[pg[ns][0]], m1 = a[ps[ns][1]] + g[pg[ns][1]];
            t[ns] = m0 > m1 ? m0 : m1;
        }
        const int32_t n0 = t[0];
        for (int ns = 0; ns < 8; ++ns) { const int32_t v = t[ns] - n0; an[ns] = v < -TD_BIG ? -TD_BIG : v; }
        a = an;
    }
    /* beta at K from the termination: three steps with input u = s1 ^ s2 (feedback zero), parity z = s0 ^ s2 */
    int32_t b[8], bn[8];
    for (int s = 0; s < 8; ++s) b[s] = s == 0 ? 0 : -TD_BIG;
    for (int t = 2; t >= 0; --t) {
        for (int s = 0; s < 8; ++s) {
            const int u = ((s >> 1) & 1) ^ ((s >> 2) & 1), z = (s & 1) ^ ((s >> 2) & 1), ns = (s << 1) & 7;
            const int32_t v = b[ns] + (u ? tail[2 * t] : 0) + (z ? tail[2 * t + 1] : 0);
            bn[s] = v < -TD_BIG ? -TD_BIG : v;
        }
        memcpy(b, bn, sizeof(b));
    }
    for (int k = (int)K - 1; k >= 0; --k) {
        const int32_t* ak = alpha + 8 * k;
        const int32_t g[4] = {0, y[k], x[k], (int32_t)x[k] + y[k]};
        int32_t m1 = -4 * TD_BIG, m0 = -4 * TD_BIG;
        for (int s = 0; s < 8; ++s) bn[s] = -4 * TD_BIG;
        for (int ns = 0; ns < 8; ++ns)
            for (int c = 0; c < 2; ++c) {
                const int s = ps[ns][c], gi = pg[ns][c];
                const int32_t bm = b[ns] + g[gi];
                if (bm > bn[s]) bn[s] = bm;
                const int32_t tot = ak[s] + bm;
                if (gi & 2) { if (tot > m1) m1 = tot; } else { if (tot > m0) m0 = tot; }
            }
        out[k] = sat16(m1 - m0);
        const int32_t n0 = bn[0];
        for (int s = 0; s < 8; ++s) { const int32_t v = bn[s] - n0; b[s] = v < -TD_BIG ? -TD_BIG : v; }
    }
}
void srsran_tdec_iteration(srsran_tdec_t* h, int16_t* input, uint8_t* output) {
    const uint32_t K = h->long_cb;
    int32_t* alpha = (int32_t*)h->alpha;
    if (h->n_iter == 0) {
        /* input: d-stream layout [3 k + s] (srsran_rm_turbo_rx_lut above) */
        for (uint32_t k = 0; k < K; ++k) {
            h->syst[k] = input[3 * k]; h->par0[k] = input[3 * k + 1]; h->par1[k] = input[3 * k + 2]; h->apri[k] = 0;
        }
        /* termination, TS 36.212 5.1.3.2.2: d0 = x_K z_K+1 x'_K z'_K+1; d1 = z_K x_K+2 z'_K x'_K+2; d2 = x_K+1 z_K+2 x'_K+1
         * z'_K+2, i.e. input[3 K ..] = x_K z_K x_K+1 z_K+1 x_K+2 z_K+2 x'_K z'_K x'_K+1 z'_K+1 x'_K+2 z'_K+2 */
        for (int i = 0; i < 6; ++i) { h->tail0[i] = input[3 * K + i]; h->tail1[i] = input[3 * K + 6 + i]; }
    }
    for (uint32_t k = 0; k < K; ++k) h->xin[k] = sat16((int)h->syst[k] + h->apri[k]);
    map_dec(h->xin, h->par0, h->tail0, K, alpha, h->ext1);
    /* a-posteriori minus a-priori: systematic + extrinsic of decoder 1, decoder 2's "systematic" input */
    for (uint32_t k = 0; k < K; ++k) h->ext1[k] = sat16((int)h->ext1[k] - h->apri[k]);
    for (uint32_t k = 0; k < K; ++k) h->xin[k] = h->ext1[h->pi[k]];
    map_dec(h->xin, h->par1, h->tail1, K, alpha, h->llr2);
    for (uint32_t k = 0; k < K; ++k) {
        const uint32_t p = h->pi[k];
        const int16_t l = h->llr2[k];
        put_bit(output, p, l > 0);
        h->apri[p] = sat16((int)l - h->ext1[p]);   /* extrinsic of decoder 2 = a-priori of the next iteration */
    }
    h->n_iter++;
}
void srsran_tdec_iteration_8bit(srsran_tdec_t* h, int8_t* input, uint8_t* output) {
    const uint32_t K = h->long_cb;
    int16_t* tmp = srsran_vec_i16_malloc(3 * K + 12);
    for (uint32_t i = 0; i < 3 * K + 12; ++i) tmp[i] = input[i];
    srsran_tdec_iteration(h, tmp, output);
    free(tmp);
}

/* ---- DFT: unnormalised mixed-radix (2, 3, 4, 5, generic) decimation in time, float data, double-precision twiddles ---------- */
typedef struct { float re, im; } c32;
static void dft_rec(const c32* in, size_t stride, c32* out, int n, int N, const c32* tw, c32* tmp) {
    if (n == 1) { out[0] = in[0]; return; }
    int p = 2;
    if (n % 4 == 0) p = 4;
    else if (n % 2 == 0) p = 2;
    else if (n % 3 == 0) p = 3;
    else if (n % 5 == 0) p = 5;
    else { p = n; for (int q = 7; q * q <= n; q += 2) if (n % q == 0) { p = q; break; } }
    const int m = n / p, step = N / n;
    for (int r = 0; r < p; ++r) dft_rec(in + (size_t)r * stride, stride * p, out + (size_t)r * m, m, N, tw, tmp);
    for (int k = 0; k < m; ++k) {
        c32 v[16];
        c32* vv = p <= 16 ? v : tmp;
        for (int r = 0; r < p; ++r) {
            const c32 w = tw[(size_t)((r * k) % n) * step], a = out[(size_t)r * m + k];
            vv[r].re = a.re * w.re - a.im * w.im;
            vv[r].im = a.re * w.im + a.im * w.re;
        }
        if (p == 2) {
            out[k].re = vv[0].re + vv[1].re; out[k].im = vv[0].im + vv[1].im;
            out[k + m].re = vv[0].re - vv[1].re; out[k + m].im = vv[0].im - vv[1].im;
        } else if (p == 4) {
            /* sign of the quarter turn follows the transform direction: tw[N/4] = -j (forward) or +j (backward) */
            const float sj = tw[(size_t)(n / 4) * step].im;
            const c32 s02 = {vv[0].re + vv[2].re, vv[0].im + vv[2].im}, d02 = {vv[0].re - vv[2].re, vv[0].im - vv[2].im};
            const c32 s13 = {vv[1].re + vv[3].re, vv[1].im + vv[3].im}, d13 = {vv[1].re - vv[3].re, vv[1].im - vv[3].im};
            const c32 jd = {-sj * d13.im, sj * d13.re};   /* (sj j) * d13 */
            out[k].re = s02.re + s13.re; out[k].im = s02.im + s13.im;
            out[k + m].re = d02.re + jd.re; out[k + m].im = d02.im + jd.im;
            out[k + 2 * m].re = s02.re - s13.re; out[k + 2 * m].im = s02.im - s13.im;
            out[k + 3 * m].re = d02.re - jd.re; out[k + 3 * m].im = d02.im - jd.im;
        } else {
            for (int q = 0; q < p; ++q) {
                float sr = 0.f, si = 0.f;
                for (int r = 0; r < p; ++r) {
                    const c32 w = tw[(size_t)((r * q * m) % n) * step];
                    sr += vv[r].re * w.re - vv[r].im * w.im;
                    si += vv[r].re * w.im + vv[r].im * w.re;
                }
                tmp[p + q].re = sr; tmp[p + q].im = si;
            }
            for (int q = 0; q < p; ++q) out[k + (size_t)q * m] = tmp[p + q];
        }
    }
}
int srsran_dft_plan_c(srsran_dft_plan_t* plan, int N, srsran_dft_dir_t dir) {
    memset(plan, 0, sizeof(*plan));
    plan->size = plan->init_size = N;
    plan->dir = dir;
    plan->forward = dir == SRSRAN_DFT_FORWARD;
    plan->mode = SRSRAN_DFT_COMPLEX;
    c32* tw = (c32*)srsran_vec_malloc(sizeof(c32) * (size_t)N);
    for (int i = 0; i < N; ++i) {
        const double a = 2.0 * M_PI * (double)i / (double)N;
        tw[i].re = (float)cos(a);
        tw[i].im = (float)(plan->forward ? -sin(a) : sin(a));
    }
    plan->p = tw;
    plan->in = srsran_vec_malloc(sizeof(c32) * (size_t)N);
    plan->out = srsran_vec_malloc(sizeof(c32) * (size_t)N);
    plan->scratch = srsran_vec_malloc(sizeof(c32) * (size_t)(2 * N + 32));
    return (plan->p && plan->in && plan->out && plan->scratch) ? SRSRAN_SUCCESS : SRSRAN_ERROR;
}
void srsran_dft_plan_free(srsran_dft_plan_t* plan) {
    free(plan->p); free(plan->in); free(plan->out); free(plan->scratch);
    memset(plan, 0, sizeof(*plan));
}
void srsran_dft_run_c_zerocopy(srsran_dft_plan_t* plan, const cf_t* in, cf_t* out) {
    const c32* src = (const c32*)in;
    if ((const void*)in == (void*)out) {   /* in place: go through the plan's buffer */
        memcpy(plan->in, in, sizeof(c32) * (size_t)plan->size);
        src = (const c32*)plan->in;
    }
    dft_rec(src, 1, (c32*)out, plan->size, plan->size, (const c32*)plan->p, (c32*)plan->scratch);
}
void srsran_dft_run_c(srsran_dft_plan_t* plan, const cf_t* in, cf_t* out) { srsran_dft_run_c_zerocopy(plan, in, out); }

/* ---- AWGN: Box-Muller on xorshift64* (srsRAN's generator is table driven; only the statistics are the same) ------------------- */
int srsran_channel_awgn_init(srsran_channel_awgn_t* q, uint32_t seed) {
    q->std_dev = 0.f;
    q->state = 0x9E3779B97F4A7C15ull ^ ((uint64_t)seed * 0xD1342543DE82EF95ull + 1);
    return SRSRAN_SUCCESS;
}
int srsran_channel_awgn_set_n0(srsran_channel_awgn_t* q, float n0_dBfs) {
    q->std_dev = powf(10.f, n0_dBfs / 20.f);
    return SRSRAN_SUCCESS;
}
static inline double u01(uint64_t* s) {
    uint64_t x = *s;
    x ^= x >> 12; x ^= x << 25; x ^= x >> 27;
    *s = x;
    return (double)((x * 0x2545F4914F6CDD1Dull) >> 11) * (1.0 / 9007199254740992.0);
}
void srsran_channel_awgn_run_c(srsran_channel_awgn_t* q, const cf_t* in, cf_t* out, uint32_t size) {
    const float sd = q->std_dev * (float)M_SQRT1_2;
    for (uint32_t i = 0; i < size; ++i) {
        double u1 = u01(&q->state), u2 = u01(&q->state);
        if (u1 < 1e-300) u1 = 1e-300;
        const double r = sqrt(-2.0 * log(u1));
        cf_t n;
        RE(n) = sd * (float)(r * cos(2.0 * M_PI * u2));
        IM(n) = sd * (float)(r * sin(2.0 * M_PI * u2));
        out[i] = in[i] + n;
    }
}
void srsran_channel_awgn_free(srsran_channel_awgn_t* q) { (void)q; }

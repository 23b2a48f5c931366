/* srsran/phy/fec/turbo/turbocoder.h stand-in: TS 36.212 section 5.1.3.2 turbo encoder on packed bytes */
#ifndef SHIM_SRSRAN_TURBOCODER_H
#define SHIM_SRSRAN_TURBOCODER_H
#include "srsran/phy/fec/crc.h"
#ifdef __cplusplus
extern "C" {
#endif
typedef struct {
    uint32_t max_long_cb;
    uint8_t* temp;
} srsran_tcod_t;
int srsran_tcod_init(srsran_tcod_t* h, uint32_t max_long_cb);
void srsran_tcod_free(srsran_tcod_t* h);
/* input: packed code block (K / 8 bytes with room for the CRCs); appends the transport block CRC when last_cb (running
 * checksum in crc_tb over the payload bytes of all code blocks) and the code block CRC when crc_cb != NULL; rewrites
 * input with the CRCs in place; parity: packed, bit 2k = first parity of step k, 2k+1 = second parity, then 12 tail bits
 * in the order of 36.212 (x_K z_K x_K+1 z_K+1 ... x'_K+2 z'_K+2) */
int srsran_tcod_encode_lut(srsran_tcod_t* h, srsran_crc_t* crc_tb, srsran_crc_t* crc_cb, uint8_t* input, uint8_t* parity,
                           uint32_t cblen_idx, bool last_cb);
#ifdef __cplusplus
}
#endif
#endif

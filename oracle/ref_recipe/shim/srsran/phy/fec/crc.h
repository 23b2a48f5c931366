/* srsran/phy/fec/crc.h stand-in: MSB-first CRC with init 0 over packed bytes (TS 36.212 section 5.1.1) */
#ifndef SHIM_SRSRAN_CRC_H
#define SHIM_SRSRAN_CRC_H
#include "srsran/config.h"
#define SRSRAN_LTE_CRC24A 0x1864CFB
#define SRSRAN_LTE_CRC24B 0X1800063
#define SRSRAN_LTE_CRC16 0x11021
#define SRSRAN_LTE_CRC8 0x19B
#ifdef __cplusplus
extern "C" {
#endif
typedef struct {
    uint64_t table[256];
    int polynom;
    int order;
    uint64_t crcinit;   /* running remainder (srsran_crc_checksum_put_byte) */
    uint64_t crcmask;
    uint64_t crchighbit;
} srsran_crc_t;
int srsran_crc_init(srsran_crc_t* h, uint32_t crc_poly, int crc_order);
int srsran_crc_set_init(srsran_crc_t* h, uint64_t init_value);
void srsran_crc_checksum_put_byte(srsran_crc_t* h, uint8_t byte);
uint64_t srsran_crc_checksum_get(srsran_crc_t* h);
/* remainder of len bits (multiple of 8) */
uint32_t srsran_crc_checksum_byte(srsran_crc_t* h, const uint8_t* data, int len);
/* appends the order / 8 checksum bytes behind len bits */
uint32_t srsran_crc_attach_byte(srsran_crc_t* h, uint8_t* data, int len);
/* recomputes the checksum of len bits and compares it with the bytes behind them */
bool srsran_crc_match_byte(srsran_crc_t* h, uint8_t* data, int len);
#ifdef __cplusplus
}
#endif
#endif

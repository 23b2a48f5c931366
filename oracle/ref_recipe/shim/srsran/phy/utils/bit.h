#ifndef SHIM_SRSRAN_BIT_H
#define SHIM_SRSRAN_BIT_H
#include "srsran/config.h"
#ifdef __cplusplus
extern "C" {
#endif
/* packed bytes (MSB first) <-> one bit per byte */
void srsran_bit_unpack_vector(const uint8_t* packed, uint8_t* unpacked, int nof_bits);
void srsran_bit_pack_vector(uint8_t* unpacked, uint8_t* packed, int nof_bits);
#ifdef __cplusplus
}
#endif
#endif

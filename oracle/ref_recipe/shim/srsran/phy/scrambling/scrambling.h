#ifndef SHIM_SRSRAN_SCRAMBLING_H
#define SHIM_SRSRAN_SCRAMBLING_H
#include "srsran/phy/common/sequence.h"
#ifdef __cplusplus
extern "C" {
#endif
/* LLR sign flip where the sequence bit is 1 (sequence read from bit `offset`) */
void srsran_scrambling_s_offset(srsran_sequence_t* s, short* data, int offset, int len);
void srsran_scrambling_sb_offset(srsran_sequence_t* s, int8_t* data, int offset, int len);
/* xor of packed bytes with the packed sequence, len in bits */
void srsran_scrambling_bytes(srsran_sequence_t* s, uint8_t* data, int len);
#ifdef __cplusplus
}
#endif
#endif

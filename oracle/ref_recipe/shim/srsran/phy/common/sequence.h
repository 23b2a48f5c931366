/* srsran/phy/common/sequence.h stand-in: TS 36.211 section 7.2 pseudo-random (Gold) sequence, Nc = 1600 */
#ifndef SHIM_SRSRAN_SEQUENCE_H
#define SHIM_SRSRAN_SEQUENCE_H
#include "srsran/config.h"
#include "srsran/phy/common/phy_common.h"
#ifdef __cplusplus
extern "C" {
#endif
typedef struct {
    uint8_t* c;        /* one bit per byte */
    uint8_t* c_bytes;  /* packed, MSB first */
    float* c_float;    /* +1 / -1 */
    short* c_short;
    int8_t* c_char;
    uint32_t cur_len;
    uint32_t max_len;
} srsran_sequence_t;
int srsran_sequence_LTE_pr(srsran_sequence_t* q, uint32_t len, uint32_t seed);
void srsran_sequence_free(srsran_sequence_t* q);
#ifdef __cplusplus
}
#endif
#endif

#ifndef SHIM_SRSRAN_DEMOD_SOFT_H
#define SHIM_SRSRAN_DEMOD_SOFT_H
#include "srsran/phy/modem/modem_table.h"
#ifdef __cplusplus
extern "C" {
#endif
/* max-log piece-wise-linear LLRs as int16 / int8, LLR > 0 <=> bit 1 (see shim_srsran.c for the scale constants) */
int srsran_demod_soft_demodulate_s(srsran_mod_t modulation, const cf_t* symbols, short* llr, int nsymbols);
int srsran_demod_soft_demodulate_b(srsran_mod_t modulation, const cf_t* symbols, int8_t* llr, int nsymbols);
#ifdef __cplusplus
}
#endif
#endif

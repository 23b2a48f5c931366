/* srsran/phy/fec/turbo/turbodecoder.h stand-in: int16 max-log-MAP turbo decoder over the FULL trellis (no windows),
 * one srsran_tdec_iteration = both constituent decoders + hard decision (see shim_srsran.c) */
#ifndef SHIM_SRSRAN_TURBODECODER_H
#define SHIM_SRSRAN_TURBODECODER_H
#include "srsran/config.h"
#include "srsran/phy/fec/cbsegm.h"
#ifdef __cplusplus
extern "C" {
#endif
typedef struct {
    uint32_t max_long_cb;
    uint32_t long_cb;
    uint32_t n_iter;
    int16_t *syst, *par0, *par1, *apri, *ext1, *llr2, *xin;
    int16_t *alpha;   /* [K + 1][8] */
    uint16_t* pi;     /* QPP permutation of the current block length */
    int16_t tail0[6], tail1[6];
    bool force_not_sb;
} srsran_tdec_t;
int srsran_tdec_init(srsran_tdec_t* h, uint32_t max_long_cb);
void srsran_tdec_free(srsran_tdec_t* h);
void srsran_tdec_force_not_sb(srsran_tdec_t* h);
int srsran_tdec_new_cb(srsran_tdec_t* h, uint32_t long_cb);
void srsran_tdec_iteration(srsran_tdec_t* h, int16_t* input, uint8_t* output);
void srsran_tdec_iteration_8bit(srsran_tdec_t* h, int8_t* input, uint8_t* output);
#ifdef __cplusplus
}
#endif
#endif

#ifndef SHIM_SRSRAN_CBSEGM_H
#define SHIM_SRSRAN_CBSEGM_H
#include "srsran/config.h"
#define SRSRAN_NOF_TC_CB_SIZES 188
typedef struct {
    uint32_t F, C, K1, K2, K1_idx, K2_idx, C1, C2, tbs, L_tb, L_cb, Z;
} srsran_cbsegm_t;
#endif

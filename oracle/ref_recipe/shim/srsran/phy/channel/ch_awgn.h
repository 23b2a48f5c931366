#ifndef SHIM_SRSRAN_CH_AWGN_H
#define SHIM_SRSRAN_CH_AWGN_H
#include "srsran/config.h"
#ifdef __cplusplus
extern "C" {
#endif
typedef struct {
    float std_dev;
    uint64_t state;   /* xorshift64* state (srsRAN uses its own table-driven generator: realisations differ) */
} srsran_channel_awgn_t;
int srsran_channel_awgn_init(srsran_channel_awgn_t* q, uint32_t seed);
int srsran_channel_awgn_set_n0(srsran_channel_awgn_t* q, float n0_dBfs);
void srsran_channel_awgn_run_c(srsran_channel_awgn_t* q, const cf_t* in, cf_t* out, uint32_t size);
void srsran_channel_awgn_free(srsran_channel_awgn_t* q);
#ifdef __cplusplus
}
#endif
#endif

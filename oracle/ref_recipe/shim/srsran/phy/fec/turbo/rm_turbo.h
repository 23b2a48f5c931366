/* srsran/phy/fec/turbo/rm_turbo.h stand-in: TS 36.212 section 5.1.4.1 rate matching, N_cb = K_w (no N_soft limit) */
#ifndef SHIM_SRSRAN_RM_TURBO_H
#define SHIM_SRSRAN_RM_TURBO_H
#include "srsran/config.h"
#ifdef __cplusplus
extern "C" {
#endif
void srsran_rm_turbo_gentables(void);
void srsran_rm_turbo_free_tables(void);
int srsran_rm_turbo_tx_lut(uint8_t* w_buff, uint8_t* systematic, uint8_t* parity, uint8_t* output, uint32_t cb_idx,
                           uint32_t out_len, uint32_t w_offset, uint32_t rv_idx);
/* accumulates (int16 wrap-around adds) the in_len soft values onto the d-stream positions of output [3 (K + 4)]:
 * output[3 k + s] = soft value of d_s[k] */
int srsran_rm_turbo_rx_lut(int16_t* input, int16_t* output, uint32_t in_len, uint32_t cb_idx, uint32_t rv_idx);
int srsran_rm_turbo_rx_lut_(int16_t* input, int16_t* output, uint32_t in_len, uint32_t cb_idx, uint32_t rv_idx,
                            bool enable_input_tdec);
int srsran_rm_turbo_rx_lut_8bit(int8_t* input, int8_t* output, uint32_t in_len, uint32_t cb_idx, uint32_t rv_idx);
#ifdef __cplusplus
}
#endif
#endif

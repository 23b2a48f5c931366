#ifndef SHIM_SRSRAN_SOFTBUFFER_H
#define SHIM_SRSRAN_SOFTBUFFER_H
#include "srsran/config.h"
#define SRSRAN_MAX_CODEBLOCKS 32
#ifdef __cplusplus
extern "C" {
#endif
typedef struct {
    uint32_t max_cb;
    uint32_t max_cb_size;
    int16_t** buffer_f;   /* per code block: de-rate-matched soft values */
    uint8_t** data;       /* per code block: bytes of blocks decoded in an earlier transmission */
    bool* cb_crc;
    bool tb_crc;
} srsran_softbuffer_rx_t;
typedef struct {
    uint32_t max_cb;
    uint32_t max_cb_size;
    uint8_t** buffer_b;   /* per code block: circular buffer of the rate matcher */
} srsran_softbuffer_tx_t;
int srsran_softbuffer_rx_init_guru(srsran_softbuffer_rx_t* q, uint32_t max_cb, uint32_t max_cb_size);
void srsran_softbuffer_rx_free(srsran_softbuffer_rx_t* q);
void srsran_softbuffer_rx_reset(srsran_softbuffer_rx_t* q);
void srsran_softbuffer_rx_reset_cb(srsran_softbuffer_rx_t* q, uint32_t nof_cb);
void srsran_softbuffer_rx_reset_cb_crc(srsran_softbuffer_rx_t* q, uint32_t nof_cb);
int srsran_softbuffer_tx_init_guru(srsran_softbuffer_tx_t* q, uint32_t max_cb, uint32_t max_cb_size);
void srsran_softbuffer_tx_free(srsran_softbuffer_tx_t* q);
void srsran_softbuffer_tx_reset(srsran_softbuffer_tx_t* q);
void srsran_softbuffer_tx_reset_cb(srsran_softbuffer_tx_t* q, uint32_t nof_cb);
#ifdef __cplusplus
}
#endif
#endif

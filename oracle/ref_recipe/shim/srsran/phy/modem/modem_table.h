/* srsran/phy/modem/modem_table.h stand-in: TS 36.211 section 7.1 constellations (unit average power) */
#ifndef SHIM_SRSRAN_MODEM_TABLE_H
#define SHIM_SRSRAN_MODEM_TABLE_H
#include "srsran/config.h"
#include "srsran/phy/common/phy_common.h"
#ifdef __cplusplus
extern "C" {
#endif
typedef struct { cf_t symbol[8]; } bpsk_packed_t;
typedef struct { cf_t symbol[4]; } qpsk_packed_t;
typedef struct { cf_t symbol[2]; } qam16_packed_t;
typedef struct {
    cf_t* symbol_table;
    uint32_t nsymbols;
    uint32_t nbits_x_symbol;
    bool byte_tables_init;
    bpsk_packed_t* symbol_table_bpsk;
    qpsk_packed_t* symbol_table_qpsk;
    qam16_packed_t* symbol_table_16qam;
} srsran_modem_table_t;
int srsran_modem_table_lte(srsran_modem_table_t* q, srsran_mod_t modulation);
void srsran_modem_table_bytes(srsran_modem_table_t* q);
void srsran_modem_table_free(srsran_modem_table_t* q);
#ifdef __cplusplus
}
#endif
#endif

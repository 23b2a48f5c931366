#ifndef SHIM_SRSRAN_MOD_H
#define SHIM_SRSRAN_MOD_H
#include "srsran/phy/modem/modem_table.h"
#ifdef __cplusplus
extern "C" {
#endif
int srsran_mod_modulate(const srsran_modem_table_t* q, uint8_t* bits, cf_t* symbols, uint32_t nbits);
#ifdef __cplusplus
}
#endif
#endif

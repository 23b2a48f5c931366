#ifndef SHIM_SRSRAN_VECTOR_SIMD_H
#define SHIM_SRSRAN_VECTOR_SIMD_H
#include "srsran/phy/utils/vector.h"
#ifdef __cplusplus
extern "C" {
#endif
void srsran_vec_sc_prod_cfc_simd(const cf_t* x, float h, cf_t* z, int n);
void srsran_vec_sc_prod_ccc_simd(const cf_t* x, cf_t h, cf_t* z, int n);
void srsran_vec_add_fff_simd(const float* x, const float* y, float* z, int n);
#ifdef __cplusplus
}
#endif
#endif

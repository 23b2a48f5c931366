/* srsran/phy/utils/vector.h stand-in: the allocation and element-wise helpers the reference calls, in plain C */
#ifndef SHIM_SRSRAN_VECTOR_H
#define SHIM_SRSRAN_VECTOR_H
#include <stdlib.h>
#include <string.h>
#include "srsran/config.h"
#ifdef __cplusplus
extern "C" {
#endif
void* srsran_vec_malloc(uint32_t size);
cf_t* srsran_vec_cf_malloc(uint32_t n);
float* srsran_vec_f_malloc(uint32_t n);
uint8_t* srsran_vec_u8_malloc(uint32_t n);
int16_t* srsran_vec_i16_malloc(uint32_t n);
void srsran_vec_cf_zero(cf_t* p, uint32_t n);
void srsran_vec_f_zero(float* p, uint32_t n);
void srsran_vec_u8_zero(uint8_t* p, uint32_t n);
void srsran_vec_cf_copy(cf_t* dst, const cf_t* src, uint32_t n);
void srsran_vec_f_copy(float* dst, const float* src, uint32_t n);
void srsran_vec_u8_copy(uint8_t* dst, const uint8_t* src, uint32_t n);
void srsran_vec_sum_ccc(const cf_t* x, const cf_t* y, cf_t* z, uint32_t n);
void srsran_vec_sub_ccc(const cf_t* x, const cf_t* y, cf_t* z, uint32_t n);
void srsran_vec_sub_fff(const float* x, const float* y, float* z, uint32_t n);
void srsran_vec_prod_fff(const float* x, const float* y, float* z, uint32_t n);
void srsran_vec_sc_prod_cfc(const cf_t* x, float h, cf_t* z, uint32_t n);
void srsran_vec_sc_prod_ccc(const cf_t* x, cf_t h, cf_t* z, uint32_t n);
void srsran_vec_xor_bbb(const uint8_t* x, const uint8_t* y, uint8_t* z, uint32_t n);
uint32_t srsran_vec_max_abs_fi(const float* x, uint32_t n);
uint32_t srsran_vec_max_abs_ci(const cf_t* x, uint32_t n);
#ifdef __cplusplus
}
#endif
#endif

/* srsran/phy/dft/dft.h stand-in (srsRAN wraps FFTW3f): unnormalised complex DFT, no mirroring, DC kept */
#ifndef SHIM_SRSRAN_DFT_H
#define SHIM_SRSRAN_DFT_H
#include "srsran/config.h"
#ifdef __cplusplus
extern "C" {
#endif
typedef enum { SRSRAN_DFT_COMPLEX, SRSRAN_DFT_REAL } srsran_dft_mode_t;
typedef enum { SRSRAN_DFT_FORWARD, SRSRAN_DFT_BACKWARD } srsran_dft_dir_t;
typedef struct {
    int init_size;
    int size;
    void* in;
    void* out;
    void* p;        /* twiddles (double precision) */
    void* scratch;
    bool is_guru, forward, mirror, db, norm, dc;
    srsran_dft_dir_t dir;
    srsran_dft_mode_t mode;
} srsran_dft_plan_t;
int srsran_dft_plan_c(srsran_dft_plan_t* plan, int dft_points, srsran_dft_dir_t dir);
void srsran_dft_plan_free(srsran_dft_plan_t* plan);
void srsran_dft_run_c(srsran_dft_plan_t* plan, const cf_t* in, cf_t* out);
void srsran_dft_run_c_zerocopy(srsran_dft_plan_t* plan, const cf_t* in, cf_t* out);
#ifdef __cplusplus
}
#endif
#endif

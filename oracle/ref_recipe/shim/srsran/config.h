/* Stand-in for srsRAN_4G's srsran/config.h (the library is not installed on this image).  Everything under shim/ restates
 * the SEMANTICS of third-party calls the reference makes (srsRAN_4G release_23_11, VOLK 3.1, Eigen 3, Boost.Outcome) so
 * that the reference's own sources compile UNMODIFIED from /root/reference into oracle/_ref/ (test infrastructure only,
 * see oracle/ref_recipe/README.md).  No reference file is copied or edited. */
#ifndef SHIM_SRSRAN_CONFIG_H
#define SHIM_SRSRAN_CONFIG_H
#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>
#define SRSRAN_API
#define SRSRAN_SUCCESS 0
#define SRSRAN_ERROR (-1)
#define SRSRAN_ERROR_INVALID_INPUTS (-2)
#define SRSRAN_CEIL(NUM, DEN) (((NUM) + ((DEN)-1)) / (DEN))
#define SRSRAN_MAX(a, b) ((a) > (b) ? (a) : (b))
#define SRSRAN_MIN(a, b) ((a) < (b) ? (a) : (b))
typedef float _Complex cf_t;
#endif

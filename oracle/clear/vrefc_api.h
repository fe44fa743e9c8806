/* TEST INFRASTRUCTURE — NOT PRODUCT CODE.
 *
 * C API of `oracle/_ref/libvamp_ref_clear.so`: the reference's fkcc / validate_motion / validate_vector run with
 * instrumented collision primitives (oracle/clear/vrefc_instr.hh) so that, beside the reference's verdict, each query
 * returns the smallest |signed clearance| in metres over every test the reference executed for it.  `envf` is the
 * `collision::Environment<float>` of a vref_env (vref_env_float_ptr() of libvamp_ref.so). */
#pragma once
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#define VREFC_API __attribute__((visibility("default")))

VREFC_API int vrefc_fkcc_clearance(const char *robot, const void *envf, const float *cfg, size_t n, uint8_t *valid,
                                   float *clearance, int threads);
/* dist == NULL: (start, goal) edges = validate_motion; else (start, vector, distance) = validate_vector */
VREFC_API int vrefc_motion_clearance(const char *robot, const void *envf, const float *a, const float *second,
                                     const float *dist, size_t n, uint8_t *valid, float *clearance, int threads);

#ifdef __cplusplus
}
#endif

/* TEST INFRASTRUCTURE — NOT PRODUCT CODE.
 *
 * C API of `oracle/_ref/libvamp_ref.so`: the UNMODIFIED mr-vamp reference headers
 * (/root/reference/src/impl/vamp/...) compiled for the host CPU (AVX2, rake = 8) behind
 * plain-C entry points, so the tests, `__graft_entry__.smoke()` and `bench.py`'s CPU
 * baseline legs can run the reference itself.  Nothing under mr_vamp_b200/ links or
 * loads this library.  Built by oracle/Makefile from the sources where they lie; no
 * reference source is copied into this repository.
 *
 * Robots: "sphere", "ur5", "fetch", "baxter", "panda_0_0", "panda_1_0", "panda_2_2"
 * ("panda" is an alias of panda_2_2, the fork's default, robots/panda_grid.hh:39-41).
 */
#pragma once
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- numerics probes (vector/interface.hh:438-469, vector/avx.hh:411-415) ---- */
void vref_trig(const float *x, size_t n, float *sin_out, float *cos_out);
void vref_sqrt(const float *x, size_t n, float *out);

/* ---- environment (collision/environment.hh, bindings/environment.cc:107-163) ---- */
typedef struct vref_env vref_env;
vref_env *vref_env_create(void);
void vref_env_destroy(vref_env *);
void vref_env_add_sphere(vref_env *, float x, float y, float z, float r);
/* c[15] = center(3) axis1(3) axis2(3) axis3(3) half(3); classified z-aligned iff axis_3_z == 1 */
void vref_env_add_cuboid(vref_env *, const float *c);
/* c[8] = x1 y1 z1 xv yv zv r rdv; classified z-aligned iff xv == 0 && yv == 0 */
void vref_env_add_capsule(vref_env *, const float *c);
void vref_env_add_heightfield(vref_env *, const float *off_scale6, size_t xd, size_t yd, const float *data);
/* returns CAPT build nanoseconds */
double vref_env_add_pointcloud(vref_env *, const float *xyz, size_t n, float r_min, float r_max, float r_point);
/* tf[7] = tx ty tz qx qy qz qw ; spheres [n][4] = x y z r */
void vref_env_attach(vref_env *, const float *tf, const float *spheres, size_t n);
void vref_env_detach(vref_env *);

/* Export in the reference's sorted order (what a device environment is built from).
 * kind: 0 spheres [n][5]=x y z r min_d ; 1 capsules, 2 z_aligned_capsules [n][9]= x1 y1 z1 xv yv zv r rdv min_d ;
 *       3 cuboids, 4 z_aligned_cuboids [n][16] = 15 shape floats + min_d.
 * Returns the count; writes when out != NULL. */
size_t vref_env_export(const vref_env *, int kind, float *out);

/* CAPT buffers of pointcloud `idx` (collision/capt.hh:588-621). */
typedef struct {
    size_t n_tests, n_aff_starts, n_aff_vec, n_aabbs;
    int nlog2;
    float r_min, r_max, r_point;
    float aabb_top[6];
    const float *tests;
    const uint32_t *aff_starts;
    const float *aff_x, *aff_y, *aff_z; /* n_aff_vec * 8 floats each */
    const float *aabbs;                 /* [n_aabbs][6] lower(3) upper(3) */
} vref_capt_view;
int vref_env_capt(const vref_env *, size_t idx, vref_capt_view *out);
/* direct CAPT queries, collides_simd with the query in lane 0.. (capt.hh:457-541) / scalar collides (:403-444) */
void vref_capt_query(const vref_env *, size_t idx, const float *xyzr, size_t n, uint8_t *hit_simd, uint8_t *hit_scalar);

/* single-sphere environment query (collision/validity.hh:46-150), one sphere broadcast over the lanes */
void vref_sphere_env(const vref_env *, const float *xyzr, size_t n, uint8_t *hit);

/* ---- robots ---- */
int vref_robot_info(const char *robot, int *dim, int *n_spheres, int *resolution);
/* [n][dim] -> [n][n_spheres][4] */
int vref_sphere_fk(const char *robot, const float *cfg, size_t n, float *out);
/* per-configuration verdict: the config broadcast to all 8 lanes (what Helper::validate does,
 * bindings/common.hh:172-182, without the joint-limit test).  valid[i] = 1 iff fkcc true. */
int vref_fkcc(const char *robot, const vref_env *, const float *cfg, size_t n, uint8_t *valid, int threads);
/* reference-native throughput mode: 8 distinct configs per block, one verdict per block (n % 8 == 0). */
int vref_fkcc_packed(const char *robot, const vref_env *, const float *cfg, size_t n, uint8_t *valid_blocks,
                     int threads);
/* fkcc_attach on a broadcast block (requires an attachment in env). */
int vref_fkcc_attach(const char *robot, const vref_env *, const float *cfg, size_t n, uint8_t *valid, int threads);
/* planning::validate_motion<Robot, 8, Robot::resolution>(a, b, env) (planning/validate.hh:67-75) */
int vref_validate_motion(const char *robot, const vref_env *, const float *a, const float *b, size_t n,
                         uint8_t *valid, int threads);
/* collision::filter_pointcloud (collision/filter.hh:176-275) -> number of points kept (written to out when != NULL) */
size_t vref_filter_pointcloud(const float *xyz, size_t n, float min_dist, float max_range, const float *origin,
                              const float *ws_min, const float *ws_max, int cull, float *out);
/* Helper::sphere_validate (bindings/common.hh:155-170) of ONE configuration: per robot sphere the obstacles
 * sphere_environment_get_collisions reports, as their insertion indices (the order of the vref_env_add_* calls) */
int vref_sphere_validity(const char *robot, const vref_env *, const float *cfg, uint32_t cap, uint32_t *hits,
                         uint32_t *counts);
/* planning::validate_vector<Robot, 8, Robot::resolution>(start, vector, distance, env) (planning/validate.hh:23-65) */
int vref_validate_vector(const char *robot, const vref_env *, const float *a, const float *v, const float *dist,
                         size_t n, uint8_t *valid, int threads);
/* the collision::Environment<float> inside a vref_env (handed to libvamp_ref_clear.so, oracle/clear/vrefc_api.h) */
const void *vref_env_float_ptr(const vref_env *);
/* Helper::validate_configuration: joint-limit test && validate_motion<Robot,8,1>(q,q) */
int vref_validate(const char *robot, const vref_env *, const float *cfg, size_t n, uint8_t *valid);
/* scale_configuration: [0,1]^d -> joint ranges, in place */
int vref_scale(const char *robot, float *cfg, size_t n);
void vref_sphere_robot_set_radius(float r);

#ifdef __cplusplus
}
#endif

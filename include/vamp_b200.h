/* vamp_b200.h — C ABI of libvamp_b200.so: the B200 (sm_100a) implementation of mr-vamp's
 * data-parallel hot path, batched forward kinematics fused with collision checking (fkcc) and
 * discretised motion validation.
 *
 * This is the whole drop-in boundary.  Plain pointers and sizes only; every pointer is a HOST
 * pointer unless its name ends in `_d`; the caller owns every buffer it passes, the library owns
 * the handles it returns.  All functions return 0 on success or a non-zero status whose text is
 * available from vb_last_error() (thread-local).  There is no CPU fallback: when no CUDA device is
 * usable every query fails with VB_ERR_CUDA.
 *
 * What each entry point replaces in the reference (paths under src/impl/vamp/):
 *
 *   vb_env_add_*            collision::Environment<float> population as done by the bindings:
 *                           bindings/environment.cc:107-163 (z-aligned classification, sort()),
 *                           collision/shapes.hh:57-73,166-192,236-239 (min_distance),
 *                           collision/environment.hh:26-66 (per-call float->vector conversion: gone,
 *                           the device copy is rebuilt only by vb_env_commit)
 *   vb_env_add_capt         collision::CAPT query buffers, collision/capt.hh:588-621
 *   vb_env_add_pointcloud   Environment.add_pointcloud, bindings/environment.cc:148-158 -> collision::CAPT(points, r_min,
 *                           r_max, r_point), collision/capt.hh:137-398: the tree is BUILT by the library, on the device
 *   vb_env_set_attachment   Environment::attachments, collision/attachments.hh:11-73
 *   vb_sphere_fk            Robot::sphere_fk, robots/panda_base.hh:67-71 (+ ur5/fetch/baxter/sphere .hh),
 *                           bindings/common.hh:132-152 (Helper::fk)
 *   vb_fkcc                 Robot::fkcc / fkcc_attach, robots/panda_base.hh:53-66 ->
 *                           robots/<robot>/fk.hh interleaved_sphere_fk[_attachment]; one verdict per
 *                           configuration (= the reference called on a broadcast block,
 *                           bindings/common.hh:172-182)
 *   vb_sphere_validity      Helper::sphere_validate, bindings/common.hh:155-170 -> sphere_environment_get_collisions,
 *                           collision/validity.hh:152-249 (per robot sphere: the obstacles it overlaps)
 *   vb_validate             Helper::validate_configuration, bindings/common.hh:172-182 (joint limits + fkcc)
 *   vb_validate_motion      planning::validate_motion / validate_vector, planning/validate.hh:23-75,
 *                           and Path::validate, planning/plan.hh:155-168 (edges of a path in one call)
 *   vb_validate_vector      planning::validate_vector(start, vector, distance, env), planning/validate.hh:23-65 --
 *                           the form the planners' inner loops call (rrtc.hh:139,180; prm.hh; fcit.hh)
 *   vb_connect_increments   RRTC's connect loop, planning/rrtc.hh:172-194 (all n_extensions increments of many
 *                           attempts in one batch, first failing increment per attempt)
 *   vb_sample_valid         the sampling loops of planning/prm.hh:111-124 and planning/fcit.hh:316-343 (draw, Robot::fkcc,
 *                           keep the valid ones in stream order) for a whole batch: the first k valid of a stream
 *   vb_validate_motion_indexed  the neighbour edges of planning/prm.hh:139 / planning/fcit.hh:230-236 as index pairs
 *                           into a device-resident vertex table
 *   vb_validate_path        Path::validate, planning/plan.hh:155-168 (first invalid edge)
 *   vb_shortcut_path        planning::shortcut_path, planning/simplify.hh:115-140 (all candidates in one batch)
 *   vb_validate_pairs       the fork's planned MultiRobotEnvironment::check_inter_robot_collisions
 *                           (MR_DEV_PLAN.md:176-198; no reference implementation exists)
 */
#ifndef VAMP_B200_H
#define VAMP_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct vb_ctx vb_ctx; /* devices, streams, scratch, optional NCCL communicator */
typedef struct vb_env vb_env; /* host description + device-resident replicas of one environment */

typedef enum
{
    VB_SPHERE = 0,
    VB_UR5 = 1,
    VB_PANDA = 2,
    VB_FETCH = 3,
    VB_BAXTER = 4
} vb_robot;

enum
{
    VB_OK = 0,
    VB_ERR_ARG = 1,   /* bad argument */
    VB_ERR_CUDA = 2,  /* CUDA runtime error or no device */
    VB_ERR_NCCL = 3,  /* NCCL error */
    VB_ERR_STATE = 4  /* e.g. environment not committed */
};

const char *vb_last_error(void);
const char *vb_version(void);

/* ---- context ------------------------------------------------------------------------------ */
/* device_ids == NULL, n_devices == 0: device 0.  With n_devices > 1 the host-pointer queries shard
 * their batch over the devices in contiguous index ranges; environments are replicated. */
int vb_ctx_create(vb_ctx **out, const int *device_ids, int n_devices);
void vb_ctx_destroy(vb_ctx *);
int vb_ctx_device_count(const vb_ctx *);
int vb_ctx_sm_count(const vb_ctx *, int slot);
int vb_sync(vb_ctx *);
/* "domain:bus:device.function" of a CUDA device: lets a host place its threads and pinned staging memory on the
 * GPU's NUMA node (/sys/bus/pci/devices/<id>/numa_node, local_cpulist) -- mr_vamp_b200/hostbind.py does */
int vb_device_pci_bus_id(int device, char *out, int len);

/* ---- robots ------------------------------------------------------------------------------- */
int vb_robot_by_name(const char *name, vb_robot *out);
int vb_robot_info(vb_robot, int *dimension, int *n_spheres, int *resolution);
/* joint ranges of scale_configuration: q = lo + u * range (robots/<robot>/fk.hh) */
int vb_robot_limits(vb_robot, float *lo, float *range);
/* configuration `index` of the synthetic stream used by the *_random entry points */
int vb_random_config(vb_robot, uint64_t seed, uint64_t index, float *q);
/* edge `index` of the synthetic edge stream: a = config(2*index), b = a + len * dir clipped to limits */
int vb_random_edge(vb_robot, uint64_t seed, uint64_t index, float len_lo, float len_hi, float *a, float *b);
/* the same streams in batches: rows first .. first+n-1 into host buffers, or generated on the device */
int vb_random_configs(vb_robot, uint64_t seed, uint64_t first, size_t n, float *q /* [n][dim] */);
int vb_random_edges(vb_robot, uint64_t seed, uint64_t first, size_t n, float len_lo, float len_hi, float *a, float *b);
/* robots::sphere global radius (robots/sphere/fk.hh:16-18); default 0.2 */
int vb_ctx_set_sphere_radius(vb_ctx *, float r);

/* ---- environment -------------------------------------------------------------------------- */
int vb_env_create(vb_ctx *, vb_env **out);
void vb_env_destroy(vb_env *);
int vb_env_clear(vb_env *);
int vb_env_add_sphere(vb_env *, float x, float y, float z, float r);
/* c[15] = centre(3) axis1(3) axis2(3) axis3(3) half-extents(3); z-aligned iff axis_3_z == 1 */
int vb_env_add_cuboid(vb_env *, const float *c);
/* c[8] = x1 y1 z1 xv yv zv r rdv (rdv = 1/|v|^2); z-aligned iff xv == 0 && yv == 0 */
int vb_env_add_capsule(vb_env *, const float *c);
/* off_scale[6] = x y z xs ys zs (offsets, already-inverted scales as collision/factory.hh:365-386) */
int vb_env_add_heightfield(vb_env *, const float *off_scale, size_t xd, size_t yd, const float *data);
/* CAPT query buffers (collision/capt.hh:588-621) built by the host; aff_* hold n_aff_vec*8 floats */
int vb_env_add_capt(vb_env *, const float *tests, size_t n_tests, const uint32_t *aff_starts, const float *aff_x,
                    const float *aff_y, const float *aff_z, size_t n_aff_vec, const float *aabbs, size_t n_aabbs,
                    const float *aabb_top, float r_point, int nlog2);
/* a point cloud as raw points [n][3]: the CAPT (split values, per-leaf affordance lists, leaf boxes) is built on the
 * context's first device exactly as collision::CAPT's constructor defines it; *build_ms = device time of the build */
int vb_env_add_pointcloud(vb_env *, const float *xyz, size_t n, float r_min, float r_max, float r_point, double *build_ms);
/* the query buffers of point cloud `idx` as the library holds them (affordances interleaved per 8-point vector:
 * x[8] y[8] z[8]); counts always, data where the pointer is non-null */
int vb_env_capt_export(const vb_env *, size_t idx, size_t *n_tests, size_t *n_aff_vec, float *tests, uint32_t *aff_starts,
                       float *aff, float *aabbs, float *aabb_top, int *nlog2);
/* tf[7] = tx ty tz qx qy qz qw; spheres [n][4] = x y z r (end-effector frame); n <= 16 */
int vb_env_set_attachment(vb_env *, const float *tf, const float *spheres, size_t n);
int vb_env_clear_attachment(vb_env *);
/* upload / replicate to every device of the context; bumps the version */
int vb_env_commit(vb_env *);
/* introspection (sorted order): kind 0 spheres[5] 1 capsules[9] 2 z-capsules[9] 3 cuboids[16] 4 z-cuboids[16];
 * rows are the shape floats followed by min_distance.  Returns the count via *n; writes when out != NULL. */
int vb_env_export(const vb_env *, int kind, float *out, size_t *n);
/* insertion sequence number (0-based over all vb_env_add_sphere/cuboid/capsule/heightfield calls since the last
 * vb_env_clear) of every obstacle, in the library's object order: sorted spheres, capsules, z-capsules, cuboids,
 * z-cuboids, then height fields.  This is how a binding maps the object numbers of vb_sphere_validity back to the
 * names it keeps (the reference's shapes carry a std::string name, collision/shapes.hh). */
int vb_env_object_ids(const vb_env *, uint32_t *ids, size_t *n);

/* ---- queries, host buffers (sharded over the context's devices) --------------------------- */
/* configs: row-major [n][dimension] float32; base_xyz: robot base offset in metres (Panda grid
 * variants, robots/panda_grid.hh:10-41), NULL = origin. */
int vb_sphere_fk(vb_ctx *, vb_robot, const float *base_xyz, const float *configs, size_t n,
                 float *spheres_xyzr /* [n][n_spheres][4] */);
/* valid_bits: ceil(n/32) words, bit i%32 of word i/32 = 1 iff configuration i is collision free.
 * margin_or_null: [n] smallest |signed clearance| in metres over the tests that decided the verdict. */
int vb_fkcc(vb_ctx *, vb_robot, const float *base_xyz, const vb_env *, const float *configs, size_t n,
            uint32_t *valid_bits, float *margin_or_null);
/* Per robot sphere (Robot::sphere_fk order) the obstacles it overlaps: counts[i * n_spheres + k] = how many, and
 * the first min(count, cap) object numbers (vb_env_object_ids order) in hits[(i * n_spheres + k) * cap ..].  Point
 * clouds are not reported (neither does the reference: validity.hh:152-249 has no CAPT loop). */
int vb_sphere_validity(vb_ctx *, vb_robot, const float *base_xyz, const vb_env *, const float *configs, size_t n,
                       uint32_t cap, uint32_t *hits, uint32_t *counts);
/* joint-limit test && fkcc (Helper::validate_configuration) */
int vb_validate(vb_ctx *, vb_robot, const float *base_xyz, const vb_env *, const float *configs, size_t n,
                uint32_t *valid_bits);
/* resolution <= 0: the robot's own (Robot::resolution) */
int vb_validate_motion(vb_ctx *, vb_robot, const float *base_xyz, const vb_env *, const float *starts,
                       const float *goals, size_t n_edges, int resolution, uint32_t *valid_bits,
                       float *margin_or_null);
/* (start, vector, distance) form of an edge: validate_vector, planning/validate.hh:23-65.  The number of passes
 * comes from `distances` exactly as in the reference (n = max(ceil(distance / 8 * resolution), 1)); the interpolants
 * are start + vector * k/8 - i * vector / (8 n).  vb_validate_motion is the special case vector = goal - start,
 * distance = |vector|. */
int vb_validate_vector(vb_ctx *, vb_robot, const float *base_xyz, const vb_env *, const float *starts,
                       const float *vectors, const float *distances, size_t n_edges, int resolution,
                       uint32_t *valid_bits, float *margin_or_null);
/* RRTC connect loop (planning/rrtc.hh:172-194) for n attempts: attempt i walks from from[i] towards to[i]
 * (distances[i] = their distance as the caller's nearest-neighbour query returned it) in n_ext = ceil(distance /
 * range) increments, prior += increment accumulated in float as the reference does.  n_valid[i] = number of
 * increments validated before the first failure (the reference's i_extension at loop exit; == n_ext iff connected). */
int vb_connect_increments(vb_ctx *, vb_robot, const float *base_xyz, const vb_env *, const float *from, const float *to,
                          const float *distances, size_t n, float range, int resolution, uint32_t *n_valid,
                          uint32_t *n_ext_or_null);
/* PRM / FCIT* batching.
 * vb_sample_valid: walks a stream of candidate configurations -- `candidates` [m][dim] (host), or when NULL rows
 *   first .. first+m-1 of the library stream (seed) -- and returns its first k valid ones (Robot::fkcc) in stream order:
 *   out_cfgs [k][dim], out_index [k] = position in the stream, *n_found <= k, *n_consumed = candidates drawn up to and
 *   including the k-th valid one (all m when fewer than k are valid), i.e. where the reference's loop would stand.
 * vb_validate_motion_indexed: edge e runs from row pairs[2e] to row pairs[2e+1] of `table_d`, a DEVICE-resident
 *   [n_rows][dim] table (vb_dev_alloc / vb_dev_upload); 8 bytes per edge cross PCIe.  Out-of-range rows: edge invalid. */
int vb_sample_valid(vb_ctx *, vb_robot, const float *base_xyz, const vb_env *, const float *candidates, uint64_t seed,
                    uint64_t first, size_t m, size_t k, float *out_cfgs, uint64_t *out_index, size_t *n_found,
                    size_t *n_consumed);
int vb_validate_motion_indexed(vb_ctx *, vb_robot, const float *base_xyz, const vb_env *, const float *table_d, size_t n_rows,
                               const uint32_t *pairs, size_t n_edges, int resolution, uint32_t *valid_bits);
/* Batch edge callers.
 * vb_validate_path: Path::validate, planning/plan.hh:155-168 -- all consecutive edges of `waypoints` [n][dim] in one
 *   batch; *first_invalid_edge = index of the first invalid edge, -1 if the path is valid.
 * vb_shortcut_path: planning::shortcut_path, planning/simplify.hh:115-140 -- all (n-1)(n-2)/2 shortcut candidates in
 *   one batch, then the reference's scan on the verdicts; keep[i] = 1 iff waypoint i survives (same waypoints
 *   as the reference leaves in the path), *n_kept = how many. */
int vb_validate_path(vb_ctx *, vb_robot, const float *base_xyz, const vb_env *, const float *waypoints, size_t n,
                     int resolution, int64_t *first_invalid_edge);
int vb_shortcut_path(vb_ctx *, vb_robot, const float *base_xyz, const vb_env *, const float *waypoints, size_t n,
                     int resolution, uint8_t *keep, size_t *n_kept_or_null);
/* composite validity of two robots sharing one environment: fkcc(a) && fkcc(b) && no sphere of a
 * overlaps a sphere of b (world frame).  Two robots of one kind (the Panda grid) run as ONE fused kernel per chunk
 * (both chains into shared memory, both walks, then the robot-robot spheres where they lie); robots of different
 * kinds take three launches per chunk (fkcc twice, sphere_fk twice, a pair kernel). */
int vb_validate_pairs(vb_ctx *, vb_robot robot_a, const float *base_a, vb_robot robot_b, const float *base_b,
                      const vb_env *, const float *cfg_a, const float *cfg_b, size_t n, uint32_t *valid_bits);

/* the robot-robot term alone (the fork's planned check_inter_robot_collisions, MR_DEV_PLAN.md:176-198): bit = 1 iff no
 * world-frame sphere of robot a overlaps a sphere of robot b; environment and self collision are not looked at */
int vb_inter_robot(vb_ctx *, vb_robot robot_a, const float *base_a, vb_robot robot_b, const float *base_b,
                   const float *cfg_a, const float *cfg_b, size_t n, uint32_t *clear_bits);

/* ---- queries, device-resident (single device slot, asynchronous on that slot's stream) ---- */
int vb_dev_alloc(vb_ctx *, int slot, size_t bytes, void **ptr_d);
int vb_dev_free(vb_ctx *, int slot, void *ptr_d);
int vb_dev_upload(vb_ctx *, int slot, void *dst_d, const void *src, size_t bytes);
int vb_dev_download(vb_ctx *, int slot, void *dst, const void *src_d, size_t bytes);
int vb_fkcc_d(vb_ctx *, int slot, vb_robot, const float *base_xyz, const vb_env *, const float *configs_d, size_t n,
              uint32_t *valid_bits_d, float *margin_d_or_null);
/* as vb_fkcc_d, and writes per configuration the algorithmic flop count W of the reference algorithm
 * (FK chain + every test the reference-ordered walk executes; cost constants in DESIGN.md) */
int vb_fkcc_work_d(vb_ctx *, int slot, vb_robot, const float *base_xyz, const vb_env *, const float *configs_d,
                   size_t n, uint32_t *valid_bits_d, float *work_d);
int vb_validate_motion_d(vb_ctx *, int slot, vb_robot, const float *base_xyz, const vb_env *, const float *starts_d,
                         const float *goals_d, size_t n_edges, int resolution, uint32_t *valid_bits_d,
                         float *margin_d_or_null);
int vb_validate_vector_d(vb_ctx *, int slot, vb_robot, const float *base_xyz, const vb_env *, const float *starts_d,
                         const float *vectors_d, const float *distances_d, size_t n_edges, int resolution,
                         uint32_t *valid_bits_d, float *margin_d_or_null);
int vb_validate_motion_indexed_d(vb_ctx *, int slot, vb_robot, const float *base_xyz, const vb_env *, const float *table_d,
                                 size_t n_rows, const uint32_t *pairs_d, size_t n_edges, int resolution,
                                 uint32_t *valid_bits_d);
/* vb_validate_motion_noalloc_d: vb_validate_motion_d (verdicts only) for callers that cannot call cudaMalloc --
 * code under CUDA graph capture, tight serving loops, harnesses that forbid allocation.  It has NO counterpart in
 * the reference.  It runs the prefix-array form of the motion
 * kernels (one thread per interpolant), whose scratch memory (interpolant counts, their prefix, the scan's temporary
 * storage) is a file-scope `static __device__` array of
 * 2 * (VB_NOALLOC_MAX_EDGES + 64) * 4 + 65536 bytes (about 2.1 MB) that every GPU loading the library holds until
 * exit; the call allocates and frees nothing.  n_edges > VB_NOALLOC_MAX_EDGES (also returned by
 * vb_noalloc_max_edges()) is rejected with VB_ERR_ARG, an environment with an attachment with VB_ERR_STATE.  All
 * calls share the one array per device: calls that overlap on the GPU (other streams, other contexts, other host
 * threads) race on it -- serialise them.  Asynchronous on the slot's stream like the other _d entry points. */
#define VB_NOALLOC_MAX_EDGES 262144
size_t vb_noalloc_max_edges(void);
int vb_validate_motion_noalloc_d(vb_ctx *, int slot, vb_robot, const float *base_xyz, const vb_env *,
                                 const float *starts_d, const float *goals_d, size_t n_edges, int resolution,
                                 uint32_t *valid_bits_d);
/* composite configurations of two robots of ONE kind (different bases): one launch, verdict words on the device */
int vb_validate_pairs_d(vb_ctx *, int slot, vb_robot, const float *base_a, const float *base_b, const vb_env *,
                        const float *cfg_a_d, const float *cfg_b_d, size_t n, uint32_t *valid_bits_d);
/* configurations first .. first+n-1 of the synthetic stream, generated in registers (0 input bytes) */
int vb_fkcc_random_d(vb_ctx *, int slot, vb_robot, const float *base_xyz, const vb_env *, uint64_t seed,
                     uint64_t first, size_t n, uint32_t *valid_bits_d);
int vb_random_configs_d(vb_ctx *, int slot, vb_robot, uint64_t seed, uint64_t first, size_t n, float *q_d);
int vb_random_edges_d(vb_ctx *, int slot, vb_robot, uint64_t seed, uint64_t first, size_t n, float len_lo, float len_hi,
                      float *a_d, float *b_d);
/* page-locked host memory for the host-buffer queries (they run at PCIe speed only from pinned memory):
 * allocate it here, or pin an existing buffer once instead of paying the staging copy on every call */
int vb_host_alloc(size_t bytes, void **ptr);
int vb_host_free(void *ptr);
int vb_host_register(void *ptr, size_t bytes);
int vb_host_unregister(void *ptr);
/* CUDA-event timer on the slot's stream */
int vb_timer_start(vb_ctx *, int slot);
int vb_timer_stop(vb_ctx *, int slot, double *milliseconds);
/* number of kernels this context has launched so far */
uint64_t vb_launch_count(const vb_ctx *);
/* counts set bits of a device bitmask with n valid bits */
int vb_popcount_d(vb_ctx *, int slot, const uint32_t *bits_d, size_t n, uint64_t *count);

/* measured FP32 FMA throughput of the slot's device (roofline denominator of the fkcc kernels);
   best of 3 timed launches of 8 independent FMA chains per thread, in TFLOP/s */
int vb_fp32_peak_d(vb_ctx *, int slot, double *tflops);

/* ---- multi-process gather (one process per GPU, NCCL over NVLink) ------------------------- */
/* id_bytes: 128-byte ncclUniqueId created by rank 0 with vb_comm_unique_id and distributed by the caller */
int vb_comm_unique_id(void *id_bytes_128);
int vb_comm_init(vb_ctx *, int n_ranks, int rank, const void *id_bytes_128);
/* all ranks contribute n_words verdict words; every rank receives n_ranks*n_words in rank order */
int vb_comm_allgather_bits_d(vb_ctx *, const uint32_t *send_d, uint32_t *recv_d, size_t n_words);
int vb_comm_destroy(vb_ctx *);

#ifdef __cplusplus
}
#endif
#endif /* VAMP_B200_H */

// Per-robot launch table: the only interface between the C ABI (vb_capi.cu) and the kernel
// instantiations of one robot trace (vb_robot.cu, compiled once per robot).
#pragma once
#include "vb_kernels.cuh"

namespace vb
{
    struct RobotOps
    {
        const char *name;
        int dim, n_spheres, resolution, has_attach, n_slots;
        int chain_flops, env_groups, env_fine, self_groups, self_fine;
        float lo[16], range[16];
        cudaError_t (*fkcc)(const LaunchArgs &, int mode, int grid);  // mode: see fkcc_kernel
        cudaError_t (*fkcc_random)(const LaunchArgs &, size_t first, int grid);
        cudaError_t (*sphere_fk)(const LaunchArgs &);
        cudaError_t (*edge_count)(const LaunchArgs &, uint32_t *counts_d);
        cudaError_t (*motion)(const LaunchArgs &, bool margin, int grid);
        cudaError_t (*motion_rake)(const LaunchArgs &, bool margin, int grid);  // one rake block per octet, no scratch
        cudaError_t (*limits)(const LaunchArgs &);
        // composite configurations of two robots of this kind: CTAs of kBlock / 2 threads; *_blocks = resident CTAs per SM
        cudaError_t (*pairs_fused)(const LaunchArgs &, int grid);
        cudaError_t (*pairs_occupancy)(const LaunchArgs &, int *blocks_per_sm, int *threads);
        cudaError_t (*occupancy)(const LaunchArgs &, int *blocks_per_sm, int *threads);  // of the fkcc kernel
        void (*random_config)(uint64_t seed, uint64_t index, float *q);
    };

    const RobotOps *robot_ops_sphere();
    const RobotOps *robot_ops_ur5();
    const RobotOps *robot_ops_panda();
    const RobotOps *robot_ops_fetch();
    const RobotOps *robot_ops_baxter();
}  // namespace vb

// Kernel instantiations for ONE robot trace.  Compiled once per robot with
//   -DVB_TRACE_HEADER="<path to _gen/<robot>_trace.cuh>" -DVB_TRACE=<Robot>Trace -DVB_OPS=robot_ops_<robot>
#include VB_TRACE_HEADER
#include "vb_robot_ops.h"

namespace vb
{
    using R = VB_TRACE;

    static size_t smem_bytes(const LaunchArgs &a)
    {
        return a.env.n_quads > kMaxSharedQuads ? 0 : sizeof(float4) * static_cast<size_t>(a.env.n_quads);
    }

    static cudaError_t launch_fkcc(const LaunchArgs &a, bool margin, int grid)
    {
        if (margin)
        {
            fkcc_kernel<R, true><<<grid, kBlock, smem_bytes(a), a.stream>>>(a);
        }
        else
        {
            fkcc_kernel<R, false><<<grid, kBlock, smem_bytes(a), a.stream>>>(a);
        }
        return cudaGetLastError();
    }

    static cudaError_t launch_fkcc_random(const LaunchArgs &a, size_t first, int grid)
    {
        fkcc_random_kernel<R><<<grid, kBlock, smem_bytes(a), a.stream>>>(a, first);
        return cudaGetLastError();
    }

    static cudaError_t launch_sphere_fk(const LaunchArgs &a)
    {
        const size_t grid = (a.n + kBlock - 1) / kBlock;
        sphere_fk_kernel<R><<<static_cast<unsigned>(grid), kBlock, 0, a.stream>>>(a);
        return cudaGetLastError();
    }

    static cudaError_t launch_edge_count(const LaunchArgs &a, uint32_t *counts)
    {
        const size_t grid = (a.n + 255) / 256;
        edge_count_kernel<R><<<static_cast<unsigned>(grid), 256, 0, a.stream>>>(a, counts);
        return cudaGetLastError();
    }

    static cudaError_t launch_motion(const LaunchArgs &a, bool margin, int grid)
    {
        if (margin)
        {
            motion_kernel<R, true><<<grid, kBlock, smem_bytes(a), a.stream>>>(a);
        }
        else
        {
            motion_kernel<R, false><<<grid, kBlock, smem_bytes(a), a.stream>>>(a);
        }
        return cudaGetLastError();
    }

    static cudaError_t occupancy(int *blocks)
    {
        return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks, fkcc_kernel<R, false>, kBlock, 0);
    }

    static void host_random_config(uint64_t seed, uint64_t index, float *q)
    {
        random_config<R>(seed, index, q);
    }

    const RobotOps *VB_OPS()
    {
        static RobotOps ops = [] {
            RobotOps o{};
            o.name = R::kName;
            o.dim = R::kDim;
            o.n_spheres = R::kSpheres;
            o.resolution = R::kResolution;
            o.has_attach = R::kHasAttach;
            o.chain_flops = R::kChainFlops;
            o.env_groups = R::kEnvGroups;
            o.env_fine = R::kEnvFine;
            o.self_groups = R::kSelfGroups;
            o.self_fine = R::kSelfFine;
            for (int j = 0; j < R::kDim; ++j)
            {
                o.lo[j] = R::lo(j);
                o.range[j] = R::range(j);
            }
            o.fkcc = &launch_fkcc;
            o.fkcc_random = &launch_fkcc_random;
            o.sphere_fk = &launch_sphere_fk;
            o.edge_count = &launch_edge_count;
            o.motion = &launch_motion;
            o.occupancy = &occupancy;
            o.random_config = &host_random_config;
            return o;
        }();
        return &ops;
    }
}  // namespace vb

// Kernel instantiations for ONE robot trace.  Compiled once per robot with
//   -DVB_TRACE_HEADER="<path to _gen/<robot>_trace.cuh>" -DVB_TRACE=<Robot>Trace -DVB_OPS=robot_ops_<robot>
#include VB_TRACE_HEADER
#include "vb_robot_ops.h"

#include <atomic>

namespace vb
{
    using R = VB_TRACE;

    constexpr int kBlock = block_threads<R>();

    // one slot column per thread (+ the per-warp work queues of the warp walk)
    static size_t smem_bytes(const LaunchArgs &a, bool queues)
    {
        (void)a;
        const size_t slot_floats = (static_cast<size_t>(3 * R::kSlots * R::kStride) + 3) & ~static_cast<size_t>(3);
        return queues ? static_cast<size_t>(walk_smem_bytes<R>()) : sizeof(float) * slot_floats;
    }

    template <class K>
    static cudaError_t allow_smem(K kernel)
    {
        // static + dynamic shared memory of a CTA may not exceed 227 KB
        cudaFuncAttributes attr;
        if (cudaError_t e = cudaFuncGetAttributes(&attr, kernel); e != cudaSuccess)
        {
            return e;
        }
        return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    227 * 1024 - static_cast<int>(attr.sharedSizeBytes));
    }

    // cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device attribute of the kernel: it is raised once for every
    // device a launch is about to run on (the caller has made that device current), not once per process.
    static cudaError_t prepare()
    {
        constexpr int kMaxDevices = 64;
        static std::atomic<int> ready[kMaxDevices];
        int dev = 0;
        if (cudaError_t e = cudaGetDevice(&dev); e != cudaSuccess)
        {
            return e;
        }
        if (dev >= 0 && dev < kMaxDevices && ready[dev].load(std::memory_order_acquire))
        {
            return cudaSuccess;
        }
        cudaError_t e = allow_smem(fkcc_kernel<R, 0>);
        if (e == cudaSuccess) e = allow_smem(fkcc_kernel<R, 1>);
        if (e == cudaSuccess) e = allow_smem(fkcc_kernel<R, 2>);
        if (e == cudaSuccess) e = allow_smem(fkcc_random_kernel<R>);
        if (e == cudaSuccess) e = allow_smem(motion_kernel<R, false>);
        if (e == cudaSuccess) e = allow_smem(motion_rake_kernel<R, false>);
        if (e == cudaSuccess) e = allow_smem(motion_rake_kernel<R, true>);
        if (e == cudaSuccess) e = allow_smem(pairs_fused_kernel<R>);
        if (e == cudaSuccess && dev >= 0 && dev < kMaxDevices)
        {
            ready[dev].store(1, std::memory_order_release);
        }
        return e;
    }

    static cudaError_t launch_fkcc(const LaunchArgs &a, int mode, int grid)
    {
        if (cudaError_t e = prepare(); e != cudaSuccess)
        {
            return e;
        }
        if (mode == 1)
        {
            fkcc_kernel<R, 1><<<grid, kBlock, smem_bytes(a, false), a.stream>>>(a);
        }
        else if (mode == 2)
        {
            fkcc_kernel<R, 2><<<grid, kBlock, smem_bytes(a, false), a.stream>>>(a);
        }
        else
        {
            fkcc_kernel<R, 0><<<grid, kBlock, smem_bytes(a, true), a.stream>>>(a);
        }
        return cudaGetLastError();
    }

    static cudaError_t launch_fkcc_random(const LaunchArgs &a, size_t first, int grid)
    {
        if (cudaError_t e = prepare(); e != cudaSuccess)
        {
            return e;
        }
        fkcc_random_kernel<R><<<grid, kBlock, smem_bytes(a, true), a.stream>>>(a, first);
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

    static cudaError_t launch_limits(const LaunchArgs &a)
    {
        const size_t grid = ((a.n + 31) / 32 * 32 + 255) / 256;
        limits_kernel<R><<<static_cast<unsigned>(grid), 256, 0, a.stream>>>(a);
        return cudaGetLastError();
    }

    static cudaError_t launch_motion(const LaunchArgs &a, bool margin, int grid)
    {
        if (cudaError_t e = prepare(); e != cudaSuccess)
        {
            return e;
        }
        (void)margin;  // the prefix-array kernel only serves the allocation-free entry point (verdicts)
        motion_kernel<R, false><<<grid, kBlock, smem_bytes(a, true), a.stream>>>(a);
        return cudaGetLastError();
    }

    static cudaError_t launch_motion_rake(const LaunchArgs &a, bool margin, int grid)
    {
        if (cudaError_t e = prepare(); e != cudaSuccess)
        {
            return e;
        }
        if (margin)
        {
            motion_rake_kernel<R, true><<<grid, kBlock, smem_bytes(a, false), a.stream>>>(a);
        }
        else
        {
            motion_rake_kernel<R, false><<<grid, kBlock, smem_bytes(a, true), a.stream>>>(a);
        }
        return cudaGetLastError();
    }

    static cudaError_t launch_pairs_fused(const LaunchArgs &a, int grid)
    {
        if (cudaError_t e = prepare(); e != cudaSuccess)
        {
            return e;
        }
        pairs_fused_kernel<R><<<grid, 2 * pair_threads<R>(), smem_bytes(a, true), a.stream>>>(a);
        return cudaGetLastError();
    }

    static cudaError_t pairs_occupancy(const LaunchArgs &a, int *blocks, int *threads)
    {
        if (cudaError_t e = prepare(); e != cudaSuccess)
        {
            return e;
        }
        *threads = pair_threads<R>();  // composites per CTA step (the CTA has twice as many threads)
        return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks, pairs_fused_kernel<R>, 2 * pair_threads<R>(), smem_bytes(a, true));
    }

    static cudaError_t occupancy(const LaunchArgs &a, int *blocks, int *threads)
    {
        if (cudaError_t e = prepare(); e != cudaSuccess)
        {
            return e;
        }
        *threads = kBlock;
        return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks, fkcc_kernel<R, 0>, kBlock, smem_bytes(a, true));
    }

    static void host_random_config(uint64_t seed, uint64_t index, float *q)
    {
        random_config<R>(seed, index, q);
    }

#ifdef VB_PHASE_CLOCKS
    extern "C" int vb_debug_phase_clocks(unsigned long long *out16, int reset)
    {
        if (reset)
        {
            unsigned long long z[16] = {};
            return cudaMemcpyToSymbol(vb_phase_clocks, z, sizeof(z));
        }
        return cudaMemcpyFromSymbol(out16, vb_phase_clocks, 16 * sizeof(unsigned long long));
    }
#endif

#ifdef VB_STATS
    extern "C" int vb_debug_walk_stats(unsigned long long *out32, int reset)
    {
        if (reset)
        {
            unsigned long long z[32] = {};
            return cudaMemcpyToSymbol(vb_stats, z, sizeof(z));
        }
        return cudaMemcpyFromSymbol(out32, vb_stats, 32 * sizeof(unsigned long long));
    }
#endif

    const RobotOps *VB_OPS()
    {
        static RobotOps ops = [] {
            RobotOps o{};
            o.name = R::kName;
            o.dim = R::kDim;
            o.n_spheres = R::kSpheres;
            o.resolution = R::kResolution;
            o.has_attach = R::kHasAttach;
            o.n_slots = R::kSlots;
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
            o.motion_rake = &launch_motion_rake;
            o.limits = &launch_limits;
            o.pairs_fused = &launch_pairs_fused;
            o.pairs_occupancy = &pairs_occupancy;
            o.occupancy = &occupancy;
            o.random_config = &host_random_config;
            return o;
        }();
        return &ops;
    }
}  // namespace vb

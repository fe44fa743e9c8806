// Kernel templates of the fkcc path, instantiated once per robot trace (see codegen/emit_cuda.py).
//
// One thread evaluates one configuration.  Phase A runs the robot's FK chain as straight-line
// register code and parks every sphere centre the checks need in that thread's column of shared
// memory; phase B decides the collision tests of the 32 configurations of a warp together
// (vb_walk.cuh: uniform guard phase, then compacted work units); a warp ballot packs 32 verdicts
// into one uint32.  The lane-serial walk below (reference test order, one copy of the environment
// query) serves the margin / work-count outputs and attachments.  Replaces Robot::fkcc / Robot::sphere_fk (robots/panda_base.hh:53-71 and the
// per-robot equivalents) and planning::validate_motion (planning/validate.hh:23-75).
#pragma once
#include "vb_sinks.cuh"
#include "vb_tables.cuh"
#include "vb_walk.cuh"

namespace vb
{
    // the warps of a CTA are kept on one instruction stream with CTA barriers (the hot code is ~75 KB of mostly
    // straight-line SASS; seven independent streams per SM thrash the instruction caches)
#ifdef VB_NO_LOCKSTEP
    constexpr bool kLockstep = false;
#else
    constexpr bool kLockstep = true;
#endif
#ifdef VB_MID_BARRIER
    constexpr bool kMidBarrier = true;
#else
    constexpr bool kMidBarrier = false;
#endif
#if defined(VB_NO_STAGE) || defined(VB_SPLIT_BARRIER)
    constexpr bool kStageAllowed = false;
#else
    constexpr bool kStageAllowed = true;
#endif
#ifndef VB_LOCKSTEP_EVERY
#define VB_LOCKSTEP_EVERY 1
#endif
    constexpr unsigned kLockstepEvery = VB_LOCKSTEP_EVERY;
    // Phase-staggered warp groups (fkcc kernels): the CTA's warps form kGroups groups that take turns in the straight-line
    // ("dense": chain, guards, filters -- issue bound, ~100 KB of SASS streamed through the instruction caches) part of a
    // step; while one group is there the others are in their sparse, latency-bound unit stages.  A ring of named barriers
    // passes the dense section from group to group: group g waits on barrier 1 + g (bar.sync) before its chain and arrives
    // at barrier 1 + (g + 1) % kGroups (bar.arrive) after its filters.  One group at a time streams the big code, so the
    // instruction caches still serve a single stream, and the schedulers always have an issue-bound group to run.
#ifndef VB_GROUPS
#define VB_GROUPS 1
#endif
    constexpr int kGroupsWanted = VB_GROUPS;
    template <class Robot>
    __host__ __device__ constexpr int walk_groups()
    {
        // groups of equal size, at least 2 warps each
        return (kGroupsWanted > 1 && (Robot::kBlock / 32) % kGroupsWanted == 0 && (Robot::kBlock / 32) / kGroupsWanted >= 2) ? kGroupsWanted : 1;
    }
    // Every fkcc / motion kernel states its resident CTAs per SM (Robot::kCtasPerSm, fixed by the shared-memory
    // columns) as the second __launch_bounds__ argument: without it ptxas plans for 4 CTAs of kBlock threads,
    // caps the kernels at 64 registers and spills the in-flight grid cells (i.e. waits for every L2 load).
    struct LaunchArgs
    {
        EnvView env;
        float base[3];         // robot base offset in metres (Panda grid variants; 0 otherwise)
        float base_b[3];       // composite configurations: base offset of the second robot (cfg_b holds its configurations)
        const float *cfg;      // [n][dim] row-major, device
        const float *cfg_b;    // edges: goals [n][dim]
        size_t n;              // configurations or edges
        uint64_t seed;         // device-side input synthesis
        uint32_t *bits;        // verdict words, bit = 1 => valid
        float *margin;         // optional per-item min |clearance| (metres)
        float *spheres;        // sphere_fk output [n][S][4]
        unsigned long long *counters;  // optional work counters
        const uint32_t *prefix; // edges: exclusive prefix of interpolant counts [n+1]
        size_t n_points;        // edges: total interpolants
        float *att_scratch;    // posed attachment spheres [3 * kMaxAttach][gridDim * blockDim] (attachment runs)
        int resolution;
        int attach;            // use the attachment trace
        float sphere_radius;   // robots::sphere global radius (robots/sphere/fk.hh:16)
        float edge_len[2];     // synthetic edges: length range
        const uint32_t *edge_idx;  // edges as index pairs into ONE configuration table: start = cfg[idx[2e]], goal = cfg[idx[2e+1]]
        size_t table_rows;         // rows of that table (out-of-range indices make the edge invalid)
        const float *dist;     // edges in (start, vector, distance) form: distances [n] (validate_vector), else null
        int vector_form;       // cfg_b holds vectors (goal - start) instead of goals
        cudaStream_t stream;
    };

    // ---- geometry of one CTA ----------------------------------------------------------------------
    template <class Robot>
    constexpr int block_threads()
    {
        return Robot::kBlock;
    }

    // Shared memory is addressed by INDEX from the start of the dynamic window (see vb_walk.cuh: the compiler
    // then emits LDS/STS instead of generic accesses).
    struct Scratch
    {
        const float4 *prims;  // obstacle records (global memory, read through the non-coherent path)
        unsigned col;         // float index of this thread's slot column
        int stride;
        float *att_col;       // this thread's column of posed attachment spheres (global memory)
        int att_stride;
        unsigned queues;      // byte offset of this warp's work queues
    };

    // dynamic shared memory: [3 * kSlots rows of (blockDim + 1) floats: one column per thread][per-warp queues]
    template <class Robot>
    VB_D Scratch stage_env(const LaunchArgs &a, bool with_queues)
    {
        Scratch s;
        s.prims = a.env.prims;
        s.col = threadIdx.x;
        s.stride = Robot::kStride;  // odd: a warp reading ONE column at 32 different slots is conflict free too
        s.att_stride = static_cast<int>(gridDim.x * blockDim.x);
        s.att_col = a.att_scratch ? a.att_scratch + static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x : nullptr;
        // 16-byte aligned start of the queues
        const unsigned slot_floats = (3u * Robot::kSlots * Robot::kStride + 3u) & ~3u;
        s.queues = with_queues ? 4u * slot_floats + (threadIdx.x >> 5) * queue_bytes<Robot>() : 0u;
        if (with_queues)
        {
            WT<Robot>::stage();  // check tables of the warp walk into shared memory (when the generator found room)
            __syncthreads();
        }
        return s;
    }

    // ---- the table walk: every collision test of one configuration -------------------------------
    template <int kStride, int kConstRow, int kLiftRow = 0>
    struct TableCtx
    {
        const EnvView &env;
        Scratch sc;
        float bx, by, bz, radius;
        const DerivedRec *ders;  // centres rebuilt from their link's basis (vb_tables.cuh)
        int lift_z_row;          // Tables::lift_z_row of the trace being walked
        VB_D Center load(const SlotRef &ref) const
        {
            Center c;
            if (ref.slot & kDerivedSlot)
            {
                const DerivedRec d = ldg_record(ders + (ref.slot & (kDerivedSlot - 1u)));
                c.x = c.y = c.z = 0.f;
                constexpr unsigned kConstOff = 3u * kConstRow * kStride;
#pragma unroll
                for (int k = 0; k < 4; ++k)
                {
                    const unsigned off = d.off[k];
                    const float *col = shm_f32() + (sc.col & ~31u) + off + (off >= kConstOff ? 0u : (sc.col & 31u));
                    float z = col[2 * kStride];
                    if (kLiftRow > 0 && off >= 3u * kLiftRow * kStride)
                    {
                        z += shm_f32()[sc.col + static_cast<unsigned>(lift_z_row) * kStride];  // a translated constant as basis centre
                    }
                    c.x += d.w[k] * col[0];
                    c.y += d.w[k] * col[kStride];
                    c.z += d.w[k] * z;
                }
            }
            else
            {
                const float *col = shm_f32() + slot_index<kStride, kConstRow, kLiftRow>(sc.col & ~31u, sc.col & 31u, ref.slot);
                c.x = col[0];
                c.y = col[kStride];
                c.z = col[2 * kStride];
                if (kLiftRow > 0 && (ref.slot & kLiftSlot))
                {
                    c.z += shm_f32()[sc.col + static_cast<unsigned>(lift_z_row) * kStride];
                }
            }
            if (ref.base & 1)
            {
                c.x += bx;
            }
            if (ref.base & 2)
            {
                c.y += by;
            }
            if (ref.base & 4)
            {
                c.z += bz;
            }
            c.r = ref.r < 0.f ? radius : ref.r;
            return c;
        }
        // posed attachment sphere i (world frame, written by SlotSink::attach_pose)
        VB_D Center load_att(int i) const
        {
            Center c;
            c.x = sc.att_col[(3 * i + 0) * static_cast<size_t>(sc.att_stride)];
            c.y = sc.att_col[(3 * i + 1) * static_cast<size_t>(sc.att_stride)];
            c.z = sc.att_col[(3 * i + 2) * static_cast<size_t>(sc.att_stride)];
            c.r = env.attach.local[i].w;
            return c;
        }
    };

    // collision/validity.hh:13-44 + sphere_sphere.hh:9-23
    template <class M>
    VB_D bool pair_hit(const Center &a, const Center &b, M &mg)
    {
        const float dx = a.x - b.x, dy = a.y - b.y, dz = a.z - b.z, rs = a.r + b.r;
        const float v = dx * dx + dy * dy + dz * dz - rs * rs;
        mg.count(11);
        mg.sq(v, rs);
        return v < 0.f;
    }

    template <class T, class M, int kStride, int kConstRow, int kLiftRow>
    VB_D bool walk_tables(const T &t, const TableCtx<kStride, kConstRow, kLiftRow> &c, M &mg)
    {
        // environment groups (robots/<robot>/fk.hh: `if (sphere_environment_in_collision(bounding)) {...}`)
        for (int g = 0; g < t.n_env; ++g)
        {
            const RefGroup grp = t.env[g];
            const Center b = c.load(grp.guard);
            if (env_hit(c.env, c.sc.prims, b.x, b.y, b.z, b.r, mg))
            {
                if (grp.count == 0)
                {
                    return false;
                }
                for (int k = 0; k < grp.count; ++k)
                {
                    const Center s = c.load(t.refs[grp.begin + k]);
                    if (env_hit(c.env, c.sc.prims, s.x, s.y, s.z, s.r, mg))
                    {
                        return false;
                    }
                }
            }
        }
        // self-collision groups
        for (int g = 0; g < t.n_self; ++g)
        {
            const SelfGroup grp = t.self[g];
            if (pair_hit(c.load(grp.ga), c.load(grp.gb), mg))
            {
                if (grp.a_count == 0)
                {
                    return false;
                }
                for (int i = 0; i < grp.a_count; ++i)
                {
                    const Center a = c.load(t.refs[grp.a_begin + i]);
                    bool hit = false;
                    for (int j = 0; j < grp.b_count; ++j)
                    {
                        hit = pair_hit(a, c.load(t.refs[grp.b_begin + j]), mg) || hit;
                    }
                    if (hit)
                    {
                        return false;
                    }
                }
            }
        }
        // attachment vs robot (validity.hh:270-293) and attachment vs environment (validity.hh:251-268)
        const AttachmentView &att = c.env.attach;
        for (int g = 0; g < t.n_att; ++g)
        {
            const RefGroup grp = t.att[g];
            const Center b = c.load(grp.guard);
            bool hit = false;
            for (int i = 0; i < att.n; ++i)
            {
                hit = pair_hit(b, c.load_att(i), mg) || hit;
            }
            if (hit)
            {
                if (grp.count == 0)
                {
                    return false;
                }
                for (int k = 0; k < grp.count; ++k)
                {
                    const Center s = c.load(t.refs[grp.begin + k]);
                    for (int i = 0; i < att.n; ++i)
                    {
                        if (pair_hit(s, c.load_att(i), mg))
                        {
                            return false;
                        }
                    }
                }
            }
        }
        if (t.att_env)
        {
            for (int i = 0; i < att.n; ++i)
            {
                const Center s = c.load_att(i);
                if (env_hit(c.env, c.sc.prims, s.x, s.y, s.z, s.r, mg))
                {
                    return false;
                }
            }
        }
        return true;
    }

    template <class Robot, class M>
    VB_D bool check_config(const LaunchArgs &a, const Scratch &sc, const float *q, bool attach, M &mg)
    {
        SlotSink sink{sc.col, sc.stride, &a.env.attach, sc.att_col, sc.att_stride};
        if (Robot::kHasAttach && attach)
        {
            const TableCtx<Robot::kStride, Robot::kConstRow, Robot::kLiftRow> ctx{
                a.env, sc, a.base[0], a.base[1], a.base[2], a.sphere_radius, Robot::ders_attach(), Robot::tables_attach().lift_z_row};
            Robot::chain_attach(sink, q);
            return walk_tables(Robot::tables_attach(), ctx, mg);
        }
        const TableCtx<Robot::kStride, Robot::kConstRow, Robot::kLiftRow> ctx{
            a.env, sc, a.base[0], a.base[1], a.base[2], a.sphere_radius, Robot::ders(), Robot::tables().lift_z_row};
        Robot::chain(sink, q);
        return walk_tables(Robot::tables(), ctx, mg);
    }

    // Verdicts of 32 configurations, one per lane (all lanes must call; `alive` = lane has a configuration).
    // Attachment traces keep the lane-serial walk (rare, first rake pass only).
    // `base`: robot base offset (default: a.base; the second robot of a composite configuration passes a.base_b)
    template <class Robot, bool kSync = false>
    VB_D bool check_warp(const LaunchArgs &a, const Scratch &sc, const float *q, bool alive, bool attach, int &cstate,
                         bool octet = false, const float *base = nullptr, int release_id = 0, int release_count = 0)
    {
        const float *bs = base ? base : a.base;
        // kSync (CTA-uniform callers, `attach` uniform): the attachment decision must not depend on the warp
        if (Robot::kHasAttach && (kSync ? attach : __any_sync(kFull, attach && alive)))
        {
            if (kSync)
            {
                __syncthreads();  // the one mid-step barrier of Walk::run
            }
            bool ok = false;
            if (alive)
            {
                NoMargin mg;
                ok = check_config<Robot>(a, sc, q, attach, mg);
            }
            __syncwarp();
            if (release_id != 0)
            {
                asm volatile("bar.arrive %0, %1;" ::"r"(release_id), "r"(release_count) : "memory");
            }
            return ok;
        }
        VB_TICK_BEGIN;
        // joint-pair tables (codegen/pair_tables.py): the load is issued ahead of the chain, which hides its latency
        uint32_t pfree[Robot::kSelfWords];
        bool phit = false;
        if (Robot::kPairTables > 0 && alive)
        {
            phit = Robot::pair_lookup(q, pfree);
        }
        else
        {
#pragma unroll
            for (int wi = 0; wi < Robot::kSelfWords; ++wi)
            {
                pfree[wi] = 0u;
            }
        }
        if (alive)
        {
            SlotSink sink{sc.col, sc.stride, &a.env.attach, sc.att_col, sc.att_stride};
            VB_TICK(0);
            Robot::chain(sink, q);
        }
        __syncwarp();
        VB_TICK(1);
        WalkCtx wc;
        wc.wcol = sc.col - (threadIdx.x & 31);
        wc.queues = sc.queues;
        wc.prims = sc.prims;
        wc.bx = bs[0];
        wc.by = bs[1];
        wc.bz = bs[2];
        wc.radius = a.sphere_radius;
        wc.octet = octet;
        wc.release_id = release_id;
        wc.release_count = release_count;
        const bool ok = Walk<Robot>::template run<kSync>(a.env, wc, !alive || phit, cstate, pfree);
        __syncwarp();  // columns are rewritten by the next configuration
        return ok;
    }

    // ---- configuration staging: TMA bulk copy (global -> shared) one grid-stride step ahead ------------
    // With ~8 resident warps per SM nothing hides the HBM latency of the next configurations, and registers
    // loaded ahead would be spilled around the chain's call.  One elected thread therefore issues a 1-D bulk
    // copy (cp.async.bulk, completion on an mbarrier) of the CTA's next kBlock configurations (contiguous,
    // kBlock * 4 * dim bytes) right after the previous ones were consumed; a full walk later they are in
    // shared memory.
    VB_D uint32_t shm_addr(unsigned byte_offset)
    {
        extern __shared__ float4 vb_shared_window[];
        return static_cast<uint32_t>(__cvta_generic_to_shared(vb_shared_window)) + byte_offset;
    }
    VB_D void mbar_init(uint32_t mbar, uint32_t count)
    {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbar), "r"(count) : "memory");
    }
    VB_D void mbar_expect_tx(uint32_t mbar, uint32_t bytes)
    {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
    }
    VB_D void bulk_global_to_shared(uint32_t dst, const void *src, uint32_t bytes, uint32_t mbar)
    {
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                     "l"(src), "r"(bytes), "r"(mbar)
                     : "memory");
    }
    VB_D void mbar_wait(uint32_t mbar, uint32_t parity)
    {
        uint32_t ok;
        do
        {
            asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                         : "=r"(ok)
                         : "r"(mbar), "r"(parity)
                         : "memory");
        } while (!ok);
    }

    // ---- K1: batched fkcc over explicit configurations -----------------------------------------
    // kMode 0: verdicts (warp-cooperative walk); 1: verdicts + margin (lane-serial reference-ordered walk);
    // 2: verdicts + algorithmic flops of the checks (same walk), written to the margin buffer
    template <class Robot, int kMode>
    __global__ void __launch_bounds__(block_threads<Robot>(), Robot::kCtasPerSm) fkcc_kernel(const __grid_constant__ LaunchArgs a)
    {
        const Scratch prims = stage_env<Robot>(a, kMode == 0);
        int cstate = -1;
        const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
        const size_t n_round = (a.n + 31) / 32 * 32;
        // configuration staging (verdict mode): [slot columns][queues][kBlock x dim floats][mbarrier]
        constexpr unsigned kCfgBytes = static_cast<unsigned>(Robot::kBlock) * Robot::kDim * 4u;
        constexpr unsigned stage_off = cfg_stage_offset<Robot>();
        constexpr int kG = kMode == 0 ? walk_groups<Robot>() : 1;
        constexpr int kGroupThreads = Robot::kBlock / kG;
        const int group = static_cast<int>(threadIdx.x) / kGroupThreads;
        const bool staged = kG == 1 && kMode == 0 && kStageAllowed && Robot::kStageCfg && kCfgBytes % 16u == 0u && (reinterpret_cast<uintptr_t>(a.cfg) & 15u) == 0u;
        const uint32_t stage_buf = shm_addr(stage_off), stage_bar = stage_buf + kCfgBytes;
        uint32_t stage_parity = 0u;
        if (staged)
        {
            if (threadIdx.x == 0)
            {
                mbar_init(stage_bar, 1u);
                asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
                const size_t first = static_cast<size_t>(blockIdx.x) * blockDim.x;
                if (first + blockDim.x <= a.n)
                {
                    mbar_expect_tx(stage_bar, kCfgBytes);
                    bulk_global_to_shared(stage_buf, a.cfg + first * Robot::kDim, kCfgBytes, stage_bar);
                }
            }
            __syncthreads();
        }
#ifdef VB_SPLIT_BARRIER
        // experiment: warps may drift one grid-stride step apart (two alternating mbarriers in the staging area)
        if (kMode == 0)
        {
            if (threadIdx.x == 0)
            {
                mbar_init(stage_buf, blockDim.x / 32u);
                mbar_init(stage_buf + 8u, blockDim.x / 32u);
                asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            }
            __syncthreads();
        }
#endif
        // CTA-uniform trip count with a barrier per iteration: the warps of a CTA walk the (large, mostly
        // straight-line) code together, so the instruction cache serves one stream instead of seven
        unsigned iter = 0;
        for (size_t base = static_cast<size_t>(blockIdx.x) * blockDim.x; base < n_round; base += stride)
        {
            const size_t i = base + threadIdx.x;
            bool ok = false;
            const bool alive = i < a.n;
            float q[Robot::kDim];
            const bool full = staged && base + blockDim.x <= a.n;  // CTA-uniform: this step's configurations were staged
            if (full)
            {
                mbar_wait(stage_bar, stage_parity);
                stage_parity ^= 1u;
                const float *row = shm_f32() + stage_off / 4u + threadIdx.x * Robot::kDim;  // odd stride: conflict free
#pragma unroll
                for (int j = 0; j < Robot::kDim; ++j)
                {
                    q[j] = row[j];
                }
            }
            else
            {
#pragma unroll
                for (int j = 0; j < Robot::kDim; ++j)
                {
                    q[j] = alive ? __ldg(a.cfg + i * Robot::kDim + j) : 0.f;
                }
                if (!staged && i + stride < a.n)
                {
                    // pull the next step's configuration towards the SM while this one is processed
                    const float *nx = a.cfg + (i + stride) * Robot::kDim;
                    asm volatile("prefetch.global.L1 [%0];" ::"l"(nx));
                    asm volatile("prefetch.global.L1 [%0];" ::"l"(nx + Robot::kDim - 1));
                }
            }
#ifdef VB_SPLIT_BARRIER
            if (kMode == 0)
            {
                const unsigned k = iter++;
                if ((threadIdx.x & 31) == 0)
                {
                    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(stage_buf + 8u * (k & 1u)) : "memory");
                }
                if (k > 0)
                {
                    mbar_wait(stage_buf + 8u * ((k - 1u) & 1u), ((k - 1u) >> 1) & 1u);
                }
            }
            else
#endif
            if (full)
            {
                // The reads of the staging buffer above are generic-proxy accesses, the next bulk copy writes it through
                // the async proxy: without a proxy fence the copy may overtake shared-memory loads that are still in
                // flight when the CTA barrier completes (observed with 2 CTAs per SM: a warp then walked the NEXT
                // step's configurations).  Every reader fences before the barrier, the issuing thread after it.
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            }
            const bool first_step = base == static_cast<size_t>(blockIdx.x) * blockDim.x, last_step = base + stride >= n_round;
            if (kG > 1)
            {
                // wait for the dense section (the first group starts the ring)
                if (!(first_step && group == 0))
                {
                    asm volatile("bar.sync %0, %1;" ::"r"(1 + group), "r"(2 * kGroupThreads) : "memory");
                }
            }
            else if (staged || (kLockstep && (iter++ % kLockstepEvery) == 0))
            {
                __syncthreads();  // (staging: every thread has taken its configuration out of the buffer)
            }
            if (staged && threadIdx.x == 0 && base + stride + blockDim.x <= a.n)
            {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                mbar_expect_tx(stage_bar, kCfgBytes);
                bulk_global_to_shared(stage_buf, a.cfg + (base + stride) * Robot::kDim, kCfgBytes, stage_bar);
            }
            if constexpr (kMode == 1)
            {
                if (alive)
                {
                    MinMargin mg;
                    ok = check_config<Robot>(a, prims, q, a.attach != 0, mg);
                    a.margin[i] = mg.m;
                }
            }
            else if constexpr (kMode == 2)
            {
                if (alive)
                {
                    WorkCount mg;
                    ok = check_config<Robot>(a, prims, q, a.attach != 0, mg);
                    a.margin[i] = mg.m + static_cast<float>(Robot::kChainFlops);
                }
            }
            else
            {
                // (the last group's last release would find nobody waiting: skipped)
                const int rel = (kG > 1 && !(last_step && group == kG - 1)) ? 1 + (group + 1) % kG : 0;
                ok = check_warp<Robot, kMidBarrier>(a, prims, q, alive, a.attach != 0, cstate, false, nullptr, rel, 2 * kGroupThreads);
            }
            const uint32_t word = __ballot_sync(0xffffffffu, ok);
            if ((threadIdx.x & 31) == 0 && i < n_round)
            {
                a.bits[i >> 5] = word;
            }
        }
    }

    // configuration i of the synthetic stream: uniform in the joint ranges (scale_configuration)
    template <class Robot>
    VB_HD void random_config(uint64_t seed, uint64_t i, float *q)
    {
#pragma unroll
        for (int j = 0; j < Robot::kDim; ++j)
        {
            q[j] = VB_ADD(Robot::lo(j), VB_MUL(u01(seed, i, j), Robot::range(j)));  // no FMA: host and device agree bit for bit
        }
    }

    // ---- K1r: fkcc over device-synthesised configurations (0 input bytes) -----------------------
    template <class Robot>
    __global__ void __launch_bounds__(block_threads<Robot>(), Robot::kCtasPerSm) fkcc_random_kernel(const __grid_constant__ LaunchArgs a, size_t first)
    {
        const Scratch prims = stage_env<Robot>(a, true);
        int cstate = -1;
        const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
        const size_t n_round = (a.n + 31) / 32 * 32;
        // CTA-uniform trip count with a barrier per iteration: the warps of a CTA walk the (large, mostly
        // straight-line) code together, so the instruction cache serves one stream instead of seven
        unsigned iter = 0;
        for (size_t base = static_cast<size_t>(blockIdx.x) * blockDim.x; base < n_round; base += stride)
        {
            const size_t i = base + threadIdx.x;
            VB_TICK_BEGIN;
            constexpr int kG = walk_groups<Robot>();
            constexpr int kGroupThreads = Robot::kBlock / kG;
            const int group = static_cast<int>(threadIdx.x) / kGroupThreads;
            const bool first_step = base == static_cast<size_t>(blockIdx.x) * blockDim.x, last_step = base + stride >= n_round;
            if (kG > 1)
            {
                if (!(first_step && group == 0))
                {
                    asm volatile("bar.sync %0, %1;" ::"r"(1 + group), "r"(2 * kGroupThreads) : "memory");
                }
            }
            else if (kLockstep && (iter++ % kLockstepEvery) == 0)
            {
                __syncthreads();
            }
            VB_TICK(10);
            const bool alive = i < a.n;
            float q[Robot::kDim];
            random_config<Robot>(a.seed, first + i, q);
            VB_TICK(11);
            const int rel = (kG > 1 && !(last_step && group == kG - 1)) ? 1 + (group + 1) % kG : 0;
            const bool ok = check_warp<Robot, kMidBarrier>(a, prims, q, alive, a.attach != 0, cstate, false, nullptr, rel, 2 * kGroupThreads);
            VB_TICK(12);
            const uint32_t word = __ballot_sync(0xffffffffu, ok);
            if ((threadIdx.x & 31) == 0 && i < n_round)
            {
                a.bits[i >> 5] = word;
            }
        }
    }

    // ---- K2: sphere_fk ---------------------------------------------------------------------------
    template <class Robot>
    __global__ void __launch_bounds__(block_threads<Robot>(), Robot::kCtasPerSm) sphere_fk_kernel(const LaunchArgs a)
    {
        const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
        if (i >= a.n)
        {
            return;
        }
        float q[Robot::kDim];
#pragma unroll
        for (int j = 0; j < Robot::kDim; ++j)
        {
            q[j] = __ldg(a.cfg + i * Robot::kDim + j);
        }
        SphereSink pol{a.base[0], a.base[1], a.base[2], a.sphere_radius,
                                          a.spheres + i * Robot::kSpheres * 4};
        Robot::sphere_fk(pol, q);
    }

    // ---- joint-limit test of Helper::validate_configuration (bindings/common.hh:172-179) ----------
    // descale_configuration then all(0 <= q' <= 1); ANDed into the verdict words.
    template <class Robot>
    __global__ void limits_kernel(const LaunchArgs a)
    {
        const size_t n_round = (a.n + 31) / 32 * 32;
        const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
        if (i >= n_round)
        {
            return;
        }
        bool ok = false;
        if (i < a.n)
        {
            ok = true;
#pragma unroll
            for (int j = 0; j < Robot::kDim; ++j)
            {
                const float d = VB_ADD(__ldg(a.cfg + i * Robot::kDim + j), -Robot::lo(j));
                const float u = Robot::kDescaleDivides ? d / Robot::inv_range(j) : VB_MUL(d, Robot::inv_range(j));
                ok = ok && (u <= 1.f) && (u >= 0.f);
            }
        }
        const uint32_t word = __ballot_sync(0xffffffffu, ok);
        if ((threadIdx.x & 31) == 0)
        {
            a.bits[i >> 5] &= word;
        }
    }

    // ---- K3: discretised motion validation --------------------------------------------------------
    // planning/validate.hh:23-75: an edge a->b is valid iff fkcc holds at a + m/(8n) (b-a), m = 1..8n,
    // n = max(ceil(|b-a| / 8 * resolution), 1).  Pass i of the reference's rake checks lanes
    // k = 0..7 at start + v * (k+1)/8 - i * backstep with backstep = v / (8n) subtracted repeatedly;
    // the same recurrence is used here so interpolants are bit-identical.  Work is flattened: thread t
    // handles interpolant t of the whole batch (prefix[e] <= t < prefix[e+1] selects the edge), so a
    // long edge simply occupies more threads; verdicts are AND-reduced with atomicAnd.
    constexpr int kRake = 8;

    // number of rake passes as a float (not yet cast: the caller rejects values the cast cannot hold)
    template <int D>
    VB_HD float edge_passes(const float *qa, const float *qb, int resolution)
    {
        // squared_l2_norm (vector/interface.hh:407-410): rows are added first, then the 8 lanes are
        // folded as ((s0+s4)+(s2+s6)) + ((s1+s5)+(s3+s7)) (vector/avx.hh:442-452)
        float s[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        for (int j = 0; j < D; ++j)
        {
            const float v = VB_ADD(qb[j], -qa[j]);
            s[j & 7] = (j < 8) ? VB_MUL(v, v) : VB_ADD(s[j & 7], VB_MUL(v, v));
        }
        const float ss = VB_ADD(VB_ADD(VB_ADD(s[0], s[4]), VB_ADD(s[2], s[6])), VB_ADD(VB_ADD(s[1], s[5]), VB_ADD(s[3], s[7])));
        const float dist = sqrtf(ss);
        // (a NaN length stays NaN: fmaxf alone would turn it into one pass of NaN interpolants, which compare "free")
        return dist <= 3.0e38f ? fmaxf(ceilf(dist / static_cast<float>(kRake) * static_cast<float>(resolution)), 1.f) : dist;
    }

    // Prefix-array form of the motion kernels (only the allocation-free entry point still uses it: at most 2^18 edges
    // per launch).  An edge of kMaxPrefixPasses passes or more (> 255 rad at resolution 32; also NaN / Inf lengths)
    // would overflow the 32-bit interpolant prefix: it gets no interpolants and is reported invalid.  The verdict
    // words must have been preset before this kernel runs.
    constexpr float kMaxPrefixPasses = 1024.f;

    template <class Robot>
    __global__ void edge_count_kernel(const LaunchArgs a, uint32_t *counts)
    {
        const size_t e = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
        if (e >= a.n)
        {
            return;
        }
        float qa[Robot::kDim], qb[Robot::kDim];
        for (int j = 0; j < Robot::kDim; ++j)
        {
            qa[j] = a.cfg[e * Robot::kDim + j];
            qb[j] = a.cfg_b[e * Robot::kDim + j];
        }
        const float nf = edge_passes<Robot::kDim>(qa, qb, a.resolution);
        if (!(nf < kMaxPrefixPasses))
        {
            counts[e] = 0u;
            atomicAnd(a.bits + (e >> 5), ~(1u << (e & 31)));
            return;
        }
        counts[e] = kRake * static_cast<uint32_t>(nf);
    }

    template <class Robot>
    VB_D void edge_point(const float *qa, const float *qb, uint32_t n_pass, uint32_t pass, uint32_t lane, float *q)
    {
        const float pct = static_cast<float>(lane + 1) / static_cast<float>(kRake);
        const float denom = static_cast<float>(kRake * n_pass);
#pragma unroll
        for (int j = 0; j < Robot::kDim; ++j)
        {
            const float v = qb[j] - qa[j];
            float x = fmaf(v, pct, qa[j]);
            const float back = v / denom;
            for (uint32_t s = 0; s < pass; ++s)
            {
                x = VB_ADD(x, -back);
            }
            q[j] = x;
        }
    }

    template <class Robot, bool kMargin>
    __global__ void __launch_bounds__(block_threads<Robot>(), Robot::kCtasPerSm) motion_kernel(const __grid_constant__ LaunchArgs a)
    {
        const Scratch prims = stage_env<Robot>(a, !kMargin);
        int cstate = -1;
        const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
        const size_t n_points = __ldg(a.prefix + a.n);  // total interpolants, left on the device by the scan
        const size_t n_round = (n_points + 31) / 32 * 32;
        // CTA-uniform trip count with a barrier per step, as in the fkcc kernels: the warps of a CTA stream the
        // (large, straight-line) walk code together
        for (size_t base = static_cast<size_t>(blockIdx.x) * blockDim.x; base < n_round; base += stride)
        {
            const size_t t = base + threadIdx.x;
            if (kLockstep && !kMargin)
            {
                __syncthreads();
            }
            bool alive = t < n_points;
            size_t e = 0;
            uint32_t pass = 0;
            float q[Robot::kDim];
            if (alive)
            {
                // binary search: last edge with prefix[e] <= t
                size_t lo = 0, hi = a.n;
                while (hi - lo > 1)
                {
                    const size_t mid = (lo + hi) >> 1;
                    if (__ldg(a.prefix + mid) <= t)
                    {
                        lo = mid;
                    }
                    else
                    {
                        hi = mid;
                    }
                }
                e = lo;
                const uint32_t begin = __ldg(a.prefix + e);
                const uint32_t count = __ldg(a.prefix + e + 1) - begin;
                const uint32_t local = static_cast<uint32_t>(t - begin);
                pass = local / kRake;
                // edge already refuted by another interpolant: nothing left to decide (verdict kernels only)
                if (!kMargin && !((*reinterpret_cast<volatile const uint32_t *>(a.bits + (e >> 5)) >> (e & 31)) & 1u))
                {
                    alive = false;
                }
                else
                {
                    float qa[Robot::kDim], qb[Robot::kDim];
#pragma unroll
                    for (int j = 0; j < Robot::kDim; ++j)
                    {
                        qa[j] = __ldg(a.cfg + e * Robot::kDim + j);
                        qb[j] = __ldg(a.cfg_b + e * Robot::kDim + j);
                    }
                    edge_point<Robot>(qa, qb, count / kRake, pass, local % kRake, q);
                }
            }
            // the attachment trace is only used for the first pass (validate.hh:43-44 vs :58)
            const bool attach = a.attach && pass == 0;
            bool ok = true;
            if constexpr (kMargin)
            {
                if (alive)
                {
                    MinMargin mg;
                    ok = check_config<Robot>(a, prims, q, attach, mg);
                    atomicMin(reinterpret_cast<int *>(a.margin + e), __float_as_int(mg.m));
                }
            }
            else
            {
                ok = check_warp<Robot>(a, prims, q, alive, attach, cstate, true) || !alive;
            }
            if (alive && !ok)
            {
                atomicAnd(a.bits + (e >> 5), ~(1u << (e & 31)));
            }
        }
    }
    // ---- K3r: motion validation, one rake block per octet, passes in the reference's order ---------------
    // planning/validate.hh:23-65 (validate_vector).  The 8 lanes of an octet are the 8 interpolants of one rake
    // block of ONE edge; the octet walks that edge's passes one after the other exactly as the reference's loop
    // does (block = start + vector * percents, then block -= backstep per pass; same recurrence, bit-identical
    // interpolants), stops at the first colliding pass and then takes the next edge of its CTA's range from a
    // shared-memory counter.  Refuted edges therefore never run their later passes, no lane waits for an edge
    // that is longer than its neighbours', and there is neither a prefix array nor a search: one launch, no
    // scratch memory.  a.cfg = starts; a.cfg_b = goals, or vectors when a.vector_form (then a.dist = distances).
    // Edges whose length is not finite (or beyond 2^31 passes) are reported invalid; the reference's cast is
    // undefined there.
    template <int D>
    VB_HD float edge_norm(const float *v)
    {
        // squared_l2_norm (vector/interface.hh:407-410): rows are added first, then the 8 lanes are
        // folded as ((s0+s4)+(s2+s6)) + ((s1+s5)+(s3+s7)) (vector/avx.hh:442-452)
        float s[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        for (int j = 0; j < D; ++j)
        {
            s[j & 7] = (j < 8) ? VB_MUL(v[j], v[j]) : VB_ADD(s[j & 7], VB_MUL(v[j], v[j]));
        }
        const float ss = VB_ADD(VB_ADD(VB_ADD(s[0], s[4]), VB_ADD(s[2], s[6])), VB_ADD(VB_ADD(s[1], s[5]), VB_ADD(s[3], s[7])));
        return sqrtf(ss);
    }

    // kMargin: the lane-serial reference-ordered walk, min |clearance| over the executed interpolants per edge
    template <class Robot, bool kMargin>
    __global__ void __launch_bounds__(block_threads<Robot>(), Robot::kCtasPerSm) motion_rake_kernel(const __grid_constant__ LaunchArgs a)
    {
        __shared__ unsigned next_edge;  // edges of this CTA's range handed out so far
        const Scratch prims = stage_env<Robot>(a, !kMargin);
        int cstate = -1;
        const int lane = threadIdx.x & 31;
        const unsigned oct_bit = 1u << (lane & ~7);
        // contiguous range of edges per CTA, boundaries on multiples of 32: the CTA owns its verdict words
        const size_t per = ((a.n + gridDim.x - 1) / gridDim.x + 31) / 32 * 32;
        const size_t e_lo = min(a.n, static_cast<size_t>(blockIdx.x) * per), e_hi = min(a.n, e_lo + per);
        const unsigned count = static_cast<unsigned>(e_hi - e_lo);
        for (size_t w = e_lo / 32 + threadIdx.x; w * 32 < e_hi; w += blockDim.x)
        {
            const size_t rem = a.n - w * 32;
            a.bits[w] = rem >= 32 ? 0xffffffffu : ((1u << rem) - 1u);
        }
        if constexpr (kMargin)
        {
            for (size_t i = e_lo + threadIdx.x; i < e_hi; i += blockDim.x)
            {
                a.margin[i] = 3.0e38f;
            }
        }
        if (threadIdx.x == 0)
        {
            next_edge = 0u;
        }
        __syncthreads();
        constexpr unsigned kNone = 0xffffffffu;
        unsigned e = kNone;  // octet-uniform: edge (offset in the CTA's range) this octet is walking
        uint32_t pass = 0, n_pass = 1;
        float x[Robot::kDim], back[Robot::kDim];
#pragma unroll
        for (int j = 0; j < Robot::kDim; ++j)
        {
            x[j] = back[j] = 0.f;
        }
        bool more = count > 0u;  // warp-uniform: the range may still hold edges nobody has taken
        const bool free_running = !kMargin && a.env.capt_heavy != 0 && !(Robot::kHasAttach && a.attach);
        const float pct = static_cast<float>((lane & 7) + 1) / static_cast<float>(kRake);
        for (;;)
        {
            const unsigned want = __ballot_sync(0xffffffffu, e == kNone) & 0x01010101u;
            if (more && want)
            {
                unsigned first = 0u;
                if (lane == 0)
                {
                    first = atomicAdd(&next_edge, static_cast<unsigned>(__popc(want)));
                }
                first = __shfl_sync(0xffffffffu, first, 0);
                more = first + static_cast<unsigned>(__popc(want)) < count;
                const unsigned mine = first + static_cast<unsigned>(__popc(want & (oct_bit - 1u)));
                if (e == kNone && mine < count)
                {
                    const size_t g = e_lo + mine;
                    // rows of the two endpoints: edge g of two arrays, or an index pair into one table (PRM / FCIT*
                    // neighbour edges, planning/prm.hh:139: 8 bytes per edge instead of two configurations)
                    size_t row_a = g, row_b = g;
                    bool in_range = true;
                    if (a.edge_idx)
                    {
                        row_a = __ldg(a.edge_idx + 2 * g);
                        row_b = __ldg(a.edge_idx + 2 * g + 1);
                        in_range = row_a < a.table_rows && row_b < a.table_rows;
                        row_a = in_range ? row_a : 0;
                        row_b = in_range ? row_b : 0;
                    }
                    float qa[Robot::kDim], v[Robot::kDim];
#pragma unroll
                    for (int j = 0; j < Robot::kDim; ++j)
                    {
                        qa[j] = __ldg(a.cfg + row_a * Robot::kDim + j);
                        const float second = __ldg(a.cfg_b + row_b * Robot::kDim + j);
                        v[j] = a.vector_form ? second : VB_ADD(second, -qa[j]);
                    }
                    const float dist = a.dist ? __ldg(a.dist + g) : edge_norm<Robot::kDim>(v);
                    const float nf = fmaxf(ceilf(dist / static_cast<float>(kRake) * static_cast<float>(a.resolution)), 1.f);
                    // (NaN / Inf lengths: the reference's pass count is undefined for them; the edge is invalid)
                    if (!(dist <= 3.0e38f) || !(nf < 2147483648.f) || !in_range)
                    {
                        if ((lane & 7) == 0)
                        {
                            atomicAnd(a.bits + (g >> 5), ~(1u << (g & 31)));
                        }
                    }
                    else
                    {
                        e = mine;
                        pass = 0u;
                        n_pass = static_cast<uint32_t>(nf);
                        const float denom = static_cast<float>(kRake * n_pass);
#pragma unroll
                        for (int j = 0; j < Robot::kDim; ++j)
                        {
                            x[j] = fmaf(v[j], pct, qa[j]);
                            back[j] = v[j] / denom;
                        }
                    }
                }
            }
            const bool alive = e != kNone;
            // CTA-uniform exit, and the barrier that keeps the warps of the CTA on one instruction stream.  With a dense
            // point cloud (long leaf scans: EnvView::capt_heavy) the step time of a warp is dominated by how many scans
            // its lanes draw, not by instruction fetch: waiting for the slowest warp every step costs more than it saves
            // (ncu: 15 warps stalled at the barrier per issue on the 100k-point cloud; 3.2e7 -> 3.8e7 edges/s without it,
            // while the reference-filtered 2 220-point cloud loses 15 %), so the warps run free there -- they share nothing
            // but the edge counter.
            if (free_running)
            {
                if (!__any_sync(0xffffffffu, alive || more))
                {
                    break;
                }
            }
            else if (!__syncthreads_or(alive || more))
            {
                break;
            }
            // the attachment trace is only used for the first pass (validate.hh:43-44 vs :58)
            const bool attach = a.attach && pass == 0u;
            bool ok = true;
            if constexpr (kMargin)
            {
                if (alive)
                {
                    MinMargin mg;
                    ok = check_config<Robot>(a, prims, x, attach, mg);
                    atomicMin(reinterpret_cast<int *>(a.margin + e_lo + e), __float_as_int(mg.m));
                }
                __syncwarp();
            }
            else
            {
                ok = check_warp<Robot>(a, prims, x, alive, attach, cstate, true) || !alive;
            }
            const unsigned okm = __ballot_sync(0xffffffffu, ok);
            if (alive)
            {
                if (((okm >> (lane & ~7)) & 0xffu) != 0xffu)
                {
                    if ((lane & 7) == 0)
                    {
                        const size_t g = e_lo + e;
                        atomicAnd(a.bits + (g >> 5), ~(1u << (g & 31)));
                    }
                    e = kNone;
                }
                else if (++pass == n_pass)
                {
                    e = kNone;
                }
                else
                {
#pragma unroll
                    for (int j = 0; j < Robot::kDim; ++j)
                    {
                        x[j] = VB_ADD(x[j], -back[j]);
                    }
                }
            }
        }
    }
    // ---- K4: composite configurations of two robots of the same kind (mr-vamp's Panda grid) ---------------
    // The fork's planned MultiRobotEnvironment::check_inter_robot_collisions (MR_DEV_PLAN.md:176-198; no reference
    // implementation exists: PARITY UNPINNED): a composite configuration (q_a, q_b) is valid iff robot A at base_a
    // and robot B at base_b are each valid (Robot::fkcc: environment + self collision) and no world-frame sphere of
    // A overlaps a sphere of B (collision/sphere_sphere.hh:9-23).  ONE launch: the CTA runs with half its usual
    // threads, thread t owns slot column t for A's chain and column t + kBlock/2 for B's, both walks run as in the
    // fkcc kernels, and the robot-robot test reads both sets of centres where they already are -- no sphere buffer
    // in HBM, no host round trip.  Robot-robot test of one lane: B's axis-aligned box from its spheres, every sphere
    // of A outside that box is done, the others meet every sphere of B.
    template <class Robot>
    VB_D bool robots_overlap(const WalkCtx &ca, const WalkCtx &cb, int lane)
    {
        using W = Walk<Robot>;
        using T = WT<Robot>;
        float lo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, hi[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
        for (int g = 0; g < Robot::kEnvGroups; ++g)
        {
            const RefGroup G = T::env(g);
            for (int k = 0; k < G.ucount; ++k)
            {
                const Center s = W::centre(cb, lane, T::ref(G.ubegin + k));
                lo[0] = fminf(lo[0], s.x - s.r), hi[0] = fmaxf(hi[0], s.x + s.r);
                lo[1] = fminf(lo[1], s.y - s.r), hi[1] = fmaxf(hi[1], s.y + s.r);
                lo[2] = fminf(lo[2], s.z - s.r), hi[2] = fmaxf(hi[2], s.z + s.r);
            }
        }
        bool hit = false;
        for (int g = 0; g < Robot::kEnvGroups && !hit; ++g)
        {
            const RefGroup G = T::env(g);
            for (int k = 0; k < G.ucount && !hit; ++k)
            {
                const Center a = W::centre(ca, lane, T::ref(G.ubegin + k));
                if (a.x + a.r < lo[0] || a.x - a.r > hi[0] || a.y + a.r < lo[1] || a.y - a.r > hi[1] || a.z + a.r < lo[2] ||
                    a.z - a.r > hi[2])
                {
                    continue;
                }
                for (int g2 = 0; g2 < Robot::kEnvGroups; ++g2)
                {
                    const RefGroup G2 = T::env(g2);
                    for (int k2 = 0; k2 < G2.ucount; ++k2)
                    {
                        const Center b = W::centre(cb, lane, T::ref(G2.ubegin + k2));
                        const float dx = a.x - b.x, dy = a.y - b.y, dz = a.z - b.z, rs = a.r + b.r;
                        hit = hit || (dx * dx + dy * dy + dz * dz - rs * rs < 0.f);
                    }
                }
            }
        }
        return hit;
    }

    // composites per step of a composite CTA: half the warps of the robot's fkcc CTA (rounded down) walk robot A, the
    // other half robot B of the same composites
    template <class Robot>
    __host__ __device__ constexpr int pair_threads()
    {
        return (Robot::kBlock / 64) * 32;
    }

    // Composite configurations (q_a, q_b) of two robots of one type at two bases.  Warps [0, W/2) walk robot A
    // of 32 composites each, warp w + W/2 walks robot B of warp w's composites AT THE SAME TIME (all W warps resident, as
    // in the fkcc kernels); after a CTA barrier the A warp combines the two verdicts and, for the composites whose robots
    // are both valid on their own, tests A's spheres against B's (both columns are in shared memory).
    template <class Robot>
    __global__ void __launch_bounds__(2 * pair_threads<Robot>(), Robot::kCtasPerSm) pairs_fused_kernel(const __grid_constant__ LaunchArgs a)
    {
        constexpr unsigned kHalf = static_cast<unsigned>(pair_threads<Robot>());
        __shared__ uint32_t b_words[kHalf / 32u];
        const Scratch sc = stage_env<Robot>(a, true);
        const bool is_b = threadIdx.x >= kHalf;  // warp-uniform
        const unsigned slot = threadIdx.x - (is_b ? kHalf : 0u);
        int cstate = -1;
        const int lane = threadIdx.x & 31;
        const size_t stride = static_cast<size_t>(gridDim.x) * kHalf;
        const size_t n_round = (a.n + 31) / 32 * 32;
        for (size_t base = static_cast<size_t>(blockIdx.x) * kHalf; base < n_round; base += stride)
        {
            const size_t i = base + slot;
            const bool alive = i < a.n;
            __syncthreads();  // lockstep; and the previous step's b_words and columns have been consumed
            float q[Robot::kDim];
            const float *src = is_b ? a.cfg_b : a.cfg;
#pragma unroll
            for (int j = 0; j < Robot::kDim; ++j)
            {
                q[j] = alive ? __ldg(src + i * Robot::kDim + j) : 0.f;
            }
            bool ok = check_warp<Robot>(a, sc, q, alive, false, cstate, false, is_b ? a.base_b : a.base);
            const uint32_t mine = __ballot_sync(0xffffffffu, ok && alive);
            if (is_b && lane == 0)
            {
                b_words[slot >> 5] = mine;
            }
            __syncthreads();
            if (!is_b)
            {
                ok = ok && alive && ((b_words[slot >> 5] >> lane) & 1u);
                if (ok)
                {
                    WalkCtx ca, cb;
                    ca.wcol = sc.col - lane, cb.wcol = sc.col - lane + kHalf;
                    ca.queues = sc.queues, cb.queues = sc.queues + (kHalf / 32u) * queue_bytes<Robot>();
                    ca.prims = cb.prims = sc.prims;
                    ca.bx = a.base[0], ca.by = a.base[1], ca.bz = a.base[2];
                    cb.bx = a.base_b[0], cb.by = a.base_b[1], cb.bz = a.base_b[2];
                    ca.radius = cb.radius = a.sphere_radius;
                    ca.octet = cb.octet = false;
                    ca.release_id = cb.release_id = 0;
                    ca.release_count = cb.release_count = 0;
                    ok = !robots_overlap<Robot>(ca, cb, lane);
                }
                __syncwarp();
                const uint32_t word = __ballot_sync(0xffffffffu, ok);
                if (lane == 0 && i < n_round)
                {
                    a.bits[i >> 5] = word;
                }
            }
        }
    }
}  // namespace vb

// Warp-level collision walk of 32 configurations (one per lane): verdicts only.
//
// The reference returns at the first colliding sphere of an 8-lane block (robots/<robot>/fk.hh); a
// literal SIMT port of that control flow leaves most lanes idle, because the lanes of a warp
// disagree on which groups fire.  Here the walk is split into a uniform phase and compacted work:
//
//   guard phase   straight-line code generated per robot (codegen/emit_cuda.py: self_guards,
//                 env_guards): every lane tests ITS configuration's guard pairs (self collision) and
//                 looks its guard spheres up in the distance grid (environment, vb_sdf.cuh).  No
//                 lane-dependent branch, every load independent of every other.
//   work items    every (configuration, group) whose guard fired becomes a 16-bit item in a per-warp
//                 shared-memory queue: lane | group << 5, appended group by group with a ballot.  The
//                 32 lanes of the warp then take one item each, whatever configuration it belongs
//                 to, and read that configuration's sphere centres from its shared-memory column.
//   stages        self collision: filter (every a_i of the item against the inflated bounding sphere of
//                 the other link) -> pair units (a_i against every b_j); environment: every fine
//                 sphere of the item against the distance grid -> exact primitive arithmetic for the
//                 few spheres the grid cannot decide, including the exact guard test the reference's
//                 `if (guard) {...}` nesting asks for.
//
// The verdict is the same boolean function as the reference's (valid iff no group has a guard hit
// and a fine hit, collision/validity.hh:13-150 + robots/<robot>/fk.hh); only the order of evaluation
// differs.  Collided lanes are recorded in a per-warp bit mask.
#pragma once
#include "vb_sdf.cuh"
#include "vb_tables.cuh"

namespace vb
{
    constexpr unsigned kFull = 0xffffffffu;
    // pair stage, reduced b-links: |d^2 - (ra + rb)^2| below this (m^2) is decided by the direct formula
    constexpr float kGramBand = 4e-5f;

    // The stages of the walk are inlined into the kernel: as separate functions every call spills the caller's
    // live state to local memory (ABI), ~170 local loads/stores per warp and iteration that also evict the
    // tables from the small L1 left beside 224 KB of shared memory.
#ifndef VB_STAGE
#define VB_STAGE __device__ __forceinline__
#endif

    // development aid (-DVB_PHASE_CLOCKS): cycles per phase of the walk, summed over warps (lane 0's clock)
#ifdef VB_PHASE_CLOCKS
    __device__ unsigned long long vb_phase_clocks[16];
#define VB_TICK_BEGIN long long vb_last_ = clock64()
#define VB_TICK(k)                                                                               \
    do                                                                                           \
    {                                                                                            \
        const long long now_ = clock64();                                                        \
        if ((threadIdx.x & 31) == 0)                                                             \
        {                                                                                        \
            atomicAdd(&vb_phase_clocks[k], static_cast<unsigned long long>(now_ - vb_last_));    \
        }                                                                                        \
        vb_last_ = clock64();                                                                    \
    } while (0)
#else
#define VB_TICK_BEGIN
#define VB_TICK(k)
#endif

    // development aid (-DVB_STATS): work counters of the walk (scripts/walk_stats.py); warp sums / maxima per stage pass
#ifdef VB_STATS
    __device__ unsigned long long vb_stats[32];
#define VB_STAT_ADD(k, v)                                                                          \
    do                                                                                             \
    {                                                                                              \
        const unsigned s_ = __reduce_add_sync(0xffffffffu, static_cast<unsigned>(v));              \
        if ((threadIdx.x & 31) == 0)                                                               \
        {                                                                                          \
            atomicAdd(&vb_stats[k], static_cast<unsigned long long>(s_));                          \
        }                                                                                          \
    } while (0)
#define VB_STAT_MAX(k, v)                                                                          \
    do                                                                                             \
    {                                                                                              \
        const unsigned s_ = __reduce_max_sync(0xffffffffu, static_cast<unsigned>(v));              \
        if ((threadIdx.x & 31) == 0)                                                               \
        {                                                                                          \
            atomicAdd(&vb_stats[k], static_cast<unsigned long long>(s_));                          \
        }                                                                                          \
    } while (0)
#define VB_STAT_ONE(k, v)                                                                          \
    do                                                                                             \
    {                                                                                              \
        if ((threadIdx.x & 31) == 0)                                                               \
        {                                                                                          \
            atomicAdd(&vb_stats[k], static_cast<unsigned long long>(v));                           \
        }                                                                                          \
    } while (0)
#else
#define VB_STAT_ADD(k, v)
#define VB_STAT_MAX(k, v)
#define VB_STAT_ONE(k, v)
#endif

    // The walk addresses its shared memory as INDICES from the start of the dynamic shared-memory window
    // (not as pointers carried through structs): the compiler then knows the address space and emits
    // LDS/STS/ATOMS instead of generic loads that resolve the window at run time.
    VB_D float *shm_f32()
    {
        extern __shared__ float4 vb_shared_window[];
        return reinterpret_cast<float *>(vb_shared_window);
    }
    VB_D uint16_t *shm_u16()
    {
        extern __shared__ float4 vb_shared_window[];
        return reinterpret_cast<uint16_t *>(vb_shared_window);
    }
    VB_D unsigned *shm_u32()
    {
        extern __shared__ float4 vb_shared_window[];
        return reinterpret_cast<unsigned *>(vb_shared_window);
    }

    // warp-uniform state of one walk (passed BY VALUE to the stages: 10 registers instead of a local-memory struct)
    struct WalkCtx
    {
        unsigned wcol;    // float index: slot column of lane 0 of this warp
        unsigned queues;  // byte offset of this warp's queue block: [item queue u16][pairq u16][gmask u32][cmask u32 x 4]
        const float4 *prims;
        float bx, by, bz, radius;
        // Motion validation: the 8 lanes of an octet are the 8 interpolants of one rake block of the
        // reference (planning/validate.hh:36-60), whose `if (guard collides in ANY lane)` opens the fine
        // tests of ALL lanes.  That coupling changes verdicts where a guard query is unreliable: CAPT point
        // clouds queried with a bounding-sphere radius beyond the tree's r_max.
        bool octet;
        // Phase-staggered warp groups (vb_kernels.cuh: kGroups): named barrier this warp's group ARRIVES at when it leaves
        // the straight-line ("dense") part of a step, handing the dense section to the next group; 0: none.
        int release_id, release_count;
    };

    VB_D void group_release(const WalkCtx &c)
    {
        if (c.release_id != 0)
        {
            __syncwarp();
            asm volatile("bar.arrive %0, %1;" ::"r"(c.release_id), "r"(c.release_count) : "memory");
        }
    }

    // ---- layout of the dynamic shared-memory window (bytes) -------------------------------------------------
    //   [3 * kSlots rows of kStride floats: one column per thread][per-warp work queues: items u16, units u16, filter tests
    //   u16, candidate masks u32 x 32 x groups, state u32 x 4][staged configurations +
    //   mbarrier][check tables]
    template <class Robot>
    __host__ __device__ constexpr int queue_bytes()
    {
        return 2 * ((Robot::kEnvQ > Robot::kSelfQ ? Robot::kEnvQ : Robot::kSelfQ) + Robot::kPairQ + Robot::kFiltQ) +
               128 * (Robot::kEnvGroups > 0 ? Robot::kEnvGroups : 1) + 16;
    }
    template <class Robot>
    __host__ __device__ constexpr unsigned cfg_stage_bytes()
    {
        return Robot::kStageCfg ? static_cast<unsigned>(Robot::kBlock) * Robot::kDim * 4u + 16u : 0u;  // + the mbarrier
    }
    template <class Robot>
    __host__ __device__ constexpr unsigned cfg_stage_offset()
    {
        return 4u * ((3u * Robot::kSlots * Robot::kStride + 3u) & ~3u) + static_cast<unsigned>(Robot::kBlock / 32) * queue_bytes<Robot>();
    }
    template <class Robot>
    __host__ __device__ constexpr unsigned tab_base()
    {
        return (cfg_stage_offset<Robot>() + cfg_stage_bytes<Robot>() + 15u) & ~15u;
    }
    template <class Robot>
    __host__ __device__ constexpr unsigned walk_smem_bytes()
    {
        return tab_base<Robot>() + (Robot::kTablesInShared ? static_cast<unsigned>(Robot::kTabBytes) : 0u);
    }

    // Check-table records of the warp walk.  Robot::kTablesInShared: the code generator found room for them behind
    // the slot columns (Robot::stage_tables() copies them in at kernel start) and they are read with LDS at fixed
    // window offsets; otherwise through the non-coherent global path.  The unit stages are chains of dependent table
    // reads, and the L1 left beside ~220 KB of shared memory is small.
    constexpr unsigned kNoTab = 0xffffffffu;  // Robot::kTab<Name> of a table that is not in shared memory

    template <class Robot>
    struct WT
    {
        template <unsigned kOff, class T>
        VB_D static T get(const T *global, unsigned i)
        {
            if constexpr (kOff != kNoTab)
            {
                extern __shared__ float4 vb_shared_window[];
                const T *base = reinterpret_cast<const T *>(reinterpret_cast<const unsigned char *>(vb_shared_window) + tab_base<Robot>() + kOff);
                T out;
                if constexpr (sizeof(T) % 16 == 0 && alignof(T) >= 16)
                {
                    const uint4 *src = reinterpret_cast<const uint4 *>(base + i);
                    uint4 *dst = reinterpret_cast<uint4 *>(&out);
#pragma unroll
                    for (unsigned k = 0; k < sizeof(T) / 16; ++k)
                    {
                        dst[k] = src[k];
                    }
                }
                else if constexpr (sizeof(T) % 8 == 0)
                {
                    const uint2 *src = reinterpret_cast<const uint2 *>(base + i);
                    uint2 *dst = reinterpret_cast<uint2 *>(&out);
#pragma unroll
                    for (unsigned k = 0; k < sizeof(T) / 8; ++k)
                    {
                        dst[k] = src[k];
                    }
                }
                else
                {
                    out = base[i];
                }
                return out;
            }
            else
            {
                if constexpr (sizeof(T) % 8 == 0)
                {
                    return ldg_record(global + i);
                }
                else
                {
                    return __ldg(global + i);
                }
            }
        }
        // all threads of the CTA: copy the tables the code generator placed behind the staged configurations (caller
        // synchronises)
        template <unsigned kOff>
        VB_D static void stage_one(const void *src, unsigned bytes)
        {
            if constexpr (kOff != kNoTab)
            {
                extern __shared__ float4 vb_shared_window[];
                copy(reinterpret_cast<unsigned char *>(vb_shared_window) + tab_base<Robot>() + kOff, src, bytes);
            }
        }
        VB_D static void stage()
        {
            if constexpr (Robot::kTablesInShared)
            {
                const WalkTables w = Robot::walk_tables();
                stage_one<Robot::kTabUnits>(w.units, Robot::kTabUnitsBytes);
                stage_one<Robot::kTabBrefs>(w.brefs, Robot::kTabBrefsBytes);
                stage_one<Robot::kTabBws>(w.bws, Robot::kTabBwsBytes);
                stage_one<Robot::kTabDers>(w.ders, Robot::kTabDersBytes);
                stage_one<Robot::kTabEnv>(w.env, Robot::kTabEnvBytes);
                stage_one<Robot::kTabRefs>(w.refs, Robot::kTabRefsBytes);
                stage_one<Robot::kTabRefGroup>(w.ref_group, Robot::kTabRefGroupBytes);
                stage_one<Robot::kTabFilts>(w.filts, Robot::kTabFiltsBytes);
                stage_one<Robot::kTabGUnits>(w.gunits, Robot::kTabGUnitsBytes);
            }
        }
        // tables are arrays of 2-byte-multiple records at 16-byte aligned window offsets
        VB_D static void copy(unsigned char *dst, const void *src, unsigned bytes)
        {
            const uint16_t *s16 = static_cast<const uint16_t *>(src);
            uint16_t *d16 = reinterpret_cast<uint16_t *>(dst);
            for (unsigned i = threadIdx.x; i < bytes / 2u; i += blockDim.x)
            {
                d16[i] = __ldg(s16 + i);
            }
        }
        VB_D static SelfUnit unit(unsigned i) { return get<Robot::kTabUnits>(Robot::walk_tables().units, i); }
        VB_D static BRef bref(unsigned i) { return get<Robot::kTabBrefs>(Robot::walk_tables().brefs, i); }
        VB_D static BWeights bw(unsigned i) { return get<Robot::kTabBws>(Robot::walk_tables().bws, i); }
        VB_D static DerivedRec der(unsigned i) { return get<Robot::kTabDers>(Robot::walk_tables().ders, i); }
        VB_D static RefGroup env(unsigned g) { return get<Robot::kTabEnv>(Robot::walk_tables().env, g); }
        VB_D static SlotRef ref(unsigned i) { return get<Robot::kTabRefs>(Robot::walk_tables().refs, i); }
        VB_D static uint16_t ref_group(unsigned i) { return get<Robot::kTabRefGroup>(Robot::walk_tables().ref_group, i); }
        VB_D static FilterRec filt(unsigned i) { return get<Robot::kTabFilts>(Robot::walk_tables().filts, i); }
        VB_D static uint32_t gunit(unsigned g) { return get<Robot::kTabGUnits>(Robot::walk_tables().gunits, g); }
    };

    template <class Robot>
    struct Walk
    {
        using T = WT<Robot>;
        static constexpr int S = Robot::kStride;
        static constexpr unsigned kGroups = Robot::kEnvGroups > 0 ? Robot::kEnvGroups : 1;

        static constexpr unsigned kConstOff = 3u * Robot::kConstRow * S;
        // translated constants (vb_tables.cuh: kLiftSlot): offsets from kLiftOff on add the z at kLiftZOff (+ lane)
        static constexpr bool kHasLift = Robot::kLiftCount > 0;
        static constexpr unsigned kLiftOff = 3u * Robot::kLiftRow * S;
        static constexpr unsigned kLiftZOff = static_cast<unsigned>(Robot::kLiftZRow) * S;
        static constexpr int kItemQ = Robot::kEnvQ > Robot::kSelfQ ? Robot::kEnvQ : Robot::kSelfQ;

        // self items and environment items are never queued at the same time: one array serves both
        VB_D static uint16_t *envq(const WalkCtx &c) { return shm_u16() + c.queues / 2; }
        VB_D static uint16_t *pairq(const WalkCtx &c) { return envq(c) + kItemQ; }
        VB_D static uint16_t *testq(const WalkCtx &c) { return pairq(c) + Robot::kPairQ; }  // (Robot::kFiltQ entries)
        VB_D static unsigned *gmask(const WalkCtx &c) { return shm_u32() + (c.queues + 2 * (kItemQ + Robot::kPairQ + Robot::kFiltQ)) / 4; }
        VB_D static unsigned *cmask(const WalkCtx &c) { return gmask(c) + 32 * kGroups; }
        // rake blocks (WalkCtx::octet): bit L of gfire[g] = the exact guard test of group g fired for configuration L.  The
        // words live in the last kFireQ entries of the item queue, which holds that many items less for rake blocks.
        static constexpr int kFireQ = 2 * ((static_cast<int>(kGroups) + 1) & ~1);
        static_assert(Robot::kEnvQ >= kFireQ + 32, "the item queue holds the fired-guard words and at least one group's items");
        VB_D static unsigned *gfire(const WalkCtx &c) { return shm_u32() + (c.queues + 2 * (kItemQ - kFireQ)) / 4; }

        VB_D static unsigned vol(const unsigned *p)
        {
            return *reinterpret_cast<volatile const unsigned *>(p);
        }

        VB_D static bool done(const WalkCtx &c, int L)
        {
            return (vol(cmask(c)) >> L) & 1u;
        }

        VB_D static Center centre(const WalkCtx &c, int L, const SlotRef &ref)
        {
            Center s;
            if (Robot::kDerived > 0 && (ref.slot & kDerivedSlot))
            {
                derived(c, L, T::der(ref.slot & (kDerivedSlot - 1u)), s.x, s.y, s.z);
            }
            else
            {
                const float *col = shm_f32() + slot_index<S, Robot::kConstRow, Robot::kLiftRow>(c.wcol, L, ref.slot);
                s.x = col[0];
                s.y = col[S];
                s.z = col[2 * S];
                if (kHasLift && (ref.slot & kLiftSlot))
                {
                    s.z += shm_f32()[c.wcol + kLiftZOff + L];
                }
            }
            s.x += (ref.base & 1) ? c.bx : 0.f;
            s.y += (ref.base & 2) ? c.by : 0.f;
            s.z += (ref.base & 4) ? c.bz : 0.f;
            s.r = ref.r < 0.f ? c.radius : ref.r;
            return s;
        }

        // centre rebuilt from its link's basis (vb_tables.cuh: DerivedRec)
        VB_D static void derived(const WalkCtx &c, int L, const DerivedRec &d, float &x, float &y, float &z)
        {
            float bx[4], by[4], bz[4];
#pragma unroll
            for (int k = 0; k < 4; ++k)
            {
                at(c, L, d.off[k], bx[k], by[k], bz[k]);
            }
            x = d.w[0] * bx[0] + d.w[1] * bx[1] + d.w[2] * bx[2] + d.w[3] * bx[3];
            y = d.w[0] * by[0] + d.w[1] * by[1] + d.w[2] * by[2] + d.w[3] * by[3];
            z = d.w[0] * bz[0] + d.w[1] * bz[1] + d.w[2] * bz[2] + d.w[3] * bz[3];
        }

        // pre-multiplied offset (SelfUnit / BRef): per-thread slots add the lane, per-warp constants do not
        VB_D static void at(const WalkCtx &c, int L, unsigned off, float &x, float &y, float &z)
        {
            const float *col = shm_f32() + c.wcol + off + (off >= kConstOff ? 0u : static_cast<unsigned>(L));
            x = col[0];
            y = col[S];
            z = col[2 * S];
            if (kHasLift && off >= kLiftOff)
            {
                z += shm_f32()[c.wcol + kLiftZOff + L];
            }
        }

        // a_i of configuration L against every b_j of its group
        VB_D static bool pair_unit(const WalkCtx &c, const WalkTables &w, int L, const SelfUnit &u)
        {
            float ax, ay, az;
            if (Robot::kDerived > 0 && u.a_derived())
            {
                derived(c, L, T::der(u.a_off), ax, ay, az);
            }
            else
            {
                at(c, L, u.a_off, ax, ay, az);
            }
            const int bc = u.b_count();
            bool hit = false;
            if (Robot::kDerived > 0 && u.b_basis() >= 0)
            {
                // reduced b-link: a_i relative to the link's basis once, then 3 FMAs per b_j (vb_tables.cuh: BWeights)
                const DerivedRec bd = T::der(u.b_basis());
                float px[4], py[4], pz[4];
#pragma unroll
                for (int k = 0; k < 4; ++k)
                {
                    at(c, L, bd.off[k], px[k], py[k], pz[k]);
                }
                const float ux = ax - px[0], uy = ay - py[0], uz = az - pz[0];
                const float s0 = ux * ux + uy * uy + uz * uz;
                float q[3];
#pragma unroll
                for (int k = 0; k < 3; ++k)
                {
                    q[k] = -2.f * (ux * (px[k + 1] - px[0]) + uy * (py[k + 1] - py[0]) + uz * (pz[k + 1] - pz[0]));
                }
                // (bitwise accumulation, four b spheres per trip with their loads issued together: `hit = hit || ...` compiles
                // to one dependent branch per sphere, i.e. a full load-to-compare latency chain each)
                bool near = false;
                for (int j = 0; j < bc; j += 4)
                {
                    BWeights bw[4];
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                    {
                        bw[k] = T::bw(u.b_begin + min(j + k, bc - 1));  // (past the end: the last sphere again)
                    }
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                    {
                        const float d2 = s0 + bw[k].w[0] * q[0] + bw[k].w[1] * q[1] + bw[k].w[2] * q[2] + bw[k].w[3];
                        const float rs = u.a_r + bw[k].r;
                        const float v = d2 - rs * rs;
                        near |= fabsf(v) < kGramBand;
                        hit |= v < -kGramBand;
                    }
                }
                if (near && !hit)
                {
                    // |u|^2 and the dot products cancel (a few 1e-6 m^2 of rounding): a unit with a sphere that close to
                    // contact is decided by rebuilding the b spheres from the basis edges and testing them directly, as
                    // the stored b-loop does
                    for (int j = 0; j < bc; ++j)
                    {
                        const BWeights bw = T::bw(u.b_begin + j);
                        const float bx = px[0] + bw.w[0] * (px[1] - px[0]) + bw.w[1] * (px[2] - px[0]) + bw.w[2] * (px[3] - px[0]);
                        const float by = py[0] + bw.w[0] * (py[1] - py[0]) + bw.w[1] * (py[2] - py[0]) + bw.w[2] * (py[3] - py[0]);
                        const float bz = pz[0] + bw.w[0] * (pz[1] - pz[0]) + bw.w[1] * (pz[2] - pz[0]) + bw.w[2] * (pz[3] - pz[0]);
                        const float dx = ax - bx, dy = ay - by, dz = az - bz, rs = u.a_r + bw.r;
                        hit = hit || (dx * dx + dy * dy + dz * dz - rs * rs < 0.f);
                    }
                }
                return hit;
            }
            for (int j = 0; j < bc; j += 4)
            {
                BRef br[4];
#pragma unroll
                for (int k = 0; k < 4; ++k)
                {
                    br[k] = T::bref(u.b_begin + min(j + k, bc - 1));  // (past the end: the last sphere again)
                }
                float bx[4], by[4], bz[4];
#pragma unroll
                for (int k = 0; k < 4; ++k)
                {
                    at(c, L, br[k].off, bx[k], by[k], bz[k]);
                }
#pragma unroll
                for (int k = 0; k < 4; ++k)
                {
                    const float dx = ax - bx[k], dy = ay - by[k], dz = az - bz[k], rs = u.a_r + br[k].r;
                    hit |= dx * dx + dy * dy + dz * dz - rs * rs < 0.f;
                }
            }
            return hit;
        }

        // ---- self collision: queued (configuration, a_i) units, one per lane -----------------------------
        VB_STAGE static void pair_stage(const WalkCtx c, int np)
        {
            const WalkTables w = Robot::walk_tables();
            const int lane = threadIdx.x & 31;
            VB_STAT_ONE(10, np);
            for (int base = 0; base < np; base += 32)
            {
                const int i = base + lane;
                int nb_ = 0;
                if (i < np)
                {
                    const unsigned code = pairq(c)[i];
                    const int L = code & 31;
                    const SelfUnit u = T::unit(code >> 5);
                    const bool skip_ = done(c, L);
                    nb_ = skip_ ? 0 : u.b_count();
                    if (!skip_ && pair_unit(c, w, L, u))
                    {
                        atomicOr(cmask(c), 1u << L);
                    }
                }
                VB_STAT_ONE(11, 1);
                VB_STAT_ADD(12, nb_);
                VB_STAT_MAX(13, nb_);
            }
            __syncwarp();
        }

        // Queues the units of this lane's bit set (bit = unit index) behind those of the lower lanes and runs
        // the pair stage.  All 32 lanes call.
        VB_STAGE static void run_units(const WalkCtx c, const uint32_t *ubits)
        {
            const int lane = threadIdx.x & 31;
            int cnt = 0;
#pragma unroll
            for (int wi = 0; wi < Robot::kUnitWords; ++wi)
            {
                cnt += __popc(ubits[wi]);
            }
            int incl = cnt;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1)
            {
                const int v = __shfl_up_sync(kFull, incl, d);
                incl += lane >= d ? v : 0;
            }
            const int total = __shfl_sync(kFull, incl, 31);
            if (total == 0)
            {
                return;
            }
            if (total <= Robot::kPairQ)
            {
                int at_ = incl - cnt;
#pragma unroll
                for (int wi = 0; wi < Robot::kUnitWords; ++wi)
                {
                    uint32_t m = ubits[wi];
                    while (m)
                    {
                        const int b = __ffs(m) - 1;
                        m &= m - 1u;
                        pairq(c)[at_++] = static_cast<uint16_t>(lane | ((wi * 32 + b) << 5));
                    }
                }
                __syncwarp();
                pair_stage(c, total);
                return;
            }
            // more units than the queue holds (rare): every lane runs its own units
            const WalkTables w = Robot::walk_tables();
            bool hit = false;
#pragma unroll
            for (int wi = 0; wi < Robot::kUnitWords; ++wi)
            {
                uint32_t m = ubits[wi];
                while (m && !hit)
                {
                    const int b = __ffs(m) - 1;
                    m &= m - 1u;
                    hit = pair_unit(c, w, lane, T::unit(wi * 32 + b));
                }
            }
            if (hit)
            {
                atomicOr(cmask(c), 1u << lane);
            }
            __syncwarp();
        }

        // centre of a slot code (SlotRef::slot: stored column, per-warp constant or rebuilt from the link basis), no base offset
        VB_D static void centre_xyz(const WalkCtx &c, int L, unsigned slot, float &x, float &y, float &z)
        {
            if (Robot::kDerived > 0 && (slot & kDerivedSlot))
            {
                derived(c, L, T::der(slot & (kDerivedSlot - 1u)), x, y, z);
            }
            else
            {
                const float *col = shm_f32() + slot_index<S, Robot::kConstRow, Robot::kLiftRow>(c.wcol, L, slot);
                x = col[0];
                y = col[S];
                z = col[2 * S];
                if (kHasLift && (slot & kLiftSlot))
                {
                    z += shm_f32()[c.wcol + kLiftZOff + L];
                }
            }
        }

        VB_D static int warp_scan(int v, int lane)
        {
#pragma unroll
            for (int d = 1; d < 32; d <<= 1)
            {
                const int o = __shfl_up_sync(kFull, v, d);
                v += lane >= d ? o : 0;
            }
            return v;
        }

        // ---- self collision, queued a-side filters (Robot::kQueuedFilters) ------------------------------------------
        // fired[]: this lane's guard hits (leaf groups, collided lanes and tabulated-free groups already cleared).  The
        // (lane, group) items of the warp are queued, expanded into one (configuration, a_i) filter test per lane --
        // a_i against the other link's bounding sphere inflated by the group's static slack, the comparison of the
        // straight-line form -- and the survivors are compacted into the pair queue.  All 32 lanes call.
        VB_STAGE static void filter_units(const WalkCtx c, uint32_t *fired)
        {
            const int lane = threadIdx.x & 31;
            const unsigned lt = (1u << lane) - 1u;
            uint16_t *iq = envq(c), *tq = testq(c), *pq = pairq(c);
            // tests per lane and trip (measured, MBM scenes: Fetch 1 / 2 / 4 -> 1.00 / 1.12 / 1.14 Gcfg/s; Baxter, whose rebuilt
            // centres make the body branchy and which is at the register ceiling, 0.92 / 0.90 / 0.87)
#ifdef VB_FILT_ILP
            constexpr int kIlp = VB_FILT_ILP;
#else
            constexpr int kIlp = Robot::kDerived > 0 ? 1 : 4;
#endif
            static_assert(Robot::kPairQ >= 32 * kIlp && Robot::kFiltQ >= 32, "a trip's survivors must fit the pair queue");
            int np = 0;  // units waiting in the pair queue
            for (;;)
            {
                int cnt = 0;
#pragma unroll
                for (int wi = 0; wi < Robot::kSelfWords; ++wi)
                {
                    cnt += __popc(fired[wi]);
                }
                const int total = __reduce_add_sync(kFull, cnt);
                if (total == 0)
                {
                    break;
                }
                // more items than the queue holds (rare): every lane queues its first kSelfQ / 32 now, the rest next round
                int take = total > Robot::kSelfQ ? min(cnt, Robot::kSelfQ / 32) : cnt;
                const int incl = warp_scan(take, lane);
                const int n_items = __shfl_sync(kFull, incl, 31);
                int at_ = incl - take;
#pragma unroll
                for (int wi = 0; wi < Robot::kSelfWords; ++wi)
                {
                    uint32_t m = fired[wi];
                    while (m != 0u && take > 0)
                    {
                        const int b = __ffs(m) - 1;
                        m &= m - 1u;
                        iq[at_++] = static_cast<uint16_t>(lane | ((wi * 32 + b) << 5));
                        --take;
                    }
                    fired[wi] = m;
                }
                __syncwarp();
                for (int ib = 0; ib < n_items;)
                {
                    // expand up to 32 items into filter tests (as many whole items as the test queue holds)
                    const int i = ib + lane;
                    int L = 0, u0 = 0, ac = 0;
                    if (i < n_items)
                    {
                        const unsigned code = iq[i];
                        L = code & 31;
                        if (!done(c, L))
                        {
                            const uint32_t gu = T::gunit(code >> 5);
                            u0 = gu & 0xffffu;
                            ac = gu >> 16;
                        }
                    }
                    const int tin = warp_scan(ac, lane);
                    const int nfit = __popc(__ballot_sync(kFull, tin <= Robot::kFiltQ));  // (>= 1: kFiltQ >= every a_count)
                    const int n_tests = __shfl_sync(kFull, tin, nfit - 1);
                    if (lane < nfit)
                    {
                        uint16_t *dst = tq + (tin - ac);
                        for (int k = 0; k < ac; ++k)
                        {
                            dst[k] = static_cast<uint16_t>(L | ((u0 + k) << 5));
                        }
                    }
                    __syncwarp();
                    ib += nfit;
                    // kIlp tests per lane and trip, their loads issued together (few resident warps: the dependent chain
                    // queue -> record -> centres -> compare is what a pass costs)
                    for (int tb = 0; tb < n_tests; tb += 32 * kIlp)
                    {
                        unsigned code[kIlp];
                        FilterRec f[kIlp];
                        bool pass[kIlp];
#pragma unroll
                        for (int k = 0; k < kIlp; ++k)
                        {
                            code[k] = tq[min(tb + 32 * k + lane, n_tests - 1)];
                        }
#pragma unroll
                        for (int k = 0; k < kIlp; ++k)
                        {
                            f[k] = T::filt(code[k] >> 5);
                        }
#pragma unroll
                        for (int k = 0; k < kIlp; ++k)
                        {
                            const int Lt = code[k] & 31;
                            float ax, ay, az, gx, gy, gz;
                            centre_xyz(c, Lt, f[k].a_slot, ax, ay, az);
                            centre_xyz(c, Lt, f[k].gb_slot, gx, gy, gz);
                            const float dx = ax - gx, dy = ay - gy, dz = az - gz;
                            pass[k] = (tb + 32 * k + lane < n_tests) & ((f[k].thr2 < 0.f) | (dx * dx + dy * dy + dz * dz < f[k].thr2));
                        }
                        if (np + 32 * kIlp > Robot::kPairQ)
                        {
                            __syncwarp();
                            pair_stage(c, np);
                            np = 0;
                        }
#pragma unroll
                        for (int k = 0; k < kIlp; ++k)
                        {
                            const unsigned m = __ballot_sync(kFull, pass[k]);
                            if (pass[k])
                            {
                                pq[np + __popc(m & lt)] = static_cast<uint16_t>(code[k]);
                            }
                            np += __popc(m);
                        }
                    }
                    __syncwarp();
                }
            }
            __syncwarp();
            pair_stage(c, np);
        }

        // one fine sphere refs[idx] of configuration L against the tight candidates of its group
        VB_D static bool fine_unit(const WalkCtx &c, const WalkTables &w, const PrimIndex &ix, int L, int idx, int *n_tested = nullptr)
        {
            const int g = T::ref_group(idx) & 0x7fff;
            uint32_t tight = gmask(c)[32 * g + L];
            const bool thigh = (tight >> 31) != 0u;
            tight &= 0x7fffffffu;
            const Center s = centre(c, L, T::ref(idx));
            bool hit = false;
            int cur = 0;
            for (int o = next_candidate(tight, thigh, ix.first[5], cur); o >= 0 && !hit;
                 o = next_candidate(tight, thigh, ix.first[5], cur))
            {
                hit = object_hit(object_dist(c.prims, ix, o, s.x, s.y, s.z), s.r);
#ifdef VB_STATS
                if (n_tested)
                {
                    ++*n_tested;
                }
#endif
            }
            return hit;
        }

        // ---- environment, stage 2: queued (configuration, fine sphere) units ---------------------------
        VB_STAGE static void fine_stage(const EnvView &env, const WalkCtx c)
        {
            const WalkTables w = Robot::walk_tables();
            const int lane = threadIdx.x & 31;
            const PrimIndex ix = prim_index(env);
            __syncwarp();
            const int nu = min(static_cast<int>(vol(cmask(c) + 1)), Robot::kPairQ);
            VB_STAT_ONE(6, nu);
            for (int base = 0; base < nu; base += 32)
            {
                const int i = base + lane;
                int nt_ = 0;
                if (i < nu)
                {
                    const unsigned code = pairq(c)[i];
                    const int L = code & 31;
                    if (!done(c, L) && fine_unit(c, w, ix, L, code >> 5, &nt_))
                    {
                        atomicOr(cmask(c), 1u << L);
                    }
                }
                VB_STAT_ONE(7, 1);
                VB_STAT_ADD(8, nt_);
                VB_STAT_MAX(9, nt_);
            }
            __syncwarp();
            if (lane == 0)
            {
                cmask(c)[1] = 0u;
            }
            __syncwarp();
        }

        // ---- environment, stage 1: one (configuration, group) item per lane ---------------------------
        // The item's guard sphere is tested exactly against the candidate obstacles of its grid cell (this
        // is the reference's `if (guard collides)`), which also narrows the candidates to the obstacles
        // that touch the guard sphere inflated by the group's static slack.  The fine spheres of a fired
        // guard become units of stage 2 (the pair queue is free at this point and is reused), where they
        // are tested, with the exact primitive arithmetic of the reference, against those few obstacles.
        // ---- environment with height fields / point clouds: the fine spheres of the queued (configuration, group) items
        // get the full query of the reference (their guards were tested exactly), one sphere per lane and step, decided by
        // the warp together (vb_env.cuh: env_hit_warp -- leaf scans of point clouds are done by all 32 lanes)
        VB_STAGE static void env_stage_other(const EnvView &env, const WalkCtx c, int ne)
        {
            const int lane = threadIdx.x & 31;
            for (int base = 0; base < ne; base += 32)
            {
                const int i = base + lane;
                int L = 0, u0 = 0, uc = 0;
                if (i < ne)
                {
                    const unsigned code = envq(c)[i];
                    L = code & 31;
                    if (!done(c, L))
                    {
                        const RefGroup G = T::env(code >> 5);
                        u0 = G.ubegin;
                        uc = G.ucount;
                    }
                }
                const int max_uc = __reduce_max_sync(kFull, uc);
                bool hit = false;
                for (int k = 0; k < max_uc; ++k)
                {
                    const bool active = k < uc && !hit;
                    Center s{0.f, 0.f, 0.f, 0.f};
                    if (active)
                    {
                        s = centre(c, L, T::ref(u0 + k));
                    }
                    hit = env_hit_warp(env, c.prims, active, s.x, s.y, s.z, s.r) || hit;
                }
                if (hit)
                {
                    atomicOr(cmask(c), 1u << L);
                }
                __syncwarp();
            }
        }

        VB_STAGE static void env_stage(const EnvView &env, const WalkCtx c, int ne)
        {
            const WalkTables w = Robot::walk_tables();
            const int lane = threadIdx.x & 31;
            const bool has_other = (env.n_heightfields | env.n_capts) != 0;
            if (has_other)
            {
                env_stage_other(env, c, ne);
                return;
            }
            const PrimIndex ix = prim_index(env);
            const int n_obj = ix.first[5];
            VB_STAT_ONE(1, ne);
            if (c.octet)
            {
                // Edges: the 8 lanes of an octet are one rake block of the reference, whose guard opens the fine tests of ALL
                // its lanes when it fires in ANY of them (planning/validate.hh:36-60).  Two sweeps over the queued items: the
                // exact guard tests (bit L of gfire[g], candidates narrowed into gmask), then the fine units of every item
                // whose block fired.  The items of one group are queued together, so both sweeps see all of them.
                if (lane < static_cast<int>(kGroups))
                {
                    gfire(c)[lane] = 0u;
                }
                __syncwarp();
                for (int base = 0; base < ne; base += 32)
                {
                    const int i = base + lane;
                    if (i < ne)
                    {
                        const unsigned code = envq(c)[i];
                        const int L = code & 31, g = code >> 5;
                        if (!done(c, L))
                        {
                            const RefGroup G = T::env(g);
                            const uint32_t mask = gmask(c)[32 * g + L];
                            const Center b = centre(c, L, G.guard);
                            uint32_t low = mask & 0x7fffffffu, tight = 0u;
                            if (n_obj < 31)
                            {
                                low &= (1u << n_obj) - 1u;
                            }
                            const bool high = (mask >> 31) != 0u;
                            bool ghit = false;
                            int cursor = 0;
                            for (int o = next_candidate(low, high, n_obj, cursor); o >= 0; o = next_candidate(low, high, n_obj, cursor))
                            {
                                const ObjDist d = object_dist(c.prims, ix, o, b.x, b.y, b.z);
                                ghit |= object_hit(d, b.r);
                                if (G.slack < 0.f || object_hit(d, b.r + G.slack))
                                {
                                    tight |= 1u << (o < 31 ? o : 31);
                                }
                            }
                            gmask(c)[32 * g + L] = tight;
                            if (ghit)
                            {
                                atomicOr(gfire(c) + g, 1u << L);
                            }
                        }
                    }
                }
                __syncwarp();
                for (int base = 0; base < ne; base += 32)
                {
                    const int i = base + lane;
                    if (i < ne)
                    {
                        const unsigned code = envq(c)[i];
                        const int L = code & 31, g = code >> 5;
                        const unsigned fired = vol(gfire(c) + g);
                        if (!done(c, L) && ((fired >> (L & ~7)) & 0xffu) != 0u)
                        {
                            const RefGroup G = T::env(g);
                            bool hit = false;
                            if (G.count == 0)
                            {
                                hit = (fired >> L) & 1u;  // leaf group: the guard is the test
                            }
                            else if (gmask(c)[32 * g + L] != 0u)
                            {
                                const int u0 = G.ubegin, uc = G.ucount;
                                const unsigned at = atomicAdd(cmask(c) + 1, static_cast<unsigned>(uc));
                                for (int k = 0; k < uc; ++k)
                                {
                                    if (at + k < static_cast<unsigned>(Robot::kPairQ))
                                    {
                                        pairq(c)[at + k] = static_cast<uint16_t>(L | ((u0 + k) << 5));
                                    }
                                    else
                                    {
                                        hit = hit || fine_unit(c, w, ix, L, u0 + k);  // queue full: run the unit here
                                    }
                                }
                            }
                            if (hit)
                            {
                                atomicOr(cmask(c), 1u << L);
                            }
                        }
                    }
                    __syncwarp();
                    if (vol(cmask(c) + 1) >= static_cast<unsigned>(Robot::kPairQ) / 2)
                    {
                        fine_stage(env, c);
                    }
                }
                fine_stage(env, c);
                return;
            }
            for (int base = 0; base < ne; base += 32)
            {
                const int i = base + lane;
                int nc_ = 0, gh_ = 0;
                if (i < ne)
                {
                    const unsigned code = envq(c)[i];
                    const int L = code & 31, g = code >> 5;
                    if (!done(c, L))
                    {
                        const RefGroup Gr = T::env(g);
                        const RefGroup *G = &Gr;
                        const int u0 = G->ubegin, uc = G->ucount;
                        bool hit = false;
                        if (!has_other)
                        {
                            const uint32_t mask = gmask(c)[32 * g + L];
                            const Center b = centre(c, L, G->guard);
                            const float slack = G->slack;
                            uint32_t low = mask & 0x7fffffffu, tight = 0u;
                            if (n_obj < 31)
                            {
                                low &= (1u << n_obj) - 1u;
                            }
                            const bool high = (mask >> 31) != 0u;
                            bool ghit = false;
                            int cursor = 0;
                            for (int o = next_candidate(low, high, n_obj, cursor); o >= 0;
                                 o = next_candidate(low, high, n_obj, cursor))
                            {
                                const ObjDist d = object_dist(c.prims, ix, o, b.x, b.y, b.z);
                                ++nc_;
                                ghit = ghit || object_hit(d, b.r);
                                if (slack < 0.f || object_hit(d, b.r + slack))
                                {
                                    tight |= 1u << (o < 31 ? o : 31);
                                }
                            }
                            if (G->count == 0)
                            {
                                hit = ghit;  // leaf group: the guard is the test
                            }
                            else if (ghit)
                            {
                                gh_ = 1;
                                gmask(c)[32 * g + L] = tight;
                                const unsigned at = atomicAdd(cmask(c) + 1, static_cast<unsigned>(uc));
                                for (int k = 0; k < uc; ++k)
                                {
                                    if (at + k < static_cast<unsigned>(Robot::kPairQ))
                                    {
                                        pairq(c)[at + k] = static_cast<uint16_t>(L | ((u0 + k) << 5));
                                    }
                                    else
                                    {
                                        hit = hit || fine_unit(c, w, ix, L, u0 + k);  // queue full: run the unit here
                                    }
                                }
                            }
                        }
                        // (height fields / point clouds never get here: env_stage_other)
                        if (hit)
                        {
                            atomicOr(cmask(c), 1u << L);
                        }
                    }
                }
                __syncwarp();
                VB_STAT_ONE(2, 1);
                VB_STAT_ADD(3, nc_);
                VB_STAT_MAX(4, nc_);
                VB_STAT_ADD(5, gh_);
                if (vol(cmask(c) + 1) >= static_cast<unsigned>(Robot::kPairQ) / 2)
                {
                    fine_stage(env, c);
                }
            }
            fine_stage(env, c);
        }

        // Must be called by all 32 lanes of a warp, converged, after the chain of every live lane has
        // written its slot column (and a __syncwarp).  `collided` enters true for lanes without a
        // configuration.  Returns this lane's verdict (true = collision free).
        // `cstate` is the warp's cached verdict of the configuration-independent environment groups
        // (-1 unknown, 0 free, 1 colliding): a property of (environment, base), evaluated once.
        // `pfree`: bit g set = self group g cannot collide at this lane's configuration (joint-pair tables).
        // kSync: the caller's loop is CTA-uniform (fkcc kernels); every path through run() then passes exactly one CTA
        // barrier between the self-collision and the environment part, which re-aligns the warps' instruction streams
        template <bool kSync>
        VB_D static bool run(const EnvView &env, const WalkCtx &c, bool collided, int &cstate, const uint32_t *pfree)
        {
            VB_TICK_BEGIN;
            const auto &t = Robot::tables();
            const int lane = threadIdx.x & 31;
            const unsigned lt = (1u << lane) - 1u;
            const float *col = shm_f32() + c.wcol + lane;
            VB_STAT_ONE(0, 1);
            VB_STAT_ADD(19, collided ? 0 : 1);
            if (lane < 2)
            {
                cmask(c)[lane] = 0u;
            }
            __syncwarp();
            if (Robot::kEnvConstAny && (env.n_quads > 0 || ((env.n_heightfields | env.n_capts) != 0)))
            {
                const unsigned alive = __ballot_sync(kFull, !collided);
                if (cstate < 0 && alive != 0u)
                {
                    // reference semantics, evaluated on the column of the first live lane (the spheres are constants)
                    const int L0 = __ffs(alive) - 1;
                    NoMargin mg;
                    bool hit = false;
                    for (int g = 0; g < t.n_env && !hit; ++g)
                    {
                        if (!((Robot::env_const_mask(g >> 5) >> (g & 31)) & 1u))
                        {
                            continue;
                        }
                        const RefGroup grp = t.env[g];
                        const Center b = centre(c, L0, grp.guard);
                        if (env_hit(env, c.prims, b.x, b.y, b.z, b.r, mg))
                        {
                            hit = grp.count == 0;
                            for (int k = 0; k < grp.count && !hit; ++k)
                            {
                                const Center s = centre(c, L0, t.refs[grp.begin + k]);
                                hit = env_hit(env, c.prims, s.x, s.y, s.z, s.r, mg);
                            }
                        }
                    }
                    cstate = hit ? 1 : 0;
                }
                if (cstate > 0)
                {
                    if (kSync)
                    {
                        __syncthreads();
                    }
                    group_release(c);
                    return false;
                }
            }
            // the grid cells of the environment guard spheres are fetched now and consumed after the
            // self-collision phase, which hides their L2 latency
            const bool has_other = (env.n_heightfields | env.n_capts) != 0;
            const bool grid = env.sdf.cell != nullptr && !has_other;
            uint2 cells[Robot::kEnvGroups > 0 ? Robot::kEnvGroups : 1];
            uint32_t inside[Robot::kEnvWords], maybe[Robot::kEnvWords], cert[Robot::kEnvWords];
            bool classified = false;
            if (grid)
            {
                Robot::env_fetch(col, c.bx, c.by, c.bz, env.sdf, env.sdf_c, c.radius, cells, inside);
                if constexpr (Robot::kEnvGroups > 16)
                {
                    // many guards (Baxter: 33): 2 registers of in-flight cell per guard held across the self-collision
                    // guards cost more in spills than the hidden latency earns (1.57 -> 1.61 Gcfg/s)
                    Robot::env_classify(cells, inside, c.radius, env.sdf, maybe, cert, gmask(c) + lane);
                    classified = true;
                }
            }

            // ---------------- self collision ----------------
            if (t.n_self > 0)
            {
                uint32_t fired[Robot::kSelfWords];
                VB_TICK(2);
                Robot::self_guards(col, fired);
                VB_TICK(3);
                if (grid && !classified)
                {
                    // consumed here, not after the stages below: 22 registers of in-flight loads held across
                    // their calls would be spilled, i.e. waited for, at the first call
                    Robot::env_classify(cells, inside, c.radius, env.sdf, maybe, cert, gmask(c) + lane);
                    classified = true;
                }
                VB_TICK(7);
                const uint32_t live = collided ? 0u : 0xffffffffu;
                if (c.octet)
                {
                    // rake coupling (planning/validate.hh:36-60 + robots/<robot>/fk.hh): a guard that fires in ANY of the 8
                    // lanes of a block opens the fine tests of ALL its lanes.  Fine spheres overhang their traced bounding
                    // spheres by up to 0.6 mm, so a lane whose own guard misses can still collide in that sliver.
#pragma unroll
                    for (int wi = 0; wi < Robot::kSelfWords; ++wi)
                    {
                        uint32_t v = fired[wi] & ~Robot::self_leaf_mask(wi);  // (a leaf group's guard IS its test: no coupling)
                        v |= __shfl_xor_sync(kFull, v, 1);
                        v |= __shfl_xor_sync(kFull, v, 2);
                        v |= __shfl_xor_sync(kFull, v, 4);
                        fired[wi] |= v;
                    }
                }
#pragma unroll
                for (int wi = 0; wi < Robot::kSelfWords; ++wi)
                {
                    fired[wi] &= live & ~pfree[wi];
                    collided = collided || (fired[wi] & Robot::self_leaf_mask(wi)) != 0u;
                    VB_STAT_ADD(14, __popc(fired[wi]));
#ifdef VB_STATS
                    {
                        const unsigned any_ = __reduce_or_sync(kFull, fired[wi]);  // (all lanes: not inside VB_STAT_ONE's lane-0 branch)
                        VB_STAT_ONE(15, __popc(any_));
                    }
#endif
                }
                VB_TICK(4);
                // a-side filters of the fired groups, lane-parallel and straight-line; the surviving (group, a_i)
                // units of all 32 configurations are then spread over the lanes
                uint32_t ubits[Robot::kUnitWords];
#pragma unroll
                for (int wi = 0; wi < Robot::kSelfWords; ++wi)
                {
                    fired[wi] = collided ? 0u : (fired[wi] & ~Robot::self_leaf_mask(wi));
                }
                if constexpr (Robot::kQueuedFilters)
                {
                    group_release(c);
                    VB_TICK(5);
                    filter_units(c, fired);
                }
                else
                {
                    Robot::self_filters(col, fired, ubits);
                    VB_TICK(5);
                    group_release(c);  // end of the straight-line part of the step
                    run_units(c, ubits);
                }
                VB_TICK(6);
                collided = collided || done(c, lane);
            }
            else
            {
                if (grid)
                {
                    Robot::env_classify(cells, inside, c.radius, env.sdf, maybe, cert, gmask(c) + lane);
                    classified = true;
                }
                group_release(c);
            }
            if (kSync)
            {
                __syncthreads();
            }
            VB_STAT_ADD(18, collided ? 1 : 0);
            if (__all_sync(kFull, collided))
            {
                return false;
            }

            // ---------------- environment ----------------
            if (env.n_quads == 0 && !has_other)
            {
                return !collided;
            }
            if (grid)
            {
                if (!classified)
                {
                    Robot::env_classify(cells, inside, c.radius, env.sdf, maybe, cert, gmask(c) + lane);
                }
            }
            else
            {
                // height fields / point clouds (or no grid): exact guard tests, lane-parallel
#pragma unroll
                for (int wi = 0; wi < Robot::kEnvWords; ++wi)
                {
                    maybe[wi] = cert[wi] = 0u;
                }
                for (int g = 0; g < t.n_env; ++g)
                {
                    const RefGroup grp = t.env[g];
                    const Center b = centre(c, lane, grp.guard);
                    const bool hit = env_hit_warp(env, c.prims, !collided, b.x, b.y, b.z, b.r);
                    gmask(c)[32 * g + lane] = 0xffffffffu;
#pragma unroll
                    for (int wi = 0; wi < Robot::kEnvWords; ++wi)
                    {
                        if (wi == (g >> 5) && hit)
                        {
                            maybe[wi] |= 1u << (g & 31);
                            cert[wi] |= 1u << (g & 31);
                        }
                    }
                }
                if (c.octet)
                {
                    // the guard of a rake block fires when it fires in any of its 8 lanes
#pragma unroll
                    for (int wi = 0; wi < Robot::kEnvWords; ++wi)
                    {
                        uint32_t v = maybe[wi];
                        v |= __shfl_xor_sync(kFull, v, 1);
                        v |= __shfl_xor_sync(kFull, v, 2);
                        v |= __shfl_xor_sync(kFull, v, 4);
                        maybe[wi] = v;
                    }
                }
            }
            int ne = 0;
#pragma unroll
            for (int wi = 0; wi < Robot::kEnvWords; ++wi)
            {
                VB_STAT_ADD(16, collided ? 0 : __popc(cert[wi] & ~Robot::env_const_mask(wi)));
                const uint32_t mine = collided ? 0u : (maybe[wi] & ~Robot::env_const_mask(wi));
                uint32_t any = __reduce_or_sync(kFull, mine);
                while (any)
                {
                    const int b = __ffs(any) - 1;
                    any &= any - 1u;
                    const bool on = (mine >> b) & 1u;
                    const unsigned m = __ballot_sync(kFull, on);
                    if (ne + 32 > Robot::kEnvQ - (c.octet ? kFireQ : 0))
                    {
                        __syncwarp();
                        env_stage(env, c, ne);
                        ne = 0;
                    }
                    if (on)
                    {
                        envq(c)[ne + __popc(m & lt)] = static_cast<uint16_t>(lane | ((wi * 32 + b) << 5));
                    }
                    ne += __popc(m);
                }
            }
            __syncwarp();
            VB_TICK(8);
            env_stage(env, c, ne);
            VB_TICK(9);
            return !(collided || done(c, lane));
        }
    };
}  // namespace vb

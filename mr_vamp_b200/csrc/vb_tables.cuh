// Check tables emitted by codegen/emit_cuda.py: which sphere slots form which collision groups.
// They restate the control structure of interleaved_sphere_fk (robots/<robot>/fk.hh): a group is
// "if (guard test) { fine tests }" and the configuration is valid iff no group has both a guard
// hit and a fine hit.
#pragma once
#include <stdint.h>

namespace vb
{
    // slot code of a configuration-independent centre: kConstSlot | j (codegen/emit_cuda.py: CONST_SLOT)
    constexpr unsigned kConstSlot = 0x4000u;

    // slot code of a centre the chain does not store: kDerivedSlot | index into the DerivedRec table.  Such a
    // centre is a constant affine combination of <= 4 stored centres of the same rigid link (codegen/emit_cuda.py:
    // plan_links); it is rebuilt where a fired group needs it.
    constexpr unsigned kDerivedSlot = 0x2000u;

    // slot code of a translated constant: kLiftSlot | j.  (x, y, dz) lives once per warp in the rows from Robot::kLiftRow;
    // the centre is (x, y, dz + z of the stored centre whose z row is Tables::lift_z_row) -- Fetch's torso and head
    // spheres ride on the prismatic torso_lift joint (codegen/emit_cuda.py: LIFT_SLOT).
    constexpr unsigned kLiftSlot = 0x1000u;

    struct alignas(8) DerivedRec
    {
        uint16_t off[4];  // pre-multiplied column offsets of the link's basis centres (as SelfUnit::a_off)
        float w[4];       // weights, sum 1; unused entries 0 (their offset repeats off[0])
    };

    struct alignas(8) SlotRef
    {
        uint16_t slot;  // sphere-centre slot written by the chain (or kConstSlot | j)
        uint16_t base;  // bit k set: add the robot base offset to coordinate k (Panda, robots/panda/fk.hh)
        float r;        // radius; < 0: the run-time radius of robots::sphere
    };

    // sphere_environment_in_collision(guard) { sphere_environment_in_collision(fine_i)... };
    // count == 0: the guard itself is the test ("leaf" group)
    // slack: every fine sphere lies within the guard sphere inflated by `slack` metres (measured over
    // sampled configurations at code-generation time, plus a 1e-4 m margin; the traced bounding spheres
    // are up to 0.6 mm too tight).  slack < 0: no such bound.
    // ubegin/ucount: the work units of the warp walk (vb_walk.cuh): refs[ubegin .. ubegin+ucount) are the
    // spheres to test once the guard may have fired (= the fine spheres; for a leaf group a copy of the
    // guard, appended behind the regular refs).
    struct alignas(8) RefGroup
    {
        SlotRef guard;
        uint16_t begin, count;
        float slack;
        uint16_t ubegin, ucount;
        uint32_t pad_;
    };

    // sphere_sphere_self_collision(ga, gb) { all pairs a_i x b_j }; a_count == 0: the guard is the test
    // slack_b: as above for the b-side fine spheres relative to the guard sphere gb
    // unit0: index of this group's first (group, a_i) work unit; dense: the guard fires for most
    // configurations (sampled at code-generation time), so the a-side filter runs lane-parallel
    struct alignas(16) SelfGroup
    {
        SlotRef ga, gb;
        uint16_t a_begin, a_count, b_begin, b_count;
        float slack_b;
        uint16_t unit0, dense;
    };

    // One (self group, a_i) work unit of the warp walk, flattened so that ONE 16-byte load gives the
    // filter and the pair stage everything they need.  Offsets are float indices from the warp's first
    // column: 3 * slot * kStride for a per-thread slot (the lane is added), >= 3 * kConstRow * kStride for a
    // per-warp constant (no lane term).
    struct alignas(16) SelfUnit
    {
        uint16_t a_off;    // column offset of a_i
        uint16_t gb_off;   // column offset of the other link's bounding sphere
        float a_r;         // radius of a_i
        float filter_r;    // a_r + r(gb) + slack_b, or < 0: no containment bound (always passes)
        uint16_t b_begin;  // the other link's fine spheres: brefs[b_begin .. b_begin + b_count()) (bws[] for a reduced link)
        uint16_t b_pack;   // bits 0-5: b count; bits 6-14: 1 + index of a DerivedRec whose off[] is the b-link's basis
                           // (0: every b_j is stored); bit 15: a_i is not stored, a_off is the index of its DerivedRec
        __host__ __device__ int b_count() const { return b_pack & 63; }
        __host__ __device__ int b_basis() const { return ((b_pack >> 6) & 511) - 1; }
        __host__ __device__ bool a_derived() const { return (b_pack >> 15) != 0; }
    };
    static_assert(sizeof(SelfUnit) == 16, "SelfUnit is read as one 16-byte load");

    // Queued a-side filter of one (self group, a_i) work unit (vb_walk.cuh: filter_units; Robot::kQueuedFilters): slot
    // codes as in SlotRef (kConstSlot / kDerivedSlot flags), thr2 = (a_r + r(gb) + slack_b)^2 or < 0: always passes.
    struct alignas(8) FilterRec
    {
        uint16_t a_slot, gb_slot;
        float thr2;
    };

    // b-side sphere of a reduced link: w[0..2] = weights over the basis EDGES e_k = c_k - c_0 of the link, w[3] =
    // w^T G w with G the (rigid-motion invariant) Gram matrix of the edges, + radius.  With u = a - c_0:
    //   |a - b|^2 = |u|^2 - 2 sum_k w_k (u . e_k) + w^T G w        (3 FMAs per b sphere once u . e_k are known)
    struct alignas(8) BWeights
    {
        float w[4];
        float r;
        uint32_t pad_;
    };

    // b-side sphere of a self group: pre-multiplied column offset + radius
    struct alignas(8) BRef
    {
        uint32_t off;
        float r;
    };

    template <int NE, int NS, int NA, int NR>
    struct Tables
    {
        int n_slots, n_env, n_self, n_att, att_env;
        int lift_z_row;  // row of the z coordinate the translated constants (kLiftSlot) add; -1: none
        RefGroup env[NE];
        SelfGroup self[NS];
        RefGroup att[NA];  // attachment_sphere_collision groups (robot sphere vs every attachment sphere)
        SlotRef refs[NR];
    };

#ifdef __CUDACC__
    // float index of a slot's x coordinate (y, z at +kStride each): `wcol` = first column of the warp, L = lane
    // whose configuration is addressed
    template <int kStride, int kConstRow, int kLiftRow = 0>
    __device__ __forceinline__ unsigned slot_index(unsigned wcol, int L, unsigned slot)
    {
        if (kLiftRow > 0 && (slot & kLiftSlot))
        {
            const unsigned j = slot & (kLiftSlot - 1u);
            return 3u * (kLiftRow + (j >> 5)) * kStride + wcol + (j & 31u);
        }
        const unsigned j = slot & (kConstSlot - 1u);
        return (slot & kConstSlot) ? 3u * (kConstRow + (j >> 5)) * kStride + wcol + (j & 31u) : 3u * slot * kStride + wcol + L;
    }

    // Table records are read with LDG through the non-coherent path in 8/16-byte pieces (a plain struct copy
    // through a pointer of unknown address space compiles to generic loads).
    template <class T>
    __device__ __forceinline__ T ldg_record(const T *p)
    {
        static_assert(sizeof(T) % 8 == 0 && alignof(T) >= 8, "table records are 8-byte multiples");
        T out;
        if constexpr (sizeof(T) % 16 == 0 && alignof(T) >= 16)
        {
            const uint4 *src = reinterpret_cast<const uint4 *>(p);
            uint4 *dst = reinterpret_cast<uint4 *>(&out);
#pragma unroll
            for (unsigned i = 0; i < sizeof(T) / 16; ++i)
            {
                dst[i] = __ldg(src + i);
            }
        }
        else
        {
            const uint2 *src = reinterpret_cast<const uint2 *>(p);
            uint2 *dst = reinterpret_cast<uint2 *>(&out);
#pragma unroll
            for (unsigned i = 0; i < sizeof(T) / 8; ++i)
            {
                dst[i] = __ldg(src + i);
            }
        }
        return out;
    }
#endif

    // Global-memory mirrors for lane-dependent indices (constant memory serialises those).
    struct WalkTables
    {
        const SlotRef *refs;        // refs + leaf-guard copies
        const uint16_t *ref_group;  // environment group of refs[i] (0xffff: not an environment unit)
        const RefGroup *env;
        const SelfGroup *self;
        const SelfUnit *units;
        const BRef *brefs;  // refs as (column offset, radius)
        const DerivedRec *ders;
        const BWeights *bws;  // refs of reduced links as (basis weights, radius)
        const FilterRec *filts;   // per work unit (queued filters)
        const uint32_t *gunits;   // per self group: first work unit | a_count << 16
    };
}  // namespace vb

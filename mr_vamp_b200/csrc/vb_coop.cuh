// Warp-cooperative collision checking of 32 configurations (one per lane).
//
// The reference returns at the first colliding sphere of an 8-lane block; a literal SIMT port of
// that control flow leaves ~9 of 32 lanes active (measured, profiles/r1_*).  Here every group of the
// robot's check tables is handled in two steps:
//
//   guard step  (lane-parallel, uniform): every lane tests ITS configuration's guard sphere against
//               all obstacles (or its guard pair for self collision).  No lane-dependent branches:
//               the obstacle loops stop only when all 32 lanes may prune (warp vote).
//   fine step   (warp-cooperative): lanes whose guard fired are served one after the other; the 32
//               lanes split that configuration's fine tests between them (lane k takes fine sphere
//               k) and read its sphere centres from its shared-memory column.  A ballot reduces the
//               result.  Obstacles are pre-filtered by "touches the guard sphere inflated by the
//               group's static slack" (vb_tables.cuh), which is conservative and exact.
//
// The verdict is the same boolean function as the reference's (valid iff no group has a guard hit
// and a fine hit, collision/validity.hh:46-150 + robots/<robot>/fk.hh); only the order of evaluation
// differs.
#pragma once
#include "vb_env.cuh"
#include "vb_tables.cuh"

namespace vb
{
    constexpr unsigned kFull = 0xffffffffu;

    struct PrimDist
    {
        float d2;    // squared distance term of the primitive test
        float objr;  // radius added to the sphere's radius (0 for boxes)
        float mind;  // min_distance of the obstacle (sorted key of the reference's pruning)
    };

    // type 0 sphere, 1 capsule, 2 z-capsule, 3 cuboid, 4 z-cuboid; p points at the record
    template <int T>
    VB_D PrimDist prim_dist(const float4 *p, float sx, float sy, float sz)
    {
        PrimDist o;
        if constexpr (T == 0)
        {
            const float4 a = p[0];
            const float dx = a.x - sx, dy = a.y - sy, dz = a.z - sz;
            o.d2 = dx * dx + dy * dy + dz * dz;
            o.objr = a.w;
            o.mind = p[1].x;
        }
        else if constexpr (T == 1)
        {
            const float4 a = p[0], b = p[1];
            const float dot = (sx - a.x) * b.x + (sy - a.y) * b.y + (sz - a.z) * b.z;
            const float cdf = fmaxf(fminf(dot * b.w, 1.f), 0.f);
            const float dx = sx - (a.x + b.x * cdf), dy = sy - (a.y + b.y * cdf), dz = sz - (a.z + b.z * cdf);
            o.d2 = dx * dx + dy * dy + dz * dz;
            o.objr = a.w;
            o.mind = p[2].x;
        }
        else if constexpr (T == 2)
        {
            const float4 a = p[0], b = p[1];
            const float dot = (sz - a.z) * b.x;
            const float cdf = fmaxf(fminf(dot * b.y, 1.f), 0.f);
            const float dx = sx - a.x, dy = sy - a.y, dz = sz - (a.z + b.x * cdf);
            o.d2 = dx * dx + dy * dy + dz * dz;
            o.objr = a.w;
            o.mind = b.z;
        }
        else if constexpr (T == 3)
        {
            const float4 c = p[0], a1 = p[1], a2 = p[2], a3 = p[3];
            const float xs = sx - c.x, ys = sy - c.y, zs = sz - c.z;
            const float d1 = fmaxf(fabsf(a1.x * xs + a1.y * ys + a1.z * zs) - a1.w, 0.f);
            const float d2 = fmaxf(fabsf(a2.x * xs + a2.y * ys + a2.z * zs) - a2.w, 0.f);
            const float d3 = fmaxf(fabsf(a3.x * xs + a3.y * ys + a3.z * zs) - a3.w, 0.f);
            o.d2 = d1 * d1 + d2 * d2 + d3 * d3;
            o.objr = 0.f;
            o.mind = c.w;
        }
        else
        {
            const float4 c = p[0], ax = p[1], hr = p[2];
            const float xs = sx - c.x, ys = sy - c.y, zs = sz - c.z;
            const float d1 = fmaxf(fabsf(ax.x * xs + ax.y * ys) - hr.x, 0.f);
            const float d2 = fmaxf(fabsf(ax.z * xs + ax.w * ys) - hr.y, 0.f);
            const float d3 = fmaxf(fabsf(zs) - hr.z, 0.f);
            o.d2 = d1 * d1 + d2 * d2 + d3 * d3;
            o.objr = 0.f;
            o.mind = c.w;
        }
        return o;
    }

    // same arithmetic as env_hit (vb_env.cuh): value = d2 - R*R with R = sr + objr (boxes: sr*sr)
    VB_D bool prim_collides(const PrimDist &d, float sr, float ext)
    {
        const float R = sr + d.objr;
        return !(d.mind >= ext) && (d.d2 - R * R < 0.f);
    }

    struct PrimIndex
    {
        int first[6];  // object index range of each type: [first[T], first[T+1])
        int quad[5];   // float4 offset of each type's first record
    };

    VB_D PrimIndex prim_index(const EnvView &e)
    {
        PrimIndex ix;
        const int n[5] = {e.n_spheres, e.n_capsules, e.n_zcapsules, e.n_cuboids, e.n_zcuboids};
        const int q[5] = {kSphereQ, kCapsuleQ, kZCapsQ, kCuboidQ, kZCubQ};
        int o = 0, f = 0;
#pragma unroll
        for (int t = 0; t < 5; ++t)
        {
            ix.first[t] = o;
            ix.quad[t] = f;
            o += n[t];
            f += n[t] * q[t];
        }
        ix.first[5] = o;
        return ix;
    }

    struct GuardResult
    {
        uint32_t mask;  // objects 0..31 touching the inflated guard sphere
        bool high;      // some object >= 32 touches it
        bool fired;     // the guard sphere itself collides (reference semantics)
    };

    // `ext_warp` is the largest (inflated) extent of the 32 guard spheres of this warp: the obstacle lists
    // are sorted by min_distance, so the first record at or beyond it ends the list for every lane, and
    // the loop bound is warp-uniform without a vote.
    template <int T, int Q>
    VB_D void guard_type(const float4 *P, const PrimIndex &ix, const Center &b, float ext, float ext_warp,
                         float slack, GuardResult &g)
    {
        const float4 *p = P + ix.quad[T];
        const float Rg = b.r + slack;
#pragma unroll 2
        for (int o = ix.first[T]; o < ix.first[T + 1]; ++o, p += Q)
        {
            const PrimDist d = prim_dist<T>(p, b.x, b.y, b.z);
            if (d.mind >= ext_warp)
            {
                break;
            }
            g.fired = g.fired || prim_collides(d, b.r, ext);
            const float Ri = Rg + d.objr;
            const bool touch = d.d2 - Ri * Ri < 0.f;
            if (o < 32)
            {
                g.mask |= touch ? (1u << o) : 0u;
            }
            else
            {
                g.high = g.high || touch;
            }
        }
    }

    // lane-parallel: this lane's guard sphere against every obstacle primitive
    VB_D GuardResult guard_env(const EnvView &e, const float4 *P, const PrimIndex &ix, const Center &b, float slack)
    {
        GuardResult g{0u, false, false};
        const float ext = sqrtf(b.x * b.x + b.y * b.y + b.z * b.z) + b.r;
        const float s = slack < 0.f ? 0.f : slack;
        // non-negative floats order like their bit patterns; NaN/inf extents disable the cut
        const float ext_lane = ext + s;
        const unsigned bits = (ext_lane >= 0.f) ? __float_as_uint(ext_lane) : 0x7f800000u;
        const float ext_warp = __uint_as_float(__reduce_max_sync(kFull, bits));
        guard_type<0, kSphereQ>(P, ix, b, ext, ext_warp, s, g);
        guard_type<1, kCapsuleQ>(P, ix, b, ext, ext_warp, s, g);
        guard_type<2, kZCapsQ>(P, ix, b, ext, ext_warp, s, g);
        guard_type<3, kCuboidQ>(P, ix, b, ext, ext_warp, s, g);
        guard_type<4, kZCubQ>(P, ix, b, ext, ext_warp, s, g);
        if (slack < 0.f)
        {
            // no containment bound for this group: every obstacle is a candidate for the fine tests
            g.mask = 0xffffffffu;
            g.high = true;
        }
        return g;
    }

    // one fine sphere against obstacle o (uniform o across the warp)
    VB_D bool fine_vs_object(const float4 *P, const PrimIndex &ix, int o, const Center &s, float ext)
    {
        if (o < ix.first[1])
        {
            return prim_collides(prim_dist<0>(P + ix.quad[0] + (o - ix.first[0]) * kSphereQ, s.x, s.y, s.z), s.r, ext);
        }
        if (o < ix.first[2])
        {
            return prim_collides(prim_dist<1>(P + ix.quad[1] + (o - ix.first[1]) * kCapsuleQ, s.x, s.y, s.z), s.r, ext);
        }
        if (o < ix.first[3])
        {
            return prim_collides(prim_dist<2>(P + ix.quad[2] + (o - ix.first[2]) * kZCapsQ, s.x, s.y, s.z), s.r, ext);
        }
        if (o < ix.first[4])
        {
            return prim_collides(prim_dist<3>(P + ix.quad[3] + (o - ix.first[3]) * kCuboidQ, s.x, s.y, s.z), s.r, ext);
        }
        return prim_collides(prim_dist<4>(P + ix.quad[4] + (o - ix.first[4]) * kZCubQ, s.x, s.y, s.z), s.r, ext);
    }
}  // namespace vb

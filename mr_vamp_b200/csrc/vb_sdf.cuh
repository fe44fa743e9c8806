// Distance / candidate grid of the primitive obstacles: an L2-resident acceleration structure for
// the per-sphere environment query (collision/validity.hh:46-150).
//
// Per voxel v (centre p_v):
//   d[v]    = min over the spheres, capsules and cuboids of the environment of the SIGNED distance from
//             p_v to the obstacle's surface.  Signed distance is 1-Lipschitz, so for a query sphere
//             (c, r) whose centre falls in voxel v
//                 d[v] - pad > r   =>  no primitive touches it              ("free")
//                 d[v] + pad < r   =>  some primitive certainly overlaps it ("certain hit")
//             with pad = half the voxel diagonal + a 1e-4 m safety margin.
//   mask[v] = bit o set iff obstacle o (sorted order, o < 31) comes within kMaskReach + pad of p_v; bit 31
//             stands for "some obstacle with index >= 31 does".  Every obstacle that can touch ANY sphere
//             lying inside a ball of radius kMaskReach - pad around a point of the voxel is in the mask.
// The guard phase of the warp walk fetches one cell per link bounding sphere; the fine spheres of a
// link are then tested with the exact primitive arithmetic of the reference, but only against the
// candidates of the mask.  The grid never changes a verdict: it only removes tests whose outcome is
// certain.  148 SMs share one copy in B200's L2 (<= 32 MB per environment).
#pragma once
#include "vb_env.cuh"

namespace vb
{
    constexpr float kSdfCover = 0.25f;   // the grid extends this far beyond the obstacles' bounding box
    constexpr float kSdfMargin = 1e-4f;  // safety margin on both classification thresholds
    constexpr float kMaskReach = 0.2f;   // candidate masks cover spheres within this distance of the voxel

    // Voxel of (x, y, z) relative to the grid's centre voxel (vb_env.cuh: SdfView); `in` = the point lies inside the grid.
    // 3 FMA + 3 compares + 3 conversions: the guard phase does one such lookup per link and configuration.
    VB_D void grid_coords(const SdfView &g, float x, float y, float z, int &ix, int &iy, int &iz, bool &in)
    {
        const float gx = fmaf(x, g.inv_h, g.cx), gy = fmaf(y, g.inv_h, g.cy), gz = fmaf(z, g.inv_h, g.cz);
        in = fabsf(gx) < g.hx && fabsf(gy) < g.hy && fabsf(gz) < g.hz;
        ix = __float2int_rd(gx);
        iy = __float2int_rd(gy);
        iz = __float2int_rd(gz);
    }
    // One level: the cell of (x, y, z), loaded only when the point is inside (a straight-line sequence of fetches keeps
    // all its loads in flight at once).
    VB_D uint2 grid_fetch(const SdfView &g, float x, float y, float z, bool &in)
    {
        int ix, iy, iz;
        grid_coords(g, x, y, z, ix, iy, iz, in);
        // (outside: the centre voxel is loaded and ignored -- a select instead of a branch around the load)
        return __ldg(g.centre + (in ? (iz * g.ny + iy) * g.nx + ix : 0));
    }
    // Two levels over the same box: the coarse cell is fetched with grid_fetch(coarse, ...); a guard it cannot call free
    // then fetches its fine cell.  `in` comes from the coarse lookup (same box; the voxel is clamped, because the two
    // roundings may disagree on a point within an ulp of the box's faces).
    VB_D uint2 grid_fetch_inside(const SdfView &g, float x, float y, float z, bool in)
    {
        const float gx = fmaf(x, g.inv_h, g.cx), gy = fmaf(y, g.inv_h, g.cy), gz = fmaf(z, g.inv_h, g.cz);
        const int hx = g.nx >> 1, hy = g.ny >> 1, hz = g.nz >> 1;
        const int ix = min(max(__float2int_rd(gx), -hx), hx - 1), iy = min(max(__float2int_rd(gy), -hy), hy - 1),
                  iz = min(max(__float2int_rd(gz), -hz), hz - 1);
        return __ldg(g.centre + (in ? (iz * g.ny + iy) * g.nx + ix : 0));
    }

    // coarse level of the two-level lookup: true iff the coarse cell alone proves that no primitive touches a sphere
    // of radius r centred in it (same test as grid_class's `free`, with the coarse grid's pad)
    VB_D bool grid_coarse_free(const SdfView &g, uint2 cell, bool in, float r)
    {
        return in ? (__uint_as_float(cell.x) - g.pad > r) : (r < g.cover);
    }

    // Classification of a sphere of radius r whose cell was fetched with grid_fetch; `reach` = radius of
    // the ball around the sphere centre that contains every sphere that will be tested against `mask`
    // (< 0: unknown).  Outside the grid every obstacle is at least `cover` away (cover == 0: unknown).
    VB_D void grid_class(const SdfView &g, uint2 cell, bool in, float r, float reach, bool &maybe, bool &cert, uint32_t &mask)
    {
        const float d = __uint_as_float(cell.x);
        const bool free_ = in ? (d - g.pad > r) : (r < g.cover);
        maybe = !free_;
        cert = in && (d + g.pad < r);
        const bool covered = reach >= 0.f && reach + g.pad < g.mask_reach;
        mask = !covered ? 0xffffffffu : (in ? cell.y : (reach < g.cover ? 0u : 0xffffffffu));
    }

    // signed distance from p to every obstacle record of `e` (same packing as env_hit): f(o, distance)
    template <class F>
    VB_D void sdf_objects(const EnvView &e, float px, float py, float pz, F f)
    {
        const float4 *p = e.prims;
        int o = 0;
        for (int i = 0; i < e.n_spheres; ++i, p += kSphereQ, ++o)
        {
            const float4 a = __ldg(p);
            const float dx = px - a.x, dy = py - a.y, dz = pz - a.z;
            f(o, sqrtf(dx * dx + dy * dy + dz * dz) - a.w);
        }
        for (int i = 0; i < e.n_capsules; ++i, p += kCapsuleQ, ++o)
        {
            const float4 a = __ldg(p), b = __ldg(p + 1);
            const float dot = (px - a.x) * b.x + (py - a.y) * b.y + (pz - a.z) * b.z;
            const float t = fmaxf(fminf(dot * b.w, 1.f), 0.f);
            const float dx = px - (a.x + b.x * t), dy = py - (a.y + b.y * t), dz = pz - (a.z + b.z * t);
            f(o, sqrtf(dx * dx + dy * dy + dz * dz) - a.w);
        }
        for (int i = 0; i < e.n_zcapsules; ++i, p += kZCapsQ, ++o)
        {
            const float4 a = __ldg(p), b = __ldg(p + 1);
            const float t = fmaxf(fminf((pz - a.z) * b.x * b.y, 1.f), 0.f);
            const float dx = px - a.x, dy = py - a.y, dz = pz - (a.z + b.x * t);
            f(o, sqrtf(dx * dx + dy * dy + dz * dz) - a.w);
        }
        for (int i = 0; i < e.n_cuboids; ++i, p += kCuboidQ, ++o)
        {
            const float4 c = __ldg(p), a1 = __ldg(p + 1), a2 = __ldg(p + 2), a3 = __ldg(p + 3);
            const float xs = px - c.x, ys = py - c.y, zs = pz - c.z;
            const float q1 = fabsf(a1.x * xs + a1.y * ys + a1.z * zs) - a1.w;
            const float q2 = fabsf(a2.x * xs + a2.y * ys + a2.z * zs) - a2.w;
            const float q3 = fabsf(a3.x * xs + a3.y * ys + a3.z * zs) - a3.w;
            const float o1 = fmaxf(q1, 0.f), o2 = fmaxf(q2, 0.f), o3 = fmaxf(q3, 0.f);
            f(o, sqrtf(o1 * o1 + o2 * o2 + o3 * o3) + fminf(fmaxf(q1, fmaxf(q2, q3)), 0.f));
        }
        for (int i = 0; i < e.n_zcuboids; ++i, p += kZCubQ, ++o)
        {
            const float4 c = __ldg(p), ax = __ldg(p + 1), hr = __ldg(p + 2);
            const float xs = px - c.x, ys = py - c.y, zs = pz - c.z;
            const float q1 = fabsf(ax.x * xs + ax.y * ys) - hr.x;
            const float q2 = fabsf(ax.z * xs + ax.w * ys) - hr.y;
            const float q3 = fabsf(zs) - hr.z;
            const float o1 = fmaxf(q1, 0.f), o2 = fmaxf(q2, 0.f), o3 = fmaxf(q3, 0.f);
            f(o, sqrtf(o1 * o1 + o2 * o2 + o3 * o3) + fminf(fmaxf(q1, fmaxf(q2, q3)), 0.f));
        }
    }

    // ---- exact test of one sphere against ONE obstacle (index o in the sorted order) ------------------
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

    // Squared-distance term d2 and added radius objr of the reference primitive for obstacle o, so that
    // its value is d2 - (sr + objr)^2 (same arithmetic as env_hit, vb_env.cuh; objr = 0 for boxes).
    struct ObjDist
    {
        float d2, objr;
    };

    VB_D ObjDist object_dist(const float4 *P, const PrimIndex &ix, int o, float sx, float sy, float sz)
    {
        ObjDist r;
        if (o >= ix.first[4])
        {
            const float4 *p = P + ix.quad[4] + (o - ix.first[4]) * kZCubQ;
            const float4 c = __ldg(p + 0), ax = __ldg(p + 1), hr = __ldg(p + 2);
            const float xs = sx - c.x, ys = sy - c.y, zs = sz - c.z;
            const float d1 = fmaxf(fabsf(ax.x * xs + ax.y * ys) - hr.x, 0.f);
            const float d2 = fmaxf(fabsf(ax.z * xs + ax.w * ys) - hr.y, 0.f);
            const float d3 = fmaxf(fabsf(zs) - hr.z, 0.f);
            r.d2 = d1 * d1 + d2 * d2 + d3 * d3;
            r.objr = 0.f;
        }
        else if (o >= ix.first[3])
        {
            const float4 *p = P + ix.quad[3] + (o - ix.first[3]) * kCuboidQ;
            const float4 c = __ldg(p + 0), a1 = __ldg(p + 1), a2 = __ldg(p + 2), a3 = __ldg(p + 3);
            const float xs = sx - c.x, ys = sy - c.y, zs = sz - c.z;
            const float d1 = fmaxf(fabsf(a1.x * xs + a1.y * ys + a1.z * zs) - a1.w, 0.f);
            const float d2 = fmaxf(fabsf(a2.x * xs + a2.y * ys + a2.z * zs) - a2.w, 0.f);
            const float d3 = fmaxf(fabsf(a3.x * xs + a3.y * ys + a3.z * zs) - a3.w, 0.f);
            r.d2 = d1 * d1 + d2 * d2 + d3 * d3;
            r.objr = 0.f;
        }
        else if (o >= ix.first[2])
        {
            const float4 *p = P + ix.quad[2] + (o - ix.first[2]) * kZCapsQ;
            const float4 a = __ldg(p + 0), b = __ldg(p + 1);
            const float dot = (sz - a.z) * b.x;
            const float cdf = fmaxf(fminf(dot * b.y, 1.f), 0.f);
            const float dx = sx - a.x, dy = sy - a.y, dz = sz - (a.z + b.x * cdf);
            r.d2 = dx * dx + dy * dy + dz * dz;
            r.objr = a.w;
        }
        else if (o >= ix.first[1])
        {
            const float4 *p = P + ix.quad[1] + (o - ix.first[1]) * kCapsuleQ;
            const float4 a = __ldg(p + 0), b = __ldg(p + 1);
            const float dot = (sx - a.x) * b.x + (sy - a.y) * b.y + (sz - a.z) * b.z;
            const float cdf = fmaxf(fminf(dot * b.w, 1.f), 0.f);
            const float dx = sx - (a.x + b.x * cdf), dy = sy - (a.y + b.y * cdf), dz = sz - (a.z + b.z * cdf);
            r.d2 = dx * dx + dy * dy + dz * dz;
            r.objr = a.w;
        }
        else
        {
            const float4 a = __ldg(P + ix.quad[0] + o * kSphereQ);
            const float dx = a.x - sx, dy = a.y - sy, dz = a.z - sz;
            r.d2 = dx * dx + dy * dy + dz * dz;
            r.objr = a.w;
        }
        return r;
    }

    VB_D bool object_hit(const ObjDist &d, float sr)
    {
        const float rs = sr + d.objr;
        return d.d2 - rs * rs < 0.f;
    }

    // next candidate of a mask walk: bits 0..30 are obstacles 0..30, bit 31 stands for every obstacle with
    // index >= 31.  `cursor` starts at 0; returns -1 when the candidates are exhausted.
    VB_D int next_candidate(uint32_t &low, bool high, int n_obj, int &cursor)
    {
        if (low)
        {
            const int o = __ffs(low) - 1;
            low &= low - 1u;
            return o;
        }
        if (high)
        {
            const int o = 31 + cursor++;
            return o < n_obj ? o : -1;
        }
        return -1;
    }
}  // namespace vb

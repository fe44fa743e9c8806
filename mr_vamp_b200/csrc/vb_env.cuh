// Device-side environment: packed obstacle records + the per-sphere environment query.
//
// Replaces collision::Environment<FloatVector<rake>> (collision/environment.hh:11-82), the shape
// structs (collision/shapes.hh) and sphere_environment_in_collision (collision/validity.hh:46-150).
// Obstacles are kept in the reference's order (each list sorted by min_distance,
// environment.hh:40-66) as float4 records so that one warp-uniform LDS.128/LDG.128 fetches a
// quarter of a box; every lane of a warp walks the same record at the same time, so the loads
// are broadcasts.
#pragma once
// inlining policy of the two large device functions (A/B switchable at build time)
#ifndef VB_CHAIN_ATTR
#define VB_CHAIN_ATTR __forceinline__
#endif
#ifndef VB_ENVHIT_ATTR
#define VB_ENVHIT_ATTR __noinline__
#endif
#include <type_traits>

#include "vb_numerics.cuh"

namespace vb
{
    // record sizes in float4 units
    constexpr int kSphereQ = 2;   // [x y z r] [min_d 0 0 0]
    constexpr int kCapsuleQ = 3;  // [x1 y1 z1 r] [xv yv zv rdv] [min_d 0 0 0]
    constexpr int kZCapsQ = 2;    // [x1 y1 z1 r] [zv rdv min_d 0]
    constexpr int kCuboidQ = 4;   // [c.xyz min_d] [ax1 r1] [ax2 r2] [ax3 r3]
    constexpr int kZCubQ = 3;     // [c.xyz min_d] [a1x a1y a2x a2y] [r1 r2 r3 0]

    struct CaptView
    {
        const float *tests;          // (2^nlog2 - 1) split values (capt.hh:588-590)
        const uint32_t *aff_starts;  // 2^nlog2 + 1
        const float4 *aff;           // per 8-point affordance vector: 6 float4 = x[0:8] y[0:8] z[0:8]
        const float *aabbs;          // [2^nlog2][6]
        float top[6];                // aabb_top lower(3) upper(3)
        float r_point;
        uint32_t n_tests;
        int nlog2;
        // occupancy bitmap over the cloud (built at vb_env_commit, vb_capi.cu: capt_occupancy_kernel): bit of voxel v set
        // iff some point of the cloud lies within occ_reach + half a voxel diagonal of v's centre.  A query sphere
        // whose voxel is clear (or that lies outside the bitmap's box) with r + r_point <= occ_reach cannot reach any
        // point, so collides_simd -- which answers true only for a point of the cloud in range -- is false without the
        // descent.  nullptr: none.
        const uint32_t *occ;
        float occ_o[3], occ_inv_h, occ_reach;
        int occ_n[3];
        // fine level: per voxel the distance from its centre to the nearest point of the cloud, capped at dist_cap
        // (capt_distance_kernel).  d - dist_pad > r + r_point proves that no point is in range (distance to a point set is
        // 1-Lipschitz; dist_pad = half a voxel diagonal + margin), which spares the leaf scan of the many spheres that
        // pass close to a surface without touching it.  Only ever used to answer "false"; a hit is always found by the
        // reference's own scan.  nullptr: none.
        const float *dist;
        float dist_o[3], dist_inv_h, dist_cap, dist_pad;
        int dist_n[3];
    };

    struct HeightFieldView
    {
        float x, y, z, xs, ys, zs;  // offset, scale (shapes.hh:249-262)
        float xd2, yd2;             // half image size, already float(size_t(xd / 2))
        float xdf, ydf;             // float(xd), float(yd)
        int xd;
        const float *data;
    };

    struct AttachmentView
    {
        float tf[7];          // tx ty tz rx ry rz rw (attachments.hh:67-73)
        int n;                // number of attachment spheres
        const float4 *local;  // [n] x y z r in the end-effector frame
    };

    constexpr int kMaxAttach = 16;

    // Distance / candidate grid over the primitive obstacles (vb_sdf.cuh); cell == nullptr: no grid.
    struct SdfView
    {
        const uint2 *cell;     // [nz][ny][nx]: x = float bits of the signed distance at the voxel centre, y = candidate mask
        float ox, oy, oz;      // lower corner
        float inv_h;           // 1 / voxel edge
        float fnx, fny, fnz;   // the largest floats below nx, ny, nz (voxel coordinates are clamped to [0, fn*])
        int nx, ny, nz;
        float pad;             // half voxel diagonal + safety margin
        float cover;           // outside the grid every obstacle is at least this far away (0: unknown)
        float mask_reach;      // candidate masks hold every obstacle within mask_reach + pad of the voxel centre
        // Lookup in CENTRED voxel coordinates g = x * inv_h + c (one FMA; c = -o * inv_h - n / 2; n is even): inside the
        // grid iff |g| < n / 2 on every axis (one compare with |.|; NaN is outside), voxel = floor(g) relative to the
        // grid's centre voxel `centre` = cell + ((nz/2) * ny + ny/2) * nx + nx/2.
        float cx, cy, cz;
        float hx, hy, hz;
        const uint2 *centre;
    };
    // the coarse level of the two-level lookup has voxels of 2^kCoarseShift fine voxels on the same lattice (the fine
    // dimensions are multiples of 2 << kCoarseShift): its voxel index is the fine one shifted
    constexpr int kCoarseShift = 3;

    // Passed by value to every kernel (lives in the parameter constant bank).
    struct EnvView
    {
        const float4 *prims;  // packed records: spheres | capsules | z-capsules | cuboids | z-cuboids
        int n_quads;          // total float4 count of `prims`
        int n_spheres, n_capsules, n_zcapsules, n_cuboids, n_zcuboids;
        int n_heightfields, n_capts;
        int capt_heavy;  // some point cloud averages more than 16 affordance vectors per leaf (long leaf scans)
        const HeightFieldView *heightfields;
        const CaptView *capts;
        AttachmentView attach;
        SdfView sdf;
        SdfView sdf_c;  // coarse level of the two-level guard lookup (cell == nullptr: none)
    };

    // Margin bookkeeping: smallest |signed clearance| (metres) over every test that decided control
    // flow.  Used by the parity reports to split disagreements into "within 1e-5 m of contact" and
    // real mismatches (BASELINE north_star).
    struct NoMargin
    {
        VB_D void sq(float, float) {}
        VB_D void lin(float) {}
        VB_D void count(int) {}
    };
    // Algorithmic-flop bookkeeping (DESIGN.md "W"): 9 flops per environment query + (1 + C_type) per
    // visited obstacle, 11 per sphere-sphere test (SURVEY.md 8(a) cost constants), counted along the
    // reference-ordered lane-serial walk.  `m` is reported through the margin output.
    struct WorkCount
    {
        float m = 0.f;
        VB_D void sq(float, float) {}
        VB_D void lin(float) {}
        VB_D void count(int flops) { m += static_cast<float>(flops); }
    };
    struct MinMargin
    {
        float m = 3.0e38f;
        // value = d^2 - R^2  ->  |d - R|
        VB_D void sq(float value, float R)
        {
            const float d = sqrtf(fmaxf(value + R * R, 0.f));
            m = fminf(m, fabsf(d - R));
        }
        VB_D void lin(float value) { m = fminf(m, fabsf(value)); }
        VB_D void count(int) {}
    };

    // First part of collides_simd (capt.hh:457-513): top box, [the library's "no point in range" grids], descent, leaf box.
    // false: the query is answered "no collision"; true: the affordance vectors [v0, v1) of the leaf must be scanned
    // (capt_scan_warp).  Verdict walks only.
    VB_D bool capt_probe(const CaptView &c, float x, float y, float z, float r, uint32_t &v0, uint32_t &v1)
    {
        const float ctr[3] = {x, y, z};
        bool inb = true;
#pragma unroll
        for (int k = 0; k < 3; ++k)
        {
            inb = inb && (ctr[k] + r >= c.top[k]) && (ctr[k] - r <= c.top[3 + k]);
        }
        if (!inb)
        {
            return false;
        }
        if (c.occ != nullptr && r + c.r_point <= c.occ_reach)
        {
            const float fx = (x - c.occ_o[0]) * c.occ_inv_h, fy = (y - c.occ_o[1]) * c.occ_inv_h, fz = (z - c.occ_o[2]) * c.occ_inv_h;
            const bool inside = fx >= 0.f && fy >= 0.f && fz >= 0.f && fx < static_cast<float>(c.occ_n[0]) &&
                                fy < static_cast<float>(c.occ_n[1]) && fz < static_cast<float>(c.occ_n[2]);
            if (!inside)
            {
                return false;
            }
            const unsigned v = (static_cast<unsigned>(fz) * c.occ_n[1] + static_cast<unsigned>(fy)) * c.occ_n[0] + static_cast<unsigned>(fx);
            if (!((__ldg(c.occ + (v >> 5)) >> (v & 31u)) & 1u))
            {
                return false;
            }
        }
        if (c.dist != nullptr && r + c.r_point < c.dist_cap - c.dist_pad)
        {
            const float fx = (x - c.dist_o[0]) * c.dist_inv_h, fy = (y - c.dist_o[1]) * c.dist_inv_h, fz = (z - c.dist_o[2]) * c.dist_inv_h;
            const bool inside = fx >= 0.f && fy >= 0.f && fz >= 0.f && fx < static_cast<float>(c.dist_n[0]) &&
                                fy < static_cast<float>(c.dist_n[1]) && fz < static_cast<float>(c.dist_n[2]);
            if (!inside)
            {
                return false;
            }
            const size_t v = (static_cast<size_t>(fz) * c.dist_n[1] + static_cast<size_t>(fy)) * c.dist_n[0] + static_cast<size_t>(fx);
            if (__ldg(c.dist + v) - c.dist_pad > r + c.r_point)
            {
                return false;
            }
        }
        uint32_t idx = 0;
        int k = 0;
        for (int i = 0; i < c.nlog2; ++i)
        {
            idx = 2u * idx + 1u + (ctr[k] >= __ldg(c.tests + idx) ? 1u : 0u);
            k = (k == 2) ? 0 : k + 1;
        }
        const uint32_t leaf = idx - c.n_tests;
        const float R = r + c.r_point;
        const float rsq = R * R;
        const float *bb = c.aabbs + 6ull * leaf;
        float dsq = 0.f;
#pragma unroll
        for (int a = 0; a < 3; ++a)
        {
            const float lo = __ldg(bb + a), hi = __ldg(bb + 3 + a);
            const float cl = fminf(fmaxf(ctr[a], lo), hi);
            const float d = ctr[a] - cl;
            dsq += d * d;
        }
        if (!(dsq <= rsq))
        {
            return false;
        }
        v0 = __ldg(c.aff_starts + leaf);
        v1 = __ldg(c.aff_starts + leaf + 1);
        return v1 > v0;
    }

    // Second part (capt.hh:516-538) for ONE query, by the whole warp: lane l tests the 8 points of vector v0 + l,
    // v0 + l + 32, ... (one coalesced 96-byte read each); true iff some point of the leaf's list is in range -- the same
    // predicate as the reference's serial scan, whose order does not matter.  All 32 lanes call with the same arguments.
    VB_D bool capt_scan_warp(const CaptView &c, uint32_t v0, uint32_t v1, float x, float y, float z, float r)
    {
        const float R = r + c.r_point, rsq = R * R;
        const uint32_t lane = threadIdx.x & 31u;
        for (uint32_t base = v0; base < v1; base += 32u)
        {
            const uint32_t i = base + lane;
            bool hit = false;
            if (i < v1)
            {
                const float4 *v = c.aff + 6ull * i;
                const float4 x0 = __ldg(v), x1 = __ldg(v + 1), y0 = __ldg(v + 2), y1 = __ldg(v + 3), z0 = __ldg(v + 4), z1 = __ldg(v + 5);
                const float px[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
                const float py[8] = {y0.x, y0.y, y0.z, y0.w, y1.x, y1.y, y1.z, y1.w};
                const float pz[8] = {z0.x, z0.y, z0.z, z0.w, z1.x, z1.y, z1.z, z1.w};
#pragma unroll
                for (int j = 0; j < 8; ++j)
                {
                    const float dx = px[j] - x, dy = py[j] - y, dz = pz[j] - z;
                    hit = hit || (dx * dx + dy * dy + dz * dz <= rsq);
                }
            }
            if (__any_sync(0xffffffffu, hit))
            {
                return true;
            }
        }
        return false;
    }

    // collision/capt.hh:457-541 (collides_simd with one query per lane)
    template <class M>
    VB_D bool capt_hit(const CaptView &c, float x, float y, float z, float r, M &mg)
    {
        const float ctr[3] = {x, y, z};
        bool inb = true;
#pragma unroll
        for (int k = 0; k < 3; ++k)
        {
            inb = inb && (ctr[k] + r >= c.top[k]) && (ctr[k] - r <= c.top[3 + k]);
        }
        if (!inb)
        {
            return false;
        }
        if constexpr (std::is_same<M, NoMargin>::value)
        {
            // (verdict walks only: the margin / work-count walks execute every test of the reference)
            if (c.occ != nullptr && r + c.r_point <= c.occ_reach)
            {
                const float fx = (x - c.occ_o[0]) * c.occ_inv_h, fy = (y - c.occ_o[1]) * c.occ_inv_h, fz = (z - c.occ_o[2]) * c.occ_inv_h;
                const bool inside = fx >= 0.f && fy >= 0.f && fz >= 0.f && fx < static_cast<float>(c.occ_n[0]) &&
                                    fy < static_cast<float>(c.occ_n[1]) && fz < static_cast<float>(c.occ_n[2]);
                if (!inside)
                {
                    return false;  // the box extends occ_reach + a voxel beyond the cloud
                }
                const unsigned v = (static_cast<unsigned>(fz) * c.occ_n[1] + static_cast<unsigned>(fy)) * c.occ_n[0] + static_cast<unsigned>(fx);
                if (!((__ldg(c.occ + (v >> 5)) >> (v & 31u)) & 1u))
                {
                    return false;
                }
            }
            if (c.dist != nullptr && r + c.r_point < c.dist_cap - c.dist_pad)
            {
                const float fx = (x - c.dist_o[0]) * c.dist_inv_h, fy = (y - c.dist_o[1]) * c.dist_inv_h, fz = (z - c.dist_o[2]) * c.dist_inv_h;
                const bool inside = fx >= 0.f && fy >= 0.f && fz >= 0.f && fx < static_cast<float>(c.dist_n[0]) &&
                                    fy < static_cast<float>(c.dist_n[1]) && fz < static_cast<float>(c.dist_n[2]);
                if (!inside)
                {
                    return false;  // the box extends dist_cap + a voxel beyond the cloud
                }
                const size_t v = (static_cast<size_t>(fz) * c.dist_n[1] + static_cast<size_t>(fy)) * c.dist_n[0] + static_cast<size_t>(fx);
                if (__ldg(c.dist + v) - c.dist_pad > r + c.r_point)
                {
                    return false;
                }
            }
        }
        uint32_t idx = 0;
        int k = 0;
        for (int i = 0; i < c.nlog2; ++i)
        {
            idx = 2u * idx + 1u + (ctr[k] >= __ldg(c.tests + idx) ? 1u : 0u);
            k = (k == 2) ? 0 : k + 1;
        }
        const uint32_t leaf = idx - c.n_tests;
        const float R = r + c.r_point;
        const float rsq = R * R;
        const float *bb = c.aabbs + 6ull * leaf;
        float dsq = 0.f;
#pragma unroll
        for (int a = 0; a < 3; ++a)
        {
            const float lo = __ldg(bb + a), hi = __ldg(bb + 3 + a);
            const float cl = fminf(fmaxf(ctr[a], lo), hi);
            const float d = ctr[a] - cl;
            dsq += d * d;
        }
        if (!(dsq <= rsq))
        {
            return false;
        }
        const uint32_t s = __ldg(c.aff_starts + leaf), e = __ldg(c.aff_starts + leaf + 1);
        for (uint32_t i = s; i < e; ++i)
        {
            const float4 *v = c.aff + 6ull * i;
            const float4 x0 = __ldg(v), x1 = __ldg(v + 1), y0 = __ldg(v + 2), y1 = __ldg(v + 3), z0 = __ldg(v + 4),
                         z1 = __ldg(v + 5);
            const float px[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
            const float py[8] = {y0.x, y0.y, y0.z, y0.w, y1.x, y1.y, y1.z, y1.w};
            const float pz[8] = {z0.x, z0.y, z0.z, z0.w, z1.x, z1.y, z1.z, z1.w};
            bool hit = false;
#pragma unroll
            for (int j = 0; j < 8; ++j)
            {
                const float dx = px[j] - x, dy = py[j] - y, dz = pz[j] - z;
                const float d2 = dx * dx + dy * dy + dz * dz;
                mg.sq(d2 - rsq, R);
                hit = hit || (d2 <= rsq);
            }
            if (hit)
            {
                return true;
            }
        }
        return false;
    }

    // collision/sphere_heightfield.hh:8-30
    VB_D float heightfield_value(const HeightFieldView &h, float x, float y, float z, float r)
    {
        const float xo = h.x - x, yo = h.y - y;
        const float xi = floorf(fminf(fmaxf(h.xs * xo + h.xd2, 0.f), h.xdf));
        const float yi = floorf(fminf(fmaxf(h.ys * yo + h.yd2, 0.f), h.ydf));
        const int index = static_cast<int>(yi * static_cast<float>(h.xd) + xi);
        const float zh = __ldg(h.data + index);
        return z - r - (h.zs * zh + h.z);
    }

    // the non-primitive obstacle kinds of sphere_environment_in_collision (validity.hh:132-147):
    // heightfields, then CAPT point clouds
    template <class M, bool kCapts = true>
    VB_D bool other_hit(const EnvView &e, float sx, float sy, float sz, float sr, M &mg)
    {
        for (int i = 0; i < e.n_heightfields; ++i)
        {
            const float v = heightfield_value(e.heightfields[i], sx, sy, sz, sr);
            mg.count(18);
            mg.lin(v);
            if (v < 0.f)
            {
                return true;
            }
        }
        if constexpr (kCapts)
        {
            for (int i = 0; i < e.n_capts; ++i)
            {
                if (capt_hit(e.capts[i], sx, sy, sz, sr, mg))
                {
                    return true;
                }
            }
        }
        return false;
    }

    struct Center
    {
        float x, y, z, r;
    };

    // collision/validity.hh:46-150.  `P` points at the packed records in global memory (read through the
    // non-coherent path: a few hundred bytes shared by every thread, L1-resident).
    // Each loop stops at the first record whose min_distance is >= the sphere's largest distance
    // from the origin, as the reference does; with an exact sqrt the cut is conservative, so the
    // verdict equals the unpruned OR over all obstacles.
    // kCapts == false: primitives and height fields only (the warp walk queries point clouds cooperatively: env_hit_warp)
    template <class M, bool kCapts = true>
    __device__ VB_ENVHIT_ATTR bool env_hit(const EnvView &e, const float4 *P, float sx, float sy, float sz, float sr, M &mg)
    {
        const float ext = sqrtf(sx * sx + sy * sy + sz * sz) + sr;
        mg.count(9);
        const float4 *p = P;
        for (int i = 0; i < e.n_spheres; ++i, p += kSphereQ)
        {
            const float4 a = __ldg(p + 0);
            if (__ldg(p + 1).x >= ext)
            {
                break;
            }
            const float dx = a.x - sx, dy = a.y - sy, dz = a.z - sz, rs = a.w + sr;
            const float v = dx * dx + dy * dy + dz * dz - rs * rs;
            mg.count(1 + 11);
            mg.sq(v, rs);
            if (v < 0.f)
            {
                return true;
            }
        }
        p = P + kSphereQ * e.n_spheres;
        for (int i = 0; i < e.n_capsules; ++i, p += kCapsuleQ)
        {
            const float4 a = __ldg(p + 0), b = __ldg(p + 1);
            if (__ldg(p + 2).x >= ext)
            {
                break;
            }
            // sphere_capsule.hh:8-22
            mg.count(1 + 28);
            const float dot = (sx - a.x) * b.x + (sy - a.y) * b.y + (sz - a.z) * b.z;
            const float cdf = fmaxf(fminf(dot * b.w, 1.f), 0.f);
            const float dx = sx - (a.x + b.x * cdf), dy = sy - (a.y + b.y * cdf), dz = sz - (a.z + b.z * cdf);
            const float rs = sr + a.w;
            const float v = dx * dx + dy * dy + dz * dz - rs * rs;
            mg.sq(v, rs);
            if (v < 0.f)
            {
                return true;
            }
        }
        p = P + kSphereQ * e.n_spheres + kCapsuleQ * e.n_capsules;
        for (int i = 0; i < e.n_zcapsules; ++i, p += kZCapsQ)
        {
            const float4 a = __ldg(p + 0), b = __ldg(p + 1);
            if (b.z >= ext)
            {
                break;
            }
            // sphere_capsule.hh:31-44
            mg.count(1 + 18);
            const float dot = (sz - a.z) * b.x;
            const float cdf = fmaxf(fminf(dot * b.y, 1.f), 0.f);
            const float dx = sx - a.x, dy = sy - a.y, dz = sz - (a.z + b.x * cdf);
            const float rs = sr + a.w;
            const float v = dx * dx + dy * dy + dz * dz - rs * rs;
            mg.sq(v, rs);
            if (v < 0.f)
            {
                return true;
            }
        }
        const float rsq = sr * sr;
        p = P + kSphereQ * e.n_spheres + kCapsuleQ * e.n_capsules + kZCapsQ * e.n_zcapsules;
        for (int i = 0; i < e.n_cuboids; ++i, p += kCuboidQ)
        {
            const float4 c = __ldg(p + 0);
            if (c.w >= ext)
            {
                break;
            }
            // sphere_cuboid.hh:8-26
            mg.count(1 + 30);
            const float4 a1 = __ldg(p + 1), a2 = __ldg(p + 2), a3 = __ldg(p + 3);
            const float xs = sx - c.x, ys = sy - c.y, zs = sz - c.z;
            const float d1 = fmaxf(fabsf(a1.x * xs + a1.y * ys + a1.z * zs) - a1.w, 0.f);
            const float d2 = fmaxf(fabsf(a2.x * xs + a2.y * ys + a2.z * zs) - a2.w, 0.f);
            const float d3 = fmaxf(fabsf(a3.x * xs + a3.y * ys + a3.z * zs) - a3.w, 0.f);
            const float v = d1 * d1 + d2 * d2 + d3 * d3 - rsq;
            mg.sq(v, sr);
            if (v < 0.f)
            {
                return true;
            }
        }
        p = P + kSphereQ * e.n_spheres + kCapsuleQ * e.n_capsules + kZCapsQ * e.n_zcapsules +
            kCuboidQ * e.n_cuboids;
        for (int i = 0; i < e.n_zcuboids; ++i, p += kZCubQ)
        {
            const float4 c = __ldg(p + 0);
            if (c.w >= ext)
            {
                break;
            }
            // sphere_cuboid.hh:35-52
            mg.count(1 + 21);
            const float4 ax = __ldg(p + 1), hr = __ldg(p + 2);
            const float xs = sx - c.x, ys = sy - c.y, zs = sz - c.z;
            const float d1 = fmaxf(fabsf(ax.x * xs + ax.y * ys) - hr.x, 0.f);
            const float d2 = fmaxf(fabsf(ax.z * xs + ax.w * ys) - hr.y, 0.f);
            const float d3 = fmaxf(fabsf(zs) - hr.z, 0.f);
            const float v = d1 * d1 + d2 * d2 + d3 * d3 - rsq;
            mg.sq(v, sr);
            if (v < 0.f)
            {
                return true;
            }
        }
        return other_hit<M, kCapts>(e, sx, sy, sz, sr, mg);
    }
    // sphere_environment_in_collision (collision/validity.hh:46-150) for one sphere per lane, decided by the warp together:
    // every active lane runs the primitive / height-field part and the descent of each point cloud on its own; the leaf
    // scans that remain are then done one query at a time by all 32 lanes (capt_scan_warp).  Same boolean per lane as
    // env_hit.  All 32 lanes must call, converged; verdict walks only.
    VB_D bool env_hit_warp(const EnvView &e, const float4 *P, bool active, float sx, float sy, float sz, float sr)
    {
        NoMargin mg;
        bool hit = active && env_hit<NoMargin, false>(e, P, sx, sy, sz, sr, mg);
        for (int i = 0; i < e.n_capts; ++i)
        {
            const CaptView c = e.capts[i];
            uint32_t v0 = 0u, v1 = 0u;
            const bool pending = active && !hit && capt_probe(c, sx, sy, sz, sr, v0, v1);
            unsigned todo = __ballot_sync(0xffffffffu, pending);
            while (todo)
            {
                const int src = __ffs(todo) - 1;
                todo &= todo - 1u;
                const float qx = __shfl_sync(0xffffffffu, sx, src), qy = __shfl_sync(0xffffffffu, sy, src),
                            qz = __shfl_sync(0xffffffffu, sz, src), qr = __shfl_sync(0xffffffffu, sr, src);
                const uint32_t b0 = __shfl_sync(0xffffffffu, v0, src), b1 = __shfl_sync(0xffffffffu, v1, src);
                const bool any = capt_scan_warp(c, b0, b1, qx, qy, qz, qr);
                if (static_cast<int>(threadIdx.x & 31u) == src)
                {
                    hit = any;
                }
            }
        }
        return hit;
    }
}  // namespace vb

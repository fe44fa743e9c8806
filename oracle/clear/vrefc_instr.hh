// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
//
// Signed-clearance instrumentation of the reference: the reference's own walk
// (collision/validity.hh:13-150,251-293 and robots/<robot>/fk.hh, UNMODIFIED, compiled from where they lie) runs
// with its collision primitives wrapped: every primitive value v the walk computes is also turned into a signed
// clearance in metres, sqrt(v + R^2) - R with R the sum of radii (sphere_sphere.hh:9-23, sphere_capsule.hh:8-44,
// sphere_cuboid.hh:8-52: value = squared gap - R^2), or v itself for height fields (sphere_heightfield.hh:8-30),
// and the smallest |clearance| over all tests the reference EXECUTED for a query is kept.  A verdict of another
// implementation can only differ from the reference's if one of these tests is that close to contact: this is the
// judge of the "+-1e-5 m band" (BASELINE.json north_star), computed by the oracle, not by the code under test.
//
// The wrapping is done with macros that rename the primitives' call sites inside validity.hh (the primitive headers
// are included first, untouched); nothing of the reference is restated.  CAPT queries return only a boolean: a query
// whose answer flips when the radius changes by +-1e-5 m is recorded as clearance 0.
//
// Lives in its own shared library (libvamp_ref_clear.so, hidden visibility): the same reference templates are
// instantiated with different bodies here and in libvamp_ref.so.
#pragma once
#include <algorithm>
#include <array>
#include <cmath>

#include <vamp/vector.hh>
#include <vamp/collision/shapes.hh>
#include <vamp/collision/math.hh>
#include <vamp/collision/environment.hh>
#include <vamp/collision/sphere_sphere.hh>
#include <vamp/collision/sphere_capsule.hh>
#include <vamp/collision/sphere_cuboid.hh>
#include <vamp/collision/sphere_heightfield.hh>

namespace vrefc
{
    constexpr float kBand = 1e-5f;
    inline thread_local float g_min_abs = 3.0e38f;

    inline void note(float clearance)
    {
        g_min_abs = std::min(g_min_abs, std::fabs(clearance));
    }

    // v = squared gap - rs^2 per lane
    template <typename V>
    inline void record_sq(const V &v, const V &rs)
    {
        const auto va = v.to_array();
        const auto ra = rs.to_array();
        for (std::size_t i = 0; i < V::num_scalars; ++i)
        {
            note(std::sqrt(std::max(va[i] + ra[i] * ra[i], 0.f)) - ra[i]);
        }
    }

    // v = squared gap - rsq per lane
    template <typename V>
    inline void record_rsq(const V &v, const V &rsq)
    {
        const auto va = v.to_array();
        const auto ra = rsq.to_array();
        for (std::size_t i = 0; i < V::num_scalars; ++i)
        {
            note(std::sqrt(std::max(va[i] + ra[i], 0.f)) - std::sqrt(std::max(ra[i], 0.f)));
        }
    }

    template <typename V>
    inline void record_lin(const V &v)
    {
        const auto va = v.to_array();
        for (std::size_t i = 0; i < V::num_scalars; ++i)
        {
            note(va[i]);
        }
    }
}  // namespace vrefc

namespace vamp::collision
{
    template <typename DataT>
    inline auto vrefc_sphere_sphere_sql2(
        const DataT &ax, const DataT &ay, const DataT &az, const DataT &ar,
        const DataT &bx, const DataT &by, const DataT &bz, const DataT &br) noexcept -> DataT
    {
        const auto v = sphere_sphere_sql2(ax, ay, az, ar, bx, by, bz, br);
        vrefc::record_sq(v, ar + br);
        return v;
    }

    template <typename DataT>
    inline auto vrefc_sphere_sphere_sql2(
        const Sphere<DataT> &a, const DataT &x, const DataT &y, const DataT &z, const DataT &r) noexcept -> DataT
    {
        const auto v = sphere_sphere_sql2(a, x, y, z, r);
        vrefc::record_sq(v, a.r + r);
        return v;
    }

    template <typename DataT>
    inline auto vrefc_sphere_capsule(
        const Capsule<DataT> &c, const DataT &x, const DataT &y, const DataT &z, const DataT &r) noexcept -> DataT
    {
        const auto v = sphere_capsule(c, x, y, z, r);
        vrefc::record_sq(v, r + c.r);
        return v;
    }

    template <typename DataT>
    inline auto vrefc_sphere_z_aligned_capsule(
        const Capsule<DataT> &c, const DataT &x, const DataT &y, const DataT &z, const DataT &r) noexcept -> DataT
    {
        const auto v = sphere_z_aligned_capsule(c, x, y, z, r);
        vrefc::record_sq(v, r + c.r);
        return v;
    }

    template <typename DataT>
    inline auto vrefc_sphere_cuboid(
        const Cuboid<DataT> &c, const DataT &x, const DataT &y, const DataT &z, const DataT &rsq) noexcept -> DataT
    {
        const auto v = sphere_cuboid(c, x, y, z, rsq);
        vrefc::record_rsq(v, rsq);
        return v;
    }

    template <typename DataT>
    inline auto vrefc_sphere_z_aligned_cuboid(
        const Cuboid<DataT> &c, const DataT &x, const DataT &y, const DataT &z, const DataT &rsq) noexcept -> DataT
    {
        const auto v = sphere_z_aligned_cuboid(c, x, y, z, rsq);
        vrefc::record_rsq(v, rsq);
        return v;
    }

    template <typename DataT>
    inline auto vrefc_sphere_heightfield(
        const HeightField<DataT> &a, const DataT &x, const DataT &y, const DataT &z, const DataT &r) noexcept -> DataT
    {
        const auto v = sphere_heightfield(a, x, y, z, r);
        vrefc::record_lin(v);
        return v;
    }
}  // namespace vamp::collision

// ---- the reference's walk, with its primitive call sites renamed --------------------------------------------
#define sphere_sphere_sql2 vrefc_sphere_sphere_sql2
#define sphere_capsule vrefc_sphere_capsule
#define sphere_z_aligned_capsule vrefc_sphere_z_aligned_capsule
#define sphere_cuboid vrefc_sphere_cuboid
#define sphere_z_aligned_cuboid vrefc_sphere_z_aligned_cuboid
#define sphere_heightfield vrefc_sphere_heightfield
#include <vamp/collision/validity.hh>
#undef sphere_sphere_sql2
#undef sphere_capsule
#undef sphere_z_aligned_capsule
#undef sphere_cuboid
#undef sphere_z_aligned_cuboid
#undef sphere_heightfield

namespace vamp
{
    // robots/<robot>/fk.hh calls sphere_environment_in_collision unqualified: renamed to this wrapper, which runs the
    // reference's query (above) and probes the point clouds at radius -+ band.
    template <typename DataT, typename ArgT1, typename ArgT2, typename ArgT3, typename ArgT4>
    inline auto vrefc_sphere_environment_in_collision(
        const collision::Environment<DataT> &e, ArgT1 sx_, ArgT2 sy_, ArgT3 sz_, ArgT4 sr_) noexcept -> bool
    {
        const bool hit = sphere_environment_in_collision(e, sx_, sy_, sz_, sr_);
        if (not e.pointclouds.empty())
        {
            const auto sr = static_cast<DataT>(sr_);
            const std::array<DataT, 3> positions = {
                static_cast<DataT>(sx_), static_cast<DataT>(sy_), static_cast<DataT>(sz_)};
            for (const auto &pc : e.pointclouds)
            {
                if (pc.collides_simd(positions, sr - vrefc::kBand) != pc.collides_simd(positions, sr + vrefc::kBand))
                {
                    vrefc::note(0.f);
                }
            }
        }
        return hit;
    }
}  // namespace vamp
#define sphere_environment_in_collision vrefc_sphere_environment_in_collision
// (robot headers are included after this point by the per-robot translation units)

// vamp_b200_planners.hh — makes mr-vamp's own planner / path code run its edge checks on the B200 WITHOUT editing it.
//
// The planners reach the hot path through two function templates of planning/validate.hh:
//
//   validate_vector<Robot, rake, resolution>(start, vector, distance, environment)     planning/validate.hh:23-65
//   validate_motion<Robot, rake, resolution>(start, goal, environment)                 planning/validate.hh:67-75
//
// called unqualified from namespace vamp::planning (Path::validate, planning/plan.hh:155-168; shortcut_path and the
// other simplifiers, planning/simplify.hh:95-140; RRTC's extend / connect loops, planning/rrtc.hh:139,180; PRM, FCIT*).
// Include this header BEFORE those headers (or compile with `-include vamp_b200_planners.hh`): it pulls in the
// reference's validate.hh first -- its CPU functions stay available as vamp::planning::cpu_validate_vector /
// cpu_validate_motion -- and then renames the two names at every later call site to device-backed functions with the SAME
// signatures.  The vector-form environment the callers pass (collision::Environment<FloatVector<rake>>) is looked up in a
// cache of device replicas keyed by its address and a fingerprint of its contents, so an environment is uploaded once
// and again only when it changes.
//
// One call = one edge = one kernel launch (~10 us): correct and source compatible, not fast.  Throughput comes from the
// batch entry points (vb_validate_motion, vb_validate_vector, vb_connect_increments, vb_validate_path, vb_shortcut_path).
#pragma once
#include <cstring>
#include <unordered_map>

#include <vamp/planning/validate.hh>

#include "vamp_b200_dropin.hh"

namespace vamp::planning
{
    // the reference's own CPU functions under names the macros below do not touch
    template <typename Robot, std::size_t rake, std::size_t resolution>
    inline auto cpu_validate_vector(
        const typename Robot::Configuration &start,
        const typename Robot::Configuration &vector,
        float distance,
        const collision::Environment<FloatVector<rake>> &environment) -> bool
    {
        return vamp::planning::validate_vector<Robot, rake, resolution>(start, vector, distance, environment);
    }

    template <typename Robot, std::size_t rake, std::size_t resolution>
    inline auto cpu_validate_motion(
        const typename Robot::Configuration &start,
        const typename Robot::Configuration &goal,
        const collision::Environment<FloatVector<rake>> &environment) -> bool
    {
        return vamp::planning::validate_motion<Robot, rake, resolution>(start, goal, environment);
    }
}  // namespace vamp::planning

namespace vamp::b200
{
    // FNV-1a over the numbers that define an environment (shape fields, sizes of the bulk buffers, attachment pose)
    struct Fingerprint
    {
        uint64_t h = 1469598103934665603ull;
        void bytes(const void *p, std::size_t n)
        {
            const auto *b = static_cast<const unsigned char *>(p);
            for (std::size_t i = 0; i < n; ++i)
            {
                h = (h ^ b[i]) * 1099511628211ull;
            }
        }
        template <typename T>
        void num(const T &v)
        {
            const float f = scalar_of(v);
            bytes(&f, sizeof(f));
        }
        void size(std::size_t n) { bytes(&n, sizeof(n)); }
    };

    template <typename DataT>
    inline uint64_t fingerprint(const collision::Environment<DataT> &e)
    {
        Fingerprint fp;
        fp.size(e.spheres.size());
        for (const auto &s : e.spheres)
        {
            fp.num(s.x), fp.num(s.y), fp.num(s.z), fp.num(s.r);
        }
        auto capsules = [&](const auto &list) {
            fp.size(list.size());
            for (const auto &c : list)
            {
                fp.num(c.x1), fp.num(c.y1), fp.num(c.z1), fp.num(c.xv), fp.num(c.yv), fp.num(c.zv), fp.num(c.r);
            }
        };
        capsules(e.capsules);
        capsules(e.z_aligned_capsules);
        auto cuboids = [&](const auto &list) {
            fp.size(list.size());
            for (const auto &c : list)
            {
                fp.num(c.x), fp.num(c.y), fp.num(c.z), fp.num(c.axis_1_x), fp.num(c.axis_1_y), fp.num(c.axis_1_z), fp.num(c.axis_2_x),
                    fp.num(c.axis_2_y), fp.num(c.axis_2_z), fp.num(c.axis_1_r), fp.num(c.axis_2_r), fp.num(c.axis_3_r);
            }
        };
        cuboids(e.cuboids);
        cuboids(e.z_aligned_cuboids);
        fp.size(e.heightfields.size());
        for (const auto &h : e.heightfields)
        {
            fp.num(h.x), fp.num(h.y), fp.num(h.z), fp.size(h.xd), fp.size(h.yd);
            fp.bytes(&h.data, sizeof(void *));  // (the buffer's identity, not its megabytes)
        }
        fp.size(e.pointclouds.size());
        for (const auto &c : e.pointclouds)
        {
            fp.size(c.tests.size()), fp.size(c.affordances[0].size());
            fp.bytes(c.tests.data(), std::min<std::size_t>(c.tests.size(), 64) * sizeof(float));
        }
        fp.size(e.attachments ? 1 + e.attachments->spheres.size() : 0);
        if (e.attachments)
        {
            const auto &a = *e.attachments;
            fp.num(a.tf_tx), fp.num(a.tf_ty), fp.num(a.tf_tz), fp.num(a.tf_rx), fp.num(a.tf_ry), fp.num(a.tf_rz), fp.num(a.tf_rw);
        }
        return fp.h;
    }

    // device replica of the environment a reference-typed call passes by reference (one cache per host thread:
    // the reference's calls are single threaded and a vb_ctx belongs to one thread)
    template <typename DataT>
    inline const DeviceEnvironment &replica_of(const collision::Environment<DataT> &e)
    {
        struct Entry
        {
            uint64_t fp;
            std::unique_ptr<DeviceEnvironment> env;
        };
        thread_local std::unordered_map<const void *, Entry> cache;
        const uint64_t fp = fingerprint(e);
        auto it = cache.find(&e);
        if (it == cache.end() || it->second.fp != fp)
        {
            if (cache.size() > 64)
            {
                cache.clear();  // addresses of short-lived environments: do not grow without bound
            }
            Entry entry{fp, std::make_unique<DeviceEnvironment>(e)};
            it = cache.insert_or_assign(&e, std::move(entry)).first;
        }
        return *it->second.env;
    }
}  // namespace vamp::b200

namespace vamp::planning
{
    // planning/validate.hh:23-65 on the device: same signature, same number of passes (from `distance`), same interpolants
    template <typename Robot, std::size_t rake, std::size_t resolution>
    inline auto b200_validate_vector(
        const typename Robot::Configuration &start,
        const typename Robot::Configuration &vector,
        float distance,
        const collision::Environment<FloatVector<rake>> &environment) -> bool
    {
        const auto &denv = b200::replica_of(environment);
        const auto a = start.to_array(), v = vector.to_array();
        uint32_t bits = 0;
        const auto base = b200::base_of<Robot>::get();
        b200::check(vb_validate_vector(
            denv.ctx(), b200::robot_id<Robot>(), base.data(), denv.get(), a.data(), v.data(), &distance, 1,
            static_cast<int>(resolution), &bits, nullptr));
        return (bits & 1u) != 0u;
    }

    // planning/validate.hh:67-75
    template <typename Robot, std::size_t rake, std::size_t resolution>
    inline auto b200_validate_motion(
        const typename Robot::Configuration &start,
        const typename Robot::Configuration &goal,
        const collision::Environment<FloatVector<rake>> &environment) -> bool
    {
        auto vector = goal - start;
        return b200_validate_vector<Robot, rake, resolution>(start, vector, vector.l2_norm(), environment);
    }
}  // namespace vamp::planning

// every later call site of the two names -- in headers of the reference included after this point -- now calls the device
#define validate_vector b200_validate_vector
#define validate_motion b200_validate_motion

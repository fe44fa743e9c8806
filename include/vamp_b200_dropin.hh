// vamp_b200_dropin.hh — C++17 host layer over the C ABI of libvamp_b200.so for code that already
// uses mr-vamp's own types.  Header only; needs the VAMP headers on the include path.
//
// It keeps the reference's signatures on the fkcc path and forwards them to the B200 kernels:
//
//   vamp::b200::DeviceEnvironment(env)            <- vamp::collision::Environment<float>
//                                                    (collision/environment.hh:11-82; replaces the per-call
//                                                    conversion to Environment<FloatVector<rake>>,
//                                                    bindings/common.hh:181)
//   vamp::b200::fkcc<Robot, rake>(denv, block)    <- Robot::fkcc<rake>(env, block)      robots/panda_base.hh:53-59
//   vamp::b200::sphere_fk<Robot, rake>(block, s)  <- Robot::sphere_fk<rake>(block, s)   robots/panda_base.hh:67-71
//   vamp::b200::validate_motion<Robot>(a, b, denv)<- planning::validate_motion<Robot, rake, res>(a, b, env)
//                                                    planning/validate.hh:67-75
//   vamp::b200::validate_batch / validate_motion_batch / validate_path / shortcut_path
//                                                 <- the batch calls README.md:308 lists as a TODO, and
//                                                    Path::validate (planning/plan.hh:155-168) in one launch
//
// Single-block calls cost a kernel launch (~10 us); planners should collect their queries and use the
// batch calls.  Errors of the C ABI become std::runtime_error; there is no CPU fallback.
#pragma once
#include <array>
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <string>
#include <string_view>
#include <type_traits>
#include <vector>

#include <vamp/collision/environment.hh>
#include <vamp/vector.hh>

#include "vamp_b200.h"

namespace vamp::b200
{
    inline void check(int rc)
    {
        if (rc != VB_OK)
        {
            throw std::runtime_error(std::string("vamp_b200: ") + vb_last_error());
        }
    }

    // One context (device set + streams) shared by every query of the process unless the caller makes its own.
    class Context
    {
    public:
        explicit Context(const std::vector<int> &devices = {})
        {
            check(vb_ctx_create(&ctx_, devices.empty() ? nullptr : devices.data(), static_cast<int>(devices.size())));
        }
        ~Context() { vb_ctx_destroy(ctx_); }
        Context(const Context &) = delete;
        Context &operator=(const Context &) = delete;
        vb_ctx *get() const { return ctx_; }
        static Context &instance()
        {
            static Context c;
            return c;
        }

    private:
        vb_ctx *ctx_ = nullptr;
    };

    // Robot type -> robot id + base offset (PandaBase<X,Y,Z>::base_x/y/z; zero for the other robots).
    template <typename Robot, typename = void>
    struct base_of
    {
        static std::array<float, 3> get() { return {0.f, 0.f, 0.f}; }
    };
    template <typename Robot>
    struct base_of<Robot, std::void_t<decltype(Robot::base_x)>>
    {
        static std::array<float, 3> get() { return {Robot::base_x, Robot::base_y, Robot::base_z}; }
    };

    template <typename Robot>
    inline vb_robot robot_id()
    {
        // every Panda grid variant (robots/panda_grid.hh) shares one trace; the base offset is an argument
        const std::string_view name(Robot::name);
        vb_robot id;
        check(vb_robot_by_name(name.substr(0, 5) == "panda" ? "panda" : Robot::name, &id));
        return id;
    }

    // scalar of an environment field: the value itself (Environment<float>) or lane 0 of its broadcast
    // (Environment<FloatVector<rake>>: every lane holds the same number, collision/environment.hh:26-38)
    template <typename T>
    inline float scalar_of(const T &v)
    {
        if constexpr (std::is_arithmetic_v<T>)
        {
            return static_cast<float>(v);
        }
        else
        {
            return v.to_array()[0];
        }
    }

    // Device-resident copy of a collision::Environment<float> (or of its vector form); rebuilt only when constructed.
    class DeviceEnvironment
    {
    public:
        template <typename DataT>
        explicit DeviceEnvironment(const collision::Environment<DataT> &e, Context &ctx = Context::instance()) : ctx_(&ctx)
        {
            check(vb_env_create(ctx.get(), &env_));
            try
            {
                auto f = [](const auto &v) { return scalar_of(v); };
                for (const auto &s : e.spheres)
                {
                    check(vb_env_add_sphere(env_, f(s.x), f(s.y), f(s.z), f(s.r)));
                }
                auto capsules = [&](const auto &list) {
                    for (const auto &c : list)
                    {
                        const float raw[8] = {f(c.x1), f(c.y1), f(c.z1), f(c.xv), f(c.yv), f(c.zv), f(c.r), f(c.rdv)};
                        check(vb_env_add_capsule(env_, raw));
                    }
                };
                capsules(e.capsules);
                capsules(e.z_aligned_capsules);
                auto cuboids = [&](const auto &list) {
                    for (const auto &c : list)
                    {
                        const float raw[15] = {f(c.x),        f(c.y),        f(c.z),        f(c.axis_1_x), f(c.axis_1_y),
                                               f(c.axis_1_z), f(c.axis_2_x), f(c.axis_2_y), f(c.axis_2_z), f(c.axis_3_x),
                                               f(c.axis_3_y), f(c.axis_3_z), f(c.axis_1_r), f(c.axis_2_r), f(c.axis_3_r)};
                        check(vb_env_add_cuboid(env_, raw));
                    }
                };
                cuboids(e.cuboids);
                cuboids(e.z_aligned_cuboids);
                for (const auto &h : e.heightfields)
                {
                    const float os[6] = {f(h.x), f(h.y), f(h.z), f(h.xs), f(h.ys), f(h.zs)};
                    check(vb_env_add_heightfield(env_, os, h.xd, h.yd, h.data.data()));
                }
                for (const auto &c : e.pointclouds)
                {
                    const float top[6] = {c.aabb_top.lower[0], c.aabb_top.lower[1], c.aabb_top.lower[2],
                                          c.aabb_top.upper[0], c.aabb_top.upper[1], c.aabb_top.upper[2]};
                    check(vb_env_add_capt(env_, c.tests.data(), c.tests.size(), c.aff_starts.data(),
                                          reinterpret_cast<const float *>(c.affordances[0].data()),
                                          reinterpret_cast<const float *>(c.affordances[1].data()),
                                          reinterpret_cast<const float *>(c.affordances[2].data()), c.affordances[0].size(),
                                          c.aabbs.empty() ? nullptr : &c.aabbs.front().lower.front(), c.aabbs.size(), top,
                                          c.r_point, static_cast<int>(c.nlog2)));
                }
                if (e.attachments)
                {
                    const auto &a = *e.attachments;
                    const float tf[7] = {f(a.tf_tx), f(a.tf_ty), f(a.tf_tz), f(a.tf_rx), f(a.tf_ry), f(a.tf_rz), f(a.tf_rw)};
                    std::vector<float> sph;
                    for (const auto &s : a.spheres)
                    {
                        sph.insert(sph.end(), {f(s.x), f(s.y), f(s.z), f(s.r)});
                    }
                    check(vb_env_set_attachment(env_, tf, sph.data(), a.spheres.size()));
                }
                check(vb_env_commit(env_));
            }
            catch (...)
            {
                vb_env_destroy(env_);
                throw;
            }
        }
        ~DeviceEnvironment() { vb_env_destroy(env_); }
        DeviceEnvironment(const DeviceEnvironment &) = delete;
        DeviceEnvironment &operator=(const DeviceEnvironment &) = delete;
        vb_env *get() const { return env_; }
        vb_ctx *ctx() const { return ctx_->get(); }

    private:
        Context *ctx_;
        vb_env *env_ = nullptr;
    };

    inline std::vector<bool> unpack(const std::vector<uint32_t> &words, std::size_t n)
    {
        std::vector<bool> out(n);
        for (std::size_t i = 0; i < n; ++i)
        {
            out[i] = (words[i >> 5] >> (i & 31)) & 1u;
        }
        return out;
    }

    // ---- batch calls: configurations row-major [n][Robot::dimension] -----------------------------------
    template <typename Robot>
    inline std::vector<bool> fkcc_batch(const DeviceEnvironment &env, const float *configs, std::size_t n)
    {
        std::vector<uint32_t> bits((n + 31) / 32);
        const auto base = base_of<Robot>::get();
        check(vb_fkcc(env.ctx(), robot_id<Robot>(), base.data(), env.get(), configs, n, bits.data(), nullptr));
        return unpack(bits, n);
    }

    // Helper::validate_configuration (bindings/common.hh:172-182): joint limits && fkcc
    template <typename Robot>
    inline std::vector<bool> validate_batch(const DeviceEnvironment &env, const float *configs, std::size_t n)
    {
        std::vector<uint32_t> bits((n + 31) / 32);
        const auto base = base_of<Robot>::get();
        check(vb_validate(env.ctx(), robot_id<Robot>(), base.data(), env.get(), configs, n, bits.data()));
        return unpack(bits, n);
    }

    template <typename Robot>
    inline std::vector<bool> validate_motion_batch(const DeviceEnvironment &env, const float *starts, const float *goals,
                                                   std::size_t n, int resolution = Robot::resolution)
    {
        std::vector<uint32_t> bits((n + 31) / 32);
        const auto base = base_of<Robot>::get();
        check(vb_validate_motion(env.ctx(), robot_id<Robot>(), base.data(), env.get(), starts, goals, n, resolution,
                                 bits.data(), nullptr));
        return unpack(bits, n);
    }

    // Path::validate (planning/plan.hh:155-168): every consecutive edge of a waypoint list in one launch
    template <typename Robot>
    inline bool validate_path(const DeviceEnvironment &env, const float *waypoints, std::size_t n_waypoints)
    {
        if (n_waypoints < 2)
        {
            return true;
        }
        int64_t first = -1;
        const auto base = base_of<Robot>::get();
        check(vb_validate_path(env.ctx(), robot_id<Robot>(), base.data(), env.get(), waypoints, n_waypoints, Robot::resolution,
                               &first));
        return first < 0;
    }

    // planning::shortcut_path (planning/simplify.hh:115-140): every candidate shortcut of the path validated in one
    // batch, the reference's scan replayed on the verdicts.  keep[i] != 0: waypoint i stays in the path.  Returns
    // true when at least one waypoint was removed (the reference's return value).
    template <typename Robot>
    inline bool shortcut_path(const DeviceEnvironment &env, const float *waypoints, std::size_t n_waypoints,
                              std::vector<uint8_t> &keep)
    {
        keep.assign(n_waypoints, 1);
        std::size_t kept = n_waypoints;
        const auto base = base_of<Robot>::get();
        check(vb_shortcut_path(env.ctx(), robot_id<Robot>(), base.data(), env.get(), waypoints, n_waypoints, Robot::resolution,
                               keep.data(), &kept));
        return kept != n_waypoints;
    }

    // ---- the reference's own signatures -------------------------------------------------------------------
    template <typename Robot, std::size_t rake>
    inline std::array<float, rake * Robot::dimension> unpack_block(const typename Robot::template ConfigurationBlock<rake> &q)
    {
        std::array<float, rake * Robot::dimension> rows{};
        for (std::size_t j = 0; j < Robot::dimension; ++j)
        {
            const auto lane = q[j].to_array();
            for (std::size_t l = 0; l < rake; ++l)
            {
                rows[l * Robot::dimension + j] = lane[l];
            }
        }
        return rows;
    }

    // Robot::fkcc<rake>: true iff every lane of the block is collision free
    template <typename Robot, std::size_t rake>
    inline bool fkcc(const DeviceEnvironment &env, const typename Robot::template ConfigurationBlock<rake> &q)
    {
        const auto rows = unpack_block<Robot, rake>(q);
        uint32_t bits[(rake + 31) / 32] = {};
        const auto base = base_of<Robot>::get();
        check(vb_fkcc(env.ctx(), robot_id<Robot>(), base.data(), env.get(), rows.data(), rake, bits, nullptr));
        for (std::size_t l = 0; l < rake; ++l)
        {
            if (!((bits[l >> 5] >> (l & 31)) & 1u))
            {
                return false;
            }
        }
        return true;
    }

    // Robot::sphere_fk<rake>
    template <typename Robot, std::size_t rake>
    inline void sphere_fk(const typename Robot::template ConfigurationBlock<rake> &q, typename Robot::template Spheres<rake> &out,
                          Context &ctx = Context::instance())
    {
        const auto rows = unpack_block<Robot, rake>(q);
        std::vector<float> s(rake * Robot::n_spheres * 4);
        const auto base = base_of<Robot>::get();
        check(vb_sphere_fk(ctx.get(), robot_id<Robot>(), base.data(), rows.data(), rake, s.data()));
        for (std::size_t k = 0; k < Robot::n_spheres; ++k)
        {
            std::array<float, rake> x{}, y{}, z{}, r{};
            for (std::size_t l = 0; l < rake; ++l)
            {
                const float *p = s.data() + (l * Robot::n_spheres + k) * 4;
                x[l] = p[0];
                y[l] = p[1];
                z[l] = p[2];
                r[l] = p[3];
            }
            out.x[k] = FloatVector<rake>(x);
            out.y[k] = FloatVector<rake>(y);
            out.z[k] = FloatVector<rake>(z);
            out.r[k] = FloatVector<rake>(r);
        }
    }

    // planning::validate_motion<Robot, rake, resolution>(a, b, env)
    template <typename Robot, std::size_t resolution = Robot::resolution>
    inline bool validate_motion(const typename Robot::Configuration &a, const typename Robot::Configuration &b,
                                const DeviceEnvironment &env)
    {
        const auto qa = a.to_array(), qb = b.to_array();
        return validate_motion_batch<Robot>(env, qa.data(), qb.data(), 1, static_cast<int>(resolution))[0];
    }
}  // namespace vamp::b200

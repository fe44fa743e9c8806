// mr_vamp_b200._core_ext — compiled Python module over the C ABI of libvamp_b200.so, on the surface of the
// reference's nanobind module `vamp._core_ext` for the fkcc path:
//
//   Sphere / Cuboid / Cylinder / HeightField / Attachment      bindings/environment.cc:18-105
//   Environment.add_sphere / add_cuboid / add_capsule / add_heightfield / add_pointcloud / attach / detach
//                                                              bindings/environment.cc:107-163
//   <robot>.validate(configuration, environment)               bindings/common.hh:172-189,685-697
//   <robot>.fk(configuration)                                  bindings/common.hh:132-152,706-710
//   <robot>.sphere_validity(configuration, environment)        bindings/common.hh:155-170,699-704
//   <robot>.dimension() / resolution() / n_spheres() / lower_bounds() / upper_bounds()
//   problem_dict_to_vamp(problem, ignore_names)                src/vamp/__init__.py:135-181 (here in C++)
//   RobotFactory / Robot / MultiRobotEnvironment               robots/robot_factory.hh:23-189, robots/robot_interface.hh and
//                                                              the planned collision/multi_robot_environment.hh + bindings
//                                                              (MR_DEV_PLAN.md:176-198,244-279) over vb_validate_pairs /
//                                                              vb_inter_robot
//   + the batch calls the reference lists as a TODO (README.md:308), zero-copy numpy in / out, GIL released:
//   <robot>.validate_batch / fkcc_batch / validate_motion_batch / validate_vector_batch / fk_batch
//
// nanobind is not in this image; pybind11 is.  The module computes nothing: every query is one call of the C
// ABI (include/vamp_b200.h) and fails with a RuntimeError when the library reports an error -- there is no CPU path.
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <algorithm>
#include <array>
#include <cmath>
#include <cstdio>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/vamp_b200.h"

namespace py = pybind11;

namespace
{
    using FArray = py::array_t<float, py::array::c_style | py::array::forcecast>;

    void check(int rc)
    {
        if (rc != VB_OK)
        {
            throw std::runtime_error(std::string("vamp_b200: [") + std::to_string(rc) + "] " + vb_last_error());
        }
    }

    struct Context
    {
        vb_ctx *h = nullptr;
        Context() { check(vb_ctx_create(&h, nullptr, 0)); }
        explicit Context(const std::vector<int> &devices) { check(vb_ctx_create(&h, devices.data(), static_cast<int>(devices.size()))); }
        ~Context() { vb_ctx_destroy(h); }
    };

    std::shared_ptr<Context> default_context()
    {
        static std::shared_ptr<Context> ctx = std::make_shared<Context>();
        return ctx;
    }

    // ---- shapes: the raw floats of collision/shapes.hh, built as collision/factory.hh does (float32) -------------
    // Rz(phi) * Ry(theta) * Rx(rho), collision/factory.hh:37-39
    std::array<float, 9> rotation_xyz(const std::array<float, 3> &e)
    {
        // trig in double, narrowed to float (so that a yaw-only box has axis_3_z == 1 exactly, bindings/environment.cc:120)
        auto c = [](float a) { return static_cast<float>(std::cos(static_cast<double>(a))); };
        auto s = [](float a) { return static_cast<float>(std::sin(static_cast<double>(a))); };
        const float cx = c(e[0]), sx = s(e[0]), cy = c(e[1]), sy = s(e[1]), cz = c(e[2]), sz = s(e[2]);
        const float rx[9] = {1, 0, 0, 0, cx, -sx, 0, sx, cx};
        const float ry[9] = {cy, 0, sy, 0, 1, 0, -sy, 0, cy};
        const float rz[9] = {cz, -sz, 0, sz, cz, 0, 0, 0, 1};
        auto mul = [](const float *a, const float *b) {
            std::array<float, 9> c{};
            for (int i = 0; i < 3; ++i)
            {
                for (int j = 0; j < 3; ++j)
                {
                    float s = 0.f;
                    for (int k = 0; k < 3; ++k)
                    {
                        s += a[3 * i + k] * b[3 * k + j];
                    }
                    c[3 * i + j] = s;
                }
            }
            return c;
        };
        const auto zy = mul(rz, ry);
        return mul(zy.data(), rx);
    }

    struct Sphere
    {
        float x, y, z, r;
        std::string name;
    };

    struct Cuboid
    {
        std::array<float, 15> raw;
        std::string name;
    };

    struct Cylinder
    {
        std::array<float, 8> raw;
        std::string name;
    };

    struct HeightField
    {
        std::array<float, 6> off_scale;
        size_t xd, yd;
        std::vector<float> data;
        std::string name;
    };

    struct Attachment
    {
        std::array<float, 7> tf;
        std::vector<Sphere> spheres;
    };

    Cuboid make_cuboid(const std::array<float, 3> &c, const std::array<float, 3> &euler, const std::array<float, 3> &half)
    {
        const auto r = rotation_xyz(euler);
        Cuboid out;
        out.raw = {c[0], c[1], c[2], r[0], r[3], r[6], r[1], r[4], r[7], r[2], r[5], r[8], half[0], half[1], half[2]};
        return out;
    }

    Cylinder cylinder_from_endpoints(const std::array<float, 3> &p1, const std::array<float, 3> &p2, float radius)
    {
        const float v[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
        const float dot = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
        Cylinder out;
        out.raw = {p1[0], p1[1], p1[2], v[0], v[1], v[2], radius, 1.0f / dot};  // collision/factory.hh:233-244
        return out;
    }

    Cylinder make_cylinder(const std::array<float, 3> &c, const std::array<float, 3> &euler, float radius, float length)
    {
        const auto r = rotation_xyz(euler);
        const float half = length / 2.f;
        const std::array<float, 3> p1 = {r[2] * half + c[0], r[5] * half + c[1], r[8] * half + c[2]};
        const std::array<float, 3> p2 = {r[2] * -half + c[0], r[5] * -half + c[1], r[8] * -half + c[2]};
        return cylinder_from_endpoints(p1, p2, radius);
    }

    // ---- environment ------------------------------------------------------------------------------------------------
    class Environment
    {
    public:
        explicit Environment(std::shared_ptr<Context> ctx = default_context()) : ctx_(std::move(ctx))
        {
            check(vb_env_create(ctx_->h, &env_));
        }
        ~Environment() { vb_env_destroy(env_); }
        Environment(const Environment &) = delete;
        Environment &operator=(const Environment &) = delete;

        void add_sphere(const Sphere &s)
        {
            check(vb_env_add_sphere(env_, s.x, s.y, s.z, s.r));
            named(s.name);
        }
        void add_cuboid(const Cuboid &c)
        {
            check(vb_env_add_cuboid(env_, c.raw.data()));
            named(c.name);
        }
        void add_capsule(const Cylinder &c)
        {
            check(vb_env_add_capsule(env_, c.raw.data()));
            named(c.name);
        }
        void add_heightfield(const HeightField &h)
        {
            check(vb_env_add_heightfield(env_, h.off_scale.data(), h.xd, h.yd, h.data.data()));
            named(h.name);
        }
        // bindings/environment.cc:148-158: the CAPT is built by the library on the device; returns nanoseconds
        double add_pointcloud(const FArray &points, float r_min, float r_max, float r_point)
        {
            if (points.ndim() != 2 || points.shape(1) != 3)
            {
                throw std::invalid_argument("add_pointcloud: float32[n, 3] expected");
            }
            double ms = 0.0;
            check(vb_env_add_pointcloud(env_, points.data(), static_cast<size_t>(points.shape(0)), r_min, r_max, r_point, &ms));
            dirty_ = true;
            return ms * 1e6;
        }
        void attach(const Attachment &a)
        {
            std::vector<float> sph;
            for (const auto &s : a.spheres)
            {
                sph.insert(sph.end(), {s.x, s.y, s.z, s.r});
            }
            check(vb_env_set_attachment(env_, a.tf.data(), sph.data(), a.spheres.size()));
            dirty_ = true;
        }
        void detach()
        {
            check(vb_env_clear_attachment(env_));
            dirty_ = true;
        }
        // device copy rebuilt lazily: one upload per change set (the reference converts on every query, common.hh:181)
        vb_env *handle()
        {
            if (dirty_)
            {
                check(vb_env_commit(env_));
                dirty_ = false;
            }
            return env_;
        }
        vb_ctx *ctx() const { return ctx_->h; }
        std::vector<std::string> object_names() const
        {
            size_t n = 0;
            check(vb_env_object_ids(env_, nullptr, &n));
            std::vector<uint32_t> ids(n);
            if (n)
            {
                check(vb_env_object_ids(env_, ids.data(), &n));
            }
            std::vector<std::string> out;
            for (uint32_t id : ids)
            {
                out.push_back(names_.at(id));
            }
            return out;
        }

    private:
        void named(const std::string &name)
        {
            names_.push_back(name);
            dirty_ = true;
        }
        std::shared_ptr<Context> ctx_;
        vb_env *env_ = nullptr;
        bool dirty_ = true;
        std::vector<std::string> names_;
    };

    // ---- robots -----------------------------------------------------------------------------------------------------
    struct Robot
    {
        vb_robot id;
        std::array<float, 3> base{0.f, 0.f, 0.f};
        int dim = 0, n_spheres = 0, resolution = 0;
        std::vector<float> lo, range;
        std::shared_ptr<Environment> empty;  // the reference's default argument: an empty Environment

        Robot(const char *name, std::array<float, 3> base_xyz) : base(base_xyz)
        {
            check(vb_robot_by_name(name, &id));
            check(vb_robot_info(id, &dim, &n_spheres, &resolution));
            lo.resize(dim);
            range.resize(dim);
            check(vb_robot_limits(id, lo.data(), range.data()));
        }

        Environment &env_or_empty(Environment *e)
        {
            if (e)
            {
                return *e;
            }
            if (!empty)
            {
                empty = std::make_shared<Environment>();
            }
            return *empty;
        }

        // [n][dim] float32, C contiguous: zero copy when the caller already holds that layout
        size_t rows(const FArray &a, const char *what) const
        {
            if (a.ndim() == 1 && a.shape(0) == dim)
            {
                return 1;
            }
            if (a.ndim() != 2 || a.shape(1) != dim)
            {
                throw std::invalid_argument(std::string(what) + ": expected float32[n, " + std::to_string(dim) + "]");
            }
            return static_cast<size_t>(a.shape(0));
        }

        static py::array_t<bool> unpack(const std::vector<uint32_t> &bits, size_t n)
        {
            py::array_t<bool> out(static_cast<py::ssize_t>(n));
            bool *o = out.mutable_data();
            for (size_t i = 0; i < n; ++i)
            {
                o[i] = (bits[i >> 5] >> (i & 31)) & 1u;
            }
            return out;
        }

        py::array_t<bool> fkcc_batch(const FArray &q, Environment *env)
        {
            const size_t n = rows(q, "fkcc_batch");
            Environment &e = env_or_empty(env);
            vb_env *eh = e.handle();
            std::vector<uint32_t> bits((n + 31) / 32);
            {
                py::gil_scoped_release nogil;
                check(vb_fkcc(e.ctx(), id, base.data(), eh, q.data(), n, bits.data(), nullptr));
            }
            return unpack(bits, n);
        }

        py::array_t<bool> validate_batch(const FArray &q, Environment *env)
        {
            const size_t n = rows(q, "validate_batch");
            Environment &e = env_or_empty(env);
            vb_env *eh = e.handle();
            std::vector<uint32_t> bits((n + 31) / 32);
            {
                py::gil_scoped_release nogil;
                check(vb_validate(e.ctx(), id, base.data(), eh, q.data(), n, bits.data()));
            }
            return unpack(bits, n);
        }

        bool validate(const FArray &q, Environment *env)
        {
            if (rows(q, "validate") != 1)
            {
                throw std::invalid_argument("validate: one configuration expected (validate_batch takes many)");
            }
            return validate_batch(q, env).at(0);
        }

        py::array_t<bool> validate_motion_batch(const FArray &a, const FArray &b, Environment *env, int res)
        {
            const size_t n = rows(a, "validate_motion_batch");
            if (rows(b, "validate_motion_batch") != n)
            {
                throw std::invalid_argument("validate_motion_batch: starts and goals differ in length");
            }
            Environment &e = env_or_empty(env);
            vb_env *eh = e.handle();
            std::vector<uint32_t> bits((n + 31) / 32);
            {
                py::gil_scoped_release nogil;
                check(vb_validate_motion(e.ctx(), id, base.data(), eh, a.data(), b.data(), n, res, bits.data(), nullptr));
            }
            return unpack(bits, n);
        }

        py::array_t<bool> validate_vector_batch(const FArray &a, const FArray &v, const FArray &d, Environment *env, int res)
        {
            const size_t n = rows(a, "validate_vector_batch");
            if (rows(v, "validate_vector_batch") != n || d.size() != static_cast<py::ssize_t>(n))
            {
                throw std::invalid_argument("validate_vector_batch: starts, vectors and distances differ in length");
            }
            Environment &e = env_or_empty(env);
            vb_env *eh = e.handle();
            std::vector<uint32_t> bits((n + 31) / 32);
            {
                py::gil_scoped_release nogil;
                check(vb_validate_vector(e.ctx(), id, base.data(), eh, a.data(), v.data(), d.data(), n, res, bits.data(), nullptr));
            }
            return unpack(bits, n);
        }

        py::array_t<float> fk_batch(const FArray &q)
        {
            const size_t n = rows(q, "fk_batch");
            py::array_t<float> out({static_cast<py::ssize_t>(n), static_cast<py::ssize_t>(n_spheres), static_cast<py::ssize_t>(4)});
            float *o = out.mutable_data();
            vb_ctx *ctx = default_context()->h;
            {
                py::gil_scoped_release nogil;
                check(vb_sphere_fk(ctx, id, base.data(), q.data(), n, o));
            }
            return out;
        }

        std::vector<Sphere> fk(const FArray &q)
        {
            if (rows(q, "fk") != 1)
            {
                throw std::invalid_argument("fk: one configuration expected (fk_batch takes many)");
            }
            std::vector<float> s(static_cast<size_t>(n_spheres) * 4);
            check(vb_sphere_fk(default_context()->h, id, base.data(), q.data(), 1, s.data()));
            std::vector<Sphere> out;
            for (int k = 0; k < n_spheres; ++k)
            {
                out.push_back(Sphere{s[4 * k], s[4 * k + 1], s[4 * k + 2], s[4 * k + 3], ""});
            }
            return out;
        }

        std::vector<std::vector<std::string>> sphere_validity(const FArray &q, Environment *env)
        {
            if (rows(q, "sphere_validity") != 1)
            {
                throw std::invalid_argument("sphere_validity: one configuration expected");
            }
            Environment &e = env_or_empty(env);
            vb_env *eh = e.handle();
            const auto names = e.object_names();
            const uint32_t cap = static_cast<uint32_t>(names.empty() ? 1 : names.size());
            std::vector<uint32_t> hits(static_cast<size_t>(n_spheres) * cap), counts(n_spheres);
            check(vb_sphere_validity(e.ctx(), id, base.data(), eh, q.data(), 1, cap, hits.data(), counts.data()));
            std::vector<std::vector<std::string>> out(n_spheres);
            for (int k = 0; k < n_spheres; ++k)
            {
                for (uint32_t i = 0; i < counts[k] && i < cap; ++i)
                {
                    out[k].push_back(names.at(hits[static_cast<size_t>(k) * cap + i]));
                }
            }
            return out;
        }
    };

    // ---- the fork's OO multi-robot layer (robots/robot_factory.hh, MR_DEV_PLAN.md) ---------------------------------------
    // RobotInterface: one Panda of the 3 x 3 grid (robots/panda_grid.hh:10-41), identified by name and base position
    struct GridRobot
    {
        std::string name;
        std::shared_ptr<Robot> robot;
        std::string get_name() const { return name; }
        std::array<float, 3> get_base_position() const { return robot->base; }
    };

    struct RobotFactory
    {
        static std::vector<std::string> variants()
        {
            std::vector<std::string> out;
            for (int x = 0; x < 3; ++x)
            {
                for (int y = 0; y < 3; ++y)
                {
                    out.push_back("panda_" + std::to_string(x) + "_" + std::to_string(y));
                }
            }
            return out;
        }
        // robots/robot_factory.hh:66-85 (create_panda_grid): "panda_X_Y" -> base (X, Y, 0.05) metres.  The generated FK the
        // kernels run is PandaBase<X*100, Y*100, 0> (robots/panda_grid.hh: z = 0; the factory's comment says 5 cm), so z = 0.
        static GridRobot create_panda_grid(const std::string &variant)
        {
            int x = -1, y = -1;
            if (std::sscanf(variant.c_str(), "panda_%d_%d", &x, &y) != 2 || x < 0 || x > 2 || y < 0 || y > 2)
            {
                throw std::runtime_error("Unknown Panda grid variant: " + variant);
            }
            return GridRobot{variant, std::make_shared<Robot>("panda", std::array<float, 3>{static_cast<float>(x), static_cast<float>(y), 0.f})};
        }
        // robots/robot_factory.hh:33-64 (create_panda): nearest grid position, error when it is not on the grid
        static GridRobot create_panda(float x, float y, float z)
        {
            const int gx = static_cast<int>(std::lround(x * 100)), gy = static_cast<int>(std::lround(y * 100)),
                      gz = static_cast<int>(std::lround(z * 100));
            if (gx % 100 != 0 || gy % 100 != 0 || gx < 0 || gx > 200 || gy < 0 || gy > 200 || (gz != 0 && gz != 5))
            {
                throw std::runtime_error("Position (" + std::to_string(x) + ", " + std::to_string(y) + ", " + std::to_string(z) +
                                         ") not supported in grid");
            }
            return create_panda_grid("panda_" + std::to_string(gx / 100) + "_" + std::to_string(gy / 100));
        }
    };

    // MR_DEV_PLAN.md:176-198: robots + one base environment; inter-robot collision checks on composite configurations
    class MultiRobotEnvironment
    {
    public:
        MultiRobotEnvironment() : env_(std::make_shared<Environment>()) {}
        void add_robot(const GridRobot &r) { robots_.push_back(r); }
        void remove_robot(size_t i)
        {
            if (i >= robots_.size())
            {
                throw std::out_of_range("remove_robot: index out of range");
            }
            robots_.erase(robots_.begin() + static_cast<std::ptrdiff_t>(i));
        }
        size_t size() const { return robots_.size(); }
        void add_obstacle(const Sphere &s) { env_->add_sphere(s); }
        void add_cuboid(const Cuboid &c) { env_->add_cuboid(c); }
        void add_capsule(const Cylinder &c) { env_->add_capsule(c); }
        std::shared_ptr<Environment> environment() const { return env_; }

        // configs float32[n][n_robots][7].  inter_only: true = no sphere of any robot overlaps a sphere of another
        // (check_inter_robot_collisions negated); false = additionally every robot is valid on its own (environment + self
        // collision): the composite validity a multi-robot planner needs.  One batch per robot pair.
        py::array_t<bool> batch(const FArray &configs, bool inter_only)
        {
            const size_t nr = robots_.size();
            if (configs.ndim() != 3 || static_cast<size_t>(configs.shape(1)) != nr || configs.shape(2) != 7)
            {
                throw std::invalid_argument("expected float32[n, " + std::to_string(nr) + ", 7]");
            }
            const size_t n = static_cast<size_t>(configs.shape(0));
            std::vector<uint32_t> all((n + 31) / 32, 0xffffffffu), bits((n + 31) / 32);
            std::vector<std::vector<float>> q(nr, std::vector<float>(n * 7));
            for (size_t i = 0; i < n; ++i)
            {
                for (size_t r = 0; r < nr; ++r)
                {
                    std::copy_n(configs.data() + (i * nr + r) * 7, 7, q[r].data() + i * 7);
                }
            }
            vb_env *eh = env_->handle();
            {
                py::gil_scoped_release nogil;
                if (nr == 1 && !inter_only)
                {
                    check(vb_fkcc(env_->ctx(), robots_[0].robot->id, robots_[0].robot->base.data(), eh, q[0].data(), n, all.data(), nullptr));
                }
                for (size_t a = 0; a < nr; ++a)
                {
                    for (size_t b = a + 1; b < nr; ++b)
                    {
                        const Robot &ra = *robots_[a].robot, &rb = *robots_[b].robot;
                        if (inter_only)
                        {
                            check(vb_inter_robot(env_->ctx(), ra.id, ra.base.data(), rb.id, rb.base.data(), q[a].data(), q[b].data(), n,
                                                 bits.data()));
                        }
                        else
                        {
                            check(vb_validate_pairs(env_->ctx(), ra.id, ra.base.data(), rb.id, rb.base.data(), eh, q[a].data(),
                                                    q[b].data(), n, bits.data()));
                        }
                        for (size_t w = 0; w < all.size(); ++w)
                        {
                            all[w] &= bits[w];
                        }
                    }
                }
            }
            return Robot::unpack(all, n);
        }

        // MR_DEV_PLAN.md:190: true iff some pair of robots collides at this composite configuration
        bool check_inter_robot_collisions(const std::vector<std::vector<float>> &configs)
        {
            if (configs.size() != robots_.size())
            {
                throw std::invalid_argument("one configuration per robot expected");
            }
            FArray flat({static_cast<py::ssize_t>(1), static_cast<py::ssize_t>(robots_.size()), static_cast<py::ssize_t>(7)});
            for (size_t r = 0; r < configs.size(); ++r)
            {
                if (configs[r].size() != 7)
                {
                    throw std::invalid_argument("7 joint values per Panda expected");
                }
                std::copy_n(configs[r].data(), 7, flat.mutable_data() + r * 7);
            }
            return !batch(flat, true).at(0);
        }

    private:
        std::vector<GridRobot> robots_;
        std::shared_ptr<Environment> env_;
    };

    // src/vamp/__init__.py:135-181 (problem_dict_to_vamp): MotionBenchMaker problem dictionary -> Environment
    std::shared_ptr<Environment> problem_dict_to_vamp(const py::dict &problem, const std::vector<std::string> &ignore)
    {
        auto env = std::make_shared<Environment>();
        auto ignored = [&](const py::handle &obj) {
            const std::string name = obj.contains("name") ? py::cast<std::string>(obj["name"]) : std::string();
            return std::find(ignore.begin(), ignore.end(), name) != ignore.end();
        };
        auto name_of = [](const py::handle &obj) { return obj.contains("name") ? py::cast<std::string>(obj["name"]) : std::string(); };
        auto vec3 = [](const py::handle &v) { return py::cast<std::array<float, 3>>(v); };
        auto list_of = [&](const char *key) { return problem.contains(key) ? py::cast<py::list>(problem[key]) : py::list(); };
        for (const auto &obj : list_of("sphere"))
        {
            if (!ignored(obj))
            {
                const auto p = vec3(obj["position"]);
                env->add_sphere(Sphere{p[0], p[1], p[2], py::cast<float>(obj["radius"]), name_of(obj)});
            }
        }
        // "box" problems approximate their cylinders by boxes (a top-down grasp of the cylinder, __init__.py:146-160)
        const bool box_problem = problem.contains("problem") && py::cast<std::string>(problem["problem"]) == "box";
        for (const auto &obj : list_of("cylinder"))
        {
            if (ignored(obj))
            {
                continue;
            }
            const float radius = py::cast<float>(obj["radius"]), length = py::cast<float>(obj["length"]);
            if (box_problem)
            {
                Cuboid c = make_cuboid(vec3(obj["position"]), vec3(obj["orientation_euler_xyz"]), {radius, radius, length / 2.f});
                c.name = name_of(obj);
                env->add_cuboid(c);
            }
            else
            {
                Cylinder c = make_cylinder(vec3(obj["position"]), vec3(obj["orientation_euler_xyz"]), radius, length);
                c.name = name_of(obj);
                env->add_capsule(c);
            }
        }
        for (const auto &obj : list_of("box"))
        {
            if (!ignored(obj))
            {
                Cuboid c = make_cuboid(vec3(obj["position"]), vec3(obj["orientation_euler_xyz"]), vec3(obj["half_extents"]));
                c.name = name_of(obj);
                env->add_cuboid(c);
            }
        }
        return env;
    }

    void bind_robot(py::module_ &m, const char *module_name, const char *robot_name, std::array<float, 3> base)
    {
        auto sub = m.def_submodule(module_name);
        auto robot = std::make_shared<Robot>(robot_name, base);
        using namespace py::literals;
        sub.def("dimension", [robot] { return robot->dim; });
        sub.def("resolution", [robot] { return robot->resolution; });
        sub.def("n_spheres", [robot] { return robot->n_spheres; });
        sub.def("lower_bounds", [robot] { return robot->lo; });
        sub.def("upper_bounds", [robot] {
            std::vector<float> hi(robot->dim);
            for (int j = 0; j < robot->dim; ++j)
            {
                hi[j] = robot->lo[j] + robot->range[j];
            }
            return hi;
        });
        sub.def(
            "validate", [robot](const FArray &q, Environment *env) { return robot->validate(q, env); }, "configuration"_a,
            "environment"_a = py::none(), "Check if a configuration is valid. Returns true if valid.");
        sub.def(
            "sphere_validity", [robot](const FArray &q, Environment *env) { return robot->sphere_validity(q, env); },
            "configuration"_a, "environment"_a = py::none(), "Check which spheres of a robot configuration are in collision.");
        sub.def(
            "fk", [robot](const FArray &q) { return robot->fk(q); }, "configuration"_a,
            "Computes the forward kinematics of the robot. Returns array of all collision sphere positions.");
        sub.def(
            "validate_batch", [robot](const FArray &q, Environment *env) { return robot->validate_batch(q, env); },
            "configurations"_a, "environment"_a = py::none(), "validate() per row of float32[n, dim] -> bool[n] (one device batch).");
        sub.def(
            "fkcc_batch", [robot](const FArray &q, Environment *env) { return robot->fkcc_batch(q, env); }, "configurations"_a,
            "environment"_a = py::none(), "Robot::fkcc per row (no joint-limit test) -> bool[n].");
        sub.def(
            "validate_motion_batch",
            [robot](const FArray &a, const FArray &b, Environment *env, int res) { return robot->validate_motion_batch(a, b, env, res); },
            "starts"_a, "goals"_a, "environment"_a = py::none(), "resolution"_a = 0,
            "planning::validate_motion per edge -> bool[n] (one device batch).");
        sub.def(
            "validate_vector_batch",
            [robot](const FArray &a, const FArray &v, const FArray &d, Environment *env, int res) {
                return robot->validate_vector_batch(a, v, d, env, res);
            },
            "starts"_a, "vectors"_a, "distances"_a, "environment"_a = py::none(), "resolution"_a = 0,
            "planning::validate_vector(start, vector, distance) per edge -> bool[n].");
        sub.def(
            "fk_batch", [robot](const FArray &q) { return robot->fk_batch(q); }, "configurations"_a,
            "Robot::sphere_fk per row -> float32[n, n_spheres, 4] (x, y, z, r).");
    }
}  // namespace

PYBIND11_MODULE(_core_ext, m)
{
    using namespace py::literals;
    m.doc() = "B200 implementation of mr-vamp's fkcc / validate_motion path on the surface of vamp._core_ext";
    m.def("version", [] { return std::string(vb_version()); });

    py::class_<Sphere>(m, "Sphere")
        .def(py::init([](std::array<float, 3> c, float r) { return Sphere{c[0], c[1], c[2], r, ""}; }), "center"_a, "radius"_a)
        .def_readonly("x", &Sphere::x)
        .def_readonly("y", &Sphere::y)
        .def_readonly("z", &Sphere::z)
        .def_readonly("r", &Sphere::r)
        .def_property_readonly("position", [](const Sphere &s) { return std::array<float, 3>{s.x, s.y, s.z}; })
        .def_readwrite("name", &Sphere::name);
    py::class_<Cuboid>(m, "Cuboid")
        .def(py::init(&make_cuboid), "center"_a, "euler_xyz"_a, "half_extents"_a)
        .def_readwrite("name", &Cuboid::name)
        .def("raw", [](const Cuboid &c) { return c.raw; });
    py::class_<Cylinder>(m, "Cylinder")
        .def(py::init(&make_cylinder), "center"_a, "euler_xyz"_a, "radius"_a, "length"_a)
        .def(py::init(&cylinder_from_endpoints), "endpoint1"_a, "endpoint2"_a, "radius"_a)
        .def_readwrite("name", &Cylinder::name)
        .def("raw", [](const Cylinder &c) { return c.raw; });
    py::class_<HeightField>(m, "HeightField")
        .def(py::init([](std::array<float, 3> center, std::array<float, 3> scaling, const FArray &data) {
                 if (data.ndim() != 2)
                 {
                     throw std::invalid_argument("HeightField: float32[yd, xd] expected");
                 }
                 HeightField h;  // collision/factory.hh:365-386: scales are stored inverted
                 h.off_scale = {center[0], center[1], center[2], 1.f / scaling[0], 1.f / scaling[1], 1.f / scaling[2]};
                 h.yd = static_cast<size_t>(data.shape(0));
                 h.xd = static_cast<size_t>(data.shape(1));
                 h.data.assign(data.data(), data.data() + data.size());
                 return h;
             }),
             "center"_a, "scaling"_a, "data"_a)
        .def_readwrite("name", &HeightField::name);
    py::class_<Attachment>(m, "Attachment")
        .def(py::init([](std::array<float, 3> p, std::array<float, 4> q) {
                 return Attachment{{p[0], p[1], p[2], q[0], q[1], q[2], q[3]}, {}};
             }),
             "position"_a, "quaternion_xyzw"_a)
        .def("add_spheres", [](Attachment &a, const std::vector<Sphere> &s) { a.spheres.insert(a.spheres.end(), s.begin(), s.end()); });

    py::class_<Environment, std::shared_ptr<Environment>>(m, "Environment")
        .def(py::init([] { return std::make_shared<Environment>(); }))
        .def("add_sphere", &Environment::add_sphere)
        .def("add_cuboid", &Environment::add_cuboid)
        .def("add_capsule", &Environment::add_capsule)
        .def("add_heightfield", &Environment::add_heightfield)
        .def("add_pointcloud", &Environment::add_pointcloud, "pointcloud"_a, "r_min"_a, "r_max"_a, "r_point"_a,
             "Builds the CAPT of the cloud on the device; returns the build time in nanoseconds.")
        .def("attach", &Environment::attach)
        .def("detach", &Environment::detach);

    m.def("problem_dict_to_vamp", &problem_dict_to_vamp, "problem"_a, "ignore_names"_a = std::vector<std::string>{},
          "MotionBenchMaker problem dictionary -> Environment (src/vamp/__init__.py:135-181)");

    py::class_<GridRobot>(m, "Robot")
        .def("get_name", &GridRobot::get_name)
        .def("get_base_position", &GridRobot::get_base_position)
        .def("fk", [](GridRobot &r, const FArray &q) { return r.robot->fk(q); })
        .def("fk_batch", [](GridRobot &r, const FArray &q) { return r.robot->fk_batch(q); })
        .def(
            "validate", [](GridRobot &r, const FArray &q, Environment *env) { return r.robot->validate(q, env); }, "configuration"_a,
            "environment"_a = py::none())
        .def(
            "validate_batch", [](GridRobot &r, const FArray &q, Environment *env) { return r.robot->validate_batch(q, env); },
            "configurations"_a, "environment"_a = py::none());
    py::class_<RobotFactory>(m, "RobotFactory")
        .def_static("create_panda", &RobotFactory::create_panda, "x"_a, "y"_a, "z"_a)
        .def_static("create_panda_grid", &RobotFactory::create_panda_grid, "variant"_a)
        .def_static("get_available_variants", &RobotFactory::variants);
    py::class_<MultiRobotEnvironment>(m, "MultiRobotEnvironment")
        .def(py::init<>())
        .def("add_robot", &MultiRobotEnvironment::add_robot)
        .def("remove_robot", &MultiRobotEnvironment::remove_robot)
        .def("__len__", &MultiRobotEnvironment::size)
        .def("add_obstacle", &MultiRobotEnvironment::add_obstacle)
        .def("add_cuboid", &MultiRobotEnvironment::add_cuboid)
        .def("add_capsule", &MultiRobotEnvironment::add_capsule)
        .def_property_readonly("environment", &MultiRobotEnvironment::environment)
        .def("check_inter_robot_collisions", &MultiRobotEnvironment::check_inter_robot_collisions, "configurations"_a,
             "True iff some pair of robots collides at this composite configuration (MR_DEV_PLAN.md:190).")
        .def(
            "inter_robot_clear_batch", [](MultiRobotEnvironment &e, const FArray &q) { return e.batch(q, true); }, "configurations"_a,
            "float32[n, n_robots, 7] -> bool[n]: no sphere of any robot overlaps a sphere of another.")
        .def(
            "validate_batch", [](MultiRobotEnvironment &e, const FArray &q) { return e.batch(q, false); }, "configurations"_a,
            "float32[n, n_robots, 7] -> bool[n]: every robot valid (environment + self collision) and no robot-robot overlap.");

    bind_robot(m, "sphere", "sphere", {0.f, 0.f, 0.f});
    bind_robot(m, "ur5", "ur5", {0.f, 0.f, 0.f});
    bind_robot(m, "fetch", "fetch", {0.f, 0.f, 0.f});
    bind_robot(m, "baxter", "baxter", {0.f, 0.f, 0.f});
    // robots/panda_grid.hh:10-41: `panda` is the fork's default PandaBase<200, 200, 0>; grid variants by name
    bind_robot(m, "panda", "panda", {2.f, 2.f, 0.f});
    bind_robot(m, "panda_0_0", "panda", {0.f, 0.f, 0.f});
    bind_robot(m, "panda_1_0", "panda", {1.f, 0.f, 0.f});
    bind_robot(m, "panda_2_2", "panda", {2.f, 2.f, 0.f});
}

// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Shared glue of the reference wrapper (see vref_api.h).
#pragma once
#include <algorithm>
#include <array>
#include <cmath>
#include <cstring>
#include <functional>
#include <map>
#include <string>
#include <thread>
#include <vector>

#include <vamp/collision/environment.hh>
#include <vamp/collision/validity.hh>
#include <vamp/planning/validate.hh>
#include <vamp/vector.hh>

#include "vref_api.h"

namespace vref
{
    constexpr std::size_t rake = vamp::FloatVectorWidth;
    static_assert(rake == 8, "the reference oracle is built for AVX2 (8 lanes)");
    using EnvF = vamp::collision::Environment<float>;
    using EnvV = vamp::collision::Environment<vamp::FloatVector<rake>>;

    struct RobotVT
    {
        int dim, n_spheres, resolution;
        void (*sphere_fk)(const float *, std::size_t, float *);
        void (*fkcc)(const EnvV &, const float *, std::size_t, std::size_t, uint8_t *);
        void (*fkcc_packed)(const EnvV &, const float *, std::size_t, std::size_t, uint8_t *);
        void (*fkcc_attach)(const EnvV &, const float *, std::size_t, std::size_t, uint8_t *);
        void (*motion)(const EnvV &, const float *, const float *, std::size_t, std::size_t, uint8_t *);
        void (*vector)(const EnvV &, const float *, const float *, const float *, std::size_t, std::size_t, uint8_t *);
        void (*validate)(const EnvV &, const float *, std::size_t, uint8_t *);
        void (*scale)(float *, std::size_t);
    };

    std::map<std::string, RobotVT> &registry();

    template <typename Robot>
    struct Impl
    {
        static constexpr int dim = Robot::dimension;
        using Block = typename Robot::template ConfigurationBlock<rake>;
        using Config = typename Robot::Configuration;

        static Config make_config(const float *q)
        {
            std::array<float, dim> a;
            std::copy_n(q, dim, a.begin());
            return Config(a);
        }

        static void sphere_fk(const float *cfg, std::size_t n, float *out)
        {
            constexpr int S = Robot::n_spheres;
            for (std::size_t i = 0; i < n; ++i)
            {
                Block b;
                for (int j = 0; j < dim; ++j)
                {
                    b[j] = cfg[i * dim + j];
                }
                typename Robot::template Spheres<rake> s;
                Robot::template sphere_fk<rake>(b, s);
                for (int k = 0; k < S; ++k)
                {
                    float *o = out + (i * S + k) * 4;
                    o[0] = s.x[{(std::size_t)k, 0}];
                    o[1] = s.y[{(std::size_t)k, 0}];
                    o[2] = s.z[{(std::size_t)k, 0}];
                    o[3] = s.r[{(std::size_t)k, 0}];
                }
            }
        }

        static void fkcc(const EnvV &env, const float *cfg, std::size_t lo, std::size_t hi, uint8_t *valid)
        {
            for (std::size_t i = lo; i < hi; ++i)
            {
                Block b;
                for (int j = 0; j < dim; ++j)
                {
                    b[j] = cfg[i * dim + j];
                }
                valid[i] = Robot::template fkcc<rake>(env, b) ? 1 : 0;
            }
        }

        static void fkcc_attach(const EnvV &env, const float *cfg, std::size_t lo, std::size_t hi, uint8_t *valid)
        {
            for (std::size_t i = lo; i < hi; ++i)
            {
                Block b;
                for (int j = 0; j < dim; ++j)
                {
                    b[j] = cfg[i * dim + j];
                }
                valid[i] = Robot::template fkcc_attach<rake>(env, b) ? 1 : 0;
            }
        }

        // blocks [lo, hi): block k holds configs 8k .. 8k+7, one per lane
        static void fkcc_packed(const EnvV &env, const float *cfg, std::size_t lo, std::size_t hi, uint8_t *valid)
        {
            for (std::size_t k = lo; k < hi; ++k)
            {
                Block b;
                for (int j = 0; j < dim; ++j)
                {
                    alignas(32) float lane[rake];
                    for (std::size_t l = 0; l < rake; ++l)
                    {
                        lane[l] = cfg[(k * rake + l) * dim + j];
                    }
                    b[j] = vamp::FloatVector<rake>(lane);
                }
                valid[k] = Robot::template fkcc<rake>(env, b) ? 1 : 0;
            }
        }

        static void
        motion(const EnvV &env, const float *a, const float *b, std::size_t lo, std::size_t hi, uint8_t *valid)
        {
            for (std::size_t i = lo; i < hi; ++i)
            {
                valid[i] = vamp::planning::validate_motion<Robot, rake, Robot::resolution>(
                               make_config(a + i * dim), make_config(b + i * dim), env) ?
                               1 :
                               0;
            }
        }

        static void vector(const EnvV &env, const float *a, const float *v, const float *dist, std::size_t lo,
                           std::size_t hi, uint8_t *valid)
        {
            for (std::size_t i = lo; i < hi; ++i)
            {
                valid[i] = vamp::planning::validate_vector<Robot, rake, Robot::resolution>(
                               make_config(a + i * dim), make_config(v + i * dim), dist[i], env) ?
                               1 :
                               0;
            }
        }

        // bindings/common.hh:172-182
        static void validate(const EnvV &env, const float *cfg, std::size_t n, uint8_t *valid)
        {
            for (std::size_t i = 0; i < n; ++i)
            {
                const auto c = make_config(cfg + i * dim);
                auto copy = c.trim();
                Robot::descale_configuration(copy);
                valid[i] = ((copy <= 1.F).all() and (copy >= 0.F).all() and
                            vamp::planning::validate_motion<Robot, rake, 1>(c, c, env)) ?
                               1 :
                               0;
            }
        }

        static void scale(float *cfg, std::size_t n)
        {
            for (std::size_t i = 0; i < n; ++i)
            {
                auto c = make_config(cfg + i * dim);
                Robot::scale_configuration(c);
                const auto arr = c.to_array();
                std::copy_n(arr.begin(), dim, cfg + i * dim);
            }
        }

        static RobotVT vt()
        {
            return RobotVT{
                dim,
                (int)Robot::n_spheres,
                (int)Robot::resolution,
                &sphere_fk,
                &fkcc,
                &fkcc_packed,
                &fkcc_attach,
                &motion,
                &vector,
                &validate,
                &scale};
        }
    };

    struct Registrar
    {
        Registrar(const char *name, RobotVT vt)
        {
            registry()[name] = vt;
        }
    };
}  // namespace vref

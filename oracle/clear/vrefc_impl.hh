// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Glue of the clearance oracle (see vrefc_instr.hh, vrefc_api.h).
// Included AFTER vrefc_instr.hh and the robot header.
#pragma once
#include <map>
#include <string>
#include <thread>
#include <vector>

#include <vamp/planning/validate.hh>

#include "vrefc_api.h"

namespace vrefc
{
    constexpr std::size_t rake = vamp::FloatVectorWidth;
    using EnvF = vamp::collision::Environment<float>;
    using EnvV = vamp::collision::Environment<vamp::FloatVector<rake>>;

    struct RobotVT
    {
        int dim;
        void (*fkcc)(const EnvV &, const float *, std::size_t, std::size_t, uint8_t *, float *);
        void (*motion)(const EnvV &, const float *, const float *, const float *, std::size_t, std::size_t, uint8_t *, float *);
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

        // one configuration broadcast over the lanes, as Helper::validate_configuration does (bindings/common.hh:172-182)
        static void fkcc(const EnvV &env, const float *cfg, std::size_t lo, std::size_t hi, uint8_t *valid, float *clear)
        {
            for (std::size_t i = lo; i < hi; ++i)
            {
                Block b;
                for (int j = 0; j < dim; ++j)
                {
                    b[j] = cfg[i * dim + j];
                }
                g_min_abs = 3.0e38f;
                const bool ok = env.attachments ? Robot::template fkcc_attach<rake>(env, b) : Robot::template fkcc<rake>(env, b);
                valid[i] = ok ? 1 : 0;
                clear[i] = g_min_abs;
            }
        }

        // validate_motion (dist == nullptr: second = goals) or validate_vector (second = vectors), planning/validate.hh
        static void motion(const EnvV &env, const float *a, const float *second, const float *dist, std::size_t lo,
                           std::size_t hi, uint8_t *valid, float *clear)
        {
            for (std::size_t i = lo; i < hi; ++i)
            {
                g_min_abs = 3.0e38f;
                const auto s = make_config(a + i * dim), t = make_config(second + i * dim);
                const bool ok = dist ? vamp::planning::validate_vector<Robot, rake, Robot::resolution>(s, t, dist[i], env) :
                                       vamp::planning::validate_motion<Robot, rake, Robot::resolution>(s, t, env);
                valid[i] = ok ? 1 : 0;
                clear[i] = g_min_abs;
            }
        }

        static RobotVT vt()
        {
            return RobotVT{dim, &fkcc, &motion};
        }
    };

    struct Registrar
    {
        Registrar(const char *name, RobotVT vt)
        {
            registry()[name] = vt;
        }
    };
}  // namespace vrefc

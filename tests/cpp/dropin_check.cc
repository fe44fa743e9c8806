// Drop-in check: the SAME mr-vamp C++ program (reference types, reference calls) evaluated by the
// reference's CPU code and by the B200 library through include/vamp_b200_dropin.hh.
// Built by __graft_entry__.build() when the reference checkout is present:
//   g++ -std=c++17 -O2 -mavx2 -mfma -Ioracle/shim -I$VAMP_SRC/src/impl -Iinclude tests/cpp/dropin_check.cc
//       -Lmr_vamp_b200/lib -lvamp_b200 -o build/dropin_check
// Prints one line per check and exits non-zero on a mismatch.
#include <cstdio>
#include <random>
#include <vector>

#include <vamp/collision/environment.hh>
#include <vamp/collision/validity.hh>
#include <vamp/planning/validate.hh>
#include <vamp/robots/panda_grid.hh>
#include <vamp/robots/ur5.hh>
#include <vamp/vector.hh>

#include "vamp_b200_dropin.hh"

namespace vc = vamp::collision;
constexpr std::size_t rake = vamp::FloatVectorWidth;

template <typename Robot>
int run(const char *label, const vc::Environment<float> &env, unsigned seed)
{
    using Block = typename Robot::template ConfigurationBlock<rake>;
    const vc::Environment<vamp::FloatVector<rake>> venv(env);  // what the reference does per call
    vamp::b200::DeviceEnvironment denv(env);                   // what the drop-in does once
    std::mt19937 rng(seed);
    std::uniform_real_distribution<float> u(0.f, 1.f);
    constexpr int dim = Robot::dimension;
    int fk_bad = 0, cc_diff = 0, mv_diff = 0;
    const int n_blocks = 400;
    std::vector<float> cfgs;
    for (int b = 0; b < n_blocks; ++b)
    {
        Block q;
        std::array<std::array<float, rake>, dim> rows;
        for (int j = 0; j < dim; ++j)
        {
            for (std::size_t l = 0; l < rake; ++l)
            {
                rows[j][l] = u(rng);
            }
            q[j] = vamp::FloatVector<rake>(rows[j]);
        }
        Robot::template scale_configuration_block<rake>(q);
        const bool want = Robot::template fkcc<rake>(venv, q);
        const bool got = vamp::b200::fkcc<Robot, rake>(denv, q);
        cc_diff += want != got;
        if (b < 20)
        {
            typename Robot::template Spheres<rake> a, c;
            Robot::template sphere_fk<rake>(q, a);
            vamp::b200::sphere_fk<Robot, rake>(q, c);
            for (std::size_t k = 0; k < Robot::n_spheres; ++k)
            {
                const auto ax = a.x[k].to_array(), cx = c.x[k].to_array(), az = a.z[k].to_array(), cz = c.z[k].to_array();
                for (std::size_t l = 0; l < rake; ++l)
                {
                    fk_bad += std::fabs(ax[l] - cx[l]) > 1e-5f || std::fabs(az[l] - cz[l]) > 1e-5f;
                }
            }
        }
        const auto lane0 = vamp::b200::unpack_block<Robot, rake>(q);
        cfgs.insert(cfgs.end(), lane0.begin(), lane0.begin() + 2 * dim);  // two configurations per block
    }
    // edges between consecutive sampled configurations, reference one by one vs one batch launch
    const std::size_t n_edges = cfgs.size() / dim - 1;
    const auto got = vamp::b200::validate_motion_batch<Robot>(denv, cfgs.data(), cfgs.data() + dim, n_edges);
    for (std::size_t e = 0; e < n_edges; ++e)
    {
        std::array<float, dim> a, b;
        std::copy_n(cfgs.data() + e * dim, dim, a.begin());
        std::copy_n(cfgs.data() + (e + 1) * dim, dim, b.begin());
        const typename Robot::Configuration ca(a), cb(b);
        const bool want = vamp::planning::validate_motion<Robot, rake, Robot::resolution>(ca, cb, venv);
        mv_diff += want != got[e];
        if (e < 8 && want != vamp::b200::validate_motion<Robot>(ca, cb, denv))
        {
            ++mv_diff;
        }
    }
    std::printf("%s: %d blocks, fkcc verdicts differing %d, sphere centres off by >1e-5 m: %d, %zu edges, edge verdicts differing %d\n",
                label, n_blocks, cc_diff, fk_bad, n_edges, mv_diff);
    // grazing contacts (|clearance| < 1e-5 m) may differ; more than a handful is a failure
    return (fk_bad == 0 && cc_diff <= 1 && mv_diff <= 2) ? 0 : 1;
}

int main()
{
    vc::Environment<float> cage;
    const float centers[14][3] = {{0.55f, 0, 0.25f},      {0.35f, 0.35f, 0.25f}, {0, 0.55f, 0.25f},    {-0.55f, 0, 0.25f},
                                  {-0.35f, -0.35f, 0.25f}, {0, -0.55f, 0.25f},    {0.35f, -0.35f, 0.25f}, {0.35f, 0.35f, 0.8f},
                                  {0, 0.55f, 0.8f},        {-0.35f, 0.35f, 0.8f}, {-0.55f, 0, 0.8f},    {-0.35f, -0.35f, 0.8f},
                                  {0, -0.55f, 0.8f},       {0.35f, -0.35f, 0.8f}};
    for (const auto &c : centers)
    {
        cage.spheres.emplace_back(c[0], c[1], c[2], 0.2f);
    }
    // one box and one capsule so every list type travels through the upload
    cage.z_aligned_cuboids.emplace_back(0.6f, 0.f, 0.05f, 1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f, 0.2f, 0.3f, 0.05f);
    cage.capsules.emplace_back(-0.5f, 0.4f, 0.2f, 0.3f, 0.1f, 0.4f, 0.05f, 1.f / (0.09f + 0.01f + 0.16f));
    cage.sort();
    int rc = 0;
    try
    {
        rc |= run<vamp::robots::Panda_0_0>("panda_0_0 / cage", cage, 1);
        rc |= run<vamp::robots::UR5>("ur5 / cage", cage, 2);
        vc::Environment<float> shifted;
        for (const auto &c : centers)
        {
            shifted.spheres.emplace_back(c[0] + 2.f, c[1] + 2.f, c[2], 0.2f);
        }
        shifted.sort();
        rc |= run<vamp::robots::Panda>("panda (base 2,2) / shifted cage", shifted, 3);
    }
    catch (const std::exception &e)
    {
        std::printf("error: %s\n", e.what());
        return 2;
    }
    std::printf(rc == 0 ? "dropin_check: OK\n" : "dropin_check: MISMATCH\n");
    return rc;
}

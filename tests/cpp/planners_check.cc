// Source-compatibility check: the reference's OWN planner-side code -- Path::validate (planning/plan.hh:155-168) and
// shortcut_path (planning/simplify.hh:115-140), compiled UNCHANGED from the checkout -- runs its edge checks on the
// B200 because include/vamp_b200_planners.hh was included first (the effect of `-include vamp_b200_planners.hh`).
// Every verdict is compared with the reference's CPU functions (vamp::planning::cpu_validate_*), and the simplified path with
// the same scan replayed on the CPU verdicts.
// Built by __graft_entry__.build() when the reference checkout is present (tests/cpp/shim holds TEST-ONLY stubs of
// Eigen / nigh declarations that planning/nn.hh needs to parse; the path under test never touches them):
//   g++ -std=c++17 -O2 -mavx2 -mfma -Itests/cpp/shim -Ioracle/shim -I$VAMP_SRC/src/impl -Iinclude
//       tests/cpp/planners_check.cc -Lmr_vamp_b200/lib -lvamp_b200 -o build/planners_check
#include "vamp_b200_planners.hh"  // FIRST: redirects validate_vector / validate_motion in what follows

#include <cstdio>
#include <random>
#include <vector>

#include <vamp/planning/plan.hh>      // unchanged reference code: Path<dim>::validate
#include <vamp/planning/simplify.hh>  // unchanged reference code: shortcut_path
#include <vamp/robots/panda_grid.hh>

namespace vc = vamp::collision;
namespace vp = vamp::planning;
constexpr std::size_t rake = vamp::FloatVectorWidth;
using Robot = vamp::robots::Panda_0_0;
using Config = Robot::Configuration;

static Config random_config(std::mt19937 &rng)
{
    std::uniform_real_distribution<float> u(0.f, 1.f);
    std::array<float, Robot::dimension> a;
    for (auto &v : a)
    {
        v = u(rng);
    }
    Config c(a);
    Robot::scale_configuration(c);
    return c;
}

int main()
{
    // scripts/sphere_cage_example.py:16-31 (a few of its spheres: some random paths stay valid)
    vc::Environment<float> env;
    const float centres[6][3] = {{0.55f, 0, 0.25f}, {0.35f, 0.35f, 0.25f}, {0, 0.55f, 0.25f},
                                 {-0.55f, 0, 0.25f}, {0.35f, 0.35f, 0.8f}, {0, 0.55f, 0.8f}};
    for (const auto &c : centres)
    {
        env.spheres.emplace_back(vc::Sphere<float>(c[0], c[1], c[2], 0.2f));
    }
    env.sort();
    const vc::Environment<vamp::FloatVector<rake>> venv(env);

    std::mt19937 rng(11);
    std::normal_distribution<float> step(0.f, 0.12f);
    int edge_diff = 0, edges = 0, path_diff = 0, shortcut_diff = 0, valid_paths = 0, shortened = 0;
    for (int p = 0; p < 40; ++p)
    {
        vp::Path<Robot::dimension> path;
        const auto first = random_config(rng).to_array();
        std::array<float, Robot::dimension> arr;
        std::copy_n(first.begin(), Robot::dimension, arr.begin());
        const int n = 4 + p % 9;
        for (int i = 0; i < n; ++i)
        {
            path.emplace_back(Config(arr));
            for (auto &v : arr)
            {
                v += step(rng);
            }
        }
        // every edge: device-backed function (what the reference code below calls) against the reference's CPU function
        bool cpu_valid = true;
        for (std::size_t i = 0; i + 1 < path.size(); ++i)
        {
            const bool want = vp::cpu_validate_motion<Robot, rake, Robot::resolution>(path[i], path[i + 1], venv);
            const bool got = vp::validate_motion<Robot, rake, Robot::resolution>(path[i], path[i + 1], venv);  // (renamed by the macro)
            edge_diff += want != got;
            ++edges;
            cpu_valid = cpu_valid && want;
        }
        // unchanged reference member function, now running on the device
        const bool gpu_valid = path.validate<Robot, rake>(venv);
        path_diff += gpu_valid != cpu_valid;
        valid_paths += cpu_valid;
        // unchanged reference shortcut_path on the device; the same scan on the CPU verdicts
        auto cpu_path = path;
        for (std::size_t i = 0; i + 2 < cpu_path.size(); ++i)
        {
            for (std::size_t j = cpu_path.size() - 1; j > i + 1; --j)
            {
                if (vp::cpu_validate_motion<Robot, rake, Robot::resolution>(cpu_path[i], cpu_path[j], venv))
                {
                    cpu_path.erase(cpu_path.begin() + i + 1, cpu_path.begin() + j);
                    break;
                }
            }
        }
        auto gpu_path = path;
        vp::shortcut_path<Robot, rake, Robot::resolution>(gpu_path, venv, vp::ShortcutSettings{});
        bool same = gpu_path.size() == cpu_path.size();
        for (std::size_t i = 0; same && i < gpu_path.size(); ++i)
        {
            same = gpu_path[i].to_array() == cpu_path[i].to_array();
        }
        shortcut_diff += !same;
        shortened += gpu_path.size() < path.size();
    }
    std::printf("planners_check: %d edges, %d differ; %d paths (%d valid), Path::validate differs on %d; shortcut_path shortened %d, "
                "differs on %d\n", edges, edge_diff, 40, valid_paths, path_diff, shortened, shortcut_diff);
    // a changed environment at the same address must be re-uploaded (fingerprint), not served from the cache
    vc::Environment<vamp::FloatVector<rake>> venv2(env);
    const auto q = random_config(rng);
    const bool before = vp::validate_motion<Robot, rake, Robot::resolution>(q, q, venv2);
    venv2.spheres.clear();
    const bool after_clear = vp::validate_motion<Robot, rake, Robot::resolution>(q, q, venv2);
    const bool want_clear = vp::cpu_validate_motion<Robot, rake, Robot::resolution>(q, q, venv2);
    std::printf("planners_check: cache refresh %s (before %d, after clearing the obstacles %d, cpu %d)\n",
                after_clear == want_clear ? "ok" : "STALE", before, after_clear, want_clear);
    if (edge_diff || path_diff || shortcut_diff || after_clear != want_clear || valid_paths == 0 || valid_paths == 40 || shortened == 0)
    {
        std::printf("planners_check: FAILED\n");
        return 1;
    }
    std::printf("planners_check: OK\n");
    return 0;
}

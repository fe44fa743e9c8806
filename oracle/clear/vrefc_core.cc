// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Dispatch of the clearance oracle (see vrefc_api.h).
#include "vrefc_instr.hh"
#include "vrefc_impl.hh"

namespace vrefc
{
    std::map<std::string, RobotVT> &registry()
    {
        static std::map<std::string, RobotVT> r;
        return r;
    }

    static const RobotVT *find(const char *name)
    {
        std::string s(name);
        if (s == "panda")
        {
            s = "panda_2_2";
        }
        auto it = registry().find(s);
        return it == registry().end() ? nullptr : &it->second;
    }

    template <typename Fn>
    static void parallel(std::size_t n, int threads, const EnvF &f, Fn fn)
    {
        const std::size_t T = std::max<std::size_t>(1, std::min<std::size_t>(threads > 0 ? threads : 1, n ? n : 1));
        std::vector<EnvV> envs;  // a private vector-form copy per worker (attachment poses are mutable state)
        for (std::size_t t = 0; t < T; ++t)
        {
            envs.emplace_back(f);
        }
        if (T == 1)
        {
            fn(envs[0], 0, n);
            return;
        }
        std::vector<std::thread> pool;
        const std::size_t chunk = (n + T - 1) / T;
        for (std::size_t t = 0; t < T; ++t)
        {
            const std::size_t lo = std::min(n, t * chunk), hi = std::min(n, lo + chunk);
            pool.emplace_back([&, t, lo, hi] { fn(envs[t], lo, hi); });
        }
        for (auto &th : pool)
        {
            th.join();
        }
    }
}  // namespace vrefc

extern "C" {

int vrefc_fkcc_clearance(const char *robot, const void *envf, const float *cfg, size_t n, uint8_t *valid, float *clearance,
                         int threads)
{
    const auto *vt = vrefc::find(robot);
    if (not vt or not envf)
    {
        return -1;
    }
    vrefc::parallel(n, threads, *static_cast<const vrefc::EnvF *>(envf),
                    [&](const vrefc::EnvV &ev, size_t lo, size_t hi) { vt->fkcc(ev, cfg, lo, hi, valid, clearance); });
    return 0;
}

int vrefc_motion_clearance(const char *robot, const void *envf, const float *a, const float *second, const float *dist,
                           size_t n, uint8_t *valid, float *clearance, int threads)
{
    const auto *vt = vrefc::find(robot);
    if (not vt or not envf)
    {
        return -1;
    }
    vrefc::parallel(n, threads, *static_cast<const vrefc::EnvF *>(envf), [&](const vrefc::EnvV &ev, size_t lo, size_t hi) {
        vt->motion(ev, a, second, dist, lo, hi, valid, clearance);
    });
    return 0;
}

}  // extern "C"

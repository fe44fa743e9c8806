// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Environment, numerics probes and dispatch of the
// reference wrapper (see vref_api.h).  Compiles the reference headers where they lie.
#include <chrono>

#include "vref_impl.hh"

#include <vamp/collision/filter.hh>

namespace vc = vamp::collision;
using FV = vamp::FloatVector<vref::rake>;

struct vref_env
{
    vref::EnvF f;
    unsigned next_name = 0;  // shapes are named by their insertion index ("0", "1", ...) for vref_sphere_validity
    mutable bool dirty = true;
    mutable std::vector<vref::EnvV> per_thread;  // private copies, one per worker thread

    const vref::EnvV &vec(std::size_t thread = 0, std::size_t n_threads = 1) const
    {
        if (dirty or per_thread.size() < n_threads)
        {
            per_thread.clear();
            for (std::size_t i = 0; i < std::max<std::size_t>(n_threads, 1); ++i)
            {
                per_thread.emplace_back(f);
            }
            dirty = false;
        }
        return per_thread[thread];
    }
};

namespace vref
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
    static void parallel(std::size_t n, int threads, const vref_env *env, Fn fn)
    {
        const std::size_t T = std::max(1, threads);
        env->vec(0, T);
        if (T == 1 or n < T)
        {
            fn(env->per_thread[0], 0, n);
            return;
        }
        std::vector<std::thread> pool;
        const std::size_t chunk = (n + T - 1) / T;
        for (std::size_t t = 0; t < T; ++t)
        {
            const std::size_t lo = std::min(n, t * chunk), hi = std::min(n, lo + chunk);
            pool.emplace_back([&, t, lo, hi] { fn(env->per_thread[t], lo, hi); });
        }
        for (auto &th : pool)
        {
            th.join();
        }
    }
}  // namespace vref

extern "C" {

void vref_trig(const float *x, size_t n, float *s, float *c)
{
    for (size_t i = 0; i < n; ++i)
    {
        FV v(x[i]);
        s[i] = v.sin()[{0, 0}];
        c[i] = v.cos()[{0, 0}];
    }
}

void vref_sqrt(const float *x, size_t n, float *out)
{
    for (size_t i = 0; i < n; ++i)
    {
        FV v(x[i]);
        out[i] = v.sqrt()[{0, 0}];
    }
}

vref_env *vref_env_create(void)
{
    return new vref_env();
}

void vref_env_destroy(vref_env *e)
{
    delete e;
}

void vref_env_add_sphere(vref_env *e, float x, float y, float z, float r)
{
    e->f.spheres.emplace_back(vc::Sphere<float>(x, y, z, r));
    e->f.spheres.back().name = std::to_string(e->next_name++);
    e->f.sort();
    e->dirty = true;
}

void vref_env_add_cuboid(vref_env *e, const float *c)
{
    vc::Cuboid<float> s(
        c[0], c[1], c[2], c[3], c[4], c[5], c[6], c[7], c[8], c[9], c[10], c[11], c[12], c[13], c[14]);
    s.name = std::to_string(e->next_name++);
    // bindings/environment.cc:120-127
    if (s.axis_3_z == 1.)
    {
        e->f.z_aligned_cuboids.emplace_back(s);
    }
    else
    {
        e->f.cuboids.emplace_back(s);
    }
    e->f.sort();
    e->dirty = true;
}

void vref_env_add_capsule(vref_env *e, const float *c)
{
    vc::Cylinder<float> s(c[0], c[1], c[2], c[3], c[4], c[5], c[6], c[7]);
    s.name = std::to_string(e->next_name++);
    // bindings/environment.cc:134-141
    if (s.xv == 0. and s.yv == 0.)
    {
        e->f.z_aligned_capsules.emplace_back(s);
    }
    else
    {
        e->f.capsules.emplace_back(s);
    }
    e->f.sort();
    e->dirty = true;
}

void vref_env_add_heightfield(vref_env *e, const float *o, size_t xd, size_t yd, const float *data)
{
    std::vector<float> d(data, data + xd * yd);
    e->f.heightfields.emplace_back(vc::HeightField<float>(o[0], o[1], o[2], o[3], o[4], o[5], xd, yd, d));
    e->f.heightfields.back().name = std::to_string(e->next_name++);
    e->dirty = true;
}

double vref_env_add_pointcloud(vref_env *e, const float *xyz, size_t n, float r_min, float r_max, float r_point)
{
    std::vector<vc::Point> pts(n);
    for (size_t i = 0; i < n; ++i)
    {
        pts[i] = {xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]};
    }
    const auto t0 = std::chrono::steady_clock::now();
    e->f.pointclouds.emplace_back(pts, r_min, r_max, r_point);
    const auto t1 = std::chrono::steady_clock::now();
    e->dirty = true;
    return std::chrono::duration<double, std::nano>(t1 - t0).count();
}

void vref_env_attach(vref_env *e, const float *tf, const float *spheres, size_t n)
{
    vc::Attachment<float> a(tf[0], tf[1], tf[2], tf[3], tf[4], tf[5], tf[6]);
    for (size_t i = 0; i < n; ++i)
    {
        a.spheres.emplace_back(vc::Sphere<float>(spheres[4 * i], spheres[4 * i + 1], spheres[4 * i + 2], spheres[4 * i + 3]));
    }
    e->f.attachments.emplace(a);
    e->dirty = true;
}

void vref_env_detach(vref_env *e)
{
    e->f.attachments.reset();
    e->dirty = true;
}

size_t vref_env_export(const vref_env *e, int kind, float *out)
{
    const auto &f = e->f;
    auto caps = [&](const std::vector<vc::Capsule<float>> &v)
    {
        if (out)
        {
            for (size_t i = 0; i < v.size(); ++i)
            {
                const auto &c = v[i];
                const float row[9] = {c.x1, c.y1, c.z1, c.xv, c.yv, c.zv, c.r, c.rdv, c.min_distance};
                std::memcpy(out + 9 * i, row, sizeof(row));
            }
        }
        return v.size();
    };
    auto cubs = [&](const std::vector<vc::Cuboid<float>> &v)
    {
        if (out)
        {
            for (size_t i = 0; i < v.size(); ++i)
            {
                const auto &c = v[i];
                const float row[16] = {
                    c.x,        c.y,        c.z,        c.axis_1_x, c.axis_1_y, c.axis_1_z, c.axis_2_x, c.axis_2_y,
                    c.axis_2_z, c.axis_3_x, c.axis_3_y, c.axis_3_z, c.axis_1_r, c.axis_2_r, c.axis_3_r, c.min_distance};
                std::memcpy(out + 16 * i, row, sizeof(row));
            }
        }
        return v.size();
    };
    switch (kind)
    {
        case 0:
            if (out)
            {
                for (size_t i = 0; i < f.spheres.size(); ++i)
                {
                    const auto &s = f.spheres[i];
                    const float row[5] = {s.x, s.y, s.z, s.r, s.min_distance};
                    std::memcpy(out + 5 * i, row, sizeof(row));
                }
            }
            return f.spheres.size();
        case 1:
            return caps(f.capsules);
        case 2:
            return caps(f.z_aligned_capsules);
        case 3:
            return cubs(f.cuboids);
        case 4:
            return cubs(f.z_aligned_cuboids);
        default:
            return 0;
    }
}

int vref_env_capt(const vref_env *e, size_t idx, vref_capt_view *o)
{
    if (idx >= e->f.pointclouds.size())
    {
        return -1;
    }
    const auto &c = e->f.pointclouds[idx];
    o->n_tests = c.tests.size();
    o->n_aff_starts = c.aff_starts.size();
    o->n_aff_vec = c.affordances[0].size();
    o->n_aabbs = c.aabbs.size();
    o->nlog2 = c.nlog2;
    o->r_min = c.r_min;
    o->r_max = c.r_max;
    o->r_point = c.r_point;
    for (int k = 0; k < 3; ++k)
    {
        o->aabb_top[k] = c.aabb_top.lower[k];
        o->aabb_top[3 + k] = c.aabb_top.upper[k];
    }
    o->tests = c.tests.data();
    o->aff_starts = c.aff_starts.data();
    o->aff_x = reinterpret_cast<const float *>(c.affordances[0].data());
    o->aff_y = reinterpret_cast<const float *>(c.affordances[1].data());
    o->aff_z = reinterpret_cast<const float *>(c.affordances[2].data());
    o->aabbs = c.aabbs.empty() ? nullptr : &c.aabbs.front().lower.front();
    return 0;
}

void vref_capt_query(const vref_env *e, size_t idx, const float *q, size_t n, uint8_t *hs, uint8_t *hc)
{
    const auto &c = e->f.pointclouds[idx];
    for (size_t i = 0; i < n; ++i)
    {
        const float *s = q + 4 * i;
        if (hs)
        {
            const std::array<FV, 3> ctr = {FV(s[0]), FV(s[1]), FV(s[2])};
            hs[i] = c.collides_simd(ctr, FV(s[3])) ? 1 : 0;
        }
        if (hc)
        {
            hc[i] = c.collides({s[0], s[1], s[2]}, s[3]) ? 1 : 0;
        }
    }
}

void vref_sphere_env(const vref_env *e, const float *q, size_t n, uint8_t *hit)
{
    const auto &ev = e->vec();
    for (size_t i = 0; i < n; ++i)
    {
        const float *s = q + 4 * i;
        hit[i] = vamp::sphere_environment_in_collision(ev, s[0], s[1], s[2], s[3]) ? 1 : 0;
    }
}

int vref_robot_info(const char *robot, int *dim, int *n_spheres, int *resolution)
{
    const auto *vt = vref::find(robot);
    if (not vt)
    {
        return -1;
    }
    *dim = vt->dim;
    *n_spheres = vt->n_spheres;
    *resolution = vt->resolution;
    return 0;
}

int vref_sphere_fk(const char *robot, const float *cfg, size_t n, float *out)
{
    const auto *vt = vref::find(robot);
    if (not vt)
    {
        return -1;
    }
    vt->sphere_fk(cfg, n, out);
    return 0;
}

int vref_fkcc(const char *robot, const vref_env *env, const float *cfg, size_t n, uint8_t *valid, int threads)
{
    const auto *vt = vref::find(robot);
    if (not vt)
    {
        return -1;
    }
    vref::parallel(
        n, threads, env, [&](const vref::EnvV &ev, size_t lo, size_t hi) { vt->fkcc(ev, cfg, lo, hi, valid); });
    return 0;
}

int vref_fkcc_attach(const char *robot, const vref_env *env, const float *cfg, size_t n, uint8_t *valid, int threads)
{
    const auto *vt = vref::find(robot);
    if (not vt or not env->f.attachments)
    {
        return -1;
    }
    vref::parallel(
        n,
        threads,
        env,
        [&](const vref::EnvV &ev, size_t lo, size_t hi) { vt->fkcc_attach(ev, cfg, lo, hi, valid); });
    return 0;
}

int vref_fkcc_packed(const char *robot, const vref_env *env, const float *cfg, size_t n, uint8_t *valid, int threads)
{
    const auto *vt = vref::find(robot);
    if (not vt or n % vref::rake)
    {
        return -1;
    }
    vref::parallel(
        n / vref::rake,
        threads,
        env,
        [&](const vref::EnvV &ev, size_t lo, size_t hi) { vt->fkcc_packed(ev, cfg, lo, hi, valid); });
    return 0;
}

int vref_validate_motion(
    const char *robot,
    const vref_env *env,
    const float *a,
    const float *b,
    size_t n,
    uint8_t *valid,
    int threads)
{
    const auto *vt = vref::find(robot);
    if (not vt)
    {
        return -1;
    }
    vref::parallel(
        n, threads, env, [&](const vref::EnvV &ev, size_t lo, size_t hi) { vt->motion(ev, a, b, lo, hi, valid); });
    return 0;
}

// collision/filter.hh:176-275 (filter_pointcloud): the reference's own point-cloud thinning, as src/vamp/pointcloud.py
// calls it before building the CAPT.  Returns the number of points kept; writes them when out != NULL.
namespace
{
    struct PointView
    {
        const float *p;
        size_t n;
        size_t shape(int) const { return n; }
        float operator()(size_t i, size_t k) const { return p[3 * i + k]; }
    };
}  // namespace

size_t vref_filter_pointcloud(const float *xyz, size_t n, float min_dist, float max_range, const float *origin,
                              const float *ws_min, const float *ws_max, int cull, float *out)
{
    const PointView pv{xyz, n};
    const auto kept = vc::filter_pointcloud(pv, min_dist, max_range, vc::Point{origin[0], origin[1], origin[2]},
                                            vc::Point{ws_min[0], ws_min[1], ws_min[2]}, vc::Point{ws_max[0], ws_max[1], ws_max[2]},
                                            cull != 0);
    if (out)
    {
        for (size_t i = 0; i < kept.size(); ++i)
        {
            out[3 * i] = kept[i][0];
            out[3 * i + 1] = kept[i][1];
            out[3 * i + 2] = kept[i][2];
        }
    }
    return kept.size();
}

// Helper::sphere_validate (bindings/common.hh:155-170): fk, then sphere_environment_get_collisions per sphere
int vref_sphere_validity(const char *robot, const vref_env *env, const float *cfg, uint32_t cap, uint32_t *hits,
                         uint32_t *counts)
{
    const auto *vt = vref::find(robot);
    if (not vt)
    {
        return -1;
    }
    std::vector<float> spheres(static_cast<size_t>(vt->n_spheres) * 4);
    vt->sphere_fk(cfg, 1, spheres.data());
    const auto &ev = env->vec();
    for (int k = 0; k < vt->n_spheres; ++k)
    {
        const float *s = spheres.data() + 4 * k;
        const auto names = vamp::sphere_environment_get_collisions<>(ev, s[0], s[1], s[2], s[3]);
        counts[k] = static_cast<uint32_t>(names.size());
        for (size_t i = 0; i < names.size() and i < cap; ++i)
        {
            hits[static_cast<size_t>(k) * cap + i] = static_cast<uint32_t>(std::stoul(names[i]));
        }
    }
    return 0;
}

int vref_validate_vector(const char *robot, const vref_env *env, const float *a, const float *v, const float *dist,
                         size_t n, uint8_t *valid, int threads)
{
    const auto *vt = vref::find(robot);
    if (not vt)
    {
        return -1;
    }
    vref::parallel(n, threads, env, [&](const vref::EnvV &ev, size_t lo, size_t hi) { vt->vector(ev, a, v, dist, lo, hi, valid); });
    return 0;
}

const void *vref_env_float_ptr(const vref_env *env)
{
    return &env->f;
}

int vref_validate(const char *robot, const vref_env *env, const float *cfg, size_t n, uint8_t *valid)
{
    const auto *vt = vref::find(robot);
    if (not vt)
    {
        return -1;
    }
    vt->validate(env->vec(), cfg, n, valid);
    return 0;
}

int vref_scale(const char *robot, float *cfg, size_t n)
{
    const auto *vt = vref::find(robot);
    if (not vt)
    {
        return -1;
    }
    vt->scale(cfg, n);
    return 0;
}

}  // extern "C"

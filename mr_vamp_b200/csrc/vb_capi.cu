// libvamp_b200.so — implementation of the C ABI declared in include/vamp_b200.h.
//
// Host side of the fkcc path: device contexts and streams, the device-resident environment
// (replaces collision::Environment<float> + its per-call vector conversion, collision/environment.hh),
// batch sharding over the devices of a context, and the launches of the per-robot kernels
// (vb_robot.cu).  No CPU implementation of any query exists here: every query either runs on a
// CUDA device or fails.
#include "../../include/vamp_b200.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdlib>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include <dlfcn.h>
#include <nccl.h>  // types only: the library is dlopen()ed by vb_comm_* so the core path does not need it

#include "vb_capt_build.h"
#include "vb_robot_ops.h"

namespace
{
    thread_local std::string g_error;

    int fail(int code, const char *fmt, ...)
    {
        char buf[512];
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(buf, sizeof(buf), fmt, ap);
        va_end(ap);
        g_error = buf;
        return code;
    }

#define VB_CUDA(expr)                                                                                \
    do                                                                                               \
    {                                                                                                \
        cudaError_t err__ = (expr);                                                                  \
        if (err__ != cudaSuccess)                                                                    \
        {                                                                                            \
            return fail(VB_ERR_CUDA, "%s: %s (%s:%d)", #expr, cudaGetErrorString(err__), __FILE__, __LINE__); \
        }                                                                                            \
    } while (0)

#define VB_NCCL(expr)                                                                                \
    do                                                                                               \
    {                                                                                                \
        ncclResult_t err__ = (expr);                                                                 \
        if (err__ != ncclSuccess)                                                                    \
        {                                                                                            \
            return fail(VB_ERR_NCCL, "%s: %s (%s:%d)", #expr, nccl().GetErrorString(err__), __FILE__, __LINE__); \
        }                                                                                            \
    } while (0)

#define VB_TRY(expr)            \
    do                          \
    {                           \
        int rc__ = (expr);      \
        if (rc__ != VB_OK)      \
        {                       \
            return rc__;        \
        }                       \
    } while (0)

    const vb::RobotOps *ops_of(vb_robot r)
    {
        switch (r)
        {
            case VB_SPHERE:
                return vb::robot_ops_sphere();
            case VB_UR5:
                return vb::robot_ops_ur5();
            case VB_PANDA:
                return vb::robot_ops_panda();
            case VB_FETCH:
                return vb::robot_ops_fetch();
            case VB_BAXTER:
                return vb::robot_ops_baxter();
        }
        return nullptr;
    }

    // A grow-only device buffer.
    struct DevBuf
    {
        void *p = nullptr;
        size_t cap = 0;
        cudaError_t reserve(size_t bytes)
        {
            if (bytes <= cap)
            {
                return cudaSuccess;
            }
            if (p)
            {
                cudaFree(p);
                p = nullptr;
                cap = 0;
            }
            const size_t want = std::max<size_t>(bytes, 256);
            cudaError_t e = cudaMalloc(&p, want);
            if (e == cudaSuccess)
            {
                cap = want;
            }
            return e;
        }
        void release()
        {
            if (p)
            {
                cudaFree(p);
            }
            p = nullptr;
            cap = 0;
        }
        template <class T>
        T *as() const
        {
            return static_cast<T *>(p);
        }
    };

    struct Slot
    {
        int device = 0;
        int sm_count = 0;
        cudaStream_t stream = nullptr;
        cudaStream_t stream2 = nullptr;  // second lane of the host-buffer pipeline (copy of chunk k+1 under the kernel of chunk k)
        cudaEvent_t ev0 = nullptr, ev1 = nullptr, evj = nullptr;
        DevBuf in_a, in_a2, in_b, in_b2, in_d, in_d2, bits, margin, out, scan_tmp, att;
    };
}  // namespace

struct vb_ctx
{
    std::vector<Slot> slots;
    float sphere_radius = 0.2f;  // robots/sphere/fk.hh:16
    std::atomic<uint64_t> launches{0};
    ncclComm_t comm = nullptr;
    int comm_ranks = 0;
};

namespace
{
    // NCCL entry points, resolved on first use (libnccl.so.2 — torch's bundled copy when the host
    // process already loaded it, the system one otherwise).
    struct NcclApi
    {
        bool ok = false;
        ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
        ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
        ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
        ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
        const char *(*GetErrorString)(ncclResult_t) = nullptr;
    };

    NcclApi &nccl()
    {
        static NcclApi api = [] {
            NcclApi a;
            void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
            if (!h)
            {
                h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
            }
            if (!h)
            {
                return a;
            }
            a.GetUniqueId = reinterpret_cast<decltype(a.GetUniqueId)>(dlsym(h, "ncclGetUniqueId"));
            a.CommInitRank = reinterpret_cast<decltype(a.CommInitRank)>(dlsym(h, "ncclCommInitRank"));
            a.AllGather = reinterpret_cast<decltype(a.AllGather)>(dlsym(h, "ncclAllGather"));
            a.CommDestroy = reinterpret_cast<decltype(a.CommDestroy)>(dlsym(h, "ncclCommDestroy"));
            a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(dlsym(h, "ncclGetErrorString"));
            a.ok = a.GetUniqueId && a.CommInitRank && a.AllGather && a.CommDestroy && a.GetErrorString;
            return a;
        }();
        return api;
    }

    int check_slot(const vb_ctx *ctx, int slot)
    {
        if (!ctx)
        {
            return fail(VB_ERR_ARG, "null context");
        }
        if (slot < 0 || slot >= static_cast<int>(ctx->slots.size()))
        {
            return fail(VB_ERR_ARG, "device slot %d out of range (context has %zu)", slot, ctx->slots.size());
        }
        return VB_OK;
    }
}  // namespace

extern "C" {

const char *vb_last_error(void)
{
    return g_error.c_str();
}

const char *vb_version(void)
{
    return "vamp_b200 0.1 (sm_100a)";
}

int vb_ctx_create(vb_ctx **out, const int *device_ids, int n_devices)
{
    if (!out || n_devices < 0 || (n_devices > 0 && !device_ids))
    {
        return fail(VB_ERR_ARG, "vb_ctx_create: bad arguments");
    }
    int available = 0;
    cudaError_t e = cudaGetDeviceCount(&available);
    if (e != cudaSuccess || available == 0)
    {
        return fail(VB_ERR_CUDA, "no usable CUDA device (%s); this library has no CPU path",
                    e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    }
    std::unique_ptr<vb_ctx> ctx(new vb_ctx());
    const int n = n_devices == 0 ? 1 : n_devices;
    for (int i = 0; i < n; ++i)
    {
        Slot s;
        s.device = n_devices == 0 ? 0 : device_ids[i];
        if (s.device < 0 || s.device >= available)
        {
            return fail(VB_ERR_ARG, "device id %d not in [0, %d)", s.device, available);
        }
        VB_CUDA(cudaSetDevice(s.device));
        cudaDeviceProp prop;
        VB_CUDA(cudaGetDeviceProperties(&prop, s.device));
        if (prop.major < 10)
        {
            return fail(VB_ERR_CUDA, "device %d is sm_%d%d; the kernels are built for sm_100a only", s.device,
                        prop.major, prop.minor);
        }
        s.sm_count = prop.multiProcessorCount;
        VB_CUDA(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
        VB_CUDA(cudaStreamCreateWithFlags(&s.stream2, cudaStreamNonBlocking));
        VB_CUDA(cudaEventCreateWithFlags(&s.evj, cudaEventDisableTiming));
        VB_CUDA(cudaEventCreate(&s.ev0));
        VB_CUDA(cudaEventCreate(&s.ev1));
        ctx->slots.push_back(s);
    }
    *out = ctx.release();
    return VB_OK;
}

void vb_ctx_destroy(vb_ctx *ctx)
{
    if (!ctx)
    {
        return;
    }
    if (ctx->comm)
    {
        nccl().CommDestroy(ctx->comm);
    }
    for (auto &s : ctx->slots)
    {
        cudaSetDevice(s.device);
        cudaStreamSynchronize(s.stream);
        cudaStreamSynchronize(s.stream2);
        cudaStreamDestroy(s.stream2);
        cudaEventDestroy(s.evj);
        for (DevBuf *b : {&s.in_a, &s.in_a2, &s.in_b, &s.in_b2, &s.in_d, &s.in_d2, &s.bits, &s.margin, &s.out, &s.scan_tmp, &s.att})
        {
            b->release();
        }
        cudaEventDestroy(s.ev0);
        cudaEventDestroy(s.ev1);
        cudaStreamDestroy(s.stream);
    }
    delete ctx;
}

int vb_ctx_device_count(const vb_ctx *ctx)
{
    return ctx ? static_cast<int>(ctx->slots.size()) : 0;
}

int vb_ctx_sm_count(const vb_ctx *ctx, int slot)
{
    return check_slot(ctx, slot) == VB_OK ? ctx->slots[slot].sm_count : 0;
}

int vb_sync(vb_ctx *ctx)
{
    if (!ctx)
    {
        return fail(VB_ERR_ARG, "null context");
    }
    for (auto &s : ctx->slots)
    {
        VB_CUDA(cudaSetDevice(s.device));
        VB_CUDA(cudaStreamSynchronize(s.stream));
        VB_CUDA(cudaStreamSynchronize(s.stream2));
    }
    return VB_OK;
}

int vb_device_pci_bus_id(int device, char *out, int len)
{
    if (!out || len < 16)
    {
        return fail(VB_ERR_ARG, "vb_device_pci_bus_id: need a buffer of at least 16 bytes");
    }
    VB_CUDA(cudaDeviceGetPCIBusId(out, len, device));
    return VB_OK;
}

int vb_ctx_set_sphere_radius(vb_ctx *ctx, float r)
{
    if (!ctx || !(r > 0.f))
    {
        return fail(VB_ERR_ARG, "vb_ctx_set_sphere_radius: bad arguments");
    }
    ctx->sphere_radius = r;
    return VB_OK;
}

uint64_t vb_launch_count(const vb_ctx *ctx)
{
    return ctx ? ctx->launches.load() : 0;
}

/* ---- robots ------------------------------------------------------------------------------- */
int vb_robot_by_name(const char *name, vb_robot *out)
{
    if (!name || !out)
    {
        return fail(VB_ERR_ARG, "vb_robot_by_name: null argument");
    }
    for (int r = VB_SPHERE; r <= VB_BAXTER; ++r)
    {
        if (std::strcmp(ops_of(static_cast<vb_robot>(r))->name, name) == 0)
        {
            *out = static_cast<vb_robot>(r);
            return VB_OK;
        }
    }
    return fail(VB_ERR_ARG, "unknown robot '%s'", name);
}

int vb_robot_info(vb_robot r, int *dimension, int *n_spheres, int *resolution)
{
    const vb::RobotOps *o = ops_of(r);
    if (!o)
    {
        return fail(VB_ERR_ARG, "unknown robot id %d", static_cast<int>(r));
    }
    if (dimension)
    {
        *dimension = o->dim;
    }
    if (n_spheres)
    {
        *n_spheres = o->n_spheres;
    }
    if (resolution)
    {
        *resolution = o->resolution;
    }
    return VB_OK;
}

int vb_robot_limits(vb_robot r, float *lo, float *range)
{
    const vb::RobotOps *o = ops_of(r);
    if (!o || !lo || !range)
    {
        return fail(VB_ERR_ARG, "vb_robot_limits: bad arguments");
    }
    std::copy_n(o->lo, o->dim, lo);
    std::copy_n(o->range, o->dim, range);
    return VB_OK;
}

int vb_random_config(vb_robot r, uint64_t seed, uint64_t index, float *q)
{
    const vb::RobotOps *o = ops_of(r);
    if (!o || !q)
    {
        return fail(VB_ERR_ARG, "vb_random_config: bad arguments");
    }
    o->random_config(seed, index, q);
    return VB_OK;
}

}  // extern "C"

/* ================================ environment ================================================= */
namespace
{
    // `id`: insertion sequence number over all vb_env_add_* calls of the environment (the caller's handle on an
    // obstacle -- e.g. its name -- once the lists are sorted by min_distance)
    struct HostSphere
    {
        float x, y, z, r, min_d;
        uint32_t id;
    };
    struct HostCapsule
    {
        float c[8];  // x1 y1 z1 xv yv zv r rdv
        float min_d;
        uint32_t id;
    };
    struct HostCuboid
    {
        float c[15];  // centre, axis1, axis2, axis3, half extents
        float min_d;
        uint32_t id;
    };
    struct HostHeightField
    {
        float os[6];
        size_t xd, yd;
        std::vector<float> data;
        uint32_t id;
    };
    using HostCapt = vb::CaptBuffers;

    float dot3(float ax, float ay, float az, float bx, float by, float bz)
    {
        return (ax * bx) + (ay * by) + (az * bz);
    }
    float clampf(float v, float lo, float hi)
    {
        return std::max(std::min(v, hi), lo);
    }

    // collision/shapes.hh:57-73
    float cuboid_min_distance(const float *c)
    {
        const float x = c[0], y = c[1], z = c[2];
        const float d1 = dot3(-x, -y, -z, c[3], c[4], c[5]);
        const float d2 = dot3(-x, -y, -z, c[6], c[7], c[8]);
        const float d3 = dot3(-x, -y, -z, c[9], c[10], c[11]);
        const float v1 = clampf(d1, -c[12], c[12]);
        const float v2 = clampf(d2, -c[13], c[13]);
        const float v3 = clampf(d3, -c[14], c[14]);
        const float xn = x + c[3] * v1 + c[6] * v2 + c[9] * v3;
        const float yn = y + c[4] * v1 + c[7] * v2 + c[10] * v3;
        const float zn = z + c[5] * v1 + c[8] * v2 + c[11] * v3;
        return std::sqrt(xn * xn + yn * yn + zn * zn);
    }

    // collision/shapes.hh:166-192
    float capsule_min_distance(const float *c)
    {
        const float x1 = c[0], y1 = c[1], z1 = c[2], xv = c[3], yv = c[4], zv = c[5], r = c[6], rdv = c[7];
        const float dot = clampf(dot3(-x1, -y1, -z1, xv, yv, zv) * rdv, 0.f, 1.f);
        const float xp = x1 + xv * dot, yp = y1 + yv * dot, zp = z1 + zv * dot;
        float xo = -xp, yo = -yp, zo = -zp;
        const float ol = std::sqrt(dot3(xo, yo, zo, xo, yo, zo));
        xo = xo / ol;
        yo = yo / ol;
        zo = zo / ol;
        const float ro = clampf(ol, 0.f, r);
        const float xn = xp + ro * xo, yn = yp + ro * yo, zn = zp + ro * zo;
        return std::sqrt(xn * xn + yn * yn + zn * zn);
    }

    struct Replica
    {
        DevBuf prims, hf_desc, capt_desc, att_local, sdf, sdf_c;
        std::vector<DevBuf> blobs;  // heightfield data and CAPT arrays
        vb::EnvView view{};
    };
}  // namespace

struct vb_env
{
    vb_ctx *ctx = nullptr;
    std::vector<HostSphere> spheres;
    std::vector<HostCapsule> capsules, zcapsules;
    std::vector<HostCuboid> cuboids, zcuboids;
    std::vector<HostHeightField> heightfields;
    std::vector<HostCapt> capts;
    bool has_attachment = false;
    float att_tf[7] = {0, 0, 0, 0, 0, 0, 1};
    std::vector<float4> att_spheres;
    std::vector<Replica> replicas;
    bool committed = false;
    uint64_t version = 0;
    uint32_t next_id = 0;
};

namespace
{
    void release_replicas(vb_env *env)
    {
        for (size_t i = 0; i < env->replicas.size(); ++i)
        {
            cudaSetDevice(env->ctx->slots[i].device);
            Replica &r = env->replicas[i];
            r.prims.release();
            r.hf_desc.release();
            r.capt_desc.release();
            r.att_local.release();
            r.sdf.release();
            r.sdf_c.release();
            for (auto &b : r.blobs)
            {
                b.release();
            }
        }
        env->replicas.clear();
        env->committed = false;
    }

    template <class T>
    void sort_by_min_distance(std::vector<T> &v)
    {
        // environment.hh:40-66
        std::stable_sort(v.begin(), v.end(), [](const T &a, const T &b) { return a.min_d < b.min_d; });
    }

    // float4 records in the order vb_env.cuh documents
    std::vector<float4> pack_records(const vb_env *e)
    {
        std::vector<float4> q;
        for (const auto &s : e->spheres)
        {
            q.push_back(make_float4(s.x, s.y, s.z, s.r));
            q.push_back(make_float4(s.min_d, 0.f, 0.f, 0.f));
        }
        for (const auto &c : e->capsules)
        {
            q.push_back(make_float4(c.c[0], c.c[1], c.c[2], c.c[6]));
            q.push_back(make_float4(c.c[3], c.c[4], c.c[5], c.c[7]));
            q.push_back(make_float4(c.min_d, 0.f, 0.f, 0.f));
        }
        for (const auto &c : e->zcapsules)
        {
            q.push_back(make_float4(c.c[0], c.c[1], c.c[2], c.c[6]));
            q.push_back(make_float4(c.c[5], c.c[7], c.min_d, 0.f));
        }
        for (const auto &c : e->cuboids)
        {
            q.push_back(make_float4(c.c[0], c.c[1], c.c[2], c.min_d));
            q.push_back(make_float4(c.c[3], c.c[4], c.c[5], c.c[12]));
            q.push_back(make_float4(c.c[6], c.c[7], c.c[8], c.c[13]));
            q.push_back(make_float4(c.c[9], c.c[10], c.c[11], c.c[14]));
        }
        for (const auto &c : e->zcuboids)
        {
            q.push_back(make_float4(c.c[0], c.c[1], c.c[2], c.min_d));
            q.push_back(make_float4(c.c[3], c.c[4], c.c[6], c.c[7]));
            q.push_back(make_float4(c.c[12], c.c[13], c.c[14], 0.f));
        }
        return q;
    }
}  // namespace

extern "C" {

int vb_env_create(vb_ctx *ctx, vb_env **out)
{
    if (!out)
    {
        return fail(VB_ERR_ARG, "vb_env_create: null argument");
    }
    *out = new vb_env();
    (*out)->ctx = ctx;
    return VB_OK;
}

void vb_env_destroy(vb_env *env)
{
    if (env)
    {
        release_replicas(env);
        delete env;
    }
}

int vb_env_clear(vb_env *env)
{
    if (!env)
    {
        return fail(VB_ERR_ARG, "null environment");
    }
    vb_ctx *ctx = env->ctx;
    release_replicas(env);
    *env = vb_env();
    env->ctx = ctx;
    return VB_OK;
}

int vb_env_add_sphere(vb_env *env, float x, float y, float z, float r)
{
    if (!env)
    {
        return fail(VB_ERR_ARG, "null environment");
    }
    // shapes.hh:236-239
    env->spheres.push_back({x, y, z, r, std::sqrt(x * x + y * y + z * z) - r, env->next_id++});
    sort_by_min_distance(env->spheres);
    env->committed = false;
    return VB_OK;
}

int vb_env_add_cuboid(vb_env *env, const float *c)
{
    if (!env || !c)
    {
        return fail(VB_ERR_ARG, "vb_env_add_cuboid: null argument");
    }
    HostCuboid h;
    std::copy_n(c, 15, h.c);
    h.min_d = cuboid_min_distance(c);
    h.id = env->next_id++;
    // bindings/environment.cc:120-127
    auto &list = (c[11] == 1.f) ? env->zcuboids : env->cuboids;
    list.push_back(h);
    sort_by_min_distance(list);
    env->committed = false;
    return VB_OK;
}

int vb_env_add_capsule(vb_env *env, const float *c)
{
    if (!env || !c)
    {
        return fail(VB_ERR_ARG, "vb_env_add_capsule: null argument");
    }
    HostCapsule h;
    std::copy_n(c, 8, h.c);
    h.min_d = capsule_min_distance(c);
    h.id = env->next_id++;
    // bindings/environment.cc:134-141
    auto &list = (c[3] == 0.f && c[4] == 0.f) ? env->zcapsules : env->capsules;
    list.push_back(h);
    sort_by_min_distance(list);
    env->committed = false;
    return VB_OK;
}

int vb_env_add_heightfield(vb_env *env, const float *off_scale, size_t xd, size_t yd, const float *data)
{
    if (!env || !off_scale || !data || xd == 0 || yd == 0)
    {
        return fail(VB_ERR_ARG, "vb_env_add_heightfield: bad arguments");
    }
    HostHeightField h;
    std::copy_n(off_scale, 6, h.os);
    h.xd = xd;
    h.yd = yd;
    h.data.assign(data, data + xd * yd);
    h.id = env->next_id++;
    env->heightfields.push_back(std::move(h));
    env->committed = false;
    return VB_OK;
}

int vb_env_add_capt(vb_env *env, const float *tests, size_t n_tests, const uint32_t *aff_starts, const float *aff_x,
                    const float *aff_y, const float *aff_z, size_t n_aff_vec, const float *aabbs, size_t n_aabbs,
                    const float *aabb_top, float r_point, int nlog2)
{
    if (!env || !aff_starts || !aabbs || !aabb_top || nlog2 < 0 || nlog2 > 30 || (n_tests && !tests) ||
        (n_aff_vec && (!aff_x || !aff_y || !aff_z)))
    {
        return fail(VB_ERR_ARG, "vb_env_add_capt: bad arguments");
    }
    const size_t leaves = size_t(1) << nlog2;
    if (n_tests != leaves - 1 || n_aabbs != leaves)
    {
        return fail(VB_ERR_ARG, "vb_env_add_capt: nlog2=%d needs %zu tests and %zu aabbs (got %zu, %zu)", nlog2,
                    leaves - 1, leaves, n_tests, n_aabbs);
    }
    if (aff_starts[leaves] > n_aff_vec)
    {
        return fail(VB_ERR_ARG, "vb_env_add_capt: aff_starts end %u exceeds %zu affordance vectors",
                    aff_starts[leaves], n_aff_vec);
    }
    HostCapt h;
    h.tests.assign(tests, tests + n_tests);
    h.aabbs.assign(aabbs, aabbs + 6 * n_aabbs);
    h.aff_starts.assign(aff_starts, aff_starts + leaves + 1);
    h.aff.resize(24 * n_aff_vec);
    for (size_t i = 0; i < n_aff_vec; ++i)
    {
        std::copy_n(aff_x + 8 * i, 8, h.aff.data() + 24 * i);
        std::copy_n(aff_y + 8 * i, 8, h.aff.data() + 24 * i + 8);
        std::copy_n(aff_z + 8 * i, 8, h.aff.data() + 24 * i + 16);
    }
    std::copy_n(aabb_top, 6, h.top);
    h.r_point = r_point;
    h.nlog2 = nlog2;
    env->capts.push_back(std::move(h));
    env->committed = false;
    return VB_OK;
}

// collision::Environment::pointclouds.emplace_back(points, r_min, r_max, r_point) (bindings/environment.cc:148-158):
// the CAPT is BUILT here, on the device (vb_capt_build.cu), from raw points.
int vb_env_add_pointcloud(vb_env *env, const float *xyz, size_t n, float r_min, float r_max, float r_point, double *build_ms)
{
    if (!env || !xyz || n == 0 || !(r_max >= r_min) || !(r_min >= 0.f) || !(r_point >= 0.f))
    {
        return fail(VB_ERR_ARG, "vb_env_add_pointcloud: bad arguments");
    }
    if (!env->ctx || env->ctx->slots.empty())
    {
        return fail(VB_ERR_STATE, "vb_env_add_pointcloud: the tree is built on the device; the environment needs a context");
    }
    Slot &s = env->ctx->slots[0];
    VB_CUDA(cudaSetDevice(s.device));
    HostCapt h;
    const char *what = "";
    VB_CUDA(cudaEventRecord(s.ev0, s.stream));
    const cudaError_t e = vb::build_capt(s.stream, xyz, n, r_min, r_max, r_point, h, &what);
    if (e != cudaSuccess)
    {
        return fail(e == cudaErrorInvalidValue ? VB_ERR_ARG : VB_ERR_CUDA, "vb_env_add_pointcloud: %s%s%s", cudaGetErrorString(e),
                    what[0] ? ": " : "", what);
    }
    VB_CUDA(cudaEventRecord(s.ev1, s.stream));
    VB_CUDA(cudaEventSynchronize(s.ev1));
    float ms = 0.f;
    VB_CUDA(cudaEventElapsedTime(&ms, s.ev0, s.ev1));
    if (build_ms)
    {
        *build_ms = ms;
    }
    env->ctx->launches += 2 * static_cast<uint64_t>(h.nlog2) + 8;
    env->capts.push_back(std::move(h));
    env->committed = false;
    return VB_OK;
}

// the buffers of point cloud `idx` as the library holds them (layout of vb_env_add_capt; affordances interleaved per
// 8-point vector: x[8] y[8] z[8]); sizes via the count arguments, data when the pointers are non-null
int vb_env_capt_export(const vb_env *env, size_t idx, size_t *n_tests, size_t *n_aff_vec, float *tests, uint32_t *aff_starts,
                       float *aff, float *aabbs, float *aabb_top, int *nlog2)
{
    if (!env || idx >= env->capts.size())
    {
        return fail(VB_ERR_ARG, "vb_env_capt_export: no such point cloud");
    }
    const HostCapt &c = env->capts[idx];
    if (n_tests) *n_tests = c.tests.size();
    if (n_aff_vec) *n_aff_vec = c.aff.size() / 24;
    if (nlog2) *nlog2 = c.nlog2;
    if (tests) std::copy(c.tests.begin(), c.tests.end(), tests);
    if (aff_starts) std::copy(c.aff_starts.begin(), c.aff_starts.end(), aff_starts);
    if (aff) std::copy(c.aff.begin(), c.aff.end(), aff);
    if (aabbs) std::copy(c.aabbs.begin(), c.aabbs.end(), aabbs);
    if (aabb_top) std::copy_n(c.top, 6, aabb_top);
    return VB_OK;
}

int vb_env_set_attachment(vb_env *env, const float *tf, const float *spheres, size_t n)
{
    if (!env || !tf || (n && !spheres) || n > static_cast<size_t>(vb::kMaxAttach))
    {
        return fail(VB_ERR_ARG, "vb_env_set_attachment: bad arguments (at most %d spheres)", vb::kMaxAttach);
    }
    std::copy_n(tf, 7, env->att_tf);
    env->att_spheres.clear();
    for (size_t i = 0; i < n; ++i)
    {
        env->att_spheres.push_back(make_float4(spheres[4 * i], spheres[4 * i + 1], spheres[4 * i + 2], spheres[4 * i + 3]));
    }
    env->has_attachment = true;
    env->committed = false;
    return VB_OK;
}

int vb_env_clear_attachment(vb_env *env)
{
    if (!env)
    {
        return fail(VB_ERR_ARG, "null environment");
    }
    env->has_attachment = false;
    env->att_spheres.clear();
    env->committed = false;
    return VB_OK;
}

int vb_env_export(const vb_env *env, int kind, float *out, size_t *n)
{
    if (!env || !n || kind < 0 || kind > 4)
    {
        return fail(VB_ERR_ARG, "vb_env_export: bad arguments");
    }
    if (kind == 0)
    {
        *n = env->spheres.size();
        for (size_t i = 0; out && i < *n; ++i)
        {
            const auto &s = env->spheres[i];
            const float row[5] = {s.x, s.y, s.z, s.r, s.min_d};
            std::copy_n(row, 5, out + 5 * i);
        }
    }
    else if (kind <= 2)
    {
        const auto &v = kind == 1 ? env->capsules : env->zcapsules;
        *n = v.size();
        for (size_t i = 0; out && i < *n; ++i)
        {
            std::copy_n(v[i].c, 8, out + 9 * i);
            out[9 * i + 8] = v[i].min_d;
        }
    }
    else
    {
        const auto &v = kind == 3 ? env->cuboids : env->zcuboids;
        *n = v.size();
        for (size_t i = 0; out && i < *n; ++i)
        {
            std::copy_n(v[i].c, 15, out + 16 * i);
            out[16 * i + 15] = v[i].min_d;
        }
    }
    return VB_OK;
}

int vb_env_object_ids(const vb_env *env, uint32_t *ids, size_t *n)
{
    if (!env || !n)
    {
        return fail(VB_ERR_ARG, "vb_env_object_ids: bad arguments");
    }
    std::vector<uint32_t> all;
    for (const auto &s : env->spheres) all.push_back(s.id);
    for (const auto &s : env->capsules) all.push_back(s.id);
    for (const auto &s : env->zcapsules) all.push_back(s.id);
    for (const auto &s : env->cuboids) all.push_back(s.id);
    for (const auto &s : env->zcuboids) all.push_back(s.id);
    for (const auto &s : env->heightfields) all.push_back(s.id);
    *n = all.size();
    if (ids)
    {
        std::copy(all.begin(), all.end(), ids);
    }
    return VB_OK;
}

}  // extern "C"

/* ================================ commit (upload) ============================================= */
namespace
{
    template <class T>
    int upload(DevBuf &buf, const std::vector<T> &host, cudaStream_t stream)
    {
        VB_CUDA(buf.reserve(std::max<size_t>(host.size() * sizeof(T), 16)));
        if (!host.empty())
        {
            VB_CUDA(cudaMemcpyAsync(buf.p, host.data(), host.size() * sizeof(T), cudaMemcpyHostToDevice, stream));
        }
        return VB_OK;
    }

    // one thread per voxel of the distance / candidate grid (vb_sdf.cuh)
    __global__ void sdf_build_kernel(const vb::EnvView e, const vb::SdfView g, uint2 *cell)
    {
        const size_t n = static_cast<size_t>(g.nx) * g.ny * g.nz;
        const float h = 1.f / g.inv_h;
        const float reach = g.mask_reach + g.pad;
        for (size_t v = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; v < n;
             v += static_cast<size_t>(gridDim.x) * blockDim.x)
        {
            const int ix = static_cast<int>(v % g.nx), iy = static_cast<int>((v / g.nx) % g.ny),
                      iz = static_cast<int>(v / (static_cast<size_t>(g.nx) * g.ny));
            float best = 3.0e38f;
            uint32_t mask = 0u;
            vb::sdf_objects(e, g.ox + (ix + 0.5f) * h, g.oy + (iy + 0.5f) * h, g.oz + (iz + 0.5f) * h,
                            [&](int o, float d) {
                                best = fminf(best, d);
                                if (d < reach)
                                {
                                    mask |= 1u << (o < 31 ? o : 31);
                                }
                            });
            cell[v] = make_uint2(__float_as_uint(best), mask);
        }
    }

    // Occupancy bitmap of a point cloud (vb_env.cuh: CaptView::occ).  Every point of the cloud is the representative of
    // exactly one leaf, stored first in that leaf's affordance list (collision/capt.hh:166-177): one thread per leaf
    // marks the voxels whose centre lies within reach + half a voxel diagonal of its representative.
    constexpr float kCaptOccVoxel = 0.04f;
    constexpr float kCaptOccReach = 0.25f;  // covers every robot's largest bounding sphere + the point radius

    __global__ void capt_occupancy_kernel(const vb::CaptView c, size_t n_leaves, uint32_t *occ)
    {
        const size_t z = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
        if (z >= n_leaves || c.aff_starts[z] == c.aff_starts[z + 1])
        {
            return;
        }
        const float4 *v = c.aff + 6ull * c.aff_starts[z];
        const float p[3] = {v[0].x, v[2].x, v[4].x};
        if (!isfinite(p[0]))
        {
            return;
        }
        const float h = 1.f / c.occ_inv_h, reach = c.occ_reach + 0.5f * 1.7320508f * h * 1.001f;
        int lo[3], hi[3];
        for (int k = 0; k < 3; ++k)
        {
            lo[k] = max(0, static_cast<int>(floorf((p[k] - reach - c.occ_o[k]) * c.occ_inv_h)));
            hi[k] = min(c.occ_n[k] - 1, static_cast<int>(floorf((p[k] + reach - c.occ_o[k]) * c.occ_inv_h)));
        }
        for (int iz = lo[2]; iz <= hi[2]; ++iz)
        {
            for (int iy = lo[1]; iy <= hi[1]; ++iy)
            {
                for (int ix = lo[0]; ix <= hi[0]; ++ix)
                {
                    const float dx = c.occ_o[0] + (ix + 0.5f) * h - p[0], dy = c.occ_o[1] + (iy + 0.5f) * h - p[1],
                                dz = c.occ_o[2] + (iz + 0.5f) * h - p[2];
                    if (dx * dx + dy * dy + dz * dz <= reach * reach)
                    {
                        const unsigned bit = (static_cast<unsigned>(iz) * c.occ_n[1] + iy) * c.occ_n[0] + ix;
                        if (!((occ[bit >> 5] >> (bit & 31u)) & 1u))
                        {
                            atomicOr(occ + (bit >> 5), 1u << (bit & 31u));
                        }
                    }
                }
            }
        }
    }

    // Fine level (vb_env.cuh: CaptView::dist): distance from every voxel centre to the nearest point, capped.  One
    // thread per leaf splats its representative point into the voxels within the cap (atomicMin on the bit pattern of
    // a non-negative float).
    constexpr float kCaptDistCap = 0.10f;
    constexpr float kCaptDistVoxel = 0.0125f;

    __global__ void fill_float_bits_kernel(uint32_t *p, size_t n, float v)
    {
        for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += static_cast<size_t>(gridDim.x) * blockDim.x)
        {
            p[i] = __float_as_uint(v);
        }
    }

    __global__ void capt_distance_kernel(const vb::CaptView c, size_t n_leaves, uint32_t *dist_bits)
    {
        const size_t z = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
        if (z >= n_leaves || c.aff_starts[z] == c.aff_starts[z + 1])
        {
            return;
        }
        const float4 *v = c.aff + 6ull * c.aff_starts[z];
        const float p[3] = {v[0].x, v[2].x, v[4].x};
        if (!isfinite(p[0]))
        {
            return;
        }
        const float h = 1.f / c.dist_inv_h, cap = c.dist_cap;
        int lo[3], hi[3];
        for (int k = 0; k < 3; ++k)
        {
            lo[k] = max(0, static_cast<int>(floorf((p[k] - cap - c.dist_o[k]) * c.dist_inv_h)));
            hi[k] = min(c.dist_n[k] - 1, static_cast<int>(floorf((p[k] + cap - c.dist_o[k]) * c.dist_inv_h)));
        }
        for (int iz = lo[2]; iz <= hi[2]; ++iz)
        {
            for (int iy = lo[1]; iy <= hi[1]; ++iy)
            {
                for (int ix = lo[0]; ix <= hi[0]; ++ix)
                {
                    const float dx = c.dist_o[0] + (ix + 0.5f) * h - p[0], dy = c.dist_o[1] + (iy + 0.5f) * h - p[1],
                                dz = c.dist_o[2] + (iz + 0.5f) * h - p[2];
                    const float d = sqrtf(dx * dx + dy * dy + dz * dz);
                    if (d < cap)
                    {
                        const size_t cell = (static_cast<size_t>(iz) * c.dist_n[1] + iy) * c.dist_n[0] + ix;
                        if (__float_as_uint(d) < dist_bits[cell])
                        {
                            atomicMin(dist_bits + cell, __float_as_uint(d));
                        }
                    }
                }
            }
        }
    }

    // Bounding box of the primitive obstacles -> grid geometry.  The box is inflated by kSdfCover so that
    // "outside the grid" means "at least kSdfCover away from every obstacle"; scenes larger than the
    // +-kSdfReach cube around the origin are clipped to it and lose that guarantee (cover = 0).
    constexpr float kSdfReach = 4.0f;
    constexpr size_t kSdfMaxVoxels = size_t(4) << 20;
    constexpr float kSdfMinVoxel = 0.015f;

    // coarse level of the two-level guard lookup (vb_sdf.cuh): 2^kCoarseShift = 8 fine voxels (0.12 m at the finest
    // pitch; voxel edges of 0.05 .. 0.3 m were swept on B200, flat optimum around 0.15), on the fine lattice
    constexpr float kSdfCoarseMaxVoxel = 0.3f;
    constexpr int kSdfDimMultiple = 2 << vb::kCoarseShift;  // fine dimensions: even coarse dimensions (centred lookup)

    void sdf_finish(vb::SdfView &g, float h)
    {
        g.inv_h = 1.f / h;
        g.fnx = std::nextafterf(static_cast<float>(g.nx), 0.f);
        g.fny = std::nextafterf(static_cast<float>(g.ny), 0.f);
        g.fnz = std::nextafterf(static_cast<float>(g.nz), 0.f);
        g.pad = 0.5f * std::sqrt(3.f) * h * 1.001f + vb::kSdfMargin;
        g.hx = 0.5f * static_cast<float>(g.nx);
        g.hy = 0.5f * static_cast<float>(g.ny);
        g.hz = 0.5f * static_cast<float>(g.nz);
        g.cx = -g.ox * g.inv_h - g.hx;
        g.cy = -g.oy * g.inv_h - g.hy;
        g.cz = -g.oz * g.inv_h - g.hz;
    }
    void sdf_attach(vb::SdfView &g, const uint2 *cells)
    {
        g.cell = cells;
        g.centre = cells + (static_cast<size_t>(g.nz / 2) * g.ny + g.ny / 2) * g.nx + g.nx / 2;
    }
    // the coarse level of `g`: same lower corner, voxels of 2^kCoarseShift fine voxels
    vb::SdfView sdf_coarse_of(const vb::SdfView &g)
    {
        vb::SdfView c = g;
        c.cell = c.centre = nullptr;
        c.nx = g.nx >> vb::kCoarseShift;
        c.ny = g.ny >> vb::kCoarseShift;
        c.nz = g.nz >> vb::kCoarseShift;
        sdf_finish(c, static_cast<float>(1 << vb::kCoarseShift) / g.inv_h);
        return c;
    }

    bool sdf_geometry(const vb_env *env, vb::SdfView &g, float min_voxel_arg = kSdfMinVoxel)
    {
        float lo[3] = {3e38f, 3e38f, 3e38f}, hi[3] = {-3e38f, -3e38f, -3e38f};
        auto grow = [&](const float *c, const float *e) {
            for (int k = 0; k < 3; ++k)
            {
                lo[k] = std::min(lo[k], c[k] - e[k]);
                hi[k] = std::max(hi[k], c[k] + e[k]);
            }
        };
        for (const auto &s : env->spheres)
        {
            const float c[3] = {s.x, s.y, s.z}, e[3] = {s.r, s.r, s.r};
            grow(c, e);
        }
        auto capsule = [&](const HostCapsule &c) {
            const float e[3] = {c.c[6], c.c[6], c.c[6]};
            const float p2[3] = {c.c[0] + c.c[3], c.c[1] + c.c[4], c.c[2] + c.c[5]};
            grow(c.c, e);
            grow(p2, e);
        };
        for (const auto &c : env->capsules) capsule(c);
        for (const auto &c : env->zcapsules) capsule(c);
        auto cuboid = [&](const HostCuboid &c) {
            float e[3];
            for (int k = 0; k < 3; ++k)
            {
                e[k] = std::fabs(c.c[3 + k]) * c.c[12] + std::fabs(c.c[6 + k]) * c.c[13] + std::fabs(c.c[9 + k]) * c.c[14];
            }
            grow(c.c, e);
        };
        for (const auto &c : env->cuboids) cuboid(c);
        for (const auto &c : env->zcuboids) cuboid(c);
        if (!(lo[0] <= hi[0]))
        {
            return false;  // no primitives (or NaNs): no grid
        }
        bool clipped = false;
        float ext[3];
        for (int k = 0; k < 3; ++k)
        {
            lo[k] -= vb::kSdfCover + 0.01f;
            hi[k] += vb::kSdfCover + 0.01f;
            if (!(lo[k] >= -kSdfReach) || !(hi[k] <= kSdfReach) || !std::isfinite(lo[k]) || !std::isfinite(hi[k]))
            {
                clipped = true;
                lo[k] = std::isfinite(lo[k]) ? std::max(lo[k], -kSdfReach) : -kSdfReach;
                hi[k] = std::isfinite(hi[k]) ? std::min(hi[k], kSdfReach) : kSdfReach;
                if (!(lo[k] < hi[k]))
                {
                    return false;
                }
            }
            ext[k] = hi[k] - lo[k];
        }
        float min_voxel = min_voxel_arg;
        if (const char *v = std::getenv("VB_SDF_VOXEL"))  // development: voxel edge of the fine grid in metres
        {
            if (min_voxel_arg == kSdfMinVoxel)
            {
                min_voxel = std::max(1e-3f, static_cast<float>(std::atof(v)));
            }
        }
        float h = std::max(min_voxel, std::cbrt(ext[0] * ext[1] * ext[2] / static_cast<float>(kSdfMaxVoxels)));
        for (;;)
        {
            auto dim = [&](float e) { return (static_cast<int>(std::ceil(e / h)) + kSdfDimMultiple - 1) / kSdfDimMultiple * kSdfDimMultiple; };
            g.nx = dim(ext[0]);
            g.ny = dim(ext[1]);
            g.nz = dim(ext[2]);
            if (static_cast<size_t>(g.nx) * g.ny * g.nz <= kSdfMaxVoxels)
            {
                break;
            }
            h *= 1.05f;
        }
        g.ox = lo[0];
        g.oy = lo[1];
        g.oz = lo[2];
        sdf_finish(g, h);
        g.cover = clipped ? 0.f : vb::kSdfCover;
        g.mask_reach = vb::kMaskReach;
        if (const char *v = std::getenv("VB_SDF_REACH"))  // development: candidate-mask reach in metres
        {
            g.mask_reach = std::max(0.05f, static_cast<float>(std::atof(v)));
        }
        return true;
    }

    int commit_replica(vb_env *env, Slot &slot, Replica &rep)
    {
        VB_CUDA(cudaSetDevice(slot.device));
        const std::vector<float4> recs = pack_records(env);
        VB_TRY(upload(rep.prims, recs, slot.stream));
        vb::EnvView &v = rep.view;
        v = vb::EnvView{};
        v.prims = rep.prims.as<float4>();
        v.n_quads = static_cast<int>(recs.size());
        v.n_spheres = static_cast<int>(env->spheres.size());
        v.n_capsules = static_cast<int>(env->capsules.size());
        v.n_zcapsules = static_cast<int>(env->zcapsules.size());
        v.n_cuboids = static_cast<int>(env->cuboids.size());
        v.n_zcuboids = static_cast<int>(env->zcuboids.size());

        rep.blobs.clear();
        std::vector<vb::HeightFieldView> hfs;
        for (const auto &h : env->heightfields)
        {
            // the reference clamps the cell index to [0, xd] x [0, yd] (sphere_heightfield.hh:18-21), one
            // past the last row/column; pad so the same cell arithmetic stays inside the allocation
            std::vector<float> padded(h.data);
            padded.resize(h.data.size() + h.xd + 2, 0.f);
            rep.blobs.emplace_back();
            VB_TRY(upload(rep.blobs.back(), padded, slot.stream));
            vb::HeightFieldView hv{};
            hv.x = h.os[0];
            hv.y = h.os[1];
            hv.z = h.os[2];
            hv.xs = h.os[3];
            hv.ys = h.os[4];
            hv.zs = h.os[5];
            hv.xd2 = static_cast<float>(h.xd / 2);
            hv.yd2 = static_cast<float>(h.yd / 2);
            hv.xdf = static_cast<float>(h.xd);
            hv.ydf = static_cast<float>(h.yd);
            hv.xd = static_cast<int>(h.xd);
            hv.data = rep.blobs.back().as<float>();
            hfs.push_back(hv);
        }
        VB_TRY(upload(rep.hf_desc, hfs, slot.stream));
        v.n_heightfields = static_cast<int>(hfs.size());
        v.heightfields = rep.hf_desc.as<vb::HeightFieldView>();

        std::vector<vb::CaptView> cvs;
        for (const auto &c : env->capts)
        {
            vb::CaptView cv{};
            rep.blobs.emplace_back();
            VB_TRY(upload(rep.blobs.back(), c.tests, slot.stream));
            cv.tests = rep.blobs.back().as<float>();
            rep.blobs.emplace_back();
            VB_TRY(upload(rep.blobs.back(), c.aff_starts, slot.stream));
            cv.aff_starts = rep.blobs.back().as<uint32_t>();
            rep.blobs.emplace_back();
            VB_TRY(upload(rep.blobs.back(), c.aff, slot.stream));
            cv.aff = rep.blobs.back().as<float4>();
            rep.blobs.emplace_back();
            VB_TRY(upload(rep.blobs.back(), c.aabbs, slot.stream));
            cv.aabbs = rep.blobs.back().as<float>();
            std::copy_n(c.top, 6, cv.top);
            cv.r_point = c.r_point;
            cv.n_tests = static_cast<uint32_t>(c.tests.size());
            cv.nlog2 = c.nlog2;
            // occupancy bitmap over aabb_top grown by the reach + one voxel (skipped for degenerate / huge clouds)
            cv.occ = nullptr;
            if (!std::getenv("VB_CAPT_NO_OCC") && std::isfinite(c.top[0]) && c.top[0] <= c.top[3])
            {
                const float grow = kCaptOccReach + 2.f * kCaptOccVoxel;
                size_t voxels = 1;
                for (int k = 0; k < 3; ++k)
                {
                    cv.occ_o[k] = c.top[k] - grow;
                    cv.occ_n[k] = static_cast<int>(std::ceil((c.top[3 + k] - c.top[k] + 2.f * grow) / kCaptOccVoxel));
                    voxels *= static_cast<size_t>(std::max(cv.occ_n[k], 1));
                }
                if (voxels <= (size_t(1) << 28))
                {
                    cv.occ_inv_h = 1.f / kCaptOccVoxel;
                    cv.occ_reach = kCaptOccReach;
                    rep.blobs.emplace_back();
                    const size_t words = (voxels + 31) / 32;
                    VB_CUDA(rep.blobs.back().reserve(words * sizeof(uint32_t)));
                    uint32_t *occ = rep.blobs.back().as<uint32_t>();
                    VB_CUDA(cudaMemsetAsync(occ, 0, words * sizeof(uint32_t), slot.stream));
                    const size_t leaves = size_t(1) << c.nlog2;
                    capt_occupancy_kernel<<<static_cast<unsigned>((leaves + 127) / 128), 128, 0, slot.stream>>>(cv, leaves, occ);
                    VB_CUDA(cudaGetLastError());
                    cv.occ = occ;
                }
            }
            cv.dist = nullptr;
            if (!std::getenv("VB_CAPT_NO_DIST") && std::isfinite(c.top[0]) && c.top[0] <= c.top[3])
            {
                float h = kCaptDistVoxel;
                size_t voxels = 0;
                for (;;)
                {
                    const float grow = kCaptDistCap + 2.f * h;
                    voxels = 1;
                    for (int k = 0; k < 3; ++k)
                    {
                        cv.dist_o[k] = c.top[k] - grow;
                        cv.dist_n[k] = std::max(1, static_cast<int>(std::ceil((c.top[3 + k] - c.top[k] + 2.f * grow) / h)));
                        voxels *= static_cast<size_t>(cv.dist_n[k]);
                    }
                    if (voxels <= (size_t(1) << 25) || h > 0.05f)
                    {
                        break;
                    }
                    h *= 1.26f;
                }
                if (voxels <= (size_t(1) << 25))
                {
                    cv.dist_inv_h = 1.f / h;
                    cv.dist_cap = kCaptDistCap;
                    cv.dist_pad = 0.5f * std::sqrt(3.f) * h * 1.001f + 1e-4f;
                    rep.blobs.emplace_back();
                    VB_CUDA(rep.blobs.back().reserve(voxels * sizeof(float)));
                    uint32_t *bits = rep.blobs.back().as<uint32_t>();
                    fill_float_bits_kernel<<<slot.sm_count * 8, 256, 0, slot.stream>>>(bits, voxels, kCaptDistCap);
                    VB_CUDA(cudaGetLastError());
                    const size_t leaves = size_t(1) << c.nlog2;
                    capt_distance_kernel<<<static_cast<unsigned>((leaves + 127) / 128), 128, 0, slot.stream>>>(cv, leaves, bits);
                    VB_CUDA(cudaGetLastError());
                    cv.dist = rep.blobs.back().as<float>();
                }
            }
            cvs.push_back(cv);
        }
        VB_TRY(upload(rep.capt_desc, cvs, slot.stream));
        v.n_capts = static_cast<int>(cvs.size());
        v.capt_heavy = 0;
        for (const auto &c : env->capts)
        {
            const size_t leaves = size_t(1) << c.nlog2;
            v.capt_heavy |= (c.aff.size() / 24 > 16 * leaves) ? 1 : 0;
        }
        v.capts = rep.capt_desc.as<vb::CaptView>();

        VB_TRY(upload(rep.att_local, env->att_spheres, slot.stream));
        std::copy_n(env->att_tf, 7, v.attach.tf);
        v.attach.n = env->has_attachment ? static_cast<int>(env->att_spheres.size()) : 0;
        v.attach.local = rep.att_local.as<float4>();
        // distance grid over the primitives (an acceleration structure: never changes a verdict)
        if (v.n_quads > 0 && sdf_geometry(env, v.sdf))
        {
            const size_t voxels = static_cast<size_t>(v.sdf.nx) * v.sdf.ny * v.sdf.nz;
            VB_CUDA(rep.sdf.reserve(voxels * sizeof(uint2)));
            sdf_attach(v.sdf, rep.sdf.as<uint2>());
            sdf_build_kernel<<<slot.sm_count * 8, 256, 0, slot.stream>>>(v, v.sdf, rep.sdf.as<uint2>());
            VB_CUDA(cudaGetLastError());
            // coarse level on the same lattice, unless the fine voxels are already coarse (very large scenes)
            v.sdf_c = vb::SdfView{};
            const bool want_coarse = !std::getenv("VB_SDF_NO_COARSE") &&
                                     static_cast<float>(1 << vb::kCoarseShift) / v.sdf.inv_h <= kSdfCoarseMaxVoxel;
            if (want_coarse)
            {
                v.sdf_c = sdf_coarse_of(v.sdf);
                const size_t cv = static_cast<size_t>(v.sdf_c.nx) * v.sdf_c.ny * v.sdf_c.nz;
                VB_CUDA(rep.sdf_c.reserve(cv * sizeof(uint2)));
                sdf_attach(v.sdf_c, rep.sdf_c.as<uint2>());
                sdf_build_kernel<<<slot.sm_count * 8, 256, 0, slot.stream>>>(v, v.sdf_c, rep.sdf_c.as<uint2>());
                VB_CUDA(cudaGetLastError());
            }
        }
        VB_CUDA(cudaStreamSynchronize(slot.stream));
        return VB_OK;
    }
}  // namespace

extern "C" int vb_env_commit(vb_env *env)
{
    if (!env)
    {
        return fail(VB_ERR_ARG, "null environment");
    }
    if (!env->ctx)
    {
        return fail(VB_ERR_STATE, "vb_env_commit: environment was created without a context (host-side description only)");
    }
    release_replicas(env);
    env->replicas.resize(env->ctx->slots.size());
    for (size_t i = 0; i < env->replicas.size(); ++i)
    {
        VB_TRY(commit_replica(env, env->ctx->slots[i], env->replicas[i]));
    }
    env->committed = true;
    ++env->version;
    return VB_OK;
}

/* ================================ queries ===================================================== */
#include <cub/device/device_scan.cuh>

namespace
{
    // verdict words preset to "valid" for the first n bits (edge verdicts are AND-reduced into them)
    __global__ void fill_bits_kernel(uint32_t *bits, size_t n)
    {
        const size_t w = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
        const size_t n_words = (n + 31) / 32;
        if (w < n_words)
        {
            const size_t rem = n - w * 32;
            bits[w] = rem >= 32 ? 0xffffffffu : ((1u << rem) - 1u);
        }
    }

    __global__ void fill_float_kernel(float *p, size_t n, float v)
    {
        const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
        if (i < n)
        {
            p[i] = v;
        }
    }

    __global__ void popcount_kernel(const uint32_t *bits, size_t n, unsigned long long *out)
    {
        const size_t n_words = (n + 31) / 32;
        unsigned long long local = 0;
        for (size_t w = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; w < n_words;
             w += static_cast<size_t>(gridDim.x) * blockDim.x)
        {
            uint32_t v = bits[w];
            const size_t rem = n - w * 32;
            if (rem < 32)
            {
                v &= (1u << rem) - 1u;
            }
            local += __popc(v);
        }
        for (int o = 16; o > 0; o >>= 1)
        {
            local += __shfl_down_sync(0xffffffffu, local, o);
        }
        if ((threadIdx.x & 31) == 0 && local)
        {
            atomicAdd(out, local);
        }
    }


    // FP32 roofline denominator: 8 independent FMA chains per thread, 2 flops per FMA
    __global__ void __launch_bounds__(256) fma_peak_kernel(float *sink, int iters, float a, float b)
    {
        float x0 = threadIdx.x * 1e-3f, x1 = x0 + 1.f, x2 = x0 + 2.f, x3 = x0 + 3.f;
        float x4 = x0 + 4.f, x5 = x0 + 5.f, x6 = x0 + 6.f, x7 = x0 + 7.f;
        for (int i = 0; i < iters; ++i)
        {
#pragma unroll
            for (int k = 0; k < 8; ++k)
            {
                x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
                x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
            }
        }
        const float v = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
        if (v == 12345.678f)
        {
            sink[0] = v;  // never true in practice: keeps the chains alive
        }
    }

    // synthetic streams on the device (same generator as vb_random_config / vb_random_edge)
    struct StreamLimits
    {
        float lo[16], range[16];
        int dim;
    };

    __global__ void random_configs_kernel(const StreamLimits lim, uint64_t seed, uint64_t first, size_t n, float *out)
    {
        const size_t total = n * lim.dim;
        for (size_t t = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < total;
             t += static_cast<size_t>(gridDim.x) * blockDim.x)
        {
            const size_t i = t / lim.dim;
            const int j = static_cast<int>(t % lim.dim);
            out[t] = VB_ADD(lim.lo[j], VB_MUL(vb::u01(seed, first + i, j), lim.range[j]));  // as random_config<Robot>
        }
    }

    __host__ __device__ void random_edge_row(const StreamLimits &lim, uint64_t seed, uint64_t index, float len_lo, float len_hi,
                                             float *a, float *b)
    {
        // a = config(2 index); direction ~ normalised uniform cube sample; length uniform in [len_lo, len_hi];
        // b clipped to the joint ranges ("RRTC-shaped" edges)
        // (every product and sum is rounded on its own -- VB_ADD / VB_MUL -- so that host and device agree bit for bit)
        float dir[16], norm = 0.f;
        for (int j = 0; j < lim.dim; ++j)
        {
            a[j] = VB_ADD(lim.lo[j], VB_MUL(vb::u01(seed, 2 * index, j), lim.range[j]));
            dir[j] = VB_ADD(VB_MUL(2.f, vb::u01(seed ^ 0xED6Eull, index, j)), -1.f);
            norm = VB_ADD(norm, VB_MUL(dir[j], dir[j]));
        }
        norm = sqrtf(fmaxf(norm, 1e-12f));
        const float len = VB_ADD(len_lo, VB_MUL(vb::u01(seed ^ 0xED6Eull, index, 63), VB_ADD(len_hi, -len_lo)));
        for (int j = 0; j < lim.dim; ++j)
        {
            const float v = VB_ADD(a[j], VB_MUL(len, dir[j]) / norm);
            b[j] = fminf(fmaxf(v, lim.lo[j]), VB_ADD(lim.lo[j], lim.range[j]));
        }
    }

    __global__ void random_edges_kernel(const StreamLimits lim, uint64_t seed, uint64_t first, size_t n, float len_lo,
                                        float len_hi, float *a, float *b)
    {
        for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
             i += static_cast<size_t>(gridDim.x) * blockDim.x)
        {
            float qa[16], qb[16];
            random_edge_row(lim, seed, first + i, len_lo, len_hi, qa, qb);
            for (int j = 0; j < lim.dim; ++j)
            {
                a[i * lim.dim + j] = qa[j];
                b[i * lim.dim + j] = qb[j];
            }
        }
    }

    // ---- first-K-valid compaction (PRM / FCIT* sampling, planning/prm.hh:111-124, planning/fcit.hh:316-343) -------
    __global__ void word_popcount_kernel(const uint32_t *bits, size_t n, uint32_t *counts)
    {
        const size_t w = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
        const size_t n_words = (n + 31) / 32;
        if (w <= n_words)
        {
            uint32_t v = w < n_words ? bits[w] : 0u;
            const size_t rem = w < n_words ? n - w * 32 : 0;
            if (rem < 32)
            {
                v &= rem ? ((1u << rem) - 1u) : 0u;
            }
            counts[w] = __popc(v);  // counts[n_words] = 0: its exclusive prefix is the total
        }
    }

    // candidate i (valid, with `have` + rank < k) -> out rows; the k-th valid overall records where the stream stops
    __global__ void compact_valid_kernel(const uint32_t *bits, const uint32_t *prefix, const float *cfg, int dim, size_t n,
                                         uint64_t stream_first, size_t have, size_t k, float *out_cfg, uint64_t *out_index,
                                         unsigned long long *consumed)
    {
        const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
        if (i >= n)
        {
            return;
        }
        const uint32_t word = bits[i >> 5];
        if (!((word >> (i & 31)) & 1u))
        {
            return;
        }
        const size_t rank = have + prefix[i >> 5] + __popc(word & ((1u << (i & 31)) - 1u));
        if (rank >= k)
        {
            return;
        }
        for (int j = 0; j < dim; ++j)
        {
            out_cfg[rank * dim + j] = cfg[i * dim + j];
        }
        out_index[rank] = stream_first + i;
        if (rank + 1 == k)
        {
            *consumed = static_cast<unsigned long long>(i + 1);
        }
    }

    StreamLimits limits_of(const vb::RobotOps *o)
    {
        StreamLimits l{};
        l.dim = o->dim;
        std::copy_n(o->lo, 16, l.lo);
        std::copy_n(o->range, 16, l.range);
        return l;
    }

    struct Query
    {
        vb_ctx *ctx;
        Slot *slot;
        const vb::RobotOps *ops;
        vb::LaunchArgs args;
    };

    int begin_query(vb_ctx *ctx, int slot, vb_robot robot, const float *base, const vb_env *env, Query &q)
    {
        VB_TRY(check_slot(ctx, slot));
        q.ctx = ctx;
        q.slot = &ctx->slots[slot];
        q.ops = ops_of(robot);
        if (!q.ops)
        {
            return fail(VB_ERR_ARG, "unknown robot id %d", static_cast<int>(robot));
        }
        q.args = vb::LaunchArgs{};
        if (env)
        {
            if (env->ctx != ctx)
            {
                return fail(VB_ERR_ARG, "environment belongs to another context");
            }
            if (!env->committed)
            {
                return fail(VB_ERR_STATE, "environment changed since the last vb_env_commit");
            }
            q.args.env = env->replicas[slot].view;
        }
        for (int k = 0; k < 3; ++k)
        {
            q.args.base[k] = base ? base[k] : 0.f;
        }
        q.args.sphere_radius = ctx->sphere_radius;
        q.args.attach = (q.ops->has_attach && q.args.env.attach.n > 0) ? 1 : 0;
        q.args.stream = q.slot->stream;
        VB_CUDA(cudaSetDevice(q.slot->device));
        return VB_OK;
    }

    // persistent grid: resident CTAs per SM x SM count, never more CTAs than there is work
    int grid_for(const Query &q, size_t items, int *grid)
    {
        int blocks = 0, threads = 0;
        VB_CUDA(q.ops->occupancy(q.args, &blocks, &threads));
        if (blocks < 1)
        {
            return fail(VB_ERR_CUDA, "%s: environment does not fit one CTA's shared memory", q.ops->name);
        }
        const size_t need = (items + threads - 1) / threads;
        *grid = static_cast<int>(std::max<size_t>(1, std::min<size_t>(need, static_cast<size_t>(blocks) * q.slot->sm_count)));
        if (q.args.attach)
        {
            // posed attachment spheres live in a global scratch column per thread
            Slot &s = *q.slot;
            VB_CUDA(s.att.reserve(sizeof(float) * 3 * vb::kMaxAttach * static_cast<size_t>(*grid) * threads));
            const_cast<Query &>(q).args.att_scratch = s.att.as<float>();
        }
        return VB_OK;
    }

    int run_fkcc_d(Query &q, const float *cfg_d, size_t n, uint32_t *bits_d, float *margin_d, int mode = -1)
    {
        if (n == 0)
        {
            return VB_OK;
        }
        q.args.cfg = cfg_d;
        q.args.n = n;
        q.args.bits = bits_d;
        q.args.margin = margin_d;
        int grid = 1;
        VB_TRY(grid_for(q, n, &grid));
        VB_CUDA(q.ops->fkcc(q.args, mode >= 0 ? mode : (margin_d != nullptr ? 1 : 0), grid));
        q.ctx->launches += 1;
        return VB_OK;
    }

    // Edges in (start, goal) form, or (start, vector, distance) form when vector_form (planning/validate.hh:23-65).
    // One launch of the rake kernel (vb_kernels.cuh: motion_rake_kernel); it presets and writes the verdict words
    // (and margins) itself and needs no scratch memory.
    int run_motion_d(Query &q, const float *a_d, const float *b_d, const float *dist_d, bool vector_form, size_t n,
                     int resolution, uint32_t *bits_d, float *margin_d)
    {
        if (n == 0)
        {
            return VB_OK;
        }
        q.args.cfg = a_d;
        q.args.cfg_b = b_d;
        q.args.dist = dist_d;
        q.args.vector_form = vector_form ? 1 : 0;
        q.args.n = n;
        q.args.bits = bits_d;
        q.args.margin = margin_d;
        q.args.resolution = resolution > 0 ? resolution : q.ops->resolution;
        int grid = 1;
        // one edge per octet of 8 lanes at a time
        VB_TRY(grid_for(q, n * 8, &grid));
        VB_CUDA(q.ops->motion_rake(q.args, margin_d != nullptr, grid));
        q.ctx->launches += 1;
        return VB_OK;
    }

    struct Shard
    {
        size_t lo, hi;
    };

    // contiguous index ranges, boundaries on multiples of 32 so verdict words are never shared
    std::vector<Shard> shards_of(size_t n, size_t n_slots)
    {
        std::vector<Shard> out;
        const size_t per = ((n + n_slots - 1) / n_slots + 31) / 32 * 32;
        for (size_t i = 0; i < n_slots; ++i)
        {
            const size_t lo = std::min(n, i * per), hi = std::min(n, lo + per);
            out.push_back({lo, hi});
        }
        return out;
    }
}  // namespace

extern "C" {

int vb_dev_alloc(vb_ctx *ctx, int slot, size_t bytes, void **ptr_d)
{
    VB_TRY(check_slot(ctx, slot));
    if (!ptr_d)
    {
        return fail(VB_ERR_ARG, "null pointer");
    }
    VB_CUDA(cudaSetDevice(ctx->slots[slot].device));
    VB_CUDA(cudaMalloc(ptr_d, std::max<size_t>(bytes, 16)));
    return VB_OK;
}

int vb_dev_free(vb_ctx *ctx, int slot, void *ptr_d)
{
    VB_TRY(check_slot(ctx, slot));
    VB_CUDA(cudaSetDevice(ctx->slots[slot].device));
    VB_CUDA(cudaFree(ptr_d));
    return VB_OK;
}

int vb_dev_upload(vb_ctx *ctx, int slot, void *dst_d, const void *src, size_t bytes)
{
    VB_TRY(check_slot(ctx, slot));
    VB_CUDA(cudaSetDevice(ctx->slots[slot].device));
    VB_CUDA(cudaMemcpyAsync(dst_d, src, bytes, cudaMemcpyHostToDevice, ctx->slots[slot].stream));
    return VB_OK;
}

int vb_dev_download(vb_ctx *ctx, int slot, void *dst, const void *src_d, size_t bytes)
{
    VB_TRY(check_slot(ctx, slot));
    VB_CUDA(cudaSetDevice(ctx->slots[slot].device));
    VB_CUDA(cudaMemcpyAsync(dst, src_d, bytes, cudaMemcpyDeviceToHost, ctx->slots[slot].stream));
    VB_CUDA(cudaStreamSynchronize(ctx->slots[slot].stream));
    return VB_OK;
}

int vb_timer_start(vb_ctx *ctx, int slot)
{
    VB_TRY(check_slot(ctx, slot));
    VB_CUDA(cudaSetDevice(ctx->slots[slot].device));
    VB_CUDA(cudaEventRecord(ctx->slots[slot].ev0, ctx->slots[slot].stream));
    return VB_OK;
}

int vb_timer_stop(vb_ctx *ctx, int slot, double *milliseconds)
{
    VB_TRY(check_slot(ctx, slot));
    Slot &s = ctx->slots[slot];
    VB_CUDA(cudaSetDevice(s.device));
    VB_CUDA(cudaEventRecord(s.ev1, s.stream));
    VB_CUDA(cudaEventSynchronize(s.ev1));
    float ms = 0.f;
    VB_CUDA(cudaEventElapsedTime(&ms, s.ev0, s.ev1));
    if (milliseconds)
    {
        *milliseconds = ms;
    }
    return VB_OK;
}

int vb_popcount_d(vb_ctx *ctx, int slot, const uint32_t *bits_d, size_t n, uint64_t *count)
{
    VB_TRY(check_slot(ctx, slot));
    Slot &s = ctx->slots[slot];
    VB_CUDA(cudaSetDevice(s.device));
    VB_CUDA(s.scan_tmp.reserve(sizeof(unsigned long long)));
    VB_CUDA(cudaMemsetAsync(s.scan_tmp.p, 0, sizeof(unsigned long long), s.stream));
    popcount_kernel<<<s.sm_count * 4, 256, 0, s.stream>>>(bits_d, n, s.scan_tmp.as<unsigned long long>());
    VB_CUDA(cudaGetLastError());
    ctx->launches += 1;
    unsigned long long host = 0;
    VB_CUDA(cudaMemcpyAsync(&host, s.scan_tmp.p, sizeof(host), cudaMemcpyDeviceToHost, s.stream));
    VB_CUDA(cudaStreamSynchronize(s.stream));
    if (count)
    {
        *count = host;
    }
    return VB_OK;
}


int vb_fp32_peak_d(vb_ctx *ctx, int slot, double *tflops)
{
    VB_TRY(check_slot(ctx, slot));
    Slot &s = ctx->slots[slot];
    VB_CUDA(cudaSetDevice(s.device));
    VB_CUDA(s.scan_tmp.reserve(sizeof(unsigned long long)));
    const int iters = 4096, ctas = s.sm_count * 8;
    double best = 0.0;
    for (int rep = 0; rep < 4; ++rep)
    {
        VB_CUDA(cudaEventRecord(s.ev0, s.stream));
        fma_peak_kernel<<<ctas, 256, 0, s.stream>>>(s.scan_tmp.as<float>(), iters, 0.999f, 1e-3f);
        VB_CUDA(cudaGetLastError());
        VB_CUDA(cudaEventRecord(s.ev1, s.stream));
        VB_CUDA(cudaEventSynchronize(s.ev1));
        ctx->launches += 1;
        float ms = 0.f;
        VB_CUDA(cudaEventElapsedTime(&ms, s.ev0, s.ev1));
        const double flops = 2.0 * 64.0 * iters * 256.0 * ctas;
        if (rep > 0 && ms > 0.f)
        {
            best = std::max(best, flops / (ms * 1e-3) / 1e12);
        }
    }
    if (tflops)
    {
        *tflops = best;
    }
    return VB_OK;
}

int vb_fkcc_d(vb_ctx *ctx, int slot, vb_robot robot, const float *base, const vb_env *env, const float *configs_d,
              size_t n, uint32_t *valid_bits_d, float *margin_d)
{
    if (!env || (n && (!configs_d || !valid_bits_d)))
    {
        return fail(VB_ERR_ARG, "vb_fkcc_d: null argument");
    }
    Query q;
    VB_TRY(begin_query(ctx, slot, robot, base, env, q));
    return run_fkcc_d(q, configs_d, n, valid_bits_d, margin_d);
}

int vb_fkcc_work_d(vb_ctx *ctx, int slot, vb_robot robot, const float *base, const vb_env *env, const float *configs_d,
                   size_t n, uint32_t *valid_bits_d, float *work_d)
{
    if (!env || (n && (!configs_d || !valid_bits_d || !work_d)))
    {
        return fail(VB_ERR_ARG, "vb_fkcc_work_d: null argument");
    }
    Query q;
    VB_TRY(begin_query(ctx, slot, robot, base, env, q));
    return run_fkcc_d(q, configs_d, n, valid_bits_d, work_d, 2);
}


/* ---- batch edge callers ---------------------------------------------------- */
// Path::validate (planning/plan.hh:155-168): every consecutive edge in one batch; the index of the first
// invalid edge (the reference's loop stops there), or -1.
int vb_validate_path(vb_ctx *ctx, vb_robot robot, const float *base, const vb_env *env, const float *waypoints, size_t n,
                     int resolution, int64_t *first_invalid_edge)
{
    if (!ctx || !env || !first_invalid_edge || (n && !waypoints))
    {
        return fail(VB_ERR_ARG, "vb_validate_path: null argument");
    }
    const vb::RobotOps *ops = ops_of(robot);
    if (!ops)
    {
        return fail(VB_ERR_ARG, "unknown robot id %d", static_cast<int>(robot));
    }
    *first_invalid_edge = -1;
    if (n < 2)
    {
        return VB_OK;
    }
    const size_t edges = n - 1;
    std::vector<uint32_t> bits((edges + 31) / 32);
    VB_TRY(vb_validate_motion(ctx, robot, base, env, waypoints, waypoints + ops->dim, edges, resolution, bits.data(), nullptr));
    for (size_t e = 0; e < edges; ++e)
    {
        if (!((bits[e >> 5] >> (e & 31)) & 1u))
        {
            *first_invalid_edge = static_cast<int64_t>(e);
            break;
        }
    }
    return VB_OK;
}

// planning::shortcut_path (planning/simplify.hh:115-140).  The reference scans i ascending, j descending and
// calls validate_motion once per candidate; here ALL candidates (a, b > a + 1) of the original path are
// validated in one batch and the same scan is replayed on the verdict matrix.  Erasing waypoints never
// creates a new candidate pair, so the surviving waypoints are exactly the reference's.
int vb_shortcut_path(vb_ctx *ctx, vb_robot robot, const float *base, const vb_env *env, const float *waypoints, size_t n,
                     int resolution, uint8_t *keep, size_t *n_kept)
{
    if (!ctx || !env || (n && (!waypoints || !keep)))
    {
        return fail(VB_ERR_ARG, "vb_shortcut_path: null argument");
    }
    const vb::RobotOps *ops = ops_of(robot);
    if (!ops)
    {
        return fail(VB_ERR_ARG, "unknown robot id %d", static_cast<int>(robot));
    }
    for (size_t i = 0; i < n; ++i)
    {
        keep[i] = 1;
    }
    if (n_kept)
    {
        *n_kept = n;
    }
    if (n < 3)
    {
        return VB_OK;
    }
    const size_t dim = static_cast<size_t>(ops->dim);
    const size_t pairs = (n - 1) * (n - 2) / 2;
    // pair (a, b), b >= a + 2, at index offset(a) + (b - a - 2), offset(a) = sum_{k<a} (n - 2 - k)
    auto offset = [n](size_t a) { return a * (n - 2) - a * (a - 1) / 2; };
    std::vector<float> starts(pairs * dim), goals(pairs * dim);
    for (size_t a = 0; a + 2 < n; ++a)
    {
        for (size_t b = a + 2; b < n; ++b)
        {
            const size_t e = offset(a) + (b - a - 2);
            std::copy(waypoints + a * dim, waypoints + (a + 1) * dim, starts.begin() + e * dim);
            std::copy(waypoints + b * dim, waypoints + (b + 1) * dim, goals.begin() + e * dim);
        }
    }
    std::vector<uint32_t> bits((pairs + 31) / 32);
    VB_TRY(vb_validate_motion(ctx, robot, base, env, starts.data(), goals.data(), pairs, resolution, bits.data(), nullptr));
    auto valid = [&](size_t a, size_t b) {
        const size_t e = offset(a) + (b - a - 2);
        return ((bits[e >> 5] >> (e & 31)) & 1u) != 0u;
    };
    std::vector<size_t> path(n);
    for (size_t i = 0; i < n; ++i)
    {
        path[i] = i;
    }
    for (size_t i = 0; i + 2 < path.size(); ++i)
    {
        for (size_t j = path.size() - 1; j > i + 1; --j)
        {
            if (valid(path[i], path[j]))
            {
                for (size_t k = i + 1; k < j; ++k)
                {
                    keep[path[k]] = 0;
                }
                path.erase(path.begin() + static_cast<std::ptrdiff_t>(i) + 1, path.begin() + static_cast<std::ptrdiff_t>(j));
                break;
            }
        }
    }
    if (n_kept)
    {
        *n_kept = path.size();
    }
    return VB_OK;
}

int vb_validate_motion_d(vb_ctx *ctx, int slot, vb_robot robot, const float *base, const vb_env *env,
                         const float *starts_d, const float *goals_d, size_t n, int resolution,
                         uint32_t *valid_bits_d, float *margin_d)
{
    if (!env || (n && (!starts_d || !goals_d || !valid_bits_d)))
    {
        return fail(VB_ERR_ARG, "vb_validate_motion_d: null argument");
    }
    Query q;
    VB_TRY(begin_query(ctx, slot, robot, base, env, q));
    return run_motion_d(q, starts_d, goals_d, nullptr, false, n, resolution, valid_bits_d, margin_d);
}

int vb_validate_vector_d(vb_ctx *ctx, int slot, vb_robot robot, const float *base, const vb_env *env,
                         const float *starts_d, const float *vectors_d, const float *distances_d, size_t n,
                         int resolution, uint32_t *valid_bits_d, float *margin_d)
{
    if (!env || (n && (!starts_d || !vectors_d || !distances_d || !valid_bits_d)))
    {
        return fail(VB_ERR_ARG, "vb_validate_vector_d: null argument");
    }
    Query q;
    VB_TRY(begin_query(ctx, slot, robot, base, env, q));
    return run_motion_d(q, starts_d, vectors_d, distances_d, true, n, resolution, valid_bits_d, margin_d);
}

}  // extern "C"

/* ---- allocation-free motion validation ---------------------------------------------------------- */
// The prefix-array form of motion validation (edge_count_kernel -> exclusive scan -> motion_kernel: one thread per
// interpolant) needs scratch memory: the interpolant counts, their exclusive prefix and the scan's temporary storage.
// Callers that must not allocate (CUDA graph capture, tight serving loops) get it over a fixed file-scope device
// array.  Every device that loads the library
// holds the array until exit; every call shares it, so calls that overlap on the GPU race on it.
namespace
{
    constexpr size_t kNoAllocMaxEdges = VB_NOALLOC_MAX_EDGES;
    constexpr size_t kNoAllocWords = (kNoAllocMaxEdges + 1 + 63) / 64 * 64;  // counts / prefix: 256-byte multiples
    constexpr size_t kNoAllocScanBytes = size_t(64) << 10;                   // cub::DeviceScan temporary storage
    constexpr size_t kNoAllocBytes = 2 * kNoAllocWords * sizeof(uint32_t) + kNoAllocScanBytes;

    static __device__ __align__(256) unsigned char vb_noalloc_scratch[kNoAllocBytes];

    // In host code `vb_noalloc_scratch` is only a handle for the runtime's symbol table; the device address is looked
    // up once per device, with that device current.
    constexpr int kNoAllocMaxDevices = 64;
    std::atomic<unsigned char *> g_noalloc_addr[kNoAllocMaxDevices];

    int noalloc_scratch(int device, unsigned char **out)
    {
        if (device < 0 || device >= kNoAllocMaxDevices)
        {
            return fail(VB_ERR_ARG, "device id %d beyond the %d the static scratch table covers", device, kNoAllocMaxDevices);
        }
        unsigned char *p = g_noalloc_addr[device].load(std::memory_order_acquire);
        if (!p)
        {
            VB_CUDA(cudaSetDevice(device));
            void *addr = nullptr;
            VB_CUDA(cudaGetSymbolAddress(&addr, vb_noalloc_scratch));
            if (!addr)
            {
                return fail(VB_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the static scratch array");
            }
            p = static_cast<unsigned char *>(addr);
            g_noalloc_addr[device].store(p, std::memory_order_release);  // same value from every thread: a benign race
        }
        *out = p;
        return VB_OK;
    }
}  // namespace

extern "C" {

size_t vb_noalloc_max_edges(void)
{
    return kNoAllocMaxEdges;
}

int vb_validate_motion_noalloc_d(vb_ctx *ctx, int slot, vb_robot robot, const float *base, const vb_env *env,
                                 const float *starts_d, const float *goals_d, size_t n, int resolution,
                                 uint32_t *valid_bits_d)
{
    if (!env || (n && (!starts_d || !goals_d || !valid_bits_d)))
    {
        return fail(VB_ERR_ARG, "vb_validate_motion_noalloc_d: null argument");
    }
    if (n > kNoAllocMaxEdges)
    {
        return fail(VB_ERR_ARG, "vb_validate_motion_noalloc_d: %zu edges exceed the static scratch array's maximum of %zu "
                                "per call (split the batch, or use vb_validate_motion_d)", n, kNoAllocMaxEdges);
    }
    Query q;
    VB_TRY(begin_query(ctx, slot, robot, base, env, q));
    if (q.args.attach)
    {
        return fail(VB_ERR_STATE, "vb_validate_motion_noalloc_d: environments with an attachment need a per-launch scratch "
                                  "column that is allocated on demand; use vb_validate_motion_d");
    }
    if (n == 0)
    {
        return VB_OK;
    }
    Slot &s = *q.slot;
    unsigned char *scratch = nullptr;
    VB_TRY(noalloc_scratch(s.device, &scratch));
    uint32_t *counts = reinterpret_cast<uint32_t *>(scratch);
    uint32_t *prefix = counts + kNoAllocWords;
    void *scan_tmp = prefix + kNoAllocWords;
    q.args.cfg = starts_d;
    q.args.cfg_b = goals_d;
    q.args.n = n;
    q.args.bits = valid_bits_d;
    q.args.margin = nullptr;
    q.args.resolution = resolution > 0 ? resolution : q.ops->resolution;
    // the array keeps the previous call's contents: the terminating count must start at zero, everything else the
    // kernels read is rewritten first (counts[0..n) by edge_count, prefix[0..n] by the scan)
    VB_CUDA(cudaMemsetAsync(counts + n, 0, sizeof(uint32_t), s.stream));
    fill_bits_kernel<<<static_cast<unsigned>(((n + 31) / 32 + 255) / 256), 256, 0, s.stream>>>(valid_bits_d, n);
    VB_CUDA(cudaGetLastError());
    VB_CUDA(q.ops->edge_count(q.args, counts));  // (clears the verdict bit of edges too long for the 32-bit prefix)
    size_t tmp_bytes = 0;
    VB_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, counts, prefix, static_cast<int>(n + 1), s.stream));
    if (tmp_bytes > kNoAllocScanBytes)
    {
        return fail(VB_ERR_STATE, "vb_validate_motion_noalloc_d: the scan wants %zu bytes of temporary storage, the static "
                                  "array reserves %zu", tmp_bytes, kNoAllocScanBytes);
    }
    VB_CUDA(cub::DeviceScan::ExclusiveSum(scan_tmp, tmp_bytes, counts, prefix, static_cast<int>(n + 1), s.stream));
    q.args.prefix = prefix;
    int grid = 1;
    VB_TRY(grid_for(q, n * 8 * 4, &grid));
    VB_CUDA(q.ops->motion(q.args, false, grid));
    ctx->launches += 4;
    return VB_OK;
}

int vb_fkcc_random_d(vb_ctx *ctx, int slot, vb_robot robot, const float *base, const vb_env *env, uint64_t seed,
                     uint64_t first, size_t n, uint32_t *valid_bits_d)
{
    if (!env || (n && !valid_bits_d))
    {
        return fail(VB_ERR_ARG, "vb_fkcc_random_d: null argument");
    }
    Query q;
    VB_TRY(begin_query(ctx, slot, robot, base, env, q));
    if (n == 0)
    {
        return VB_OK;
    }
    q.args.n = n;
    q.args.seed = seed;
    q.args.bits = valid_bits_d;
    int grid = 1;
    VB_TRY(grid_for(q, n, &grid));
    VB_CUDA(q.ops->fkcc_random(q.args, first, grid));
    ctx->launches += 1;
    return VB_OK;
}

/* ---- host-buffer entry points: shard -> H2D -> kernels -> D2H of the verdict words ---------- */
int vb_sphere_fk(vb_ctx *ctx, vb_robot robot, const float *base, const float *configs, size_t n, float *spheres)
{
    if (!ctx || (n && (!configs || !spheres)))
    {
        return fail(VB_ERR_ARG, "vb_sphere_fk: null argument");
    }
    const vb::RobotOps *ops = ops_of(robot);
    if (!ops)
    {
        return fail(VB_ERR_ARG, "unknown robot id %d", static_cast<int>(robot));
    }
    const auto shards = shards_of(n, ctx->slots.size());
    for (size_t i = 0; i < shards.size(); ++i)
    {
        const size_t m = shards[i].hi - shards[i].lo;
        if (m == 0)
        {
            continue;
        }
        Query q;
        VB_TRY(begin_query(ctx, static_cast<int>(i), robot, base, nullptr, q));
        Slot &s = *q.slot;
        const size_t out_floats = m * ops->n_spheres * 4;
        VB_CUDA(s.in_a.reserve(m * ops->dim * sizeof(float)));
        VB_CUDA(s.out.reserve(out_floats * sizeof(float)));
        VB_CUDA(cudaMemcpyAsync(s.in_a.p, configs + shards[i].lo * ops->dim, m * ops->dim * sizeof(float),
                                cudaMemcpyHostToDevice, s.stream));
        q.args.cfg = s.in_a.as<float>();
        q.args.n = m;
        q.args.spheres = s.out.as<float>();
        VB_CUDA(ops->sphere_fk(q.args));
        ctx->launches += 1;
        VB_CUDA(cudaMemcpyAsync(spheres + shards[i].lo * ops->n_spheres * 4, s.out.p, out_floats * sizeof(float),
                                cudaMemcpyDeviceToHost, s.stream));
    }
    return vb_sync(ctx);
}

// Host buffers are processed in chunks on two streams per device, so that the host->device copy of
// chunk k+1 and the device->host copy of chunk k-1 run under the kernel of chunk k.
static constexpr size_t kHostChunk = size_t(1) << 22;  // configurations per chunk (multiple of 32)

static int fkcc_host(vb_ctx *ctx, vb_robot robot, const float *base, const vb_env *env, const float *configs, size_t n,
                     uint32_t *valid_bits, float *margin, bool limits)
{
    if (!ctx || !env || (n && (!configs || !valid_bits)))
    {
        return fail(VB_ERR_ARG, "vb_fkcc: null argument");
    }
    const vb::RobotOps *ops = ops_of(robot);
    if (!ops)
    {
        return fail(VB_ERR_ARG, "unknown robot id %d", static_cast<int>(robot));
    }
    const auto shards = shards_of(n, ctx->slots.size());
    for (size_t i = 0; i < shards.size(); ++i)
    {
        const size_t m = shards[i].hi - shards[i].lo;
        if (m == 0)
        {
            continue;
        }
        Query q;
        VB_TRY(begin_query(ctx, static_cast<int>(i), robot, base, env, q));
        Slot &s = *q.slot;
        const size_t chunk = std::min(m, kHostChunk);
        VB_CUDA(s.in_a.reserve(chunk * ops->dim * sizeof(float)));
        if (m > chunk)
        {
            VB_CUDA(s.in_a2.reserve(chunk * ops->dim * sizeof(float)));
        }
        VB_CUDA(s.bits.reserve((m + 31) / 32 * sizeof(uint32_t)));
        if (margin)
        {
            VB_CUDA(s.margin.reserve(m * sizeof(float)));
        }
        // the second stream starts after whatever the caller queued on the first
        VB_CUDA(cudaEventRecord(s.evj, s.stream));
        VB_CUDA(cudaStreamWaitEvent(s.stream2, s.evj, 0));
        size_t k = 0;
        for (size_t lo = 0; lo < m; lo += chunk, ++k)
        {
            const size_t c = std::min(chunk, m - lo);
            const bool second = (k & 1) && !q.args.attach;  // attachment runs share one scratch buffer: one stream
            cudaStream_t st = second ? s.stream2 : s.stream;
            float *in = (second ? s.in_a2 : s.in_a).as<float>();
            uint32_t *bits_d = s.bits.as<uint32_t>() + lo / 32;
            float *margin_d = margin ? s.margin.as<float>() + lo : nullptr;
            q.args.stream = st;
            VB_CUDA(cudaMemcpyAsync(in, configs + (shards[i].lo + lo) * ops->dim, c * ops->dim * sizeof(float),
                                    cudaMemcpyHostToDevice, st));
            VB_TRY(run_fkcc_d(q, in, c, bits_d, margin_d));
            if (limits)
            {
                VB_CUDA(ops->limits(q.args));
                ctx->launches += 1;
            }
            VB_CUDA(cudaMemcpyAsync(valid_bits + (shards[i].lo + lo) / 32, bits_d, (c + 31) / 32 * sizeof(uint32_t),
                                    cudaMemcpyDeviceToHost, st));
            if (margin)
            {
                VB_CUDA(cudaMemcpyAsync(margin + shards[i].lo + lo, margin_d, c * sizeof(float), cudaMemcpyDeviceToHost, st));
            }
        }
    }
    return vb_sync(ctx);
}

int vb_fkcc(vb_ctx *ctx, vb_robot robot, const float *base, const vb_env *env, const float *configs, size_t n,
            uint32_t *valid_bits, float *margin_or_null)
{
    return fkcc_host(ctx, robot, base, env, configs, n, valid_bits, margin_or_null, false);
}

int vb_validate(vb_ctx *ctx, vb_robot robot, const float *base, const vb_env *env, const float *configs, size_t n,
                uint32_t *valid_bits)
{
    return fkcc_host(ctx, robot, base, env, configs, n, valid_bits, nullptr, true);
}

// Host-buffer edges: chunks on two streams per device like fkcc_host, so that the copies of chunk k+1 run under the
// kernel of chunk k.  second = goals, or vectors when dist != nullptr.
static constexpr size_t kEdgeChunk = size_t(1) << 19;  // edges per chunk (multiple of 32)

static int motion_host(vb_ctx *ctx, vb_robot robot, const float *base, const vb_env *env, const float *starts,
                       const float *second, const float *dist, size_t n, int resolution, uint32_t *valid_bits, float *margin,
                       const char *who)
{
    if (!ctx || !env || (n && (!starts || !second || !valid_bits)))
    {
        return fail(VB_ERR_ARG, "%s: null argument", who);
    }
    const vb::RobotOps *ops = ops_of(robot);
    if (!ops)
    {
        return fail(VB_ERR_ARG, "unknown robot id %d", static_cast<int>(robot));
    }
    const auto shards = shards_of(n, ctx->slots.size());
    for (size_t i = 0; i < shards.size(); ++i)
    {
        const size_t m = shards[i].hi - shards[i].lo;
        if (m == 0)
        {
            continue;
        }
        Query q;
        VB_TRY(begin_query(ctx, static_cast<int>(i), robot, base, env, q));
        Slot &s = *q.slot;
        const size_t chunk = std::min(m, kEdgeChunk), row = ops->dim * sizeof(float);
        const bool two = m > chunk && !q.args.attach;  // attachment runs share one scratch buffer: one stream
        VB_CUDA(s.in_a.reserve(chunk * row));
        VB_CUDA(s.in_b.reserve(chunk * row));
        if (dist)
        {
            VB_CUDA(s.in_d.reserve(chunk * sizeof(float)));
        }
        if (two)
        {
            VB_CUDA(s.in_a2.reserve(chunk * row));
            VB_CUDA(s.in_b2.reserve(chunk * row));
            if (dist)
            {
                VB_CUDA(s.in_d2.reserve(chunk * sizeof(float)));
            }
        }
        VB_CUDA(s.bits.reserve((m + 31) / 32 * sizeof(uint32_t)));
        if (margin)
        {
            VB_CUDA(s.margin.reserve(m * sizeof(float)));
        }
        VB_CUDA(cudaEventRecord(s.evj, s.stream));
        VB_CUDA(cudaStreamWaitEvent(s.stream2, s.evj, 0));
        size_t k = 0;
        for (size_t lo = 0; lo < m; lo += chunk, ++k)
        {
            const size_t c = std::min(chunk, m - lo), g0 = shards[i].lo + lo;
            const bool alt = two && (k & 1);
            cudaStream_t st = alt ? s.stream2 : s.stream;
            float *a_d = (alt ? s.in_a2 : s.in_a).as<float>(), *b_d = (alt ? s.in_b2 : s.in_b).as<float>();
            float *d_d = dist ? (alt ? s.in_d2 : s.in_d).as<float>() : nullptr;
            uint32_t *bits_d = s.bits.as<uint32_t>() + lo / 32;
            float *margin_d = margin ? s.margin.as<float>() + lo : nullptr;
            q.args.stream = st;
            VB_CUDA(cudaMemcpyAsync(a_d, starts + g0 * ops->dim, c * row, cudaMemcpyHostToDevice, st));
            VB_CUDA(cudaMemcpyAsync(b_d, second + g0 * ops->dim, c * row, cudaMemcpyHostToDevice, st));
            if (dist)
            {
                VB_CUDA(cudaMemcpyAsync(d_d, dist + g0, c * sizeof(float), cudaMemcpyHostToDevice, st));
            }
            VB_TRY(run_motion_d(q, a_d, b_d, d_d, dist != nullptr, c, resolution, bits_d, margin_d));
            VB_CUDA(cudaMemcpyAsync(valid_bits + g0 / 32, bits_d, (c + 31) / 32 * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
            if (margin)
            {
                VB_CUDA(cudaMemcpyAsync(margin + g0, margin_d, c * sizeof(float), cudaMemcpyDeviceToHost, st));
            }
        }
    }
    return vb_sync(ctx);
}

int vb_validate_motion(vb_ctx *ctx, vb_robot robot, const float *base, const vb_env *env, const float *starts,
                       const float *goals, size_t n, int resolution, uint32_t *valid_bits, float *margin)
{
    return motion_host(ctx, robot, base, env, starts, goals, nullptr, n, resolution, valid_bits, margin, "vb_validate_motion");
}

int vb_validate_vector(vb_ctx *ctx, vb_robot robot, const float *base, const vb_env *env, const float *starts,
                       const float *vectors, const float *distances, size_t n, int resolution, uint32_t *valid_bits,
                       float *margin)
{
    if (n && !distances)
    {
        return fail(VB_ERR_ARG, "vb_validate_vector: null distances");
    }
    return motion_host(ctx, robot, base, env, starts, vectors, distances, n, resolution, valid_bits, margin, "vb_validate_vector");
}

// RRTC's connect loop (planning/rrtc.hh:172-194): from `from` towards `to` in n_ext = ceil(dist / range) equal
// increments, each validated with validate_vector(prior, increment, increment_length) and prior += increment
// accumulated in float exactly as the reference does; stops at the first failing increment.  All increments of all
// attempts go to the device in ONE batch; n_valid[i] = the reference's i_extension when its loop ends (== n_ext[i]
// iff the trees connect).  The reference's free_index < max_samples condition is the caller's to apply.
int vb_connect_increments(vb_ctx *ctx, vb_robot robot, const float *base, const vb_env *env, const float *from,
                          const float *to, const float *distances, size_t n, float range, int resolution,
                          uint32_t *n_valid, uint32_t *n_ext_or_null)
{
    if (!ctx || !env || (n && (!from || !to || !distances || !n_valid)) || !(range > 0.f))
    {
        return fail(VB_ERR_ARG, "vb_connect_increments: bad arguments");
    }
    const vb::RobotOps *ops = ops_of(robot);
    if (!ops)
    {
        return fail(VB_ERR_ARG, "unknown robot id %d", static_cast<int>(robot));
    }
    const size_t dim = static_cast<size_t>(ops->dim);
    std::vector<size_t> first(n + 1, 0);
    for (size_t i = 0; i < n; ++i)
    {
        const float ne = std::ceil(distances[i] / range);
        if (!(ne < 1048576.f))
        {
            return fail(VB_ERR_ARG, "vb_connect_increments: attempt %zu needs %g extensions", i, static_cast<double>(ne));
        }
        first[i + 1] = first[i] + static_cast<size_t>(ne < 0.f ? 0.f : ne);
    }
    const size_t total = first[n];
    std::vector<float> priors(total * dim), incs(total * dim), lens(total);
    for (size_t i = 0; i < n; ++i)
    {
        const size_t ne = first[i + 1] - first[i];
        if (ne == 0)
        {
            continue;
        }
        const float inv = 1.0F / static_cast<float>(ne), len = distances[i] / static_cast<float>(ne);
        float prior[16], inc[16];
        for (size_t j = 0; j < dim; ++j)
        {
            prior[j] = from[i * dim + j];
            inc[j] = (to[i * dim + j] - from[i * dim + j]) * inv;
        }
        for (size_t k = 0; k < ne; ++k)
        {
            float *p = priors.data() + (first[i] + k) * dim, *v = incs.data() + (first[i] + k) * dim;
            for (size_t j = 0; j < dim; ++j)
            {
                p[j] = prior[j];
                v[j] = inc[j];
                prior[j] = prior[j] + inc[j];
            }
            lens[first[i] + k] = len;
        }
    }
    std::vector<uint32_t> bits((total + 31) / 32);
    VB_TRY(motion_host(ctx, robot, base, env, priors.data(), incs.data(), lens.data(), total, resolution, bits.data(), nullptr,
                       "vb_connect_increments"));
    for (size_t i = 0; i < n; ++i)
    {
        size_t k = first[i];
        while (k < first[i + 1] && ((bits[k >> 5] >> (k & 31)) & 1u))
        {
            ++k;
        }
        n_valid[i] = static_cast<uint32_t>(k - first[i]);
        if (n_ext_or_null)
        {
            n_ext_or_null[i] = static_cast<uint32_t>(first[i + 1] - first[i]);
        }
    }
    return VB_OK;
}

}  // extern "C"

namespace
{
    // Robot-robot sphere pairs of composite configurations (the fork's planned
    // MultiRobotEnvironment::check_inter_robot_collisions, MR_DEV_PLAN.md:176-198; primitive:
    // collision/sphere_sphere.hh:9-23 on world-frame spheres of Robot::sphere_fk).  One warp per composite
    // configuration: robot A's spheres are parked in shared memory, every lane owns spheres of robot B and
    // sweeps all of A against them; an axis-aligned box test on the two sphere sets skips configurations
    // whose robots are nowhere near each other.  ANDs "no pair overlaps" into the verdict words.
    constexpr int kPairWarps = 8;
    constexpr int kPairMaxSpheres = 128;

    __global__ void __launch_bounds__(kPairWarps * 32)
        pairs_kernel(const float4 *__restrict__ sa, int n_a, const float4 *__restrict__ sb, int n_b, size_t n, uint32_t *bits)
    {
        __shared__ float4 park[kPairWarps][kPairMaxSpheres];
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        const size_t n_warps = static_cast<size_t>(gridDim.x) * kPairWarps;
        for (size_t i = static_cast<size_t>(blockIdx.x) * kPairWarps + warp; i < n; i += n_warps)
        {
            const float4 *A = sa + i * n_a, *B = sb + i * n_b;
            float lo_a[3] = {3e38f, 3e38f, 3e38f}, hi_a[3] = {-3e38f, -3e38f, -3e38f};
            float lo_b[3] = {3e38f, 3e38f, 3e38f}, hi_b[3] = {-3e38f, -3e38f, -3e38f};
            for (int k = lane; k < n_a; k += 32)
            {
                const float4 s = A[k];
                park[warp][k] = s;
                const float c[3] = {s.x, s.y, s.z};
#pragma unroll
                for (int d = 0; d < 3; ++d)
                {
                    lo_a[d] = fminf(lo_a[d], c[d] - s.w);
                    hi_a[d] = fmaxf(hi_a[d], c[d] + s.w);
                }
            }
            for (int k = lane; k < n_b; k += 32)
            {
                const float4 s = B[k];
                const float c[3] = {s.x, s.y, s.z};
#pragma unroll
                for (int d = 0; d < 3; ++d)
                {
                    lo_b[d] = fminf(lo_b[d], c[d] - s.w);
                    hi_b[d] = fmaxf(hi_b[d], c[d] + s.w);
                }
            }
            bool apart = false;
#pragma unroll
            for (int d = 0; d < 3; ++d)
            {
                for (int o = 16; o > 0; o >>= 1)
                {
                    lo_a[d] = fminf(lo_a[d], __shfl_xor_sync(0xffffffffu, lo_a[d], o));
                    hi_a[d] = fmaxf(hi_a[d], __shfl_xor_sync(0xffffffffu, hi_a[d], o));
                    lo_b[d] = fminf(lo_b[d], __shfl_xor_sync(0xffffffffu, lo_b[d], o));
                    hi_b[d] = fmaxf(hi_b[d], __shfl_xor_sync(0xffffffffu, hi_b[d], o));
                }
                apart = apart || lo_a[d] > hi_b[d] || lo_b[d] > hi_a[d];
            }
            __syncwarp();
            bool hit = false;
            if (!apart)
            {
                for (int k = lane; k < n_b; k += 32)
                {
                    const float4 b = B[k];
#pragma unroll 4
                    for (int j = 0; j < n_a; ++j)
                    {
                        const float4 a = park[warp][j];
                        const float dx = a.x - b.x, dy = a.y - b.y, dz = a.z - b.z, rs = a.w + b.w;
                        hit = hit || (dx * dx + dy * dy + dz * dz - rs * rs < 0.f);
                    }
                }
            }
            const bool any = __any_sync(0xffffffffu, hit);
            if (any && lane == 0)
            {
                atomicAnd(bits + (i >> 5), ~(1u << (i & 31)));
            }
            __syncwarp();
        }
    }
}  // namespace

extern "C" {

// composite configurations of two robots of the SAME kind (the Panda grid): one launch of the fused kernel
static int run_pairs_fused_d(Query &q, const float *base_b, const float *cfg_a_d, const float *cfg_b_d, size_t n, uint32_t *bits_d)
{
    if (n == 0)
    {
        return VB_OK;
    }
    q.args.cfg = cfg_a_d;
    q.args.cfg_b = cfg_b_d;
    q.args.n = n;
    q.args.bits = bits_d;
    for (int k = 0; k < 3; ++k)
    {
        q.args.base_b[k] = base_b ? base_b[k] : 0.f;
    }
    int blocks = 0, threads = 0;
    VB_CUDA(q.ops->pairs_occupancy(q.args, &blocks, &threads));
    if (blocks < 1)
    {
        return fail(VB_ERR_CUDA, "%s: the composite kernel does not fit one CTA's shared memory", q.ops->name);
    }
    const size_t need = (n + threads - 1) / threads;
    const int grid = static_cast<int>(std::max<size_t>(1, std::min<size_t>(need, static_cast<size_t>(blocks) * q.slot->sm_count)));
    VB_CUDA(q.ops->pairs_fused(q.args, grid));
    q.ctx->launches += 1;
    return VB_OK;
}

int vb_validate_pairs_d(vb_ctx *ctx, int slot, vb_robot robot, const float *base_a, const float *base_b, const vb_env *env,
                        const float *cfg_a_d, const float *cfg_b_d, size_t n, uint32_t *valid_bits_d)
{
    if (!env || (n && (!cfg_a_d || !cfg_b_d || !valid_bits_d)))
    {
        return fail(VB_ERR_ARG, "vb_validate_pairs_d: null argument");
    }
    Query q;
    VB_TRY(begin_query(ctx, slot, robot, base_a, env, q));
    if (q.args.attach)
    {
        return fail(VB_ERR_STATE, "vb_validate_pairs_d: environments with an attachment are not supported");
    }
    return run_pairs_fused_d(q, base_b, cfg_a_d, cfg_b_d, n, valid_bits_d);
}

static int pairs_host_fused(vb_ctx *ctx, vb_robot robot, const float *base_a, const float *base_b, const vb_env *env,
                            const float *cfg_a, const float *cfg_b, size_t n, uint32_t *valid_bits)
{
    const vb::RobotOps *ops = ops_of(robot);
    const auto shards = shards_of(n, ctx->slots.size());
    for (size_t i = 0; i < shards.size(); ++i)
    {
        const size_t m = shards[i].hi - shards[i].lo;
        if (m == 0)
        {
            continue;
        }
        Query q;
        VB_TRY(begin_query(ctx, static_cast<int>(i), robot, base_a, env, q));
        Slot &s = *q.slot;
        const size_t chunk = std::min(m, kHostChunk / 2), row = ops->dim * sizeof(float);
        const bool two = m > chunk;
        VB_CUDA(s.in_a.reserve(chunk * row));
        VB_CUDA(s.in_b.reserve(chunk * row));
        if (two)
        {
            VB_CUDA(s.in_a2.reserve(chunk * row));
            VB_CUDA(s.in_b2.reserve(chunk * row));
        }
        VB_CUDA(s.bits.reserve((m + 31) / 32 * sizeof(uint32_t)));
        VB_CUDA(cudaEventRecord(s.evj, s.stream));
        VB_CUDA(cudaStreamWaitEvent(s.stream2, s.evj, 0));
        size_t k = 0;
        for (size_t lo = 0; lo < m; lo += chunk, ++k)
        {
            const size_t c = std::min(chunk, m - lo), g0 = shards[i].lo + lo;
            const bool alt = two && (k & 1);
            cudaStream_t st = alt ? s.stream2 : s.stream;
            float *a_d = (alt ? s.in_a2 : s.in_a).as<float>(), *b_d = (alt ? s.in_b2 : s.in_b).as<float>();
            uint32_t *bits_d = s.bits.as<uint32_t>() + lo / 32;
            q.args.stream = st;
            VB_CUDA(cudaMemcpyAsync(a_d, cfg_a + g0 * ops->dim, c * row, cudaMemcpyHostToDevice, st));
            VB_CUDA(cudaMemcpyAsync(b_d, cfg_b + g0 * ops->dim, c * row, cudaMemcpyHostToDevice, st));
            VB_TRY(run_pairs_fused_d(q, base_b, a_d, b_d, c, bits_d));
            VB_CUDA(cudaMemcpyAsync(valid_bits + g0 / 32, bits_d, (c + 31) / 32 * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
        }
    }
    return vb_sync(ctx);
}

static int pairs_host_unfused(vb_ctx *ctx, vb_robot robot_a, const float *base_a, vb_robot robot_b, const float *base_b,
                              const vb_env *env, const float *cfg_a, const float *cfg_b, size_t n, uint32_t *valid_bits);

int vb_validate_pairs(vb_ctx *ctx, vb_robot robot_a, const float *base_a, vb_robot robot_b, const float *base_b,
                      const vb_env *env, const float *cfg_a, const float *cfg_b, size_t n, uint32_t *valid_bits)
{
    if (!ctx || !env || (n && (!cfg_a || !cfg_b || !valid_bits)))
    {
        return fail(VB_ERR_ARG, "vb_validate_pairs: null argument");
    }
    const vb::RobotOps *oa = ops_of(robot_a), *ob = ops_of(robot_b);
    if (!oa || !ob)
    {
        return fail(VB_ERR_ARG, "vb_validate_pairs: unknown robot id");
    }
    if (robot_a == robot_b && !(oa->has_attach && env->has_attachment) && !std::getenv("VB_PAIRS_UNFUSED"))
    {
        // two robots of one kind (mr-vamp's Panda grid): the fused kernel, one launch per chunk
        return pairs_host_fused(ctx, robot_a, base_a, base_b, env, cfg_a, cfg_b, n, valid_bits);
    }
    return pairs_host_unfused(ctx, robot_a, base_a, robot_b, base_b, env, cfg_a, cfg_b, n, valid_bits);
}

// The robot-robot term alone (MR_DEV_PLAN.md:176-198 check_inter_robot_collisions): bit = 1 iff NO world-frame sphere
// of robot A overlaps a sphere of robot B; environment and self collision are not looked at.
int vb_inter_robot(vb_ctx *ctx, vb_robot robot_a, const float *base_a, vb_robot robot_b, const float *base_b, const float *cfg_a,
                   const float *cfg_b, size_t n, uint32_t *clear_bits)
{
    if (!ctx || (n && (!cfg_a || !cfg_b || !clear_bits)))
    {
        return fail(VB_ERR_ARG, "vb_inter_robot: null argument");
    }
    return pairs_host_unfused(ctx, robot_a, base_a, robot_b, base_b, nullptr, cfg_a, cfg_b, n, clear_bits);
}

// env == nullptr: the robot-robot term only (verdict words start as all ones)
static int pairs_host_unfused(vb_ctx *ctx, vb_robot robot_a, const float *base_a, vb_robot robot_b, const float *base_b,
                              const vb_env *env, const float *cfg_a, const float *cfg_b, size_t n, uint32_t *valid_bits)
{
    const vb::RobotOps *oa = ops_of(robot_a), *ob = ops_of(robot_b);
    if (!oa || !ob)
    {
        return fail(VB_ERR_ARG, "unknown robot id");
    }
    if (oa->n_spheres > kPairMaxSpheres)
    {
        return fail(VB_ERR_ARG, "vb_validate_pairs: robot A has more than %d spheres", kPairMaxSpheres);
    }
    // robots of different kinds: each robot's own validity (environment + self collision) on all devices, then the
    // robot-robot pairs from both sphere sets, chunked so that the two sphere buffers stay below ~1 GB per device
    if (env)
    {
        std::vector<uint32_t> bits_b((n + 31) / 32);
        VB_TRY(vb_fkcc(ctx, robot_a, base_a, env, cfg_a, n, valid_bits, nullptr));
        VB_TRY(vb_fkcc(ctx, robot_b, base_b, env, cfg_b, n, bits_b.data(), nullptr));
        for (size_t w = 0; w < bits_b.size(); ++w)
        {
            valid_bits[w] &= bits_b[w];
        }
    }
    else
    {
        for (size_t w = 0; w < (n + 31) / 32; ++w)
        {
            const size_t rem = n - w * 32;
            valid_bits[w] = rem >= 32 ? 0xffffffffu : ((1u << rem) - 1u);
        }
    }
    const size_t per_cfg = static_cast<size_t>(oa->n_spheres + ob->n_spheres) * sizeof(float4);
    const size_t chunk = std::max<size_t>(32, (size_t(1) << 30) / per_cfg / 32 * 32);
    const auto shards = shards_of(n, ctx->slots.size());
    for (size_t i = 0; i < shards.size(); ++i)
    {
        const size_t m = shards[i].hi - shards[i].lo;
        if (m == 0)
        {
            continue;
        }
        Query qa, qb;
        VB_TRY(begin_query(ctx, static_cast<int>(i), robot_a, base_a, nullptr, qa));
        VB_TRY(begin_query(ctx, static_cast<int>(i), robot_b, base_b, nullptr, qb));
        Slot &s = *qa.slot;
        const size_t c_max = std::min(m, chunk);
        VB_CUDA(s.in_a.reserve(c_max * oa->dim * sizeof(float)));
        VB_CUDA(s.in_b.reserve(c_max * ob->dim * sizeof(float)));
        VB_CUDA(s.out.reserve(c_max * per_cfg));
        VB_CUDA(s.bits.reserve((c_max + 31) / 32 * sizeof(uint32_t)));
        for (size_t lo = 0; lo < m; lo += chunk)
        {
            const size_t c = std::min(chunk, m - lo), g0 = shards[i].lo + lo, words = (c + 31) / 32;
            float4 *sph_a = s.out.as<float4>(), *sph_b = sph_a + c * oa->n_spheres;
            VB_CUDA(cudaMemcpyAsync(s.in_a.p, cfg_a + g0 * oa->dim, c * oa->dim * sizeof(float), cudaMemcpyHostToDevice, s.stream));
            VB_CUDA(cudaMemcpyAsync(s.in_b.p, cfg_b + g0 * ob->dim, c * ob->dim * sizeof(float), cudaMemcpyHostToDevice, s.stream));
            VB_CUDA(cudaMemcpyAsync(s.bits.p, valid_bits + g0 / 32, words * sizeof(uint32_t), cudaMemcpyHostToDevice, s.stream));
            qa.args.cfg = s.in_a.as<float>();
            qa.args.n = c;
            qa.args.spheres = reinterpret_cast<float *>(sph_a);
            VB_CUDA(oa->sphere_fk(qa.args));
            qb.args.cfg = s.in_b.as<float>();
            qb.args.n = c;
            qb.args.spheres = reinterpret_cast<float *>(sph_b);
            VB_CUDA(ob->sphere_fk(qb.args));
            const unsigned grid = static_cast<unsigned>(std::min<size_t>((c + kPairWarps - 1) / kPairWarps, size_t(s.sm_count) * 8));
            pairs_kernel<<<grid, kPairWarps * 32, 0, s.stream>>>(sph_a, oa->n_spheres, sph_b, ob->n_spheres, c, s.bits.as<uint32_t>());
            VB_CUDA(cudaGetLastError());
            ctx->launches += 3;
            VB_CUDA(cudaMemcpyAsync(valid_bits + g0 / 32, s.bits.p, words * sizeof(uint32_t), cudaMemcpyDeviceToHost, s.stream));
            VB_CUDA(cudaStreamSynchronize(s.stream));
        }
    }
    return vb_sync(ctx);
}

}  // extern "C"

namespace
{
    // Helper::sphere_validate (bindings/common.hh:155-170) -> sphere_environment_get_collisions
    // (collision/validity.hh:152-249): for ONE robot sphere, every primitive obstacle and height field it overlaps.
    // One thread per (configuration, sphere); object numbering = the library's sorted order (vb_env_object_ids).
    // The reference's min_distance break only skips obstacles that cannot touch the sphere, so all are tested.
    __global__ void sphere_objects_kernel(const vb::EnvView env, const float4 *__restrict__ spheres, size_t n_spheres, uint32_t cap,
                                          uint32_t *__restrict__ hits, uint32_t *__restrict__ counts)
    {
        const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
        if (i >= n_spheres)
        {
            return;
        }
        const float4 s = spheres[i];
        const vb::PrimIndex ix = vb::prim_index(env);
        uint32_t c = 0;
        for (int o = 0; o < ix.first[5]; ++o)
        {
            if (vb::object_hit(vb::object_dist(env.prims, ix, o, s.x, s.y, s.z), s.w))
            {
                if (c < cap)
                {
                    hits[i * cap + c] = static_cast<uint32_t>(o);
                }
                ++c;
            }
        }
        for (int k = 0; k < env.n_heightfields; ++k)
        {
            if (vb::heightfield_value(env.heightfields[k], s.x, s.y, s.z, s.w) < 0.f)
            {
                if (c < cap)
                {
                    hits[i * cap + c] = static_cast<uint32_t>(ix.first[5] + k);
                }
                ++c;
            }
        }
        counts[i] = c;
    }
}  // namespace

extern "C" {

int vb_sphere_validity(vb_ctx *ctx, vb_robot robot, const float *base, const vb_env *env, const float *configs, size_t n,
                       uint32_t cap, uint32_t *hits, uint32_t *counts)
{
    if (!ctx || !env || (n && (!configs || !counts || (cap && !hits))))
    {
        return fail(VB_ERR_ARG, "vb_sphere_validity: null argument");
    }
    if (n == 0)
    {
        return VB_OK;
    }
    Query q;
    VB_TRY(begin_query(ctx, 0, robot, base, env, q));
    Slot &s = *q.slot;
    const size_t total = n * static_cast<size_t>(q.ops->n_spheres);
    VB_CUDA(s.in_a.reserve(n * q.ops->dim * sizeof(float)));
    VB_CUDA(s.out.reserve(total * sizeof(float4)));
    VB_CUDA(s.margin.reserve(total * (static_cast<size_t>(cap) + 1) * sizeof(uint32_t)));
    VB_CUDA(cudaMemcpyAsync(s.in_a.p, configs, n * q.ops->dim * sizeof(float), cudaMemcpyHostToDevice, s.stream));
    q.args.cfg = s.in_a.as<float>();
    q.args.n = n;
    q.args.spheres = s.out.as<float>();
    VB_CUDA(q.ops->sphere_fk(q.args));
    uint32_t *counts_d = s.margin.as<uint32_t>(), *hits_d = counts_d + total;
    sphere_objects_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, s.stream>>>(q.args.env, s.out.as<float4>(), total, cap,
                                                                                          hits_d, counts_d);
    VB_CUDA(cudaGetLastError());
    ctx->launches += 2;
    VB_CUDA(cudaMemcpyAsync(counts, counts_d, total * sizeof(uint32_t), cudaMemcpyDeviceToHost, s.stream));
    if (cap)
    {
        VB_CUDA(cudaMemcpyAsync(hits, hits_d, total * cap * sizeof(uint32_t), cudaMemcpyDeviceToHost, s.stream));
    }
    VB_CUDA(cudaStreamSynchronize(s.stream));
    return VB_OK;
}

/* ---- PRM / FCIT* batching ------------------------------------------------------ */
// The first k valid configurations of a stream, in stream order -- what the sampling loops of planning/prm.hh:111-124
// and planning/fcit.hh:316-343 keep (sample -> Robot::fkcc -> continue when invalid), for a whole batch at once.
int vb_sample_valid(vb_ctx *ctx, vb_robot robot, const float *base, const vb_env *env, const float *candidates, uint64_t seed,
                    uint64_t first, size_t m, size_t k, float *out_cfgs, uint64_t *out_index, size_t *n_found, size_t *n_consumed)
{
    if (!ctx || !env || !n_found || !n_consumed || (k && (!out_cfgs || !out_index)))
    {
        return fail(VB_ERR_ARG, "vb_sample_valid: null argument");
    }
    *n_found = 0;
    *n_consumed = 0;
    if (k == 0 || m == 0)
    {
        return VB_OK;
    }
    Query q;
    VB_TRY(begin_query(ctx, 0, robot, base, env, q));
    Slot &s = *q.slot;
    const int dim = q.ops->dim;
    const size_t chunk_max = std::min<size_t>(m, std::max<size_t>(4096, std::min<size_t>(size_t(1) << 22, 4 * k)));
    const size_t words_max = (chunk_max + 31) / 32;
    VB_CUDA(s.in_a.reserve(chunk_max * dim * sizeof(float)));
    VB_CUDA(s.bits.reserve(words_max * sizeof(uint32_t)));
    VB_CUDA(s.in_d.reserve((words_max + 1) * sizeof(uint32_t)));   // word popcounts
    VB_CUDA(s.in_d2.reserve((words_max + 1) * sizeof(uint32_t)));  // their exclusive prefix
    VB_CUDA(s.out.reserve(k * dim * sizeof(float)));
    VB_CUDA(s.margin.reserve(k * sizeof(uint64_t) + 16));
    uint64_t *index_d = s.margin.as<uint64_t>();
    unsigned long long *consumed_d = reinterpret_cast<unsigned long long *>(index_d + k);
    size_t tmp_bytes = 0;
    VB_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, s.in_d.as<uint32_t>(), s.in_d2.as<uint32_t>(),
                                          static_cast<int>(words_max + 1), s.stream));
    VB_CUDA(s.scan_tmp.reserve(tmp_bytes));
    const StreamLimits lim = limits_of(q.ops);
    size_t have = 0, done = 0;
    while (have < k && done < m)
    {
        const size_t c = std::min(chunk_max, m - done), words = (c + 31) / 32;
        if (candidates)
        {
            VB_CUDA(cudaMemcpyAsync(s.in_a.p, candidates + done * dim, c * dim * sizeof(float), cudaMemcpyHostToDevice, s.stream));
        }
        else
        {
            random_configs_kernel<<<s.sm_count * 8, 256, 0, s.stream>>>(lim, seed, first + done, c, s.in_a.as<float>());
            VB_CUDA(cudaGetLastError());
            ctx->launches += 1;
        }
        VB_TRY(run_fkcc_d(q, s.in_a.as<float>(), c, s.bits.as<uint32_t>(), nullptr));
        word_popcount_kernel<<<static_cast<unsigned>((words + 1 + 255) / 256), 256, 0, s.stream>>>(s.bits.as<uint32_t>(), c,
                                                                                                   s.in_d.as<uint32_t>());
        VB_CUDA(cudaGetLastError());
        VB_CUDA(cub::DeviceScan::ExclusiveSum(s.scan_tmp.p, tmp_bytes, s.in_d.as<uint32_t>(), s.in_d2.as<uint32_t>(),
                                              static_cast<int>(words + 1), s.stream));
        VB_CUDA(cudaMemsetAsync(consumed_d, 0, sizeof(unsigned long long), s.stream));
        compact_valid_kernel<<<static_cast<unsigned>((c + 255) / 256), 256, 0, s.stream>>>(
            s.bits.as<uint32_t>(), s.in_d2.as<uint32_t>(), s.in_a.as<float>(), dim, c, first + done, have, k, s.out.as<float>(), index_d,
            consumed_d);
        VB_CUDA(cudaGetLastError());
        ctx->launches += 4;
        uint32_t found = 0;
        unsigned long long consumed = 0;
        VB_CUDA(cudaMemcpyAsync(&found, s.in_d2.as<uint32_t>() + words, sizeof(found), cudaMemcpyDeviceToHost, s.stream));
        VB_CUDA(cudaMemcpyAsync(&consumed, consumed_d, sizeof(consumed), cudaMemcpyDeviceToHost, s.stream));
        VB_CUDA(cudaStreamSynchronize(s.stream));
        if (have + found >= k)
        {
            *n_consumed = done + static_cast<size_t>(consumed);
            have = k;
            break;
        }
        have += found;
        done += c;
        *n_consumed = done;
    }
    *n_found = have;
    if (have)
    {
        VB_CUDA(cudaMemcpyAsync(out_cfgs, s.out.p, have * dim * sizeof(float), cudaMemcpyDeviceToHost, s.stream));
        VB_CUDA(cudaMemcpyAsync(out_index, index_d, have * sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
        VB_CUDA(cudaStreamSynchronize(s.stream));
    }
    return VB_OK;
}

// Edges as index pairs into a device-resident configuration table (roadmap vertices): the neighbour edges of
// planning/prm.hh:139 / planning/fcit.hh:230-236 upload 8 bytes each instead of two configurations.
int vb_validate_motion_indexed_d(vb_ctx *ctx, int slot, vb_robot robot, const float *base, const vb_env *env, const float *table_d,
                                 size_t n_rows, const uint32_t *pairs_d, size_t n_edges, int resolution, uint32_t *valid_bits_d)
{
    if (!env || (n_edges && (!table_d || !pairs_d || !valid_bits_d)))
    {
        return fail(VB_ERR_ARG, "vb_validate_motion_indexed_d: null argument");
    }
    Query q;
    VB_TRY(begin_query(ctx, slot, robot, base, env, q));
    q.args.edge_idx = pairs_d;
    q.args.table_rows = n_rows;
    return run_motion_d(q, table_d, table_d, nullptr, false, n_edges, resolution, valid_bits_d, nullptr);
}

int vb_validate_motion_indexed(vb_ctx *ctx, vb_robot robot, const float *base, const vb_env *env, const float *table_d, size_t n_rows,
                               const uint32_t *pairs, size_t n_edges, int resolution, uint32_t *valid_bits)
{
    if (!ctx || !env || (n_edges && (!table_d || !pairs || !valid_bits)))
    {
        return fail(VB_ERR_ARG, "vb_validate_motion_indexed: null argument");
    }
    if (n_edges == 0)
    {
        return VB_OK;
    }
    VB_TRY(check_slot(ctx, 0));
    Slot &s = ctx->slots[0];
    VB_CUDA(cudaSetDevice(s.device));
    const size_t words = (n_edges + 31) / 32;
    VB_CUDA(s.in_d.reserve(n_edges * 2 * sizeof(uint32_t)));
    VB_CUDA(s.bits.reserve(words * sizeof(uint32_t)));
    VB_CUDA(cudaMemcpyAsync(s.in_d.p, pairs, n_edges * 2 * sizeof(uint32_t), cudaMemcpyHostToDevice, s.stream));
    VB_TRY(vb_validate_motion_indexed_d(ctx, 0, robot, base, env, table_d, n_rows, s.in_d.as<uint32_t>(), n_edges, resolution,
                                        s.bits.as<uint32_t>()));
    VB_CUDA(cudaMemcpyAsync(valid_bits, s.bits.p, words * sizeof(uint32_t), cudaMemcpyDeviceToHost, s.stream));
    VB_CUDA(cudaStreamSynchronize(s.stream));
    return VB_OK;
}

int vb_random_edge(vb_robot r, uint64_t seed, uint64_t index, float len_lo, float len_hi, float *a, float *b)
{
    return vb_random_edges(r, seed, index, 1, len_lo, len_hi, a, b);
}

int vb_random_configs(vb_robot r, uint64_t seed, uint64_t first, size_t n, float *q)
{
    const vb::RobotOps *o = ops_of(r);
    if (!o || (n && !q))
    {
        return fail(VB_ERR_ARG, "vb_random_configs: bad arguments");
    }
    for (size_t i = 0; i < n; ++i)
    {
        o->random_config(seed, first + i, q + i * o->dim);
    }
    return VB_OK;
}

int vb_random_edges(vb_robot r, uint64_t seed, uint64_t first, size_t n, float len_lo, float len_hi, float *a, float *b)
{
    const vb::RobotOps *o = ops_of(r);
    if (!o || (n && (!a || !b)))
    {
        return fail(VB_ERR_ARG, "vb_random_edges: bad arguments");
    }
    const StreamLimits lim = limits_of(o);
    for (size_t i = 0; i < n; ++i)
    {
        random_edge_row(lim, seed, first + i, len_lo, len_hi, a + i * o->dim, b + i * o->dim);
    }
    return VB_OK;
}

int vb_random_configs_d(vb_ctx *ctx, int slot, vb_robot r, uint64_t seed, uint64_t first, size_t n, float *q_d)
{
    VB_TRY(check_slot(ctx, slot));
    const vb::RobotOps *o = ops_of(r);
    if (!o || (n && !q_d))
    {
        return fail(VB_ERR_ARG, "vb_random_configs_d: bad arguments");
    }
    Slot &s = ctx->slots[slot];
    VB_CUDA(cudaSetDevice(s.device));
    if (n)
    {
        random_configs_kernel<<<s.sm_count * 8, 256, 0, s.stream>>>(limits_of(o), seed, first, n, q_d);
        VB_CUDA(cudaGetLastError());
        ctx->launches += 1;
    }
    return VB_OK;
}

int vb_random_edges_d(vb_ctx *ctx, int slot, vb_robot r, uint64_t seed, uint64_t first, size_t n, float len_lo, float len_hi,
                      float *a_d, float *b_d)
{
    VB_TRY(check_slot(ctx, slot));
    const vb::RobotOps *o = ops_of(r);
    if (!o || (n && (!a_d || !b_d)))
    {
        return fail(VB_ERR_ARG, "vb_random_edges_d: bad arguments");
    }
    Slot &s = ctx->slots[slot];
    VB_CUDA(cudaSetDevice(s.device));
    if (n)
    {
        random_edges_kernel<<<s.sm_count * 8, 256, 0, s.stream>>>(limits_of(o), seed, first, n, len_lo, len_hi, a_d, b_d);
        VB_CUDA(cudaGetLastError());
        ctx->launches += 1;
    }
    return VB_OK;
}

/* ---- pinned host memory for the host-buffer entry points ------------------------------------- */
int vb_host_alloc(size_t bytes, void **ptr)
{
    if (!ptr)
    {
        return fail(VB_ERR_ARG, "vb_host_alloc: null argument");
    }
    VB_CUDA(cudaHostAlloc(ptr, std::max<size_t>(bytes, 16), cudaHostAllocPortable));
    return VB_OK;
}

int vb_host_free(void *ptr)
{
    VB_CUDA(cudaFreeHost(ptr));
    return VB_OK;
}

int vb_host_register(void *ptr, size_t bytes)
{
    if (!ptr || !bytes)
    {
        return fail(VB_ERR_ARG, "vb_host_register: bad arguments");
    }
    VB_CUDA(cudaHostRegister(ptr, bytes, cudaHostRegisterPortable));
    return VB_OK;
}

int vb_host_unregister(void *ptr)
{
    VB_CUDA(cudaHostUnregister(ptr));
    return VB_OK;
}

/* ---- multi-process gather -------------------------------------------------------------------- */
int vb_comm_unique_id(void *id_bytes_128)
{
    if (!id_bytes_128)
    {
        return fail(VB_ERR_ARG, "null id buffer");
    }
    if (!nccl().ok)
    {
        return fail(VB_ERR_NCCL, "libnccl.so.2 could not be loaded");
    }
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    VB_NCCL(nccl().GetUniqueId(&id));
    std::memcpy(id_bytes_128, &id, sizeof(id));
    return VB_OK;
}

int vb_comm_init(vb_ctx *ctx, int n_ranks, int rank, const void *id_bytes_128)
{
    if (!ctx || !id_bytes_128 || n_ranks < 1 || rank < 0 || rank >= n_ranks)
    {
        return fail(VB_ERR_ARG, "vb_comm_init: bad arguments");
    }
    if (!nccl().ok)
    {
        return fail(VB_ERR_NCCL, "libnccl.so.2 could not be loaded");
    }
    if (ctx->comm)
    {
        return fail(VB_ERR_STATE, "communicator already initialised");
    }
    ncclUniqueId id;
    std::memcpy(&id, id_bytes_128, sizeof(id));
    VB_CUDA(cudaSetDevice(ctx->slots[0].device));
    VB_NCCL(nccl().CommInitRank(&ctx->comm, n_ranks, id, rank));
    ctx->comm_ranks = n_ranks;
    return VB_OK;
}

int vb_comm_allgather_bits_d(vb_ctx *ctx, const uint32_t *send_d, uint32_t *recv_d, size_t n_words)
{
    if (!ctx || !ctx->comm)
    {
        return fail(VB_ERR_STATE, "vb_comm_init has not been called");
    }
    VB_CUDA(cudaSetDevice(ctx->slots[0].device));
    VB_NCCL(nccl().AllGather(send_d, recv_d, n_words, ncclUint32, ctx->comm, ctx->slots[0].stream));
    return VB_OK;
}

int vb_comm_destroy(vb_ctx *ctx)
{
    if (ctx && ctx->comm)
    {
        VB_NCCL(nccl().CommDestroy(ctx->comm));
        ctx->comm = nullptr;
    }
    return VB_OK;
}

}  // extern "C"

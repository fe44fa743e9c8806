// Kernel templates of the fkcc path, instantiated once per robot trace (see codegen/emit_cuda.py).
//
// One thread evaluates one configuration: the whole FK chain and every sphere centre of that
// configuration live in that thread's registers (the trace is straight-line code), obstacles are
// read as warp-uniform broadcasts from shared memory, and a warp ballot packs 32 verdicts into one
// uint32 store.  Replaces Robot::fkcc / Robot::sphere_fk (robots/panda_base.hh:53-71 and the
// per-robot equivalents) and planning::validate_motion (planning/validate.hh:23-75).
#pragma once
#include "vb_env.cuh"

namespace vb
{
    constexpr int kBlock = 128;           // threads per CTA
    constexpr int kMaxSharedQuads = 3072; // obstacle records staged in shared memory (48 KB)

    struct LaunchArgs
    {
        EnvView env;
        float base[3];         // robot base offset in metres (Panda grid variants; 0 otherwise)
        const float *cfg;      // [n][dim] row-major, device
        const float *cfg_b;    // edges: goals [n][dim]
        size_t n;              // configurations or edges
        uint64_t seed;         // device-side input synthesis
        uint32_t *bits;        // verdict words, bit = 1 => valid
        float *margin;         // optional per-item min |clearance| (metres)
        float *spheres;        // sphere_fk output [n][S][4]
        unsigned long long *counters;  // optional work counters
        const uint32_t *prefix; // edges: exclusive prefix of interpolant counts [n+1]
        size_t n_points;        // edges: total interpolants
        int resolution;
        int attach;            // use the attachment trace
        float sphere_radius;   // robots::sphere global radius (robots/sphere/fk.hh:16)
        float edge_len[2];     // synthetic edges: length range
        cudaStream_t stream;
    };

    // ---- policy handed to the generated traces -------------------------------------------------
    template <class M>
    struct CheckPolicy
    {
        const EnvView &env;
        const float4 *prims;
        float bx_, by_, bz_;
        float radius;  // robots::sphere
        M mg;
        float4 posed[kMaxAttach];

        VB_D CheckPolicy(const EnvView &e, const float4 *p, const float *base, float rad)
          : env(e), prims(p), bx_(base[0]), by_(base[1]), bz_(base[2]), radius(rad)
        {
        }
        VB_D float bx() const { return bx_; }
        VB_D float by() const { return by_; }
        VB_D float bz() const { return bz_; }
        // `literal + base_*` is evaluated in double and narrowed, as the C++ expression is
        VB_D float cbx(double c) const { return static_cast<float>(c + static_cast<double>(bx_)); }
        VB_D float cby(double c) const { return static_cast<float>(c + static_cast<double>(by_)); }
        VB_D float cbz(double c) const { return static_cast<float>(c + static_cast<double>(bz_)); }
        VB_D void ops(int) {}

        VB_D bool env_test(float x, float y, float z, float r) { return env_hit(env, prims, x, y, z, r, mg); }
        // collision/validity.hh:13-44 + sphere_sphere.hh:9-23
        VB_D bool self_test(float ax, float ay, float az, float ar, float bx, float by, float bz, float br)
        {
            const float dx = ax - bx, dy = ay - by, dz = az - bz, rs = ar + br;
            const float v = dx * dx + dy * dy + dz * dz - rs * rs;
            mg.sq(v, rs);
            return v < 0.f;
        }
        // collision/attachments.hh:75-122
        VB_D void attach_pose(float p_tx, float p_ty, float p_tz, float p_rx, float p_ry, float p_rz, float p_rw)
        {
            const float *t = env.attach.tf;
            const float tf_tx = t[0], tf_ty = t[1], tf_tz = t[2], tf_rx = t[3], tf_ry = t[4], tf_rz = t[5], tf_rw = t[6];
            const float rx = p_rw * tf_rx + p_rx * tf_rw + p_ry * tf_rz - p_rz * tf_ry;
            const float ry = p_rw * tf_ry - p_rx * tf_rz + p_ry * tf_rw + p_rz * tf_rx;
            const float rz = p_rw * tf_rz + p_rx * tf_ry - p_ry * tf_rx + p_rz * tf_rw;
            const float rw = p_rw * tf_rw - p_rx * tf_rx - p_ry * tf_ry - p_rz * tf_rz;
            const float x0 = p_ry * tf_tz - p_rz * tf_ty;
            const float x1 = p_rx * tf_ty - p_ry * tf_tx;
            const float x2 = p_rx * tf_tz - p_rz * tf_tx;
            const float tx = p_tx + 2.0f * (p_rw * x0 + p_ry * x1 + p_rz * x2) + tf_tx;
            const float ty = p_ty + 2.0f * (-p_rw * x2 - p_rx * x1 + p_rz * x0) + tf_ty;
            const float tz = p_tz + 2.0f * (p_rw * x1 - p_rx * x2 - p_ry * x0) + tf_tz;
            const float b0 = ry * ry, b1 = rz * rz, b2 = rw * rz, b3 = rw * ry, b4 = rx * rx, b5 = rw * rx,
                        b6 = rx * ry, b7 = rx * rz, b8 = ry * rz;
            const float xx = -2.0f * (b0 + b1) + 1.0f, xy = 2.0f * (b6 + b2), xz = 2.0f * (b7 - b3);
            const float yx = 2.0f * (b6 - b2), yy = -2.0f * (b1 + b4) + 1.0f, yz = 2.0f * (b8 + b5);
            const float zx = 2.0f * (b7 + b3), zy = 2.0f * (b8 - b5), zz = -2.0f * (b0 + b4) + 1.0f;
#pragma unroll
            for (int i = 0; i < kMaxAttach; ++i)
            {
                if (i < env.attach.n)
                {
                    const float4 s = env.attach.local[i];
                    posed[i] = make_float4(
                        s.x * xx + s.y * yx + s.z * zx + tx,
                        s.x * xy + s.y * yy + s.z * zy + ty,
                        s.x * xz + s.y * yz + s.z * zz + tz,
                        s.w);
                }
            }
        }
        // validity.hh:270-293
        VB_D bool attach_self(float x, float y, float z, float r)
        {
            bool hit = false;
#pragma unroll
            for (int i = 0; i < kMaxAttach; ++i)
            {
                if (i < env.attach.n)
                {
                    hit = hit || self_test(x, y, z, r, posed[i].x, posed[i].y, posed[i].z, posed[i].w);
                }
            }
            return hit;
        }
        // validity.hh:251-268
        VB_D bool attach_env()
        {
            bool hit = false;
#pragma unroll
            for (int i = 0; i < kMaxAttach; ++i)
            {
                if (i < env.attach.n && !hit)
                {
                    hit = env_test(posed[i].x, posed[i].y, posed[i].z, posed[i].w);
                }
            }
            return hit;
        }
    };

    template <int S>
    struct SpherePolicy
    {
        float bx_, by_, bz_, radius;
        float *out;  // this configuration's [S][4]
        VB_D float bx() const { return bx_; }
        VB_D float by() const { return by_; }
        VB_D float bz() const { return bz_; }
        VB_D float cbx(double c) const { return static_cast<float>(c + static_cast<double>(bx_)); }
        VB_D float cby(double c) const { return static_cast<float>(c + static_cast<double>(by_)); }
        VB_D float cbz(double c) const { return static_cast<float>(c + static_cast<double>(bz_)); }
        VB_D void ops(int) {}
        VB_D void out_x(int i, float v) { out[4 * i + 0] = v; }
        VB_D void out_y(int i, float v) { out[4 * i + 1] = v; }
        VB_D void out_z(int i, float v) { out[4 * i + 2] = v; }
        VB_D void out_r(int i, float v) { out[4 * i + 3] = v; }
    };

    // ---- shared-memory staging of the obstacle records -----------------------------------------
    VB_D const float4 *stage_env(const EnvView &env, float4 *smem)
    {
        if (env.n_quads > kMaxSharedQuads)
        {
            return env.prims;  // large scenes: L1/L2-resident broadcasts
        }
        for (int i = threadIdx.x; i < env.n_quads; i += blockDim.x)
        {
            smem[i] = env.prims[i];
        }
        __syncthreads();
        return smem;
    }

    template <class Robot, class M>
    VB_D bool check_config(const LaunchArgs &a, const float4 *prims, const float *q, bool attach, M &mg_out)
    {
        CheckPolicy<M> pol(a.env, prims, a.base, a.sphere_radius);
        bool ok;
        if (Robot::kHasAttach && attach)
        {
            ok = Robot::fkcc_attach(pol, q);
        }
        else
        {
            ok = Robot::fkcc(pol, q);
        }
        mg_out = pol.mg;
        return ok;
    }

    // ---- K1: batched fkcc over explicit configurations -----------------------------------------
    template <class Robot, bool kMargin>
    __global__ void __launch_bounds__(kBlock) fkcc_kernel(const LaunchArgs a)
    {
        extern __shared__ float4 smem[];
        const float4 *prims = stage_env(a.env, smem);
        const size_t stride = static_cast<size_t>(gridDim.x) * kBlock;
        const size_t n_round = (a.n + 31) / 32 * 32;
        for (size_t i = static_cast<size_t>(blockIdx.x) * kBlock + threadIdx.x; i < n_round; i += stride)
        {
            bool ok = false;
            if (i < a.n)
            {
                float q[Robot::kDim];
#pragma unroll
                for (int j = 0; j < Robot::kDim; ++j)
                {
                    q[j] = __ldg(a.cfg + i * Robot::kDim + j);
                }
                if constexpr (kMargin)
                {
                    MinMargin mg;
                    ok = check_config<Robot>(a, prims, q, a.attach != 0, mg);
                    a.margin[i] = mg.m;
                }
                else
                {
                    NoMargin mg;
                    ok = check_config<Robot>(a, prims, q, a.attach != 0, mg);
                }
            }
            const uint32_t word = __ballot_sync(0xffffffffu, ok);
            if ((threadIdx.x & 31) == 0)
            {
                a.bits[i >> 5] = word;
            }
        }
    }

    // configuration i of the synthetic stream: uniform in the joint ranges (scale_configuration)
    template <class Robot>
    VB_HD void random_config(uint64_t seed, uint64_t i, float *q)
    {
#pragma unroll
        for (int j = 0; j < Robot::kDim; ++j)
        {
            q[j] = Robot::lo(j) + u01(seed, i, j) * Robot::range(j);
        }
    }

    // ---- K1r: fkcc over device-synthesised configurations (0 input bytes) -----------------------
    template <class Robot>
    __global__ void __launch_bounds__(kBlock) fkcc_random_kernel(const LaunchArgs a, size_t first)
    {
        extern __shared__ float4 smem[];
        const float4 *prims = stage_env(a.env, smem);
        const size_t stride = static_cast<size_t>(gridDim.x) * kBlock;
        const size_t n_round = (a.n + 31) / 32 * 32;
        for (size_t i = static_cast<size_t>(blockIdx.x) * kBlock + threadIdx.x; i < n_round; i += stride)
        {
            bool ok = false;
            if (i < a.n)
            {
                float q[Robot::kDim];
                random_config<Robot>(a.seed, first + i, q);
                NoMargin mg;
                ok = check_config<Robot>(a, prims, q, a.attach != 0, mg);
            }
            const uint32_t word = __ballot_sync(0xffffffffu, ok);
            if ((threadIdx.x & 31) == 0)
            {
                a.bits[i >> 5] = word;
            }
        }
    }

    // ---- K2: sphere_fk ---------------------------------------------------------------------------
    template <class Robot>
    __global__ void __launch_bounds__(kBlock) sphere_fk_kernel(const LaunchArgs a)
    {
        const size_t i = static_cast<size_t>(blockIdx.x) * kBlock + threadIdx.x;
        if (i >= a.n)
        {
            return;
        }
        float q[Robot::kDim];
#pragma unroll
        for (int j = 0; j < Robot::kDim; ++j)
        {
            q[j] = __ldg(a.cfg + i * Robot::kDim + j);
        }
        SpherePolicy<Robot::kSpheres> pol{a.base[0], a.base[1], a.base[2], a.sphere_radius,
                                          a.spheres + i * Robot::kSpheres * 4};
        Robot::sphere_fk(pol, q);
    }

    // ---- K3: discretised motion validation --------------------------------------------------------
    // planning/validate.hh:23-75: an edge a->b is valid iff fkcc holds at a + m/(8n) (b-a), m = 1..8n,
    // n = max(ceil(|b-a| / 8 * resolution), 1).  Pass i of the reference's rake checks lanes
    // k = 0..7 at start + v * (k+1)/8 - i * backstep with backstep = v / (8n) subtracted repeatedly;
    // the same recurrence is used here so interpolants are bit-identical.  Work is flattened: thread t
    // handles interpolant t of the whole batch (prefix[e] <= t < prefix[e+1] selects the edge), so a
    // long edge simply occupies more threads; verdicts are AND-reduced with atomicAnd.
    constexpr int kRake = 8;

    template <int D>
    VB_HD uint32_t edge_passes(const float *qa, const float *qb, int resolution)
    {
        // squared_l2_norm (vector/interface.hh:407-410): rows are added first, then the 8 lanes are
        // folded as ((s0+s4)+(s2+s6)) + ((s1+s5)+(s3+s7)) (vector/avx.hh:442-452)
        float s[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        for (int j = 0; j < D; ++j)
        {
            const float v = VB_ADD(qb[j], -qa[j]);
            s[j & 7] = (j < 8) ? VB_MUL(v, v) : VB_ADD(s[j & 7], VB_MUL(v, v));
        }
        const float ss = VB_ADD(VB_ADD(VB_ADD(s[0], s[4]), VB_ADD(s[2], s[6])), VB_ADD(VB_ADD(s[1], s[5]), VB_ADD(s[3], s[7])));
        const float dist = sqrtf(ss);
        const float nf = fmaxf(ceilf(dist / static_cast<float>(kRake) * static_cast<float>(resolution)), 1.f);
        return static_cast<uint32_t>(nf);
    }

    template <class Robot>
    __global__ void edge_count_kernel(const LaunchArgs a, uint32_t *counts)
    {
        const size_t e = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
        if (e >= a.n)
        {
            return;
        }
        float qa[Robot::kDim], qb[Robot::kDim];
        for (int j = 0; j < Robot::kDim; ++j)
        {
            qa[j] = a.cfg[e * Robot::kDim + j];
            qb[j] = a.cfg_b[e * Robot::kDim + j];
        }
        counts[e] = kRake * edge_passes<Robot::kDim>(qa, qb, a.resolution);
    }

    template <class Robot>
    VB_D void edge_point(const float *qa, const float *qb, uint32_t n_pass, uint32_t pass, uint32_t lane, float *q)
    {
        const float pct = static_cast<float>(lane + 1) / static_cast<float>(kRake);
        const float denom = static_cast<float>(kRake * n_pass);
#pragma unroll
        for (int j = 0; j < Robot::kDim; ++j)
        {
            const float v = qb[j] - qa[j];
            float x = fmaf(v, pct, qa[j]);
            const float back = v / denom;
            for (uint32_t s = 0; s < pass; ++s)
            {
                x = VB_ADD(x, -back);
            }
            q[j] = x;
        }
    }

    template <class Robot, bool kMargin>
    __global__ void __launch_bounds__(kBlock) motion_kernel(const LaunchArgs a)
    {
        extern __shared__ float4 smem[];
        const float4 *prims = stage_env(a.env, smem);
        const size_t stride = static_cast<size_t>(gridDim.x) * kBlock;
        for (size_t t = static_cast<size_t>(blockIdx.x) * kBlock + threadIdx.x; t < a.n_points; t += stride)
        {
            // binary search: last edge with prefix[e] <= t
            size_t lo = 0, hi = a.n;
            while (hi - lo > 1)
            {
                const size_t mid = (lo + hi) >> 1;
                if (__ldg(a.prefix + mid) <= t)
                {
                    lo = mid;
                }
                else
                {
                    hi = mid;
                }
            }
            const size_t e = lo;
            const uint32_t begin = __ldg(a.prefix + e);
            const uint32_t count = __ldg(a.prefix + e + 1) - begin;
            const uint32_t local = static_cast<uint32_t>(t - begin);
            const uint32_t pass = local / kRake, lane = local % kRake;
            if (!kMargin && !((*reinterpret_cast<volatile const uint32_t *>(a.bits + (e >> 5)) >> (e & 31)) & 1u))
            {
                continue;  // edge already refuted by another interpolant
            }
            // the attachment trace is only used for the first pass (validate.hh:43-44 vs :58)
            const bool attach = a.attach && pass == 0;
            float qa[Robot::kDim], qb[Robot::kDim], q[Robot::kDim];
#pragma unroll
            for (int j = 0; j < Robot::kDim; ++j)
            {
                qa[j] = __ldg(a.cfg + e * Robot::kDim + j);
                qb[j] = __ldg(a.cfg_b + e * Robot::kDim + j);
            }
            edge_point<Robot>(qa, qb, count / kRake, pass, lane, q);
            bool ok;
            if constexpr (kMargin)
            {
                MinMargin mg;
                ok = check_config<Robot>(a, prims, q, attach, mg);
                atomicMin(reinterpret_cast<int *>(a.margin + e), __float_as_int(mg.m));
            }
            else
            {
                NoMargin mg;
                ok = check_config<Robot>(a, prims, q, attach, mg);
            }
            if (!ok)
            {
                atomicAnd(a.bits + (e >> 5), ~(1u << (e & 31)));
            }
        }
    }
}  // namespace vb

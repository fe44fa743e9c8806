// Numerics contract of the fkcc path (host + device).
//
// The reference evaluates joint half-angle sines/cosines with a low-order odd polynomial, not
// libm (vector/interface.hh:438-469).  Sphere centres only agree with the reference to 1e-5 m if
// the same polynomial, constants and the same "cos(x) = sin(wrap(x + pi/2))" folding are used, so
// they are restated here in scalar form.  Explicit *_rn intrinsics keep nvcc from re-associating
// or contracting the polynomial itself; the FK chain around it is free to use FMA.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#ifdef __CUDA_ARCH__
#define VB_MUL(a, b) __fmul_rn((a), (b))
#define VB_ADD(a, b) __fadd_rn((a), (b))
#else
#define VB_MUL(a, b) ((a) * (b))
#define VB_ADD(a, b) ((a) + (b))
#endif

#define VB_HD __host__ __device__ __forceinline__
#define VB_D __device__ __forceinline__

namespace vb
{
    // vector/interface.hh:438-456
    VB_HD float vamp_sin(float v)
    {
        const float v_sq = VB_MUL(v, fabsf(v));
        const float s_vsq = VB_MUL(-0.478637850138f, v_sq);
        const float s_v = VB_MUL(1.503684069359f, v);
        const float p = VB_ADD(s_vsq, s_v);
        const float abs_p = fabsf(p);
        const float p_2 = VB_MUL(p, abs_p);
        const float p_3 = VB_MUL(p_2, abs_p);
        const float d_3 = VB_MUL(0.011596870476f, p_3);
        const float d_2 = VB_MUL(0.140024078368f, p_2);
        const float d_1 = VB_MUL(0.665200679751f, p);
        return VB_ADD(VB_ADD(d_3, d_2), d_1);
    }

    // vector/interface.hh:458-469 : PI is the float 3.14159265359f, PI/2. and 2*PI are formed in
    // double resp. float and narrowed to float before use.
    VB_HD float vamp_cos(float v)
    {
        constexpr float PI = 3.14159265359f;
        constexpr float HALF_PI = static_cast<float>(PI / 2.);
        constexpr float TWO_PI = static_cast<float>(2 * PI);
        const float w = VB_ADD(v, HALF_PI);
        const float wrapped = (w >= PI) ? VB_ADD(w, -TWO_PI) : w;
        return vamp_sin(wrapped);
    }

    // Counter-based generator shared by the device-side input synthesis (vb_fkcc_random,
    // vb_validate_motion_random) and by the host so that a CPU checker can regenerate sample i.
    VB_HD uint64_t splitmix64(uint64_t x)
    {
        x += 0x9E3779B97F4A7C15ull;
        x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
        x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
        return x ^ (x >> 31);
    }

    // uniform in [0, 1) with 24 random bits
    VB_HD float u01(uint64_t seed, uint64_t index, uint32_t stream)
    {
        const uint64_t h = splitmix64(seed ^ splitmix64(index * 64ull + stream));
        return static_cast<float>(static_cast<uint32_t>(h >> 40)) * (1.0f / 16777216.0f);
    }
}  // namespace vb

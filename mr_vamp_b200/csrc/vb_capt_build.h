// Host-side view of a CAPT's query buffers (collision/capt.hh:588-621) and the device builder (vb_capt_build.cu).
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>

#include <cuda_runtime.h>

namespace vb
{
    struct CaptBuffers
    {
        std::vector<float> tests, aabbs;
        std::vector<uint32_t> aff_starts;
        std::vector<float> aff;  // interleaved per 8-point vector: x[8] y[8] z[8]
        float top[6];
        float r_point = 0.f, r_min = 0.f, r_max = 0.f;
        int nlog2 = 0;
    };

    // collision::CAPT(points, r_min, r_max, r_point) (collision/capt.hh:326-398) built on the current device; `what`
    // names the problem when the return value is cudaErrorInvalidValue
    cudaError_t build_capt(cudaStream_t stream, const float *xyz, size_t n_points, float r_min, float r_max, float r_point,
                           CaptBuffers &out, const char **what);
}  // namespace vb

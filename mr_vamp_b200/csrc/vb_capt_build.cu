// CAPT construction on the device: builds the query buffers of collision::CAPT
// (collision/capt.hh:137-398: tests, aff_starts, affordances, aabbs, aabb_top) from a point cloud, so that a host can
// hand the library raw points instead of a tree built elsewhere.
//
// The reference builds recursively: a median k-d split per level (dimension cycling x, y, z; capt.hh:137-153,225-232)
// and, per cell, the list of points that could collide with a sphere centred in it ("affordances"), propagated down
// the recursion (capt.hh:234-283) and cut by the final distance test at the leaf (capt.hh:181-188).  Here:
//
//   1. one stable radix sort per level orders the points inside every node by the split coordinate (key = node index |
//      coordinate); the split value of node j is the mean of the two middle coordinates, exactly as median_partition;
//   2. the propagation rule is evaluated per (leaf, point) pair in closed form.  A point q enters the candidate list of
//      leaf z at the level where their paths part: when z lies in the low half, q (in the high half) enters iff
//      q[d] <= test + r_max; when z lies in the high half, the reference scans the low half FROM ITS SMALLEST element
//      while p[d] >= test - r_max (capt.hh:262-267), so q enters iff the whole low half satisfies the bound -- reproduced
//      as it is, because it decides what a query can see.  Below that level q must pass z's own splits
//      (q[d] <= test + r_max on the low side, >= test - r_max on the high side, capt.hh:246-257), and at the leaf the
//      distance test volume.affords(q, (r_max + r_point)^2), unless the cell lies inside the ball of radius
//      r_min + r_point around its point (capt.hh:181);
//   3. candidates come from a uniform grid over the points (cell = r_max + r_point), one thread per leaf: a counting
//      pass, an exclusive scan of the vector counts, a filling pass.
//
// The resulting buffers answer every query as the reference's do: same split values, same affordance SET per leaf (the
// order inside a leaf and the placement of equal coordinates around a median are free in the reference too: pdqsort is
// not stable), same leaf boxes.  tests/test_capt_build_gpu.py compares them with the reference's buffers.
#include "vb_capt_build.h"

#include <cmath>
#include <cstdint>
#include <limits>

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

namespace
{
#define CB_CUDA(expr)              \
    do                             \
    {                              \
        cudaError_t e__ = (expr);  \
        if (e__ != cudaSuccess)    \
        {                          \
            return e__;            \
        }                          \
    } while (0)

    struct Buf
    {
        void *p = nullptr;
        cudaError_t alloc(size_t bytes) { return cudaMalloc(&p, bytes < 16 ? 16 : bytes); }
        ~Buf()
        {
            if (p)
            {
                cudaFree(p);
            }
        }
        template <class T>
        T *as() const
        {
            return static_cast<T *>(p);
        }
    };

    __device__ __forceinline__ uint32_t sortable(float f)
    {
        const uint32_t b = __float_as_uint(f);
        return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
    }

    __global__ void level_keys_kernel(const float *coord, const uint32_t *idx, uint32_t n, uint32_t seg_shift, uint64_t *keys)
    {
        const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i < n)
        {
            keys[i] = (static_cast<uint64_t>(i >> seg_shift) << 32) | sortable(coord[idx[i]]);
        }
    }

    // node j of this level: split value and the smallest coordinate of its low half (capt.hh:146-153, 262-267)
    __global__ void level_tests_kernel(const float *coord, const uint32_t *idx, uint32_t nodes, uint32_t seg, uint32_t first_node,
                                       float *tests, float *lo_first)
    {
        const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
        if (j < nodes)
        {
            const uint32_t begin = j * seg, mid = begin + seg / 2;
            tests[first_node + j] = __fadd_rn(coord[idx[mid - 1]], coord[idx[mid]]) * 0.5f;
            lo_first[first_node + j] = coord[idx[begin]];
        }
    }

    struct BuildView
    {
        const float *x, *y, *z;  // points (padded with +inf to n = 2^nlog2)
        const uint32_t *leaf_point;  // leaf -> point index
        const uint32_t *point_leaf;  // point index -> leaf
        const float *tests, *lo_first;
        const uint32_t *cell_start;  // uniform grid over the finite points: [cells + 1]
        const uint32_t *cell_points;
        float g_o[3], g_inv_h;
        int g_n[3];
        uint32_t n, nlog2;
        float r_max, max_l1, max_l2, min_l2;
    };

    __device__ __forceinline__ float coord_of(const BuildView &b, uint32_t p, int d)
    {
        return d == 0 ? b.x[p] : (d == 1 ? b.y[p] : b.z[p]);
    }

    // the cell of leaf z from its ancestors' split values (capt.hh:233-237)
    __device__ void leaf_volume(const BuildView &b, uint32_t z, float *lo, float *hi)
    {
        for (int k = 0; k < 3; ++k)
        {
            lo[k] = -INFINITY;
            hi[k] = INFINITY;
        }
        for (uint32_t L = 0; L < b.nlog2; ++L)
        {
            const uint32_t node = ((1u << L) - 1u) + (z >> (b.nlog2 - L));
            const float t = b.tests[node];
            const int d = static_cast<int>(L % 3u);
            if ((z >> (b.nlog2 - L - 1u)) & 1u)
            {
                lo[d] = t;
            }
            else
            {
                hi[d] = t;
            }
        }
    }

    // does point q belong to the affordance list of leaf z (q != the leaf's own point)?
    __device__ bool affords(const BuildView &b, uint32_t z, const float *lo, const float *hi, uint32_t q)
    {
        const float qc[3] = {b.x[q], b.y[q], b.z[q]};
        // Volume::affords / distsq_to (capt.hh:79-91): the reference's LAST test, evaluated first because it rejects most
        // candidates without walking the tree
        {
            float dsq = 0.f;
            for (int k = 0; k < 3; ++k)
            {
                const float c = fminf(fmaxf(qc[k], lo[k]), hi[k]);
                const float dd = __fadd_rn(qc[k], -c);
                dsq = __fadd_rn(dsq, __fmul_rn(dd, dd));
            }
            if (!(dsq <= b.max_l2))
            {
                return false;
            }
        }
        // the level where the two paths part
        const uint32_t zq = b.point_leaf[q];
        const uint32_t h = 31u - __clz(z ^ zq);  // highest differing bit
        const uint32_t Ls = b.nlog2 - 1u - h;
        {
            const uint32_t node = ((1u << Ls) - 1u) + (z >> (h + 1u));
            const float t = b.tests[node];
            const int d = static_cast<int>(Ls % 3u);
            if (!isfinite(qc[d]))
            {
                return false;
            }
            if ((z >> h) & 1u)
            {
                // z in the high half: the reference's scan of the low half starts at its smallest element
                const float first = b.lo_first[node];
                if (!(first >= __fadd_rn(t, -b.r_max)) || !isfinite(first))
                {
                    return false;
                }
            }
            else if (!(qc[d] <= __fadd_rn(t, b.r_max)))
            {
                return false;
            }
        }
        for (uint32_t L = Ls + 1u; L < b.nlog2; ++L)
        {
            const uint32_t node = ((1u << L) - 1u) + (z >> (b.nlog2 - L));
            const float t = b.tests[node];
            const int d = static_cast<int>(L % 3u);
            if ((z >> (b.nlog2 - L - 1u)) & 1u)
            {
                if (!(qc[d] >= __fadd_rn(t, -b.r_max)))
                {
                    return false;
                }
            }
            else if (!(qc[d] <= __fadd_rn(t, b.r_max)))
            {
                return false;
            }
        }
        return true;
    }

    // kFill == false: count the affordances of every leaf (vector counts out); true: write them
    template <bool kFill>
    __global__ void leaf_affordances_kernel(const BuildView b, uint32_t *n_vec, const uint32_t *aff_starts, float *aff, float *aabbs)
    {
        const uint32_t z = blockIdx.x * blockDim.x + threadIdx.x;
        if (z >= b.n)
        {
            return;
        }
        const uint32_t rep = b.leaf_point[z];
        const float rc[3] = {b.x[rep], b.y[rep], b.z[rep]};
        float bb_lo[3] = {rc[0], rc[1], rc[2]}, bb_hi[3] = {rc[0], rc[1], rc[2]};
        if (!isfinite(rc[0]))
        {
            if (kFill)
            {
                for (int k = 0; k < 3; ++k)
                {
                    aabbs[6 * z + k] = rc[k];
                    aabbs[6 * z + 3 + k] = rc[k];
                }
            }
            else
            {
                n_vec[z] = 0u;
            }
            return;
        }
        float lo[3], hi[3];
        leaf_volume(b, z, lo, hi);
        uint32_t count = 1u;  // the cell's own point comes first
        float *out = kFill ? aff + 24ull * aff_starts[z] : nullptr;
        if (kFill)
        {
            out[0] = rc[0];
            out[8] = rc[1];
            out[16] = rc[2];
        }
        // contained_by_internal_ball (capt.hh:70-77)
        const float d0 = fmaxf(rc[0] - lo[0], hi[0] - rc[0]), d1 = fmaxf(rc[1] - lo[1], hi[1] - rc[1]), d2 = fmaxf(rc[2] - lo[2], hi[2] - rc[2]);
        const bool inside_ball = (d0 * d0 + d1 * d1 + d2 * d2) <= b.min_l2;
        if (!inside_ball)
        {
            int c_lo[3], c_hi[3];
            for (int k = 0; k < 3; ++k)
            {
                const float a = (fmaxf(lo[k], b.g_o[k]) - b.max_l1 - b.g_o[k]) * b.g_inv_h;
                const float e = (fminf(hi[k], b.g_o[k] + static_cast<float>(b.g_n[k]) / b.g_inv_h) + b.max_l1 - b.g_o[k]) * b.g_inv_h;
                c_lo[k] = max(0, static_cast<int>(floorf(a)));
                c_hi[k] = min(b.g_n[k] - 1, static_cast<int>(floorf(e)));
            }
            for (int iz = c_lo[2]; iz <= c_hi[2]; ++iz)
            {
                for (int iy = c_lo[1]; iy <= c_hi[1]; ++iy)
                {
                    const uint32_t row = (static_cast<uint32_t>(iz) * b.g_n[1] + iy) * b.g_n[0];
                    const uint32_t s = b.cell_start[row + c_lo[0]], e = b.cell_start[row + c_hi[0] + 1];
                    for (uint32_t i = s; i < e; ++i)
                    {
                        const uint32_t q = b.cell_points[i];
                        if (q == rep || !affords(b, z, lo, hi, q))
                        {
                            continue;
                        }
                        if (kFill)
                        {
                            const uint32_t v = count >> 3, j = count & 7u;
                            out[24u * v + j] = b.x[q];
                            out[24u * v + 8u + j] = b.y[q];
                            out[24u * v + 16u + j] = b.z[q];
                            bb_lo[0] = fminf(bb_lo[0], b.x[q]), bb_hi[0] = fmaxf(bb_hi[0], b.x[q]);
                            bb_lo[1] = fminf(bb_lo[1], b.y[q]), bb_hi[1] = fmaxf(bb_hi[1], b.y[q]);
                            bb_lo[2] = fminf(bb_lo[2], b.z[q]), bb_hi[2] = fmaxf(bb_hi[2], b.z[q]);
                        }
                        ++count;
                    }
                }
            }
        }
        if (kFill)
        {
            // "ensure that extra garbage does not result in a false collision" (capt.hh:208-214)
            for (uint32_t c = count; (c & 7u) != 0u; ++c)
            {
                const uint32_t v = c >> 3, j = c & 7u;
                out[24u * v + j] = INFINITY;
                out[24u * v + 8u + j] = INFINITY;
                out[24u * v + 16u + j] = INFINITY;
            }
            for (int k = 0; k < 3; ++k)
            {
                aabbs[6 * z + k] = bb_lo[k];
                aabbs[6 * z + 3 + k] = bb_hi[k];
            }
        }
        else
        {
            n_vec[z] = (count + 7u) / 8u;
        }
    }

    __global__ void invert_kernel(const uint32_t *leaf_point, uint32_t n, uint32_t *point_leaf)
    {
        const uint32_t z = blockIdx.x * blockDim.x + threadIdx.x;
        if (z < n)
        {
            point_leaf[leaf_point[z]] = z;
        }
    }

    __global__ void iota_kernel(uint32_t *p, uint32_t n)
    {
        const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i < n)
        {
            p[i] = i;
        }
    }

    __global__ void grid_cell_kernel(const float *x, const float *y, const float *z, uint32_t n_real, const float *o, float inv_h,
                                     const int *gn, uint32_t *cell_of, uint32_t *ids)
    {
        const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i < n_real)
        {
            const int cx = min(gn[0] - 1, max(0, static_cast<int>(floorf((x[i] - o[0]) * inv_h))));
            const int cy = min(gn[1] - 1, max(0, static_cast<int>(floorf((y[i] - o[1]) * inv_h))));
            const int cz = min(gn[2] - 1, max(0, static_cast<int>(floorf((z[i] - o[2]) * inv_h))));
            cell_of[i] = (static_cast<uint32_t>(cz) * gn[1] + cy) * gn[0] + cx;
            ids[i] = i;
        }
    }

    // cell_start[c] = first position (in the cell-sorted order) whose cell is >= c
    __global__ void cell_start_kernel(const uint32_t *sorted_cells, uint32_t n_real, uint32_t n_cells, uint32_t *cell_start)
    {
        const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i > n_real)
        {
            return;
        }
        const uint32_t here = i < n_real ? sorted_cells[i] : n_cells, prev = i > 0 ? sorted_cells[i - 1] + 1 : 0;
        for (uint32_t c = prev; c <= here; ++c)
        {
            cell_start[c] = i;
        }
    }
}  // namespace

namespace vb
{
    cudaError_t build_capt(cudaStream_t stream, const float *xyz, size_t n_points, float r_min, float r_max, float r_point,
                           CaptBuffers &out, const char **what)
    {
        *what = "";
        if (n_points == 0 || n_points > (size_t(1) << 26))
        {
            *what = "point count must be in [1, 2^26]";
            return cudaErrorInvalidValue;
        }
        uint32_t nlog2 = 0;
        while ((size_t(1) << nlog2) < n_points)
        {
            ++nlog2;
        }
        const uint32_t n = 1u << nlog2, n_real = static_cast<uint32_t>(n_points);
        const float inf = std::numeric_limits<float>::infinity();
        // host SoA + padding, bounding box of the cloud
        std::vector<float> hx(n, inf), hy(n, inf), hz(n, inf);
        float lo[3] = {inf, inf, inf}, hi[3] = {-inf, -inf, -inf};
        for (uint32_t i = 0; i < n_real; ++i)
        {
            const float p[3] = {xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]};
            if (!std::isfinite(p[0]) || !std::isfinite(p[1]) || !std::isfinite(p[2]))
            {
                *what = "non-finite point";
                return cudaErrorInvalidValue;
            }
            hx[i] = p[0], hy[i] = p[1], hz[i] = p[2];
            for (int k = 0; k < 3; ++k)
            {
                lo[k] = std::min(lo[k], p[k]);
                hi[k] = std::max(hi[k], p[k]);
            }
        }
        Buf dx, dy, dz, idx_a, idx_b, keys_a, keys_b, tests, lo_first, sort_tmp, point_leaf;
        CB_CUDA(dx.alloc(4ull * n));
        CB_CUDA(dy.alloc(4ull * n));
        CB_CUDA(dz.alloc(4ull * n));
        CB_CUDA(idx_a.alloc(4ull * n));
        CB_CUDA(idx_b.alloc(4ull * n));
        CB_CUDA(keys_a.alloc(8ull * n));
        CB_CUDA(keys_b.alloc(8ull * n));
        CB_CUDA(tests.alloc(4ull * n));
        CB_CUDA(lo_first.alloc(4ull * n));
        CB_CUDA(point_leaf.alloc(4ull * n));
        CB_CUDA(cudaMemcpyAsync(dx.p, hx.data(), 4ull * n, cudaMemcpyHostToDevice, stream));
        CB_CUDA(cudaMemcpyAsync(dy.p, hy.data(), 4ull * n, cudaMemcpyHostToDevice, stream));
        CB_CUDA(cudaMemcpyAsync(dz.p, hz.data(), 4ull * n, cudaMemcpyHostToDevice, stream));
        const unsigned tb = 256, nb = (n + tb - 1) / tb;
        iota_kernel<<<nb, tb, 0, stream>>>(idx_a.as<uint32_t>(), n);
        size_t tmp_bytes = 0;
        CB_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys_a.as<uint64_t>(), keys_b.as<uint64_t>(), idx_a.as<uint32_t>(),
                                                idx_b.as<uint32_t>(), static_cast<int>(n), 0, 64, stream));
        CB_CUDA(sort_tmp.alloc(tmp_bytes));
        uint32_t *cur = idx_a.as<uint32_t>(), *nxt = idx_b.as<uint32_t>();
        for (uint32_t L = 0; L < nlog2; ++L)
        {
            const float *coord = (L % 3 == 0) ? dx.as<float>() : (L % 3 == 1 ? dy.as<float>() : dz.as<float>());
            const uint32_t seg_shift = nlog2 - L, seg = 1u << seg_shift, nodes = 1u << L;
            level_keys_kernel<<<nb, tb, 0, stream>>>(coord, cur, n, seg_shift, keys_a.as<uint64_t>());
            CB_CUDA(cub::DeviceRadixSort::SortPairs(sort_tmp.p, tmp_bytes, keys_a.as<uint64_t>(), keys_b.as<uint64_t>(), cur, nxt,
                                                    static_cast<int>(n), 0, static_cast<int>(32 + L + 1 > 64 ? 64 : 32 + L + 1), stream));
            std::swap(cur, nxt);
            level_tests_kernel<<<(nodes + tb - 1) / tb, tb, 0, stream>>>(coord, cur, nodes, seg, nodes - 1u, tests.as<float>(),
                                                                         lo_first.as<float>());
        }
        invert_kernel<<<nb, tb, 0, stream>>>(cur, n, point_leaf.as<uint32_t>());
        CB_CUDA(cudaGetLastError());

        // uniform grid over the real points, cell edge = r_max + r_point
        const float max_l1 = r_max + r_point;
        float g_o[3];
        int g_n[3];
        size_t cells = 1;
        float h = max_l1;
        for (;;)
        {
            cells = 1;
            for (int k = 0; k < 3; ++k)
            {
                g_o[k] = lo[k] - h;
                g_n[k] = std::max(1, static_cast<int>(std::ceil((hi[k] - lo[k] + 2.f * h) / h)));
                cells *= static_cast<size_t>(g_n[k]);
            }
            if (cells <= (size_t(1) << 26))
            {
                break;
            }
            h *= 1.5f;
        }
        Buf cell_of, cell_sorted, ids, ids_sorted, cell_start, g_od, g_nd, n_vec, aff_starts;
        CB_CUDA(cell_of.alloc(4ull * n_real));
        CB_CUDA(cell_sorted.alloc(4ull * n_real));
        CB_CUDA(ids.alloc(4ull * n_real));
        CB_CUDA(ids_sorted.alloc(4ull * n_real));
        CB_CUDA(cell_start.alloc(4ull * (cells + 2)));
        CB_CUDA(g_od.alloc(16));
        CB_CUDA(g_nd.alloc(16));
        CB_CUDA(n_vec.alloc(4ull * (n + 1)));
        CB_CUDA(aff_starts.alloc(4ull * (n + 1)));
        CB_CUDA(cudaMemcpyAsync(g_od.p, g_o, sizeof(g_o), cudaMemcpyHostToDevice, stream));
        CB_CUDA(cudaMemcpyAsync(g_nd.p, g_n, sizeof(g_n), cudaMemcpyHostToDevice, stream));
        const unsigned nbr = (n_real + tb - 1) / tb;
        grid_cell_kernel<<<nbr, tb, 0, stream>>>(dx.as<float>(), dy.as<float>(), dz.as<float>(), n_real, g_od.as<float>(), 1.f / h,
                                                 g_nd.as<int>(), cell_of.as<uint32_t>(), ids.as<uint32_t>());
        size_t tmp2 = 0;
        CB_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp2, cell_of.as<uint32_t>(), cell_sorted.as<uint32_t>(), ids.as<uint32_t>(),
                                                ids_sorted.as<uint32_t>(), static_cast<int>(n_real), 0, 32, stream));
        Buf sort_tmp2;
        CB_CUDA(sort_tmp2.alloc(tmp2));
        CB_CUDA(cub::DeviceRadixSort::SortPairs(sort_tmp2.p, tmp2, cell_of.as<uint32_t>(), cell_sorted.as<uint32_t>(), ids.as<uint32_t>(),
                                                ids_sorted.as<uint32_t>(), static_cast<int>(n_real), 0, 32, stream));
        cell_start_kernel<<<(n_real + 1 + tb - 1) / tb, tb, 0, stream>>>(cell_sorted.as<uint32_t>(), n_real, static_cast<uint32_t>(cells),
                                                                         cell_start.as<uint32_t>());
        CB_CUDA(cudaGetLastError());

        BuildView b{};
        b.x = dx.as<float>(), b.y = dy.as<float>(), b.z = dz.as<float>();
        b.leaf_point = cur;
        b.point_leaf = point_leaf.as<uint32_t>();
        b.tests = tests.as<float>();
        b.lo_first = lo_first.as<float>();
        b.cell_start = cell_start.as<uint32_t>();
        b.cell_points = ids_sorted.as<uint32_t>();
        for (int k = 0; k < 3; ++k)
        {
            b.g_o[k] = g_o[k];
            b.g_n[k] = g_n[k];
        }
        b.g_inv_h = 1.f / h;
        b.n = n;
        b.nlog2 = nlog2;
        b.r_max = r_max;
        b.max_l1 = max_l1;
        b.max_l2 = max_l1 * max_l1;
        b.min_l2 = (r_min + r_point) * (r_min + r_point);
        CB_CUDA(cudaMemsetAsync(n_vec.as<uint32_t>() + n, 0, 4, stream));
        leaf_affordances_kernel<false><<<nb, tb, 0, stream>>>(b, n_vec.as<uint32_t>(), nullptr, nullptr, nullptr);
        size_t tmp3 = 0;
        CB_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp3, n_vec.as<uint32_t>(), aff_starts.as<uint32_t>(), static_cast<int>(n + 1), stream));
        Buf scan_tmp;
        CB_CUDA(scan_tmp.alloc(tmp3));
        CB_CUDA(cub::DeviceScan::ExclusiveSum(scan_tmp.p, tmp3, n_vec.as<uint32_t>(), aff_starts.as<uint32_t>(), static_cast<int>(n + 1), stream));
        uint32_t total_vec = 0;
        CB_CUDA(cudaMemcpyAsync(&total_vec, aff_starts.as<uint32_t>() + n, 4, cudaMemcpyDeviceToHost, stream));
        CB_CUDA(cudaStreamSynchronize(stream));
        Buf aff, aabbs;
        CB_CUDA(aff.alloc(96ull * total_vec));
        CB_CUDA(aabbs.alloc(24ull * n));
        leaf_affordances_kernel<true><<<nb, tb, 0, stream>>>(b, nullptr, aff_starts.as<uint32_t>(), aff.as<float>(), aabbs.as<float>());
        CB_CUDA(cudaGetLastError());

        out.tests.resize(n - 1);
        out.aff_starts.resize(n + 1);
        out.aabbs.resize(6ull * n);
        out.aff.resize(24ull * total_vec);
        if (n > 1)
        {
            CB_CUDA(cudaMemcpyAsync(out.tests.data(), tests.p, 4ull * (n - 1), cudaMemcpyDeviceToHost, stream));
        }
        CB_CUDA(cudaMemcpyAsync(out.aff_starts.data(), aff_starts.p, 4ull * (n + 1), cudaMemcpyDeviceToHost, stream));
        CB_CUDA(cudaMemcpyAsync(out.aabbs.data(), aabbs.p, 24ull * n, cudaMemcpyDeviceToHost, stream));
        if (total_vec)
        {
            CB_CUDA(cudaMemcpyAsync(out.aff.data(), aff.p, 96ull * total_vec, cudaMemcpyDeviceToHost, stream));
        }
        CB_CUDA(cudaStreamSynchronize(stream));
        for (int k = 0; k < 3; ++k)
        {
            out.top[k] = lo[k];
            out.top[3 + k] = hi[k];
        }
        out.r_point = r_point;
        out.r_min = r_min;
        out.r_max = r_max;
        out.nlog2 = static_cast<int>(nlog2);
        return cudaSuccess;
    }
}  // namespace vb

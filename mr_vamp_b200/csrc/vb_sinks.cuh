// Sinks the generated robot traces write to (see codegen/emit_cuda.py).  Concrete (non-template)
// types on purpose: a 1000-statement trace inside a template costs the CUDA front end minutes.
#pragma once
#include "vb_env.cuh"

namespace vb
{
    // ---- chain output: sphere-centre slots in shared memory, one column per thread ---------------
    // slot s, coordinate c of thread t lives at col[(3 s + c) * stride] with col = slots + t: a warp
    // reading "its" slot touches 32 consecutive words (conflict free).  Centres that do not depend on the
    // configuration are not replicated per thread: they live once per warp in the rows behind the
    // per-thread slots (put_c; slot code kConstSlot | j in the tables).
    // `col` is a float INDEX from the start of the dynamic shared-memory window, so that the compiler emits
    // STS (a generic store would resolve the shared window at run time).
    struct SlotSink
    {
        unsigned col;
        int stride;
        const AttachmentView *att;
        float *att_col;  // this thread's column of posed attachment spheres (global scratch)
        int att_stride;

        VB_D static float *window()
        {
            extern __shared__ float4 vb_shared_window[];
            return reinterpret_cast<float *>(vb_shared_window);
        }
        VB_D void put(int slot, float x, float y, float z) const
        {
            float *c = window() + col;
            c[(3 * slot + 0) * stride] = x;
            c[(3 * slot + 1) * stride] = y;
            c[(3 * slot + 2) * stride] = z;
        }
        // compile-time stride: the stores of the straight-line chain get immediate offsets
        template <int kStride>
        VB_D void put_s(int slot, float x, float y, float z) const
        {
            float *c = window() + col;
            c[(3 * slot + 0) * kStride] = x;
            c[(3 * slot + 1) * kStride] = y;
            c[(3 * slot + 2) * kStride] = z;
        }
        // configuration-independent centre (e.g. the base link): one copy per WARP, constant j of row-triple `row`
        // lives in column (warp's first column + j); every lane stores the same value
        template <int kStride>
        VB_D void put_c(int row, int j, float x, float y, float z) const
        {
            float *c = window() + (col & ~31u) + j;
            c[(3 * row + 0) * kStride] = x;
            c[(3 * row + 1) * kStride] = y;
            c[(3 * row + 2) * kStride] = z;
        }
        // collision/attachments.hh:75-122 (Attachment::pose); the posed centres go to the global scratch column
        VB_D void attach_pose(float p_tx, float p_ty, float p_tz, float p_rx, float p_ry, float p_rz, float p_rw) const
        {
            const float *t = att->tf;
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
            for (int i = 0; i < att->n; ++i)
            {
                const float4 s = att->local[i];
                att_col[(3 * i + 0) * static_cast<size_t>(att_stride)] = s.x * xx + s.y * yx + s.z * zx + tx;
                att_col[(3 * i + 1) * static_cast<size_t>(att_stride)] = s.x * xy + s.y * yy + s.z * zy + ty;
                att_col[(3 * i + 2) * static_cast<size_t>(att_stride)] = s.x * xz + s.y * yz + s.z * zz + tz;
            }
        }
    };

    struct SphereSink
    {
        float bx_, by_, bz_, radius;
        float *out;  // this configuration's [S][4]
        VB_D float bx() const { return bx_; }
        VB_D float by() const { return by_; }
        VB_D float bz() const { return bz_; }
        // `literal + base_*` is evaluated in double and narrowed, as the C++ expression is
        VB_D float cbx(double c) const { return static_cast<float>(c + static_cast<double>(bx_)); }
        VB_D float cby(double c) const { return static_cast<float>(c + static_cast<double>(by_)); }
        VB_D float cbz(double c) const { return static_cast<float>(c + static_cast<double>(bz_)); }
        VB_D void out_x(int i, float v) const { out[4 * i + 0] = v; }
        VB_D void out_y(int i, float v) const { out[4 * i + 1] = v; }
        VB_D void out_z(int i, float v) const { out[4 * i + 2] = v; }
        VB_D void out_r(int i, float v) const { out[4 * i + 3] = v; }
    };

}  // namespace vb

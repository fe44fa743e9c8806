"""Python face of ``libvamp_b200.so`` (ctypes over the C ABI of ``include/vamp_b200.h``).

Mirrors the part of the reference's nanobind module ``vamp._core_ext`` that sits on the fkcc path:
shape constructors and ``Environment.add_*`` (``bindings/environment.cc:18-250``), and per robot
``validate`` / ``fk`` (``bindings/common.hh:132-189,685-710``) — plus the batch entry points the
reference lists as a TODO (``README.md:308``): ``validate_batch``, ``validate_motion_batch``,
``fk_batch``.  Nothing here computes a verdict: every query goes to the CUDA library, and the
import fails loudly when the library is missing.
"""
from __future__ import annotations

import ctypes as C
import math
import os
from typing import Iterable, Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("VAMP_B200_LIB") or os.path.join(_HERE, "lib", "libvamp_b200.so")  # override: development A/B builds

_f32p = C.POINTER(C.c_float)
_u32p = C.POINTER(C.c_uint32)
_vp = C.c_void_p

ROBOT_IDS = {"sphere": 0, "ur5": 1, "panda": 2, "fetch": 3, "baxter": 4}

# every symbol include/vamp_b200.h declares: (name, restype, argtypes)
_SIGNATURES = [
    ("vb_last_error", C.c_char_p, []),
    ("vb_version", C.c_char_p, []),
    ("vb_ctx_create", C.c_int, [C.POINTER(_vp), C.POINTER(C.c_int), C.c_int]),
    ("vb_ctx_destroy", None, [_vp]),
    ("vb_ctx_device_count", C.c_int, [_vp]),
    ("vb_ctx_sm_count", C.c_int, [_vp, C.c_int]),
    ("vb_sync", C.c_int, [_vp]),
    ("vb_device_pci_bus_id", C.c_int, [C.c_int, C.c_char_p, C.c_int]),
    ("vb_robot_by_name", C.c_int, [C.c_char_p, C.POINTER(C.c_int)]),
    ("vb_robot_info", C.c_int, [C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    ("vb_robot_limits", C.c_int, [C.c_int, _f32p, _f32p]),
    ("vb_random_config", C.c_int, [C.c_int, C.c_uint64, C.c_uint64, _f32p]),
    ("vb_random_edge", C.c_int, [C.c_int, C.c_uint64, C.c_uint64, C.c_float, C.c_float, _f32p, _f32p]),
    ("vb_random_configs", C.c_int, [C.c_int, C.c_uint64, C.c_uint64, C.c_size_t, _f32p]),
    ("vb_random_edges", C.c_int, [C.c_int, C.c_uint64, C.c_uint64, C.c_size_t, C.c_float, C.c_float, _f32p, _f32p]),
    ("vb_random_configs_d", C.c_int, [_vp, C.c_int, C.c_int, C.c_uint64, C.c_uint64, C.c_size_t, _vp]),
    ("vb_random_edges_d", C.c_int, [_vp, C.c_int, C.c_int, C.c_uint64, C.c_uint64, C.c_size_t, C.c_float, C.c_float, _vp, _vp]),
    ("vb_host_alloc", C.c_int, [C.c_size_t, C.POINTER(_vp)]),
    ("vb_host_free", C.c_int, [_vp]),
    ("vb_host_register", C.c_int, [_vp, C.c_size_t]),
    ("vb_host_unregister", C.c_int, [_vp]),
    ("vb_ctx_set_sphere_radius", C.c_int, [_vp, C.c_float]),
    ("vb_env_create", C.c_int, [_vp, C.POINTER(_vp)]),
    ("vb_env_destroy", None, [_vp]),
    ("vb_env_clear", C.c_int, [_vp]),
    ("vb_env_add_sphere", C.c_int, [_vp, C.c_float, C.c_float, C.c_float, C.c_float]),
    ("vb_env_add_cuboid", C.c_int, [_vp, _f32p]),
    ("vb_env_add_capsule", C.c_int, [_vp, _f32p]),
    ("vb_env_add_heightfield", C.c_int, [_vp, _f32p, C.c_size_t, C.c_size_t, _f32p]),
    ("vb_env_add_capt", C.c_int, [_vp, _f32p, C.c_size_t, _u32p, _f32p, _f32p, _f32p, C.c_size_t, _f32p,
                                  C.c_size_t, _f32p, C.c_float, C.c_int]),
    ("vb_env_add_pointcloud", C.c_int, [_vp, _f32p, C.c_size_t, C.c_float, C.c_float, C.c_float, C.POINTER(C.c_double)]),
    ("vb_env_capt_export", C.c_int, [_vp, C.c_size_t, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t), _f32p, _u32p, _f32p, _f32p, _f32p,
                                     C.POINTER(C.c_int)]),
    ("vb_env_set_attachment", C.c_int, [_vp, _f32p, _f32p, C.c_size_t]),
    ("vb_env_clear_attachment", C.c_int, [_vp]),
    ("vb_env_commit", C.c_int, [_vp]),
    ("vb_env_export", C.c_int, [_vp, C.c_int, _f32p, C.POINTER(C.c_size_t)]),
    ("vb_env_object_ids", C.c_int, [_vp, _u32p, C.POINTER(C.c_size_t)]),
    ("vb_sphere_validity", C.c_int, [_vp, C.c_int, _f32p, _vp, _f32p, C.c_size_t, C.c_uint32, _u32p, _u32p]),
    ("vb_sphere_fk", C.c_int, [_vp, C.c_int, _f32p, _f32p, C.c_size_t, _f32p]),
    ("vb_fkcc", C.c_int, [_vp, C.c_int, _f32p, _vp, _f32p, C.c_size_t, _u32p, _f32p]),
    ("vb_validate", C.c_int, [_vp, C.c_int, _f32p, _vp, _f32p, C.c_size_t, _u32p]),
    ("vb_validate_motion", C.c_int, [_vp, C.c_int, _f32p, _vp, _f32p, _f32p, C.c_size_t, C.c_int, _u32p, _f32p]),
    ("vb_validate_vector", C.c_int, [_vp, C.c_int, _f32p, _vp, _f32p, _f32p, _f32p, C.c_size_t, C.c_int, _u32p, _f32p]),
    ("vb_connect_increments", C.c_int, [_vp, C.c_int, _f32p, _vp, _f32p, _f32p, _f32p, C.c_size_t, C.c_float, C.c_int,
                                        _u32p, _u32p]),
    ("vb_sample_valid", C.c_int, [_vp, C.c_int, _f32p, _vp, _f32p, C.c_uint64, C.c_uint64, C.c_size_t, C.c_size_t, _f32p,
                                  C.POINTER(C.c_uint64), C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]),
    ("vb_validate_motion_indexed", C.c_int, [_vp, C.c_int, _f32p, _vp, _vp, C.c_size_t, _u32p, C.c_size_t, C.c_int, _u32p]),
    ("vb_validate_pairs", C.c_int, [_vp, C.c_int, _f32p, C.c_int, _f32p, _vp, _f32p, _f32p, C.c_size_t, _u32p]),
    ("vb_inter_robot", C.c_int, [_vp, C.c_int, _f32p, C.c_int, _f32p, _f32p, _f32p, C.c_size_t, _u32p]),
    ("vb_validate_pairs_d", C.c_int, [_vp, C.c_int, C.c_int, _f32p, _f32p, _vp, _vp, _vp, C.c_size_t, _vp]),
    ("vb_dev_alloc", C.c_int, [_vp, C.c_int, C.c_size_t, C.POINTER(_vp)]),
    ("vb_dev_free", C.c_int, [_vp, C.c_int, _vp]),
    ("vb_dev_upload", C.c_int, [_vp, C.c_int, _vp, _vp, C.c_size_t]),
    ("vb_dev_download", C.c_int, [_vp, C.c_int, _vp, _vp, C.c_size_t]),
    ("vb_fkcc_d", C.c_int, [_vp, C.c_int, C.c_int, _f32p, _vp, _vp, C.c_size_t, _vp, _vp]),
    ("vb_fkcc_work_d", C.c_int, [_vp, C.c_int, C.c_int, _f32p, _vp, _vp, C.c_size_t, _vp, _vp]),
    ("vb_validate_motion_d", C.c_int, [_vp, C.c_int, C.c_int, _f32p, _vp, _vp, _vp, C.c_size_t, C.c_int, _vp, _vp]),
    ("vb_validate_vector_d", C.c_int, [_vp, C.c_int, C.c_int, _f32p, _vp, _vp, _vp, _vp, C.c_size_t, C.c_int, _vp, _vp]),
    ("vb_validate_motion_indexed_d", C.c_int, [_vp, C.c_int, C.c_int, _f32p, _vp, _vp, C.c_size_t, _vp, C.c_size_t, C.c_int, _vp]),
    ("vb_noalloc_max_edges", C.c_size_t, []),
    ("vb_validate_motion_noalloc_d", C.c_int, [_vp, C.c_int, C.c_int, _f32p, _vp, _vp, _vp, C.c_size_t, C.c_int, _vp]),
    ("vb_fkcc_random_d", C.c_int, [_vp, C.c_int, C.c_int, _f32p, _vp, C.c_uint64, C.c_uint64, C.c_size_t, _vp]),
    ("vb_timer_start", C.c_int, [_vp, C.c_int]),
    ("vb_timer_stop", C.c_int, [_vp, C.c_int, C.POINTER(C.c_double)]),
    ("vb_launch_count", C.c_uint64, [_vp]),
    ("vb_validate_path", C.c_int, [_vp, C.c_int, _f32p, _vp, _f32p, C.c_size_t, C.c_int, C.POINTER(C.c_int64)]),
    ("vb_shortcut_path", C.c_int, [_vp, C.c_int, _f32p, _vp, _f32p, C.c_size_t, C.c_int, C.POINTER(C.c_uint8),
                          C.POINTER(C.c_size_t)]),
    ("vb_fp32_peak_d", C.c_int, [_vp, C.c_int, C.POINTER(C.c_double)]),
    ("vb_popcount_d", C.c_int, [_vp, C.c_int, _vp, C.c_size_t, C.POINTER(C.c_uint64)]),
    ("vb_comm_unique_id", C.c_int, [_vp]),
    ("vb_comm_init", C.c_int, [_vp, C.c_int, C.c_int, _vp]),
    ("vb_comm_allgather_bits_d", C.c_int, [_vp, _vp, _vp, C.c_size_t]),
    ("vb_comm_destroy", C.c_int, [_vp]),
]

_lib = None


class VampB200Error(RuntimeError):
    pass


def lib():
    """Load the CUDA library.  There is no fallback: a missing library is an error."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise VampB200Error(f"{LIB_PATH} is missing: build it with `make` (nvcc, sm_100a); there is no CPU path")
        L = C.CDLL(LIB_PATH)
        for name, res, args in _SIGNATURES:
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def exported_symbols() -> list[str]:
    return [s[0] for s in _SIGNATURES]


def _check(rc: int):
    if rc != 0:
        raise VampB200Error(f"[{rc}] {lib().vb_last_error().decode()}")


def _f32(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.float32)


def _ptr(a: Optional[np.ndarray], typ=_f32p):
    return None if a is None else a.ctypes.data_as(typ)


def unpack_bits(words: np.ndarray, n: int) -> np.ndarray:
    """uint32 verdict words -> bool[n] (bit i%32 of word i//32)."""
    return np.unpackbits(words.view(np.uint8), bitorder="little")[:n].astype(bool)


def pinned_empty(shape, dtype) -> np.ndarray:
    """numpy array over page-locked host memory (vb_host_alloc): what the host-buffer queries want for PCIe-speed copies.
    The memory is released when the array (and every view of it) is garbage collected."""
    dtype = np.dtype(dtype)
    n = int(np.prod(shape))
    p = _vp()
    _check(lib().vb_host_alloc(max(1, n * dtype.itemsize), C.byref(p)))
    buf = (C.c_char * max(1, n * dtype.itemsize)).from_address(p.value)

    class _Owner:
        def __init__(self, ptr):
            self.ptr = ptr

        def __del__(self):
            try:
                lib().vb_host_free(self.ptr)
            except Exception:
                pass

    arr = np.frombuffer(buf, dtype=dtype, count=n).reshape(shape)
    _PINNED_OWNERS[id(buf)] = (buf, _Owner(p))
    return arr


_PINNED_OWNERS: dict = {}


# ---- shapes (collision/factory.hh; arithmetic in float32 like the reference's Eigen::*f types) ----
def _rotation_xyz(euler_xyz: Sequence[float]) -> np.ndarray:
    """Rz(phi) * Ry(theta) * Rx(rho) (collision/factory.hh:37-39)."""
    rho, theta, phi = (np.float32(v) for v in euler_xyz)
    cx, sx = np.float32(math.cos(rho)), np.float32(math.sin(rho))
    cy, sy = np.float32(math.cos(theta)), np.float32(math.sin(theta))
    cz, sz = np.float32(math.cos(phi)), np.float32(math.sin(phi))
    rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]], dtype=np.float32)
    ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]], dtype=np.float32)
    rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]], dtype=np.float32)
    return (rz @ ry @ rx).astype(np.float32)


class Sphere:
    """bindings/environment.cc:20-36."""

    def __init__(self, center: Sequence[float], radius: float):
        self.x, self.y, self.z = (float(np.float32(v)) for v in center)
        self.r = float(np.float32(radius))
        self.name = ""

    @property
    def position(self):
        return [self.x, self.y, self.z]


class Cuboid:
    """bindings/environment.cc:70-98; raw() = the 15 floats of collision/shapes.hh:34-50."""

    def __init__(self, center: Sequence[float], euler_xyz: Sequence[float], half_extents: Sequence[float]):
        rot = _rotation_xyz(euler_xyz)
        self._raw = np.concatenate(
            [_f32(center), rot[:, 0], rot[:, 1], rot[:, 2], _f32(half_extents)]
        ).astype(np.float32)
        self.name = ""

    @classmethod
    def from_raw(cls, raw15: Sequence[float]) -> "Cuboid":
        c = cls.__new__(cls)
        c._raw = _f32(raw15).copy()
        assert c._raw.shape == (15,)
        c.name = ""
        return c

    def raw(self) -> np.ndarray:
        return self._raw


class Cylinder:
    """bindings/environment.cc:38-68: (center, euler_xyz, radius, length) or (endpoint1, endpoint2, radius)."""

    def __init__(self, a: Sequence[float], b: Sequence[float], radius: float, length: Optional[float] = None):
        if length is not None:
            rot = _rotation_xyz(b)
            half = np.float32(length) / np.float32(2)
            ctr = _f32(a)
            p1 = rot @ np.array([0, 0, half], dtype=np.float32) + ctr
            p2 = rot @ np.array([0, 0, -half], dtype=np.float32) + ctr
        else:
            p1, p2 = _f32(a), _f32(b)
        v = (p2 - p1).astype(np.float32)
        dot = np.float32(v[0] * v[0]) + np.float32(v[1] * v[1]) + np.float32(v[2] * v[2])
        rdv = np.float32(1.0 / float(dot))  # collision/factory.hh:233-244
        self._raw = np.array([p1[0], p1[1], p1[2], v[0], v[1], v[2], np.float32(radius), rdv], dtype=np.float32)
        self.name = ""

    @classmethod
    def from_raw(cls, raw8: Sequence[float]) -> "Cylinder":
        c = cls.__new__(cls)
        c._raw = _f32(raw8).copy()
        assert c._raw.shape == (8,)
        c.name = ""
        return c

    def raw(self) -> np.ndarray:
        return self._raw


Capsule = Cylinder


class HeightField:
    """collision/factory.hh:365-386 (make_heightfield): scales are stored inverted."""

    def __init__(self, center: Sequence[float], scaling: Sequence[float], data: np.ndarray):
        data = _f32(data)
        assert data.ndim == 2
        self.yd, self.xd = data.shape
        inv = [np.float32(1.0) / np.float32(v) for v in scaling]
        self.off_scale = np.array([center[0], center[1], center[2], inv[0], inv[1], inv[2]], dtype=np.float32)
        self.data = data.reshape(-1)

    @classmethod
    def from_raw(cls, off_scale6, xd: int, yd: int, data) -> "HeightField":
        h = cls.__new__(cls)
        h.off_scale = _f32(off_scale6).copy()
        h.xd, h.yd = int(xd), int(yd)
        h.data = _f32(data).reshape(-1)
        return h


class Attachment:
    """collision/attachments.hh:11-73: transform (translation + xyzw quaternion) and local spheres."""

    def __init__(self, position: Sequence[float], quaternion_xyzw: Sequence[float]):
        self.tf = np.concatenate([_f32(position), _f32(quaternion_xyzw)]).astype(np.float32)
        self.spheres: list[Sphere] = []

    def add_spheres(self, spheres: Iterable[Sphere]):
        self.spheres.extend(spheres)


# ---- context ---------------------------------------------------------------------------------
class Context:
    def __init__(self, devices: Optional[Sequence[int]] = None):
        self._h = _vp()
        if devices is None:
            _check(lib().vb_ctx_create(C.byref(self._h), None, 0))
        else:
            ids = (C.c_int * len(devices))(*devices)
            _check(lib().vb_ctx_create(C.byref(self._h), ids, len(devices)))

    def close(self):
        if self._h:
            lib().vb_ctx_destroy(self._h)
            self._h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def handle(self):
        return self._h

    @property
    def n_devices(self) -> int:
        return lib().vb_ctx_device_count(self._h)

    def sm_count(self, slot: int = 0) -> int:
        return lib().vb_ctx_sm_count(self._h, slot)

    def sync(self):
        _check(lib().vb_sync(self._h))

    def launch_count(self) -> int:
        return int(lib().vb_launch_count(self._h))

    def set_sphere_radius(self, r: float):
        _check(lib().vb_ctx_set_sphere_radius(self._h, r))


_default_ctx: Optional[Context] = None


def default_context() -> Context:
    global _default_ctx
    if _default_ctx is None:
        _default_ctx = Context()
    return _default_ctx


# ---- environment -----------------------------------------------------------------------------
class Environment:
    """Mirror of ``vamp.Environment`` (bindings/environment.cc:107-163).  Mutations are recorded on
    the host; the device copy is rebuilt lazily (one upload per change set instead of the reference's
    float->vector deep copy on every query, bindings/common.hh:181)."""

    def __init__(self, ctx: Optional[Context] = None, host_only: bool = False):
        """host_only: a host-side description without a device context (population, classification,
        sorting and export work; queries against it fail) -- used by tools and the CPU tests."""
        self.ctx = None if host_only else (ctx or default_context())
        self._h = _vp()
        _check(lib().vb_env_create(self.ctx.handle if self.ctx else None, C.byref(self._h)))
        self._dirty = True
        self._names: list[str] = []  # by insertion id (vb_env_object_ids)

    def __del__(self):
        try:
            if self._h and (self.ctx is None or self.ctx._h):
                lib().vb_env_destroy(self._h)
        except Exception:
            pass

    def add_sphere(self, s: Sphere):
        _check(lib().vb_env_add_sphere(self._h, s.x, s.y, s.z, s.r))
        self._names.append(getattr(s, "name", ""))
        self._dirty = True

    def add_cuboid(self, c: Cuboid):
        _check(lib().vb_env_add_cuboid(self._h, _ptr(c.raw())))
        self._names.append(getattr(c, "name", ""))
        self._dirty = True

    def add_capsule(self, c: Cylinder):
        _check(lib().vb_env_add_capsule(self._h, _ptr(c.raw())))
        self._names.append(getattr(c, "name", ""))
        self._dirty = True

    def add_heightfield(self, h: HeightField):
        _check(lib().vb_env_add_heightfield(self._h, _ptr(h.off_scale), h.xd, h.yd, _ptr(h.data)))
        self._names.append(getattr(h, "name", ""))
        self._dirty = True

    def add_capt_buffers(self, tests, aff_starts, aff_x, aff_y, aff_z, aabbs, aabb_top, r_point: float, nlog2: int):
        """CAPT query buffers built on the host (collision/capt.hh:588-621)."""
        tests, aff_x, aff_y, aff_z, aabbs = (_f32(a).reshape(-1) for a in (tests, aff_x, aff_y, aff_z, aabbs))
        aff_starts = np.ascontiguousarray(aff_starts, dtype=np.uint32)
        top = _f32(aabb_top)
        _check(
            lib().vb_env_add_capt(
                self._h, _ptr(tests), tests.size, _ptr(aff_starts, _u32p), _ptr(aff_x), _ptr(aff_y), _ptr(aff_z),
                aff_x.size // 8, _ptr(aabbs), aabbs.size // 6, _ptr(top), r_point, nlog2,
            )
        )
        self._dirty = True

    def add_pointcloud(self, points, r_min: float, r_max: float, r_point: float) -> float:
        """Environment.add_pointcloud (bindings/environment.cc:148-158): builds the CAPT from raw points ON THE DEVICE
        (collision/capt.hh:137-398); returns the build time in nanoseconds like the reference's binding."""
        pts = _f32(points).reshape(-1, 3)
        ms = C.c_double()
        _check(lib().vb_env_add_pointcloud(self._h, _ptr(pts), pts.shape[0], r_min, r_max, r_point, C.byref(ms)))
        self._dirty = True
        return ms.value * 1e6

    def capt(self, idx: int = 0) -> dict:
        """Query buffers of point cloud idx (as ref.Env.capt of the oracle returns them; affordances split per axis)."""
        nt, nv, nl = C.c_size_t(), C.c_size_t(), C.c_int()
        _check(lib().vb_env_capt_export(self._h, idx, C.byref(nt), C.byref(nv), None, None, None, None, None, C.byref(nl)))
        leaves = 1 << nl.value
        tests = np.zeros(nt.value, np.float32)
        starts = np.zeros(leaves + 1, np.uint32)
        aff = np.zeros((nv.value, 3, 8), np.float32)
        aabbs = np.zeros((leaves, 6), np.float32)
        top = np.zeros(6, np.float32)
        _check(lib().vb_env_capt_export(self._h, idx, C.byref(nt), C.byref(nv), _ptr(tests), _ptr(starts, _u32p), _ptr(aff),
                                        _ptr(aabbs), _ptr(top), C.byref(nl)))
        return {"nlog2": nl.value, "tests": tests, "aff_starts": starts, "aff_x": aff[:, 0, :].reshape(-1).copy(),
                "aff_y": aff[:, 1, :].reshape(-1).copy(), "aff_z": aff[:, 2, :].reshape(-1).copy(), "aabbs": aabbs, "aabb_top": top}

    def attach(self, a: Attachment):
        sph = _f32([[s.x, s.y, s.z, s.r] for s in a.spheres]).reshape(-1, 4)
        _check(lib().vb_env_set_attachment(self._h, _ptr(a.tf), _ptr(sph), len(a.spheres)))
        self._dirty = True

    def detach(self):
        _check(lib().vb_env_clear_attachment(self._h))
        self._dirty = True

    def object_names(self) -> list[str]:
        """Obstacle names in the library's object order (sorted lists, then height fields)."""
        n = C.c_size_t()
        _check(lib().vb_env_object_ids(self._h, None, C.byref(n)))
        ids = np.zeros(n.value, dtype=np.uint32)
        if n.value:
            _check(lib().vb_env_object_ids(self._h, _ptr(ids, _u32p), C.byref(n)))
        return [self._names[i] for i in ids]

    def export(self, kind: int) -> np.ndarray:
        width = {0: 5, 1: 9, 2: 9, 3: 16, 4: 16}[kind]
        n = C.c_size_t()
        _check(lib().vb_env_export(self._h, kind, None, C.byref(n)))
        out = np.zeros((n.value, width), dtype=np.float32)
        if n.value:
            _check(lib().vb_env_export(self._h, kind, _ptr(out), C.byref(n)))
        return out

    @property
    def handle(self):
        if self._dirty:
            _check(lib().vb_env_commit(self._h))
            self._dirty = False
        return self._h


# ---- robots ----------------------------------------------------------------------------------
class Robot:
    """One robot submodule of the reference's Python module (``vamp.panda`` etc., bindings/common.hh:355-728),
    restricted to the fkcc path, plus the batch calls."""

    def __init__(self, name: str, base: Sequence[float] = (0.0, 0.0, 0.0), ctx: Optional[Context] = None):
        self.name = name
        self.robot_id = ROBOT_IDS[name]
        self._ctx = ctx
        self.base = _f32(base)
        d, s, r = C.c_int(), C.c_int(), C.c_int()
        _check(lib().vb_robot_info(self.robot_id, C.byref(d), C.byref(s), C.byref(r)))
        self.dimension, self.n_spheres, self.resolution = d.value, s.value, r.value
        lo = np.zeros(self.dimension, dtype=np.float32)
        rg = np.zeros(self.dimension, dtype=np.float32)
        _check(lib().vb_robot_limits(self.robot_id, _ptr(lo), _ptr(rg)))
        self.lower_bounds, self.ranges = lo, rg

    @property
    def ctx(self) -> Context:
        return self._ctx or default_context()

    def _env(self, env: Optional[Environment]) -> Environment:
        if env is None:
            if not hasattr(self, "_empty_env") or self._empty_env.ctx is not self.ctx:
                self._empty_env = Environment(self.ctx)
            return self._empty_env
        return env

    def _configs(self, q) -> np.ndarray:
        q = _f32(q)
        if q.ndim == 1:
            q = q[None, :]
        if q.ndim != 2 or q.shape[1] != self.dimension:
            raise ValueError(f"{self.name}: expected [n, {self.dimension}] configurations, got {q.shape}")
        return q

    # -- batch API (README.md:308 "Batch configuration validation") ------------------------------
    def fkcc_batch(self, configs, env: Optional[Environment] = None, margin: bool = False):
        """Robot::fkcc per configuration -> bool[n] (True = collision free) [, float32[n] margins]."""
        q = self._configs(configs)
        n = q.shape[0]
        bits = np.zeros((n + 31) // 32, dtype=np.uint32)
        mg = np.zeros(n, dtype=np.float32) if margin else None
        _check(lib().vb_fkcc(self.ctx.handle, self.robot_id, _ptr(self.base), self._env(env).handle, _ptr(q), n,
                             _ptr(bits, _u32p), _ptr(mg)))
        ok = unpack_bits(bits, n)
        return (ok, mg) if margin else ok

    def validate_batch(self, configs, env: Optional[Environment] = None) -> np.ndarray:
        """Helper::validate_configuration per configuration: joint limits && fkcc."""
        q = self._configs(configs)
        n = q.shape[0]
        bits = np.zeros((n + 31) // 32, dtype=np.uint32)
        _check(lib().vb_validate(self.ctx.handle, self.robot_id, _ptr(self.base), self._env(env).handle, _ptr(q), n,
                                 _ptr(bits, _u32p)))
        return unpack_bits(bits, n)

    def validate_motion_batch(self, starts, goals, env: Optional[Environment] = None, resolution: int = 0,
                              margin: bool = False):
        """planning::validate_motion per edge -> bool[n]."""
        a, b = self._configs(starts), self._configs(goals)
        if a.shape != b.shape:
            raise ValueError("starts and goals differ in shape")
        n = a.shape[0]
        bits = np.zeros((n + 31) // 32, dtype=np.uint32)
        mg = np.zeros(n, dtype=np.float32) if margin else None
        _check(lib().vb_validate_motion(self.ctx.handle, self.robot_id, _ptr(self.base), self._env(env).handle,
                                        _ptr(a), _ptr(b), n, resolution, _ptr(bits, _u32p), _ptr(mg)))
        ok = unpack_bits(bits, n)
        return (ok, mg) if margin else ok

    def validate_vector_batch(self, starts, vectors, distances, env: Optional[Environment] = None, resolution: int = 0,
                              margin: bool = False):
        """planning::validate_vector(start, vector, distance, env) per edge (planning/validate.hh:23-65) -> bool[n]."""
        a, v = self._configs(starts), self._configs(vectors)
        d = _f32(distances).reshape(-1)
        if a.shape != v.shape or d.shape[0] != a.shape[0]:
            raise ValueError("starts, vectors and distances differ in shape")
        n = a.shape[0]
        bits = np.zeros((n + 31) // 32, dtype=np.uint32)
        mg = np.zeros(n, dtype=np.float32) if margin else None
        _check(lib().vb_validate_vector(self.ctx.handle, self.robot_id, _ptr(self.base), self._env(env).handle,
                                        _ptr(a), _ptr(v), _ptr(d), n, resolution, _ptr(bits, _u32p), _ptr(mg)))
        ok = unpack_bits(bits, n)
        return (ok, mg) if margin else ok

    def connect_increments(self, from_, to, distances, range_: float, env: Optional[Environment] = None, resolution: int = 0):
        """RRTC's connect loop (planning/rrtc.hh:172-194) for many attempts in one batch ->
        (n_valid uint32[n], n_ext uint32[n]); attempt i connects iff n_valid[i] == n_ext[i]."""
        a, b = self._configs(from_), self._configs(to)
        d = _f32(distances).reshape(-1)
        n = a.shape[0]
        n_valid = np.zeros(n, dtype=np.uint32)
        n_ext = np.zeros(n, dtype=np.uint32)
        _check(lib().vb_connect_increments(self.ctx.handle, self.robot_id, _ptr(self.base), self._env(env).handle, _ptr(a),
                                           _ptr(b), _ptr(d), n, range_, resolution, _ptr(n_valid, _u32p), _ptr(n_ext, _u32p)))
        return n_valid, n_ext

    def sample_valid(self, k: int, env: Optional[Environment] = None, candidates=None, seed: int = 0, first: int = 0,
                     max_candidates: int = 0):
        """The first k valid configurations of a stream, in stream order (the sampling loops of planning/prm.hh:111-124,
        planning/fcit.hh:316-343 for a whole batch) -> (configs float32[found, dim], stream index uint64[found], consumed)."""
        cand = None if candidates is None else self._configs(candidates)
        m = cand.shape[0] if cand is not None else (max_candidates or 64 * k)
        out = np.zeros((k, self.dimension), dtype=np.float32)
        idx = np.zeros(k, dtype=np.uint64)
        found, consumed = C.c_size_t(), C.c_size_t()
        _check(lib().vb_sample_valid(self.ctx.handle, self.robot_id, _ptr(self.base), self._env(env).handle, _ptr(cand), seed, first,
                                     m, k, _ptr(out), idx.ctypes.data_as(C.POINTER(C.c_uint64)), C.byref(found), C.byref(consumed)))
        return out[: found.value], idx[: found.value], int(consumed.value)

    def fk_batch(self, configs) -> np.ndarray:
        """Robot::sphere_fk -> float32[n, n_spheres, 4] (x, y, z, r), world frame."""
        q = self._configs(configs)
        out = np.zeros((q.shape[0], self.n_spheres, 4), dtype=np.float32)
        _check(lib().vb_sphere_fk(self.ctx.handle, self.robot_id, _ptr(self.base), _ptr(q), q.shape[0], _ptr(out)))
        return out

    # -- the reference's single-query signatures -------------------------------------------------
    def validate(self, configuration, environment: Optional[Environment] = None) -> bool:
        """vamp.<robot>.validate(configuration, environment) (bindings/common.hh:685-697)."""
        return bool(self.validate_batch(configuration, environment)[0])

    def fk(self, configuration) -> list[Sphere]:
        """vamp.<robot>.fk(configuration) -> list[Sphere] (bindings/common.hh:706-710)."""
        return [Sphere(s[:3], s[3]) for s in self.fk_batch(configuration)[0]]

    def sphere_validity(self, configuration, environment: Optional[Environment] = None) -> list[list[str]]:
        """vamp.<robot>.sphere_validity(configuration, environment) (bindings/common.hh:155-170,699-704): for every
        collision sphere of the robot, the names of the obstacles it overlaps."""
        q = self._configs(configuration)[:1]
        env = self._env(environment)
        names = env.object_names()
        cap = max(1, len(names))
        counts = np.zeros(self.n_spheres, dtype=np.uint32)
        hits = np.zeros((self.n_spheres, cap), dtype=np.uint32)
        _check(lib().vb_sphere_validity(self.ctx.handle, self.robot_id, _ptr(self.base), env.handle, _ptr(q), 1, cap,
                                        _ptr(hits, _u32p), _ptr(counts, _u32p)))
        return [[names[o] for o in hits[k, : counts[k]]] for k in range(self.n_spheres)]

    def validate_path(self, waypoints, environment: Optional[Environment] = None) -> bool:
        """Path.validate(env) (planning/plan.hh:155-168): every consecutive edge in ONE batch call."""
        return self.first_invalid_edge(waypoints, environment) < 0

    def first_invalid_edge(self, waypoints, environment: Optional[Environment] = None, resolution: int = 0) -> int:
        """Index of the edge at which Path::validate's loop returns false (planning/plan.hh:155-168), or -1."""
        w = self._configs(waypoints)
        first = C.c_int64(-1)
        _check(lib().vb_validate_path(self.ctx.handle, self.robot_id, _ptr(self.base), self._env(environment).handle,
                                      _ptr(w), w.shape[0], resolution, C.byref(first)))
        return int(first.value)

    def shortcut_path(self, waypoints, environment: Optional[Environment] = None, resolution: int = 0) -> np.ndarray:
        """planning::shortcut_path (planning/simplify.hh:115-140): the waypoints the reference's greedy scan keeps,
        with all (n-1)(n-2)/2 candidate shortcuts validated in ONE batch call."""
        w = self._configs(waypoints)
        keep = np.ones(w.shape[0], dtype=np.uint8)
        kept = C.c_size_t(0)
        _check(lib().vb_shortcut_path(self.ctx.handle, self.robot_id, _ptr(self.base), self._env(environment).handle,
                                      _ptr(w), w.shape[0], resolution, keep.ctypes.data_as(C.POINTER(C.c_uint8)),
                                      C.byref(kept)))
        assert int(keep.sum()) == kept.value
        return w[keep.astype(bool)]

    # -- synthetic streams shared with the device-side generators ---------------------------------
    def random_configs(self, seed: int, first: int, n: int) -> np.ndarray:
        out = np.zeros((n, self.dimension), dtype=np.float32)
        _check(lib().vb_random_configs(self.robot_id, seed, first, n, _ptr(out)))
        return out

    def random_edges(self, seed: int, first: int, n: int, len_lo: float, len_hi: float):
        a = np.zeros((n, self.dimension), dtype=np.float32)
        b = np.zeros((n, self.dimension), dtype=np.float32)
        _check(lib().vb_random_edges(self.robot_id, seed, first, n, len_lo, len_hi, _ptr(a), _ptr(b)))
        return a, b


class DeviceTable:
    """A device-resident [n][dim] configuration table (roadmap vertices) for index-pair edge validation."""

    def __init__(self, robot: Robot, rows):
        self.robot = robot
        rows = robot._configs(rows)
        self.n = rows.shape[0]
        self._p = _vp()
        _check(lib().vb_dev_alloc(robot.ctx.handle, 0, max(16, rows.nbytes), C.byref(self._p)))
        _check(lib().vb_dev_upload(robot.ctx.handle, 0, self._p, rows.ctypes.data_as(_vp), rows.nbytes))
        _check(lib().vb_sync(robot.ctx.handle))

    def validate_edges(self, pairs, env: Optional[Environment] = None, resolution: int = 0) -> np.ndarray:
        """pairs uint32[n_edges, 2] (row of the start, row of the goal) -> bool[n_edges]."""
        pairs = np.ascontiguousarray(pairs, dtype=np.uint32).reshape(-1, 2)
        n = pairs.shape[0]
        bits = np.zeros((n + 31) // 32, dtype=np.uint32)
        r = self.robot
        _check(lib().vb_validate_motion_indexed(r.ctx.handle, r.robot_id, _ptr(r.base), r._env(env).handle, self._p, self.n,
                                                _ptr(pairs, _u32p), n, resolution, _ptr(bits, _u32p)))
        return unpack_bits(bits, n)

    def close(self):
        if self._p:
            lib().vb_dev_free(self.robot.ctx.handle, 0, self._p)
            self._p = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def validate_pairs_batch(robot_a: Robot, cfg_a, robot_b: Robot, cfg_b, env: Optional[Environment] = None) -> np.ndarray:
    """Composite validity of two robots sharing one environment (the fork's planned
    MultiRobotEnvironment::check_inter_robot_collisions, MR_DEV_PLAN.md:176-198): fkcc(a) and fkcc(b) and
    no world-frame sphere of a overlaps a sphere of b.  -> bool[n]."""
    qa, qb = robot_a._configs(cfg_a), robot_b._configs(cfg_b)
    if qa.shape[0] != qb.shape[0]:
        raise ValueError("both robots need the same number of configurations")
    n = qa.shape[0]
    bits = np.zeros((n + 31) // 32, dtype=np.uint32)
    _check(lib().vb_validate_pairs(robot_a.ctx.handle, robot_a.robot_id, _ptr(robot_a.base), robot_b.robot_id, _ptr(robot_b.base),
                                   robot_a._env(env).handle, _ptr(qa), _ptr(qb), n, _ptr(bits, _u32p)))
    return unpack_bits(bits, n)


def panda_grid(x100: int, y100: int, z100: int = 0, ctx: Optional[Context] = None) -> Robot:
    """PandaBase<X100, Y100, Z100> (robots/panda_base.hh:15-51): base offset in centimetres."""
    base = [np.float32(x100) / np.float32(100.0), np.float32(y100) / np.float32(100.0), np.float32(z100) / np.float32(100.0)]
    return Robot("panda", base, ctx)


def robot(name: str, ctx: Optional[Context] = None) -> Robot:
    """Robots by the reference's names: sphere, ur5, fetch, baxter, panda (= the fork's default
    PandaBase<200,200,0>, robots/panda_grid.hh:39-41), panda_X_Y (grid variants, metres)."""
    if name == "panda":
        return panda_grid(200, 200, 0, ctx)
    if name.startswith("panda_"):
        x, y = name.split("_")[1:3]
        return panda_grid(int(x) * 100, int(y) * 100, 0, ctx)
    return Robot(name, ctx=ctx)

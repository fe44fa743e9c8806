"""TEST INFRASTRUCTURE — NOT PRODUCT CODE.

ctypes front end of ``oracle/_ref/libvamp_ref.so`` — the unmodified mr-vamp reference
headers compiled for the host CPU (see ``oracle/ref/vref_api.h`` and ``oracle/Makefile``).
Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs import this.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_ref", "libvamp_ref.so")
_SO_CLEAR = os.path.join(_HERE, "_ref", "libvamp_ref_clear.so")  # instrumented primitives (oracle/clear/)
_lib = None
_clear = None

_f32p = C.POINTER(C.c_float)
_u8p = C.POINTER(C.c_uint8)


def build(force: bool = False) -> bool:
    """Compile the library when the reference checkout is present (CPU container only)."""
    src = os.environ.get("VAMP_SRC", "/root/reference")
    if not os.path.isdir(os.path.join(src, "src", "impl", "vamp")):
        return os.path.exists(_SO)
    args = ["make", "-C", _HERE, "-j8", f"VAMP_SRC={src}", "all"]
    if force:
        args.insert(1, "-B")
    subprocess.run(args, check=True, stdout=subprocess.DEVNULL)
    return True


def available() -> bool:
    return os.path.exists(_SO) and os.path.exists(_SO_CLEAR)


class CaptView(C.Structure):
    _fields_ = [
        ("n_tests", C.c_size_t),
        ("n_aff_starts", C.c_size_t),
        ("n_aff_vec", C.c_size_t),
        ("n_aabbs", C.c_size_t),
        ("nlog2", C.c_int),
        ("r_min", C.c_float),
        ("r_max", C.c_float),
        ("r_point", C.c_float),
        ("aabb_top", C.c_float * 6),
        ("tests", _f32p),
        ("aff_starts", C.POINTER(C.c_uint32)),
        ("aff_x", _f32p),
        ("aff_y", _f32p),
        ("aff_z", _f32p),
        ("aabbs", _f32p),
    ]


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            raise FileNotFoundError(f"{_SO} missing: run `make -C oracle` where the reference checkout is present")
        L = C.CDLL(_SO)
        L.vref_env_create.restype = C.c_void_p
        L.vref_env_destroy.argtypes = [C.c_void_p]
        L.vref_env_add_sphere.argtypes = [C.c_void_p] + [C.c_float] * 4
        L.vref_env_add_cuboid.argtypes = [C.c_void_p, _f32p]
        L.vref_env_add_capsule.argtypes = [C.c_void_p, _f32p]
        L.vref_env_add_heightfield.argtypes = [C.c_void_p, _f32p, C.c_size_t, C.c_size_t, _f32p]
        L.vref_env_add_pointcloud.argtypes = [C.c_void_p, _f32p, C.c_size_t, C.c_float, C.c_float, C.c_float]
        L.vref_env_add_pointcloud.restype = C.c_double
        L.vref_env_attach.argtypes = [C.c_void_p, _f32p, _f32p, C.c_size_t]
        L.vref_env_detach.argtypes = [C.c_void_p]
        L.vref_env_export.argtypes = [C.c_void_p, C.c_int, _f32p]
        L.vref_env_export.restype = C.c_size_t
        L.vref_env_capt.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(CaptView)]
        L.vref_capt_query.argtypes = [C.c_void_p, C.c_size_t, _f32p, C.c_size_t, _u8p, _u8p]
        L.vref_sphere_env.argtypes = [C.c_void_p, _f32p, C.c_size_t, _u8p]
        L.vref_trig.argtypes = [_f32p, C.c_size_t, _f32p, _f32p]
        L.vref_sqrt.argtypes = [_f32p, C.c_size_t, _f32p]
        L.vref_robot_info.argtypes = [C.c_char_p] + [C.POINTER(C.c_int)] * 3
        L.vref_sphere_fk.argtypes = [C.c_char_p, _f32p, C.c_size_t, _f32p]
        for fn in (L.vref_fkcc, L.vref_fkcc_packed, L.vref_fkcc_attach):
            fn.argtypes = [C.c_char_p, C.c_void_p, _f32p, C.c_size_t, _u8p, C.c_int]
        L.vref_validate_motion.argtypes = [C.c_char_p, C.c_void_p, _f32p, _f32p, C.c_size_t, _u8p, C.c_int]
        L.vref_filter_pointcloud.argtypes = [_f32p, C.c_size_t, C.c_float, C.c_float, _f32p, _f32p, _f32p, C.c_int, _f32p]
        L.vref_filter_pointcloud.restype = C.c_size_t
        L.vref_sphere_validity.argtypes = [C.c_char_p, C.c_void_p, _f32p, C.c_uint32, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
        L.vref_validate_vector.argtypes = [C.c_char_p, C.c_void_p, _f32p, _f32p, _f32p, C.c_size_t, _u8p, C.c_int]
        L.vref_env_float_ptr.argtypes = [C.c_void_p]
        L.vref_env_float_ptr.restype = C.c_void_p
        L.vref_validate.argtypes = [C.c_char_p, C.c_void_p, _f32p, C.c_size_t, _u8p]
        L.vref_scale.argtypes = [C.c_char_p, _f32p, C.c_size_t]
        L.vref_sphere_robot_set_radius.argtypes = [C.c_float]
        _lib = L
    return _lib


def clear_lib():
    """libvamp_ref_clear.so: the reference with instrumented primitives (min |signed clearance| per query)."""
    global _clear
    if _clear is None:
        if not os.path.exists(_SO_CLEAR):
            raise FileNotFoundError(f"{_SO_CLEAR} missing: run `make -C oracle` where the reference checkout is present")
        L = C.CDLL(_SO_CLEAR)
        L.vrefc_fkcc_clearance.argtypes = [C.c_char_p, C.c_void_p, _f32p, C.c_size_t, _u8p, _f32p, C.c_int]
        L.vrefc_motion_clearance.argtypes = [C.c_char_p, C.c_void_p, _f32p, _f32p, _f32p, C.c_size_t, _u8p, _f32p, C.c_int]
        _clear = L
    return _clear


def _f(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a, a.ctypes.data_as(_f32p)


def trig(x):
    x, px = _f(x)
    s = np.empty_like(x)
    c = np.empty_like(x)
    lib().vref_trig(px, x.size, s.ctypes.data_as(_f32p), c.ctypes.data_as(_f32p))
    return s, c


def approx_sqrt(x):
    x, px = _f(x)
    o = np.empty_like(x)
    lib().vref_sqrt(px, x.size, o.ctypes.data_as(_f32p))
    return o


def robot_info(robot: str):
    d, s, r = C.c_int(), C.c_int(), C.c_int()
    if lib().vref_robot_info(robot.encode(), C.byref(d), C.byref(s), C.byref(r)):
        raise KeyError(robot)
    return d.value, s.value, r.value


class Env:
    """The reference ``collision::Environment<float>`` built the way bindings/environment.cc does."""

    _KIND_W = {0: 5, 1: 9, 2: 9, 3: 16, 4: 16}

    def __init__(self):
        self.h = C.c_void_p(lib().vref_env_create())

    def __del__(self):
        try:
            if self.h:
                lib().vref_env_destroy(self.h)
                self.h = None
        except Exception:
            pass

    def add_sphere(self, x, y, z, r):
        lib().vref_env_add_sphere(self.h, x, y, z, r)

    def add_cuboid(self, c15):
        a, p = _f(c15)
        assert a.size == 15
        lib().vref_env_add_cuboid(self.h, p)

    def add_capsule(self, c8):
        a, p = _f(c8)
        assert a.size == 8
        lib().vref_env_add_capsule(self.h, p)

    def add_heightfield(self, off_scale6, data2d):
        o, po = _f(off_scale6)
        d, pd = _f(data2d)
        yd, xd = d.shape
        lib().vref_env_add_heightfield(self.h, po, xd, yd, pd)

    def add_pointcloud(self, xyz, r_min, r_max, r_point) -> float:
        a, p = _f(xyz)
        return lib().vref_env_add_pointcloud(self.h, p, a.shape[0], r_min, r_max, r_point)

    def attach(self, tf7, spheres_xyzr):
        t, pt = _f(tf7)
        s, ps = _f(spheres_xyzr)
        lib().vref_env_attach(self.h, pt, ps, s.shape[0])

    def detach(self):
        lib().vref_env_detach(self.h)

    def export(self, kind: int) -> np.ndarray:
        n = lib().vref_env_export(self.h, kind, None)
        out = np.zeros((n, self._KIND_W[kind]), dtype=np.float32)
        if n:
            lib().vref_env_export(self.h, kind, out.ctypes.data_as(_f32p))
        return out

    def capt(self, idx: int = 0) -> dict:
        v = CaptView()
        if lib().vref_env_capt(self.h, idx, C.byref(v)):
            raise IndexError(idx)
        arr = lambda p, n, dt=np.float32: np.ctypeslib.as_array(p, shape=(n,)).astype(dt, copy=True)
        return {
            "nlog2": v.nlog2,
            "r_min": v.r_min,
            "r_max": v.r_max,
            "r_point": v.r_point,
            "aabb_top": np.array(list(v.aabb_top), dtype=np.float32),
            "tests": arr(v.tests, v.n_tests),
            "aff_starts": arr(v.aff_starts, v.n_aff_starts, np.uint32),
            "aff_x": arr(v.aff_x, v.n_aff_vec * 8),
            "aff_y": arr(v.aff_y, v.n_aff_vec * 8),
            "aff_z": arr(v.aff_z, v.n_aff_vec * 8),
            "aabbs": arr(v.aabbs, v.n_aabbs * 6).reshape(-1, 6),
        }

    def capt_query(self, xyzr, idx: int = 0):
        a, p = _f(xyzr)
        n = a.shape[0]
        hs = np.zeros(n, dtype=np.uint8)
        hc = np.zeros(n, dtype=np.uint8)
        lib().vref_capt_query(self.h, idx, p, n, hs.ctypes.data_as(_u8p), hc.ctypes.data_as(_u8p))
        return hs.astype(bool), hc.astype(bool)

    def sphere_hits(self, xyzr):
        a, p = _f(xyzr)
        n = a.shape[0]
        h = np.zeros(n, dtype=np.uint8)
        lib().vref_sphere_env(self.h, p, n, h.ctypes.data_as(_u8p))
        return h.astype(bool)


def sphere_fk(robot: str, cfg) -> np.ndarray:
    a, p = _f(cfg)
    dim, ns, _ = robot_info(robot)
    a = a.reshape(-1, dim)
    out = np.empty((a.shape[0], ns, 4), dtype=np.float32)
    lib().vref_sphere_fk(robot.encode(), p, a.shape[0], out.ctypes.data_as(_f32p))
    return out


def _verdicts(fn, robot, env, cfg, threads, packed=False):
    a, p = _f(cfg)
    dim, _, _ = robot_info(robot)
    a = a.reshape(-1, dim)
    n = a.shape[0]
    out = np.zeros(n // 8 if packed else n, dtype=np.uint8)
    rc = fn(robot.encode(), env.h, p, n, out.ctypes.data_as(_u8p), threads)
    if rc:
        raise RuntimeError(f"reference call failed ({rc})")
    return out.astype(bool)


def fkcc(robot, env, cfg, threads=1):
    return _verdicts(lib().vref_fkcc, robot, env, cfg, threads)


def fkcc_packed(robot, env, cfg, threads=1):
    return _verdicts(lib().vref_fkcc_packed, robot, env, cfg, threads, packed=True)


def fkcc_attach(robot, env, cfg, threads=1):
    return _verdicts(lib().vref_fkcc_attach, robot, env, cfg, threads)


def validate_motion(robot, env, a, b, threads=1):
    a, pa = _f(a)
    b, pb = _f(b)
    dim, _, _ = robot_info(robot)
    n = a.reshape(-1, dim).shape[0]
    out = np.zeros(n, dtype=np.uint8)
    rc = lib().vref_validate_motion(robot.encode(), env.h, pa, pb, n, out.ctypes.data_as(_u8p), threads)
    if rc:
        raise RuntimeError(f"reference call failed ({rc})")
    return out.astype(bool)


def filter_pointcloud(xyz, min_dist, max_range, origin, ws_min, ws_max, cull=True) -> np.ndarray:
    """collision::filter_pointcloud (collision/filter.hh:176-275), as src/vamp/pointcloud.py:150-160 calls it."""
    a, p = _f(xyz)
    o, po = _f(origin)
    lo, plo = _f(ws_min)
    hi, phi = _f(ws_max)
    out = np.zeros_like(a)
    n = lib().vref_filter_pointcloud(p, a.shape[0], min_dist, max_range, po, plo, phi, int(cull), out.ctypes.data_as(_f32p))
    return out[:n].copy()


def sphere_validity(robot, env, cfg, cap=64):
    """Helper::sphere_validate of one configuration -> per robot sphere the insertion indices of the obstacles it overlaps."""
    a, p = _f(cfg)
    _, ns, _ = robot_info(robot)
    hits = np.zeros((ns, cap), dtype=np.uint32)
    counts = np.zeros(ns, dtype=np.uint32)
    rc = lib().vref_sphere_validity(robot.encode(), env.h, p, cap, hits.ctypes.data_as(C.POINTER(C.c_uint32)),
                                    counts.ctypes.data_as(C.POINTER(C.c_uint32)))
    if rc:
        raise RuntimeError(f"reference call failed ({rc})")
    return [sorted(int(o) for o in hits[k, : counts[k]]) for k in range(ns)]


def validate_vector(robot, env, a, v, dist, threads=1):
    """planning::validate_vector(start, vector, distance, env) per row."""
    a, pa = _f(a)
    v, pv = _f(v)
    d, pd = _f(dist)
    dim, _, _ = robot_info(robot)
    n = a.reshape(-1, dim).shape[0]
    out = np.zeros(n, dtype=np.uint8)
    rc = lib().vref_validate_vector(robot.encode(), env.h, pa, pv, pd, n, out.ctypes.data_as(_u8p), threads)
    if rc:
        raise RuntimeError(f"reference call failed ({rc})")
    return out.astype(bool)


def fkcc_clearance(robot, env, cfg, threads=1):
    """Reference verdict per configuration AND the smallest |signed clearance| (metres) over every collision test the
    reference executed for it (fkcc_attach when the environment carries an attachment)."""
    a, p = _f(cfg)
    dim, _, _ = robot_info(robot)
    n = a.reshape(-1, dim).shape[0]
    ok = np.zeros(n, dtype=np.uint8)
    clr = np.zeros(n, dtype=np.float32)
    rc = clear_lib().vrefc_fkcc_clearance(robot.encode(), lib().vref_env_float_ptr(env.h), p, n, ok.ctypes.data_as(_u8p),
                                          clr.ctypes.data_as(_f32p), threads)
    if rc:
        raise RuntimeError(f"clearance oracle call failed ({rc})")
    return ok.astype(bool), clr


def motion_clearance(robot, env, a, second, dist=None, threads=1):
    """validate_motion (dist None: second = goals) or validate_vector (second = vectors): verdict and min |clearance|."""
    a, pa = _f(a)
    b, pb = _f(second)
    dim, _, _ = robot_info(robot)
    n = a.reshape(-1, dim).shape[0]
    pd = None
    if dist is not None:
        d, pd = _f(dist)
    ok = np.zeros(n, dtype=np.uint8)
    clr = np.zeros(n, dtype=np.float32)
    rc = clear_lib().vrefc_motion_clearance(robot.encode(), lib().vref_env_float_ptr(env.h), pa, pb, pd, n,
                                            ok.ctypes.data_as(_u8p), clr.ctypes.data_as(_f32p), threads)
    if rc:
        raise RuntimeError(f"clearance oracle call failed ({rc})")
    return ok.astype(bool), clr


def validate(robot, env, cfg):
    a, p = _f(cfg)
    dim, _, _ = robot_info(robot)
    n = a.reshape(-1, dim).shape[0]
    out = np.zeros(n, dtype=np.uint8)
    lib().vref_validate(robot.encode(), env.h, p, n, out.ctypes.data_as(_u8p))
    return out.astype(bool)


def scale(robot, unit_cfg):
    a = np.array(unit_cfg, dtype=np.float32, order="C", copy=True)
    dim, _, _ = robot_info(robot)
    lib().vref_scale(robot.encode(), a.ctypes.data_as(_f32p), a.reshape(-1, dim).shape[0])
    return a

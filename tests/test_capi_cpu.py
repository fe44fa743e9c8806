"""CPU tests of the C-ABI shared library (no GPU): it loads, exports every symbol include/vamp_b200.h
declares, its host-side logic (robot metadata, synthetic streams, environment classification / sorting /
min_distance) matches the reference, and every query fails loudly without a CUDA device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from mr_vamp_b200 import api, scenes
from conftest import apply_ref

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "vamp_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(vb_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = C.CDLL(api.LIB_PATH)
    names = _header_symbols()
    assert len(names) >= 45
    for name in names:
        assert hasattr(lib, name), f"{name} declared in include/vamp_b200.h but not exported"
    # and the ctypes layer binds exactly the declared set
    assert sorted(api.exported_symbols()) == names


def test_no_product_path_imports_the_oracle():
    # the product package must never import, link or execute anything under oracle/
    for dirpath, _, files in os.walk(os.path.join(ROOT, "mr_vamp_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")) and "_gen" not in dirpath:
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "from oracle" not in text and "libvamp_ref" not in text, f


def test_robot_metadata_matches_reference(ref):
    # robots/{sphere,ur5,panda_base,fetch,baxter}.hh: dimension, n_spheres, resolution
    for name in ["sphere", "ur5", "panda", "fetch", "baxter"]:
        d, s, res = C.c_int(), C.c_int(), C.c_int()
        api._check(api.lib().vb_robot_info(api.ROBOT_IDS[name], C.byref(d), C.byref(s), C.byref(res)))
        rd, rs, rr = ref.robot_info("panda_0_0" if name == "panda" else name)
        assert (d.value, s.value, res.value) == (rd, rs, rr), name


@pytest.mark.parametrize("name", ["ur5", "panda", "fetch", "baxter"])
def test_limits_match_scale_configuration(ref, name):
    # scale_configuration: q = lo + u * range (robots/<robot>/fk.hh)
    rid = api.ROBOT_IDS[name]
    dim = ref.robot_info("panda_0_0" if name == "panda" else name)[0]
    lo = np.zeros(dim, np.float32)
    rg = np.zeros(dim, np.float32)
    api._check(api.lib().vb_robot_limits(rid, api._ptr(lo), api._ptr(rg)))
    rn = "panda_0_0" if name == "panda" else name
    assert np.allclose(ref.scale(rn, np.zeros((1, dim), np.float32))[0], lo, atol=1e-6)
    assert np.allclose(ref.scale(rn, np.ones((1, dim), np.float32))[0], lo + rg, atol=1e-5)


def test_random_stream_is_reproducible_and_in_limits():
    rid = api.ROBOT_IDS["panda"]
    lo = np.zeros(7, np.float32)
    rg = np.zeros(7, np.float32)
    api._check(api.lib().vb_robot_limits(rid, api._ptr(lo), api._ptr(rg)))
    q = np.zeros((64, 7), np.float32)
    q2 = np.zeros((64, 7), np.float32)
    for i in range(64):
        api._check(api.lib().vb_random_config(rid, 42, i, api._ptr(q[i])))
        api._check(api.lib().vb_random_config(rid, 42, i, api._ptr(q2[i])))
    assert np.array_equal(q, q2) and len({tuple(r) for r in q}) == 64
    assert (q >= lo - 1e-6).all() and (q <= lo + rg + 1e-5).all()
    a = np.zeros(7, np.float32)
    b = np.zeros(7, np.float32)
    api._check(api.lib().vb_random_edge(rid, 42, 3, 0.05, 1.0, api._ptr(a), api._ptr(b)))
    assert 0.0 < np.linalg.norm(b - a) <= 1.0 + 1e-5


def test_environment_host_logic_matches_reference(ref):
    """z-aligned classification (bindings/environment.cc:120-141), min_distance (shapes.hh:57-73,166-192)
    and the per-list sort (environment.hh:40-66): vb_env_export against the reference's own lists."""
    mbm = scenes.load_mbm(os.path.join(scenes.GOLDEN_DIR, "mbm_panda.json.gz"))
    specs = [scenes.random_scene(7), scenes.random_scene(11), scenes.cage(),
             scenes.problem_to_spec(mbm["problems"]["table_pick"][0]), scenes.problem_to_spec(mbm["problems"]["cage"][3])]
    for spec in specs:
        env = scenes.apply(spec, api.Environment(host_only=True))
        renv = apply_ref(spec, ref.Env())
        for kind in range(5):
            got, want = env.export(kind), renv.export(kind)
            assert got.shape == want.shape, (kind, got.shape, want.shape)
            assert np.allclose(got, want, rtol=0, atol=1e-6), kind
    # table_pick scene 1: 10 z-aligned cuboids + 2 z-aligned capsules
    env = scenes.apply(specs[3], api.Environment(host_only=True))
    assert [env.export(k).shape[0] for k in range(5)] == [0, 0, 2, 0, 10]


def test_queries_fail_loudly_without_a_device():
    import torch

    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    h = api._vp()
    rc = api.lib().vb_ctx_create(C.byref(h), None, 0)
    assert rc == 2, "vb_ctx_create must return VB_ERR_CUDA when no device is usable"
    assert api.lib().vb_last_error()
    with pytest.raises(api.VampB200Error):
        api.Robot("panda").fkcc_batch(np.zeros((1, 7), np.float32))
    env = api.Environment(host_only=True)
    with pytest.raises(api.VampB200Error):
        _ = env.handle  # commit needs a device context


def test_compiled_python_module_loads_and_exposes_the_reference_surface():
    """mr_vamp_b200._core_ext (pybind11): importable without a device, robot metadata from the library, shape math equal
    to the ctypes layer's; the reference's names for this path are all there (bindings/common.hh:685-710)."""
    import mr_vamp_b200._core_ext as m

    for sub in ("sphere", "ur5", "panda", "fetch", "baxter", "panda_0_0", "panda_1_0", "panda_2_2"):
        mod = getattr(m, sub)
        for fn in ("validate", "fk", "sphere_validity", "validate_batch", "validate_motion_batch", "validate_vector_batch",
                   "fk_batch", "fkcc_batch", "dimension", "resolution", "n_spheres"):
            assert hasattr(mod, fn), (sub, fn)
    for cls in ("Environment", "Sphere", "Cuboid", "Cylinder", "HeightField", "Attachment"):
        assert hasattr(m, cls)
    assert (m.baxter.dimension(), m.baxter.n_spheres(), m.fetch.dimension()) == (14, 75, 8)
    rng = np.random.default_rng(0)
    for _ in range(20):
        c, e, h = rng.uniform(-1, 1, 3), rng.uniform(-3, 3, 3), rng.uniform(0.05, 0.4, 3)
        # (three float32 matrix products: the last bit depends on the summation order)
        assert np.allclose(np.asarray(m.Cuboid(c, e, h).raw(), np.float32), api.Cuboid(c, e, h).raw(), rtol=0, atol=1e-6)
        yaw = [0.0, 0.0, e[2]]
        assert np.array_equal(np.asarray(m.Cuboid(c, yaw, h).raw(), np.float32), api.Cuboid(c, yaw, h).raw())
        assert np.allclose(np.asarray(m.Cylinder(c, e, 0.05, 0.3).raw(), np.float32), api.Cylinder(c, e, 0.05, 0.3).raw(), rtol=0, atol=1e-6)
    assert np.asarray(m.Cuboid([0, 0, 0], [0, 0, 1.234], [1, 1, 1]).raw())[11] == 1.0  # yaw only: z-aligned

import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


BAND = 1e-5  # metres (BASELINE.json north_star): verdicts may differ only where the decisive clearance is within it
_PARITY_RECORDS = []


def judge(what, ok_gpu, ok_ref, clr_ref, margin_gpu=None):
    """Parity of GPU verdicts with the reference's.  The band is judged by the ORACLE: clr_ref is the smallest
    |signed clearance| over every test the reference executed for the query (oracle/clear/, ref.fkcc_clearance /
    ref.motion_clearance).  margin_gpu (the library's own margin output) is only cross-checked against it.
    Every call is recorded; the records are written to gpurun_out/parity_records.json at the end of the session."""
    ok_gpu, ok_ref, clr_ref = np.asarray(ok_gpu, bool), np.asarray(ok_ref, bool), np.asarray(clr_ref, np.float32)
    diff = ok_gpu != ok_ref
    hard = diff & (clr_ref > BAND)
    rec = {"what": what, "n": int(ok_ref.size), "valid_ref": float(ok_ref.mean()) if ok_ref.size else 0.0,
           "differ": int(diff.sum()), "within_band": int((diff & ~hard).sum()), "mismatch_hard": int(hard.sum()),
           "ref_queries_within_band": int((clr_ref <= BAND).sum())}
    if margin_gpu is not None:
        mg = np.asarray(margin_gpu, np.float32)
        # same tests, same order, same early exits on both sides: the two figures are the same quantity, except where
        # the reference's approximate-sqrt pruning (validity.hh:59) skips an obstacle the exact pruning
        # visits or the other way round (rare; seen 1 in 176 at base (2, 2), 5e-4 m) -- hence a 1 % allowance
        both = ok_gpu & ok_ref
        gap = np.abs(mg - clr_ref)[both]
        rec["margin_cross_check"] = {"compared": int(both.sum()), "max_abs_diff": float(gap.max()) if gap.size else 0.0,
                                     "over_1e-5": int((gap > BAND).sum())}
    _PARITY_RECORDS.append(rec)
    print(f"{what}: n={rec['n']} valid_ref={rec['valid_ref']:.3f} differ={rec['differ']} within_band={rec['within_band']} "
          f"hard={rec['mismatch_hard']}" + (f" margin|gpu-oracle|max={rec['margin_cross_check']['max_abs_diff']:.2e}"
                                           if margin_gpu is not None else ""))
    assert rec["mismatch_hard"] == 0, f"{what}: {rec['mismatch_hard']} verdict mismatches outside the {BAND} m band (oracle clearance)"
    if margin_gpu is not None:
        assert rec["margin_cross_check"]["over_1e-5"] <= 0.01 * max(rec["margin_cross_check"]["compared"], 100), \
            f"{what}: the library's margin differs from the oracle's clearance by more than {BAND} m on " \
            f"{rec['margin_cross_check']['over_1e-5']} queries (max {rec['margin_cross_check']['max_abs_diff']:.3e})"
    return rec


def pytest_sessionfinish(session, exitstatus):
    if _PARITY_RECORDS:
        import json

        out = os.path.join(ROOT, "gpurun_out")
        os.makedirs(out, exist_ok=True)
        with open(os.path.join(out, "parity_records.json"), "w") as f:
            json.dump({"band_m": BAND, "judge": "oracle clearance (oracle/clear)", "records": _PARITY_RECORDS}, f, indent=1)


def apply_ref(spec, renv):
    """Apply a scene spec (mr_vamp_b200.scenes) to the reference oracle environment."""
    for item in spec:
        kind = item[0]
        if kind == "sphere":
            renv.add_sphere(*[float(np.float32(v)) for v in item[1:5]])
        elif kind == "cuboid":
            renv.add_cuboid(item[1])
        elif kind == "capsule":
            renv.add_capsule(item[1])
        elif kind == "heightfield":
            renv.add_heightfield(item[1], item[2])
        else:
            raise ValueError(kind)
    return renv


@pytest.fixture(scope="session")
def ref():
    from oracle import ref as r

    if not r.available():
        r.build()
    if not r.available():
        pytest.skip("oracle/_ref/libvamp_ref.so not built (needs the reference checkout)")
    return r


@pytest.fixture(scope="session")
def vb():
    import mr_vamp_b200.api as api

    api.lib()
    return api


@pytest.fixture(scope="session")
def gpu_ctx(vb):
    return vb.default_context()

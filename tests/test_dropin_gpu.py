"""The C++ drop-in layer (include/vamp_b200_dropin.hh): a program written against mr-vamp's own types
(collision::Environment<float>, Robot::ConfigurationBlock<8>, planning::validate_motion) runs every
query twice -- through the reference's CPU code and through the B200 library -- and compares.
The binary is built by __graft_entry__.build() where the reference checkout is present."""
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_program_through_the_dropin_layer():
    exe = os.path.join(ROOT, "build", "dropin_check")
    if not os.path.exists(exe):
        pytest.skip("build/dropin_check not built (needs the reference checkout at build time)")
    out = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    print(out.stdout, out.stderr)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "dropin_check: OK" in out.stdout


def test_reference_planner_code_runs_on_the_device_unchanged():
    """include/vamp_b200_planners.hh: the reference's Path::validate (planning/plan.hh:155-168) and shortcut_path
    (planning/simplify.hh:115-140), compiled unchanged, call the device through redirected validate_motion /
    validate_vector; verdicts and simplified paths equal the reference's CPU functions'."""
    exe = os.path.join(ROOT, "build", "planners_check")
    if not os.path.exists(exe):
        pytest.skip("build/planners_check not built (needs the reference checkout at build time)")
    out = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    print(out.stdout, out.stderr)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "planners_check: OK" in out.stdout

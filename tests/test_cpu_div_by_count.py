"""Phys::div_by_count (the sub-step division of the sedimentation kernels through a table of correctly rounded
reciprocals, scream_b200/csrc/p3_physics.cuh) against the IEEE division it replaces: bit for bit, for every count the
table holds, on time-step-like operands, arbitrary exponents in the admitted range and exact / nearly halfway
quotients.  The GPU side is covered by the all-columns parity tests (sub-step counts and fields bit-identical to the
reference); this pins the arithmetic identity itself, on the CPU, with the same fma sequence."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_fma_division_by_a_small_count_is_the_ieee_quotient(tmp_path):
    exe = str(tmp_path / "div_by_count_check")
    subprocess.check_call(["gcc", "-O2", "-mfma", "-ffp-contract=off", "-o", exe,
                           os.path.join(ROOT, "tests", "cpp", "div_by_count_check.c"), "-lm"])
    out = subprocess.run([exe, "1023", "30000"], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout[-2000:]
    checked, bad = (int(x) for x in out.stdout.split()[-3::2])
    assert checked > 2.5e7 and bad == 0, out.stdout[-500:]

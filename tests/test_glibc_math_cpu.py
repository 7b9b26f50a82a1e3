"""The device's sin / cos / pow are glibc's float routines restated (csrc/spk_glibc_math.cuh) from tables read out of the
image's libm. On the CPU: (1) the committed table header is what tools/gen_glibc_tables.py reads out of THIS libm (a new
glibc would show up here, not as a parity drift), (2) the host twin of the restatement (tools/glibc_math_probe.cpp, the same
operations with the same fused multiply-adds) equals libm's sinf / cosf / powf on every one of 2e6 random arguments each.
The device functions themselves are checked against libm by tests/test_libm_gpu.py."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBM = "/lib/x86_64-linux-gnu/libm.so.6"


@pytest.mark.skipif(not os.path.exists(LIBM), reason="no x86-64 glibc libm at the usual place")
def test_committed_tables_are_this_libms(tmp_path):
    out = tmp_path / "tables.h"
    subprocess.run([sys.executable, os.path.join(ROOT, "tools", "gen_glibc_tables.py"), LIBM, str(out)], check=True,
                   capture_output=True)
    want = open(os.path.join(ROOT, "sparkle_b200", "csrc", "spk_glibc_tables.h")).read()
    assert out.read_text() == want


@pytest.mark.skipif("fma" not in open("/proc/cpuinfo").read(), reason="libm runs its non-FMA build on this CPU")
def test_host_twin_of_the_device_routines_equals_libm(tmp_path):
    exe = tmp_path / "probe"
    subprocess.run(["g++", "-O2", "-ffp-contract=off", "-mfma", "-DFMA=1", "-o", str(exe),
                    os.path.join(ROOT, "tools", "glibc_math_probe.cpp"), "-lm"], check=True)
    r = subprocess.run([str(exe), "2000000"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout
    assert "sin(0..2pi) 0  cos(0..2pi) 0  sincos(+-100) 0  pow(x,2.4) 0  pow(x,0.45) 0 differ" in r.stdout, r.stdout

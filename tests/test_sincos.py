"""CPU: the libm restatement the device uses for Vector3::random_direction (math/vector.rs:36-45) —
rust-raytracing_b200/csrc/rtx_sincos.cuh — must return the bits of this machine's glibc sin/cos (the functions
the oracle and, through Rust's f64::sin/cos, the reference call).  The header is host+device; here its host
build is compared with the live libm on millions of arguments per range, and the committed table is compared
with the one inside libm.so."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_table_matches_libm_and_first_principles():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "gen_sincostab.py"), "--check"],
                         capture_output=True, text=True, timeout=300)
    print(out.stderr[-1500:])
    assert out.returncode == 0, out.stderr[-1500:]
    assert "heads bit-identical to the computed ones: 220/220" in out.stderr


def test_host_build_is_bit_identical_to_libm(libm_check):
    code, text = libm_check
    print(text)
    if code != 0 and "fma" not in open("/proc/cpuinfo").read():
        pytest.skip("this CPU has no FMA: glibc selected a non-FMA sin/cos variant, which rounds differently")
    assert code == 0, text

"""fmt_g6 (csrc/hf2d_fmt.cuh): the "%g" the device-side Tecplot formatter prints with -- the default `ostream << double` of the
reference's SaveData2D (deeps2d_core.cpp:2589-2673) -- compiled for the HOST with g++ and compared with glibc printf on
millions of values incl. every kind of tie (the same header compiles for the device)."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_fmt_g6_equals_printf(tmp_path):
    exe = str(tmp_path / "fmt_g6_check")
    src = os.path.join(ROOT, "tests", "native", "fmt_g6_check.cpp")
    inc = os.path.join(ROOT, "openhyperflow2d_b200", "csrc")
    subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-w", "-I", inc, "-o", exe, src])
    r = subprocess.run([exe], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:]
    assert r.stdout.strip().splitlines()[-1].startswith("0 mismatches of ")

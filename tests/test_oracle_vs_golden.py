"""Pins the CPU oracle (oracle/deeps2d_oracle.c) against the reference's own output: for every golden case the
oracle is started from the reference's pre-step AoS image and must reproduce the reference's post-step image
BIT-EXACTLY in every field the step owns (both were compiled with -O2 -ffp-contract=off, same libm)."""
import numpy as np
import pytest

import oracle_py as O
from golden_util import STEP_FIELDS, bc_case_names, case_names, geometry_case_names, load_case, r2_case_names, restart_case_names


@pytest.mark.parametrize("name", case_names() + geometry_case_names() + bc_case_names() + restart_case_names() + r2_case_names())
def test_oracle_matches_reference(name):
    c = load_case(name)
    cfg = c["cfg"]
    nodes = O.nodes_from_ref(c["state0"])
    dt_min, last_iter, t = cfg["dt"], cfg["last_iter"], cfg.get("GlobalTime", 0.0)
    for _ in range(c["ncycles"]):
        rc, dt_min, rep, tp = O.run_cycle(nodes, cfg, last_iter, c["nsteps"], dt_min, t)
        assert rc == 0
        last_iter += c["nsteps"]
        t += tp
    assert t == c["global_time"]
    assert dt_min == c["dt_final"]
    ref = c["stateN"]
    for f in STEP_FIELDS:
        a, b = nodes[f], ref[f]
        same = (a == b) | ((a != a) & (b != b))
        assert same.all(), f"{name}: field {f} differs at {int((~same).sum())} entries"
    # the reference's RMS file holds sum_RMS of the last step (6 significant digits)
    last = c["rms_file"].strip().splitlines()[-1].split()
    got = [float(f"{v:.6g}") for v in rep.sum_RMS]
    want = [float(v) for v in last[1:1 + len(got)]]
    assert np.allclose(got[:len(want)], want, rtol=2e-5, atol=0), (got, want)


def test_table_getval_edges():
    import ctypes as C
    x = (C.c_double * 3)(3.0, 2.0, 1.0)   # descending table: always extrapolates from rows 0-1
    y = (C.c_double * 3)(30.0, 20.0, 10.0)
    t = O.Table(3, C.cast(x, C.POINTER(C.c_double)), C.cast(y, C.POINTER(C.c_double)))
    L = O.lib()
    assert L.ohf_table_getval(C.byref(t), 2.5) == 25.0
    assert L.ohf_table_getval(C.byref(t), 0.0) == 0.0
    t1 = O.Table(1, C.cast(x, C.POINTER(C.c_double)), C.cast(y, C.POINTER(C.c_double)))
    assert L.ohf_table_getval(C.byref(t1), 99.0) == 30.0
    t0 = O.Table(0, None, None)
    assert L.ohf_table_getval(C.byref(t0), 1.0) == 0.0

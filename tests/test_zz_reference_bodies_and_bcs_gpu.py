"""GPU parity on bodies drawn by the reference's OWN rasterisers (run with -m gpu on the B200 box; sorted last on purpose).

tests/golden/geometry/*.npz were written by the unmodified reference (tests/golden/make_golden.py, geometry_cases()):
three cylinders (BoundCircle2D; BASELINE config 4 in miniature), a NACA 2412 airfoil (SolidBoundAirfoil2D; config 3), a
solid rectangle (SolidBoundRect2D).  Their staircase no-slip contours hold every corner type the rasterisers emit --
wall nodes with two solid neighbours, one-cell notches, diagonal steps -- which the straight ramp of the other golden
cases and the synthetic grids of gridgen.py do not all contain.  The CPU oracle reproduces all four bit for bit
(tests/test_oracle_vs_golden.py); the CUDA path is held to the same bar as in test_gpu_parity.py: bit-exact for the
laminar cases, rel-L_inf <= 1e-10 where the k-eps wall rewrite calls pow().

tests/golden/bc/*.npz (bc_cases()) do the same for boundary-condition branches of the step that the other cases do not
reach: soft (second-derivative) outflow / far boundaries, the wall-law wall, the second-derivative eps wall rewrites, the
integral turbulence model, single Neumann / Dirichlet flags instead of the macro node types.

tests/golden/restart/*.npz (restart_cases()) are runs RESTARTED from the swap file of an earlier run: the reference's
start-up then derives y_plus from real wall gradients (it is 0 everywhere in a fresh run), so these are the cases in
which the y+ damping functions (van Driest exp, Chien / JL / LSY) see non-trivial input."""
import numpy as np
import pytest

from golden_util import LAYOUT, bc_case_names, geometry_case_names, load_case, restart_case_names
from test_gpu_parity import check_against_reference

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", geometry_case_names())
def test_gpu_matches_reference_on_rasterised_bodies(name):
    assert geometry_case_names(), "tests/golden/geometry is empty"
    check_against_reference(name, exact="keps" not in name)


@pytest.mark.parametrize("name", restart_case_names())
def test_gpu_matches_reference_after_restart(name):
    c = load_case(name)
    assert float(np.abs(c["state0"]["y_plus"]).max()) > 1.0        # the point of these cases
    check_against_reference(name, exact="laminar" in name)


@pytest.mark.parametrize("name", bc_case_names())
def test_gpu_matches_reference_on_bc_branches(name):
    check_against_reference(name, exact=not any(t in name for t in ("keps", "ns_sa")))     # pow() / exp() closures


def assert_rms_close(got, want, cfg, S_before, S_after):
    """Residual norms of one step, GPU vs oracle.  With isAlternateRMS = 1 the reference's serial build sums SIGNED
    differences, RMS[k] = sqrt(sum(S_next - S) / sum(S^2)) (deeps2d_core.cpp:1138-1143): the numerator cancels, so a sum in
    another order (atomics vs the sweep) is only accurate relative to sum|S_next - S|, not to its own value -- and when it
    comes out <= 0 the norm is reported as 0.  The tolerance is therefore set on the numerator: 1e-14 * sum|dS| (~100 ulp; double
    rounding over <= 1e5 terms stays far below that), on top of 1e-9 relative."""
    got, want = np.asarray(list(got)), np.asarray(list(want))
    if not cfg["isAlternateRMS"]:
        np.testing.assert_allclose(got, want, rtol=1e-9, atol=1e-300)
        return
    for k in range(len(want)):
        if np.isnan(want[k]) or np.isnan(got[k]):      # (NaN states: the Chien restart case)
            assert np.isnan(want[k]) and np.isnan(got[k]), (k, got[k], want[k])
            continue
        den = float(np.nansum(S_before[..., k] ** 2))
        slack = 1e-14 * float(np.nansum(np.abs(S_after[..., k] - S_before[..., k]))) / den if den > 0 else 0.0
        assert abs(got[k] ** 2 - want[k] ** 2) <= 1e-9 * want[k] ** 2 + slack, (k, got[k], want[k], slack)


@pytest.mark.parametrize("name", ["geometry/cylinders_keps_n8", "geometry/airfoil_keps_n8", "bc/euler_soft_n8", "bc/ns_keps_soft_mud2k_n8",
                                  "restart/ns_keps_chien_r12_n6", "restart/ns_prandtl_vd_r12_n6"])
def test_gpu_matches_oracle_stepwise_on_rasterised_bodies(name):
    """dt and the residual norms agree with the oracle after every single step (not only at the end of the cycle)."""
    import oracle_py as O
    from openhyperflow2d_b200 import cabi
    c = load_case(name)
    cfg = c["cfg"]
    nodes = O.nodes_from_ref(c["state0"])
    J = np.array(c["state0"], copy=True)
    s = cabi.Solver(cfg, LAYOUT)
    s.upload_aos(J)
    dt_min = cfg["dt"]
    for it in range(c["nsteps"]):
        S_before = np.array(nodes["S"], dtype=float, copy=True)
        rc, dt_min, orep = O.step(nodes, cfg, cfg["last_iter"] + it, 0, dt_min)
        assert rc == 0
        grep = s.step(1)
        assert grep.dt == orep.dt
        assert abs(grep.dt_min - orep.dt_min) <= 1e-13 * orep.dt_min
        assert_rms_close(grep.sum_RMS, orep.sum_RMS, cfg, S_before, np.asarray(nodes["S"], float))
        np.testing.assert_allclose(list(grep.DD_max), list(orep.DD_max), rtol=1e-9, atol=1e-300)
    s.close()


GRAD_MEMBERS = ["droYdx", "droYdy", "dUdx", "dUdy", "dVdx", "dVdy", "dTdx", "dTdy", "dkdx", "dkdy", "depsdx", "depsdy"]


@pytest.mark.parametrize("name", ["ns_laminar_n8", "ns_keps_std_n8", "ns_keps_std_n4x2", "geometry/cylinders_keps_n8",
                                  "geometry/airfoil_keps_n8", "geometry/rect_laminar_n8", "restart/ns_keps_chien_r12_n6",
                                  "bc/ns_sa_subsonic_n4", "bc/ns_laminar_mixed_flags_n8", "bc/ns_keps_wall_law_n2"])
def test_gradient_members_come_back_for_a_restart(name):
    """The 18 gradient members of the records are intermediates of a step, but a run restarted from the swap file re-runs
    FillNode2D(0,1) on every node with them (InitDEEPS2D, deeps2d_core.cpp:4565) and Recalc_y_plus reads the wall nodes'
    dUdy, dVdx (:2364-2386).  The last step of a step call keeps them (hf2d_gpu_keep_gradients, default on) and
    hf2d_gpu_download_aos writes exactly the members the reference assigns: the whole image must then agree with the
    reference's in every gradient member, on every node."""
    from test_gpu_parity import REL_TOL, rel_linf, run_gpu
    c = load_case(name)
    J, _, _ = run_gpu(c)
    ref = c["stateN"]
    ct = np.ascontiguousarray(ref["CT"], dtype=np.uint64)
    wall = ((ct & 0x040000000) == 0) & (((ct & 0x04000000) != 0) | ((ct & 0x08000000) != 0))
    assert wall.sum() > 0
    assert np.nanmax(np.abs(ref["dUdy"][wall])) > 0, name
    for f in GRAD_MEMBERS:
        a, b = np.asarray(J[f], float), np.asarray(ref[f], float)
        if "laminar" in name:
            same = (a == b) | ((a != a) & (b != b))
            assert same.all(), (name, f, int((~same).sum()))
        else:       # (the Chien restart case holds NaN where the reference itself produces them: same entries required)
            na, nb = np.isnan(a), np.isnan(b)
            assert (na == nb).all(), (name, f, int((na != nb).sum()))
            a, b = np.where(nb, 0.0, a), np.where(nb, 0.0, b)
            assert rel_linf(a, b) <= REL_TOL, (name, f, rel_linf(a, b))


@pytest.mark.parametrize("case,exact,gpu_startup", [("ns_laminar", True, False), ("ns_prandtl_vd", False, False),
                                                    ("ns_laminar", True, True), ("ns_prandtl_vd", False, True),
                                                    ("ns_keps_std", False, False), ("ns_keps_jl", False, False)])
def test_restart_chain_through_the_reference_program(case, exact, gpu_startup, tmp_path, monkeypatch):
    """Swap-file checkpoint / restart, end to end: the reference program with its time loop on the GPU
    (oracle/_ref/ohf_gpu) runs 8 steps, leaves its swap file, is started AGAIN from it and runs 6 more; the unmodified
    program does the same.  The second start-up recomputes y_plus from the wall gradients of the first run's image
    (Recalc_y_plus) -- so y_plus, which the second run's damping functions read, agrees only if the GPU run's swap file
    carries the right dUdy / dVdx on the wall nodes (van Driest: the mixing length is a function of y_plus).
    The k-eps decks set isTurbulenceReset = 1: their restart re-runs FillNode2D(0,1) on every node with the gradient
    members of the records (ScanArea, deeps2d_core.cpp:2186-2205) -- all 18 of them come back with the download since
    the last step of every step call writes them (hf2d_gpu_keep_gradients, the default)."""
    import os
    import subprocess
    import test_shim_dropin as D
    from openhyperflow2d_b200 import aos
    if not (os.path.exists(D.REF) and os.path.exists(D.GPU)):
        pytest.skip("oracle/_ref not built (make -C oracle ref)")

    def chain(binary, wd):
        os.makedirs(wd)
        outs = []
        for nmax in (8, 6):
            deck = D._deck(case, nmax, 1, 1e-30)            # exit after one output cycle
            with open(os.path.join(wd, case + ".dat"), "w") as f:
                f.write(deck)
            r = subprocess.run([binary, case + ".dat"], cwd=wd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
            assert r.returncode == 0, r.stdout[-3000:]
            outs.append(r.stdout)
        assert "Use old 2D-swap file" in outs[1] and "Use old 2D-swap file" not in outs[0]
        nx = int([l for l in deck.splitlines() if l.startswith("<data/MaxX=")][0].split("=")[1][:-1])
        ny = int([l for l in deck.splitlines() if l.startswith("<data/MaxY=")][0].split("=")[1][:-1])
        return np.fromfile(os.path.join(wd, case + ".hf2d"), dtype=aos.ref_dtype(LAYOUT)).reshape(nx, ny), outs

    Jr, _ = chain(D.REF, str(tmp_path / "ref"))
    if gpu_startup:     # wall distance and y+ of BOTH start-ups on the GPU as well (HF2D_GPU_WALL_DISTANCE)
        monkeypatch.setenv("HF2D_GPU_WALL_DISTANCE", "1")
    Jg, outs = chain(D.GPU, str(tmp_path / "gpu"))
    if gpu_startup:     # second start-up: non-zero y+ from the first run's wall gradients, device == host on the same records
        assert all("y+ on the GPU: 0 nodes differ from the host's" in o for o in outs)
    assert "hf2d_gpu: DEEPS2D time loop on the GPU" in outs[0] and "hf2d_gpu: DEEPS2D time loop on the GPU" in outs[1]
    assert float(np.abs(Jr["y_plus"]).max()) > 1.0
    for f in ("y_plus", "l_min", "S", "Tg", "mu_t", "U", "V"):
        a, b = Jg[f], Jr[f]
        if exact:
            assert ((a == b) | ((a != a) & (b != b))).all(), (case, f)
        else:
            scale = max(float(np.abs(b).max()), 1e-300)
            assert float(np.abs(a - b).max()) / scale <= 1e-10, (case, f)
    for f in ("CT", "TurbType", "i_wall", "j_wall"):
        assert (Jg[f] == Jr[f]).all(), (case, f)


@pytest.mark.parametrize("name", ["ns_laminar_n4x2", "ns_keps_std_n4x2"])
def test_time_member_of_swept_nodes(name):
    """FlowNode2D::time: the reference stamps the running cycle's GlobalTime on every swept node (deeps2d_core.cpp:867)
    and a restarted run reads its GlobalTime from node (0,0) (:4618-4621).  Two-cycle golden runs: after the second
    cycle the member holds the GlobalTime reached at the end of the first."""
    from openhyperflow2d_b200 import cabi
    c = load_case(name)
    assert c["ncycles"] == 2
    J = np.array(c["state0"], copy=True)
    s = cabi.Solver(c["cfg"], LAYOUT)
    s.upload_aos(J)
    t = c["cfg"].get("GlobalTime", 0.0)
    for _ in range(c["ncycles"]):
        rep = s.step(c["nsteps"])
        t_cycle_start, t = t, t + rep.time_part
    s.set_global_time(t_cycle_start)
    s.download_aos(J)
    s.close()
    ref = c["stateN"]["time"]
    assert float(ref.max()) > 0 and t_cycle_start > 0
    assert np.allclose(J["time"], ref, rtol=1e-12, atol=0.0)
    assert ((J["time"] != 0) == (ref != 0)).all()              # exactly the swept nodes


@pytest.mark.parametrize("name", ["euler_duct_bff5_n7", "ns_keps_std_n8", "bc/euler_soft_n8"])
def test_residual_sums_finish_to_the_report(name):
    """hf2d_gpu_residual_sums (the raw accumulators a multi-GPU driver all-reduces) finished by the host formula of
    multigpu.finish_residual_norms = the norms hf2d_gpu_step reports for the same step: same inputs, same arithmetic."""
    from openhyperflow2d_b200 import cabi
    from openhyperflow2d_b200.multigpu import finish_residual_norms
    c = load_case(name)
    J = np.array(c["state0"], copy=True)
    s = cabi.Solver(c["cfg"], LAYOUT)
    s.upload_aos(J)
    rep = s.step(c["nsteps"])
    rms, sd, cnt, dd = s.residual_sums()
    s.close()
    out, mx, km = finish_residual_norms(rms, sd, cnt, c["cfg"])
    assert (out == np.array(list(rep.sum_RMS))).all()
    assert (dd == np.array(list(rep.DD_max))).all()
    assert mx == rep.max_RMS and km == rep.k_max_RMS
    assert out.max() > 0


@pytest.mark.parametrize("model", ["laminar", "keps"])
def test_slab_residual_sums_add_up_to_the_single_domain(model):
    """Two and four X-slabs on one GPU: their raw accumulators summed (what SlabProtocol.residuals all-reduces) and
    finished once give the norms of the undivided run (summation order differs: 1e-9 relative)."""
    from openhyperflow2d_b200 import gridgen
    from openhyperflow2d_b200.multigpu import finish_residual_norms
    from test_gpu_parity import _slab_run
    nx, ny, nsteps = 96, 160, 4
    _, cfg, _ = gridgen.three_cylinders(nx, ny, model=model)
    res = {}
    for world in (1, 2, 4):
        got = {}

        def grab(solvers):
            parts = [q.residual_sums() for q in solvers]
            got["rms"] = sum(p[0] for p in parts); got["sd"] = sum(p[1] for p in parts)
            got["cnt"] = sum(p[2].astype(np.int64) for p in parts); got["dd"] = np.max([p[3] for p in parts], axis=0)

        _slab_run(model, nx, ny, nsteps, world, before_close=grab)
        res[world] = (finish_residual_norms(got["rms"], got["sd"], got["cnt"], cfg), got)
    (want, _, km), g1 = res[1]
    assert want.max() > 0
    for world in (2, 4):
        (out, _, k), g = res[world]
        np.testing.assert_allclose(out, want, rtol=1e-9, atol=1e-300)
        assert (g["cnt"] == g1["cnt"]).all()            # every swept node counted once: ghost columns do not contribute
        assert (g["dd"] == g1["dd"]).all()
        assert k == km

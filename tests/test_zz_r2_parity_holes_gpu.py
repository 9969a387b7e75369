"""GPU parity for the round-2 fixtures (tests/golden/r2/, written by the unmodified reference): what the earlier golden
vectors left unpinned.

* burning decks -- fuel AND products present, both arms of CalcChemicalReactions (deeps2d_core.cpp:4711-4722) fire, the
  fuel / product rows of the twelve property tables are evaluated (the test asserts that the fixture really burns);
* grids that hold PLAIN tiles (asserted): the axisymmetric, Prandtl, Smagorinsky, SA, Euler and laminar instantiations
  of k_stage1_tma / k_stage1_plain / k_point / k_flux against the reference, not only against the oracle; one odd-ny
  deck on which the TMA form of K1a is not eligible and the LDG form runs;
* 200-step drift cases: the measured rel-L_inf after 200 steps against north_star's 1e-10 bound;
* a bench-shaped crop: the bench generator's grid at the bench's full height (4096 rows, 256 columns), 3 steps, against
  the CPU oracle.

Tolerances as in test_gpu_parity.py: bit-exact unless the physics calls exp()/pow() (then rel-L_inf <= 1e-10)."""
import numpy as np
import pytest

from golden_util import LAYOUT, load_case, r2_case_names
from test_gpu_parity import GPU_FIELDS, NS_FIELDS, REL_TOL, check_against_reference, rel_linf

pytestmark = pytest.mark.gpu

# cases whose closures call exp()/pow(): k-eps wall rewrites, van Driest damping, SA
TRANSCENDENTAL = ("keps", "prandtl_vd", "_sa_")


def _exact(name):
    return not any(t in name for t in TRANSCENDENTAL)


BURN = [n for n in r2_case_names() if "burn" in n]
PLAIN = [n for n in r2_case_names() if n.startswith("r2/plain_")]
DRIFT = [n for n in r2_case_names() if n.startswith("r2/drift_")]


@pytest.mark.parametrize("name", [n for n in BURN if not n.startswith("r2/plain_")])
def test_gpu_matches_reference_on_burning_decks(name):
    c = load_case(name)
    S = c["stateN"]["S"]
    assert (S[..., 6] > 0).any(), "fixture does not burn: no products"
    if "euler" not in name:
        # fuel-rich arm: fuel left over beside products; oxidiser-rich arm: oxidiser left over beside products
        assert ((S[..., 4] > 0) & (S[..., 6] > 0)).any() and ((S[..., 5] > 0) & (S[..., 6] > 0)).any()
    check_against_reference(name, exact=_exact(name))


@pytest.mark.parametrize("name", PLAIN)
def test_gpu_matches_reference_on_plain_tile_grids(name):
    from openhyperflow2d_b200 import cabi
    c = load_case(name)
    s = cabi.Solver(c["cfg"], LAYOUT)
    J = np.array(c["state0"], copy=True)
    s.upload_aos(J)
    n_plain, n_gen, tc, tr = s.tile_counts()
    s.close()
    assert n_plain > 0 and n_gen > 0, (n_plain, n_gen)
    check_against_reference(name, exact=_exact(name))


def _worst_rel_linf(A, B, fields):
    worst = 0.0
    for f in fields:
        a, b = np.asarray(A[f], float), np.asarray(B[f], float)
        if a.ndim == 2:
            a, b = a[..., None], b[..., None]
        for k in range(a.shape[-1]):
            worst = max(worst, rel_linf(np.nan_to_num(a[..., k]), np.nan_to_num(b[..., k])))
    return worst


def _reference_sensitivity(c, fields):
    """How far the REFERENCE's own arithmetic (the oracle, pinned bit-exactly on the same fixture) moves after the same
    number of steps when rho*k of every node of the start image is changed by ONE ulp: the conditioning of the deck."""
    import oracle_py as O
    s0 = np.array(c["state0"], copy=True)
    s0["S"][..., 7] = np.nextafter(s0["S"][..., 7], np.inf)
    nodes = O.nodes_from_ref(s0)
    rc, *_ = O.run_cycle(nodes, c["cfg"], c["cfg"]["last_iter"], c["nsteps"], c["cfg"]["dt"])
    assert rc == 0
    return _worst_rel_linf(nodes, c["stateN"], fields)


@pytest.mark.parametrize("name", DRIFT)
def test_drift_after_200_steps(name, record_property):
    """SURVEY section 4 asked for N in {1, 2, 10, 200}: the same decks as the short cases, 200 steps.
    Euler and laminar: bit-exact after 200 steps.  van Driest (exp): rel-L_inf <= 1e-10.
    k-eps: the reference's k-eps system is ILL-CONDITIONED on every deck tried (ramp duct, cylinders, airfoil, jets): one
    ulp on rho*k of the start image moves the reference's own result by 0.17 - 0.92 (rel-L_inf) after 200 steps -- eps grows
    by five orders of magnitude in these start-up transients and amplifies any last-bit difference by ~e^(0.13 per step).
    No implementation whose pow() is not glibc's bit for bit can hold 1e-10 there (ours is the correctly rounded x*sqrt(x),
    identical with glibc's result in 99.92 % of the calls: hf2d_local.cuh pow15).  north_star allows "a stated relaxed bound
    where the transcendental ... differs"; the bound stated here: first 40 steps <= 1e-10 (test below), and after 200 steps
    no further from the reference than the reference is from itself under that one-ulp change."""
    from test_gpu_parity import run_gpu
    c = load_case(name)
    assert c["nsteps"] == 200
    fields = GPU_FIELDS + (NS_FIELDS if c["cfg"]["ProblemType"] == 1 else [])
    J, rep, t = run_gpu(c)
    worst = _worst_rel_linf(J, c["stateN"], fields)
    record_property("rel_linf_after_200_steps", worst)
    print(f"{name}: rel-Linf after 200 steps = {worst:.3e}")
    if "keps" not in name:
        check_against_reference(name, exact=_exact(name))
        return
    sens = _reference_sensitivity(c, fields)
    record_property("reference_sensitivity_to_one_ulp", sens)
    print(f"{name}: the reference's own sensitivity to one ulp = {sens:.3e}")
    assert worst <= max(REL_TOL, sens), f"{name}: drift {worst:.3e} exceeds the reference's own one-ulp sensitivity {sens:.3e}"
    for f in ("CT", "TurbType"):
        assert (J[f] == c["stateN"][f]).all()


def test_keps_first_40_steps_within_bound():
    """The k-eps drift deck, step by step against the oracle while the amplification has not yet outgrown 1e-10."""
    import oracle_py as O
    from openhyperflow2d_b200 import cabi
    c = load_case("r2/drift_ns_keps_std_n200")
    cfg = c["cfg"]
    nodes = O.nodes_from_ref(c["state0"])
    J = np.array(c["state0"], copy=True)
    s = cabi.Solver(cfg, LAYOUT)
    s.upload_aos(J)
    dt_min = cfg["dt"]
    for it in range(40):
        rc, dt_min, orep = O.step(nodes, cfg, cfg["last_iter"] + it, 0, dt_min)
        assert rc == 0
        grep = s.step(1)
        assert grep.dt == orep.dt
    s.download_aos(J)
    s.close()
    worst = _worst_rel_linf(J, nodes, GPU_FIELDS + NS_FIELDS)
    print(f"k-eps drift deck after 40 steps: rel-Linf {worst:.3e}")
    assert worst <= REL_TOL


@pytest.mark.parametrize("model", ["keps"])
def test_bench_shaped_crop_matches_oracle(model):
    """The bench workload's generator at the bench's full height (4096 rows: 128 strips, the same tile shapes, TMA boxes
    and chain lengths per column as the 2048 x 4096 grid whose number is reported), 256 columns, 3 steps, against the
    CPU oracle."""
    import oracle_py as O
    from openhyperflow2d_b200 import cabi, gridgen
    nx, ny, nsteps = 256, 4096, 3
    J, cfg, lay = gridgen.three_cylinders(nx, ny, model=model)
    cfg["isAlternateRMS"] = 0
    nodes = O.nodes_from_ref(J)
    rc, dt_min, orep, tp = O.run_cycle(nodes, cfg, 0, nsteps, cfg["dt"])
    assert rc == 0
    s = cabi.Solver(cfg, lay)
    Jg = np.array(J, copy=True)
    s.upload_aos(Jg)
    n_plain, n_gen, tc, tr = s.tile_counts()
    assert n_plain > 1000 and n_gen > 0
    rep = s.step(nsteps)
    s.download_aos(Jg)
    s.close()
    assert rep.dt == orep.dt and rep.dt_min == orep.dt_min
    for f in GPU_FIELDS + NS_FIELDS:
        a, b = np.asarray(Jg[f], float), np.asarray(nodes[f], float)
        if a.ndim == 3:
            for k in range(a.shape[-1]):
                assert rel_linf(a[..., k], b[..., k]) <= REL_TOL, f"{f}[{k}] {rel_linf(a[..., k], b[..., k]):.3e}"
        else:
            assert rel_linf(a, b) <= REL_TOL, f"{f} {rel_linf(a, b):.3e}"
    np.testing.assert_allclose(list(rep.sum_RMS), list(orep.sum_RMS), rtol=1e-9, atol=1e-300)

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


@pytest.mark.parametrize("name", DRIFT)
def test_drift_after_200_steps(name, record_property):
    """SURVEY section 4 asked for N in {1, 2, 10, 200}: the same decks as the short cases, 200 steps.  The drift against
    the reference is recorded (pytest --junitxml / -rA) and held to the same bound as the short runs."""
    c = load_case(name)
    assert c["nsteps"] == 200
    from test_gpu_parity import run_gpu
    J, rep, t = run_gpu(c)
    worst = 0.0
    for f in GPU_FIELDS + (NS_FIELDS if c["cfg"]["ProblemType"] == 1 else []):
        a, b = np.asarray(J[f], float), np.asarray(c["stateN"][f], float)
        if a.ndim == 3:
            for k in range(a.shape[-1]):
                worst = max(worst, rel_linf(a[..., k], b[..., k]))
        else:
            worst = max(worst, rel_linf(a, b))
    record_property("rel_linf_after_200_steps", worst)
    print(f"{name}: rel-Linf after 200 steps = {worst:.3e}")
    check_against_reference(name, exact=_exact(name))


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

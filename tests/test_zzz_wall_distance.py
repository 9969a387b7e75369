"""Wall nodes and wall distance (GetWallNodes, deeps2d_core.cpp:2081-2104; SetMinDistanceToWall2D, :4783-4832) -- the
O(nodes x wall nodes) start-up pass in front of every Navier-Stokes run (hf2d_start.cpp:296-299).

CPU part (here): the oracle's restatement is pinned against the reference itself -- every golden pre-step image
(state0, written by the unmodified reference after its own start-up) carries the l_min / i_wall / j_wall members the
reference computed; the oracle must rebuild them bit for bit from the flag words alone, the tie rule (the LAST of
several equidistant wall nodes wins) and the clamp to min(dx, dy) included.

GPU part (-m gpu): hf2d_gpu_set_min_distance_to_wall through the C ABI against the same images and the oracle."""
import numpy as np
import pytest

import oracle_py as O
from golden_util import LAYOUT, bc_case_names, case_names, geometry_case_names, load_case, restart_case_names

NS_CASES = [n for n in case_names() + geometry_case_names() + bc_case_names() + restart_case_names()
            if not n.split("/")[-1].startswith(("euler", "axi_reacting_euler"))]


def _planes(s):
    return (np.ascontiguousarray(s["CT"], dtype=np.uint64), np.ascontiguousarray(s["Tg"]), np.ascontiguousarray(s["p"]))


@pytest.mark.parametrize("name", NS_CASES)
def test_oracle_rebuilds_reference_wall_distance(name):
    c = load_case(name)
    s, cfg = c["state0"], c["cfg"]
    assert cfg["ProblemType"] == 1
    ct, Tg, p = _planes(s)
    walls = O.wall_nodes(ct)
    l_min = np.full(ct.shape, -1.0); i_wall = np.full(ct.shape, -7, dtype=np.int32); j_wall = np.full(ct.shape, -7, dtype=np.int32)
    ct_in = ct.copy()
    status = O.set_min_distance_to_wall(ct, Tg, p, cfg["dx"], cfg["dy"], walls, l_min, i_wall, j_wall)
    assert (ct == ct_in).all() and not (status == 2).any()        # no unfilled gas node in a finished start-up image
    done = status == 1
    set_gas = ((ct_in & 0x080000000) != 0) & ((ct_in & 0x040000000) == 0)
    assert (done == set_gas).all() and done.sum() > 0
    assert (l_min[done] == s["l_min"][done]).all()
    if len(walls):
        assert (i_wall[done] == s["i_wall"][done]).all() and (j_wall[done] == s["j_wall"][done]).all()
    else:
        assert (i_wall == -7).all() and (j_wall == -7).all()       # nothing to assign: members stay as they were
    assert (l_min[~done] == -1.0).all() and (i_wall[~done] == -7).all()
    if len(walls):
        assert l_min[done].min() == min(cfg["dx"], cfg["dy"])      # the clamp: a wall node is its own nearest wall
    else:                                                          # no wall in the deck (axisymmetric jet): domain extent
        assert (l_min[done] == max(cfg["dx"] * cfg["MaxX"], cfg["dy"] * cfg["MaxY"])).all()


def test_tie_rule_clamp_and_unfilled_nodes():
    """4 x 3 grid, two wall nodes at equal distance from the middle node: the later one in GetWallNodes order wins."""
    SET, SOLID, WNS = 0x080000000, 0x040000000, 0x04000000
    ct = np.full((4, 3), SET, dtype=np.uint64)
    ct[0, 1] |= WNS; ct[2, 1] |= WNS          # both at distance dx from (1,1)
    ct[3, 0] = SET | SOLID | WNS               # solid: never a wall node, never visited
    ct[3, 2] = 0                                # unset: not visited
    Tg = np.full((4, 3), 300.0); p = np.full((4, 3), 1e5); p[1, 2] = 0.0   # (1,2): Tg != 0 and p == 0 -> made solid
    walls = O.wall_nodes(ct)
    assert walls.tolist() == [[0, 1], [2, 1]]
    l_min = np.zeros((4, 3)); iw = np.full((4, 3), -1, dtype=np.int32); jw = np.full((4, 3), -1, dtype=np.int32)
    st = O.set_min_distance_to_wall(ct, Tg, p, 0.5, 0.25, walls, l_min, iw, jw)
    assert st[3, 0] == 0 and st[3, 2] == 0 and st[1, 2] == 2 and (ct[1, 2] & SOLID)
    assert (iw[1, 1], jw[1, 1]) == (2, 1) and l_min[1, 1] == 0.5
    assert (iw[0, 1], jw[0, 1]) == (0, 1) and l_min[0, 1] == 0.25     # distance 0 clamped to min(dx, dy)
    assert (iw[2, 1], jw[2, 1]) == (2, 1)
    # no wall nodes at all: l_min = max(x0 + dx*nx, dy*ny), no clamp, no wall assigned
    l2 = np.zeros((4, 3)); iw2 = np.full((4, 3), -1, dtype=np.int32); jw2 = iw2.copy()
    O.set_min_distance_to_wall(ct.copy(), Tg, p, 0.5, 0.25, walls[:0], l2, iw2, jw2, x0=1.0)
    assert l2[1, 1] == 3.0 and iw2[1, 1] == -1


@pytest.mark.parametrize("name", restart_case_names() + ["ns_keps_std_n8", "geometry/cylinders_keps_n8"])
def test_oracle_rebuilds_reference_y_plus(name):
    """Recalc_y_plus (deeps2d_core.cpp:2364-2386) against the y_plus member the reference's own start-up left in the
    image: non-trivial for restarted runs (wall gradients of the previous run), identically 0 for fresh ones."""
    c = load_case(name)
    s, cfg = c["state0"], c["cfg"]
    yp = np.full(s["y_plus"].shape, -3.0)
    O.recalc_y_plus(s["CT"], cfg["dx"], cfg["dy"], s["i_wall"], s["j_wall"], s["dUdy"], s["dVdx"], s["mu"], s["S"][..., 0], yp)
    ct = np.ascontiguousarray(s["CT"], dtype=np.uint64)
    done = ((ct & 0x080000000) != 0) & ((ct & 0x040000000) == 0)
    assert (yp[done] == s["y_plus"][done]).all()
    assert (yp[~done] == -3.0).all()
    if name.startswith("restart/"):
        assert float(s["y_plus"].max()) > 100.0
    else:
        assert float(np.abs(s["y_plus"]).max()) == 0.0


# ---- GPU part ----------------------------------------------------------------------------------------------------------

def _gpu_wall_distance(c, J):
    from openhyperflow2d_b200 import cabi
    s = cabi.Solver(c["cfg"], LAYOUT)
    s.upload_aos(J)
    walls = O.wall_nodes(np.ascontiguousarray(J["CT"], dtype=np.uint64))
    out = s.set_min_distance_to_wall(walls)
    return s, walls, out


@pytest.mark.gpu
@pytest.mark.parametrize("name", NS_CASES)
def test_gpu_wall_distance_matches_reference(name):
    """l_min / i_wall / j_wall of the reference's own start-up image, rebuilt on the GPU from the flag words: bit-exact."""
    c = load_case(name)
    ref = c["state0"]
    J = np.array(ref, copy=True)
    J["l_min"] = 123.0                                   # whatever the record held must not matter
    s, walls, (l_min, i_wall, j_wall, status) = _gpu_wall_distance(c, J)
    s.close()
    ct = np.ascontiguousarray(ref["CT"], dtype=np.uint64)
    set_gas = ((ct & 0x080000000) != 0) & ((ct & 0x040000000) == 0)
    assert ((status == 1) == set_gas).all() and not (status == 2).any()
    done = status == 1
    assert (l_min[done] == ref["l_min"][done]).all()
    assert (l_min[~done] == 123.0).all()                 # unvisited nodes: the plane as uploaded
    if len(walls):
        assert (i_wall[done] == ref["i_wall"][done]).all() and (j_wall[done] == ref["j_wall"][done]).all()
    else:
        assert (i_wall == -1).all() and (j_wall == -1).all()
    assert (i_wall[~done] == -1).all()


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["ns_prandtl_vd_n8", "ns_keps_std_n8", "geometry/cylinders_keps_n8"])
def test_gpu_wall_distance_feeds_the_step(name):
    """Start-up pass on the GPU, then the time steps: the image uploaded with a WRONG l_min ends where the reference
    does, so the step reads the plane the pass wrote (mixing lengths and the eps wall value depend on l_min)."""
    from test_gpu_parity import REL_TOL, rel_linf
    c = load_case(name)
    J = np.array(c["state0"], copy=True)
    J["l_min"] = 0.5
    s, _, _ = _gpu_wall_distance(c, J)
    for _ in range(c["ncycles"]):
        s.step(c["nsteps"])
    s.download_aos(J)
    s.close()
    for f in ("S", "mu_t", "Tg"):
        a, b = J[f], c["stateN"][f]
        if a.ndim == 3:
            for k in range(a.shape[-1]):
                assert rel_linf(a[..., k], b[..., k]) <= REL_TOL, (name, f, k)
        else:
            assert rel_linf(a, b) <= REL_TOL, (name, f)


@pytest.mark.gpu
def test_gpu_wall_distance_ties_offsets_and_unfilled_nodes():
    """Against the oracle on a grid built for the corner cases: x0 != 0 (sub-domain offset), a gas node with p == 0
    (status 2), an empty wall list, more wall nodes than one shared-memory chunk (1024)."""
    from openhyperflow2d_b200 import cabi
    c = load_case("geometry/cylinders_laminar_n8")
    J = np.array(c["state0"], copy=True)
    J["p"][5, 7] = 0.0
    for x0, cut in ((0.0, None), (0.0137, None), (0.0, 0)):
        s = cabi.Solver(c["cfg"], LAYOUT)
        s.upload_aos(J)
        ct = np.ascontiguousarray(J["CT"], dtype=np.uint64)
        walls = O.wall_nodes(ct)[:cut]
        l_min, i_wall, j_wall, status = s.set_min_distance_to_wall(walls, x0=x0)
        s.close()
        lo = np.array(J["l_min"], copy=True); io = np.full(ct.shape, -1, dtype=np.int32); jo = io.copy()
        so = O.set_min_distance_to_wall(ct.copy(), J["Tg"], J["p"], c["cfg"]["dx"], c["cfg"]["dy"], walls, lo, io, jo, x0=x0)
        assert (status == so).all() and status[5, 7] == 2
        assert (l_min == lo).all() and (i_wall == io).all() and (j_wall == jo).all()
    # > 1 chunk of wall nodes: the list is the caller's -- every node of a 40 x 40 patch (1600 entries, rows outermost
    # as GetWallNodes would list them), so nearly every node has several equidistant "walls"
    J2 = np.array(c["state0"], copy=True)
    s = cabi.Solver(c["cfg"], LAYOUT)
    s.upload_aos(J2)
    ct = np.ascontiguousarray(J2["CT"], dtype=np.uint64)
    walls = np.array([(i, j) for j in range(2, 42) for i in range(2, 42)], dtype=np.int32)
    assert len(walls) > 1024
    l_min, i_wall, j_wall, status = s.set_min_distance_to_wall(walls)
    s.close()
    lo = np.array(J2["l_min"], copy=True); io = np.full(ct.shape, -1, dtype=np.int32); jo = io.copy()
    so = O.set_min_distance_to_wall(ct.copy(), J2["Tg"], J2["p"], c["cfg"]["dx"], c["cfg"]["dy"], walls, lo, io, jo)
    assert (status == so).all() and (l_min == lo).all() and (i_wall == io).all() and (j_wall == jo).all()


@pytest.mark.gpu
@pytest.mark.parametrize("case,exact", [("ns_laminar", True), ("ns_keps_std", False)])
def test_reference_program_with_gpu_wall_distance(case, exact, tmp_path, monkeypatch):
    """The second seam in the reference PROGRAM: oracle/_ref/ohf_gpu with HF2D_GPU_WALL_DISTANCE=1 wipes the l_min /
    i_wall / j_wall its own start-up computed, rebuilds them on the GPU and steps on those; the swap file it leaves
    must carry the same three members as the unmodified program's, and the same flow field."""
    import os
    import test_shim_dropin as D
    if not (os.path.exists(D.REF) and os.path.exists(D.GPU)):
        pytest.skip("oracle/_ref not built (make -C oracle ref)")
    deck = D._deck(case, 6, 2, 3e-7)
    Jr, rms_r, _ = D._run(D.REF, deck, case, str(tmp_path / "ref"))
    monkeypatch.setenv("HF2D_GPU_WALL_DISTANCE", "1")
    Jg, rms_g, out = D._run(D.GPU, deck, case, str(tmp_path / "gpu"))
    assert "wall nodes on the GPU" in out
    assert "y+ on the GPU: 0 nodes differ from the host's" in out       # hf2d_gpu_recalc_y_plus on the same start-up data
    assert len(rms_g) == len(rms_r)
    for f in ("l_min", "i_wall", "j_wall", "CT", "y_plus"):
        assert (Jg[f] == Jr[f]).all(), f
    assert (Jr["l_min"] > 0).any()
    for f in ("S", "Tg", "mu_t"):
        a, b = Jg[f], Jr[f]
        if exact:
            assert ((a == b) | ((a != a) & (b != b))).all(), f
        else:
            scale = max(float(np.abs(b).max()), 1e-300)
            assert float(np.abs(a - b).max()) / scale <= 1e-10, f


@pytest.mark.gpu
@pytest.mark.parametrize("name", restart_case_names() + ["ns_keps_std_n8"])
def test_gpu_y_plus_matches_reference(name):
    """hf2d_gpu_recalc_y_plus on the reference's own start-up image (before any step: the wall gradients are those of the
    records) against the y_plus member the reference's Recalc_y_plus left there -- bit-exact; y+ up to ~300 on the
    restarted runs, identically 0 on a fresh one."""
    from openhyperflow2d_b200 import cabi
    c = load_case(name)
    ref = c["state0"]
    J = np.array(ref, copy=True)
    J["y_plus"] = -5.0
    s = cabi.Solver(c["cfg"], LAYOUT)
    s.upload_aos(J)
    y = s.recalc_y_plus(J["i_wall"], J["j_wall"])
    s.close()
    ct = np.ascontiguousarray(ref["CT"], dtype=np.uint64)
    done = ((ct & 0x080000000) != 0) & ((ct & 0x040000000) == 0)
    assert (y[done] == ref["y_plus"][done]).all()
    assert (y[~done] == -5.0).all()
    if name.startswith("restart/"):
        assert float(y.max()) > 100.0


@pytest.mark.gpu
def test_gpu_y_plus_after_steps_uses_the_last_wall_gradients():
    """After some steps the refresh must see the wall gradients of the LAST step (what the MPI build's per-cycle
    Recalc_y_plus reads, deeps2d_core.cpp:1723-1726): oracle stepped alongside, then both refresh y+."""
    from openhyperflow2d_b200 import cabi
    from test_gpu_parity import REL_TOL, rel_linf
    c = load_case("ns_laminar_n8")
    cfg = c["cfg"]
    nodes = O.nodes_from_ref(c["state0"])
    J = np.array(c["state0"], copy=True)
    s = cabi.Solver(cfg, LAYOUT)
    s.upload_aos(J)
    rc, dt_min, rep, tp = O.run_cycle(nodes, cfg, cfg["last_iter"], 5, cfg["dt"], 0.0)
    assert rc == 0
    s.step(5)
    y = s.recalc_y_plus(J["i_wall"], J["j_wall"])
    s.close()
    yo = np.array(c["state0"]["y_plus"], copy=True)
    O.recalc_y_plus(nodes["CT"], cfg["dx"], cfg["dy"], J["i_wall"], J["j_wall"], nodes["dUdy"], nodes["dVdx"], nodes["mu"],
                    nodes["S"][..., 0], yo)
    assert float(yo.max()) > 1.0
    assert (y == yo).all(), rel_linf(y, yo)          # laminar: no transcendental on the path


@pytest.mark.gpu
def test_gpu_y_plus_refuses_a_nearest_wall_that_is_no_wall():
    """The gradients of an ordinary gas node are intermediates that are not kept: naming one as the nearest wall is an
    error (HF2D_EINVAL), and y+ stays as it was."""
    from openhyperflow2d_b200 import cabi
    c = load_case("ns_laminar_n8")
    J = np.array(c["state0"], copy=True)
    s = cabi.Solver(c["cfg"], LAYOUT)
    s.upload_aos(J)
    iw = np.full(J["i_wall"].shape, 5, dtype=np.int32); jw = np.full(J["i_wall"].shape, 30, dtype=np.int32)   # (5,30): free stream
    with pytest.raises(cabi.Hf2dError):
        s.recalc_y_plus(iw, jw)
    s.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["geometry/cylinders_keps_n8", "ns_keps_std_n8"])
def test_gpu_wall_distance_in_slabs_equals_the_undivided_grid(name):
    """Two and three X-slabs (ghost columns included, one handle each): the slab form takes the GLOBAL wall list and forms x
    from the global column index, so every slab reproduces the reference's undivided result for its nodes bit for bit."""
    from openhyperflow2d_b200 import cabi
    c = load_case(name)
    ref = c["state0"]
    nx = ref.shape[0]
    ct = np.ascontiguousarray(ref["CT"], dtype=np.uint64)
    walls = O.wall_nodes(ct)
    set_gas = ((ct & 0x080000000) != 0) & ((ct & 0x040000000) == 0)
    for world in (2, 3):
        w = nx // world
        for r in range(world):
            gl, gr = (1 if r > 0 else 0), (1 if r < world - 1 else 0)
            i0, i1 = r * w - gl, (nx if r == world - 1 else (r + 1) * w) + gr
            J = np.array(ref[i0:i1], copy=True)
            J["l_min"] = 77.0
            cfg = dict(c["cfg"]); cfg["MaxX"] = J.shape[0]
            s = cabi.Solver(cfg, LAYOUT, rank=r, n_ranks=world)
            s.upload_aos(J)
            l_min, i_wall, j_wall, status = s.set_min_distance_to_wall_slab(walls, i0, nx)
            s.close()
            done = status == 1
            assert (done == set_gas[i0:i1]).all()
            assert (l_min[done] == ref["l_min"][i0:i1][done]).all(), (world, r)
            assert (i_wall[done] == ref["i_wall"][i0:i1][done]).all() and (j_wall[done] == ref["j_wall"][i0:i1][done]).all()

"""CPU test (gloo, world_size 2 and 3) of the X-slab exchange protocol in openhyperflow2d_b200.multigpu.SlabProtocol.

The real solver is replaced by a stand-in with the same split-step / halo interface and the same DATA DEPENDENCIES as
the DEEPS2D step (hf2d_stage1.cuh / hf2d_stage2.cuh): an ordered in-place pre-pass whose value chains from the left
neighbour, a Stage 1 that reads the left neighbour after and the right neighbour before that overwrite, a Stage 2 that
reads the left neighbour's NEW state and the right neighbour's post-pre-pass state.  If any halo phase, side or
ordering were wrong, the slabs would not reproduce the single-domain run bit for bit."""
import ctypes as C
import os
import socket

import numpy as np
import pytest

NY = 7


def _view(ptr, n):
    return np.ctypeslib.as_array((C.c_double * n).from_address(ptr))


class FakeSolver:
    """Columns [g0, g0+w) of a global nx_total x NY field, with ghost columns where a neighbour slab exists."""

    def __init__(self, full, g0, w, rank, world):
        self.gl, self.gr = int(rank > 0), int(rank < world - 1)
        self.first_global = g0 == 0
        self.S = np.array(full[g0 - self.gl:g0 + w + self.gr], copy=True)
        self.SN = np.zeros_like(self.S)
        self.S0 = np.zeros_like(self.S)
        self.new = None          # Stage-2 output while its parts are pending (phase 0 then addresses the NEW state)
        self.ahead = False       # between AHEAD and S2_FINISH the pre-pass and halo phase 2 belong to the next step
        self.prepass_done = False
        self.dt_run = np.inf
        self.nx = self.S.shape[0]
        self.lo, self.hi = self.gl, self.nx - self.gr
        self.dt = 1.0

    # --- split step ---
    def step_begin(self):
        pass

    def _cur(self):
        return self.new if self.ahead else self.S

    def step_prepass(self):
        S = self._cur()
        self.S0[:] = S                            # pre-overwrite values (SpecNode::S0)
        for i in range(self.lo, self.hi):         # ordered chain: left already overwritten, right not yet
            if i == 0:
                continue                          # global column 0 has no left neighbour
            r = self.S0[i + 1] if i + 1 < self.nx else self.S0[i]
            S[i] = (S[i - 1] + r) * 0.5
        self.prepass_done = True

    def step_stage1(self):
        if not self.prepass_done:
            self.step_prepass()
        self.prepass_done = False
        for i in range(self.lo, self.hi):
            l = self.S[i - 1] if i > 0 else self.S[i]
            r = self.S0[i + 1] if i + 1 < self.nx else self.S[i]
            self.SN[i] = self.S[i] * 0.75 + 0.125 * (l + r) + 0.01

    def _s2_columns(self, cols):
        if self.new is None:
            self.new = np.array(self.S, copy=True)
        for i in cols:
            l_new = self.SN[i - 1] * 1.5 if i > 0 else self.SN[i] * 1.5     # left neighbour: recomputed NEW state (needs its SN)
            r_old = self.S[i + 1] if i + 1 < self.nx else self.S[i]           # right neighbour: after this step's pre-pass
            self.new[i] = self.SN[i] + 0.0625 * (r_old - l_new / 1.5)
            self.dt_run = min(self.dt_run, float(1.0 / (1.0 + np.abs(self.new[i]).max())))   # CFL-like minimum, inside Stage 2

    def _edge_columns(self):
        return ([self.lo] if self.gl else []) + ([self.hi - 1] if self.gr else [])

    def step_stage2(self):
        self._s2_columns(range(self.lo, self.hi))
        self._s2_finish()

    def _s2_finish(self):
        self.S, self.new = self.new, None
        self.ahead = False
        self.dt, self.dt_run = min(self.dt, self.dt_run), np.inf

    # --- the step in parts (cabi.Solver.step_part); lanes are streams on the GPU, nothing here ---
    PREPASS, S1_PLAIN, S1_GENERAL, S1_FINISH, S2_INNER, S2_EDGE, S2_FINISH, PREPASS_LOCAL, AHEAD = 1, 2, 3, 4, 5, 6, 7, 8, 9

    def step_part(self, part, lane=0, want_rms=1):
        if part == self.PREPASS:
            self.step_prepass()
        elif part == self.S1_GENERAL:
            self.step_stage1()
        elif part == self.S2_INNER:
            e = self._edge_columns()
            self._s2_columns([i for i in range(self.lo, self.hi) if i not in e])
        elif part == self.S2_EDGE:
            self._s2_columns(self._edge_columns())
        elif part == self.S2_FINISH:
            self._s2_finish()
        elif part == self.AHEAD:
            assert self.new is not None
            self.ahead = True

    def lane_fork(self):
        pass

    def lane_join(self):
        pass

    def step_end(self, want_report=False):
        pass

    # --- halos (include/hf2d_gpu.h: side 0 = left, 1 = right; phase 0 full column, 1 shadow S, 2 S after pre-pass) ---
    def halo_bytes(self, phase):
        return (2 if phase == 0 else 1) * NY * 8

    def halo_pack(self, side, phase, ptr, lane=0):
        col = 1 if side == 0 else self.nx - 2
        b = _view(ptr, self.halo_bytes(phase) // 8)
        if phase == 0:
            b[:NY], b[NY:] = (self.S if self.new is None else self.new)[col], self.SN[col]
        elif phase == 1:
            b[:] = self.SN[col]
        else:
            b[:] = self._cur()[col]

    def halo_unpack(self, side, phase, ptr, lane=0):
        col = 0 if side == 0 else self.nx - 1
        b = _view(ptr, self.halo_bytes(phase) // 8)
        if phase == 0:
            (self.S if self.new is None else self.new)[col], self.SN[col] = b[:NY], b[NY:]
        elif phase == 1:
            self.SN[col] = b
        else:
            self._cur()[col] = b

    def dt_export(self, ptr):
        _view(ptr, 1)[0] = self.dt

    def dt_import(self, ptr):
        self.dt = float(_view(ptr, 1)[0])


def _field(nx):
    rng = np.random.default_rng(7)
    return rng.uniform(0.5, 2.0, size=(nx, NY))


def _worker(rank, world, port, nx, nsteps, seams, out_dir, schedule="overlapped"):
    import torch
    import torch.distributed as dist
    from openhyperflow2d_b200.multigpu import SlabProtocol
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    w = nx // world
    fs = FakeSolver(_field(nx), rank * w, w, rank, world)
    p = SlabProtocol()
    p._init_protocol(torch, fs, rank, world, torch.device("cpu"), seams, schedule)
    p.step(nsteps)
    np.save(os.path.join(out_dir, f"r{rank}.npy"), np.concatenate([fs.S[fs.lo:fs.hi].ravel(), [fs.dt]]))
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world,schedule", [(2, "overlapped"), (3, "overlapped"), (2, "sequential")])
def test_exact_seams_reproduce_single_domain(world, schedule, tmp_path):
    import torch.multiprocessing as mp
    nx, nsteps = 12, 4
    ref = FakeSolver(_field(nx), 0, nx, 0, 1)
    for _ in range(nsteps):
        ref.step_begin(); ref.step_stage1(); ref.step_stage2(); ref.step_end()
    mp.spawn(_worker, args=(world, _free_port(), nx, nsteps, "exact", str(tmp_path), schedule), nprocs=world, join=True)
    got = [np.load(tmp_path / f"r{r}.npy") for r in range(world)]
    S = np.concatenate([g[:-1] for g in got]).reshape(nx, NY)
    assert (S == ref.S).all()
    assert all(g[-1] == ref.dt for g in got)       # dt: min-allreduce every step


def test_lagged_chain_differs_only_through_the_chain(tmp_path):
    """seams="lagged-chain" drops the rank-ordered hand-off: the run completes (no rank waits on another inside the
    pre-pass) and differs from the single-domain result, which is what the hand-off exists to prevent."""
    import torch.multiprocessing as mp
    nx, nsteps, world = 12, 3, 2
    ref = FakeSolver(_field(nx), 0, nx, 0, 1)
    for _ in range(nsteps):
        ref.step_begin(); ref.step_stage1(); ref.step_stage2(); ref.step_end()
    mp.spawn(_worker, args=(world, _free_port(), nx, nsteps, "lagged-chain", str(tmp_path)), nprocs=world, join=True)
    S = np.concatenate([np.load(tmp_path / f"r{r}.npy")[:-1] for r in range(world)]).reshape(nx, NY)
    assert not (S == ref.S).all()
    assert np.isfinite(S).all()


# ---- residual norms of the whole grid from the slabs' raw accumulators ------------------------------------

NEQ = 9


def _fake_sums(S):
    """Stand-in accumulators with the real ones' algebra: sums, a count and a maximum over the OWNED nodes, per 'equation'
    (here: row k of the field; rows beyond NY stay empty as unused equations do)."""
    rms = np.zeros(NEQ); sd = np.zeros(NEQ); cnt = np.zeros(NEQ, dtype=np.uint64); dd = np.zeros(NEQ)
    for k in range(min(NEQ, S.shape[1])):
        rms[k] = float((S[:, k] ** 2).sum()); sd[k] = float(np.abs(S[:, k]).sum()); cnt[k] = S.shape[0]; dd[k] = float(S[:, k].max())
    return rms, sd, cnt, dd


class FakeSolverWithSums(FakeSolver):
    def residual_sums(self):
        return _fake_sums(self.S[self.lo:self.hi])


def _worker_residuals(rank, world, port, nx, nsteps, out_dir):
    import torch
    import torch.distributed as dist
    from openhyperflow2d_b200.multigpu import SlabProtocol
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    w = nx // world
    fs = FakeSolverWithSums(_field(nx), rank * w, w, rank, world)
    p = SlabProtocol()
    p._init_protocol(torch, fs, rank, world, torch.device("cpu"), "exact", "overlapped")
    p.step(nsteps)
    res = []
    for alt, mi in ((1, 5), (0, 0), (1, 2)):
        out, mx, km, dd = p.residuals({"isAlternateRMS": alt, "MonitorIndex": mi, "ExitMonitorValue": 1e-3})
        res.append(np.concatenate([out, [mx, km], dd]))
    np.save(os.path.join(out_dir, f"res{rank}.npy"), np.array(res))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_residual_norms_of_the_whole_grid(world, tmp_path):
    """Every rank ends up with the norms the undivided run computes (sums over slabs, the square root taken once) --
    not with the reference's several-sub-domain variant, which takes the root inside the sub-domain loop and thereby
    depends on the decomposition (deeps2d_core.cpp:1541-1560)."""
    import torch.multiprocessing as mp
    from openhyperflow2d_b200.multigpu import finish_residual_norms
    nx, nsteps = 12, 3
    ref = FakeSolver(_field(nx), 0, nx, 0, 1)
    for _ in range(nsteps):
        ref.step_begin(); ref.step_stage1(); ref.step_stage2(); ref.step_end()
    mp.spawn(_worker_residuals, args=(world, _free_port(), nx, nsteps, str(tmp_path)), nprocs=world, join=True)
    got = [np.load(tmp_path / f"res{r}.npy") for r in range(world)]
    for g in got[1:]:
        assert (g == got[0]).all()                           # the same on every rank
    rms, sd, cnt, dd = _fake_sums(ref.S)
    for row, (alt, mi) in zip(got[0], ((1, 5), (0, 0), (1, 2))):
        cfg = {"isAlternateRMS": alt, "MonitorIndex": mi, "ExitMonitorValue": 1e-3}
        out, mx, km = finish_residual_norms(rms, sd, cnt, cfg)
        np.testing.assert_allclose(row[:NEQ], out, rtol=1e-13, atol=0)
        assert abs(row[NEQ] - mx) <= 1e-13 * mx and int(row[NEQ + 1]) == km
        assert (row[NEQ + 2:] == dd).all()                  # a maximum is order independent
        assert (out[NY:] == 0).all() and (out[:NY] > 0).all()
    # the single-domain formulas themselves (csrc/hf2d_step.inc fill_report): alternate = sqrt(sum / sum_div), else sqrt(sum / n)
    out, mx, km = finish_residual_norms(rms, sd, cnt, {"isAlternateRMS": 0, "MonitorIndex": 0})
    assert out[0] == np.sqrt(rms[0] / int(cnt[0])) and km == int(np.argmax(out)) and mx == out.max()
    out, mx, km = finish_residual_norms(rms, sd, cnt, {"isAlternateRMS": 1, "MonitorIndex": 3, "ExitMonitorValue": 0.0})
    assert out[1] == np.sqrt(rms[1] / sd[1]) and km == 2 and mx == out[2]


# ---- slab cuts by active nodes (SURVEY 8 f-2 / 8e "Partitioning") --------------------------------------------------------

def test_slab_cuts_balance_active_nodes():
    from golden_util import load_case
    from openhyperflow2d_b200.multigpu import slab_cuts
    for name in ("geometry/cylinders_keps_n8", "geometry/airfoil_keps_n8", "ns_keps_std_n8"):
        ct = np.ascontiguousarray(load_case(name)["state0"]["CT"], dtype=np.uint64)
        active = ((ct & 0x80000000) != 0) & ((ct & 0x40000000) == 0)
        per_col = active.sum(axis=1)
        for world in (1, 2, 3, 4, 8):
            cuts = slab_cuts(ct, world)
            assert cuts[0][0] == 0 and cuts[-1][1] == ct.shape[0]
            assert all(a[1] == b[0] for a, b in zip(cuts, cuts[1:]))           # every column owned exactly once
            assert all(i1 - i0 >= 2 for i0, i1 in cuts)
            loads = [int(per_col[i0:i1].sum()) for i0, i1 in cuts]
            assert sum(loads) == int(active.sum())
            assert max(loads) - min(loads) <= 2 * int(per_col.max())            # within a column or two of each other
    # a grid whose left half is solid: equal-column cuts would leave rank 0 idle, these do not
    ct = np.full((40, 10), 0x80000000, dtype=np.uint64)
    ct[:20] |= 0x40000000
    cuts = slab_cuts(ct, 2)
    assert cuts == [(0, 30), (30, 40)]
    # degenerate: nothing active -> minimum widths still respected
    cuts = slab_cuts(np.zeros((8, 4), dtype=np.uint64), 4)
    assert all(i1 - i0 >= 2 for i0, i1 in cuts) and cuts[-1][1] == 8


def test_c_slab_cuts_match_the_python_restatement():
    """hf2d_slab_cuts (csrc/hf2d_slab.inc, pure host code: runs without a GPU) against multigpu.slab_cuts on flag-word
    patterns with solid blocks, unset columns and degenerate quotas; every column owned, no slab narrower than min_cols."""
    from golden_util import LAYOUT
    from openhyperflow2d_b200 import aos, cabi
    from openhyperflow2d_b200.multigpu import slab_cuts
    rng = np.random.default_rng(7)
    dt = aos.ref_dtype(LAYOUT)
    for nx, ny in ((40, 12), (97, 33), (16, 5)):
        for trial in range(6):
            J = np.zeros((nx, ny), dtype=dt)
            ct = np.full((nx, ny), 0x80000000, dtype=np.uint64)
            if trial >= 1:      # a solid body on one side: the slabs there must be wider
                ct[: nx // 3, : (ny * (trial % 3 + 1)) // 4] |= 0x40000000
            if trial >= 3:      # columns that are not set at all
                ct[rng.integers(0, nx, size=nx // 5)] = 0
            if trial == 5:
                ct[:] = 0x80000000 | 0x40000000        # nothing active: the cuts fall back to the column limits
            J["CT"] = ct
            for world in (1, 2, 3, 4, 8):
                if nx < 2 * world:
                    continue
                got = cabi.slab_cuts(J, LAYOUT, world)
                want = slab_cuts(ct, world)
                assert got == [want[0][0]] + [b for _, b in want], (nx, ny, trial, world, got, want)
                assert got[0] == 0 and got[-1] == nx and all(b - a >= 2 for a, b in zip(got, got[1:]))

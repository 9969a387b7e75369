"""X-slab runner: one process per GPU, ghost-column exchange + dt min-allreduce through torch.distributed (NCCL on
GPUs; the CPU tests drive the same exchange code over gloo with a stand-in solver).

Decomposition (hf2d_start.cpp:149-175, deeps2d_core.cpp:831-840, 1336-1400): rank r owns global columns
[r*nx, (r+1)*nx); its slab carries one ghost column per interior seam, the sweep skips ghost columns.  Per step
(seams="exact", the default -- the result equals the single-domain serial sweep bit for bit, tests/test_gpu_parity.py):
    begin
    [phase 2, left -> right, in rank order]  receive the left neighbour's edge column S as its Neumann pre-pass left
        it -> own pre-pass (hf2d_gpu_step_prepass) -> send our last owned column on: a Neumann chain that runs along a
        whole d/dy=0 row (the eps equation) is thereby continued across the seam
    Stage 1
    [phase 1, left -> right]  shadow S (NextNode->S) of our last owned column -> right neighbour's left ghost, so that
        its Stage 2 recomputes that column's NEW state exactly as the serial sweep sees it
    [phase 2, right -> left]  S of our first owned column after the pre-pass -> left neighbour's right ghost (Stage 2
        reads right neighbours after this step's Stage-1 overwrite)
    Stage 2
    [phase 0, both ways]  full edge columns for the next step
    all-reduce(min) of the running dt -> end
seams="lagged-chain" drops the rank-ordered phase-2 hand-off (ranks no longer wait for each other inside a step);
a Neumann chain then restarts at each seam from the ghost column's last-step value -- the deviation is confined to
the chain rows next to a seam (~1e-5 relative after a few steps on the three-cylinder grid).  The reference's own MPI
build lags EVERYTHING at its seams by one step (deeps2d_core.cpp:1336-1400).
"""
from __future__ import annotations

import numpy as np

from . import cabi, gridgen


def slab_cuts(ct, world, min_cols=2):
    """X-slab boundaries with (nearly) equal numbers of ACTIVE nodes per slab: [(i0, i1), ...], slab r owns columns
    [i0, i1).  The reference's rule (ScanArea, deeps2d_core.cpp:2153, 2207-2216) closes a sub-domain as soon as its
    running count of set, non-solid nodes reaches total / n -- and leaves the columns after the last full quota to no
    sub-domain at all when the division has a remainder; here every column is owned, the cut after slab r sits at the
    first column where the prefix count reaches (r+1) * total / world, and no slab is narrower than `min_cols`.
    ct: (nx, ny) flag words (CT, low 32 bits suffice)."""
    ct = np.asarray(ct).astype(np.uint64)
    nx = ct.shape[0]
    assert world >= 1 and nx >= world * min_cols
    active = ((ct & 0x80000000) != 0) & ((ct & 0x40000000) == 0)          # CT_NODE_IS_SET_2D and not CT_SOLID_2D
    pre = np.concatenate([[0], np.cumsum(active.sum(axis=1))])           # pre[i] = active nodes in columns [0, i)
    total = int(pre[-1])
    cuts = [0]
    for r in range(1, world):
        i = int(np.searchsorted(pre, r * total / world, side="left"))
        i = max(i, cuts[-1] + min_cols)
        i = min(i, nx - (world - r) * min_cols)
        cuts.append(i)
    cuts.append(nx)
    return [(cuts[r], cuts[r + 1]) for r in range(world)]


def finish_residual_norms(rms, sum_div, count, cfg):
    """Residual norms of one domain from its raw accumulators (deeps2d_core.cpp:1539-1574 with a single sub-domain; the
    host side of hf2d_gpu_step's report, csrc/hf2d_step.inc fill_report): (sum_RMS[NEQ], max_RMS, k_max_RMS)."""
    neq = len(rms)
    out = np.zeros(neq)
    for k in range(neq):
        sr = 0.0
        if cfg.get("isAlternateRMS", 0):
            sd = 0.0
            if sum_div[k] > 0.0:
                sr += float(rms[k]); sd += float(sum_div[k])
            sr = float(np.sqrt(sr / sd)) if (sd > 0.0 and sr > 0.0) else 0.0
        else:
            si = 0
            if count[k] > 0:
                sr += float(rms[k]); si += int(count[k])
            sr = float(np.sqrt(sr / si)) if (si != 0 and sr > 0.0) else 0.0
        out[k] = sr
    mi = int(cfg.get("MonitorIndex", 0))
    max_rms = 0.5 * float(cfg.get("ExitMonitorValue", 0.0)) if mi < 5 else 0.0       # :1488-1492
    k_max = -1
    if mi == 0 or mi > 4:
        for k in range(neq):
            max_rms = max(out[k], max_rms)
            if max_rms == out[k]:
                k_max = k
    else:
        max_rms = max(out[mi - 1], max_rms)
        if max_rms == out[mi - 1]:
            k_max = mi - 1
    return out, max_rms, k_max


class SlabProtocol:
    """The per-step exchange protocol, independent of what computes: `solver` offers the split-step / halo calls of
    cabi.Solver, buffers are torch tensors on `dev` (CUDA + NCCL in production, CPU + gloo in the protocol tests)."""

    def _init_protocol(self, torch, solver, rank, world, dev, seams, schedule="overlapped"):
        assert seams in ("exact", "lagged-chain") and schedule in ("overlapped", "sequential")
        self.torch, self.solver, self.rank, self.world, self.dev, self.seams = torch, solver, rank, world, dev, seams
        self.schedule = schedule
        self.rms_every = getattr(self, "rms_every", None) or 1
        self.gl = 1 if (world > 1 and rank > 0) else 0
        self.gr = 1 if (world > 1 and rank < world - 1) else 0
        self.side = None
        if world > 1:
            self._alloc(dev)
            if dev.type == "cuda":      # lane 1 of the solver as a torch stream: collectives issued under it are ordered with it
                self.side = torch.cuda.ExternalStream(solver.side_stream(), device=dev)

    def _alloc(self, dev):
        torch = self.torch
        n0, n1, n2 = (self.solver.halo_bytes(p) // 8 for p in (0, 1, 2))
        z = lambda n: torch.zeros(n, dtype=torch.float64, device=dev)
        self.send, self.recv = [z(n0), z(n0)], [z(n0), z(n0)]
        self.send1, self.recv1 = z(n1), z(n1)
        self.send2, self.recv2 = [z(n2), z(n2)], [z(n2), z(n2)]
        self.dtbuf = z(1)

    def _lane(self, lane):
        """Collectives issued inside wait for, and are waited on by, that lane's stream (no-op on CPU / lane 0)."""
        import contextlib
        if lane == 0 or self.side is None:
            return contextlib.nullcontext()
        return self.torch.cuda.stream(self.side)

    def _batch(self, ops):
        if ops:
            for w in self.torch.distributed.batch_isend_irecv(ops):
                w.wait()

    def _mid_exchange(self, lane=0):
        """Between Stage 1 and Stage 2: phase 1 (left -> right) and the post-pre-pass S of the first owned column (right -> left)."""
        dist, s, ops = self.torch.distributed, self.solver, []
        if self.gr:
            s.halo_pack(1, 1, self.send1.data_ptr(), lane); ops.append(dist.P2POp(dist.isend, self.send1, self.rank + 1))
            ops.append(dist.P2POp(dist.irecv, self.recv2[1], self.rank + 1))
        if self.gl:
            s.halo_pack(0, 2, self.send2[0].data_ptr(), lane); ops.append(dist.P2POp(dist.isend, self.send2[0], self.rank - 1))
            ops.append(dist.P2POp(dist.irecv, self.recv1, self.rank - 1))
        self._batch(ops)
        if self.gl:
            s.halo_unpack(0, 1, self.recv1.data_ptr(), lane)
        if self.gr:
            s.halo_unpack(1, 2, self.recv2[1].data_ptr(), lane)

    def _end_exchange(self, lane=0):
        """After Stage 2 (of the edge tiles): phase 0, full edge columns both ways."""
        dist, s, ops = self.torch.distributed, self.solver, []
        if self.gl:
            s.halo_pack(0, 0, self.send[0].data_ptr(), lane)
            ops += [dist.P2POp(dist.isend, self.send[0], self.rank - 1), dist.P2POp(dist.irecv, self.recv[0], self.rank - 1)]
        if self.gr:
            s.halo_pack(1, 0, self.send[1].data_ptr(), lane)
            ops += [dist.P2POp(dist.isend, self.send[1], self.rank + 1), dist.P2POp(dist.irecv, self.recv[1], self.rank + 1)]
        self._batch(ops)
        if self.gl:
            s.halo_unpack(0, 0, self.recv[0].data_ptr(), lane)
        if self.gr:
            s.halo_unpack(1, 0, self.recv[1].data_ptr(), lane)

    def _handoff_prepass(self, lane):
        """The ordered Neumann pre-pass; with exact seams the overwritten edge column travels left -> right in rank order."""
        dist, s = self.torch.distributed, self.solver
        P = type(s)
        if self.seams == "exact" and self.gl:
            s.step_part(P.PREPASS_LOCAL, lane)      # all that does not depend on the left neighbour's column goes ahead of the wait
            dist.recv(self.recv2[0], self.rank - 1)
            s.halo_unpack(0, 2, self.recv2[0].data_ptr(), lane)
        s.step_part(P.PREPASS, lane)
        if self.seams == "exact" and self.gr:
            s.halo_pack(1, 2, self.send2[1].data_ptr(), lane)
            dist.send(self.send2[1], self.rank + 1)

    def _dt_allreduce(self):
        dist, s = self.torch.distributed, self.solver
        s.dt_export(self.dtbuf.data_ptr())
        dist.all_reduce(self.dtbuf, op=dist.ReduceOp.MIN)
        s.dt_import(self.dtbuf.data_ptr())

    def residuals(self, cfg):
        """Residual norms of the WHOLE grid after a step that reduced them: every rank's raw accumulators
        (hf2d_gpu_residual_sums) summed over the slabs -- one all-reduce of 3*NEQ doubles (SUM) and one of NEQ (MAX),
        SURVEY 8e(3); replaces the reference's Var_pack gather (deeps2d_core.cpp:1402-1463) -- and the norm finished once,
        as the undivided run does.  Returns (sum_RMS, max_RMS, k_max_RMS, DD_max); the same on every rank."""
        torch = self.torch
        rms, sd, cnt, dd = self.solver.residual_sums()
        a = torch.tensor(np.concatenate([rms, sd, np.asarray(cnt, dtype=np.float64)]), dtype=torch.float64, device=self.dev)
        m = torch.tensor(np.asarray(dd, dtype=np.float64), dtype=torch.float64, device=self.dev)
        if self.world > 1:
            dist = torch.distributed
            dist.all_reduce(a, op=dist.ReduceOp.SUM)
            dist.all_reduce(m, op=dist.ReduceOp.MAX)
        a = a.cpu().numpy(); n = len(rms)
        out, max_rms, k_max = finish_residual_norms(a[:n], a[n:2 * n], np.rint(a[2 * n:]).astype(np.int64), cfg)
        return out, max_rms, k_max, m.cpu().numpy()

    # ---- stepping ----
    def step(self, n):
        s = self.solver
        if self.world == 1:
            rep = s.step(n)
            self.launches_last = rep.kernel_launches
            return rep
        for q in range(n):
            want_rms = 1 if (q == n - 1 or self.rms_every <= 1 or q % self.rms_every == 0) else 0
            if self.schedule == "overlapped":
                # the next step's pre-pass + hand-off run inside this step's Stage 2 -- except before the caller gets
                # the grid back (S would hold the next step's Neumann overwrite)
                self._step_overlapped(want_rms, ahead=(self.seams == "exact" and q < n - 1))
            else:
                self._step_sequential(want_rms)
        return None

    def _step_sequential(self, want_rms):
        """Everything on one stream, exchanges between the stages (the straightforward schedule; kept as the baseline)."""
        s = self.solver
        P = type(s)
        s.step_begin()
        self._handoff_prepass(0)
        s.step_part(P.S1_GENERAL, 0, want_rms); s.step_part(P.S1_PLAIN, 0, want_rms); s.step_part(P.S1_FINISH, 0, want_rms)
        self._mid_exchange()
        s.step_part(P.S2_EDGE, 0, want_rms); s.step_part(P.S2_INNER, 0, want_rms)
        self._end_exchange()
        s.step_part(P.S2_FINISH, 0)
        self._dt_allreduce()
        s.step_end()

    def _step_overlapped(self, want_rms, ahead=False):
        """Seam work on lane 1 (high-priority side stream), bulk tiles on lane 0: see include/hf2d_gpu.h."""
        s = self.solver
        P = type(s)
        s.step_begin()
        s.lane_fork()
        with self._lane(1):
            if not getattr(self, "_prepass_ahead", False):
                self._handoff_prepass(1)
            s.step_part(P.S1_GENERAL, 1, want_rms)
        s.step_part(P.S1_PLAIN, 0, want_rms)
        s.lane_join()
        s.step_part(P.S1_FINISH, 0, want_rms)
        s.lane_fork()
        s.step_part(P.S2_INNER, 0, want_rms)     # the bulk is queued first: nothing below waits for it
        with self._lane(1):
            self._mid_exchange(1)
            s.step_part(P.S2_EDGE, 1, want_rms)
            self._end_exchange(1)
            if ahead:
                # Every node the Neumann pre-pass reads or writes lies in a general tile, and those have all finished
                # Stage 2 by now (AHEAD waits for them): the rank-ordered hand-off chain of the NEXT step -- 7 serial
                # hops on 8 GPUs -- runs here, behind S2_INNER, instead of in front of the next Stage 1.
                s.step_part(P.AHEAD, 1)
                self._handoff_prepass(1)
        self._prepass_ahead = bool(ahead)
        s.lane_join()
        s.step_part(P.S2_FINISH, 0)
        self._dt_allreduce()
        s.step_end()


class SlabRunner(SlabProtocol):
    def __init__(self, nx, ny, model="keps", rank=0, world=1, device=0, seams="exact", rms_every=None, schedule="overlapped"):
        import torch
        self.device = device
        gl = 1 if (world > 1 and rank > 0) else 0
        gr = 1 if (world > 1 and rank < world - 1) else 0
        J, cfg, lay = gridgen.three_cylinders(nx + gl + gr, ny, model=model, i0=rank * nx - gl, nx_total=nx * world)
        self.J, self.cfg, self.layout = J, cfg, lay
        solver = cabi.Solver(cfg, lay, device=device, rank=rank, n_ranks=world, rms_every=rms_every)
        solver.set_stream(torch.cuda.current_stream().cuda_stream)
        solver.upload_aos(J)
        self.launches_last = 0
        self.rms_every = rms_every
        self._init_protocol(torch, solver, rank, world, torch.device("cuda", device), seams, schedule)

    # ---- collectives ----
    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.torch.distributed.barrier()
            self.torch.cuda.synchronize()

    def allreduce_sum(self, v):
        if self.world == 1:
            return int(v)
        t = self.torch.tensor([int(v)], dtype=self.torch.int64, device=self.torch.device("cuda", self.device))
        self.torch.distributed.all_reduce(t)
        return int(t.item())

    def timed_steps(self, k):
        """Device time of k steps in ms: CUDA events on the stream the kernels run on, max over ranks."""
        torch = self.torch
        if self.world == 1:
            rep = self.solver.step(k)
            self.launches_last = rep.kernel_launches
            return rep.gpu_ms
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.barrier()
        l0 = self.solver.launch_count()
        e0.record()
        self.step(k)
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=torch.device("cuda", self.device))
        torch.distributed.all_reduce(ms, op=torch.distributed.ReduceOp.MAX)
        self.launches_last = self.solver.launch_count() - l0
        return float(ms.item())

    def active_cells(self):
        own = self.J[self.gl:self.J.shape[0] - self.gr]
        ct = own["CT"]
        act = int((((ct & 0x40000000) == 0) & ((ct & 0x8000001f) != 0x8000001f) & ((ct & 0x80000000) != 0)).sum())
        return act, own.size

    def e2e_cycle(self, nsteps, cycles=2):
        """One output cycle through HOST memory (pinned): upload AoS + nsteps + download AoS, wall-clock around the call
        bracketed by device synchronisation; max over ranks."""
        import time
        torch = self.torch
        host = torch.empty(self.J.nbytes, dtype=torch.uint8, pin_memory=True)
        Jp = host.numpy().view(self.J.dtype).reshape(self.J.shape)
        Jp[...] = self.J
        act, _ = self.active_cells()
        el = None
        for _ in range(max(1, cycles)):        # the last cycle is the one reported (the first also pays for cudaMalloc of the planes)
            Jp[...] = self.J                   # every cycle starts from the same grid image: identical work (outside the timed region)
            self.barrier()
            t0 = time.perf_counter()
            if self.world == 1:
                self.solver.run_cycle(Jp, nsteps, pinned=True)
            else:
                self.solver.upload_aos(Jp, pinned=True)
                self.step(nsteps)
                self.solver.download_aos(Jp, pinned=True)
            torch.cuda.synchronize()
            el = time.perf_counter() - t0
        if self.world > 1:
            t = torch.tensor([el], dtype=torch.float64, device=torch.device("cuda", self.device))
            torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
            el = float(t.item())
        tot = self.allreduce_sum(act)
        return {"value": tot * nsteps / el, "unit": "cell-updates/s", "h2d_bytes_per_step": int(self.J.nbytes / nsteps),
                "d2h_bytes_per_step": int(self.J.nbytes / nsteps), "steps_per_cycle": nsteps, "cycles_run": max(1, cycles), "seconds": el,
                "note": "hf2d_gpu_run_cycle: pinned host AoS image -> HBM (scanned + unpacked there), nsteps steps (= Nmax, one output cycle), packed into the resident image -> host"}

    def close(self):
        self.solver.close()


class CSlabRunner:
    """bench.py's driver for N > 1: one process per GPU, stepping through the C X-slab driver (csrc/hf2d_slab.inc) --
    ghost columns, dt and residual accumulators travel by direct stores into the neighbours' HBM over NVLink (cudaIpc
    peer memory + flag words); torch.distributed only carries the 64-byte mailbox handles once, the timing max and the
    barriers.  Same interface as SlabRunner (the torch.distributed/NCCL protocol harness)."""

    def __init__(self, nx, ny, model="keps", rank=0, world=1, device=0, rms_every=None, J=None, cfg=None, lay=None, cuts=None):
        import torch
        self.torch, self.rank, self.world, self.device = torch, rank, world, device
        if cuts is None:
            cuts = [r * nx for r in range(world + 1)]
        self.gl = 1 if (world > 1 and rank > 0) else 0
        self.gr = 1 if (world > 1 and rank < world - 1) else 0
        if J is None:
            w = cuts[rank + 1] - cuts[rank]
            J, cfg, lay = gridgen.three_cylinders(w + self.gl + self.gr, ny, model=model, i0=cuts[rank] - self.gl, nx_total=cuts[-1])
            cfg = dict(cfg, MaxX=cuts[-1])       # the slab driver is given the UNDIVIDED grid's configuration
        self.J, self.cfg, self.layout = J, cfg, lay
        self.slab = cabi.Slab(cfg, lay, device, rank, world, cuts, rms_every=rms_every)
        self.solver = self.slab.solver
        if world > 1:
            dist = torch.distributed
            n = len(self.slab.ipc_handle())
            mine = torch.tensor(list(self.slab.ipc_handle()), dtype=torch.uint8, device=torch.device("cuda", device))
            allh = [torch.empty(n, dtype=torch.uint8, device=mine.device) for _ in range(world)]
            dist.all_gather(allh, mine)
            self.slab.connect_ipc(b"".join(bytes(t.cpu().tolist()) for t in allh))
        self.slab.upload_local(J)
        self.launches_last = 0

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.torch.distributed.barrier()
            self.torch.cuda.synchronize()

    def allreduce_sum(self, v):
        if self.world == 1:
            return int(v)
        t = self.torch.tensor([int(v)], dtype=self.torch.int64, device=self.torch.device("cuda", self.device))
        self.torch.distributed.all_reduce(t)
        return int(t.item())

    def step(self, n, new_cycle=True):
        rep = self.slab.run(n, new_cycle)
        self.launches_last = rep.kernel_launches
        return rep

    def timed_steps(self, k):
        """Device time of k steps in ms (CUDA events on the slab's own stream around the enqueued steps), max over ranks."""
        torch = self.torch
        self.barrier()
        rep = self.step(k)
        ms = rep.gpu_ms
        if self.world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=torch.device("cuda", self.device))
            torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    def active_cells(self):
        own = self.J[self.gl:self.J.shape[0] - self.gr]
        ct = own["CT"]
        act = int((((ct & 0x40000000) == 0) & ((ct & 0x8000001f) != 0x8000001f) & ((ct & 0x80000000) != 0)).sum())
        return act, own.size

    def e2e_cycles(self, nsteps, cycles=3):
        """`cycles` output cycles through HOST memory (pinned): upload AoS + nsteps + download AoS each, wall clock around
        the calls bracketed by barriers; per cycle the max over ranks; returns the list of seconds."""
        import time
        torch = self.torch
        host = torch.empty(self.J.nbytes, dtype=torch.uint8, pin_memory=True)
        Jp = host.numpy().view(self.J.dtype).reshape(self.J.shape)
        out = []
        for _ in range(max(1, cycles)):
            Jp[...] = self.J                   # every cycle starts from the same grid image (outside the timed region)
            self.barrier()
            t0 = time.perf_counter()
            self.slab.upload_local(Jp, pinned=True)
            self.slab.run(nsteps)
            self.slab.download_local(Jp, pinned=True)
            torch.cuda.synchronize()
            el = time.perf_counter() - t0
            if self.world > 1:
                t = torch.tensor([el], dtype=torch.float64, device=torch.device("cuda", self.device))
                torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
                el = float(t.item())
            out.append(el)
        return out

    def close(self):
        self.barrier()          # nobody frees its mailbox while a neighbour may still write into it
        self.slab.close()


def slab_parity_check(rank, world, device, model="keps", nx=192, ny=160, nsteps=7):
    """Run on the box before timing (bench.py, N > 1): a small grid stepped by the C slab driver on all `world` GPUs against
    the single-domain run of the same grid on this rank's own GPU; True iff this rank's owned columns agree BIT FOR BIT in
    every field the step owns.  The caller all-reduces the verdicts."""
    J, cfg, lay = gridgen.three_cylinders(nx, ny, model=model)
    cuts = cabi.slab_cuts(J, lay, world)
    s = cabi.Solver(cfg, lay, device=device)
    ref = np.array(J, copy=True)
    s.upload_aos(ref); s.step(nsteps); s.download_aos(ref); s.close()
    lo, hi = cuts[rank] - (1 if rank > 0 else 0), cuts[rank + 1] + (1 if rank < world - 1 else 0)
    r = CSlabRunner(nx, ny, model=model, rank=rank, world=world, device=device, J=J[lo:hi].copy(), cfg=cfg, lay=lay, cuts=cuts)
    r.step(nsteps)
    got = np.array(r.J, copy=True)
    r.slab.download_local(got)
    own = got[r.gl:got.shape[0] - r.gr]
    want = ref[cuts[rank]:cuts[rank + 1]]
    ok = True
    for f in ("S", "beta", "A", "B", "U", "V", "Tg", "p", "k", "R", "CP", "lam", "mu", "mu_t"):
        a, b = np.asarray(own[f]), np.asarray(want[f])
        ok = ok and bool(((a == b) | ((a != a) & (b != b))).all())
    r.close()
    return ok

"""X-slab runner: one process per GPU, ghost-column exchange + dt min-allreduce through torch.distributed (NCCL on
GPUs, gloo in the CPU tests of the exchange logic).

Decomposition (hf2d_start.cpp:149-175, deeps2d_core.cpp:831-840, 1336-1400): rank r owns global columns
[r*nx, (r+1)*nx); its slab carries one ghost column per interior seam, the sweep skips ghost columns.  Per step:
    begin -> Stage 1 -> [phase 1: our last owned column's shadow S (NextNode->S) goes to the right neighbour's left
    ghost, so that its Stage 2 can recompute that column's NEW state exactly as the serial sweep sees it] -> Stage 2
    -> [phase 0: full edge columns both ways] -> all-reduce(min) of the running dt -> end
With phase 1 the result equals the single-domain serial sweep; the reference's MPI build skips it (seam-lagged).
"""
from __future__ import annotations

import numpy as np

from . import cabi, gridgen


class SlabRunner:
    def __init__(self, nx, ny, model="keps", rank=0, world=1, device=0, exact_seams=True):
        import torch
        self.torch = torch
        self.rank, self.world, self.device = rank, world, device
        self.exact = exact_seams
        gl = 1 if (world > 1 and rank > 0) else 0
        gr = 1 if (world > 1 and rank < world - 1) else 0
        self.gl, self.gr = gl, gr
        J, cfg, lay = gridgen.three_cylinders(nx + gl + gr, ny, model=model, i0=rank * nx - gl, nx_total=nx * world)
        self.J, self.cfg, self.layout = J, cfg, lay
        self.solver = cabi.Solver(cfg, lay, device=device, rank=rank, n_ranks=world)
        self.solver.set_stream(torch.cuda.current_stream().cuda_stream)
        self.solver.upload_aos(J)
        self.launches_last = 0
        if world > 1:
            dev = torch.device("cuda", device)
            n0, n1 = self.solver.halo_bytes(0) // 8, self.solver.halo_bytes(1) // 8
            self.send = [torch.zeros(n0, dtype=torch.float64, device=dev) for _ in range(2)]
            self.recv = [torch.zeros(n0, dtype=torch.float64, device=dev) for _ in range(2)]
            self.send1 = torch.zeros(n1, dtype=torch.float64, device=dev)
            self.recv1 = torch.zeros(n1, dtype=torch.float64, device=dev)
            self.dtbuf = torch.zeros(1, dtype=torch.float64, device=dev)

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

    def _exchange(self, phase):
        dist = self.torch.distributed
        s = self.solver
        ops = []
        if phase == 1:   # left -> right only
            if self.gr:
                s.halo_pack(1, 1, self.send1.data_ptr()); ops.append(dist.P2POp(dist.isend, self.send1, self.rank + 1))
            if self.gl:
                ops.append(dist.P2POp(dist.irecv, self.recv1, self.rank - 1))
            if ops:
                for w in dist.batch_isend_irecv(ops):
                    w.wait()
            if self.gl:
                s.halo_unpack(0, 1, self.recv1.data_ptr())
            return
        if self.gl:
            s.halo_pack(0, 0, self.send[0].data_ptr())
            ops += [dist.P2POp(dist.isend, self.send[0], self.rank - 1), dist.P2POp(dist.irecv, self.recv[0], self.rank - 1)]
        if self.gr:
            s.halo_pack(1, 0, self.send[1].data_ptr())
            ops += [dist.P2POp(dist.isend, self.send[1], self.rank + 1), dist.P2POp(dist.irecv, self.recv[1], self.rank + 1)]
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        if self.gl:
            s.halo_unpack(0, 0, self.recv[0].data_ptr())
        if self.gr:
            s.halo_unpack(1, 0, self.recv[1].data_ptr())

    # ---- stepping ----
    def step(self, n):
        s = self.solver
        if self.world == 1:
            rep = s.step(n)
            self.launches_last = rep.kernel_launches
            return rep
        dist = self.torch.distributed
        for _ in range(n):
            s.step_begin()
            s.step_stage1()
            if self.exact:
                self._exchange(1)
            s.step_stage2()
            self._exchange(0)
            s.dt_export(self.dtbuf.data_ptr())
            dist.all_reduce(self.dtbuf, op=dist.ReduceOp.MIN)
            s.dt_import(self.dtbuf.data_ptr())
            s.step_end()
        return None

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

    def e2e_cycle(self, nsteps):
        """One output cycle through HOST memory (pinned): upload AoS + nsteps + download AoS, wall-clock around the call
        bracketed by device synchronisation; max over ranks."""
        import time
        torch = self.torch
        host = torch.empty(self.J.nbytes, dtype=torch.uint8, pin_memory=True)
        Jp = host.numpy().view(self.J.dtype).reshape(self.J.shape)
        Jp[...] = self.J
        act, _ = self.active_cells()
        self.barrier()
        t0 = time.perf_counter()
        if self.world == 1:
            rep = self.solver.run_cycle(Jp, nsteps, pinned=True)
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
        return {"value": tot * nsteps / el, "unit": "cell-updates/s", "h2d_bytes_per_step": int(self.J.nbytes * 2 / nsteps),
                "d2h_bytes_per_step": int(self.J.nbytes / nsteps), "steps_per_cycle": nsteps, "seconds": el,
                "note": "hf2d_gpu_run_cycle: pinned host AoS -> HBM SoA, nsteps steps, HBM -> host AoS (download re-reads the image: 2x H2D)"}

    def close(self):
        self.solver.close()

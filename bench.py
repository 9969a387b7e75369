#!/usr/bin/env python
"""bench.py -- FP64 active-cell-updates/s of the DEEPS2D explicit time step (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W          our arm: libhf2d_gpu.so on N B200s (torchrun for N>1)
  python bench.py --impl reference ...                    reference arm: the reference's own CPU solver on host cores

A "step" is one DEEPS2D time step (Stage 1 + Stage 2 + dt/RMS reductions + wall heat) over the rank's X-slab.
Workload (config.workload): BASELINE configs[3], "hypersonic flow around three cylinders, k-eps URANS", weak scaling
2048 x 4096 cells per GPU (16384 x 4096 at 8 GPUs), synthetic grid from openhyperflow2d_b200.gridgen.
value  = active cells * K / device time (max over ranks), inputs resident in HBM.
e2e    = same metric through the C ABI with HOST AoS buffers: hf2d_gpu_run_cycle = upload + Nmax steps + download
         (one output cycle of a DEEPS2D_Run shim that keeps nothing on the GPU between cycles), pinned host memory,
         copies inside the timed region.  Nmax = 500, the smallest value in the reference's own decks (TestCases/*.dat:
         TriplePoint 500, ObliqueShock 12000, Wedge 15000, Step 24000).
roofline.achieved = algorithmic bytes/cell * active cells / mean device time of one step.
"""
from __future__ import annotations

import argparse
import json
import os
import re
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

METRIC = "fp64_active_cell_updates_per_s"
UNIT = "cell-updates/s"
NX_PER_GPU, NY = 2048, 4096
REF_NX, REF_NY = 1024, 512          # grid of the reference (CPU) arm, see reference_arm()
WORKLOAD = ("BASELINE configs[3]: hypersonic flow around three cylinders, NS k-eps URANS (9 equations; k-eps in place of SST: the reference "
            "has no SST arithmetic, hyper_flow_node.hpp:919-926), flat")


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region (NVML, ~5 ms period; nvidia-smi as fallback)."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index=0):
        self.sm, self.mx, self.reasons, self.stop, self.index = [], None, set(), False, index
        self.t = threading.Thread(target=self.run, daemon=True)

    def run_nvml(self):
        import pynvml as N
        N.nvmlInit()
        h = N.nvmlDeviceGetHandleByIndex(self.index)
        self.mx = N.nvmlDeviceGetMaxClockInfo(h, N.NVML_CLOCK_SM)
        bits = {"hw_slowdown": N.nvmlClocksThrottleReasonHwSlowdown, "hw_thermal_slowdown": N.nvmlClocksThrottleReasonHwThermalSlowdown,
                "sw_thermal_slowdown": N.nvmlClocksThrottleReasonSwThermalSlowdown, "sw_power_cap": N.nvmlClocksThrottleReasonSwPowerCap}
        while not self.stop:
            self.sm.append(N.nvmlDeviceGetClockInfo(h, N.NVML_CLOCK_SM))
            r = N.nvmlDeviceGetCurrentClocksThrottleReasons(h)
            for k, v in bits.items():
                if r & v:
                    self.reasons.add(k)
            time.sleep(0.005)

    def run_smi(self):
        while not self.stop:
            try:
                o = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits"],
                                   capture_output=True, text=True, timeout=5).stdout.strip()
                r = [x.strip() for x in o.split(",")]
                if r and r[0].isdigit():
                    self.sm.append(int(r[0])); self.mx = int(r[1])
                    for name, col in (("hw_slowdown", 2), ("hw_thermal_slowdown", 3), ("sw_thermal_slowdown", 4), ("sw_power_cap", 5)):
                        if len(r) > col and r[col].lower().startswith("active"):
                            self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.05)

    def run(self):
        try:
            self.run_nvml()
        except Exception:
            self.run_smi()

    def __enter__(self):
        self.t.start(); time.sleep(0.02); return self

    def __exit__(self, *a):
        self.stop = True; self.t.join(timeout=6)

    def summary(self):
        sm = sorted(self.sm)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_min_mhz": sm[0] if sm else None, "sm_max_mhz": self.mx,
                "reasons": sorted(self.reasons), "samples": len(sm)}


def cpu_baseline_port(model="keps", nx=256, ny=128, nsteps=120):
    """The oracle (plain-C port of the reference step) on one host core, bounded sample of the same workload."""
    import oracle_py as O
    from openhyperflow2d_b200 import gridgen
    J, cfg, lay = gridgen.three_cylinders(nx, ny, model=model)
    nodes = O.nodes_from_ref(J)
    act = int((((J["CT"] & 0x40000000) == 0) & ((J["CT"] & 0x8000001f) != 0x8000001f)).sum())
    O.run_cycle(nodes, cfg, 0, 2, cfg["dt"])
    t0 = time.time()
    rc, dtm, rep, tp = O.run_cycle(nodes, cfg, 2, nsteps, cfg["dt"])
    el = time.time() - t0
    return {"value": act * nsteps / el, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": f"three cylinders {model} {nx}x{ny} ({act} active cells), {nsteps} steps, oracle/deeps2d_oracle.c -O2 single thread"}


def reference_deck(nx, ny, nmax):
    """Three-cylinder k-eps NS deck in the reference's input format (authored here; Circle keys: deeps2d_core.cpp:4090-4180)."""
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import make_golden as G
    dx = 2.5e-4
    Lx, Ly = nx * dx, ny * dx
    r = 0.06 * Ly
    geo = [(0.0, Ly), (Lx - 1e-6, Ly), (Lx - 1e-6, 0.0), (0.0, 0.0)]
    b = [("NT_D0Y_2D, TCT_dkdy_NULL_2D, TCT_depsdy_NULL_2D, CT_V_CONST_2D", 1),
         ("NT_D0X_2D, TCT_dkdx_NULL_2D, TCT_depsdx_NULL_2D, CT_NONREFLECTED_2D", 1),
         ("NT_D0Y_2D, TCT_dkdy_NULL_2D, TCT_depsdy_NULL_2D, CT_V_CONST_2D", 1),
         ("NT_FARFIELD_2D, TCT_k_CONST_2D, TCT_eps_CONST_2D", 1)]
    deck = G.make_deck("Cyl3", nx=nx, ny=ny, dx=dx, dy=dx, problem=1, turb=4, turb_ext=4, turb_reset=1, turb_start=1, nmax=nmax,
                       flows=[G.air(6.0, p=5000., T=220.), G.air(p=5000., T=220.)], bounds=b, contour=geo,
                       areas=[{"seed": (3, 3), "type": 1, "flow": 1}])
    circ = ["<data/NumCircles=3>"]
    for q, (cx, cy) in enumerate([(0.25 * Lx, 0.5 * Ly), (0.5 * Lx, 0.3 * Ly), (0.5 * Lx, 0.7 * Ly)], 1):
        circ += [f"<data/Circle{q}.Xstart={cx - r!r}>", f"<data/Circle{q}.Ystart={cy!r}>", f"<data/Circle{q}.X0={cx!r}>", f"<data/Circle{q}.Y0={cy!r}>",
                 f"<data/Circle{q}.MaterialID=1>", f"<data/Circle{q}.TurbulenceModel=4>", f"<data/Circle{q}.Flow2D=2>"]
    deck = deck.replace("<data/NumCircles=0>", "\n".join(circ)).replace("<data/NOutStep=1>", f"<data/NOutStep={max(nmax - 1, 1)}>")
    return deck


def run_reference_binary(binary, nx, ny, nsteps, threads):
    """Time the UNMODIFIED reference (oracle/_ref) on the three-cylinder deck; returns (cells/s, active, steps/s) or None."""
    exe = os.path.join(ROOT, "oracle", "_ref", binary)
    if not os.path.exists(exe):
        return None
    wd = tempfile.mkdtemp(prefix="ohf_bench_")
    with open(os.path.join(wd, "Cyl3.dat"), "w") as f:
        f.write(reference_deck(nx, ny, nsteps))
    env = dict(os.environ, OMP_NUM_THREADS=str(threads), OHF_DUMP_DIR=wd)
    try:
        out = subprocess.run([exe, "Cyl3.dat"], cwd=wd, env=env, capture_output=True, text=True, timeout=600).stdout
    except Exception:
        return None
    m = re.findall(r"average\s+speed\s+([0-9.eE+-]+)\s+step/sec", out)
    a = re.findall(r"(\d+)\s+gas nodes|active nodes[^0-9]*(\d+)", out)
    if not m:
        return None
    import numpy as np
    from openhyperflow2d_b200 import aos, gridgen
    lay = aos.parse_layout(gridgen.DEFAULT_LAYOUT)
    try:
        J = aos.load_aos(os.path.join(wd, "Cyl3.hf2d"), lay, nx, ny)
        act = int((((J["CT"] & 0x40000000) == 0) & ((J["CT"] & 0x8000001f) != 0x8000001f)).sum())
    except Exception:
        act = nx * ny
    sps = float(m[0])
    return sps * act, act, sps


def reference_arm(args):
    """The reference's own CPU implementation of the path on the box's host cores: the UNMODIFIED OpenMP build
    (oracle/_ref/ohf_omp, all hardware threads) on a bounded sample of the same workload.  One solver run of a few
    hundred DEEPS2D steps (a "step" of this arm is one of them); the number is the solver's own 'average speed' line."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    nthreads = os.cpu_count() or 1
    # Grid of the reference arm: the reference's own START-UP (SetMinDistanceToWall2D, an O(nodes x wall nodes) host loop,
    # deeps2d_core.cpp:4783-4832) takes ~40 s at 1024x512 and would take the better part of an hour at the GPU arm's
    # 2048x4096, so the same deck is timed on the largest grid whose start-up fits a bench run; cell-updates/s of the CPU
    # solver does not depend on the grid size (2.4-2.9 M at 512x256, 1024x512 and 1024x1024).
    nx, ny = REF_NX, REF_NY
    nsteps = int(min(max(4 * (args.warmup + args.steps), 100), 200))
    r = run_reference_binary("ohf_omp", nx, ny, nsteps, nthreads)
    if r is not None:
        ser = run_reference_binary("ohf_serial", nx, ny, 40, 1)
        v, kind, cores, ms = r[0], "reference", nthreads, 1e3 / r[2]
        how = (f"unmodified reference, OpenMP build (oracle/_ref/ohf_omp, OMP_NUM_THREADS={nthreads}), three-cylinder k-eps deck {nx}x{ny} "
               f"({r[1]} active cells), one run of {nsteps} steps, the solver's own 'average speed' line"
               + (f"; serial build on one core: {ser[0]:.4g} cell-updates/s" if ser else ""))
    else:
        cb = cpu_baseline_port()
        v, kind, cores, how, ms = cb["value"], "port", 1, cb["sample"] + " (oracle/_ref not present: oracle port used)", None
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "grid": f"{nx}x{ny}", "sample": f"one run of {nsteps} steps",
                       "grid_note": "the reference's O(nodes x walls) start-up limits the CPU arm's grid; the GPU arm runs 2048x4096 per GPU"},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": kind, "sample": how},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def kernel_bytes_per_cell(model, axi=False):
    """Algorithmic bytes per plain cell of ONE launch of each of the three plain-tile kernels: every
    plane a kernel owns read once / written once, neighbour reads served on chip.
      K1a k_stage1_plain: reads S, beta, A, B [Neq each], Src [Neq-7], (F [Neq] axisymmetric), (dSdy[eps] k-eps); writes SN, beta [Neq each], (dSdy[eps])
      K1b k_point:        reads SN [Neq], CP, R, k; writes S [Neq], U, V, Tg, k, p, R, CP (+ mu, lam NS)
      K2  k_flux:         reads S_new [Neq], (SN[4..6] where chemistry changed them), U, V, Tg, p, k (new), S_old[0,4..Neq-1], U, V, Tg (old), CP, R, k (old)
                          (+ mu, lam, mu_t, l_min NS); writes A, B [Neq each], (F), Src [Neq-7], mu_t"""
    neq = {"keps": 9, "sa": 8}.get(model, 7)
    ns = model != "euler"
    ke = 1 if model == "keps" else 0
    k1a = (4 * neq + (neq - 7) + (neq if axi else 0) + ke) + (2 * neq + ke)
    k1b = (neq + 3) + (neq + 7 + (2 if ns else 0))
    # (SN[4..6] -- the species densities before chemistry -- are read only on tiles where chemistry changed them, Params::tile_inert:
    #  nowhere on the air-only bench grid, so they are not counted)
    k2 = (neq + 0 + 5 + (neq - 3) + 3 + 3 + (4 if ns else 0)) + (2 * neq + (neq if axi else 0) + (neq - 7) + (1 if ns else 0))
    return {"k1a": 8.0 * k1a, "k1b": 8.0 * k1b, "k2": 8.0 * k2}


def cpu_baseline(model):
    """CPU baseline beside the GPU number: the UNMODIFIED reference (oracle/_ref/ohf_serial, one core -- its OpenMP
    build does not scale) on a bounded sample; the oracle port when the reference binary is absent."""
    if model == "keps":
        r = run_reference_binary("ohf_serial", 512, 256, 60, 1)       # (one core: a quarter of the reference arm's grid keeps the default run short)
        if r is not None:
            return {"value": r[0], "unit": UNIT, "cores": 1, "kind": "reference",
                    "sample": f"unmodified reference, serial build (oracle/_ref/ohf_serial), three-cylinder k-eps deck 512x256 "
                              f"({r[1]} active cells), 60 steps, the solver's own 'average speed' line ({r[2]:.3g} step/s)"}
    return cpu_baseline_port(model)


# ---- the other BASELINE configs: grids drawn by the reference's OWN deck parser and rasterisers -------------------------------
# oracle/_ref/ohf_dump is the unmodified reference + a hook that writes the grid image and the step's scalars right before
# the first step (oracle/ref_hook.cpp); it travels to the GPU box with the snapshot.  Decks come from the fixtures' generator
# (tests/golden/make_golden.py), scaled up by `scale` with the geometry kept (dx = 1e-3 / scale).  Sizes are what the
# reference's start-up allows in a bench run (its wall-distance pass is O(nodes x walls)); each line states its own grid.
def _deck_cases():
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import make_golden as G
    C, GEO, R2 = G.cases(), G.geometry_cases(), G.r2_cases()
    return G, {
        # name: (BASELINE config, deck spec, base nx, base ny, scale, what differs from the named config)
        "c1": ("configs[0]: supersonic duct with narrowing, flat, inviscid, single component", C["euler_duct"]["deck"], 60, 40, 7, "420x280 (named ~400x100)"),
        "c2": ("configs[1]: shock reflection over a ramp in a duct, flat, viscous laminar", C["ns_laminar"]["deck"], 60, 40, 34, "2040x1360 (named ~2k x 500)"),
        "c3": ("configs[2]: NACA airfoil, flat, k-eps RANS in place of SST (no SST arithmetic in the reference)", GEO["airfoil_keps"]["deck"], 96, 64, 16,
               "1536x1024 (named ~4k x 2k: the reference's O(nodes x walls) start-up does not fit a bench run there)"),
        "c5": ("configs[4]: axisymmetric hydrogen jets burning in hot oxidiser, multicomponent reacting, k-eps in place of SST", R2["plain_axi_burn_keps"]["deck"], 40, 100, 40,
               "1600x4000 (named ~8k x 2k)"),
    }


def config_grid(name):
    """(J, cfg, layout, label) of BASELINE config `name`, drawn by the reference program itself."""
    import numpy as np
    from openhyperflow2d_b200 import aos, config
    G, T = _deck_cases()
    label, spec, bx, by, sc, note = T[name]
    nx, ny = bx * sc, by * sc
    spec = dict(spec, nx=nx, ny=ny, dx=1e-3 / sc, dy=1e-3 / sc, nmax=1)
    spec["delta_bl"] = spec.get("delta_bl", 1e-3)
    # area seeds are node indices: keep them where they were in physical space
    spec["areas"] = [dict(a, seed=(min(a["seed"][0] * sc, nx - 4) if a["seed"][0] < bx - 8 else nx - 4 * sc, a["seed"][1] * sc)) for a in spec["areas"]]
    exe = os.path.join(ROOT, "oracle", "_ref", "ohf_dump")
    if not os.path.exists(exe):
        raise SystemExit("bench.py --config needs oracle/_ref/ohf_dump (built by __graft_entry__.build() where the reference tree is present)")
    wd = tempfile.mkdtemp(prefix="ohf_cfg_")
    with open(os.path.join(wd, "cfg.dat"), "w") as f:
        f.write(G.make_deck("cfg", **spec))
    env = dict(os.environ, OHF_DUMP_DIR=wd, OHF_EXIT_TIME="1e-300", OHF_NSTEP="1", OHF_DUMP_ONLY="1")
    t0 = time.time()
    with open(os.path.join(wd, "log.txt"), "w") as log:
        subprocess.run([exe, "cfg.dat"], cwd=wd, env=env, stdout=log, stderr=subprocess.STDOUT, timeout=1500)
    if not os.path.exists(os.path.join(wd, "state0.hf2d")):
        raise SystemExit("ohf_dump left no grid image:\n" + open(os.path.join(wd, "log.txt")).read()[-2000:])
    lay = aos.parse_layout(os.path.join(wd, "layout.txt"))
    cfg = config.parse_cfg(os.path.join(wd, "cfg.txt"))
    J = np.fromfile(os.path.join(wd, "state0.hf2d"), dtype=aos.ref_dtype(lay)).reshape(nx, ny)
    for fn in os.listdir(wd):
        os.remove(os.path.join(wd, fn))
    return J, cfg, lay, f"BASELINE {label}; grid {note}; drawn by the reference's own deck parser and rasterisers in {time.time() - t0:.0f} s"


def run_config(args):
    """One JSON line for one of the other BASELINE configs on ONE GPU: device-timed steps on the reference-drawn grid."""
    import numpy as np
    import torch
    from openhyperflow2d_b200 import cabi
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- this framework has no CPU fallback")
    J, cfg, lay, label = config_grid(args.config)
    cfg["NOutStep"] = args.rms_every
    s = cabi.Solver(cfg, lay, device=0, rms_every=args.rms_every)
    s.upload_aos(J)
    act = s.active_cells()[0]
    W, K = max(args.warmup, 3), args.steps
    s.step(W)
    s.set_timing(True)
    with ClockSampler(0) as cs:
        rep = s.step_continue(K)
    kt = s.get_timing()
    n_plain, n_gen, tc, tr = s.tile_counts()
    bpc = s.bytes_per_cell()
    peak, peak_src = peaks()
    ms = rep.gpu_ms
    gbs = bpc * act * K / (ms * 1e-3) / 1e9
    line = {"metric": METRIC, "value": act * K / (ms * 1e-3), "unit": UNIT, "n_gpus": 1, "steps": K, "warmup": W, "ms_per_step": ms / K,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": label, "grid": f"{J.shape[0]}x{J.shape[1]}", "active_cells": act, "ProblemType": cfg["ProblemType"], "FT": cfg["FT"],
                       "TurbExtModel": cfg["TurbExtModel"], "tiles": {"plain": n_plain, "general": n_gen, "cols": tc, "rows": tr}, "rms_every": args.rms_every},
            "roofline": {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak, "traffic": None, "peak_source": peak_src,
                         "kernel": "whole step (algorithmic bytes per active cell, DESIGN.md)", "bytes_per_active_cell": bpc,
                         "plain_kernels_ms": {"k1a": kt["k1a_plain_ms"], "k1b": kt["k1b_plain_ms"], "k2": kt["k2_plain_ms"]}},
            "gpu_launches": int(rep.kernel_launches), "clocks": cs.summary(), "dt": rep.dt}
    s.close()
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--nx", type=int, default=NX_PER_GPU)
    ap.add_argument("--ny", type=int, default=NY)
    ap.add_argument("--model", default="keps")
    ap.add_argument("--seams", default="exact", choices=["exact", "lagged-chain"])
    ap.add_argument("--schedule", default="overlapped", choices=["overlapped", "sequential"],
                    help="N>1: seam exchange on a side stream beside the bulk tiles, or everything in sequence on one stream")
    ap.add_argument("--rms-every", type=int, default=100, help="NOutStep: residual norms reduced every n-th step and on the last one")
    ap.add_argument("--e2e-steps", type=int, default=500,
                    help="steps per output cycle (Nmax) of the end-to-end leg; the reference's own decks use 500 (TriplePoint.dat) to 24000 (Step.dat)")
    ap.add_argument("--config", default=None, choices=["c1", "c2", "c3", "c5"],
                    help="one of the other BASELINE configs on one GPU, grid drawn by the reference program (oracle/_ref/ohf_dump); default: configs[3] (c4)")
    ap.add_argument("--driver", default="c", choices=["c", "python"],
                    help="N>1: the C X-slab driver (direct stores into the neighbours' HBM; default) or the torch.distributed/NCCL protocol harness")
    ap.add_argument("--e2e-cycles", type=int, default=3, help="end-to-end cycles timed; the median is reported")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm(args)
    if args.config:
        return run_config(args)

    import numpy as np
    import torch
    from openhyperflow2d_b200 import cabi, gridgen
    from openhyperflow2d_b200.multigpu import SlabRunner

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- this framework has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        opts = dist.ProcessGroupNCCL.Options(is_high_priority_stream=True)   # seam traffic must not queue behind the bulk kernels
        dist.init_process_group("nccl", device_id=torch.device("cuda", local), pg_options=opts)
    parity = None
    use_c = world > 1 and args.driver == "c"
    if use_c:
        from openhyperflow2d_b200.multigpu import CSlabRunner, slab_parity_check
        # driver-side evidence that the multi-GPU path is exact: a small grid on all N GPUs vs the single-domain run, before timing
        ok = torch.tensor([1 if slab_parity_check(rank, world, local, model=args.model) else 0], dtype=torch.int64, device=torch.device("cuda", local))
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        parity = ("bit-exact" if int(ok.item()) == 1 else "MISMATCH") + f" (192x160 {args.model}, 7 steps, {world} X-slabs on {world} GPUs through the C slab driver vs the single-domain run; every rank's owned columns)"
        runner = CSlabRunner(args.nx, args.ny, model=args.model, rank=rank, world=world, device=local, rms_every=args.rms_every)
    else:
        runner = SlabRunner(args.nx, args.ny, model=args.model, rank=rank, world=world, device=local, seams=args.seams, rms_every=args.rms_every, schedule=args.schedule)
    act_local, tot_local = runner.active_cells()
    W, K = max(args.warmup, 3), args.steps
    runner.step(W)
    runner.barrier()
    runner.solver.set_timing(True)          # event pairs around the two plain-tile launches, inside the timed region
    with ClockSampler(local) as cs:
        ms = runner.timed_steps(K)          # device time (CUDA events on the solver's stream), max over ranks
    clocks = cs.summary()
    kt = runner.solver.get_timing()
    runner.solver.set_timing(False)
    if os.environ.get("HF2D_BENCH_RANKS") == "1":   # per-rank view: tile mix and the two timed launches
        print(f"[rank {rank}] tiles {runner.solver.tile_counts()} edge {runner.solver.edge_tile_counts()} active {act_local} "
              f"k1a {kt['k1a_plain_ms']:.3f} ms  k1b {kt['k1b_plain_ms']:.3f} ms  k2 {kt['k2_plain_ms']:.3f} ms", file=sys.stderr, flush=True)
    act = runner.allreduce_sum(act_local)
    value = act * K / (ms * 1e-3)
    bpc_step = runner.solver.bytes_per_cell()
    n_plain, n_gen, tc, tr = runner.solver.tile_counts()
    if world > 1:      # the timed launch covers the plain tiles that read no ghost column (hf2d_gpu_step_part S2_INNER)
        n_plain -= runner.solver.edge_tile_counts()[0]
    plain_cells = n_plain * tc * tr
    peak, peak_src = peaks()
    kb = kernel_bytes_per_cell(args.model)
    kms = {"k1a": kt["k1a_plain_ms"], "k1b": kt["k1b_plain_ms"], "k2": kt["k2_plain_ms"]}
    kname = {"k1a": "k_stage1_tma<plain tiles> (K1a: Stage 1 + residual + blending factor, TMA-staged)", "k1b": "k_point<plain tiles> (EOS, chemistry, tables, dt)",
             "k2": "k_flux<plain tiles> (gradients, turbulence, fluxes)"}
    kfrac = {q: (kb[q] * plain_cells / (kms[q] * 1e-3) / 1e9 / peak if kms[q] > 0 else 0.0) for q in kb}
    top = max(kms, key=lambda q: kms[q])          # the dominant kernel = the longest of the three
    achieved = kfrac[top] * peak
    traffic = None
    try:   # DRAM bytes per plain cell of that kernel, from the committed ncu capture (profiles/): a property of the profile, not of this run
        with open(os.path.join(ROOT, "profiles", "kernel_traffic.json")) as f:
            traffic = json.load(f).get(args.model, {}).get(top + "_dram_bytes_per_plain_cell") * plain_cells
    except Exception:
        pass
    step_gbs = bpc_step * act_local * K / (ms * 1e-3) / 1e9
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms / K,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD if args.model == "keps" else f"three cylinders, model {args.model}, flat",
                       "grid": f"{args.nx}x{args.ny} cells per GPU ({args.nx * world}x{args.ny} total), X-slabs", "model": args.model,
                       "active_cells": act, "total_cells": args.nx * args.ny * world, "seams": args.seams, "schedule": args.schedule if world > 1 else None, "rms_every": args.rms_every,
                       "multi_gpu_driver": (None if world == 1 else "C slab driver: peer-memory mailboxes over NVLink (hf2d_slab_run)" if use_c else "python SlabProtocol over NCCL"),
                       "tiles_per_gpu": {"plain": n_plain, "general": n_gen, "cols": tc, "rows": tr},
                       "l2": "per-step working set (~6 GB/GPU) >> 126 MB L2: no flush needed", "parallelism": f"x-slab{world}"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                         "traffic_source": "profiles/kernel_traffic.json (ncu capture of the same command, not this run)" if traffic else None,
                         "kernel": kname[top], "bytes_per_cell": kb[top], "cells_per_launch": plain_cells,
                         "kernel_ms": kms[top], "kernel_launches_timed": kt["launches"][0], "peak_source": peak_src,
                         "kernels": {q: {"ms": kms[q], "bytes_per_cell": kb[q], "frac": kfrac[q]} for q in kb},
                         "whole_step": {"bytes_per_active_cell": bpc_step, "achieved": step_gbs, "frac": step_gbs / peak}},
            "gpu_launches": int(runner.launches_last), "clocks": clocks}
    if rank == 0 and not args.no_cpu:
        line["cpu_baseline"] = cpu_baseline(args.model)
    if world == 1 and not args.no_cpu and args.model == "keps":
        # like for like with the reference arm: the same deck on the reference arm's grid (its start-up limits that grid,
        # reference_arm()), device-timed; 0.5 M cells fill a B200 only partly (launch-latency bound)
        Js, cs_, ls = gridgen.three_cylinders(REF_NX, REF_NY, model=args.model)
        ss = cabi.Solver(cs_, ls, device=local, rms_every=args.rms_every)
        ss.upload_aos(Js)
        ss.step(20)
        rs = ss.step_continue(200)
        a_s = ss.active_cells()[0]
        ss.close()
        line["same_grid_as_reference_arm"] = {"grid": f"{REF_NX}x{REF_NY}", "active_cells": a_s, "steps": 200, "ms_per_step": rs.gpu_ms / 200,
                                              "value": a_s * 200 / (rs.gpu_ms * 1e-3), "unit": UNIT}
    if world == 1 and not args.no_e2e:
        # the output step behind the hot loop: the body of the reference's Tecplot snapshot (SaveData2D, 16 "%g"
        # columns per node) formatted on the device from the state the timed steps left, and copied to the host
        try:
            import ctypes as C
            n = C.c_size_t()
            dxo, dyo = runner.cfg["dx"] * args.nx / (args.nx - 1), runner.cfg["dy"] * args.ny / (args.ny - 1)
            t0 = time.perf_counter()
            cabi._ck(cabi.lib().hf2d_gpu_tecplot_format(runner.solver._h, C.byref(runner.solver.layout), 0, dxo, dyo, C.byref(n)))
            t1 = time.perf_counter()
            buf = C.create_string_buffer(max(n.value, 1))
            t2 = time.perf_counter()
            cabi._ck(cabi.lib().hf2d_gpu_tecplot_fetch(runner.solver._h, buf, n.value))
            t3 = time.perf_counter()
            line["tecplot_snapshot"] = {"lines": args.nx * args.ny, "text_bytes": n.value, "format_ms": (t1 - t0) * 1e3, "fetch_ms": (t3 - t2) * 1e3,
                                        "note": "hf2d_gpu_tecplot_format + _fetch, FIRST call (includes allocating the line slots and loading the kernels; 33 ms repeated, profiles/r2_tecplot.json); byte-identical with the reference SaveData2D body (tests/test_zz_tecplot_gpu.py)"}
            del buf
        except Exception as e:      # never fails the bench line
            line["tecplot_snapshot"] = {"error": str(e)[:200]}
    if parity is not None:
        line["parity"] = parity
    if not args.no_e2e:
        if use_c:
            secs = runner.e2e_cycles(args.e2e_steps, cycles=args.e2e_cycles)
        else:
            secs = [runner.e2e_cycle(args.e2e_steps, cycles=1)["seconds"] for _ in range(max(1, args.e2e_cycles))]
        med = sorted(secs)[len(secs) // 2]
        line["e2e"] = {"value": act * args.e2e_steps / med, "unit": UNIT, "h2d_bytes_per_step": int(runner.J.nbytes / args.e2e_steps),
                       "d2h_bytes_per_step": int(runner.J.nbytes / args.e2e_steps), "steps_per_cycle": args.e2e_steps, "cycles_run": len(secs),
                       "seconds_per_cycle": secs, "seconds": med,
                       "note": "median cycle: pinned host AoS image -> HBM (scanned + unpacked there), nsteps steps (= Nmax, one output cycle), packed into the resident image -> host; per-GPU bytes"}
    if rank == 0:
        print(json.dumps(line))
    runner.close()
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

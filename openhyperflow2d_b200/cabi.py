"""ctypes binding of include/hf2d_gpu.h (libhf2d_gpu.so) -- the Python mirror of the reference-side shim.

Host language note: the reference is C++; its toolchain (g++) is present, but the reference has no FFI of its own,
so the seam is the C ABI in include/hf2d_gpu.h.  INTEGRATION.md shows the C++ shim a maintainer adds to
DEEPS2D_Run; this module is the same thing for Python callers (tests, bench) and never touches oracle/.
There is NO CPU fallback: every call fails loudly if the CUDA library is missing or no device is present.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import build as _build
from .config import TABLE_ORDER

NEQ, NTAB = 9, 14

LAYOUT_FIELDS = ["S", "dSdx", "dSdy", "TurbType", "l_min", "y_plus", "mu_t", "lam_t", "dkdx", "dkdy", "depsdx", "depsdy", "y", "p",
                 "idXl", "idYu", "idXr", "idYd", "CT", "beta", "Q_conv", "time", "k", "R", "lam", "mu", "CP", "Diff", "Tf",
                 "A", "B", "F", "Src", "SrcAdd", "Tg", "U", "V", "Y", "Uw", "Vw", "droYdx", "droYdy", "dUdx", "dUdy", "dVdx", "dVdy",
                 "dTdx", "dTdy", "BGX", "BGY"]


class Table(C.Structure):
    _fields_ = [("n", C.c_int), ("x", C.POINTER(C.c_double)), ("y", C.POINTER(C.c_double))]


class Cfg(C.Structure):
    _fields_ = [
        ("nx", C.c_int), ("ny", C.c_int), ("dx", C.c_double), ("dy", C.c_double), ("dt", C.c_double),
        ("CFL", C.c_double), ("beta0", C.c_double), ("nrbc_beta0", C.c_double),
        ("bFF", C.c_int), ("problem_type", C.c_int), ("flow_type", C.c_int), ("Hu", C.c_double * 4),
        ("TurbStartIter", C.c_int), ("TurbExtModel", C.c_int), ("isTurbulenceReset", C.c_int),
        ("SigW", C.c_double), ("SigF", C.c_double), ("delta_bl", C.c_double), ("turb_intensity", C.c_double),
        ("isAdiabaticWall", C.c_int), ("isAlternateRMS", C.c_int), ("AddSrcStartIter", C.c_uint), ("last_iter", C.c_uint),
        ("MonitorIndex", C.c_int), ("ExitMonitorValue", C.c_double), ("K0", C.c_double), ("Rgas", C.c_double * 4),
        ("tab", Table * NTAB), ("dt_rule", C.c_int), ("rank", C.c_int), ("n_ranks", C.c_int), ("NOutStep", C.c_int),
    ]


class NodeLayout(C.Structure):
    _fields_ = [("node_bytes", C.c_size_t)] + [(f, C.c_size_t) for f in LAYOUT_FIELDS]


class StepReport(C.Structure):
    _fields_ = [
        ("dt", C.c_double), ("dt_min", C.c_double), ("time_part", C.c_double), ("sum_RMS", C.c_double * NEQ),
        ("max_RMS", C.c_double), ("k_max_RMS", C.c_int), ("DD_max", C.c_double * NEQ), ("iter", C.c_uint),
        ("err", C.c_int), ("err_i", C.c_int), ("err_j", C.c_int), ("gpu_ms", C.c_double), ("kernel_launches", C.c_uint64),
    ]


class Hf2dError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"hf2d_gpu error {code}: {msg}")
        self.code = code


_lib = None
EXPORTS = ["hf2d_gpu_last_error", "hf2d_gpu_device_count", "hf2d_gpu_create", "hf2d_gpu_destroy", "hf2d_gpu_upload_aos",
           "hf2d_gpu_download_aos", "hf2d_gpu_pin_host", "hf2d_gpu_unpin_host", "hf2d_gpu_step", "hf2d_gpu_step_continue", "hf2d_gpu_read_node", "hf2d_gpu_run_cycle", "hf2d_gpu_field_ptr", "hf2d_gpu_active_cells",
           "hf2d_gpu_bytes_per_cell", "hf2d_gpu_tile_counts", "hf2d_gpu_halo_bytes", "hf2d_gpu_halo_pack", "hf2d_gpu_halo_unpack", "hf2d_gpu_step_begin",
           "hf2d_gpu_step_prepass", "hf2d_gpu_step_stage1", "hf2d_gpu_step_stage2", "hf2d_gpu_step_end", "hf2d_gpu_dt_ptr", "hf2d_gpu_stream",
           "hf2d_gpu_set_stream", "hf2d_gpu_set_timing", "hf2d_gpu_get_timing", "hf2d_gpu_dt_export", "hf2d_gpu_dt_import", "hf2d_gpu_launch_count",
           "hf2d_gpu_step_part", "hf2d_gpu_lane_fork", "hf2d_gpu_lane_join", "hf2d_gpu_side_stream", "hf2d_gpu_halo_pack_lane",
           "hf2d_gpu_halo_unpack_lane", "hf2d_gpu_edge_tile_counts", "hf2d_gpu_set_min_distance_to_wall", "hf2d_gpu_set_global_time", "hf2d_gpu_recalc_y_plus", "hf2d_gpu_residual_sums", "hf2d_gpu_set_min_distance_to_wall_slab", "hf2d_gpu_keep_gradients",
           "hf2d_gpu_wall_distance_noalloc", "hf2d_slab_cuts", "hf2d_slab_create", "hf2d_slab_destroy", "hf2d_slab_solver", "hf2d_slab_columns",
           "hf2d_slab_ipc_handle_bytes", "hf2d_slab_ipc_handle", "hf2d_slab_connect_ipc", "hf2d_slab_connect_local", "hf2d_slab_upload",
           "hf2d_slab_download", "hf2d_slab_upload_local", "hf2d_slab_download_local", "hf2d_slab_run", "hf2d_slab_read_node", "hf2d_gpu_tecplot_format", "hf2d_gpu_tecplot_fetch", "hf2d_slab_tecplot"]


def lib():
    global _lib
    if _lib is None:
        so = _build.build()
        L = C.CDLL(so)
        L.hf2d_gpu_last_error.restype = C.c_char_p
        L.hf2d_gpu_create.argtypes = [C.POINTER(Cfg), C.c_int, C.POINTER(C.c_void_p)]
        L.hf2d_gpu_destroy.argtypes = [C.c_void_p]
        L.hf2d_gpu_upload_aos.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(NodeLayout), C.c_int]
        L.hf2d_gpu_download_aos.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(NodeLayout), C.c_int]
        L.hf2d_gpu_pin_host.argtypes = [C.c_void_p, C.c_size_t]
        L.hf2d_gpu_unpin_host.argtypes = [C.c_void_p]
        L.hf2d_gpu_step.argtypes = [C.c_void_p, C.c_int, C.POINTER(StepReport)]
        L.hf2d_gpu_step_continue.argtypes = [C.c_void_p, C.c_int, C.POINTER(StepReport)]
        L.hf2d_gpu_read_node.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.hf2d_gpu_run_cycle.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(NodeLayout), C.c_int, C.c_int, C.POINTER(StepReport)]
        L.hf2d_gpu_field_ptr.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.POINTER(C.c_void_p)]
        L.hf2d_gpu_active_cells.argtypes = [C.c_void_p, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong)]
        L.hf2d_gpu_tile_counts.argtypes = [C.c_void_p] + [C.POINTER(C.c_int)] * 4
        L.hf2d_gpu_bytes_per_cell.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        L.hf2d_gpu_halo_bytes.argtypes = [C.c_void_p, C.c_int]
        L.hf2d_gpu_halo_bytes.restype = C.c_size_t
        L.hf2d_gpu_halo_pack.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.hf2d_gpu_halo_unpack.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.hf2d_gpu_halo_pack_lane.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.hf2d_gpu_halo_unpack_lane.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.hf2d_gpu_step_part.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.hf2d_gpu_lane_fork.argtypes = [C.c_void_p]
        L.hf2d_gpu_lane_join.argtypes = [C.c_void_p]
        L.hf2d_gpu_side_stream.argtypes = [C.c_void_p]
        L.hf2d_gpu_side_stream.restype = C.c_void_p
        L.hf2d_gpu_edge_tile_counts.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        for f in ("hf2d_gpu_step_begin", "hf2d_gpu_step_prepass", "hf2d_gpu_step_stage1", "hf2d_gpu_step_stage2"):
            getattr(L, f).argtypes = [C.c_void_p]
        L.hf2d_gpu_step_end.argtypes = [C.c_void_p, C.POINTER(StepReport)]
        L.hf2d_gpu_dt_ptr.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
        L.hf2d_gpu_stream.argtypes = [C.c_void_p]
        L.hf2d_gpu_stream.restype = C.c_void_p
        L.hf2d_gpu_set_stream.argtypes = [C.c_void_p, C.c_void_p]
        L.hf2d_gpu_dt_export.argtypes = [C.c_void_p, C.c_void_p]
        L.hf2d_gpu_dt_import.argtypes = [C.c_void_p, C.c_void_p]
        L.hf2d_gpu_set_timing.argtypes = [C.c_void_p, C.c_int]
        L.hf2d_gpu_get_timing.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_int)]
        L.hf2d_gpu_launch_count.argtypes = [C.c_void_p]
        L.hf2d_gpu_set_global_time.argtypes = [C.c_void_p, C.c_double]
        L.hf2d_gpu_set_min_distance_to_wall_slab.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                                             C.c_void_p, C.c_void_p, C.c_void_p]
        L.hf2d_gpu_residual_sums.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.hf2d_gpu_recalc_y_plus.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.hf2d_gpu_set_min_distance_to_wall.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_double, C.c_void_p, C.c_void_p,
                                                        C.c_void_p, C.c_void_p]
        L.hf2d_gpu_launch_count.restype = C.c_ulonglong
        L.hf2d_gpu_keep_gradients.argtypes = [C.c_void_p, C.c_int]
        L.hf2d_gpu_wall_distance_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.hf2d_gpu_tecplot_format.argtypes = [C.c_void_p, C.POINTER(NodeLayout), C.c_int, C.c_double, C.c_double, C.POINTER(C.c_size_t)]
        L.hf2d_gpu_tecplot_fetch.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.hf2d_slab_tecplot.argtypes = [C.c_void_p, C.c_int, C.POINTER(NodeLayout), C.c_int, C.c_double, C.c_double, C.c_void_p, C.POINTER(C.c_size_t)]
        L.hf2d_slab_cuts.argtypes = [C.c_void_p, C.POINTER(NodeLayout), C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.hf2d_slab_create.argtypes = [C.POINTER(Cfg), C.c_int, C.c_int, C.c_int, C.c_void_p, C.POINTER(C.c_void_p)]
        L.hf2d_slab_destroy.argtypes = [C.c_void_p]
        L.hf2d_slab_solver.argtypes = [C.c_void_p]; L.hf2d_slab_solver.restype = C.c_void_p
        L.hf2d_slab_columns.argtypes = [C.c_void_p] + [C.POINTER(C.c_int)] * 4
        L.hf2d_slab_ipc_handle_bytes.restype = C.c_size_t
        L.hf2d_slab_ipc_handle.argtypes = [C.c_void_p, C.c_void_p]
        L.hf2d_slab_connect_ipc.argtypes = [C.c_void_p, C.c_void_p]
        L.hf2d_slab_connect_local.argtypes = [C.c_void_p, C.c_int]
        for f in (L.hf2d_slab_upload, L.hf2d_slab_download, L.hf2d_slab_upload_local, L.hf2d_slab_download_local):
            f.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(NodeLayout), C.c_int]
        L.hf2d_slab_run.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(StepReport)]
        L.hf2d_slab_read_node.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        _lib = L
    return _lib


def _ck(rc):
    if rc != 0:
        raise Hf2dError(rc, lib().hf2d_gpu_last_error().decode())


def make_layout(layout: dict) -> NodeLayout:
    nl = NodeLayout()
    nl.node_bytes = layout["node_bytes"]
    for f in LAYOUT_FIELDS:
        setattr(nl, f, layout["fields"][f][0])
    return nl


def make_cfg(cfg: dict, nx=None, ny=None, rank=0, n_ranks=1, dt_rule=0, rms_every=None):
    c = Cfg()
    keep = []
    c.nx, c.ny = nx or cfg["MaxX"], ny or cfg["MaxY"]
    c.dx, c.dy, c.dt = cfg["dx"], cfg["dy"], cfg["dt"]
    c.CFL, c.beta0, c.nrbc_beta0 = cfg["CFL"], cfg["beta0"], cfg["nrbc_beta0"]
    c.bFF, c.problem_type, c.flow_type = cfg["bFF"], cfg["ProblemType"], cfg["FT"]
    for i in range(4):
        c.Hu[i] = cfg["Hu"][i]
    c.TurbStartIter, c.TurbExtModel, c.isTurbulenceReset = cfg["TurbStartIter"], cfg["TurbExtModel"], cfg["isTurbulenceReset"]
    c.SigW, c.SigF, c.delta_bl, c.turb_intensity = cfg["SigW"], cfg["SigF"], cfg["delta_bl"], cfg.get("TurbIntensity", 0.005)
    c.isAdiabaticWall, c.isAlternateRMS = cfg["isAdiabaticWall"], cfg["isAlternateRMS"]
    c.AddSrcStartIter, c.last_iter = cfg.get("AddSrcStartIter", 0), cfg.get("last_iter", 0)
    c.MonitorIndex, c.ExitMonitorValue, c.K0 = cfg.get("MonitorIndex", 5), cfg.get("ExitMonitorValue", 0.0), cfg["K0"]
    for i, k in enumerate(["R_Fuel", "R_OX", "R_cp", "R_air"]):
        c.Rgas[i] = cfg[k]
    for i, name in enumerate(TABLE_ORDER):
        x, y = cfg["tables"].get(name, ([], []))
        xa = (C.c_double * max(len(x), 1))(*x)
        ya = (C.c_double * max(len(y), 1))(*y)
        keep += [xa, ya]
        c.tab[i].n = len(x)
        c.tab[i].x = C.cast(xa, C.POINTER(C.c_double))
        c.tab[i].y = C.cast(ya, C.POINTER(C.c_double))
    c.dt_rule, c.rank, c.n_ranks = dt_rule, rank, n_ranks
    c.NOutStep = int(cfg.get("NOutStep", 1)) if rms_every is None else int(rms_every)
    return c, keep


class Solver:
    """One X-slab of the DEEPS2D step on one GPU (hf2d_gpu handle)."""

    def __init__(self, cfg: dict, layout: dict, device: int = 0, **kw):
        self._h = C.c_void_p()
        self.cfg = cfg
        self.layout = make_layout(layout)
        c, self._keep = make_cfg(cfg, **kw)
        self.nx, self.ny = c.nx, c.ny
        _ck(lib().hf2d_gpu_create(C.byref(c), device, C.byref(self._h)))

    def close(self):
        if self._h:
            lib().hf2d_gpu_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @staticmethod
    def _ptr(J):
        if isinstance(J, np.ndarray):
            assert J.flags.c_contiguous
            return J.ctypes.data
        return int(J)

    def upload_aos(self, J, pinned=False):
        _ck(lib().hf2d_gpu_upload_aos(self._h, self._ptr(J), C.byref(self.layout), int(pinned)))

    def download_aos(self, J, pinned=False):
        _ck(lib().hf2d_gpu_download_aos(self._h, self._ptr(J), C.byref(self.layout), int(pinned)))

    def step(self, nsteps: int) -> StepReport:
        rep = StepReport()
        _ck(lib().hf2d_gpu_step(self._h, nsteps, C.byref(rep)))
        return rep

    def step_continue(self, nsteps: int) -> StepReport:
        rep = StepReport()
        _ck(lib().hf2d_gpu_step_continue(self._h, nsteps, C.byref(rep)))
        return rep

    def run_cycle(self, J, nsteps: int, pinned=False) -> StepReport:
        rep = StepReport()
        _ck(lib().hf2d_gpu_run_cycle(self._h, self._ptr(J), C.byref(self.layout), int(pinned), nsteps, C.byref(rep)))
        return rep

    def field_ptr(self, name: str, k: int = 0) -> int:
        p = C.c_void_p()
        _ck(lib().hf2d_gpu_field_ptr(self._h, name.encode(), k, C.byref(p)))
        return p.value

    def active_cells(self):
        a, t = C.c_longlong(), C.c_longlong()
        _ck(lib().hf2d_gpu_active_cells(self._h, C.byref(a), C.byref(t)))
        return a.value, t.value

    def tile_counts(self):
        v = [C.c_int() for _ in range(4)]
        _ck(lib().hf2d_gpu_tile_counts(self._h, *[C.byref(x) for x in v]))
        return tuple(x.value for x in v)

    def bytes_per_cell(self) -> float:
        b = C.c_double()
        _ck(lib().hf2d_gpu_bytes_per_cell(self._h, C.byref(b)))
        return b.value

    # --- split step / halo API (multi-GPU drivers) ---
    def set_stream(self, cuda_stream: int):
        _ck(lib().hf2d_gpu_set_stream(self._h, C.c_void_p(cuda_stream)))

    def halo_bytes(self, phase: int) -> int:
        return lib().hf2d_gpu_halo_bytes(self._h, phase)

    def halo_pack(self, side: int, phase: int, dev_ptr: int, lane: int = 0):
        _ck(lib().hf2d_gpu_halo_pack_lane(self._h, side, phase, C.c_void_p(dev_ptr), lane))

    def halo_unpack(self, side: int, phase: int, dev_ptr: int, lane: int = 0):
        _ck(lib().hf2d_gpu_halo_unpack_lane(self._h, side, phase, C.c_void_p(dev_ptr), lane))

    # one step in parts on two stream lanes (include/hf2d_gpu.h "overlapped X-slab step")
    PREPASS, S1_PLAIN, S1_GENERAL, S1_FINISH, S2_INNER, S2_EDGE, S2_FINISH, PREPASS_LOCAL, AHEAD = 1, 2, 3, 4, 5, 6, 7, 8, 9

    def step_part(self, part: int, lane: int = 0, want_rms: int = 1):
        _ck(lib().hf2d_gpu_step_part(self._h, part, lane, want_rms))

    def lane_fork(self):
        _ck(lib().hf2d_gpu_lane_fork(self._h))

    def lane_join(self):
        _ck(lib().hf2d_gpu_lane_join(self._h))

    def side_stream(self) -> int:
        return int(lib().hf2d_gpu_side_stream(self._h) or 0)

    def edge_tile_counts(self):
        a, b = C.c_int(), C.c_int()
        _ck(lib().hf2d_gpu_edge_tile_counts(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def step_begin(self):
        _ck(lib().hf2d_gpu_step_begin(self._h))

    def step_prepass(self):
        _ck(lib().hf2d_gpu_step_prepass(self._h))

    def step_stage1(self):
        _ck(lib().hf2d_gpu_step_stage1(self._h))

    def step_stage2(self):
        _ck(lib().hf2d_gpu_step_stage2(self._h))

    def step_end(self, want_report=False):
        if want_report:
            rep = StepReport()
            _ck(lib().hf2d_gpu_step_end(self._h, C.byref(rep)))
            return rep
        _ck(lib().hf2d_gpu_step_end(self._h, None))

    def dt_export(self, dev_ptr: int):
        _ck(lib().hf2d_gpu_dt_export(self._h, C.c_void_p(dev_ptr)))

    def dt_import(self, dev_ptr: int):
        _ck(lib().hf2d_gpu_dt_import(self._h, C.c_void_p(dev_ptr)))

    def set_timing(self, enable=True):
        _ck(lib().hf2d_gpu_set_timing(self._h, int(enable)))

    def get_timing(self):
        ms, cnt = (C.c_double * 3)(), (C.c_int * 3)()
        _ck(lib().hf2d_gpu_get_timing(self._h, ms, cnt))
        return {"k1a_plain_ms": ms[0], "k1b_plain_ms": ms[1], "k2_plain_ms": ms[2], "launches": (cnt[0], cnt[1], cnt[2])}

    def launch_count(self) -> int:
        return int(lib().hf2d_gpu_launch_count(self._h))

    def set_global_time(self, global_time: float):
        """GlobalTime of the running cycle: download_aos then writes FlowNode2D::time of the swept nodes (deeps2d_core.cpp:867)."""
        _ck(lib().hf2d_gpu_set_global_time(self._h, float(global_time)))

    def keep_gradients(self, enable=True):
        """The last step of every step call keeps the 18 gradient members for the download (restart fidelity); default on."""
        _ck(lib().hf2d_gpu_keep_gradients(self._h, int(enable)))

    def residual_sums(self):
        """Raw residual accumulators of the last reducing step: (rms, sum_div, count, dd_max), NEQ entries each."""
        rms = np.zeros(NEQ); sd = np.zeros(NEQ); cnt = np.zeros(NEQ, dtype=np.uint64); dd = np.zeros(NEQ)
        _ck(lib().hf2d_gpu_residual_sums(self._h, rms.ctypes.data, sd.ctypes.data, cnt.ctypes.data, dd.ctypes.data))
        return rms, sd, cnt, dd

    def recalc_y_plus(self, i_wall, j_wall):
        """Recalc_y_plus(J, WallNodes) on the uploaded grid (deeps2d_core.cpp:2364-2386): returns the y_plus plane (nx, ny);
        the device plane the step reads is updated in place."""
        i_wall = np.ascontiguousarray(i_wall, dtype=np.int32); j_wall = np.ascontiguousarray(j_wall, dtype=np.int32)
        assert i_wall.shape == (self.nx, self.ny) and j_wall.shape == (self.nx, self.ny)
        y = np.empty((self.nx, self.ny), dtype=np.float64)
        _ck(lib().hf2d_gpu_recalc_y_plus(self._h, i_wall.ctypes.data, j_wall.ctypes.data, y.ctypes.data))
        return y

    def set_min_distance_to_wall_slab(self, walls, i_off: int, nx_total: int):
        """The slab form: `walls` and the returned i_wall are GLOBAL columns, i_off = global column of local column 0."""
        walls = np.ascontiguousarray(walls, dtype=np.int32).reshape(-1, 2)
        nx, ny = self.nx, self.ny
        l_min = np.empty((nx, ny), dtype=np.float64)
        i_wall = np.empty((nx, ny), dtype=np.int32); j_wall = np.empty((nx, ny), dtype=np.int32)
        status = np.empty((nx, ny), dtype=np.uint8)
        _ck(lib().hf2d_gpu_set_min_distance_to_wall_slab(self._h, walls.ctypes.data, len(walls), int(i_off), int(nx_total),
                                                         l_min.ctypes.data, i_wall.ctypes.data, j_wall.ctypes.data, status.ctypes.data))
        return l_min, i_wall, j_wall, status

    def wall_distance_noalloc(self, walls, dev_l_min: int, dev_i_wall: int, dev_j_wall: int, dev_status: int, x0: float = 0.0):
        """hf2d_gpu_wall_distance_noalloc: the same pass without any allocation; the four result planes are DEVICE pointers
        (nx*ny doubles / int32 / int32 / uint8) of the caller.  At most 16384 wall nodes."""
        walls = np.ascontiguousarray(walls, dtype=np.int32).reshape(-1, 2)
        _ck(lib().hf2d_gpu_wall_distance_noalloc(self._h, walls.ctypes.data, len(walls), float(x0), dev_l_min, dev_i_wall, dev_j_wall, dev_status))

    def tecplot_body(self, type_: int, dx_out: float, dy_out: float) -> bytes:
        """hf2d_gpu_tecplot_format + _fetch: the body of SaveData2D for the current state, formatted on the device."""
        n = C.c_size_t()
        _ck(lib().hf2d_gpu_tecplot_format(self._h, C.byref(self.layout), int(type_), float(dx_out), float(dy_out), C.byref(n)))
        buf = C.create_string_buffer(max(n.value, 1))
        _ck(lib().hf2d_gpu_tecplot_fetch(self._h, buf, n.value))
        return buf.raw[:n.value]

    def set_min_distance_to_wall(self, walls, x0: float = 0.0):
        """SetMinDistanceToWall2D(J, WallNodes, x0) on the uploaded grid (deeps2d_core.cpp:4783-4832).  walls: (n, 2) int
        array of (i, j) in GetWallNodes order.  Returns (l_min, i_wall, j_wall, status) planes of shape (nx, ny); the
        device l_min plane is updated in place."""
        walls = np.ascontiguousarray(walls, dtype=np.int32).reshape(-1, 2)
        nx, ny = self.nx, self.ny
        l_min = np.empty((nx, ny), dtype=np.float64)
        i_wall = np.empty((nx, ny), dtype=np.int32); j_wall = np.empty((nx, ny), dtype=np.int32)
        status = np.empty((nx, ny), dtype=np.uint8)
        _ck(lib().hf2d_gpu_set_min_distance_to_wall(self._h, walls.ctypes.data, len(walls), float(x0), l_min.ctypes.data,
                                                    i_wall.ctypes.data, j_wall.ctypes.data, status.ctypes.data))
        return l_min, i_wall, j_wall, status


def slab_cuts(J, layout: dict, world: int, min_cols: int = 2):
    """hf2d_slab_cuts: X-slab boundaries of the host image J (nx, ny records) with equal numbers of active nodes."""
    nx, ny = J.shape
    cuts = (C.c_int * (world + 1))()
    lay = make_layout(layout)
    _ck(lib().hf2d_slab_cuts(J.ctypes.data, C.byref(lay), nx, ny, world, min_cols, cuts))
    return list(cuts)


class _BorrowedSolver(Solver):
    """The hf2d_gpu handle inside a slab: same calls as Solver, owned (and destroyed) by the slab."""

    def __init__(self, handle, cfg, layout, nx, ny):
        self._h = C.c_void_p(handle)
        self.cfg, self.layout, self.nx, self.ny = cfg, layout, nx, ny

    def close(self):
        self._h = C.c_void_p()


class Slab:
    """One rank of the C X-slab driver (csrc/hf2d_slab.inc): ghost columns by direct stores into the neighbour's HBM."""

    def __init__(self, cfg_global: dict, layout: dict, device: int, rank: int, world: int, cuts, **kw):
        self._s = C.c_void_p()
        self.layout = make_layout(layout)
        c, self._keep = make_cfg(cfg_global, **kw)
        self.rank, self.world, self.device = rank, world, device
        arr = (C.c_int * (world + 1))(*[int(v) for v in cuts])
        _ck(lib().hf2d_slab_create(C.byref(c), device, rank, world, arr, C.byref(self._s)))
        v = [C.c_int() for _ in range(4)]
        _ck(lib().hf2d_slab_columns(self._s, *[C.byref(x) for x in v]))
        self.c0, self.c1, self.gl, self.gr = (x.value for x in v)
        self.solver = _BorrowedSolver(lib().hf2d_slab_solver(self._s), cfg_global, self.layout, self.c1 - self.c0 + self.gl + self.gr, c.ny)

    def close(self):
        if self._s:
            self.solver.close()
            lib().hf2d_slab_destroy(self._s)
            self._s = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def ipc_handle(self) -> bytes:
        n = lib().hf2d_slab_ipc_handle_bytes()
        buf = C.create_string_buffer(n)
        _ck(lib().hf2d_slab_ipc_handle(self._s, buf))
        return buf.raw

    def connect_ipc(self, handles: bytes):
        assert len(handles) == self.world * lib().hf2d_slab_ipc_handle_bytes()
        _ck(lib().hf2d_slab_connect_ipc(self._s, handles))

    @staticmethod
    def connect_local(slabs):
        arr = (C.c_void_p * len(slabs))(*[s._s for s in slabs])
        _ck(lib().hf2d_slab_connect_local(arr, len(slabs)))

    def upload(self, J_global, pinned=False):
        _ck(lib().hf2d_slab_upload(self._s, Solver._ptr(J_global), C.byref(self.layout), int(pinned)))

    def download(self, J_global, pinned=False):
        _ck(lib().hf2d_slab_download(self._s, Solver._ptr(J_global), C.byref(self.layout), int(pinned)))

    def upload_local(self, J_slab, pinned=False):
        _ck(lib().hf2d_slab_upload_local(self._s, Solver._ptr(J_slab), C.byref(self.layout), int(pinned)))

    def download_local(self, J_slab, pinned=False):
        _ck(lib().hf2d_slab_download_local(self._s, Solver._ptr(J_slab), C.byref(self.layout), int(pinned)))

    @staticmethod
    def run_all(slabs, nsteps: int, new_cycle: bool = True) -> StepReport:
        """hf2d_slab_run on the slabs this process drives (all ranks in one process, or [mine])."""
        arr = (C.c_void_p * len(slabs))(*[s._s for s in slabs])
        rep = StepReport()
        _ck(lib().hf2d_slab_run(arr, len(slabs), nsteps, int(new_cycle), C.byref(rep)))
        return rep

    def run(self, nsteps: int, new_cycle: bool = True) -> StepReport:
        return Slab.run_all([self], nsteps, new_cycle)

    @staticmethod
    def tecplot_body(slabs, type_: int, dx_out: float, dy_out: float) -> bytes:
        """hf2d_slab_tecplot: the body of SaveData2D for the whole grid, every slab formatted on its own GPU."""
        arr = (C.c_void_p * len(slabs))(*[s._s for s in slabs])
        n = C.c_size_t()
        _ck(lib().hf2d_slab_tecplot(arr, len(slabs), C.byref(slabs[0].layout), int(type_), float(dx_out), float(dy_out), None, C.byref(n)))
        buf = C.create_string_buffer(max(n.value, 1))
        _ck(lib().hf2d_slab_tecplot(arr, len(slabs), C.byref(slabs[0].layout), int(type_), float(dx_out), float(dy_out), buf, C.byref(n)))
        return buf.raw[:n.value]

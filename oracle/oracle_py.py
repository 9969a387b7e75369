"""TEST INFRASTRUCTURE ONLY: ctypes wrapper of oracle/libdeeps2d_oracle.so (the CPU restatement).

Imported only by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
The product package (openhyperflow2d_b200/) must never import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
NEQ, NCOMP, NTAB = 9, 3, 14

# mirrors `struct ohf_node` (deeps2d_oracle.h): every member is 8 bytes, so there is no padding
_NODE_FIELDS = [
    ("S", "<f8", NEQ), ("dSdx", "<f8", NEQ), ("dSdy", "<f8", NEQ), ("SN", "<f8", NEQ),
    ("TurbType", "<u8", 1),
    ("l_min", "<f8", 1), ("y_plus", "<f8", 1), ("Re_local", "<f8", 1), ("mu_t", "<f8", 1), ("lam_t", "<f8", 1),
    ("dkdx", "<f8", 1), ("dkdy", "<f8", 1), ("depsdx", "<f8", 1), ("depsdy", "<f8", 1),
    ("y", "<f8", 1), ("p", "<f8", 1),
    ("idXl", "<i8", 1), ("idYu", "<i8", 1), ("idXr", "<i8", 1), ("idYd", "<i8", 1),
    ("CT", "<u8", 1),
    ("beta", "<f8", NEQ),
    ("Q_conv", "<f8", 1), ("time", "<f8", 1), ("k", "<f8", 1), ("R", "<f8", 1), ("lam", "<f8", 1), ("mu", "<f8", 1),
    ("CP", "<f8", 1), ("Diff", "<f8", 1), ("Tf", "<f8", 1),
    ("A", "<f8", NEQ), ("B", "<f8", NEQ), ("F", "<f8", NEQ), ("RX", "<f8", NEQ), ("RY", "<f8", NEQ),
    ("Src", "<f8", NEQ), ("SrcAdd", "<f8", NEQ),
    ("Tg", "<f8", 1), ("U", "<f8", 1), ("V", "<f8", 1), ("Y", "<f8", NCOMP + 1),
    ("Uw", "<f8", 1), ("Vw", "<f8", 1),
    ("droYdx", "<f8", NCOMP + 1), ("droYdy", "<f8", NCOMP + 1),
    ("dUdx", "<f8", 1), ("dUdy", "<f8", 1), ("dVdx", "<f8", 1), ("dVdy", "<f8", 1), ("dTdx", "<f8", 1), ("dTdy", "<f8", 1),
    ("BGX", "<f8", 1), ("BGY", "<f8", 1),
]
NODE_DTYPE = np.dtype([(n, t) if c == 1 else (n, t, (c,)) for n, t, c in _NODE_FIELDS])


class Table(C.Structure):
    _fields_ = [("n", C.c_int), ("x", C.POINTER(C.c_double)), ("y", C.POINTER(C.c_double))]


class Cfg(C.Structure):
    _fields_ = [
        ("nx", C.c_int), ("ny", C.c_int), ("dx", C.c_double), ("dy", C.c_double),
        ("CFL", C.c_double), ("beta0", C.c_double), ("nrbc_beta0", C.c_double),
        ("bFF", C.c_int), ("problem_type", C.c_int), ("flow_type", C.c_int),
        ("Hu", C.c_double * 4),
        ("TurbStartIter", C.c_int), ("TurbExtModel", C.c_int), ("isTurbulenceReset", C.c_int),
        ("SigW", C.c_double), ("SigF", C.c_double), ("delta_bl", C.c_double), ("turb_intensity", C.c_double),
        ("isAdiabaticWall", C.c_int), ("isAlternateRMS", C.c_int),
        ("AddSrcStartIter", C.c_uint), ("MonitorIndex", C.c_int),
        ("ExitMonitorValue", C.c_double), ("K0", C.c_double), ("Rgas", C.c_double * 4),
        ("tab", Table * NTAB),
    ]


class StepReport(C.Structure):
    _fields_ = [
        ("dt", C.c_double), ("dt_min", C.c_double), ("sum_RMS", C.c_double * NEQ), ("max_RMS", C.c_double),
        ("k_max_RMS", C.c_int), ("DD_max", C.c_double * NEQ), ("i_c", C.c_int), ("j_c", C.c_int),
        ("err", C.c_int), ("err_i", C.c_int), ("err_j", C.c_int),
    ]


_lib = None


def build(force: bool = False) -> str:
    so = os.path.join(HERE, "libdeeps2d_oracle.so")
    src = os.path.join(HERE, "deeps2d_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", HERE, "oracle"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.ohf_node_bytes.restype = C.c_int
        assert _lib.ohf_node_bytes() == NODE_DTYPE.itemsize, (_lib.ohf_node_bytes(), NODE_DTYPE.itemsize)
        _lib.ohf_table_getval.restype = C.c_double
        _lib.ohf_table_getval.argtypes = [C.POINTER(Table), C.c_double]
        _lib.ohf_step.restype = C.c_int
        _lib.ohf_step.argtypes = [C.c_void_p, C.POINTER(Cfg), C.c_uint, C.c_int, C.c_double,
                                  C.POINTER(C.c_double), C.POINTER(StepReport)]
        _lib.ohf_run_cycle.restype = C.c_int
        _lib.ohf_run_cycle.argtypes = [C.c_void_p, C.POINTER(Cfg), C.c_uint, C.c_int, C.c_double,
                                       C.POINTER(C.c_double), C.POINTER(StepReport), C.POINTER(C.c_double)]
    return _lib


TABLE_ORDER = ["Cp_Fuel", "Cp_OX", "Cp_cp", "Cp_air", "lam_Fuel", "lam_OX", "lam_cp", "lam_air",
               "mu_Fuel", "mu_OX", "mu_cp", "mu_air", "CFL_Scenario", "beta_Scenario"]


def make_cfg(cfg: dict):
    """cfg dict (openhyperflow2d_b200.config.parse_cfg) -> (Cfg, keepalive)."""
    c = Cfg()
    keep = []
    c.nx, c.ny, c.dx, c.dy = cfg["MaxX"], cfg["MaxY"], cfg["dx"], cfg["dy"]
    c.CFL, c.beta0, c.nrbc_beta0 = cfg["CFL"], cfg["beta0"], cfg["nrbc_beta0"]
    c.bFF, c.problem_type, c.flow_type = cfg["bFF"], cfg["ProblemType"], cfg["FT"]
    for i in range(4):
        c.Hu[i] = cfg["Hu"][i]
    c.TurbStartIter, c.TurbExtModel, c.isTurbulenceReset = cfg["TurbStartIter"], cfg["TurbExtModel"], cfg["isTurbulenceReset"]
    c.SigW, c.SigF, c.delta_bl = cfg["SigW"], cfg["SigF"], cfg["delta_bl"]
    c.turb_intensity = cfg.get("TurbIntensity", 0.005)
    c.isAdiabaticWall, c.isAlternateRMS = cfg["isAdiabaticWall"], cfg["isAlternateRMS"]
    c.AddSrcStartIter, c.MonitorIndex = cfg.get("AddSrcStartIter", 0), cfg.get("MonitorIndex", 5)
    c.ExitMonitorValue, c.K0 = cfg.get("ExitMonitorValue", 0.0), cfg["K0"]
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
    return c, keep


def nodes_from_ref(aos: np.ndarray) -> np.ndarray:
    """reference AoS record array (nx, ny) -> oracle node array (nx, ny).

    SN (the FlowNodeCore2D shadow, "NextNode") starts as zeros: hf2d_start.cpp:318 allocates it with
    UMatrix2D's zero-filling constructor (utl/umatrix2d.hpp:178) and it then persists across steps, so
    an equation whose Stage-1 update is skipped but whose Stage-2 commit is not receives 0 (or the
    last value written)."""
    out = np.zeros(aos.shape, dtype=NODE_DTYPE)
    for name in NODE_DTYPE.names:
        if name == "SN":
            continue
        else:
            out[name] = aos[name]
    return out


def step(nodes: np.ndarray, cfg: dict, global_iter: int, is_src_add: int, dt_min: float, global_time: float = 0.0):
    c, keep = make_cfg(cfg)
    assert nodes.dtype == NODE_DTYPE and nodes.flags.c_contiguous
    dm = C.c_double(dt_min)
    rep = StepReport()
    rc = lib().ohf_step(nodes.ctypes.data, C.byref(c), global_iter, is_src_add, global_time, C.byref(dm), C.byref(rep))
    return rc, dm.value, rep


def run_cycle(nodes: np.ndarray, cfg: dict, last_iter: int, nsteps: int, dt_min: float, global_time: float = 0.0):
    c, keep = make_cfg(cfg)
    assert nodes.dtype == NODE_DTYPE and nodes.flags.c_contiguous
    dm = C.c_double(dt_min)
    tp = C.c_double(0.0)
    rep = StepReport()
    rc = lib().ohf_run_cycle(nodes.ctypes.data, C.byref(c), last_iter, nsteps, global_time, C.byref(dm), C.byref(rep), C.byref(tp))
    return rc, dm.value, rep, tp.value


def wall_nodes(ct: np.ndarray) -> np.ndarray:
    """GetWallNodes (deeps2d_core.cpp:2081-2104): (n, 2) int32 array of (i, j), rows outermost."""
    ct = np.ascontiguousarray(ct, dtype=np.uint64)
    nx, ny = ct.shape
    L = lib()
    L.ohf_get_wall_nodes.restype = C.c_long
    L.ohf_get_wall_nodes.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
    n = L.ohf_get_wall_nodes(ct.ctypes.data, nx, ny, None)
    out = np.zeros((max(n, 1), 2), dtype=np.int32)
    L.ohf_get_wall_nodes(ct.ctypes.data, nx, ny, out.ctypes.data)
    return out[:n]


def set_min_distance_to_wall(ct, Tg, p, dx, dy, walls, l_min, i_wall, j_wall, x0=0.0):
    """SetMinDistanceToWall2D (deeps2d_core.cpp:4783-4832) on x-major planes.  ct / l_min / i_wall / j_wall are updated in
    place (contiguous uint64 / float64 / int32 / int32); returns the status plane (0 untouched, 1 computed, 2 made solid)."""
    nx, ny = ct.shape
    for a, t in ((ct, np.uint64), (l_min, np.float64), (i_wall, np.int32), (j_wall, np.int32)):
        assert a.dtype == t and a.flags.c_contiguous and a.shape == (nx, ny)
    Tg = np.ascontiguousarray(Tg, dtype=np.float64); p = np.ascontiguousarray(p, dtype=np.float64)
    walls = np.ascontiguousarray(walls, dtype=np.int32)
    status = np.zeros((nx, ny), dtype=np.uint8)
    L = lib()
    L.ohf_set_min_distance_to_wall.restype = None
    L.ohf_set_min_distance_to_wall.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_double,
                                               C.c_double, C.c_void_p, C.c_long, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.ohf_set_min_distance_to_wall(ct.ctypes.data, Tg.ctypes.data, p.ctypes.data, nx, ny, dx, dy, x0, walls.ctypes.data,
                                   len(walls), l_min.ctypes.data, i_wall.ctypes.data, j_wall.ctypes.data, status.ctypes.data)
    return status


def recalc_y_plus(ct, dx, dy, i_wall, j_wall, dUdy, dVdx, mu, rho, y_plus):
    """Recalc_y_plus (deeps2d_core.cpp:2364-2386) on x-major planes; y_plus (float64, contiguous) is updated in place."""
    ct = np.ascontiguousarray(ct, dtype=np.uint64)
    nx, ny = ct.shape
    a = [np.ascontiguousarray(v, dtype=t) for v, t in ((i_wall, np.int32), (j_wall, np.int32), (dUdy, np.float64),
                                                       (dVdx, np.float64), (mu, np.float64), (rho, np.float64))]
    assert y_plus.dtype == np.float64 and y_plus.flags.c_contiguous and y_plus.shape == (nx, ny)
    L = lib()
    L.ohf_recalc_y_plus.restype = None
    L.ohf_recalc_y_plus.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_double] + [C.c_void_p] * 7
    L.ohf_recalc_y_plus(ct.ctypes.data, nx, ny, dx, dy, *[v.ctypes.data for v in a], y_plus.ctypes.data)
    return y_plus

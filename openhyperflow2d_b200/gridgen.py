"""Synthetic grids for benchmarks and smoke tests: builds the reference's AoS grid image directly (numpy), without
the reference's init code (InitDEEPS2D is init-time, host-only and out of scope -- SURVEY.md section 2; its
O(N*N_wall) wall-distance pass alone takes minutes at 2048x4096).

The node classes and their CT / TurbType words are the ones the reference's own init produces for the same
boundary macros (checked against the golden images: interior 0x80000000, NT_FARFIELD inflow, NT_D0X|NONREFLECTED
outflow, NT_D0Y|V_CONST floor/ceiling, NT_WNS walls around CT_SOLID bodies, hyper_flow_node.hpp:103-128), the
neighbour flags follow deeps2d_core.cpp:4516-4543, node y = (j+0.5)*dy (:3878).

`three_cylinders` is BASELINE.json configs[3] ("hypersonic flow around three cylinders, k-eps URANS").
"""
from __future__ import annotations

import os

import numpy as np

from . import aos

HERE = os.path.dirname(os.path.abspath(__file__))
DEFAULT_LAYOUT = os.path.join(HERE, "ref_layout.txt")   # FlowNode2D<double,3>, g++ x86-64 (1248 bytes)

# CondType2D / NodeType2D (hyper_flow_node.hpp:63-128)
CT_U_CONST, CT_V_CONST = 0x02, 0x04
CT_NONREFLECTED, CT_WALL_NO_SLIP, CT_SOLID, CT_NODE_IS_SET = 0x02000000, 0x04000000, 0x40000000, 0x80000000
NT_FC = 0x01 | 0x02 | 0x04 | 0x10 | 0x08 | CT_NODE_IS_SET
NT_D0X = CT_NODE_IS_SET | 0x20 | 0x40 | 0x80 | 0x100 | 0x200
NT_D0Y = CT_NODE_IS_SET | 0x400 | 0x800 | 0x1000 | 0x2000 | 0x4000
NT_WNS = CT_NODE_IS_SET | CT_WALL_NO_SLIP | CT_U_CONST | CT_V_CONST
NT_FARFIELD = NT_FC | CT_NONREFLECTED
TCT_k_CONST, TCT_eps_CONST, TCT_dkdx, TCT_depsdx, TCT_dkdy, TCT_depsdy = 0x01, 0x02, 0x04, 0x08, 0x10, 0x20
TCT_k_eps, TCT_SA, TCT_eps_wall = 0x400, 0x10000, 0x8000

# authored property tables (species: fuel, oxidiser, products, air)
TABLES = {
    "Cp_Fuel": ([20., 400., 1500.], [10000., 14500., 16050.]),
    "Cp_OX": ([1700., 1000., 600., 300., 230.], [1398., 1240., 1130., 1053., 1052.]),
    "Cp_cp": ([800., 1200., 2000., 3400.], [1540., 1620., 1770., 1930.]),
    "Cp_air": ([200., 300., 400., 1500., 3500.], [1005.8, 1004.5, 1009.7, 1210.8, 1307.4]),
    "lam_Fuel": ([200., 400., 1000.], [0.14, 0.226, 0.45]),
    "lam_OX": ([1700., 1000., 600., 300., 230.], [0.115, 0.0755, 0.0465, 0.0262, 0.021]),
    "lam_cp": ([800., 1200., 1600., 2400., 3400.], [0.12, 0.17, 0.22, 0.30, 0.41]),
    "lam_air": ([90., 300., 1000., 2000., 3500.], [0.015, 0.0267, 0.07, 0.115, 0.173]),
    "mu_Fuel": ([20., 100., 1000., 2000.], [2.5e-6, 4.2e-6, 20.1e-6, 33.6e-6]),
    "mu_OX": ([1700., 1000., 600., 300., 230.], [5.9e-5, 4.2e-5, 2.9e-5, 1.75e-5, 1.4e-5]),
    "mu_cp": ([800., 1200., 2000., 3400.], [3.3e-5, 4.4e-5, 6.5e-5, 9.3e-5]),
    "mu_air": ([100., 300., 1000., 2000., 3500.], [7.7e-6, 18.3e-6, 48.5e-6, 64.7e-6, 93.1e-6]),
}


def table_getval(tab, x):
    """Table::GetVal (obj_data.cpp:1822-1860) for scalars, host side (init-time use only)."""
    xs, ys = tab
    n = len(xs)
    if n == 0:
        return 0.0
    if n == 1:
        return ys[0]
    if x <= xs[0]:
        i = 1
    elif x >= xs[-1]:
        i = n - 1
    else:
        i = next(q for q in range(1, n) if xs[q - 1] <= x < xs[q])
    return ys[i] + (ys[i - 1] - ys[i]) * (x - xs[i]) / (xs[i - 1] - xs[i])


def make_cfg(nx, ny, dx, dy, *, problem=1, flow_type=0, tem=4, turb_start=1, turb_reset=1, adiabatic=1, bff=4, dt=1e-9,
             cfl=0.1, beta=0.985):
    return {
        "MaxX": nx, "MaxY": ny, "dx": dx, "dy": dy, "dt": dt, "CFL": cfl, "beta0": beta, "nrbc_beta0": 0.25, "bFF": bff,
        "ProblemType": problem, "FT": flow_type, "Hu": [0.0, 50.0e3, -1.2e7, 0.0], "TurbStartIter": turb_start,
        "TurbExtModel": tem, "isTurbulenceReset": turb_reset, "SigW": 1.0, "SigF": 1.0, "delta_bl": 1e-3, "TurbIntensity": 0.005,
        "isAdiabaticWall": adiabatic, "isAlternateRMS": 1, "AddSrcStartIter": 0, "last_iter": 0, "Nstep": 100, "NOutStep": 100,
        "MonitorIndex": 5, "ExitMonitorValue": 1e30, "GlobalTime": 0.0, "K0": 34.0, "Tf": 1000.0,
        "R_Fuel": 4157.0, "R_OX": 290.0, "R_cp": 428.59, "R_air": 287.0, "H_Fuel": 0.0, "H_OX": 50.0e3, "H_cp": -1.2e7, "H_air": 0.0,
        "tables": dict(TABLES, CFL_Scenario=([0., 1000., 200000.], [0.05, cfl, cfl]),
                       beta_Scenario=([0., 20000., 100000.], [beta - 0.03, beta - 0.001, beta])),
    }


def three_cylinders(nx, ny, *, layout=None, model="keps", mach=6.0, p_inf=5000.0, T_inf=220.0, Ts0=300.0, dx=2.5e-4, i0=0, nx_total=None, bl_cells=24):
    """AoS image (nx, ny) of columns [i0, i0+nx) of an nx_total x ny domain with three solid cylinders in a Mach-6 stream.

    Returns (J, cfg, layout).  model: "keps" (9 equations), "sa" (8), "laminar" or "euler" (7)."""
    layout = layout or aos.parse_layout(DEFAULT_LAYOUT)
    nx_total = nx_total or nx
    dy = dx
    ns = model != "euler"
    cfg = make_cfg(nx_total, ny, dx, dy, problem=1 if ns else 0, tem={"keps": 4, "sa": 10}.get(model, 0))
    J = np.zeros((nx, ny), dtype=aos.ref_dtype(layout))
    ii = (np.arange(nx) + i0)[:, None]
    jj = np.arange(ny)[None, :]
    X, Y = (ii + 0.5) * dx, (jj + 0.5) * dy
    Lx, Ly = nx_total * dx, ny * dy
    r = 0.06 * Ly
    centres = [(0.25 * Lx, 0.5 * Ly), (0.5 * Lx, 0.3 * Ly), (0.5 * Lx, 0.7 * Ly)]

    def solid_at(x, y):
        m = np.zeros(np.broadcast(x, y).shape, bool)
        for cx, cy in centres:
            m |= (x - cx) ** 2 + (y - cy) ** 2 < r * r
        return m
    solid = solid_at(X, Y)
    sl = solid_at(X - dx, Y) | (ii == 0)          # left neighbour is solid / missing
    sr = solid_at(X + dx, Y) | (ii == nx_total - 1)
    sd = solid_at(X, Y - dy) | (jj == 0)
    su = solid_at(X, Y + dy) | (jj == ny - 1)
    wall = ~solid & (solid_at(X - dx, Y) | solid_at(X + dx, Y) | solid_at(X, Y - dy) | solid_at(X, Y + dy))
    tmod = {"keps": TCT_k_eps, "sa": TCT_SA}.get(model, 0)
    ct = np.full((nx, ny), CT_NODE_IS_SET, np.uint64)
    tt = np.full((nx, ny), tmod, np.uint64)
    top_bot = (jj == 0) | (jj == ny - 1)
    ct[np.broadcast_to(top_bot, ct.shape)] = NT_D0Y | CT_V_CONST
    tt[np.broadcast_to(top_bot, ct.shape)] = tmod | ((TCT_dkdy | TCT_depsdy) if model == "keps" else TCT_dkdy if model == "sa" else 0)
    right = np.broadcast_to(ii == nx_total - 1, ct.shape)
    ct[right] = NT_D0X | CT_NONREFLECTED
    tt[right] = tmod | ((TCT_dkdx | TCT_depsdx) if model == "keps" else TCT_dkdx if model == "sa" else 0)
    left = np.broadcast_to(ii == 0, ct.shape)
    ct[left] = NT_FARFIELD
    tt[left] = tmod | ((TCT_k_CONST | TCT_eps_CONST) if model == "keps" else TCT_k_CONST if model == "sa" else 0)
    ct[wall] = NT_WNS
    tt[wall] = tmod | (TCT_eps_wall if model == "keps" else 0)
    ct[solid] = CT_SOLID | CT_NODE_IS_SET
    if not ns:
        tt[:] = 0
    J["CT"], J["TurbType"] = ct, tt
    J["idXl"], J["idXr"], J["idYd"], J["idYu"] = (~sl).astype(np.int32), (~sr).astype(np.int32), (~sd).astype(np.int32), (~su).astype(np.int32)
    # free stream (FlowNode2D ctor arithmetic, hyper_flow_node.hpp:323-351)
    T = cfg["tables"]
    R = cfg["R_air"]
    CP = table_getval(T["Cp_air"], T_inf)
    kg = CP / (CP - R)
    rho = p_inf / (R * T_inf)
    U0 = mach * np.sqrt(kg * R * T_inf)
    # boundary-layer start profile around the bodies (the reference's SetInitBoundaryLayer, deeps2d_core.cpp:2243, does
    # the same job for its decks): without it the impulsively started stream one cell off a no-slip wall drives k < 0
    # within two steps.  U ramps from 0 at the surface to U0 over `bl_cells` cells.
    dsurf = np.full((nx, ny), max(Lx, Ly))
    for cx, cy in centres:
        dsurf = np.minimum(dsurf, np.sqrt((X - cx) ** 2 + (Y - cy) ** 2) - r)
    ramp = np.clip(dsurf / (bl_cells * dx), 0.0, 1.0)
    U = np.where(wall | solid, 0.0, U0 * np.sin(0.5 * np.pi * ramp) ** 2) * np.ones((nx, ny))
    J["U"], J["V"] = U, 0.0
    J["p"] = p_inf
    J["Tg"] = np.where(solid, Ts0, T_inf)
    J["k"], J["R"], J["CP"] = kg, R, CP
    J["lam"], J["mu"] = table_getval(T["lam_air"], T_inf), table_getval(T["mu_air"], T_inf)
    J["Tf"] = cfg["Tf"]
    S = J["S"]
    S[..., 0] = rho
    S[..., 1] = rho * U
    S[..., 3] = p_inf / (kg - 1.0) + rho * U * U * 0.5
    if model == "keps":
        tmpi = 0.005 * np.sqrt(U * U + 1e-30)
        kk = 1.5 * tmpi * tmpi * rho
        S[..., 7] = kk
    elif model == "sa":
        S[..., 7] = J["mu"] / rho / 100.0
    J["S"] = S
    A, B = J["A"], J["B"]
    A[..., 0] = S[..., 1]; A[..., 1] = p_inf + S[..., 1] * U; A[..., 3] = (S[..., 3] + p_inf) * U
    B[..., 2] = p_inf
    J["A"], J["B"] = A, B
    J["beta"] = cfg["beta0"]
    J["Y"] = np.array([0.0, 0.0, 0.0, 1.0])
    J["BGX"], J["BGY"] = 1.0, 1.0
    J["y"] = np.broadcast_to(Y, (nx, ny))
    J["x"] = np.broadcast_to(X, (nx, ny))
    # distance to the nearest cylinder surface, floored like SetMinDistanceToWall2D (deeps2d_core.cpp:4783-4832)
    lmin = np.full((nx, ny), max(Lx, Ly))
    for cx, cy in centres:
        lmin = np.minimum(lmin, np.abs(np.sqrt((X - cx) ** 2 + (Y - cy) ** 2) - r))
    J["l_min"] = np.maximum(lmin, min(dx, dy))
    J["y_plus"] = 30.0
    if model == "keps":
        l = np.maximum(J["l_min"], min(dx, dy)) * 0.41
        S = J["S"]
        S[..., 8] = 0.09 ** 0.75 * (S[..., 7] / rho) ** 1.5 / l
        J["S"] = S
    # a stable first dt: CFL * min(dx/(a+|U|)) of the free stream (deeps2d_core.cpp:3845-3857 does the same from the Flow list)
    a = np.sqrt(kg * R * T_inf)
    cfg["dt"] = float(cfg["CFL"] * 0.5 * min(dx, dy) / (a + U0))
    cfg["MaxX"] = nx
    return J, cfg, layout

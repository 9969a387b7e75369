#!/usr/bin/env python
"""Generate the golden vectors under tests/golden/ by running the UNMODIFIED reference solver.

Run HERE (build container, /root/reference present):   python tests/golden/make_golden.py
Needs oracle/_ref/ohf_dump (make -C oracle ref).  For every case below it writes an input deck in
the reference's own format (authored here, small grids), runs the reference for N steps (one or two
output cycles), and stores
    tests/golden/<case>.npz :  state0 / stateN = raw AoS grid images (uint8, zlib-compressed by npz),
                               cfg = text of every scalar/table the step reads, nx, ny, nsteps, ncycles,
                               global_time, dt_final, rms_file (the reference's RMS-*.plt text)
    tests/golden/layout.txt :  FlowNode2D member offsets (offsetof on the reference's header)
The GPU box has no /root/reference: tests only read the committed .npz files.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
DUMP = os.path.join(ROOT, "oracle", "_ref", "ohf_dump")

# ---- property tables (authored; loosely air / H2 / O2 / H2O like, deliberately including a
# ---- descending-x table and tables whose range does not cover every temperature met) ----------
TABLES = {
    "lam_cp": [(800., 0.12), (1200., 0.17), (1600., 0.22), (2400., 0.30), (3400., 0.41)],
    "mu_cp": [(800., 3.3e-5), (1200., 4.4e-5), (2000., 6.5e-5), (3400., 9.3e-5)],
    "Cp_cp": [(800., 1540.), (1200., 1620.), (2000., 1770.), (3400., 1930.)],
    "lam_Fuel": [(200., 0.14), (400., 0.226), (1000., 0.45)],
    "mu_Fuel": [(20., 2.5e-6), (100., 4.2e-6), (1000., 20.1e-6), (2000., 33.6e-6)],
    "Cp_Fuel": [(20., 10000.), (400., 14500.), (1500., 16050.)],
    "lam_OX": [(1700., 0.115), (1000., 0.0755), (600., 0.0465), (300., 0.0262), (230., 0.021)],   # descending
    "mu_OX": [(1700., 5.9e-5), (1000., 4.2e-5), (600., 2.9e-5), (300., 1.75e-5), (230., 1.4e-5)],  # descending
    "Cp_OX": [(1700., 1398.), (1000., 1240.), (600., 1130.), (300., 1053.), (230., 1052.)],       # descending
    "lam_air": [(90., 0.015), (300., 0.0267), (1000., 0.07), (2000., 0.115), (3500., 0.173)],
    "mu_air": [(100., 7.7e-6), (300., 18.3e-6), (1000., 48.5e-6), (2000., 64.7e-6), (3500., 93.1e-6)],
    "Cp_air": [(200., 1005.8), (300., 1004.5), (400., 1009.7), (1500., 1210.8), (3500., 1307.4)],
}


def table(name, rows):
    return [f"<table={name}/{len(rows)}>"] + [f"{x!r} {y!r}" for x, y in rows] + ["<endtable>"]


def make_deck(name, *, nx=60, ny=40, dx=1e-3, dy=1e-3, problem=0, flow_type=0, turb=0, turb_ext=0, bff=4,
              adiabatic=1, alt_rms=1, nmax=2, tf=1000., k0=34., turb_reset=0, turb_start=0, cfl=0.1, beta=0.985,
              flows=None, bounds=None, contour=None, areas=None, single_bounds=(), monitor_index=5, delta_bl=1e-3,
              sig_w=1.0, sig_f=1.0, beta_nrbc=0.25, circles=(), rects=(), airfoils=(), cx_body=None):
    """Return deck text.  `flows`: list of dicts (Flow2D-n.*), `bounds`: list of (cond string, flow2d idx)."""
    L = [f"<start/{name}>"]
    d = lambda k, v: L.append(f"<data/{k}={v}>")
    d("ProjectName", name); d("OutputFile", ".plt"); d("ErrorFile", "-err.plt"); d("GasSwapFile", ".hf2d")
    d("isAdiabaticWall", adiabatic); d("isOutHeatFluxX", 0); d("isOutHeatFluxY", 0); d("is_p_asterisk_out", 0)
    d("isIgnoreUnsetNodes", 0); d("ProblemType", problem); d("FlowType", flow_type)
    d("MaxX", nx); d("MaxY", ny); d("dx", repr(dx)); d("dy", repr(dy))
    d("isAlternateRMS", alt_rms); d("MonitorIndex", monitor_index); d("ExitMonitorValue", "1.e-30"); d("NumMonitorPoints", 0)
    d("CFL", repr(cfl))
    L += table("CFL_Scenario", [(0., cfl * 0.5), (4., cfl), (200000., cfl)])
    d("beta", repr(beta)); d("beta_NonReflectedBC", repr(beta_nrbc))
    L += table("beta_Scenario", [(0., beta - 0.03), (6., beta - 0.001), (100000., beta)])
    d("Nmax", nmax); d("NOutStep", 1); d("NSaveStep", 1); d("isVerboseOutput", 1); d("InitTime", "0.")
    d("BFF", bff); d("TurbulenceModel", turb); d("isTurbulenceReset", turb_reset); d("ZeroEqFactor", 0)
    d("SigW", repr(sig_w)); d("SigF", repr(sig_f)); d("delta_bl", repr(delta_bl)); d("TurbStartIter", turb_start); d("TurbExtModel", turb_ext)
    d("Ts0", "320."); d("K0", repr(k0)); d("gamma", "0."); d("Tf", repr(tf)); d("NumSrc", 0)
    d("H_cp", "-1.2e7"); d("R_cp", "428.59"); d("H_Fuel", "0.0"); d("R_Fuel", "4157.0")
    d("H_OX", "50.0e3"); d("R_OX", "290.0"); d("H_air", "0.0"); d("R_air", "287.0")
    for tn, rows in TABLES.items():
        L += table(tn, rows)
    d("NumFlow", 0)
    d("NumFlow2D", len(flows))
    for i, f in enumerate(flows, 1):
        for k, v in f.items():
            d(f"Flow2D-{i}.{k}", v)
    # solid bodies drawn by the reference's own rasterisers (deeps2d_core.cpp:4013-4282): no-slip contour nodes
    # (NT_WNS_2D); the inside is filled by a solid Area seeded in it
    d("NumRects", len(rects))
    for i, r in enumerate(rects, 1):
        for k in ("Xstart", "Ystart", "DX", "DY"):
            d(f"Rect{i}.{k}", repr(r[k]))
        d(f"Rect{i}.Flow2D", r.get("flow", 1)); d(f"Rect{i}.TurbulenceModel", turb)
    d("NumCircles", len(circles))
    for i, c in enumerate(circles, 1):
        for k in ("Xstart", "Ystart", "X0", "Y0"):
            d(f"Circle{i}.{k}", repr(c[k]))
        d(f"Circle{i}.MaterialID", c.get("mat", 1)); d(f"Circle{i}.TurbulenceModel", turb); d(f"Circle{i}.Flow2D", c.get("flow", 1))
    d("NumAirfoils", len(airfoils))
    for i, a in enumerate(airfoils, 1):
        d(f"Airfoil{i}.Xstart", repr(a["Xstart"])); d(f"Airfoil{i}.Ystart", repr(a["Ystart"])); d(f"Airfoil{i}.Type", 0)
        d(f"Airfoil{i}.pp", repr(a["pp"])); d(f"Airfoil{i}.mm", repr(a["mm"])); d(f"Airfoil{i}.thick", repr(a["thick"]))
        d(f"Airfoil{i}.scale", repr(a["scale"])); d(f"Airfoil{i}.attack_angle", repr(a["attack_angle"]))
        d(f"Airfoil{i}.Flow2D", a.get("flow", 1)); d(f"Airfoil{i}.TurbulenceModel", turb)
    d("NumSingleBounds", len(single_bounds))
    for i, sb in enumerate(single_bounds, 1):
        L += table(f"SingleBound{i}.Points", sb["points"])
        d(f"SingleBound{i}.Cond", sb["cond"]); d(f"SingleBound{i}.Flow2D", sb["flow"])
        d(f"SingleBound{i}.TurbulenceModel", turb); d(f"SingleBound{i}.isReset", 0); d(f"SingleBound{i}.MaterialID", 0)
    d("NumContour", 1)
    L += table("Contour1", contour)
    d("Contour1.MaterialID", 0)
    for i, (cond, fl) in enumerate(bounds, 1):
        d(f"Contour1.Bound{i}.Cond", cond); d(f"Contour1.Bound{i}.Flow2D", fl)
        d(f"Contour1.Bound{i}.TurbulenceModel", turb); d(f"Contour1.Bound{i}.isReset", 0)
    d("NumArea", len(areas))
    for i, a in enumerate(areas, 1):
        L += [f"<table=Area{i}/1>", f"{a['seed'][0]} {a['seed'][1]}", "<endtable>"]
        d(f"Area{i}.Type", a["type"]); d(f"Area{i}.MaterialID", a.get("mat", 0))
        if a["type"] == 1:
            d(f"Area{i}.Flow2D", a["flow"]); d(f"Area{i}.TurbulenceModel", turb)
    # airfoils need is_Cx_calc = 1: their constructor call reads Flow2D[Cx_Flow_Index - 1] (deeps2d_core.cpp:3972-3982,4284)
    xb, yb, dxb, dyb = cx_body or (0.0, 0.0, 0.0, 0.0)
    d("is_Cx_calc", 1 if cx_body else 0); d("is_Cd_calc", 0); d("x_body", repr(xb)); d("y_body", repr(yb)); d("dx_body", repr(dxb)); d("dy_body", repr(dyb))
    d("Cx_Flow_Index", 1); d("NumXCut", 0)
    L.append(f"<end/{name}>")
    return "\n".join(L) + "\n"


def air(mach=None, p=101350., T=288.9, comp=3, angle=0.0):
    if mach is None:
        return {"CompIndex": comp, "Mode": 0, "p": p, "T": T, "U": 0.0, "V": 0.0}
    return {"CompIndex": comp, "Mode": 2, "Mach": mach, "Angle": angle, "p": p, "T": T, "U": 0.0, "V": 0.0}


def wedge_geometry(nx, ny, dx, dy, ramp_x, ramp_h):
    X, Y = nx * dx, ny * dy
    xr = X - 1e-5
    return [(0.0, Y), (xr, Y), (xr, ramp_h), (ramp_x, 0.0), (0.0, 0.0)]


def cases():
    C = {}
    nx, ny, dx, dy = 60, 40, 1e-3, 1e-3
    geo = wedge_geometry(nx, ny, dx, dy, 0.02, 0.011)
    areas = [{"seed": (3, 3), "type": 1, "flow": 1}, {"seed": (nx - 4, 2), "type": 0, "mat": 1}]
    # 1. Euler duct with ramp (BASELINE config 1 in miniature): FC inflow, D0X outflow, symmetric floor, slip ramp
    euler_b = [("NT_FC_2D", 1), ("NT_D0X_2D", 1), ("NT_D0Y_2D, CT_V_CONST_2D", 2), ("NT_D0Y_2D, CT_V_CONST_2D", 2), ("NT_FC_2D", 1)]
    C["euler_duct"] = dict(deck=dict(problem=0, flows=[air(2.5), air()], bounds=euler_b, contour=geo, areas=areas, bff=4), steps=[1, 10])
    C["euler_duct_bff5"] = dict(deck=dict(problem=0, flows=[air(2.5), air()], bounds=euler_b, contour=geo, areas=areas, bff=5, alt_rms=0,
                                          monitor_index=0), steps=[7])
    # 2. laminar NS with a no-slip ramp, adiabatic
    ns_b = [("NT_FARFIELD_2D", 1), ("NT_D0X_2D, CT_NONREFLECTED_2D", 1), ("NT_WNS_2D", 2), ("NT_D0Y_2D, CT_V_CONST_2D", 2), ("NT_FARFIELD_2D", 1)]
    C["ns_laminar"] = dict(deck=dict(problem=1, turb=0, flows=[air(2.5), air()], bounds=ns_b, contour=geo, areas=areas), steps=[1, 8], two_cycle=4)
    # 3. k-eps (std) with non-adiabatic wall -> CalcHeatOnWallSources
    ke_b = [("NT_FARFIELD_2D, TCT_k_CONST_2D, TCT_eps_CONST_2D", 1),
            ("NT_D0X_2D, TCT_dkdx_NULL_2D, TCT_depsdx_NULL_2D,  CT_NONREFLECTED_2D", 1),
            ("NT_WNS_2D, TCT_eps_Cmk2kXn_WALL_2D", 2),
            ("NT_D0Y_2D, TCT_dkdy_NULL_2D, TCT_depsdy_NULL_2D, CT_V_CONST_2D", 2),
            ("NT_FARFIELD_2D, TCT_k_CONST_2D, TCT_eps_CONST_2D", 1)]
    for tag, tem in (("std", 4), ("chien", 5), ("jl", 6), ("lsy", 7), ("rng", 8)):
        C[f"ns_keps_{tag}"] = dict(deck=dict(problem=1, turb=4, turb_ext=tem, turb_reset=1, turb_start=1, adiabatic=0,
                                             flows=[air(2.5), air()], bounds=ke_b, contour=geo, areas=areas),
                                   steps=[8] if tag != "std" else [1, 8], two_cycle=4 if tag == "std" else None)
    # 4. Spalart-Allmaras
    sa_b = [("NT_FARFIELD_2D, TCT_k_CONST_2D", 1), ("NT_D0X_2D, TCT_dkdx_NULL_2D, CT_NONREFLECTED_2D", 1), ("NT_WNS_2D", 2),
            ("NT_D0Y_2D, TCT_dkdy_NULL_2D, CT_V_CONST_2D", 2), ("NT_FARFIELD_2D, TCT_k_CONST_2D", 1)]
    # (on the ramp geometry the reference's SA source term blows up within 3 steps -- Tg<0 abort -- because its
    #  compressibility term divides by fv1 ~ 1e-9 right after the reset; a flat plate keeps dUdy*dVdx ~ 0.)
    plate = [(0.0, ny * dy), (nx * dx - 1e-5, ny * dy), (nx * dx - 1e-5, 0.0), (0.0, 0.0)]
    sa_b = [("NT_FARFIELD_2D, TCT_k_CONST_2D", 1), ("NT_D0X_2D, TCT_dkdx_NULL_2D, CT_NONREFLECTED_2D", 1), ("NT_WNS_2D", 2),
            ("NT_FARFIELD_2D, TCT_k_CONST_2D", 1)]
    C["ns_sa"] = dict(deck=dict(problem=1, turb=3, turb_ext=10, turb_reset=1, turb_start=1, flows=[air(2.5), air()], bounds=sa_b,
                                contour=plate, areas=[{"seed": (5, 20), "type": 1, "flow": 1}]), steps=[2])  # step 3 aborts (Tg<0)
    # 5. zero-equation (van Driest) and Smagorinsky
    C["ns_prandtl_vd"] = dict(deck=dict(problem=1, turb=2, turb_ext=1, turb_start=0, flows=[air(2.5), air()], bounds=ns_b, contour=geo,
                                        areas=areas, bff=5), steps=[8])
    C["ns_klebanoff"] = dict(deck=dict(problem=1, turb=2, turb_ext=3, turb_start=0, flows=[air(2.5), air()], bounds=ns_b, contour=geo,
                                       areas=areas, delta_bl=4e-3), steps=[6])
    C["ns_smagorinsky"] = dict(deck=dict(problem=1, turb=5, turb_ext=15, turb_start=0, flows=[air(2.5), air()], bounds=ns_b, contour=geo,
                                         areas=areas), steps=[8])
    # 6. axisymmetric, multicomponent, reacting: hot oxidiser co-flow (Flow2D-1), cold fuel jet on the lower part of the
    #    inflow face (SingleBound, Flow2D-3), ignition temperature below the co-flow temperature so the burn branch fires.
    ax_geo = [(0.0, ny * dy), (nx * dx - 1e-5, ny * dy), (nx * dx - 1e-5, 0.0), (0.0, 0.0)]
    ax_b = [("NT_D0Y_2D", 1), ("NT_D0X_2D", 1), ("NT_AX_2D", 1), ("NT_FC_2D", 1)]
    jet = [{"points": [(0.0, 0.0), (0.0, 0.012)], "cond": "NT_FC_2D", "flow": 3}]
    ax_flows = [air(1.5, p=101350., T=1400., comp=1), air(comp=1, T=1400.), air(1.2, p=101350., T=400., comp=0)]
    C["axi_reacting_euler"] = dict(deck=dict(problem=0, flow_type=1, flows=ax_flows, bounds=ax_b, contour=ax_geo,
                                             areas=[{"seed": (5, 20), "type": 1, "flow": 1}], single_bounds=jet, tf=900., k0=8.0), steps=[10])
    ax_kb = [("NT_D0Y_2D, TCT_dkdy_NULL_2D, TCT_depsdy_NULL_2D", 1), ("NT_D0X_2D, TCT_dkdx_NULL_2D, TCT_depsdx_NULL_2D", 1),
             ("NT_AX_2D, TCT_dkdy_NULL_2D, TCT_depsdy_NULL_2D", 1), ("NT_FC_2D, TCT_k_CONST_2D, TCT_eps_CONST_2D", 1)]
    jet_k = [{"points": [(0.0, 0.0), (0.0, 0.012)], "cond": "NT_FC_2D, TCT_k_CONST_2D, TCT_eps_CONST_2D", "flow": 3}]
    C["axi_reacting_keps"] = dict(deck=dict(problem=1, flow_type=1, turb=4, turb_ext=4, turb_reset=1, turb_start=1, flows=ax_flows,
                                            bounds=ax_kb, contour=ax_geo, areas=[{"seed": (5, 20), "type": 1, "flow": 1}],
                                            single_bounds=jet_k, tf=900., k0=8.0), steps=[8])
    return C


def geometry_cases():
    """Bodies drawn by the reference's own rasterisers (BoundCircle2D, SolidBoundRect2D, SolidBoundAirfoil2D): staircase
    no-slip walls with every corner type, in miniature of BASELINE configs 3 and 4.  Written to tests/golden/geometry/."""
    G = {}
    nx, ny, dx, dy = 96, 64, 1e-3, 1e-3
    X, Y = nx * dx, ny * dy
    box = [(0.0, Y), (X - 1e-5, Y), (X - 1e-5, 0.0), (0.0, 0.0)]
    ke_b = [("NT_D0Y_2D, TCT_dkdy_NULL_2D, TCT_depsdy_NULL_2D", 1),
            ("NT_D0X_2D, TCT_dkdx_NULL_2D, TCT_depsdx_NULL_2D, CT_NONREFLECTED_2D", 1),
            ("NT_D0Y_2D, TCT_dkdy_NULL_2D, TCT_depsdy_NULL_2D", 1),
            ("NT_FARFIELD_2D, TCT_k_CONST_2D, TCT_eps_CONST_2D", 1)]
    ns_b = [("NT_D0Y_2D", 1), ("NT_D0X_2D, CT_NONREFLECTED_2D", 1), ("NT_D0Y_2D", 1), ("NT_FARFIELD_2D", 1)]
    # three cylinders (BASELINE config 4 in miniature): radius 7, 6 and 5 cells, staggered
    cyl = [(0.030, 0.032, 0.007), (0.058, 0.018, 0.006), (0.062, 0.047, 0.005)]
    # (the body constructors fill their own inside with solid nodes and zero the velocity of the Flow2D they are given:
    #  bodies get the still-air state, Flow2D-2)
    circles = [{"X0": x0, "Y0": y0, "Xstart": x0 + r, "Ystart": y0, "flow": 2} for x0, y0, r in cyl]
    c_areas = [{"seed": (3, 3), "type": 1, "flow": 1}]
    G["cylinders_keps"] = dict(deck=dict(nx=nx, ny=ny, problem=1, turb=4, turb_ext=4, turb_reset=1, turb_start=1, adiabatic=0,
                                         flows=[air(2.5), air()], bounds=ke_b, contour=box, areas=c_areas, circles=circles), steps=[8])
    G["cylinders_laminar"] = dict(deck=dict(nx=nx, ny=ny, problem=1, turb=0, flows=[air(2.5), air()], bounds=ns_b, contour=box,
                                            areas=c_areas, circles=circles), steps=[8])
    # NACA 2412, zero incidence (BASELINE config 3 in miniature), chord 50 cells
    foil = [{"Xstart": 0.020, "Ystart": 0.032, "pp": 0.4, "mm": 0.02, "thick": 0.12, "scale": 0.050, "attack_angle": 0.0, "flow": 2}]
    f_areas = [{"seed": (3, 3), "type": 1, "flow": 1}]
    G["airfoil_keps"] = dict(deck=dict(nx=nx, ny=ny, problem=1, turb=4, turb_ext=4, turb_reset=1, turb_start=1,
                                       flows=[air(0.8), air()], bounds=ke_b, contour=box, areas=f_areas, airfoils=foil, cx_body=(0.015, 0.020, 0.060, 0.024)), steps=[8])
    # solid rectangle (backward/forward facing steps), laminar, non-adiabatic wall
    rect = [{"Xstart": 0.030, "Ystart": 0.020, "DX": 0.020, "DY": 0.012, "flow": 2}]
    r_areas = [{"seed": (3, 3), "type": 1, "flow": 1}]
    G["rect_laminar"] = dict(deck=dict(nx=nx, ny=ny, problem=1, turb=0, adiabatic=0, flows=[air(2.5), air()], bounds=ns_b, contour=box,
                                       areas=r_areas, rects=rect), steps=[8])
    return G


def bc_cases():
    """Boundary-condition branches of the step that no case above reaches: soft (second-derivative) outflow and far
    boundaries, the wall-law wall, the second-derivative eps wall rewrites, the integral turbulence model, single Neumann
    flags.  Same ramp duct as cases(); written to tests/golden/bc/."""
    B = {}
    nx, ny, dx, dy = 60, 40, 1e-3, 1e-3
    geo = wedge_geometry(nx, ny, dx, dy, 0.02, 0.011)
    areas = [{"seed": (3, 3), "type": 1, "flow": 1}, {"seed": (nx - 4, 2), "type": 0, "mat": 1}]
    base = dict(flows=[air(2.5), air()], contour=geo, areas=areas)
    # soft outflow (d2/dx2 = 0) and soft top (d2/dy2 = 0)
    soft_e = [("NT_D2Y_2D", 1), ("NT_D2X_2D", 1), ("NT_D0Y_2D, CT_V_CONST_2D", 2), ("NT_D0Y_2D, CT_V_CONST_2D", 2), ("NT_FC_2D", 1)]
    B["euler_soft"] = dict(deck=dict(problem=0, bounds=soft_e, bff=2, **base), steps=[8])
    soft_k = [("NT_D2Y_2D, TCT_d2kdy2_NULL_2D, TCT_d2epsdy2_NULL_2D", 1),
              ("NT_D2X_2D, TCT_d2kdx2_NULL_2D, TCT_d2epsdx2_NULL_2D", 1),
              ("NT_WNS_2D, TCT_eps_mud2kdx2_WALL_2D", 2),
              ("NT_WNS_2D, TCT_eps_mud2kdy2_WALL_2D", 2),
              ("NT_FARFIELD_2D, TCT_k_CONST_2D, TCT_eps_CONST_2D", 1)]
    B["ns_keps_soft_mud2k"] = dict(deck=dict(problem=1, turb=4, turb_ext=4, turb_reset=1, turb_start=1, bounds=soft_k, bff=3, **base), steps=[8])
    # wall-law wall (slip velocity written into the momentum slot)
    wl = [("NT_FARFIELD_2D, TCT_k_CONST_2D, TCT_eps_CONST_2D", 1),
          ("NT_D0X_2D, TCT_dkdx_NULL_2D, TCT_depsdx_NULL_2D, CT_NONREFLECTED_2D", 1),
          ("NT_WALL_LAW_2D, TCT_eps_Cmk2kXn_WALL_2D", 2),
          ("NT_WALL_LAW_2D, TCT_eps_Cmk2kXn_WALL_2D", 2),
          ("NT_FARFIELD_2D, TCT_k_CONST_2D, TCT_eps_CONST_2D", 1)]
    B["ns_keps_wall_law"] = dict(deck=dict(problem=1, turb=4, turb_ext=4, turb_reset=1, turb_start=1, adiabatic=0, bounds=wl, **base), steps=[2])   # the reference itself aborts (Tg < 0) on step 3
    wl_l = [("NT_FARFIELD_2D", 1), ("NT_D0X_2D", 1), ("NT_WALL_LAW_2D", 2), ("NT_WALL_LAW_2D", 2), ("NT_FARFIELD_2D", 1)]
    B["ns_laminar_wall_law"] = dict(deck=dict(problem=1, turb=0, bounds=wl_l, bff=0, **base), steps=[2])
    # integral model f(Re)
    ns_b = [("NT_FARFIELD_2D", 1), ("NT_D0X_2D, CT_NONREFLECTED_2D", 1), ("NT_WNS_2D", 2), ("NT_D0Y_2D, CT_V_CONST_2D", 2), ("NT_FARFIELD_2D", 1)]
    B["ns_integral"] = dict(deck=dict(problem=1, turb=1, turb_ext=0, turb_start=0, bounds=ns_b, bff=1, **base), steps=[8])
    # single flags instead of the macro types: Neumann on some equations only, Dirichlet temperature on the wall
    mixed = [("CT_dRhody_NULL_2D, CT_dTdy_NULL_2D, CT_dUdy_NULL_2D, CT_V_CONST_2D", 1),
             ("CT_dRhodx_NULL_2D, CT_dUdx_NULL_2D, CT_d2Vdx2_NULL_2D, CT_d2Tdx2_NULL_2D, CT_dYdx_NULL_2D", 1),
             ("NT_WNS_2D, CT_T_CONST_2D", 2), ("NT_D0Y_2D, CT_V_CONST_2D", 2), ("NT_FC_2D", 1)]
    B["ns_laminar_mixed_flags"] = dict(deck=dict(problem=1, turb=0, bounds=mixed, **base), steps=[8])
    # Spalart-Allmaras beyond two steps: the reference's SA source blows up on step 3 at Mach 2.5 (cases(): ns_sa) and on
    # every body tried; the subsonic flat plate survives five steps, so four of them are pinned here
    C = cases()
    B["ns_sa_subsonic"] = dict(deck=dict(C["ns_sa"]["deck"], flows=[air(0.5), air()]), steps=[4])
    return B


def r2_cases():
    """Round-2 additions (tests/golden/r2/): what the earlier fixtures left unpinned.
    * burning decks: the earlier "reacting" decks put their fuel jet on a SingleBound, which the reference applies BEFORE
      the contour bounds (deeps2d_core.cpp:3417-3517 vs :3520-3800) -- the contour's inflow face then overwrites it and no
      fuel ever enters.  Here the inflow face of the contour itself is split into a hot fuel segment and a hot oxidiser
      segment, so both arms of CalcChemicalReactions (:4711-4722) fire, products form and the fuel / product rows of the
      property tables are evaluated;
    * grids tall enough (ny >= 100) to hold PLAIN tiles (32-row strips with an unflagged ring), for the kernel
      instantiations the 60x40 / 96x64 cases never reach: axisymmetric, Prandtl, Smagorinsky, Euler, laminar, SA;
    * N = 200 drift cases, one per physics family that survives 200 steps in the reference."""
    R = {}
    # ---- burning jets, 60 x 40 (general tiles) ----
    nx, ny, dx, dy = 60, 40, 1e-3, 1e-3
    X, Y = nx * dx, ny * dy
    split = [(0.0, Y), (X - 1e-5, Y), (X - 1e-5, 0.0), (0.0, 0.0), (0.0, 0.012)]
    fl = [air(1.5, p=101350., T=1400., comp=1), air(comp=1, T=1400.), air(1.2, p=101350., T=1100., comp=0)]
    eb = [("NT_D0Y_2D", 1), ("NT_D0X_2D", 1), ("NT_AX_2D", 1), ("NT_FC_2D", 3), ("NT_FC_2D", 1)]
    R["axi_burn_euler"] = dict(deck=dict(problem=0, flow_type=1, flows=fl, bounds=eb, contour=split,
                                         areas=[{"seed": (5, 20), "type": 1, "flow": 1}], tf=900., k0=8.0), steps=[12])
    kb = [("NT_D0Y_2D, TCT_dkdy_NULL_2D, TCT_depsdy_NULL_2D", 1), ("NT_D0X_2D, TCT_dkdx_NULL_2D, TCT_depsdx_NULL_2D", 1),
          ("NT_AX_2D, TCT_dkdy_NULL_2D, TCT_depsdy_NULL_2D", 1), ("NT_FC_2D, TCT_k_CONST_2D, TCT_eps_CONST_2D", 3),
          ("NT_FC_2D, TCT_k_CONST_2D, TCT_eps_CONST_2D", 1)]
    R["axi_burn_keps"] = dict(deck=dict(problem=1, flow_type=1, turb=4, turb_ext=4, turb_reset=1, turb_start=1, flows=fl, bounds=kb,
                                        contour=split, areas=[{"seed": (5, 20), "type": 1, "flow": 1}], tf=900., k0=2000.0), steps=[10])
    # flat, laminar, fuel-rich core initialised in the field (Area with the fuel state) burning against an air co-flow
    flf = [air(2.0, p=101350., T=1300., comp=1), air(comp=1), air(1.5, p=101350., T=1200., comp=0)]
    nb = [("NT_D0Y_2D", 1), ("NT_D0X_2D", 1), ("NT_D0Y_2D", 1), ("NT_FC_2D", 3), ("NT_FC_2D", 1)]
    R["flat_burn_laminar"] = dict(deck=dict(problem=1, turb=0, flows=flf, bounds=nb, contour=split,
                                            areas=[{"seed": (5, 20), "type": 1, "flow": 1}], tf=1000., k0=500.), steps=[10])
    # ---- grids that hold PLAIN tiles: 40 x 100 (two 32-row strips x three 8-column tiles with an unflagged ring) ----
    nx, ny = 40, 100
    X, Y = nx * dx, ny * dy
    geo = wedge_geometry(nx, ny, dx, dy, 0.024, 0.011)
    areas = [{"seed": (3, 3), "type": 1, "flow": 1}, {"seed": (nx - 3, 2), "type": 0, "mat": 1}]
    big = dict(nx=nx, ny=ny, contour=geo, areas=areas)
    ns_b = [("NT_FARFIELD_2D", 1), ("NT_D0X_2D, CT_NONREFLECTED_2D", 1), ("NT_WNS_2D", 2), ("NT_D0Y_2D, CT_V_CONST_2D", 2), ("NT_FARFIELD_2D", 1)]
    euler_b = [("NT_FC_2D", 1), ("NT_D0X_2D", 1), ("NT_D0Y_2D, CT_V_CONST_2D", 2), ("NT_D0Y_2D, CT_V_CONST_2D", 2), ("NT_FC_2D", 1)]
    R["plain_euler"] = dict(deck=dict(problem=0, flows=[air(2.5), air()], bounds=euler_b, bff=2, **big), steps=[8])
    R["plain_ns_laminar"] = dict(deck=dict(problem=1, turb=0, flows=[air(2.5), air()], bounds=ns_b, **big), steps=[8])
    R["plain_ns_prandtl_vd"] = dict(deck=dict(problem=1, turb=2, turb_ext=1, turb_start=0, flows=[air(2.5), air()], bounds=ns_b, bff=5, **big), steps=[8])
    R["plain_ns_smagorinsky"] = dict(deck=dict(problem=1, turb=5, turb_ext=15, turb_start=0, flows=[air(2.5), air()], bounds=ns_b, **big), steps=[8])
    plate = [(0.0, Y), (X - 1e-5, Y), (X - 1e-5, 0.0), (0.0, 0.0)]
    sa_b = [("NT_FARFIELD_2D, TCT_k_CONST_2D", 1), ("NT_D0X_2D, TCT_dkdx_NULL_2D, CT_NONREFLECTED_2D", 1), ("NT_WNS_2D", 2),
            ("NT_FARFIELD_2D, TCT_k_CONST_2D", 1)]
    R["plain_ns_sa_subsonic"] = dict(deck=dict(nx=nx, ny=ny, problem=1, turb=3, turb_ext=10, turb_reset=1, turb_start=1, flows=[air(0.5), air()],
                                               bounds=sa_b, contour=plate, areas=[{"seed": (5, 20), "type": 1, "flow": 1}]), steps=[4])
    # odd ny: the TMA path is not eligible (16-byte stride rule), the LDG form of K1a runs
    geo_o = wedge_geometry(nx, 101, dx, dy, 0.024, 0.011)
    ke_b = [("NT_FARFIELD_2D, TCT_k_CONST_2D, TCT_eps_CONST_2D", 1),
            ("NT_D0X_2D, TCT_dkdx_NULL_2D, TCT_depsdx_NULL_2D,  CT_NONREFLECTED_2D", 1),
            ("NT_WNS_2D, TCT_eps_Cmk2kXn_WALL_2D", 2),
            ("NT_D0Y_2D, TCT_dkdy_NULL_2D, TCT_depsdy_NULL_2D, CT_V_CONST_2D", 2),
            ("NT_FARFIELD_2D, TCT_k_CONST_2D, TCT_eps_CONST_2D", 1)]
    R["plain_ns_keps_odd_ny"] = dict(deck=dict(nx=nx, ny=101, problem=1, turb=4, turb_ext=4, turb_reset=1, turb_start=1, adiabatic=0,
                                               flows=[air(2.5), air()], bounds=ke_b, contour=geo_o, areas=areas), steps=[8])
    # axisymmetric burning jets on the tall grid: config C5's plain-tile kernels (KCfg<*, 1, *>)
    split_b = [(0.0, Y), (X - 1e-5, Y), (X - 1e-5, 0.0), (0.0, 0.0), (0.0, 0.040)]
    R["plain_axi_burn_euler"] = dict(deck=dict(nx=nx, ny=ny, problem=0, flow_type=1, flows=fl, bounds=eb, contour=split_b,
                                               areas=[{"seed": (5, 50), "type": 1, "flow": 1}], tf=900., k0=2000.0), steps=[10])
    R["plain_axi_burn_keps"] = dict(deck=dict(nx=nx, ny=ny, problem=1, flow_type=1, turb=4, turb_ext=4, turb_reset=1, turb_start=1, flows=fl,
                                              bounds=kb, contour=split_b, areas=[{"seed": (5, 50), "type": 1, "flow": 1}], tf=900., k0=2000.0), steps=[10])
    R["plain_axi_burn_laminar"] = dict(deck=dict(nx=nx, ny=ny, problem=1, flow_type=1, turb=0, flows=fl, bounds=eb, contour=split_b,
                                                 areas=[{"seed": (5, 50), "type": 1, "flow": 1}], tf=900., k0=40.0), steps=[10])
    # ---- drift over 200 steps: the 60 x 40 ramp duct of cases() ----
    C = cases()
    for name in ("euler_duct", "ns_laminar", "ns_keps_std", "ns_prandtl_vd"):
        R["drift_" + name] = dict(deck=C[name]["deck"], steps=[200])
    return R


def restart_cases():
    """Runs RESTARTED from the swap file of an earlier run of the same deck (`restart_after` steps): the reference's
    start-up then computes y_plus from real wall gradients (Recalc_y_plus, hf2d_start.cpp:299) -- in a fresh run they are
    zero and y_plus = 0 everywhere -- so these are the cases in which the y+ damping functions (van Driest, Chien, JL,
    LSY) and the restored-wall-node start-up path see non-trivial input.  Written to tests/golden/restart/."""
    C = cases()
    R = {}
    for name in ("ns_prandtl_vd", "ns_keps_std", "ns_keps_chien", "ns_keps_jl", "ns_keps_lsy", "ns_laminar"):
        R[name] = dict(deck=C[name]["deck"], steps=[6], restart_after=12)
    return R


def run_ref(deck_text, name, nstep, exit_time, wd, clean=True):
    os.makedirs(wd, exist_ok=True)
    for f in os.listdir(wd) if clean else ():
        os.remove(os.path.join(wd, f))
    with open(os.path.join(wd, name + ".dat"), "w") as f:
        f.write(deck_text)
    env = dict(os.environ, OHF_DUMP_DIR=wd, OHF_EXIT_TIME=repr(exit_time), OHF_NSTEP=str(nstep))
    with open(os.path.join(wd, "log.txt"), "w") as log:
        subprocess.run([DUMP, name + ".dat"], cwd=wd, env=env, stdout=log, stderr=subprocess.STDOUT, check=True, timeout=600)
    fin = dict(l.split() for l in open(os.path.join(wd, "final.txt")))
    return fin


def main(only=None):
    assert os.path.exists(DUMP), "build oracle/_ref first: make -C oracle ref"
    tmp = tempfile.mkdtemp(prefix="ohf_golden_")
    first = True
    os.makedirs(os.path.join(HERE, "geometry"), exist_ok=True)
    os.makedirs(os.path.join(HERE, "bc"), exist_ok=True)
    todo = ([(c, s, HERE) for c, s in cases().items()] + [(c, s, os.path.join(HERE, "geometry")) for c, s in geometry_cases().items()]
            + [(c, s, os.path.join(HERE, "bc")) for c, s in bc_cases().items()]
            + [(c, s, os.path.join(HERE, "restart")) for c, s in restart_cases().items()]
            + [(c, s, os.path.join(HERE, "r2")) for c, s in r2_cases().items()])
    os.makedirs(os.path.join(HERE, "r2"), exist_ok=True)
    os.makedirs(os.path.join(HERE, "restart"), exist_ok=True)
    for cname, spec, outdir in todo:
        if only and cname not in only and os.path.basename(outdir) + "/" + cname not in only:
            continue
        if only and outdir != HERE and cname in cases() and os.path.basename(outdir) + "/" + cname not in only:
            continue            # a restart case shares its name with the fresh one: ask for it as restart/<name>
        runs = [(n, 1) for n in spec["steps"]]
        if spec.get("two_cycle"):
            runs.append((spec["two_cycle"], 2))
        for nstep, ncyc in runs:
            tag = f"{cname}_n{nstep}" + ("x2" if ncyc == 2 else "")
            deck = make_deck(cname, **spec["deck"])
            wd = os.path.join(tmp, tag)
            if spec.get("restart_after"):
                tag = f"{cname}_r{spec['restart_after']}_n{nstep}"
                run_ref(deck, cname, spec["restart_after"], 1e-300, wd)       # leaves <case>.hf2d, the swap file
                fin = run_ref(deck, cname, nstep, 1e-300, wd, clean=False)    # "Use old 2D-swap file"
                assert "Use old 2D-swap file" in open(os.path.join(wd, "log.txt")).read()
            else:
                fin = run_ref(deck, cname, nstep, 1e-300, wd)
            if ncyc == 2:  # second pass: stop after the 2nd cycle (GlobalTime after 1 cycle < T <= after 2)
                t1 = float(fin["GlobalTime"])
                fin = run_ref(deck, cname, nstep, 1.5 * t1, wd)
            assert int(fin["last_iter"]) == nstep * ncyc, (tag, fin)
            if not os.path.exists(os.path.join(wd, "state0.hf2d")):
                raise RuntimeError(open(os.path.join(wd, "log.txt")).read()[-2000:])
            if first and outdir == HERE:
                shutil.copy(os.path.join(wd, "layout.txt"), os.path.join(HERE, "layout.txt"))
                first = False
            cfg = open(os.path.join(wd, "cfg.txt")).read()
            s0 = np.fromfile(os.path.join(wd, "state0.hf2d"), dtype=np.uint8)
            sN = np.fromfile(os.path.join(wd, cname + ".hf2d"), dtype=np.uint8)
            rms = open(os.path.join(wd, "RMS-" + cname + ".plt")).read()
            np.savez_compressed(os.path.join(outdir, tag + ".npz"), state0=s0, stateN=sN, cfg=np.array(cfg), nsteps=nstep, ncycles=ncyc,
                                global_time=float(fin["GlobalTime"]), dt_final=float(fin["dt"]), rms_file=np.array(rms))
            print(f"{tag}: {os.path.getsize(os.path.join(outdir, tag + '.npz')) / 1024:.0f} KiB  t={fin['GlobalTime']} dt={fin['dt']}")
    shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    main(set(sys.argv[1:]) or None)

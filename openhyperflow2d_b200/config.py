"""Solver-step configuration: every scalar and table the DEEPS2D step reads from the reference's globals.

Reference: the file-scope globals of libDEEPS2D/deeps2d_core.cpp:20-158 filled by `InitSharedData`
(:160-499) and `InitDEEPS2D` (:2835-4682); the list is SURVEY.md §8(b) "Inputs".  The text form read here is
the one a reference-side binding writes right before it calls `DEEPS2D_Run` (one `key value...` per line,
tables as `table <name> <n> x0 y0 x1 y1 ...`).
"""
from __future__ import annotations

TABLE_ORDER = ["Cp_Fuel", "Cp_OX", "Cp_cp", "Cp_air", "lam_Fuel", "lam_OX", "lam_cp", "lam_air",
               "mu_Fuel", "mu_OX", "mu_cp", "mu_air", "CFL_Scenario", "beta_Scenario"]

INT_KEYS = {"MaxX", "MaxY", "bFF", "ProblemType", "FT", "TurbStartIter", "TurbExtModel", "isTurbulenceReset",
            "isAdiabaticWall", "isAlternateRMS", "AddSrcStartIter", "last_iter", "Nstep", "NOutStep", "MonitorIndex"}


def parse_cfg(path_or_text: str) -> dict:
    try:
        with open(path_or_text) as f:
            text = f.read()
    except (FileNotFoundError, OSError):
        text = path_or_text
    cfg = {"tables": {}}
    for line in text.splitlines():
        t = line.split()
        if not t:
            continue
        if t[0] == "table":
            n = int(t[2])
            v = [float(x) for x in t[3:3 + 2 * n]]
            cfg["tables"][t[1]] = (v[0::2], v[1::2])
        elif t[0] == "Hu":
            cfg["Hu"] = [float(x) for x in t[1:5]]
        elif t[0] in INT_KEYS:
            cfg[t[0]] = int(t[1])
        else:
            cfg[t[0]] = float(t[1])
    return cfg


def dump_cfg(cfg: dict) -> str:
    lines = []
    for k, v in cfg.items():
        if k == "tables":
            continue
        if k == "Hu":
            lines.append("Hu " + " ".join(repr(float(x)) for x in v))
        elif k in INT_KEYS:
            lines.append(f"{k} {int(v)}")
        else:
            lines.append(f"{k} {float(v)!r}")
    for name, (x, y) in cfg["tables"].items():
        flat = " ".join(f"{float(a)!r} {float(b)!r}" for a, b in zip(x, y))
        lines.append(f"table {name} {len(x)} {flat}".rstrip())
    return "\n".join(lines) + "\n"

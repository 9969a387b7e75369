"""Helpers shared by the parity tests: load the committed golden vectors (tests/golden/*.npz, produced by the
unmodified reference solver through tests/golden/make_golden.py)."""
import glob
import os

import numpy as np

from openhyperflow2d_b200 import aos, config

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
LAYOUT = aos.parse_layout(os.path.join(GOLDEN, "layout.txt"))


def case_names():
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "*.npz")))


def geometry_case_names():
    """Bodies drawn by the reference's own rasterisers (circles, rectangle, NACA airfoil): tests/golden/geometry/."""
    return sorted("geometry/" + os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "geometry", "*.npz")))


def bc_case_names():
    """Boundary-condition branches (soft BCs, wall law, eps wall rewrites, integral model, single flags): tests/golden/bc/."""
    return sorted("bc/" + os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "bc", "*.npz")))


def restart_case_names():
    """Runs restarted from a swap file: non-zero y_plus, restored wall nodes (tests/golden/restart/)."""
    return sorted("restart/" + os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "restart", "*.npz")))


def r2_case_names():
    """Round-2 fixtures (tests/golden/r2/): burning decks with fuel and products present, grids that hold plain tiles
    (axisymmetric / Prandtl / Smagorinsky / SA / odd ny), 200-step drift cases."""
    return sorted("r2/" + os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "r2", "*.npz")))


def load_case(name):
    z = np.load(os.path.join(GOLDEN, *name.split("/")) + ".npz")
    cfg = config.parse_cfg(str(z["cfg"]))
    nx, ny = cfg["MaxX"], cfg["MaxY"]
    dt = aos.ref_dtype(LAYOUT)
    s0 = np.frombuffer(z["state0"].tobytes(), dtype=dt).reshape(nx, ny)
    sN = np.frombuffer(z["stateN"].tobytes(), dtype=dt).reshape(nx, ny)
    return dict(cfg=cfg, state0=s0, stateN=sN, nsteps=int(z["nsteps"]), ncycles=int(z["ncycles"]),
                global_time=float(z["global_time"]), dt_final=float(z["dt_final"]), rms_file=str(z["rms_file"]))


# fields of the reference record that the DEEPS2D step owns (everything but scratch RX/RY and bookkeeping)
STEP_FIELDS = ["S", "dSdx", "dSdy", "beta", "A", "B", "F", "Src", "SrcAdd", "U", "V", "Tg", "p", "k", "R", "CP", "lam", "mu",
               "mu_t", "lam_t", "Y", "Diff", "Q_conv", "droYdx", "droYdy", "dUdx", "dUdy", "dVdx", "dVdy", "dTdx", "dTdy",
               "dkdx", "dkdy", "depsdx", "depsdy", "CT", "TurbType", "idXl", "idXr", "idYu", "idYd"]

#!/usr/bin/env python
"""Times hf2d_gpu_set_min_distance_to_wall (k_wall_distance) and hf2d_gpu_recalc_y_plus on the bench grid
(three cylinders, k-eps, 2048 x 4096 by default).  Run on the B200 box:   python tools/bench_wall_distance.py [nx ny]
Prints one JSON line: node x wall pairs per second and the FP64 operation rate (5 operations + 1 compare per pair,
DESIGN.md 4.3b) of the whole call (kernel + 17 B/node of result copies), best of 3."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from openhyperflow2d_b200 import cabi, gridgen  # noqa: E402


def wall_list(ct):
    """(i, j) of the non-solid wall-law / no-slip nodes, rows outermost (GetWallNodes order, deeps2d_core.cpp:2081-2104)."""
    ct = np.asarray(ct).astype(np.uint64)
    m = ((ct & 0x40000000) == 0) & (((ct & 0x04000000) != 0) | ((ct & 0x08000000) != 0))
    ji = np.argwhere(m.T)                      # sorted by j, then i
    return np.ascontiguousarray(ji[:, ::-1], dtype=np.int32)


def main():
    nx = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
    ny = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
    J, cfg, lay = gridgen.three_cylinders(nx, ny, model="keps")
    walls = wall_list(J["CT"])
    s = cabi.Solver(cfg, lay)
    s.upload_aos(J)
    best = None
    for _ in range(3):
        t0 = time.perf_counter()
        l_min, i_wall, j_wall, status = s.set_min_distance_to_wall(walls)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    n_done = int((status == 1).sum())
    pairs = n_done * len(walls)
    ty = None
    # nodes farther from every wall than the domain extent keep no nearest wall (-1): give them the first wall node, as a
    # record whose members were never assigned would point somewhere arbitrary too
    iw = np.where(i_wall < 0, walls[0, 0], i_wall).astype(np.int32); jw = np.where(i_wall < 0, walls[0, 1], j_wall).astype(np.int32)
    for _ in range(3):
        t0 = time.perf_counter()
        s.recalc_y_plus(iw, jw)
        d = time.perf_counter() - t0
        ty = d if ty is None else min(ty, d)
    s.close()
    print(json.dumps({"grid": [nx, ny], "wall_nodes": int(len(walls)), "nodes": n_done, "pairs": pairs,
                      "wall_distance_s": best, "pairs_per_s": pairs / best, "fp64_ops_per_s": 5 * pairs / best,
                      "recalc_y_plus_s": ty, "l_min_range": [float(l_min[status == 1].min()), float(l_min[status == 1].max())]}))


if __name__ == "__main__":
    main()

"""Diagnostic: where does the GPU path differ from the oracle on the plain-tile grid (one model, few steps)."""
import sys
sys.path.insert(0, "tests"); sys.path.insert(0, "oracle"); sys.path.insert(0, ".")
import numpy as np
import oracle_py as O
from openhyperflow2d_b200 import cabi, gridgen
model, nsteps = (sys.argv[1], int(sys.argv[2])) if len(sys.argv) > 2 else ("euler", 1)
J, cfg, lay = gridgen.three_cylinders(176, 420, model=model)
cfg["isAlternateRMS"] = 0
nodes = O.nodes_from_ref(J)
O.run_cycle(nodes, cfg, 0, nsteps, cfg["dt"])
s = cabi.Solver(cfg, lay); Jg = np.array(J, copy=True); s.upload_aos(Jg)
print("tiles", s.tile_counts()); s.step(nsteps); s.download_aos(Jg); s.close()
for f in ("S", "A", "B", "beta"):
    a, b = np.asarray(Jg[f], float), np.asarray(nodes[f], float)
    for k in range(a.shape[-1]):
        bad = (a[..., k] != b[..., k]) & ~((a[..., k] != a[..., k]) & (b[..., k] != b[..., k]))
        if bad.any():
            idx = np.argwhere(bad)
            print(f, k, "bad", len(idx), "i range", idx[:, 0].min(), idx[:, 0].max(), "j%31 hist", np.bincount(idx[:, 1] % 32, minlength=32).tolist(),
                  "i%16 hist", np.bincount(idx[:, 0] % 16, minlength=16).tolist(), "first", idx[:5].tolist(),
                  "worst", [(int(i), int(j), float(a[i, j, k]), float(b[i, j, k])) for i, j in
                            np.argwhere(np.abs(a[..., k] - b[..., k]) > 0.1 * np.abs(a[..., k] - b[..., k]).max())[:6]],
                  "maxrel", float(np.abs(a[..., k] - b[..., k]).max() / max(np.abs(b[..., k]).max(), 1e-300)))

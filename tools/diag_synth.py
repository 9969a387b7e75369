"""Diagnostic: GPU vs oracle on the synthetic three-cylinder grid, per-field mismatch report."""
import sys
sys.path.insert(0, "oracle")
import numpy as np
import oracle_py as O
from openhyperflow2d_b200 import cabi, gridgen
model = sys.argv[1] if len(sys.argv) > 1 else "keps"
nsteps = int(sys.argv[2]) if len(sys.argv) > 2 else 1
nx, ny = (int(sys.argv[3]), int(sys.argv[4])) if len(sys.argv) > 4 else (96, 64)
J, cfg, lay = gridgen.three_cylinders(nx, ny, model=model)
nodes = O.nodes_from_ref(J)
rc, dtm, orep, tp = O.run_cycle(nodes, cfg, 0, nsteps, cfg["dt"])
s = cabi.Solver(cfg, lay)
Jg = np.array(J, copy=True)
rep = s.run_cycle(Jg, nsteps)
print("oracle rc", rc, "dt", orep.dt, rep.dt, "launches", rep.kernel_launches)
for f in ["S", "beta", "A", "B", "Src", "SrcAdd", "U", "V", "Tg", "p", "k", "R", "CP", "lam", "mu", "mu_t"]:
    a, b = np.asarray(Jg[f], float), np.asarray(nodes[f], float)
    bad = (a != b) & ~((a != a) & (b != b))
    if bad.any():
        idx = tuple(int(v) for v in np.argwhere(bad)[0])
        i, j = idx[0], idx[1]
        print(f, int(bad.sum()), "first", idx, "gpu", a[idx], "oracle", b[idx], "CT", hex(int(J["CT"][i, j])), "TT", hex(int(J["TurbType"][i, j])),
              "ids", int(J["idXl"][i, j]), int(J["idXr"][i, j]), int(J["idYu"][i, j]), int(J["idYd"][i, j]),
              "per-k", [int(bad[..., k].sum()) for k in range(a.shape[-1])] if a.ndim == 3 else "")
print("done")

"""Diagnostic (not a test): per-field mismatch statistics of the GPU path against every golden case."""
import sys
sys.path.insert(0, "tests"); sys.path.insert(0, "oracle")
import numpy as np
from golden_util import LAYOUT, case_names, load_case
from test_gpu_parity import run_gpu

names = sys.argv[1:] or case_names()
for name in names:
    c = load_case(name)
    try:
        J, rep, t = run_gpu(c)
    except Exception as e:
        print(f"{name}: FAILED {e}")
        continue
    ref = c["stateN"]
    out = []
    for f in ["S", "beta", "A", "B", "F", "Src", "SrcAdd", "U", "V", "Tg", "p", "k", "R", "CP", "lam", "mu", "mu_t", "Q_conv"]:
        a, b = np.asarray(J[f], float), np.asarray(ref[f], float)
        bad = (a != b) & ~((a != a) & (b != b))
        if bad.any():
            if a.ndim == 3:
                per = [int(bad[..., k].sum()) for k in range(a.shape[-1])]
                rel = max(np.abs(a[..., k] - b[..., k]).max() / max(np.abs(b[..., k]).max(), 1e-300) for k in range(a.shape[-1]))
                idx = np.argwhere(bad)[0]
                out.append(f"{f}:{per} rel={rel:.2e} first={tuple(idx)}")
            else:
                rel = np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)
                out.append(f"{f}:{int(bad.sum())} rel={rel:.2e} first={tuple(np.argwhere(bad)[0])}")
    print(f"{name}: dt_min {rep.dt_min!r} vs {c['dt_final']!r} t {t!r} vs {c['global_time']!r} | " + ("; ".join(out) if out else "ALL EXACT"))

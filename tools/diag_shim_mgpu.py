#!/usr/bin/env python
"""ohf_gpu with HF2D_GPUS=1 vs N on a shim-test deck: which members / nodes differ (run on the GPU box)."""
import os, sys, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import test_shim_dropin as T
case, nmax, nout, exit_time = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), float(sys.argv[4])
n = int(sys.argv[5]) if len(sys.argv) > 5 else 2
deck = T._deck(case, nmax, nout, exit_time)
wd = tempfile.mkdtemp()
J1, r1, o1 = T._run(T.GPU, deck, case, wd + "/g1")
Jn, rn, on = T._run(T.GPU, deck, case, wd + "/gn", env={"HF2D_GPUS": str(n), "CUDA_DEVICE_MAX_CONNECTIONS": "32"})
print([l for l in on.splitlines() if "X-slabs" in l])
print("records", len(r1), len(rn))
for f in Jn.dtype.names:
    a, b = np.asarray(Jn[f]), np.asarray(J1[f])
    same = (a == b) | ((a != a) & (b != b)) if a.dtype.kind == "f" else (a == b)
    if not same.all():
        idx = np.argwhere(~same)
        cols = sorted(set(int(i[0]) for i in idx))
        ks = sorted(set(int(i[2]) for i in idx)) if a.ndim == 3 else None
        rel = np.abs(np.nan_to_num(a.astype(float)) - np.nan_to_num(b.astype(float))).max() / max(np.abs(np.nan_to_num(b.astype(float))).max(), 1e-300)
        print(f"{f}: {len(idx)} entries differ, columns {cols[:12]}{'...' if len(cols) > 12 else ''} ({len(cols)} cols) k={ks} rel {rel:.3e}")

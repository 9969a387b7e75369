"""Diagnostic: Stage-1 output (SN planes) after n full steps + begin + stage1, saved for an A/B between two builds."""
import ctypes as C, sys
sys.path.insert(0, ".")
import numpy as np
import torch
from openhyperflow2d_b200 import cabi, gridgen
model, nsteps, out = sys.argv[1], int(sys.argv[2]), sys.argv[3]
J, cfg, lay = gridgen.three_cylinders(176, 420, model=model)
s = cabi.Solver(cfg, lay); s.upload_aos(np.array(J, copy=True))
if nsteps:
    s.step(nsteps)
s.step_begin(); s.step_stage1(); torch.cuda.synchronize()
N = J.size
rt = C.CDLL("libcudart.so.12")
planes = [("SN", k) for k in range(9)] + [("S", k) for k in range(9)] + [("A", k) for k in range(9)] + [("B", k) for k in range(9)] + \
         [("beta", k) for k in range(9)] + [("Src", k) for k in range(9)] + [(f, 0) for f in ("U", "V", "Tg", "p", "k", "R", "CP", "lam", "mu", "mu_t", "l_min", "y_plus")]
res = np.zeros((len(planes), N))
for q, (f, k) in enumerate(planes):
    p = s.field_ptr(f, k)
    rt.cudaMemcpy(C.c_void_p(res[q].ctypes.data), C.c_void_p(p), C.c_size_t(N * 8), 2)
np.save(out, res.reshape(len(planes), *J.shape))
open(out + ".names", "w").write(repr(planes))

"""Where the end-to-end cycle spends its time: upload / steps / download of the bench workload (run on the GPU box)."""
import sys, time
import torch
sys.path.insert(0, ".")
from openhyperflow2d_b200.multigpu import SlabRunner

nx, ny = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (2048, 4096)
r = SlabRunner(nx, ny, model="keps")
host = torch.empty(r.J.nbytes, dtype=torch.uint8, pin_memory=True)
Jp = host.numpy().view(r.J.dtype).reshape(r.J.shape)
Jp[...] = r.J
print("image GB", r.J.nbytes / 1e9, "node bytes", r.J.dtype.itemsize)
for rep in range(2):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    r.solver.upload_aos(Jp, pinned=True); torch.cuda.synchronize(); t1 = time.perf_counter()
    r.solver.step(100); torch.cuda.synchronize(); t2 = time.perf_counter()
    r.solver.download_aos(Jp, pinned=True); torch.cuda.synchronize(); t3 = time.perf_counter()
    print(f"upload {t1 - t0:.3f} s  steps {t2 - t1:.3f} s  download {t3 - t2:.3f} s")
# raw PCIe speed for the same image
d = torch.empty(r.J.nbytes, dtype=torch.uint8, device="cuda")
for _ in range(2):
    torch.cuda.synchronize(); t0 = time.perf_counter(); d.copy_(host, non_blocking=True); torch.cuda.synchronize(); t1 = time.perf_counter()
    host.copy_(d, non_blocking=True); torch.cuda.synchronize(); t2 = time.perf_counter()
    print(f"H2D {r.J.nbytes / (t1 - t0) / 1e9:.1f} GB/s  D2H {r.J.nbytes / (t2 - t1) / 1e9:.1f} GB/s")

#!/usr/bin/env python
"""Times the device-side Tecplot formatter (hf2d_gpu_tecplot_format + _fetch) on the bench grid (2048 x 4096 k-eps by default).
Run on the B200 box:   python tools/bench_tecplot.py [nx ny]      Prints one JSON line."""
import ctypes as C
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from openhyperflow2d_b200 import cabi, gridgen  # noqa: E402


def main():
    nx = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
    ny = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
    J, cfg, lay = gridgen.three_cylinders(nx, ny, model="keps")
    s = cabi.Solver(cfg, lay)
    s.upload_aos(J)
    s.step(3)
    L = cabi.lib()
    best_f = best_c = None
    nbytes = 0
    for _ in range(3):
        n = C.c_size_t()
        t0 = time.perf_counter()
        cabi._ck(L.hf2d_gpu_tecplot_format(s._h, C.byref(s.layout), 0, cfg["dx"] * nx / (nx - 1), cfg["dy"] * ny / (ny - 1), C.byref(n)))
        t1 = time.perf_counter()
        buf = C.create_string_buffer(n.value)
        t2 = time.perf_counter()
        cabi._ck(L.hf2d_gpu_tecplot_fetch(s._h, buf, n.value))
        t3 = time.perf_counter()
        nbytes = n.value
        best_f = (t1 - t0) if best_f is None else min(best_f, t1 - t0)
        best_c = (t3 - t2) if best_c is None else min(best_c, t3 - t2)
    first = buf.raw[:200].decode(errors="replace")
    s.close()
    print(json.dumps({"grid": [nx, ny], "lines": nx * ny, "text_bytes": nbytes, "format_s": best_f, "fetch_s": best_c,
                      "lines_per_s": nx * ny / best_f, "first_line": first.split("\n")[0]}))


if __name__ == "__main__":
    main()

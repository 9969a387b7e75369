"""Host-side view of the reference's AoS grid image (`UMatrix2D<FlowNode2D<double,3>>`).

The reference keeps the whole grid as one array of 1248-byte `FlowNode2D` records, `J[x*nY + y]`
(utl/umatrix2d.hpp:237-240; record: libOpenHyperFLOW2D/hyper_flow_node.hpp:138-191).  The record layout is
compiler/ABI specific, so it is never hand-copied here: a *layout file* (written with `offsetof` on the
reference's own header by the reference-side binding; the fixture
`tests/golden/layout.txt` comes from oracle/ref_hook.cpp) names each member's offset, element size and count.
"""
from __future__ import annotations

import numpy as np

INT_FIELDS = {"ix", "iy", "idXl", "idYu", "idXr", "idYd", "NGX", "NGY", "i_wall", "j_wall"}
UINT_FIELDS = {"CT", "TurbType"}


def parse_layout(path_or_text: str) -> dict:
    """Parse a layout file -> {"node_bytes": int, "num_eq": int, "fields": {name: (offset, itemsize, count)}}."""
    try:
        with open(path_or_text) as f:
            text = f.read()
    except (FileNotFoundError, OSError):
        text = path_or_text
    out = {"fields": {}}
    counts = {}
    for line in text.splitlines():
        t = line.split()
        if not t:
            continue
        if t[0] == "field":
            out["fields"][t[1]] = [int(t[2]), int(t[3]), 1]
        elif t[0] == "count":
            counts[t[1]] = int(t[2])
        else:
            out[t[0]] = int(t[1])
    for name, c in counts.items():
        out["fields"][name][2] = c
    out["fields"] = {k: tuple(v) for k, v in out["fields"].items()}
    return out


def ref_dtype(layout: dict) -> np.dtype:
    """numpy structured dtype that overlays the reference's record exactly (padding included)."""
    names, formats, offsets = [], [], []
    for name, (off, size, cnt) in layout["fields"].items():
        if name in UINT_FIELDS:
            base = {8: "<u8", 4: "<u4"}[size]
        elif name in INT_FIELDS:
            base = {8: "<i8", 4: "<i4"}[size]
        else:
            base = {8: "<f8", 4: "<f4"}[size]
        names.append(name)
        formats.append(base if cnt == 1 else (base, (cnt,)))
        offsets.append(off)
    return np.dtype({"names": names, "formats": formats, "offsets": offsets, "itemsize": layout["node_bytes"]})


def load_aos(path: str, layout: dict, nx: int, ny: int) -> np.ndarray:
    """Read a raw grid image (swap file `<Project>.hf2d`, obj_data.cpp:117-320) as a (nx, ny) record array."""
    dt = ref_dtype(layout)
    a = np.fromfile(path, dtype=dt)
    if a.size != nx * ny:
        raise ValueError(f"{path}: {a.size} records, expected {nx}*{ny}")
    return a.reshape(nx, ny)

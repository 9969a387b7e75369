"""hf2d_gpu_wall_distance_noalloc (include/hf2d_gpu.h): the wall-distance pass for callers that cannot call cudaMalloc --
its scratch is a fixed file-scope device array.  No counterpart in the reference.
Checked through the C ABI against the existing allocating entry point hf2d_gpu_set_min_distance_to_wall, bit
for bit (the tolerance every wall-distance test here uses), on two calls in a row with DIFFERENT wall lists, the second
shorter than the first: the shared array keeps the first call's tail and the second call must not see it.

Collected last on purpose (a failure here must not hide another test)."""
import numpy as np
import pytest

from golden_util import LAYOUT, load_case

pytestmark = pytest.mark.gpu


def _planes(torch, nx, ny):
    dev = torch.device("cuda", 0)
    return (torch.full((nx, ny), -7.0, dtype=torch.float64, device=dev), torch.full((nx, ny), -7, dtype=torch.int32, device=dev),
            torch.full((nx, ny), -7, dtype=torch.int32, device=dev), torch.full((nx, ny), 9, dtype=torch.uint8, device=dev))


@pytest.mark.parametrize("name", ["ns_keps_std_n8", "geometry/cylinders_keps_n8"])
def test_no_alloc_wall_distance_equals_the_allocating_entry_point(name):
    import torch
    import oracle_py as O
    from openhyperflow2d_b200 import cabi
    c = load_case(name)
    cfg = c["cfg"]
    nx, ny = cfg["MaxX"], cfg["MaxY"]
    walls_all = O.wall_nodes(np.ascontiguousarray(c["state0"]["CT"], dtype=np.uint64))
    assert len(walls_all) > 8
    ref = cabi.Solver(cfg, LAYOUT)            # the existing entry point, on a handle of its own
    J = np.array(c["state0"], copy=True)
    ref.upload_aos(J)
    s = cabi.Solver(cfg, LAYOUT)
    s.upload_aos(np.array(c["state0"], copy=True))
    outs = _planes(torch, nx, ny)             # allocated by the CALLER, before the calls
    torch.cuda.synchronize()
    for walls, x0 in ((walls_all, 0.0), (walls_all[: len(walls_all) // 2], 0.003)):   # two calls in a row
        l_ref, i_ref, j_ref, st_ref = ref.set_min_distance_to_wall(walls, x0=x0)
        s.wall_distance_noalloc(walls, outs[0].data_ptr(), outs[1].data_ptr(), outs[2].data_ptr(), outs[3].data_ptr(), x0=x0)
        l, iw, jw, st = (t.cpu().numpy() for t in outs)
        assert (st == st_ref).all()
        assert (iw == i_ref).all() and (jw == j_ref).all()
        assert (l == l_ref).all() or ((l == l_ref) | ((l != l) & (l_ref != l_ref))).all()
        # the handle's own l_min plane was updated in place, as the allocating entry point does
        a = _dev_to_host(s.field_ptr("l_min"), nx, ny); b = _dev_to_host(ref.field_ptr("l_min"), nx, ny)
        assert ((a == b) | ((a != a) & (b != b))).all()
    s.close(); ref.close()


def _cudart():
    import ctypes
    import glob
    import os
    import torch
    cands = ["libcudart.so.12", "libcudart.so"] + glob.glob(os.path.join(os.path.dirname(os.path.dirname(torch.__file__)), "nvidia", "cuda_runtime", "lib", "libcudart.so*")) \
        + glob.glob("/usr/local/cuda/lib64/libcudart.so*")
    for n in cands:
        try:
            rt = ctypes.CDLL(n)
            rt.cudaMemcpy.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int]
            return rt
        except OSError:
            continue
    pytest.skip("no libcudart to read a raw device pointer")


def _dev_to_host(ptr, nx, ny):
    """nx*ny doubles at device address `ptr` (a plane of the handle) -> numpy."""
    import torch
    torch.cuda.synchronize()
    out = np.empty((nx, ny), dtype=np.float64)
    assert _cudart().cudaMemcpy(out.ctypes.data, ptr, out.nbytes, 2) == 0      # cudaMemcpyDeviceToHost
    return out


def test_no_alloc_wall_distance_refuses_what_it_cannot_hold():
    import torch
    from openhyperflow2d_b200 import cabi
    c = load_case("ns_laminar_n8")
    cfg = c["cfg"]
    nx, ny = cfg["MaxX"], cfg["MaxY"]
    s = cabi.Solver(cfg, LAYOUT)
    s.upload_aos(np.array(c["state0"], copy=True))
    outs = _planes(torch, nx, ny)
    too_many = np.zeros((16385, 2), dtype=np.int32)
    with pytest.raises(cabi.Hf2dError, match="exceed the fixed scratch"):
        s.wall_distance_noalloc(too_many, *(t.data_ptr() for t in outs))
    with pytest.raises(cabi.Hf2dError):
        s.wall_distance_noalloc(too_many[:4], 0, outs[1].data_ptr(), outs[2].data_ptr(), outs[3].data_ptr())
    s.close()

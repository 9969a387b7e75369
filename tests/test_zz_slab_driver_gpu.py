"""The C X-slab driver (csrc/hf2d_slab.inc, hf2d_slab_* in include/hf2d_gpu.h): slabs cut from ONE host image by
active-node count, ghost columns stored straight into the neighbour's mailbox (peer memory + flag word), dt and residual
all-to-all, the overlapped schedule with the next step's hand-off hidden inside Stage 2 -- against the single-GPU run of
the undivided grid, BIT FOR BIT.  Replaces for the GPU path hf2d_start.cpp:149-175 + deeps2d_core.cpp:1336-1463.

On a 1-GPU box every rank runs on device 0 (the mailboxes are then ordinary device memory; the kernels, flags and the
schedule are the same); with >= 2 GPUs the ranks are spread over them and the stores cross NVLink."""
import numpy as np
import pytest

from test_gpu_parity import GPU_FIELDS, NS_FIELDS, rel_linf

pytestmark = pytest.mark.gpu


def _single(J, cfg, lay, chunks):
    from openhyperflow2d_b200 import cabi
    s = cabi.Solver(cfg, lay)
    Jg = np.array(J, copy=True)
    s.upload_aos(Jg)
    reps = [s.step(n) if q == 0 else s.step_continue(n) for q, n in enumerate(chunks)]
    s.download_aos(Jg)
    s.close()
    return Jg, reps


def _devices(world):
    import torch
    nd = torch.cuda.device_count()
    return [r % nd for r in range(world)]


def _same(a, b):
    a, b = np.asarray(a), np.asarray(b)
    return ((a == b) | ((a != a) & (b != b))).all()


@pytest.mark.parametrize("model,world", [("keps", 2), ("keps", 3), ("laminar", 4), ("sa", 2)])
def test_slab_driver_reproduces_the_undivided_grid(model, world):
    from openhyperflow2d_b200 import cabi, gridgen
    nx, ny = 104, 160
    J, cfg, lay = gridgen.three_cylinders(nx, ny, model=model)
    cfg["isAlternateRMS"] = 0
    cfg["NOutStep"] = 3
    chunks = [5, 2] if model != "sa" else [2]
    ref, ref_reps = _single(J, cfg, lay, chunks)
    cuts = cabi.slab_cuts(J, lay, world)
    assert cuts[0] == 0 and cuts[-1] == nx and all(b - a >= 2 for a, b in zip(cuts, cuts[1:]))
    act = ((J["CT"] & 0x80000000) != 0) & ((J["CT"] & 0x40000000) == 0)
    per = [int(act[a:b].sum()) for a, b in zip(cuts, cuts[1:])]
    assert max(per) - min(per) <= 2 * ny, per            # equal active nodes up to one column either side
    slabs = [cabi.Slab(cfg, lay, dev, r, world, cuts) for r, dev in enumerate(_devices(world))]
    cabi.Slab.connect_local(slabs)
    Jg = np.array(J, copy=True)
    for s in slabs:
        s.upload(Jg)
    reps = [cabi.Slab.run_all(slabs, n, new_cycle=(q == 0)) for q, n in enumerate(chunks)]
    for s in slabs:
        s.download(Jg)
    for s in slabs:
        s.close()
    for f in GPU_FIELDS + NS_FIELDS + ["CT", "TurbType"]:
        assert _same(Jg[f], ref[f]), f"{model} x{world}: {f} differs, rel-Linf {rel_linf(np.nan_to_num(Jg[f]), np.nan_to_num(ref[f])):.3e}"
    for a, b in zip(reps, ref_reps):
        assert a.dt == b.dt and a.dt_min == b.dt_min and a.iter == b.iter and a.time_part == b.time_part
        np.testing.assert_allclose(list(a.sum_RMS), list(b.sum_RMS), rtol=1e-12, atol=1e-300)     # summed per slab, then over slabs
        assert list(a.DD_max) == list(b.DD_max)
        assert a.k_max_RMS == b.k_max_RMS


def test_slab_driver_on_a_reference_image_with_unequal_slabs():
    """A golden start image of the unmodified reference (ramp duct with a solid block: active nodes are not uniform in x),
    three slabs, 8 steps: equals the reference's own result as the single-GPU path does."""
    from golden_util import LAYOUT, load_case
    from openhyperflow2d_b200 import cabi
    c = load_case("ns_keps_std_n8")
    cfg = c["cfg"]
    J = np.array(c["state0"], copy=True)
    world = 3
    cuts = cabi.slab_cuts(J, LAYOUT, world)
    assert len(set(b - a for a, b in zip(cuts, cuts[1:]))) > 1, cuts       # the solid block makes the slabs unequal
    slabs = [cabi.Slab(cfg, LAYOUT, dev, r, world, cuts) for r, dev in enumerate(_devices(world))]
    cabi.Slab.connect_local(slabs)
    for s in slabs:
        s.upload(J)
    rep = cabi.Slab.run_all(slabs, c["nsteps"])
    for s in slabs:
        s.download(J)
        s.close()
    single, _ = _single(c["state0"], cfg, LAYOUT, [c["nsteps"]])
    for f in GPU_FIELDS + NS_FIELDS + ["Q_conv"]:
        assert _same(J[f], single[f]), f
    assert rep.dt_min == c["dt_final"] or abs(rep.dt_min - c["dt_final"]) <= 1e-12 * c["dt_final"]
    for f in ("S", "U", "V", "Tg", "p"):
        assert rel_linf(J[f], c["stateN"][f]) <= 1e-10, f


def _ipc_worker(rank, world, model, nx, ny, nsteps, out_dir):
    import os
    import time
    import torch
    from openhyperflow2d_b200 import cabi, gridgen
    dev = rank % torch.cuda.device_count()
    torch.cuda.set_device(dev)
    J, cfg, lay = gridgen.three_cylinders(nx, ny, model=model)
    cfg["isAlternateRMS"] = 0
    cuts = cabi.slab_cuts(J, lay, world)
    s = cabi.Slab(cfg, lay, dev, rank, world, cuts)
    with open(f"{out_dir}/h{rank}.tmp", "wb") as f:
        f.write(s.ipc_handle())
    os.rename(f"{out_dir}/h{rank}.tmp", f"{out_dir}/h{rank}.bin")
    t0 = time.time()
    while not all(os.path.exists(f"{out_dir}/h{r}.bin") for r in range(world)):
        assert time.time() - t0 < 120
        time.sleep(0.01)
    s.connect_ipc(b"".join(open(f"{out_dir}/h{r}.bin", "rb").read() for r in range(world)))
    s.upload(J)
    rep = s.run(nsteps)
    s.download(J)
    np.save(f"{out_dir}/r{rank}.npy", J[s.c0:s.c1].view(np.uint8))
    np.save(f"{out_dir}/rep{rank}.npy", np.array([rep.dt, rep.dt_min] + list(rep.sum_RMS)))
    # nobody closes its mailbox while a neighbour may still write into it
    open(f"{out_dir}/done{rank}", "w").close()
    while not all(os.path.exists(f"{out_dir}/done{r}") for r in range(world)):
        assert time.time() - t0 < 240
        time.sleep(0.01)
    s.close()


def test_slab_driver_one_process_per_gpu_over_cuda_ipc(tmp_path):
    """One process per rank, mailboxes exchanged as cudaIpc handles (what bench.py does under torchrun).  Needs >= 2 GPUs:
    two processes time-slicing ONE GPU cannot overlap a waiting kernel with the kernel it waits for."""
    import torch
    import torch.multiprocessing as mp
    from openhyperflow2d_b200 import gridgen
    world = min(torch.cuda.device_count(), 4)
    if world < 2:
        pytest.skip("needs at least 2 GPUs")
    model, nx, ny, nsteps = "keps", 192, 160, 7
    J, cfg, lay = gridgen.three_cylinders(nx, ny, model=model)
    cfg["isAlternateRMS"] = 0
    ref, reps = _single(J, cfg, lay, [nsteps])
    mp.spawn(_ipc_worker, args=(world, model, nx, ny, nsteps, str(tmp_path)), nprocs=world, join=True)
    from openhyperflow2d_b200 import cabi
    cuts = cabi.slab_cuts(J, lay, world)
    for r in range(world):
        got = np.load(tmp_path / f"r{r}.npy").view(ref.dtype).reshape(cuts[r + 1] - cuts[r], ny)
        for f in GPU_FIELDS + NS_FIELDS:
            assert _same(got[f], ref[cuts[r]:cuts[r + 1]][f]), f"rank {r}: {f}"
        rp = np.load(tmp_path / f"rep{r}.npy")
        assert rp[0] == reps[0].dt and rp[1] == reps[0].dt_min
        np.testing.assert_allclose(rp[2:], list(reps[0].sum_RMS), rtol=1e-12, atol=1e-300)

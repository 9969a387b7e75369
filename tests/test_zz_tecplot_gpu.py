"""Device-side Tecplot formatter (csrc/hf2d_tecplot.inc: hf2d_gpu_tecplot_format / _fetch) -- the body of the reference's
SaveData2D (deeps2d_core.cpp:2589-2673), SURVEY section 8 f-1.

(1) through the reference PROGRAM: ohf_gpu with HF2D_GPU_TECPLOT=1 writes <Project>.plt and tp-<Project>.plt from the device
    text; on decks whose fields are bit-exact they must equal the files of the UNMODIFIED program byte for byte -- the check
    against the real writer (operator<< of libstdc++);
(2) through the C ABI against a Python restatement of SaveData2D on the downloaded records ('%g' of CPython is correctly
    rounded like glibc's), incl. k-eps, solid nodes, both file modes."""
import os

import numpy as np
import pytest

from golden_util import LAYOUT, load_case

pytestmark = pytest.mark.gpu


def _save_data_2d(J, cfg, type_):
    """SaveData2D's body (deeps2d_core.cpp:2621-2672) from a record image; is_Cx_calc = 0, is_p_asterisk_out = 0."""
    nx, ny = J.shape
    dx_out = (cfg["dx"] * nx) / (nx - 1)
    dy_out = (cfg["dy"] * ny) / (ny - 1)
    import math

    def g(v):            # printf("%g"): a NaN prints with its sign (x86 produces the NEGATIVE default NaN for 0/0)
        v = float(v)
        if v != v:
            return "-nan" if math.copysign(1.0, v) < 0 else "nan"
        return "%g" % v
    out = []
    with np.errstate(all="ignore"):
        for j in range(ny):
            for i in range(nx):
                r = J[i, j]
                s = g(i * dx_out * 1.e3) + "  " + g(dy_out * j * 1.e3) + "  "
                solid = (int(r["CT"]) & 0x40000000) != 0
                mach = 0.0
                if not solid:
                    U, V, Tg, p, S0 = float(r["U"]), float(r["V"]), float(r["Tg"]), float(r["p"]), float(r["S"][0])
                    s += g(U) + "  " + g(V) + "  " + g(Tg) + "  " + g(p) + "  " + g(S0) + "  "
                    A = np.sqrt(np.float64(r["k"]) * np.float64(r["R"]) * np.float64(Tg) + 1.e-30)
                    W = np.sqrt(np.float64(U) * U + np.float64(V) * V + 1.e-30)
                    mach = float(W / A)
                    if S0 != 0.0:
                        a, b, c = (float(np.float64(r["S"][q]) / np.float64(S0)) for q in (4, 5, 6))
                        s += g(a) + "  " + g(b) + "  " + g(c) + "  " + g(abs(1 - a - b - c)) + "  "
                        s += g(float(np.float64(r["mu_t"]) / np.float64(r["mu"]))) + "  "
                    else:
                        s += " +0. +0  +0  +0  +0  "
                else:
                    s += "  0  0  " + g(float(r["Tg"])) + "  0  0  0  0  0  0  0"
                if not solid and mach > 1.e-30:
                    s += g(mach) + "  " + g(float(r["l_min"])) + " " + g(float(r["y_plus"]))
                else:
                    s += "  0  0  0  "
                out.append(s + " 0\n")
            if type_:
                out.append("\n")
    return "".join(out).encode()


@pytest.mark.parametrize("name,type_", [("ns_keps_std_n8", 0), ("ns_laminar_n8", 1), ("geometry/cylinders_keps_n8", 0), ("r2/axi_burn_keps_n10", 1)])
def test_device_tecplot_body_equals_the_host_writer(name, type_):
    from openhyperflow2d_b200 import cabi
    c = load_case(name)
    cfg = c["cfg"]
    J = np.array(c["state0"], copy=True)
    s = cabi.Solver(cfg, LAYOUT)
    s.upload_aos(J)
    s.step(c["nsteps"])
    nx, ny = J.shape
    text = s.tecplot_body(type_, (cfg["dx"] * nx) / (nx - 1), (cfg["dy"] * ny) / (ny - 1))
    s.download_aos(J)
    s.close()
    want = _save_data_2d(J, cfg, type_)
    if text != want:
        a, b = text.split(b"\n"), want.split(b"\n")
        bad = [(q, x, y) for q, (x, y) in enumerate(zip(a, b)) if x != y][:3]
        raise AssertionError(f"{name}: {len(text)} vs {len(want)} bytes, first differing lines {bad}")


@pytest.mark.parametrize("n_gpus", [1, 3])
@pytest.mark.parametrize("case,nmax,nout,exit_time", [("euler_duct", 7, 3, 5e-7), ("ns_laminar", 6, 4, 4e-7)])
def test_reference_program_writes_identical_plt_files_from_the_device(case, nmax, nout, exit_time, n_gpus, tmp_path):
    import test_shim_dropin as D
    if not (os.path.exists(D.REF) and os.path.exists(D.GPU)):
        pytest.skip("oracle/_ref not built (make -C oracle ref)")
    deck = D._deck(case, nmax, nout, exit_time)
    D._run(D.REF, deck, case, str(tmp_path / "ref"))
    env = {"HF2D_GPU_TECPLOT": "1"}
    if n_gpus > 1:          # three X-slabs (on as many GPUs as the box has): each formats its columns, the host interleaves the rows
        env.update(HF2D_GPUS=str(n_gpus), CUDA_DEVICE_MAX_CONNECTIONS="32")
    _, _, out = D._run(D.GPU, deck, case, str(tmp_path / "gpu"), env=env)
    assert "hf2d_gpu: Tecplot snapshots formatted on the GPU" in out
    for fn in (case + ".plt", "tp-" + case + ".plt"):
        a = open(tmp_path / "ref" / fn, "rb").read()
        b = open(tmp_path / "gpu" / fn, "rb").read()
        assert len(a) > 1000
        if a != b:
            la, lb = a.split(b"\n"), b.split(b"\n")
            bad = [(q, x, y) for q, (x, y) in enumerate(zip(la, lb)) if x != y][:3]
            raise AssertionError(f"{fn}: {len(a)} vs {len(b)} bytes, first differing lines {bad}")


@pytest.mark.parametrize("name,world,type_", [("ns_keps_std_n8", 3, 1), ("geometry/cylinders_keps_n8", 4, 0)])
def test_slab_tecplot_body_equals_the_undivided_one(name, world, type_):
    """hf2d_slab_tecplot: every slab formats its owned columns, the host interleaves the row segments -- the same bytes as the
    single-GPU formatter on the undivided grid (the slabs reproduce its fields bit for bit)."""
    import torch
    from openhyperflow2d_b200 import cabi
    c = load_case(name)
    cfg = c["cfg"]
    J = np.array(c["state0"], copy=True)
    nx, ny = J.shape
    dxo, dyo = (cfg["dx"] * nx) / (nx - 1), (cfg["dy"] * ny) / (ny - 1)
    s = cabi.Solver(cfg, LAYOUT)
    s.upload_aos(J)
    s.step(c["nsteps"])
    want = s.tecplot_body(type_, dxo, dyo)
    s.close()
    cuts = cabi.slab_cuts(J, LAYOUT, world)
    nd = torch.cuda.device_count()
    slabs = [cabi.Slab(cfg, LAYOUT, r % nd, r, world, cuts) for r in range(world)]
    cabi.Slab.connect_local(slabs)
    for sl in slabs:
        sl.upload(J)
    cabi.Slab.run_all(slabs, c["nsteps"])
    got = cabi.Slab.tecplot_body(slabs, type_, dxo, dyo)
    for sl in slabs:
        sl.close()
    assert len(got) == len(want) and got == want

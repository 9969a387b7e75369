"""The drop-in itself, end to end (run with -m gpu on the B200 box): the reference PROGRAM with its time loop replaced.

oracle/_ref/ohf_gpu is the reference's own objects (deck parser, grid builder, writers -- hf2d_start.cpp,
deeps2d_core.cpp with its DEEPS2D_Run renamed away) linked with openhyperflow2d_b200/shim/deeps2d_run_gpu.cpp and
libhf2d_gpu.so; oracle/_ref/ohf_serial is the unmodified reference.  Both read the same input deck from their own
working directory; the swap file (<Project>.hf2d, the raw grid image the reference restarts from,
deeps2d_core.cpp:1725-1741) and the residual history (RMS-<Project>.plt) they leave behind must agree:
bit for bit where the physics has no exp()/pow(), rel-L_inf <= 1e-10 otherwise (see test_gpu_parity.py).

The decks exercise what the shim adds on top of hf2d_gpu_step: several output cycles (hf2d_gpu_step restarts the
cycle, the grid stays resident), NOutStep > 1 (hf2d_gpu_step_continue between output steps, residuals only where the
reference looks at them), the exit monitor, and the per-cycle writers running on the downloaded image."""
import os
import subprocess
import sys

import numpy as np
import pytest

from golden_util import GOLDEN, LAYOUT
from openhyperflow2d_b200 import aos

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref", "ohf_serial")
GPU = os.path.join(ROOT, "oracle", "_ref", "ohf_gpu")

sys.path.insert(0, GOLDEN)
import make_golden  # noqa: E402  (deck generator only; nothing here touches /root/reference)

FIELDS = ["S", "beta", "U", "V", "Tg", "p", "k", "R", "CP", "lam", "mu", "mu_t", "CT", "TurbType"]


def _deck(case, nmax, nout, exit_time):
    spec = make_golden.cases()[case]
    text = make_golden.make_deck(case, **dict(spec["deck"], nmax=nmax))
    text = text.replace("<data/NOutStep=1>", f"<data/NOutStep={nout}>")
    assert "<data/ExitMonitorValue=1.e-30>" in text
    return text.replace("<data/ExitMonitorValue=1.e-30>", f"<data/ExitMonitorValue={exit_time!r}>")


def _run(binary, deck, case, wd, env=None):
    os.makedirs(wd)
    with open(os.path.join(wd, case + ".dat"), "w") as f:
        f.write(deck)
    with open(os.path.join(wd, "log.txt"), "w") as log:
        rc = subprocess.run([binary, case + ".dat"], cwd=wd, stdout=log, stderr=subprocess.STDOUT, timeout=600,
                            env=dict(os.environ, **(env or {}))).returncode
    out = open(os.path.join(wd, "log.txt")).read()
    assert rc == 0, out[-3000:]
    nx = int([l for l in deck.splitlines() if l.startswith("<data/MaxX=")][0].split("=")[1][:-1])
    ny = int([l for l in deck.splitlines() if l.startswith("<data/MaxY=")][0].split("=")[1][:-1])
    J = np.fromfile(os.path.join(wd, case + ".hf2d"), dtype=aos.ref_dtype(LAYOUT)).reshape(nx, ny)
    rms = []
    for l in open(os.path.join(wd, "RMS-" + case + ".plt")):
        try:
            rms.append([float(x) for x in l.split()])
        except ValueError:
            pass            # header lines
    return J, rms, out


CASES = [
    # case, Nmax, NOutStep, exit time (GlobalTime >= value ends the run), bit-exact?
    ("euler_duct", 7, 1, 1e-30, True),          # one cycle, every step an output step
    ("euler_duct", 7, 3, 5e-7, True),           # NOutStep 3: chunks of 1+3+3 steps, several cycles
    ("ns_laminar", 6, 4, 4e-7, True),
    ("ns_keps_std", 6, 2, 3e-7, False),
    ("axi_reacting_euler", 5, 2, 2e-7, True),
]


@pytest.mark.skipif(not (os.path.exists(REF) and os.path.exists(GPU)), reason="oracle/_ref not built (make -C oracle ref)")
@pytest.mark.parametrize("case,nmax,nout,exit_time,exact", CASES)
def test_reference_program_with_gpu_time_loop(case, nmax, nout, exit_time, exact, tmp_path):
    deck = _deck(case, nmax, nout, exit_time)
    Jr, rms_r, _ = _run(REF, deck, case, str(tmp_path / "ref"))
    Jg, rms_g, out = _run(GPU, deck, case, str(tmp_path / "gpu"))
    assert "hf2d_gpu: DEEPS2D time loop on the GPU" in out
    # same number of residual records = same number of steps, cycles and the same exit decision
    assert len(rms_g) == len(rms_r) and len(rms_r) >= 1, (len(rms_g), len(rms_r))
    for a, b in zip(rms_g, rms_r):
        assert len(a) == len(b)
        # the file holds 6 significant digits: a last-digit flip is 1e-6 relative
        assert np.allclose(a, b, rtol=2e-5, atol=1e-300), (a, b)
    for f in FIELDS:
        a, b = Jg[f], Jr[f]
        if exact or a.dtype.kind in "iu":
            same = (a == b) | ((a != a) & (b != b))
            assert same.all(), f"{case}: {f} differs in {int((~same).sum())} entries"
        else:
            scale = max(float(np.abs(b).max()), 1e-300)
            assert float(np.abs(a - b).max()) / scale <= 1e-10, f



@pytest.mark.skipif(not (os.path.exists(REF) and os.path.exists(GPU)), reason="oracle/_ref not built (make -C oracle ref)")
@pytest.mark.parametrize("n_gpus", [2, 4])
@pytest.mark.parametrize("case,nmax,nout,exit_time,exact", [CASES[2], CASES[3]])
def test_reference_program_on_several_gpus(case, nmax, nout, exit_time, exact, n_gpus, tmp_path):
    """HF2D_GPUS=N: the reference program's grid cut into N X-slabs by active-node count and stepped by the C slab driver
    (csrc/hf2d_slab.inc; ghost columns stored straight into the neighbour's HBM).  The swap file it leaves is the single-GPU
    program's BIT FOR BIT (exact seams), hence the unmodified program's within the usual bound.  With fewer than N GPUs on
    the box the slabs share devices (r mod the number present): same kernels, flags and schedule."""
    deck = _deck(case, nmax, nout, exit_time)
    Jr, rms_r, _ = _run(REF, deck, case, str(tmp_path / "ref"))
    J1, rms_1, _ = _run(GPU, deck, case, str(tmp_path / "gpu1"))
    Jn, rms_n, out = _run(GPU, deck, case, str(tmp_path / "gpun"), env={"HF2D_GPUS": str(n_gpus), "CUDA_DEVICE_MAX_CONNECTIONS": "32"})
    assert f"hf2d_gpu: {n_gpus} X-slabs" in out
    assert len(rms_n) == len(rms_r) == len(rms_1)
    for a, b in zip(rms_n, rms_r):
        assert np.allclose(a, b, rtol=2e-5, atol=1e-300), (a, b)
    for f in Jn.dtype.names:            # every member of every record, not only the step's
        a, b = Jn[f], J1[f]
        same = (a == b) | ((a != a) & (b != b)) if a.dtype.kind == "f" else (a == b)
        assert same.all(), f"{case} x{n_gpus}: member {f} differs from the single-GPU program in {int((~same).sum())} entries"
    for f in FIELDS:
        a, b = Jn[f], Jr[f]
        if exact or a.dtype.kind in "iu":
            assert ((a == b) | ((a != a) & (b != b))).all(), f
        else:
            assert float(np.abs(a - b).max()) / max(float(np.abs(b).max()), 1e-300) <= 1e-10, f

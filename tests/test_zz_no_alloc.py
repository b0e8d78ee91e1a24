"""pdec_burgers_pair_loss_static (csrc/noalloc.cu): the allocation-free entry point, checked through the C-ABI against the existing
pdec_burgers_pair_loss on the same inputs, two calls in a row (the static scratch keeps the first call's contents).

Collected last on purpose (file name): a failure here must not hide any other test.
The entry point has no counterpart in the reference."""
import os

import numpy as np
import pytest
import torch

from pde_control_b200 import _lib

DT, NU = 0.03, 0.003
LOSS_TOL, GRAD_TOL = 1e-4, 1e-3      # the tolerances of tests/test_burgers_gpu.py (losses, control-loss gradients)


def _frames(T, B, n, seed):
    rng = np.random.RandomState(seed)
    x = (np.arange(n, dtype=np.float32) + 0.5) / n
    loc = rng.uniform(0.2, 0.8, size=(T + 1, B, 1)).astype(np.float32)
    amp = rng.uniform(-2.0, 2.0, size=(T + 1, B, 1)).astype(np.float32)
    sig = rng.uniform(0.05, 0.2, size=(T + 1, B, 1)).astype(np.float32)
    return (amp * np.exp(-0.5 * (x - loc) ** 2 / sig ** 2)).astype(np.float32)


def _existing(s, mode, gl2, gl1):
    L = _lib.lib()
    T1, B, n = s.shape
    nbytes = L.pdec_burgers_pair_loss_workspace_bytes(B, n, T1 - 1)
    ws = torch.empty((nbytes + 3) // 4, device=s.device, dtype=torch.float32)
    g = torch.full_like(s, float("nan"))
    losses = torch.full((2,), float("nan"), device=s.device)
    rc = L.pdec_burgers_pair_loss(_lib.ptr(s), _lib.ptr(g), _lib.ptr(losses), T1 - 1, B, n, mode, DT, NU, 0.0, 1.0, 1, 0, gl2, gl1,
                                  _lib.ptr(ws), nbytes, _lib.stream_ptr())
    _lib.check(rc, "pdec_burgers_pair_loss")
    torch.cuda.synchronize()
    return g, losses


def _static(s, g, losses, mode, gl2, gl1):
    T1, B, n = s.shape
    return _lib.lib().pdec_burgers_pair_loss_static(_lib.ptr(s), _lib.ptr(g), _lib.ptr(losses), T1 - 1, B, n, mode, DT, NU, 0.0, 1.0,
                                                    1, 0, gl2, gl1, _lib.stream_ptr())


def _rel(a, b):
    return float((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30))


@pytest.mark.gpu
def test_static_scratch_entry_matches_existing_entry_twice_in_a_row():
    dev = torch.device("cuda:0")
    with torch.cuda.device(dev):
        # two calls in a row with DIFFERENT shapes and data: a large one first (leaves 5120 partials and a used counter behind), then a
        # small one; then the fused unit-box kernel again, and a row length that takes the general kernel + separate reduction
        cases = [(32, 160, 128, _lib.MODE_CONSTANT, 1.0, 0.0, 1), (5, 24, 128, _lib.MODE_CONSTANT, 1.0, 0.5, 2),
                 (7, 33, 128, _lib.MODE_CIRCULAR, 0.7, 0.0, 3), (4, 19, 100, _lib.MODE_REPLICATE, 1.0, 1.0, 4)]
        inputs = [torch.tensor(_frames(T, B, n, seed), device=dev) for (T, B, n, _, _, _, seed) in cases]
        expected = [_existing(s, c[3], c[4], c[5]) for s, c in zip(inputs, cases)]
        outs = [(torch.full_like(s, float("nan")), torch.full((2,), float("nan"), device=dev)) for s in inputs]
        torch.cuda.synchronize()
        allocated, reserved = torch.cuda.memory_allocated(dev), torch.cuda.memory_reserved(dev)
        free_before = None
        for s, c, (g, losses) in zip(inputs, cases, outs):
            rc = _static(s, g, losses, c[3], c[4], c[5])
            assert rc == 0, _lib.lib().pdec_error_string(rc)
            if free_before is None:      # the first call on a device looks the array's address up (set-up; the driver may load the module's data then)
                torch.cuda.synchronize()
                free_before = torch.cuda.mem_get_info(dev)[0]
        torch.cuda.synchronize()
        # nothing was allocated by the calls: torch's allocator did not move and the device's free memory did not shrink
        assert torch.cuda.memory_allocated(dev) == allocated and torch.cuda.memory_reserved(dev) == reserved
        assert torch.cuda.mem_get_info(dev)[0] >= free_before
        for (g, losses), (g_ref, losses_ref), c in zip(outs, expected, cases):
            assert torch.isfinite(g).all() and torch.isfinite(losses).all()
            assert abs(losses[0].item() - losses_ref[0].item()) <= LOSS_TOL * abs(losses_ref[0].item()), c
            assert abs(losses[1].item() - losses_ref[1].item()) <= LOSS_TOL * abs(losses_ref[1].item()), c
            assert _rel(g, g_ref) <= GRAD_TOL, c
            # same kernels, same launch geometry, deterministic reductions: in fact bit-identical
            assert torch.equal(g, g_ref) and torch.equal(losses, losses_ref), c


@pytest.mark.gpu
def test_static_scratch_entry_rejects_sizes_beyond_its_maximum():
    dev = torch.device("cuda:0")
    L = _lib.lib()
    max_pairs = _lib.STATIC_MAX_PAIRS
    header = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "pdec.h")).read()
    assert "#define PDEC_STATIC_MAX_PAIRS %d\n" % max_pairs in header
    with torch.cuda.device(dev):
        # the limit itself is accepted: B * T == max_pairs at a short row length (keeps the arrays small: 33 * 4096 * 8 floats)
        T, B, n = 32, max_pairs // 32, 8
        s = torch.tensor(_frames(T, B, n, 5), device=dev)
        g, losses = torch.full_like(s, float("nan")), torch.full((2,), float("nan"), device=dev)
        assert _static(s, g, losses, _lib.MODE_CONSTANT, 1.0, 0.0) == 0
        g_ref, losses_ref = _existing(s, _lib.MODE_CONSTANT, 1.0, 0.0)
        torch.cuda.synchronize()
        assert torch.equal(g, g_ref) and torch.equal(losses, losses_ref)
        # one row more: refused before anything is enqueued, outputs untouched, with a message that names the limit
        s2 = torch.zeros(T + 1, B + 1, n, device=dev)
        g2, losses2 = torch.full_like(s2, 7.0), torch.full((2,), 7.0, device=dev)
        rc = _static(s2, g2, losses2, _lib.MODE_CONSTANT, 1.0, 0.0)
        torch.cuda.synchronize()
        assert rc == -7
        assert b"PDEC_STATIC_MAX_PAIRS" in L.pdec_error_string(rc)
        assert bool((g2 == 7.0).all()) and bool((losses2 == 7.0).all())
        with pytest.raises(_lib.PdecError):
            _lib.check(rc, "pdec_burgers_pair_loss_static")

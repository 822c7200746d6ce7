"""spanon_fcae_fwd_static: the allocation-free variant of spanon_fcae_fwd (its pooling scratch is a static device
array inside the library, include/spanon.h).  Called through the C ABI and checked against the workspace entry point
on two calls in a row (the shared array keeps the previous call's contents, so the second call proves it is cleared).
Collected last on purpose; passes alone in a fresh process and after the other tests.

Tolerance: the tests' FC-AE bar (1e-4 rel. on the reconstruction, 1e-4 abs. on the log-probabilities; the two entry
points run the same kernels, only the order of the fp64 pooling atomics differs)."""
import ctypes as C

import numpy as np
import pytest
import torch

import recipe
from conftest import rel_max

import speech_anonymization_b200 as sab
from speech_anonymization_b200 import _lib

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _run(fn_name, m, x, noise, with_ws):
    L = _lib.lib()
    B, T, _ = x.shape
    recon = torch.full((B, T, 80), float("nan"), dtype=torch.float32, device=DEV)
    logp = torch.full((B, 2), float("nan"), dtype=torch.float32, device=DEV)
    w = m.weight_struct()
    st = _lib.stream_ptr(torch.device(DEV))
    if with_ws:
        nbytes = L.spanon_fcae_workspace_bytes(B, T)
        ws = torch.empty((nbytes,), dtype=torch.uint8, device=DEV)
        rc = L.spanon_fcae_fwd(_lib.ptr(x), C.byref(w), _lib.ptr(noise), None, None, None, 0.0, _lib.ptr(recon),
                               _lib.ptr(logp), _lib.ptr(ws), nbytes, B, T, st)
    else:
        rc = L.spanon_fcae_fwd_static(_lib.ptr(x), C.byref(w), _lib.ptr(noise), None, None, None, 0.0, _lib.ptr(recon),
                                      _lib.ptr(logp), B, T, st)
    _lib.check(rc, fn_name)
    torch.cuda.synchronize()
    return recon.cpu().numpy(), logp.cpu().numpy()


def test_fcae_fwd_static_matches_workspace_entry_point():
    L = _lib.lib()
    m = sab.FullyConnectedAutoencoder(80, 8)
    wts = recipe.make_weights(recipe.FCAE_SHAPES, 8886)
    m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in wts.items()})
    m = m.to(DEV).eval()
    assert L.spanon_fcae_static_max_batch() == 256
    # two calls in a row on different inputs and batch sizes (the second batch is smaller: stale accumulators of the
    # first call stay in the array beyond it, and the ones it reuses must have been cleared)
    for seed, B, T in ((91, 8, 101), (92, 3, 257)):
        x = torch.from_numpy(recipe.make_feats(seed, B, T)).to(DEV)
        noise = torch.from_numpy(recipe.make_noise(seed, B, 40)).to(DEV)
        r_ws, l_ws = _run("fcae_fwd", m, x, noise, True)
        r_st, l_st = _run("fcae_fwd_static", m, x, noise, False)
        assert np.isfinite(r_st).all() and np.isfinite(l_st).all()
        assert rel_max(r_st, r_ws) < 1e-4
        assert np.max(np.abs(l_st - l_ws)) < 1e-4
    # same input twice in a row through the static entry point: identical within the same tolerance
    r_a, l_a = _run("fcae_fwd_static", m, x, noise, False)
    r_b, l_b = _run("fcae_fwd_static", m, x, noise, False)
    assert rel_max(r_b, r_a) < 1e-4 and np.max(np.abs(l_b - l_a)) < 1e-4
    # sizes beyond the built-in scratch are rejected with a clear error, before anything is launched
    w = m.weight_struct()
    rc = L.spanon_fcae_fwd_static(None, C.byref(w), None, None, None, None, 0.0, None, None, 257, 10, None)
    assert rc == -8
    assert "static scratch" in L.spanon_strerror(rc).decode()

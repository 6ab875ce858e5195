"""fm_score_noalloc: the scorer for hosts that cannot give the library a workspace.

Some callers of this library are not allowed to allocate GPU memory on our behalf: an
inference server whose framework took the whole device for its own memory pool at
start-up (so a cudaMalloc issued from a plugin fails or breaks the host's rules), or a
scoring thread with a latency budget that cannot afford the device-wide lock and the
implicit synchronisation that cudaMalloc / cudaFree carry.  For them the per-block scratch
of the scorer lives in a fixed device array inside libfm_b200.so, shared by every call.

This file is named so that pytest collects it last.  It calls the new entry point through
the C ABI and holds it to the ordinary fm_score on the same inputs, twice in a row -- the
second call runs on a smaller batch over the first call's left-overs in the shared array.
"""
import numpy as np
import pytest
import torch

from fast_tffm_b200 import _lib
from oracle import oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
RTOL = 1e-5          # the tolerance of tests/test_gpu_parity.py
MAX_BATCH = 1048576  # FM_SCORE_NOALLOC_MAX_BATCH in include/fm_b200.h


def _case(seed, batch, k, rows, max_nnz, ld=None):
    """ld: row stride in floats (default K+1: rows are not 16-byte aligned, the scalar kernels run;
    a multiple of 4 >= K+1 with zero pad columns takes the float4 kernels, like the model's tables)"""
    rng = np.random.default_rng(seed)
    sizes = rng.integers(0, max_nnz + 1, batch)
    poses = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    n = int(poses[-1])
    t = lambda a, dt: torch.as_tensor(a, dtype=dt).to(DEV)
    params = np.zeros((rows, ld or k + 1), np.float32)
    params[:, :k + 1] = rng.uniform(-0.3, 0.3, (rows, k + 1))
    return dict(
        batch=batch, k=k,
        fids=t(rng.integers(0, rows, n), torch.int32), vals=t(rng.normal(size=n), torch.float32),
        poses=t(poses, torch.int32), params=t(params, torch.float32),
        pred_ws=torch.empty(batch, device=DEV), reg_ws=torch.empty((), device=DEV),
        pred_na=torch.empty(batch, device=DEV), reg_na=torch.empty((), device=DEV))


def _score_with_workspace(lib, c, lf, lb):
    ws = _lib.workspace(lib.fm_score_workspace(c["batch"]), torch.device(DEV))
    _lib.check(lib.fm_score(_lib.ptr(c["fids"]), _lib.ptr(c["params"]), c["params"].stride(0), _lib.ptr(c["vals"]),
                            _lib.ptr(c["poses"]), c["batch"], c["k"], lf, lb, _lib.ptr(c["pred_ws"]),
                            _lib.ptr(c["reg_ws"]), _lib.ptr(ws), ws.numel(), _lib.stream()), "fm_score")
    torch.cuda.synchronize()


def _score_noalloc(lib, c, lf, lb):
    _lib.check(lib.fm_score_noalloc(_lib.ptr(c["fids"]), _lib.ptr(c["params"]), c["params"].stride(0),
                                    _lib.ptr(c["vals"]), _lib.ptr(c["poses"]), c["batch"], c["k"], lf, lb,
                                    _lib.ptr(c["pred_na"]), _lib.ptr(c["reg_na"]), _lib.stream()), "fm_score_noalloc")
    torch.cuda.synchronize()


def _assert_same(c):
    np.testing.assert_allclose(c["pred_na"].cpu().numpy(), c["pred_ws"].cpu().numpy(), rtol=RTOL, atol=1e-6)
    np.testing.assert_allclose(c["reg_na"].item(), c["reg_ws"].item(), rtol=RTOL)
    assert np.isfinite(c["reg_na"].item())


def _assert_matches_oracle(c, lf, lb):
    """the arithmetic itself, not only the scratch plumbing: the CPU oracle's scorer on the same inputs, with the
    absolute floor of tests/test_gpu_parity.py (the second-order term cancels, so the error scales with the terms)"""
    fids, params = c["fids"].cpu().numpy(), c["params"].cpu().numpy()
    vals, poses, k = c["vals"].cpu().numpy(), c["poses"].cpu().numpy(), c["k"]
    e_pred, e_reg = O.score(fids, params, vals, poses, lf, lb, k=k)
    r = np.abs(params[fids])
    terms = np.abs(vals) * r[:, 0]
    fac = np.abs(vals)[:, None] * r[:, 1:k + 1]
    scale = np.zeros(c["batch"])
    for i in range(c["batch"]):
        sl = slice(poses[i], poses[i + 1])
        scale[i] = terms[sl].sum() + 0.5 * (fac[sl].sum(0) ** 2).sum()
    got = c["pred_na"].cpu().numpy().astype(np.float64)
    assert (np.abs(got - e_pred) <= RTOL * (np.abs(e_pred) + scale)).all()
    np.testing.assert_allclose(c["reg_na"].item(), e_reg, rtol=RTOL)


def test_noalloc_scorer_matches_fm_score_on_two_calls_in_a_row():
    lib = _lib.load()
    torch.cuda.set_device(0)
    big = _case(1, 20000, 16, 5000, 45)      # 2500 blocks of scratch
    small = _case(2, 777, 8, 300, 12)        # 98 blocks: reads only what this launch wrote
    for c in (big, small):
        _score_with_workspace(lib, c, 0.02, 0.3)
    # first call on this device: also resolves the scratch address
    _score_noalloc(lib, big, 0.02, 0.3)
    _assert_same(big)
    # second call, straight after, on the smaller batch; no device memory may appear or vanish
    free_before, _ = torch.cuda.mem_get_info()
    _score_noalloc(lib, small, 0.02, 0.3)
    free_after, _ = torch.cuda.mem_get_info()
    _assert_same(small)
    assert free_after == free_before
    # and the big one again: results do not depend on what the array held before
    big["pred_na"].zero_()
    _score_noalloc(lib, big, 0.02, 0.3)
    _assert_same(big)
    _assert_matches_oracle(big, 0.02, 0.3)
    _score_noalloc(lib, small, 0.02, 0.3)
    _assert_matches_oracle(small, 0.02, 0.3)


def test_noalloc_scorer_on_16_byte_aligned_rows():
    """tables as the model lays them out (row stride a multiple of 4 floats, zero pad columns): both entry points
    take the float4 kernels there; again two calls in a row, against fm_score and against the oracle"""
    lib = _lib.load()
    torch.cuda.set_device(0)
    a = _case(5, 9000, 16, 4000, 45, ld=32)
    b = _case(6, 501, 64, 700, 20, ld=68)
    for c in (a, b):
        _score_with_workspace(lib, c, 0.02, 0.3)
        _score_noalloc(lib, c, 0.02, 0.3)
        _assert_same(c)
        _assert_matches_oracle(c, 0.02, 0.3)


def test_noalloc_scorer_refuses_batches_beyond_its_scratch():
    lib = _lib.load()
    c = _case(3, 4, 4, 10, 3)
    rc = lib.fm_score_noalloc(_lib.ptr(c["fids"]), _lib.ptr(c["params"]), c["params"].stride(0), _lib.ptr(c["vals"]),
                              _lib.ptr(c["poses"]), MAX_BATCH + 1, c["k"], 0.0, 0.0, _lib.ptr(c["pred_na"]),
                              _lib.ptr(c["reg_na"]), _lib.stream())
    assert rc == -4
    msg = lib.fm_last_error().decode()
    assert "FM_SCORE_NOALLOC_MAX_BATCH" in msg and "fm_score" in msg
    torch.cuda.synchronize()   # nothing was launched, nothing faulted

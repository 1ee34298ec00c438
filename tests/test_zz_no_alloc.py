"""GPU: spif_sparse_ffn_noalloc -- the single-token fused sparse FFN whose scratch is a static __device__ array inside the
library (no reference counterpart; added for callers that cannot allocate).

Through the C-ABI (ctypes): its output against the existing entry point spif_sparse_ffn (caller-owned workspace) on the same
inputs, within the tolerance of the other F16 parity tests (1e-3 relative; in fact the bits are identical), on TWO calls in a
row with different inputs -- the second call finds the first one's contents in the shared array.  Also: the size limit is a
clear error, not an overrun.  Collected last (file name) so that a failure here hides no other test under `pytest -x`;
passes alone in a fresh process and after the rest of the suite."""
import ctypes

import numpy as np
import pytest

import synth
from oracle_lib import ARCH_PROSPARSE, F16, Q8_0

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

TOL = 1e-3  # tests/test_gpu_parity.py TOL_DEFAULT[F16]


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def rel_err(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def _args(L, wtype, up, gate, down, x, p, y, n_embd, n_ff, stats=None):
    from sparkinfer_b200._lib import FfnArgs
    ptr = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None
    return FfnArgs(arch=ARCH_PROSPARSE, type=wtype, n_embd=n_embd, n_ff=n_ff, m=n_ff, n_tok=1, flags=0, thr=0.5,
                   w_up=ptr(up), w_gate=ptr(gate), w_down_t=ptr(down), b_up=None, b_gate=None, b_down=None, x=ptr(x),
                   sparse_idx=ptr(p), neuron_idx=None, y=ptr(y), hidden_out=None, stats_out=ptr(stats))


@pytest.mark.parametrize("wtype,n_embd,n_ff", [(F16, 1024, 4096), (F16, 5120, 13824), (Q8_0, 1024, 2048)])
def test_noalloc_matches_workspace_entry_point_twice(oracle, wtype, n_embd, n_ff):
    import sparkinfer_b200 as sp
    from sparkinfer_b200 import _lib
    L = _lib.load()
    assert L.spif_noalloc_max_bytes() == 12 << 20
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    seed = synth.SEED0 + 3 * n_embd + n_ff + wtype
    Wu, Wg, Wd = synth.ffn_weights(n_embd, n_ff, seed)
    du, dg, dd = (dev(oracle.quantize(wtype, w)) for w in (Wu, Wg, Wd))
    ws = sp.Workspace(n_embd, n_ff, 1)
    for call in range(2):  # two calls in a row: different x and mask, the static array keeps the first call's contents
        x = dev(synth.activations(n_embd, 1, seed + 10 + call)[0])
        p_host = synth.direct_mask(n_ff, 1, 0.10 + 0.05 * call, seed + 20 + call)
        p = dev(p_host[0])
        y_ws = torch.full((n_embd,), float("nan"), device="cuda")
        y_na = torch.full((n_embd,), float("nan"), device="cuda")
        st_ws = torch.zeros((4,), dtype=torch.int32, device="cuda")
        st_na = torch.zeros((4,), dtype=torch.int32, device="cuda")
        a_ws = _args(L, wtype, du, dg, dd, x, p, y_ws, n_embd, n_ff, st_ws)
        a_na = _args(L, wtype, du, dg, dd, x, p, y_na, n_embd, n_ff, st_na)
        _lib.check(L.spif_sparse_ffn(ctypes.byref(a_ws), ws.ptr, ws.nbytes, stream), "spif_sparse_ffn")
        _lib.check(L.spif_sparse_ffn_noalloc(ctypes.byref(a_na), stream), "spif_sparse_ffn_noalloc")
        torch.cuda.synchronize()
        got, want = y_na.cpu().numpy(), y_ws.cpu().numpy()
        assert np.all(np.isfinite(got)) and np.all(np.isfinite(want))
        assert rel_err(got, want) < TOL, (call, rel_err(got, want))
        assert np.array_equal(got, want), "same kernels, same fixed summation order: the bits must agree"
        assert np.array_equal(st_na.cpu().numpy(), st_ws.cpu().numpy())
        assert int(st_na[0]) == int((~(p_host[0] < 0.5)).sum())
        if wtype == F16:  # and both against the oracle, like every other parity test
            y_ref = oracle.sparse_ffn(ARCH_PROSPARSE, wtype, oracle.quantize(wtype, Wu), oracle.quantize(wtype, Wg),
                                      oracle.quantize(wtype, Wd), x.cpu().numpy()[None], p_host, n_embd, n_ff)[0]
            assert rel_err(got, y_ref[0]) < TOL


def test_noalloc_rejects_what_does_not_fit():
    """beyond the static array: SPIF_E_STATIC_LIMIT with a message, nothing launched; n_tok > 1: SPIF_E_UNSUPPORTED"""
    from sparkinfer_b200 import _lib
    L = _lib.load()
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    n_embd, n_ff = 16384, 65536  # 320 partial rows of 64 KB = 20 MiB > 12 MiB
    assert L.spif_workspace_size(n_embd, n_ff, 1) > L.spif_noalloc_max_bytes()
    t = torch.zeros(16, device="cuda")
    a = _args(L, F16, t, t, t, t, t, t, n_embd, n_ff)
    rc = L.spif_sparse_ffn_noalloc(ctypes.byref(a), stream)
    assert rc == -7 and b"static array" in L.spif_error_string(rc)
    a = _args(L, F16, t, t, t, t, t, t, 1024, 4096)
    a.n_tok = 2
    assert L.spif_sparse_ffn_noalloc(ctypes.byref(a), stream) == -5
    assert L.spif_sparse_ffn_noalloc(None, stream) == -3
    torch.cuda.synchronize()

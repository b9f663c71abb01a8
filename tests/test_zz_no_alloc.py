"""gpe_mlp_jet_fwd_noalloc (include/gpe_b200.h): the forward with its scratch in a static device array of the library, for
callers that cannot allocate.  Collected last on purpose (file name).

Called through the C ABI with plain ctypes and torch device pointers, twice in a row, against gpe_mlp_jet_fwd with a caller
workspace: the two entry points run the same kernels on the same packed weights, so the outputs are held bit-equal (tighter
than the 1e-5 the parity tests use).  Works alone in a fresh process and after the other tests."""
import ctypes

import pytest
import torch

from test_jet_kernels import make_net

pytestmark = pytest.mark.gpu


def _fwd_ref(lib, _lib, desc, Wp, bp, x, n, d):
    fwd, bwd = ctypes.c_size_t(), ctypes.c_size_t()
    assert lib.gpe_workspace_bytes(ctypes.byref(desc), n, ctypes.byref(fwd), ctypes.byref(bwd)) == 0
    ws = torch.empty(max(fwd.value, 16), dtype=torch.uint8, device="cuda")
    u, g, h = torch.empty(n, device="cuda"), torch.empty(n, d, device="cuda"), torch.empty(n, 1, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    rc = lib.gpe_mlp_jet_fwd(ctypes.byref(desc), Wp, bp, x.data_ptr(), n, None, u.data_ptr(), g.data_ptr(), h.data_ptr(),
                             ws.data_ptr(), fwd.value, ctypes.c_void_p(stream))
    assert rc == 0, lib.gpe_last_error()
    torch.cuda.synchronize()
    return u, g, h


@pytest.mark.parametrize("d,W,L,prec", [(2, 128, 4, 2), (1, 64, 3, 0)])   # tensor-core path, FFMA path
def test_noalloc_forward_matches_the_workspace_forward_twice(d, W, L, prec):
    from gpe_b200 import _lib
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for f in ("gpe_workspace_bytes", "gpe_mlp_jet_fwd", "gpe_mlp_jet_fwd_noalloc"):
        getattr(lib, f).argtypes = getattr(_lib.lib, f).argtypes
        getattr(lib, f).restype = ctypes.c_int
    lib.gpe_last_error.restype = ctypes.c_char_p
    jet = _lib.JET_LAPLACIAN if d == 2 else _lib.JET_DIAG
    desc = _lib.MlpDesc(d, L, W, _lib.ACT_TANH, jet, prec)
    stream = torch.cuda.current_stream().cuda_stream
    # two calls in a row with DIFFERENT weights and sizes: the static array keeps the first call's contents, the second call
    # must not see them
    for seed, n in ((11, 5003), (12, 20011)):
        Ws, bs = make_net(d, W, L, seed=seed, act=0)
        Ws, bs = [w.cuda() for w in Ws], [b.cuda() for b in bs]
        g0 = torch.Generator(device="cuda").manual_seed(seed)
        x = torch.rand(n, d, device="cuda", generator=g0) * 4 - 2
        Wp, keep1 = _lib.ptr_array([w.data_ptr() for w in Ws])
        bp, keep2 = _lib.ptr_array([b.data_ptr() for b in bs])
        ur, gr, hr = _fwd_ref(lib, _lib, desc, Wp, bp, x, n, d)
        u, g, h = torch.full((n,), float("nan"), device="cuda"), torch.empty(n, d, device="cuda"), torch.empty(n, 1, device="cuda")
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        rc = lib.gpe_mlp_jet_fwd_noalloc(ctypes.byref(desc), Wp, bp, x.data_ptr(), n, None, u.data_ptr(), g.data_ptr(), h.data_ptr(),
                                         ctypes.c_void_p(stream))
        assert rc == 0, lib.gpe_last_error()
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        assert bool(torch.isfinite(u).all())
        assert torch.equal(u, ur) and torch.equal(g, gr) and torch.equal(h, hr)


def test_noalloc_rejects_sizes_beyond_the_maximum():
    from gpe_b200 import _lib
    lib = ctypes.CDLL(_lib.LIB_PATH)
    lib.gpe_mlp_jet_fwd_noalloc.argtypes = _lib.lib.gpe_mlp_jet_fwd_noalloc.argtypes
    lib.gpe_mlp_jet_fwd_noalloc.restype = ctypes.c_int
    lib.gpe_last_error.restype = ctypes.c_char_p
    d, W, L = 2, 128, 4
    Ws, bs = make_net(d, W, L, seed=1, act=0)
    Ws, bs = [w.cuda() for w in Ws], [b.cuda() for b in bs]
    Wp, keep1 = _lib.ptr_array([w.data_ptr() for w in Ws])
    bp, keep2 = _lib.ptr_array([b.data_ptr() for b in bs])
    desc = _lib.MlpDesc(d, L, W, _lib.ACT_TANH, _lib.JET_LAPLACIAN, 2)
    x = torch.zeros(8, d, device="cuda")
    u, g, h = torch.zeros(8, device="cuda"), torch.zeros(8, d, device="cuda"), torch.zeros(8, 1, device="cuda")
    n_big = (1 << 24) + 1                      # GPE_STATIC_MAX_POINTS + 1: refused before any pointer is touched
    rc = lib.gpe_mlp_jet_fwd_noalloc(ctypes.byref(desc), Wp, bp, x.data_ptr(), n_big, None, u.data_ptr(), g.data_ptr(), h.data_ptr(), None)
    assert rc != 0 and b"GPE_STATIC_MAX_POINTS" in lib.gpe_last_error()
    torch.cuda.synchronize()

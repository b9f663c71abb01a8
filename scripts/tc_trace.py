"""Debug probe: clock64 stamps of the MMA thread / epilogue warp 0 of cluster 0 in the two-tile forward kernel (needs the
trace build: libgpe_b200_trace.so built with -DGPE_TC_TRACE; loaded through raw ctypes, no torch shim)."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "gross-pitaevskii-eigenvalue-problem_b200"))
import torch
from gpe_b200 import _lib
from test_jet_kernels import make_net
lib = ctypes.CDLL(os.path.join(ROOT, "gross-pitaevskii-eigenvalue-problem_b200", "gpe_b200", "libgpe_b200_trace.so"), mode=os.RTLD_NOW | os.RTLD_DEEPBIND)   # its own copies of the launchers, not the ones libgpe_b200.so exported globally
lib.gpe_workspace_bytes.argtypes = _lib.lib.gpe_workspace_bytes.argtypes
lib.gpe_mlp_jet_fwd.argtypes = _lib.lib.gpe_mlp_jet_fwd.argtypes
n, d, W, L, jet = 148 * 32 * 6, 2, 128, 4, 3
Ws, bs = make_net(d, W, L, seed=1, act=0)
Ws = [w.cuda() for w in Ws]; bs = [b.cuda() for b in bs]
x = torch.rand(n, d, device="cuda") * 4 - 2
desc = _lib.MlpDesc(d, L, W, 0, jet, 2)
fwd, bwd = ctypes.c_size_t(), ctypes.c_size_t()
lib.gpe_workspace_bytes(ctypes.byref(desc), n, ctypes.byref(fwd), ctypes.byref(bwd))
ws = torch.empty(fwd.value, dtype=torch.uint8, device="cuda")
u = torch.empty(n, device="cuda"); g = torch.empty(n, d, device="cuda"); h = torch.empty(n, 1, device="cuda")
Wp, _k1 = _lib.ptr_array([w.data_ptr() for w in Ws]); bp, _k2 = _lib.ptr_array([b.data_ptr() for b in bs])
for _ in range(3):
    rc = lib.gpe_mlp_jet_fwd(ctypes.byref(desc), Wp, bp, x.data_ptr(), n, None, u.data_ptr(), g.data_ptr(), h.data_ptr(), ws.data_ptr(), fwd.value, None)
    torch.cuda.synchronize()
assert rc == 0
buf = (ctypes.c_longlong * 4096)()
rc2 = lib.gpe_debug_trace(buf, 4096); print('memcpy rc', rc2, 'u finite', bool(torch.isfinite(u).all()), float(u.abs().mean()), 'nonzero stamps', sum(1 for v in buf if v))
t0 = min(v for v in buf if v) if any(buf) else 0
print("chunk  mma_woke  mma_committed  epi_got_slot  epi_arrived   (cycles since chunk 0's slot)")
for i in range(40):
    print(i, *[buf[4 * i + k] - t0 for k in range(4)])

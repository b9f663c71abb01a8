"""ctypes binding of libgpe_b200.so (C ABI declared in include/gpe_b200.h).

The shared library is the product: there is NO CPU or PyTorch fallback.  Importing this module
without the built library raises, and every compute entry point fails loudly without a CUDA
device (the library returns a negative code, turned into RuntimeError here).
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgpe_b200.so")

ACT_TANH, ACT_SHIFTED_TANH, ACT_LEAKY_RELU, ACT_RESBLOCK_TANH = 0, 1, 2, 3
JET_VALUE, JET_DIAG, JET_FULL, JET_LAPLACIAN = 0, 1, 2, 3
MAX_LAYERS = 8
NFEAT, NMOM, MOM_GRAD2, MOM_NN2, MOM_SYM2 = 7, 32, 28, 29, 30
F_ONE, F_U, F_LAP, F_V, F_VU, F_NL, F_Q = range(7)
NL_NONE, NL_POW, NL_ABSPOW = 0, 1, 2


def mom(i, j):
    """Index of sum_n f_i f_j in the moment vector (row-major upper triangle of the 7x7 Gram matrix)."""
    a, b = (i, j) if i <= j else (j, i)
    return a * NFEAT - a * (a - 1) // 2 + (b - a)


class MlpDesc(ctypes.Structure):
    """gpe_mlp_desc (include/gpe_b200.h)."""
    _fields_ = [("d", ctypes.c_int), ("n_hidden", ctypes.c_int), ("width", ctypes.c_int),
                ("act", ctypes.c_int), ("jet", ctypes.c_int), ("precision", ctypes.c_int)]


class LossDesc(ctypes.Structure):
    """gpe_loss_desc (include/gpe_b200.h)."""
    _fields_ = [("d", ctypes.c_int), ("nl_kind", ctypes.c_int), ("nl_power", ctypes.c_float), ("scale", ctypes.c_float),
                ("bfactor", ctypes.c_int), ("sym_sign", ctypes.c_int), ("h_cols", ctypes.c_int), ("lap_cols", ctypes.c_int)]


class XForm(ctypes.Structure):
    """gpe_xform: x_eff = a*x + b per coordinate."""
    _fields_ = [("a", ctypes.c_float * 2), ("b", ctypes.c_float * 2)]


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `make -C gross-pitaevskii-eigenvalue-problem_b200/csrc` "
            "(or `python -c 'import __graft_entry__ as g; g.build()'`).  gpe_b200 has no CPU fallback.")
    return ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)


lib = _load()

_vp, _i64, _sz = ctypes.c_void_p, ctypes.c_int64, ctypes.c_size_t
_pp = ctypes.POINTER(ctypes.c_void_p)
_desc_p = ctypes.POINTER(MlpDesc)
_xf_p = ctypes.POINTER(XForm)

lib.gpe_version.restype = ctypes.c_char_p
lib.gpe_last_error.restype = ctypes.c_char_p
lib.gpe_workspace_bytes.argtypes = [_desc_p, _i64, ctypes.POINTER(_sz), ctypes.POINTER(_sz)]
lib.gpe_mlp_jet_fwd.argtypes = [_desc_p, _pp, _pp, _vp, _i64, _xf_p, _vp, _vp, _vp, _vp, _sz, _vp]
lib.gpe_mlp_jet_bwd.argtypes = [_desc_p, _pp, _pp, _vp, _i64, _xf_p, _vp, _vp, _vp, _pp, _pp, _vp, _sz, _vp]
lib.gpe_mlp_jet_fwd_noalloc.argtypes = [_desc_p, _pp, _pp, _vp, _i64, _xf_p, _vp, _vp, _vp, _vp]
_loss_p = ctypes.POINTER(LossDesc)
lib.gpe_gram_moments.argtypes = [_loss_p, _i64] + [_vp] * 8 + [_vp, _vp]
lib.gpe_gram_adjoints.argtypes = [_loss_p, _i64] + [_vp] * 8 + [_vp] + [_vp] * 4 + [_vp]
lib.gpe_ffma_probe.argtypes = [ctypes.c_int, ctypes.c_int, _vp, ctypes.POINTER(ctypes.c_double), _vp]
EXPORTS = ("gpe_version", "gpe_last_error", "gpe_workspace_bytes", "gpe_mlp_jet_fwd", "gpe_mlp_jet_bwd",
           "gpe_gram_moments", "gpe_gram_adjoints", "gpe_ffma_probe", "gpe_set_tensor_path", "gpe_tc_profile",
           "gpe_tc_timings", "gpe_stash_bytes", "gpe_mlp_jet_fwd_stash", "gpe_mlp_jet_bwd_stashed", "gpe_scalar_program",
           "gpe_mlp_jet_fwd_noalloc", "gpe_stage_bookkeeping")
for _f in EXPORTS[2:]:
    getattr(lib, _f).restype = ctypes.c_int


def check(rc, what):
    if rc != 0:
        raise RuntimeError(f"{what} failed (code {rc}): {lib.gpe_last_error().decode()}")


def ptr_array(ptrs):
    arr = (ctypes.c_void_p * len(ptrs))(*ptrs)
    return ctypes.cast(arr, _pp), arr   # keep `arr` alive for the duration of the call


def version():
    return lib.gpe_version().decode()

"""phantom-fhe-boot_b200 — B200-native RNS polynomial hot path for CKKS (NTT, hybrid key switching, HMult,
rotation, rescale), exposed through the C ABI in include/b200fhe.h (libb200fhe.so, hand-written sm_100a CUDA).

This Python module is only plumbing for tests and bench.py: ctypes over the C ABI, torch for device memory and
streams.  It never imports anything from oracle/ and has no CPU fallback: compute calls raise if the CUDA
library or a GPU is missing.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libb200fhe.so")
_u64p = C.POINTER(C.c_uint64)
_u32p = C.POINTER(C.c_uint32)
_lib = None


class B200FheError(RuntimeError):
    pass


def build(verbose=False):
    """Compile the CUDA library in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    subprocess.check_call(["make", "-C", os.path.join(_HERE, "csrc"), "-j8"] + ([] if verbose else ["-s"]))
    return _SO


_SIG = {
    "b200fhe_coeff_modulus_create": [C.c_size_t, C.POINTER(C.c_int), C.c_size_t, _u64p],
    "b200fhe_ctx_create": [C.c_int, _u64p, C.c_size_t, C.c_size_t, C.c_int, C.POINTER(C.c_void_p)],
    "b200fhe_ctx_create_host": [C.c_int, _u64p, C.c_size_t, C.c_size_t, C.POINTER(C.c_void_p)],
    "b200fhe_ctx_destroy": [C.c_void_p],
    "b200fhe_ctx_export_table": [C.c_void_p, C.c_int, _u64p],
    "b200fhe_level_export": [C.c_void_p, C.c_size_t, C.c_int, _u64p, C.c_size_t, C.POINTER(C.c_size_t)],
    "b200fhe_level_beta": [C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)],
    "b200fhe_ntt_forward": [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t, C.c_void_p],
    "b200fhe_ntt_inverse": [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t, C.c_void_p],
    "b200fhe_ntt_inverse_scale": [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p],
    "b200fhe_ntt_forward_fuse_moddown": [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p],
    "b200fhe_bconv_P_to_Ql": [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p],
    "b200fhe_modup": [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p],
    "b200fhe_inner_prod": [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p), C.c_void_p],
    "b200fhe_moddown_from_ntt": [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p],
    "b200fhe_keyswitch": [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p), C.c_void_p],
    "b200fhe_tensor": [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p],
    "b200fhe_hmult_relin": [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p), C.c_void_p],
    "b200fhe_rescale": [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p],
    "b200fhe_automap_table": [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p],
    "b200fhe_galois_table_host": [C.c_int, C.c_uint32, _u32p],
    "b200fhe_apply_galois_ntt": [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_void_p],
    "b200fhe_rotate_fused": [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_uint32, C.POINTER(C.c_void_p), C.c_void_p],
    "b200fhe_eltwise": [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p],
    "b200fhe_ntt_forward_host": [C.c_void_p, _u64p, C.c_size_t, C.c_size_t, C.c_size_t],
    "b200fhe_ntt_inverse_host": [C.c_void_p, _u64p, C.c_size_t, C.c_size_t, C.c_size_t],
    "b200fhe_ntt_host_batch": [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int],
}
EXPORTS = sorted(list(_SIG) + ["b200fhe_last_error", "b200fhe_version", "b200fhe_launch_count"])


def lib():
    """Load libb200fhe.so; fails loudly if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            raise B200FheError("libb200fhe.so is missing: run __graft_entry__.build() (there is no CPU fallback)")
        L = C.CDLL(_SO)
        for name, args in _SIG.items():
            f = getattr(L, name)
            f.argtypes = args
            f.restype = C.c_int
        L.b200fhe_last_error.restype = C.c_char_p
        L.b200fhe_version.restype = C.c_char_p
        L.b200fhe_launch_count.restype = C.c_uint64
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise B200FheError("b200fhe error %d: %s" % (rc, lib().b200fhe_last_error().decode()))


def launch_count():
    return int(lib().b200fhe_launch_count())


def coeff_modulus_create(N, bits):
    out = np.zeros(len(bits), dtype=np.uint64)
    _check(lib().b200fhe_coeff_modulus_create(N, (C.c_int * len(bits))(*bits), len(bits), out.ctypes.data_as(_u64p)))
    return out


def galois_table(logN, elt):
    t = np.zeros(1 << logN, dtype=np.uint32)
    _check(lib().b200fhe_galois_table_host(logN, elt, t.ctypes.data_as(_u32p)))
    return t


def to_dev(a, device="cuda:0"):
    """numpy uint64 array -> torch int64 tensor on the GPU (same bits)."""
    import torch
    a = np.ascontiguousarray(a, dtype=np.uint64)
    return torch.from_numpy(a.view(np.int64)).to(device)


def to_host(t):
    return t.detach().cpu().contiguous().numpy().view(np.uint64)


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _stream():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


class Context:
    """PhantomContext-like owner of the NTT tables and per-level RNS constants (reference src/context.cu:121)."""

    def __init__(self, logN, primes_QP, size_P, device=0, host_only=False):
        self.logN, self.N = logN, 1 << logN
        self.primes = np.ascontiguousarray(primes_QP, dtype=np.uint64)
        self.size_QP, self.size_P = len(self.primes), size_P
        self.size_Q = self.size_QP - size_P
        self.host_only = host_only
        h = C.c_void_p()
        if host_only:
            _check(lib().b200fhe_ctx_create_host(logN, self.primes.ctypes.data_as(_u64p), self.size_QP, size_P, C.byref(h)))
        else:
            _check(lib().b200fhe_ctx_create(logN, self.primes.ctypes.data_as(_u64p), self.size_QP, size_P, device, C.byref(h)))
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            lib().b200fhe_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- tables / constants (host) ----
    def table(self, which):
        n = self.size_QP * (self.N if which < 4 else 1)
        out = np.zeros(n, dtype=np.uint64)
        _check(lib().b200fhe_ctx_export_table(self.h, which, out.ctypes.data_as(_u64p)))
        return out.reshape(self.size_QP, -1)

    def beta(self, size_Ql):
        b = C.c_size_t()
        _check(lib().b200fhe_level_beta(self.h, size_Ql, C.byref(b)))
        return b.value

    def level_const(self, size_Ql, which):
        n = C.c_size_t()
        _check(lib().b200fhe_level_export(self.h, size_Ql, which, None, 0, C.byref(n)))
        out = np.zeros(max(n.value, 1), dtype=np.uint64)
        _check(lib().b200fhe_level_export(self.h, size_Ql, which, out.ctypes.data_as(_u64p), n.value, C.byref(n)))
        return out[:n.value]

    # ---- device ops (torch int64 tensors carrying uint64 residues) ----
    def ntt_forward(self, x, limbs, start=0, size_Ql=0, npoly=1, poly_stride=0):
        _check(lib().b200fhe_ntt_forward(self.h, _ptr(x), limbs, start, size_Ql, npoly, poly_stride, _stream()))

    def ntt_inverse(self, x, limbs, start=0, size_Ql=0, npoly=1, poly_stride=0):
        _check(lib().b200fhe_ntt_inverse(self.h, _ptr(x), limbs, start, size_Ql, npoly, poly_stride, _stream()))

    def ntt_inverse_scale(self, out, x, limbs, start=0, size_Ql=0, scale=None, scale_shoup=None):
        _check(lib().b200fhe_ntt_inverse_scale(self.h, _ptr(out), _ptr(x), limbs, start, size_Ql, _ptr(scale),
                                               _ptr(scale_shoup), _stream()))

    def ntt_forward_fuse_moddown(self, ct, cx, pinv, pinv_shoup, delta, limbs, start=0):
        _check(lib().b200fhe_ntt_forward_fuse_moddown(self.h, _ptr(ct), _ptr(cx), _ptr(pinv), _ptr(pinv_shoup),
                                                      _ptr(delta), limbs, start, _stream()))

    @staticmethod
    def _evk(digits):
        return (C.c_void_p * len(digits))(*[d.data_ptr() for d in digits])

    def bconv_P_to_Ql(self, size_Ql, dst, src):
        _check(lib().b200fhe_bconv_P_to_Ql(self.h, size_Ql, _ptr(dst), _ptr(src), _stream()))

    def modup(self, size_Ql, dst, c2):
        _check(lib().b200fhe_modup(self.h, size_Ql, _ptr(dst), _ptr(c2), _stream()))

    def inner_prod(self, size_Ql, cx, up, evk):
        _check(lib().b200fhe_inner_prod(self.h, size_Ql, _ptr(cx), _ptr(up), self._evk(evk), _stream()))

    def moddown_from_ntt(self, size_Ql, ct, cx):
        _check(lib().b200fhe_moddown_from_ntt(self.h, size_Ql, _ptr(ct), _ptr(cx), _stream()))

    def keyswitch(self, size_Ql, ct, c2, evk):
        _check(lib().b200fhe_keyswitch(self.h, size_Ql, _ptr(ct), _ptr(c2), self._evk(evk), _stream()))

    def tensor(self, L, out, a, b):
        _check(lib().b200fhe_tensor(self.h, L, _ptr(out), _ptr(a), _ptr(b), _stream()))

    def hmult_relin(self, size_Ql, out, a, b, rlk):
        _check(lib().b200fhe_hmult_relin(self.h, size_Ql, _ptr(out), _ptr(a), _ptr(b), self._evk(rlk), _stream()))

    def rescale(self, size_Ql, dst, src, polys=2):
        _check(lib().b200fhe_rescale(self.h, size_Ql, _ptr(dst), _ptr(src), polys, _stream()))

    def automap_table(self, k, perm):
        _check(lib().b200fhe_automap_table(self.h, k, _ptr(perm), _stream()))

    def apply_galois_ntt(self, dst, src, perm, limbs, npoly=1, poly_stride=0):
        _check(lib().b200fhe_apply_galois_ntt(self.h, _ptr(dst), _ptr(src), _ptr(perm), limbs, npoly, poly_stride, _stream()))

    def rotate_fused(self, size_Ql, out, ct, k, glk):
        _check(lib().b200fhe_rotate_fused(self.h, size_Ql, _ptr(out), _ptr(ct), k, self._evk(glk), _stream()))

    def eltwise(self, op, out, a, b, limbs, size_Ql=0):
        _check(lib().b200fhe_eltwise(self.h, op, _ptr(out), _ptr(a), _ptr(b), limbs, size_Ql, _stream()))

    # ---- host-buffer entry points (H2D + compute + D2H inside) ----
    def ntt_forward_host(self, a, limbs, start=0, size_Ql=0):
        _check(lib().b200fhe_ntt_forward_host(self.h, a.ctypes.data_as(_u64p), limbs, start, size_Ql))

    def ntt_inverse_host(self, a, limbs, start=0, size_Ql=0):
        _check(lib().b200fhe_ntt_inverse_host(self.h, a.ctypes.data_as(_u64p), limbs, start, size_Ql))

    def ntt_host_batch(self, host_ptr, limbs, npoly, ops=3, size_Ql=0):
        """host_ptr: address of [npoly][limbs][N] uint64 in (pinned) host memory"""
        _check(lib().b200fhe_ntt_host_batch(self.h, C.c_void_p(host_ptr), limbs, size_Ql, npoly, ops))

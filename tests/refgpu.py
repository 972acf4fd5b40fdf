"""ctypes binding of oracle/_ref/libphantom_ref.so — the UNMODIFIED reference GPU library behind oracle/ref_shim.cu.
TEST / BASELINE INFRASTRUCTURE (tests and bench.py only)."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "oracle", "_ref", "libphantom_ref.so")
_lib = None
vp, sz, i32, u64p = C.c_void_p, C.c_size_t, C.c_int, C.POINTER(C.c_uint64)


def available():
    return os.path.exists(SO)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(SO)
        L.rref_last_error.restype = C.c_char_p
        for n in ("rref_create", "rref_gen_relinkey", "rref_gen_rotkeys", "rref_rotkey_get", "rref_encrypt",
                  "rref_ct_from_raw", "rref_ct_clone", "rref_ct_data"):
            getattr(L, n).restype = vp
        for n in ("rref_ct_size", "rref_ct_limbs", "rref_ct_chain_index"):
            getattr(L, n).restype = sz
            getattr(L, n).argtypes = [vp]
        L.rref_ct_scale.restype = C.c_double
        L.rref_ct_scale.argtypes = [vp]
        L.rref_auto_index.restype = C.c_uint32
        L.rref_auto_index.argtypes = [vp, C.c_int32]
        L.rref_create.argtypes = [i32, C.POINTER(i32), sz, sz]
        L.rref_destroy.argtypes = [vp]
        L.rref_primes.argtypes = [vp, u64p]
        for n in ("rref_ntt_forward", "rref_ntt_inverse", "rref_ntt_forward_special", "rref_ntt_inverse_special"):
            getattr(L, n).argtypes = [vp, vp, sz, sz]
        L.rref_modup.argtypes = [vp, sz, vp, vp]
        L.rref_moddown_from_ntt.argtypes = [vp, sz, vp, vp]
        L.rref_inner_prod.argtypes = [vp, sz, vp, vp, vp]
        L.rref_gen_relinkey.argtypes = [vp]
        L.rref_free_relinkey.argtypes = [vp]
        L.rref_relinkey_digits.argtypes = [vp, vp, C.POINTER(vp), sz]
        L.rref_gen_rotkeys.argtypes = [vp, C.POINTER(C.c_int32), sz]
        L.rref_free_rotkeys.argtypes = [vp]
        L.rref_rotkey_get.argtypes = [vp, C.c_uint32]
        L.rref_encrypt.argtypes = [vp, C.POINTER(C.c_double), sz, C.c_double, sz]
        L.rref_ct_from_raw.argtypes = [vp, sz, sz, vp, C.c_double]
        L.rref_ct_clone.argtypes = [vp]
        L.rref_ct_free.argtypes = [vp]
        L.rref_ct_data.argtypes = [vp]
        L.rref_encode_decode.argtypes = [vp, C.POINTER(C.c_double), sz, C.c_double, sz, C.POINTER(C.c_double)]
        L.rref_decrypt_raw.argtypes = [vp, vp, vp]
        L.rref_decrypt_decode.argtypes = [vp, vp, C.POINTER(C.c_double), sz]
        L.rref_multiply_and_relin.argtypes = [vp, vp, vp, vp]
        L.rref_multiply.argtypes = [vp, vp, vp]
        L.rref_relinearize.argtypes = [vp, vp, vp]
        L.rref_rescale.argtypes = [vp, vp]
        L.rref_rotate_fused.argtypes = [vp, vp, vp, vp, C.c_int32]
        L.rref_keyswitch.argtypes = [vp, vp, vp, vp]
        L.rref_time_ntt.argtypes = [vp, vp, sz, sz, sz, i32, C.POINTER(C.c_float), C.POINTER(C.c_float)]
        L.rref_time_hmult.argtypes = [vp, vp, vp, vp, i32, C.POINTER(C.c_float)]
        L.rref_time_rotate.argtypes = [vp, vp, vp, C.c_int32, i32, C.POINTER(C.c_float)]
        L.rref_time_rescale.argtypes = [vp, vp, i32, C.POINTER(C.c_float)]
        _lib = L
    return _lib


def ck(rc):
    if rc != 0:
        raise RuntimeError("reference: " + lib().rref_last_error().decode())


class _Ext:
    """minimal __cuda_array_interface__ holder to view a raw device pointer as a torch tensor"""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<i8", "data": (int(ptr), False), "version": 2}


def as_tensor(ptr, n):
    import torch
    return torch.as_tensor(_Ext(ptr, n), device="cuda")


class Ref:
    def __init__(self, logN, bits, nP):
        self.logN, self.N, self.nP, self.nQP = logN, 1 << logN, nP, len(bits)
        self.h = lib().rref_create(logN, (i32 * len(bits))(*bits), len(bits), nP)
        if not self.h:
            raise RuntimeError("reference context: " + lib().rref_last_error().decode())
        self.primes = np.zeros(len(bits), dtype=np.uint64)
        lib().rref_primes(self.h, self.primes.ctypes.data_as(u64p))

    def sync(self):
        lib().rref_sync()

    def relinkey_digits(self, key):
        """-> list of torch int64 tensors [2*QP*N] COPIED from the reference key (both sides own their bytes)"""
        import torch
        ptrs = (vp * 64)()
        n = lib().rref_relinkey_digits(self.h, key, ptrs, 64)
        assert n > 0
        self.sync()
        out = [as_tensor(ptrs[d], 2 * self.nQP * self.N).clone() for d in range(n)]
        torch.cuda.synchronize()
        return out

    def ct_tensor(self, ct):
        """copy of the ciphertext limbs as a torch int64 tensor [size][limbs][N]"""
        import torch
        L = lib()
        size, limbs = L.rref_ct_size(ct), L.rref_ct_limbs(ct)
        self.sync()
        t = as_tensor(L.rref_ct_data(ct), size * limbs * self.N).clone().view(size, limbs, self.N)
        torch.cuda.synchronize()
        return t

    def decrypt_slots(self, ct, o):
        """decrypt with the reference's secret key, decode on the host (tests/ckks_host.py) -> complex slots"""
        import torch
        import ckks_host
        L = lib()
        limbs = L.rref_ct_limbs(ct)
        buf = torch.zeros(limbs * self.N, dtype=torch.int64, device="cuda")
        n = L.rref_decrypt_raw(self.h, ct, buf.data_ptr())
        if n != limbs:
            raise RuntimeError("reference decrypt: " + L.rref_last_error().decode())
        pt = buf.cpu().numpy().view(np.uint64).reshape(limbs, self.N)
        return ckks_host.decode(o, pt, L.rref_ct_scale(ct))

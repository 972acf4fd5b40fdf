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
        L.rref_encode_raw.argtypes = [vp, C.POINTER(C.c_double), sz, C.c_double, sz, C.c_int, vp]
        L.rref_decode_raw.argtypes = [vp, vp, sz, C.c_double, C.POINTER(C.c_double)]
        L.rref_decrypt_decode.argtypes = [vp, vp, C.POINTER(C.c_double), sz]
        L.rref_multiply_and_relin.argtypes = [vp, vp, vp, vp]
        L.rref_multiply.argtypes = [vp, vp, vp]
        L.rref_relinearize.argtypes = [vp, vp, vp]
        L.rref_rescale.argtypes = [vp, vp]
        L.rref_rotate_fused.argtypes = [vp, vp, vp, vp, C.c_int32]
        L.rref_keyswitch.argtypes = [vp, vp, vp, vp]
        L.rref_time_ntt.argtypes = [vp, vp, sz, sz, sz, i32, C.POINTER(C.c_float), C.POINTER(C.c_float)]
        if hasattr(L, "rref_time_ntt_wide"):
            L.rref_time_ntt_wide.argtypes = [vp, vp, sz, sz, i32, C.POINTER(C.c_float), C.POINTER(C.c_float)]
        L.rref_time_hmult.argtypes = [vp, vp, vp, vp, i32, C.POINTER(C.c_float)]
        L.rref_time_rotate.argtypes = [vp, vp, vp, C.c_int32, i32, C.POINTER(C.c_float)]
        L.rref_time_rescale.argtypes = [vp, vp, i32, C.POINTER(C.c_float)]
        # ---- FLEXIBLEAUTO layer + bootstrap (oracle/ref_shim.cu, second half)
        i32p, dblp, szp = C.POINTER(C.c_int32), C.POINTER(C.c_double), C.POINTER(sz)
        for n in ("rref_boot_create", "rref_boot_mult_key", "rref_boot_galois_keys", "rref_boot_eval", "rref_boot_c2s",
                  "rref_boot_s2c", "rref_boot_mult_auto", "rref_boot_square", "rref_boot_chebyshev", "rref_boot_raise_mod",
                  "rref_boot_conj", "rref_boot_fast_rotation_ext", "rref_boot_keyswitch_down"):
            getattr(L, n).restype = vp
        L.rref_ct_noise.restype = C.c_uint32
        L.rref_ct_noise.argtypes = [vp]
        L.rref_ct_set_noise.argtypes = [vp, C.c_uint32]
        L.rref_ct_set_noise.restype = None
        L.rref_ct_set_scale.argtypes = [vp, C.c_double]
        L.rref_ct_set_scale.restype = None
        L.rref_boot_create.argtypes = [vp, C.c_uint32, C.c_uint32, C.c_double, C.c_uint32, i32]
        L.rref_boot_destroy.argtypes = [vp]
        L.rref_boot_destroy.restype = None
        L.rref_boot_scaling_factors.argtypes = [vp, i32, dblp, sz]
        L.rref_boot_correction_factor.argtypes = [vp]
        L.rref_boot_correction_factor.restype = C.c_uint32
        L.rref_boot_params.argtypes = [vp, i32, i32p]
        L.rref_boot_rotation_indices.argtypes = [vp, i32p, sz]
        L.rref_boot_diag.argtypes = [vp, i32, sz, sz, C.POINTER(vp), szp, dblp, szp]
        L.rref_boot_nstages.argtypes = [vp, i32]
        L.rref_boot_mult_key.argtypes = [vp]
        L.rref_boot_galois_keys.argtypes = [vp]
        L.rref_rotkey_has.argtypes = [vp, C.c_uint32]
        L.rref_boot_eval.argtypes = [vp, vp]
        L.rref_boot_mod_switch_level.argtypes = [vp, vp, sz]
        L.rref_boot_eval_iter.restype = vp
        L.rref_boot_eval_iter.argtypes = [vp, vp, C.c_uint32, C.c_uint32]
        L.rref_boot_c2s.argtypes = [vp, vp]
        L.rref_boot_s2c.argtypes = [vp, vp]
        L.rref_boot_mult_const.argtypes = [vp, vp, C.c_double]
        L.rref_boot_add_const.argtypes = [vp, vp, C.c_double]
        L.rref_boot_mod_reduce.argtypes = [vp, vp, sz]
        L.rref_boot_add_auto.argtypes = [vp, vp, vp, i32]
        L.rref_boot_mult_auto.argtypes = [vp, vp, vp]
        L.rref_boot_square.argtypes = [vp, vp]
        L.rref_boot_chebyshev.argtypes = [vp, vp, dblp, sz, C.c_double, C.c_double]
        L.rref_boot_raise_mod.argtypes = [vp, vp]
        L.rref_boot_mult_monomial.argtypes = [vp, vp, C.c_uint32]
        L.rref_boot_mult_integer.argtypes = [vp, vp, C.c_uint64]
        L.rref_boot_conj.argtypes = [vp, vp]
        L.rref_boot_fast_rotation_ext.argtypes = [vp, vp, C.c_int32, i32]
        L.rref_boot_keyswitch_down.argtypes = [vp, vp, i32]
        L.rref_time_bootstrap.argtypes = [vp, vp, i32, C.POINTER(C.c_float)]
        # ---- key material exchange (tests/test_keygen_gpu.py)
        L.rref_secret_key.argtypes = [vp, vp]
        L.rref_relinkey_from_raw.argtypes = [C.POINTER(vp), sz]
        L.rref_relinkey_from_raw.restype = vp
        L.rref_rotkeys_new.restype = vp
        L.rref_rotkeys_set_raw.argtypes = [vp, C.c_uint32, C.POINTER(vp), sz]
        L.rref_save.argtypes = [i32, vp, vp, sz]
        L.rref_save.restype = C.c_long
        L.rref_load.argtypes = [i32, vp, vp, sz]
        L.rref_load.restype = vp
        # ---- phantom::DNN (oracle/ref_shim.cu, third part)
        for n in ("rref_dnn_create", "rref_dnn_galois_keys", "rref_dnn_mult_key", "rref_tensor_new", "rref_tensor_ct",
                  "rref_dnn_enc_tensor", "rref_dnn_conv", "rref_dnn_batch_norm", "rref_dnn_add", "rref_dnn_relu",
                  "rref_dnn_relu_deg_two", "rref_dnn_sign", "rref_dnn_relu_composite", "rref_dnn_bootstrap",
                  "rref_dnn_avgpool_fullcon", "rref_dnn_softmax"):
            getattr(L, n).restype = vp
        for n in ("rref_tensor_size", "rref_tensor_width", "rref_tensor_slotstr", "rref_tensor_num_ch"):
            getattr(L, n).restype = sz
            getattr(L, n).argtypes = [vp]
        L.rref_dnn_create.argtypes = [vp, C.c_double, i32p, sz]
        L.rref_dnn_destroy.argtypes = [vp]
        L.rref_dnn_destroy.restype = None
        L.rref_dnn_galois_keys.argtypes = [vp]
        L.rref_dnn_mult_key.argtypes = [vp]
        L.rref_dnn_rotation_indices.argtypes = [vp, i32p, sz, sz, sz, sz, sz, i32]
        L.rref_tensor_new.argtypes = [C.POINTER(vp), sz, sz, sz, sz]
        L.rref_tensor_free.argtypes = [vp]
        L.rref_tensor_free.restype = None
        L.rref_tensor_ct.argtypes = [vp, sz]
        L.rref_dnn_enc_tensor.argtypes = [vp, dblp, sz, sz]
        L.rref_dnn_dec_tensor.argtypes = [vp, vp, dblp]
        L.rref_dnn_conv.argtypes = [vp, vp, dblp, sz, sz, sz, sz]
        L.rref_dnn_batch_norm.argtypes = [vp, vp, dblp, dblp, dblp, dblp]
        L.rref_dnn_add.argtypes = [vp, vp, vp]
        L.rref_dnn_relu.argtypes = [vp, vp, i32, i32, i32]
        L.rref_dnn_relu_deg_two.argtypes = [vp, vp]
        L.rref_dnn_sign.argtypes = [vp, vp, sz]
        L.rref_dnn_relu_composite.argtypes = [vp, vp]
        L.rref_dnn_bootstrap.argtypes = [vp, vp]
        L.rref_dnn_avgpool_fullcon.argtypes = [vp, vp, dblp, sz, sz, dblp]
        L.rref_dnn_softmax.argtypes = [vp, vp, sz, sz, sz, sz]
        L.rref_dnn_time_conv.argtypes = [vp, vp, dblp, sz, sz, sz, sz, i32, C.POINTER(C.c_float)]
        # ---- SEAL-style surface (oracle/ref_shim.cu, fourth part)
        for n in ("rref_gen_galois_keys", "rref_galois_key_get", "rref_encode_plain", "rref_plain_data", "rref_add_many"):
            getattr(L, n).restype = vp
        L.rref_gen_galois_keys.argtypes = [vp]
        L.rref_free_galois_keys.argtypes = [vp]
        L.rref_free_galois_keys.restype = None
        L.rref_galois_elts.argtypes = [vp, C.POINTER(C.c_uint32), sz]
        L.rref_galois_key_get.argtypes = [vp, sz]
        L.rref_encode_plain.argtypes = [vp, dblp, sz, C.c_double, sz]
        L.rref_plain_free.argtypes = [vp]
        L.rref_plain_free.restype = None
        L.rref_plain_data.argtypes = [vp]
        L.rref_seal_op.argtypes = [vp, i32, vp, vp, vp, i32, C.POINTER(i32), sz]
        L.rref_add_many.argtypes = [vp, C.POINTER(vp), sz]
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


class RefBoot:
    """The reference's FHECKKSRNS with its own keys and precomputed diagonals (bootstrapping_example.cu:121-160)."""

    def __init__(self, ref, budget=(2, 2), scale=2.0 ** 59, num_slots=0, keygen=True):
        self.ref = ref
        self.h = lib().rref_boot_create(ref.h, budget[0], budget[1], scale, num_slots, 1 if keygen else 0)
        if not self.h:
            raise RuntimeError("reference bootstrap setup: " + lib().rref_last_error().decode())

    def scaling_factors(self, which=0):
        buf = (C.c_double * 128)()
        n = lib().rref_boot_scaling_factors(self.h, which, buf, 128)
        return list(buf)[:n]

    def params(self, which):
        out = (C.c_int32 * 9)()
        lib().rref_boot_params(self.h, which, out)
        return list(out)

    def rotation_indices(self):
        out = (C.c_int32 * 4096)()
        n = lib().rref_boot_rotation_indices(self.h, out, 4096)
        assert 0 <= n <= 4096
        return list(out)[:n]

    def diagonals(self, which):
        """-> list over stages of lists of (ptr or None, chain_index, scale, limbs)"""
        L, res = lib(), []
        for s in range(L.rref_boot_nstages(self.h, which)):
            n = L.rref_boot_diag(self.h, which, s, 1 << 30, None, None, None, None)
            row = []
            for i in range(n):
                p, ch, sc, lm = vp(), sz(), C.c_double(), sz()
                L.rref_boot_diag(self.h, which, s, i, C.byref(p), C.byref(ch), C.byref(sc), C.byref(lm))
                row.append((p.value, ch.value, sc.value, lm.value))
            res.append(row)
        return res

    def key_digit_ptrs(self, key):
        ptrs = (vp * 64)()
        n = lib().rref_relinkey_digits(self.ref.h, key, ptrs, 64)
        assert n > 0
        return [ptrs[d] for d in range(n)]

    def install_keys(self, ev, rotations):
        """zero-copy: our evaluator reads the reference's own key memory"""
        L = lib()
        ev.set_relin_key(self.key_digit_ptrs(L.rref_boot_mult_key(self.h)))
        gk = L.rref_boot_galois_keys(self.h)
        ks = sorted({L.rref_auto_index(self.ref.h, r) for r in rotations} | {2 * self.ref.N - 1})
        for k in ks:
            assert L.rref_rotkey_has(gk, k), "reference has no key for automorphism %d" % k
            ev.set_galois_key(k, self.key_digit_ptrs(L.rref_rotkey_get(gk, k)))
        return ks

    def ct(self, f, *args):
        o = f(self.h, *args)
        if not o:
            raise RuntimeError("reference: " + lib().rref_last_error().decode())
        return o


def _dp(a):
    a = np.ascontiguousarray(np.asarray(a, dtype=np.float64))
    return a, a.ctypes.data_as(C.POINTER(C.c_double))


class RefDnn:
    """The reference's phantom::DNN on a RefBoot's context, with its OWN relinearisation key and rotation keys
    (DNN::RelinKeyGen / BuildGaloisKey) as in Resnet/resnet.cu."""

    def __init__(self, rb, scale, rotations):
        self.rb, self.ref = rb, rb.ref
        arr = (C.c_int32 * len(rotations))(*rotations)
        self.h = lib().rref_dnn_create(rb.h, scale, arr, len(rotations))
        if not self.h:
            raise RuntimeError("reference DNN: " + lib().rref_last_error().decode())
        self.rotations = list(rotations)

    def add_rotation_indices(self, have, width, kh, slotstr, avgpool=False):
        buf = (C.c_int32 * (len(have) + 64))(*have)
        n = lib().rref_dnn_rotation_indices(self.h, buf, len(have), len(have) + 64, width, kh, slotstr, 1 if avgpool else 0)
        return list(buf)[:n]

    def install_keys(self, dnn):
        """zero-copy: our DNN reads the reference DNN's own key memory"""
        L = lib()
        dnn.set_relin_key(self.rb.key_digit_ptrs(L.rref_dnn_mult_key(self.h)))
        gk = L.rref_dnn_galois_keys(self.h)
        for k in sorted({L.rref_auto_index(self.ref.h, r) for r in self.rotations}):
            assert L.rref_rotkey_has(gk, k), "reference DNN has no key for automorphism %d" % k
            dnn.set_galois_key(k, self.rb.key_digit_ptrs(L.rref_rotkey_get(gk, k)))

    def t(self, f, *args):
        o = f(self.h, *args)
        if not o:
            raise RuntimeError("reference DNN: " + lib().rref_last_error().decode())
        return o

    def enc_tensor(self, image):
        a, p = _dp(image)
        return self.t(lib().rref_dnn_enc_tensor, p, a.shape[0], a.shape[2])

    def dec_tensor(self, t):
        L = lib()
        w, ch = L.rref_tensor_width(t), L.rref_tensor_num_ch(t)
        out = np.zeros((w, w, ch))
        ck(L.rref_dnn_dec_tensor(self.h, t, out.ctypes.data_as(C.POINTER(C.c_double))))
        return out

    def conv(self, t, weight, stride=1):
        w, p = _dp(weight)
        return self.t(lib().rref_dnn_conv, t, p, w.shape[0], w.shape[2], w.shape[3], stride)

    def batch_norm(self, t, weight, bias, mean, var):
        keep = [_dp(x) for x in (weight, bias, mean, var)]
        return self.t(lib().rref_dnn_batch_norm, t, *[k[1] for k in keep])

    def avgpool_fullcon(self, t, weight, bias):
        w, p = _dp(weight)
        b, pb = _dp(bias)
        return self.t(lib().rref_dnn_avgpool_fullcon, t, p, w.shape[0], w.shape[1], pb)

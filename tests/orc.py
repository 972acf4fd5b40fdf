"""ctypes binding of oracle/liboracle.so (the CPU checker) for tests, smoke() and bench.py's CPU legs.

TEST INFRASTRUCTURE: nothing under phantom-fhe-boot_b200/ may import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
_u64p = C.POINTER(C.c_uint64)
_u32p = C.POINTER(C.c_uint32)


def _p(a):
    assert a.dtype == np.uint64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_u64p)


def _p32(a):
    assert a.dtype == np.uint32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_u32p)


def build_oracle():
    so = os.path.join(ORACLE_DIR, "liboracle.so")
    src = os.path.join(ORACLE_DIR, "fhe_oracle.c")
    if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "-s"])
    return so


class Level(C.Structure):
    _fields_ = [("size_Ql", C.c_size_t), ("size_P", C.c_size_t), ("size_Q", C.c_size_t),
                ("alpha", C.c_size_t), ("beta", C.c_size_t),
                ("partQlHatInv_mod_Ql", _u64p), ("partQlHatInv_mod_Ql_shoup", _u64p),
                ("modup_QHatModp", _u64p), ("modup_off", C.POINTER(C.c_size_t)),
                ("bigP_mod_q", _u64p), ("bigP_mod_q_shoup", _u64p),
                ("bigPInv_mod_q", _u64p), ("bigPInv_mod_q_shoup", _u64p),
                ("PHatInv_mod_p", _u64p), ("PHatInv_mod_p_shoup", _u64p),
                ("PHat_mod_q", _u64p),
                ("inv_q_last_mod_q", _u64p), ("inv_q_last_mod_q_shoup", _u64p)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(build_oracle())
        L.orc_mulmod.restype = C.c_uint64
        L.orc_mulmod.argtypes = [C.c_uint64] * 3
        L.orc_powmod.restype = C.c_uint64
        L.orc_powmod.argtypes = [C.c_uint64] * 3
        L.orc_invmod.restype = C.c_uint64
        L.orc_invmod.argtypes = [C.c_uint64] * 2
        L.orc_shoup.restype = C.c_uint64
        L.orc_shoup.argtypes = [C.c_uint64] * 2
        L.orc_minimal_primitive_root.restype = C.c_uint64
        L.orc_minimal_primitive_root.argtypes = [C.c_uint64] * 2
        L.orc_ctx_create.restype = C.c_void_p
        L.orc_ctx_create.argtypes = [C.c_int, _u64p, C.c_size_t, C.c_size_t]
        L.orc_ctx_destroy.argtypes = [C.c_void_p]
        L.orc_ctx_tables.restype = _u64p
        L.orc_ctx_tables.argtypes = [C.c_void_p, C.c_int]
        L.orc_level_create.restype = C.POINTER(Level)
        L.orc_level_create.argtypes = [C.c_void_p, C.c_size_t]
        L.orc_level_destroy.argtypes = [C.POINTER(Level)]
        L.orc_set_threads.restype = C.c_int
        L.orc_set_threads.argtypes = [C.c_int]
        _lib = L
    return _lib


def coeff_modulus_create(N, bits):
    b = (C.c_int * len(bits))(*bits)
    out = np.zeros(len(bits), dtype=np.uint64)
    rc = lib().orc_coeff_modulus_create(C.c_size_t(N), b, C.c_size_t(len(bits)), _p(out))
    if rc:
        raise ValueError("orc_coeff_modulus_create failed: %d" % rc)
    return out


def ntt_tables(logN, q):
    N = 1 << logN
    arrs = [np.zeros(N, dtype=np.uint64) for _ in range(4)]
    ninv = np.zeros(2, dtype=np.uint64)
    lib().orc_ntt_tables(C.c_int(logN), C.c_uint64(q), *[_p(a) for a in arrs],
                         _p(ninv), _p(ninv[1:]))
    return arrs + [int(ninv[0]), int(ninv[1])]


def ntt_fwd_naive(a, logN, q, psi):
    out = np.zeros_like(a)
    lib().orc_ntt_fwd_naive(_p(a), _p(out), C.c_int(logN), C.c_uint64(q), C.c_uint64(psi))
    return out


def negacyclic_mul_naive(a, b, q):
    c = np.zeros_like(a)
    lib().orc_negacyclic_mul_naive(_p(a), _p(b), _p(c), C.c_size_t(len(a)), C.c_uint64(q))
    return c


def bconv(src, ibase, obase):
    N = src.shape[1]
    ibase = np.ascontiguousarray(ibase, dtype=np.uint64)
    obase = np.ascontiguousarray(obase, dtype=np.uint64)
    dst = np.zeros((len(obase), N), dtype=np.uint64)
    lib().orc_bconv(_p(dst), _p(np.ascontiguousarray(src)), C.c_size_t(N), _p(ibase), C.c_size_t(len(ibase)),
                    _p(obase), C.c_size_t(len(obase)))
    return dst


def galois_table(logN, elt):
    t = np.zeros(1 << logN, dtype=np.uint32)
    lib().orc_galois_table(C.c_int(logN), C.c_uint32(elt), _p32(t))
    return t


def automap_table(logN, k):
    t = np.zeros(1 << logN, dtype=np.uint32)
    lib().orc_automap_table(C.c_int(logN), C.c_uint32(k), _p32(t))
    return t


class Ctx:
    """orc_ctx + per-level constants.  primes = Q primes then P primes (PhantomContext order)."""

    def __init__(self, logN, primes_QP, size_P):
        self.logN, self.N = logN, 1 << logN
        self.primes = np.ascontiguousarray(primes_QP, dtype=np.uint64)
        self.size_QP, self.size_P = len(self.primes), size_P
        self.size_Q = self.size_QP - size_P
        self.h = C.c_void_p(lib().orc_ctx_create(logN, _p(self.primes), self.size_QP, size_P))
        self._levels = {}

    def __del__(self):
        try:
            for l in self._levels.values():
                lib().orc_level_destroy(l)
            lib().orc_ctx_destroy(self.h)
        except Exception:
            pass

    def table(self, which):
        n = self.size_QP * (self.N if which < 4 else 1)
        ptr = lib().orc_ctx_tables(self.h, which)
        return np.ctypeslib.as_array(ptr, shape=(n,)).reshape(self.size_QP, -1).copy()

    def level(self, size_Ql):
        if size_Ql not in self._levels:
            self._levels[size_Ql] = lib().orc_level_create(self.h, size_Ql)
        return self._levels[size_Ql]

    def level_consts(self, size_Ql):
        l = self.level(size_Ql).contents
        P, Ql = self.size_P, size_Ql

        def arr(p, n):
            return np.ctypeslib.as_array(p, shape=(n,)).copy() if n else np.zeros(0, dtype=np.uint64)
        off = [l.modup_off[d] for d in range(l.beta + 1)]
        mats = []
        allm = arr(l.modup_QHatModp, off[-1])
        for d in range(l.beta):
            ni = (Ql - l.alpha * (l.beta - 1)) if d == l.beta - 1 else l.alpha
            mats.append(allm[off[d]:off[d + 1]].reshape(Ql + P - ni, ni))
        return dict(alpha=l.alpha, beta=l.beta,
                    partQlHatInv=arr(l.partQlHatInv_mod_Ql, Ql), partQlHatInv_shoup=arr(l.partQlHatInv_mod_Ql_shoup, Ql),
                    modup_mats=mats, bigP_mod_q=arr(l.bigP_mod_q, Ql), bigPInv_mod_q=arr(l.bigPInv_mod_q, Ql),
                    bigPInv_mod_q_shoup=arr(l.bigPInv_mod_q_shoup, Ql),
                    PHatInv_mod_p=arr(l.PHatInv_mod_p, P), PHatInv_mod_p_shoup=arr(l.PHatInv_mod_p_shoup, P),
                    PHat_mod_q=arr(l.PHat_mod_q, Ql * P).reshape(Ql, P),
                    inv_q_last_mod_q=arr(l.inv_q_last_mod_q, max(Ql - 1, 0)),
                    inv_q_last_mod_q_shoup=arr(l.inv_q_last_mod_q_shoup, max(Ql - 1, 0)))

    # transforms over limbs; size_Ql=0 -> limb t uses prime t
    def ntt_fwd(self, data, start=0, size_Ql=0):
        d = np.ascontiguousarray(data).copy()
        limbs = d.size // self.N - start
        lib().orc_ntt_fwd_limbs(self.h, _p(d), C.c_size_t(limbs), C.c_size_t(start), C.c_size_t(size_Ql))
        return d

    def ntt_inv(self, data, start=0, size_Ql=0):
        d = np.ascontiguousarray(data).copy()
        limbs = d.size // self.N - start
        lib().orc_ntt_inv_limbs(self.h, _p(d), C.c_size_t(limbs), C.c_size_t(start), C.c_size_t(size_Ql))
        return d

    def _evk(self, evk):
        arrs = [np.ascontiguousarray(e, dtype=np.uint64) for e in evk]
        ptrs = (_u64p * len(arrs))(*[_p(a) for a in arrs])
        return arrs, ptrs

    def modup(self, size_Ql, c2):
        l = self.level(size_Ql)
        beta = l.contents.beta
        dst = np.zeros((beta, size_Ql + self.size_P, self.N), dtype=np.uint64)
        lib().orc_modup(self.h, l, _p(dst), _p(np.ascontiguousarray(c2)))
        return dst

    def inner_prod(self, size_Ql, t_mod_up, evk):
        l = self.level(size_Ql)
        keep, ptrs = self._evk(evk)
        cx = np.zeros((2, size_Ql + self.size_P, self.N), dtype=np.uint64)
        lib().orc_inner_prod(self.h, l, _p(cx), _p(np.ascontiguousarray(t_mod_up)), ptrs)
        return cx

    def moddown(self, size_Ql, cx):
        l = self.level(size_Ql)
        cxc = np.ascontiguousarray(cx).copy()
        ct = np.zeros((size_Ql, self.N), dtype=np.uint64)
        lib().orc_moddown_from_ntt(self.h, l, _p(ct), _p(cxc))
        return ct

    def keyswitch(self, size_Ql, ct, c2, evk):
        l = self.level(size_Ql)
        keep, ptrs = self._evk(evk)
        out = np.ascontiguousarray(ct).copy()
        lib().orc_keyswitch(self.h, l, _p(out), _p(np.ascontiguousarray(c2)), ptrs)
        return out

    def tensor(self, L, a, b):
        out = np.zeros((3, L, self.N), dtype=np.uint64)
        lib().orc_tensor_prod(self.h, C.c_size_t(L), _p(np.ascontiguousarray(a)), _p(np.ascontiguousarray(b)), _p(out))
        return out

    def tensor_square(self, L, a):
        out = np.zeros((3, L, self.N), dtype=np.uint64)
        lib().orc_tensor_square(self.h, C.c_size_t(L), _p(np.ascontiguousarray(a)), _p(out))
        return out

    def rescale(self, size_Ql, src):
        l = self.level(size_Ql)
        s = np.ascontiguousarray(src).copy()
        polys = s.shape[0]
        dst = np.zeros((polys, size_Ql - 1, self.N), dtype=np.uint64)
        lib().orc_rescale(self.h, l, _p(s), C.c_size_t(polys), _p(dst))
        return dst

    def hmult_relin(self, size_Ql, a, b, rlk):
        l = self.level(size_Ql)
        keep, ptrs = self._evk(rlk)
        out = np.zeros((2, size_Ql, self.N), dtype=np.uint64)
        aa, bb = np.ascontiguousarray(a), np.ascontiguousarray(b)
        lib().orc_hmult_relin(self.h, l, _p(aa), _p(bb), ptrs, _p(out))
        return out

    def rotate_fused(self, size_Ql, ct, glk, table):
        l = self.level(size_Ql)
        keep, ptrs = self._evk(glk)
        out = np.zeros((2, size_Ql, self.N), dtype=np.uint64)
        lib().orc_rotate_fused(self.h, l, _p(np.ascontiguousarray(ct)), ptrs, _p32(table), _p(out))
        return out

    def binop(self, name, L, a, b):
        out = np.zeros((L, self.N), dtype=np.uint64)
        getattr(lib(), "orc_" + name)(self.h, C.c_size_t(L), _p(np.ascontiguousarray(a)), _p(np.ascontiguousarray(b)), _p(out))
        return out


def rand_residues(rng, primes, N, lead=()):
    """uniform residues in [0, q_i) per limb; shape lead + (len(primes), N)"""
    primes = np.asarray(primes, dtype=np.uint64)
    shape = tuple(lead) + (len(primes), N)
    x = rng.integers(0, 1 << 63, size=shape, dtype=np.uint64)
    return x % primes.reshape((1,) * len(lead) + (-1, 1))

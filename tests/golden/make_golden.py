"""Generates tests/golden/ref_host_golden.npz from the REFERENCE's own host code
(oracle/_ref/libphantom_hostref.so = /root/reference/src/host/*.cu compiled unmodified; `make -C oracle ref`).
Run in the build container (the reference tree is not on the GPU box):  python tests/golden/make_golden.py
The vectors pin oracle/fhe_oracle.c and the product's hostmath.cpp to the reference's primes, NTT tables
and base-conversion constants (tests/test_golden.py)."""
import ctypes as C
import hashlib
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
href = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libphantom_hostref.so"))
u64p = C.POINTER(C.c_uint64)

CONFIGS = {
    # name: (logN, bit sizes of Q..., number of special primes at the end)
    "T50": (12, [50, 50, 50], 0),                       # test/ntt_test.cu primes
    "T60": (12, [60, 60, 60], 1),
    "S14": (14, [60] + [59] * 7 + [60] * 4, 4),
    "K": (16, [60] + [59] * 29 + [60] * 10, 10),        # bootstrapping_example.cu chain
    "Kp": (16, [60] + [59] * 35 + [60] * 12, 12),
    "KB15": (15, [60] + [50] * 29 + [60] * 15, 15),     # benchmark/keyswitch_bench.cu
}


def primes(N, bits):
    out = np.zeros(len(bits), dtype=np.uint64)
    assert href.href_coeff_modulus_create(C.c_size_t(N), (C.c_int * len(bits))(*bits), C.c_size_t(len(bits)), out.ctypes.data_as(u64p)) == 0
    return out


def tables(logN, q):
    N = 1 << logN
    arrs = [np.zeros(N, dtype=np.uint64) for _ in range(4)]
    root, ni, nis = C.c_uint64(), C.c_uint64(), C.c_uint64()
    assert href.href_ntt_tables(logN, C.c_uint64(int(q)), C.byref(root), *[a.ctypes.data_as(u64p) for a in arrs], C.byref(ni), C.byref(nis)) == 0
    return root.value, arrs, ni.value, nis.value


def converter(ib, ob):
    ib, ob = np.ascontiguousarray(ib, dtype=np.uint64), np.ascontiguousarray(ob, dtype=np.uint64)
    hi, his = np.zeros(len(ib), dtype=np.uint64), np.zeros(len(ib), dtype=np.uint64)
    M = np.zeros((len(ob), len(ib)), dtype=np.uint64)
    assert href.href_base_converter(ib.ctypes.data_as(u64p), C.c_size_t(len(ib)), ob.ctypes.data_as(u64p), C.c_size_t(len(ob)),
                                    hi.ctypes.data_as(u64p), his.ctypes.data_as(u64p), M.ctypes.data_as(u64p)) == 0
    return hi, his, M


def main():
    out = {}
    for name, (logN, bits, nP) in CONFIGS.items():
        pr = primes(1 << logN, bits)
        out[name + "_primes"] = pr
        roots, ninv, ninvs, digests, ratios = [], [], [], [], []
        for i, q in enumerate(pr):
            if logN == 16 and i not in (0, 1, len(pr) - 1):   # keep generation fast: 3 primes of the big chains
                continue
            root, arrs, ni, nis = tables(logN, q)
            roots.append(root); ninv.append(ni); ninvs.append(nis)
            digests.append(np.frombuffer(b"".join(hashlib.sha256(a.tobytes()).digest() for a in arrs), dtype=np.uint8))
            r = np.zeros(2, dtype=np.uint64)
            assert href.href_const_ratio(C.c_uint64(int(q)), r.ctypes.data_as(u64p)) == 0
            ratios.append(r)
            if name == "T60" and i == 0:
                for k, a in zip(("tw", "tw_shoup", "itw", "itw_shoup"), arrs):
                    out["T60_q0_" + k] = a
        out[name + "_roots"] = np.array(roots, dtype=np.uint64)
        out[name + "_ninv"] = np.array(ninv, dtype=np.uint64)
        out[name + "_ninv_shoup"] = np.array(ninvs, dtype=np.uint64)
        out[name + "_table_sha256"] = np.stack(digests)
        out[name + "_ratio"] = np.stack(ratios)
        if nP:
            Q = len(pr) - nP
            # top level: digit 0 converter (part Ql -> complement of QlP) and P -> Ql
            hi, his, M = converter(pr[:nP], np.concatenate([pr[nP:Q], pr[Q:]]))
            out[name + "_d0_hatinv"], out[name + "_d0_hatinv_shoup"], out[name + "_d0_mat"] = hi, his, M
            hi, his, M = converter(pr[Q:], pr[:Q])
            out[name + "_P_hatinv"], out[name + "_P_hatinv_shoup"], out[name + "_P_mat"] = hi, his, M
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_host_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()

import sys, os, ctypes as C
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np, torch
import refgpu
def dp(v): return v.ctypes.data_as(C.POINTER(C.c_double))
for logN, bits, nP, sc in [(13, [60, 50, 50, 50, 50, 50, 60, 60], 2, 50), (13, [60,40,40,60], 1, 40), (15, [60]+[40]*15+[60,60], 2, 40), (12, [60,40,60],1,40)]:
    ref = refgpu.Ref(logN, bits, nP); L = refgpu.lib(); N = 1 << logN
    rng = np.random.default_rng(23)
    va = rng.uniform(-1, 1, N // 2)
    scale = 2.0 ** sc
    dec = np.zeros(N // 2)
    for chain in (1, 2):
        refgpu.ck(L.rref_encode_decode(ref.h, dp(va), N//2, scale, chain, dp(dec)))
        print(logN, bits[:3], nP, "encode/decode chain", chain, "err", np.abs(dec - va).max())
    a = L.rref_encrypt(ref.h, dp(va), N // 2, scale, 1)
    refgpu.ck(L.rref_decrypt_decode(ref.h, a, dp(dec), N // 2))
    print("   fresh enc/dec err", np.abs(dec - va).max())

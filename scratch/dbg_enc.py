import sys, ctypes as C
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np, torch
import phantom_fhe_boot_b200 as fhe, refgpu, orc, ckks_host
for logN in (12, 14):
    bits = [60, 50, 50, 50, 50, 60, 60]
    ref = refgpu.Ref(logN, bits, 2)
    ctx = fhe.Context(logN, ref.primes, 2)
    enc = fhe.Encoder(ctx)
    o = orc.Ctx(logN, ref.primes, 2)
    slots = (1 << logN) // 2
    rng = np.random.default_rng(logN)
    v = rng.uniform(-1, 1, slots) + 1j * rng.uniform(-1, 1, slots)
    scale = 2.0 ** 50
    ours = enc.encode(v, 5, scale)
    out = torch.zeros((5, ref.N), dtype=torch.int64, device="cuda")
    f = np.ascontiguousarray(v).view(np.float64)
    n = refgpu.lib().rref_encode_raw(ref.h, f.ctypes.data_as(C.POINTER(C.c_double)), v.size, scale, 1, 0, out.data_ptr())
    a = ours.cpu().numpy().view(np.uint64); b = out.cpu().numpy().view(np.uint64)
    da = ckks_host.decode(o, a, scale); db = ckks_host.decode(o, b, scale)
    print(logN, "equal", np.array_equal(a, b), "ours err", np.abs(da - v).max(), "ref err", np.abs(db - v).max())
    ca = o.ntt_inv(a); cb = o.ntt_inv(b)
    q0 = int(ref.primes[0])
    d = (ca[0].astype(np.int64) - cb[0].astype(np.int64))
    nz = np.nonzero(d)[0]
    print("   coefficient diffs:", len(nz), "of", ref.N, "first idx", nz[:8], "vals", d[nz[:8]])

"""GPU parity of the encrypted CNN layers against the REFERENCE's phantom::DNN (src/dnn.cu,
built unmodified into oracle/_ref/libphantom_ref.so): same context, the reference DNN's own relinearisation / rotation
keys and the reference bootstrapper's keys read zero-copy, the same input tensor -> every limb of every output
ciphertext of Conv (stride 1 and 2, NoiseScaleDeg 1 and 2 inputs), BatchNorm, Add, Relu, ReluDegTwo, Sign,
AvgPoolFullCon, BootStrap and ReluComposite must be bit-identical, with equal level / scale / NoiseScaleDeg.
Decrypted values are checked against a plain numpy model of the layer with the tolerance stated at each test."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

import refgpu

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
import resnet_model  # noqa: E402

pytestmark = pytest.mark.gpu

LOGN = 12
W = 8                       # image width: W*W = 64 sparse slots (Resnet/resnet.cu uses 32 -> 1024 at N = 2^16)
LEVELS_AFTER = 11           # Resnet/resnet.cu: levelsAvailableAfterBootstrap
DEPTH = LEVELS_AFTER + 18   # + FHECKKSRNS::GetBootstrapDepth({2,2})
NP = 10
BITS = [60] + [59] * DEPTH + [60] * NP
SCALE = 2.0 ** 59


class Env:
    pass


@pytest.fixture(scope="module")
def env(fhe):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    if not refgpu.available():
        pytest.skip("oracle/_ref/libphantom_ref.so not built")
    e = Env()
    e.torch, e.fhe, e.L = torch, fhe, refgpu.lib()
    e.ref = refgpu.Ref(LOGN, BITS, NP)
    e.N, e.Q = 1 << LOGN, len(BITS) - NP
    e.ctx = fhe.Context(LOGN, e.ref.primes, NP)
    e.rb = refgpu.RefBoot(e.ref, (2, 2), SCALE, num_slots=W * W, keygen=True)   # also set_sparse_encode(W*W*2)
    e.boot = fhe.Bootstrapper(e.ctx, (2, 2), num_slots=W * W)
    e.ev = e.boot.ev
    e.enc = fhe.Encoder(e.ctx)
    e.boot.precompute(e.enc)
    brots = e.boot.rotation_indices()
    assert brots == e.rb.rotation_indices()
    e.rb.install_keys(e.ev, brots)
    # the layer rotations of a 3-stage network on a WxW image (model_resnet20.cu PrePareResNet20, scaled down)
    rots = fhe.DNN.add_rotation_indices([], W, 3, 0)
    rots = fhe.DNN.add_rotation_indices(rots, W // 2, 3, 1)
    rots = fhe.DNN.add_rotation_indices(rots, W // 4, 3, 2)
    rots = fhe.DNN.add_avgpool_rotations(rots, W // 4, 2)
    rots = fhe.DNN.add_avgpool_rotations(rots, W // 2, 1)
    assert rots[:len(fhe.resnet20_rotation_indices(W))] == fhe.resnet20_rotation_indices(W)
    e.rd = refgpu.RefDnn(e.rb, SCALE, rots)
    r2 = e.rd.add_rotation_indices([], W, 3, 0)
    r2 = e.rd.add_rotation_indices(r2, W // 2, 3, 1)
    r2 = e.rd.add_rotation_indices(r2, W // 4, 3, 2)
    r2 = e.rd.add_rotation_indices(r2, W // 4, 0, 2, avgpool=True)
    r2 = e.rd.add_rotation_indices(r2, W // 2, 0, 1, avgpool=True)
    assert r2 == rots, "rotation index helpers differ from the reference"
    e.rots = rots
    e.dnn = fhe.DNN(e.ev, e.enc, SCALE)
    e.rd.install_keys(e.dnn)
    # decryption for the value checks: OUR decryptor / decoder holding the reference's secret key (the reference's
    # multi-limb decoder is miscompiled by this toolchain)
    e.kg = fhe.KeyGenerator(e.ctx, seed=bytes(range(32)))
    sk = torch.zeros(e.ref.nQP * e.N, dtype=torch.int64, device="cuda")
    refgpu.ck(e.L.rref_secret_key(e.ref.h, sk.data_ptr()))
    e.kg.set_secret(sk)
    return e


def dec(e, t):
    return e.dnn.dec_tensor(e.kg, t)


def ours_tensor(e, rt):
    """reference TensorCT -> ours (copies of the limbs + metadata)"""
    L = e.L
    cts = []
    for i in range(L.rref_tensor_size(rt)):
        rct = L.rref_tensor_ct(rt, i)
        cts.append(e.ev.import_ct(e.ref.ct_tensor(rct), L.rref_ct_limbs(rct), polys=L.rref_ct_size(rct),
                                  scale=L.rref_ct_scale(rct), noise=L.rref_ct_noise(rct)))
    return e.fhe.TensorCT(cts, L.rref_tensor_width(rt), L.rref_tensor_slotstr(rt), L.rref_tensor_num_ch(rt))


def same_tensor(e, ours, rt, what):
    L = e.L
    e.ref.sync()
    e.torch.cuda.synchronize()
    assert (ours.width, ours.slotstr, ours.num_ch) == (L.rref_tensor_width(rt), L.rref_tensor_slotstr(rt), L.rref_tensor_num_ch(rt)), what
    assert len(ours.cts) == L.rref_tensor_size(rt), what + ": ciphertext count"
    for i, ct in enumerate(ours.cts):
        rct = L.rref_tensor_ct(rt, i)
        d = ct.desc()
        tag = "%s channel %d" % (what, i)
        assert d.limbs == L.rref_ct_limbs(rct), "%s: limbs %d vs reference %d" % (tag, d.limbs, L.rref_ct_limbs(rct))
        assert d.polys == L.rref_ct_size(rct), tag
        assert d.scale == L.rref_ct_scale(rct), "%s: scale %r vs %r" % (tag, d.scale, L.rref_ct_scale(rct))
        assert d.noise_scale_deg == L.rref_ct_noise(rct), tag + ": NoiseScaleDeg"
        a, b = ct.tensor(), e.ref.ct_tensor(rct)
        e.torch.cuda.synchronize()
        if not bool((a.reshape(-1) == b.reshape(-1)).all()):
            bad = (a != b).reshape(d.polys, d.limbs, -1).any(dim=2).nonzero().tolist()
            raise AssertionError("%s: limbs differ from the reference at (poly, limb) %s" % (tag, bad[:8]))


def image(seed, w=W, ch=2, lo=-1.0, hi=1.0):
    return np.random.default_rng(seed).uniform(lo, hi, (w, w, ch))


def conv_plain(x, wt, stride):
    """the layer DNN::Conv computes on the slots: 'same' correlation, then subsampling by `stride`"""
    w, kh, half = x.shape[0], wt.shape[0], wt.shape[0] // 2
    out = np.zeros((w, w, wt.shape[3]))
    for i in range(kh):
        for j in range(kh):
            for ip in range(w):
                for jp in range(w):
                    a, b = ip + i - half, jp + j - half
                    if 0 <= a < w and 0 <= b < w:
                        out[ip, jp, :] += x[a, b, :] @ wt[i, j, :, :]
    return out[::stride, ::stride, :]


def test_enc_dec_tensor(env):
    """EncTensor of each side decrypts to the image within 2^-30 (fresh ciphertext at scale 2^59)"""
    e = env
    img = image(1)
    rt = e.rd.enc_tensor(img)
    assert np.abs(dec(e, ours_tensor(e, rt)) - img).max() < 2.0 ** -30
    # our EncTensor under our public key of the same secret
    pk = e.kg.public_key()
    ot = e.dnn.enc_tensor(e.kg, pk, img)
    assert (ot.width, ot.slotstr, ot.num_ch) == (W, 0, 2)
    assert np.abs(dec(e, ot) - img).max() < 2.0 ** -30
    # ... and the reference evaluates on it: one Conv on each side of OUR ciphertexts stays bit-identical
    arr = (C.c_void_p * 2)()
    keep = []
    for i, ct in enumerate(ot.cts):
        d = ct.desc()
        t = ct.tensor()
        keep.append(t)
        arr[i] = e.L.rref_ct_from_raw(e.ref.h, 1, 2, t.data_ptr(), d.scale)
    rt2 = e.L.rref_tensor_new(arr, 2, W, 0, 2)
    wt = np.random.default_rng(5).uniform(-0.5, 0.5, (3, 3, 2, 2))
    same_tensor(e, e.dnn.conv(ot, wt, 1), e.rd.conv(rt2, wt, 1), "Conv on our EncTensor")


@pytest.mark.parametrize("stride,in_ch,out_ch", [(1, 2, 3), (2, 2, 5), (1, 1, 1)])
def test_conv_fresh_input(env, stride, in_ch, out_ch):
    """Conv on a fresh (NoiseScaleDeg 1) tensor: bit-exact; decrypted |err| < 2^-25 of the plain convolution"""
    e = env
    img = image(2 + stride, ch=in_ch)
    wt = np.random.default_rng(7).uniform(-0.5, 0.5, (3, 3, in_ch, out_ch))
    rt = e.rd.enc_tensor(img)
    ot = ours_tensor(e, rt)
    rout = e.rd.conv(rt, wt, stride)
    oout = e.dnn.conv(ot, wt, stride)
    same_tensor(e, oout, rout, "Conv stride %d" % stride)
    if stride == 1:  # (a strided output keeps its pixels on the fine grid: compared in the chain test below)
        assert np.abs(dec(e, oout) - conv_plain(img, wt, 1)).max() < 2.0 ** -25


def test_conv_noise2_input_and_chain(env):
    """Conv -> BatchNorm -> Conv (stride 2 first, so the second runs at slotstr 1 on a NoiseScaleDeg-2 input) -> Add"""
    e = env
    img = image(11, ch=2)
    rng = np.random.default_rng(12)
    w1 = rng.uniform(-0.5, 0.5, (3, 3, 2, 3))
    w2 = rng.uniform(-0.5, 0.5, (3, 3, 3, 2))
    bn = [rng.uniform(0.5, 1.5, 3), rng.uniform(-0.2, 0.2, 3), rng.uniform(-0.1, 0.1, 3), rng.uniform(0.5, 1.5, 3)]
    rt = e.rd.enc_tensor(img)
    ot = ours_tensor(e, rt)
    r1, o1 = e.rd.conv(rt, w1, 2), e.dnn.conv(ot, w1, 2)
    same_tensor(e, o1, r1, "Conv stride 2")
    r2, o2 = e.rd.batch_norm(r1, *bn), e.dnn.batch_norm(o1, *bn)
    same_tensor(e, o2, r2, "BatchNorm")
    r3, o3 = e.rd.conv(r2, w2, 1), e.dnn.conv(o2, w2, 1)
    same_tensor(e, o3, r3, "Conv on a NoiseScaleDeg-2 input at slotstr 1")
    r4 = e.rd.t(e.L.rref_dnn_add, r3, r3)
    o4 = e.dnn.add(o3, o3)
    same_tensor(e, o4, r4, "Add")
    # Add of tensors at different levels (residual connection): adjusts the shallower one
    w3 = rng.uniform(-0.5, 0.5, (3, 3, 2, 2))
    r5, o5 = e.rd.conv(r3, w3, 1), e.dnn.conv(o3, w3, 1)
    same_tensor(e, o5, r5, "third Conv")
    r6 = e.rd.t(e.L.rref_dnn_add, r5, r3)
    o6 = e.dnn.add(o5, o3)
    same_tensor(e, o6, r6, "Add across levels")
    # values: plain model of the first two layers on the strided grid
    y = conv_plain(img, w1, 1)
    a = bn[0] / np.sqrt(bn[3] + 1e-5)
    y = y * a + (bn[1] - a * bn[2])
    # pixel (i, j) of the stride-2 output lives in slot ((i*W + j) << 1) of the fine W x W grid: read the whole grid
    # (descriptor of the unstrided tensor) and compare on the kept positions, |err| < 2^-20
    raw = dec(e, e.fhe.TensorCT(o2.cts, W, 0, 3))
    assert np.abs(raw[::2, ::2, :] - y[::2, ::2, :]).max() < 2.0 ** -20


def test_relu_variants_and_sign(env):
    e = env
    L = e.L
    img = image(21, ch=2, lo=-0.9, hi=0.9)
    rt = e.rd.enc_tensor(img)
    ot = ours_tensor(e, rt)
    same_tensor(e, e.dnn.relu(ot, -1, 1, 7), e.rd.t(L.rref_dnn_relu, rt, -1, 1, 7), "Relu deg 7")
    same_tensor(e, e.dnn.relu_deg_two(ot), e.rd.t(L.rref_dnn_relu_deg_two, rt), "ReluDegTwo")
    for k in (0, 1, 2):
        same_tensor(e, e.dnn.sign(ot, k), e.rd.t(L.rref_dnn_sign, rt, k), "Sign stage %d" % k)


def test_avgpool_fullcon(env):
    e = env
    rng = np.random.default_rng(31)
    img = image(32, ch=2)
    w1 = rng.uniform(-0.5, 0.5, (3, 3, 2, 4))
    rt = e.rd.enc_tensor(img)
    ot = ours_tensor(e, rt)
    r1, o1 = e.rd.conv(rt, w1, 2), e.dnn.conv(ot, w1, 2)       # width W/2, slotstr 1, 4 channels
    fc_w, fc_b = rng.uniform(-1, 1, (3, 4)), rng.uniform(-1, 1, 3)
    r2, o2 = e.rd.avgpool_fullcon(r1, fc_w, fc_b), e.dnn.avgpool_fullcon(o1, fc_w, fc_b)
    same_tensor(e, o2, r2, "AvgPoolFullCon")


def test_bootstrap_and_relu_composite(env):
    """BootStrap and ReluComposite (3 x Sign + 4 sparse bootstraps per channel) on a 2-channel tensor: bit-exact;
    decrypted |relu(x) - out| < 0.02 for |x| <= 1 (the reference's own approximation error of the composite sign
    polynomial plus four bootstraps: same bits on both sides)"""
    e = env
    L = e.L
    img = image(41, ch=2)
    rng = np.random.default_rng(42)
    wt = rng.uniform(-0.25, 0.25, (3, 3, 2, 2))
    rt = e.rd.enc_tensor(img)
    ot = ours_tensor(e, rt)
    r1, o1 = e.rd.conv(rt, wt, 1), e.dnn.conv(ot, wt, 1)
    r2 = e.rd.t(L.rref_dnn_relu_composite, r1)
    o2 = e.dnn.relu_composite(o1, e.boot)
    same_tensor(e, o2, r2, "ReluComposite")
    r3 = e.rd.t(L.rref_dnn_bootstrap, r2)
    o3 = e.dnn.bootstrap(o2, e.boot)
    same_tensor(e, o3, r3, "BootStrap")
    pre = conv_plain(img, wt, 1)
    err = np.abs(dec(e, o2) - np.maximum(pre, 0.0))
    for lo, hi in ((0, 0.25), (0.25, 0.5), (0.5, 1.0), (1.0, 2.0), (2.0, 10.0)):
        m = (np.abs(pre) >= lo) & (np.abs(pre) < hi)
        if m.any():
            print("ReluComposite |x| in [%g, %g): %d slots, max |err| = %.3g" % (lo, hi, m.sum(), err[m].max()))
    # relu(x) = x/2 (1 + sign(x/10)) followed by a bootstrap, whose EvalMod only covers messages of magnitude < 1
    # (a slot at 3.4 came back 0.37 off on BOTH sides in an earlier run of this test): tolerance 0.02 on |x| <= 1
    assert err[np.abs(pre) <= 1.0].max() < 0.02
    # a second convolution on the bootstrapped tensor (the ResNet block pattern)
    w2 = rng.uniform(-0.3, 0.3, (3, 3, 2, 2))
    same_tensor(e, e.dnn.conv(o2, w2, 1), e.rd.conv(r2, w2, 1), "Conv after ReluComposite")


@pytest.mark.parametrize("lanes,batch", [(0, 1), (0, 4), (3, 1), (3, 2)])
def test_resnet_small_network(env, lanes, batch):
    """ResNet20_infer (Resnet/models/model_resnet20.cu:44-153) on a scaled-down network (8 x 8 image, 2 input channels,
    base width 2, one block per stage, 3 classes: 7 ReluComposites = 90 sparse bootstraps): our driver against the same
    sequence of the reference's own layer calls -- every limb of the result bit-identical; the decrypted pooled
    activations agree with the plain numpy network (exact ReLU) within 0.05"""
    e = env
    L = e.L
    cfg = dict(in_ch=2, base_ch=2, blocks=1, classes=3)
    net, flat = resnet_model.synthetic_params(3, **cfg)
    assert flat.size == e.fhe.resnet20_param_count(**cfg)
    img = image(51, ch=2)
    rt = e.rd.enc_tensor(img)
    ot = ours_tensor(e, rt)
    e.dnn.set_lanes(e.boot, lanes)   # lanes > 0: the channels of every ReluComposite run on concurrent streams
    e.dnn.set_batch(batch)           # batch > 1: that many channels go through a ReluComposite as one batched ciphertext
    out, ms = e.fhe.resnet20_infer(e.dnn, e.boot, ot, flat, base_ch=2, blocks=1, classes=3, timings=True)
    e.dnn.set_lanes(e.boot, 0)
    e.dnn.set_batch(4)
    print("our layer times (ms), %d lanes, batch %d:" % (lanes, batch), ms)
    rd = e.rd
    relu = lambda t: rd.t(L.rref_dnn_relu_composite, t)   # noqa: E731
    x = relu(rd.batch_norm(rd.conv(rt, net["conv1"], 1), *net["bn1"]))
    for blk in net["blocks"]:
        skip = x
        y = relu(rd.batch_norm(rd.conv(x, blk["w1"], 2 if blk["down"] else 1), *blk["bn1"]))
        y = rd.batch_norm(rd.conv(y, blk["w2"], 1), *blk["bn2"])
        if blk["down"]:
            skip = rd.batch_norm(rd.conv(skip, blk["ws"], 2), *blk["bns"])
        x = relu(rd.t(L.rref_dnn_add, y, skip))
    rout = rd.avgpool_fullcon(x, net["fc_w"], net["fc_b"])
    same_tensor(e, out, rout, "ResNet20_infer")
    # values: the last feature map before pooling, on the kept positions of the fine grid
    feat, _ = resnet_model.forward_plain(net, img)
    ox = ours_tensor(e, x)
    raw = dec(e, e.fhe.TensorCT(ox.cts, W, 0, ox.num_ch))
    err = np.abs(raw[::4, ::4, :] - feat).max()
    print("last feature map: max |err| vs the plain network =", err)
    assert err < 0.05


def test_resnet_one_image_on_two_gpus():
    """ONE encrypted image on two GPUs (b200fhe_dnn_set_peer_ranks): the channels of every ReluComposite are dealt to the
    ranks and the results all-gathered over NVLink peer memory; the result ciphertexts of the scaled-down network must be the
    single-GPU ones (checksum over every limb) on both ranks.  Skipped on a box with one GPU."""
    import json
    import subprocess
    import torch
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    args = [os.path.join(root, "tools", "resnet_bench.py"), "--mode", "ours", "--one-image", "--logn", "12", "--width", "8", "--in-ch", "2",
            "--base-ch", "2", "--blocks", "1", "--classes", "3", "--lanes", "2", "--batch", "2"]

    def run(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
        assert r.returncode == 0, r.stderr[-2000:]
        return json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
    one = run([sys.executable] + args)
    two = run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
               "--master-port", "29651"] + args)
    assert two["n_gpus"] == 2 and two["same_limbs_on_every_rank"] is True
    assert two["output_checksum"] == one["output_checksum"]
    assert two["max_abs_err_vs_plain"] == one["max_abs_err_vs_plain"]

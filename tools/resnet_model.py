"""Synthetic ResNet-20 parameters in the flat order b200fhe_resnet20_infer reads (include/b200fhe.h) and a plain numpy
forward pass of the same network (exact ReLU) for value checks.  The reference's trained weights and CIFAR file are not
in its tree (Resnet/resnet.cu:71,100), hence random weights scaled so that activations stay inside the bootstrap's
message range |x| < 1."""
import numpy as np


def architecture(in_ch=3, base_ch=16, blocks=3):
    """-> list of blocks: (stage, block, cin, cout, downsample)"""
    out, ch = [], base_ch
    for stage in range(3):
        cout = ch if stage == 0 else 2 * ch
        for b in range(blocks):
            out.append((stage, b, ch if b == 0 else cout, cout, stage > 0 and b == 0))
        ch = cout
    return out


def synthetic_params(seed, in_ch=3, base_ch=16, blocks=3, classes=10, gain=0.6):
    rng = np.random.default_rng(seed)
    parts = []

    def conv(kh, cin, cout):
        a = gain * np.sqrt(3.0 / (kh * kh * cin))
        w = rng.uniform(-a, a, (kh, kh, cin, cout))
        parts.append(w)
        return w

    def bn(c):
        p = [rng.uniform(0.8, 1.2, c), rng.uniform(-0.05, 0.05, c), rng.uniform(-0.05, 0.05, c), rng.uniform(0.8, 1.2, c)]
        parts.extend(p)
        return p

    net = {"conv1": conv(3, in_ch, base_ch), "bn1": bn(base_ch), "blocks": []}
    for stage, b, cin, cout, down in architecture(in_ch, base_ch, blocks):
        blk = {"down": down, "w1": conv(3, cin, cout), "bn1": bn(cout), "w2": conv(3, cout, cout), "bn2": bn(cout)}
        if down:
            blk["ws"], blk["bns"] = conv(1, cin, cout), bn(cout)
        net["blocks"].append(blk)
    ch = base_ch * 4
    net["fc_w"] = rng.uniform(-1, 1, (classes, ch)) / ch
    net["fc_b"] = rng.uniform(-0.1, 0.1, classes)
    parts += [net["fc_w"], net["fc_b"]]
    flat = np.concatenate([np.asarray(p, dtype=np.float64).reshape(-1) for p in parts])
    return net, flat


def conv_plain(x, wt, stride):
    """DNN::Conv on the slots: 'same' correlation (zero padding), subsampled by `stride`"""
    w, kh = x.shape[0], wt.shape[0]
    half = kh // 2
    xp = np.zeros((w + 2 * half, w + 2 * half, x.shape[2]))
    xp[half:half + w, half:half + w, :] = x
    out = np.zeros((w, w, wt.shape[3]))
    for i in range(kh):
        for j in range(kh):
            out += xp[i:i + w, j:j + w, :] @ wt[i, j, :, :]
    return out[::stride, ::stride, :]


def bn_plain(x, p):
    a = p[0] / np.sqrt(p[3] + 1e-5)
    return x * a + (p[1] - a * p[2])


def forward_plain(net, img, track=None):
    """-> pooled-and-connected outputs as DNN::AvgPoolFullCon WOULD give them if it accumulated every row (sum pooling
    over the final grid, no division, as in src/dnn.cu:397-447); `track` collects max |activation| before each ReLU"""
    def relu(x):
        if track is not None:
            track.append(float(np.abs(x).max()))
        return np.maximum(x, 0.0)
    x = relu(bn_plain(conv_plain(img, net["conv1"], 1), net["bn1"]))
    for blk in net["blocks"]:
        skip = x
        y = relu(bn_plain(conv_plain(x, blk["w1"], 2 if blk["down"] else 1), blk["bn1"]))
        y = bn_plain(conv_plain(y, blk["w2"], 1), blk["bn2"])
        if blk["down"]:
            skip = bn_plain(conv_plain(skip, blk["ws"], 2), blk["bns"])
        x = relu(y + skip)
    pooled = x.sum(axis=(0, 1))
    return x, pooled

"""Multi-GPU sharding of the path (host side; one process per GPU, torch.distributed for the plumbing only).

The path shards by independent units — ciphertexts of a batch (bootstraps, HMults, NTT polynomial batches) — with the
context, evaluation keys and plaintext diagonals replicated on every GPU and NO data-path collective (DESIGN.md 6,
SURVEY.md 8e mode 1).  The second mode spreads the RNS limbs of ONE ciphertext's key switch over the GPUs
(`LimbShardedKeySwitch`): per-limb work stays local, the rows a base conversion needs from the other ranks are exchanged
over NCCL / NVLink at the two conversion boundaries.  Timings of a sharded job are device times reduced with MAX over
the ranks."""
import ctypes as C


def shard_units(n_units, rank, world):
    """contiguous, balanced block of unit indices for `rank`: sizes differ by at most one, every unit exactly once"""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad rank/world")
    base, extra = divmod(n_units, world)
    start = rank * base + min(rank, extra)
    return range(start, start + base + (1 if rank < extra else 0))


def limb_ranges(n_limbs, world):
    """contiguous limb ownership [(lo, hi)] per rank for the limb-sharded mode (the kernels take limb ranges)"""
    out = []
    for r in range(world):
        u = shard_units(n_limbs, r, world)
        out.append((u.start, u.stop))
    return out


def exchange_rows(dist, rows, ranges):
    """all-gather of row blocks in place: block ranges[r] = (lo, hi) of the 2-D contiguous tensor `rows` is owned by
    rank r and is broadcast from it (blocks may differ in size, which a plain NCCL all-gather does not take)"""
    if dist is None or dist.get_world_size() == 1:
        return
    sizes = {hi - lo for lo, hi in ranges}
    if len(sizes) == 1 and ranges[0][0] == 0 and ranges[-1][1] == rows.shape[0] and dist.get_backend() == "nccl":
        # equal blocks covering the tensor: ONE in-place all-gather (ncclAllGather with sendbuff = recvbuff + rank * count)
        lo, hi = ranges[dist.get_rank()]
        dist.all_gather_into_tensor(rows, rows[lo:hi])
        return
    me, world = dist.get_rank(), dist.get_world_size()
    if dist.get_backend() == "nccl":
        # uneven blocks: every rank sends its block to every other rank and receives theirs, as ONE NCCL group
        mlo, mhi = ranges[me]
        ops = []
        for r, (lo, hi) in enumerate(ranges):
            if r == me:
                continue
            if mhi > mlo:
                ops.append(dist.P2POp(dist.isend, rows[mlo:mhi], r))
            if hi > lo:
                ops.append(dist.P2POp(dist.irecv, rows[lo:hi], r))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        return
    for r, (lo, hi) in enumerate(ranges):
        if hi > lo:
            dist.broadcast(rows[lo:hi], src=r)


class LimbShardedKeySwitch:
    """One key switch spread over `world` GPUs by RNS limb (SURVEY.md 8e mode 2; include/b200fhe.h b200fhe_shard_ks_*).
    Rank g owns ciphertext limbs q[g] of Q_l and special limbs p[g] of P.  Buffers keep the single-GPU layout; the rows
    a base conversion needs from the other ranks are exchanged with one broadcast per owner (rows are contiguous blocks,
    possibly of different sizes).  dist = None runs every rank's share one after the other on this GPU ("virtual
    ranks": same arithmetic, no exchange needed because the buffers are shared) -- used by the single-GPU parity test."""

    def __init__(self, ctx, size_Ql, world=1, rank=0, dist=None):
        import torch
        self.ctx, self.Ql, self.world, self.rank, self.dist = ctx, size_Ql, world, rank, dist
        self.P, self.N = ctx.size_P, ctx.N
        self.q, self.p = limb_ranges(size_Ql, world), limb_ranges(self.P, world)
        self.tcks = torch.zeros((size_Ql, self.N), dtype=torch.int64, device="cuda")
        self.cx = torch.zeros((2, size_Ql + self.P, self.N), dtype=torch.int64, device="cuda")

    def _ranks(self):
        return range(self.world) if self.dist is None else [self.rank]

    def _exchange(self, rows, ranges):
        if self.dist is not None and self.world > 1:
            exchange_rows(self.dist, rows, ranges)

    def run(self, c2, evk, acc=None, acc_polys=0, out=None, gather_out=True):
        """c2: [Ql][N] (NTT form; own rows valid), evk: digit tensors, acc: [2][Ql][N] or None -> out [2][Ql][N]"""
        import torch
        from . import lib, _check, _ptr, _stream
        L, Ql = lib(), self.Ql
        if out is None:
            out = torch.zeros((2, Ql, self.N), dtype=torch.int64, device="cuda")
        ptrs = (C.c_void_p * len(evk))(*[d if isinstance(d, int) else d.data_ptr() for d in evk])
        st = _stream()
        for g in self._ranks():
            _check(L.b200fhe_shard_ks_begin(self.ctx.h, Ql, self.q[g][0], self.q[g][1], _ptr(c2), _ptr(self.tcks), st))
        self._exchange(self.tcks, self.q)
        for g in self._ranks():
            _check(L.b200fhe_shard_ks_middle(self.ctx.h, Ql, self.q[g][0], self.q[g][1], self.p[g][0], self.p[g][1],
                                             _ptr(self.tcks), _ptr(c2), ptrs, _ptr(self.cx), st))
        for poly in (0, 1):
            self._exchange(self.cx[poly, Ql:], self.p)
        for g in self._ranks():
            _check(L.b200fhe_shard_ks_end(self.ctx.h, Ql, self.q[g][0], self.q[g][1], _ptr(self.cx),
                                          _ptr(acc) if acc is not None else None, acc_polys, _ptr(out), st))
        if gather_out:
            for poly in (0, 1):
                self._exchange(out[poly], self.q)
        return out


class PeerLimbShardedKeySwitch(LimbShardedKeySwitch):
    """LimbShardedKeySwitch whose row exchanges go over NVLink peer memory (b200fhe_peer_exchange) instead of NCCL: tcks,
    cx and the result sit at fixed offsets of ONE symmetric arena (torch.distributed._symmetric_memory maps every peer's
    arena into this process); at each boundary one kernel stores the own rows into the peers' arenas and doubles as
    the inter-GPU barrier.  No collective call, no host synchronisation on the data path, and the whole key switch can be
    captured in a CUDA graph (the exchanges are numbered on the device).  run() returns a view of the arena's result
    block (valid until the next run)."""

    def __init__(self, ctx, size_Ql, world, rank, dist, group=None, fused=False, multicast=False):
        """fused: the inverse transforms that produce the exchanged rows store them into the peers' arenas themselves
        (b200fhe_shard_ks_*_mirrored) and the boundary kernel only exchanges flags; False (default): a separate store
        kernel per boundary.  Measured equal at 2 GPUs and 4 % slower at 4 (the 8-byte column stores of the transform's
        last pass are poor NVLink writes: the pass grows by what the store kernel took), hence not the default."""
        import torch
        import torch.distributed._symmetric_memory as symm
        self.fused = bool(fused) and ctx.N >= 4096
        self.ctx, self.Ql, self.world, self.rank, self.dist = ctx, size_Ql, world, rank, dist
        self.P, self.N = ctx.size_P, ctx.N
        self.q, self.p = limb_ranges(size_Ql, world), limb_ranges(self.P, world)
        Ql, QlP, N = size_Ql, size_Ql + self.P, self.N
        self._off = {"tcks": 0, "cx": Ql * N, "out": Ql * N + 2 * QlP * N, "flags": Ql * N + 2 * QlP * N + 2 * Ql * N}
        words = self._off["flags"] + 16
        self.arena = symm.empty(words, dtype=torch.int64, device=torch.device("cuda", torch.cuda.current_device()))
        self.arena.zero_()
        self.hdl = symm.rendezvous(self.arena, group if group is not None else dist.group.WORLD)
        bases = [int(b) for b in self.hdl.buffer_ptrs]
        if len(bases) != world or bases[rank] != self.arena.data_ptr():
            raise RuntimeError("symmetric memory rendezvous returned unexpected peer pointers")
        self._bases = (C.c_void_p * world)(*bases)
        self._flags = (C.c_void_p * world)(*[b + 8 * self._off["flags"] for b in bases])
        others = [g for g in range(world) if g != rank]
        self._peer_tcks = (C.c_void_p * max(1, len(others)))(*[bases[g] + 8 * self._off["tcks"] for g in others])
        self._peer_cx = (C.c_void_p * max(1, len(others)))(*[bases[g] + 8 * self._off["cx"] for g in others])
        self._npeers = len(others)
        # multicast: the rows go ONCE to the arena's NVSwitch multicast address and the switch replicates them
        # (measured: 2 % faster at 4 GPUs, 6 % slower at 2 -- the own copy comes back through the switch as well)
        self._mc = 0
        if multicast:
            try:
                self._mc = int(self.hdl.multicast_ptr or 0)
            except Exception:
                self._mc = 0
        self.multicast = bool(self._mc)
        self._counter = torch.zeros(4, dtype=torch.int32, device="cuda")
        o = self._off
        self.tcks = self.arena[o["tcks"]:o["cx"]].view(Ql, N)
        self.cx = self.arena[o["cx"]:o["out"]].view(2, QlP, N)
        self.out = self.arena[o["out"]:o["flags"]].view(2, Ql, N)
        torch.cuda.synchronize()
        dist.barrier()        # every arena zeroed before anybody's first flag arrives

    def _put(self, segs, st):
        from . import lib, _check
        n = len(segs)
        off = (C.c_size_t * max(n, 1))(*[8 * o for o, _ in segs])
        ln = (C.c_size_t * max(n, 1))(*[8 * w for _, w in segs])
        _check(lib().b200fhe_peer_exchange_mc(self.arena.data_ptr(), self._bases, C.c_void_p(self._mc), self._flags,
                                              self._counter.data_ptr(), self.world, self.rank, off, ln, n, 0, st))

    def run(self, c2, evk, acc=None, acc_polys=0, out=None, gather_out=True, mark=None):
        """mark: optional callable(name) invoked after each phase has been enqueued (phase timing with CUDA events)"""
        from . import lib, _check, _ptr, _stream
        L, Ql, QlP, N, g, o = lib(), self.Ql, self.Ql + self.P, self.N, self.rank, self._off
        mark = mark or (lambda name: None)
        (q0, q1), (p0, p1) = self.q[g], self.p[g]
        ptrs = (C.c_void_p * len(evk))(*[d if isinstance(d, int) else d.data_ptr() for d in evk])
        st = _stream()
        if self.fused:
            _check(L.b200fhe_shard_ks_begin_mirrored(self.ctx.h, Ql, q0, q1, _ptr(c2), _ptr(self.tcks), self._peer_tcks,
                                                     self._npeers, st))
            mark("begin")
            self._put([], st)
            mark("exchange_tcks")
            _check(L.b200fhe_shard_ks_middle_mirrored(self.ctx.h, Ql, q0, q1, p0, p1, _ptr(self.tcks), _ptr(c2), ptrs,
                                                      _ptr(self.cx), self._peer_cx, self._npeers, st))
            mark("middle")
            self._put([], st)
            mark("exchange_cx")
        else:
            _check(L.b200fhe_shard_ks_begin(self.ctx.h, Ql, q0, q1, _ptr(c2), _ptr(self.tcks), st))
            mark("begin")
            self._put([(o["tcks"] + q0 * N, (q1 - q0) * N)], st)
            mark("exchange_tcks")
            _check(L.b200fhe_shard_ks_middle(self.ctx.h, Ql, q0, q1, p0, p1, _ptr(self.tcks), _ptr(c2), ptrs, _ptr(self.cx), st))
            mark("middle")
            self._put([(o["cx"] + (poly * QlP + Ql + p0) * N, (p1 - p0) * N) for poly in (0, 1)], st)
            mark("exchange_cx")
        _check(L.b200fhe_shard_ks_end(self.ctx.h, Ql, q0, q1, _ptr(self.cx), _ptr(acc) if acc is not None else None, acc_polys,
                                      _ptr(self.out), st))
        mark("end")
        if gather_out:
            self._put([(o["out"] + (poly * Ql + q0) * N, (q1 - q0) * N) for poly in (0, 1)], st)
            mark("exchange_out")
        if out is not None:
            out.copy_(self.out)
            return out
        return self.out


class PeerShardedBootstrap:
    """ONE bootstrap on `world` GPUs: every rank runs Bootstrapper.eval on the same ciphertext; the hoisted rotations and the
    plaintext-matrix product of every CoeffsToSlots / SlotsToCoeffs stage compute the rank's own limb rows only and the
    inner sums are exchanged once per stage over NVLink peer memory (b200fhe_boot_set_peer_shard).  This class allocates
    the symmetric arena (torch.distributed._symmetric_memory maps every peer's arena into this process) and hands the
    mapped bases to the bootstrapper.  No NCCL call on the data path."""

    def __init__(self, boot, world, rank, dist, group=None):
        import torch
        import torch.distributed._symmetric_memory as symm
        self.boot, self.world, self.rank = boot, world, rank
        nbytes = boot.peer_arena_bytes()
        self.arena = symm.empty((nbytes + 7) // 8, dtype=torch.int64, device=torch.device("cuda", torch.cuda.current_device()))
        self.arena.zero_()
        self.hdl = symm.rendezvous(self.arena, group if group is not None else dist.group.WORLD)
        bases = [int(b) for b in self.hdl.buffer_ptrs]
        if len(bases) != world or bases[rank] != self.arena.data_ptr():
            raise RuntimeError("symmetric memory rendezvous returned unexpected peer pointers")
        self._counter = torch.zeros(4, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        dist.barrier()        # every arena zeroed before anybody's first flag arrives
        boot.set_peer_shard(world, rank, bases, nbytes, self._counter.data_ptr())

    def close(self):
        import torch
        torch.cuda.synchronize()
        self.boot.set_peer_shard(1)
        if int(self._counter[1].item()) != 0:
            raise RuntimeError("a peer exchange of the limb-sharded bootstrap timed out (a rank did not arrive)")


class PeerShardedDNN:
    """ONE encrypted image on `world` GPUs: every rank runs the same network on the same ciphertexts with the same keys; the
    channels of ReluComposite / BootStrap are dealt to the ranks and the results all-gathered over NVLink peer memory, and a
    convolution's input channels are dealt to the ranks and the partial sums reduced (b200fhe_dnn_set_peer_ranks).
    half_bytes: twice the widest layer's results, 2 x channels x 2 x |Q_l| x N x 8 (partial-sum slots + sums)."""

    def __init__(self, dnn, world, rank, dist, half_bytes, group=None):
        import torch
        import torch.distributed._symmetric_memory as symm
        self.dnn = dnn
        half_bytes = (half_bytes + 15) // 16 * 16
        nbytes = 2 * half_bytes + 256
        self.arena = symm.empty(nbytes // 8, dtype=torch.int64, device=torch.device("cuda", torch.cuda.current_device()))
        self.arena.zero_()
        self.hdl = symm.rendezvous(self.arena, group if group is not None else dist.group.WORLD)
        bases = [int(b) for b in self.hdl.buffer_ptrs]
        if len(bases) != world or bases[rank] != self.arena.data_ptr():
            raise RuntimeError("symmetric memory rendezvous returned unexpected peer pointers")
        self._counter = torch.zeros(4, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        dist.barrier()        # every arena zeroed before anybody's first flag arrives
        dnn.set_peer_ranks(world, rank, bases, nbytes, self._counter.data_ptr())

    def close(self):
        import torch
        torch.cuda.synchronize()
        self.dnn.set_peer_ranks(1)
        if int(self._counter[1].item()) != 0:
            raise RuntimeError("a peer exchange of the rank-sharded layers timed out (a rank did not arrive)")


def max_over_ranks(value, dist=None, device=None):
    """MAX-reduce a per-rank scalar (a device-side duration in ms); identity without a process group"""
    if dist is None or not dist.is_available() or not dist.is_initialized() or dist.get_world_size() == 1:
        return float(value)
    import torch
    t = torch.tensor([float(value)], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def aggregate_throughput(units_per_rank, ms, world):
    """whole-job units/s when every rank processed `units_per_rank` units in `ms` (already the max over ranks)"""
    return world * units_per_rank / (ms * 1e-3)

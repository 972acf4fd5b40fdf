"""Multi-GPU sharding of the path (host side; one process per GPU, torch.distributed for the plumbing only).

The path shards by independent units — ciphertexts of a batch (bootstraps, HMults, NTT polynomial batches) — with the
context, evaluation keys and plaintext diagonals replicated on every GPU and NO data-path collective (DESIGN.md 6,
SURVEY.md 8e mode 1).  `limb_ranges` describes the second mode (RNS limbs of one ciphertext spread over the GPUs,
an all-gather of the alpha-limb digit at each base conversion); it is the partition the limb-sharded mode will use
and is only bookkeeping here.  Timings of a sharded job are device times reduced with MAX over the ranks."""


def shard_units(n_units, rank, world):
    """contiguous, balanced block of unit indices for `rank`: sizes differ by at most one, every unit exactly once"""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad rank/world")
    base, extra = divmod(n_units, world)
    start = rank * base + min(rank, extra)
    return range(start, start + base + (1 if rank < extra else 0))


def limb_ranges(size_QlP, world):
    """round-robin limb ownership for the limb-sharded mode: limb t of a Q_l P buffer lives on rank t % world
    (keeps every digit's alpha limbs spread over all GPUs, so each base conversion all-gathers equal pieces)"""
    return [list(range(r, size_QlP, world)) for r in range(world)]


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

"""Multi-GPU: independent-sequence sharding.

Tracking is causal within a sequence and sequences share nothing but read-only weights, so sequence s goes to
rank s mod world_size, one process per GPU, and there is NO collective on the data path.  The only exchange is
a host-side gather of each sequence's (T, 3, 3) float64 homographies (and a max over ranks of the elapsed time
for reporting), done with torch.distributed object collectives on whatever backend the job was launched with.
The reference itself has no distributed code at all (grep for distributed|nccl|all_reduce: no hits).
"""
import torch
import torch.distributed as dist


def shard_sequences(n_sequences, rank, world_size):
    """Indices of the sequences rank `rank` tracks: s = rank, rank + world, ..."""
    return list(range(rank, n_sequences, world_size))


def _ready():
    return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1


def gather_results(per_rank):
    """dict {sequence index: array} on every rank -> merged dict on rank 0 (None elsewhere)."""
    if not _ready():
        return dict(per_rank)
    bucket = [None] * dist.get_world_size() if dist.get_rank() == 0 else None
    dist.gather_object(per_rank, bucket, dst=0)
    if dist.get_rank() != 0:
        return None
    merged = {}
    for part in bucket:
        overlap = set(merged) & set(part)
        assert not overlap, f"sequences {sorted(overlap)} were tracked by more than one rank"
        merged.update(part)
    return merged


def max_over_ranks(value, device=None):
    """Largest `value` (e.g. device-timed seconds) over all ranks, on every rank."""
    if not _ready():
        return float(value)
    dev = device if device is not None else ("cuda" if dist.get_backend() == "nccl" else "cpu")
    t = torch.tensor([float(value)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


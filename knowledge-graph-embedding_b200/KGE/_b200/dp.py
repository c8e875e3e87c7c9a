"""Host side of multi-GPU runs: who owns which positives / entities, and the
one collective of a data-parallel step.  torch.distributed is the plumbing (NCCL on GPUs; the
same code runs over gloo on CPU tensors in tests/test_dp_gloo.py).

  * evaluation: every rank scores all queries against its own slice of the entity table
    (`entity_shard`) and the int32 "better candidate" counts are summed (`allreduce_counts`);
  * training: every rank holds all tables and `batch_shard` of the step's positives; the
    per-table summed gradients and the five scalars (norm^2 x 4, loss) live in ONE flat fp32
    buffer (`GradBuffer`) so a step needs a single all-reduce; each rank then applies the
    identical update, so replicas stay bit-identical.
"""
import torch
import torch.distributed as dist


def world():
    """(rank, world_size) of the default process group, (0, 1) when not initialised."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def split_even(n, parts, i):
    """[lo, hi) of part i when n items are split into `parts` contiguous near-equal ranges
    (the first n % parts ranges get one more)."""
    if not (0 <= i < parts):
        raise ValueError(f"part {i} out of range [0, {parts})")
    q, r = divmod(int(n), int(parts))
    lo = i * q + min(i, r)
    return lo, lo + q + (1 if i < r else 0)


def entity_shard(n_ent, rank=None, world_size=None):
    r, w = world()
    return split_even(n_ent, w if world_size is None else world_size, r if rank is None else rank)


def batch_shard(B_global, rank=None, world_size=None):
    """Positives [i_base, i_base + B) of a global batch owned by `rank`."""
    r, w = world()
    return split_even(B_global, w if world_size is None else world_size, r if rank is None else rank)


def allreduce_counts(counts, group=None):
    """Sum of the per-shard integer counts; exact and order independent."""
    if world()[1] > 1:
        dist.all_reduce(counts, op=dist.ReduceOp.SUM, group=group)
    return counts


class GradBuffer:
    """One flat fp32 buffer [table 0 | table 1 | table 2 | table 3 | scal(8)].
    `tables` = {slot: (rows, ld)}; `view(slot)` is that table's (rows, ld) window and `scal`
    the trailing 8 floats (norm^2 of slots 0..3, loss share, 3 unused)."""

    N_SCAL = 8

    def __init__(self, tables, device):
        self.offsets = {}
        o = 0
        for s in sorted(tables):
            rows, ld = tables[s]
            if ld % 4:
                raise ValueError("row stride must be a multiple of 4 floats")
            self.offsets[s] = (o, rows, ld)
            o += rows * ld
        self.scal_off = o
        self.flat = torch.zeros(o + self.N_SCAL, dtype=torch.float32, device=device)

    def view(self, slot):
        o, rows, ld = self.offsets[slot]
        return self.flat[o:o + rows * ld].view(rows, ld)

    @property
    def scal(self):
        return self.flat[self.scal_off:self.scal_off + self.N_SCAL]

    def allreduce(self, group=None):
        """The step's single collective.  NCCL / gloo sum in a rank-independent order, so
        every rank ends with the same bits."""
        if world()[1] > 1:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group)
        return self

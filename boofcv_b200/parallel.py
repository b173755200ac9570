"""Multi-GPU use of the SURF path: one process per GPU, torch.distributed for the plumbing.

Detect/describe shards by frame: frames are independent, every rank runs its own frames, there is NO collective on the
data path (`frame_shard`).  Greedy association shards the source rows: every rank scores its rows against all destination
rows; backward validation (AssociateGreedy.java:98-117) needs the minimum of every destination column over ALL source
rows, which is the one real exchange step: an all-gather of three M-entry arrays per rank, merged on the device.

The compute steps are the C-ABI calls bs_assoc_forward_device / bs_assoc_merge_columns_device / bs_assoc_validate_device
(`GpuAssocBackend`).  `ShardedAssociate` only orchestrates, so that the same code runs over NCCL with CUDA tensors and, in
the CPU tests, over gloo with a checker backend supplied by the test.
"""
import ctypes as C

import numpy as np

from . import _native as N


def shard_range(total, rank, world):
    """Contiguous balanced split of range(total): the first total % world ranks get one extra item."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad rank/world")
    base, extra = divmod(int(total), world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def frame_shard(n_frames, rank, world):
    """Frames [begin, end) of a batch that `rank` detects and describes."""
    return shard_range(n_frames, rank, world)


class GpuAssocBackend:
    """The three device steps of a sharded association, on torch CUDA tensors, through the C ABI."""

    def __init__(self, assoc):
        self.assoc = assoc                      # AssociateGreedyGPU

    def _stream(self, t):
        import torch
        return C.c_void_p(torch.cuda.current_stream(t.device).cuda_stream or 1)

    def forward(self, src, row_offset, dst):
        import torch
        n, dof = src.shape; m = dst.shape[0]
        dev = src.device
        pairs = torch.empty(n, dtype=torch.int32, device=dev); fit = torch.empty(n, dtype=torch.float64, device=dev)
        cmin = torch.empty(m, dtype=torch.float64, device=dev)
        cfirst = torch.empty(m, dtype=torch.int32, device=dev); clast = torch.empty(m, dtype=torch.int32, device=dev)
        p = lambda t: C.c_void_p(t.data_ptr())
        N.check(N.lib().bs_assoc_forward_device(self.assoc._h, p(src), n, int(row_offset), p(dst), m, dof, p(pairs), p(fit), p(cmin),
                                                p(cfirst), p(clast), self._stream(src)))
        return pairs, fit, cmin, cfirst, clast

    def merge(self, all_min, all_first, all_last):
        import torch
        ranks, m = all_min.shape
        dev = all_min.device
        cmin = torch.empty(m, dtype=torch.float64, device=dev)
        cfirst = torch.empty(m, dtype=torch.int32, device=dev); clast = torch.empty(m, dtype=torch.int32, device=dev)
        p = lambda t: C.c_void_p(t.data_ptr())
        N.check(N.lib().bs_assoc_merge_columns_device(self.assoc._h, ranks, m, p(all_min), p(all_first), p(all_last), p(cmin), p(cfirst),
                                                      p(clast), self._stream(all_min)))
        return cmin, cfirst, clast

    def validate(self, pairs, fit, row_offset, cmin, cfirst, clast):
        p = lambda t: C.c_void_p(t.data_ptr())
        N.check(N.lib().bs_assoc_validate_device(self.assoc._h, pairs.shape[0], int(row_offset), p(pairs), p(fit), p(cmin), p(cfirst),
                                                 p(clast), self._stream(pairs)))
        return pairs, fit


class ShardedAssociate:
    """AssociateGreedy over source rows split across the ranks of `group`.

    associate(src_local, dst) -> (pairs, fit) for the local rows; pairs index the full dst list, exactly as a
    single-process AssociateGreedy over the concatenated source list would produce them."""

    def __init__(self, backend, backwards_validation, group=None):
        self.backend = backend
        self.backwards = bool(backwards_validation)
        self.group = group

    def associate(self, src_local, dst):
        import torch
        import torch.distributed as dist
        world = dist.get_world_size(self.group) if dist.is_initialized() else 1
        n_local = int(src_local.shape[0])
        # global numbering of the local rows: exclusive prefix of the shard sizes
        if world > 1:
            sizes = [torch.zeros(1, dtype=torch.int64, device=src_local.device) for _ in range(world)]
            dist.all_gather(sizes, torch.tensor([n_local], dtype=torch.int64, device=src_local.device), group=self.group)
            rank = dist.get_rank(self.group)
            row_offset = int(sum(int(s.item()) for s in sizes[:rank]))
        else:
            row_offset = 0
        pairs, fit, cmin, cfirst, clast = self.backend.forward(src_local, row_offset, dst)
        if not self.backwards:
            return pairs, fit
        if world > 1:
            # the exchange step: 16 bytes per destination column and rank
            g_min = [torch.empty_like(cmin) for _ in range(world)]
            g_first = [torch.empty_like(cfirst) for _ in range(world)]
            g_last = [torch.empty_like(clast) for _ in range(world)]
            dist.all_gather(g_min, cmin, group=self.group)
            dist.all_gather(g_first, cfirst, group=self.group)
            dist.all_gather(g_last, clast, group=self.group)
            cmin, cfirst, clast = self.backend.merge(torch.stack(g_min), torch.stack(g_first), torch.stack(g_last))
        return self.backend.validate(pairs, fit, row_offset, cmin, cfirst, clast)


def gather_matches(pairs, fit, group=None):
    """All ranks' (pairs, fit) in global source order, as numpy arrays on every rank (small: 12 bytes per source row)."""
    import torch
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return pairs.cpu().numpy(), fit.cpu().numpy()
    out = [None] * dist.get_world_size(group)
    dist.all_gather_object(out, (pairs.cpu().numpy(), fit.cpu().numpy()), group=group)
    return np.concatenate([o[0] for o in out]), np.concatenate([o[1] for o in out])

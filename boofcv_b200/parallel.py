"""Multi-GPU use of the SURF path: one process per GPU, torch.distributed for the plumbing.

Detect/describe shards by frame: frames are independent, every rank runs its own frames, there is NO collective on the
data path (`frame_shard`).  Greedy association shards the source rows: every rank scores its rows against all destination
rows; backward validation (AssociateGreedy.java:98-117) needs the minimum of every destination column over ALL source
rows, which is the one real exchange step: ONE all-gather of a 16*M-byte record per rank, merged on the device.

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
        self._packed = None                     # the 16*m-byte record the last forward() wrote its column partials into

    def _stream(self, t):
        import torch
        return C.c_void_p(torch.cuda.current_stream(t.device).cuda_stream or 1)

    @staticmethod
    def _check(t, name, dtype):
        import torch
        if t.dtype != dtype or not t.is_contiguous() or not t.is_cuda:
            raise ValueError("%s must be a contiguous CUDA tensor of %s (got %s, contiguous=%s, device=%s)"
                             % (name, dtype, t.dtype, t.is_contiguous(), t.device))

    def forward(self, src, row_offset, dst, want_columns=True):
        """pairs, fit of the local rows + (want_columns) the column partials over them, written into ONE 16*m-byte record
        [col_min f64 | col_first i32 | col_last i32] so that a single all-gather moves them (`packed()`)."""
        import torch
        self._check(src, "src", torch.float64); self._check(dst, "dst", torch.float64)
        if src.device != dst.device:
            raise ValueError("src and dst live on different devices")
        n, dof = src.shape; m = dst.shape[0]
        dev = src.device
        pairs = torch.full((n,), -1, dtype=torch.int32, device=dev)
        fit = torch.full((n,), self.assoc.maxFitError(), dtype=torch.float64, device=dev)
        rec = torch.empty(16 * m, dtype=torch.uint8, device=dev)
        cmin = rec[:8 * m].view(torch.float64); cfirst = rec[8 * m:12 * m].view(torch.int32); clast = rec[12 * m:].view(torch.int32)
        self._packed = rec
        if n == 0 or m == 0:
            # an empty shard (more ranks than rows, an empty frame) contributes nothing: no minimum, no row (AssociateGreedy
            # leaves pairs = -1 and the running best untouched); it must still take part in the collectives
            cmin.fill_(float("inf")); cfirst.fill_(-1); clast.fill_(-1)
            return pairs, fit, cmin, cfirst, clast
        p = lambda t: C.c_void_p(t.data_ptr())
        if want_columns:
            N.check(N.lib().bs_assoc_forward_device(self.assoc._h, p(src), n, int(row_offset), p(dst), m, dof, p(pairs), p(fit), p(cmin),
                                                    p(cfirst), p(clast), self._stream(src)))
        else:                                   # no backward validation: the plain forward association, no column pass
            N.check(N.lib().bs_assoc_forward_only_device(self.assoc._h, p(src), n, p(dst), m, dof, p(pairs), p(fit), self._stream(src)))
        return pairs, fit, cmin, cfirst, clast

    def packed(self):
        return self._packed

    def merge(self, all_min, all_first, all_last):
        import torch
        ranks, m = all_min.shape
        dev = all_min.device
        cmin = torch.empty(m, dtype=torch.float64, device=dev)
        cfirst = torch.empty(m, dtype=torch.int32, device=dev); clast = torch.empty(m, dtype=torch.int32, device=dev)
        p = lambda t: C.c_void_p(t.data_ptr())
        N.check(N.lib().bs_assoc_merge_columns_device(self.assoc._h, ranks, m, p(all_min.contiguous()), p(all_first.contiguous()),
                                                      p(all_last.contiguous()), p(cmin), p(cfirst), p(clast), self._stream(all_min)))
        return cmin, cfirst, clast

    def merge_packed(self, gathered, m):
        """gathered: [ranks, 16*m] uint8, one record per rank as forward() writes it (what all_gather_into_tensor returns)."""
        import torch
        ranks = gathered.shape[0]
        dev = gathered.device
        cmin = torch.empty(m, dtype=torch.float64, device=dev)
        cfirst = torch.empty(m, dtype=torch.int32, device=dev); clast = torch.empty(m, dtype=torch.int32, device=dev)
        p = lambda t: C.c_void_p(t.data_ptr())
        N.check(N.lib().bs_assoc_merge_packed_device(self.assoc._h, ranks, m, p(gathered), p(cmin), p(cfirst), p(clast), self._stream(gathered)))
        return cmin, cfirst, clast

    def validate(self, pairs, fit, row_offset, cmin, cfirst, clast):
        if pairs.shape[0] == 0:
            return pairs, fit
        p = lambda t: C.c_void_p(t.data_ptr())
        N.check(N.lib().bs_assoc_validate_device(self.assoc._h, pairs.shape[0], int(row_offset), p(pairs), p(fit), p(cmin), p(cfirst),
                                                 p(clast), self._stream(pairs)))
        return pairs, fit


class ShardedAssociate:
    """AssociateGreedy over source rows split across the ranks of `group`.

    associate(src_local, dst) -> (pairs, fit) for the local rows; pairs index the full dst list, exactly as a
    single-process AssociateGreedy over the concatenated source list would produce them.

    The one exchange step is ONE all-gather of a 16*m-byte record per rank (column minimum, first and last global row
    attaining it); `row_offset` (the global index of the first local row) saves the second, tiny all-gather of the shard
    sizes when the caller knows it (shard_range gives it)."""

    def __init__(self, backend, backwards_validation, group=None):
        self.backend = backend
        self.backwards = bool(backwards_validation)
        self.group = group

    def associate(self, src_local, dst, row_offset=None):
        import torch
        import torch.distributed as dist
        world = dist.get_world_size(self.group) if dist.is_initialized() else 1
        n_local = int(src_local.shape[0])
        if row_offset is None:
            row_offset = 0
            if world > 1:                       # global numbering of the local rows: exclusive prefix of the shard sizes
                sizes = torch.empty(world, dtype=torch.int64, device=src_local.device)
                dist.all_gather_into_tensor(sizes, torch.tensor([n_local], dtype=torch.int64, device=src_local.device), group=self.group)
                rank = dist.get_rank(self.group)
                row_offset = int(sizes[:rank].sum().item()) if rank > 0 else 0
        if not self.backwards:
            try:
                pairs, fit, *_ = self.backend.forward(src_local, row_offset, dst, want_columns=False)
            except TypeError:                   # a backend without the switch
                pairs, fit, *_ = self.backend.forward(src_local, row_offset, dst)
            return pairs, fit
        pairs, fit, cmin, cfirst, clast = self.backend.forward(src_local, row_offset, dst)
        if world > 1:
            m = int(cmin.shape[0])
            rec = self.backend.packed() if hasattr(self.backend, "packed") else None
            if rec is None:                     # a backend that keeps three arrays: pack them here
                rec = torch.cat([cmin.contiguous().view(torch.uint8), cfirst.contiguous().view(torch.uint8), clast.contiguous().view(torch.uint8)])
            flat = torch.empty(world * 16 * m, dtype=torch.uint8, device=rec.device)
            dist.all_gather_into_tensor(flat, rec, group=self.group)           # the exchange step: 16 bytes per column and rank
            gathered = flat.view(world, 16 * m)
            if hasattr(self.backend, "merge_packed"):
                cmin, cfirst, clast = self.backend.merge_packed(gathered, m)
            else:
                cmin, cfirst, clast = self.backend.merge(gathered[:, :8 * m].contiguous().view(torch.float64),
                                                         gathered[:, 8 * m:12 * m].contiguous().view(torch.int32),
                                                         gathered[:, 12 * m:].contiguous().view(torch.int32))
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

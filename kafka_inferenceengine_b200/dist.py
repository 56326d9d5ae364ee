"""Multi-GPU plumbing: one process per GPU, the image sharded by pixel.

Every matrix of the filter is block diagonal per pixel, so ranks never exchange
state.  The only coupling in the reference is the scalar convergence norm
``||x - x_prev||_2 / len(x)`` over ALL pixels (kafka/linear_kf.py:293): each rank
contributes its local sum of squares and element count to one 2-element
all-reduce per iteration (NCCL over NVLink on the GPU box, gloo in CPU tests),
after which every rank takes the same stop decision as a single process holding
the whole image.  Results are gathered (or written per rank) at the end.

The reference's own spatial partitioning is ``get_chunks``
(kafka/input_output/utils.py:12-40) + one ``LinearKalman`` per chunk
(kafka_test_Py36.py:241-255); there every chunk has its OWN norm, so chunked and
unchunked reference runs can stop at different iterations.  Sharding here keeps
the unchunked semantics.
"""
import numpy as np
import torch


def shard_rows(state_mask, rank, world_size):
    """Row range [r0, r1) of the raster owned by ``rank``: contiguous row bands
    with (nearly) equal numbers of True pixels.  Row-major compaction order is
    preserved, so concatenating the shards' compacted pixels in rank order gives
    the unsharded compaction."""
    mask = np.asarray(state_mask, dtype=bool)
    per_row = mask.sum(axis=1)
    cum = np.concatenate([[0], np.cumsum(per_row)])
    total = cum[-1]
    bounds = [0]
    for r in range(1, world_size):
        target = total * r / world_size
        bounds.append(int(np.searchsorted(cum, target, side="left")))
    bounds.append(mask.shape[0])
    bounds = np.maximum.accumulate(np.minimum(bounds, mask.shape[0]))
    return int(bounds[rank]), int(bounds[rank + 1])


def shard_pixels(n_pix, rank, world_size):
    """Compacted pixel range [i0, i1) of ``rank``: ceil(N/G) pixels per rank
    (SURVEY.md section 8e)."""
    per = -(-int(n_pix) // int(world_size))
    i0 = min(n_pix, rank * per)
    return i0, min(n_pix, i0 + per)


class NativeComm(object):
    """The multi-GPU step behind the C ABI (include/kafka_b200.h: kafka_ctx_create,
    kafka_allreduce_norm, kafka_gather_outputs): an NCCL communicator owned by
    libkafka_b200.so, one rank per GPU, usable by a binder that has no
    torch.distributed.  ``LinearKalman(comm=NativeComm(...))`` then takes its
    convergence all-reduce through it.

    ``NativeComm(rank, n_ranks, unique_id, device)`` is collective; ``unique_id``
    is the KAFKA_COMM_ID_BYTES blob rank 0 got from ``NativeComm.unique_id()`` and
    passed to the others by any means (``from_process_group`` uses
    torch.distributed for exactly that and nothing else)."""

    def __init__(self, rank, n_ranks, unique_id, device=None):
        import ctypes as C
        from . import _lib
        self._lib = _lib
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.rank, self.n_ranks = int(rank), int(n_ranks)
        if len(unique_id) != _lib.COMM_ID_BYTES:
            raise ValueError("unique_id must be %d bytes" % _lib.COMM_ID_BYTES)
        buf = C.create_string_buffer(bytes(unique_id), _lib.COMM_ID_BYTES)
        ctx = C.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(_lib.get_lib().kafka_ctx_create(self.rank, self.n_ranks, buf, C.byref(ctx)), "kafka_ctx_create")
        self._ctx = ctx

    @staticmethod
    def unique_id():
        import ctypes as C
        from . import _lib
        buf = C.create_string_buffer(_lib.COMM_ID_BYTES)
        _lib.check(_lib.get_lib().kafka_comm_unique_id(buf), "kafka_comm_unique_id")
        return buf.raw

    @classmethod
    def from_process_group(cls, group=None, device=None):
        """One NativeComm per rank of an initialised torch.distributed group (which only
        carries the unique id from rank 0 to the others)."""
        import torch.distributed as dist
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        box = [cls.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        return cls(rank, world, box[0], device)

    def all_reduce_sums(self, sums):
        """In-place SUM over ranks of a small contiguous float64 CUDA tensor, on the current stream."""
        if not (sums.is_cuda and sums.dtype == torch.float64 and sums.is_contiguous()):
            raise ValueError("all_reduce_sums needs a contiguous float64 CUDA tensor")
        with torch.cuda.device(self.device):
            self._lib.check(self._lib.get_lib().kafka_allreduce_norm(self._ctx, self._lib.ptr(sums), sums.numel(),
                                                                     self._lib.stream_ptr()), "kafka_allreduce_norm")
        return sums

    def gather(self, local, counts):
        """Concatenation in rank order of every rank's 1-D float64 CUDA tensor (``counts[r]``
        elements from rank r) on every rank."""
        import ctypes as C
        counts = [int(c) for c in counts]
        if len(counts) != self.n_ranks or counts[self.rank] != local.numel():
            raise ValueError("counts must list every rank's element count")
        local = local.contiguous()
        out = torch.empty(sum(counts), dtype=torch.float64, device=self.device)
        arr = (C.c_int64 * self.n_ranks)(*counts)
        with torch.cuda.device(self.device):
            self._lib.check(self._lib.get_lib().kafka_gather_outputs(self._ctx, self._lib.ptr(local), arr,
                                                                     self._lib.ptr(out), self._lib.stream_ptr()),
                            "kafka_gather_outputs")
        return out

    def close(self):
        if getattr(self, "_ctx", None) is not None and self._ctx.value:
            ctx, self._ctx = self._ctx, None
            self._lib.check(self._lib.get_lib().kafka_ctx_destroy(ctx), "kafka_ctx_destroy")


def global_norm(sum_sq_local, n_elems_local, comm=None, n_bad_local=None):
    """``(sqrt(sum of squares over all ranks) / total element count)`` and the
    total count.  ``sum_sq_local`` may be a python float or a 0-d/1-element
    tensor on any device.  With ``n_bad_local`` (same kinds; a tensor may hold
    several counters, e.g. [pixels left at the forecast, pixels solved by pivoted
    elimination]) the counts of the solve ride in the same all-reduce and their
    totals are returned as a third value (an int, or a tuple of ints)."""
    import torch.distributed as dist

    def counts_on(dev):
        if n_bad_local is None:
            return []
        if isinstance(n_bad_local, torch.Tensor):
            return [n_bad_local.reshape(-1).to(device=dev, dtype=torch.float64)]
        return [torch.tensor([float(n_bad_local)], dtype=torch.float64, device=dev)]

    def result(tot):
        res = (float(np.sqrt(tot[0]) / tot[1]), float(tot[1]))
        if n_bad_local is None:
            return res
        counts = tuple(int(round(float(v))) for v in tot[2:])
        return res + (counts[0] if len(counts) == 1 else counts,)

    if isinstance(comm, NativeComm):
        from . import _lib
        dev = comm.device
        head = [sum_sq_local.reshape(-1)[0:1].to(device=dev, dtype=torch.float64)
                if isinstance(sum_sq_local, torch.Tensor) else torch.tensor([float(sum_sq_local)], dtype=torch.float64, device=dev),
                torch.tensor([float(n_elems_local)], dtype=torch.float64, device=dev)]
        buf = comm.all_reduce_sums(torch.cat(head + counts_on(dev)))
        with torch.cuda.device(dev):
            tot = _lib.read_small(buf)
        return result(tot)
    if comm is None and not (dist.is_available() and dist.is_initialized()):
        s = float(sum_sq_local)
        res = (np.sqrt(s) / float(n_elems_local), float(n_elems_local))
        if n_bad_local is None:
            return res
        if isinstance(n_bad_local, torch.Tensor) and n_bad_local.numel() > 1:
            return res + (tuple(int(v) for v in n_bad_local.reshape(-1).tolist()),)
        return res + (int(n_bad_local),)
    if isinstance(sum_sq_local, torch.Tensor):
        dev = sum_sq_local.device
        buf = torch.cat([sum_sq_local.reshape(-1)[0:1].to(torch.float64),
                         torch.tensor([float(n_elems_local)], dtype=torch.float64, device=dev)] + counts_on(dev))
    else:
        buf = torch.cat([torch.tensor([float(sum_sq_local), float(n_elems_local)], dtype=torch.float64)]
                        + counts_on(torch.device("cpu")))
    backend = dist.get_backend(comm)
    if backend == "gloo" and buf.is_cuda:
        buf = buf.cpu()
    dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=comm)
    if buf.is_cuda:
        from . import _lib
        with torch.cuda.stream(torch.cuda.current_stream(buf.device)):
            tot = _lib.read_small(buf)           # no device->host memcpy: see kafka_read_small
    else:
        tot = buf.numpy()
    return result(tot)


def gather_interleaved(x_local, comm=None, dst=None):
    """Concatenate per-rank 1-D tensors in rank order (all ranks, or only ``dst``)."""
    import torch.distributed as dist
    if isinstance(comm, NativeComm):
        cnt = torch.zeros(comm.n_ranks, dtype=torch.float64, device=comm.device)
        cnt[comm.rank] = float(x_local.numel())
        counts = [int(c) for c in comm.all_reduce_sums(cnt).cpu().tolist()]
        return comm.gather(x_local.to(device=comm.device, dtype=torch.float64).reshape(-1), counts)
    if comm is None and not (dist.is_available() and dist.is_initialized()):
        return x_local
    world = dist.get_world_size(comm)
    sizes = [torch.zeros(1, dtype=torch.int64, device=x_local.device) for _ in range(world)]
    dist.all_gather(sizes, torch.tensor([x_local.numel()], dtype=torch.int64, device=x_local.device), group=comm)
    sizes = [int(s.item()) for s in sizes]
    m = max(sizes)
    pad = torch.zeros(m, dtype=x_local.dtype, device=x_local.device)
    pad[:x_local.numel()] = x_local
    out = [torch.empty(m, dtype=x_local.dtype, device=x_local.device) for _ in range(world)]
    dist.all_gather(out, pad, group=comm)
    return torch.cat([o[:s] for o, s in zip(out, sizes)])

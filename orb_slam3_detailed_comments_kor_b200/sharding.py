"""Multi-GPU plumbing (one process per GPU, torch.distributed): how the two shardable parts of the path are split.

* Frames / stereo pairs are independent: rank r processes pairs r, r+R, ... with its own extractors — no collective.
* The map-descriptor database is split into R contiguous row ranges; every rank searches its range with
  index_base = first global row, the per-shard top-2 records are all-gathered (NCCL on GPUs, gloo in the CPU tests)
  and merged on the GPU with the same strict-< rule a single scan applies (k_top2_merge).
The reference is single-process (SURVEY.md §5 'Distributed communication backend: none'); this file is new."""


def shard_rows(ndb, world, rank):
    """Contiguous row range [lo, hi) of the database owned by `rank` (the last rank takes the remainder)."""
    base = ndb // world
    lo = rank * base
    hi = ndb if rank == world - 1 else lo + base
    return lo, hi


def frames_for_rank(nframes, world, rank):
    return list(range(rank, nframes, world))


def gather_top2(local, world, group=None):
    """local: torch int32 tensor [nq, 4] (best1, best2, idx, accepted) on the backend's device.
    Returns the gathered [world, nq, 4] tensor in rank (= ascending global row) order."""
    import torch
    import torch.distributed as dist
    if world == 1:
        return local.unsqueeze(0)
    out = torch.empty((world * local.shape[0],) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, local.contiguous(), group=group)   # concatenated along dim 0 (gloo and NCCL agree)
    return out.view((world,) + tuple(local.shape))


def merge_gathered(gathered, th, ratio, strict_lt, device=None, stream=None):
    """Merge [world, nq, 4] per-shard records on the GPU (k_top2_merge via orbfe_top2_merge_device).  There is no CPU
    path: a host tensor is an error."""
    import torch
    from . import matcher as M
    world, nq = gathered.shape[0], gathered.shape[1]
    if not gathered.is_cuda:
        raise RuntimeError("merge_gathered needs CUDA tensors: the product has no CPU path")
    out = torch.empty((nq, 4), dtype=torch.int32, device=gathered.device)
    M.merge_top2_device(gathered.device.index if device is None else device, gathered.data_ptr(), world, nq, th, ratio,
                        strict_lt, out.data_ptr(), stream)
    return out


def create_nccl_comm(rank, world, device):
    """An ncclComm_t for orbfe_hamming_top2_sharded (the C entry point takes the communicator of its C++ host; this harness has
    torch.distributed, whose communicator is not reachable from Python, so it bootstraps a second one): rank 0 draws the unique
    id, torch.distributed broadcasts the 128 bytes, every rank joins.  Returns the handle as an int (0 when world == 1)."""
    import ctypes as C
    import torch
    import torch.distributed as dist
    from . import _lib
    if world == 1:
        return 0
    L = _lib.lib()
    buf = (C.c_uint8 * 128)()
    if rank == 0:
        _lib.check(L.orbfe_nccl_unique_id(buf))
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    t = torch.tensor(list(buf), dtype=torch.uint8, device=dev)
    dist.broadcast(t, 0)
    raw = bytes(t.cpu().tolist())
    ident = (C.c_uint8 * 128).from_buffer_copy(raw)
    comm = C.c_void_p()
    _lib.check(L.orbfe_nccl_comm_create(device, world, rank, ident, C.byref(comm)))
    return comm.value


def destroy_nccl_comm(comm):
    from . import _lib
    if comm:
        _lib.check(_lib.lib().orbfe_nccl_comm_destroy(comm))

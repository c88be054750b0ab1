"""Multi-GPU use of the codec (BASELINE config 3(ii) and config 5).

Two things live here, both host-side plumbing over `torch.distributed` (one process per GPU):

* `partition` / `compress_sharded`: batch elements are independent (reference README.md:110), so a batch is
  split by element index into contiguous ranges, one per rank, and every rank runs the single-GPU path on its
  range. No data-path collective; the archives are byte-identical to the single-GPU ones.
* `all_gather_compressed` / `all_to_all_compressed`: the reference's stated purpose (README.md:92-96,
  :125-127) — compress, move only the compressed bytes with NCCL over NVLink, decompress from the receive
  buffer. rANS archives are not summable, so there is no compressed all-reduce.

The codec is injected (`codec=`) so the world_size-2 gloo tests can drive the exact same exchange logic on CPU
tensors with a stand-in codec; the default is the CUDA codec and nothing here falls back to a CPU path.
"""
import torch
import torch.distributed as dist


def partition(num_elems, world, rank):
    """Contiguous, balanced range [lo, hi) of batch elements owned by `rank` (first `num % world` ranks get
    one extra)."""
    base, extra = divmod(num_elems, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


class CudaFloatCodec:
    """floatCompress / floatDecompress through the torch-op mirror (libdietgpu_b200.so)."""

    def __init__(self, checksum=False, temp_mem=None):
        from . import torch_ops
        self._ops = torch_ops
        self.checksum = checksum
        self.temp_mem = temp_mem

    def compress(self, ts, out=None, out_sizes=None):
        comp, sizes, _ = self._ops.compress_data(True, ts, self.checksum, self.temp_mem, out, out_sizes)
        return comp, sizes

    def decompress(self, rows, outs):
        self._ops.decompress_data(True, rows, outs, self.checksum, self.temp_mem)


def compress_sharded(ts, codec=None, group=None):
    """Compress this rank's slice of a batch that every rank holds the description of.
    -> (lo, hi, compressed matrix, sizes) for elements [lo, hi) (matrix rows are in that order)."""
    codec = codec or CudaFloatCodec()
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    lo, hi = partition(len(ts), world, rank)
    if hi == lo:
        return lo, hi, None, None
    comp, sizes = codec.compress(list(ts[lo:hi]))
    return lo, hi, comp, sizes


def _round_up(x, m):
    return (x + m - 1) // m * m


def _chunks(flat, num_chunks):
    n = flat.numel()
    num_chunks = max(1, min(num_chunks, n)) if n else 1
    per = _round_up(-(-n // num_chunks), 8) if n else 0  # keep every chunk start 16-byte aligned for 2-byte words
    out, off = [], 0
    while off < n:
        out.append(flat.narrow(0, off, min(per, n - off)))
        off += per
    return out or [flat]


def all_gather_compressed(t, out=None, group=None, num_chunks=4, codec=None):
    """All-gather of equal-sized float shards that moves compressed bytes only.

    Every rank passes a tensor of the same shape and dtype; returns `out` of shape [world, *t.shape] with
    out[r] == rank r's `t`, bit for bit. The shard is cut into `num_chunks` batch elements: their row gathers
    are issued asynchronously, so chunk k+1 is on the wire while chunk k is being decompressed."""
    codec = codec or CudaFloatCodec()
    world = dist.get_world_size(group)
    flat = t.contiguous().view(-1)
    chunks = _chunks(flat, num_chunks)
    nchunk = len(chunks)
    if out is None:
        out = torch.empty((world,) + tuple(t.shape), dtype=t.dtype, device=t.device)
    out_flat = out.view(world, -1)

    comp, sizes = codec.compress(chunks)
    # sizes exchange: nchunk u32 per rank
    all_sizes = torch.empty((world * nchunk,), dtype=torch.int32, device=t.device)
    dist.all_gather_into_tensor(all_sizes, sizes[:nchunk].contiguous(), group=group)
    all_sizes_h = all_sizes.cpu().view(world, nchunk)
    row_bytes = [min(_round_up(int(all_sizes_h[:, c].max()), 16), comp.size(1)) for c in range(nchunk)]

    recv, work = [], []
    for c in range(nchunk):
        buf = torch.empty((world, row_bytes[c]), dtype=torch.uint8, device=t.device)
        recv.append(buf)
        work.append(dist.all_gather_into_tensor(buf.view(-1), comp[c, :row_bytes[c]], group=group, async_op=True))
    off = 0
    for c in range(nchunk):
        work[c].wait()
        n = chunks[c].numel()
        rows = [recv[c][r, :int(all_sizes_h[r, c])] for r in range(world)]
        outs = [out_flat[r].narrow(0, off, n) for r in range(world)]
        codec.decompress(rows, outs)
        off += n
    return out


def all_to_all_compressed(t, out=None, group=None, codec=None):
    """All-to-all of float chunks through their compressed form.

    `t` has shape [world, ...]: t[j] is sent to rank j; returns `out` of the same shape with out[i] == rank i's
    t[my rank]. One batched compress of the `world` outgoing chunks, a sizes all-to-all, grouped point-to-point
    sends of exactly the compressed bytes (all-to-all-v), one batched decompress."""
    codec = codec or CudaFloatCodec()
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    assert t.size(0) == world, "first dimension must be the group size"
    t = t.contiguous()
    if out is None:
        out = torch.empty_like(t)
    send_chunks = [t[j].view(-1) for j in range(world)]
    comp, sizes = codec.compress(send_chunks)

    in_sizes = torch.empty((world,), dtype=torch.int32, device=t.device)
    dist.all_to_all_single(in_sizes, sizes[:world].contiguous(), group=group)
    send_h = sizes[:world].cpu().tolist()
    recv_h = in_sizes.cpu().tolist()

    recv = torch.empty((world, _round_up(max(recv_h), 16)), dtype=torch.uint8, device=t.device)
    recv[rank, :send_h[rank]].copy_(comp[rank, :send_h[rank]])
    ops = []
    for step in range(1, world):  # rank r talks to r+step / r-step in step order: no two sends target one peer
        to, frm = (rank + step) % world, (rank - step) % world
        ops.append(dist.P2POp(dist.isend, comp[to, :send_h[to]], _global_rank(group, to), group))
        ops.append(dist.P2POp(dist.irecv, recv[frm, :recv_h[frm]], _global_rank(group, frm), group))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    rows = [recv[i, :recv_h[i]] for i in range(world)]
    outs = [out[i].view(-1) for i in range(world)]
    codec.decompress(rows, outs)
    return out


def _global_rank(group, group_rank):
    return group_rank if group is None else dist.get_global_rank(group, group_rank)

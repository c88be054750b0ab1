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


# ---- peer-memory collectives: the exchange fused into the decompress kernel --------------------------------
# Inside one NVLink / NVSwitch node the compressed bytes need not be moved by a collective at all: every rank
# compresses into a buffer its peers can map (CUDA IPC), a stream-ordered barrier says "archives are complete",
# and every rank's decompress kernel reads the peers' archives in place - its asynchronous copies pull exactly the
# bytes it decodes over NVLink, tile by tile, while it decodes. No staging buffer, no size exchange (the archives
# are self-describing), no host synchronisation anywhere in the call.
class _Raw:
    """pointer-only stand-in for a tensor (the codec layer only asks for data_ptr())"""

    def __init__(self, ptr):
        self._p = ptr

    def data_ptr(self):
        return self._p


class PeerBuffers:
    """One archive buffer per rank, mapped by every rank of the group (all on one node). Two halves are used in
    turn by successive collectives, which is what lets one barrier per call suffice: a rank reaches the barrier
    of call k + 1 only after its own decode of call k, so when any rank passes that barrier and later overwrites
    half (k + 2) % 2 == k % 2, nobody is still reading it.

    Like any collective, this requires every rank of the group to issue the SAME sequence of peer collectives on one
    PeerBuffers object (all-gather and all-to-all calls in the same order, on the same stream): the half a call uses
    is chosen by a local call counter, not negotiated. tests/multigpu_worker.py runs five consecutive mixed calls
    with one rank artificially delayed to exercise the reuse."""

    def __init__(self, half_bytes, group=None, device=None):
        import ctypes as C

        from . import _capi
        self._capi, self._C = _capi, C
        self.group = group
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.device = device if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.half_bytes = (int(half_bytes) + 255) // 256 * 256
        ptr, handle = C.c_void_p(), (C.c_uint8 * 64)()
        _capi.check(_capi.lib.dgb_peer_alloc(2 * self.half_bytes, C.byref(ptr), handle), "dgb_peer_alloc")
        self.local = ptr.value
        mine = torch.tensor(list(handle), dtype=torch.uint8, device=self.device)
        handles = torch.empty(self.world * 64, dtype=torch.uint8, device=self.device)
        dist.all_gather_into_tensor(handles, mine, group=group)
        handles = handles.cpu().view(self.world, 64)
        self.ptrs, self._opened = [], []
        for r in range(self.world):
            if r == self.rank:
                self.ptrs.append(self.local)
                continue
            p = C.c_void_p()
            h = (C.c_uint8 * 64)(*handles[r].tolist())
            _capi.check(_capi.lib.dgb_peer_open(h, C.byref(p)), "dgb_peer_open")
            self.ptrs.append(p.value)
            self._opened.append(p.value)
        self._flag = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._sizes = None
        self.calls = 0
        # pull mode (see all_gather_compressed_peer): local staging for the peers' archives, a copy stream, one event
        # per source, and cached device-side pointer tables for the pull kernel (they depend only on the layout)
        self._staging = None
        self.copy_stream = torch.cuda.Stream(device=self.device)
        self._events = [torch.cuda.Event() for _ in range(max(1, self.world))]
        self._plans = {}
        dist.barrier(group=group)

    def barrier(self):
        """Stream-ordered barrier: a one-word all-reduce. It completes on a rank's stream only after every rank's
        stream has reached it; the host does not wait."""
        dist.all_reduce(self._flag, group=self.group)

    def next_half(self):
        off = (self.calls & 1) * self.half_bytes
        self.calls += 1
        return off

    def staging(self, nbytes):
        """Local buffer the peers' archives are pulled into (grown on demand, then reused by every call)."""
        if self._staging is None or self._staging.numel() < nbytes:
            self._staging = torch.empty(int(nbytes), dtype=torch.uint8, device=self.device)
        return self._staging

    def pull(self, plan):
        """Enqueues one pull (dgb_peer_pull_float) on the copy stream. plan = (src_dev, dst_dev, cap_dev, count,
        max_bytes): device tables of archive addresses in peer memory, local destinations and their capacities."""
        src, dst, cap, count, max_bytes = plan
        C = self._C
        rc = self._capi.lib.dgb_peer_pull_float(count, C.c_void_p(src.data_ptr()), C.c_void_p(dst.data_ptr()),
                                                C.c_void_p(cap.data_ptr()), int(max_bytes),
                                                C.c_void_p(self.copy_stream.cuda_stream))
        self._capi.check(rc, "dgb_peer_pull_float")

    def plan(self, key, srcs, dsts, cap_bytes):
        """Cached device tables for one pull: srcs / dsts are lists of raw addresses."""
        got = self._plans.get(key)
        if got is None:
            dev = self.device
            got = (torch.tensor(srcs, dtype=torch.int64, device=dev), torch.tensor(dsts, dtype=torch.int64, device=dev),
                   torch.full((len(srcs),), int(cap_bytes), dtype=torch.int32, device=dev), len(srcs), int(cap_bytes))
            self._plans[key] = got
        return got

    def sizes_scratch(self, n):
        if self._sizes is None or self._sizes.numel() < n:
            self._sizes = torch.zeros(n, dtype=torch.int32, device=self.device)
        return self._sizes

    def close(self):
        torch.cuda.synchronize()
        dist.barrier(group=self.group)
        for p in self._opened:
            self._capi.lib.dgb_peer_close(self._C.c_void_p(p))
        self._opened = []
        if self.local:
            self._capi.lib.dgb_peer_free(self._C.c_void_p(self.local))
            self.local = 0


def peer_bytes_needed(dtype, numel, num_chunks=1):
    """Bytes of one PeerBuffers half for collectives over tensors of `numel` elements cut into `num_chunks`."""
    from . import codec
    per = _round_up(-(-numel // max(1, num_chunks)), 8)
    return max(1, num_chunks) * _round_up(codec.get_max_float_compressed_size(codec.FLOAT_TYPE_OF_DTYPE[dtype], per), 256)


def all_gather_compressed_peer(t, peers, out=None, num_chunks=1, temp=None, prob_bits=10, mode="inplace", stages=None):
    """All-gather of equal-sized float shards through peer memory: every rank compresses its shard into its
    peer-visible buffer; after a stream-ordered barrier every rank fetches and decodes every other rank's archives.
    Returns `out` of shape [world, *t.shape], bit-identical to the uncompressed all-gather. No sizes travel (the
    archives are self-describing) and the host never waits.

    mode "inplace" (default): one batched decompress whose inputs are the peers' buffers; the decoder's own
    asynchronous copies pull exactly the bytes it consumes over NVLink while it decodes. Each 4096-symbol block
    starts with dependent remote reads (325 GB/s per archive against 1.9 TB/s locally), but it is one launch.
    mode "pull": a copy kernel streams a peer's archives into local staging (dgb_peer_pull_float: lengths are read
    from the archive headers on the device) on a copy stream, and the local decode of source k runs while source
    k + 1 is being pulled; the own shard is decoded during the first pull. `stages` groups the world - 1 sources into
    that many pull / decode steps (default: one per source).
    Measured on 8 x B200, 64 MiB of bf16 per rank (profiles/r2_bench_ours_8gpu_pullmode.json): raw NCCL 0.817 ms,
    in place 0.905 ms, pull with 7 stages 1.104 ms, with 2 stages 0.920 ms - the per-stage launches cost more than the
    overlap returns, so in place stays the default. (Bound: at ratio 0.67 compression saves 0.27 ms of wire time per
    call while compress + decode cost 0.41 ms; with perfect overlap the compressed form could reach about 1.1 x raw.)"""
    from . import codec
    ft = codec.FLOAT_TYPE_OF_DTYPE[t.dtype]
    world, rank = peers.world, peers.rank
    flat = t.contiguous().view(-1)
    chunks = _chunks(flat, num_chunks)
    nch = len(chunks)
    stride = _round_up(codec.get_max_float_compressed_size(ft, chunks[0].numel()), 256)
    assert stride * nch <= peers.half_bytes, "PeerBuffers too small (see peer_bytes_needed)"
    if out is None:
        out = torch.empty((world,) + tuple(t.shape), dtype=t.dtype, device=t.device)
    out_flat = out.view(world, -1)
    half = peers.next_half()
    rows = [_Raw(peers.local + half + c * stride) for c in range(nch)]
    codec.float_compress(ft, chunks, rows, peers.sizes_scratch(nch), prob_bits, False, temp=temp)
    peers.barrier()

    def outs_of(r):
        res, off = [], 0
        for ch in chunks:
            res.append(out_flat[r].narrow(0, off, ch.numel()))
            off += ch.numel()
        return res

    # Source order = read order: rank i starts with rank i + 1's archives, then i + 2's, ..., so that at any moment
    # every source is read by one peer only instead of all ranks pulling from the same GPU.
    order = [(rank + k) % world for k in range(1, world)]
    if mode == "inplace":
        ins, outs = [], []
        for r in order + [rank]:
            ins += [_Raw(peers.ptrs[r] + half + c * stride) for c in range(nch)]
            outs += outs_of(r)
        codec.float_decompress(ft, ins, outs, None, None, prob_bits, False, temp=temp)
        return out

    main = torch.cuda.current_stream()
    nstage = max(1, min(len(order), stages if stages else len(order))) if order else 0
    groups = [order[len(order) * g // nstage:len(order) * (g + 1) // nstage] for g in range(nstage)] if order else []
    staging = peers.staging(max(1, world - 1) * nch * stride)
    sbase = staging.data_ptr()
    peers.copy_stream.wait_stream(main)  # archives complete (barrier) and the previous call's decodes done with staging
    slot = 0
    staged = []
    for g, grp in enumerate(groups):
        srcs, dsts = [], []
        for r in grp:
            for c in range(nch):
                srcs.append(peers.ptrs[r] + half + c * stride)
                dsts.append(sbase + (slot * nch + c) * stride)
            slot += 1
        plan = peers.plan(("ag", half, stride, nch, nstage, g, sbase), srcs, dsts, stride)
        with torch.cuda.stream(peers.copy_stream):
            peers.pull(plan)
            peers._events[g].record(peers.copy_stream)
        staged.append(dsts)
    codec.float_decompress(ft, rows, outs_of(rank), None, None, prob_bits, False, temp=temp)  # own shard, during pull 0
    for g, grp in enumerate(groups):
        main.wait_event(peers._events[g])
        outs = []
        for r in grp:
            outs += outs_of(r)
        codec.float_decompress(ft, [_Raw(p) for p in staged[g]], outs, None, None, prob_bits, False, temp=temp)
    return out


def all_to_all_compressed_peer(t, peers, out=None, temp=None, prob_bits=10, mode="inplace"):
    """All-to-all of float chunks (t[j] goes to rank j) through peer memory: one batched compress of the outgoing
    chunks into the own buffer, a barrier, then row `rank` of every peer's buffer is fetched and decoded: in "pull"
    mode one copy kernel streams the world - 1 archives into local staging (while the own chunk is decoded) and one
    batched decompress follows; in "inplace" mode the decompress reads the peers' buffers directly."""
    from . import codec
    ft = codec.FLOAT_TYPE_OF_DTYPE[t.dtype]
    world, rank = peers.world, peers.rank
    assert t.size(0) == world, "first dimension must be the group size"
    t = t.contiguous()
    if out is None:
        out = torch.empty_like(t)
    send = [t[j].view(-1) for j in range(world)]
    stride = _round_up(codec.get_max_float_compressed_size(ft, send[0].numel()), 256)
    assert stride * world <= peers.half_bytes, "PeerBuffers too small (see peer_bytes_needed)"
    half = peers.next_half()
    rows = [_Raw(peers.local + half + j * stride) for j in range(world)]
    codec.float_compress(ft, send, rows, peers.sizes_scratch(world), prob_bits, False, temp=temp)
    peers.barrier()
    order = [(rank + k) % world for k in range(1, world)]  # staggered sources (see all_gather_compressed_peer)
    if mode == "inplace" or not order:
        ins = [_Raw(peers.ptrs[i] + half + rank * stride) for i in order + [rank]]
        outs = [out[i].view(-1) for i in order + [rank]]
        codec.float_decompress(ft, ins, outs, None, None, prob_bits, False, temp=temp)
        return out
    main = torch.cuda.current_stream()
    staging = peers.staging((world - 1) * stride)
    sbase = staging.data_ptr()
    srcs = [peers.ptrs[i] + half + rank * stride for i in order]
    dsts = [sbase + k * stride for k in range(len(order))]
    plan = peers.plan(("aa", half, stride, sbase), srcs, dsts, stride)
    peers.copy_stream.wait_stream(main)
    with torch.cuda.stream(peers.copy_stream):
        peers.pull(plan)
        peers._events[0].record(peers.copy_stream)
    codec.float_decompress(ft, [rows[rank]], [out[rank].view(-1)], None, None, prob_bits, False, temp=temp)
    main.wait_event(peers._events[0])
    codec.float_decompress(ft, [_Raw(p) for p in dsts], [out[i].view(-1) for i in order], None, None, prob_bits, False,
                           temp=temp)
    return out

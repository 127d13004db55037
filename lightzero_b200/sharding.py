"""Multi-GPU sharding of one search batch.

Trees never interact (every loop of cbatch_traverse / cbatch_backpropagate is over independent roots,
ctree_muzero/lib/cnode.cpp:495-499,783-823), so the env batch is cut into contiguous slices, one per rank, each
with its own forest, latent pool and model replica.  There is NO collective inside the simulation loop; the only
exchange is one all-gather of the root outputs at the end of a search.  The tie-break stream is keyed by the
GLOBAL root index (Roots.root_offset), so results do not depend on the number of ranks.
"""
import ctypes as C

import torch
import torch.distributed as dist


def shard_bounds(num_roots, world_size, rank):
    """Contiguous slice [lo, hi) of the env batch owned by `rank` (first `num_roots % world_size` ranks get one more)."""
    base, extra = divmod(int(num_roots), int(world_size))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def all_gather_rows(local, counts=None, group=None):
    """Concatenate per-rank [n_r, ...] tensors along dim 0 on every rank.  Equal n_r: one all_gather_into_tensor
    (NCCL over NVLink / NVSwitch on GPUs, gloo on CPU); ragged: padded to max(counts) and trimmed."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local
    world = dist.get_world_size(group)
    local = local.contiguous()
    if counts is None or len(set(counts)) == 1:
        out = torch.empty((world * local.shape[0],) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        dist.all_gather_into_tensor(out, local, group=group)
        return out
    m = max(counts)
    pad = torch.zeros((m,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[:local.shape[0]] = local
    out = torch.empty((world * m,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, pad, group=group)
    return torch.cat([out[r * m:r * m + c] for r, c in enumerate(counts)], dim=0)


class PeerTimeout(RuntimeError):
    """A rank's shard did not reach this rank's receive buffer in time (PeerExchange.gather)."""


class _RawCuda:
    """A device allocation owned by liblzb200 as something torch.as_tensor can alias (no copy)."""

    def __init__(self, ptr, shape, typestr="<i4"):
        self.__cuda_array_interface__ = dict(shape=tuple(shape), typestr=typestr, data=(int(ptr), False), version=2)


class PeerExchange:
    """The end-of-search all-gather as ONE kernel per rank over NVLink / NVSwitch peer memory (include/lzb200.h,
    lzb_peer_allgather): rows are packed and stored straight into every peer's receive buffer, then a flag barrier over
    the same mapping.  One object serves every search of a fixed shard shape (rows_per_rank x row_words 32-bit words per
    row, e.g. A + 1 for visit counts + value).  Ranks must call ``gather`` in lockstep, like any collective.

    ``PeerExchange(rows, row_words, group)``: one rank per process, buffers shared through CUDA IPC handles moved with
    ``all_gather_object``.  ``PeerExchange.local_group(world, rows, row_words)``: ``world`` ranks inside ONE process on
    one device (tests, single-GPU smoke): same kernel, same protocol, each rank on its own stream."""

    def __init__(self, rows_per_rank, row_words, group=None, device=None, _local=None):
        from . import _lib
        self.L, self.check = _lib.lib(), _lib.check
        self.rows, self.words = int(rows_per_rank), int(row_words)
        self.epoch, self.imported = 0, []
        if _local is not None:
            self.rank, self.world, self.device, self.buf, self.bufs = _local
            self.owner = False
            return
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.buf, self.bufs, self.owner, err = None, [], True, None
        # every rank takes part in every collective below whether or not its own CUDA calls succeed, and all ranks
        # raise together, so a caller can fall back to NCCL without the ranks drifting apart
        with torch.cuda.device(self.device):
            handle = (C.c_ubyte * 64)()
            try:
                self.buf = self._alloc(self.world)
                self.check(self.L.lzb_peer_export(self.buf, handle))
            except Exception as exc:
                err = exc
            handles = [None] * self.world
            dist.all_gather_object(handles, (err is None, bytes(handle)), group=group)
            if all(ok for ok, _ in handles):
                try:
                    for r, (_, h) in enumerate(handles):
                        if r == self.rank:
                            self.bufs.append(self.buf)
                            continue
                        p = C.c_void_p()
                        self.check(self.L.lzb_peer_import((C.c_ubyte * 64).from_buffer_copy(h), C.byref(p)))
                        self.imported.append(p)
                        self.bufs.append(p)
                except Exception as exc:
                    err = exc
            elif err is None:
                err = RuntimeError("a peer could not allocate / export its receive buffer")
            ok = torch.tensor([0.0 if err is not None else 1.0], device=self.device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)  # also: every buffer is zeroed and mapped before first use
        if float(ok) == 0.0:
            self.close()
            raise RuntimeError("peer-memory exchange unavailable: %s" % (err if err is not None else "a peer failed to map the buffers"))

    def _alloc(self, world):
        n = C.c_size_t()
        self.check(self.L.lzb_peer_bytes(world, self.rows, self.words, C.byref(n)))
        p = C.c_void_p()
        self.check(self.L.lzb_peer_alloc(n.value, C.byref(p)))
        return p

    @classmethod
    def local_group(cls, world, rows_per_rank, row_words, device=None):
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        first = cls(rows_per_rank, row_words, _local=(0, world, dev, None, None))
        with torch.cuda.device(dev):
            bufs = [first._alloc(world) for _ in range(world)]
        ranks = [first] + [cls(rows_per_rank, row_words, _local=(r, world, dev, None, None)) for r in range(1, world)]
        for r, x in enumerate(ranks):
            x.buf, x.bufs, x.owner = bufs[r], bufs, True
        return ranks

    def gather(self, pieces, timeout_ms=30000, check=True):
        """pieces: 32-bit tensors [rows_per_rank, ...] of this rank -> the same tensors for all ranks, [world * rows,
        ...] (copied out of this rank's receive buffer, which the next-but-one gather reuses).

        ``check`` (default): read the kernel's error word back (one stream synchronisation - the caller reads the
        results on the host next anyway) and RAISE if a peer did not arrive within ``timeout_ms``: its rows would
        otherwise be consumed as visit counts.  ``check=False`` leaves that to the caller (``error()``); the rows of a
        missing peer are poisoned with INT_MIN by the kernel either way."""
        out = self.collect(self.launch(pieces, timeout_ms))
        if check:
            e = self.error()
            if e:
                raise PeerTimeout("peer exchange: the shard of rank %d did not arrive within %d ms (epoch %d)" % (e - 1, timeout_ms, self.epoch))
        return out

    def launch(self, pieces, timeout_ms=30000):
        """First half of ``gather``: enqueue the exchange kernel on the current stream."""
        rows = self.rows
        flat = [p.reshape(rows, -1).contiguous() for p in pieces]
        if any(p.element_size() != 4 or not p.is_cuda for p in flat) or sum(p.shape[1] for p in flat) != self.words:
            raise ValueError("pieces must be 32-bit device tensors with %d words per row in total" % self.words)
        self.epoch += 1
        n = len(flat)
        ptrs = (C.c_void_p * n)(*[p.data_ptr() for p in flat])
        words = (C.c_int32 * n)(*[p.shape[1] for p in flat])
        bufs = (C.c_void_p * self.world)(*[b.value if isinstance(b, C.c_void_p) else b for b in self.bufs])
        out = C.c_void_p()
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        self.check(self.L.lzb_peer_allgather(bufs, self.rank, self.world, rows, ptrs, words, n, self.epoch & 0xffffffff,
                                             int(timeout_ms), C.byref(out), stream))
        return out.value, [(p.dtype, tuple(p.shape[1:]), f.shape[1]) for p, f in zip(pieces, flat)], flat

    def collect(self, launched):
        """Second half of ``gather``: the gathered pieces as tensors (stream-ordered after the exchange kernel)."""
        ptr, metas, _keep = launched
        full = torch.as_tensor(_RawCuda(ptr, (self.world * self.rows, self.words)), device=self.device)
        res, at = [], 0
        for dtype, tail, w in metas:
            res.append(full[:, at:at + w].contiguous().view(dtype).reshape((self.world * self.rows,) + tail))
            at += w
        return res

    def error(self):
        """0, or 1 + the rank whose shard did not arrive within the timeout (synchronises the stream)."""
        e = C.c_int()
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        self.check(self.L.lzb_peer_error(self.buf, self.world, self.rows, self.words, C.byref(e), stream))
        return e.value

    def close(self):
        for p in self.imported:
            self.L.lzb_peer_unimport(p)
        self.imported = []
        if getattr(self, "owner", False) and self.buf is not None:
            self.L.lzb_peer_free(self.buf)
        self.buf = None


def gather_search_outputs(visits, values, counts=None, group=None, extras=(), peer=None, check=True):
    """End-of-search exchange: visit counts [n_r, A] int32 and root values [n_r] f32 (plus Gumbel policies /
    Sampled action tables in `extras`) -> the full batch on every rank.  Everything is 32-bit, so the pieces travel
    together: through ``peer`` (a PeerExchange: one NVLink kernel per rank) when given and the shards are equal,
    else as ONE [n_r, A + 1 + ...] int32 NCCL / gloo all-gather (the exchange is launch-latency bound: 0.6 MB per rank
    at the headline shape); they are split again on arrival.  ``check`` (peer route): see PeerExchange.gather."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return [visits, values] + list(extras)
    parts = [visits, values] + list(extras)
    if peer is not None and (counts is None or len(set(counts)) == 1) and all(p.element_size() == 4 for p in parts):
        return peer.gather(parts, check=check)  # raises PeerTimeout when a rank's shard did not arrive (check=True)
    if all(p.element_size() == 4 for p in parts):
        n = visits.shape[0]
        cols = [p.reshape(n, -1).view(torch.int32) for p in parts]
        widths = [c.shape[1] for c in cols]
        full = all_gather_rows(torch.cat(cols, dim=1), counts, group)
        out, at = [], 0
        for p, w in zip(parts, widths):
            piece = full[:, at:at + w].contiguous().view(p.dtype)
            out.append(piece.reshape((full.shape[0],) + tuple(p.shape[1:])))
            at += w
        return out
    return [all_gather_rows(p, counts, group) for p in parts]

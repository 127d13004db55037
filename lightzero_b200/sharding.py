"""Multi-GPU sharding of one search batch.

Trees never interact (every loop of cbatch_traverse / cbatch_backpropagate is over independent roots,
ctree_muzero/lib/cnode.cpp:495-499,783-823), so the env batch is cut into contiguous slices, one per rank, each
with its own forest, latent pool and model replica.  There is NO collective inside the simulation loop; the only
exchange is one all-gather of the root outputs at the end of a search.  The tie-break stream is keyed by the
GLOBAL root index (Roots.root_offset), so results do not depend on the number of ranks.
"""
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


def gather_search_outputs(visits, values, counts=None, group=None, extras=()):
    """End-of-search exchange: visit counts [n_r, A] int32 and root values [n_r] f32 (plus Gumbel policies /
    Sampled action tables in `extras`) -> the full batch on every rank.  Everything is 32-bit, so the pieces travel
    as ONE [n_r, A + 1 + ...] int32 all-gather (the exchange is launch-latency bound: 0.6 MB per rank at the headline
    shape) and are split again on arrival."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return [visits, values] + list(extras)
    parts = [visits, values] + list(extras)
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

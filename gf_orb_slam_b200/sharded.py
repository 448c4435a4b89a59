"""Multi-GPU layer of the front end: one process per GPU, `torch.distributed` for the plumbing.

* Frames / frame pairs / selection problems are independent units: `frame_shard` gives every rank a contiguous chunk of
  a sequence plus a 1-frame halo (the last frame of the previous chunk is re-extracted locally, so frame-to-frame
  matching needs no transfer).  No collective on that path.
* The one exchange step (BASELINE.json config 5): all-pairs Hamming top-2 of a query block against a descriptor
  database whose rows are sharded across ranks.  Queries are replicated (all-gathered when every rank extracted its own
  frames), each rank runs `gforb_allpairs_top2` on its shard, the packed (dist << 32 | global row) keys are all-gathered
  (16 B per query and rank) and merged with `gforb_top2_merge` (min key = min distance, then min global index).

The CUDA operators are the defaults and there is no CPU fallback; `local_top2` / `merge` are injectable only so that the
world_size-2 `gloo` test can drive the plumbing with the oracle on CPU."""
from __future__ import annotations

from typing import Callable, Optional, Tuple


def frame_shard(n_frames: int, world_size: int, rank: int, halo: int = 1) -> Tuple[int, int, int]:
    """Contiguous chunk of a sequence for `rank`: returns (first_frame, n_own, n_halo).  The rank extracts frames
    [first_frame - n_halo, first_frame + n_own) and reports results for its own frames only; chunk sizes differ by at
    most one frame and cover [0, n_frames) exactly."""
    if world_size <= 0 or not (0 <= rank < world_size) or n_frames < 0:
        raise ValueError("bad shard request")
    base, rem = divmod(n_frames, world_size)
    first = rank * base + min(rank, rem)
    own = base + (1 if rank < rem else 0)
    return first, own, (min(halo, first) if own > 0 else 0)


def db_shard(n_rows: int, world_size: int, rank: int) -> Tuple[int, int]:
    """Row range [row0, row0 + n) of the descriptor database held by `rank`."""
    first, own, _ = frame_shard(n_rows, world_size, rank, halo=0)
    return first, own


class ShardedDescriptorDB:
    """Descriptor database sharded by rows over the ranks of a process group; `top2` is the global brute-force top-2."""

    def __init__(self, shard, row0: int, *, group=None, matcher=None, local_top2: Optional[Callable] = None,
                 merge: Optional[Callable] = None):
        import torch.distributed as dist
        self.shard, self.row0, self.group = shard, int(row0), group
        self.dist = dist
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        if local_top2 is None or merge is None:
            if matcher is None:
                raise RuntimeError("ShardedDescriptorDB needs an ORBmatcher (CUDA); there is no CPU path")
            local_top2 = local_top2 or (lambda q, db, base: matcher.AllPairsTop2(q, db, idx_base=base))
            merge = merge or (lambda gathered: matcher.Top2Merge(gathered))
        self.local_top2, self.merge = local_top2, merge

    def gather_queries(self, local_queries):
        """all ranks extracted their own frames: replicate every rank's query block (n_q x 32 u8, equal n_q per rank)"""
        import torch
        if self.world == 1:
            return local_queries
        out = torch.empty((self.world * local_queries.shape[0],) + tuple(local_queries.shape[1:]), dtype=local_queries.dtype,
                          device=local_queries.device)                       # concatenated along dim 0 (rank-major)
        self.dist.all_gather_into_tensor(out, local_queries.contiguous(), group=self.group)
        return out

    def top2(self, queries):
        """queries: n_q x 32 u8, identical on every rank.  Returns n_q x 2 int64 keys (dist << 32 | global row), best
        then second best, identical on every rank."""
        import torch
        keys = self.local_top2(queries, self.shard, self.row0)                  # n_q x 2 (local shard)
        if self.world == 1:
            return self.merge(keys.reshape(1, *keys.shape))
        gathered = torch.empty((self.world * keys.shape[0],) + tuple(keys.shape[1:]), dtype=keys.dtype, device=keys.device)
        self.dist.all_gather_into_tensor(gathered, keys.contiguous(), group=self.group)
        return self.merge(gathered.reshape((self.world,) + tuple(keys.shape)))


def unpack_keys(keys):
    """(dist, global row) from packed keys; rows of queries with fewer than 2 database rows are -1 (key = all ones)"""
    dist = (keys >> 32) & 0xFFFFFFFF
    row = keys & 0xFFFFFFFF
    return dist, row

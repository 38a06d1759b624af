"""Host-side sharding of the path across GPUs (one process per GPU).

* DAG generation: rank g builds 256-byte items [g*share, min(M,(g+1)*share)), share = ceil(M/world);
  one in-place all-gather follows (the only collective on the path).
* Nonce search: the nonce space is cut into super-batches of B nonces; rank g of N scans batch k*N+g.
  The answer is the LOWEST qualifying nonce (reference progpow.cpp:572-577 returns the first hit of an
  ascending scan), so ranks are reduced with MIN, round by round.

No data-path collective is needed for the search; the reduction moves one int64 per rank.
"""
from __future__ import annotations

from dataclasses import dataclass

NOT_FOUND = (1 << 63) - 1  # int64 max: "no hit" in the MIN reduction (nonce offsets stay below 2^63)


def dag_share_items(num_items_256: int, world: int) -> int:
    return (num_items_256 + world - 1) // world


def dag_slice(num_items_256: int, rank: int, world: int):
    """[first, last) of the items rank builds; empty for trailing ranks of tiny DAGs."""
    share = dag_share_items(num_items_256, world)
    first = min(rank * share, num_items_256)
    return first, min(num_items_256, first + share)


def super_batch_start(start_nonce: int, round_index: int, rank: int, world: int, batch: int) -> int:
    return (start_nonce + (round_index * world + rank) * batch) & (2**64 - 1)


@dataclass
class ShardedSearchPlan:
    start_nonce: int
    iterations: int
    batch: int
    world: int

    @property
    def rounds(self) -> int:
        per_round = self.batch * self.world
        return (self.iterations + per_round - 1) // per_round

    def window(self, round_index: int, rank: int):
        """(start, count) rank scans in this round; count 0 when the range is exhausted."""
        off = (round_index * self.world + rank) * self.batch
        if off >= self.iterations:
            return (self.start_nonce + off) & (2**64 - 1), 0
        return (self.start_nonce + off) & (2**64 - 1), min(self.batch, self.iterations - off)


def reduce_lowest_offset(local_offset: int | None, dist=None, device="cpu") -> int | None:
    """MIN over ranks of the offset (nonce - start_nonce) of each rank's hit; None when nobody hit."""
    v = NOT_FOUND if local_offset is None else int(local_offset)
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        import torch
        t = torch.tensor([v], dtype=torch.int64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        v = int(t.item())
    return None if v == NOT_FOUND else v


def search_sharded(ctx, block_number: int, header_hash: bytes, boundary: bytes, start_nonce: int, iterations: int,
                   rank: int, world: int, batch: int = 1 << 24, dist=None, device="cpu"):
    """progpow::search over [start_nonce, start_nonce+iterations) split across `world` GPUs.
    Returns the same SearchResult on every rank (lowest qualifying nonce, or the all-zero result)."""
    from . import progpow
    if (start_nonce + iterations) % 2**64 <= start_nonce:  # uint64_t wrap: the reference's loop scans nothing
        return progpow.SearchResult()
    plan = ShardedSearchPlan(start_nonce, iterations, batch, world)
    for k in range(plan.rounds):
        s, n = plan.window(k, rank)
        local = progpow.search(ctx, block_number, header_hash, boundary, s, n) if n else progpow.SearchResult()
        off = ((local.nonce - start_nonce) & (2**64 - 1)) if local.solution_found else None
        best = reduce_lowest_offset(off, dist, device)
        if best is not None:
            nonce = (start_nonce + best) & (2**64 - 1)
            if local.solution_found and local.nonce == nonce:
                return local
            r = progpow.hash(ctx, block_number, header_hash, nonce)  # every rank holds the full DAG
            return progpow.SearchResult(True, nonce, r.final_hash, r.mix_hash)
    return progpow.SearchResult()

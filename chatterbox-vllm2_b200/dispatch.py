"""Replica dispatcher: whole requests go to the replica with the least outstanding KV tokens.

The reference has no distributed code (``LLM(...)`` runs with tensor_parallel_size = 1, tts.py:150-171) and requests are
independent (no cross-request state; the cond / uncond rows of a request stay together), so multi-GPU = one
``T3B200Engine`` per GPU and a host-side assignment of requests; nothing crosses NVLink on the decode path.

Two forms:
* ``assign_requests`` - the static form used when the request list is known up front (bench.py --workload cfg5, one process per
  GPU under torchrun): every rank computes the same assignment from the same list, so no message is needed at all.
* ``ReplicaDispatcher`` - the online form: requests arrive one at a time and go to the replica whose outstanding KV tokens
  (reported by ``T3B200Engine.kv_tokens_outstanding`` or estimated from the requests in flight) are lowest.
"""
import heapq
from typing import Callable, Dict, List, Optional, Sequence


def kv_token_cost(prompt_tokens: int, max_tokens: int) -> int:
    """KV token-rows a request writes over its life: both CFG rows hold prompt + generated tokens."""
    return 2 * (prompt_tokens + max_tokens)


def kv_traffic_cost(prompt_tokens: int, max_tokens: int) -> int:
    """KV token-rows a request READS over its life (what bounds tokens/s at large batch):
    sum over its steps of the context length, both rows."""
    n = max_tokens
    return 2 * (n * prompt_tokens + n * (n + 1) // 2)


def assign_requests(costs: Sequence[int], world_size: int) -> List[List[int]]:
    """Least-outstanding assignment of requests (given in arrival order) to ``world_size`` replicas: request i goes to the replica
    whose assigned cost so far is lowest (ties -> lowest rank).  Deterministic; returns per-rank index lists (arrival order kept
    inside a rank)."""
    if world_size < 1:
        raise ValueError("world_size must be >= 1")
    heap = [(0, r) for r in range(world_size)]
    out: List[List[int]] = [[] for _ in range(world_size)]
    for i, c in enumerate(costs):
        load, r = heapq.heappop(heap)
        out[r].append(i)
        heapq.heappush(heap, (load + int(c), r))
    return out


class ReplicaDispatcher:
    """Online least-outstanding-KV-tokens dispatcher over N replicas.  ``submit`` picks the replica, ``complete`` returns the
    request's tokens to its replica's budget.  ``probe`` (optional) reads the true load of a replica, e.g. over an RPC to
    ``T3B200Engine.kv_tokens_outstanding``; without it the dispatcher's own bookkeeping is used."""

    def __init__(self, n_replicas: int, probe: Optional[Callable[[int], int]] = None):
        if n_replicas < 1:
            raise ValueError("n_replicas must be >= 1")
        self.load = [0] * n_replicas
        self.where: Dict[int, tuple] = {}
        self.probe = probe

    def submit(self, request_id: int, prompt_tokens: int, max_tokens: int) -> int:
        cost = kv_token_cost(prompt_tokens, max_tokens)
        loads = [self.probe(r) for r in range(len(self.load))] if self.probe else self.load
        r = min(range(len(self.load)), key=lambda i: (loads[i], i))
        self.load[r] += cost
        self.where[request_id] = (r, cost)
        return r

    def complete(self, request_id: int) -> None:
        r, cost = self.where.pop(request_id)
        self.load[r] -= cost

    def outstanding(self) -> List[int]:
        return list(self.load)

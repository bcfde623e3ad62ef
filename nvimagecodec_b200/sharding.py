"""Host-side sharding of a batch across the GPUs of one box.

Images are independent, so there is no data-path collective: every device gets its own
bitstreams over PCIe and writes its own outputs.  The reference has no multi-GPU notion at all
(one decoder == one exec_params.device_id, include/nvimgcodec.h:1101); this thin layer sits
above the decoder ABI and only decides who decodes what.
"""
import heapq


def partition_by_bytes(sizes, n_parts, weights=None):
    """Greedy longest-processing-time partition of items by compressed byte count.

    Returns ``n_parts`` lists of indices.  Deterministic (ties broken by index), so every rank
    of a torchrun job computes the same answer without communicating.
    """
    if n_parts <= 0:
        raise ValueError("n_parts must be positive")
    cost = list(sizes) if weights is None else [s * w for s, w in zip(sizes, weights)]
    order = sorted(range(len(cost)), key=lambda i: (-cost[i], i))
    heap = [(0, p) for p in range(n_parts)]
    heapq.heapify(heap)
    parts = [[] for _ in range(n_parts)]
    for i in order:
        load, p = heapq.heappop(heap)
        parts[p].append(i)
        heapq.heappush(heap, (load + cost[i], p))
    for p in parts:
        p.sort()
    return parts


def shard_for_rank(sizes, rank, world_size):
    return partition_by_bytes(sizes, world_size)[rank]


def imbalance(sizes, parts):
    loads = [sum(sizes[i] for i in p) for p in parts]
    mean = sum(loads) / max(len(loads), 1)
    return (max(loads) / mean) if mean else 1.0

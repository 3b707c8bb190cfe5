"""Multi-GPU plumbing of the loop-closure path: one process per GPU.

The (node, submap) searches of ConstraintBuilder2D are independent (reference:
cartographer/mapping/internal/constraints/constraint_builder_2d.cc:102-111), so the
N-GPU path is a partition, not a collective:

  * submaps are assigned to ranks by greedy longest-processing-time on an estimated
    cost (every search against a submap runs on the rank that holds its stack);
  * each rank runs its shard through the C ABI on its own GPU;
  * results return to rank 0 in SUBMISSION order (the order of the reference's
    `constraints_` deque, constraint_builder_2d.cc:97-99,285-288).

The only collective is the broadcast of a submap grid (uint16 cells) from the rank
that produced it to the ranks that replicate it; `torch.distributed` carries it
(NCCL over NVLink on GPUs, gloo in the CPU tests).  Nothing here computes scores.
"""
import numpy as np


def assign_submaps(costs, world_size):
    """Greedy LPT: returns owner[rank] per submap for the given per-submap costs."""
    costs = np.asarray(costs, dtype=np.float64)
    owner = np.zeros(len(costs), dtype=np.int64)
    load = np.zeros(world_size, dtype=np.float64)
    for s in np.argsort(-costs, kind="stable"):
        r = int(np.argmin(load))
        owner[s] = r
        load[r] += costs[s]
    return owner


def shard_searches(search_submap, owner, rank):
    """Indices (into the submission order) of the searches `rank` runs."""
    search_submap = np.asarray(search_submap, dtype=np.int64)
    return np.nonzero(owner[search_submap] == rank)[0]


def estimate_costs(search_submap, points_per_search, num_submaps):
    """Per-submap cost ~ sum of points of the searches against it."""
    cost = np.zeros(num_submaps, dtype=np.float64)
    np.add.at(cost, np.asarray(search_submap, dtype=np.int64), np.asarray(points_per_search, float))
    return cost


def gather_in_submission_order(local_indices, local_results, num_searches, group=None, dst=0):
    """Collects every rank's (index, result) pairs on `dst`; returns the list of
    results in submission order there (None elsewhere)."""
    import torch.distributed as dist
    world = dist.get_world_size(group)
    payload = (np.asarray(local_indices, dtype=np.int64).tolist(), list(local_results))
    gathered = [None] * world if dist.get_rank(group) == dst else None
    dist.gather_object(payload, gathered, dst=dst, group=group)
    if gathered is None:
        return None
    out = [None] * num_searches
    seen = 0
    for idx, res in gathered:
        for i, r in zip(idx, res):
            assert out[i] is None, "search %d ran on two ranks" % i
            out[i] = r
            seen += 1
    assert seen == num_searches, "%d of %d searches returned" % (seen, num_searches)
    return out


def broadcast_grid(cells, src, device=None, group=None):
    """Broadcasts a submap's uint16 cell array [ny, nx] from `src` to all ranks.
    `cells` may be None on the receivers.  Returns a numpy uint16 array everywhere."""
    import torch
    import torch.distributed as dist
    rank = dist.get_rank(group)
    shape = torch.zeros(2, dtype=torch.int64, device=device)
    if rank == src:
        cells = np.ascontiguousarray(cells, dtype=np.uint16)
        shape = torch.tensor(cells.shape, dtype=torch.int64, device=device)
    dist.broadcast(shape, src=src, group=group)
    ny, nx = int(shape[0]), int(shape[1])
    # Sent as raw bytes: every backend carries uint8.
    if rank == src:
        buf = torch.from_numpy(cells.view(np.uint8).reshape(-1).copy())
        if device is not None:
            buf = buf.to(device)
    else:
        buf = torch.empty(ny * nx * 2, dtype=torch.uint8, device=device)
    dist.broadcast(buf, src=src, group=group)
    return buf.cpu().numpy().view(np.uint16).reshape(ny, nx)

"""Multi-GPU plumbing of the loop-closure path: one process per GPU.

The (node, submap) searches of ConstraintBuilder2D are independent (reference:
cartographer/mapping/internal/constraints/constraint_builder_2d.cc:102-111), so the
N-GPU path is a partition, not a collective:

  * searches are assigned to ranks by submap affinity (every search against a submap
    runs on a rank that holds its stack) with greedy longest-processing-time on an
    estimated cost; a submap whose searches would unbalance the ranks is REPLICATED
    and its node list split (the "new submap x all old nodes" burst of
    mapping/internal/2d/pose_graph_2d.cc:383-393);
  * each rank runs its shard through the C ABI on its own GPU;
  * results return to rank 0 in SUBMISSION order (the order of the reference's
    `constraints_` deque, constraint_builder_2d.cc:97-99,285-288).

The only collective on the data path is the broadcast of a submap grid (uint16 cells)
from the rank that produced it to the ranks that replicate it; `torch.distributed`
carries it (NCCL over NVLink on GPUs, gloo in the CPU tests).  Nothing here computes
scores.
"""
import ctypes

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


def search_costs(points_per_search, scans_per_search=None):
    """Estimated cost of a search ~ points x angular slices (the lowest-resolution sweep
    scores every slice's lattice against every point; FCSM2D.cc:264-333)."""
    c = np.asarray(points_per_search, dtype=np.float64)
    if scans_per_search is not None:
        c = c * np.asarray(scans_per_search, dtype=np.float64)
    return c


def estimate_costs(search_submap, points_per_search, num_submaps, scans_per_search=None):
    """Per-submap cost = sum of the estimated costs of the searches against it."""
    cost = np.zeros(num_submaps, dtype=np.float64)
    np.add.at(cost, np.asarray(search_submap, dtype=np.int64),
              search_costs(points_per_search, scans_per_search))
    return cost


def partition_searches(search_submap, cost_per_search, num_submaps, world_size, tolerance=0.02):
    """Owner rank of every search.  Whole submaps first (LPT on the per-submap cost), then
    the most loaded rank hands the tail of one submap's node list to the least loaded rank
    until the spread is within `tolerance` of the mean load: that submap is then replicated
    on both ranks (its grid is broadcast, each builds the stack).  Deterministic; with
    world_size 1 everything stays on rank 0."""
    search_submap = np.asarray(search_submap, dtype=np.int64)
    cost = np.asarray(cost_per_search, dtype=np.float64)
    sub_cost = np.zeros(num_submaps, dtype=np.float64)
    np.add.at(sub_cost, search_submap, cost)
    owner_of_submap = assign_submaps(sub_cost, world_size)
    owner = owner_of_submap[search_submap].copy()
    if world_size == 1 or len(cost) == 0:
        return owner
    load = np.zeros(world_size, dtype=np.float64)
    np.add.at(load, owner, cost)
    mean = load.sum() / world_size
    for _ in range(4 * world_size):
        hi, lo = int(np.argmax(load)), int(np.argmin(load))
        want = 0.5 * (load[hi] - load[lo])
        if load[hi] - load[lo] <= tolerance * mean:
            break
        mine = np.nonzero(owner == hi)[0]
        # the submap of `hi` with the most cost left on it gives up the tail of its node list
        subs, inv = np.unique(search_submap[mine], return_inverse=True)
        per = np.zeros(len(subs))
        np.add.at(per, inv, cost[mine])
        pick = subs[int(np.argmax(per))]
        cand = mine[search_submap[mine] == pick][::-1]  # from the tail, submission order
        moved = 0.0
        for i in cand:
            if moved + cost[i] > want + 1e-12 and moved > 0.0:
                break
            if moved + cost[i] > 2.0 * want:
                break
            owner[i] = lo
            moved += cost[i]
        if moved == 0.0:
            break
        load[hi] -= moved
        load[lo] += moved
    return owner


def replicas(search_submap, owner, num_submaps):
    """ranks_of[s] = sorted ranks that run at least one search against submap s."""
    search_submap = np.asarray(search_submap, dtype=np.int64)
    out = [[] for _ in range(num_submaps)]
    for s, r in sorted(set(zip(search_submap.tolist(), np.asarray(owner).tolist()))):
        out[s].append(r)
    return out


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


# One search result as it travels between ranks: 12 x int32 =
# (index, found, scan_index, x_index_offset, y_index_offset, score bits, pose x / y / theta as
# three float64 = 6 x int32).  Fixed size, so the gather is ONE collective on a device tensor.
RESULT_WORDS = 12

# csmh_result2d (host/host_capi.cc) as a numpy record: 48 bytes, naturally aligned
RESULT2D_DTYPE = np.dtype([("found", np.int32), ("scan_index", np.int32), ("x_index_offset", np.int32),
                           ("y_index_offset", np.int32), ("score", np.float32), ("tie_resolution", np.int32),
                           ("pose_x", np.float64), ("pose_y", np.float64), ("pose_theta", np.float64),
                           ("candidates_scored", np.uint64)])


def pack_results(indices, results):
    """Result2D-like objects (found, scan_index, x/y_index_offset, score, pose_x/y/theta) ->
    int32 [n, RESULT_WORDS]."""
    n = len(indices)
    out = np.zeros((n, RESULT_WORDS), dtype=np.int32)
    out[:, 0] = np.asarray(indices, dtype=np.int32)
    if isinstance(results, ctypes.Array) and ctypes.sizeof(results) == n * RESULT2D_DTYPE.itemsize:
        # the array csmh_match_batch filled: one structured view, no per-result Python work
        r = np.frombuffer(results, dtype=RESULT2D_DTYPE, count=n)
        out[:, 1], out[:, 2], out[:, 3], out[:, 4] = r["found"], r["scan_index"], r["x_index_offset"], r["y_index_offset"]
        out[:, 5] = r["score"].view(np.int32)
        pose = np.stack([r["pose_x"], r["pose_y"], r["pose_theta"]], axis=1)
        out[:, 6:12] = np.ascontiguousarray(pose).view(np.int32).reshape(n, 6)
        return out
    pose = np.zeros((n, 3), dtype=np.float64)
    score = np.zeros(n, dtype=np.float32)
    for k, r in enumerate(results):
        out[k, 1:5] = (r.found, r.scan_index, r.x_index_offset, r.y_index_offset)
        score[k] = r.score
        pose[k] = (r.pose_x, r.pose_y, r.pose_theta)
    out[:, 5] = score.view(np.int32)
    out[:, 6:12] = pose.view(np.int32).reshape(n, 6)
    return out


def gather_packed(packed, shard_sizes, num_searches, device=None, group=None):
    """all_gather of every rank's packed results (padded to the largest shard); returns the
    int32 [num_searches, RESULT_WORDS] array in submission order on every rank.  With
    `device` set the payload travels as a device tensor (NCCL), else on the CPU (gloo)."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    cap = int(max(shard_sizes))
    buf = torch.zeros((cap, RESULT_WORDS), dtype=torch.int32)
    buf[:len(packed)] = torch.from_numpy(np.ascontiguousarray(packed))
    if device is not None:
        buf = buf.to(device, non_blocking=True)
    everything = torch.empty((world, cap, RESULT_WORDS), dtype=torch.int32, device=buf.device)
    dist.all_gather_into_tensor(everything, buf, group=group)
    everything = everything.cpu().numpy()
    out = np.zeros((num_searches, RESULT_WORDS), dtype=np.int32)
    seen = np.zeros(num_searches, dtype=bool)
    for r in range(world):
        rows = everything[r, :int(shard_sizes[r])]
        idx = rows[:, 0]
        assert not seen[idx].any(), "a search ran on two ranks"
        seen[idx] = True
        out[idx] = rows
    assert seen.all(), "%d of %d searches returned" % (int(seen.sum()), num_searches)
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

#!/usr/bin/env python
"""csm_stack2d_broadcast / csm_grid2d_broadcast_dev on N GPUs (one process per GPU, launched by
torchrun): rank 0 owns a finished submap's cells, every rank ends up with a precomputation stack
byte-identical to one built locally from the same cells.  Also times the in-place device broadcast.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
      --master-port 29511 tools/broadcast_check.py
"""
import ctypes as C
import glob
import json
import os
import site
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


class UniqueId(C.Structure):
    _fields_ = [("internal", C.c_char * 128)]  # ncclUniqueId (nccl.h: NCCL_UNIQUE_ID_BYTES = 128)


def nccl_library():
    for sp in site.getsitepackages():
        hits = glob.glob(os.path.join(sp, "nvidia", "nccl", "lib", "libnccl.so.2"))
        if hits:
            return C.CDLL(hits[0], mode=C.RTLD_GLOBAL)
    return C.CDLL("libnccl.so.2", mode=C.RTLD_GLOBAL)


def main():
    import torch
    import torch.distributed as dist
    import synth
    from __graft_entry__ import _load_package
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")  # only carries the 128-byte NCCL id
    nccl = nccl_library()
    uid = UniqueId()
    if rank == 0:
        assert nccl.ncclGetUniqueId(C.byref(uid)) == 0
    box = [C.string_at(C.byref(uid), 128)] if rank == 0 else [None]  # all 128 bytes (the id contains NULs)
    dist.broadcast_object_list(box, src=0)
    assert len(box[0]) == 128
    C.memmove(C.byref(uid), box[0], 128)
    comm = C.c_void_p()
    nccl.ncclCommInitRank.argtypes = [C.POINTER(C.c_void_p), C.c_int, UniqueId, C.c_int]
    torch.zeros(1, device="cuda")  # a CUDA context on this rank's device before NCCL starts
    rc = nccl.ncclCommInitRank(C.byref(comm), world, uid, rank)
    assert rc == 0, "ncclCommInitRank -> %d" % rc

    pkg = _load_package()
    abi = pkg.abi()
    ctx = pkg.Context(local)
    raw = C.c_void_p(ctx.raw)
    out = {}
    for size, depth in ((400, 5), (1000, 7)):
        w = synth.World(size, size, seed=500 + size)  # every rank can rebuild the truth locally
        grid = pkg.Grid(w.cells, w.res, w.max_x, w.max_y)
        lut = grid.quantization_table()
        cells = np.ascontiguousarray(w.cells, np.uint16) if rank == 0 else None
        stack = C.c_void_p()
        abi.csm_stack2d_broadcast.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int,
                                              C.c_void_p, C.c_int, C.POINTER(C.c_void_p)]
        rc = abi.csm_stack2d_broadcast(raw, comm, 0, 1 if rank == 0 else 0,
                                       cells.ctypes.data_as(C.c_void_p) if rank == 0 else None, size, size,
                                       lut.ctypes.data_as(C.c_void_p), depth, C.byref(stack))
        assert rc == 0, abi.csm_last_error()
        local_matcher = pkg.Matcher(ctx, grid, 7.0, 0.5, depth)
        abi.csm_stack2d_level.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.POINTER(C.c_int),
                                          C.POINTER(C.c_int)]
        for lvl in range(depth):
            want = local_matcher.level(lvl)
            got = np.empty_like(want)
            wx, wy = C.c_int(), C.c_int()
            assert abi.csm_stack2d_level(raw, stack, lvl, got.ctypes.data_as(C.c_void_p), C.byref(wx), C.byref(wy)) == 0
            assert got.tobytes() == want.tobytes(), "rank %d size %d level %d differs" % (rank, size, lvl)
        abi.csm_stack2d_destroy.argtypes = [C.c_void_p, C.c_void_p]
        abi.csm_stack2d_destroy(raw, stack)
        # the in-place device broadcast alone, device-timed
        dev = torch.zeros(size * size, dtype=torch.int16, device="cuda")
        if rank == 0:
            dev.copy_(torch.from_numpy(w.cells.view(np.int16).reshape(-1)))
        stream = torch.cuda.ExternalStream(abi.csm_ctx_stream(raw))
        abi.csm_grid2d_broadcast_dev.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_size_t]
        torch.cuda.synchronize()
        for it in range(13):
            if it == 3:
                dist.barrier()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
            assert abi.csm_grid2d_broadcast_dev(raw, comm, 0, C.c_void_p(dev.data_ptr()), size * size) == 0
        e1.record(stream)
        torch.cuda.synchronize()
        assert np.array_equal(dev.cpu().numpy().view(np.uint16).reshape(size, size), w.cells)
        us = torch.tensor([e0.elapsed_time(e1) * 100.0])  # ms over 10 calls -> us per call
        dist.all_reduce(us, op=dist.ReduceOp.MAX)
        out["%dx%d" % (size, size)] = {"bytes": 2 * size * size, "broadcast_us": float(us[0]), "levels_identical": depth}
    nccl.ncclCommDestroy.argtypes = [C.c_void_p]
    nccl.ncclCommDestroy(comm)
    if rank == 0:
        print(json.dumps({"csm_stack2d_broadcast": "ok", "ranks": world, "sizes": out}))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

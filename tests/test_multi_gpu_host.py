"""CPU tests (gloo, world_size 2) of the N > 1 host logic: sharding by cell with HARQ affinity, and the barrier /
max-over-ranks reduction bench.py uses. The data path has no collective (codeblocks are independent)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    import __graft_entry__ as ge
    fec = ge.load_package().fec
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n_cells, cbs_per_tb, slots_per_gpu = 32, 152, 32 * 152 * 4
    # BASELINE config 5: one 152-codeblock TB per cell and slot, 4 HARQ transmissions (rv 0,2,3,1) of each.
    tbs = [dict(cell=c, harq_tx=t, cb_ids=[((c % world) * slots_per_gpu) + (c // world) * cbs_per_tb + i
                                            for i in range(cbs_per_tb)])
           for t in range(4) for c in range(n_cells)]
    mine = fec.shard_transport_blocks(tbs, world, rank)
    assert sorted({tb["cell"] for tb in mine}) == fec.cells_of_rank(n_cells, world, rank)
    # HARQ affinity: every transmission of a codeblock lands on this rank and on the same local slot.
    slots = {}
    for tb in mine:
        for cb in tb["cb_ids"]:
            owner, local = fec.harq_owner(cb, slots_per_gpu, world)
            assert owner == rank
            slots.setdefault(cb, local)
            assert slots[cb] == local
    n_cbs = sum(len(tb["cb_ids"]) for tb in mine)
    # bench.py's reduction: barrier, then MAX of the per-rank device time, SUM of the units processed.
    dist.barrier()
    t = torch.tensor([0.010 * (rank + 1)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    n = torch.tensor([n_cbs], dtype=torch.int64)
    dist.all_reduce(n, op=dist.ReduceOp.SUM)
    ret[rank] = (float(t.item()), int(n.item()), n_cbs)
    dist.destroy_process_group()


def test_sharding_and_reduction_world_size_2():
    world, port = 2, _free_port()
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, port, ret), nprocs=world, join=True)
    assert len(ret) == world
    total = 32 * 152 * 4
    for rank in range(world):
        tmax, nsum, nlocal = ret[rank]
        assert abs(tmax - 0.020) < 1e-12       # max over ranks, not this rank's own time
        assert nsum == total and nlocal == total // world


def test_harq_owner_rejects_out_of_range(fec):
    import pytest
    assert fec.harq_owner(5, 100, 2) == (0, 5)
    assert fec.harq_owner(105, 100, 2) == (1, 5)
    with pytest.raises(fec.FecError):
        fec.harq_owner(200, 100, 2)

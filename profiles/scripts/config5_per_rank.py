# config 5 as rank r of W sees it, on one GPU: per-slot time of the rank's fec_tb_run_host call
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import __graft_entry__ as ge
import bench
fec = ge.load_package().fec
class A: no_cpu_baseline = True
dev = torch.device("cuda", 0)
for W in (1, 2, 4, 8):
    r = bench.bench_config5(fec, torch, dev, A(), 0, W, lambda: None, lambda s: s) if W == 1 else None
    if W > 1:
        # same code path without torch.distributed: shard as rank 0 of W
        import types
        saved = bench.bench_config5.__globals__.get("__dist_stub__")
        mine = fec.cells_of_rank(32, W, 0)
        # reuse the function body by faking world=1 over the rank's cells: simplest is to monkey-patch cells_of_rank
        orig = fec.cells_of_rank
        fec.cells_of_rank = lambda n, w, k, _m=mine: _m
        r = bench.bench_config5(fec, torch, dev, A(), 0, 1, lambda: None, lambda s: s)
        fec.cells_of_rank = orig
    print("W=%d zero_copy=%s ms_per_call %.3f cells %d" % (W, os.environ.get("FEC_B200_HOST_ZEROCOPY", "1"), r["ms_per_slot"], len(fec.cells_of_rank(32, W, 0))))

"""Replays the UL-SCH demultiplexer launches of bench.py (for ncu: kernel durations without the host's launch rate).
Usage: python profiles/scripts/ulsch_only.py [steps]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.argv = sys.argv[:1] + sys.argv[1:]
import importlib.util  # noqa: E402

import torch  # noqa: E402

import __graft_entry__ as ge  # noqa: E402

spec = importlib.util.spec_from_file_location("bench", os.path.join(ROOT, "bench.py"))
bench = importlib.util.module_from_spec(spec)
spec.loader.exec_module(bench)
fec = ge.load_package().fec
eng = fec.FecEngine(device=0, harq_slots=64, max_batch_cbs=64)
dev = torch.device("cuda", 0)
stream = torch.cuda.ExternalStream(eng.stream, device=dev)
print(bench.bench_ulsch_demux(fec, eng, torch, dev, stream))
eng.close()

"""Runs only the soft-demodulator launch of bench.py (for ncu captures)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import bench, __graft_entry__ as ge
fec = ge.load_package().fec
eng = fec.FecEngine(device=0, harq_slots=64, max_batch_cbs=64)
qm, G = 8, 40 * 8960 + 112 * 8992
nsym = G // qm
blocks = [3276 * 4] * (nsym // (3276 * 4)) + [nsym % (3276 * 4)]
n_tb = 46
rng = np.random.default_rng(1)
sym = bench.qam_modulate(bench.gold_bits(12345, G), qm) + (0.02 * rng.standard_normal((nsym, 2))).astype(np.float32)
dm = fec.make_demod_descs(n_tb * len(blocks))
for t in range(n_tb):
    dm[t * len(blocks):(t + 1) * len(blocks)] = fec.codeword_demod_descs(blocks, qm, 12345, sym_offset=t * nsym, llr_offset=t * G)
dev = torch.device("cuda", 0)
d_sym = torch.from_numpy(np.tile(sym, (n_tb, 1))).to(dev)
d_nv = torch.full((n_tb * nsym,), 0.001, dtype=torch.float32, device=dev)
d_llr = torch.zeros(n_tb * G, dtype=torch.int8, device=dev)
d_dm = torch.from_numpy(dm.view(np.uint8).copy()).to(dev)
torch.cuda.synchronize()
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 5):
    eng.demod_launch_device(d_dm.data_ptr(), dm.size, max(blocks), d_sym.data_ptr(), d_nv.data_ptr(), d_llr.data_ptr())
torch.cuda.synchronize()
print("ok", int((d_llr != 0).sum()))

# experiment: one config-3 transport block (152 codeblocks): copy + kernel from HBM  vs  kernel reading pinned host memory
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import __graft_entry__ as ge
import bench
fec = ge.load_package().fec
K = bench.K
n_cb, fill, qm, n_short = 152, 16, 8, 40
Es = np.array([8960] * n_short + [8992] * (n_cb - n_short), np.uint32)
eng = fec.FecEngine(device=0, harq_slots=1024, max_batch_cbs=1024, nof_queues=2)
rng = np.random.default_rng(77)
sigma2 = 0.5 * 10.0 ** (-8.0 / 10.0)
offs = np.concatenate([[0], np.cumsum((Es + 15) // 16 * 16)]).astype(np.uint64)
tb_bytes = int(offs[-1])
y = 1.0 + np.sqrt(sigma2) * rng.standard_normal(tb_bytes, dtype=np.float32)
llr = np.round(np.clip(2.0 * y / sigma2, -20.0, 20.0) * 6.0).astype(np.int8)
d = fec.make_descs(n_cb)
d["base_graph"], d["lifting_size"], d["qm"], d["rv"] = 1, 384, qm, 0
d["E"] = Es
d["nof_filler_bits"], d["crc_type"], d["max_iter"], d["mode"] = fill, fec.CRC24B, 6, fec.MODE_CHAIN
d["flags"] = fec.FLAG_NEW_DATA | fec.FLAG_EARLY_STOP
d["harq_slot"] = np.arange(n_cb)
d["llr_offset"] = offs[:n_cb]
d["out_offset"] = np.arange(n_cb) * (K // 8)
fec.desc_prepare(d)
t_llr = torch.from_numpy(llr).pin_memory()
d_llr = torch.zeros(tb_bytes, dtype=torch.int8, device="cuda")
d_descs = torch.from_numpy(d.view(np.uint8).copy()).cuda()
d_out = torch.zeros(n_cb * (K // 8), dtype=torch.uint8, device="cuda")
h_out = torch.zeros(n_cb * (K // 8), dtype=torch.uint8).pin_memory()
h_out2 = torch.zeros(n_cb * (K // 8), dtype=torch.uint8).pin_memory()
d_res = torch.zeros(n_cb * 8, dtype=torch.uint8, device="cuda")
st = torch.cuda.ExternalStream(eng.stream)
def timed(fn, reps=30):
    ts = []
    for i in range(reps):
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record(st)
        fn()
        b.record(st)
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts = np.array(ts[5:]); return "p50 %.1f us  min %.1f  max %.1f" % (np.median(ts), ts.min(), ts.max())
def copy_then_kernel():
    with torch.cuda.stream(st):
        d_llr.copy_(t_llr, non_blocking=True)
        eng.launch_device(d_descs.data_ptr(), n_cb, d_llr.data_ptr(), d_out.data_ptr(), d_res.data_ptr(), 384, True, even_z=True)
        h_out.copy_(d_out, non_blocking=True)
def kernel_only():
    eng.launch_device(d_descs.data_ptr(), n_cb, d_llr.data_ptr(), d_out.data_ptr(), d_res.data_ptr(), 384, True, even_z=True)
def zero_copy():
    eng.launch_device(d_descs.data_ptr(), n_cb, t_llr.data_ptr(), h_out2.data_ptr(), d_res.data_ptr(), 384, True, even_z=True)
print("H2D + kernel + D2H (device events):", timed(copy_then_kernel))
print("kernel alone, inputs in HBM        :", timed(kernel_only))
print("kernel on pinned host in/out       :", timed(zero_copy))
print("same bytes:", bool((h_out == h_out2).all()))

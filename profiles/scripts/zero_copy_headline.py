# experiment: the codeblock kernel reading its soft bits straight from pinned host memory (zero copy) vs from HBM
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import __graft_entry__ as ge
import bench
fec = ge.load_package().fec
n = 7104
eng = fec.FecEngine(device=0, harq_slots=n, max_batch_cbs=n, nof_queues=2)
llr_h = bench.make_llrs(n, -2.0, seed=1234)
descs = fec.make_descs(n)
bench.fill_descs(fec, descs, n, early_stop=False)
fec.desc_prepare(descs)
t_llr = torch.from_numpy(llr_h).pin_memory()
d_llr = t_llr.cuda()
d_descs = torch.from_numpy(descs.view(np.uint8).copy()).cuda()
d_out = torch.zeros(n * (bench.K // 8), dtype=torch.uint8, device="cuda")
h_out = torch.zeros(n * (bench.K // 8), dtype=torch.uint8).pin_memory()
d_res = torch.zeros(n * 8, dtype=torch.uint8, device="cuda")
st = torch.cuda.ExternalStream(eng.stream)
def run(llr_ptr, out_ptr, reps=5):
    ts = []
    for i in range(reps):
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record(st)
        eng.launch_device(d_descs.data_ptr(), n, llr_ptr, out_ptr, d_res.data_ptr(), 384, True, even_z=True)
        b.record(st)
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return ts
print("HBM in, HBM out      ", [round(x, 3) for x in run(d_llr.data_ptr(), d_out.data_ptr())])
print("pinned in, HBM out   ", [round(x, 3) for x in run(t_llr.data_ptr(), d_out.data_ptr())])
print("pinned in, pinned out", [round(x, 3) for x in run(t_llr.data_ptr(), h_out.data_ptr())])
res = np.frombuffer(d_res.cpu().numpy().tobytes(), fec.RESULT_DTYPE)
print("crc pass fraction", float(((res["status"] & fec.RES_CRC_PASS) != 0).mean()))
ref = d_out.cpu()
print("pinned out equals HBM out:", bool((h_out == ref).all()))

"""Per-codeblock phase clocks of fec_cb_packed_kernel (FEC_PROFILE build: clock64 around set-up / dematch / decode / tail,
packed into fec_cb_result.reserved). Usage on the GPU box:  python profiles/scripts/phase_clocks.py
Builds srsran-5g_b200/csrc with -DFEC_PROFILE into build/lib_prof.so first (nvcc, no GPU needed for that step)."""
import os
import subprocess
import sys

import ctypes as C

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
lib = os.path.join(ROOT, "build", "lib_prof.so")
if "--build" in sys.argv or not os.path.exists(lib):
    os.makedirs(os.path.dirname(lib), exist_ok=True)
    csrc = os.path.join(ROOT, "srsran-5g_b200", "csrc")
    subprocess.check_call(["nvcc", "-DFEC_PROFILE", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-shared",
                           "-Xcompiler", "-fPIC", "-o", lib, "fec_engine.cu", "fec_harq_pool.cpp"], cwd=csrc)
    if "--build" in sys.argv:
        sys.exit(0)
os.environ["FEC_B200_LIB"] = lib
import __graft_entry__ as ge  # noqa: E402

fec = ge.load_package().fec
eng = fec.FecEngine(device=0, harq_slots=8192, max_batch_cbs=8192)
rng = np.random.default_rng(1)


def run(name, n, E, qm, fill, early, esn0, bg=1):
    eng.lib.fec_debug_profile_read((C.c_uint64 * 16)(), 1)
    K, N = fec.dims(bg, 384)
    sigma2 = 0.5 * 10.0 ** (-esn0 / 10.0)
    pad = (E + 15) // 16 * 16
    y = 1.0 + np.sqrt(sigma2) * rng.standard_normal((n, pad), dtype=np.float32)
    llr = np.round(np.clip(2.0 * y / sigma2, -20.0, 20.0) * 6.0).astype(np.int8)
    d = fec.make_descs(n)
    d["base_graph"], d["lifting_size"], d["qm"], d["E"], d["nof_filler_bits"] = bg, 384, qm, E, fill
    d["crc_type"], d["max_iter"], d["mode"] = (fec.CRC24B if bg == 1 else fec.CRC24A), 6, fec.MODE_CHAIN
    d["flags"] = fec.FLAG_NEW_DATA | (fec.FLAG_EARLY_STOP if early else 0)
    d["harq_slot"], d["llr_offset"], d["out_offset"] = np.arange(n), np.arange(n, dtype=np.uint64) * pad, np.arange(n) * (K // 8)
    out = np.zeros(n * (K // 8), np.uint8)
    for _ in range(2):
        res = eng.run_host(d, llr, out)
    r = res["reserved"]
    ph = np.stack([(r & 0xff) << 8, ((r >> 8) & 0xff) << 10, ((r >> 16) & 0xff) << 12, ((r >> 24) & 0xff) << 8], 1)
    # one device-resident launch of the same batch: phase clocks, CTA lifetimes, and the launch as CUDA events see it
    import torch
    fec.desc_prepare(d)
    d_llr = torch.from_numpy(llr.reshape(-1)).cuda()
    d_descs = torch.from_numpy(d.view(np.uint8).copy()).cuda()
    d_out = torch.zeros(out.size, dtype=torch.uint8, device="cuda")
    d_res = torch.zeros(n * 8, dtype=torch.uint8, device="cuda")
    stream = torch.cuda.ExternalStream(eng.stream)
    for _ in range(2):
        eng.launch_device(d_descs.data_ptr(), n, d_llr.data_ptr(), d_out.data_ptr(), d_res.data_ptr(), 384, True, even_z=True)
    torch.cuda.synchronize()
    eng.lib.fec_debug_profile_read((C.c_uint64 * 16)(), 1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    eng.launch_device(d_descs.data_ptr(), n, d_llr.data_ptr(), d_out.data_ptr(), d_res.data_ptr(), 384, True, even_z=True)
    e1.record(stream)
    stream.synchronize()
    acc = (C.c_uint64 * 16)()
    eng.lib.fec_debug_profile_read(acc, 1)
    cnt = max(1, acc[15])
    names = ["setup", "stage", "groups+sig", "fill", "state clear", "rows", "hard bits", "crc", "tail"]
    print("  fine: " + "  ".join("%s %.0f" % (nm, acc[k] / cnt) for k, nm in enumerate(names)) +
          "  = %.0f clocks per codeblock" % (sum(acc[k] for k in range(9)) / cnt))
    nct = max(1, acc[14])
    print("  CTA: alloc %.0f  dealloc+exit %.0f clocks; lifetime %.2f us mean; first start -> last exit %.1f us; CUDA events "
          "%.1f us; sum of lifetimes / 444 slots = %.1f us (%d CTAs)" %
          (acc[9] / nct, acc[10] / nct, acc[13] / nct / 1e3, (acc[12] - acc[11]) / 1e3, e0.elapsed_time(e1) * 1e3,
           acc[13] / 444 / 1e3, nct))
    print("%-28s n=%d  setup %6.0f  dematch+fill %6.0f  decode %7.0f  tail %5.0f  clocks/codeblock (mean); iterations %.2f layers %d" %
          (name, n, *ph.mean(0), res["nof_ldpc_iterations"].mean(), res["nof_layers"][0]))


run("headline (E=N, forced 6)", 444 * 4, 25344, 2, 0, False, -2.0)
run("config 3 (E=8992, early stop, 8 dB)", 444 * 4, 8992, 8, 16, True, 8.0)
run("headline early stop", 444 * 4, 25344, 2, 0, True, -2.0)
run("config 5: 64QAM R873 (E=9816, 45 fillers)", 444 * 4, 9816, 6, 45, True, 6.0)
run("config 5: QPSK R120, BG2 with repetition (E=30976)", 444 * 4, 30976, 2, 184, True, -4.0, bg=2)
eng.close()

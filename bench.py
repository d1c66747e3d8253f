#!/usr/bin/env python3
"""bench.py -- decoded info Gbit/s of the PUSCH FEC decode hot path (rate dematch + HARQ write -> LDPC -> CRC).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--cbs B] [--esn0 dB]

A "step" is one pass of the hot path over one batch of B synthetic codeblocks (BASELINE.json metric shape: BG1, Z=384,
rate ~1/3 i.e. E = N = 25344 LLRs per codeblock, QPSK interleaving, new data, 6 LDPC iterations FORCED (no early stop: the
work is data independent and nothing is skipped), CRC24B check). One JSON line on stdout; see DESIGN.md "Measurement".

  value        device-resident: LLRs already in HBM, one fec_launch_device per step, CUDA events on the engine's stream
  e2e          the same batch through fec_run_host from pinned HOST buffers (H2D + kernels + D2H inside the timed region)
  roofline     the decode kernel against the measured HBM peak (algorithmic bytes per codeblock: DESIGN.md), plus the
               integer-ALU view in "alu"
  cpu_baseline the reference's own best CPU decoder (oracle/_ref: AVX-512 / AVX2 chosen by its own factory) on all host
               cores, bounded sample, rank 0 at N=1 only
  --impl reference   the reference CPU implementation alone, same config / metric / unit

The product path never touches oracle/: only cpu_baseline and --impl reference load oracle/pyoracle.py (the checker).
Inputs are an all-zero codeword (a valid codeword of every linear code, CRC included) through a BPSK-AWGN channel,
quantised like log_likelihood_ratio::quantize(2y/sigma^2, 20) (lib/phy/upper/log_likelihood_ratio.cpp:90-99).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

BG, Z, QM, MAX_ITER = 1, 384, 2, 6
K, N = 22 * Z, 66 * Z
E = N
INFO_BITS = K - 24  # decoded information bits per codeblock: no fillers, the 24-bit codeblock CRC is not information
# Algorithmic HBM bytes per codeblock of this workload: E LLRs in + N soft-buffer bytes out + K/8 decoded
# bytes + 8 result bytes + 48 descriptor bytes.
ALGO_BYTES_PER_CB = E + N + K // 8 + 8 + 48
# Edge-message updates per codeblock: iterations x Z x sum of row degrees (316 for the 46 rows of BG1).
EDGE_UPDATES_PER_CB = MAX_ITER * Z * 316
LANE_OPS_PER_UPDATE = 26  # canonical scalar cost of the reference's generic code per update


def make_llrs(n, esn0_db, seed):
    """n x E int8 LLRs: all-zero codeword, y = 1 + sigma*n, llr = quantize(2y/sigma^2, range 20) -> [-120, 120]."""
    rng = np.random.default_rng(seed)
    sigma2 = 0.5 * 10.0 ** (-esn0_db / 10.0)
    out = np.empty((n, E), np.int8)
    chunk = 256
    for i in range(0, n, chunk):
        m = min(chunk, n - i)
        y = 1.0 + np.sqrt(sigma2) * rng.standard_normal((m, E), dtype=np.float32)
        v = np.clip(2.0 * y / sigma2, -20.0, 20.0) / 20.0 * 120.0
        out[i:i + m] = np.round(v).astype(np.int8)
    return out


def fill_descs(fec, d, n, early_stop, slot0=0):
    d["base_graph"], d["lifting_size"], d["qm"], d["E"], d["rv"] = BG, Z, QM, E, 0
    d["crc_type"], d["max_iter"], d["mode"] = fec.CRC24B, MAX_ITER, fec.MODE_CHAIN
    d["flags"] = fec.FLAG_NEW_DATA | (fec.FLAG_EARLY_STOP if early_stop else 0)
    d["harq_slot"] = slot0 + np.arange(n)
    d["llr_offset"] = np.arange(n, dtype=np.uint64) * E
    d["out_offset"] = np.arange(n) * (K // 8)
    fec.desc_prepare(d)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    FIELDS = ["clocks.sm", "clocks.max.sm", "power.draw", "clocks_event_reasons.hw_slowdown",
              "clocks_event_reasons.hw_thermal_slowdown", "clocks_event_reasons.sw_thermal_slowdown",
              "clocks_event_reasons.sw_power_cap"]

    def __init__(self, device):
        self.device, self.rows, self.proc = device, [], None
        self.nvml_rows, self.nvml_stop, self.nvml_thread = [], None, None

    def _start_nvml(self):
        """NVML in-process, one sample every 2 ms: a 50 ms timed region still gets a couple of dozen samples under load
        (nvidia-smi -lms 100 gets one at best; it stays as the fallback)."""
        import pynvml
        pynvml.nvmlInit()
        try:
            import torch
            uuid = "GPU-" + str(torch.cuda.get_device_properties(self.device).uuid)
            h = pynvml.nvmlDeviceGetHandleByUUID(uuid)
        except Exception:
            h = pynvml.nvmlDeviceGetHandleByIndex(self.device)
        smax = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
        bits = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}
        self.nvml_stop = threading.Event()

        def loop():
            while not self.nvml_stop.is_set():
                try:
                    r = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                    self.nvml_rows.append((time.time(), float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)), smax,
                                           pynvml.nvmlDeviceGetPowerUsage(h) / 1e3, [k for k, b in bits.items() if r & b]))
                except Exception:
                    pass
                time.sleep(0.002)

        self.nvml_thread = threading.Thread(target=loop, daemon=True)
        self.nvml_thread.start()

    def start(self):
        try:
            self._start_nvml()
            return
        except Exception:
            self.nvml_thread = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.device), "--query-gpu=" + ",".join(self.FIELDS),
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def stop(self, t0, t1):
        if self.nvml_thread is not None:
            self.nvml_stop.set()
            self.nvml_thread.join(timeout=1.0)
            rows = [r for r in self.nvml_rows if t0 <= r[0] <= t1] or self.nvml_rows[-3:]
            if rows:
                return {"sm_mhz": float(np.median([r[1] for r in rows])), "sm_max_mhz": rows[0][2],
                        "power_w_max": max(r[3] for r in rows), "samples": len(rows),
                        "reasons": sorted({x for r in rows for x in r[4]}), "source": "NVML, one sample per 2 ms inside the timed region"}
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for t, r in self.rows if t0 <= t <= t1 + 0.1] or [r for _, r in self.rows[-3:]]
        sm, smax, reasons, power = [], [], set(), []
        for r in rows:
            try:
                sm.append(float(r[0]))
                smax.append(float(r[1]))
                power.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}




def reference_cpu(llr, n_threads, seconds, early_stop, impl=None):
    """Times the reference's CPU chain (dematch + LDPC + CRC; oracle/_ref = unmodified reference sources) on a bounded
    sample: about `seconds` of wall clock; the throughput counts the parallel decode time reported by the shim (threads
    started and decoders constructed before its clock starts). Returns (Gbit/s, description dict)."""
    orc = ge.load_oracle()
    if not orc.Ref.available():
        return None, {"unavailable": "oracle/_ref/libref_fec.so was not shipped"}
    ref = orc.Ref()
    if impl is None:
        impl = orc.Ref.AVX512 if ref.impl_available(orc.Ref.AVX512) else (
            orc.Ref.AVX2 if ref.impl_available(orc.Ref.AVX2) else orc.Ref.GENERIC)
    name = {orc.Ref.GENERIC: "generic", orc.Ref.AVX2: "avx2", orc.Ref.AVX512: "avx512"}[impl]
    args_ = (True, 0, QM, 0, 0, orc.CRC24B, MAX_ITER, early_stop)
    ref.batch_chain(impl, n_threads, BG, Z, llr[:n_threads], *args_)  # warm-up: decoder objects, page faults
    total_s, total_n, n_pass, t_start = 0.0, 0, 0, time.time()
    while time.time() - t_start < seconds:  # wall clock, so that the step stays bounded whatever the call overhead
        sec, iters, _ = ref.batch_chain(impl, n_threads, BG, Z, llr, *args_)
        total_s += sec
        total_n += len(llr)
        n_pass += int((iters > 0).sum())
    gbps = total_n * INFO_BITS / total_s / 1e9
    return gbps, {"kind": "reference", "cores": n_threads, "impl": "ldpc_decoder_" + name + " + ldpc_rate_dematcher_" + name,
                  "sample": "%d codeblocks (%d distinct) of the same workload (BG1 Z=384 E=%d, %d iterations, "
                            "early_stop=%s) in %.2f s of parallel decode time" % (total_n, len(llr), E, MAX_ITER,
                                                                                  early_stop, total_s),
                  "crc_pass_fraction": n_pass / max(total_n, 1)}


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def run_reference(args):
    """--impl reference: the reference's own CPU implementation, all host threads, same config / metric / unit."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    threads = host_threads()
    per_step = min(20.0, 150.0 / max(1, args.steps + args.warmup))
    llr = make_llrs(max(32 * threads, 256), args.esn0, seed=1)
    vals, desc = [], None
    for s in range(args.warmup + args.steps):
        t0 = time.time()
        gbps, desc = reference_cpu(llr, threads, per_step, early_stop=False)
        if gbps is None:
            print(json.dumps({"impl": "reference", "unavailable": desc["unavailable"]}))
            return 0
        if s >= args.warmup:
            vals.append((gbps, time.time() - t0))
    value = float(np.mean([v for v, _ in vals]))
    desc["value"] = value
    desc["unit"] = "Gbit/s"
    line = {"impl": "reference", "metric": "decoded_info_gbps", "value": value, "unit": "Gbit/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": float(np.mean([t for _, t in vals]) * 1e3),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int8", "data": "synthetic",
            "config": workload_config(args.cbs), "cpu_baseline": desc,
            "e2e": {"value": value, "unit": "Gbit/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))
    return 0


def workload_config(n_cbs):
    return {"workload": "PUSCH FEC decode chain, BG1 Z=384 K=8448, E=N=25344 (rate ~1/3), Qm=2, rv0 new data, "
                        "6 LDPC iterations forced (no early stop), CRC24B, scaled min-sum 0.8",
            "codeblocks_per_step_per_gpu": n_cbs, "max_iterations": MAX_ITER, "early_stop": False,
            "l2_policy": "inputs larger than L2 (%d MB of LLRs + as much soft buffer per step)"
                         % ((n_cbs or 0) * E // 1000000)}


def bench_transport_block(fec, eng, torch, dev, stream, args, rank):
    n_cb, fill, qm, n_short = 152, 16, 8, 40
    Es = np.array([8960] * n_short + [8992] * (n_cb - n_short), np.uint32)
    info_bits = n_cb * (K - fill - 24) - 24  # TBS = payload of all CBs minus the TB CRC
    n_tb = max(1, min(args.cbs // n_cb, 46))
    rng = np.random.default_rng(77 + rank)
    sigma2 = 0.5 * 10.0 ** (-args.tb_esn0 / 10.0)
    offs = np.concatenate([[0], np.cumsum((Es + 15) // 16 * 16)]).astype(np.uint64)
    tb_bytes = int(offs[-1])
    llr = np.zeros((n_tb, tb_bytes), np.int8)
    for t in range(n_tb):
        y = 1.0 + np.sqrt(sigma2) * rng.standard_normal(tb_bytes, dtype=np.float32)
        llr[t] = np.round(np.clip(2.0 * y / sigma2, -20.0, 20.0) * 6.0).astype(np.int8)
    d = fec.make_descs(n_tb * n_cb)
    d["base_graph"], d["lifting_size"], d["qm"], d["rv"] = BG, Z, qm, 0
    d["E"] = np.tile(Es, n_tb)
    d["nof_filler_bits"], d["crc_type"], d["max_iter"], d["mode"] = fill, fec.CRC24B, MAX_ITER, fec.MODE_CHAIN
    d["flags"] = fec.FLAG_NEW_DATA | fec.FLAG_EARLY_STOP
    d["harq_slot"] = np.arange(n_tb * n_cb)
    d["llr_offset"] = (np.arange(n_tb, dtype=np.uint64)[:, None] * np.uint64(tb_bytes) + offs[None, :n_cb]).ravel()
    d["out_offset"] = np.arange(n_tb * n_cb) * (K // 8)
    fec.desc_prepare(d)
    t_llr = torch.from_numpy(llr).pin_memory()
    t_out = torch.zeros(n_tb * n_cb * (K // 8), dtype=torch.uint8).pin_memory()
    res = np.zeros(n_tb * n_cb, fec.RESULT_DTYPE)
    # latency of ONE transport block, host buffers, synchronous call
    one = d[:n_cb].copy()
    lat = []
    for i in range(220):
        t = i % n_tb
        one["llr_offset"] = d["llr_offset"][t * n_cb:(t + 1) * n_cb]
        t0 = time.perf_counter()
        eng.lib.fec_run_host(eng.h, one.ctypes.data, n_cb, t_llr.data_ptr(), n_tb * tb_bytes, t_out.data_ptr(),
                             n_cb * (K // 8), res.ctypes.data, None)
        lat.append((time.perf_counter() - t0) * 1e6)
    lat = np.array(lat[20:])
    r1 = res[:n_cb].copy()
    # the same transport block through the transport-block entry point: one descriptor, segmentation in the engine,
    # TB assembly + TB CRC on the device, the TB bytes (not 152 codeblock buffers) come back
    tbd = fec.make_tb_descs(1)
    tbd["tbs"], tbd["nof_llrs"], tbd["base_graph"], tbd["qm"], tbd["nof_layers"] = info_bits, int(Es.sum()), BG, qm, 4
    tbd["max_iter"], tbd["flags"], tbd["harq_slot"] = MAX_ITER, fec.FLAG_NEW_DATA | fec.FLAG_EARLY_STOP, 0
    # (the synthetic codeword is all-zero: a zero payload has zero CRCs, so the TB CRC passes like the codeblock CRCs)
    t_tb = torch.zeros(info_bits // 8, dtype=torch.uint8).pin_memory()
    tb_llr = torch.from_numpy(np.concatenate([llr[0][offs[r]:offs[r] + Es[r]] for r in range(n_cb)])).pin_memory()
    lat_tb, tb_ok = [], 0
    for i in range(220):
        t0 = time.perf_counter()
        rr_tb = eng.tb_run_host(tbd, (tb_llr.data_ptr(), tb_llr.numel()), (t_tb.data_ptr(), t_tb.numel()))
        lat_tb.append((time.perf_counter() - t0) * 1e6)
        tb_ok += int(rr_tb["tb_crc_ok"][0])
    lat_tb = np.array(lat_tb[20:])
    # a slot's worth of transport blocks (n_tb of them, e.g. many cells) in ONE fec_tb_run_host call from pinned host memory
    tbn = fec.make_tb_descs(n_tb)
    tbn["tbs"], tbn["nof_llrs"], tbn["base_graph"], tbn["qm"], tbn["nof_layers"] = info_bits, int(Es.sum()), BG, qm, 4
    tbn["max_iter"], tbn["flags"] = MAX_ITER, fec.FLAG_NEW_DATA | fec.FLAG_EARLY_STOP
    tbn["harq_slot"] = np.arange(n_tb) * n_cb
    tbn["llr_offset"] = np.arange(n_tb, dtype=np.uint64) * np.uint64(tb_bytes)
    tbn["out_offset"] = np.arange(n_tb, dtype=np.uint64) * np.uint64(info_bits // 8)
    t_tbn = torch.zeros(n_tb * (info_bits // 8), dtype=torch.uint8).pin_memory()
    batch_s, batch_ok = [], 0
    for i in range(8):
        t0 = time.perf_counter()
        rb = eng.tb_run_host(tbn, (t_llr.data_ptr(), t_llr.numel()), (t_tbn.data_ptr(), t_tbn.numel()))
        batch_s.append(time.perf_counter() - t0)
        batch_ok = int(rb["tb_crc_ok"].sum())
    batch_s = float(np.median(batch_s[2:]))
    # device-resident throughput of n_tb transport blocks per launch
    d_llr = t_llr.to(dev)
    d_descs = torch.from_numpy(d.view(np.uint8).copy()).to(dev)
    d_out = torch.zeros(n_tb * n_cb * (K // 8), dtype=torch.uint8, device=dev)
    d_res = torch.zeros(n_tb * n_cb * 8, dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()

    def step():
        eng.launch_device(d_descs.data_ptr(), n_tb * n_cb, d_llr.data_ptr(), d_out.data_ptr(), d_res.data_ptr(), Z, True, even_z=True)

    for _ in range(3):
        step()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        e0.record(stream)
        for _ in range(10):
            step()
        e1.record(stream)
    stream.synchronize()
    ms = e0.elapsed_time(e1) / 10
    rr = np.frombuffer(d_res.cpu().numpy().tobytes(), fec.RESULT_DTYPE)
    return {"workload": "273 PRB x 4 layers 256QAM R948: TBS %d bits, 152 CBs BG1 Z=384 E=8960/8992 Qm=8, early stop, max 6 "
                        "iterations, Es/N0 %.1f dB (BPSK-equivalent per coded bit)" % (info_bits, args.tb_esn0),
            "tb_latency_us": {"p50": float(np.percentile(lat, 50)), "p99": float(np.percentile(lat, 99)),
                              "max": float(lat.max()), "samples": int(lat.size),
                              "api": "fec_run_host, 152 codeblocks per call, pinned host buffers, H2D+kernel+D2H"},
            "tb_api_latency_us": {"p50": float(np.percentile(lat_tb, 50)), "p99": float(np.percentile(lat_tb, 99)),
                                  "max": float(lat_tb.max()), "samples": int(lat_tb.size), "tb_crc_ok_fraction": tb_ok / 220,
                                  "api": "fec_tb_run_host: one descriptor per transport block, segmentation in the engine, "
                                         "codeblock kernel + TB assembly/CRC24A kernel, TB bytes back"},
            "device_resident": {"value": n_tb * info_bits / (ms * 1e-3) / 1e9, "unit": "Gbit/s", "tbs_per_launch": n_tb,
                                "ms_per_launch": ms,
                                # this shape against its two roofs: HBM with the algorithmic bytes of SURVEY 8(d) (E in, N soft
                                # buffer out, K/8 out, descriptor and result), and instruction issue with the edge updates the
                                # rows actually ran (iterations x layers x 19 edges of the first base rows x Z)
                                "roofline": config3_roofline(fec, rr, Es, n_tb, ms, torch, dev)},
            "batch_e2e": {"value": n_tb * info_bits / batch_s / 1e9, "unit": "Gbit/s", "transport_blocks_per_call": n_tb,
                          "ms_per_call": batch_s * 1e3, "tb_crc_ok": batch_ok, "h2d_bytes_per_call": int(n_tb * tb_bytes),
                          "d2h_bytes_per_call": int(n_tb * (info_bits // 8)),
                          "api": "fec_tb_run_host, pinned host buffers: copies, codeblock and gather kernels pipelined in "
                                 "groups of whole transport blocks of about one wave"},
            "cb_crc_pass_fraction": float(((rr["status"] & fec.RES_CRC_PASS) != 0).mean()),
            "mean_iterations": float(rr["nof_ldpc_iterations"].mean()), "layers": int(rr["nof_layers"].max()),
            "first_tb_pass_fraction": float(((r1["status"] & fec.RES_CRC_PASS) != 0).mean())}


def config3_roofline(fec, rr, Es, n_tb, ms, torch, dev):
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak_gbs = float(peaks.get("hbm_gbs", 6650.0))
    n_cb = Es.size
    algo = float(n_tb * (Es.astype(np.float64) + N + K // 8 + 8 + 48).sum())
    sm = torch.cuda.get_device_properties(dev).multi_processor_count
    ipe = sass_instructions_per_edge() or 34.7
    # BG1 rows 0..3 have 19 edges each; rows beyond have fewer, counted exactly from the layer count of each codeblock
    deg = np.array([19, 19, 19, 19, 3, 8, 9, 7, 10, 9, 7, 8, 7, 6, 7, 7, 6, 6, 6, 6, 6, 6, 5, 5, 6, 5, 5, 4, 5, 5, 5, 5, 5, 5, 5,
                    5, 5, 4, 5, 5, 4, 5, 4, 5, 5, 4])
    cum = np.concatenate([[0], np.cumsum(deg)])
    upd = float((rr["nof_ldpc_iterations"].astype(np.float64) * cum[np.minimum(rr["nof_layers"], 46)] * Z).sum())
    issue_peak = sm * 4 * 1965e6 * 64.0 / ipe
    return {"bound": "issue", "achieved": upd / (ms * 1e-3) / 1e9, "peak": issue_peak / 1e9, "unit": "G edge updates/s",
            "frac": upd / (ms * 1e-3) / issue_peak,
            "note": "the row bodies are 45 % of this shape's instructions (profiles/r02_phase_clocks.txt): dematcher, hard decision "
                    "and CRC per iteration are the rest",
            "hbm": {"achieved": algo / (ms * 1e-3) / 1e9, "peak": peak_gbs, "unit": "GB/s", "frac": algo / (ms * 1e-3) / 1e9 / peak_gbs,
                    "algorithmic_bytes_per_cb": algo / (n_tb * n_cb)}}


def bench_pdsch_encode(fec, eng, torch, dev, stream, args):
    """Downlink twin: LDPC encoding + rate matching of PDSCH codeblocks, one CTA per codeblock.
    Two shapes: the headline one (BG1 Z=384, E=25344, Qm=2) and the config-3 one (E=8960, Qm=8, 16 fillers). Reported as
    encoded INFO Gbit/s (K - fillers bits per codeblock), device-resident and through fec_encode_run_host (pinned host)."""
    out = {}
    n = args.cbs
    rng = np.random.default_rng(4242)
    for name, e_bits, qm, fill in (("rate_1_3", E, QM, 0), ("config3", 8960, 8, 16)):
        seg_bytes, out_bytes = (K - fill + 7) // 8, (e_bits + 7) // 8
        out_stride = (out_bytes + 3) & ~3
        d = fec.make_enc_descs(n)
        d["base_graph"], d["lifting_size"], d["qm"], d["E"], d["rv"], d["nof_filler_bits"] = BG, Z, qm, e_bits, 0, fill
        d["in_offset"] = np.arange(n) * seg_bytes
        d["out_offset"] = np.arange(n, dtype=np.uint64) * out_stride
        fec.enc_desc_prepare(d)
        t_in = torch.from_numpy(rng.integers(0, 256, n * seg_bytes, dtype=np.uint8)).pin_memory()
        t_out = torch.zeros(n * out_stride, dtype=torch.uint8).pin_memory()
        d_in, d_o = t_in.to(dev), torch.zeros(n * out_stride, dtype=torch.uint8, device=dev)
        d_descs = torch.from_numpy(d.view(np.uint8).copy()).to(dev)
        torch.cuda.synchronize()

        def step():
            eng.encode_launch_device(d_descs.data_ptr(), n, d_in.data_ptr(), d_o.data_ptr(), Z)

        for _ in range(3):
            step()
        reps = 20
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            e0.record(stream)
            for _ in range(reps):
                step()
            e1.record(stream)
        stream.synchronize()
        ms = e0.elapsed_time(e1) / reps
        for _ in range(2):
            eng.encode_run_host(d, (t_in.data_ptr(), t_in.numel()), (t_out.data_ptr(), t_out.numel()))
        t0 = time.perf_counter()
        for _ in range(5):
            eng.encode_run_host(d, (t_in.data_ptr(), t_in.numel()), (t_out.data_ptr(), t_out.numel()))
        e2e_s = (time.perf_counter() - t0) / 5
        assert (t_out.numpy() == d_o.cpu().numpy()).all()
        info = n * (K - fill)
        algo = n * (seg_bytes + out_bytes + 40)
        out[name] = {"workload": "BG1 Z=384, %d fillers, E=%d, Qm=%d, rv0, packed output, %d codeblocks per launch"
                                 % (fill, e_bits, qm, n),
                     "value": info / (ms * 1e-3) / 1e9, "unit": "Gbit/s (info bits encoded)",
                     "coded_gbps": n * e_bits / (ms * 1e-3) / 1e9, "ms_per_launch": ms,
                     "e2e": {"value": info / e2e_s / 1e9, "unit": "Gbit/s", "h2d_bytes_per_step": n * (seg_bytes + 40),
                             "d2h_bytes_per_step": n * out_stride, "api": "fec_encode_run_host (pinned host buffers)"},
                     "roofline": {"bound": "hbm", "achieved": algo / (ms * 1e-3) / 1e9, "unit": "GB/s",
                                  "algorithmic_bytes_per_cb": seg_bytes + out_bytes + 40,
                                  "note": "bit gather out of shared memory: compute/LSU-bound, far from the HBM roof"}}
    if not args.no_cpu_baseline:
        try:
            orc = ge.load_oracle()
            if orc.Ref.available():
                ref = orc.Ref()
                msg = rng.integers(0, 2, K).astype(np.uint8)
                ref.rate_match(BG, Z, ref.encode(BG, Z, msg), E, 0, QM, 0, 0)
                t0, cnt = time.perf_counter(), 0
                while time.perf_counter() - t0 < 2.0:
                    ref.rate_match(BG, Z, ref.encode(BG, Z, msg), E, 0, QM, 0, 0)
                    cnt += 1
                out["cpu_baseline"] = {"value": cnt * K / (time.perf_counter() - t0) / 1e9, "unit": "Gbit/s", "cores": 1,
                                       "kind": "reference", "sample": "%d codeblocks of the rate_1_3 shape through the "
                                       "reference's ldpc_encoder + ldpc_rate_matcher (oracle/_ref shim, one thread, "
                                       "bit pack/unpack of the shim included)" % cnt}
        except Exception as ex:  # the checker is optional here
            out["cpu_baseline"] = {"unavailable": str(ex)[:100]}
    return out


def measured_hbm_peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    except (OSError, KeyError, ValueError):
        return 6650.0, "fallback 6650 GB/s"


def gold_bits(c_init, nbits):
    """TS 38.211 5.2.1 sequence, 28 bits per step on Python integers (input generation only)."""
    x1, x2, out = 1, c_init & 0x7fffffff, []
    m28 = (1 << 28) - 1

    def step(x, taps):
        nxt = 0
        for t in taps:
            nxt ^= x >> t
        return ((x >> 28) | ((nxt & m28) << 3)) & 0x7fffffff

    bits = np.zeros(((1600 + nbits) // 28 + 2) * 28, np.uint8)
    for k in range(bits.size // 28):
        c = (x1 ^ x2) & m28
        bits[k * 28:(k + 1) * 28] = (c >> np.arange(28)) & 1
        x1, x2 = step(x1, (0, 3)), step(x2, (0, 1, 2, 3))
    return bits[1600:1600 + nbits]


def qam_modulate(bits, qm):
    """TS 38.211 5.1 mapper (input generation only)."""
    s = 1 - 2 * bits.reshape(-1, qm).astype(np.int64)
    m = qm // 2
    re, im = s[:, qm - 2], s[:, qm - 1]
    for i in range(m - 2, -1, -1):
        re, im = s[:, 2 * i] * ((1 << (m - 1 - i)) - re), s[:, 2 * i + 1] * ((1 << (m - 1 - i)) - im)
    return np.stack([re, im], 1).astype(np.float32) / np.float32(np.sqrt({2: 2.0, 4: 10.0, 6: 42.0, 8: 170.0}[qm]))


def bench_pusch_demod(fec, eng, torch, dev, stream, args):
    """The step in front of the path: soft demodulation + descrambling on the device. Workload: the
    config-3 codeword (1,365,504 soft bits of 256QAM = 170,688 equalised symbols, one demodulation call per OFDM symbol of
    273 PRB x 4 layers), all-zero transport block. Device-resident throughput of the kernel, and the latency of one
    transport block from symbols (fec_tb_run_host_symbols) next to the LLR-input call of bench_transport_block."""
    n_cb, fill, qm = 152, 16, 8
    Es = np.array([8960] * 40 + [8992] * (n_cb - 40), np.uint32)
    G = int(Es.sum())
    info_bits = n_cb * (K - fill - 24) - 24
    nsym = G // qm
    blocks = [3276 * 4] * (nsym // (3276 * 4)) + ([nsym % (3276 * 4)] if nsym % (3276 * 4) else [])
    c_init = 0x4601 * 32768 + 41
    rng = np.random.default_rng(99)
    nv = 10.0 ** (-args.demod_esn0 / 10.0)
    sym = qam_modulate(gold_bits(c_init, G), qm)  # the all-zero codeword, scrambled
    sym = sym + (np.sqrt(nv / 2) * rng.standard_normal(sym.shape)).astype(np.float32)
    nvs = np.full(nsym, nv, np.float32)
    n_tb = max(1, min(args.cbs // n_cb, 46))
    dm1 = fec.codeword_demod_descs(blocks, qm, c_init)
    dm = fec.make_demod_descs(n_tb * len(blocks))
    for t in range(n_tb):
        dm[t * len(blocks):(t + 1) * len(blocks)] = fec.codeword_demod_descs(blocks, qm, c_init, sym_offset=t * nsym,
                                                                             llr_offset=t * G)
    d_sym = torch.from_numpy(np.tile(sym, (n_tb, 1))).to(dev)
    d_nv = torch.from_numpy(np.tile(nvs, n_tb)).to(dev)
    d_llr = torch.zeros(n_tb * G, dtype=torch.int8, device=dev)
    d_dm = torch.from_numpy(dm.view(np.uint8).copy()).to(dev)
    torch.cuda.synchronize()

    def step():
        eng.demod_launch_device(d_dm.data_ptr(), dm.size, max(blocks), d_sym.data_ptr(), d_nv.data_ptr(), d_llr.data_ptr())

    for _ in range(3):
        step()
    reps = 20
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        e0.record(stream)
        for _ in range(reps):
            step()
        e1.record(stream)
    stream.synchronize()
    ms = e0.elapsed_time(e1) / reps
    # symbols -> decoded codeblocks with everything resident in HBM: the demodulator launch followed by the codeblock launch
    # that reads its soft bits (n_tb x 152 codeblocks, early stop as in bench_transport_block)
    cbd = fec.make_descs(n_tb * n_cb)
    cbd["base_graph"], cbd["lifting_size"], cbd["qm"], cbd["rv"] = BG, Z, qm, 0
    cbd["E"] = np.tile(Es, n_tb)
    cbd["nof_filler_bits"], cbd["crc_type"], cbd["max_iter"], cbd["mode"] = fill, fec.CRC24B, MAX_ITER, fec.MODE_CHAIN
    cbd["flags"] = fec.FLAG_NEW_DATA | fec.FLAG_EARLY_STOP
    cbd["harq_slot"] = np.arange(n_tb * n_cb)
    cw_off = np.concatenate([[0], np.cumsum(Es)[:-1]]).astype(np.uint64)
    cbd["llr_offset"] = (np.arange(n_tb, dtype=np.uint64)[:, None] * np.uint64(G) + cw_off[None, :]).ravel()
    cbd["out_offset"] = np.arange(n_tb * n_cb) * (K // 8)
    fec.desc_prepare(cbd)
    d_cbd = torch.from_numpy(cbd.view(np.uint8).copy()).to(dev)
    d_out = torch.zeros(n_tb * n_cb * (K // 8), dtype=torch.uint8, device=dev)
    d_res = torch.zeros(n_tb * n_cb * 8, dtype=torch.uint8, device=dev)

    def fused():
        step()
        eng.launch_device(d_cbd.data_ptr(), n_tb * n_cb, d_llr.data_ptr(), d_out.data_ptr(), d_res.data_ptr(), Z, True, even_z=True)

    for _ in range(3):
        fused()
    with torch.cuda.stream(stream):
        e0.record(stream)
        for _ in range(10):
            fused()
        e1.record(stream)
    stream.synchronize()
    ms_fused = e0.elapsed_time(e1) / 10
    rr = np.frombuffer(d_res.cpu().numpy().tobytes(), fec.RESULT_DTYPE)
    # one transport block from symbols, pinned host buffers
    tbd = fec.make_tb_descs(1)
    tbd["tbs"], tbd["nof_llrs"], tbd["base_graph"], tbd["qm"], tbd["nof_layers"] = info_bits, G, BG, qm, 4
    tbd["max_iter"], tbd["flags"], tbd["harq_slot"] = MAX_ITER, fec.FLAG_NEW_DATA | fec.FLAG_EARLY_STOP, 0
    t_sym, t_nv = torch.from_numpy(sym).pin_memory(), torch.from_numpy(nvs).pin_memory()
    t_tb = torch.zeros(info_bits // 8, dtype=torch.uint8).pin_memory()
    lat, ok = [], 0
    for i in range(220):
        t0 = time.perf_counter()
        r = eng.tb_run_host_symbols(tbd, dm1, (t_sym.data_ptr(), nsym), (t_nv.data_ptr(), nsym),
                                    (t_tb.data_ptr(), t_tb.numel()))
        lat.append((time.perf_counter() - t0) * 1e6)
        ok += int(r["tb_crc_ok"][0])
    lat = np.array(lat[20:])
    algo = n_tb * nsym * (12 + qm)
    peak, peak_src = measured_hbm_peak()
    out = {"workload": "config-3 codeword: %d symbols of 256QAM -> %d soft bits, %d demodulation calls (one per OFDM symbol "
                       "of 273 PRB x 4 layers), descrambling included, reference flavour avx512, Es/N0 %.1f dB"
                       % (nsym, G, len(blocks), args.demod_esn0),
           "value": n_tb * G / (ms * 1e-3) / 1e9, "unit": "G soft bits/s", "codewords_per_launch": n_tb, "ms_per_launch": ms,
           "roofline": {"bound": "hbm", "achieved": algo / (ms * 1e-3) / 1e9, "peak": peak, "peak_source": peak_src,
                        "unit": "GB/s", "frac": algo / (ms * 1e-3) / 1e9 / peak,
                        "algorithmic_bytes_per_symbol": 12 + qm,
                        # dram__bytes_read.sum + dram__bytes_write.sum of one launch of this workload (ncu --set full,
                        # profiles/r01_ncu_demod_full_metrics.csv: 94.4 MB read + 30.6 MB written for 7 851 648 symbols; the
                        # rest of the 62.8 MB of soft bits is still in L2 when the kernel ends), scaled to this launch
                        "traffic": n_tb * nsym * 15.9, "traffic_unit": "bytes/launch",
                        "kernel": "fec_demod_kernel (1 launch/step)",
                        "note": "issue-bound: 66 % of the issue slots busy, ALU pipe 54 %; exact emulation of the "
                                "reference's float pipeline costs ~330 instructions per 256QAM symbol"},
           "symbols_to_codeblocks_device_resident": {
               "value": n_tb * info_bits / (ms_fused * 1e-3) / 1e9, "unit": "Gbit/s (decoded info bits)", "ms_per_step": ms_fused,
               "launches_per_step": 2, "cb_crc_pass_fraction": float(((rr["status"] & fec.RES_CRC_PASS) != 0).mean()),
               "mean_iterations": float(rr["nof_ldpc_iterations"].mean()),
               "note": "demodulator launch + codeblock launch on the same stream, %d transport blocks of 152 codeblocks" % n_tb},
           "tb_from_symbols_latency_us": {"p50": float(np.percentile(lat, 50)), "p99": float(np.percentile(lat, 99)),
                                          "samples": int(lat.size), "tb_crc_ok_fraction": ok / 220,
                                          "h2d_bytes": nsym * 12, "d2h_bytes": info_bits // 8,
                                          "api": "fec_tb_run_host_symbols: symbols + noise variances in, demodulation, "
                                                 "descrambling, segmentation, decoding, TB CRC on the device, TB bytes back"}}
    if not args.no_cpu_baseline:
        try:
            orc = ge.load_oracle()
            if orc.RefDemod.available(2):
                ref = orc.RefDemod(2)
                llr = np.zeros(G, np.int8)
                t0, cnt = time.perf_counter(), 0
                while time.perf_counter() - t0 < 2.0:
                    pos = 0
                    for b in blocks:
                        llr[pos * qm:(pos + b) * qm] = ref.demodulate_soft(qm, sym[pos:pos + b], nvs[pos:pos + b])
                        pos += b
                    ref.apply_xor_llr(c_init, 0, llr)
                    cnt += 1
                out["cpu_baseline"] = {"value": cnt * G / (time.perf_counter() - t0) / 1e9, "unit": "G soft bits/s", "cores": 1,
                                       "kind": "reference", "sample": "%d codewords through the reference's "
                                       "demodulation_mapper_impl (AVX-512 build) + pseudo_random_generator apply_xor, "
                                       "one thread (oracle/_ref/libref_demod_avx512.so)" % cnt}
        except Exception as ex:  # the checker is optional here
            out["cpu_baseline"] = {"unavailable": str(ex)[:100]}
    return out


def bench_ulsch_demux(fec, eng, torch, dev, stream):
    """UL-SCH demultiplexing (fec_ulsch_demux_launch_device) on the config-3 allocation (273 PRB x 4 layers 256QAM, 13 data
    symbols + 1 DM-RS symbol: 1.36 M soft bits) carrying a 2-bit HARQ-ACK on reserved elements, CSI Part 1 and CSI Part 2:
    46 codewords back to back, device-resident. HBM-bound byte moving: reads the codeword once, writes every soft bit once."""
    cfg = fec.ulsch_cfg(qm=8, nof_layers=4, nof_prb=273, start_symbol=0, nof_symbols=14, dmrs_type=1, dmrs_mask=0b100,
                        nof_cdm_groups_without_data=2, nof_harq_ack_rvd=64 * 32, nof_harq_ack_bits=2, nof_enc_harq_ack_bits=48 * 32,
                        nof_csi_part1_bits=7, nof_enc_csi_part1_bits=300 * 32, nof_csi_part2_bits=20,
                        nof_enc_csi_part2_bits=800 * 32)
    z = fec.ulsch_demux_sizes(cfg)
    n, reps = z["nof_codeword_bits"], 46
    rng = np.random.default_rng(3)
    d_cw = torch.from_numpy(rng.integers(-100, 101, (reps, n)).astype(np.int8)).to(dev)
    d_seq = torch.from_numpy(np.packbits(rng.integers(0, 2, n).astype(np.uint8))).to(dev)
    d_sch = torch.zeros((reps, z["nof_sch_bits"] + 64), dtype=torch.int8, device=dev)
    d_uci = [torch.zeros(max(z[k], 1), dtype=torch.int8, device=dev) for k in ("nof_harq_ack_bits", "nof_csi_part1_bits",
                                                                             "nof_csi_part2_bits")]
    torch.cuda.synchronize()

    def step():
        for r in range(reps):
            eng.ulsch_demux_launch_device(cfg, d_cw[r].data_ptr(), d_seq.data_ptr(), d_sch[r].data_ptr(),
                                          *[u.data_ptr() for u in d_uci])

    step()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        e0.record(stream)
        for _ in range(5):
            step()
        e1.record(stream)
    stream.synchronize()
    ms = e0.elapsed_time(e1) / 5
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    moved = reps * (n + sum(z[k] for k in ("nof_sch_bits", "nof_harq_ack_bits", "nof_csi_part1_bits", "nof_csi_part2_bits")))
    return {"workload": "273 PRB x 4 layers 256QAM, 2-bit HARQ-ACK on reserved elements + CSI Part 1 + CSI Part 2, %d soft bits per "
                        "codeword, %d codewords (one launch each) per step" % (n, reps),
            "value": reps * n / (ms * 1e-3) / 1e9, "unit": "G soft bits/s", "ms_per_step": ms, "us_per_codeword": ms * 1e3 / reps,
            "sizes": z,
            "roofline": {"bound": "hbm", "achieved": moved / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": moved / (ms * 1e-3) / 1e9 / peak,
                         "note": "one launch per codeword of 84 CTAs: launch-bound at this size, not bandwidth-bound"}}


def ncu_capture_numbers():
    """DRAM traffic per codeblock and issue statistics of the committed ncu --set full capture of the decode kernel
    (profiles/r02_ncu_packed_full_metrics.csv: raw page of the .ncu-rep, filtered), read at run time so that the line can be
    recomputed from profiles/. Returns {} if the file is not there."""
    path = os.path.join(ROOT, "profiles", "r02_ncu_packed_full_metrics.csv")
    out = {}
    try:
        import csv
        rows = list(csv.reader(open(path)))
        vals = dict(zip(rows[0], rows[2]))
        units = dict(zip(rows[0], rows[1]))

        def num(k):
            v = float(vals[k].replace(",", ""))
            return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}.get(units.get(k, ""), 1)

        # the kernel is persistent (the grid is the resident CTAs, not the codeblocks): the codeblocks of the captured launch
        # are recorded beside the capture
        meta = json.load(open(os.path.join(ROOT, "profiles", "r02_ncu_packed_full_meta.json")))
        out = {"dram_bytes": num("dram__bytes_read.sum") + num("dram__bytes_write.sum"),
               "codeblocks": float(meta["codeblocks_per_launch"]),
               "issue_active_pct": num("smsp__issue_active.avg.pct_of_peak_sustained_active"),
               "warp_instructions": num("smsp__inst_executed.sum"), "source": os.path.relpath(path, ROOT)}
    except Exception:
        pass
    return out


def sass_instructions_per_edge():
    """Instructions per edge and thread (two checks) of the row bodies, from the committed SASS histogram."""
    try:
        for line in open(os.path.join(ROOT, "profiles", "r02_sass_row_hist.txt")):
            if "instructions per edge" in line:
                return float(line.split("->")[1].split()[0])
    except Exception:
        pass
    return None


def bench_config4(fec, eng, torch, dev, stream, args, rank):
    """BASELINE config 4: HARQ retransmissions with soft-buffer combining below the waterfall, 10 iterations at most, early
    stop. n codeblocks (BG1 Z=384, E=11268: rate 0.75 at the first transmission, 64QAM interleaving) go through rv 0, 2, 3, 1;
    every pass rate-dematches and COMBINES into the soft buffers in HBM and decodes all of them again (device-resident, one
    launch per pass). Goodput counts the information bits of the codeblocks whose CRC has passed by the last pass."""
    n, E4, qm, iters = 148 * 3 * 8, 11268, 6, 10
    rng = np.random.default_rng(4 + rank)
    sigma2 = 0.5 * 10.0 ** (-args.harq_esn0 / 10.0)
    llr = np.empty((4, n, E4), np.int8)
    for t in range(4):
        for i in range(0, n, 512):
            y = 1.0 + np.sqrt(sigma2) * rng.standard_normal((min(512, n - i), E4), dtype=np.float32)
            llr[t, i:i + 512] = np.round(np.clip(2.0 * y / sigma2, -20.0, 20.0) * 6.0).astype(np.int8)
    d_llr = [torch.from_numpy(llr[t]).to(dev) for t in range(4)]
    d_descs = []
    for t, rv in enumerate((0, 2, 3, 1)):
        d = fec.make_descs(n)
        d["base_graph"], d["lifting_size"], d["qm"], d["E"], d["rv"] = BG, Z, qm, E4, rv
        d["crc_type"], d["max_iter"], d["mode"] = fec.CRC24B, iters, fec.MODE_CHAIN
        d["flags"] = (fec.FLAG_NEW_DATA if t == 0 else 0) | fec.FLAG_EARLY_STOP
        d["harq_slot"], d["llr_offset"], d["out_offset"] = np.arange(n), np.arange(n, dtype=np.uint64) * E4, np.arange(n) * (K // 8)
        fec.desc_prepare(d)
        d_descs.append(torch.from_numpy(d.view(np.uint8).copy()).to(dev))
    d_out = torch.zeros(n * (K // 8), dtype=torch.uint8, device=dev)
    d_res = [torch.zeros(n * 8, dtype=torch.uint8, device=dev) for _ in range(4)]
    torch.cuda.synchronize()

    def sequence():
        for t in range(4):
            eng.launch_device(d_descs[t].data_ptr(), n, d_llr[t].data_ptr(), d_out.data_ptr(), d_res[t].data_ptr(), Z, True, even_z=True)

    sequence()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    reps = 3
    torch.cuda.synchronize()
    with torch.cuda.stream(stream):
        evs[0].record(stream)
        for _ in range(reps):
            sequence()
        evs[1].record(stream)
    stream.synchronize()
    sec = evs[0].elapsed_time(evs[1]) / 1e3 / reps
    res = [np.frombuffer(r.cpu().numpy().tobytes(), fec.RESULT_DTYPE) for r in d_res]
    ok = [(r["status"] & fec.RES_CRC_PASS) != 0 for r in res]
    passed_by = np.cumsum([o.mean() for o in ok])  # a codeblock the bench keeps decoding may pass more than once: use any()
    any_ok = np.zeros(n, bool)
    frac = []
    for o in ok:
        any_ok |= o
        frac.append(float(any_ok.mean()))
    out = {"workload": "HARQ rv 0,2,3,1 with combining in HBM, BG1 Z=384 E=%d Qm=%d, Es/N0 %.1f dB, max %d iterations, early "
                       "stop, %d codeblocks per pass, device-resident" % (E4, qm, args.harq_esn0, iters, n),
           "ms_per_pass": sec / 4 * 1e3, "codeblock_passes_per_s": 4 * n / sec,
           "goodput_gbps": float(any_ok.sum()) * INFO_BITS / sec / 1e9,
           "crc_pass_fraction_after_pass": frac, "mean_iterations_per_pass": [float(r["nof_ldpc_iterations"].mean()) for r in res]}
    del passed_by
    if not args.no_cpu_baseline and rank == 0:
        try:
            import ctypes as C
            orc = ge.load_oracle()
            ref = orc.Ref()
            impl = orc.Ref.AVX512 if ref.impl_available(orc.Ref.AVX512) else orc.Ref.AVX2
            m, th = 32 * host_threads(), host_threads()
            soft = np.zeros((m, N), np.int8)
            outb = np.zeros((m, K // 8), np.uint8)
            it = np.zeros(m, np.int32)
            ref.lib.ref_batch_chain.restype = C.c_double
            tot, cnt, t0 = 0.0, 0, time.time()
            while time.time() - t0 < 4.0:
                for t, rv in enumerate((0, 2, 3, 1)):
                    x = np.ascontiguousarray(llr[t, :m])
                    tot += ref.lib.ref_batch_chain(impl, th, C.c_uint(m), BG, Z, soft.ctypes.data_as(C.c_void_p),
                                                   x.ctypes.data_as(C.c_void_p), C.c_uint(E4), int(t == 0), C.c_uint(rv),
                                                   C.c_uint(qm), C.c_uint(0), C.c_uint(0), orc.CRC24B, C.c_uint(iters), 1,
                                                   outb.ctypes.data_as(C.c_void_p), it.ctypes.data_as(C.c_void_p))
                cnt += 4 * m
            out["cpu_baseline"] = {"codeblock_passes_per_s": cnt / tot, "cores": th, "kind": "reference",
                                   "sample": "%d codeblock passes of the same HARQ sequence through the reference's AVX "
                                             "dematcher + decoder + CRC on all host threads" % cnt}
        except Exception as ex:
            out["cpu_baseline"] = {"unavailable": str(ex)[:120]}
    return out


def config5_cpu_baseline(fec, d, buf):
    """The same slot on the host cores with the reference's own code (oracle/_ref): every codeblock of every cell through the
    AVX-512 (or AVX2) rate dematcher + LDPC decoder + CRC, all host threads, codeblocks grouped by their parameters. The
    reference's segmentation, TB assembly and TB CRC (microseconds per block) are not in the sample. N = 1 only."""
    orc = ge.load_oracle()
    if not orc.Ref.available():
        return {"unavailable": "oracle/_ref/libref_fec.so was not shipped"}
    ref = orc.Ref()
    impl = orc.Ref.AVX512 if ref.impl_available(orc.Ref.AVX512) else (
        orc.Ref.AVX2 if ref.impl_available(orc.Ref.AVX2) else orc.Ref.GENERIC)
    nthr = host_threads()
    groups = {}
    for i in range(d.size):
        cbs, _ = fec.tb_segment(d[i:i + 1])
        for c in cbs:
            key = (int(c["base_graph"]), int(c["lifting_size"]), int(c["E"]), int(c["qm"]), int(c["nof_filler_bits"]),
                   int(c["crc_type"]))
            off = int(d["llr_offset"][i]) + int(c["llr_offset"])
            groups.setdefault(key, []).append(buf[off:off + int(c["E"])])
    hal_to_orc = {0: orc.CRC16, 1: orc.CRC24B, 2: orc.CRC24A}
    sec, n_cb, n_pass = 0.0, 0, 0
    for rep in range(3):  # first pass warms the decoder objects of every thread
        tot = 0.0
        for (bg, z, e, qm, fill, crc), rows in groups.items():
            s_, iters, _ = ref.batch_chain(impl, nthr, bg, z, np.stack(rows), True, 0, qm, 0, fill, hal_to_orc[crc], MAX_ITER, True)
            tot += s_
            if rep:
                n_cb += len(rows)
                n_pass += int((iters > 0).sum())
        if rep:
            sec += tot
    bits = float(d["tbs"].astype(np.float64).sum())
    return {"kind": "reference", "cores": nthr, "impl": {orc.Ref.GENERIC: "generic", orc.Ref.AVX2: "avx2", orc.Ref.AVX512: "avx512"}[impl],
            "value": bits * 2 / sec / 1e9, "unit": "Gbit/s", "ms_per_slot": sec / 2 * 1e3, "crc_pass_fraction": n_pass / max(n_cb, 1),
            "sample": "2 slots of the same 32 cells (%d codeblocks each, %d parameter groups), parallel decode time of the "
                      "reference's dematcher + decoder + CRC per codeblock" % (n_cb // 2, len(groups))}


def bench_config5(fec, torch, dev, args, rank, world, barrier, max_over_ranks):
    """BASELINE config 5: one uplink slot of 32 cells (100 MHz, 273 PRB, 4 layers; MCS mix of tests/test_gpu_tb.py: 4 cells
    256QAM R948 = 152 codeblocks, 12 cells 64QAM R873, 16 cells QPSK R120 = BG2 with repetition), cells sharded `cell mod N`
    over the ranks, each rank's cells in ONE fec_tb_run_host call per slot from pinned host memory (copies inside the timed
    region). No collective on the data path; value = information bits of all cells / max over ranks of the time."""
    nsym = 42588 * 4
    mcs = [dict(tbs=1277992, bg=1, qm=8, nl=4, G=nsym * 8, snr=8.0), dict(tbs=871384, bg=1, qm=6, nl=4, G=nsym * 6, snr=6.0),
           dict(tbs=39920, bg=2, qm=2, nl=4, G=nsym * 2, snr=-4.0)]
    n_cells = 32
    cell_mcs = [0 if c % 8 == 0 else (1 if c % 2 == 0 else 2) for c in range(n_cells)]
    mine = fec.cells_of_rank(n_cells, world, rank)
    rng = np.random.default_rng(500 + rank)
    d = fec.make_tb_descs(len(mine))
    lo, oo, parts, slot, bits = 0, 0, [], 0, 0
    for i, c in enumerate(mine):
        p = mcs[cell_mcs[c]]
        sigma2 = 0.5 * 10.0 ** (-p["snr"] / 10.0)
        y = 1.0 + np.sqrt(sigma2) * rng.standard_normal(p["G"], dtype=np.float32)
        parts.append(np.round(np.clip(2.0 * y / sigma2, -20.0, 20.0) * 6.0).astype(np.int8))
        d["tbs"][i], d["nof_llrs"][i], d["base_graph"][i], d["qm"][i], d["nof_layers"][i] = p["tbs"], p["G"], p["bg"], p["qm"], p["nl"]
        d["max_iter"][i], d["flags"][i], d["harq_slot"][i] = MAX_ITER, fec.FLAG_NEW_DATA | fec.FLAG_EARLY_STOP, slot
        d["llr_offset"][i], d["out_offset"][i] = lo, oo
        lo += (p["G"] + 63) // 64 * 64
        oo += p["tbs"] // 8
        slot += 162
        bits += p["tbs"]
    buf = np.zeros(lo, np.int8)
    for i, x in enumerate(parts):
        buf[int(d["llr_offset"][i]):int(d["llr_offset"][i]) + x.size] = x
    eng = fec.FecEngine(device=dev.index, harq_slots=slot + 162, max_batch_cbs=max(slot, 512), max_batch_llrs=lo + 4096)
    t_llr = torch.from_numpy(buf).pin_memory()
    t_out = torch.zeros(oo, dtype=torch.uint8).pin_memory()
    ok = 0
    for _ in range(2):
        r = eng.tb_run_host(d, (t_llr.data_ptr(), t_llr.numel()), (t_out.data_ptr(), t_out.numel()))
    barrier()
    reps = 10
    t0 = time.perf_counter()
    for _ in range(reps):
        r = eng.tb_run_host(d, (t_llr.data_ptr(), t_llr.numel()), (t_out.data_ptr(), t_out.numel()))
        ok += int(r["tb_crc_ok"].sum())
    sec = time.perf_counter() - t0
    barrier()
    sec = max_over_ranks(sec)
    total_bits = bits
    if world > 1:
        import torch.distributed as dist
        t = torch.tensor([float(bits), float(ok), float(len(mine) * reps)], dtype=torch.float64, device=dev)
        dist.all_reduce(t)
        total_bits, ok_all, n_all = float(t[0]), float(t[1]), float(t[2])
    else:
        ok_all, n_all = float(ok), float(len(mine) * reps)
    eng.close()
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        cpu = config5_cpu_baseline(fec, d, buf)
    return {"cpu_baseline": cpu,
            "workload": "32 cells x one full-slot transport block (4 x 256QAM R948, 12 x 64QAM R873, 16 x QPSK R120), cell mod N "
                        "sharding, one fec_tb_run_host call per rank and slot from pinned host memory",
            "value": total_bits * reps / sec / 1e9, "unit": "Gbit/s", "ms_per_slot": sec / reps * 1e3, "cells": n_cells,
            "tb_crc_ok_fraction": ok_all / max(n_all, 1.0), "h2d_bytes_per_slot_rank0": int(lo), "d2h_bytes_per_slot_rank0": int(oo),
            "slot_budget_ms": 0.5}


def bench_hal(args):
    """Throughput and LOADED latency through the reference's interfaces: tests/cpp/_build/bench_hal (C++; the reference's
    unmodified pusch_decoder_hw_impl over the CUDA HAL adapter, and pusch_decoder_cuda synchronous / asynchronous), T threads
    decoding config-3 transport blocks back to back. Built by __graft_entry__.build() where the reference tree is present."""
    exe = os.path.join(ROOT, "tests", "cpp", "_build", "bench_hal")
    if not os.path.exists(exe):
        return {"unavailable": "tests/cpp/_build/bench_hal is not built (needs the reference headers at build time)"}
    try:
        p = subprocess.run([exe, str(args.hal_seconds), args.hal_threads, "hal,tb,tb_async", str(args.tb_esn0)],
                           capture_output=True, text=True, timeout=240)
        rows = [json.loads(x) for x in p.stdout.splitlines() if x.startswith("{")]
        best = {}
        for r in rows:
            if r["mode"] not in best or r["gbps"] > best[r["mode"]]["gbps"]:
                best[r["mode"]] = r
        return {"host_threads": host_threads(), "points": rows,
                "best_gbps": {m: {"gbps": r["gbps"], "threads": r["threads"], "lat_us_p99": r["lat_us_p99"]} for m, r in best.items()},
                "note": "latency = on_end_softbits to notification while all threads load the GPU; exit code %d" % p.returncode}
    except Exception as ex:
        return {"unavailable": str(ex)[:160]}


def run_ours(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the FEC engine has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    fec = ge.load_package().fec
    n = args.cbs
    eng = fec.FecEngine(device=local, harq_slots=n, max_batch_cbs=n, nof_queues=2)
    dev = torch.device("cuda", local)

    # ---- synthetic inputs: pinned host copies (e2e) and device-resident copies (value) --------------------------------
    llr_h = make_llrs(n, args.esn0, seed=1234 + rank)
    descs = fec.make_descs(n)
    fill_descs(fec, descs, n, early_stop=False)
    t_llr_pin = torch.from_numpy(llr_h).pin_memory()
    t_out_pin = torch.zeros(n * (K // 8), dtype=torch.uint8).pin_memory()
    d_llr = t_llr_pin.to(dev)
    d_descs = torch.from_numpy(descs.view(np.uint8).copy()).to(dev)
    d_out = torch.zeros(n * (K // 8), dtype=torch.uint8, device=dev)
    d_res = torch.zeros(n * 8, dtype=torch.uint8, device=dev)
    stream = torch.cuda.ExternalStream(eng.stream, device=dev)  # events must sit on the stream the kernels run on
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_device(dd=d_descs):
        eng.launch_device(dd.data_ptr(), n, d_llr.data_ptr(), d_out.data_ptr(), d_res.data_ptr(), Z, True, even_z=True)

    def timed(fn, steps, per_step_events=False):
        """Barrier + sync both sides, CUDA events on the engine stream; returns (seconds, per-step ms list)."""
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
        barrier()
        t0 = time.time()
        with torch.cuda.stream(stream):
            evs[0].record(stream)
            for i in range(steps):
                fn()
                evs[i + 1].record(stream)
        stream.synchronize()
        barrier()
        t1 = time.time()
        per = [evs[i].elapsed_time(evs[i + 1]) for i in range(steps)]
        return evs[0].elapsed_time(evs[-1]) / 1e3, per, t0, t1

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- value: device-resident ------------------------------------------------------------------------------------------
    for _ in range(args.warmup):
        step_device()
    sampler = ClockSampler(local)
    sampler.start()
    time.sleep(0.25)
    l0 = eng.launch_count
    sec, per_step, t0, t1 = timed(step_device, args.steps)
    launches = eng.launch_count - l0
    clocks = sampler.stop(t0, t1)
    sec = max_over_ranks(sec)
    total_cbs = n * args.steps * world
    value = total_cbs * INFO_BITS / sec / 1e9
    kernel_s = float(np.mean(per_step)) / 1e3  # one launch per step: the step time IS the kernel's launch duration
    res = np.frombuffer(d_res.cpu().numpy().tobytes(), fec.RESULT_DTYPE)
    out_h = d_out.cpu().numpy().reshape(n, K // 8)
    passed = (res["status"] & fec.RES_CRC_PASS) != 0
    assert (res["status"] & fec.RES_DONE).all(), "kernel did not complete every codeblock"
    assert (out_h[passed] == 0).all(), "a CRC-passing codeblock does not carry the transmitted (all-zero) message"

    # ---- early-stop variant of the same batch (reported beside the headline, not instead of it) ------------------------
    descs_es = descs.copy()
    descs_es["flags"] = fec.FLAG_NEW_DATA | fec.FLAG_EARLY_STOP
    d_descs_es = torch.from_numpy(descs_es.view(np.uint8).copy()).to(dev)
    step_device(d_descs_es)
    sec_es, _, _, _ = timed(lambda: step_device(d_descs_es), max(2, args.steps // 2))
    sec_es = max_over_ranks(sec_es)
    res_es = np.frombuffer(d_res.cpu().numpy().tobytes(), fec.RESULT_DTYPE)
    early = {"value": n * max(2, args.steps // 2) * world * INFO_BITS / sec_es / 1e9, "unit": "Gbit/s",
             "esn0_db": args.esn0, "mean_iterations": float(res_es["nof_ldpc_iterations"].mean()),
             "crc_pass_fraction": float(((res_es["status"] & fec.RES_CRC_PASS) != 0).mean())}

    # ---- BASELINE config 3: one full-slot transport block (273 PRB, 4 layers, 256QAM R=948/1024: TBS 1 277 992 bits,
    # 152 codeblocks of BG1 Z=384 with 16 fillers, E = 8960 / 8992, Qm = 8), early stop, through fec_run_host from
    # pinned host memory: latency of one TB (call to return) and device-resident throughput of many TBs.
    tb = None
    if not args.no_tb:
        tb = bench_transport_block(fec, eng, torch, dev, stream, args, rank)

    # the side legs (downlink twin, demodulator, HAL bench) run at N = 1 only: in a multi-GPU run the other ranks would sit in
    # a barrier while rank 0 works alone; there the time goes to the multi-cell workload (config 5) instead
    enc = None
    if not args.no_enc and world == 1:
        enc = bench_pdsch_encode(fec, eng, torch, dev, stream, args)
    dmd = None
    if not args.no_demod and world == 1:
        dmd = bench_pusch_demod(fec, eng, torch, dev, stream, args)
    if dmd is not None:
        try:
            dmd["ulsch_demux"] = bench_ulsch_demux(fec, eng, torch, dev, stream)
        except Exception as ex:  # a side leg must not take the line down
            dmd["ulsch_demux"] = {"error": str(ex)}
    cfg4 = None
    if not args.no_config4 and world == 1:
        cfg4 = bench_config4(fec, eng, torch, dev, stream, args, rank)

    # ---- e2e: host buffers through fec_run_host (H2D + kernels + D2H inside the timed region) --------------------------
    res_h = np.zeros(n, fec.RESULT_DTYPE)
    llr_ptr, out_ptr = t_llr_pin.data_ptr(), t_out_pin.data_ptr()

    def step_host():
        eng.lib.fec_run_host(eng.h, descs.ctypes.data, n, llr_ptr, n * E, out_ptr, n * (K // 8), res_h.ctypes.data, None)

    for _ in range(2):
        step_host()
    barrier()
    e2e_steps = max(2, min(args.steps, 10))
    tw0 = time.perf_counter()
    for _ in range(e2e_steps):
        step_host()  # synchronous: returns after the D2H copies have landed
    barrier()
    e2e_sec = max_over_ranks(time.perf_counter() - tw0)
    e2e_val = n * e2e_steps * world * INFO_BITS / e2e_sec / 1e9
    assert (res_h["status"] & fec.RES_DONE).all()

    # ---- what the host can feed this GPU while every rank copies at once: pinned -> device, the ceiling of `e2e` ---------
    # (profiles/ubench/h2d_bw.cu measures the same outside Python; here it is part of the line so that pcie_frac recomputes)
    h2d_peak = None
    try:
        nb = min(t_llr_pin.numel(), 128 << 20)
        src, dst = t_llr_pin.view(-1)[:nb], d_llr.view(-1)[:nb]
        cs = torch.cuda.Stream(device=dev)
        with torch.cuda.stream(cs):
            dst.copy_(src, non_blocking=True)
        cs.synchronize()
        barrier()
        tc0 = time.perf_counter()
        with torch.cuda.stream(cs):
            for _ in range(8):
                dst.copy_(src, non_blocking=True)
        cs.synchronize()
        gbs = 8 * nb / (time.perf_counter() - tc0) / 1e9
        barrier()
        h2d_peak = -max_over_ranks(-gbs)  # the slowest rank's rate
    except Exception:
        pass
    stats = eng.stats()
    cfg5 = None
    if not args.no_config5:
        cfg5 = bench_config5(fec, torch, dev, args, rank, world, barrier, max_over_ranks)
    hal = None
    if not args.no_hal and world == 1 and rank == 0:
        eng.close()  # the C++ bench creates its own engines
        hal = bench_hal(args)
    line = None
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        peak_gbs = float(peaks.get("hbm_gbs", 6650.0))
        achieved = n * ALGO_BYTES_PER_CB / kernel_s / 1e9
        sm_count = torch.cuda.get_device_properties(local).multi_processor_count
        sm_mhz = clocks.get("sm_mhz") or 1965.0
        upd_per_s = n * EDGE_UPDATES_PER_CB / kernel_s
        alu_peak = sm_count * 128 * sm_mhz * 1e6 * 4  # packed int8 lane-ops/s if every INT32 lane did 4 bytes per clock
        # What binds this kernel is instruction ISSUE (integer / half2 work on shared memory, one warp instruction per clock and
        # scheduler at best): the roof is the edge updates per second the SMs could issue if every issue slot carried an
        # instruction of the row bodies as compiled (instructions per edge and thread of two checks: SASS histogram in profiles/).
        ipe = sass_instructions_per_edge() or 34.7
        issue_peak = sm_count * 4 * sm_mhz * 1e6 * 64.0 / ipe  # 32 threads x 2 checks per warp instruction
        cap = ncu_capture_numbers()
        traffic = (cap["dram_bytes"] / cap["codeblocks"] * n) if cap.get("codeblocks") else None
        tb_lat = None
        if tb:
            tb_lat = {"unloaded": dict(tb["tb_api_latency_us"], what="one config-3 transport block per fec_tb_run_host call, idle GPU")}
            if hal and hal.get("points"):
                tb_lat["loaded"] = [{"consumer": r["mode"], "threads": r["threads"], "p50": r["lat_us_p50"], "p99": r["lat_us_p99"],
                                     "gbps": r["gbps"]} for r in hal["points"]]
        line = {"metric": "decoded_info_gbps", "value": value, "unit": "Gbit/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": sec / args.steps * 1e3, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "int8", "data": "synthetic",
                "config": workload_config(n), "per_gpu_gbps": value / world,
                "crc_pass_fraction": float(passed.mean()),
                "e2e": {"value": e2e_val, "unit": "Gbit/s", "h2d_bytes_per_step": n * (E + 48),
                        "d2h_bytes_per_step": n * (K // 8 + 8), "api": "fec_run_host (pinned host buffers)",
                        # host -> device rate of the slowest rank with all ranks copying at once (128 MB pinned copies), the
                        # rate the e2e steps achieved per rank, and their ratio
                        "h2d_peak_gbs_per_gpu": h2d_peak,
                        "h2d_achieved_gbs_per_gpu": n * (E + 48) * e2e_steps / e2e_sec / 1e9,
                        "pcie_frac": (n * (E + 48) * e2e_steps / e2e_sec / 1e9 / h2d_peak) if h2d_peak else None},
                "gpu_launches": int(launches), "clocks": clocks, "early_stop": early,
                "tb_latency_us": tb_lat,
                # the second half of BASELINE's metric as plain numbers: one config-3 transport block per call on an idle GPU
                # (fec_tb_run_host: whole block in, block bytes + CRC out; fec_run_host: its 152 codeblocks in one call)
                "tb_latency_p50_us": tb["tb_api_latency_us"]["p50"] if tb else None,
                "tb_latency_p99_us": tb["tb_api_latency_us"]["p99"] if tb else None,
                "tb_codeblocks_call_latency_p99_us": tb["tb_latency_us"]["p99"] if tb else None,
                "transport_block": tb, "config4_harq": cfg4, "config5_multicell": cfg5, "hal": hal,
                "pdsch_encode": enc, "pusch_demod": dmd, "stats": stats,
                "roofline": {"bound": "issue", "achieved": upd_per_s / 1e9, "peak": issue_peak / 1e9,
                             "unit": "G edge updates/s", "frac": upd_per_s / issue_peak,
                             "peak_definition": "SMs x 4 schedulers x sampled SM clock x 64 checks per warp instruction / "
                                                "%.1f instructions per edge (profiles/r02_sass_row_hist.txt)" % ipe,
                             # dram__bytes_read.sum + dram__bytes_write.sum of one launch of the committed ncu --set full capture,
                             # per codeblock, scaled to this launch
                             "traffic": traffic, "traffic_unit": "bytes/launch", "traffic_source": cap.get("source"),
                             "kernel": "fec_cb_packed_kernel (1 launch/step)", "kernel_ms": kernel_s * 1e3,
                             "ncu_issue_active_pct": cap.get("issue_active_pct"),
                             "alu": {"edge_updates_per_s": upd_per_s, "lane_ops_per_s": upd_per_s * LANE_OPS_PER_UPDATE,
                                     "peak_lane_ops_per_s": alu_peak, "frac": upd_per_s * LANE_OPS_PER_UPDATE / alu_peak,
                                     "sm_count": sm_count,
                                     "note": "SURVEY 8(d) form: canonical 26 int8 lane-ops per edge update vs SMs x 128 "
                                             "lanes x 4 bytes x sampled SM clock"},
                             "hbm": {"achieved": achieved, "peak": peak_gbs, "unit": "GB/s", "frac": achieved / peak_gbs,
                                     "algorithmic_bytes_per_cb": ALGO_BYTES_PER_CB,
                                     "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650 GB/s",
                                     "note": "secondary: the decode needs 9 % of HBM at the 100 Gbit/s target"}}}
        if world == 1 and not args.no_cpu_baseline:
            gbps, desc = reference_cpu(llr_h[:max(32 * host_threads(), 256)], host_threads(), args.cpu_seconds, False)
            if gbps is not None:
                desc["value"], desc["unit"] = gbps, "Gbit/s"
            line["cpu_baseline"] = desc
    if hal is None:
        eng.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(line))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cbs", type=int, default=148 * 48, help="codeblocks per step per GPU")
    ap.add_argument("--esn0", type=float, default=-2.0, help="BPSK Es/N0 in dB of the synthetic channel")
    ap.add_argument("--tb-esn0", type=float, default=8.0, help="Es/N0 (dB) of the config-3 transport-block measurement")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="bounded CPU baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-tb", action="store_true", help="skip the config-3 transport-block latency measurement")
    ap.add_argument("--no-enc", action="store_true", help="skip the PDSCH encoder (downlink twin) measurement")
    ap.add_argument("--no-demod", action="store_true", help="skip the soft demodulator measurement")
    ap.add_argument("--demod-esn0", type=float, default=30.0, help="Es/N0 of the 256QAM symbols of the demodulator bench")
    ap.add_argument("--no-config4", action="store_true", help="skip the HARQ retransmission workload (BASELINE config 4)")
    ap.add_argument("--no-config5", action="store_true", help="skip the 32-cell slot workload (BASELINE config 5)")
    ap.add_argument("--no-hal", action="store_true", help="skip the multi-threaded C++ bench through the reference's interfaces")
    ap.add_argument("--harq-esn0", type=float, default=0.5, help="Es/N0 (dB) of the config-4 HARQ workload (rate 0.75: below the waterfall)")
    ap.add_argument("--hal-seconds", type=float, default=1.5)
    ap.add_argument("--hal-threads", default="1,4,16,32")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    return run_reference(args) if args.impl == "reference" else run_ours(args)


if __name__ == "__main__":
    sys.exit(main())

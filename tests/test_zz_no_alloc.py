"""fec_tb_assemble_device_noalloc (allocation-free transport-block assembly + TB CRC, scratch in a static __device__ array)
against the existing entry point fec_tb_run_host, through the C ABI, on two calls in a row. Bit-exact: integer path.

Named test_zz_* so that pytest collects it last: a failure here must not hide the
other tests. It passes alone in a fresh process and after the other tests (it builds its own engine)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _tb_desc(fec, p, slot, llr_offset, out_offset):
    d = fec.make_tb_descs(1)
    d["tbs"], d["nof_llrs"], d["base_graph"], d["qm"], d["nof_layers"] = p["tbs"], p["G"], p["bg"], p["qm"], p["nl"]
    d["rv"], d["max_iter"], d["harq_slot"] = 0, 6, slot
    d["flags"] = fec.FLAG_NEW_DATA | fec.FLAG_EARLY_STOP
    d["llr_offset"], d["out_offset"] = llr_offset, out_offset
    return d


def _one_call(fec, eng, orc, oracle, params, seed):
    import torch
    rng = np.random.default_rng(seed)
    n = len(params)
    tbs_ = [rng.integers(0, 256, p["tbs"] // 8, dtype=np.uint8) for p in params]
    llrs = [orc.tb_transmit(oracle, orc.tb_encode(oracle, tb, p["bg"]), p["tbs"], p["bg"], 0, p["qm"], 0, p["nl"], p["G"],
                            p["snr"], seed=seed * 10 + i) for i, (tb, p) in enumerate(zip(tbs_, params))]
    lo = np.concatenate([[0], np.cumsum([(x.size + 63) // 64 * 64 for x in llrs])])
    oo = np.concatenate([[0], np.cumsum([p["tbs"] // 8 for p in params])])
    buf = np.zeros(int(lo[-1]), np.int8)
    d = fec.make_tb_descs(n)
    for j, p in enumerate(params):
        buf[lo[j]:lo[j] + llrs[j].size] = llrs[j]
        d[j:j + 1] = _tb_desc(fec, p, 200 * j, int(lo[j]), int(oo[j]))
    # the existing entry point: whole transport blocks from host buffers
    out_ref = np.full(int(oo[-1]), 0xAA, np.uint8)
    res_ref = eng.tb_run_host(d, buf, out_ref)
    # the same codeblocks decoded one by one (other HARQ slots), their messages and results placed in device memory
    cbs = []
    for j in range(n):
        c, _ = fec.tb_segment(d[j:j + 1])
        c["llr_offset"] += np.uint64(lo[j])
        c["harq_slot"] = 1000 + len(cbs) * 32 + np.arange(c.size)  # disjoint slots: at most 11 codeblocks per block here
        cbs.append(c)
    cb = np.concatenate(cbs)
    cb["out_offset"] = np.arange(cb.size) * fec.MAX_CB_OUT_BYTES
    cb_out = np.zeros(cb.size * fec.MAX_CB_OUT_BYTES, np.uint8)
    cb_res = eng.run_host(cb, buf, cb_out)
    d_msgs = torch.from_numpy(cb_out).cuda()
    d_cbres = torch.from_numpy(cb_res.view(np.uint8).copy()).cuda()
    d_out = torch.full((int(oo[-1]),), 0xAA, dtype=torch.uint8, device="cuda")
    d_tbres = torch.zeros(n * fec.TB_RESULT_DTYPE.itemsize, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    eng.tb_assemble_device_noalloc(d, d_msgs.data_ptr(), d_cbres.data_ptr(), d_out.data_ptr(), d_out.numel(),
                                   d_tbres.data_ptr())
    torch.cuda.synchronize()
    res = d_tbres.cpu().numpy().view(fec.TB_RESULT_DTYPE)
    for f in fec.TB_RESULT_DTYPE.names:
        assert (res[f] == res_ref[f]).all(), (f, res[f], res_ref[f])
    assert (d_out.cpu().numpy() == out_ref).all()
    for j, (tb, p) in enumerate(zip(tbs_, params)):
        if res_ref["tb_crc_ok"][j]:
            assert (out_ref[oo[j]:oo[j + 1]] == tb).all()
    return res_ref


def test_noalloc_tb_assembly_matches_tb_run_host_twice(fec, orc, oracle):
    eng = fec.FecEngine(device=0, harq_slots=2048, max_batch_cbs=512)
    try:
        good = dict(tbs=19992, bg=1, qm=2, nl=1, G=40000, snr=4.0)       # 3 codeblocks, decodes (TS 38.214 sizes: tbs + 24 divides by 8 C)
        big = dict(tbs=90000, bg=1, qm=4, nl=2, G=130000, snr=6.0)        # 11 codeblocks
        weak = dict(tbs=29992, bg=1, qm=2, nl=1, G=48000, snr=-6.0)       # fails: not delivered, output untouched
        small = dict(tbs=3000, bg=2, qm=2, nl=1, G=9000, snr=3.0)         # one codeblock, CRC16
        r1 = _one_call(fec, eng, orc, oracle, [good, big, weak], seed=1)
        assert r1["tb_crc_ok"].tolist() == [1, 1, 0] and int(r1["nof_codeblocks"][1]) > 8
        # second call in a row: the static scratch still holds the first call's jobs, map and accumulators
        r2 = _one_call(fec, eng, orc, oracle, [small, weak, good, big], seed=2)
        assert r2["tb_crc_ok"].tolist() == [1, 0, 1, 1]
    finally:
        eng.close()


def test_noalloc_rejects_sizes_beyond_the_static_scratch(fec):
    import torch
    eng = fec.FecEngine(device=0, harq_slots=64, max_batch_cbs=64)
    try:
        buf = torch.zeros(4096, dtype=torch.uint8, device="cuda")
        d = fec.make_tb_descs(fec.NOALLOC_MAX_TBS + 1)
        d["tbs"], d["nof_llrs"], d["base_graph"], d["qm"], d["nof_layers"], d["max_iter"] = 800, 2400, 2, 2, 1, 6
        with pytest.raises(fec.FecError, match="static scratch"):
            eng.tb_assemble_device_noalloc(d, buf.data_ptr(), buf.data_ptr(), buf.data_ptr(), 4096, buf.data_ptr())
        # too many codeblocks: 14 blocks of 152 codeblocks
        d = fec.make_tb_descs(14)
        d["tbs"], d["nof_llrs"], d["base_graph"], d["qm"], d["nof_layers"], d["max_iter"] = 1277992, 1362816, 1, 8, 4, 6
        with pytest.raises(fec.FecError, match="codeblocks in one call"):
            eng.tb_assemble_device_noalloc(d, buf.data_ptr(), buf.data_ptr(), buf.data_ptr(), 1 << 40, buf.data_ptr())
    finally:
        eng.close()

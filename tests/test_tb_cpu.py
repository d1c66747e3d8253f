"""Transport-block layer, CPU side: the segmentation restatement against the reference, the host-only segmentation entry
point of the C ABI against the restatement, and the TB oracle against the reference's pusch_decoder_impl and the fixtures."""
import os

import numpy as np
import pytest

import cases

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "tb_golden.npz")


def _random_tb(rng):
    p = cases.random_tb_params(rng, 200000, 24000)
    # any codeword length the reference accepts, not only sensible rates
    if rng.random() < 0.5:
        p["G"] = int(rng.integers(1, 40000)) * p["nl"] * p["qm"]
    return p


def test_tb_geometry_vs_reference(orc, ref):
    rng = np.random.default_rng(11)
    for _ in range(400):
        p = _random_tb(rng)
        g = orc.tb_geometry(p["tbs"], p["bg"], p["qm"], p["nl"], p["G"])
        r = ref.segment(p["tbs"], p["bg"], 0, p["qm"], 0, p["nl"], p["G"] // p["qm"])
        assert r["n"] == g["C"] and r["Z"] == g["Z"] and r["fillers"] == g["fillers"] and r["N"] == g["N"], p
        assert list(r["rm_len"]) == g["rm"] and list(r["cw_off"]) == [int(x) for x in g["cw_off"]], p


def test_config3_segmentation(orc):
    """BASELINE config 3: 273 PRB x 4 layers 256QAM -> TBS 1 277 992, 152 codeblocks, Z = 384, 16 fillers."""
    g = orc.tb_geometry(1277992, 1, 8, 4, 1362816)
    assert (g["C"], g["Z"], g["fillers"]) == (152, 384, 16)
    assert sorted(set(g["rm"])) == [8960, 8992] and sum(g["rm"]) == 1362816


def test_fec_tb_segment_vs_oracle(fec, orc):
    """The C ABI's host-side segmentation (no GPU involved) against the restatement, field by field."""
    rng = np.random.default_rng(12)
    hal_crc = {orc.CRC24A: fec.CRC24A, orc.CRC24B: fec.CRC24B, orc.CRC16: fec.CRC16}
    for i in range(300):
        p = _random_tb(rng)
        g = orc.tb_geometry(p["tbs"], p["bg"], p["qm"], p["nl"], p["G"])
        K = g["K"]
        if g["N"] * g["C"] < 1 or min(g["rm"]) == 0:
            continue
        d = fec.make_tb_descs(1)
        d["tbs"], d["nof_llrs"], d["base_graph"], d["qm"], d["nof_layers"] = p["tbs"], p["G"], p["bg"], p["qm"], p["nl"]
        d["rv"], d["max_iter"], d["harq_slot"], d["flags"] = i % 4, 6, 100, fec.FLAG_NEW_DATA | fec.FLAG_EARLY_STOP
        cbs, payload = fec.tb_segment(d)
        assert cbs.size == g["C"]
        assert (cbs["E"] == g["rm"]).all() and (cbs["llr_offset"] == g["cw_off"]).all()
        assert (cbs["lifting_size"] == g["Z"]).all() and (cbs["nof_filler_bits"] == g["fillers"]).all()
        assert (cbs["crc_type"] == hal_crc[g["crc"]]).all() and (cbs["harq_slot"] == 100 + np.arange(g["C"])).all()
        assert (cbs["mode"] == fec.MODE_CHAIN).all() and (cbs["rv"] == i % 4).all() and (cbs["qm"] == p["qm"]).all()
        assert list(payload) == g["payload"] and K - g["fillers"] == -(-(p["tbs"] + g["tb_crc"] + (24 * g["C"] if g["C"] > 1 else 0)) // g["C"])


def test_fec_tb_segment_rejects_bad_input(fec):
    d = fec.make_tb_descs(1)
    d["tbs"], d["nof_llrs"], d["base_graph"], d["qm"], d["nof_layers"], d["max_iter"] = 1000, 3004, 1, 2, 4, 6
    with pytest.raises(fec.FecError):
        fec.tb_segment(d)  # 1502 channel symbols are not a multiple of 4 layers (ldpc_segmenter_rx_impl.cpp:49-50)
    d["nof_llrs"] = 3200
    fec.tb_segment(d)
    d["tbs"] = 1001  # transport blocks are whole bytes
    with pytest.raises(fec.FecError):
        fec.tb_segment(d)
    d["tbs"], d["base_graph"] = 1000, 3
    with pytest.raises(fec.FecError):
        fec.tb_segment(d)


def test_tb_oracle_vs_reference(orc, oracle, ref):
    """TbOracle (restatement over the C oracle) against the reference's pusch_decoder_impl + rx buffer pool over HARQ
    sequences: delivered bytes, tb_crc_ok, codeblock CRC flags and soft buffers after every transmission."""
    rt, to = orc.RefTb(ref), orc.TbOracle(oracle)
    rng = np.random.default_rng(5)
    n_fail = n_partial = 0
    for case in range(40):
        p = cases.random_tb_params(rng)
        g = orc.tb_geometry(p["tbs"], p["bg"], p["qm"], p["nl"], p["G"])
        tb = rng.integers(0, 256, p["tbs"] // 8, dtype=np.uint8)
        cbs = orc.tb_encode(oracle, tb, p["bg"])
        bid = (case % 8 + 1) << 8
        for t, rv in enumerate([0, 2, 3, 1]):
            llr = orc.tb_transmit(oracle, cbs, p["tbs"], p["bg"], rv, p["qm"], 0, p["nl"], p["G"], p["snr"] - 1.0, case * 16 + t)
            a = np.full(p["tbs"] // 8, 0xAA, np.uint8)
            b = a.copy()
            ra = rt.decode(bid, p["tbs"], p["bg"], rv, p["qm"], 0, p["nl"], t == 0, 6, True, llr, a)
            rb = to.decode(bid, p["tbs"], p["bg"], rv, p["qm"], 0, p["nl"], t == 0, 6, True, llr, b)
            assert ra["tb_crc_ok"] == rb["tb_crc_ok"] and ra["nof_codeblocks"] == rb["nof_codeblocks"] == g["C"]
            assert (a == b).all()
            if ra["tb_crc_ok"]:
                assert (a == tb).all()
                break
            n_fail += 1
            flags, soft = rt.peek(bid, g["C"], g["N"])
            mine = to.peek(bid)
            assert (flags == mine["crc"]).all() and (soft == mine["soft"]).all()
            n_partial += int(flags.any() and not flags.all())
    assert n_fail >= 10 and n_partial >= 1
    rt.close()


def test_tb_oracle_vs_golden(orc, oracle):
    """Same, against the committed fixture (generated from the reference by tests/golden/make_tb_golden.py)."""
    G = np.load(GOLDEN)
    to = orc.TbOracle(oracle)
    seqs = sorted({k[:3] for k in G.files})
    assert len(seqs) == 10
    for k in seqs:
        tbs, bg, qm, nl, nllr, C_, ntx = [int(x) for x in G[k + "_params"]]
        for t in range(ntx):
            rv, ok = [int(x) for x in G["%s_t%d_rv" % (k, t)]]
            out = np.full(tbs // 8, 0xAA, np.uint8)
            r = to.decode(1, tbs, bg, rv, qm, 0, nl, t == 0, 6, True, G["%s_t%d_llr" % (k, t)], out)
            assert r["tb_crc_ok"] == bool(ok) and (out == G["%s_t%d_out" % (k, t)]).all()
            if not ok:
                mine = to.peek(1)
                assert (mine["crc"] == G["%s_t%d_flags" % (k, t)]).all() and (mine["soft"] == G["%s_t%d_soft" % (k, t)]).all()

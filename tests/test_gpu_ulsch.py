"""UL-SCH demultiplexing on the device (fec_ulsch_demux_*; reference ulsch_demultiplex_impl, TS 38.212 6.2.7) through the C ABI:
the vectors generated from the reference build, random configurations against the C oracle (every stream, bit for bit), and the
device-resident chain demultiplexer -> codeblock decoder on a transport block that carries HARQ-ACK and CSI."""
import os

import numpy as np
import pytest

import ulsch_cases

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ulsch_golden.npz")
NAMES = ("sch", "harq", "csi1", "csi2")


def test_golden_vectors_from_the_reference(fec, orc, engine):
    g = np.load(GOLDEN)
    for i in range(int(g["n"])):
        cfg = {k: int(v) for k, v in zip(orc.ULSCH_CFG_FIELDS, g["c%d_cfg" % i])}
        got = engine.ulsch_demux_run_host(fec.ulsch_cfg(**cfg), g["c%d_cw" % i], g["c%d_seq" % i])
        for name, x in zip(NAMES, got):
            want = g["c%d_%s" % (i, name)]
            assert x.size == want.size and (x == want).all(), (i, name, cfg)


def test_random_configurations_vs_oracle(fec, orc, oracle, engine):
    rng = np.random.default_rng(77)
    seen = np.zeros(4, int)
    for t in range(400):
        cfg = ulsch_cases.random_ulsch_cfg(rng, big=(t % 10 == 0))
        n = orc.ulsch_codeword_len(cfg)
        cw, seq = ulsch_cases.random_codeword(rng, n)
        rc, want, _ = orc.oracle_ulsch_demux(oracle, cfg, cw, seq)
        assert rc == 0
        got = engine.ulsch_demux_run_host(fec.ulsch_cfg(**cfg), cw, seq)
        for name, a, b in zip(NAMES, got, want):
            assert a.size == b.size and (a == b).all(), (t, name, cfg)
        seen += [x.size > 0 for x in want]
    assert (seen > 50).all(), seen


def test_rejects_bad_configurations_and_missing_sequence(fec, engine):
    good = dict(qm=4, nof_layers=2, nof_prb=10, start_symbol=0, nof_symbols=14, dmrs_type=1, dmrs_mask=0b100, nof_cdm_groups_without_data=2)
    assert fec.ulsch_demux_sizes(fec.ulsch_cfg(**good))["nof_sch_bits"] == 13 * 120 * 8
    for bad in (dict(qm=3), dict(nof_layers=5), dict(nof_prb=0), dict(nof_symbols=15), dict(dmrs_mask=0), dict(nof_cdm_groups_without_data=3)):
        with pytest.raises(fec.FecError):
            fec.ulsch_demux_sizes(fec.ulsch_cfg(**{**good, **bad}))
    cfg = fec.ulsch_cfg(**good, nof_harq_ack_bits=1, nof_enc_harq_ack_bits=80, nof_harq_ack_rvd=80)
    n = fec.ulsch_demux_sizes(cfg)["nof_codeword_bits"]
    with pytest.raises(fec.FecError, match="scrambling sequence"):
        engine.ulsch_demux_run_host(cfg, np.zeros(n, np.int8), None)


def _sch_source_positions(orc, oracle, cfg, n):
    """For every soft bit of the SCH stream, the codeword position it comes from (-1: punctured by HARQ-ACK). Found with the
    oracle itself: three passes whose soft bits spell the position in base 120 (the SCH stream is copied unchanged or zeroed)."""
    idx = np.arange(n)
    digits = []
    for k in range(3):
        cw = ((idx // 120 ** k) % 120 + 1).astype(np.int8)
        rc, streams, _ = orc.oracle_ulsch_demux(oracle, cfg, cw, np.zeros(n, np.uint8))
        assert rc == 0
        digits.append(streams[0].astype(np.int64))
    punct = digits[0] == 0
    pos = (digits[0] - 1) + 120 * (digits[1] - 1) + 14400 * (digits[2] - 1)
    pos[punct] = -1
    return pos


def test_device_chain_demultiplex_then_decode(fec, orc, oracle, engine):
    """A 106-PRB, 2-layer 64QAM transport block with a 2-bit HARQ-ACK on reserved elements (punctures the data), CSI Part 1 and
    CSI Part 2: the codeword goes to the device once, the demultiplexer writes the SCH stream in HBM, the codeblock kernel
    decodes from there (fec_launch_device) -- no soft bit returns to the host in between -- and the UCI streams come back."""
    import torch
    cfg = dict(qm=6, nof_layers=2, nof_prb=106, start_symbol=0, nof_symbols=14, nof_harq_ack_rvd=36 * 12, dmrs_type=1,
               dmrs_mask=0b00100000000100, nof_cdm_groups_without_data=2, nof_harq_ack_bits=2, nof_enc_harq_ack_bits=30 * 12,
               nof_csi_part1_bits=7, nof_enc_csi_part1_bits=80 * 12, nof_csi_part2_bits=12, nof_enc_csi_part2_bits=150 * 12,
               csi2_after_csi1=0)
    c = fec.ulsch_cfg(**cfg)
    z = fec.ulsch_demux_sizes(c)
    n, G = z["nof_codeword_bits"], z["nof_sch_bits"]
    assert n == orc.ulsch_codeword_len(cfg) and G % 12 == 0
    # a transport block whose rate-matched length is the SCH stream
    tbs = 53288
    tb = np.random.default_rng(9).integers(0, 256, tbs // 8, dtype=np.uint8)
    llr_sch = orc.tb_transmit(oracle, orc.tb_encode(oracle, tb, 1), tbs, 1, 0, 6, 0, 2, G, 9.0, seed=4)
    assert llr_sch.size == G
    src = _sch_source_positions(orc, oracle, cfg, n)
    rng = np.random.default_rng(10)
    cw = rng.integers(-100, 101, n).astype(np.int8)  # UCI positions: anything
    cw[src[src >= 0]] = llr_sch[src >= 0]
    seq = rng.integers(0, 2, n).astype(np.uint8)
    rc, want, _ = orc.oracle_ulsch_demux(oracle, cfg, cw, seq)
    assert rc == 0 and 0 < (src < 0).sum() <= 30 * 12  # HARQ-ACK elements over data read as zeros in the SCH stream
    # device: codeword and sequence in, four streams in HBM
    d_cw = torch.from_numpy(cw).cuda()
    d_seq = torch.from_numpy(np.packbits(seq)).cuda()
    d_sch = torch.zeros(G + 64, dtype=torch.int8, device="cuda")
    d_uci = [torch.zeros(max(z[k], 1), dtype=torch.int8, device="cuda") for k in ("nof_harq_ack_bits", "nof_csi_part1_bits",
                                                                                 "nof_csi_part2_bits")]
    # codeblock descriptors of the transport block, reading the SCH stream where the demultiplexer writes it
    tbd = fec.make_tb_descs(1)
    tbd["tbs"], tbd["nof_llrs"], tbd["base_graph"], tbd["qm"], tbd["nof_layers"] = tbs, G, 1, 6, 2
    tbd["max_iter"], tbd["flags"], tbd["harq_slot"] = 6, fec.FLAG_NEW_DATA | fec.FLAG_EARLY_STOP, 0
    cbs, payload_bits = fec.tb_segment(tbd)
    Zc = int(cbs["lifting_size"][0])
    K = 22 * Zc
    cbs["out_offset"] = np.arange(cbs.size) * (K // 8)
    cbs["harq_slot"] = 300 + np.arange(cbs.size)
    for s in cbs["harq_slot"]:
        engine.harq_free(int(s))
    engine.harq_flush()
    fec.desc_prepare(cbs)
    d_descs = torch.from_numpy(cbs.view(np.uint8).copy()).cuda()
    d_out = torch.zeros(cbs.size * (K // 8), dtype=torch.uint8, device="cuda")
    d_res = torch.zeros(cbs.size * 8, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    engine.ulsch_demux_launch_device(c, d_cw.data_ptr(), d_seq.data_ptr(), d_sch.data_ptr(), *[u.data_ptr() for u in d_uci])
    engine.launch_device(d_descs.data_ptr(), cbs.size, d_sch.data_ptr(), d_out.data_ptr(), d_res.data_ptr(), Zc, True, even_z=True)
    torch.cuda.synchronize()
    assert (d_sch.cpu().numpy()[:G] == want[0]).all()
    for u, w in zip(d_uci, want[1:]):
        assert (u.cpu().numpy()[:w.size] == w).all()
    res = np.frombuffer(d_res.cpu().numpy().tobytes(), fec.RESULT_DTYPE)
    assert ((res["status"] & fec.RES_CRC_PASS) != 0).all(), res
    # the decoded codeblocks carry the transport block (payload bits of each codeblock, in order)
    out = d_out.cpu().numpy().reshape(cbs.size, K // 8)
    bits = np.concatenate([np.unpackbits(out[i])[:int(payload_bits[i])] for i in range(cbs.size)])
    assert (np.packbits(bits)[:tbs // 8] == tb).all()

"""UL-SCH demultiplexer (TS 38.212 6.2.7; reference ulsch_demultiplex_impl): the C oracle against the reference itself
(oracle/_ref, the unmodified ulsch_demultiplex_impl.cpp) on random configurations, and against the committed vectors that were
generated from that build. The reference's own golden files for this block are not in the checkout."""
import os

import numpy as np
import pytest

import ulsch_cases

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ulsch_golden.npz")


def _same(a, b):
    return all(x.size == y.size and (x == y).all() for x, y in zip(a, b))


def test_oracle_vs_reference_random(orc, oracle, ref):
    rng = np.random.default_rng(2025)
    seen = np.zeros(4, int)
    dynamic = placeholders = 0
    for t in range(1500):
        cfg = ulsch_cases.random_ulsch_cfg(rng, big=(t % 25 == 0))
        n = orc.ulsch_codeword_len(cfg)
        cw, seq = ulsch_cases.random_codeword(rng, n)
        rc, a, ea = orc.oracle_ulsch_demux(oracle, cfg, cw, seq)
        rc2, b, eb = orc.ref_ulsch_demux(ref, cfg, cw, seq, int(rng.choice([64, 777, 4096, 1 << 20])))
        assert rc == 0 and rc2 == 0, (t, cfg)
        assert _same(a, b), (t, cfg, [x.size for x in a], [y.size for y in b])
        assert (ea[1:] == eb[1:]).all(), (t, cfg)
        assert sum(x.size for x in a) >= n  # punctured elements are delivered twice (HARQ-ACK and the stream underneath)
        seen += [x.size > 0 for x in a]
        dynamic += cfg["csi2_after_csi1"] and a[3].size > 0
        placeholders += any(cfg[k] in (1, 2) for k in ("nof_harq_ack_bits", "nof_csi_part1_bits", "nof_csi_part2_bits"))
    assert (seen > 200).all() and dynamic > 50 and placeholders > 300, (seen, dynamic, placeholders)


def test_no_uci_is_a_copy(orc, oracle):
    rng = np.random.default_rng(3)
    cfg = ulsch_cases.random_ulsch_cfg(rng)
    for k in list(cfg):
        if k.startswith("nof_harq") or k.startswith("nof_csi") or k.startswith("nof_enc"):
            cfg[k] = 0
    n = orc.ulsch_codeword_len(cfg)
    cw, seq = ulsch_cases.random_codeword(rng, n)
    rc, a, _ = orc.oracle_ulsch_demux(oracle, cfg, cw, seq)
    assert rc == 0 and (a[0] == cw).all() and all(x.size == 0 for x in a[1:])


def test_oracle_vs_golden(orc, oracle):
    g = np.load(GOLDEN)
    ncase = int(g["n"])
    assert ncase >= 24
    for i in range(ncase):
        cfg = {k: int(v) for k, v in zip(orc.ULSCH_CFG_FIELDS, g["c%d_cfg" % i])}
        rc, a, ended = orc.oracle_ulsch_demux(oracle, cfg, g["c%d_cw" % i], g["c%d_seq" % i])
        assert rc == 0
        for name, x in zip(("sch", "harq", "csi1", "csi2"), a):
            assert (x == g["c%d_%s" % (i, name)]).all(), (i, name)

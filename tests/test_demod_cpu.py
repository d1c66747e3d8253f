"""CPU tests of the soft-demodulation checker (oracle/fec_demod.c): pinned against the reference compiled here for three
instruction sets (oracle/_ref/libref_demod_*.so, live where they exist) and against the committed vectors those libraries
produced (tests/golden/demod_golden.npz). Bit-exact: the LLRs are int8."""
import os

import numpy as np
import pytest

import cases

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "demod_golden.npz")


@pytest.fixture(scope="module")
def dorc(orc, oracle):
    return orc.DemodOracle(oracle)


def by_blocks(fn, blocks, sym, nv):
    out, pos = [], 0
    for b in blocks:
        out.append(fn(sym[pos:pos + int(b)], nv[pos:pos + int(b)]))
        pos += int(b)
    return np.concatenate(out)


def test_oracle_vs_golden(dorc):
    g = np.load(GOLDEN)
    for qm in (1, 2, 4, 6, 8):
        sym, nv, blocks = g[f"q{qm}_sym"], g[f"q{qm}_noise"], g[f"q{qm}_blocks"]
        outs = []
        for f in range(3):
            got = by_blocks(lambda s, n: dorc.demodulate_soft(f, qm, s, n), blocks, sym, nv)
            assert (got == g[f"q{qm}_llr_f{f}"]).all(), (qm, f)
            outs.append(got)
        if qm > 1:
            assert (outs[0] != outs[1]).any()  # the vectors do separate the scalar code from the SIMD code
    assert (dorc.demodulate_soft(0, 1, g["q1_sym"], g["q1_noise"], pi2=True) == g["q1_llr_pi2"]).all()
    for i, (c_init, off, nbits) in enumerate(g["seq_meta"]):
        assert (dorc.gold_sequence(int(c_init), int(off), int(nbits)) == g[f"seq{i}"]).all(), i


@pytest.mark.parametrize("flavour", [0, 1, 2])
def test_oracle_vs_reference_live(orc, dorc, flavour):
    if not orc.RefDemod.available(flavour):
        pytest.skip("oracle/_ref/libref_demod_%s.so not built" % orc.DEMOD_FLAVOURS[flavour])
    ref = orc.RefDemod(flavour)
    rng = np.random.default_rng(77 + flavour)
    for qm in (1, 2, 4, 6, 8):
        for n in (1, 3, 7, 16, 33, 100, 257, 3276, 20000):
            sym, nv = cases.demod_random_symbols(rng, qm, n)
            for pi2 in ((False, True) if qm == 1 else (False,)):
                assert (dorc.demodulate_soft(flavour, qm, sym, nv, pi2) == ref.demodulate_soft(qm, sym, nv, pi2)).all(), (qm, n)
        sym, nv = cases.demod_extreme_symbols(rng, qm)
        assert (dorc.demodulate_soft(flavour, qm, sym, nv) == ref.demodulate_soft(qm, sym, nv)).all(), ("extreme", qm)
        for nvv in (1.0, 0.5, 0.1, 0.3):
            for shift in range(3):  # moves every crafted symbol through the positions of the SIMD block
                sym = np.concatenate([np.full((shift, 2), 0.3, np.float32), cases.demod_crafted_symbols(qm, nvv)])
                nv = np.full(len(sym), nvv, np.float32)
                assert (dorc.demodulate_soft(flavour, qm, sym, nv) == ref.demodulate_soft(qm, sym, nv)).all(), (qm, nvv)
    for c_init, off, nbits in ((0x12345, 0, 1000), (1, 5, 77), (0x7fffffff, 12345, 40000), (0, 3, 100)):
        seq = dorc.gold_sequence(c_init, off, nbits)
        assert (seq == ref.gold_sequence(c_init, off, nbits)).all()
        llr = rng.integers(-120, 121, nbits).astype(np.int8)
        assert (dorc.descramble(seq, llr) == ref.apply_xor_llr(c_init, off, llr)).all()


def test_block_layout(dorc):
    assert dorc.blocks(0, 8) == (0, 0) and dorc.blocks(1, 8) == (0, 4) and dorc.blocks(2, 8) == (8, 4)
    assert dorc.blocks(2, 6) == (32, 16) and dorc.blocks(2, 4) == (0, 8) and dorc.blocks(2, 2) == (0, 16)


def test_desc_prepare_lfsr_states(fec, dorc):
    """fec_demod_desc_prepare (host only): the LFSR states it stores reproduce the next 31 sequence bits at any position."""
    rng = np.random.default_rng(3)
    for _ in range(40):
        d = fec.make_demod_descs(1)
        d["c_init"], d["qm"], d["nof_symbols"] = int(rng.integers(0, 1 << 31)), 2, 10
        d["seq_offset"] = int(rng.choice([0, 1, 1599, 1600, 40000, int(rng.integers(0, 1 << 22))]))
        fec.demod_desc_prepare(d)
        got = ((int(d["x1_state"][0]) ^ int(d["x2_state"][0])) >> np.arange(31)) & 1
        want = np.unpackbits(dorc.gold_sequence(int(d["c_init"][0]), int(d["seq_offset"][0]), 31))[:31]
        assert (got == want).all()
    bad = fec.make_demod_descs(1)
    bad["qm"] = 3
    with pytest.raises(fec.FecError):
        fec.demod_desc_prepare(bad)
    bad["qm"], bad["seq_offset"], bad["nof_symbols"] = 8, 0xffffff00, 1000
    with pytest.raises(fec.FecError):
        fec.demod_desc_prepare(bad)

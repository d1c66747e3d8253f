"""GPU parity tests of the soft demodulator + descrambler, through the C ABI: bit-exact against the
oracle (oracle/fec_demod.c, pinned to the reference build) and against the vectors the reference produced
(tests/golden/demod_golden.npz), for the three reference builds the kernel can stand in for."""
import os

import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "demod_golden.npz")
MODS = {1: "BPSK", 2: "QPSK", 4: "16QAM", 6: "64QAM", 8: "256QAM"}


@pytest.fixture(scope="module")
def dorc(orc, oracle):
    return orc.DemodOracle(oracle)


def plain_descs(fec, blocks, qm, flags=0):
    d = fec.codeword_demod_descs(blocks, qm, 0, flags=flags | fec.DEMOD_FLAG_NO_DESCRAMBLE)
    return d


def test_golden_vectors_all_flavours(fec, engine):
    g = np.load(GOLDEN)
    for qm in (1, 2, 4, 6, 8):
        sym, nv, blocks = g[f"q{qm}_sym"], g[f"q{qm}_noise"], g[f"q{qm}_blocks"]
        for f in range(3):
            llr = np.zeros(len(nv) * qm, np.int8)
            engine.demod_run_host(plain_descs(fec, blocks, qm), sym, nv, llr, flavour=f)
            assert (llr == g[f"q{qm}_llr_f{f}"]).all(), (qm, f)
    llr = np.zeros(len(g["q1_noise"]), np.int8)
    engine.demod_run_host(plain_descs(fec, [len(llr)], 1, fec.DEMOD_FLAG_PI2_BPSK), g["q1_sym"], g["q1_noise"], llr, flavour=0)
    assert (llr == g["q1_llr_pi2"]).all()


def test_tables_derived_from_the_constellation(fec, engine, dorc):
    """The engine derives the piecewise-linear LLR tables from the TS 38.211 constellations; with crafted symbols on every
    interval edge and every rounding tie (cases.demod_crafted_symbols) any table difference shows."""
    for qm in (2, 4, 6, 8):
        for nvv in (1.0, 0.25, 0.3):
            sym = cases.demod_crafted_symbols(qm, nvv)
            nv = np.full(len(sym), nvv, np.float32)
            for f in range(3):
                for shift in range(2):
                    s = np.concatenate([np.full((shift, 2), 0.3, np.float32), sym])
                    n = np.full(len(s), nvv, np.float32)
                    llr = np.zeros(len(s) * qm, np.int8)
                    engine.demod_run_host(plain_descs(fec, [len(s)], qm), s, n, llr, flavour=f)
                    assert (llr == dorc.demodulate_soft(f, qm, s, n)).all(), (qm, nvv, f, shift)


def test_mapper_interface_random_lengths(fec, engine, dorc):
    """demodulation_mapper::demodulate_soft mirror: every modulation, lengths around the SIMD block sizes, special inputs."""
    rng = np.random.default_rng(11)
    for flavour in range(3):
        dm = fec.DemodulationMapper(engine, flavour)
        for qm in (1, 2, 4, 6, 8):
            for n in (1, 3, 7, 8, 15, 16, 31, 32, 33, 255, 256, 257, 3276, 10001):
                sym, nv = cases.demod_random_symbols(rng, qm, n)
                llr = np.zeros(n * qm, np.int8)
                dm.demodulate_soft(llr, sym.view(np.complex64).reshape(-1), nv, MODS[qm])
                assert (llr == dorc.demodulate_soft(flavour, qm, sym, nv)).all(), (flavour, qm, n)
        llr = np.zeros(101, np.int8)
        sym, nv = cases.demod_random_symbols(rng, 1, 101)
        dm.demodulate_soft(llr, sym, nv, "pi/2-BPSK")
        assert (llr == dorc.demodulate_soft(flavour, 1, sym, nv, pi2=True)).all()
    with pytest.raises(fec.FecError):
        fec.DemodulationMapper(engine).demodulate_soft(np.zeros(7, np.int8), np.zeros((2, 2), np.float32), np.ones(2), "QPSK")


def test_codewords_with_descrambling_one_launch(fec, engine, dorc):
    """Several PUSCH codewords in ONE launch: per-OFDM-symbol blocks (DM-RS symbols shorter), own c_init each, unaligned
    LLR offsets, the scrambling sequence continuing across the blocks and across 256-symbol tiles."""
    rng = np.random.default_rng(12)
    descs, syms, nvs, want, pos_s, pos_l = [], [], [], [], 0, 3
    for cw in range(10):
        qm = int(rng.choice([1, 2, 4, 6, 8]))
        nre = int(rng.integers(1, 274)) * 12 * int(rng.integers(1, 3))
        blocks = [nre if rng.random() < 0.8 else nre // 2 for _ in range(int(rng.integers(1, 14)))]
        c_init = int(rng.integers(1, 65536)) * 32768 + int(rng.integers(0, 1024))
        sym, nv = cases.demod_random_symbols(rng, qm, sum(blocks), specials=False)
        d = fec.codeword_demod_descs(blocks, qm, c_init, sym_offset=pos_s, llr_offset=pos_l)
        descs.append(d), syms.append(sym), nvs.append(nv)
        want.append((pos_l, dorc.codeword(2, qm, blocks, sym, nv, c_init)))
        pos_s += sum(blocks)
        pos_l += sum(blocks) * qm
    llr = np.full(pos_l + 5, 99, np.int8)
    alld = fec.make_demod_descs(sum(len(d) for d in descs))  # (np.concatenate would drop the struct padding)
    k = 0
    for d in descs:
        alld[k:k + len(d)] = d
        k += len(d)
    engine.demod_run_host(alld, np.concatenate(syms), np.concatenate(nvs), llr)
    assert (llr[:3] == 99).all() and (llr[pos_l:] == 99).all()  # nothing outside the span is touched
    for off, w in want:
        assert (llr[off:off + w.size] == w).all()


def test_long_sequence_offsets(fec, engine, dorc):
    """Sequence positions far into the Gold sequence (every bit of the jump is exercised)."""
    rng = np.random.default_rng(13)
    for seq_offset in (0, 1, 31, 32, 1599, 1600, 65535, (1 << 20) + 12345, (1 << 27) + 7, (1 << 30) + (1 << 29) + 3, (1 << 31) + 5):
        n, qm = 517, 2
        sym, nv = cases.demod_random_symbols(rng, qm, n, specials=False)
        d = fec.make_demod_descs(1)
        d["nof_symbols"], d["qm"], d["c_init"], d["seq_offset"] = n, qm, 0x5a5a5a5, seq_offset
        llr = np.zeros(n * qm, np.int8)
        engine.demod_run_host(d, sym, nv, llr)
        seq = dorc.gold_sequence(0x5a5a5a5, seq_offset, n * qm)
        assert (llr == dorc.descramble(seq, dorc.demodulate_soft(2, qm, sym, nv))).all(), seq_offset

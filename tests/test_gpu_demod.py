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
            # the fully general instantiation: huge / infinite / NaN symbols, denormal and huge noise variances
            sym, nv = cases.demod_extreme_symbols(rng, qm)
            llr = np.zeros(len(nv) * qm, np.int8)
            dm.demodulate_soft(llr, sym, nv, MODS[qm])
            assert (llr == dorc.demodulate_soft(flavour, qm, sym, nv)).all(), ("extreme", flavour, qm)
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


def test_transport_blocks_from_symbols(fec, engine, oracle, orc, dorc):
    """fec_tb_run_host_symbols: equalised symbols in, transport blocks out, the LLRs only ever on the device. Several
    transport blocks per call (every modulation, 1-4 layers, per-OFDM-symbol demodulation calls), a first transmission
    below the waterfall and a retransmission that combines; against the oracle chain demodulate -> descramble -> decode."""
    rng = np.random.default_rng(21)
    tbo = orc.TbOracle(oracle)
    tbl = []
    for i, (qm, nl, bg, tbs) in enumerate([(8, 4, 1, 60000), (6, 2, 1, 25000), (4, 1, 2, 3000), (2, 1, 2, 640), (1, 1, 2, 200),
                                            (8, 1, 1, 9000)]):
        B = tbs + (16 if tbs <= 3824 else 24)
        nsym = -(-int(B / 0.6) // (qm * nl)) * nl
        tb = rng.integers(0, 256, tbs // 8).astype(np.uint8)
        tbl.append(dict(qm=qm, nl=nl, bg=bg, tbs=tbs, G=nsym * qm, tb=tb, cbs=orc.tb_encode(oracle, tb, bg),
                        c_init=(1000 + i) * 32768 + 17 * i, slot=200 * i))
    # Es/N0 a little under what the rate needs so that most blocks fail the first time and pass after combining
    need = {1: -1.5, 2: 1.0, 4: 6.5, 6: 11.0, 8: 15.5}
    delivered = 0
    for t, rv in enumerate([0, 2]):
        d = fec.make_tb_descs(len(tbl))
        dm, syms, nvs, want_llr, pos_s, pos_l = [], [], [], [], 0, 0
        for i, T in enumerate(tbl):
            sym, nv = orc.tb_transmit_symbols(oracle, dorc, T["cbs"], T["tbs"], T["bg"], rv, T["qm"], 0, T["nl"], T["G"],
                                              need[T["qm"]] - 1.0, T["c_init"], seed=100 * t + i)
            nsym = T["G"] // T["qm"]
            per = -(-nsym // 13)
            blocks = [min(per, nsym - k) for k in range(0, nsym, per)]
            dm.append(fec.codeword_demod_descs(blocks, T["qm"], T["c_init"], sym_offset=pos_s, llr_offset=pos_l))
            syms.append(sym), nvs.append(nv)
            want_llr.append(dorc.codeword(2, T["qm"], blocks, sym, nv, T["c_init"]))
            d["tbs"][i], d["nof_llrs"][i], d["base_graph"][i], d["qm"][i], d["nof_layers"][i] = T["tbs"], T["G"], T["bg"], T["qm"], T["nl"]
            d["rv"][i], d["max_iter"][i], d["harq_slot"][i] = rv, 6, T["slot"]
            d["flags"][i] = (fec.FLAG_NEW_DATA if t == 0 else 0) | fec.FLAG_EARLY_STOP
            d["llr_offset"][i], d["out_offset"][i] = pos_l, sum(x["tbs"] // 8 for x in tbl[:i])
            pos_s += nsym
            pos_l += T["G"]
        alld = fec.make_demod_descs(sum(len(x) for x in dm))
        k = 0
        for x in dm:
            alld[k:k + len(x)] = x
            k += len(x)
        out = np.zeros(sum(x["tbs"] // 8 for x in tbl), np.uint8)
        res = engine.tb_run_host_symbols(d, alld, np.concatenate(syms), np.concatenate(nvs), out)
        for i, T in enumerate(tbl):
            if T.get("done"):
                continue
            want = np.zeros(T["tbs"] // 8, np.uint8)
            w = tbo.decode(i, T["tbs"], T["bg"], rv, T["qm"], 0, T["nl"], t == 0, 6, True, want_llr[i], want)
            assert bool(res["tb_crc_ok"][i]) == w["tb_crc_ok"], (t, i)
            assert int(res["nof_cb_decoded"][i]) == len(w["iters"]) and int(res["nof_codeblocks"][i]) == w["nof_codeblocks"]
            o0 = int(d["out_offset"][i])
            if w["tb_crc_ok"]:
                assert (out[o0:o0 + T["tbs"] // 8] == want).all() and (want == T["tb"]).all(), (t, i)
                T["done"] = True
                delivered += 1
    assert delivered >= 4
    # a gap in the demodulation calls is refused (the transport blocks would read stale LLRs)
    with pytest.raises(fec.FecError):
        engine.tb_run_host_symbols(d, alld[:-1], np.concatenate(syms), np.concatenate(nvs), out)


def test_argument_checks(fec, engine):
    """Invalid calls are refused with an error, never executed: unknown flavour, unsupported modulation order, descriptors
    that index outside the symbol or LLR buffers, calls that do not tile one contiguous LLR span."""
    sym, nv = np.zeros((64, 2), np.float32), np.ones(64, np.float32)
    llr = np.zeros(64 * 2, np.int8)
    d = fec.codeword_demod_descs([64], 2, 1)
    with pytest.raises(fec.FecError):
        engine.demod_run_host(d, sym, nv, llr, flavour=3)
    bad = d.copy()
    bad["qm"] = 3
    with pytest.raises(fec.FecError):
        engine.demod_run_host(bad, sym, nv, llr)
    bad = d.copy()
    bad["nof_symbols"] = 65  # one symbol more than the buffers hold
    with pytest.raises(fec.FecError):
        engine.demod_run_host(bad, sym, nv, np.zeros(130, np.int8))
    bad = d.copy()
    bad["llr_offset"] = 8  # writes past the end of the LLR buffer
    with pytest.raises(fec.FecError):
        engine.demod_run_host(bad, sym, nv, llr)
    two = fec.make_demod_descs(2)
    two[0:1], two[1:2] = fec.codeword_demod_descs([16], 2, 1), fec.codeword_demod_descs([16], 2, 1, sym_offset=16, llr_offset=40)
    with pytest.raises(fec.FecError):  # a gap between the two calls
        engine.demod_run_host(two, sym, nv, llr)
    engine.demod_run_host(d, sym, nv, llr)  # and the engine is still usable afterwards
    assert (llr == 0).all()  # zero symbols give zero soft bits

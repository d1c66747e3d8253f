"""Parity in depth on the shapes the bench is quoted on, through the C ABI, against the C oracle (bit-exact):
  * 1024 seeded AWGN codeblocks of the headline shape (BG1, Z = 384, E = N = 25344) across the waterfall
    (Es/N0 -4.5 .. +1.5 dB: failures, late and early convergence), max_iter 6 and 10, early stop on and off;
  * 512 codeblocks of BASELINE config 3 (BG1, Z = 384, 16 filler bits, E = 8960 / 8992, 256QAM: rate 0.94, the last
    systematic columns are punctured by the rate matcher and the row trimming of the decoder is data-dependent);
  * normalisation factors other than the reference's default 0.8 (ldpc_decoder_generic/avx2/... take it as a constructor
    argument: ldpc_decoder_impl.h:50-58): 0.5 and 0.75 run on the packed kernel, 0.9 has a rounding table the packed
    kernel's 8-bit multiply-add cannot reproduce and must fall back to the generic kernel.
Decoded bytes, iteration counts, CRC flags and the HARQ soft buffer are compared for every codeblock."""
import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu


def _run(fec, eng, items, align=16):
    """One fec_run_host call over `items` (dict(p, llr, slot)); new data everywhere. Returns (res, out bytes list). The soft
    bits of each codeblock start at a multiple of `align` bytes (1: back to back, any alignment)."""
    n = len(items)
    d = fec.make_descs(n)
    offs = np.cumsum([0] + [((it["llr"].size + align - 1) // align) * align for it in items])
    llr = np.zeros(int(offs[-1]) + 16, np.int8)
    nbytes = [(cases.dims(it["p"]["bg"], it["p"]["Z"])[0] + 7) // 8 for it in items]
    out_off = np.cumsum([0] + nbytes)
    for i, it in enumerate(items):
        eng.harq_free(it["slot"])
        cases.fill_chain_desc(fec, d, i, it["p"], True, it["slot"], int(offs[i]), int(out_off[i]))
        llr[offs[i]:offs[i] + it["llr"].size] = it["llr"]
    out = np.full(int(out_off[-1]), 0x5A, np.uint8)
    res = eng.run_host(d, llr, out)
    assert (res["status"] & fec.RES_DONE).all()
    return res, [out[out_off[i]:out_off[i + 1]] for i in range(n)]


def _check(fec, eng, oracle, items, tag, align=16):
    res, outs = _run(fec, eng, items, align)
    n_pass = n_fail = 0
    its = []
    for i, it in enumerate(items):
        p = it["p"]
        K, N = cases.dims(p["bg"], p["Z"])
        soft = np.zeros(N, np.int8)
        want_it, want_out = oracle.cb_chain(p["bg"], p["Z"], soft, it["llr"], True, p["rv"], p["qm"], p["Nref"], p["fill"],
                                            p["poly"], p["max_iter"], p["early"])
        got_it = int(res["nof_ldpc_iterations"][i]) if res["status"][i] & fec.RES_CRC_PASS else -1
        assert got_it == want_it, (tag, i, got_it, want_it)
        assert (outs[i] == want_out).all(), (tag, i)
        assert (eng.harq_read(it["slot"], N) == soft).all(), (tag, i)
        n_pass += want_it >= 0
        n_fail += want_it < 0
        its.append(want_it)
    return n_pass, n_fail, its


@pytest.mark.parametrize("max_iter,early", [(6, False), (6, True), (10, False), (10, True)])
def test_headline_shape_awgn_256_codeblocks(fec, engine, oracle, orc, max_iter, early):
    """256 codeblocks per parameter set, 1024 in all. Es/N0 is drawn over the waterfall of the rate-1/3 mother code so that
    the set holds failed blocks, blocks that converge in the last iterations and blocks that stop early."""
    bg, Z = 1, 384
    K, N = cases.dims(bg, Z)
    rng = np.random.default_rng(100 * max_iter + early)
    items = []
    for i in range(256):
        p = dict(bg=bg, Z=Z, E=N, rv=0, qm=int(rng.choice([2, 4, 6, 8])), Nref=0, fill=0, poly=orc.CRC24B, max_iter=max_iter,
                 early=early)
        p["E"] = N // p["qm"] * p["qm"]
        snr = float(rng.uniform(-4.5, 1.5))
        llr, _ = oracle.make_case(bg, Z, p["E"], 0, p["qm"], 0, 0, orc.CRC24B, snr, seed=70000 + 1000 * max_iter + 300 * early + i)
        items.append(dict(p=p, llr=llr, slot=i))
    n_pass, n_fail, its = _check(fec, engine, oracle, items, (max_iter, early))
    assert n_pass >= 40 and n_fail >= 10, (n_pass, n_fail)
    if early:
        assert len({x for x in its if x >= 0}) >= 3, sorted(set(its))  # several distinct stopping iterations


def test_config3_shape_512_codeblocks(fec, engine, oracle, orc):
    """BASELINE config 3: 100 MHz / 4 layers / 256QAM / rate 0.93: K' = 8432 of K = 8448 (16 fillers), E = 8960 or 8992
    per codeblock (the segmenter gives the last codeblocks one more symbol per layer), 10 iterations, early stop."""
    bg, Z, fill = 1, 384, 16
    rng = np.random.default_rng(33)
    items = []
    for i in range(512):
        E = 8960 if i % 3 else 8992
        p = dict(bg=bg, Z=Z, E=E, rv=0, qm=8, Nref=0, fill=fill, poly=orc.CRC24B, max_iter=int(rng.choice([6, 10])),
                 early=bool(i % 4))
        rate = (8448 - fill) / E
        snr = 10 * np.log10(2 ** (2 * rate) - 1) - 0.5 + float(rng.uniform(-1.5, 4.0))
        llr, _ = oracle.make_case(bg, Z, E, 0, 8, 0, fill, orc.CRC24B, snr, seed=90000 + i)
        items.append(dict(p=p, llr=llr, slot=i))
    n_pass, n_fail, _ = _check(fec, engine, oracle, items, "config3")
    assert n_pass >= 100 and n_fail >= 30, (n_pass, n_fail)


def test_config5_shapes(fec, engine, oracle, orc):
    """The other two codeblock shapes of BASELINE config 5: 64QAM R873 (45 fillers: not a multiple of four, so the run behind
    the filler gap is read with unaligned word loads) and QPSK R120 on BG2 with repetition (E > N: the transmission wraps
    around the circular buffer and the passes are summed in order)."""
    rng = np.random.default_rng(55)
    items = []
    for i in range(192):
        E = 9816 if i % 2 else 9840
        p = dict(bg=1, Z=384, E=E, rv=0, qm=6, Nref=0, fill=45, poly=orc.CRC24B, max_iter=6, early=bool(i % 3))
        snr = 10 * np.log10(2 ** (2 * (8448 - 45) / E) - 1) - 0.5 + float(rng.uniform(-1.5, 3.0))
        llr, _ = oracle.make_case(1, 384, E, 0, 6, 0, 45, orc.CRC24B, snr, seed=120000 + i)
        items.append(dict(p=p, llr=llr, slot=i))
    for i in range(48):
        E = 30976 if i % 2 else 30968
        p = dict(bg=2, Z=384, E=E, rv=0, qm=2, Nref=0, fill=184, poly=orc.CRC24B, max_iter=6, early=bool(i % 3))
        llr, _ = oracle.make_case(2, 384, E, 0, 2, 0, 184, orc.CRC24B, float(rng.uniform(-8.5, -4.0)), seed=130000 + i)
        items.append(dict(p=p, llr=llr, slot=200 + i))
    n_pass, n_fail, _ = _check(fec, engine, oracle, items, "config5")
    assert n_pass >= 60 and n_fail >= 10, (n_pass, n_fail)


@pytest.mark.parametrize("sf,packed", [(0.5, True), (0.75, True), (0.9, False)])
def test_scaling_factors_other_than_default(fec, oracle, orc, sf, packed):
    """The decoder alone (LdpcDecoder mirror, MODE_DECODE) with another normalisation factor against the oracle's decoder
    with the same factor. `generic_launches` of the engine statistics tells which kernel ran."""
    eng = fec.FecEngine(device=0, harq_slots=64, max_batch_cbs=64, scaling_factor=sf)
    try:
        dec = fec.LdpcDecoder(eng)
        rng = np.random.default_rng(int(sf * 1000))
        n_pass = n_fail = n_odd = 0
        for trial in range(60):
            bg = int(rng.choice([1, 2]))
            Z = int(rng.choice([4, 6, 9, 16, 28, 52, 104, 224, 384, 384]))
            n_odd += 2 * (Z % 2)  # odd lifting sizes always take the generic kernel (launched beside the packed one)
            K, N = cases.dims(bg, Z)
            n_in = int(rng.integers(K + 4 * Z, N + 1))
            snr = 10 * np.log10(2 ** (2 * K / n_in) - 1) - 0.5 + float(rng.uniform(-2.0, 3.0))
            llr, _ = oracle.make_case(bg, Z, N, 0, 1, 0, 0, orc.CRC24B, snr, seed=int(sf * 1e4) + trial)
            llr = llr[:n_in].copy()
            for crc, ocrc in ((None, -1), (fec.CRC24B, orc.CRC24B)):
                mi = int(rng.choice([2, 6, 10]))
                it, ref_out = oracle.decode(bg, Z, llr, 0, ocrc, mi, scaling=sf)
                out = np.zeros((K + 7) // 8, np.uint8)
                got = dec.decode(out, llr, crc, fec.CodeblockMetadata(base_graph=bg, lifting_size=Z), max_iterations=mi)
                assert (got if got is not None else -1) == it, (sf, trial, bg, Z, n_in)
                assert (out == ref_out).all(), (sf, trial, bg, Z, n_in)
                if crc is not None:
                    n_pass += it >= 0
                    n_fail += it < 0
        assert n_pass >= 10 and n_fail >= 3, (n_pass, n_fail)
        st = eng.stats()
        assert st["cb_launches"] == 120
        assert st["generic_launches"] == (n_odd if packed else 120), st
    finally:
        eng.close()


def test_scaling_factor_changes_the_result(fec, oracle, orc):
    """Guard against a test that would pass with the factor ignored: at 0.5 and 0.8 the oracle itself must disagree on
    at least one block of this set (iterations or bytes)."""
    differ = 0
    for trial in range(30):
        bg, Z = 1, 52
        K, N = cases.dims(bg, Z)
        llr, _ = oracle.make_case(bg, Z, N, 0, 1, 0, 0, orc.CRC24B, -3.2 + 0.05 * trial, seed=4000 + trial)
        a = oracle.decode(bg, Z, llr, 0, orc.CRC24B, 10, scaling=0.5)
        b = oracle.decode(bg, Z, llr, 0, orc.CRC24B, 10, scaling=0.8)
        differ += (a[0] != b[0]) or not (a[1] == b[1]).all()
    assert differ > 0


def test_dematcher_paths_first_transmissions(fec, engine, oracle, orc):
    """The round-2 dematcher paths, 640 random first transmissions from the start of the buffer (rv 0) at the lifting sizes that
    take them (Z a multiple of 8): the direct path with and without fillers that are a multiple of four, E below / at / above
    the systematic part, limited buffers (Nref), BPSK (no transposed staging), and repetition (E > Ncb - fillers: the
    word-wise wrap path), inputs aligned to 16 bytes and packed back to back (any alignment), in four launches and against the
    oracle: bytes, iterations, CRC flags, soft buffers."""
    rng = np.random.default_rng(808)
    zs = [8, 16, 24, 32, 40, 48, 56, 64, 80, 96, 128, 192, 256, 384]
    items = []
    for i in range(640):
        bg = int(rng.integers(1, 3))
        Z = int(zs[i % len(zs)]) if i % 5 else 384
        K, N = cases.dims(bg, Z)
        Kb = 22 if bg == 1 else 10
        poly = int(rng.choice([0, 1, 2]))
        max_fill = min((Kb - 2) * Z - 1, 2 * Z, K - cases.ORC_CRC_LEN[poly] - 1)
        fill = int(rng.integers(0, max_fill + 1)) if i % 3 else 0
        qm = int(rng.choice([1, 2, 4, 6, 8]))
        Nref = 0 if rng.random() < 0.6 else int(rng.integers(K + 2 * Z + 1, N + 10))
        Ncb = min(Nref, N) if Nref else N
        kind = i % 4
        if kind == 0:    # high rate: ends inside or just behind the systematic part
            E = int(rng.integers(K - fill - 3 * Z, K - fill + 4 * Z))
        elif kind == 1:  # ordinary
            E = int(rng.integers(K - fill + 2 * Z, Ncb - fill))
        elif kind == 2:  # around the end of the buffer
            E = int(rng.integers(Ncb - fill - 40, Ncb - fill + 40))
        else:            # repetition: one to three passes
            E = int(rng.integers(Ncb - fill + 1, 3 * (Ncb - fill)))
        E = max(qm, E // qm * qm)
        if E > 50000:
            E = 50000 // qm * qm
        p = dict(bg=bg, Z=Z, E=E, rv=0, qm=qm, Nref=Nref, fill=fill, poly=poly, max_iter=int(rng.choice([2, 6])),
                 early=bool(i % 2))
        rate = (K - fill) / E
        snr = 10 * np.log10(max(2 ** (2 * min(rate, 1.2)) - 1, 1e-3)) + float(rng.uniform(0.0, 7.0))
        llr, _ = oracle.make_case(bg, Z, E, 0, qm, Nref, fill, poly, snr, seed=700000 + i)
        items.append(dict(p=p, llr=llr, slot=i))
    n_pass = n_fail = 0
    for c0 in range(0, len(items), 160):
        a, b, _ = _check(fec, engine, oracle, items[c0:c0 + 160], ("dematch", c0), align=16 if c0 < 320 else 1)
        n_pass += a
        n_fail += b
    assert n_pass >= 200 and n_fail >= 60, (n_pass, n_fail)


def test_pinned_buffers_are_read_in_place_and_give_the_same_answer(fec, oracle, orc):
    """fec_run_host / fec_tb_run_host on pinned host buffers (small call, alone on the engine: the kernel reads the soft bits
    and writes the bytes in the caller's memory, no copies) against the same calls on pageable numpy buffers (copied): same
    results, same bytes (blocks that never pass included), same soft buffers."""
    import ctypes as C
    import torch
    eng = fec.FecEngine(device=0, harq_slots=512, max_batch_cbs=512)
    try:
        rng = np.random.default_rng(4242)
        items = []
        for i in range(96):
            bg = int(rng.integers(1, 3))
            Z = int(rng.choice([32, 96, 208, 384, 384]))
            K, N = cases.dims(bg, Z)
            qm = int(rng.choice([2, 4, 6, 8]))
            E = int(rng.integers(K, N)) // qm * qm
            p = dict(bg=bg, Z=Z, E=E, rv=0, qm=qm, Nref=0, fill=0, poly=orc.CRC24B, max_iter=6, early=True)
            snr = 10 * np.log10(2 ** (2 * K / E) - 1) + float(rng.uniform(-3.0, 3.0))
            llr, _ = oracle.make_case(bg, Z, E, 0, qm, 0, 0, orc.CRC24B, snr, seed=99000 + i)
            items.append(dict(p=p, llr=llr))
        n = len(items)
        d = fec.make_descs(n)
        offs = np.cumsum([0] + [it["llr"].size for it in items])  # back to back: any alignment
        nbytes = [(cases.dims(it["p"]["bg"], it["p"]["Z"])[0] + 7) // 8 for it in items]
        out_off = np.cumsum([0] + nbytes)
        llr = np.zeros(int(offs[-1]) + 16, np.int8)
        for i, it in enumerate(items):
            cases.fill_chain_desc(fec, d, i, it["p"], True, i, int(offs[i]), int(out_off[i]))
            llr[offs[i]:offs[i] + it["llr"].size] = it["llr"]
        # pageable buffers: the copy path
        for i in range(n):
            eng.harq_free(i)
        out_a = np.full(int(out_off[-1]), 0x5A, np.uint8)
        res_a = eng.run_host(d.copy(), llr, out_a)
        soft_a = [eng.harq_read(i, cases.dims(items[i]["p"]["bg"], items[i]["p"]["Z"])[1]).copy() for i in range(n)]
        # pinned buffers: read and written in place
        for i in range(n):
            eng.harq_free(i)
        t_llr = torch.from_numpy(llr.copy()).pin_memory()
        t_out = torch.full((int(out_off[-1]),), 0x5A, dtype=torch.uint8).pin_memory()
        res_b = np.zeros(n, fec.RESULT_DTYPE)
        db = d.copy()
        rc = eng.lib.fec_run_host(eng.h, db.ctypes.data, n, C.c_void_p(t_llr.data_ptr()), t_llr.numel(),
                                  C.c_void_p(t_out.data_ptr()), t_out.numel(), res_b.ctypes.data, None)
        assert rc == 0
        assert (res_a["status"] == res_b["status"]).all() and (res_a["nof_ldpc_iterations"] == res_b["nof_ldpc_iterations"]).all()
        assert (t_out.numpy() == out_a).all()
        for i in range(n):
            N = cases.dims(items[i]["p"]["bg"], items[i]["p"]["Z"])[1]
            assert (eng.harq_read(i, N) == soft_a[i]).all(), i
        ok = (res_a["status"] & fec.RES_CRC_PASS) != 0
        assert 10 <= ok.sum() <= n - 5, ok.sum()
    finally:
        eng.close()

"""GPU parity tests proper: the CUDA path, called through the C ABI (include/fec_b200.h), against the oracle on the same
seeded inputs and against the committed golden fixtures (generated from the reference build). Bit-exact: decoded bytes,
iteration counts, CRC pass/fail and soft-buffer contents."""
import os

import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "fec_golden.npz")


def run_chain_batch(fec, eng, items, mode=None):
    """items: list of dict(p=params, llr=int8[E], new_data=bool, slot=int). One launch through fec_run_host.
    Returns (results, out[n, 1056], soft[n, 25344])."""
    n = len(items)
    d = fec.make_descs(n)
    offs = np.cumsum([0] + [((it["llr"].size + 15) // 16) * 16 for it in items])
    llr = np.zeros(int(offs[-1]) + 16, np.int8)
    for i, it in enumerate(items):
        if it["new_data"]:
            eng.harq_free(it["slot"])  # the reference starts a new HARQ process from a zeroed buffer (trap T8)
        cases.fill_chain_desc(fec, d, i, it["p"], it["new_data"], it["slot"], int(offs[i]), i * fec.MAX_CB_OUT_BYTES,
                              mode=mode)
        llr[offs[i]:offs[i] + it["llr"].size] = it["llr"]
    out = np.zeros((n, fec.MAX_CB_OUT_BYTES), np.uint8)
    soft = np.zeros((n, fec.HARQ_SLOT_BYTES), np.int8)
    res = eng.run_host(d, llr, out, soft_out=soft)
    assert (res["status"] & fec.RES_DONE).all()
    return res, out, soft


def got_iterations(fec, res, i):
    return int(res["nof_ldpc_iterations"][i]) if res["status"][i] & fec.RES_CRC_PASS else -1


@pytest.fixture(params=["auto", "generic"])
def eng(request, engine, engine_generic):
    return engine if request.param == "auto" else engine_generic


def test_golden_harq_sequences(fec, eng):
    """Golden fixtures: every transmission t of every case, run as one batch per t so HARQ state carries in HBM."""
    g = np.load(GOLDEN)
    meta = g["meta"]
    for t in range(int(meta[:, 1].max()) + 1):
        rows = [r for r in meta if r[1] == t]
        items = []
        for r in rows:
            ci, _, bg, Z, E, rv, qm, Nref, fill, poly, max_iter, early, _ = (int(x) for x in r)
            p = dict(bg=bg, Z=Z, E=E, rv=rv, qm=qm, Nref=Nref, fill=fill, poly=poly, max_iter=max_iter, early=bool(early))
            items.append(dict(p=p, llr=g[f"c{ci}_t{t}_llr"], new_data=(t == 0), slot=ci))
        res, out, soft = run_chain_batch(fec, eng, items)
        for i, r in enumerate(rows):
            ci, bg, Z, it_ref = int(r[0]), int(r[2]), int(r[3]), int(r[-1])
            K, N = cases.dims(bg, Z)
            assert got_iterations(fec, res, i) == it_ref, (ci, t)
            assert (out[i, :(K + 7) // 8] == g[f"c{ci}_t{t}_out"]).all(), (ci, t)
            assert (soft[i, :N] == g[f"c{ci}_t{t}_soft"]).all(), (ci, t)


@pytest.mark.parametrize("bg", [1, 2])
def test_sweep_all_lifting_sizes_vs_oracle(fec, eng, oracle, bg):
    """BASELINE config 2: every Z, random rate / Qm / rv / LBRM / fillers / CRC / iterations, mixed in ONE launch."""
    rng = np.random.default_rng(1000 + bg)
    items, want = [], []
    for Z in cases.ALL_Z:
        for rep in range(4 if Z <= 104 else 2):
            p = cases.random_chain_params(rng, bg=bg, Z=Z)
            K, N = cases.dims(bg, Z)
            p["E"] = max(p["qm"], min(p["E"], 2 * N + 64) // p["qm"] * p["qm"])
            llr, _ = oracle.make_case(bg, Z, p["E"], p["rv"], p["qm"], p["Nref"], p["fill"], p["poly"], p["snr"],
                                      seed=Z * 10 + rep)
            soft = np.zeros(N, np.int8)
            it, ref_out = oracle.cb_chain(bg, Z, soft, llr, True, p["rv"], p["qm"], p["Nref"], p["fill"], p["poly"],
                                          p["max_iter"], p["early"])
            items.append(dict(p=p, llr=llr, new_data=True, slot=len(items)))
            want.append((it, ref_out, soft))
    res, out, soft = run_chain_batch(fec, eng, items)
    n_pass = 0
    for i, (it, ref_out, ref_soft) in enumerate(want):
        p = items[i]["p"]
        K, N = cases.dims(bg, p["Z"])
        assert got_iterations(fec, res, i) == it, p
        assert (out[i, :(K + 7) // 8] == ref_out).all(), p
        assert (soft[i, :N] == ref_soft).all(), p
        n_pass += it > 0
    assert 0 < n_pass < len(want)  # the sweep must exercise both converging and non-converging blocks


def test_harq_retransmissions_low_snr(fec, eng, oracle, orc):
    """BASELINE config 4: rv sequence {0,2,3,1}, new_data only first, 10 iterations with early stop, below waterfall."""
    rng = np.random.default_rng(4)
    shapes = [(1, 384, 8960, 8), (1, 384, 25344, 2), (2, 208, 5000, 4), (1, 96, 3000, 6), (2, 384, 30968 // 2 * 2, 2)]
    cbs = []
    for si, (bg, Z, E, qm) in enumerate(shapes):
        for rep in range(3):
            K, N = cases.dims(bg, Z)
            E = E // qm * qm
            fill = [0, 16, 40][rep]
            msg = oracle.make_message(bg, Z, fill, orc.CRC24B, seed=si * 10 + rep)
            cbs.append(dict(bg=bg, Z=Z, E=E, qm=qm, fill=fill, cb=oracle.encode(bg, Z, msg), soft=np.zeros(N, np.int8),
                            snr=10 * np.log10(max(2 ** (2 * (K - fill) / E) - 1, 1e-3)) - 4.5 + rep))
    passed_any = False
    for t, rv in enumerate([0, 2, 3, 1]):
        items, want = [], []
        for i, c in enumerate(cbs):
            tx = oracle.rate_match(c["bg"], c["Z"], c["cb"], c["E"], rv, c["qm"], 0, c["fill"])
            llr = oracle.awgn_llr(tx, c["snr"], seed=i * 4 + t)
            p = dict(bg=c["bg"], Z=c["Z"], E=c["E"], rv=rv, qm=c["qm"], Nref=0, fill=c["fill"], poly=orc.CRC24B,
                     max_iter=10, early=True)
            it, ref_out = oracle.cb_chain(c["bg"], c["Z"], c["soft"], llr, t == 0, rv, c["qm"], 0, c["fill"], orc.CRC24B,
                                          10, True)
            items.append(dict(p=p, llr=llr, new_data=(t == 0), slot=100 + i))
            want.append((it, ref_out))
        res, out, soft = run_chain_batch(fec, eng, items)
        for i, c in enumerate(cbs):
            K, N = cases.dims(c["bg"], c["Z"])
            assert got_iterations(fec, res, i) == want[i][0], (t, i)
            assert (out[i, :K // 8] == want[i][1]).all(), (t, i)
            assert (soft[i, :N] == c["soft"]).all(), (t, i)
            passed_any |= want[i][0] > 0
    assert passed_any


def test_decoder_interface_noiseless_all_Z(fec, engine, oracle):
    """ldpc_enc_dec_test.cpp:285-315 through the LdpcDecoder mirror: one iteration, crc = None, trimmed lengths."""
    dec = fec.LdpcDecoder(engine)
    for bg in (1, 2):
        for Z in cases.ALL_Z[::3] + [384]:
            K, N = cases.dims(bg, Z)
            Kb = K // Z
            msg = oracle.random_bits(K, seed=Z + bg)
            cb = oracle.encode(bg, Z, msg)
            llr = (10 - 20 * cb.astype(np.int16)).astype(np.int8)
            cfg = fec.CodeblockMetadata(base_graph=bg, lifting_size=Z)
            for n_cols in (Kb + 2, (Kb + 2 + N // Z) // 2, N // Z):
                out = np.zeros((K + 7) // 8, np.uint8)
                assert dec.decode(out, llr[:n_cols * Z], None, cfg, max_iterations=1) is None
                assert (np.unpackbits(out)[:K] == msg).all(), (bg, Z, n_cols)


def test_decoder_interface_corner_cases(fec, engine, oracle, orc):
    """Zero LLRs (nullopt + all ones / untouched), unaligned input tail (unclamped), data-dependent layer trimming."""
    dec = fec.LdpcDecoder(engine)
    bg, Z = 1, 52
    K, N = cases.dims(bg, Z)
    cfg = fec.CodeblockMetadata(base_graph=bg, lifting_size=Z)
    out = np.zeros(K // 8, np.uint8)
    assert dec.decode(out, np.zeros(N, np.int8), None, cfg) is None and (out == 0xFF).all()
    out = np.full(K // 8, 0x5A, np.uint8)
    assert dec.decode(out, np.zeros(N, np.int8), fec.CRC24B, cfg) is None and (out == 0x5A).all()
    rng = np.random.default_rng(5)
    for trial in range(40):
        bg = int(rng.choice([1, 2]))
        Z = int(rng.choice([3, 7, 16, 30, 52, 104, 208]))
        K, N = cases.dims(bg, Z)
        n_in = int(rng.integers(K + 2 * Z, N + 1))  # not a multiple of Z in general
        llr, _ = oracle.make_case(bg, Z, N, 0, 1, 0, 0, orc.CRC24B, float(rng.uniform(-3, 3)), seed=trial)
        llr = llr[:n_in].copy()
        llr[rng.integers(0, n_in, 5)] = 127  # infinities inside the span, also in the unclamped tail
        if trial % 3 == 0:
            llr[int(rng.integers(K, n_in)):] = 0  # trailing zeros -> fewer layers
        for crc, ocrc in ((None, -1), (fec.CRC24B, orc.CRC24B)):
            mi = int(rng.choice([1, 3, 6, 10]))
            it, ref_out = oracle.decode(bg, Z, llr, 0, ocrc, mi)
            out = np.zeros((K + 7) // 8, np.uint8)
            got = dec.decode(out, llr, crc, fec.CodeblockMetadata(base_graph=bg, lifting_size=Z), max_iterations=mi)
            assert (got if got is not None else -1) == it, (trial, bg, Z, n_in)
            assert (out == ref_out).all(), (trial, bg, Z, n_in)


def test_rate_dematcher_interface(fec, engine, oracle):
    """ldpc_rate_dematcher::rate_dematch mirror vs oracle: new data and combining, wrap-around, LBRM, fillers."""
    dm = fec.LdpcRateDematcher(engine)
    rng = np.random.default_rng(6)
    for trial in range(120):
        p = cases.random_chain_params(rng, z_pool=[2, 5, 8, 13, 22, 36, 64, 112, 384])
        bg, Z = p["bg"], p["Z"]
        K, N = cases.dims(bg, Z)
        E = int(rng.integers(max(K // 2, p["qm"]), 2 * N + 200)) // p["qm"] * p["qm"]
        E = max(E, p["qm"])
        cfg = fec.CodeblockMetadata(base_graph=bg, lifting_size=Z, rv=p["rv"], mod_order=p["qm"], Nref=p["Nref"],
                                    nof_filler_bits=p["fill"])
        a, b = np.zeros(N, np.int8), np.zeros(N, np.int8)
        for t in range(3):
            llr = rng.integers(-120, 121, E).astype(np.int8)
            oracle.rate_dematch(bg, Z, a, llr, t == 0, p["rv"], p["qm"], p["Nref"], p["fill"])
            dm.rate_dematch(b, llr, t == 0, cfg)
            assert (a == b).all(), (trial, t, p, E)


def test_large_batch_chunked_pipeline_vs_oracle(fec, engine, oracle, orc):
    """Batches of two kernel waves or more take the chunked three-stream pipeline of fec_run_host. 700 small codeblocks,
    once without early stop (every output byte is overwritten: the caller's bytes are not sent to the device) and once
    with early stop at an Es/N0 where some blocks fail, plus three blocks whose input is all zeros (the decoder returns
    without iterating and their output bytes must stay the caller's), against the oracle."""
    rng = np.random.default_rng(2024)
    for early in (False, True):
        items, want = [], []
        for i in range(700):
            bg = int(rng.integers(1, 3))
            Z = int(rng.choice([8, 12, 16, 20, 28, 36]))
            p = cases.random_chain_params(rng, bg=bg, Z=Z)
            K, N = cases.dims(bg, Z)
            p["E"] = max(p["qm"], min(p["E"], N + 64) // p["qm"] * p["qm"])
            p["early"], p["max_iter"] = early, 4
            llr, _ = oracle.make_case(bg, Z, p["E"], p["rv"], p["qm"], p["Nref"], p["fill"], p["poly"], p["snr"] - 1.0, seed=5000 + i)
            if early and i in (5, 300, 650):
                llr[:] = 0
            soft = np.zeros(N, np.int8)
            it, ref_out = oracle.cb_chain(bg, Z, soft, llr, True, p["rv"], p["qm"], p["Nref"], p["fill"], p["poly"], 4, early)
            items.append(dict(p=p, llr=llr, new_data=True, slot=i))
            want.append((it, ref_out, soft))
        n = len(items)
        d = fec.make_descs(n)
        offs = np.cumsum([0] + [((it["llr"].size + 15) // 16) * 16 for it in items])
        llr = np.zeros(int(offs[-1]) + 16, np.int8)
        out_off = np.cumsum([0] + [(cases.dims(it["p"]["bg"], it["p"]["Z"])[0] + 7) // 8 for it in items])
        for i, it in enumerate(items):
            engine.harq_free(it["slot"])
            cases.fill_chain_desc(fec, d, i, it["p"], True, it["slot"], int(offs[i]), int(out_off[i]))
            llr[offs[i]:offs[i] + it["llr"].size] = it["llr"]
        out = np.full(int(out_off[-1]), 0x5A, np.uint8)
        res = engine.run_host(d, llr, out)
        n_fail = 0
        for i, (it, ref_out, ref_soft) in enumerate(want):
            K, N = cases.dims(items[i]["p"]["bg"], items[i]["p"]["Z"])
            assert got_iterations(fec, res, i) == it, (early, i)
            got = out[out_off[i]:out_off[i + 1]]
            n_fail += it < 0
            if early and i in (5, 300, 650):
                assert it < 0 and (got == 0x5A).all(), (early, i)  # nothing to decode: the caller's bytes are left alone
                assert res["status"][i] & fec.RES_TOO_SHORT
            else:
                assert (got == ref_out).all(), (early, i)
            assert (engine.harq_read(items[i]["slot"], N) == ref_soft).all(), (early, i)
        assert not early or n_fail > 10

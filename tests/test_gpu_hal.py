"""GPU tests of the hal::hw_accelerator_pusch_dec mirror (queue entry points of the C ABI), driven with the call sequence
pusch_decoder_hw_impl uses (lib/phy/upper/channel_processors/pusch/pusch_decoder_hw_impl.cpp:149-358), and size-independent
properties at the full BASELINE size."""
import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu


def _decode_tb_through_hal(fec, acc, cfgs, llrs, K):
    """One TB: reserve, configure+enqueue every CB, then poll dequeue / read outputs. Returns (outs, results)."""
    acc.reserve_queue()
    n = len(cfgs)
    enq = 0
    outs = [np.zeros((K + 7) // 8, np.uint8) for _ in range(n)]
    results = [None] * n
    deq = 0
    spins = 0
    while deq < n:
        while enq < n:
            acc.configure_operation(cfgs[enq], enq)
            if not acc.enqueue_operation(llrs[enq], None, enq):
                break
            enq += 1
        while deq < enq:
            if not acc.dequeue_operation(outs[deq], None, deq):
                break
            o = fec.HwPuschDecoderOutputs()
            acc.read_operation_outputs(o, deq, cfgs[deq].absolute_cb_id)
            results[deq] = o
            deq += 1
        spins += 1
        assert spins < 2_000_000, "accelerator never completed"
    acc.free_queue()
    return outs, results


def test_hal_tb_harq_sequence(fec, engine, oracle, orc):
    """A 12-CB TB (BG1, Z=384, 256QAM-like E) over rv {0,2,3,1}; external HARQ keyed by absolute_cb_id."""
    assert fec.HwAcceleratorPuschDec(engine).is_harq_external()
    acc = fec.HwAcceleratorPuschDec(engine)
    bg, Z, qm, fill, n_cb = 1, 384, 8, 16, 12
    K, N = cases.dims(bg, Z)
    Es = [8960 if i < 5 else 8992 for i in range(n_cb)]
    cbs = [oracle.encode(bg, Z, oracle.make_message(bg, Z, fill, orc.CRC24B, seed=300 + i)) for i in range(n_cb)]
    softs = [np.zeros(N, np.int8) for _ in range(n_cb)]
    base_id = 500
    seen_pass = seen_fail = False
    for t, rv in enumerate([0, 2, 3, 1]):
        cfgs, llrs, want = [], [], []
        for i in range(n_cb):
            tx = oracle.rate_match(bg, Z, cbs[i], Es[i], rv, qm, 0, fill)
            llr = oracle.awgn_llr(tx, 4.6 + 0.25 * i, seed=i * 8 + t)
            cfgs.append(fec.HwPuschDecoderConfiguration(
                base_graph_index=bg, modulation=qm, nof_segments=n_cb, rv=rv, cw_length=Es[i], lifting_size=Z,
                Ncb=N, Nref=0, nof_segment_bits=K - fill - 24, nof_filler_bits=fill, max_nof_ldpc_iterations=6,
                use_early_stop=True, new_data=(t == 0), cb_crc_len=24, cb_crc_type=fec.CRC24B,
                absolute_cb_id=base_id + i))
            llrs.append(llr)
            want.append(oracle.cb_chain(bg, Z, softs[i], llr, t == 0, rv, qm, 0, fill, orc.CRC24B, 6, True))
        outs, results = _decode_tb_through_hal(fec, acc, cfgs, llrs, K)
        for i in range(n_cb):
            it, ref_out = want[i]
            assert results[i].CRC_pass == (it > 0), (t, i)
            if it > 0:
                assert results[i].nof_ldpc_iterations == it, (t, i)
            assert (outs[i] == ref_out).all(), (t, i)
            assert (engine.harq_read(base_id + i, N) == softs[i]).all(), (t, i)
            seen_pass |= it > 0
            seen_fail |= it < 0
    assert seen_pass and seen_fail
    for i in range(n_cb):
        acc.free_harq_context_entry(base_id + i)
    assert (engine.harq_read(base_id, N) == 0).all()


def test_hal_mixed_tbs_and_back_pressure(fec, oracle, orc):
    """Small staging (max_batch_cbs = 4) forces several launches and FEC_ERR_QUEUE_FULL retries; CRC16 / CRC24A CBs."""
    eng = fec.FecEngine(device=0, harq_slots=64, max_batch_cbs=4, nof_queues=2)
    try:
        acc = fec.HwAcceleratorPuschDec(eng)
        rng = np.random.default_rng(11)
        cfgs, llrs, want, Ks = [], [], [], []
        bg, Z = 2, 36
        K, N = cases.dims(bg, Z)
        for i in range(23):
            poly = [orc.CRC16, orc.CRC24A, orc.CRC24B][i % 3]
            qm = int(rng.choice([2, 4, 6]))
            E = int(rng.integers(K + 3 * Z, 2 * N)) // qm * qm
            llr, _ = oracle.make_case(bg, Z, E, 0, qm, 0, 0, poly, float(rng.uniform(-6, 2)), seed=700 + i)
            cfgs.append(fec.HwPuschDecoderConfiguration(
                base_graph_index=bg, modulation=qm, rv=0, cw_length=E, lifting_size=Z, Ncb=N,
                nof_filler_bits=0, max_nof_ldpc_iterations=8, use_early_stop=bool(i % 2), new_data=True,
                cb_crc_len=orc.CRC_LEN[poly], cb_crc_type=cases.ORC_TO_HAL_CRC[poly], absolute_cb_id=i))
            llrs.append(llr)
            want.append(oracle.cb_chain(bg, Z, np.zeros(N, np.int8), llr, True, 0, qm, 0, 0, poly, 8, bool(i % 2)))
        outs, results = _decode_tb_through_hal(fec, acc, cfgs, llrs, K)
        for i in range(len(cfgs)):
            assert results[i].CRC_pass == (want[i][0] > 0), i
            assert (outs[i] == want[i][1]).all(), i
    finally:
        eng.close()


def test_full_size_batch_properties(fec, oracle, orc):
    """BASELINE headline shape at bench scale: 1184 CBs of BG1 Z=384 rate 1/3. Properties: every CB that reports CRC pass
    carries exactly the transmitted message; forced-6-iteration and early-stop runs agree on which CBs pass when the
    early-stop run converged within 6; a sample is compared bit-exactly with the oracle."""
    eng = fec.FecEngine(device=0, harq_slots=1184, max_batch_cbs=1184)
    try:
        bg, Z, n_distinct, n = 1, 384, 37, 1184
        K, N = cases.dims(bg, Z)
        base, msgs = [], []
        for i in range(n_distinct):
            llr, msg = oracle.make_case(bg, Z, N, 0, 2, 0, 0, orc.CRC24B, -4.2 + 0.1 * i, seed=9000 + i)
            base.append(llr)
            msgs.append(np.packbits(msg))
        llr = np.stack([base[i % n_distinct] for i in range(n)])
        runs = {}
        for early in (True, False):
            d = fec.make_descs(n)
            d["base_graph"], d["lifting_size"], d["qm"], d["E"] = bg, Z, 2, N
            d["crc_type"], d["max_iter"], d["mode"] = fec.CRC24B, 6, fec.MODE_CHAIN
            d["flags"] = fec.FLAG_NEW_DATA | (fec.FLAG_EARLY_STOP if early else 0)
            d["harq_slot"] = np.arange(n)
            d["llr_offset"] = np.arange(n, dtype=np.uint64) * N
            d["out_offset"] = np.arange(n) * (K // 8)
            out = np.zeros((n, K // 8), np.uint8)
            res = eng.run_host(d, llr, out)
            runs[early] = (res.copy(), out)
            ok = (res["status"] & fec.RES_CRC_PASS) != 0
            assert 0 < ok.sum() < n
            for i in np.nonzero(ok)[0]:
                assert (out[i] == msgs[i % n_distinct]).all(), (early, i)
            # identical inputs give identical results wherever they sit in the batch
            for i in range(n_distinct, n):
                assert res[i] == res[i % n_distinct] and (out[i] == out[i % n_distinct]).all()
        for i in range(n_distinct):
            soft = np.zeros(N, np.int8)
            for early in (True, False):
                it, ref_out = oracle.cb_chain(bg, Z, soft, base[i], True, 0, 2, 0, 0, orc.CRC24B, 6, early)
                res, out = runs[early]
                got = int(res["nof_ldpc_iterations"][i]) if res["status"][i] & fec.RES_CRC_PASS else -1
                assert got == it and (out[i] == ref_out).all(), (early, i)
            assert (eng.harq_read(i, N) == soft).all()
    finally:
        eng.close()


def test_hal_concurrent_instances(fec, oracle, orc):
    """Four accelerator instances (one per thread, each with its own engine queue) share one engine and decode different
    transport blocks at the same time, while a fifth thread runs fec_run_host batches: every result matches the oracle."""
    import threading
    eng = fec.FecEngine(device=0, harq_slots=256, max_batch_cbs=16, nof_queues=4)
    try:
        bg, Z, qm, n_cb, n_thr, n_rounds = 1, 96, 4, 10, 4, 3
        K, N = cases.dims(bg, Z)
        E = (K + 8 * Z) // qm * qm
        work = {}
        for th in range(n_thr + 1):
            for rnd in range(n_rounds):
                items = []
                for i in range(n_cb):
                    llr, _ = oracle.make_case(bg, Z, E, 0, qm, 0, 0, orc.CRC24B, -1.0 + 0.6 * i, seed=5000 + th * 100 + rnd * 10 + i)
                    items.append((llr, oracle.cb_chain(bg, Z, np.zeros(N, np.int8), llr, True, 0, qm, 0, 0, orc.CRC24B, 6, True)))
                work[th, rnd] = items
        errors = []

        def hal_worker(th):
            try:
                acc = fec.HwAcceleratorPuschDec(eng)
                for rnd in range(n_rounds):
                    items = work[th, rnd]
                    cfgs = [fec.HwPuschDecoderConfiguration(
                        base_graph_index=bg, modulation=qm, nof_segments=n_cb, rv=0, cw_length=E, lifting_size=Z, Ncb=N,
                        nof_filler_bits=0, max_nof_ldpc_iterations=6, use_early_stop=True, new_data=True, cb_crc_len=24,
                        cb_crc_type=fec.CRC24B, absolute_cb_id=th * 40 + i) for i in range(n_cb)]
                    outs, results = _decode_tb_through_hal(fec, acc, cfgs, [x[0] for x in items], K)
                    for i, (_, (it, ref_out)) in enumerate(items):
                        if results[i].CRC_pass != (it > 0) or (it > 0 and results[i].nof_ldpc_iterations != it) or \
                                not (outs[i] == ref_out).all():
                            errors.append((th, rnd, i))
            except Exception as ex:  # noqa: BLE001
                errors.append((th, repr(ex)))

        def batch_worker(th):
            try:
                for rnd in range(n_rounds):
                    items = work[th, rnd]
                    d = fec.make_descs(n_cb)
                    d["base_graph"], d["lifting_size"], d["qm"], d["E"] = bg, Z, qm, E
                    d["crc_type"], d["max_iter"], d["mode"] = fec.CRC24B, 6, fec.MODE_CHAIN
                    d["flags"] = fec.FLAG_NEW_DATA | fec.FLAG_EARLY_STOP
                    d["harq_slot"] = th * 40 + np.arange(n_cb)
                    d["llr_offset"] = np.arange(n_cb, dtype=np.uint64) * E
                    d["out_offset"] = np.arange(n_cb) * (K // 8)
                    out = np.zeros((n_cb, K // 8), np.uint8)
                    res = eng.run_host(d, np.stack([x[0] for x in items]), out)
                    for i, (_, (it, ref_out)) in enumerate(items):
                        ok = bool(res["status"][i] & fec.RES_CRC_PASS)
                        if ok != (it > 0) or not (out[i] == ref_out).all():
                            errors.append((th, rnd, i))
            except Exception as ex:  # noqa: BLE001
                errors.append((th, repr(ex)))

        threads = [threading.Thread(target=hal_worker, args=(t,)) for t in range(n_thr)]
        threads.append(threading.Thread(target=batch_worker, args=(n_thr,)))
        for t in threads:
            t.start()
        for t in threads:
            t.join(timeout=120)
        assert not any(t.is_alive() for t in threads), "a worker hung"
        assert not errors, errors[:5]
    finally:
        eng.close()

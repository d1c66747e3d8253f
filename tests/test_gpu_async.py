"""GPU tests of the concurrent / asynchronous host paths of the C ABI: host-call contexts (several threads inside
fec_tb_run_host / fec_run_host at once), fec_tb_enqueue + poll / wait / callback, back-pressure, the engine counters and the
fault path (a refused launch must come back as failed codeblocks / blocks, never as a crash, and must leave the HARQ state
usable). Expected values come from the synchronous entry points, which the other test files check against the oracle."""
import threading
import time

import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu


def _tb_desc(fec, p, slot, early=True):
    d = fec.make_tb_descs(1)
    d["tbs"], d["nof_llrs"], d["base_graph"], d["qm"], d["nof_layers"] = p["tbs"], p["G"], p["bg"], p["qm"], p["nl"]
    d["rv"], d["max_iter"], d["harq_slot"] = 0, 6, slot
    d["flags"] = fec.FLAG_NEW_DATA | (fec.FLAG_EARLY_STOP if early else 0)
    return d


def _make_blocks(orc, oracle, n, seed):
    """n transport blocks with TS 38.214-like sizes (whole-byte payloads), some decodable, some not."""
    rng = np.random.default_rng(seed)
    shapes = [dict(tbs=19992, bg=1, qm=2, nl=1, G=40000, snr=4.0), dict(tbs=90000, bg=1, qm=4, nl=2, G=130000, snr=6.0),
              dict(tbs=29992, bg=1, qm=2, nl=1, G=48000, snr=-6.0), dict(tbs=3000, bg=2, qm=2, nl=1, G=9000, snr=3.0)]
    out = []
    for i in range(n):
        p = shapes[i % len(shapes)]
        tb = rng.integers(0, 256, p["tbs"] // 8, dtype=np.uint8)
        llr = orc.tb_transmit(oracle, orc.tb_encode(oracle, tb, p["bg"]), p["tbs"], p["bg"], 0, p["qm"], 0, p["nl"], p["G"],
                              p["snr"], seed=seed * 100 + i)
        out.append((p, tb, llr))
    return out


@pytest.fixture(scope="module")
def blocks(orc, oracle):
    return _make_blocks(orc, oracle, 8, seed=5)


@pytest.fixture(scope="module")
def async_engine(fec):
    eng = fec.FecEngine(device=0, harq_slots=4096, max_batch_cbs=512, host_contexts=3)
    yield eng
    eng.close()


def _sync_reference(fec, eng, blocks, slot0):
    ref = []
    for i, (p, tb, llr) in enumerate(blocks):
        out = np.full(p["tbs"] // 8, 0xAA, np.uint8)
        res, cbr = eng.tb_run_host(_tb_desc(fec, p, slot0 + 32 * i), llr, out, cb_results=True)
        ref.append((res.copy(), cbr.copy(), out))
    return ref


def test_tb_enqueue_poll_matches_the_synchronous_call(fec, async_engine, blocks):
    eng = async_engine
    ref = _sync_reference(fec, eng, blocks, 0)
    assert [int(r[0]["tb_crc_ok"][0]) for r in ref[:4]] == [1, 1, 0, 1]
    outs = [np.full(p["tbs"] // 8, 0xAA, np.uint8) for p, _, _ in blocks]
    pending, done, nxt, full_seen = {}, {}, 0, 0
    deadline = time.time() + 60
    while len(done) < len(blocks):
        assert time.time() < deadline
        while nxt < len(blocks):
            p, _, llr = blocks[nxt]
            try:
                pending[nxt] = eng.tb_enqueue(_tb_desc(fec, p, 1024 + 32 * nxt), llr, outs[nxt])
            except fec.FecError as err:
                assert "-5" in str(err)  # FEC_ERR_QUEUE_FULL: every context in flight, retry after collecting
                full_seen += 1
                break
            nxt += 1
        for i in list(pending):
            r = eng.tb_poll(pending[i], 1, cb_results=True)
            if r is not None:
                done[i] = r
                del pending[i]
    assert full_seen > 0  # eight blocks through three contexts
    for i, (res, cbr) in done.items():
        for f in fec.TB_RESULT_DTYPE.names:
            assert res[f][0] == ref[i][0][f][0], (i, f)
        n_cb = int(res["nof_codeblocks"][0])
        assert (cbr[0, :n_cb] == ref[i][1][0, :n_cb]).all()
        assert (outs[i] == ref[i][2]).all()
    # a ticket is good for one collection only
    with pytest.raises(fec.FecError):
        eng.tb_poll(next(iter(done)) + 999, 1)


def test_tb_enqueue_with_callback(fec, async_engine, blocks):
    eng = async_engine
    ref = _sync_reference(fec, eng, blocks[:4], 2048)
    got, ev = {}, threading.Event()

    def on_done(user, ticket, status, results, n):
        r = np.ctypeslib.as_array((fec.C.c_uint8 * (16 * n)).from_address(results)).view(fec.TB_RESULT_DTYPE).copy()
        got[ticket] = (status, r)
        if len(got) == 4:
            ev.set()

    cb = fec.TB_CALLBACK(on_done)
    outs = [np.full(p["tbs"] // 8, 0xAA, np.uint8) for p, _, _ in blocks[:4]]
    tickets = []
    for i, (p, _, llr) in enumerate(blocks[:4]):
        while True:
            try:
                tickets.append(eng.tb_enqueue(_tb_desc(fec, p, 3072 + 32 * i), llr, outs[i], callback=cb))
                break
            except fec.FecError:
                time.sleep(0.0005)
    assert ev.wait(30)
    for i, t in enumerate(tickets):
        status, r = got[t]
        assert status == 0
        for f in fec.TB_RESULT_DTYPE.names:
            assert r[f][0] == ref[i][0][f][0], (i, f)
        assert (outs[i] == ref[i][2]).all()


def test_threads_inside_the_synchronous_calls_run_concurrently(fec, async_engine, blocks):
    """Four threads call fec_tb_run_host / fec_run_host at the same time (ctypes drops the GIL): same results as one by one."""
    eng = async_engine
    ref = _sync_reference(fec, eng, blocks, 0)
    results, errors = {}, []

    def worker(k):
        try:
            for rep in range(3):
                for i in range(k, len(blocks), 4):
                    p, _, llr = blocks[i]
                    out = np.full(p["tbs"] // 8, 0xAA, np.uint8)
                    res = eng.tb_run_host(_tb_desc(fec, p, 1024 + 32 * i), llr, out)
                    results[(i, rep)] = (res.copy(), out)
        except Exception as err:  # noqa: BLE001
            errors.append(err)

    threads = [threading.Thread(target=worker, args=(k,)) for k in range(4)]
    [t.start() for t in threads]
    [t.join() for t in threads]
    assert not errors, errors
    for (i, rep), (res, out) in results.items():
        for f in fec.TB_RESULT_DTYPE.names:
            assert res[f][0] == ref[i][0][f][0], (i, rep, f)
        assert (out == ref[i][2]).all()


def test_two_queues_from_two_threads_mixed_lifting_sizes(fec, oracle, orc):
    """launches of small-Z and Z=384 batches from different queues / threads at the same time (the shared-memory
    attribute of the kernel is set once, at engine creation)."""
    eng = fec.FecEngine(device=0, harq_slots=512, max_batch_cbs=64, nof_queues=4)
    try:
        rng = np.random.default_rng(3)
        work = {}
        for k, Z in enumerate((384, 8)):
            items = []
            for i in range(24):
                p = cases.random_chain_params(rng, bg=1, Z=Z)
                p["rv"], p["Nref"] = 0, 0
                p["E"] = min(p["E"], 25344)
                llr, _ = oracle.make_case(1, Z, p["E"], 0, p["qm"], 0, p["fill"], p["poly"], p["snr"], seed=1000 * k + i)
                items.append((p, llr))
            work[k] = items
        expect = {}
        for k, items in work.items():
            d = fec.make_descs(len(items))
            Kb = cases.dims(1, items[0][0]["Z"])[0] // 8
            off = 0
            buf = []
            for i, (p, llr) in enumerate(items):
                cases.fill_chain_desc(fec, d, i, p, True, 256 * k + i, off, i * 1056)
                buf.append(np.pad(llr, (0, (-llr.size) % 16)))
                off += buf[-1].size
            out = np.zeros(len(items) * 1056, np.uint8)
            res = eng.run_host(d, np.concatenate(buf), out)
            expect[k] = (res.copy(), out, d)
        got, errors = {}, []

        def worker(k):
            try:
                q = eng.queue_reserve()
                items = work[k]
                d = expect[k][2]
                for rep in range(20):
                    for i, (p, llr) in enumerate(items):
                        one = d[i:i + 1].copy()
                        assert eng.queue_enqueue(q, one, llr, i)
                    eng.queue_flush(q)
                    seen = {}
                    t0 = time.time()
                    while len(seen) < len(items):
                        assert time.time() - t0 < 30
                        tags, rs = eng.queue_poll(q)
                        for tag, r in zip(tags, rs):
                            seen[int(tag)] = r.copy()
                    got[(k, rep)] = seen
                eng.queue_free(q)
            except Exception as err:  # noqa: BLE001
                errors.append(err)

        threads = [threading.Thread(target=worker, args=(k,)) for k in work]
        [t.start() for t in threads]
        [t.join() for t in threads]
        assert not errors, errors
        for (k, rep), seen in got.items():
            res = expect[k][0]
            for i in range(len(work[k])):
                assert seen[i]["status"] == res["status"][i] and seen[i]["nof_ldpc_iterations"] == res["nof_ldpc_iterations"][i], (k, rep, i)
    finally:
        eng.close()


def test_harq_zero_flag_stays_out_of_the_callers_descriptors(fec, engine, oracle, orc):
    """a descriptor array reused for a retransmission after its slot was freed must not carry a stale
    FEC_FLAG_HARQ_ZERO (the engine applies it to its own copy only)."""
    bg, Z, E, qm = 1, 384, 9000, 2
    K, N = cases.dims(bg, Z)
    msg = oracle.make_message(bg, Z, 0, orc.CRC24B, seed=77)
    cw = oracle.encode(bg, Z, msg)
    slot = 900
    engine.harq_write(slot, np.full(N, 5, np.int8))
    engine.harq_free(slot)
    d = fec.make_descs(1)
    d["base_graph"], d["lifting_size"], d["qm"], d["E"], d["crc_type"], d["max_iter"] = bg, Z, qm, E, fec.CRC24B, 6
    d["mode"], d["harq_slot"], d["flags"] = fec.MODE_CHAIN, slot, fec.FLAG_NEW_DATA | fec.FLAG_EARLY_STOP
    soft = np.zeros(N, np.int8)
    out = np.zeros(K // 8, np.uint8)
    for t, rv in enumerate((0, 2)):
        llr = oracle.awgn_llr(oracle.rate_match(bg, Z, cw, E, rv, qm, 0, 0), -3.5, seed=50 + t)
        d["rv"] = rv
        d["flags"] = (fec.FLAG_NEW_DATA if t == 0 else 0) | fec.FLAG_EARLY_STOP  # only NEW_DATA is edited, as a caller would
        res = engine.run_host(d, llr, out)
        assert not (int(d["flags"][0]) & fec.FLAG_HARQ_ZERO)
        it, _ = oracle.cb_chain(bg, Z, soft, llr, t == 0, rv, qm, 0, 0, orc.CRC24B, 6, True)
        assert (engine.harq_read(slot, N) == soft).all(), t  # rv2 was COMBINED with rv0, not with zeros
        assert (int(res["nof_ldpc_iterations"][0]) if res["status"][0] & fec.RES_CRC_PASS else -1) == it


def test_refused_launch_comes_back_as_failed_codeblocks(fec, orc, oracle, blocks):
    eng = fec.FecEngine(device=0, harq_slots=512, max_batch_cbs=256, host_contexts=2)
    try:
        p, tb, llr = blocks[0]
        # transport-block call: error code AND "nothing delivered"; the next call works and the freed slots still read as zeros
        out = np.full(p["tbs"] // 8, 0xAA, np.uint8)
        res = eng.tb_run_host(_tb_desc(fec, p, 0), llr, out)
        assert res["tb_crc_ok"][0] == 1  # releases (frees) the slots
        eng.debug_fail_launches(1)
        out2 = np.full(p["tbs"] // 8, 0xAA, np.uint8)
        with pytest.raises(fec.FecError, match="injected"):
            eng.tb_run_host(_tb_desc(fec, p, 0), llr, out2)
        assert (out2 == 0xAA).all()
        d = _tb_desc(fec, p, 0)
        d["flags"] = fec.FLAG_EARLY_STOP  # a retransmission into the freed slots: must combine with zeros, i.e. decode like new data
        res3 = eng.tb_run_host(d, llr, out2)
        assert res3["tb_crc_ok"][0] == 1 and (out2 == tb).all()
        # asynchronous call: the failure is reported by enqueue, the context is free again
        eng.debug_fail_launches(1)
        with pytest.raises(fec.FecError, match="injected"):
            eng.tb_enqueue(_tb_desc(fec, p, 64), llr, out2)
        t = eng.tb_enqueue(_tb_desc(fec, p, 64), llr, out2)
        assert eng.tb_poll(t, 1, wait=True)["tb_crc_ok"][0] == 1
        # codeblock call: every operation of the call reported as failed
        cbs, _ = fec.tb_segment(_tb_desc(fec, p, 128))
        cbs["out_offset"] = np.arange(cbs.size) * 1056
        cb_out = np.zeros(cbs.size * 1056, np.uint8)
        results = np.zeros(cbs.size, fec.RESULT_DTYPE)
        results["status"] = 0xFF
        eng.debug_fail_launches(1)
        with pytest.raises(fec.FecError, match="injected"):
            eng.run_host(cbs, llr, cb_out, results=results)
        assert (results["status"] == fec.RES_DONE).all() and (results["nof_ldpc_iterations"] == 6).all()
        r = eng.run_host(cbs, llr, cb_out, results=results)
        assert (r["status"] & fec.RES_CRC_PASS).all()
        st = eng.stats()
        assert st["faults"] == 3 and st["tb_crc_ok"] == 3 and st["transport_blocks"] >= 3
        assert st["results"] == sum(st["iterations_hist"]) and st["crc_pass"] <= st["results"]
        assert st["codeblocks"] >= st["results"] and st["cb_launches"] >= 5
    finally:
        eng.close()


@pytest.mark.parametrize("max_wait_us", [0, 200])
def test_cross_queue_batcher_many_threads(fec, oracle, orc, max_wait_us):
    """Eight threads, one queue each, batches of 6 codeblocks and batch enqueue: the batcher coalesces what several queues
    hand over into shared launches; every result equals the one-call reference; back-pressure is reported, never blocks."""
    eng = fec.FecEngine(device=0, harq_slots=2048, max_batch_cbs=256, nof_queues=8, queue_batch_cbs=6, queue_batches=3,
                        max_wait_us=max_wait_us)
    try:
        bg, Z, qm, fill, E = 1, 384, 8, 16, 8960
        K, N = cases.dims(bg, Z)
        n_cb = 20
        llrs = []
        for i in range(n_cb):
            llr, _ = oracle.make_case(bg, Z, E, 0, qm, 0, fill, orc.CRC24B, 3.0 + 0.4 * i, seed=900 + i)
            llrs.append(llr)
        d = fec.make_descs(n_cb)
        d["base_graph"], d["lifting_size"], d["qm"], d["E"], d["nof_filler_bits"] = bg, Z, qm, E, fill
        d["crc_type"], d["max_iter"], d["mode"] = fec.CRC24B, 6, fec.MODE_CHAIN
        d["flags"] = fec.FLAG_NEW_DATA | fec.FLAG_EARLY_STOP
        d["harq_slot"] = 1900 + np.arange(n_cb)
        d["llr_offset"], d["out_offset"] = np.arange(n_cb, dtype=np.uint64) * E, np.arange(n_cb) * 1056
        ref_out = np.zeros(n_cb * 1056, np.uint8)
        ref = eng.run_host(d, np.concatenate(llrs), ref_out).copy()
        assert 0 < int((ref["status"] & fec.RES_CRC_PASS).sum()) < n_cb  # both outcomes present
        errors, full = [], [0]

        def worker(k):
            try:
                q = eng.queue_reserve()
                mine = d.copy()
                mine["harq_slot"] = 200 * k + np.arange(n_cb)
                for rep in range(12):
                    seen, nxt = {}, 0
                    t0 = time.time()
                    while len(seen) < n_cb:
                        assert time.time() - t0 < 30
                        if nxt < n_cb:
                            took = eng.queue_enqueue_batch(q, mine[nxt:], llrs[nxt:], np.arange(nxt, n_cb))
                            full[0] += (took < n_cb - nxt)
                            nxt += took
                            if nxt == n_cb:
                                eng.queue_flush(q)
                        tags, rs = eng.queue_poll(q)
                        for tag, r in zip(tags, rs):
                            out = np.zeros(K // 8, np.uint8)
                            eng.queue_fetch(q, tag, out)
                            seen[int(tag)] = (r.copy(), out)
                    for i in range(n_cb):
                        r, out = seen[i]
                        assert r["status"] == ref["status"][i] and r["nof_ldpc_iterations"] == ref["nof_ldpc_iterations"][i]
                        if r["status"] & fec.RES_CRC_PASS:
                            assert (out == ref_out[i * 1056:i * 1056 + K // 8]).all()
                eng.queue_free(q)
            except Exception as err:  # noqa: BLE001
                errors.append(err)

        threads = [threading.Thread(target=worker, args=(k,)) for k in range(8)]
        [t.start() for t in threads]
        [t.join() for t in threads]
        assert not errors, errors
        st = eng.stats()
        assert full[0] > 0 and st["queue_full"] > 0  # 20 codeblocks against a ring of 3 x 6
        assert st["dispatches"] >= 1 and st["dispatch_batches"] >= st["dispatches"]
        assert st["max_queue_depth"] <= 18
    finally:
        eng.close()

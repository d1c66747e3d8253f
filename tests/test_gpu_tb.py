"""Transport-block layer on the GPU, through the C ABI (fec_tb_run_host) and the PuschDecoder mirror: parity with the TB
oracle (pinned against the reference's pusch_decoder_impl in tests/test_tb_cpu.py) and with the committed fixture."""
import os

import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "tb_golden.npz")


@pytest.fixture(scope="module")
def tb_engine(fec):
    eng = fec.FecEngine(device=0, harq_slots=2048, max_batch_cbs=1024)
    yield eng
    eng.close()


def _tb_desc(fec, p, rv, new_data, slot, llr_offset=0, out_offset=0, max_iter=6, early=True):
    d = fec.make_tb_descs(1)
    d["tbs"], d["nof_llrs"], d["base_graph"], d["qm"], d["nof_layers"] = p["tbs"], p["G"], p["bg"], p["qm"], p["nl"]
    d["rv"], d["max_iter"], d["harq_slot"] = rv, max_iter, slot
    d["flags"] = (fec.FLAG_NEW_DATA if new_data else 0) | (fec.FLAG_EARLY_STOP if early else 0)
    d["llr_offset"], d["out_offset"] = llr_offset, out_offset
    return d


def test_tb_golden_sequences(fec, tb_engine):
    """The reference's own outputs (tests/golden/tb_golden.npz): delivered bytes, tb_crc_ok, soft buffers."""
    G = np.load(GOLDEN)
    for i, k in enumerate(sorted({k[:3] for k in G.files})):
        tbs, bg, qm, nl, nllr, C_, ntx = [int(x) for x in G[k + "_params"]]
        p = dict(tbs=tbs, bg=bg, qm=qm, nl=nl, G=nllr)
        slot = 40 * i
        for t in range(ntx):
            rv, ok = [int(x) for x in G["%s_t%d_rv" % (k, t)]]
            out = np.full(tbs // 8, 0xAA, np.uint8)
            res, cbr = tb_engine.tb_run_host(_tb_desc(fec, p, rv, t == 0, slot), G["%s_t%d_llr" % (k, t)], out, cb_results=True)
            assert bool(res["tb_crc_ok"][0]) == bool(ok) and int(res["nof_codeblocks"][0]) == C_, (k, t)
            assert (out == G["%s_t%d_out" % (k, t)]).all(), (k, t)
            if not ok:
                flags, soft = G["%s_t%d_flags" % (k, t)], G["%s_t%d_soft" % (k, t)]
                for r in range(C_):
                    assert (tb_engine.harq_read(slot + r, soft.shape[1]) == soft[r]).all(), (k, t, r)
                assert int(res["nof_cb_crc_ok"][0]) == int(flags.sum()), (k, t)


def test_tb_random_harq_vs_oracle(fec, tb_engine, orc, oracle):
    """Random transport blocks (both base graphs, 1-4 codeblocks, every Qm, 1-4 layers) over rv {0,2,3,1}, several per
    call with explicit slot lists; everything compared with the TB oracle after every transmission."""
    rng = np.random.default_rng(99)
    to = orc.TbOracle(oracle)
    n_fail = n_skip = 0
    for rnd in range(8):
        n_tb = 4
        ps = [cases.random_tb_params(rng) for _ in range(n_tb)]
        gs = [orc.tb_geometry(p["tbs"], p["bg"], p["qm"], p["nl"], p["G"]) for p in ps]
        tbs_ = [rng.integers(0, 256, p["tbs"] // 8, dtype=np.uint8) for p in ps]
        enc = [orc.tb_encode(oracle, tb, p["bg"]) for tb, p in zip(tbs_, ps)]
        # scattered, non-contiguous slots
        perm = rng.permutation(2048)[:sum(g["C"] for g in gs)].astype(np.uint32)
        first = np.concatenate([[0], np.cumsum([g["C"] for g in gs])])
        alive = list(range(n_tb))
        for t, rv in enumerate([0, 2, 3, 1]):
            if not alive:
                break
            llrs = [orc.tb_transmit(oracle, enc[i], ps[i]["tbs"], ps[i]["bg"], rv, ps[i]["qm"], 0, ps[i]["nl"], ps[i]["G"],
                                    ps[i]["snr"] - 1.5, seed=rnd * 1000 + i * 16 + t) for i in alive]
            lo = np.concatenate([[0], np.cumsum([(x.size + 63) // 64 * 64 for x in llrs])])
            oo = np.concatenate([[0], np.cumsum([ps[i]["tbs"] // 8 for i in alive])])
            buf = np.zeros(int(lo[-1]), np.int8)
            d = fec.make_tb_descs(len(alive))
            for j, i in enumerate(alive):
                buf[lo[j]:lo[j] + llrs[j].size] = llrs[j]
                d[j:j + 1] = _tb_desc(fec, ps[i], rv, t == 0, 0, int(lo[j]), int(oo[j]))
                d["slot_list"][j] = first[i]
            out = np.full(int(oo[-1]), 0xAA, np.uint8)
            res, cbr = tb_engine.tb_run_host(d, buf, out, slots=perm, cb_results=True)
            still = []
            for j, i in enumerate(alive):
                want = np.full(ps[i]["tbs"] // 8, 0xAA, np.uint8)
                w = to.decode(i, ps[i]["tbs"], ps[i]["bg"], rv, ps[i]["qm"], 0, ps[i]["nl"], t == 0, 6, True, llrs[j], want)
                got = out[oo[j]:oo[j + 1]]
                assert bool(res["tb_crc_ok"][j]) == w["tb_crc_ok"] and bool(res["tb_delivered"][j]) == w["delivered"], (rnd, t, i)
                assert int(res["nof_codeblocks"][j]) == gs[i]["C"] and (got == want).all(), (rnd, t, i)
                dec = [r for r in range(gs[i]["C"]) if not (cbr[j, r]["status"] & fec.RES_SKIPPED)]
                n_skip += gs[i]["C"] - len(dec)
                assert int(res["nof_cb_decoded"][j]) == len(dec) == len(w["iters"])
                its = [int(cbr[j, r]["nof_ldpc_iterations"]) if (cbr[j, r]["status"] & fec.RES_CRC_PASS) else 6 for r in dec]
                assert its == w["iters"], (rnd, t, i)
                assert int(res["iter_sum"][j]) == sum(its) and int(res["iter_min"][j]) == min(its) and int(res["iter_max"][j]) == max(its)
                if w["tb_crc_ok"]:
                    assert (got == tbs_[i]).all()
                else:
                    n_fail += 1
                    still.append(i)
                    mine = to.peek(i)
                    for r in range(gs[i]["C"]):
                        assert (tb_engine.harq_read(int(perm[first[i] + r]), gs[i]["N"]) == mine["soft"][r]).all(), (rnd, t, i, r)
            alive = still
    assert n_fail >= 8  # partially decoded blocks (skipped codeblocks) are forced in test_tb_config3_full_size


def test_tb_config3_full_size(fec, orc, oracle):
    """BASELINE config 3: one 273-PRB, 4-layer, 256QAM transport block (TBS 1 277 992, 152 codeblocks, Z = 384)."""
    eng = fec.FecEngine(device=0, harq_slots=160, max_batch_cbs=160)
    p = dict(tbs=1277992, bg=1, qm=8, nl=4, G=1362816)
    g = orc.tb_geometry(p["tbs"], 1, 8, 4, p["G"])
    rng = np.random.default_rng(3)
    tb = rng.integers(0, 256, p["tbs"] // 8, dtype=np.uint8)
    enc = orc.tb_encode(oracle, tb, 1)
    llr = orc.tb_transmit(oracle, enc, p["tbs"], 1, 0, 8, 0, 4, p["G"], 8.0, seed=1)
    out = np.full(p["tbs"] // 8, 0xAA, np.uint8)
    res, cbr = eng.tb_run_host(_tb_desc(fec, p, 0, True, 0), llr, out, cb_results=True)
    assert int(res["nof_codeblocks"][0]) == 152 and bool(res["tb_crc_ok"][0]) and (out == tb).all()
    # per-codeblock iteration counts against the oracle on a sample of the codeblocks
    for r in (0, 39, 40, 151):
        soft = np.zeros(g["N"], np.int8)
        it, _ = oracle.cb_chain(1, 384, soft, llr[g["cw_off"][r]:g["cw_off"][r] + g["rm"][r]], True, 0, 8, 0, g["fillers"],
                                orc.CRC24B, 6, True)
        assert it == int(cbr[0, r]["nof_ldpc_iterations"]) and (cbr[0, r]["status"] & fec.RES_CRC_PASS)
    # a corrupted codeword fails the block and leaves the output alone
    bad = llr.copy()
    bad[g["cw_off"][7]:g["cw_off"][7] + g["rm"][7]] = 0
    out2 = np.full(p["tbs"] // 8, 0xAA, np.uint8)
    res2 = eng.tb_run_host(_tb_desc(fec, p, 0, True, 0), bad, out2)
    assert not res2["tb_crc_ok"][0] and not res2["tb_delivered"][0] and int(res2["nof_cb_crc_ok"][0]) == 151
    assert (out2 == 0xAA).all()
    # the retransmission only decodes the missing codeblock and completes the block
    # (rv 0 again: at this code rate rv 2 alone is not self-decodable and codeblock 7 has nothing to combine with)
    llr3 = orc.tb_transmit(oracle, enc, p["tbs"], 1, 0, 8, 0, 4, p["G"], 8.0, seed=2)
    res3 = eng.tb_run_host(_tb_desc(fec, p, 0, False, 0), llr3, out2)
    assert int(res3["nof_cb_decoded"][0]) == 1 and bool(res3["tb_crc_ok"][0]) and (out2 == tb).all()
    eng.close()


def test_tb_crc_false_positive_resets_flags(fec, tb_engine, orc, oracle):
    """All codeblock CRCs good but a bad TB CRC: the block is delivered, flagged bad, and every codeblock is decoded again
    on the next transmission (rx_buffer::reset_codeblocks_crc)."""
    p = dict(tbs=20000, bg=1, qm=2, nl=1, G=40000)
    g = orc.tb_geometry(p["tbs"], 1, 2, 1, p["G"])
    assert g["C"] == 3
    rng = np.random.default_rng(8)
    tb = rng.integers(0, 256, p["tbs"] // 8, dtype=np.uint8)
    # corrupt the TB CRC before segmentation: re-encode a stream whose last byte differs
    stream_tb = tb.copy()
    enc = orc.tb_encode(oracle, stream_tb, 1)
    # flip one payload bit of codeblock 1 and fix up its CRC24B so that only the TB CRC can notice
    K = g["K"]
    msg = np.zeros(K + 32, np.uint8)
    per = g["data_bits"]
    bits = np.unpackbits(tb)
    msg[:per] = bits[per:2 * per]
    msg[5] ^= 1
    oracle.lib.orc_crc_attach_bits(orc.CRC24B, msg.ctypes.data_as(__import__("ctypes").c_void_p), per)
    enc[1] = oracle.encode(1, g["Z"], msg[:K].copy())
    to = orc.TbOracle(oracle)
    slot = 1500
    for t, rv in enumerate([0, 2]):
        llr = orc.tb_transmit(oracle, enc, p["tbs"], 1, rv, 2, 0, 1, p["G"], 6.0, seed=50 + t)
        out = np.full(p["tbs"] // 8, 0xAA, np.uint8)
        want = out.copy()
        res = tb_engine.tb_run_host(_tb_desc(fec, p, rv, t == 0, slot), llr, out)
        w = to.decode(7, p["tbs"], 1, rv, 2, 0, 1, t == 0, 6, True, llr, want)
        assert w["delivered"] and not w["tb_crc_ok"]
        assert bool(res["tb_delivered"][0]) and not res["tb_crc_ok"][0] and (out == want).all()
        assert int(res["nof_cb_decoded"][0]) == 3  # flags were reset after the first transmission


def test_pusch_decoder_mirror(fec, tb_engine, orc, oracle):
    """pusch_decoder call sequence: new_data -> on_new_softbits (in pieces) -> on_end_softbits -> notifier."""
    p = dict(tbs=12000, bg=1, qm=4, nl=2, G=18000)
    rng = np.random.default_rng(21)
    tb = rng.integers(0, 256, p["tbs"] // 8, dtype=np.uint8)
    enc = orc.tb_encode(oracle, tb, 1)
    dec = fec.PuschDecoder(tb_engine)
    got = []
    for t, rv in enumerate([0, 2, 3, 1]):
        llr = orc.tb_transmit(oracle, enc, p["tbs"], 1, rv, 4, 0, 2, p["G"], -0.5, seed=300 + t)
        out = np.zeros(p["tbs"] // 8, np.uint8)
        cfg = fec.PuschDecoderConfiguration(base_graph=1, rv=rv, mod_order=4, nof_layers=2, new_data=(t == 0))
        buf = dec.new_data(out, 1700, got.append, cfg)
        buf.on_new_softbits(llr[:7000])
        buf.on_new_softbits(llr[7000:])
        buf.on_end_softbits()
        assert got[-1].nof_codeblocks_total == 2
        if got[-1].tb_crc_ok:
            assert (out == tb).all()
            break
    assert got[-1].tb_crc_ok and len(got) >= 2


def test_multi_cell_batch_config5_sharded(fec, orc, oracle):
    """BASELINE config 5: one uplink slot of 32 cells with an MCS mix (256QAM R948 full-slot block of 152 codeblocks, 64QAM
    R873, QPSK R120 = BG2 with repetition), decoded (a) as one batch on one engine and (b) sharded by cell_id mod n over
    n = 2, 4, 8 engines-worth of calls (one GPU here: the shards run one after the other). No data-path collective exists,
    so every shard must reproduce the unsharded results bit for bit; every block must come back intact; a sample of the
    codeblocks is compared with the oracle."""
    nsym = 42588 * 4  # 273 PRB x 13 data symbols x 12 subcarriers x 4 layers
    mcs = [dict(tbs=1277992, bg=1, qm=8, nl=4, G=nsym * 8, snr=8.0),
           dict(tbs=871384, bg=1, qm=6, nl=4, G=nsym * 6, snr=6.0),
           dict(tbs=39920, bg=2, qm=2, nl=4, G=nsym * 2, snr=-4.0)]
    n_cells = 32
    cell_mcs = [0 if c % 8 == 0 else (1 if c % 2 == 0 else 2) for c in range(n_cells)]  # 4 + 12 + 16 cells
    rng = np.random.default_rng(55)
    tbs, llrs, geos, slot0, nslots = [], [], [], [], 0
    enc_cache = {}
    for c in range(n_cells):
        p = mcs[cell_mcs[c]]
        g = orc.tb_geometry(p["tbs"], p["bg"], p["qm"], p["nl"], p["G"])
        tb = rng.integers(0, 256, p["tbs"] // 8, dtype=np.uint8)
        enc = orc.tb_encode(oracle, tb, p["bg"])
        llrs.append(orc.tb_transmit(oracle, enc, p["tbs"], p["bg"], 0, p["qm"], 0, p["nl"], p["G"], p["snr"], seed=900 + c))
        tbs.append(tb)
        geos.append(g)
        slot0.append(nslots)
        nslots += g["C"]
    assert any(g["rm"][0] > g["N"] for g in geos), "the QPSK R120 cells must exercise repetition (E > N)"
    eng = fec.FecEngine(device=0, harq_slots=nslots, max_batch_cbs=nslots, max_batch_llrs=sum(x.size for x in llrs) + 64)

    def run(cells):
        d = fec.make_tb_descs(len(cells))
        lo, oo, parts = 0, 0, []
        for i, c in enumerate(cells):
            p = mcs[cell_mcs[c]]
            d[i:i + 1] = _tb_desc(fec, p, 0, True, slot0[c], lo, oo)
            parts.append(llrs[c])
            lo += llrs[c].size
            oo += p["tbs"] // 8
        out = np.zeros(oo, np.uint8)
        res, cbr = eng.tb_run_host(d, np.concatenate(parts), out, cb_results=True)
        outs, o = [], 0
        for c in cells:
            n = mcs[cell_mcs[c]]["tbs"] // 8
            outs.append(out[o:o + n].copy())
            o += n
        return res, cbr, outs

    res0, cbr0, out0 = run(list(range(n_cells)))
    for c in range(n_cells):
        assert bool(res0["tb_crc_ok"][c]) and (out0[c] == tbs[c]).all(), c
        assert int(res0["nof_codeblocks"][c]) == geos[c]["C"]
    # oracle sample: first and last codeblock of one cell per MCS
    for c in (0, 2, 1):
        p, g = mcs[cell_mcs[c]], geos[c]
        for r in (0, g["C"] - 1):
            soft = np.zeros(g["N"], np.int8)
            it, _ = oracle.cb_chain(p["bg"], g["Z"], soft, llrs[c][g["cw_off"][r]:g["cw_off"][r] + g["rm"][r]], True, 0,
                                    p["qm"], 0, g["fillers"], orc.CRC24B if g["C"] > 1 else (orc.CRC24A if p["tbs"] > 3824 else orc.CRC16),
                                    6, True)
            assert it == int(cbr0[c, r]["nof_ldpc_iterations"]), (c, r)
            assert (eng.harq_read(slot0[c] + r, g["N"]) == soft).all() or bool(res0["tb_crc_ok"][c])  # slots are freed on success
    for world in (2, 4, 8):
        for rank in range(world):
            cells = fec.cells_of_rank(n_cells, world, rank)
            res, cbr, outs = run(cells)
            for i, c in enumerate(cells):
                assert res[i].tobytes() == res0[c].tobytes(), (world, rank, c)
                assert (outs[i] == out0[c]).all(), (world, rank, c)
                assert cbr[i].tobytes() == cbr0[c].tobytes(), (world, rank, c)
    eng.close()


def test_harq_pool_drives_the_tb_interface(fec, orc, oracle):
    """HARQ processes of several UEs through the engine's buffer pool (fec_harq_pool_*), the way
    pusch_processor uses rx_buffer_pool: reserve(new data) -> decode -> unlock on a failed CRC / release on a good one,
    retransmissions find their buffer by identifier, freed slots are reused by other processes and read as zeros, a
    process that is never retransmitted expires. Results against the TB oracle (one oracle buffer per HARQ process)."""
    rng = np.random.default_rng(314)
    eng = fec.FecEngine(device=0, harq_slots=64, max_batch_cbs=64)
    pool = fec.HarqPool(eng, nof_buffers=6, nof_codeblocks=64, expire_timeout_slots=8)
    to = orc.TbOracle(oracle)
    procs = []
    for i in range(6):
        p = cases.random_tb_params(rng, max_bits_bg1=26000, max_bits_bg2=9000)
        tb = rng.integers(0, 256, p["tbs"] // 8, dtype=np.uint8)
        procs.append(dict(p=p, g=orc.tb_geometry(p["tbs"], p["bg"], p["qm"], p["nl"], p["G"]), tb=tb, ident=(0x4600 + i) << 8 | (i % 4),
                          enc=orc.tb_encode(oracle, tb, p["bg"]), tx=0, done=False, obuf=i))
    delivered = expired_checked = 0
    slot = 0
    for rnd in range(5):
        for i, P in enumerate(procs):
            if P["done"] or (i == 5 and rnd > 0):  # process 5 sends once, fails, and is left to expire
                continue
            p, rv = P["p"], [0, 2, 3, 1][P["tx"] % 4]
            st, buf = pool.reserve(slot, P["ident"], P["g"]["C"], P["tx"] == 0)
            assert st == fec.POOL_OK, (rnd, i, st)
            slots = pool.codeblocks(buf)
            snr = p["snr"] - (9.0 if i == 5 else 2.0)
            llr = orc.tb_transmit(oracle, P["enc"], p["tbs"], p["bg"], rv, p["qm"], 0, p["nl"], p["G"], snr, seed=rnd * 50 + i)
            d = fec.make_tb_descs(1)
            d[0:1] = _tb_desc(fec, p, rv, P["tx"] == 0, 0, 0, 0)
            d["slot_list"][0] = 0
            out = np.zeros(p["tbs"] // 8, np.uint8)
            res = eng.tb_run_host(d, llr, out, slots=slots)
            want = np.zeros(p["tbs"] // 8, np.uint8)
            w = to.decode(P["obuf"], p["tbs"], p["bg"], rv, p["qm"], 0, p["nl"], P["tx"] == 0, 6, True, llr, want)
            assert bool(res["tb_crc_ok"][0]) == w["tb_crc_ok"], (rnd, i)
            if w["tb_crc_ok"]:
                assert (out == P["tb"]).all()
                pool.release(buf)
                P["done"] = True
                delivered += 1
            else:
                mine = to.peek(P["obuf"])
                for r in range(P["g"]["C"]):
                    assert (eng.harq_read(int(slots[r]), P["g"]["N"]) == mine["soft"][r]).all(), (rnd, i, r)
                pool.unlock(buf)
                P["tx"] += 1
        slot += 3
        pool.run_slot(slot)
        if rnd == 3:
            # process 5 was reserved at slot 0 with a timeout of 8 slots: gone by now, its retransmission is refused and its
            # slots are back in the free list, zeroed for the next owner
            assert pool.reserve(slot, procs[5]["ident"], procs[5]["g"]["C"], False)[0] == fec.POOL_RETX_UNKNOWN
            expired_checked += 1
    assert delivered >= 4 and expired_checked == 1
    assert pool.nof_free_codeblocks() == 64 - sum(P["g"]["C"] for P in procs[:5] if not P["done"])
    # a new process on recycled slots starts from zeros: decoding a clean block from them must match a fresh oracle buffer
    P = procs[0]
    st, buf = pool.reserve(slot, 0x7777 << 8, P["g"]["C"], True)
    assert st == fec.POOL_OK
    slots = pool.codeblocks(buf)
    for s in slots:
        assert (eng.harq_read(int(s), 64) == 0).all()
    pool.release(buf)
    pool.close()
    eng.close()

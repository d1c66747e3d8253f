"""Small batches of every kernel for compute-sanitizer (memcheck / racecheck / initcheck), each checked against the oracle:
packed decode at Z in {4, 52, 384} with HARQ combining, generic path at odd Z, dematch-only and decoder-only operations,
transport blocks (gather + bit-gather assembly), PDSCH encoder, soft demodulator.
    compute-sanitizer --tool racecheck python profiles/scripts/sanitizer_case.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge  # noqa: E402
import cases  # noqa: E402

fec = ge.load_package().fec
orc = ge.load_oracle()
o = orc.Oracle()
eng = fec.FecEngine(device=0, harq_slots=256, max_batch_cbs=64, host_contexts=2)
rng = np.random.default_rng(11)
checked = 0
for Z in (4, 52, 384, 7, 15):
    for bg in (1, 2):
        K, N = cases.dims(bg, Z)
        p = cases.random_chain_params(rng, bg=bg, Z=Z)
        p["early"], p["max_iter"], p["Nref"], p["rv"] = True, 4, 0, 0
        p["E"] = min(p["E"], N)
        p["E"] -= p["E"] % p["qm"]
        msg = o.make_message(bg, Z, p["fill"], p["poly"], seed=Z)
        cw = o.encode(bg, Z, msg)
        soft = np.zeros(N, np.int8)
        for t, rv in enumerate((0, 2)):
            llr = o.awgn_llr(o.rate_match(bg, Z, cw, p["E"], rv, p["qm"], 0, p["fill"]), p["snr"] - 3.0, seed=Z * 4 + t)
            d = fec.make_descs(1)
            q = dict(p, rv=rv)
            cases.fill_chain_desc(fec, d, 0, q, t == 0, 3, 0, 0)
            out = np.zeros((K + 7) // 8, np.uint8)
            res = eng.run_host(d, llr, out)
            it, ref_out = o.cb_chain(bg, Z, soft, llr, t == 0, rv, p["qm"], 0, p["fill"], p["poly"], 4, True)
            got = int(res["nof_ldpc_iterations"][0]) if res["status"][0] & fec.RES_CRC_PASS else -1
            assert got == it and (eng.harq_read(3, N) == soft).all(), (Z, bg, t)
            checked += 1
# decoder-only with an unaligned tail, dematch-only
K, N = cases.dims(1, 52)
llr = rng.integers(-60, 60, K + 2 * 52 + 17).astype(np.int8)
out = np.zeros(K // 8, np.uint8)
assert fec.LdpcDecoder(eng) is not None
d = fec.make_descs(1)
d["base_graph"], d["lifting_size"], d["E"], d["max_iter"], d["mode"], d["crc_type"], d["flags"] = 1, 52, llr.size, 3, fec.MODE_DECODE, fec.CRC24B, fec.FLAG_NO_CRC
eng.run_host(d, llr, out)
d["mode"], d["E"], d["qm"], d["flags"], d["harq_slot"] = fec.MODE_DEMATCH, 2000, 2, fec.FLAG_NEW_DATA, 9
eng.run_host(d, rng.integers(-60, 60, 2000).astype(np.int8), None)
# transport blocks: whole-byte payloads (gather kernel) and an odd size (bit-gather assembly kernel)
for tbs, G in ((19992, 40000), (20000, 40000)):
    tb = rng.integers(0, 256, tbs // 8, dtype=np.uint8)
    tl = orc.tb_transmit(o, orc.tb_encode(o, tb, 1), tbs, 1, 0, 2, 0, 1, G, 4.0, seed=tbs)
    td = fec.make_tb_descs(1)
    td["tbs"], td["nof_llrs"], td["base_graph"], td["qm"], td["nof_layers"], td["max_iter"] = tbs, G, 1, 2, 1, 6
    td["harq_slot"], td["flags"] = 32, fec.FLAG_NEW_DATA | fec.FLAG_EARLY_STOP
    tout = np.zeros(tbs // 8, np.uint8)
    r = eng.tb_run_host(td, tl, tout)
    assert r["tb_crc_ok"][0] == 1 and (tout == tb).all(), tbs
    checked += 1
# PDSCH encoder
ed = fec.make_enc_descs(2)
ed["base_graph"], ed["lifting_size"], ed["qm"], ed["E"], ed["rv"] = 1, 384, 8, 8960, np.array([0, 2])
ed["nof_filler_bits"], ed["in_offset"], ed["out_offset"] = 16, np.arange(2) * 1056, np.arange(2) * 1120
msgs = [o.make_message(1, 384, 16, orc.CRC24B, 40 + i) for i in range(2)]
eout = np.zeros(2 * 1120, np.uint8)
eng.encode_run_host(ed, np.concatenate([np.pad(np.packbits(m[:8448 - 16]), (0, 2)) for m in msgs]), eout)
for i in range(2):
    want = np.packbits(o.rate_match(1, 384, o.encode(1, 384, msgs[i]), 8960, int(ed["rv"][i]), 8, 0, 16))
    assert (eout[i * 1120:(i + 1) * 1120] == want).all()
# soft demodulator
dorc = orc.DemodOracle(o)
sym = rng.uniform(-1.2, 1.2, (700, 2)).astype(np.float32)
nv = rng.uniform(0.01, 0.2, 700).astype(np.float32)
dd = fec.codeword_demod_descs([400, 300], 6, 0x1234567)
sllr = np.zeros(700 * 6, np.int8)
eng.demod_run_host(dd, sym, nv, sllr)
assert (sllr == dorc.codeword(2, 6, [400, 300], sym, nv, 0x1234567)).all()
eng.close()
print("sanitizer case ok:", checked, "checked operations")

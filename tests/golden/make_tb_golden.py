"""Generates tests/golden/tb_golden.npz: whole-transport-block HARQ sequences decoded by the REFERENCE's own software
PUSCH decoder (pusch_decoder_impl + rx buffer pool, compiled from /root/reference into oracle/_ref by oracle/Makefile).

Run in the build container (needs /root/reference):   python tests/golden/make_tb_golden.py
Per transmission the fixture holds the codeword LLRs and what the reference produced: transport block bytes (0xAA where
the reference left them untouched), tb_crc_ok, number of codeblocks, codeblock CRC flags and soft buffers afterwards."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, ".."))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
import cases  # noqa: E402
import __graft_entry__ as ge  # noqa: E402


def main():
    orc = ge.load_oracle()
    assert orc.build_ref(), "needs /root/reference"
    o, ref = orc.Oracle(), orc.Ref()
    rt = orc.RefTb(ref)
    rng = np.random.default_rng(2024)
    out = {}
    n_seq, n_fail = 0, 0
    while n_seq < 10:
        p = cases.random_tb_params(rng, 14000, 6000)
        if n_seq < 6:
            p["snr"] -= 2.5  # make sure retransmissions happen
        g = orc.tb_geometry(p["tbs"], p["bg"], p["qm"], p["nl"], p["G"])
        tb = rng.integers(0, 256, p["tbs"] // 8, dtype=np.uint8)
        cbs = orc.tb_encode(o, tb, p["bg"])
        bid = (n_seq % 8 + 1) << 8
        seq = []
        for t, rv in enumerate([0, 2, 3, 1]):
            llr = orc.tb_transmit(o, cbs, p["tbs"], p["bg"], rv, p["qm"], 0, p["nl"], p["G"], p["snr"], seed=n_seq * 16 + t)
            got = np.full(p["tbs"] // 8, 0xAA, np.uint8)
            r = rt.decode(bid, p["tbs"], p["bg"], rv, p["qm"], 0, p["nl"], t == 0, 6, True, llr, got)
            flags = np.ones(g["C"], bool)
            soft = np.zeros((g["C"], g["N"]), np.int8)
            if not r["tb_crc_ok"]:
                flags, soft = rt.peek(bid, g["C"], g["N"])
                n_fail += 1
            seq.append((rv, llr, got, r["tb_crc_ok"], flags, soft))
            if r["tb_crc_ok"]:
                assert (got == tb).all()
                break
        k = "s%02d" % n_seq
        out[k + "_params"] = np.array([p["tbs"], p["bg"], p["qm"], p["nl"], p["G"], g["C"], len(seq)], np.int64)
        out[k + "_tb"] = tb
        for t, (rv, llr, got, ok, flags, soft) in enumerate(seq):
            out["%s_t%d_rv" % (k, t)] = np.array([rv, int(ok)], np.int64)
            out["%s_t%d_llr" % (k, t)] = llr
            out["%s_t%d_out" % (k, t)] = got
            out["%s_t%d_flags" % (k, t)] = flags
            out["%s_t%d_soft" % (k, t)] = soft
        n_seq += 1
    np.savez_compressed(os.path.join(HERE, "tb_golden.npz"), **out)
    print("sequences", n_seq, "failed transmissions", n_fail, "bytes", os.path.getsize(os.path.join(HERE, "tb_golden.npz")))


if __name__ == "__main__":
    main()

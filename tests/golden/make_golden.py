#!/usr/bin/env python3
"""Generates tests/golden/fec_golden.npz from the REFERENCE ITSELF (oracle/_ref/libref_fec.so = unmodified reference
sources compiled by oracle/Makefile). Run in the build container (needs /root/reference); the .npz is committed.

Every case is a HARQ sequence of 1-3 transmissions of one codeblock through the reference's
ldpc_rate_dematcher_impl + ldpc_decoder_generic + CRC (composed as pusch_codeblock_decoder::decode). Stored per
transmission: input LLRs, soft buffer after dematching, decoded bytes, iteration count (-1 = CRC fail).
Inputs come from oracle/fec_tx.c (encode -> rate match -> AWGN), which tests/test_oracle_vs_ref.py checks against
the reference encoder / rate matcher.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, ".."))
sys.path.insert(0, os.path.join(HERE, "..", "..", "oracle"))
import cases  # noqa: E402
import pyoracle  # noqa: E402


def main():
    o, r = pyoracle.Oracle(), pyoracle.Ref()
    rng = np.random.default_rng(20261104)
    small = [2, 3, 5, 6, 7, 8, 11, 13, 16, 22, 30, 36, 52, 64, 104]
    plan = [(None, None, small)] * 44 + [(1, 384, None), (1, 384, None), (2, 384, None), (1, 352, None), (2, 208, None)]
    meta, blobs = [], {}
    for ci, (bg, Z, pool) in enumerate(plan):
        p = cases.random_chain_params(rng, bg=bg, Z=Z, z_pool=pool)
        K, N = cases.dims(p["bg"], p["Z"])
        p["E"] = min(p["E"], 2 * N) // p["qm"] * p["qm"]
        ntx = int(rng.choice([1, 2, 3]))
        msg = o.make_message(p["bg"], p["Z"], p["fill"], p["poly"], seed=ci)
        cb = o.encode(p["bg"], p["Z"], msg)
        soft = np.zeros(N, np.int8)
        rvs = [p["rv"]] + [int(x) for x in rng.choice([0, 1, 2, 3], ntx - 1)]
        for t, rv in enumerate(rvs):
            tx = o.rate_match(p["bg"], p["Z"], cb, p["E"], rv, p["qm"], p["Nref"], p["fill"])
            llr = o.awgn_llr(tx, p["snr"] - 1.5 * (ntx - 1), seed=1000 * ci + t)
            it, out = r.cb_chain(p["bg"], p["Z"], soft, llr, t == 0, rv, p["qm"], p["Nref"], p["fill"], p["poly"],
                                 p["max_iter"], p["early"], impl=pyoracle.Ref.GENERIC)
            blobs[f"c{ci}_t{t}_llr"] = llr
            blobs[f"c{ci}_t{t}_soft"] = soft.copy()
            blobs[f"c{ci}_t{t}_out"] = out
            meta.append([ci, t, p["bg"], p["Z"], p["E"], rv, p["qm"], p["Nref"], p["fill"], p["poly"], p["max_iter"],
                         int(p["early"]), it])
    blobs["meta"] = np.array(meta, np.int64)
    dst = os.path.join(HERE, "fec_golden.npz")
    np.savez_compressed(dst, **blobs)
    its = blobs["meta"][:, -1]
    print("wrote", dst, os.path.getsize(dst), "bytes;", len(meta), "transmissions; pass", int((its > 0).sum()), "fail",
          int((its < 0).sum()))


if __name__ == "__main__":
    main()

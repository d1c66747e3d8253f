"""Generates tests/golden/enc_golden.npz: LDPC encoding + rate matching vectors produced by the REFERENCE itself
(oracle/_ref/libref_fec.so = the unmodified ldpc_encoder / ldpc_rate_matcher of /root/reference, see oracle/Makefile).
Run in the container that has /root/reference:  python tests/golden/make_enc_golden.py
Each case: packed segment (K - fillers bits), parameters, packed rate-matched output."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge  # noqa: E402
import cases  # noqa: E402


def enc_cases(seed=20261105):
    rng = np.random.default_rng(seed)
    out = []
    # fixed corner cases first: headline shape, config-3 shape, tiny Z, Z not a multiple of 32, limited buffer, repetition
    fixed = [dict(bg=1, Z=384, E=25344, rv=0, qm=2, Nref=0, fill=0), dict(bg=1, Z=384, E=8960, rv=0, qm=8, Nref=0, fill=16),
             dict(bg=2, Z=2, E=40, rv=0, qm=1, Nref=0, fill=0), dict(bg=2, Z=3, E=198, rv=2, qm=6, Nref=0, fill=2),
             dict(bg=1, Z=36, E=3000, rv=3, qm=4, Nref=1500, fill=40), dict(bg=2, Z=208, E=31200, rv=1, qm=8, Nref=0, fill=100),
             dict(bg=1, Z=7, E=462 * 3, rv=2, qm=2, Nref=300, fill=5), dict(bg=2, Z=384, E=3840, rv=0, qm=6, Nref=0, fill=383)]
    for p in fixed:
        out.append(p)
    for _ in range(24):
        p = cases.random_chain_params(rng)
        out.append({k: p[k] for k in ("bg", "Z", "E", "rv", "qm", "Nref", "fill")})
    return out


def main():
    orc = ge.load_oracle()
    assert orc.Ref.available() or orc.build_ref()
    ref = orc.Ref()
    oracle = orc.Oracle()
    data = {}
    meta = []
    for ci, p in enumerate(enc_cases()):
        K, N = cases.dims(p["bg"], p["Z"])
        kp = K - p["fill"]
        msg = np.zeros(K, np.uint8)
        msg[:kp] = oracle.random_bits(kp, 1000 + ci)
        cb = ref.encode(p["bg"], p["Z"], msg)
        tx = ref.rate_match(p["bg"], p["Z"], cb, p["E"], p["rv"], p["qm"], p["Nref"], p["fill"])
        data[f"c{ci}_in"] = np.packbits(msg[:kp])
        data[f"c{ci}_out"] = np.packbits(tx)
        meta.append([ci, p["bg"], p["Z"], p["E"], p["rv"], p["qm"], p["Nref"], p["fill"]])
    data["meta"] = np.array(meta, np.int64)
    np.savez_compressed(os.path.join(HERE, "enc_golden.npz"), **data)
    print("wrote", len(meta), "cases")


if __name__ == "__main__":
    main()

"""CPU tests of the downlink twin: the oracle's encoder + rate matcher against the committed
vectors generated from the reference build (tests/golden/make_enc_golden.py), and the host-only parts of the C ABI."""
import os

import numpy as np
import pytest

import cases

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "enc_golden.npz")


def golden_cases():
    g = np.load(GOLDEN)
    for r in g["meta"]:
        ci, bg, Z, E, rv, qm, Nref, fill = (int(x) for x in r)
        yield dict(ci=ci, bg=bg, Z=Z, E=E, rv=rv, qm=qm, Nref=Nref, fill=fill), g[f"c{ci}_in"], g[f"c{ci}_out"]


def test_oracle_encoder_matches_reference_vectors(oracle):
    n = 0
    for p, seg, want in golden_cases():
        K, N = cases.dims(p["bg"], p["Z"])
        kp = K - p["fill"]
        msg = np.zeros(K, np.uint8)
        msg[:kp] = np.unpackbits(seg)[:kp]
        cb = oracle.encode(p["bg"], p["Z"], msg)
        tx = oracle.rate_match(p["bg"], p["Z"], cb, p["E"], p["rv"], p["qm"], p["Nref"], p["fill"])
        assert (np.packbits(tx) == want).all(), p
        n += 1
    assert n == 32


def test_oracle_encoder_vs_reference_live(oracle, ref):
    """Fresh random cases against the reference build itself (skipped where /root/reference is absent)."""
    rng = np.random.default_rng(77)
    for _ in range(40):
        p = cases.random_chain_params(rng)
        K, N = cases.dims(p["bg"], p["Z"])
        msg = np.zeros(K, np.uint8)
        msg[:K - p["fill"]] = rng.integers(0, 2, K - p["fill"])
        cb = oracle.encode(p["bg"], p["Z"], msg)
        assert (cb == ref.encode(p["bg"], p["Z"], msg)).all(), p
        tx = oracle.rate_match(p["bg"], p["Z"], cb, p["E"], p["rv"], p["qm"], p["Nref"], p["fill"])
        assert (tx == ref.rate_match(p["bg"], p["Z"], cb, p["E"], p["rv"], p["qm"], p["Nref"], p["fill"])).all(), p


def test_enc_desc_prepare_host_only(fec):
    """fec_enc_desc_prepare needs no GPU: k0 / graph_index agree with the decode-side descriptor, bad fields are refused."""
    rng = np.random.default_rng(5)
    for _ in range(200):
        p = cases.random_chain_params(rng)
        e = fec.make_enc_descs(1)
        e["base_graph"], e["lifting_size"], e["rv"], e["qm"] = p["bg"], p["Z"], p["rv"], p["qm"]
        e["Nref"], e["nof_filler_bits"], e["E"] = p["Nref"], p["fill"], p["E"]
        fec.enc_desc_prepare(e)
        d = fec.make_descs(1)
        cases.fill_chain_desc(fec, d, 0, p, True, 0, 0, 0)
        fec.desc_prepare(d)
        assert int(e["k0"][0]) == int(d["k0"][0]) and int(e["graph_index"][0]) == int(d["graph_index"][0])
    for field, bad in (("base_graph", 3), ("lifting_size", 17), ("rv", 4), ("qm", 3), ("E", 0)):
        e = fec.make_enc_descs(1)
        e["base_graph"], e["lifting_size"], e["rv"], e["qm"], e["E"] = 1, 384, 0, 2, 25344
        e[field] = bad
        with pytest.raises(fec.FecError):
            fec.enc_desc_prepare(e)
    assert fec.ENC_DESC_DTYPE.itemsize == 40

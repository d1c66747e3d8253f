"""CPU tests of the checker itself: the C restatement (oracle/) against the committed golden fixtures (generated from the
reference build, tests/golden/make_golden.py), against the reference itself when oracle/_ref is present, and against
the file-free properties the reference's own unit tests assert."""
import os

import numpy as np
import pytest

import cases

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "fec_golden.npz")


@pytest.fixture(scope="module")
def golden():
    return np.load(GOLDEN)


def test_oracle_matches_golden(oracle, golden):
    """Every transmission of every golden HARQ sequence: soft buffer, decoded bytes and iteration count."""
    meta = golden["meta"]
    soft = None
    for row in meta:
        ci, t, bg, Z, E, rv, qm, Nref, fill, poly, max_iter, early, it_ref = (int(x) for x in row)
        K, N = cases.dims(bg, Z)
        if t == 0:
            soft = np.zeros(N, np.int8)
        llr = golden[f"c{ci}_t{t}_llr"]
        it, out = oracle.cb_chain(bg, Z, soft, llr, t == 0, rv, qm, Nref, fill, poly, max_iter, bool(early))
        assert it == it_ref, (ci, t)
        assert (soft == golden[f"c{ci}_t{t}_soft"]).all(), (ci, t)
        assert (out == golden[f"c{ci}_t{t}_out"]).all(), (ci, t)


def _crc_bitwise(poly_full, nbits_crc, bits):
    """In-test bitwise reference, as in the reference's crc_calculator_test.cpp:32-105."""
    r = 0
    top = 1 << nbits_crc
    for b in bits:
        r = (r << 1) | int(b)
        if r & top:
            r ^= poly_full
    for _ in range(nbits_crc):
        r <<= 1
        if r & top:
            r ^= poly_full
    return r


@pytest.mark.parametrize("nbits", [8, 16, 32, 257, 997, 6012])
def test_crc_against_bitwise(oracle, orc, nbits):
    """crc_calculator_test.cpp:225 sizes; CRC(data) == bitwise long division, and CRC(data || crc) == 0."""
    rng = np.random.default_rng(nbits)
    bits = rng.integers(0, 2, nbits).astype(np.uint8)
    for poly, full, ln in ((orc.CRC24A, 0x1864CFB, 24), (orc.CRC24B, 0x1800063, 24), (orc.CRC16, 0x11021, 16)):
        want = _crc_bitwise(full, ln, bits)
        got = oracle.crc(poly, np.packbits(bits), nbits)
        assert got == want
        crc_bits = np.array([(want >> (ln - 1 - i)) & 1 for i in range(ln)], np.uint8)
        assert oracle.crc(poly, np.packbits(np.concatenate([bits, crc_bits])), nbits + ln) == 0


@pytest.mark.parametrize("bg", [1, 2])
def test_noiseless_decodes_in_one_iteration_all_Z(oracle, orc, bg):
    """ldpc_enc_dec_test.cpp:285-315: LLR = +-10, max_iterations = 1, crc = nullptr, several codeblock lengths."""
    for Z in cases.ALL_Z:
        K, N = cases.dims(bg, Z)
        Kb = K // Z
        msg = oracle.random_bits(K, seed=Z * 3 + bg)
        cb = oracle.encode(bg, Z, msg)
        llr_full = (10 - 20 * cb.astype(np.int16)).astype(np.int8)
        for n_cols in sorted({Kb + 2, Kb + 6, (Kb + 2 + N // Z) // 2, N // Z}):
            it, out = oracle.decode(bg, Z, llr_full[: n_cols * Z], 0, -1, 1)
            assert it == -1
            assert (np.unpackbits(out)[:K] == msg).all(), (bg, Z, n_cols)


def test_zero_and_almost_zero_llrs(oracle, orc):
    """ldpc_enc_dec_test.cpp:332-356: all-zero input -> nullopt and all-ones output (crc == nullptr)."""
    bg, Z = 1, 52
    K, N = cases.dims(bg, Z)
    it, out = oracle.decode(bg, Z, np.zeros(N, np.int8), 0, -1, 6)
    assert it == -1 and (np.unpackbits(out)[:K] == 1).all()
    # With a CRC the output is left untouched.
    keep = np.full(K // 8, 0x5A, np.uint8)
    it, out = oracle.decode(bg, Z, np.zeros(N, np.int8), 0, orc.CRC24B, 6, out=keep.copy())
    assert it == -1 and (out == keep).all()
    # Almost zero: only the message part is significant -> decodes (the trimmed length is clamped to (Kb + 4) Z).
    msg = oracle.random_bits(K, seed=9)
    cb = oracle.encode(bg, Z, msg)
    llr = np.zeros(N, np.int8)
    llr[:K - 2 * Z] = (10 - 20 * cb[:K - 2 * Z].astype(np.int16)).astype(np.int8)
    it, out = oracle.decode(bg, Z, llr, 0, -1, 1)
    assert it == -1


RM_ROWS = [(1, q, rv, nref, f) for q in (1, 2, 4, 6, 8) for rv in (0, 1, 2, 3) for nref, f in ((0, 0), (700, 12))][:25]


def test_rate_match_dematch_roundtrip(oracle):
    """ldpc_rm_test.cpp:145-227 without the MATLAB files: match -> dematch(new data, LLR = 1 - 2b) -> hard bits
    (fillers = +127) -> re-match reproduces the matched bits."""
    rng = np.random.default_rng(7)
    for bg in (1, 2):
        for Z in (2, 7, 16, 52):
            K, N = cases.dims(bg, Z)
            Kb = K // Z
            for qm in (1, 2, 4, 6, 8):
                for rv in (0, 1, 2, 3):
                    for lbrm in (False, True):
                        fill = int(rng.integers(0, Z)) if rng.random() < 0.5 else 0
                        Nref = int(rng.integers(K + 2 * Z + 1, N)) if lbrm else 0
                        E = int(rng.integers(K // 2, 2 * N)) // qm * qm
                        msg = oracle.random_bits(K, seed=int(rng.integers(1 << 30)))
                        msg[K - fill:] = 0
                        cb = oracle.encode(bg, Z, msg)
                        tx = oracle.rate_match(bg, Z, cb, E, rv, qm, Nref, fill)
                        soft = np.zeros(N, np.int8)
                        oracle.rate_dematch(bg, Z, soft, (1 - 2 * tx.astype(np.int16)).astype(np.int8), True, rv, qm,
                                            Nref, fill)
                        # Positions that were transmitted must carry the right sign; fillers are +127.
                        sys_end = (Kb - 2) * Z
                        assert (soft[sys_end - fill:sys_end] == 127).all()
                        hard = (soft < 0).astype(np.uint8)
                        known = soft != 0
                        assert (hard[known] == cb[known]).all()
                        cb2 = np.where(known, hard, cb).astype(np.uint8)
                        assert (oracle.rate_match(bg, Z, cb2, E, rv, qm, Nref, fill) == tx).all()


def test_oracle_vs_reference_fuzz(oracle, ref, orc):
    """The restatement against the reference's own generic kernels on seeded random chains (HARQ sequences of 1-3)."""
    rng = np.random.default_rng(123)
    pool = [2, 3, 4, 5, 6, 7, 8, 9, 10, 13, 15, 16, 18, 22, 26, 30, 32, 36, 44, 52, 64, 72, 104, 128, 176]
    for ci in range(250):
        p = cases.random_chain_params(rng, z_pool=pool)
        bg, Z = p["bg"], p["Z"]
        K, N = cases.dims(bg, Z)
        E = min(p["E"], 2 * N) // p["qm"] * p["qm"]
        msg = oracle.make_message(bg, Z, p["fill"], p["poly"], seed=ci)
        cb = oracle.encode(bg, Z, msg)
        assert (cb == ref.encode(bg, Z, msg)).all()
        s1, s2 = np.zeros(N, np.int8), np.zeros(N, np.int8)
        for t in range(int(rng.integers(1, 4))):
            rv = p["rv"] if t == 0 else int(rng.integers(0, 4))
            tx = oracle.rate_match(bg, Z, cb, E, rv, p["qm"], p["Nref"], p["fill"])
            assert (tx == ref.rate_match(bg, Z, cb, E, rv, p["qm"], p["Nref"], p["fill"])).all()
            llr = oracle.awgn_llr(tx, p["snr"] - 2.0, seed=ci * 10 + t)
            a = oracle.cb_chain(bg, Z, s1, llr, t == 0, rv, p["qm"], p["Nref"], p["fill"], p["poly"], p["max_iter"],
                                p["early"])
            b = ref.cb_chain(bg, Z, s2, llr, t == 0, rv, p["qm"], p["Nref"], p["fill"], p["poly"], p["max_iter"],
                             p["early"])
            assert a[0] == b[0] and (a[1] == b[1]).all() and (s1 == s2).all(), (ci, t, p)


def test_oracle_vs_reference_big(oracle, ref, orc):
    """BG1 Z=384 (the headline shape) across the waterfall, decode-only and chain."""
    bg, Z = 1, 384
    K, N = cases.dims(bg, Z)
    for i, snr in enumerate([-3.0, -2.0, -1.5, -1.0, 0.0, 2.0]):
        llr, _ = oracle.make_case(bg, Z, N, 0, 2, 0, 0, orc.CRC24B, snr, seed=40 + i)
        for crc, mi in ((orc.CRC24B, 10), (-1, 6)):
            a = oracle.decode(bg, Z, llr, 0, crc, mi)
            b = ref.decode(bg, Z, llr, 0, crc, mi)
            assert a[0] == b[0] and (a[1] == b[1]).all(), (snr, crc)

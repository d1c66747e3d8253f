"""GPU parity tests of the downlink twin: LDPC encoding + rate matching on the device, through the C ABI
(fec_encode_run_host) and the hal::hw_accelerator_pdsch_enc mirror, bit-exact against the oracle and against the vectors
generated from the reference build."""
import numpy as np
import pytest

import cases
from test_encode_cpu import golden_cases

pytestmark = pytest.mark.gpu


def oracle_tx(oracle, p, msg):
    cb = oracle.encode(p["bg"], p["Z"], msg)
    return oracle.rate_match(p["bg"], p["Z"], cb, p["E"], p["rv"], p["qm"], p["Nref"], p["fill"])


def run_encode_batch(fec, eng, items, unpacked=False, misalign=0):
    """items: list of (params, packed segment). One launch; returns the list of outputs (packed bytes or unpacked bits)."""
    n = len(items)
    d = fec.make_enc_descs(n)
    ins, in_off, out_off, outs = [], 0, misalign, []
    for i, (p, seg) in enumerate(items):
        d[i]["base_graph"], d[i]["lifting_size"], d[i]["rv"], d[i]["qm"] = p["bg"], p["Z"], p["rv"], p["qm"]
        d[i]["Nref"], d[i]["nof_filler_bits"], d[i]["E"] = p["Nref"], p["fill"], p["E"]
        d[i]["flags"] = fec.ENC_FLAG_UNPACKED if unpacked else 0
        d[i]["in_offset"], d[i]["out_offset"] = in_off, out_off
        ins.append(seg)
        in_off += len(seg)
        olen = p["E"] if unpacked else (p["E"] + 7) // 8
        outs.append((out_off, olen))
        out_off += olen if misalign else (olen + 3) & ~3
    out = np.full(out_off + 8, 0xAA, np.uint8)
    eng.encode_run_host(d, np.concatenate(ins), out)
    return [out[o:o + ln].copy() for o, ln in outs], out


def test_encode_golden_vectors(fec, engine):
    items, want = [], []
    for p, seg, w in golden_cases():
        items.append((p, seg))
        want.append(w)
    got, _ = run_encode_batch(fec, engine, items)
    for (p, _), g, w in zip(items, got, want):
        assert (g == w).all(), p
    got, _ = run_encode_batch(fec, engine, items, unpacked=True)
    for (p, _), g, w in zip(items, got, want):
        assert (g == np.unpackbits(w)[:p["E"]]).all(), p


@pytest.mark.parametrize("bg", [1, 2])
def test_encode_sweep_all_lifting_sizes(fec, engine, oracle, bg):
    """Every lifting size, random rate / Qm / rv / limited buffer / fillers, mixed in one launch (BASELINE config 2 spirit)."""
    rng = np.random.default_rng(100 + bg)
    items, want = [], []
    for Z in cases.ALL_Z:
        for _ in range(3):
            p = cases.random_chain_params(rng, bg=bg, Z=Z)
            K, _ = cases.dims(bg, Z)
            kp = K - p["fill"]
            msg = np.zeros(K, np.uint8)
            msg[:kp] = rng.integers(0, 2, kp)
            items.append((p, np.packbits(msg[:kp])))
            want.append(oracle_tx(oracle, p, msg))
    for unpacked, misalign in ((False, 0), (True, 0), (False, 1), (True, 3)):
        got, _ = run_encode_batch(fec, engine, items, unpacked=unpacked, misalign=misalign)
        for (p, _), g, w in zip(items, got, want):
            ref_out = w if unpacked else np.packbits(w)
            assert g.size == ref_out.size and (g == ref_out).all(), (p, unpacked, misalign)


def test_encode_does_not_write_outside_its_output(fec, engine, oracle):
    rng = np.random.default_rng(9)
    p = dict(bg=1, Z=384, E=25344 - 6, rv=0, qm=2, Nref=0, fill=0)  # E not a multiple of 8 or 32
    K, _ = cases.dims(1, 384)
    msg = rng.integers(0, 2, K).astype(np.uint8)
    for unpacked in (False, True):
        d = fec.make_enc_descs(2)
        ln = p["E"] if unpacked else (p["E"] + 7) // 8
        for i in range(2):
            d[i]["base_graph"], d[i]["lifting_size"], d[i]["rv"], d[i]["qm"], d[i]["E"] = 1, 384, 0, 2, p["E"]
            d[i]["flags"] = fec.ENC_FLAG_UNPACKED if unpacked else 0
            d[i]["out_offset"] = 5 + i * (ln + 11)  # unaligned, 11 bytes between the two outputs
        raw = np.full(5 + 2 * ln + 11 + 9, 0xAA, np.uint8)
        engine.encode_run_host(d, np.packbits(msg), raw)
        w = oracle_tx(oracle, p, msg)
        want = w if unpacked else np.packbits(w)
        assert (raw[:5] == 0xAA).all() and (raw[5 + 2 * ln + 11:] == 0xAA).all()  # outside the span: untouched
        assert (raw[5 + ln:5 + ln + 11] == 0).all()                              # between the outputs: zeros
        assert (raw[5:5 + ln] == want).all() and (raw[5 + ln + 11:5 + 2 * ln + 11] == want).all()


def test_hw_accelerator_pdsch_enc_call_sequence(fec, engine, oracle):
    """The call sequence pdsch_encoder_hw_impl drives in CB mode (pdsch_encoder_hw_impl.cpp:31-186): all codeblocks of a
    transport block are configured + enqueued, then dequeued in order into the codeword."""
    rng = np.random.default_rng(31)
    acc = fec.HwAcceleratorPdschEnc(engine)
    assert acc.is_cb_mode_supported() and acc.get_max_supported_buff_size() == 1056
    acc.reserve_queue()
    nof_cbs, Z, bg, qm = 7, 384, 1, 8
    K, N = cases.dims(bg, Z)
    rm = [8960] * 4 + [8992] * 3
    msgs = []
    for cb in range(nof_cbs):
        fill = 16
        msg = np.zeros(K, np.uint8)
        msg[:K - fill] = rng.integers(0, 2, K - fill)
        msgs.append(msg)
        cfg = fec.HwPdschEncoderConfiguration(base_graph_index=bg, modulation=qm, rv=0, lifting_size=Z, Ncb=N, Nref=0,
                                              nof_segment_bits=K, nof_filler_bits=fill, rm_length=rm[cb])
        acc.configure_operation(cfg, cb)
        assert acc.enqueue_operation(np.packbits(msg[:K - fill]), None, cb)
    codeword = np.zeros(sum(rm), np.uint8)
    off = 0
    for cb in range(nof_cbs):
        packed = np.zeros((rm[cb] + 7) // 8, np.uint8)
        assert acc.dequeue_operation(codeword[off:off + rm[cb]], packed, cb)
        p = dict(bg=bg, Z=Z, E=rm[cb], rv=0, qm=qm, Nref=0, fill=16)
        w = oracle_tx(oracle, p, msgs[cb])
        assert (codeword[off:off + rm[cb]] == w).all() and (packed == np.packbits(w)).all()
        off += rm[cb]
    assert not acc.dequeue_operation(codeword, None, 99)
    acc.free_queue()


def test_encoder_rate_matcher_mirror(fec, engine, oracle):
    rng = np.random.default_rng(3)
    enc = fec.LdpcEncoderRateMatcher(engine)
    for _ in range(10):
        p = cases.random_chain_params(rng)
        K, _ = cases.dims(p["bg"], p["Z"])
        msg = np.zeros(K, np.uint8)
        msg[:K - p["fill"]] = rng.integers(0, 2, K - p["fill"])
        cfg = fec.CodeblockMetadata(base_graph=p["bg"], lifting_size=p["Z"], rv=p["rv"], mod_order=p["qm"], Nref=p["Nref"],
                                    nof_filler_bits=p["fill"])
        assert (enc.encode_rate_match(msg, p["E"], cfg) == oracle_tx(oracle, p, msg)).all(), p


def test_full_size_encode_decode_round_trip(fec, engine, oracle):
    """BASELINE-size property: 1024 codeblocks of the headline shape encoded on the device, sent through a noiseless BPSK
    channel and decoded on the device give back every message (encode -> decode round trip), and the first codeblocks
    agree with the oracle bit for bit."""
    rng = np.random.default_rng(12)
    n, bg, Z, E = 1024, 1, 384, 25344
    K, _ = cases.dims(bg, Z)
    msgs = np.zeros((n, K), np.uint8)
    for i in range(n):
        msgs[i] = oracle.make_message(bg, Z, 0, 1, 5000 + i)
    p = dict(bg=bg, Z=Z, E=E, rv=0, qm=2, Nref=0, fill=0)
    got, _ = run_encode_batch(fec, engine, [(p, np.packbits(m)) for m in msgs], unpacked=True)
    for i in range(4):
        assert (got[i] == oracle_tx(oracle, p, msgs[i])).all()
    d = fec.make_descs(n)
    llr = np.zeros(n * E, np.int8)
    for i in range(n):
        pp = dict(p, poly=1, max_iter=6, early=True)
        cases.fill_chain_desc(fec, d, i, pp, True, i, i * E, i * fec.MAX_CB_OUT_BYTES)
        engine.harq_free(i)
        llr[i * E:(i + 1) * E] = 20 - 40 * got[i].astype(np.int8)
    out = np.zeros((n, fec.MAX_CB_OUT_BYTES), np.uint8)
    res = engine.run_host(d, llr, out)
    assert (res["status"] & fec.RES_CRC_PASS).all()
    assert (out[:, :K // 8] == np.packbits(msgs, axis=1)).all()


def test_encode_large_batch_is_pipelined(fec, engine, oracle):
    """Batches of two kernel waves or more go through the chunked copy / kernel / copy-back pipeline of
    fec_encode_run_host: 2600 small codeblocks with mixed parameters and ragged output offsets, every one against the
    oracle, gaps between outputs read as zeros and the bytes around the span stay untouched."""
    rng = np.random.default_rng(77)
    items, want = [], []
    for i in range(2600):
        bg = int(rng.integers(1, 3))
        Z = int(rng.choice([2, 3, 5, 8, 11, 16, 20, 26, 36]))
        p = cases.random_chain_params(rng, bg=bg, Z=Z)
        K, N = cases.dims(bg, Z)
        p["E"] = max(p["qm"], min(p["E"], N + 40) // p["qm"] * p["qm"])
        kp = K - p["fill"]
        msg = np.zeros(K, np.uint8)
        msg[:kp] = rng.integers(0, 2, kp)
        items.append((p, np.packbits(msg[:kp])))
        want.append(np.packbits(oracle_tx(oracle, p, msg)))
    got, out = run_encode_batch(fec, engine, items)
    for (p, _), g, w in zip(items, got, want):
        assert (g == w).all(), p
    assert (out[-8:] == 0xAA).all()

"""oracle/pyoracle.py -- TEST INFRASTRUCTURE. ctypes bindings to the two CPU checkers:

  * `Oracle`  : oracle/_build/libfec_oracle.so, the plain-C restatement (fec_oracle.c / fec_tx.c); always buildable.
  * `Ref`     : oracle/_ref/libref_fec.so, the UNMODIFIED reference sources + ref_shim.cpp; built only where
                /root/reference exists (the build container) and shipped to the GPU box as a prebuilt file.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "_build", "libfec_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libref_fec.so")

ALL_Z = [2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 18, 20, 22, 24, 26, 28, 30, 32, 36, 40, 44, 48, 52, 56, 60,
         64, 72, 80, 88, 96, 104, 112, 120, 128, 144, 160, 176, 192, 208, 224, 240, 256, 288, 320, 352, 384]
CRC24A, CRC24B, CRC16 = 0, 1, 2
CRC_LEN = {CRC24A: 24, CRC24B: 24, CRC16: 16}


def build_oracle():
    subprocess.check_call(["make", "-s", "-C", HERE, "oracle"])


def build_ref():
    """Builds oracle/_ref when the reference tree is present; returns True if the .so exists afterwards."""
    if os.path.isdir("/root/reference/lib"):
        subprocess.check_call(["make", "-s", "-j8", "-C", HERE, "ref"])
    return os.path.exists(REF_SO)


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def dims(bg, Z):
    kb, nb = (22, 66) if bg == 1 else (10, 50)
    return kb * Z, nb * Z


class Oracle:
    """The C restatement (kind "port")."""

    def __init__(self):
        if not os.path.exists(ORACLE_SO):
            build_oracle()
        self.lib = C.CDLL(ORACLE_SO)
        L = self.lib
        L.orc_crc.restype = C.c_uint32
        L.orc_quantize.restype = C.c_int8
        L.orc_quantize.argtypes = [C.c_float, C.c_float]

    def crc(self, poly, packed, nbits):
        packed = np.ascontiguousarray(packed, np.uint8)
        return int(self.lib.orc_crc(poly, _p(packed), C.c_uint32(nbits)))

    def rate_dematch(self, bg, Z, soft, llr, new_data, rv, qm, Nref, fillers):
        assert soft.dtype == np.int8 and soft.flags.c_contiguous and soft.size == dims(bg, Z)[1]
        llr = np.ascontiguousarray(llr, np.int8)
        r = self.lib.orc_rate_dematch(bg, Z, _p(soft), _p(llr), C.c_uint32(llr.size), int(new_data), C.c_uint32(rv),
                                      C.c_uint32(qm), C.c_uint32(Nref), C.c_uint32(fillers))
        assert r == 0
        return soft

    def decode(self, bg, Z, llr, fillers=0, crc_poly=-1, max_iter=6, scaling=0.8, out=None):
        K, _ = dims(bg, Z)
        llr = np.ascontiguousarray(llr, np.int8)
        if out is None:
            out = np.zeros((K + 7) // 8, np.uint8)
        it = self.lib.orc_decode(bg, Z, _p(out), _p(llr), C.c_uint32(llr.size), C.c_uint32(fillers), crc_poly,
                                 C.c_uint32(max_iter), C.c_float(scaling))
        return it, out

    def cb_chain(self, bg, Z, soft, llr, new_data, rv, qm, Nref, fillers, crc_poly, max_iter, early_stop):
        K, _ = dims(bg, Z)
        llr = np.ascontiguousarray(llr, np.int8)
        out = np.zeros((K + 7) // 8, np.uint8)
        it = self.lib.orc_cb_chain(bg, Z, _p(soft), _p(llr), C.c_uint32(llr.size), int(new_data), C.c_uint32(rv),
                                   C.c_uint32(qm), C.c_uint32(Nref), C.c_uint32(fillers), crc_poly,
                                   C.c_uint32(max_iter), int(early_stop), _p(out))
        return it, out

    # ---- transmit side (input generation) ----
    def random_bits(self, n, seed):
        b = np.zeros(n, np.uint8)
        self.lib.orc_random_bits(_p(b), C.c_uint32(n), C.c_uint64(seed))
        return b

    def make_message(self, bg, Z, fillers, crc_poly, seed):
        """Random payload + CB CRC + zero fillers -> K unpacked bits."""
        K, _ = dims(bg, Z)
        n_pay = K - fillers - CRC_LEN[crc_poly]
        msg = np.zeros(K + 32, np.uint8)
        msg[:n_pay] = self.random_bits(n_pay, seed)
        self.lib.orc_crc_attach_bits(crc_poly, _p(msg), C.c_uint32(n_pay))
        return msg[:K].copy()

    def encode(self, bg, Z, msg):
        _, N = dims(bg, Z)
        msg = np.ascontiguousarray(msg, np.uint8)
        cb = np.zeros(N, np.uint8)
        r = self.lib.orc_encode(bg, Z, _p(msg), _p(cb))
        assert r == 0, r
        return cb

    def rate_match(self, bg, Z, cb, E, rv, qm, Nref, fillers):
        cb = np.ascontiguousarray(cb, np.uint8)
        out = np.zeros(E, np.uint8)
        r = self.lib.orc_rate_match(bg, Z, _p(cb), _p(out), C.c_uint32(E), C.c_uint32(rv), C.c_uint32(qm),
                                    C.c_uint32(Nref), C.c_uint32(fillers))
        assert r == 0
        return out

    def awgn_llr(self, bits, esn0_db, seed):
        bits = np.ascontiguousarray(bits, np.uint8)
        llr = np.zeros(bits.size, np.int8)
        self.lib.orc_awgn_llr(_p(bits), _p(llr), C.c_uint32(bits.size), C.c_float(esn0_db), C.c_uint64(seed))
        return llr

    def make_case(self, bg, Z, E, rv=0, qm=1, Nref=0, fillers=0, crc_poly=CRC24B, esn0_db=2.0, seed=0):
        """message -> encode -> rate match -> AWGN LLRs. Returns (llr[E], msg_bits[K])."""
        msg = self.make_message(bg, Z, fillers, crc_poly, seed)
        cb = self.encode(bg, Z, msg)
        tx = self.rate_match(bg, Z, cb, E, rv, qm, Nref, fillers)
        return self.awgn_llr(tx, esn0_db, seed + 7919), msg


# ---- soft demodulation + descrambling ------------------------------------------------------------
DEMOD_FLAVOURS = ("generic", "avx2", "avx512")  # what the reference was compiled for; index = flavour id


def _demod_args(sym, noise):
    sym = np.ascontiguousarray(sym, np.float32).reshape(-1, 2)
    noise = np.ascontiguousarray(noise, np.float32).reshape(-1)
    assert sym.shape[0] == noise.size
    return sym, noise


class DemodOracle:
    """fec_demod.c: the C restatement of demodulate_soft (three compile-time flavours), the Gold sequence and descrambling."""

    def __init__(self, oracle):
        self.lib = oracle.lib
        self.lib.orc_demodulate_soft.argtypes = [C.c_int, C.c_int, C.c_int, C.c_uint, C.c_void_p, C.c_void_p, C.c_void_p]
        self.lib.orc_gold_sequence.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p]
        self.lib.orc_descramble.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p]

    def demodulate_soft(self, flavour, qm, sym, noise, pi2=False):
        sym, noise = _demod_args(sym, noise)
        out = np.zeros(noise.size * qm, np.int8)
        self.lib.orc_demodulate_soft(flavour, qm, int(pi2), noise.size, _p(sym), _p(noise), _p(out))
        return out

    def blocks(self, flavour, qm):
        a, b = C.c_uint(), C.c_uint()
        self.lib.orc_demod_blocks(flavour, qm, C.byref(a), C.byref(b))
        return a.value, b.value

    def tables(self):
        t = np.zeros((7, 36), np.float32)
        self.lib.orc_demod_tables(_p(t))
        return t

    def gold_sequence(self, c_init, offset, nbits):
        packed = np.zeros((nbits + 7) // 8, np.uint8)
        self.lib.orc_gold_sequence(c_init, offset, nbits, _p(packed))
        return packed

    def descramble(self, packed_seq, llr):
        llr = np.ascontiguousarray(llr, np.int8)
        out = np.zeros_like(llr)
        self.lib.orc_descramble(_p(packed_seq), llr.size, _p(llr), _p(out))
        return out

    def codeword(self, flavour, qm, blocks, sym, noise, c_init):
        """pusch_demodulator_impl.cpp:262-400 without the equaliser: one demodulate_soft call per block of symbols (one
        OFDM symbol each), the scrambling sequence running on across the blocks."""
        sym, noise = _demod_args(sym, noise)
        out, pos = [], 0
        for n in blocks:
            out.append(self.demodulate_soft(flavour, qm, sym[pos:pos + n], noise[pos:pos + n]))
            pos += n
        llr = np.concatenate(out) if out else np.zeros(0, np.int8)
        return self.descramble(self.gold_sequence(c_init, 0, llr.size), llr)


class RefDemod:
    """The reference's demodulation_mapper_impl / pseudo_random_generator_impl compiled for one flavour."""

    def __init__(self, flavour):
        path = os.path.join(HERE, "_ref", "libref_demod_%s.so" % DEMOD_FLAVOURS[flavour])
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        self.lib = C.CDLL(path)
        self.lib.ref_demodulate_soft.argtypes = [C.c_int, C.c_int, C.c_uint, C.c_void_p, C.c_void_p, C.c_void_p]
        self.lib.ref_prg_generate.argtypes = [C.c_uint, C.c_uint, C.c_uint, C.c_void_p]
        self.lib.ref_prg_apply_xor_llr.argtypes = [C.c_uint, C.c_uint, C.c_uint, C.c_void_p, C.c_void_p]

    @staticmethod
    def available(flavour):
        return os.path.exists(os.path.join(HERE, "_ref", "libref_demod_%s.so" % DEMOD_FLAVOURS[flavour]))

    def demodulate_soft(self, qm, sym, noise, pi2=False):
        sym, noise = _demod_args(sym, noise)
        out = np.zeros(noise.size * qm, np.int8)
        self.lib.ref_demodulate_soft(qm, int(pi2), noise.size, _p(sym), _p(noise), _p(out))
        return out

    def gold_sequence(self, c_init, offset, nbits):
        packed = np.zeros((nbits + 7) // 8, np.uint8)
        self.lib.ref_prg_generate(c_init, offset, nbits, _p(packed))
        return packed

    def apply_xor_llr(self, c_init, offset, llr):
        llr = np.ascontiguousarray(llr, np.int8)
        out = np.zeros_like(llr)
        self.lib.ref_prg_apply_xor_llr(c_init, offset, llr.size, _p(llr), _p(out))
        return out


class Ref:
    """The reference itself (kind "reference"). impl: 0 generic, 1 avx2, 2 avx512, 3 auto."""
    GENERIC, AVX2, AVX512, AUTO = 0, 1, 2, 3

    def __init__(self):
        if not os.path.exists(REF_SO):
            raise FileNotFoundError(REF_SO)
        self.lib = C.CDLL(REF_SO)
        self.lib.ref_crc.restype = C.c_uint
        self.lib.ref_batch_chain.restype = C.c_double
        self.lib.ref_tbs.restype = C.c_uint
        self.lib.ref_tbs.argtypes = [C.c_uint, C.c_uint, C.c_uint, C.c_uint, C.c_float, C.c_uint, C.c_uint]

    @staticmethod
    def available():
        return os.path.exists(REF_SO)

    def impl_available(self, impl):
        return bool(self.lib.ref_impl_available(impl))

    def graph(self, bg, Z):
        cols = np.zeros((46, 20), np.uint16)
        sh = np.zeros((46, 20), np.uint16)
        deg = np.zeros(46, np.uint16)
        M = self.lib.ref_graph(bg, Z, _p(cols), _p(sh), _p(deg))
        return M, cols, sh, deg

    def crc(self, poly, packed, nbits):
        packed = np.ascontiguousarray(packed, np.uint8)
        return int(self.lib.ref_crc(poly, _p(packed), C.c_uint(nbits)))

    def encode(self, bg, Z, msg):
        _, N = dims(bg, Z)
        msg = np.ascontiguousarray(msg, np.uint8)
        cb = np.zeros(N, np.uint8)
        self.lib.ref_encode(bg, Z, _p(msg), _p(cb))
        return cb

    def rate_match(self, bg, Z, cb, E, rv, qm, Nref, fillers):
        cb = np.ascontiguousarray(cb, np.uint8)
        out = np.zeros(E, np.uint8)
        self.lib.ref_rate_match(bg, Z, _p(cb), _p(out), C.c_uint(E), C.c_uint(rv), C.c_uint(qm), C.c_uint(Nref),
                                C.c_uint(fillers))
        return out

    def rate_dematch(self, bg, Z, soft, llr, new_data, rv, qm, Nref, fillers, impl=0):
        llr = np.ascontiguousarray(llr, np.int8)
        assert soft.dtype == np.int8 and soft.flags.c_contiguous
        r = self.lib.ref_rate_dematch(impl, bg, Z, _p(soft), _p(llr), C.c_uint(llr.size), int(new_data), C.c_uint(rv),
                                      C.c_uint(qm), C.c_uint(Nref), C.c_uint(fillers))
        assert r == 0
        return soft

    def decode(self, bg, Z, llr, fillers=0, crc_poly=-1, max_iter=6, scaling=0.8, out=None, impl=0):
        K, _ = dims(bg, Z)
        llr = np.ascontiguousarray(llr, np.int8)
        if out is None:
            out = np.zeros((K + 7) // 8, np.uint8)
        it = self.lib.ref_decode(impl, bg, Z, _p(out), _p(llr), C.c_uint(llr.size), C.c_uint(fillers), crc_poly,
                                 C.c_uint(max_iter), C.c_float(scaling))
        return it, out

    def cb_chain(self, bg, Z, soft, llr, new_data, rv, qm, Nref, fillers, crc_poly, max_iter, early_stop, impl=0):
        K, _ = dims(bg, Z)
        llr = np.ascontiguousarray(llr, np.int8)
        out = np.zeros((K + 7) // 8, np.uint8)
        it = self.lib.ref_cb_chain(impl, bg, Z, _p(soft), _p(llr), C.c_uint(llr.size), int(new_data), C.c_uint(rv),
                                   C.c_uint(qm), C.c_uint(Nref), C.c_uint(fillers), crc_poly, C.c_uint(max_iter),
                                   int(early_stop), _p(out))
        return it, out

    def batch_chain(self, impl, n_threads, bg, Z, llr, new_data, rv, qm, Nref, fillers, crc_poly, max_iter,
                    early_stop):
        """llr: [n, E] int8. Returns (seconds, iters[n], out[n, K/8])."""
        K, N = dims(bg, Z)
        llr = np.ascontiguousarray(llr, np.int8)
        n, E = llr.shape
        soft = np.zeros((n, N), np.int8)
        out = np.zeros((n, (K + 7) // 8), np.uint8)
        iters = np.zeros(n, np.int32)
        sec = self.lib.ref_batch_chain(impl, n_threads, C.c_uint(n), bg, Z, _p(soft), _p(llr), C.c_uint(E),
                                       int(new_data), C.c_uint(rv), C.c_uint(qm), C.c_uint(Nref), C.c_uint(fillers),
                                       crc_poly, C.c_uint(max_iter), int(early_stop), _p(out), _p(iters))
        return sec, iters, out

    def tbs(self, nof_symb, nof_dmrs_prb, nof_oh_prb, qm, rate_x1024, nof_layers, n_prb):
        return int(self.lib.ref_tbs(nof_symb, nof_dmrs_prb, nof_oh_prb, qm, float(rate_x1024), nof_layers, n_prb))

    def segment(self, tbs, bg, rv, qm, Nref, nof_layers, nof_ch_symbols):
        rm = np.zeros(162, np.uint32)
        off = np.zeros(162, np.uint32)
        sc = np.zeros(8, np.uint32)
        n = self.lib.ref_segment(C.c_uint(tbs), bg, C.c_uint(rv), C.c_uint(qm), C.c_uint(Nref), C.c_uint(nof_layers),
                                 C.c_uint(nof_ch_symbols), _p(rm), _p(off), _p(sc))
        return dict(n=n, rm_len=rm[:n].copy(), cw_off=off[:n].copy(), bg=int(sc[0]), Z=int(sc[1]), fillers=int(sc[2]),
                    N=int(sc[3]), crc_bits=int(sc[4]))


# ---- transport-block layer -----------------------------------------------------------------------------------------
def tb_geometry(tbs, bg, qm, nof_layers, nof_llrs):
    """Receive-side segmentation of a transport block: restatement of ldpc::compute_nof_codeblocks / compute_lifting_size
    (include/srsran/phy/upper/channel_coding/ldpc/ldpc.h:128-208), ldpc_segmenter_rx_impl::segment
    (lib/phy/upper/channel_coding/ldpc/ldpc_segmenter_rx_impl.cpp:53-106) and compute_rm_length
    (ldpc_segmenter_helpers.h:82-94). Returns a dict."""
    tb_crc = 16 if tbs <= 3824 else 24
    B = tbs + tb_crc
    max_seg = 8448 if bg == 1 else 3840
    C_ = 1 if B <= max_seg else -(-B // (max_seg - 24))
    B_out = B + (24 * C_ if C_ > 1 else 0)
    if bg == 1:
        ref = 22
    else:
        ref = 10 if B > 640 else 9 if B > 560 else 8 if B > 192 else 6
    Z = next(z for z in ALL_Z if z * C_ * ref >= B_out)
    K, N = dims(bg, Z)
    msg_bits = -(-B_out // C_)
    sym = nof_llrs // qm // nof_layers
    n_short = C_ - sym % C_
    rm = [((sym // C_) if r < n_short else -(-sym // C_)) * nof_layers * qm for r in range(C_)]
    data_bits = msg_bits - (tb_crc if C_ == 1 else 24)
    stream_bits = tbs + (24 if C_ > 1 else 0)
    payload, left = [], stream_bits
    for r in range(C_):
        payload.append(min(left, data_bits))
        left -= payload[-1]
    return dict(C=C_, Z=Z, K=K, N=N, fillers=K - msg_bits, rm=rm, cw_off=list(np.cumsum([0] + rm[:-1])),
                crc=(CRC24B if C_ > 1 else (CRC24A if tbs > 3824 else CRC16)), data_bits=data_bits,
                stream_bits=stream_bits, payload=payload, tb_crc=tb_crc)


def _bits_of(packed, nbits):
    return np.unpackbits(np.ascontiguousarray(packed, np.uint8))[:nbits]


def tb_encode(oracle, tb_bytes, bg):
    """Transmit side of a transport block (input generation for the tests): TB CRC attachment, segmentation with
    codeblock CRCs (TS 38.212 5.2.2 as ldpc_segmenter_tx does it, zero padding in the last codeblock), LDPC encoding.
    Returns the list of encoded codeblocks (unpacked bits, N each)."""
    tbs = tb_bytes.size * 8
    g = tb_geometry(tbs, bg, 1, 1, 1)
    tb_crc_poly = CRC24A if tbs > 3824 else CRC16
    stream = np.zeros(tbs + 32, np.uint8)
    stream[:tbs] = np.unpackbits(tb_bytes)
    oracle.lib.orc_crc_attach_bits(tb_crc_poly, _p(stream), C.c_uint32(tbs))
    stream = stream[:tbs + g["tb_crc"]]
    cbs, pos = [], 0
    per = g["data_bits"] + (g["tb_crc"] if g["C"] == 1 else 0)  # bits of the TB+CRC stream per codeblock
    for r in range(g["C"]):
        msg = np.zeros(g["K"] + 32, np.uint8)
        take = min(per, stream.size - pos)
        msg[:take] = stream[pos:pos + take]
        pos += take
        if g["C"] > 1:
            oracle.lib.orc_crc_attach_bits(CRC24B, _p(msg), C.c_uint32(per))
        cbs.append(oracle.encode(bg, g["Z"], msg[:g["K"]].copy()))
    assert pos == stream.size
    return cbs


def tb_transmit(oracle, cbs, tbs, bg, rv, qm, Nref, nof_layers, nof_llrs, esn0_db, seed):
    """Rate matching of every codeblock + concatenation + AWGN: the codeword LLRs of one (re)transmission."""
    g = tb_geometry(tbs, bg, qm, nof_layers, nof_llrs)
    tx = np.concatenate([oracle.rate_match(bg, g["Z"], cbs[r], g["rm"][r], rv, qm, Nref, g["fillers"])
                         for r in range(g["C"])])
    return oracle.awgn_llr(tx, esn0_db, seed)


def modulate(bits, qm):
    """TS 38.211 5.1.2-5.1.6 modulation mapper (input generation for the tests): bits -> complex64 symbols."""
    b = np.asarray(bits, np.int64).reshape(-1, qm)
    s = 1 - 2 * b
    if qm == 1:
        return ((s[:, 0] + 1j * s[:, 0]) / np.sqrt(2)).astype(np.complex64)
    m = qm // 2
    re, im = s[:, qm - 2], s[:, qm - 1]
    for i in range(m - 2, -1, -1):
        re = s[:, 2 * i] * ((1 << (m - 1 - i)) - re)
        im = s[:, 2 * i + 1] * ((1 << (m - 1 - i)) - im)
    norm = {1: 2.0, 2: 2.0, 4: 10.0, 6: 42.0, 8: 170.0}[qm]
    return ((re + 1j * im) / np.sqrt(norm)).astype(np.complex64)


def tb_transmit_symbols(oracle, demod, cbs, tbs, bg, rv, qm, Nref, nof_layers, nof_llrs, esn0_db, c_init, seed):
    """Rate matching + concatenation + scrambling + modulation + complex AWGN: the equalised symbols (float32 re, im) and
    the noise variance of each symbol for one (re)transmission."""
    g = tb_geometry(tbs, bg, qm, nof_layers, nof_llrs)
    tx = np.concatenate([oracle.rate_match(bg, g["Z"], cbs[r], g["rm"][r], rv, qm, Nref, g["fillers"])
                         for r in range(g["C"])])
    seq = np.unpackbits(demod.gold_sequence(c_init, 0, tx.size))[:tx.size]
    x = modulate(tx ^ seq, qm)
    rng = np.random.default_rng(seed)
    nv = 10.0 ** (-esn0_db / 10.0)
    noise = (rng.normal(0, np.sqrt(nv / 2), x.size) + 1j * rng.normal(0, np.sqrt(nv / 2), x.size)).astype(np.complex64)
    # per-symbol noise variance estimates scatter around the truth, as a channel estimator's would
    nvs = (nv * rng.uniform(0.8, 1.25, x.size)).astype(np.float32)
    return np.ascontiguousarray((x + noise).astype(np.complex64)).view(np.float32).reshape(-1, 2), nvs


class TbOracle:
    """Whole transport blocks over the C restatement: the flow of pusch_decoder_impl (new_data :88-130,
    fork_codeblock_task :308-372, join_and_notify :374-437, concatenate_codeblocks :439-493) with the rx buffer (soft bits,
    codeblock CRC flags, decoded messages) kept per buffer id. Test infrastructure."""

    def __init__(self, oracle):
        self.o = oracle
        self.buffers = {}

    def decode(self, buffer_id, tbs, bg, rv, qm, Nref, nof_layers, new_data, max_iter, early_stop, llrs, tb_out):
        """Returns dict(tb_crc_ok, nof_codeblocks, iters=[per decoded CB], delivered). tb_out is written as the reference
        writes transport_block."""
        g = tb_geometry(tbs, bg, qm, nof_layers, len(llrs))
        C_, Z, K, N = g["C"], g["Z"], g["K"], g["N"]
        buf = self.buffers.get(buffer_id)
        if buf is None or new_data:
            if buf is None:
                buf = dict(soft=np.zeros((C_, N), np.int8), msg=np.zeros((C_, (K + 7) // 8), np.uint8))
            buf["crc"] = np.zeros(C_, bool)
            self.buffers[buffer_id] = buf
        iters = []
        for r in range(C_):
            cb = np.ascontiguousarray(llrs[g["cw_off"][r]:g["cw_off"][r] + g["rm"][r]], np.int8)
            if buf["crc"][r]:
                self.o.rate_dematch(bg, Z, buf["soft"][r], cb, new_data, rv, qm, Nref, g["fillers"])
                continue
            it, out = self.o.cb_chain(bg, Z, buf["soft"][r], cb, new_data, rv, qm, Nref, g["fillers"], g["crc"],
                                      max_iter, early_stop)
            buf["msg"][r] = out
            if it > 0:
                buf["crc"][r] = True
                iters.append(it)
            else:
                iters.append(max_iter)
        res = dict(tb_crc_ok=False, nof_codeblocks=C_, iters=iters, delivered=False)
        if C_ == 1:
            res["tb_crc_ok"] = bool(buf["crc"][0])
            if res["tb_crc_ok"]:
                tb_out[:] = buf["msg"][0][:tbs // 8]
                res["delivered"] = True
        elif buf["crc"].all():
            stream = np.concatenate([_bits_of(buf["msg"][r], g["payload"][r]) for r in range(C_)])
            assert stream.size == tbs + 24
            tb_out[:] = np.packbits(stream[:tbs])
            res["delivered"] = True
            res["tb_crc_ok"] = self.o.crc(CRC24A, np.packbits(stream), stream.size) == 0
            if not res["tb_crc_ok"]:
                buf["crc"][:] = False
        if res["tb_crc_ok"]:
            del self.buffers[buffer_id]  # unique_rm_buffer.release()
        return res

    def peek(self, buffer_id):
        return self.buffers.get(buffer_id)


class RefPool:
    """The reference's rx_buffer_pool_impl alone (oracle/_ref): reserve returns a handle or None."""

    def __init__(self, ref, nof_buffers, nof_codeblocks, expire_timeout_slots, numerology=0):
        self.lib = ref.lib
        self.lib.ref_pool_create.restype = C.c_void_p
        self.lib.ref_pool_create.argtypes = [C.c_uint, C.c_uint, C.c_uint, C.c_uint]
        self.lib.ref_pool_destroy.argtypes = [C.c_void_p]
        self.lib.ref_pool_reserve.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.c_uint, C.c_int]
        self.lib.ref_pool_codeblocks.argtypes = [C.c_void_p, C.c_uint, C.c_void_p]
        for f in (self.lib.ref_pool_unlock, self.lib.ref_pool_release, self.lib.ref_pool_run_slot):
            f.argtypes = [C.c_void_p, C.c_uint]
            f.restype = None
        self.h = self.lib.ref_pool_create(nof_buffers, nof_codeblocks, expire_timeout_slots, numerology)

    def close(self):
        if self.h:
            self.lib.ref_pool_destroy(self.h)
            self.h = None

    def reserve(self, slot, identifier, nof_codeblocks, new_data):
        h = self.lib.ref_pool_reserve(self.h, slot, identifier, nof_codeblocks, int(new_data))
        return None if h < 0 else h

    def codeblocks(self, handle):
        ids = np.zeros(256, np.uint32)
        n = self.lib.ref_pool_codeblocks(self.h, handle, _p(ids))
        return ids[:n].copy()

    def unlock(self, handle):
        self.lib.ref_pool_unlock(self.h, handle)

    def release(self, handle):
        self.lib.ref_pool_release(self.h, handle)

    def run_slot(self, slot):
        self.lib.ref_pool_run_slot(self.h, slot)


class RefTb:
    """The reference's pusch_decoder_impl + rx buffer pool (oracle/_ref, generic LDPC kernels)."""

    def __init__(self, ref, nof_buffers=16, nof_codeblocks=2048):
        self.lib = ref.lib
        self.lib.ref_tb_create.restype = C.c_void_p
        self.lib.ref_tb_destroy.argtypes = [C.c_void_p]
        self.h = C.c_void_p(self.lib.ref_tb_create(C.c_uint(nof_buffers), C.c_uint(nof_codeblocks)))

    def close(self):
        if self.h:
            self.lib.ref_tb_destroy(self.h)
            self.h = None

    def decode(self, buffer_id, tbs, bg, rv, qm, Nref, nof_layers, new_data, max_iter, early_stop, llrs, tb_out):
        llrs = np.ascontiguousarray(llrs, np.int8)
        info = np.zeros(8, np.uint32)
        rc = self.lib.ref_tb_decode(self.h, C.c_uint(buffer_id), C.c_uint(tbs), bg, C.c_uint(rv), C.c_uint(qm),
                                    C.c_uint(Nref), C.c_uint(nof_layers), int(new_data), C.c_uint(max_iter),
                                    int(early_stop), _p(llrs), C.c_uint(llrs.size), _p(tb_out), _p(info))
        if rc != 0:
            raise RuntimeError("ref_tb_decode failed: %d" % rc)
        return dict(tb_crc_ok=bool(info[0]), nof_codeblocks=int(info[1]), iter_min=int(info[2]), iter_max=int(info[3]),
                    nof_obs=int(info[4]))

    def peek(self, buffer_id, nof_cbs, cb_len):
        flags = np.zeros(nof_cbs, np.uint8)
        soft = np.zeros((nof_cbs, cb_len), np.int8)
        n = self.lib.ref_tb_peek(self.h, C.c_uint(buffer_id), C.c_uint(nof_cbs), C.c_uint(cb_len), _p(flags), _p(soft))
        return None if n < 0 else (flags.astype(bool), soft)


# ---- UL-SCH demultiplexer (last pointer: ulsch_demultiplex_impl) --------------------------------------
ULSCH_CFG_FIELDS = ("qm", "nof_layers", "nof_prb", "start_symbol", "nof_symbols", "nof_harq_ack_rvd", "dmrs_type", "dmrs_mask",
                    "nof_cdm_groups_without_data", "nof_harq_ack_bits", "nof_enc_harq_ack_bits", "nof_csi_part1_bits",
                    "nof_enc_csi_part1_bits", "nof_csi_part2_bits", "nof_enc_csi_part2_bits", "csi2_after_csi1")


def ulsch_cfg_array(cfg):
    return np.array([int(cfg.get(k, 0)) for k in ULSCH_CFG_FIELDS], np.uint32)


def ulsch_codeword_len(cfg):
    """Soft bits of the allocation: every RE of every OFDM symbol that is not a DM-RS RE, times Qm x layers."""
    per_prb_dmrs = 12 - cfg["nof_cdm_groups_without_data"] * (6 if cfg["dmrs_type"] == 1 else 4)
    n = 0
    for l in range(cfg["start_symbol"], cfg["start_symbol"] + cfg["nof_symbols"]):
        n += (per_prb_dmrs if (cfg["dmrs_mask"] >> l) & 1 else 12) * cfg["nof_prb"]
    return n * cfg["qm"] * cfg["nof_layers"]


def _ulsch_call(fn, cfg, codeword, scrambling, extra=()):
    codeword = np.ascontiguousarray(codeword, np.int8)
    scrambling = np.ascontiguousarray(scrambling, np.uint8)
    cap = codeword.size
    outs = [np.zeros(cap, np.int8) for _ in range(4)]
    counts = np.zeros(8, np.uint32)
    c = ulsch_cfg_array(cfg)
    rc = fn(_p(c), _p(codeword), C.c_uint32(codeword.size), _p(scrambling), *extra, *[_p(o) for o in outs], C.c_uint32(cap),
            _p(counts))
    return rc, [o[:int(k)] for o, k in zip(outs, counts[:4])], counts[4:].astype(bool)


def oracle_ulsch_demux(oracle, cfg, codeword, scrambling):
    """oracle/fec_ulsch.c. Returns (rc, [sch, harq_ack, csi1, csi2], ended flags)."""
    return _ulsch_call(oracle.lib.orc_ulsch_demux, cfg, codeword, scrambling)


def ref_ulsch_demux(ref, cfg, codeword, scrambling, block=4096):
    """The reference's ulsch_demultiplex_impl (oracle/_ref), fed `block` soft bits per on_new_block call."""
    return _ulsch_call(ref.lib.ref_ulsch_demux, cfg, codeword, scrambling, extra=(C.c_uint32(block),))

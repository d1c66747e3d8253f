"""Seeded case generators shared by the golden-vector script and the parity tests."""
import numpy as np

ALL_Z = [2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 18, 20, 22, 24, 26, 28, 30, 32, 36, 40, 44, 48, 52, 56, 60,
         64, 72, 80, 88, 96, 104, 112, 120, 128, 144, 160, 176, 192, 208, 224, 240, 256, 288, 320, 352, 384]
ORC_CRC_LEN = {0: 24, 1: 24, 2: 16}  # oracle numbering: 0=CRC24A 1=CRC24B 2=CRC16
ORC_TO_HAL_CRC = {0: 2, 1: 1, 2: 0}


def dims(bg, Z):
    kb, nb = (22, 66) if bg == 1 else (10, 50)
    return kb * Z, nb * Z


def random_chain_params(rng, bg=None, Z=None, z_pool=None):
    """Random (but valid) per-CB parameters in the spirit of BASELINE config 2."""
    bg = int(rng.choice([1, 2])) if bg is None else bg
    Z = int(rng.choice(z_pool or ALL_Z)) if Z is None else Z
    K, N = dims(bg, Z)
    Kb = 22 if bg == 1 else 10
    qm = int(rng.choice([1, 2, 4, 6, 8]))
    poly = int(rng.choice([0, 1, 2]))
    if K - ORC_CRC_LEN[poly] < 4:
        poly = 2  # tiny codeblocks (BG2, Z = 2) only have room for CRC16
    max_fill = min((Kb - 2) * Z - 1, 2 * Z, K - ORC_CRC_LEN[poly] - 1)
    fill = int(rng.integers(0, max_fill + 1)) if (rng.random() < 0.5 and max_fill > 0) else 0
    lo = K - fill + 2 * Z
    E = int(rng.integers(lo, 3 * N + 1))
    E = max(qm, (E // qm) * qm)
    rv = int(rng.choice([0, 0, 0, 1, 2, 3]))
    Nref = 0 if rng.random() < 0.6 else int(rng.integers(K + 2 * Z + 1, N + 10))
    rate = (K - fill) / E
    # Es/N0 around the waterfall of that rate (BPSK-per-bit channel model of oracle/fec_tx.c)
    snr = 10 * np.log10(max(2 ** (2 * rate) - 1, 1e-3)) - 0.5 + float(rng.uniform(-2.5, 5.0))
    return dict(bg=bg, Z=Z, E=E, rv=rv, qm=qm, Nref=Nref, fill=fill, poly=poly, snr=snr,
                max_iter=int(rng.choice([1, 2, 6, 6, 10, 10])), early=bool(rng.random() < 0.7))


def fill_chain_desc(fec, d, i, p, new_data, harq_slot, llr_offset, out_offset, mode=None):
    d["base_graph"][i], d["lifting_size"][i], d["rv"][i], d["qm"][i] = p["bg"], p["Z"], p["rv"], p["qm"]
    d["Nref"][i], d["nof_filler_bits"][i], d["E"][i] = p["Nref"], p["fill"], p["E"]
    d["crc_type"][i], d["max_iter"][i] = ORC_TO_HAL_CRC[p["poly"]], p["max_iter"]
    d["mode"][i] = fec.MODE_CHAIN if mode is None else mode
    d["flags"][i] = (fec.FLAG_NEW_DATA if new_data else 0) | (fec.FLAG_EARLY_STOP if p["early"] else 0)
    d["harq_slot"][i], d["llr_offset"][i], d["out_offset"][i] = harq_slot, llr_offset, out_offset


def random_tb_params(rng, max_bits_bg1=30000, max_bits_bg2=9000):
    """A random transport block: (tbs, bg, qm, nof_layers, nof_llrs G, Es/N0) with a code rate in (0.35, 0.93)."""
    bg = int(rng.integers(1, 3))
    qm = int(rng.choice([1, 2, 4, 6, 8]))
    nl = int(rng.integers(1, 5))
    tbs = int(rng.integers(3, (max_bits_bg1 if bg == 1 else max_bits_bg2) // 8)) * 8
    B = tbs + (16 if tbs <= 3824 else 24)
    seg = 8448 if bg == 1 else 3840
    C_ = 1 if B <= seg else -(-B // (seg - 24))
    rate = float(rng.uniform(0.35, 0.93))
    nsym = max(1, int((B + 24 * C_) / rate / (qm * nl))) * nl
    snr = float(rng.uniform(-1.0, 6.0)) + 10 * np.log10(rate)
    return dict(tbs=tbs, bg=bg, qm=qm, nl=nl, G=nsym * qm, snr=snr)


# ---- soft demodulation ---------------------------------------------------------------------------
def _around(x, k):
    x = np.float32(x)
    out, a, b = [x], x, x
    for _ in range(k):
        a = np.nextafter(a, np.float32(np.inf))
        b = np.nextafter(b, np.float32(-np.inf))
        out += [a, b]
    return out


def demod_crafted_symbols(qm, noise_var, dense=1):
    """Symbols whose LLRs sit on the rounding decisions where the reference's code paths differ: the edges of the
    piecewise-linear intervals (+- a few ulp), and values whose scaled LLR is k + 0.5 (round-half-even vs half-away,
    divide vs reciprocal-multiply, fused vs separate multiply-add)."""
    M = {1: 1 / np.sqrt(2), 2: 1 / np.sqrt(2), 4: 1 / np.sqrt(10), 6: 1 / np.sqrt(42), 8: 1 / np.sqrt(170)}[qm]
    xs = []
    for k in range(-18, 19):
        xs += _around(np.float32(k) * np.float32(M), 4)
    for s in (4, 8, 12, 16, 20, 24, 28, 32, 2.8284271):
        for k in range(-121, 122, 3 * dense):
            xs += _around((k + 0.5) / (6.0 * s * M) * noise_var if qm > 2 else (k + 0.5) / (5.0 * s) * noise_var, 2)
    xs = np.array(xs, np.float32)
    return np.stack([xs, np.roll(xs, 7)], 1)


def demod_random_symbols(rng, qm, n, specials=True):
    """Random equalised symbols and noise variances, with the special values the reference handles explicitly: zero,
    negative and NaN noise variances (LLR = 0), exact-zero and near-zero symbols."""
    amp = {1: 1.0, 2: 1.0, 4: 1.2, 6: 1.3, 8: 1.4}[qm]
    sym = rng.uniform(-amp, amp, (n, 2)).astype(np.float32)
    nv = (rng.choice([0.001, 0.01, 0.05, 0.2, 1.0], n) * rng.uniform(0.5, 1.5, n)).astype(np.float32)
    if specials and n >= 16:
        nv[rng.integers(0, n, 3)] = [0.0, -1.0, np.nan]
        sym[rng.integers(0, n, 2)] = 0.0
        sym[rng.integers(0, n, 1), 0] = 1e-10
        sym[rng.integers(0, n, 1)] = [1e-6, -1e-6]
    return sym, nv


def demod_extreme_symbols(rng, qm, n=4099):
    """Inputs outside anything an equaliser produces, where the reference's behaviour comes from x86 conversion rules and
    IEEE special values rather than from its formulas: huge, infinite and NaN symbols (float -> int32 out of range gives
    0x80000000 and the interval index clamps to 0), denormal noise variances (1 / var = inf, 0 * inf = NaN -> LLR 0)."""
    sym, nv = demod_random_symbols(rng, qm, n)
    special = np.array([1e5, -1e5, 3e9, -3e9, 16384.0, -16383.9, np.inf, -np.inf, np.nan, 1e30, -1e30, 2.5e8, -2.5e8],
                       np.float32)
    idx = rng.choice(n, 60, replace=False)
    sym[idx[:40], rng.integers(0, 2, 40)] = rng.choice(special, 40)
    nv[idx[40:]] = rng.choice(np.array([1e-42, 1e-39, 1e38, 3e38, np.inf, 1e-30], np.float32), 20)
    return sym, nv

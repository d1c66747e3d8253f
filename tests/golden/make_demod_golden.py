"""Generates tests/golden/demod_golden.npz: soft-demodulation and scrambling-sequence vectors produced by the REFERENCE
itself (oracle/_ref/libref_demod_{generic,avx2,avx512}.so = the unmodified demodulation_mapper_impl and
pseudo_random_generator_impl of /root/reference compiled for three instruction sets, see oracle/Makefile).
Run in the container that has /root/reference:  python tests/golden/make_demod_golden.py
Per modulation: symbols, noise variances, block lengths (one demodulate_soft call each) and the LLRs of each flavour."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge  # noqa: E402
import cases  # noqa: E402


def main():
    orc = ge.load_oracle()
    assert orc.build_ref()
    refs = [orc.RefDemod(f) for f in range(3)]
    rng = np.random.default_rng(20261105)
    data = {}
    for qm in (1, 2, 4, 6, 8):
        sym_r, nv_r = cases.demod_random_symbols(rng, qm, 1500)
        sym_c = cases.demod_crafted_symbols(qm, 0.25, dense=3)
        sym = np.concatenate([sym_r, sym_c]).astype(np.float32)
        nv = np.concatenate([nv_r, np.full(len(sym_c), 0.25, np.float32)])
        # ragged blocks so that every path (AVX-512 loop, AVX2 loop, scalar tail) is taken at many positions
        blocks, left = [], len(sym)
        while left:
            b = min(left, int(rng.choice([1, 5, 12, 37, 100, 399])))
            blocks.append(b)
            left -= b
        data[f"q{qm}_sym"], data[f"q{qm}_noise"], data[f"q{qm}_blocks"] = sym, nv, np.array(blocks, np.uint32)
        for f, ref in enumerate(refs):
            out, pos = [], 0
            for b in blocks:
                out.append(ref.demodulate_soft(qm, sym[pos:pos + b], nv[pos:pos + b]))
                pos += b
            data[f"q{qm}_llr_f{f}"] = np.concatenate(out)
        if qm == 1:
            data["q1_llr_pi2"] = refs[0].demodulate_soft(1, sym, nv, pi2=True)
    seq_meta = []
    for i, (c_init, off, nbits) in enumerate([(0x12345, 0, 1000), (1, 5, 77), (0x7fffffff, 12345, 40000),
                                               (30000 * 32768 + 500, 0, 3000), (0, 0, 64), (0x2aaaaaaa, 1 << 20, 4096)]):
        seq_meta.append((c_init, off, nbits))
        data[f"seq{i}"] = refs[2].gold_sequence(c_init, off, nbits)
    data["seq_meta"] = np.array(seq_meta, np.uint32)
    path = os.path.join(HERE, "demod_golden.npz")
    np.savez_compressed(path, **data)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()

"""Generates tests/golden/ulsch_golden.npz from the REFERENCE (oracle/_ref/libref_fec.so: the unmodified
ulsch_demultiplex_impl.cpp behind oracle/ref_ulsch_shim.cpp). Run in the container that has /root/reference:
    python tests/golden/make_ulsch_golden.py
28 seeded configurations: every combination of HARQ-ACK (none / 1-2 bits on reserved elements / more), CSI Part 1 and CSI Part
2 (known at configuration time or when CSI Part 1 ends), all modulations, 1-4 layers, both DM-RS types."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))
import __graft_entry__ as ge  # noqa: E402
import ulsch_cases  # noqa: E402

orc = ge.load_oracle()
assert orc.Ref.available() or orc.build_ref()
ref = orc.Ref()
rng = np.random.default_rng(38212)
out, i, kinds, tries = {}, 0, set(), 0
while i < 28:
    tries += 1
    cfg = ulsch_cases.random_ulsch_cfg(rng)
    n = orc.ulsch_codeword_len(cfg)
    if n > 16000:
        continue
    cw, seq = ulsch_cases.random_codeword(rng, n)
    rc, streams, _ = orc.ref_ulsch_demux(ref, cfg, cw, seq, 4096)
    assert rc == 0
    kind = tuple(s.size > 0 for s in streams) + (cfg["csi2_after_csi1"], min(cfg["nof_harq_ack_bits"], 3))
    if kind in kinds and i < 20 and tries < 2000:
        continue
    kinds.add(kind)
    out["c%d_cfg" % i] = orc.ulsch_cfg_array(cfg)
    out["c%d_cw" % i], out["c%d_seq" % i] = cw, seq
    for name, s in zip(("sch", "harq", "csi1", "csi2"), streams):
        out["c%d_%s" % (i, name)] = s
    i += 1
out["n"] = np.array(i)
np.savez_compressed(os.path.join(HERE, "ulsch_golden.npz"), **out)
print("wrote", i, "cases,", len(kinds), "distinct kinds")

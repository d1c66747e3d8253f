"""Seeded UL-SCH multiplexing configurations (TS 38.212 6.2.7) shared by the CPU / GPU tests and the golden-vector script."""
import numpy as np


def random_ulsch_cfg(rng, big=False):
    qm = int(rng.choice([1, 2, 4, 6, 8]))
    nl = int(rng.integers(1, 5))
    nprb = int(rng.integers(1, 275 if big else 40))
    dmrs_type = int(rng.choice([1, 2]))
    ncdm = int(rng.integers(1, 3 if dmrs_type == 1 else 4))
    start = int(rng.integers(0, 3))
    nsym = int(rng.integers(6, 15 - start))
    masks = [0b00000000000100, 0b00100000000100, 0b00100010000100, 0b00100100100100, 0b00000000001100, 0b00000000000001]
    dmrs_mask = int(rng.choice(masks))
    if not any((dmrs_mask >> l) & 1 for l in range(start, start + nsym)):
        dmrs_mask = 1 << start
    B = qm * nl
    per_prb_dmrs = 12 - ncdm * (6 if dmrs_type == 1 else 4)
    re_tot = sum((per_prb_dmrs if (dmrs_mask >> l) & 1 else 12) * nprb for l in range(start, start + nsym))
    re_uci = sum(0 if (dmrs_mask >> l) & 1 else 12 * nprb for l in range(start, start + nsym))
    kind = int(rng.integers(0, 6))
    cfg = dict(qm=qm, nof_layers=nl, nof_prb=nprb, start_symbol=start, nof_symbols=nsym, nof_harq_ack_rvd=0, dmrs_type=dmrs_type,
               dmrs_mask=dmrs_mask, nof_cdm_groups_without_data=ncdm, nof_harq_ack_bits=0, nof_enc_harq_ack_bits=0,
               nof_csi_part1_bits=0, nof_enc_csi_part1_bits=0, nof_csi_part2_bits=0, nof_enc_csi_part2_bits=0, csi2_after_csi1=0)
    budget = max(1, re_uci // 3)  # resource elements the UCI may take in all
    def enc(max_re):
        return int(rng.integers(1, max(2, max_re))) * B
    if kind in (1, 2, 4, 5):  # HARQ-ACK
        o_ack = int(rng.choice([1, 2, 3, 5, 11, 20]))
        cfg["nof_harq_ack_bits"] = o_ack
        cfg["nof_enc_harq_ack_bits"] = enc(budget // 3)
        if o_ack <= 2:
            cfg["nof_harq_ack_rvd"] = cfg["nof_enc_harq_ack_bits"] + int(rng.integers(0, 4)) * B
    elif rng.random() < 0.3:
        cfg["nof_harq_ack_rvd"] = enc(budget // 3)  # reserved elements without HARQ-ACK in this slot
    if kind in (2, 3, 4, 5):  # CSI part 1
        cfg["nof_csi_part1_bits"] = int(rng.choice([1, 2, 4, 7, 20]))
        cfg["nof_enc_csi_part1_bits"] = enc(budget // 3)
    if kind in (4, 5) :  # CSI part 2
        cfg["nof_csi_part2_bits"] = int(rng.choice([1, 2, 3, 9, 30]))
        cfg["nof_enc_csi_part2_bits"] = enc(budget // 3)
        cfg["csi2_after_csi1"] = int(kind == 5)
    return cfg


def random_codeword(rng, n):
    cw = rng.integers(-120, 121, n).astype(np.int8)
    cw[rng.integers(0, n, max(1, n // 50))] = 0
    return cw, rng.integers(0, 2, n).astype(np.uint8)

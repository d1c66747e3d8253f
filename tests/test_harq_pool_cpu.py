"""The engine's HARQ buffer pool (include/fec_b200.h fec_harq_pool_*) against the reference's own
rx_buffer_pool_impl (oracle/_ref), operation for operation: random sequences of reserve (new data and retransmissions,
changing codeblock counts, unknown identifiers), unlock, release and run_slot, with slot counts that wrap around the
hyper system frame. Compared: whether each reservation is granted, and the absolute codeblock identifiers it holds.
Host only (the pool runs without an engine here)."""
import numpy as np
import pytest


def differential_run(fec, orc, ref, seed, nof_buffers, nof_codeblocks, timeout, numerology, start_slot, steps=3000):
    rng = np.random.default_rng(seed)
    ours = fec.HarqPool(None, nof_buffers, nof_codeblocks, timeout, numerology=numerology)
    theirs = orc.RefPool(ref, nof_buffers, nof_codeblocks, timeout, numerology)
    wrap = 10240 << numerology
    slot = start_slot
    held = []  # (our buffer, their handle)
    ncb_of = {}
    granted = refused = 0
    reasons = set()
    try:
        for step in range(steps):
            op = rng.random()
            if op < 0.45:
                ident = (int(rng.integers(1, 9)) << 8) | int(rng.integers(0, 4))
                new_data = bool(rng.random() < 0.6)
                if new_data or ident not in ncb_of or rng.random() < 0.1:
                    ncb = int(rng.integers(1, 12))
                else:
                    ncb = ncb_of[ident]
                st, buf = ours.reserve(slot, ident, ncb, new_data)
                h = theirs.reserve(slot, ident, ncb, new_data)
                assert (st == fec.POOL_OK) == (h is not None), (step, st, h, ident, ncb, new_data)
                reasons.add(st)
                if h is not None:
                    assert (ours.codeblocks(buf) == theirs.codeblocks(h)).all(), step
                    assert len(ours.codeblocks(buf)) == ncb
                    held.append((buf, h))
                    ncb_of[ident] = ncb
                    granted += 1
                else:
                    refused += 1
            elif op < 0.75 and held:
                buf, h = held.pop(int(rng.integers(0, len(held))))
                if rng.random() < 0.5:
                    ours.unlock(buf), theirs.unlock(h)
                else:
                    ours.release(buf), theirs.release(h)
            else:
                slot = (slot + int(rng.choice([1, 1, 1, 2, 5, timeout]))) % wrap
                ours.run_slot(slot), theirs.run_slot(slot)
        for buf, h in held:
            ours.unlock(buf), theirs.unlock(h)
    finally:
        ours.close(), theirs.close()
    return granted, refused, reasons


@pytest.mark.parametrize("seed", range(6))
def test_pool_matches_reference(fec, orc, ref, seed):
    cfgs = [(4, 24, 8, 0, 0), (8, 40, 20, 1, 20470), (3, 64, 4, 0, 10230), (16, 30, 100, 2, 40950), (2, 12, 3, 0, 5000),
            (6, 18, 10, 1, 0)]
    nb, ncb, timeout, mu, start = cfgs[seed]
    granted, refused, reasons = differential_run(fec, orc, ref, seed, nb, ncb, timeout, mu, start)
    assert granted > 200 and refused > 50
    assert fec.POOL_OK in reasons and len(reasons) >= 4  # several distinct refusal reasons were exercised


def test_pool_lifecycle(fec):
    """The flow of one HARQ process as pusch_processor drives it: reserve(new) -> decode fails -> unlock; reserve(retx) ->
    decode passes -> release; identifiers linger until run_slot; expiry frees a reservation that is not renewed."""
    pool = fec.HarqPool(None, nof_buffers=2, nof_codeblocks=10, expire_timeout_slots=4)
    st, b = pool.reserve(0, 0x4601 << 8, 3, True)
    assert st == fec.POOL_OK and list(pool.codeblocks(b)) == [0, 1, 2] and pool.nof_free_codeblocks() == 7
    assert pool.reserve(0, 0x4601 << 8, 3, False)[0] == fec.POOL_IN_USE  # still locked
    pool.unlock(b)
    assert pool.reserve(1, 0x4601 << 8, 4, False)[0] == fec.POOL_CB_MISMATCH
    assert pool.reserve(1, 0x9999 << 8, 3, False)[0] == fec.POOL_RETX_UNKNOWN
    st, b2 = pool.reserve(1, 0x4601 << 8, 3, False)
    assert st == fec.POOL_OK and b2 == b and list(pool.codeblocks(b)) == [0, 1, 2]
    pool.release(b)
    assert pool.nof_free_codeblocks() == 10
    st, c = pool.reserve(2, 0x4602 << 8, 9, True)   # the freed identifiers went to the back of the FIFO
    assert st == fec.POOL_OK and list(pool.codeblocks(c)) == [3, 4, 5, 6, 7, 8, 9, 0, 1]
    assert pool.reserve(2, 0x4603 << 8, 1, True)[0] == fec.POOL_NO_BUFFER  # the released buffer keeps its identifier ...
    pool.run_slot(3)                                                        # ... until the next run_slot
    assert pool.reserve(3, 0x4603 << 8, 5, True)[0] == fec.POOL_NO_CODEBLOCKS
    pool.unlock(c)
    pool.run_slot(5)
    assert pool.nof_free_codeblocks() == 1   # not expired yet (renewed while locked at slot 3: expires at 7)
    pool.run_slot(7)
    assert pool.nof_free_codeblocks() == 10
    pool.stop()
    assert pool.reserve(8, 1 << 8, 1, True)[0] == fec.POOL_STOPPED
    pool.close()

"""Drop-in acceptance through the reference's own C++ interfaces: tests/cpp/dropin_pusch_decoder (built by
__graft_entry__.build() where the reference tree is present, shipped to the GPU box as a binary) runs
srsran::b200::pusch_decoder_cuda and the reference's pusch_decoder_impl side by side on the same HARQ transmissions."""
import os
import struct
import subprocess

import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu
BIN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "cpp", "_build", "dropin_pusch_decoder")


def test_pusch_decoder_cuda_is_a_drop_in(orc, oracle, tmp_path):
    if not os.path.exists(BIN):
        pytest.skip("tests/cpp/_build/dropin_pusch_decoder not built (needs the reference headers)")
    rng = np.random.default_rng(77)
    blocks = []
    for b in range(8):
        p = cases.random_tb_params(rng, 40000, 9000)
        if b == 0:
            p = dict(tbs=1277992, bg=1, qm=8, nl=4, G=1362816, snr=8.0)  # BASELINE config 3, 152 codeblocks
        tb = rng.integers(0, 256, p["tbs"] // 8, dtype=np.uint8)
        enc = orc.tb_encode(oracle, tb, p["bg"])
        txs = [(rv, orc.tb_transmit(oracle, enc, p["tbs"], p["bg"], rv, p["qm"], 0, p["nl"], p["G"],
                                    p["snr"] - (0.0 if b == 0 else 1.5), seed=b * 16 + t))
               for t, rv in enumerate([0, 2, 3, 1] if b else [0])]
        blocks.append((p, txs))
    path = tmp_path / "dropin.bin"
    with open(path, "wb") as f:
        f.write(struct.pack("<I", len(blocks)))
        for p, txs in blocks:
            f.write(struct.pack("<6I", p["tbs"], p["bg"], p["qm"], p["nl"], p["G"], len(txs)))
            for rv, llr in txs:
                f.write(struct.pack("<I", rv))
                f.write(llr.tobytes())
    r = subprocess.run([BIN, str(path), "100"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    print(r.stdout)
    assert "OK blocks=8" in r.stdout and r.stdout.count("LATENCY") == 3, r.stdout
    failed = int(r.stdout.split("failed_transmissions=")[1])
    assert failed >= 2, r.stdout  # the retransmission path was exercised


ENC_BIN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "cpp", "_build", "dropin_pdsch_encoder")


def test_pdsch_encoder_hw_impl_over_the_cuda_hal():
    """The reference's unmodified pdsch_encoder_hw_impl over srsran::b200::cuda_pdsch_enc_accelerator gives the codeword
    of the reference's software pdsch_encoder_impl, for the 152-codeblock config-3 block and 40 random blocks."""
    if not os.path.exists(ENC_BIN):
        pytest.skip("tests/cpp/_build/dropin_pdsch_encoder not built (needs the reference headers)")
    r = subprocess.run([ENC_BIN, "40"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    print(r.stdout)
    assert "OK transport_blocks=42" in r.stdout, r.stdout


DEMOD_BIN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "cpp", "_build", "dropin_demodulation_mapper")


def test_demodulation_mapper_over_the_cuda_engine():
    """The reference's own demodulation_mapper_impl (built into the test binary for AVX-512, the flavour of the hosts used
    here) against srsran::b200::cuda_demodulation_mapper obtained through the demodulation_mapper_factory interface:
    every modulation, lengths around the SIMD block sizes, zero / negative / NaN noise variances, bit for bit."""
    if not os.path.exists(DEMOD_BIN):
        pytest.skip("tests/cpp/_build/dropin_demodulation_mapper not built (needs the reference sources)")
    r = subprocess.run([DEMOD_BIN, "120"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    print(r.stdout)
    assert "OK cases=120" in r.stdout, r.stdout


ULSCH_BIN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "cpp", "_build", "dropin_ulsch_demultiplex")


def test_ulsch_demultiplex_cuda_is_a_drop_in():
    """The reference's ulsch_demultiplex_impl and srsran::b200::cuda_ulsch_demultiplex behind the same interfaces on 150 random
    configurations, CSI Part 2 announced at the start or from inside the CSI Part 1 buffer's on_end_softbits: every stream
    and every end-of-stream notification identical."""
    if not os.path.exists(ULSCH_BIN):
        pytest.skip("tests/cpp/_build/dropin_ulsch_demultiplex not built (needs the reference headers)")
    r = subprocess.run([ULSCH_BIN, "150"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    print(r.stdout)
    assert "OK cases=150" in r.stdout, r.stdout
    assert int(r.stdout.split("csi2_after_csi1=")[1]) >= 10, r.stdout

"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol include/fec_b200.h declares, the
host-only validation behaves like the reference's asserts, and the product refuses to run without a CUDA device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    names = set()
    for fn in os.listdir(os.path.join(ROOT, "include")):
        if fn.endswith(".h"):
            src = open(os.path.join(ROOT, "include", fn)).read()
            src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
            names |= set(re.findall(r"\b(fec_[a-z0-9_]+)\s*\(", src))
    return names


def test_library_exports_every_declared_symbol(fec):
    lib = fec.load_library()
    declared = _declared_symbols()
    assert len(declared) >= 20
    for name in sorted(declared):
        assert hasattr(lib, name), name
    assert set(fec.EXPORTS) <= declared
    assert lib.fec_abi_version() == fec.ABI_VERSION


def test_struct_layouts_match_header(fec):
    assert fec.DESC_DTYPE.itemsize == 48
    assert fec.RESULT_DTYPE.itemsize == 8
    assert C.sizeof(fec.EngineCfg) == 48
    assert C.sizeof(fec.EngineStats) == 8 * (14 + fec.STATS_ITER_BINS)
    assert C.sizeof(fec.UlschCfg) == 40 and C.sizeof(fec.UlschSizes) == 20


def test_desc_prepare_fills_k0_like_the_reference(fec):
    """k0 = floor(f_rv * Ncb / N) * Z (ldpc_rate_dematcher_impl.cpp:104-105)."""
    d = fec.make_descs(1)
    d["base_graph"], d["lifting_size"], d["qm"], d["E"], d["max_iter"], d["crc_type"] = 1, 384, 2, 8960, 6, fec.CRC24B
    for rv, f in enumerate((0, 17, 33, 56)):
        d["rv"] = rv
        fec.desc_prepare(d)
        assert d["k0"][0] == f * 384 and d["graph_index"][0] == 50
    d["base_graph"], d["lifting_size"], d["Nref"], d["rv"] = 2, 30, 1119, 3
    fec.desc_prepare(d)
    assert d["k0"][0] == int(43 * 1119 / 1500) * 30 and d["graph_index"][0] == 51 + fec.ALL_Z.index(30)


@pytest.mark.parametrize("field,value", [("base_graph", 3), ("lifting_size", 17), ("qm", 3), ("rv", 4), ("max_iter", 0),
                                         ("crc_type", 3), ("E", 7), ("nof_filler_bits", 384 * 20)])
def test_desc_prepare_rejects_bad_fields(fec, field, value):
    d = fec.make_descs(1)
    d["base_graph"], d["lifting_size"], d["qm"], d["E"], d["max_iter"], d["crc_type"] = 1, 384, 2, 8960, 6, fec.CRC24B
    d[field] = value
    with pytest.raises(fec.FecError):
        fec.desc_prepare(d)


def test_no_cpu_fallback(fec):
    """Without a CUDA device engine creation must fail loudly (FEC_ERR_NO_DEVICE), never decode on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(fec.FecError, match="no CPU fallback"):
        fec.FecEngine(device=0, harq_slots=4, max_batch_cbs=4)


def test_product_does_not_touch_the_oracle():
    """Nothing under the package may import, link or open anything under oracle/."""
    pkg = os.path.join(ROOT, "srsran-5g_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp")) and fn != "ldpc_tables.h":
                src = open(os.path.join(dirpath, fn), errors="ignore").read()
                assert "pyoracle" not in src and "fec_oracle" not in src and "libref_fec" not in src, fn

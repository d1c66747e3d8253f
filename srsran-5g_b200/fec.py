"""ctypes binding of the C ABI in include/fec_b200.h plus Python mirrors of the reference interfaces it serves.

The classes below keep the reference's names, argument meaning and error behaviour for this path so that the parity
tests read like the reference's own tests (paths relative to the reference tree):

  LdpcDecoder.decode            <- ldpc_decoder::decode            include/srsran/phy/upper/channel_coding/ldpc/ldpc_decoder.h:73
  LdpcRateDematcher.rate_dematch<- ldpc_rate_dematcher::rate_dematch include/srsran/phy/upper/channel_coding/ldpc/ldpc_rate_dematcher.h:52
  HwAcceleratorPuschDec         <- hal::hw_accelerator_pusch_dec   include/srsran/hal/phy/upper/channel_processors/pusch/hw_accelerator_pusch_dec.h:83
  HwAcceleratorPdschEnc         <- hal::hw_accelerator_pdsch_enc   include/srsran/hal/phy/upper/channel_processors/hw_accelerator_pdsch_enc.h:81
  LdpcEncoderRateMatcher        <- ldpc_encoder::encode + ldpc_rate_matcher::rate_match (channel_coding/ldpc/ldpc_encoder.h, ldpc_rate_matcher.h)

There is no CPU fallback: constructing FecEngine without the compiled extension or without a CUDA device raises.
"""
import ctypes as C
import os

import numpy as np

from . import build as _build

CRC16, CRC24B, CRC24A = 0, 1, 2  # hal::hw_dec_cb_crc_type numbering
MODE_CHAIN, MODE_DECODE, MODE_DEMATCH = 0, 1, 2
FLAG_NEW_DATA, FLAG_EARLY_STOP, FLAG_NO_CRC, FLAG_HARQ_ZERO = 1, 2, 4, 8
RES_CRC_PASS, RES_TOO_SHORT, RES_SKIPPED, RES_DONE = 1, 2, 4, 0x80
MAX_TB_CBS = 162
TB_SLOTS_CONTIGUOUS = 0xFFFFFFFF
HARQ_SLOT_BYTES = 25344
MAX_CB_OUT_BYTES = 1056
ABI_VERSION = 9
LAUNCH_EVEN_Z = 2  # fec_launch_device: any_bg1 flag bit
NOALLOC_MAX_TBS, NOALLOC_MAX_CBS = 64, 2048

ALL_Z = [2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 18, 20, 22, 24, 26, 28, 30, 32, 36, 40, 44, 48, 52, 56, 60,
         64, 72, 80, 88, 96, 104, 112, 120, 128, 144, 160, 176, 192, 208, 224, 240, 256, 288, 320, 352, 384]

# numpy mirror of struct fec_cb_desc (48 bytes) and fec_cb_result (8 bytes)
DESC_DTYPE = np.dtype({
    "names": ["llr_offset", "out_offset", "harq_slot", "E", "Nref", "nof_filler_bits", "k0", "lifting_size",
              "graph_index", "base_graph", "rv", "qm", "crc_type", "max_iter", "mode", "flags", "reserved0",
              "reserved1"],
    "formats": ["<u8", "<u4", "<u4", "<u4", "<u4", "<u4", "<u4", "<u2", "<u2", "u1", "u1", "u1", "u1", "u1", "u1",
                "u1", "u1", "<u4"],
    "offsets": [0, 8, 12, 16, 20, 24, 28, 32, 34, 36, 37, 38, 39, 40, 41, 42, 43, 44],
    "itemsize": 48})
RESULT_DTYPE = np.dtype({"names": ["nof_ldpc_iterations", "status", "nof_layers", "reserved"],
                         "formats": ["<u2", "u1", "u1", "<u4"], "offsets": [0, 2, 3, 4], "itemsize": 8})


# numpy mirrors of struct fec_tb_desc (48 bytes) and fec_tb_result (16 bytes)
TB_DESC_DTYPE = np.dtype({
    "names": ["llr_offset", "out_offset", "tbs", "nof_llrs", "Nref", "harq_slot", "slot_list", "base_graph", "rv", "qm",
              "nof_layers", "max_iter", "flags"],
    "formats": ["<u8", "<u8", "<u4", "<u4", "<u4", "<u4", "<u4", "u1", "u1", "u1", "u1", "u1", "u1"],
    "offsets": [0, 8, 16, 20, 24, 28, 32, 36, 37, 38, 39, 40, 41],
    "itemsize": 48})
TB_RESULT_DTYPE = np.dtype({
    "names": ["nof_codeblocks", "nof_cb_decoded", "nof_cb_crc_ok", "tb_crc_ok", "tb_delivered", "iter_sum", "iter_min",
              "iter_max"],
    "formats": ["<u2", "<u2", "<u2", "u1", "u1", "<u4", "<u2", "<u2"],
    "offsets": [0, 2, 4, 6, 7, 8, 12, 14], "itemsize": 16})


# numpy mirror of struct fec_enc_desc (40 bytes)
ENC_FLAG_UNPACKED = 1
MAX_CB_IN_BYTES = 1056
ENC_DESC_DTYPE = np.dtype({
    "names": ["out_offset", "in_offset", "E", "Nref", "k0", "lifting_size", "nof_filler_bits", "graph_index",
              "base_graph", "rv", "qm", "flags"],
    "formats": ["<u8", "<u4", "<u4", "<u4", "<u4", "<u2", "<u2", "<u2", "u1", "u1", "u1", "u1"],
    "offsets": [0, 8, 12, 16, 20, 24, 26, 28, 30, 31, 32, 33],
    "itemsize": 40})


# numpy mirror of struct fec_demod_desc (40 bytes)
DEMOD_GENERIC, DEMOD_AVX2, DEMOD_AVX512 = 0, 1, 2
DEMOD_FLAG_PI2_BPSK, DEMOD_FLAG_NO_DESCRAMBLE = 1, 2
DEMOD_DESC_DTYPE = np.dtype({
    "names": ["sym_offset", "llr_offset", "nof_symbols", "c_init", "seq_offset", "x1_state", "x2_state", "qm", "flags"],
    "formats": ["<u8", "<u8", "<u4", "<u4", "<u4", "<u4", "<u4", "u1", "u1"],
    "offsets": [0, 8, 16, 20, 24, 28, 32, 36, 37],
    "itemsize": 40})


class EngineCfg(C.Structure):
    _fields_ = [("device", C.c_int32), ("harq_slots", C.c_uint32), ("max_batch_cbs", C.c_uint32),
                ("max_batch_llrs", C.c_uint32), ("nof_queues", C.c_uint32), ("scaling_factor", C.c_float),
                ("kernel_variant", C.c_uint32), ("reserved", C.c_uint32), ("host_contexts", C.c_uint32),
                ("queue_batch_cbs", C.c_uint32), ("queue_batches", C.c_uint32), ("max_wait_us", C.c_uint32)]


STATS_ITER_BINS = 16


class UlschCfg(C.Structure):
    """fec_ulsch_cfg: mirror of ulsch_demultiplex::configuration + the arguments of set_csi_part2."""
    _fields_ = [("qm", C.c_uint8), ("nof_layers", C.c_uint8), ("start_symbol_index", C.c_uint8), ("nof_symbols", C.c_uint8),
                ("dmrs_type", C.c_uint8), ("nof_cdm_groups_without_data", C.c_uint8), ("csi2_after_csi1", C.c_uint8),
                ("reserved0", C.c_uint8), ("nof_prb", C.c_uint16), ("dmrs_symbol_mask", C.c_uint16),
                ("nof_harq_ack_rvd", C.c_uint32), ("nof_harq_ack_bits", C.c_uint32), ("nof_enc_harq_ack_bits", C.c_uint32),
                ("nof_csi_part1_bits", C.c_uint32), ("nof_enc_csi_part1_bits", C.c_uint32), ("nof_csi_part2_bits", C.c_uint32),
                ("nof_enc_csi_part2_bits", C.c_uint32)]


class UlschSizes(C.Structure):
    _fields_ = [("nof_codeword_bits", C.c_uint32), ("nof_sch_bits", C.c_uint32), ("nof_harq_ack_bits", C.c_uint32),
                ("nof_csi_part1_bits", C.c_uint32), ("nof_csi_part2_bits", C.c_uint32)]


def ulsch_cfg(**kw):
    """Builds a UlschCfg from the field names of ulsch_demultiplex::configuration (start_symbol / dmrs_mask accepted as short
    forms, as tests/ulsch_cases.py writes them)."""
    alias = {"start_symbol": "start_symbol_index", "dmrs_mask": "dmrs_symbol_mask"}
    c = UlschCfg()
    for k, v in kw.items():
        setattr(c, alias.get(k, k), int(v))
    return c


def ulsch_demux_sizes(cfg):
    """fec_ulsch_demux_sizes (host only): soft bits per stream as a dict."""
    lib = load_library()
    z = UlschSizes()
    _check(lib, lib.fec_ulsch_demux_sizes(C.byref(cfg), C.byref(z)))
    return {name: int(getattr(z, name)) for name, _ in UlschSizes._fields_}


class EngineStats(C.Structure):
    _fields_ = [("launches", C.c_uint64), ("cb_launches", C.c_uint64), ("codeblocks", C.c_uint64), ("results", C.c_uint64),
                ("crc_pass", C.c_uint64), ("iterations_hist", C.c_uint64 * STATS_ITER_BINS), ("queue_full", C.c_uint64),
                ("max_queue_depth", C.c_uint64), ("dispatches", C.c_uint64), ("dispatch_batches", C.c_uint64),
                ("tb_calls", C.c_uint64), ("transport_blocks", C.c_uint64), ("tb_crc_ok", C.c_uint64), ("faults", C.c_uint64),
                ("generic_launches", C.c_uint64)]


TB_CALLBACK = C.CFUNCTYPE(None, C.c_void_p, C.c_uint64, C.c_int, C.c_void_p, C.c_uint32)


class HarqPoolCfg(C.Structure):
    _fields_ = [("nof_buffers", C.c_uint32), ("first_slot", C.c_uint32), ("nof_codeblocks", C.c_uint32),
                ("expire_timeout_slots", C.c_uint32), ("numerology", C.c_uint32), ("reserved", C.c_uint32 * 3)]


POOL_OK, POOL_STOPPED, POOL_RETX_UNKNOWN, POOL_NO_BUFFER, POOL_CB_MISMATCH, POOL_IN_USE, POOL_NO_CODEBLOCKS = range(7)


EXPORTS = ["fec_abi_version", "fec_last_error", "fec_engine_default_cfg", "fec_engine_create", "fec_engine_destroy",
           "fec_engine_stream", "fec_harq_slot_ptr", "fec_engine_launch_count", "fec_desc_prepare",
           "fec_launch_device", "fec_run_host", "fec_host_alloc", "fec_host_free", "fec_harq_free", "fec_harq_read",
           "fec_harq_write", "fec_harq_flush", "fec_queue_reserve", "fec_queue_free", "fec_queue_enqueue", "fec_queue_flush",
           "fec_queue_poll", "fec_queue_fetch", "fec_tb_segment", "fec_tb_run_host", "fec_tb_reset_slots",
           "fec_enc_desc_prepare", "fec_encode_launch_device", "fec_encode_run_host", "fec_demod_desc_prepare",
           "fec_demod_launch_device", "fec_demod_run_host", "fec_tb_run_host_symbols", "fec_harq_pool_create",
           "fec_harq_pool_destroy", "fec_harq_pool_reserve", "fec_harq_pool_codeblocks", "fec_harq_pool_unlock",
           "fec_harq_pool_release", "fec_harq_pool_run_slot", "fec_harq_pool_stop", "fec_harq_pool_nof_free_codeblocks",
           "fec_tb_assemble_device_noalloc", "fec_engine_get_stats", "fec_engine_debug_fail_launches", "fec_tb_enqueue",
           "fec_tb_poll", "fec_tb_wait", "fec_queue_enqueue_batch"]

_lib = None


class FecError(RuntimeError):
    pass


def load_library():
    """Loads libfec_b200.so (building it first if sources are newer). Raises if it cannot be had: no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    # FEC_B200_LIB: an explicitly chosen build of the same library (A/B runs of kernel variants); still no fallback.
    path = os.environ.get("FEC_B200_LIB") or _build.build()
    if not os.path.exists(path):
        raise FecError("libfec_b200.so is missing: build the CUDA extension (__graft_entry__.build())")
    lib = C.CDLL(path)
    lib.fec_last_error.restype = C.c_char_p
    lib.fec_abi_version.restype = C.c_uint32
    lib.fec_engine_stream.restype = C.c_void_p
    lib.fec_engine_stream.argtypes = [C.c_void_p]
    lib.fec_harq_slot_ptr.restype = C.c_void_p
    lib.fec_harq_slot_ptr.argtypes = [C.c_void_p, C.c_uint32]
    lib.fec_engine_launch_count.restype = C.c_uint64
    lib.fec_engine_launch_count.argtypes = [C.c_void_p]
    lib.fec_host_alloc.restype = C.c_void_p
    lib.fec_host_alloc.argtypes = [C.c_size_t]
    lib.fec_host_free.argtypes = [C.c_void_p]
    lib.fec_engine_create.argtypes = [C.POINTER(EngineCfg), C.POINTER(C.c_void_p)]
    lib.fec_engine_destroy.argtypes = [C.c_void_p]
    lib.fec_desc_prepare.argtypes = [C.c_void_p]
    lib.fec_launch_device.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.c_uint32, C.c_int]
    lib.fec_run_host.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t,
                                 C.c_void_p, C.c_void_p]
    lib.fec_harq_free.argtypes = [C.c_void_p, C.c_uint32]
    lib.fec_harq_flush.argtypes = [C.c_void_p]
    lib.fec_harq_read.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32]
    lib.fec_harq_write.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32]
    lib.fec_queue_reserve.argtypes = [C.c_void_p, C.POINTER(C.c_uint32)]
    lib.fec_queue_free.argtypes = [C.c_void_p, C.c_uint32]
    lib.fec_queue_enqueue.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint64]
    lib.fec_queue_flush.argtypes = [C.c_void_p, C.c_uint32]
    lib.fec_queue_enqueue_batch.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32,
                                            C.POINTER(C.c_uint32)]
    lib.fec_queue_poll.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint32]
    lib.fec_queue_fetch.argtypes = [C.c_void_p, C.c_uint32, C.c_uint64, C.c_void_p, C.c_uint32]
    lib.fec_tb_segment.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint32), C.c_void_p]
    lib.fec_tb_run_host.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p,
                                    C.c_size_t, C.c_void_p, C.c_void_p]
    lib.fec_tb_reset_slots.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32]
    lib.fec_engine_get_stats.argtypes = [C.c_void_p, C.POINTER(EngineStats)]
    lib.fec_engine_debug_fail_launches.argtypes = [C.c_void_p, C.c_uint32]
    lib.fec_ulsch_demux_sizes.argtypes = [C.POINTER(UlschCfg), C.POINTER(UlschSizes)]
    lib.fec_ulsch_demux_launch_device.argtypes = [C.c_void_p, C.POINTER(UlschCfg)] + [C.c_void_p] * 6
    lib.fec_ulsch_demux_run_host.argtypes = [C.c_void_p, C.POINTER(UlschCfg)] + [C.c_void_p] * 6
    lib.fec_tb_enqueue.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p,
                                   C.c_size_t, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64)]
    lib.fec_tb_poll.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]
    lib.fec_tb_wait.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]
    lib.fec_tb_assemble_device_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p,
                                                   C.c_size_t, C.c_void_p]
    lib.fec_enc_desc_prepare.argtypes = [C.c_void_p]
    lib.fec_encode_launch_device.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint32]
    lib.fec_encode_run_host.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_size_t, C.c_void_p,
                                        C.c_size_t]
    lib.fec_harq_pool_create.argtypes = [C.c_void_p, C.POINTER(HarqPoolCfg), C.POINTER(C.c_void_p)]
    lib.fec_harq_pool_destroy.argtypes = [C.c_void_p]
    lib.fec_harq_pool_destroy.restype = None
    lib.fec_harq_pool_reserve.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, C.POINTER(C.c_uint32)]
    lib.fec_harq_pool_codeblocks.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32]
    for f in (lib.fec_harq_pool_unlock, lib.fec_harq_pool_release, lib.fec_harq_pool_run_slot):
        f.argtypes = [C.c_void_p, C.c_uint32]
    lib.fec_harq_pool_stop.argtypes = [C.c_void_p]
    lib.fec_harq_pool_nof_free_codeblocks.argtypes = [C.c_void_p]
    lib.fec_demod_desc_prepare.argtypes = [C.c_void_p]
    lib.fec_demod_launch_device.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p,
                                            C.c_void_p, C.c_uint32]
    lib.fec_demod_run_host.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p,
                                       C.c_size_t, C.c_uint32]
    lib.fec_tb_run_host_symbols.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint32,
                                            C.c_void_p, C.c_void_p, C.c_size_t, C.c_uint32, C.c_void_p, C.c_size_t,
                                            C.c_void_p, C.c_void_p]
    _lib = lib
    return lib


def _check(lib, rc):
    if rc < 0:
        raise FecError("fec error %d: %s" % (rc, lib.fec_last_error().decode()))
    return rc


def dims(bg, Z):
    kb, nb = (22, 66) if bg == 1 else (10, 50)
    return kb * Z, nb * Z


def make_descs(n):
    return np.zeros(n, DESC_DTYPE)


def make_enc_descs(n):
    return np.zeros(n, ENC_DESC_DTYPE)


def enc_desc_prepare(desc):
    """Validates encode descriptors in place (fills k0 / graph_index); raises FecError on invalid fields. Host only."""
    lib = load_library()
    arr = np.atleast_1d(desc)
    for i in range(arr.size):
        _check(lib, lib.fec_enc_desc_prepare(arr[i:i + 1].ctypes.data))
    return desc


def make_demod_descs(n):
    return np.zeros(n, DEMOD_DESC_DTYPE)


def demod_desc_prepare(descs):
    """Validates demodulation descriptors in place and fills the LFSR states; raises FecError on invalid fields. Host only."""
    lib = load_library()
    arr = np.atleast_1d(descs)
    for i in range(arr.size):
        _check(lib, lib.fec_demod_desc_prepare(arr[i:i + 1].ctypes.data))
    return descs


def codeword_demod_descs(blocks, qm, c_init, sym_offset=0, llr_offset=0, flags=0):
    """The demodulation calls of one PUSCH codeword as pusch_demodulator_impl issues them (pusch_demodulator_impl.cpp:
    262-400): one per block of symbols (one OFDM symbol each), the scrambling sequence running on across the blocks."""
    d = make_demod_descs(len(blocks))
    pos = 0
    for i, n in enumerate(blocks):
        d["sym_offset"][i], d["llr_offset"][i], d["nof_symbols"][i] = sym_offset + pos, llr_offset + pos * qm, n
        d["seq_offset"][i] = pos * qm
        pos += int(n)
    d["c_init"], d["qm"], d["flags"] = c_init, qm, flags
    return demod_desc_prepare(d)


def make_tb_descs(n):
    d = np.zeros(n, TB_DESC_DTYPE)
    d["slot_list"] = TB_SLOTS_CONTIGUOUS
    return d


def tb_segment(tb, slots=None):
    """fec_tb_segment: receive-side segmentation of ONE transport block descriptor, on the host (no GPU).
    Returns (codeblock descriptors, payload bits per codeblock)."""
    lib = load_library()
    tb = np.atleast_1d(tb)
    cbs = make_descs(MAX_TB_CBS)
    payload = np.zeros(MAX_TB_CBS, np.uint32)
    n = C.c_uint32()
    sl = None if slots is None else np.ascontiguousarray(slots, np.uint32).ctypes.data
    _check(lib, lib.fec_tb_segment(tb.ctypes.data, sl, cbs.ctypes.data, C.byref(n), payload.ctypes.data))
    return cbs[:n.value].copy(), payload[:n.value].copy()


def desc_prepare(desc):
    """Validates descriptors in place (fills k0 / graph_index); raises FecError on invalid fields. Host only."""
    lib = load_library()
    arr = np.atleast_1d(desc)
    for i in range(arr.size):
        _check(lib, lib.fec_desc_prepare(arr[i:i + 1].ctypes.data))
    return desc


class FecEngine:
    """One engine per GPU (fec_engine_create). All decode work happens in the CUDA kernels behind the C ABI."""

    def __init__(self, device=0, harq_slots=4096, max_batch_cbs=2048, max_batch_llrs=0, nof_queues=8,
                 scaling_factor=0.8, kernel_variant=0, host_contexts=0, queue_batch_cbs=0, queue_batches=0, max_wait_us=0):
        self.lib = load_library()
        cfg = EngineCfg()
        self.lib.fec_engine_default_cfg(C.byref(cfg))
        cfg.device = device
        cfg.harq_slots = harq_slots
        cfg.max_batch_cbs = max_batch_cbs
        cfg.max_batch_llrs = max_batch_llrs or min(max_batch_cbs * 25344, 0xFFFFFF00)
        cfg.nof_queues = nof_queues
        cfg.scaling_factor = scaling_factor
        cfg.kernel_variant = kernel_variant
        cfg.host_contexts, cfg.queue_batch_cbs, cfg.queue_batches = host_contexts, queue_batch_cbs, queue_batches
        cfg.max_wait_us = max_wait_us
        self.cfg = cfg
        self.h = C.c_void_p()
        _check(self.lib, self.lib.fec_engine_create(C.byref(cfg), C.byref(self.h)))

    def close(self):
        if getattr(self, "h", None):
            self.lib.fec_engine_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def stream(self):
        return self.lib.fec_engine_stream(self.h)

    @property
    def launch_count(self):
        return int(self.lib.fec_engine_launch_count(self.h))

    def harq_slot_ptr(self, slot):
        return self.lib.fec_harq_slot_ptr(self.h, slot)

    def launch_device(self, d_descs_ptr, n, d_llr_ptr, d_out_ptr, d_res_ptr, max_Z, any_bg1, even_z=False):
        """Asynchronous launch on self.stream; all pointers are device addresses (ints). even_z: every lifting size of the
        batch is even and >= 4 (FEC_LAUNCH_EVEN_Z: the packed kernel is launched alone)."""
        flags = (1 if any_bg1 else 0) | (LAUNCH_EVEN_Z if even_z else 0)
        _check(self.lib, self.lib.fec_launch_device(self.h, d_descs_ptr, n, d_llr_ptr, d_out_ptr, d_res_ptr, max_Z, flags))

    def run_host(self, descs, llr, out, results=None, soft_out=None):
        """Synchronous H2D -> kernels -> D2H over host numpy buffers (or raw pinned pointers given as ints)."""
        n = descs.size
        assert descs.dtype == DESC_DTYPE and descs.flags.c_contiguous
        if results is None:
            results = np.zeros(n, RESULT_DTYPE)

        def ptr(a):
            return a if isinstance(a, int) else (a.ctypes.data if a is not None else None)

        def size(a, given=None):
            return a.nbytes if hasattr(a, "nbytes") else (given or 0)

        _check(self.lib, self.lib.fec_run_host(self.h, descs.ctypes.data, n, ptr(llr), size(llr), ptr(out), size(out),
                                               results.ctypes.data, ptr(soft_out)))
        return results

    def encode_launch_device(self, d_descs_ptr, n, d_in_ptr, d_out_ptr, max_Z):
        """Asynchronous encode launch on self.stream; all pointers are device addresses (ints)."""
        _check(self.lib, self.lib.fec_encode_launch_device(self.h, d_descs_ptr, n, d_in_ptr, d_out_ptr, max_Z))

    def encode_run_host(self, descs, data, out):
        """fec_encode_run_host: LDPC encoding + rate matching of n codeblocks, H2D -> kernel -> D2H, synchronous.
        `data` / `out` are host numpy uint8 buffers (or (pointer, nbytes) tuples of pinned memory)."""
        assert descs.dtype == ENC_DESC_DTYPE and descs.flags.c_contiguous

        def ps(a):
            return (a[0], a[1]) if isinstance(a, tuple) else (a.ctypes.data, a.nbytes)

        (ip, ib), (op, ob) = ps(data), ps(out)
        _check(self.lib, self.lib.fec_encode_run_host(self.h, descs.ctypes.data, descs.size, ip, ib, op, ob))
        return out

    def tb_run_host(self, tbs, llr, out, slots=None, cb_results=False):
        """fec_tb_run_host: whole transport blocks (segmentation, codeblock decoding, assembly, TB CRC), synchronous.
        llr / out are host numpy buffers or raw pinned pointers given as (ptr, nbytes)."""
        assert tbs.dtype == TB_DESC_DTYPE and tbs.flags.c_contiguous
        n = tbs.size
        res = np.zeros(n, TB_RESULT_DTYPE)
        cbr = np.zeros(n * MAX_TB_CBS, RESULT_DTYPE) if cb_results else None

        def ptr(a):
            return a[0] if isinstance(a, tuple) else a.ctypes.data

        def size(a):
            return a[1] if isinstance(a, tuple) else a.nbytes

        sl = None if slots is None else np.ascontiguousarray(slots, np.uint32)
        _check(self.lib, self.lib.fec_tb_run_host(self.h, tbs.ctypes.data, n, None if sl is None else sl.ctypes.data,
                                                  ptr(llr), size(llr), ptr(out), size(out), res.ctypes.data,
                                                  None if cbr is None else cbr.ctypes.data))
        return (res, cbr.reshape(n, MAX_TB_CBS)) if cb_results else res

    def stats(self):
        """fec_engine_get_stats as a dict (iterations_hist as a list)."""
        st = EngineStats()
        _check(self.lib, self.lib.fec_engine_get_stats(self.h, C.byref(st)))
        out = {name: int(getattr(st, name)) for name, _ in EngineStats._fields_ if name != "iterations_hist"}
        out["iterations_hist"] = [int(x) for x in st.iterations_hist]
        return out

    def debug_fail_launches(self, n):
        _check(self.lib, self.lib.fec_engine_debug_fail_launches(self.h, n))

    def tb_enqueue(self, tbs, llr, out, slots=None, callback=None):
        """fec_tb_enqueue: asynchronous transport-block call. llr / out: numpy buffers or (ptr, nbytes) of pinned memory; they
        (and `callback`, a TB_CALLBACK instance) must stay alive until completion. Returns the ticket; raises FecError with
        code FEC_ERR_QUEUE_FULL (-5) in the message when every host-call context is in flight."""
        assert tbs.dtype == TB_DESC_DTYPE and tbs.flags.c_contiguous

        def ps(a):
            return (a[0], a[1]) if isinstance(a, tuple) else (a.ctypes.data, a.nbytes)

        (lp, lb), (op, ob) = ps(llr), ps(out)
        sl = None if slots is None else np.ascontiguousarray(slots, np.uint32)
        ticket = C.c_uint64()
        _check(self.lib, self.lib.fec_tb_enqueue(self.h, tbs.ctypes.data, tbs.size, None if sl is None else sl.ctypes.data,
                                                 lp, lb, op, ob, C.cast(callback, C.c_void_p) if callback else None, None,
                                                 C.byref(ticket)))
        return ticket.value

    def tb_poll(self, ticket, n, cb_results=False, wait=False):
        """fec_tb_poll / fec_tb_wait: None while the call is still running, else the results (and codeblock results)."""
        res = np.zeros(n, TB_RESULT_DTYPE)
        cbr = np.zeros(n * MAX_TB_CBS, RESULT_DTYPE) if cb_results else None
        f = self.lib.fec_tb_wait if wait else self.lib.fec_tb_poll
        rc = _check(self.lib, f(self.h, ticket, res.ctypes.data, None if cbr is None else cbr.ctypes.data))
        if not wait and rc == 0:
            return None
        return (res, cbr.reshape(n, MAX_TB_CBS)) if cb_results else res

    def tb_assemble_device_noalloc(self, tbs, d_cb_msgs_ptr, d_cb_results_ptr, d_out_ptr, out_bytes, d_tb_results_ptr):
        """fec_tb_assemble_device_noalloc: allocation-free TB assembly + TB CRC over device buffers (pointers as ints),
        asynchronous on self.stream."""
        assert tbs.dtype == TB_DESC_DTYPE and tbs.flags.c_contiguous
        _check(self.lib, self.lib.fec_tb_assemble_device_noalloc(self.h, tbs.ctypes.data, tbs.size, d_cb_msgs_ptr,
                                                                 d_cb_results_ptr, d_out_ptr, out_bytes, d_tb_results_ptr))

    def demod_launch_device(self, d_descs_ptr, n, max_symbols, d_sym_ptr, d_noise_ptr, d_llr_ptr, flavour=DEMOD_AVX512):
        """Asynchronous soft-demodulation launch on self.stream; all pointers are device addresses (ints)."""
        _check(self.lib, self.lib.fec_demod_launch_device(self.h, d_descs_ptr, n, max_symbols, d_sym_ptr, d_noise_ptr,
                                                          d_llr_ptr, flavour))

    def demod_run_host(self, descs, symbols, noise_vars, llr, flavour=DEMOD_AVX512):
        """fec_demod_run_host: soft demodulation + descrambling, H2D -> kernel -> D2H of the LLRs, synchronous."""
        assert descs.dtype == DEMOD_DESC_DTYPE and descs.flags.c_contiguous
        symbols = np.ascontiguousarray(symbols, np.float32).reshape(-1, 2)
        noise_vars = np.ascontiguousarray(noise_vars, np.float32).reshape(-1)
        assert symbols.shape[0] == noise_vars.size and llr.dtype == np.int8 and llr.flags.c_contiguous
        _check(self.lib, self.lib.fec_demod_run_host(self.h, descs.ctypes.data, descs.size, symbols.ctypes.data,
                                                     noise_vars.ctypes.data, noise_vars.size, llr.ctypes.data, llr.nbytes,
                                                     flavour))
        return llr

    def ulsch_demux_run_host(self, cfg, codeword, scrambling_bits=None):
        """fec_ulsch_demux_run_host: the soft bits of a codeword with UCI -> (sch, harq_ack, csi_part1, csi_part2) int8 arrays.
        scrambling_bits: the unpacked scrambling sequence of the codeword (0/1 per soft bit), packed MSB first here as the
        reference's bit_buffer holds it; needed only when a UCI field has a payload of one or two bits."""
        z = ulsch_demux_sizes(cfg)
        cw = np.ascontiguousarray(codeword, np.int8)
        if cw.size != z["nof_codeword_bits"]:
            raise FecError("codeword of %d soft bits, the allocation holds %d" % (cw.size, z["nof_codeword_bits"]))
        seq = None if scrambling_bits is None else np.packbits(np.ascontiguousarray(scrambling_bits, np.uint8))
        outs = [np.zeros(max(z[k], 1), np.int8) for k in ("nof_sch_bits", "nof_harq_ack_bits", "nof_csi_part1_bits",
                                                            "nof_csi_part2_bits")]
        _check(self.lib, self.lib.fec_ulsch_demux_run_host(self.h, C.byref(cfg), cw.ctypes.data,
                                                           None if seq is None else seq.ctypes.data,
                                                           *[o.ctypes.data for o in outs]))
        return [o[:z[k]] for o, k in zip(outs, ("nof_sch_bits", "nof_harq_ack_bits", "nof_csi_part1_bits", "nof_csi_part2_bits"))]

    def ulsch_demux_launch_device(self, cfg, d_codeword, d_scrambling_seq, d_sch, d_harq_ack, d_csi_part1, d_csi_part2):
        """Asynchronous on self.stream; device addresses (ints, 0 / None for streams that receive nothing)."""
        _check(self.lib, self.lib.fec_ulsch_demux_launch_device(self.h, C.byref(cfg), d_codeword, d_scrambling_seq or None,
                                                                d_sch or None, d_harq_ack or None, d_csi_part1 or None,
                                                                d_csi_part2 or None))

    def tb_run_host_symbols(self, tbs, demod, symbols, noise_vars, out, slots=None, cb_results=False,
                            flavour=DEMOD_AVX512):
        """fec_tb_run_host_symbols: equalised symbols in, transport blocks out; the LLRs only exist on the device.
        symbols / noise_vars are float32 numpy arrays or (ptr, nof_symbols) tuples of pinned memory."""
        assert tbs.dtype == TB_DESC_DTYPE and tbs.flags.c_contiguous
        assert demod.dtype == DEMOD_DESC_DTYPE and demod.flags.c_contiguous
        n = tbs.size
        res = np.zeros(n, TB_RESULT_DTYPE)
        cbr = np.zeros(n * MAX_TB_CBS, RESULT_DTYPE) if cb_results else None
        if isinstance(symbols, tuple):
            sp, nsym, np_ = symbols[0], symbols[1], noise_vars[0]
        else:
            symbols = np.ascontiguousarray(symbols, np.float32).reshape(-1, 2)
            noise_vars = np.ascontiguousarray(noise_vars, np.float32).reshape(-1)
            sp, nsym, np_ = symbols.ctypes.data, noise_vars.size, noise_vars.ctypes.data
        op, ob = (out[0], out[1]) if isinstance(out, tuple) else (out.ctypes.data, out.nbytes)
        sl = None if slots is None else np.ascontiguousarray(slots, np.uint32)
        _check(self.lib, self.lib.fec_tb_run_host_symbols(self.h, tbs.ctypes.data, n, None if sl is None else sl.ctypes.data,
                                                          demod.ctypes.data, demod.size, sp, np_, nsym, flavour, op, ob,
                                                          res.ctypes.data, None if cbr is None else cbr.ctypes.data))
        return (res, cbr.reshape(n, MAX_TB_CBS)) if cb_results else res

    def tb_reset_slots(self, first_slot, n):
        _check(self.lib, self.lib.fec_tb_reset_slots(self.h, first_slot, n))

    def harq_free(self, slot):
        _check(self.lib, self.lib.fec_harq_free(self.h, slot))

    def harq_flush(self):
        _check(self.lib, self.lib.fec_harq_flush(self.h))

    def harq_read(self, slot, n=HARQ_SLOT_BYTES):
        buf = np.zeros(n, np.int8)
        _check(self.lib, self.lib.fec_harq_read(self.h, slot, buf.ctypes.data, n))
        return buf

    def harq_write(self, slot, data):
        data = np.ascontiguousarray(data, np.int8)
        _check(self.lib, self.lib.fec_harq_write(self.h, slot, data.ctypes.data, data.size))

    # ---- queue interface ----
    def queue_reserve(self):
        q = C.c_uint32()
        _check(self.lib, self.lib.fec_queue_reserve(self.h, C.byref(q)))
        return q.value

    def queue_free(self, q):
        _check(self.lib, self.lib.fec_queue_free(self.h, q))

    def queue_enqueue(self, q, desc, llr, tag):
        """Returns False when the staging ring is full (caller retries), True when accepted."""
        llr = np.ascontiguousarray(llr, np.int8)
        rc = self.lib.fec_queue_enqueue(self.h, q, desc.ctypes.data, llr.ctypes.data, tag)
        if rc == -5:
            return False
        _check(self.lib, rc)
        return True

    def queue_enqueue_batch(self, q, descs, llrs, tags):
        """fec_queue_enqueue_batch: descs (DESC_DTYPE array), llrs (list of int8 arrays), tags. Returns how many were taken."""
        llrs = [np.ascontiguousarray(x, np.int8) for x in llrs]
        ptrs = (C.c_void_p * len(llrs))(*[x.ctypes.data for x in llrs])
        tg = np.ascontiguousarray(tags, np.uint64)
        acc = C.c_uint32()
        rc = self.lib.fec_queue_enqueue_batch(self.h, q, descs.ctypes.data, ptrs, tg.ctypes.data, len(llrs), C.byref(acc))
        if rc == -5:
            return 0
        _check(self.lib, rc)
        return acc.value

    def queue_flush(self, q):
        _check(self.lib, self.lib.fec_queue_flush(self.h, q))

    def queue_poll(self, q, max_ops=256):
        tags = np.zeros(max_ops, np.uint64)
        res = np.zeros(max_ops, RESULT_DTYPE)
        n = _check(self.lib, self.lib.fec_queue_poll(self.h, q, tags.ctypes.data, res.ctypes.data, max_ops))
        return tags[:n], res[:n]

    def queue_fetch(self, q, tag, dst):
        _check(self.lib, self.lib.fec_queue_fetch(self.h, q, int(tag), dst.ctypes.data, dst.size))


# ------------------------------------------------------------------------------------------------------------------
# Mirrors of the reference's operator interfaces for this path (same names, argument meaning, error behaviour).
# ------------------------------------------------------------------------------------------------------------------
class CodeblockMetadata:
    """codeblock_metadata (include/srsran/phy/upper/codeblock_metadata.h:42-80), flattened."""

    def __init__(self, base_graph=1, lifting_size=2, rv=0, mod_order=1, Nref=0, nof_filler_bits=0, nof_crc_bits=24,
                 rm_length=0):
        self.base_graph = base_graph
        self.lifting_size = lifting_size
        self.rv = rv
        self.mod_order = mod_order  # bits per symbol of tb_common.mod
        self.Nref = Nref
        self.nof_filler_bits = nof_filler_bits
        self.nof_crc_bits = nof_crc_bits
        self.rm_length = rm_length


class HarqPool:
    """rx_buffer_pool over the engine's HARQ slots (include/srsran/phy/upper/rx_buffer_pool.h:60-127): reserve returns
    (status, buffer); a reserved buffer is locked until unlock() or release(). `engine` may be None (bookkeeping only)."""

    def __init__(self, engine, nof_buffers, nof_codeblocks, expire_timeout_slots, first_slot=0, numerology=0):
        self.lib = load_library()
        cfg = HarqPoolCfg(nof_buffers, first_slot, nof_codeblocks, expire_timeout_slots, numerology)
        self.h = C.c_void_p()
        _check(self.lib, self.lib.fec_harq_pool_create(None if engine is None else engine.h, C.byref(cfg), C.byref(self.h)))

    def close(self):
        if self.h:
            self.lib.fec_harq_pool_destroy(self.h)
            self.h = None

    def reserve(self, slot, identifier, nof_codeblocks, new_data):
        buf = C.c_uint32()
        st = _check(self.lib, self.lib.fec_harq_pool_reserve(self.h, slot, identifier, nof_codeblocks, int(new_data), C.byref(buf)))
        return st, (buf.value if st == POOL_OK else None)

    def codeblocks(self, buffer):
        ids = np.zeros(MAX_TB_CBS, np.uint32)
        n = _check(self.lib, self.lib.fec_harq_pool_codeblocks(self.h, buffer, ids.ctypes.data, ids.size))
        return ids[:n].copy()

    def unlock(self, buffer):
        _check(self.lib, self.lib.fec_harq_pool_unlock(self.h, buffer))

    def release(self, buffer):
        _check(self.lib, self.lib.fec_harq_pool_release(self.h, buffer))

    def run_slot(self, slot):
        _check(self.lib, self.lib.fec_harq_pool_run_slot(self.h, slot))

    def stop(self):
        _check(self.lib, self.lib.fec_harq_pool_stop(self.h))

    def nof_free_codeblocks(self):
        return _check(self.lib, self.lib.fec_harq_pool_nof_free_codeblocks(self.h))


class DemodulationMapper:
    """demodulation_mapper (include/srsran/phy/upper/channel_modulation/demodulation_mapper.h:47-69):
    demodulate_soft(llrs, symbols, noise_vars, mod). `mod` is the modulation's name as the reference prints it
    ("BPSK", "pi/2-BPSK", "QPSK", "16QAM", "64QAM", "256QAM"); `flavour` the reference build to be identical to."""

    BITS = {"BPSK": 1, "pi/2-BPSK": 1, "QPSK": 2, "16QAM": 4, "64QAM": 6, "256QAM": 8}

    def __init__(self, engine, flavour=DEMOD_AVX512):
        self.engine = engine
        self.flavour = flavour

    def demodulate_soft(self, llrs, symbols, noise_vars, mod):
        qm = self.BITS[mod]
        symbols = np.ascontiguousarray(symbols).view(np.float32).reshape(-1, 2)
        if symbols.shape[0] != np.size(noise_vars):
            raise FecError("Inputs symbols and noise_vars must have the same length.")
        if symbols.shape[0] * qm != llrs.size:
            raise FecError("Input and output lengths are incompatible.")
        d = make_demod_descs(1)
        d["nof_symbols"], d["qm"] = symbols.shape[0], qm
        d["flags"] = DEMOD_FLAG_NO_DESCRAMBLE | (DEMOD_FLAG_PI2_BPSK if mod == "pi/2-BPSK" else 0)
        return self.engine.demod_run_host(d, symbols, noise_vars, llrs, self.flavour)


class LdpcRateDematcher:
    """ldpc_rate_dematcher (ldpc_rate_dematcher.h:35-58): rate_dematch(output, input, new_data, cfg), in place."""

    SCRATCH_SLOT = 0

    def __init__(self, engine):
        self.engine = engine

    def rate_dematch(self, output, input, new_data, cfg):
        K, N = dims(cfg.base_graph, cfg.lifting_size)
        if output.size != N or output.dtype != np.int8:
            raise FecError("rate_dematch: output must be the full %d-LLR int8 codeblock" % N)
        d = make_descs(1)
        d["base_graph"], d["lifting_size"], d["rv"], d["qm"] = cfg.base_graph, cfg.lifting_size, cfg.rv, cfg.mod_order
        d["Nref"], d["nof_filler_bits"], d["E"] = cfg.Nref, cfg.nof_filler_bits, input.size
        d["crc_type"], d["max_iter"], d["mode"] = CRC24B, 1, MODE_DEMATCH
        d["flags"] = FLAG_NEW_DATA if new_data else 0
        d["harq_slot"] = self.SCRATCH_SLOT
        self.engine.harq_write(self.SCRATCH_SLOT, output)
        llr = np.ascontiguousarray(input, np.int8)
        self.engine.run_host(d, llr, None)
        output[:] = self.engine.harq_read(self.SCRATCH_SLOT, N)
        return output


class LdpcDecoder:
    """ldpc_decoder (ldpc_decoder.h:37-75): decode(output, input, crc, cfg) -> iterations or None.

    `crc` conveys only the polynomial (CRC16 / CRC24A / CRC24B or None); the check itself runs on the device.
    `output` is a uint8 array of ceil(K/8) bytes, MSB-first packed like bit_buffer."""

    def __init__(self, engine):
        self.engine = engine

    def decode(self, output, input, crc, cfg, max_iterations=6):
        K, N = dims(cfg.base_graph, cfg.lifting_size)
        if output.size != (K + 7) // 8:
            raise FecError("decode: output size %d does not match the message length %d bits" % (output.size, K))
        d = make_descs(1)
        d["base_graph"], d["lifting_size"] = cfg.base_graph, cfg.lifting_size
        d["nof_filler_bits"], d["E"] = cfg.nof_filler_bits, input.size
        d["max_iter"], d["mode"] = max_iterations, MODE_DECODE
        if crc is None:
            d["crc_type"], d["flags"] = CRC16, FLAG_NO_CRC
        else:
            d["crc_type"], d["flags"] = crc, FLAG_EARLY_STOP
        llr = np.ascontiguousarray(input, np.int8)
        res = self.engine.run_host(d, llr, output)
        if crc is not None and (res["status"][0] & RES_CRC_PASS):
            return int(res["nof_ldpc_iterations"][0])
        return None


class HwPuschDecoderConfiguration:
    """hal::hw_pusch_decoder_configuration (hw_accelerator_pusch_dec.h:39-72)."""

    def __init__(self, **kw):
        self.base_graph_index = 1
        self.modulation = 2
        self.nof_segments = 1
        self.rv = 0
        self.cw_length = 0
        self.lifting_size = 2
        self.Ncb = 0
        self.Nref = 0
        self.nof_segment_bits = 0
        self.nof_filler_bits = 0
        self.max_nof_ldpc_iterations = 6
        self.use_early_stop = True
        self.new_data = True
        self.cb_crc_len = 24
        self.cb_crc_type = CRC24B
        self.absolute_cb_id = 0
        self.__dict__.update(kw)


class HwPuschDecoderOutputs:
    def __init__(self):
        self.CRC_pass = False
        self.nof_ldpc_iterations = 0


class HwAcceleratorPuschDec:
    """hal::hw_accelerator_pusch_dec (hw_accelerator_pusch_dec.h:83-115) over the queue entry points of the C ABI.

    Same call sequence as pusch_decoder_hw_impl drives (lib/phy/upper/channel_processors/pusch/pusch_decoder_hw_impl.cpp:149-358):
    reserve_queue, {configure_operation, enqueue_operation}*, {dequeue_operation (polled), read_operation_outputs}*,
    free_harq_context_entry*, free_queue. HARQ is external: soft buffers stay in HBM keyed by absolute_cb_id."""

    MAX_NOF_SEGMENTS = 162

    def __init__(self, engine):
        self.engine = engine
        self.queue = None
        self.cfgs = {}
        self.done = {}

    def reserve_queue(self):
        if self.queue is None:
            self.queue = self.engine.queue_reserve()

    def free_queue(self):
        if self.queue is not None:
            self.engine.queue_free(self.queue)
            self.queue = None
        self.cfgs.clear()
        self.done.clear()

    def configure_operation(self, config, cb_index=0):
        self.cfgs[cb_index] = config

    def enqueue_operation(self, data, aux_data=None, cb_index=0):
        c = self.cfgs[cb_index]
        d = make_descs(1)
        d["base_graph"], d["lifting_size"], d["rv"], d["qm"] = c.base_graph_index, c.lifting_size, c.rv, c.modulation
        d["Nref"], d["nof_filler_bits"], d["E"] = c.Nref, c.nof_filler_bits, c.cw_length
        d["crc_type"], d["max_iter"], d["mode"] = c.cb_crc_type, c.max_nof_ldpc_iterations, MODE_CHAIN
        d["flags"] = (FLAG_NEW_DATA if c.new_data else 0) | (FLAG_EARLY_STOP if c.use_early_stop else 0)
        d["harq_slot"] = c.absolute_cb_id
        if len(data) != c.cw_length:
            raise FecError("enqueue_operation: %d LLRs given, cw_length is %d" % (len(data), c.cw_length))
        return self.engine.queue_enqueue(self.queue, d, data, cb_index)

    def dequeue_operation(self, data, aux_data=None, segment_index=0):
        """Non-blocking: False until the operation has completed; then fills `data` with the decoded bytes."""
        if segment_index not in self.done:
            self.engine.queue_flush(self.queue)
            tags, res = self.engine.queue_poll(self.queue)
            for t, r in zip(tags, res):
                self.done[int(t)] = r.copy()
        if segment_index not in self.done:
            return False
        self.engine.queue_fetch(self.queue, segment_index, data)
        return True

    def read_operation_outputs(self, out, cb_index=0, absolute_cb_id=0):
        r = self.done.pop(cb_index)
        out.CRC_pass = bool(r["status"] & RES_CRC_PASS)
        out.nof_ldpc_iterations = int(r["nof_ldpc_iterations"])

    def free_harq_context_entry(self, absolute_cb_id):
        self.engine.harq_free(absolute_cb_id)

    def is_harq_external(self):
        return True


class HwPdschEncoderConfiguration:
    """hal::hw_pdsch_encoder_configuration (hw_accelerator_pdsch_enc.h:37-78), the fields the CB mode uses."""

    def __init__(self, **kw):
        self.base_graph_index = 1
        self.modulation = 2          # bits per symbol
        self.rv = 0
        self.lifting_size = 2
        self.Ncb = 0
        self.Nref = 0
        self.nof_segment_bits = 0
        self.nof_filler_bits = 0
        self.rm_length = 0
        self.cb_mode = True
        self.attach_cb_crc = False
        self.attach_tb_crc = False
        self.do_unpack = True
        for k, v in kw.items():
            if not hasattr(self, k):
                raise AttributeError(k)
            setattr(self, k, v)


class HwAcceleratorPdschEnc:
    """hal::hw_accelerator_pdsch_enc (hw_accelerator_pdsch_enc.h:81-106) in CB mode, as pdsch_encoder_hw_impl drives it
    (lib/phy/upper/channel_processors/pdsch/pdsch_encoder_hw_impl.cpp:31-186): reserve_queue, {configure_operation,
    enqueue_operation(packed segment)}*, {dequeue_operation(codeblock, codeblock_packed)}*, free_queue. Enqueued
    operations are batched and encoded in one launch at the first dequeue."""

    def __init__(self, engine, max_ops=162):
        self.engine = engine
        self.max_ops = max_ops
        self.cfgs = {}
        self.pending = []   # (cb_index, config, packed bytes)
        self.done = {}

    def reserve_queue(self):
        pass

    def free_queue(self):
        self.cfgs.clear()
        self.pending.clear()
        self.done.clear()

    def is_cb_mode_supported(self):
        return True

    def get_max_supported_buff_size(self):
        return MAX_CB_IN_BYTES

    def configure_operation(self, config, cb_index=0):
        if not config.cb_mode or config.attach_cb_crc or config.attach_tb_crc:
            raise FecError("configure_operation: only CB mode without CRC attachment is served (as pdsch_encoder_hw_impl "
                           "configures it when is_cb_mode_supported())")
        self.cfgs[cb_index] = config

    def enqueue_operation(self, data, aux_data=None, cb_index=0):
        if len(self.pending) >= self.max_ops:
            return False
        c = self.cfgs[cb_index]
        kb = 22 if c.base_graph_index == 1 else 10
        need = (kb * c.lifting_size - c.nof_filler_bits + 7) // 8
        if len(data) != need:
            raise FecError("enqueue_operation: %d bytes given, the segment has %d" % (len(data), need))
        self.pending.append((cb_index, c, np.array(data, np.uint8, copy=True)))
        return True

    def _run(self):
        n = len(self.pending)
        d = make_enc_descs(n)
        ins, in_off, out_off = [], 0, 0
        for i, (idx, c, seg) in enumerate(self.pending):
            d[i]["base_graph"], d[i]["lifting_size"], d[i]["rv"], d[i]["qm"] = c.base_graph_index, c.lifting_size, c.rv, c.modulation
            d[i]["Nref"], d[i]["nof_filler_bits"], d[i]["E"] = c.Nref, c.nof_filler_bits, c.rm_length
            d[i]["in_offset"], d[i]["out_offset"] = in_off, out_off
            ins.append(seg)
            pad = (-len(seg)) % 4
            if pad:
                ins.append(np.zeros(pad, np.uint8))
            in_off += len(seg) + pad
            out_off += ((c.rm_length + 7) // 8 + 3) & ~3
        out = np.zeros(out_off, np.uint8)
        self.engine.encode_run_host(d, np.concatenate(ins), out)
        for i, (idx, c, seg) in enumerate(self.pending):
            o = int(d[i]["out_offset"])
            self.done[idx] = (out[o:o + (c.rm_length + 7) // 8].copy(), c)
        self.pending.clear()

    def dequeue_operation(self, data, aux_data=None, segment_index=0):
        """Fills aux_data (codeblock_packed) with the packed rate-matched bits and, if do_unpack, data with one bit per byte."""
        if segment_index not in self.done:
            if not any(p[0] == segment_index for p in self.pending):
                return False
            self._run()
        packed, c = self.done.pop(segment_index)
        if aux_data is not None:
            aux_data[:len(packed)] = packed
        if c.do_unpack and data is not None:
            data[:c.rm_length] = np.unpackbits(packed)[:c.rm_length]
        return True


class LdpcEncoderRateMatcher:
    """ldpc_encoder::encode followed by ldpc_rate_matcher::rate_match for one codeblock
    (lib/phy/upper/channel_coding/ldpc/ldpc_encoder_impl.cpp:64-120, ldpc_rate_matcher_impl.cpp:36-160), as
    pdsch_encoder_impl chains them (lib/phy/upper/channel_processors/pdsch/pdsch_encoder_impl.cpp)."""

    def __init__(self, engine):
        self.engine = engine

    def encode_rate_match(self, message_bits, rm_length, cfg):
        """message_bits: K unpacked bits with the filler positions at the end (their values are ignored);
        cfg: CodeblockMetadata. Returns rm_length unpacked bits."""
        K, _ = dims(cfg.base_graph, cfg.lifting_size)
        kp = K - cfg.nof_filler_bits
        d = make_enc_descs(1)
        d["base_graph"], d["lifting_size"], d["rv"], d["qm"] = cfg.base_graph, cfg.lifting_size, cfg.rv, cfg.mod_order
        d["Nref"], d["nof_filler_bits"], d["E"], d["flags"] = cfg.Nref, cfg.nof_filler_bits, rm_length, ENC_FLAG_UNPACKED
        seg = np.packbits(np.asarray(message_bits[:kp], np.uint8) & 1)
        out = np.zeros(rm_length, np.uint8)
        self.engine.encode_run_host(d, seg, out)
        return out


class PuschDecoderConfiguration:
    """pusch_decoder::configuration (include/srsran/phy/upper/channel_processors/pusch/pusch_decoder.h)."""

    def __init__(self, base_graph=1, rv=0, mod_order=2, Nref=0, nof_layers=1, nof_ldpc_iterations=6, use_early_stop=True,
                 new_data=True):
        self.base_graph = base_graph
        self.rv = rv
        self.mod_order = mod_order  # bits per symbol of `mod`
        self.Nref = Nref
        self.nof_layers = nof_layers
        self.nof_ldpc_iterations = nof_ldpc_iterations
        self.use_early_stop = use_early_stop
        self.new_data = new_data


class PuschDecoderResult:
    """pusch_decoder_result: tb_crc_ok, nof_codeblocks_total, ldpc_decoder_stats (min / max / mean iterations)."""

    def __init__(self, r):
        self.tb_crc_ok = bool(r["tb_crc_ok"])
        self.nof_codeblocks_total = int(r["nof_codeblocks"])
        n = int(r["nof_cb_decoded"])
        self.ldpc_decoder_stats = {"nof_observations": n, "min": int(r["iter_min"]), "max": int(r["iter_max"]),
                                   "mean": (int(r["iter_sum"]) / n) if n else 0.0}


class PuschDecoder:
    """pusch_decoder + pusch_decoder_buffer (pusch_decoder.h, pusch_decoder_buffer.h) over the transport-block entry
    point: new_data(transport_block, first_absolute_cb_id | list of ids, notifier, cfg) -> self (the buffer);
    on_new_softbits(llrs)*; on_end_softbits() -> notifier(PuschDecoderResult). The rx buffer is represented by the
    absolute codeblock ids it would hand out (rx_buffer::get_absolute_codeblock_id); soft bits, codeblock CRC flags and
    decoded messages of earlier transmissions stay in the engine."""

    def __init__(self, engine):
        self.engine = engine

    def new_data(self, transport_block, absolute_cb_ids, notifier, cfg):
        self.tb, self.ids, self.notifier, self.cfg, self.soft = transport_block, absolute_cb_ids, notifier, cfg, []
        return self

    def on_new_softbits(self, softbits):
        self.soft.append(np.ascontiguousarray(softbits, np.int8))

    def on_end_softbits(self):
        llr = np.concatenate(self.soft) if len(self.soft) != 1 else self.soft[0]
        c = self.cfg
        d = make_tb_descs(1)
        d["tbs"], d["nof_llrs"], d["Nref"] = self.tb.size * 8, llr.size, c.Nref
        d["base_graph"], d["rv"], d["qm"], d["nof_layers"] = c.base_graph, c.rv, c.mod_order, c.nof_layers
        d["max_iter"] = c.nof_ldpc_iterations
        d["flags"] = (FLAG_NEW_DATA if c.new_data else 0) | (FLAG_EARLY_STOP if c.use_early_stop else 0)
        slots = None
        if np.isscalar(self.ids):
            d["harq_slot"] = self.ids
        else:
            d["slot_list"], slots = 0, np.asarray(self.ids, np.uint32)
        r = self.engine.tb_run_host(d, llr, self.tb, slots)
        self.notifier(PuschDecoderResult(r[0]))


# ------------------------------------------------------------------------------------------------------------------
# Multi-GPU sharding (one process and one engine per GPU). Codeblocks are independent: there is no data-path collective.
# The only constraint is HARQ affinity -- every (re)transmission of a codeblock must reach the GPU that holds its soft
# buffer -- so work is routed by cell (BASELINE config 5) or by absolute codeblock id range, never round-robin.
# ------------------------------------------------------------------------------------------------------------------
def cells_of_rank(n_cells, world_size, rank):
    """Cells served by `rank`: cell_id mod world_size == rank."""
    return [c for c in range(n_cells) if c % world_size == rank]


def harq_owner(absolute_cb_id, slots_per_gpu, world_size):
    """(rank, local slot) owning a soft buffer: each GPU owns a contiguous slice of the rx-buffer codeblock pool
    (lib/phy/upper/rx_buffer_impl.h:170-177 hands out absolute ids from one global pool)."""
    rank = absolute_cb_id // slots_per_gpu
    if rank >= world_size:
        raise FecError("absolute_cb_id %d outside the %d x %d soft-buffer slots" % (absolute_cb_id, world_size,
                                                                                    slots_per_gpu))
    return rank, absolute_cb_id - rank * slots_per_gpu


def shard_transport_blocks(tbs, world_size, rank):
    """tbs: iterable of dicts with 'cell' and 'cb_ids' (absolute ids). Returns the TBs this rank decodes, with cb ids
    rewritten to local HARQ slots; asserts that a TB never straddles two GPUs."""
    mine = []
    for tb in tbs:
        if tb["cell"] % world_size != rank:
            continue
        mine.append(dict(tb))
    return mine

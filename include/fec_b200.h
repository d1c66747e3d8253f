/* include/fec_b200.h -- C ABI of the B200 PUSCH FEC decode engine (rate dematch + HARQ combine -> layered scaled
 * min-sum LDPC -> CB CRC), the drop-in boundary for the reference's hardware-accelerator plug-in point.
 *
 * Plain C types only: no STL, no exceptions, no torch types. Every entry point returns 0 (FEC_OK) or a negative
 * fec_status; fec_last_error() gives a thread-local message. There is NO CPU fallback behind this interface: without
 * a CUDA device fec_engine_create() fails with FEC_ERR_NO_DEVICE.
 *
 * Reference interfaces replaced / served (paths relative to the reference tree):
 *   hal::hw_accelerator_pusch_dec         include/srsran/hal/phy/upper/channel_processors/pusch/hw_accelerator_pusch_dec.h:83-115
 *     enqueue_operation / dequeue_operation  include/srsran/hal/hw_accelerator.h:48,56        -> fec_queue_enqueue / fec_queue_poll
 *     configure_operation                    hw_accelerator_pusch_dec.h:98                     -> fec_cb_desc (one per CB)
 *     read_operation_outputs                 hw_accelerator_pusch_dec.h:105-106                -> fec_cb_result
 *     free_harq_context_entry                hw_accelerator_pusch_dec.h:110                    -> fec_harq_free
 *     reserve_queue / free_queue             hw_accelerator_pusch_dec.h:90,93                  -> fec_queue_reserve / fec_queue_free
 *   hal::hw_pusch_decoder_configuration   hw_accelerator_pusch_dec.h:39-72                     -> fec_cb_desc
 *   hal::hw_pusch_decoder_outputs         hw_accelerator_pusch_dec.h:75-80                     -> fec_cb_result
 *   ldpc_decoder::decode                  include/srsran/phy/upper/channel_coding/ldpc/ldpc_decoder.h:73-74      -> FEC_MODE_DECODE
 *   ldpc_rate_dematcher::rate_dematch     include/srsran/phy/upper/channel_coding/ldpc/ldpc_rate_dematcher.h:52-57 -> FEC_MODE_DEMATCH
 *   pusch_codeblock_decoder::decode       lib/phy/upper/channel_processors/pusch/pusch_codeblock_decoder.cpp:34-71  -> FEC_MODE_CHAIN
 *   hal::hw_accelerator_pdsch_enc (downlink twin)
 *                                         include/srsran/hal/phy/upper/channel_processors/hw_accelerator_pdsch_enc.h:81-106
 *     configure_operation (CB mode)          hw_accelerator_pdsch_enc.h:95                      -> fec_enc_desc (one per CB)
 *     enqueue_operation / dequeue_operation  include/srsran/hal/hw_accelerator.h:48,56         -> fec_encode_run_host
 *   hal::hw_pdsch_encoder_configuration   hw_accelerator_pdsch_enc.h:37-78                      -> fec_enc_desc
 *   ldpc_encoder::encode + ldpc_rate_matcher::rate_match
 *                                         lib/phy/upper/channel_coding/ldpc/ldpc_encoder_impl.cpp, ldpc_rate_matcher_impl.cpp:36-160
 *                                                                                               -> fec_encode_launch_device
 * The reference-side bindings (C++ adapter classes over this ABI) are shown in INTEGRATION.md and live in
 * srsran-5g_b200/host/.
 */
#ifndef FEC_B200_H
#define FEC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FEC_B200_ABI_VERSION 9

typedef enum {
  FEC_OK             = 0,
  FEC_ERR_NO_DEVICE  = -1, /* no CUDA device / driver: the engine never falls back to the CPU */
  FEC_ERR_BAD_ARG    = -2,
  FEC_ERR_CUDA       = -3,
  FEC_ERR_NO_MEMORY  = -4,
  FEC_ERR_QUEUE_FULL = -5,
  FEC_ERR_BAD_QUEUE  = -6
} fec_status;

/* CB CRC selector: same numbering as hal::hw_dec_cb_crc_type (hw_accelerator_pusch_dec.h:36). */
enum { FEC_CRC16 = 0, FEC_CRC24B = 1, FEC_CRC24A = 2 };

/* What one descriptor asks the engine to do. */
enum {
  FEC_MODE_CHAIN   = 0, /* rate dematch (+HARQ combine) into the soft-buffer slot, then LDPC decode + CRC from it */
  FEC_MODE_DECODE  = 1, /* ldpc_decoder::decode on n_in LLRs given directly (no soft buffer involved)            */
  FEC_MODE_DEMATCH = 2  /* ldpc_rate_dematcher::rate_dematch only (soft-buffer slot updated, nothing decoded)    */
};

/* fec_cb_desc.flags */
#define FEC_FLAG_NEW_DATA   0x01u /* first HARQ transmission: copy instead of combine (hw cfg .new_data)             */
#define FEC_FLAG_EARLY_STOP 0x02u /* check CRC after every iteration (hw cfg .use_early_stop; crc != nullptr)        */
#define FEC_FLAG_NO_CRC     0x04u /* FEC_MODE_DECODE with crc == nullptr: never report crc_pass                      */
#define FEC_FLAG_HARQ_ZERO  0x08u /* read the soft-buffer slot as all zeros (slot was freed; set by the engine for   */
                                  /* host-driven batches, by the caller for fec_launch_device)                        */

/* One codeblock operation: POD mirror of hal::hw_pusch_decoder_configuration + buffer placement. 48 bytes.
 * The same struct is read by the host engine and by the kernels. */
typedef struct fec_cb_desc {
  uint64_t llr_offset;      /* byte offset of this CB's input LLRs inside the batch LLR buffer                        */
  uint32_t out_offset;      /* byte offset of this CB's ceil(K/8) decoded bytes inside the batch output buffer        */
  uint32_t harq_slot;       /* soft-buffer slot in HBM (hw cfg .absolute_cb_id); ignored by FEC_MODE_DECODE            */
  uint32_t E;               /* rate-matched length (hw cfg .cw_length); FEC_MODE_DECODE: number of input LLRs n_in    */
  uint32_t Nref;            /* limited-buffer rate matching length, 0 = unlimited (hw cfg .Nref)                       */
  uint32_t nof_filler_bits; /* hw cfg .nof_filler_bits                                                                */
  uint32_t k0;              /* filled in by the engine (starting position in the circular buffer); callers pass 0     */
  uint16_t lifting_size;    /* Z (hw cfg .lifting_size)                                                               */
  uint16_t graph_index;     /* filled in by the engine: (bg-1)*51 + index of Z among the 51 lifting sizes             */
  uint8_t  base_graph;      /* 1 or 2 (hw cfg .base_graph_index)                                                      */
  uint8_t  rv;              /* redundancy version 0..3                                                                */
  uint8_t  qm;              /* bits per modulation symbol 1,2,4,6,8 (hw cfg .modulation)                              */
  uint8_t  crc_type;        /* FEC_CRC16 / FEC_CRC24B / FEC_CRC24A (hw cfg .cb_crc_type)                              */
  uint8_t  max_iter;        /* hw cfg .max_nof_ldpc_iterations (1..255)                                               */
  uint8_t  mode;            /* FEC_MODE_*                                                                             */
  uint8_t  flags;           /* FEC_FLAG_*                                                                             */
  uint8_t  reserved0;
  uint32_t reserved1;       /* bits 32..63 of out_offset (0 for output buffers below 4 GiB; the engine's batcher uses it)  */
} fec_cb_desc;

/* fec_cb_result.status bits */
#define FEC_RES_CRC_PASS  0x01u /* hw outputs .CRC_pass                                                              */
#define FEC_RES_TOO_SHORT 0x02u /* fewer than K significant LLRs: decoder returned nullopt without iterating; with     */
                                /* FEC_FLAG_EARLY_STOP the output bytes were left untouched, otherwise set to all ones */
#define FEC_RES_SKIPPED   0x04u /* transport-block interface: not decoded in this call, its CRC passed in an earlier one  */
#define FEC_RES_DONE      0x80u /* written last by the kernel                                                         */

/* Per-CB outcome: mirror of hal::hw_pusch_decoder_outputs. 8 bytes. */
typedef struct fec_cb_result {
  uint16_t nof_ldpc_iterations; /* iterations run: it+1 on early stop, else max_iter                                   */
  uint8_t  status;              /* FEC_RES_*                                                                          */
  uint8_t  nof_layers;          /* base-graph rows actually processed per iteration (trimmed codeblock length)          */
  uint32_t reserved;
} fec_cb_result;

typedef struct fec_engine_cfg {
  int32_t  device;          /* CUDA device ordinal                                                                    */
  uint32_t harq_slots;      /* number of soft-buffer slots to reserve in HBM (each FEC_HARQ_SLOT_BYTES)               */
  uint32_t max_batch_cbs;   /* capacity of one launch (descriptor / result arrays, queue staging)                     */
  uint32_t max_batch_llrs;  /* staging capacity in LLRs (bytes) for the host-buffer entry points                      */
  uint32_t nof_queues;      /* number of independent submission queues (one per hw_accelerator_pusch_dec instance)    */
  float    scaling_factor;  /* normalised min-sum factor, 0 < sf < 1 (reference default 0.8)                          */
  uint32_t kernel_variant;  /* 0 = auto; 1 = force the generic one-thread-per-check kernel (debug / A-B comparison)   */
  uint32_t reserved;
  uint32_t host_contexts;   /* host-buffer calls (fec_run_host, fec_tb_run_host) that run concurrently, each on its own  */
                            /* streams and staging = asynchronous transport-block calls in flight; 0 = 4               */
  uint32_t queue_batch_cbs; /* codeblocks per queue batch (a batch is dispatched when it is full or on flush);          */
                            /* 0 = max_batch_cbs. Small values start the GPU while the consumer is still enqueuing.     */
  uint32_t queue_batches;   /* batches a queue can have staged / in flight (ring depth); 0 = 4                          */
  uint32_t max_wait_us;     /* cross-queue batcher: a dispatch that holds less than one kernel wave of codeblocks waits  */
                            /* up to this long for batches of other queues to join its launch; 0 = never wait           */
} fec_engine_cfg;

#define FEC_HARQ_SLOT_BYTES 25344u /* 66 * 384, the largest codeblock; slots are contiguous in HBM                    */
#define FEC_MAX_CB_OUT_BYTES 1056u /* 22 * 384 / 8                                                                    */

typedef struct fec_engine fec_engine;

/* ---- lifecycle ---------------------------------------------------------------------------------------------------- */
uint32_t    fec_abi_version(void);
const char* fec_last_error(void);
void        fec_engine_default_cfg(fec_engine_cfg* cfg);
int         fec_engine_create(const fec_engine_cfg* cfg, fec_engine** out);
void        fec_engine_destroy(fec_engine* e);
/* The CUDA stream (cudaStream_t) every launch of this engine is issued on, for event timing by the caller. */
void*       fec_engine_stream(fec_engine* e);
/* Device pointer to soft-buffer slot `slot` (FEC_HARQ_SLOT_BYTES bytes), NULL if out of range. */
void*       fec_harq_slot_ptr(fec_engine* e, uint32_t slot);
/* Number of kernels this engine has launched so far (bench.py's gpu_launches). */
uint64_t    fec_engine_launch_count(fec_engine* e);

/* Engine counters since creation (the reference's per-decoder metrics -- include/srsran/phy/metrics/phy_metrics_reports.h:70-79,
 * lib/phy/metrics/phy_metrics_ldpc_decoder_decorator.h:34-71 -- kept inside the engine, plus what only a batching engine has). */
#define FEC_STATS_ITER_BINS 16
typedef struct fec_engine_stats {
  uint64_t launches;         /* kernels of any kind                                                                     */
  uint64_t cb_launches;      /* launches of the codeblock kernel                                                        */
  uint64_t codeblocks;       /* codeblock operations launched; codeblocks / cb_launches = mean batch occupancy          */
  uint64_t results;          /* codeblock results harvested by the host entry points                                   */
  uint64_t crc_pass;         /* of those, CRC passed                                                                    */
  uint64_t iterations_hist[FEC_STATS_ITER_BINS]; /* their nof_ldpc_iterations (last bin: that many or more)              */
  uint64_t queue_full;       /* enqueue calls refused with FEC_ERR_QUEUE_FULL                                           */
  uint64_t max_queue_depth;  /* largest number of operations staged or in flight in one queue                           */
  uint64_t dispatches;       /* cross-queue batcher: launches, and the queue batches they carried                       */
  uint64_t dispatch_batches;
  uint64_t tb_calls;         /* transport-block calls, blocks in them, blocks whose CRC passed                          */
  uint64_t transport_blocks;
  uint64_t tb_crc_ok;
  uint64_t faults;           /* CUDA failures turned into failed codeblocks / blocks                                    */
  uint64_t generic_launches; /* of cb_launches, those that took the generic kernel (kernel_variant 1, or a scaling_factor   */
                             /* whose rounding the packed kernel's 8-bit multiply-add cannot reproduce, e.g. 0.9)          */
} fec_engine_stats;
int fec_engine_get_stats(fec_engine* e, fec_engine_stats* out);
/* Test hook: the next n codeblock-kernel launches are refused by the driver (a launch error, not sticky). */
int fec_engine_debug_fail_launches(fec_engine* e, uint32_t n);

/* ---- validation helper (host only, no GPU work) -------------------------------------------------------------------- */
/* Checks a descriptor and fills in k0 / graph_index exactly as the engine will. Returns FEC_OK or FEC_ERR_BAD_ARG. */
int fec_desc_prepare(fec_cb_desc* d);

/* ---- device-resident batch: everything already in HBM (bench.py "value") ------------------------------------------ */
/* d_descs: n prepared descriptors in device memory; d_llr / d_out / d_results: device buffers the descriptors index
 * into. Asynchronous on fec_engine_stream(); max_Z / any_bg1 size the launch: the largest lifting size in the batch, and
 * bit 0 of any_bg1 = some codeblock uses base graph 1. FEC_LAUNCH_EVEN_Z in any_bg1: the caller guarantees that every
 * lifting size of the batch is even and >= 4 (true for every codeblock of 3 800 bits or more); the engine then launches the
 * packed kernel alone. Without it a second, small kernel covers the odd lifting sizes and Z = 2 (it finds nothing to do
 * otherwise; the results are the same either way). */
#define FEC_LAUNCH_EVEN_Z 0x2
int fec_launch_device(fec_engine* e, const fec_cb_desc* d_descs, uint32_t n, const int8_t* d_llr, uint8_t* d_out,
                      fec_cb_result* d_results, uint32_t max_Z, int any_bg1);

/* ---- host-buffer batch: H2D + kernels + D2H, synchronous (bench.py "e2e", unit-test surface) ----------------------- */
/* h_descs need not be prepared. llr_bytes / out_bytes are the sizes of the two host buffers. h_soft_out, if not NULL,
 * receives the n soft buffers (n * FEC_HARQ_SLOT_BYTES) after the operation. Pinned buffers (fec_host_alloc, or memory the
 * caller registered with cudaHostRegister) make the copies asynchronous DMA, and a small call (less than two kernel waves)
 * that is alone on the engine is not copied at all: the kernel reads h_llr and writes h_out in place. Either way the
 * buffers belong to the engine until the call returns. */
int fec_run_host(fec_engine* e, fec_cb_desc* h_descs, uint32_t n, const int8_t* h_llr, size_t llr_bytes, uint8_t* h_out,
                 size_t out_bytes, fec_cb_result* h_results, int8_t* h_soft_out);

void* fec_host_alloc(size_t bytes); /* pinned host memory */
void  fec_host_free(void* p);

/* ---- HARQ soft buffers --------------------------------------------------------------------------------------------- */
/* free_harq_context_entry: the slot reads as zeros from now on. Lazy and non-blocking: the engine marks the slot and the
 * next host-driven operation on it (fec_run_host / fec_queue_enqueue) carries FEC_FLAG_HARQ_ZERO; no GPU work here. */
int fec_harq_free(fec_engine* e, uint32_t slot);
/* Materialises the zeros of every freed slot in HBM and synchronises (needed only before fec_launch_device batches
 * whose descriptors do not carry FEC_FLAG_HARQ_ZERO, or before reading fec_harq_slot_ptr memory directly). */
int fec_harq_flush(fec_engine* e);
int fec_harq_read(fec_engine* e, uint32_t slot, int8_t* h_dst, uint32_t n);  /* D2H copy of the first n soft bits     */
int fec_harq_write(fec_engine* e, uint32_t slot, const int8_t* h_src, uint32_t n); /* H2D (tests, HARQ migration)    */

/* ---- queue interface: the shape hal::hw_accelerator_pusch_dec needs ----------------------------------------------- */
/* One queue per accelerator instance / submitting thread. enqueue copies the LLRs into pinned staging and returns
 * immediately (FEC_ERR_QUEUE_FULL when the staging ring is full: the caller retries, it never blocks). A batch is handed
 * to the engine's cross-queue batcher when it reaches fec_engine_cfg.queue_batch_cbs or on fec_queue_flush(): whichever
 * thread finds the batcher free launches EVERY batch that is waiting, of every queue, as one kernel (many codeblocks of
 * many UEs and cells per wave); fec_engine_cfg.max_wait_us lets a launch that is less than one wave wait for company.
 * poll is non-blocking and returns the number of completed operations written to tags/results (at most `max`); decoded
 * bytes are copied out with fec_queue_fetch(). */
int fec_queue_reserve(fec_engine* e, uint32_t* queue_id);
int fec_queue_free(fec_engine* e, uint32_t queue_id);
int fec_queue_enqueue(fec_engine* e, uint32_t queue_id, const fec_cb_desc* desc, const int8_t* llr, uint64_t tag);
/* Several operations in one call (one lock, parameters shared by consecutive descriptors checked once): the codeblocks of a
 * transport block. *accepted = how many were taken, in order; fewer than n means the ring filled up (FEC_OK if any was
 * taken, FEC_ERR_QUEUE_FULL if none): poll, then come back with the rest. */
int fec_queue_enqueue_batch(fec_engine* e, uint32_t queue_id, const fec_cb_desc* descs, const int8_t* const* llrs,
                            const uint64_t* tags, uint32_t n, uint32_t* accepted);
int fec_queue_flush(fec_engine* e, uint32_t queue_id);
int fec_queue_poll(fec_engine* e, uint32_t queue_id, uint64_t* tags, fec_cb_result* results, uint32_t max);
/* Copies the decoded bytes of a completed (polled) operation; valid until the next fec_queue_enqueue after a poll
 * that returned it. nbytes <= ceil(K/8). */
int fec_queue_fetch(fec_engine* e, uint32_t queue_id, uint64_t tag, uint8_t* dst, uint32_t nbytes);

/* ---- transport-block interface: segmentation, codeblock decoding, TB assembly and TB CRC ----- */
/* One call carries whole transport blocks: the engine splits the codeword into codeblocks as ldpc_segmenter_rx_impl::segment
 * does (lib/phy/upper/channel_coding/ldpc/ldpc_segmenter_rx_impl.cpp:53-106, ldpc_segmenter_helpers.h:82-118, ldpc.h:128-208),
 * decodes every codeblock whose CRC has not passed in an earlier transmission (the others only have the new LLRs combined
 * into their soft buffers, as pusch_decoder_impl.cpp:333-344 does), concatenates the codeblock payloads and checks the
 * transport-block CRC on the device -- the work of pusch_decoder_impl::fork_codeblock_task / join_and_notify /
 * concatenate_codeblocks (lib/phy/upper/channel_processors/pusch/pusch_decoder_impl.cpp:308-493) and of
 * pusch_decoder_hw_impl::run_asynch_hw_decoder / copy_tb_and_notify (pusch_decoder_hw_impl.cpp:149-445) without the
 * per-codeblock host loop. Per-codeblock state that the reference keeps in the rx buffer (CRC flags, decoded messages of earlier
 * transmissions) lives in the engine, keyed by the codeblock's HARQ slot. */
#define FEC_MAX_TB_CBS 162u /* MAX_NOF_SEGMENTS */

typedef struct fec_tb_desc {
  uint64_t llr_offset;      /* first of the nof_llrs codeword LLRs in the LLR buffer                                  */
  uint64_t out_offset;      /* transport block bytes (tbs / 8) in the output buffer                                   */
  uint32_t tbs;             /* transport block size in bits (a multiple of 8), TB CRC not included                     */
  uint32_t nof_llrs;        /* codeword length G = nof_ch_symbols * Qm                                                 */
  uint32_t Nref;            /* limited-buffer rate matching, 0 = off                                                   */
  uint32_t harq_slot;       /* HARQ slot of codeblock 0; codeblock r uses harq_slot + r unless slot_list is given      */
  uint32_t slot_list;       /* index into the `slots` array of the call (first of nof_codeblocks entries), or          */
                            /* FEC_TB_SLOTS_CONTIGUOUS                                                                 */
  uint8_t  base_graph;      /* 1 or 2                                                                                  */
  uint8_t  rv;              /* 0..3                                                                                    */
  uint8_t  qm;              /* 1, 2, 4, 6, 8                                                                           */
  uint8_t  nof_layers;      /* 1..4                                                                                    */
  uint8_t  max_iter;        /* nof_ldpc_iterations                                                                     */
  uint8_t  flags;           /* FEC_FLAG_NEW_DATA | FEC_FLAG_EARLY_STOP                                                 */
  uint8_t  reserved[6];
} fec_tb_desc; /* 48 bytes */
#define FEC_TB_SLOTS_CONTIGUOUS 0xffffffffu

/* Mirror of pusch_decoder_result (include/srsran/phy/upper/channel_processors/pusch/pusch_decoder_result.h). */
typedef struct fec_tb_result {
  uint16_t nof_codeblocks;       /* nof_codeblocks_total                                                               */
  uint16_t nof_cb_decoded;       /* codeblocks decoded in this call (the others had passed their CRC earlier)          */
  uint16_t nof_cb_crc_ok;        /* codeblocks whose CRC is good after this call                                       */
  uint8_t  tb_crc_ok;            /* tb_crc_ok: the transport block bytes were delivered and their CRC is good          */
  uint8_t  tb_delivered;         /* transport block bytes were written (all codeblock CRCs good), CRC good or not      */
  uint32_t iter_sum;             /* ldpc_decoder_stats over the decoded codeblocks: sum, min, max of the iterations    */
  uint16_t iter_min;             /* (a codeblock that fails counts max_iter, as in check_hw_results)                   */
  uint16_t iter_max;
} fec_tb_result; /* 16 bytes */

/* Segmentation only, on the host (no GPU needed): fills cbs[0 .. *nof_cbs) with prepared CHAIN descriptors whose
 * llr_offset is relative to tb->llr_offset + codeword offset, out_offset = r * FEC_MAX_CB_OUT_BYTES, harq_slot as the
 * call would use it (slots may be NULL). payload_bits[r] (may be NULL) = bits of codeblock r that belong to the
 * transport block + TB CRC stream (nof_new_bits in pusch_decoder_hw_impl.cpp:228). */
int fec_tb_segment(const fec_tb_desc* tb, const uint32_t* slots, fec_cb_desc* cbs, uint32_t* nof_cbs,
                   uint32_t* payload_bits);

/* Decodes n transport blocks from host buffers, synchronously. h_out receives the bytes of every transport block that
 * is delivered (left untouched otherwise, as the reference leaves `transport_block`). h_cb_results (may be NULL) receives
 * FEC_MAX_TB_CBS entries per transport block; entries of codeblocks skipped in this call have status FEC_RES_SKIPPED.
 * After a good TB CRC the HARQ slots of the block are freed (fec_harq_free) and its codeblock flags cleared; after a bad TB
 * CRC with all codeblock CRCs good the flags are reset (rx_buffer reset_codeblocks_crc). */
int fec_tb_run_host(fec_engine* e, const fec_tb_desc* tbs, uint32_t n, const uint32_t* slots, const int8_t* h_llr,
                    size_t llr_bytes, uint8_t* h_out, size_t out_bytes, fec_tb_result* h_results,
                    fec_cb_result* h_cb_results);

/* Asynchronous form (what pusch_decoder_impl does with its executor, lib/phy/upper/channel_processors/pusch/
 * pusch_decoder_impl.cpp:372-381: on_end_softbits returns, the notifier fires later): fec_tb_enqueue validates, segments and
 * puts every copy and kernel of the call on the streams of a free host-call context (fec_engine_cfg.host_contexts), then
 * returns without waiting for the GPU; FEC_ERR_QUEUE_FULL when every context is in flight (back-pressure, never blocks).
 * Calls of different threads run concurrently on the GPU. h_llr and h_out must stay valid until the call has completed; tbs
 * and slots are copied. Completion, either way:
 *   callback == NULL: the caller collects with fec_tb_poll (1 = done and results written, 0 = still running, < 0 = error;
 *                     non-blocking) or fec_tb_wait (blocking); both take the ticket exactly once;
 *   callback != NULL: an engine thread waits for the GPU, hands the blocks over and calls
 *                     callback(user, ticket, status, results, n) -- status FEC_OK or the error; on a device fault the
 *                     results say "nothing delivered" for the blocks that were lost. */
typedef void (*fec_tb_callback)(void* user, uint64_t ticket, int status, const fec_tb_result* results, uint32_t n);
int fec_tb_enqueue(fec_engine* e, const fec_tb_desc* tbs, uint32_t n, const uint32_t* slots, const int8_t* h_llr, size_t llr_bytes,
                   uint8_t* h_out, size_t out_bytes, fec_tb_callback callback, void* user, uint64_t* ticket);
int fec_tb_poll(fec_engine* e, uint64_t ticket, fec_tb_result* h_results, fec_cb_result* h_cb_results);
int fec_tb_wait(fec_engine* e, uint64_t ticket, fec_tb_result* h_results, fec_cb_result* h_cb_results);

/* Forgets the codeblock CRC flags of a slot range (rx buffer released / expired). */
int fec_tb_reset_slots(fec_engine* e, uint32_t first_slot, uint32_t n);

/* Allocation-free transport-block assembly + TB CRC over DEVICE buffers. This entry point has no counterpart in the
 * reference; it was added for callers that cannot call cudaMalloc (code under CUDA
 * graph capture, tight serving loops, harnesses that forbid it). It runs the same one-CTA-per-codeblock gather kernel as
 * fec_tb_run_host, with the kernel's scratch (jobs, codeblock map, per-block CRC accumulators and arrival counters) in a
 * file-scope static __device__ array of FEC_NOALLOC_MAX_TBS blocks / FEC_NOALLOC_MAX_CBS codeblocks (about 36 KB, held by every
 * GPU that loads the library until exit). It calls no cudaMalloc / cudaFree / allocating routine; larger calls are rejected
 * with FEC_ERR_BAD_ARG. Inputs: tbs = n HOST descriptors (tbs, nof_llrs, base_graph, qm, nof_layers, max_iter, out_offset are
 * read; payloads must be whole bytes, as every TS 38.214 transport block size gives); d_cb_msgs = the decoded codeblock
 * messages, FEC_MAX_CB_OUT_BYTES apart, the codeblocks of block 0 first, in order; d_cb_results = their fec_cb_result, same
 * order. Outputs: tbs/8 bytes per delivered block at d_out + out_offset (untouched otherwise), d_tb_results[n].
 * Asynchronous on fec_engine_stream(). Every call shares the one array (it keeps the last call's contents; the call clears
 * what must start at zero): calls are serialised on the host, but two calls whose GPU work overlaps -- two engines on the
 * same device, whose streams differ -- race on it. */
#define FEC_NOALLOC_MAX_TBS 64u
#define FEC_NOALLOC_MAX_CBS 2048u
int fec_tb_assemble_device_noalloc(fec_engine* e, const fec_tb_desc* tbs, uint32_t n, const uint8_t* d_cb_msgs,
                                   const fec_cb_result* d_cb_results, uint8_t* d_out, size_t out_bytes,
                                   fec_tb_result* d_tb_results);

/* ---- HARQ buffer pool: the reference's rx_buffer_pool over the engine's HARQ slots ---------------- */
/* Mirror of rx_buffer_pool / unique_rx_buffer (include/srsran/phy/upper/rx_buffer_pool.h:60-127, unique_rx_buffer.h:36-170;
 * lib/phy/upper/rx_buffer_pool_impl.cpp:36-155, rx_buffer_impl.h:40-258, rx_buffer_codeblock_pool.h:62-103): a fixed number of
 * buffers, each bound to an identifier (RNTI, HARQ process) while it is in use, each holding the HARQ slots ("absolute
 * codeblock identifiers") of one transport block, taken from and returned to a FIFO free list. A reservation comes back
 * LOCKED; the owner unlocks it (more transmissions may follow) or releases it (transport block delivered); run_slot expires
 * reservations that were not renewed. Freed slots are handed to fec_harq_free / fec_tb_reset_slots. The slot lists go into
 * fec_tb_run_host (fec_tb_desc.slot_list). */
typedef struct fec_harq_pool fec_harq_pool;

typedef struct fec_harq_pool_cfg {
  uint32_t nof_buffers;          /* rx_buffer_pool_config::nof_buffers                                                      */
  uint32_t first_slot;           /* the pool owns engine HARQ slots [first_slot, first_slot + nof_codeblocks)               */
  uint32_t nof_codeblocks;       /* rx_buffer_pool_config::nof_codeblocks                                                   */
  uint32_t expire_timeout_slots; /* rx_buffer_pool_config::expire_timeout_slots                                             */
  uint32_t numerology;           /* 0..4: slot counts wrap at 10240 << numerology (slot_point)                              */
  uint32_t reserved[3];
} fec_harq_pool_cfg;

/* fec_harq_pool_reserve results: 0 = reserved; > 0 = the reason the reference returns an invalid unique_rx_buffer for. */
#define FEC_POOL_OK            0
#define FEC_POOL_STOPPED       1 /* the pool was stopped                                                                    */
#define FEC_POOL_RETX_UNKNOWN  2 /* "identifier for retransmissions not found"                                              */
#define FEC_POOL_NO_BUFFER     3 /* "insufficient buffers in the pool"                                                      */
#define FEC_POOL_CB_MISMATCH   4 /* "number of codeblocks for retransmission do not match"                                  */
#define FEC_POOL_IN_USE        5 /* "HARQ already in use" (the buffer is locked)                                            */
#define FEC_POOL_NO_CODEBLOCKS 6 /* "exhausted CB pool"                                                                     */
#define FEC_POOL_INVALID_BUFFER 0xffffffffu

/* e may be NULL: the pool then only keeps the books (no engine notifications; host-only use and tests). */
int  fec_harq_pool_create(fec_engine* e, const fec_harq_pool_cfg* cfg, fec_harq_pool** out);
void fec_harq_pool_destroy(fec_harq_pool* p);
/* rx_buffer_pool::reserve(slot, id, nof_codeblocks, new_data). slot = slot_point::system_slot(); id != 0xffffffff. */
int fec_harq_pool_reserve(fec_harq_pool* p, uint32_t slot, uint32_t id, uint32_t nof_codeblocks, int new_data, uint32_t* buffer);
/* unique_rx_buffer::get_absolute_codeblock_id for every codeblock: returns their number, copies up to max of them. */
int fec_harq_pool_codeblocks(fec_harq_pool* p, uint32_t buffer, uint32_t* slots, uint32_t max);
int fec_harq_pool_unlock(fec_harq_pool* p, uint32_t buffer);  /* unique_rx_buffer::unlock / destructor                      */
int fec_harq_pool_release(fec_harq_pool* p, uint32_t buffer); /* unique_rx_buffer::release                                  */
int fec_harq_pool_run_slot(fec_harq_pool* p, uint32_t slot);  /* rx_buffer_pool::run_slot                                   */
int fec_harq_pool_stop(fec_harq_pool* p);                     /* rx_buffer_pool_controller::stop (no more reservations)     */
int fec_harq_pool_nof_free_codeblocks(fec_harq_pool* p);

/* ---- downlink twin: LDPC encoding + rate matching of PDSCH codeblocks ---------------------------- */
/* One operation = one codeblock in the accelerator's CB mode (pdsch_encoder_hw_impl.cpp:96-111): the segment the
 * reference's ldpc_segmenter_tx produced (payload + TB/CB CRC, K - nof_filler_bits bits, packed MSB first as in bit_buffer)
 * is LDPC-encoded (ldpc_encoder_impl.cpp / ldpc_encoder_generic.cpp: systematic bits, four core parity blocks through the
 * double-diagonal structure, extension parity blocks), E bits are selected from the circular buffer starting at k0(rv),
 * skipping the filler bits, and interleaved over Qm rows (ldpc_rate_matcher_impl.cpp:36-160). */
#define FEC_ENC_FLAG_UNPACKED 0x01u /* output one bit per byte (hw cfg .do_unpack) instead of packed MSB-first bytes      */

typedef struct fec_enc_desc {
  uint64_t out_offset;      /* byte offset of the output: ceil(E/8) packed bytes (4-byte aligned offsets are fastest) or  */
                            /* E bytes with FEC_ENC_FLAG_UNPACKED                                                        */
  uint32_t in_offset;       /* byte offset of the packed segment, ceil((K - nof_filler_bits)/8) bytes                     */
  uint32_t E;               /* rate-matched length in bits (hw cfg .rm_length), a multiple of qm                          */
  uint32_t Nref;            /* limited-buffer rate matching length, 0 = unlimited (hw cfg .Nref)                          */
  uint32_t k0;              /* filled in by the engine                                                                   */
  uint16_t lifting_size;    /* Z (hw cfg .lifting_size)                                                                  */
  uint16_t nof_filler_bits; /* hw cfg .nof_filler_bits                                                                   */
  uint16_t graph_index;     /* filled in by the engine                                                                   */
  uint8_t  base_graph;      /* 1 or 2 (hw cfg .base_graph_index)                                                         */
  uint8_t  rv;              /* 0..3                                                                                      */
  uint8_t  qm;              /* 1, 2, 4, 6, 8 (hw cfg .modulation)                                                        */
  uint8_t  flags;           /* FEC_ENC_FLAG_*                                                                            */
  uint8_t  reserved[6];
} fec_enc_desc; /* 40 bytes */

#define FEC_MAX_CB_IN_BYTES 1056u /* 22 * 384 / 8 */

/* Checks an encode descriptor and fills in k0 / graph_index (host only). */
int fec_enc_desc_prepare(fec_enc_desc* d);
/* Device-resident batch: n prepared descriptors, packed segments and output buffer already in HBM; asynchronous on
 * fec_engine_stream(). */
int fec_encode_launch_device(fec_engine* e, const fec_enc_desc* d_descs, uint32_t n, const uint8_t* d_in, uint8_t* d_out,
                             uint32_t max_Z);
/* Host buffers: H2D + kernel + D2H, synchronous. Descriptors need not be prepared. Only the span of h_out from the first
 * output to the end of the last one is written; bytes between two outputs inside that span are set to zero. */
int fec_encode_run_host(fec_engine* e, fec_enc_desc* h_descs, uint32_t n, const uint8_t* h_in, size_t in_bytes,
                        uint8_t* h_out, size_t out_bytes);

/* ---- the step in front of the path: soft demodulation + descrambling on the device ---------------- */
/* One descriptor = one demodulation_mapper::demodulate_soft call (include/srsran/phy/upper/channel_modulation/
 * demodulation_mapper.h:47-69; pusch_demodulator_impl issues one per OFDM symbol, lib/phy/upper/channel_processors/pusch/
 * pusch_demodulator_impl.cpp:355-375) followed by the descrambling of its soft bits with the TS 38.211 5.2.1 sequence
 * initialised with c_init = rnti * 2^15 + n_id (ibid. 272-274, 377-382; pseudo_random_generator.h:46-101). Symbols are
 * equalised complex samples (re, im float32), noise_vars the post-equalisation noise variance of each symbol.
 *
 * The reference picks its demodulator arithmetic when it is COMPILED (AVX-512 / AVX2 / scalar code, selected with #ifdef in
 * lib/phy/upper/channel_modulation/demodulation_mapper_*.cpp) and the variants round differently in the last place; the
 * `flavour` argument names the build the results must be identical to. */
#define FEC_DEMOD_FLAVOUR_GENERIC 0u /* scalar code, no fused multiply-add (baseline x86-64 build)                          */
#define FEC_DEMOD_FLAVOUR_AVX2    1u /* -mavx2 -mfma                                                                        */
#define FEC_DEMOD_FLAVOUR_AVX512  2u /* -mavx512f -mavx512bw -mavx512vl -mavx2 -mfma: -march=native on the hosts used here  */

#define FEC_DEMOD_FLAG_PI2_BPSK      0x01u /* qm == 1: pi/2-BPSK instead of BPSK                                           */
#define FEC_DEMOD_FLAG_NO_DESCRAMBLE 0x02u /* plain demodulation_mapper::demodulate_soft                                   */

typedef struct fec_demod_desc {
  uint64_t sym_offset;   /* first symbol of the call: index into the symbol (re, im pairs) and noise-variance arrays       */
  uint64_t llr_offset;   /* first of the nof_symbols * qm soft bits in the LLR buffer                                       */
  uint32_t nof_symbols;
  uint32_t c_init;       /* scrambling sequence initial state                                                              */
  uint32_t seq_offset;   /* position of the call's first soft bit in the sequence (soft bits of the codeword before it)    */
  uint32_t x1_state;     /* filled by fec_demod_desc_prepare: the two LFSR states at that position                         */
  uint32_t x2_state;
  uint8_t  qm;           /* 1, 2, 4, 6, 8                                                                                  */
  uint8_t  flags;        /* FEC_DEMOD_FLAG_*                                                                               */
  uint8_t  reserved[2];
} fec_demod_desc; /* 40 bytes */

/* Checks a demodulation descriptor and fills in x1_state / x2_state (host only, no GPU needed). */
int fec_demod_desc_prepare(fec_demod_desc* d);

/* Device-resident batch: n prepared descriptors, symbols, noise variances and the LLR buffer already in HBM; asynchronous
 * on fec_engine_stream(). max_symbols >= the largest nof_symbols of the batch. */
int fec_demod_launch_device(fec_engine* e, const fec_demod_desc* d_descs, uint32_t n, uint32_t max_symbols,
                            const float* d_symbols, const float* d_noise_vars, int8_t* d_llr, uint32_t flavour);
/* Host buffers: H2D + kernel + D2H of the LLR span the descriptors cover, synchronous (the mirror of demodulate_soft;
 * the production path is fec_tb_run_host_symbols, where the LLRs stay on the device). Descriptors need not be prepared. */
int fec_demod_run_host(fec_engine* e, const fec_demod_desc* h_descs, uint32_t n, const float* h_symbols,
                       const float* h_noise_vars, size_t nof_symbols, int8_t* h_llr, size_t llr_bytes, uint32_t flavour);
/* fec_tb_run_host with the codeword LLRs produced on the device: the demodulation descriptors write the same LLR index space
 * the transport blocks read (fec_tb_desc.llr_offset); together they must cover every LLR of every transport block. */
int fec_tb_run_host_symbols(fec_engine* e, const fec_tb_desc* tbs, uint32_t n, const uint32_t* slots,
                            const fec_demod_desc* h_demod, uint32_t n_demod, const float* h_symbols,
                            const float* h_noise_vars, size_t nof_symbols, uint32_t flavour, uint8_t* h_out, size_t out_bytes,
                            fec_tb_result* h_results, fec_cb_result* h_cb_results);

/* ---- UL-SCH demultiplexing on the device (last pointer) ------------------------------------------------
 * Replaces ulsch_demultiplex (include/srsran/phy/upper/channel_processors/pusch/ulsch_demultiplex.h:44-101, implementation
 * lib/phy/upper/channel_processors/pusch/ulsch_demultiplex_impl.cpp): the descrambled soft bits of a PUSCH codeword that
 * carries UCI are sorted into the SCH data stream (what the decoder's on_new_softbits receives) and the HARQ-ACK / CSI Part 1
 * / CSI Part 2 streams (what the UCI decoders receive), TS 38.212 6.2.7, bit for bit as the reference does it: allocation
 * order and spacing, HARQ-ACK of one or two bits on reserved elements puncturing what lies underneath (those soft bits read
 * as zeros), placeholder bits of one- and two-bit payloads with their scrambling undone. fec_ulsch_cfg mirrors
 * ulsch_demultiplex::configuration plus the arguments of set_csi_part2. */
typedef struct fec_ulsch_cfg {
  uint8_t  qm;                          /* bits per modulation symbol: 1 (pi/2-BPSK), 2, 4, 6, 8                           */
  uint8_t  nof_layers;                  /* 1..4                                                                             */
  uint8_t  start_symbol_index;          /* first OFDM symbol of the allocation                                              */
  uint8_t  nof_symbols;
  uint8_t  dmrs_type;                   /* 1 or 2                                                                           */
  uint8_t  nof_cdm_groups_without_data;
  uint8_t  csi2_after_csi1;             /* 0: CSI Part 2 is known from the start (set_csi_part2 right after demultiplex, as  */
                                        /* the reference's unit test does); 1: it becomes known when the last CSI Part 1   */
                                        /* soft bit has been delivered (pusch_processor_impl.cpp:73-101), and is placed from */
                                        /* that OFDM symbol on                                                              */
  uint8_t  reserved0;
  uint16_t nof_prb;
  uint16_t dmrs_symbol_mask;            /* bit l = OFDM symbol l of the slot carries DM-RS                                  */
  uint32_t nof_harq_ack_rvd;            /* G^HARQ-ACK_rvd                                                                   */
  uint32_t nof_harq_ack_bits;           /* O_HARQ-ACK                                                                       */
  uint32_t nof_enc_harq_ack_bits;       /* G^HARQ-ACK                                                                       */
  uint32_t nof_csi_part1_bits;          /* O_CSI-1                                                                          */
  uint32_t nof_enc_csi_part1_bits;      /* G^CSI-1                                                                          */
  uint32_t nof_csi_part2_bits;          /* O_CSI-2                                                                          */
  uint32_t nof_enc_csi_part2_bits;      /* G^CSI-2                                                                          */
} fec_ulsch_cfg; /* 40 bytes */

typedef struct fec_ulsch_sizes {
  uint32_t nof_codeword_bits; /* soft bits of the allocation (input)                                                          */
  uint32_t nof_sch_bits;      /* soft bits delivered to each stream (a punctured element counts for HARQ-ACK AND underneath)  */
  uint32_t nof_harq_ack_bits;
  uint32_t nof_csi_part1_bits;
  uint32_t nof_csi_part2_bits;
} fec_ulsch_sizes;

/* Host only (no GPU): validates the configuration and tells how many soft bits each stream receives. */
int fec_ulsch_demux_sizes(const fec_ulsch_cfg* cfg, fec_ulsch_sizes* out);
/* Device-resident: codeword, packed scrambling sequence (MSB first like bit_buffer, nof_codeword_bits bits; may be NULL when
 * no UCI field has a payload of one or two bits) and the four outputs (each with room for its stream) already in HBM;
 * asynchronous on fec_engine_stream(). d_sch is what fec_tb_desc.llr_offset / fec_launch_device then reads. */
int fec_ulsch_demux_launch_device(fec_engine* e, const fec_ulsch_cfg* cfg, const int8_t* d_codeword,
                                  const uint8_t* d_scrambling_seq, int8_t* d_sch, int8_t* d_harq_ack, int8_t* d_csi_part1,
                                  int8_t* d_csi_part2);
/* Host buffers: H2D + kernel + D2H, synchronous (the mirror of feeding the reference's pusch_codeword_buffer). */
int fec_ulsch_demux_run_host(fec_engine* e, const fec_ulsch_cfg* cfg, const int8_t* h_codeword, const uint8_t* h_scrambling_seq,
                             int8_t* h_sch, int8_t* h_harq_ack, int8_t* h_csi_part1, int8_t* h_csi_part2);

#ifdef __cplusplus
}
#endif
#endif /* FEC_B200_H */

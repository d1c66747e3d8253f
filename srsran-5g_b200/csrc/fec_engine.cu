// srsran-5g_b200/csrc/fec_engine.cu -- host side of the C ABI declared in include/fec_b200.h.
//
// One engine per GPU: graph/CRC tables and the HARQ soft-buffer pool live in HBM for the engine's lifetime; batches of
// codeblock descriptors are launched on the engine's stream (device-resident entry point) or on per-queue streams fed
// from pinned staging (host entry points). There is no CPU decode path in this file: every entry point that produces
// decoded bits does so by launching the kernels in fec_kernel_*.cuh.
#include "fec_device.cuh"
#include <cstdlib>
#include "fec_kernel_generic.cuh"
#include "fec_kernel_packed.cuh"
// launch() flags, accumulated over the descriptors of a batch
#define FEC_LF_BG1 1      // some codeblock uses base graph 1 (sizes the shared-memory layout)
#define FEC_LF_SMALL_Z 2  // some lifting size is outside the packed kernel (odd, or 2): the generic kernel runs beside it
static inline int launch_flags_of(uint32_t base_graph, uint32_t Z)
{
  return ((base_graph == 1) ? FEC_LF_BG1 : 0) | (fec_packed_covers(Z) ? 0 : FEC_LF_SMALL_Z);
}
#define FEC_WORK_SLOTS 1024u // counter pairs of the persistent codeblock kernel: one per launch that can be in flight
#include "fec_tb.cuh"
#include "fec_encode.cuh"
#include "fec_demod.cuh"
#include "fec_ulsch.cuh"
#include "ldpc_tables.h"
#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <thread>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <new>
#include <vector>

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...)
{
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

#define CU_TRY(expr)                                                                                                   \
  do {                                                                                                                 \
    cudaError_t err__ = (expr);                                                                                        \
    if (err__ != cudaSuccess) {                                                                                        \
      return fail(FEC_ERR_CUDA, "%s failed: %s", #expr, cudaGetErrorString(err__));                                    \
    }                                                                                                                  \
  } while (0)

// Index of a lifting size in the table of the 51 NR lifting sizes, -1 if it is not one. Table look-up: this runs once per
// descriptor of every host-side call (7104 times per step of the headline batch).
int z_index(unsigned Z)
{
  static const struct lut_t {
    int8_t idx[385];
    lut_t()
    {
      for (int8_t& v : idx) {
        v = -1;
      }
      for (int i = 0; i < 51; ++i) {
        idx[FEC_ALL_Z[i]] = (int8_t)i;
      }
    }
  } lut;
  return (Z <= 384) ? lut.idx[Z] : -1;
}

const uint32_t CRC_POLY[3] = {0x11021u, 0x1800063u, 0x1864CFBu}; // FEC_CRC16, FEC_CRC24B, FEC_CRC24A
const uint32_t CRC_LEN[3]  = {16, 24, 24};

void build_tables(fec_dev_tables& t, float sf)
{
  memset(&t, 0, sizeof(t));
  for (int bg = 0; bg < 2; ++bg) {
    const uint8_t* deg = bg ? FEC_BG2_DEG : FEC_BG1_DEG;
    const uint8_t* col = bg ? FEC_BG2_COL : FEC_BG1_COL;
    const int      M   = bg ? FEC_BG2_M : FEC_BG1_M;
    const int      ne  = bg ? FEC_BG2_EDGES : FEC_BG1_EDGES;
    int            e   = 0;
    for (int m = 0; m < M; ++m) {
      t.deg[bg][m]       = deg[m];
      t.row_start[bg][m] = (uint16_t)e;
      e += deg[m];
    }
    t.row_start[bg][M] = (uint16_t)e;
    memcpy(t.col[bg], col, ne);
    for (int zi = 0; zi < 51; ++zi) {
      const unsigned  Z   = FEC_ALL_Z[zi];
      const uint16_t* raw = bg ? FEC_BG2_SHIFT[FEC_Z_SET[zi]] : FEC_BG1_SHIFT[FEC_Z_SET[zi]];
      for (int i = 0; i < ne; ++i) {
        t.shift[bg * 51 + zi][i] = (uint16_t)(raw[i] % Z);
      }
    }
  }
  // Same expression as the reference's generic scaling: round(float(x) * sf), +-127 pass through.
  for (int x = 0; x < 128; ++x) {
    t.scale_lut[x] = (x > 120) ? (uint8_t)x : (uint8_t)(int)std::round((float)x * sf);
  }
  for (int p = 0; p < 3; ++p) {
    t.crc_poly[p]   = CRC_POLY[p];
    t.crc_len[p]    = CRC_LEN[p];
    const uint32_t top = 1u << CRC_LEN[p];
    uint32_t       r   = 1; // x^0
    for (int j = 0; j < FEC_CRC_X32_ENTRIES; ++j) {
      t.crc_x32[p][j] = r;
      for (int i = 0; i < 32; ++i) {
        r <<= 1;
        if (r & top) {
          r ^= CRC_POLY[p];
        }
      }
    }
    r = 1;
    for (int j = 0; j < FEC_CRC_X8_ENTRIES; ++j) {
      t.crc_x8[p][j] = r;
      for (int i = 0; i < 8; ++i) {
        r <<= 1;
        if (r & top) {
          r ^= CRC_POLY[p];
        }
      }
    }
  }
}

// One batch of a queue: pinned host staging, its own slice of device memory, and the state machine
//   FREE -> FILLING (enqueue) -> PENDING (handed to the batcher) -> INFLIGHT (kernel issued) -> DONE (the dispatch's event seen).
// The owner of the queue touches a batch under queue_t::mtx; the batcher (whichever thread holds fec_engine::disp_mtx) only
// moves PENDING batches to INFLIGHT, publishing with a release store of `state`.
struct dispatch_res;
struct batch_slot {
  enum state_t { FREE, FILLING, PENDING, INFLIGHT, DONE };
  std::atomic<int> state{FREE};
  std::atomic<bool> failed{false}; // the dispatch that carried this batch failed: every operation reports a failed codeblock
  // The results of a dispatch come back in one piece (dispatch_res::h_res); whoever needs this batch's part first -- its
  // consumer, or the batcher when it is about to reuse the set -- copies it into h_res: 0 not yet, 1 being copied, 2 done.
  std::atomic<int> harvest{2};
  dispatch_res* disp   = nullptr; // the dispatch that carries this batch (valid while INFLIGHT and harvest != 2)
  uint32_t      res_at = 0;       // first result of this batch in its array
  uint32_t      n        = 0;
  uint32_t      llr_used = 0;
  uint32_t      out_used = 0;
  uint32_t      polled   = 0;
  uint32_t      max_Z    = 0;
  int           any_bg1  = 0;
  uint64_t      seq      = 0;
  fec_cb_desc*  h_descs  = nullptr; // pinned; llr_offset / out_offset relative to this batch's d_llr / d_out
  uint64_t*     tags     = nullptr;
  int8_t*       h_llr    = nullptr; // pinned
  uint8_t*      h_out    = nullptr; // pinned
  fec_cb_result* h_res   = nullptr; // pinned
  int8_t*       d_llr    = nullptr;
  uint8_t*      d_out    = nullptr;
  std::chrono::steady_clock::time_point t_pending;
};

struct queue_t {
  std::mutex   mtx;
  bool         reserved = false;
  bool         inited   = false;
  std::vector<batch_slot*> slots;
  int          filling  = -1;
  uint64_t     next_seq = 1;
  uint32_t     depth    = 0; // operations staged or in flight (not yet polled)
};

// Resources of one dispatch of the cross-queue batcher: the combined descriptor array and the results of one launch.
struct dispatch_res {
  fec_cb_desc*   h_descs = nullptr; // pinned
  fec_cb_desc*   d_descs = nullptr;
  fec_cb_result* d_res   = nullptr;
  fec_cb_result* h_res   = nullptr; // pinned: the results of the launch, copied back in one piece
  cudaEvent_t    done    = nullptr; // the launch and every copy of the dispatch have finished: the set can be reused
  bool           used    = false;
  std::vector<batch_slot*> batches; // what the last dispatch with this set carried
};
constexpr int DISPATCH_RING = 16;
constexpr int DISPATCH_STREAMS = 8;

// One host-call context: the stream set, device staging and pinned bounce buffers one fec_run_host / transport-block call
// works in. An engine owns several, so that calls from different threads run concurrently (each on its own streams) instead
// of queueing behind one mutex; the asynchronous transport-block interface keeps a context from fec_tb_enqueue to completion.
// State of a submitted transport-block call, from tb_submit to tb_complete (one per context).
struct tb_pending {
  std::vector<fec_tb_desc> tbs; // the call's descriptors (the caller's array need not outlive fec_tb_enqueue)
  uint32_t                 n = 0, n_map = 0;
  size_t                   off_map = 0, off_tbres = 0;
  bool                     pipelined = false;
  uint8_t*                 h_out     = nullptr;
  std::vector<uint32_t>    taken; // "reads as zeros" marks this call consumed: put back if the call fails
  uint64_t                 ticket = 0;
  fec_tb_callback          cb     = nullptr;
  void*                    user   = nullptr;
  std::vector<fec_tb_result> res; // callback mode: where the results are assembled
  std::chrono::steady_clock::time_point t_in, t_seg, t_enq;
};

void tb_pending_free(tb_pending* p)
{
  delete p;
}

struct host_ctx {
  uint32_t              index        = 0;
  bool                  busy         = false; // guarded by fec_engine::ctx_mtx
  cudaStream_t          stream       = nullptr;
  cudaStream_t          pipe[3]      = {nullptr, nullptr, nullptr}; // chunk pipeline (copy/compute overlap)
  cudaEvent_t           pipe_done[3] = {nullptr, nullptr, nullptr};
  cudaStream_t          copy_in      = nullptr; // all input copies of a chunked call, in chunk order (one FIFO)
  cudaEvent_t           done         = nullptr; // everything the call put on its streams has finished
  std::vector<cudaEvent_t> chunk_in;            // chunk c's inputs have arrived
  std::vector<cudaEvent_t> chunk_out;           // chunk c's results and transport block bytes are back on the host
  std::vector<uint32_t>    tb_chunk_end;        // pipelined transport-block call: one-past-last block of each chunk
  // staging for fec_run_host
  fec_cb_desc*   d_descs = nullptr;
  int8_t*        d_llr   = nullptr;
  uint8_t*       d_out   = nullptr;
  fec_cb_result* d_res   = nullptr;
  fec_cb_result* h_res_pin   = nullptr; // pinned bounce buffers: async copies from/to pageable memory would block
  fec_cb_desc*   h_descs_pin = nullptr; // the host and serialise the chunk pipeline
  // transport-block layer (allocated on first use)
  uint8_t*       h_meta  = nullptr; // pinned: jobs | codeblock map | scratch zeros | codeblock descriptors
  uint8_t*       d_meta  = nullptr;
  uint8_t*       h_rblk  = nullptr; // pinned: codeblock results | transport block results
  uint8_t*       d_rblk  = nullptr;
  uint8_t*       h_tbout = nullptr; // pinned bounce buffer for the transport block bytes
  std::vector<fec_tb_geometry> tb_geo;
  // weights of the last transport-block geometry seen (the same allocation usually repeats slot after slot)
  uint32_t              wc_key[3] = {0, 0, 0}; // nof_cbs, data_bits, stream_bits
  std::vector<uint32_t> wc_weights;
  tb_pending*           pending = nullptr;
};

} // namespace

struct fec_engine {
  fec_engine_cfg        cfg;
  int                   sm_count     = 0;
  cudaStream_t          stream       = nullptr; // device-resident entry points (fec_launch_device, encoder, demodulator)
  fec_dev_tables*       d_tab        = nullptr;
  fec_packed_tables*    d_ptab       = nullptr;
  uint32_t*             d_work       = nullptr; // FEC_WORK_SLOTS x {next, done} counters of the persistent kernel
  std::atomic<uint32_t> work_slot{0};
  bool                  packed_ok    = false;
  std::atomic<int>      ctx_busy{0}; // host-call contexts in use (a call that is alone on the engine may read its input in place)
  int8_t*               d_harq       = nullptr;
  std::atomic<uint64_t> launches{0};
  std::atomic<uint8_t>* fresh = nullptr; // per HARQ slot: freed, reads as zeros until the next operation rewrites it
  // host-call contexts: ctxs[0] also serves the encoder / demodulator host paths (their staging is engine-wide)
  std::vector<host_ctx*>  ctxs;
  // UL-SCH demultiplexing plans (fec_ulsch.cuh), by configuration: flags and segments in HBM
  struct ulsch_entry {
    fec_ulsch_cfg  cfg;
    fec_ulsch_plan plan;
    fec_ulsch_seg* d_segs  = nullptr;
    uint8_t*       d_flags = nullptr;
    uint64_t       stamp   = 0;
  };
  std::mutex               ulsch_mtx;
  std::vector<ulsch_entry> ulsch_cache;
  uint64_t                 ulsch_clock = 0;
  std::mutex              ctx_mtx;
  std::condition_variable ctx_cv;
  std::vector<queue_t>* queues = nullptr;
  std::mutex            queues_mtx;
  // cross-queue batcher: batches handed over by the queues, dispatched by whichever thread gets disp_mtx (no thread of its own)
  std::mutex               pend_mtx;
  std::vector<batch_slot*> pending;
  std::atomic<uint32_t>    pending_count{0};
  std::mutex               disp_mtx;
  dispatch_res             disp[DISPATCH_RING];
  uint32_t                 disp_next = 0;
  cudaStream_t             disp_stream[DISPATCH_STREAMS] = {}; // dispatches rotate over them: that many launches in flight
  uint32_t                 disp_cap = 0; // codeblocks one dispatch can carry
  // transport-block layer (allocated on first use)
  std::mutex            tb_alloc_mtx;
  std::atomic<uint8_t>* cb_ok    = nullptr; // per HARQ slot: the codeblock's CRC passed in an earlier transmission
  uint8_t*              d_msg    = nullptr; // per HARQ slot: decoded message of the codeblock (FEC_MAX_CB_OUT_BYTES)
  // downlink encoder staging (allocated on first use, grown on demand)
  fec_enc_desc* d_enc_descs = nullptr;
  uint8_t*      d_enc_in    = nullptr;
  uint8_t*      d_enc_out   = nullptr;
  size_t        enc_descs_cap = 0, enc_in_cap = 0, enc_out_cap = 0;
  // soft demodulator: tables and staging (allocated on first use, grown on demand)
  fec_demod_tables* d_dtab     = nullptr;
  fec_demod_desc*   d_dm_descs = nullptr;
  float*            d_dm_sym   = nullptr;
  float*            d_dm_noise = nullptr;
  int8_t*           d_dm_llr   = nullptr;
  size_t            dm_descs_cap = 0, dm_sym_cap = 0, dm_noise_cap = 0, dm_llr_cap = 0;
  fec_demod_desc*   h_dm_descs   = nullptr; // pinned copy of the call's descriptors, prepared and rebased
  size_t            h_dm_descs_cap = 0;
  // counters (fec_engine_get_stats)
  std::atomic<uint64_t> st_codeblocks{0}, st_batches{0}, st_results{0}, st_crc_pass{0}, st_queue_full{0}, st_faults{0}, st_generic{0},
      st_tb_calls{0}, st_tbs{0}, st_tb_crc_ok{0}, st_max_depth{0}, st_dispatches{0}, st_dispatch_slots{0};
  std::atomic<uint64_t> st_iter_hist[FEC_STATS_ITER_BINS];
  std::atomic<uint32_t> debug_fail_launches{0}; // fec_engine_debug_fail_launches: the next n codeblock launches fail
  // asynchronous transport-block interface: completion thread for calls that carry a callback
  std::thread             tb_thread;
  std::mutex              tb_mtx;
  std::condition_variable tb_cv;
  std::vector<host_ctx*>  tb_fifo; // contexts in flight that carry a callback, in submission order
  bool                    tb_stop = false;
  std::atomic<uint64_t>   tb_ticket{1};
};

namespace {

// Counters of harvested codeblock results (fec_engine_get_stats).
inline void stats_result(fec_engine* e, const fec_cb_result& r)
{
  e->st_results.fetch_add(1, std::memory_order_relaxed);
  if (r.status & FEC_RES_CRC_PASS) {
    e->st_crc_pass.fetch_add(1, std::memory_order_relaxed);
  }
  const uint32_t it = r.nof_ldpc_iterations < FEC_STATS_ITER_BINS ? r.nof_ldpc_iterations : FEC_STATS_ITER_BINS - 1;
  e->st_iter_hist[it].fetch_add(1, std::memory_order_relaxed);
}

// The same for a batch: counted locally, then a handful of atomic adds (7104 results cost 100 us one by one).
struct stats_batch {
  uint64_t results = 0, crc_pass = 0, hist[FEC_STATS_ITER_BINS] = {};
  void     add(const fec_cb_result& r)
  {
    ++results;
    crc_pass += (r.status & FEC_RES_CRC_PASS) ? 1 : 0;
    ++hist[r.nof_ldpc_iterations < FEC_STATS_ITER_BINS ? r.nof_ldpc_iterations : FEC_STATS_ITER_BINS - 1];
  }
  void flush(fec_engine* e)
  {
    e->st_results.fetch_add(results, std::memory_order_relaxed);
    e->st_crc_pass.fetch_add(crc_pass, std::memory_order_relaxed);
    for (int b = 0; b < FEC_STATS_ITER_BINS; ++b) {
      if (hist[b]) {
        e->st_iter_hist[b].fetch_add(hist[b], std::memory_order_relaxed);
      }
    }
  }
};

int launch(fec_engine*          e,
           cudaStream_t         st,
           const fec_cb_desc*   d_descs,
           uint32_t             n,
           const int8_t*        d_llr,
           uint8_t*             d_out,
           fec_cb_result*       d_res,
           uint32_t             max_Z,
           int                  any_bg1,
           uint32_t             grid_div = 1) // > 1: use 1/grid_div of the resident CTAs (chunks that overlap each other)
{
  if (n == 0) {
    return FEC_OK;
  }
  if (z_index(max_Z) < 0) {
    return fail(FEC_ERR_BAD_ARG, "max_Z=%u is not a lifting size", max_Z);
  }
  CU_TRY(cudaSetDevice(e->cfg.device));
  e->st_batches.fetch_add(1, std::memory_order_relaxed);
  e->st_codeblocks.fetch_add(n, std::memory_order_relaxed);
  if (e->debug_fail_launches.load(std::memory_order_relaxed) != 0) {
    // fec_engine_debug_fail_launches: a launch the driver refuses (more dynamic shared memory than the SM has), to exercise
    // the fault paths. The error is not sticky.
    uint32_t left = e->debug_fail_launches.load();
    while (left != 0 && !e->debug_fail_launches.compare_exchange_weak(left, left - 1)) {
    }
    if (left != 0) {
      fec_cb_generic_kernel<<<n, 32, 1u << 20, st>>>(d_descs, d_llr, d_out, d_res, e->d_harq, e->d_tab, 0);
      const cudaError_t err = cudaGetLastError();
      return fail(FEC_ERR_CUDA, "codeblock kernel launch failed (injected): %s", cudaGetErrorString(err));
    }
  }
  const bool packed_ok = (e->cfg.kernel_variant != 1) && e->packed_ok;
  if (packed_ok) {
    // one counter pair per launch in flight (the kernel leaves its pair zeroed): far more slots than streams
    uint32_t* work = e->d_work + 2 * (e->work_slot.fetch_add(1, std::memory_order_relaxed) % FEC_WORK_SLOTS);
    int r = fec_packed_launch(st, d_descs, n, d_llr, d_out, d_res, e->d_harq, e->d_tab, e->d_ptab, max_Z, any_bg1 & FEC_LF_BG1,
                              (uint32_t)e->sm_count, work, (any_bg1 & FEC_LF_SMALL_Z) != 0, grid_div);
    if ((any_bg1 & FEC_LF_SMALL_Z) != 0) {
      e->st_generic.fetch_add(1, std::memory_order_relaxed);
    }
    if (r != FEC_OK) {
      return fail(r, "packed kernel launch failed: %s", cudaGetErrorString(cudaGetLastError()));
    }
  } else {
    const fec_generic_smem_layout L = fec_generic_layout(max_Z, any_bg1 & FEC_LF_BG1);
    unsigned                      threads = ((max_Z + 31) / 32) * 32;
    threads                               = threads > 384 ? 384 : threads;
    fec_cb_generic_kernel<<<n, threads, L.total, st>>>(d_descs, d_llr, d_out, d_res, e->d_harq, e->d_tab, 0);
    CU_TRY(cudaGetLastError());
    e->st_generic.fetch_add(1, std::memory_order_relaxed);
  }
  e->launches.fetch_add(1);
  return FEC_OK;
}

// Same operation parameters (everything fec_desc_prepare looks at), placement fields aside.
inline bool same_shape(const fec_cb_desc& a, const fec_cb_desc& b)
{
  return a.E == b.E && a.Nref == b.Nref && a.nof_filler_bits == b.nof_filler_bits && a.lifting_size == b.lifting_size &&
         a.base_graph == b.base_graph && a.rv == b.rv && a.qm == b.qm && a.crc_type == b.crc_type && a.max_iter == b.max_iter &&
         a.mode == b.mode && a.flags == b.flags;
}

// Freed slots are zeroed lazily: the first host-driven operation on one carries FEC_FLAG_HARQ_ZERO.
void apply_fresh(fec_engine* e, fec_cb_desc& d, std::vector<uint32_t>* taken = nullptr)
{
  if (d.mode != FEC_MODE_DECODE && d.harq_slot < e->cfg.harq_slots && e->fresh[d.harq_slot].load(std::memory_order_relaxed) &&
      e->fresh[d.harq_slot].exchange(0)) {
    d.flags |= FEC_FLAG_HARQ_ZERO;
    if (taken) {
      taken->push_back(d.harq_slot);
    }
  }
}

// Error exits of the host-buffer entry points: no copy into or out of the caller's buffers may still be in flight, and the
// "reads as zeros" marks this call consumed are put back (the operation that would have rewritten the slots did not run).
struct run_guard {
  fec_engine*           e;
  host_ctx*             c;
  std::vector<uint32_t> taken;
  bool                  ok = false;
  // codeblock calls: on failure every operation of the call is reported as a failed codeblock (a device fault must
  // surface as failed codeblocks, never as a hang or a crash), beside the error code
  fec_cb_result*     fail_results = nullptr;
  const fec_cb_desc* fail_descs   = nullptr;
  uint32_t           fail_n       = 0;
  run_guard(fec_engine* e_, host_ctx* c_) : e(e_), c(c_) {}
  ~run_guard()
  {
    if (ok) {
      return;
    }
    e->st_faults.fetch_add(1, std::memory_order_relaxed);
    for (uint32_t i = 0; fail_results && i < fail_n; ++i) {
      fail_results[i]                     = fec_cb_result();
      fail_results[i].status              = FEC_RES_DONE;
      fail_results[i].nof_ldpc_iterations = fail_descs[i].max_iter;
    }
    cudaStreamSynchronize(c->copy_in);
    for (int k = 0; k < 3; ++k) {
      cudaStreamSynchronize(c->pipe[k]);
    }
    cudaStreamSynchronize(c->stream);
    cudaGetLastError();
    for (uint32_t s : taken) {
      e->fresh[s].store(1);
    }
  }
};

// ---- host-call contexts ----------------------------------------------------------------------------------------------------
int ctx_create(fec_engine* e, uint32_t index, host_ctx** out)
{
  host_ctx* c = new (std::nothrow) host_ctx();
  if (!c) {
    return fail(FEC_ERR_NO_MEMORY, "host allocation failed");
  }
  *out     = c;
  c->index = index;
  CU_TRY(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  CU_TRY(cudaStreamCreateWithFlags(&c->copy_in, cudaStreamNonBlocking));
  CU_TRY(cudaEventCreateWithFlags(&c->done, cudaEventDisableTiming));
  for (int i = 0; i < 3; ++i) {
    CU_TRY(cudaStreamCreateWithFlags(&c->pipe[i], cudaStreamNonBlocking));
    CU_TRY(cudaEventCreateWithFlags(&c->pipe_done[i], cudaEventDisableTiming));
  }
  const uint32_t nc = e->cfg.max_batch_cbs;
  CU_TRY(cudaMalloc((void**)&c->d_descs, nc * sizeof(fec_cb_desc)));
  CU_TRY(cudaMalloc((void**)&c->d_llr, e->cfg.max_batch_llrs));
  CU_TRY(cudaMalloc((void**)&c->d_out, (size_t)nc * FEC_MAX_CB_OUT_BYTES));
  CU_TRY(cudaMalloc((void**)&c->d_res, nc * sizeof(fec_cb_result)));
  CU_TRY(cudaHostAlloc((void**)&c->h_res_pin, nc * sizeof(fec_cb_result), cudaHostAllocDefault));
  CU_TRY(cudaHostAlloc((void**)&c->h_descs_pin, nc * sizeof(fec_cb_desc), cudaHostAllocDefault));
  return FEC_OK;
}

void ctx_destroy(host_ctx* c)
{
  if (!c) {
    return;
  }
  cudaFree(c->d_descs);
  cudaFree(c->d_llr);
  cudaFree(c->d_out);
  cudaFree(c->d_res);
  cudaFreeHost(c->h_res_pin);
  cudaFreeHost(c->h_descs_pin);
  cudaFree(c->d_meta);
  cudaFree(c->d_rblk);
  cudaFreeHost(c->h_meta);
  cudaFreeHost(c->h_rblk);
  cudaFreeHost(c->h_tbout);
  for (cudaEvent_t ev : c->chunk_in) {
    cudaEventDestroy(ev);
  }
  for (cudaEvent_t ev : c->chunk_out) {
    cudaEventDestroy(ev);
  }
  for (int i = 0; i < 3; ++i) {
    if (c->pipe[i]) {
      cudaStreamDestroy(c->pipe[i]);
    }
    if (c->pipe_done[i]) {
      cudaEventDestroy(c->pipe_done[i]);
    }
  }
  if (c->done) {
    cudaEventDestroy(c->done);
  }
  if (c->copy_in) {
    cudaStreamDestroy(c->copy_in);
  }
  if (c->stream) {
    cudaStreamDestroy(c->stream);
  }
  tb_pending_free(c->pending);
  delete c;
}

// A free context: `index` >= 0 asks for that one (the encoder / demodulator paths, whose staging is engine-wide, use 0).
// Blocks until one is free unless `block` is false (then nullptr).
host_ctx* ctx_acquire(fec_engine* e, int index, bool block)
{
  std::unique_lock<std::mutex> lk(e->ctx_mtx);
  for (;;) {
    if (index >= 0) {
      if (!e->ctxs[index]->busy) {
        e->ctxs[index]->busy = true;
        e->ctx_busy.fetch_add(1, std::memory_order_relaxed);
        return e->ctxs[index];
      }
    } else {
      // highest index first, so that context 0 stays free for the callers that need it
      for (size_t i = e->ctxs.size(); i-- > 0;) {
        if (!e->ctxs[i]->busy) {
          e->ctxs[i]->busy = true;
          e->ctx_busy.fetch_add(1, std::memory_order_relaxed);
          return e->ctxs[i];
        }
      }
    }
    if (!block) {
      return nullptr;
    }
    e->ctx_cv.wait(lk);
  }
}

void ctx_release(fec_engine* e, host_ctx* c)
{
  {
    std::lock_guard<std::mutex> lk(e->ctx_mtx);
    c->busy = false;
    e->ctx_busy.fetch_sub(1, std::memory_order_relaxed);
  }
  e->ctx_cv.notify_all();
}

struct ctx_lease {
  fec_engine* e;
  host_ctx*   c;
  ctx_lease(fec_engine* e_, int index) : e(e_), c(ctx_acquire(e_, index, true)) {}
  ~ctx_lease()
  {
    if (c) {
      ctx_release(e, c);
    }
  }
  host_ctx* keep() // the context outlives the lease (asynchronous calls)
  {
    host_ctx* k = c;
    c           = nullptr;
    return k;
  }
};

uint32_t queue_batch_cbs(const fec_engine* e)
{
  return e->cfg.queue_batch_cbs ? e->cfg.queue_batch_cbs : e->cfg.max_batch_cbs;
}

size_t queue_batch_llrs(const fec_engine* e)
{
  // staging bytes of one batch: every codeblock of it at the largest rate-matched length the engine stages, capped by the
  // engine-wide staging limit
  const size_t all = (size_t)queue_batch_cbs(e) * 25344u;
  return all < e->cfg.max_batch_llrs ? all : e->cfg.max_batch_llrs;
}

int slot_alloc(fec_engine* e, batch_slot& s)
{
  const uint32_t nc = queue_batch_cbs(e);
  const size_t   nl = queue_batch_llrs(e);
  CU_TRY(cudaHostAlloc((void**)&s.h_descs, nc * sizeof(fec_cb_desc), cudaHostAllocDefault));
  CU_TRY(cudaHostAlloc((void**)&s.h_llr, nl, cudaHostAllocDefault));
  CU_TRY(cudaHostAlloc((void**)&s.h_out, (size_t)nc * FEC_MAX_CB_OUT_BYTES, cudaHostAllocDefault));
  CU_TRY(cudaHostAlloc((void**)&s.h_res, nc * sizeof(fec_cb_result), cudaHostAllocDefault));
  s.tags = new (std::nothrow) uint64_t[nc];
  CU_TRY(cudaMalloc((void**)&s.d_llr, nl));
  CU_TRY(cudaMalloc((void**)&s.d_out, (size_t)nc * FEC_MAX_CB_OUT_BYTES));
  return FEC_OK;
}

// The batch's part of its dispatch's results, copied into the batch's own array exactly once (see batch_slot::harvest).
// Only called when the dispatch is known to have finished.
void harvest_batch(batch_slot& s)
{
  int expected = 0;
  if (s.harvest.compare_exchange_strong(expected, 1, std::memory_order_acq_rel)) {
    memcpy(s.h_res, s.disp->h_res + s.res_at, s.n * sizeof(fec_cb_result));
    s.harvest.store(2, std::memory_order_release);
    return;
  }
  while (s.harvest.load(std::memory_order_acquire) != 2) {
    std::this_thread::yield();
  }
}

void slot_free(batch_slot* s)
{
  if (!s) {
    return;
  }
  cudaFreeHost(s->h_descs);
  cudaFreeHost(s->h_llr);
  cudaFreeHost(s->h_out);
  cudaFreeHost(s->h_res);
  delete[] s->tags;
  cudaFree(s->d_llr);
  cudaFree(s->d_out);
  delete s;
}

// Device-side address of host memory the device can address directly (cudaHostAlloc / fec_host_alloc / cudaHostRegister),
// nullptr for pageable memory. Small calls let the kernel read its soft bits -- and write its bytes -- there instead of
// copying first: the transfer then overlaps the decoding codeblock by codeblock inside one launch (one config-3 transport
// block: 61 us instead of 78 us of device time). FEC_B200_HOST_ZEROCOPY=0 (measurements) keeps the copies.
const void* pinned_device_ptr(const void* h)
{
  static const bool on = !(getenv("FEC_B200_HOST_ZEROCOPY") && atoi(getenv("FEC_B200_HOST_ZEROCOPY")) == 0);
  if (!on || h == nullptr) {
    return nullptr;
  }
  cudaPointerAttributes a{};
  if (cudaPointerGetAttributes(&a, h) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  return (a.type == cudaMemoryTypeHost) ? a.devicePointer : nullptr;
}

// True when the packed kernel reads the input of this operation exactly once, with the 16-byte loads of the dematcher's
// staging loop (the decoder-only mode reads its input byte by byte: fine from HBM, wasteful over PCIe).
inline bool reads_input_once(const fec_cb_desc& d)
{
  return d.mode != FEC_MODE_DECODE && fec_packed_covers(d.lifting_size) &&
         d.E <= ((d.base_graph == 1) ? 68u : 52u) * d.lifting_size * 2u;
}

// ---- cross-queue batcher ---------------------------------------------------------------------------------------------------
// Queues hand their full (or flushed) batches over; whichever thread gets disp_mtx dispatches EVERYTHING that is pending, of
// every queue, as one launch: one combined descriptor array (buffer positions as absolute device addresses: the kernel is given
// null base pointers), one kernel over all the codeblocks -- which reads the batches' pinned staging buffers and writes their
// output buffers in place where it can, and is framed by the batches' copies where it cannot --, one copy for all the results
// and one event. A thread that finds the lock taken leaves its batch to the holder and returns at once; progress is guaranteed
// because every poll tries again. With max_wait_us > 0 a dispatch that holds less than one kernel wave waits that long
// (since its oldest batch was handed over) for other queues to join: fewer, fuller launches under load, bounded added latency.
int disp_resources(fec_engine* e)
{
  if (e->disp_cap != 0) {
    return FEC_OK;
  }
  CU_TRY(cudaSetDevice(e->cfg.device));
  const uint32_t cap = queue_batch_cbs(e) * 8u > e->cfg.max_batch_cbs ? queue_batch_cbs(e) * 8u : e->cfg.max_batch_cbs;
  for (dispatch_res& r : e->disp) {
    CU_TRY(cudaHostAlloc((void**)&r.h_descs, (size_t)cap * sizeof(fec_cb_desc), cudaHostAllocDefault));
    CU_TRY(cudaMalloc((void**)&r.d_descs, (size_t)cap * sizeof(fec_cb_desc)));
    CU_TRY(cudaMalloc((void**)&r.d_res, (size_t)cap * sizeof(fec_cb_result)));
    CU_TRY(cudaHostAlloc((void**)&r.h_res, (size_t)cap * sizeof(fec_cb_result), cudaHostAllocDefault));
    CU_TRY(cudaEventCreateWithFlags(&r.done, cudaEventDisableTiming));
  }
  for (cudaStream_t& st : e->disp_stream) {
    CU_TRY(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  }
  e->disp_cap = cap;
  return FEC_OK;
}

// Issues one launch for the given batches (all PENDING). On failure they are marked failed and still become INFLIGHT / done.
void dispatch_batches(fec_engine* e, batch_slot** b, size_t nb)
{
  cudaSetDevice(e->cfg.device);
  dispatch_res& r = e->disp[e->disp_next % DISPATCH_RING];
  cudaStream_t  st = e->disp_stream[e->disp_next % DISPATCH_STREAMS];
  ++e->disp_next;
  bool ok = true;
  if (r.used) {
    ok = cudaEventSynchronize(r.done) == cudaSuccess; // sixteen dispatches ago: normally long done
    // Batches of that dispatch whose consumer has not looked yet still read their results from r.h_res: hand them over
    // before the array is written again.
    for (batch_slot* old : r.batches) {
      if (old->disp == &r && old->harvest.load(std::memory_order_acquire) != 2) {
        harvest_batch(*old);
      }
    }
  }
  r.batches.assign(b, b + nb);
  uint32_t n = 0, max_Z = 0;
  int      any1 = 0;
  // Zero copy: the staging buffers of a batch are pinned host memory, which the device addresses directly. The kernel then
  // reads the soft bits over PCIe itself (each is read exactly once, with coalesced 16-byte loads) and writes the decoded
  // bytes straight into the batch's host buffer: two copies per batch less for this thread to issue -- the dispatcher is
  // serial, and with 16 consumer threads its CUDA calls (about 4 per batch, 3-4 us each) were what bounded the queue path.
  // Not for operations the packed kernel reads byte by byte (inputs longer than its staging area) or leaves to the generic
  // kernel. FEC_B200_QUEUE_ZEROCOPY=0 (measurements) keeps the copies.
  static const bool zc_on = !(getenv("FEC_B200_QUEUE_ZEROCOPY") && atoi(getenv("FEC_B200_QUEUE_ZEROCOPY")) == 0);
  bool              zc    = zc_on && e->packed_ok && e->cfg.kernel_variant != 1;
  for (size_t k = 0; k < nb && zc; ++k) {
    const batch_slot& s = *b[k];
    zc                  = (s.any_bg1 & FEC_LF_SMALL_Z) == 0;
    for (uint32_t i = 0; i < s.n && zc; ++i) {
      zc = reads_input_once(s.h_descs[i]);
    }
  }
  for (size_t k = 0; k < nb; ++k) {
    batch_slot& s = *b[k];
    for (uint32_t i = 0; i < s.n; ++i) {
      fec_cb_desc d = s.h_descs[i];
      const uint64_t out = (uint64_t)(uintptr_t)(zc ? s.h_out : s.d_out) + d.out_offset;
      // absolute addresses: the launch passes null base pointers
      d.llr_offset += (uint64_t)(uintptr_t)(zc ? s.h_llr : s.d_llr);
      d.out_offset = (uint32_t)out;
      d.reserved1  = (uint32_t)(out >> 32);
      r.h_descs[n + i] = d;
    }
    n += s.n;
    max_Z = s.max_Z > max_Z ? s.max_Z : max_Z;
    any1 |= s.any_bg1;
  }
  ok = ok && cudaMemcpyAsync(r.d_descs, r.h_descs, (size_t)n * sizeof(fec_cb_desc), cudaMemcpyHostToDevice, st) == cudaSuccess;
  for (size_t k = 0; k < nb && ok && !zc; ++k) {
    ok = cudaMemcpyAsync(b[k]->d_llr, b[k]->h_llr, b[k]->llr_used, cudaMemcpyHostToDevice, st) == cudaSuccess;
  }
  ok = ok && launch(e, st, r.d_descs, n, nullptr, nullptr, r.d_res, max_Z, any1) == FEC_OK;
  uint32_t at = 0;
  for (size_t k = 0; k < nb; ++k) {
    batch_slot& s = *b[k];
    if (ok && !zc) {
      ok = cudaMemcpyAsync(s.h_out, s.d_out, s.out_used, cudaMemcpyDeviceToHost, st) == cudaSuccess;
    }
    s.disp   = &r;
    s.res_at = at;
    at += s.n;
  }
  // One copy for the results of every batch and one event for the whole dispatch (a copy and an event per batch were
  // two more driver calls per batch on the batcher's serial path).
  ok = ok && cudaMemcpyAsync(r.h_res, r.d_res, (size_t)n * sizeof(fec_cb_result), cudaMemcpyDeviceToHost, st) == cudaSuccess;
  if (!ok) {
    cudaGetLastError();
  }
  // The event is recorded BEFORE the batches are published: a consumer that sees INFLIGHT must never find the completed
  // record of this set's previous dispatch.
  cudaEventRecord(r.done, st);
  r.used = true;
  for (size_t k = 0; k < nb; ++k) {
    batch_slot& s = *b[k];
    if (!ok) {
      s.failed.store(true);
    }
    s.harvest.store(0, std::memory_order_release);
    s.state.store(batch_slot::INFLIGHT, std::memory_order_release);
  }
  if (!ok) {
    e->st_faults.fetch_add(1, std::memory_order_relaxed);
  }
  e->st_dispatches.fetch_add(1, std::memory_order_relaxed);
  e->st_dispatch_slots.fetch_add(nb, std::memory_order_relaxed);
}

void dispatch_try(fec_engine* e)
{
  while (e->pending_count.load(std::memory_order_acquire) != 0) {
    if (!e->disp_mtx.try_lock()) {
      return; // the holder looks at the list again before it lets go, and every poll comes back here
    }
    std::vector<batch_slot*> take;
    {
      std::lock_guard<std::mutex> lk(e->pend_mtx);
      take.swap(e->pending);
      e->pending_count.store(0, std::memory_order_release);
    }
    if (!take.empty() && e->cfg.max_wait_us != 0) {
      const uint32_t wave = FEC_PACKED_CTAS_PER_SM * (uint32_t)e->sm_count;
      const auto     until = take.front()->t_pending + std::chrono::microseconds(e->cfg.max_wait_us);
      for (;;) {
        uint32_t n = 0;
        for (batch_slot* s : take) {
          n += s->n;
        }
        if (n >= wave || std::chrono::steady_clock::now() >= until) {
          break;
        }
        if (e->pending_count.load(std::memory_order_acquire) != 0) {
          std::lock_guard<std::mutex> lk(e->pend_mtx);
          take.insert(take.end(), e->pending.begin(), e->pending.end());
          e->pending.clear();
          e->pending_count.store(0, std::memory_order_release);
        } else {
          std::this_thread::yield();
        }
      }
    }
    // as many launches as the combined descriptor array asks for (normally one)
    size_t k0 = 0;
    while (k0 < take.size()) {
      size_t   k1 = k0;
      uint32_t n  = 0;
      while (k1 < take.size() && (k1 == k0 || n + take[k1]->n <= e->disp_cap)) {
        n += take[k1]->n;
        ++k1;
      }
      dispatch_batches(e, take.data() + k0, k1 - k0);
      k0 = k1;
    }
    e->disp_mtx.unlock();
  }
}

// Hands a filled batch to the batcher and tries to dispatch.
void queue_submit_slot(fec_engine* e, batch_slot& s)
{
  s.t_pending = std::chrono::steady_clock::now();
  s.failed.store(false);
  s.state.store(batch_slot::PENDING, std::memory_order_release);
  {
    std::lock_guard<std::mutex> lk(e->pend_mtx);
    e->pending.push_back(&s);
    e->pending_count.fetch_add(1, std::memory_order_release);
  }
  dispatch_try(e);
}

} // namespace

extern "C" {

uint32_t fec_abi_version(void)
{
  return FEC_B200_ABI_VERSION;
}

const char* fec_last_error(void)
{
  return g_err;
}

void fec_engine_default_cfg(fec_engine_cfg* cfg)
{
  memset(cfg, 0, sizeof(*cfg));
  cfg->device         = 0;
  cfg->harq_slots     = 4096;
  cfg->max_batch_cbs  = 2048;
  cfg->max_batch_llrs = 2048u * 25344u;
  cfg->nof_queues     = 8;
  cfg->scaling_factor = 0.8f;
  cfg->kernel_variant = 0;
}

int fec_desc_prepare(fec_cb_desc* d)
{
  if (d == nullptr) {
    return fail(FEC_ERR_BAD_ARG, "null descriptor");
  }
  if (d->base_graph != 1 && d->base_graph != 2) {
    return fail(FEC_ERR_BAD_ARG, "base_graph=%u", d->base_graph);
  }
  const int zi = z_index(d->lifting_size);
  if (zi < 0) {
    return fail(FEC_ERR_BAD_ARG, "lifting_size=%u is not one of the 51 NR lifting sizes", d->lifting_size);
  }
  const uint32_t Z = d->lifting_size, Kb = (d->base_graph == 1) ? 22 : 10, nb = (d->base_graph == 1) ? 66 : 50;
  const uint32_t N = nb * Z, K = Kb * Z;
  d->graph_index   = (uint16_t)((d->base_graph - 1) * 51 + zi);
  if (d->max_iter == 0 || d->crc_type > 2 || d->mode > 2) {
    return fail(FEC_ERR_BAD_ARG, "max_iter=%u crc_type=%u mode=%u", d->max_iter, d->crc_type, d->mode);
  }
  if (d->nof_filler_bits >= (Kb - 2) * Z ||
      (!(d->flags & FEC_FLAG_NO_CRC) && d->mode != FEC_MODE_DEMATCH && K - d->nof_filler_bits <= CRC_LEN[d->crc_type])) {
    return fail(FEC_ERR_BAD_ARG, "nof_filler_bits=%u out of range", d->nof_filler_bits);
  }
  if (d->mode == FEC_MODE_DECODE) {
    if (d->E < K + 2 * Z || d->E > N) {
      return fail(FEC_ERR_BAD_ARG, "decode input length %u outside [%u, %u]", d->E, K + 2 * Z, N);
    }
    d->k0 = 0;
    return FEC_OK;
  }
  if (d->rv > 3 || (d->qm != 1 && d->qm != 2 && d->qm != 4 && d->qm != 6 && d->qm != 8) || d->E == 0 ||
      (d->E % d->qm) != 0) {
    return fail(FEC_ERR_BAD_ARG, "rv=%u qm=%u E=%u", d->rv, d->qm, d->E);
  }
  const uint32_t Ncb = (d->Nref > 0) ? (d->Nref < N ? d->Nref : N) : N;
  if (Ncb <= (Kb - 2) * Z) {
    return fail(FEC_ERR_BAD_ARG, "Nref=%u leaves no parity in the circular buffer", d->Nref);
  }
  // k0 as the reference computes it: floor(f * Ncb / N) * Z in double precision, truncated to 16 bits.
  static const double F1[4] = {0, 17, 33, 56};
  static const double F2[4] = {0, 13, 25, 43};
  const double        f     = (d->base_graph == 1) ? F1[d->rv] : F2[d->rv];
  d->k0                     = (uint32_t)(uint16_t)std::floor(f * (double)Ncb / (double)N) * Z;
  return FEC_OK;
}

int fec_engine_create(const fec_engine_cfg* cfg, fec_engine** out)
{
  if (cfg == nullptr || out == nullptr) {
    return fail(FEC_ERR_BAD_ARG, "null argument");
  }
  *out = nullptr;
  if (!(cfg->scaling_factor > 0.f && cfg->scaling_factor < 1.f) || cfg->max_batch_cbs == 0 || cfg->harq_slots == 0) {
    return fail(FEC_ERR_BAD_ARG, "scaling_factor must be in (0,1); max_batch_cbs and harq_slots must be > 0");
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= cfg->device || cfg->device < 0) {
    cudaGetLastError();
    return fail(FEC_ERR_NO_DEVICE, "CUDA device %d not available (%d devices): this engine has no CPU fallback",
                cfg->device, ndev);
  }
  CU_TRY(cudaSetDevice(cfg->device));
  fec_engine* e = new (std::nothrow) fec_engine();
  if (!e) {
    return fail(FEC_ERR_NO_MEMORY, "host allocation failed");
  }
  e->cfg = *cfg;
  if (e->cfg.max_batch_llrs == 0) {
    e->cfg.max_batch_llrs = e->cfg.max_batch_cbs * 25344u;
  }
  cudaDeviceProp prop;
  CU_TRY(cudaGetDeviceProperties(&prop, cfg->device));
  e->sm_count = prop.multiProcessorCount;
  CU_TRY(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking));
  for (auto& h : e->st_iter_hist) {
    h.store(0);
  }
  {
    // Largest dynamic shared-memory layouts (BG1, Z = 384), set once and never lowered: the attribute is global per function
    // and queues / host entry points launch concurrently from several threads.
    const uint32_t g = fec_generic_layout(384, 1).total, p = fec_packed_layout(384, 1).total;
    CU_TRY(cudaFuncSetAttribute(fec_cb_packed_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(g > p ? g : p)));
    CU_TRY(cudaFuncSetAttribute(fec_cb_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g));
  }
  fec_dev_tables* h = new (std::nothrow) fec_dev_tables();
  build_tables(*h, cfg->scaling_factor);
  CU_TRY(cudaMalloc((void**)&e->d_tab, sizeof(fec_dev_tables)));
  CU_TRY(cudaMemcpy(e->d_tab, h, sizeof(fec_dev_tables), cudaMemcpyHostToDevice));
  {
    fec_packed_tables* hp = new (std::nothrow) fec_packed_tables();
    e->packed_ok          = fec_packed_build_tables(*hp, *h, FEC_ALL_Z);
    // The row schedule lives in constant memory, i.e. it is shared by every engine of this process on this device. It is a
    // function of the base graphs alone, so every engine writes the same words -- but only a schedule that was actually built
    // may be written: an engine whose scaling factor the packed kernel cannot represent (packed_ok false above) used to skip
    // the builder and then upload the uninitialised array, which broke the packed launches of the engines created before it.
    uint32_t   rowinfo[2][48] = {};
    const bool rows_ok        = fec_packed_build_rowinfo(rowinfo, *h, *hp);
    e->packed_ok              = e->packed_ok && rows_ok;
    if (rows_ok) {
      CU_TRY(cudaMemcpyToSymbol(fec_c_rowinfo, rowinfo, sizeof(rowinfo)));
    }
    CU_TRY(cudaMalloc((void**)&e->d_ptab, sizeof(fec_packed_tables)));
    CU_TRY(cudaMalloc((void**)&e->d_work, FEC_WORK_SLOTS * 2 * sizeof(uint32_t)));
    CU_TRY(cudaMemset(e->d_work, 0, FEC_WORK_SLOTS * 2 * sizeof(uint32_t)));
    CU_TRY(cudaMemcpy(e->d_ptab, hp, sizeof(fec_packed_tables), cudaMemcpyHostToDevice));
    delete hp;
  }
  delete h;
  const size_t harq_bytes = (size_t)cfg->harq_slots * FEC_HARQ_SLOT_BYTES;
  if (cudaMalloc((void**)&e->d_harq, harq_bytes) != cudaSuccess) {
    cudaGetLastError();
    fec_engine_destroy(e);
    return fail(FEC_ERR_NO_MEMORY, "cannot reserve %zu bytes of HBM for %u HARQ slots", harq_bytes, cfg->harq_slots);
  }
  CU_TRY(cudaMemset(e->d_harq, 0, harq_bytes));
  e->fresh = new (std::nothrow) std::atomic<uint8_t>[cfg->harq_slots];
  for (uint32_t i = 0; i < cfg->harq_slots; ++i) {
    e->fresh[i].store(0);
  }
  if (e->cfg.host_contexts == 0) {
    e->cfg.host_contexts = 4;
  }
  for (uint32_t i = 0; i < e->cfg.host_contexts; ++i) {
    host_ctx* c  = nullptr;
    const int rc = ctx_create(e, i, &c);
    if (c) {
      e->ctxs.push_back(c);
    }
    if (rc != FEC_OK) {
      fec_engine_destroy(e);
      return rc;
    }
  }
  e->queues = new std::vector<queue_t>(cfg->nof_queues);
  *out      = e;
  return FEC_OK;
}

void fec_engine_destroy(fec_engine* e)
{
  if (!e) {
    return;
  }
  cudaSetDevice(e->cfg.device);
  if (e->tb_thread.joinable()) {
    {
      std::lock_guard<std::mutex> lk(e->tb_mtx);
      e->tb_stop = true;
    }
    e->tb_cv.notify_all();
    e->tb_thread.join(); // finishes what is in flight first
  }
  cudaDeviceSynchronize();
  if (e->queues) {
    for (auto& q : *e->queues) {
      for (batch_slot* s : q.slots) {
        slot_free(s);
      }
    }
    delete e->queues;
  }
  for (dispatch_res& r : e->disp) {
    cudaFreeHost(r.h_descs);
    cudaFree(r.d_descs);
    cudaFree(r.d_res);
    cudaFreeHost(r.h_res);
    if (r.done) {
      cudaEventDestroy(r.done);
    }
  }
  for (cudaStream_t st : e->disp_stream) {
    if (st) {
      cudaStreamDestroy(st);
    }
  }
  for (host_ctx* c : e->ctxs) {
    ctx_destroy(c);
  }
  cudaFree(e->d_harq);
  delete[] e->fresh;
  delete[] e->cb_ok;
  cudaFree(e->d_msg);
  cudaFree(e->d_enc_descs);
  cudaFree(e->d_enc_in);
  cudaFree(e->d_enc_out);
  cudaFree(e->d_dtab);
  cudaFree(e->d_dm_descs);
  cudaFree(e->d_dm_sym);
  cudaFree(e->d_dm_noise);
  cudaFree(e->d_dm_llr);
  cudaFreeHost(e->h_dm_descs);
  cudaFree(e->d_tab);
  cudaFree(e->d_ptab);
  cudaFree(e->d_work);
  for (auto& en : e->ulsch_cache) {
    cudaFree(en.d_segs);
    cudaFree(en.d_flags);
  }
  if (e->stream) {
    cudaStreamDestroy(e->stream);
  }
  delete e;
}

void* fec_engine_stream(fec_engine* e)
{
  return e ? (void*)e->stream : nullptr;
}

void* fec_harq_slot_ptr(fec_engine* e, uint32_t slot)
{
  if (!e || slot >= e->cfg.harq_slots) {
    return nullptr;
  }
  return e->d_harq + (size_t)slot * FEC_HARQ_SLOT_BYTES;
}

uint64_t fec_engine_launch_count(fec_engine* e)
{
  return e ? e->launches.load() : 0;
}

int fec_launch_device(fec_engine* e, const fec_cb_desc* d_descs, uint32_t n, const int8_t* d_llr, uint8_t* d_out,
                      fec_cb_result* d_results, uint32_t max_Z, int any_bg1)
{
  if (!e || !d_descs || !d_results) {
    return fail(FEC_ERR_BAD_ARG, "null argument");
  }
  // The descriptors are in device memory: whether a lifting size outside the packed kernel is among them is the caller's
  // knowledge (FEC_LAUNCH_EVEN_Z); without it the generic kernel is launched beside the packed one, to be safe.
  const int flags = ((any_bg1 & 1) ? FEC_LF_BG1 : 0) | ((any_bg1 & FEC_LAUNCH_EVEN_Z) ? 0 : FEC_LF_SMALL_Z);
  return launch(e, e->stream, d_descs, n, d_llr, d_out, d_results, max_Z, flags);
}

int fec_run_host(fec_engine* e, fec_cb_desc* h_descs, uint32_t n, const int8_t* h_llr, size_t llr_bytes, uint8_t* h_out,
                 size_t out_bytes, fec_cb_result* h_results, int8_t* h_soft_out)
{
  if (!e || !h_descs || !h_results || (llr_bytes && !h_llr) || (out_bytes && !h_out)) {
    return fail(FEC_ERR_BAD_ARG, "null argument");
  }
  if (n == 0) {
    return FEC_OK;
  }
  if (n > e->cfg.max_batch_cbs || llr_bytes > e->cfg.max_batch_llrs ||
      out_bytes > (size_t)e->cfg.max_batch_cbs * FEC_MAX_CB_OUT_BYTES) {
    return fail(FEC_ERR_BAD_ARG, "batch of %u CBs / %zu LLR bytes / %zu output bytes exceeds the engine capacity", n,
                llr_bytes, out_bytes);
  }
  static const bool trace = getenv("FEC_B200_TRACE") != nullptr; // host timestamps of the call's phases on stderr
  static std::vector<cudaEvent_t> trace_ev;                      // and, per chunk, when its kernel started and ended
  static std::vector<cudaEvent_t> trace_cp;                      // and when its input copy was complete
  const auto        t_in  = std::chrono::steady_clock::now();
  uint32_t max_Z = 0;
  int      any1  = 0;
  bool     once  = e->packed_ok && e->cfg.kernel_variant != 1; // every input is read once (zero copy is possible)
  for (uint32_t i = 0; i < n; ++i) {
    // A descriptor with the parameters of its predecessor (the usual case: the codeblocks of one transport block) inherits
    // its checked and derived fields instead of going through fec_desc_prepare again.
    fec_cb_desc& cur = h_descs[i];
    if (i > 0 && same_shape(cur, h_descs[i - 1])) {
      cur.k0          = h_descs[i - 1].k0;
      cur.graph_index = h_descs[i - 1].graph_index;
    } else {
      int r = fec_desc_prepare(&cur);
      if (r != FEC_OK) {
        return r;
      }
    }
    const fec_cb_desc& d  = h_descs[i];
    const uint32_t     kb = (((d.base_graph == 1) ? 22u : 10u) * d.lifting_size + 7) / 8;
    if (d.reserved1 != 0 || d.llr_offset + d.E > llr_bytes || (d.mode != FEC_MODE_DEMATCH && (size_t)d.out_offset + kb > out_bytes)) {
      return fail(FEC_ERR_BAD_ARG, "descriptor %u indexes outside the batch buffers", i);
    }
    if (d.mode != FEC_MODE_DECODE && d.harq_slot >= e->cfg.harq_slots) {
      return fail(FEC_ERR_BAD_ARG, "descriptor %u: harq_slot %u >= %u", i, d.harq_slot, e->cfg.harq_slots);
    }
    max_Z = d.lifting_size > max_Z ? d.lifting_size : max_Z;
    any1 |= launch_flags_of(d.base_graph, d.lifting_size);
    once = once && reads_input_once(d);
  }
  ctx_lease lease(e, -1);
  host_ctx& c      = *lease.c;
  const auto t_prep = std::chrono::steady_clock::now();
  CU_TRY(cudaSetDevice(e->cfg.device));
  // Large batches are cut into chunks of one kernel wave (the resident CTAs of the packed kernel: 3 per SM) that travel on three streams, so that the H2D copy of one
  // chunk overlaps the kernels of the previous one and the D2H copy of the one before (PCIe is full duplex).
  const uint32_t wave  = FEC_PACKED_CTAS_PER_SM * (uint32_t)e->sm_count;
  uint32_t       chunk = (n >= 2 * wave) ? wave : n; // measured on B200: 1 wave beats 2, 3, 4 and 6 (e2e 11.4 vs 10.8 Gbit/s at 3)
  // (also measured: 1/2, 3/4, 5/4, 3/2 and 2 waves per chunk, and a finer-cut first wave: all 1-3 % slower end to end; the last
  // wave cut 1/2, 1/4, 1/4 so that less work is left when the last input byte arrives: no change, 15.70 vs 15.71 Gbit/s -- the
  // SMs still hold a chunk and a half of backlog at that moment, FEC_B200_TRACE shows it)
  // The first wave is cut finer (1/4, 1/4, 1/2 of a wave): the first kernel starts after a quarter of the first wave's copy
  // instead of all of it; from then on the copy engine stays ahead of the SMs.
  // FEC_B200_CHUNK_WAVES (measurements): waves per chunk after the first wave
  static const double chunk_waves  = getenv("FEC_B200_CHUNK_WAVES") ? atof(getenv("FEC_B200_CHUNK_WAVES")) : 1.0;
  const uint32_t      chunk_steady = std::max<uint32_t>(1u, (uint32_t)(chunk_waves * wave));
  const auto chunk_len = [&](uint32_t c0) -> uint32_t {
    if (chunk == n) {
      return n;
    }
    const uint32_t q   = wave / 4;
    const uint32_t len = (c0 < 2 * q) ? q : (c0 < wave) ? wave - 2 * q : chunk_steady;
    return (n - c0 < len) ? (n - c0) : len;
  };
  if (chunk != n) {
    // Chunks may only run concurrently if their slices of the batch buffers do not interleave.
    size_t prev_l1 = 0, prev_o1 = 0;
    for (uint32_t c0 = 0, cn = 0; c0 < n && chunk != n; c0 += cn) {
      cn = chunk_len(c0);
      size_t         l0 = ~(size_t)0, l1 = 0, o0 = ~(size_t)0, o1 = 0;
      for (uint32_t i = c0; i < c0 + cn; ++i) {
        const fec_cb_desc& d = h_descs[i];
        l0                   = d.llr_offset < l0 ? (size_t)d.llr_offset : l0;
        l1                   = d.llr_offset + d.E > l1 ? (size_t)(d.llr_offset + d.E) : l1;
        const uint32_t     kb = (((d.base_graph == 1) ? 22u : 10u) * d.lifting_size + 7) / 8;
        o0                   = d.out_offset < o0 ? (size_t)d.out_offset : o0;
        o1                   = (size_t)d.out_offset + kb > o1 ? (size_t)d.out_offset + kb : o1;
      }
      if (l0 < prev_l1 || o0 < prev_o1) {
        chunk = n;
      }
      prev_l1 = l1;
      prev_o1 = o1;
    }
  }
  memcpy(c.h_descs_pin, h_descs, n * sizeof(fec_cb_desc));
  // FEC_FLAG_HARQ_ZERO goes into the engine's copy only: the caller's array gets k0 / graph_index written back and nothing
  // else, so it can be reused for a retransmission (rv, flags edited) without carrying a stale "slot reads as zeros".
  run_guard guard(e, &c);
  guard.fail_results = h_results;
  guard.fail_descs   = h_descs;
  guard.fail_n       = n;
  for (uint32_t i = 0; i < n; ++i) {
    apply_fresh(e, c.h_descs_pin[i], &guard.taken);
  }
  CU_TRY(cudaMemcpyAsync(c.d_descs, c.h_descs_pin, n * sizeof(fec_cb_desc), cudaMemcpyHostToDevice, c.stream));
  CU_TRY(cudaEventRecord(c.pipe_done[0], c.stream));
  for (int k = 0; k < 3; ++k) {
    CU_TRY(cudaStreamWaitEvent(c.pipe[k], c.pipe_done[0], 0));
  }
  uint32_t nchunks = 0;
  for (uint32_t c0 = 0, cn = 0; c0 < n; c0 += cn, ++nchunks) {
    cn              = chunk_len(c0);
    cudaStream_t st = (chunk == n) ? c.stream : c.pipe[nchunks % 3];
    // Chunked calls send every input through ONE copy stream, in chunk order: copies issued on the three kernel streams
    // were observed to complete out of order (chunk 0 third), which delays the first kernel by two chunk copies.
    cudaStream_t in = (chunk == n) ? c.stream : c.copy_in;
    // byte ranges of the batch buffers this chunk touches
    size_t   l0 = ~(size_t)0, l1 = 0, o0 = ~(size_t)0, o1 = 0;
    uint32_t cz = 0;
    int      c1 = 0;
    // Only an early-stop operation can leave its output bytes as they were; when none is in the chunk and the outputs tile
    // the span without gaps, every byte of it is overwritten and the caller's bytes need not travel to the device first.
    bool   keep     = false;
    size_t next_out = h_descs[c0].out_offset;
    for (uint32_t i = c0; i < c0 + cn; ++i) {
      const fec_cb_desc& d  = h_descs[i];
      const uint32_t     kb = (((d.base_graph == 1) ? 22u : 10u) * d.lifting_size + 7) / 8;
      l0                    = d.llr_offset < l0 ? (size_t)d.llr_offset : l0;
      l1                    = d.llr_offset + d.E > l1 ? (size_t)(d.llr_offset + d.E) : l1;
      if (d.mode != FEC_MODE_DEMATCH) {
        o0 = d.out_offset < o0 ? (size_t)d.out_offset : o0;
        o1 = (size_t)d.out_offset + kb > o1 ? (size_t)d.out_offset + kb : o1;
        keep |= (d.out_offset != next_out);
        next_out = (size_t)d.out_offset + kb;
      }
      cz = d.lifting_size > cz ? d.lifting_size : cz;
      c1 |= launch_flags_of(d.base_graph, d.lifting_size);
      keep |= (d.flags & FEC_FLAG_EARLY_STOP) != 0;
    }
    // A call that is not chunked (less than two kernel waves) reads pinned inputs and writes pinned outputs in place.
    // Only when no other host call is in flight: measured with several calls at once, kernels that wait for PCIe reads hold
    // the SMs' codeblock slots and the copy engine's pipeline is the faster one (4 asynchronous callers: 32 -> 11 Gbit/s).
    const bool    alone  = e->ctx_busy.load(std::memory_order_relaxed) <= 1;
    const int8_t* zc_llr = (chunk == n && once && alone) ? static_cast<const int8_t*>(pinned_device_ptr(h_llr)) : nullptr;
    // (chunked calls with their outputs written in place as well: measured, no change -- 15.65 vs 15.65 Gbit/s)
    uint8_t*      zc_out = (chunk == n && once && alone && out_bytes)
                               ? static_cast<uint8_t*>(const_cast<void*>(pinned_device_ptr(h_out)))
                               : nullptr;
    if (!zc_llr) {
      CU_TRY(cudaMemcpyAsync(c.d_llr + l0, h_llr + l0, l1 - l0, cudaMemcpyHostToDevice, in));
    }
    if (!zc_out && out_bytes && o1 > o0 && keep) {
      // Early-stop operations that bail out leave their output untouched: start from the caller's bytes.
      CU_TRY(cudaMemcpyAsync(c.d_out + o0, h_out + o0, o1 - o0, cudaMemcpyHostToDevice, in));
    }
    if (trace && nchunks < 32) {
      if (trace_cp.empty()) {
        trace_cp.resize(32);
        for (cudaEvent_t& ev : trace_cp) {
          cudaEventCreate(&ev);
        }
      }
      cudaEventRecord(trace_cp[nchunks], in); // this chunk's input has arrived
    }
    if (in != st) {
      if (c.chunk_in.size() <= nchunks) {
        cudaEvent_t ev = nullptr;
        CU_TRY(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        c.chunk_in.push_back(ev);
      }
      CU_TRY(cudaEventRecord(c.chunk_in[nchunks], in));
      CU_TRY(cudaStreamWaitEvent(st, c.chunk_in[nchunks], 0));
    }
    if (trace && nchunks < 32) {
      if (trace_ev.empty()) {
        trace_ev.resize(65);
        for (cudaEvent_t& ev : trace_ev) {
          cudaEventCreate(&ev);
        }
      }
      if (nchunks == 0) {
        cudaEventRecord(trace_ev[64], c.stream); // after the descriptor copy
      }
      cudaEventRecord(trace_ev[2 * nchunks], st);
    }
    // FEC_B200_CHUNK_GRID_DIV (measurements): chunk kernels on fewer CTAs so that consecutive chunks interleave on the SMs
    static const uint32_t chunk_grid_div = getenv("FEC_B200_CHUNK_GRID_DIV") ? (uint32_t)atoi(getenv("FEC_B200_CHUNK_GRID_DIV")) : 1u;
    int r = launch(e, st, c.d_descs + c0, cn, zc_llr ? zc_llr : c.d_llr, zc_out ? zc_out : c.d_out, c.d_res + c0, cz, c1,
                   (chunk != n) ? chunk_grid_div : 1u);
    if (r != FEC_OK) {
      return r;
    }
    if (trace && nchunks < 32) {
      cudaEventRecord(trace_ev[2 * nchunks + 1], st);
    }
    if (!zc_out && out_bytes && o1 > o0) {
      CU_TRY(cudaMemcpyAsync(h_out + o0, c.d_out + o0, o1 - o0, cudaMemcpyDeviceToHost, st));
    }
    CU_TRY(cudaMemcpyAsync(c.h_res_pin + c0, c.d_res + c0, cn * sizeof(fec_cb_result), cudaMemcpyDeviceToHost, st));
  }
  if (chunk != n) {
    for (int k = 0; k < 3; ++k) {
      CU_TRY(cudaEventRecord(c.pipe_done[k], c.pipe[k]));
      CU_TRY(cudaStreamWaitEvent(c.stream, c.pipe_done[k], 0));
    }
  }
  if (h_soft_out) {
    for (uint32_t i = 0; i < n; ++i) {
      if (h_descs[i].mode != FEC_MODE_DECODE) {
        CU_TRY(cudaMemcpyAsync(h_soft_out + (size_t)i * FEC_HARQ_SLOT_BYTES,
                               e->d_harq + (size_t)h_descs[i].harq_slot * FEC_HARQ_SLOT_BYTES, FEC_HARQ_SLOT_BYTES,
                               cudaMemcpyDeviceToHost, c.stream));
      }
    }
  }
  const auto t_enq = std::chrono::steady_clock::now();
  CU_TRY(cudaStreamSynchronize(c.stream));
  const auto t_sync = std::chrono::steady_clock::now();
  memcpy(h_results, c.h_res_pin, n * sizeof(fec_cb_result));
  guard.ok = true;
  {
    stats_batch sb;
    for (uint32_t i = 0; i < n; ++i) {
      sb.add(h_results[i]);
    }
    sb.flush(e);
  }
  if (trace) {
    const auto us = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
      return std::chrono::duration<double, std::micro>(b - a).count();
    };
    fprintf(stderr, "fec_b200 trace: run_host n=%u chunks=%u prepare %.1f us, enqueue %.1f us, wait %.1f us, results %.1f us\n", n,
            nchunks, us(t_in, t_prep), us(t_prep, t_enq), us(t_enq, t_sync), us(t_sync, std::chrono::steady_clock::now()));
    if (nchunks > 1 && !trace_ev.empty()) {
      fprintf(stderr, "fec_b200 trace: chunk kernels [start, end] in us after the descriptor copy:");
      for (uint32_t c = 0; c < nchunks && c < 32; ++c) {
        float a = 0, b = 0;
        cudaEventElapsedTime(&a, trace_ev[64], trace_ev[2 * c]);
        cudaEventElapsedTime(&b, trace_ev[64], trace_ev[2 * c + 1]);
        fprintf(stderr, " [%.0f, %.0f]", a * 1e3f, b * 1e3f);
      }
      fprintf(stderr, "\n");
      fprintf(stderr, "fec_b200 trace: chunk inputs complete at:");
      for (uint32_t c = 0; c < nchunks && c < 32 && !trace_cp.empty(); ++c) {
        float a = 0;
        cudaEventElapsedTime(&a, trace_ev[64], trace_cp[c]);
        fprintf(stderr, " %.0f", a * 1e3f);
      }
      fprintf(stderr, "\n");
    }
  }
  return FEC_OK;
}

void* fec_host_alloc(size_t bytes)
{
  void* p = nullptr;
  if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  return p;
}

void fec_host_free(void* p)
{
  if (p) {
    cudaFreeHost(p);
  }
}

int fec_harq_free(fec_engine* e, uint32_t slot)
{
  if (!e || slot >= e->cfg.harq_slots) {
    return fail(FEC_ERR_BAD_ARG, "bad HARQ slot %u", slot);
  }
  e->fresh[slot].store(1);
  return FEC_OK;
}

int fec_harq_flush(fec_engine* e)
{
  if (!e) {
    return fail(FEC_ERR_BAD_ARG, "null engine");
  }
  CU_TRY(cudaSetDevice(e->cfg.device));
  CU_TRY(cudaDeviceSynchronize());
  for (uint32_t s = 0; s < e->cfg.harq_slots; ++s) {
    if (e->fresh[s].exchange(0)) {
      CU_TRY(cudaMemsetAsync(e->d_harq + (size_t)s * FEC_HARQ_SLOT_BYTES, 0, FEC_HARQ_SLOT_BYTES, e->stream));
    }
  }
  CU_TRY(cudaStreamSynchronize(e->stream));
  return FEC_OK;
}

int fec_harq_read(fec_engine* e, uint32_t slot, int8_t* h_dst, uint32_t n)
{
  if (!e || slot >= e->cfg.harq_slots || n > FEC_HARQ_SLOT_BYTES || !h_dst) {
    return fail(FEC_ERR_BAD_ARG, "bad HARQ read");
  }
  if (e->fresh[slot].load()) {
    memset(h_dst, 0, n);
    return FEC_OK;
  }
  CU_TRY(cudaSetDevice(e->cfg.device));
  CU_TRY(cudaDeviceSynchronize());
  CU_TRY(cudaMemcpy(h_dst, e->d_harq + (size_t)slot * FEC_HARQ_SLOT_BYTES, n, cudaMemcpyDeviceToHost));
  return FEC_OK;
}

int fec_harq_write(fec_engine* e, uint32_t slot, const int8_t* h_src, uint32_t n)
{
  if (!e || slot >= e->cfg.harq_slots || n > FEC_HARQ_SLOT_BYTES || !h_src) {
    return fail(FEC_ERR_BAD_ARG, "bad HARQ write");
  }
  CU_TRY(cudaSetDevice(e->cfg.device));
  CU_TRY(cudaDeviceSynchronize());
  if (e->fresh[slot].exchange(0)) {
    CU_TRY(cudaMemset(e->d_harq + (size_t)slot * FEC_HARQ_SLOT_BYTES, 0, FEC_HARQ_SLOT_BYTES));
  }
  CU_TRY(cudaMemcpy(e->d_harq + (size_t)slot * FEC_HARQ_SLOT_BYTES, h_src, n, cudaMemcpyHostToDevice));
  return FEC_OK;
}

/* ---------------------------------------------------------------------------------------------------- queues */

int fec_queue_reserve(fec_engine* e, uint32_t* queue_id)
{
  if (!e || !queue_id) {
    return fail(FEC_ERR_BAD_ARG, "null argument");
  }
  std::lock_guard<std::mutex> lk(e->queues_mtx);
  int rc = disp_resources(e);
  if (rc != FEC_OK) {
    return rc;
  }
  for (uint32_t i = 0; i < e->queues->size(); ++i) {
    queue_t& q = (*e->queues)[i];
    if (!q.reserved) {
      if (!q.inited) {
        CU_TRY(cudaSetDevice(e->cfg.device));
        const uint32_t nb = e->cfg.queue_batches ? e->cfg.queue_batches : 4u;
        for (uint32_t k = 0; k < nb; ++k) {
          batch_slot* s = new (std::nothrow) batch_slot();
          if (!s) {
            return fail(FEC_ERR_NO_MEMORY, "host allocation failed");
          }
          q.slots.push_back(s);
          rc = slot_alloc(e, *s);
          if (rc != FEC_OK) {
            return rc;
          }
        }
        q.inited = true;
      }
      q.reserved = true;
      *queue_id  = i;
      return FEC_OK;
    }
  }
  return fail(FEC_ERR_QUEUE_FULL, "all %zu queues are reserved", e->queues->size());
}

int fec_queue_free(fec_engine* e, uint32_t queue_id)
{
  if (!e || queue_id >= e->queues->size()) {
    return fail(FEC_ERR_BAD_QUEUE, "bad queue %u", queue_id);
  }
  queue_t&                    q = (*e->queues)[queue_id];
  std::lock_guard<std::mutex> lk(q.mtx);
  for (batch_slot* s : q.slots) {
    // whatever was handed to the batcher gets dispatched and finishes before the batch is recycled
    while (s->state.load(std::memory_order_acquire) == batch_slot::PENDING) {
      dispatch_try(e);
      std::this_thread::yield();
    }
    if (s->state.load(std::memory_order_acquire) == batch_slot::INFLIGHT && s->harvest.load(std::memory_order_acquire) != 2) {
      cudaEventSynchronize(s->disp->done); // (a later record of the same event only waits longer)
      if (!s->failed.load()) {
        harvest_batch(*s);
      }
    }
    s->state.store(batch_slot::FREE);
    s->n = s->llr_used = s->out_used = s->polled = 0;
  }
  q.filling  = -1;
  q.depth    = 0;
  q.reserved = false;
  return FEC_OK;
}

namespace {
// Under q.mtx. Stages one prepared operation; FEC_ERR_QUEUE_FULL when every batch of the ring is taken.
int queue_stage(fec_engine* e, uint32_t queue_id, queue_t& q, fec_cb_desc& d, const int8_t* llr, uint64_t tag)
{
  const uint32_t kb      = (((d.base_graph == 1) ? 22u : 10u) * d.lifting_size + 7) / 8;
  const uint32_t llr_pad = (d.E + 15u) & ~15u;
  const uint32_t cap_cbs = queue_batch_cbs(e);
  const size_t   cap_llr = queue_batch_llrs(e);
  if (llr_pad > cap_llr) {
    return fail(FEC_ERR_BAD_ARG, "E=%u exceeds the staging capacity of a queue batch", d.E);
  }
  for (int attempt = 0; attempt < 2; ++attempt) {
    if (q.filling < 0) {
      // Recycle the oldest fully polled batch, or take a free one.
      int      pick = -1;
      uint64_t best = ~0ull;
      for (size_t i = 0; i < q.slots.size(); ++i) {
        batch_slot& s  = *q.slots[i];
        const int   st = s.state.load(std::memory_order_acquire);
        if (st == batch_slot::FREE) {
          pick = (int)i;
          break;
        }
        if (st == batch_slot::DONE && s.polled == s.n && s.seq < best) {
          best = s.seq;
          pick = (int)i;
        }
      }
      if (pick < 0) {
        e->st_queue_full.fetch_add(1, std::memory_order_relaxed);
        return fail(FEC_ERR_QUEUE_FULL, "queue %u: all batches are staged or in flight", queue_id);
      }
      batch_slot& s = *q.slots[pick];
      s.state.store(batch_slot::FILLING);
      s.n = s.llr_used = s.out_used = s.polled = 0;
      s.max_Z                                  = 0;
      s.any_bg1                                = 0;
      s.seq                                    = q.next_seq++;
      q.filling                                = pick;
    }
    batch_slot& s = *q.slots[q.filling];
    if (s.n < cap_cbs && s.llr_used + llr_pad <= cap_llr) {
      d.llr_offset = s.llr_used;
      d.out_offset = s.out_used;
      d.reserved1  = 0;
      memcpy(s.h_llr + s.llr_used, llr, d.E);
      apply_fresh(e, d); // only once the operation is certain to be accepted
      s.h_descs[s.n] = d;
      s.tags[s.n]    = tag;
      s.h_res[s.n].status = 0;
      s.n++;
      s.llr_used += llr_pad;
      s.out_used += (kb + 15u) & ~15u;
      s.max_Z = d.lifting_size > s.max_Z ? d.lifting_size : s.max_Z;
      s.any_bg1 |= launch_flags_of(d.base_graph, d.lifting_size);
      ++q.depth;
      uint64_t seen = e->st_max_depth.load(std::memory_order_relaxed);
      while (q.depth > seen && !e->st_max_depth.compare_exchange_weak(seen, q.depth)) {
      }
      if (s.n == cap_cbs) {
        q.filling = -1;
        queue_submit_slot(e, s); // a full batch goes out while the consumer is still enqueuing
      }
      return FEC_OK;
    }
    // Current batch cannot take it: hand it over and try a fresh one.
    q.filling = -1;
    queue_submit_slot(e, s);
  }
  e->st_queue_full.fetch_add(1, std::memory_order_relaxed);
  return fail(FEC_ERR_QUEUE_FULL, "queue %u: no room", queue_id);
}
} // namespace

int fec_queue_enqueue_batch(fec_engine* e, uint32_t queue_id, const fec_cb_desc* descs, const int8_t* const* llrs,
                            const uint64_t* tags, uint32_t n, uint32_t* accepted)
{
  if (accepted) {
    *accepted = 0;
  }
  if (!e || queue_id >= e->queues->size() || !descs || !llrs || !tags || !accepted) {
    return fail(FEC_ERR_BAD_QUEUE, "bad queue %u or null argument", queue_id);
  }
  queue_t&                    q = (*e->queues)[queue_id];
  std::lock_guard<std::mutex> lk(q.mtx);
  if (!q.reserved) {
    return fail(FEC_ERR_BAD_QUEUE, "queue %u is not reserved", queue_id);
  }
  fec_cb_desc prev = {};
  for (uint32_t i = 0; i < n; ++i) {
    fec_cb_desc d = descs[i];
    // the codeblocks of one transport block share their parameters: checked and derived fields are inherited
    if (i > 0 && same_shape(d, descs[i - 1])) {
      d.k0          = prev.k0;
      d.graph_index = prev.graph_index;
    } else {
      const int r = fec_desc_prepare(&d);
      if (r != FEC_OK) {
        return r;
      }
    }
    prev = d;
    if (d.mode != FEC_MODE_DECODE && d.harq_slot >= e->cfg.harq_slots) {
      return fail(FEC_ERR_BAD_ARG, "operation %u: harq_slot %u >= %u", i, d.harq_slot, e->cfg.harq_slots);
    }
    if (!llrs[i]) {
      return fail(FEC_ERR_BAD_ARG, "operation %u: null LLR pointer", i);
    }
    const int r = queue_stage(e, queue_id, q, d, llrs[i], tags[i]);
    if (r == FEC_ERR_QUEUE_FULL) {
      return (*accepted != 0) ? FEC_OK : r; // the caller drains and comes back with the rest
    }
    if (r != FEC_OK) {
      return r;
    }
    ++*accepted;
  }
  return FEC_OK;
}

int fec_queue_enqueue(fec_engine* e, uint32_t queue_id, const fec_cb_desc* desc, const int8_t* llr, uint64_t tag)
{
  uint32_t  accepted = 0;
  const int r        = fec_queue_enqueue_batch(e, queue_id, desc, &llr, &tag, 1, &accepted);
  return (r == FEC_OK && accepted == 0) ? FEC_ERR_QUEUE_FULL : r;
}

int fec_queue_flush(fec_engine* e, uint32_t queue_id)
{
  if (!e || queue_id >= e->queues->size()) {
    return fail(FEC_ERR_BAD_QUEUE, "bad queue %u", queue_id);
  }
  queue_t& q = (*e->queues)[queue_id];
  {
    std::lock_guard<std::mutex> lk(q.mtx);
    if (q.filling >= 0) {
      batch_slot& s = *q.slots[q.filling];
      q.filling     = -1;
      if (s.n > 0) {
        queue_submit_slot(e, s);
        return FEC_OK;
      }
      s.state.store(batch_slot::FREE);
    }
  }
  dispatch_try(e);
  return FEC_OK;
}

int fec_queue_poll(fec_engine* e, uint32_t queue_id, uint64_t* tags, fec_cb_result* results, uint32_t max)
{
  if (!e || queue_id >= e->queues->size() || !tags || !results) {
    return fail(FEC_ERR_BAD_QUEUE, "bad queue %u or null argument", queue_id);
  }
  dispatch_try(e); // batches left with a busy batcher (this queue's or another's) go out now
  queue_t&                    q = (*e->queues)[queue_id];
  std::lock_guard<std::mutex> lk(q.mtx);
  uint32_t                    got = 0;
  // Report in submission order.
  for (size_t pass = 0; pass < q.slots.size() && got < max; ++pass) {
    int      pick = -1;
    uint64_t best = ~0ull;
    for (size_t i = 0; i < q.slots.size(); ++i) {
      batch_slot& s  = *q.slots[i];
      const int   st = s.state.load(std::memory_order_acquire);
      if ((st == batch_slot::PENDING || st == batch_slot::INFLIGHT || (st == batch_slot::DONE && s.polled < s.n)) && s.seq < best) {
        best = s.seq;
        pick = (int)i;
      }
    }
    if (pick < 0) {
      break;
    }
    batch_slot& s = *q.slots[pick];
    if (s.state.load(std::memory_order_acquire) == batch_slot::PENDING) {
      break; // not dispatched yet
    }
    if (s.state.load(std::memory_order_acquire) == batch_slot::INFLIGHT) {
      cudaError_t st = cudaSuccess;
      if (s.harvest.load(std::memory_order_acquire) != 2) {
        // (If the batcher reuses the set meanwhile it hands the results over first and only then records the event again:
        // "not ready" from the newer record just means one more look, at which harvest reads 2.)
        st = cudaEventQuery(s.disp->done);
        if (st == cudaErrorNotReady) {
          break;
        }
        if (st == cudaSuccess && !s.failed.load()) {
          harvest_batch(s);
        }
      }
      if (st != cudaSuccess || s.failed.load()) {
        // A device fault must surface as failed codeblocks, never as a hang (the consumer's poll loop has no timeout).
        cudaGetLastError();
        // (the batcher must not hand this batch "results" of the failed dispatch later: take the hand-over ourselves)
        int expected = 0;
        if (!s.harvest.compare_exchange_strong(expected, 1, std::memory_order_acq_rel)) {
          while (s.harvest.load(std::memory_order_acquire) != 2) {
            std::this_thread::yield();
          }
        }
        for (uint32_t i = 0; i < s.n; ++i) {
          s.h_res[i].status              = FEC_RES_DONE;
          s.h_res[i].nof_ldpc_iterations = s.h_descs[i].max_iter;
          s.h_res[i].nof_layers          = 0;
          s.h_res[i].reserved            = 0;
        }
        s.harvest.store(2, std::memory_order_release);
      }
      s.state.store(batch_slot::DONE);
    }
    while (s.polled < s.n && got < max) {
      tags[got]    = s.tags[s.polled];
      results[got] = s.h_res[s.polled];
      stats_result(e, results[got]);
      ++got;
      ++s.polled;
      --q.depth;
    }
  }
  return (int)got;
}

int fec_queue_fetch(fec_engine* e, uint32_t queue_id, uint64_t tag, uint8_t* dst, uint32_t nbytes)
{
  if (!e || queue_id >= e->queues->size() || !dst) {
    return fail(FEC_ERR_BAD_QUEUE, "bad queue %u or null argument", queue_id);
  }
  queue_t&                    q = (*e->queues)[queue_id];
  std::lock_guard<std::mutex> lk(q.mtx);
  // Newest batch first: tags are caller-chosen (the HAL uses the codeblock index) and repeat from one transport block to
  // the next, while an older, fully polled batch may still sit in its slot.
  std::vector<char> seen(q.slots.size(), 0);
  for (size_t pass = 0; pass < q.slots.size(); ++pass) {
    int      pick = -1;
    uint64_t best = 0;
    for (size_t i = 0; i < q.slots.size(); ++i) {
      if (!seen[i] && q.slots[i]->state.load() == batch_slot::DONE && q.slots[i]->seq >= best) {
        best = q.slots[i]->seq;
        pick = (int)i;
      }
    }
    if (pick < 0) {
      break;
    }
    seen[pick]    = 1;
    batch_slot& s = *q.slots[pick];
    for (uint32_t i = s.polled; i-- > 0;) {
      if (s.tags[i] == tag) {
        const fec_cb_desc& d = s.h_descs[i];
        if ((s.h_res[i].status & FEC_RES_TOO_SHORT) && (d.flags & FEC_FLAG_EARLY_STOP)) {
          return FEC_OK; // reference leaves the output untouched in this case
        }
        memcpy(dst, s.h_out + d.out_offset, nbytes);
        return FEC_OK;
      }
    }
  }
  return fail(FEC_ERR_BAD_ARG, "tag not found among completed operations of queue %u", queue_id);
}

} // extern "C"

// ---- transport-block layer ----------------------------------------------------------------------------------------------
namespace {

int tb_state_alloc(fec_engine* e, host_ctx& hx)
{
  CU_TRY(cudaSetDevice(e->cfg.device));
  {
    std::lock_guard<std::mutex> lk(e->tb_alloc_mtx);
    if (e->d_msg == nullptr) {
      if ((size_t)e->cfg.harq_slots * FEC_MAX_CB_OUT_BYTES > 0xffffffffull) {
        return fail(FEC_ERR_BAD_ARG, "transport-block interface: %u HARQ slots exceed the 32-bit message-store offsets",
                    e->cfg.harq_slots);
      }
      e->cb_ok = new (std::nothrow) std::atomic<uint8_t>[e->cfg.harq_slots];
      for (uint32_t i = 0; i < e->cfg.harq_slots; ++i) {
        e->cb_ok[i].store(0);
      }
      CU_TRY(cudaMalloc((void**)&e->d_msg, (size_t)e->cfg.harq_slots * FEC_MAX_CB_OUT_BYTES));
    }
  }
  if (hx.h_meta != nullptr) {
    return FEC_OK;
  }
  const uint32_t nc   = e->cfg.max_batch_cbs;
  const size_t   meta = (size_t)nc * (sizeof(fec_tb_job) + sizeof(fec_tb_cb_ref) + 8 + sizeof(fec_cb_desc)) + 64;
  const size_t   rblk = (size_t)nc * (sizeof(fec_cb_result) + sizeof(fec_tb_result)) + 64;
  CU_TRY(cudaMalloc((void**)&hx.d_meta, meta));
  CU_TRY(cudaMalloc((void**)&hx.d_rblk, rblk));
  CU_TRY(cudaHostAlloc((void**)&hx.h_rblk, rblk, cudaHostAllocDefault));
  CU_TRY(cudaHostAlloc((void**)&hx.h_tbout, (size_t)nc * FEC_MAX_CB_OUT_BYTES, cudaHostAllocDefault));
  CU_TRY(cudaHostAlloc((void**)&hx.h_meta, meta, cudaHostAllocDefault)); // last: marks the context as allocated
  return FEC_OK;
}

uint32_t tb_slot_of(const fec_tb_desc& tb, const uint32_t* slots, uint32_t r)
{
  return (tb.slot_list != FEC_TB_SLOTS_CONTIGUOUS && slots != nullptr) ? slots[tb.slot_list + r] : tb.harq_slot + r;
}

// Codeblock r of a transport block as a CHAIN descriptor (not yet prepared). cw_off = sum of the rate-matched lengths
// of the codeblocks before it.
void tb_cb_desc(const fec_tb_desc& tb, const fec_tb_geometry& g, uint32_t r, uint32_t cw_off, uint32_t slot, fec_cb_desc& d)
{
  memset(&d, 0, sizeof(d));
  d.llr_offset      = tb.llr_offset + cw_off;
  d.out_offset      = slot * FEC_MAX_CB_OUT_BYTES;
  d.harq_slot       = slot;
  d.E               = fec_tb_rm_length(tb, g, r);
  d.Nref            = tb.Nref;
  d.nof_filler_bits = g.nof_filler_bits;
  d.lifting_size    = (uint16_t)g.lifting_size;
  d.base_graph      = tb.base_graph;
  d.rv              = tb.rv;
  d.qm              = tb.qm;
  d.crc_type        = (uint8_t)g.crc_type;
  d.max_iter        = tb.max_iter;
  d.mode            = FEC_MODE_CHAIN;
  d.flags           = tb.flags & (FEC_FLAG_NEW_DATA | FEC_FLAG_EARLY_STOP);
}

} // namespace

extern "C" {

int fec_tb_segment(const fec_tb_desc* tb, const uint32_t* slots, fec_cb_desc* cbs, uint32_t* nof_cbs, uint32_t* payload_bits)
{
  if (!tb || !cbs || !nof_cbs) {
    return fail(FEC_ERR_BAD_ARG, "null argument");
  }
  fec_tb_geometry g;
  if (!fec_tb_geometry_of(*tb, g)) {
    return fail(FEC_ERR_BAD_ARG, "transport block: tbs=%u bg=%u rv=%u qm=%u layers=%u llrs=%u", tb->tbs, tb->base_graph,
                tb->rv, tb->qm, tb->nof_layers, tb->nof_llrs);
  }
  uint32_t cw_off = 0, stream = 0;
  for (uint32_t r = 0; r < g.nof_cbs; ++r) {
    tb_cb_desc(*tb, g, r, cw_off, tb_slot_of(*tb, slots, r), cbs[r]);
    cbs[r].llr_offset = cw_off; // relative to the codeword
    cbs[r].out_offset = r * FEC_MAX_CB_OUT_BYTES;
    cw_off += cbs[r].E;
    int rc = fec_desc_prepare(&cbs[r]);
    if (rc != FEC_OK) {
      return rc;
    }
    if (payload_bits) {
      const uint32_t left = g.stream_bits - stream;
      payload_bits[r]     = left < g.data_bits ? left : g.data_bits;
      stream += payload_bits[r];
    }
  }
  if (cw_off != tb->nof_llrs) {
    return fail(FEC_ERR_BAD_ARG, "transport block: rate-matched lengths sum to %u, codeword has %u LLRs", cw_off, tb->nof_llrs);
  }
  *nof_cbs = g.nof_cbs;
  return FEC_OK;
}

int fec_tb_reset_slots(fec_engine* e, uint32_t first_slot, uint32_t n)
{
  if (!e || first_slot > e->cfg.harq_slots || n > e->cfg.harq_slots - first_slot) {
    return fail(FEC_ERR_BAD_ARG, "bad slot range");
  }
  if (e->cb_ok) {
    for (uint32_t i = 0; i < n; ++i) {
      e->cb_ok[first_slot + i].store(0);
    }
  }
  return FEC_OK;
}

namespace {
// Where the codeword LLRs of a transport-block call come from when they are not host LLRs: demodulated on the device.
struct demod_source {
  const fec_demod_desc* descs;
  uint32_t              n;
  const float*          h_symbols;
  const float*          h_noise;
  size_t                nof_symbols;
  uint32_t              flavour;
};
int demod_stage(fec_engine* e, cudaStream_t st, const demod_source& dm, size_t l0, size_t l1, int8_t* d_llr);

// First half of a transport-block call: validation, segmentation, every copy and kernel put on the context's streams,
// hx.done recorded behind all of it. Nothing here waits for the GPU.
int tb_submit(fec_engine* e, host_ctx& hx, const fec_tb_desc* tbs, uint32_t n, const uint32_t* slots, const int8_t* h_llr,
              size_t llr_bytes, const demod_source* dm, uint8_t* h_out, size_t out_bytes)
{
  if (!e || !tbs || (!h_llr && !dm) || !h_out) {
    return fail(FEC_ERR_BAD_ARG, "null argument");
  }
  if (n == 0 || n > e->cfg.max_batch_cbs) {
    return fail(FEC_ERR_BAD_ARG, "%u transport blocks in one call (the engine takes 1 .. %u)", n, e->cfg.max_batch_cbs);
  }
  int rc = tb_state_alloc(e, hx);
  if (rc != FEC_OK) {
    return rc;
  }
  if (hx.pending == nullptr) {
    hx.pending = new tb_pending();
  }
  tb_pending& P = *hx.pending;
  // FEC_B200_TRACE=1: host timestamps of the call's phases on stderr (diagnostic; off by default).
  static const bool trace = getenv("FEC_B200_TRACE") != nullptr;
  const auto        now   = [] { return std::chrono::steady_clock::now(); };
  const auto        t_in  = now();
  auto              t_seg = t_in;
  // Pass 0: geometry of every block, so that the sizes of everything that follows are known.
  hx.tb_geo.resize(n);
  uint32_t n_map = 0, max_Z = 0;
  int      any1  = 0;
  bool     aligned = true; // every payload is whole bytes: the one-CTA-per-codeblock gather applies
  bool     once    = e->packed_ok && e->cfg.kernel_variant != 1; // every input is read exactly once (zero copy is possible)
  for (uint32_t t = 0; t < n; ++t) {
    const fec_tb_desc& tb = tbs[t];
    fec_tb_geometry&   g  = hx.tb_geo[t];
    if (!fec_tb_geometry_of(tb, g)) {
      return fail(FEC_ERR_BAD_ARG, "transport block %u: tbs=%u bg=%u rv=%u qm=%u layers=%u llrs=%u", t, tb.tbs,
                  tb.base_graph, tb.rv, tb.qm, tb.nof_layers, tb.nof_llrs);
    }
    if (tb.llr_offset + tb.nof_llrs > llr_bytes || tb.out_offset + tb.tbs / 8 > out_bytes ||
        tb.out_offset + tb.tbs / 8 > (size_t)e->cfg.max_batch_cbs * FEC_MAX_CB_OUT_BYTES) {
      return fail(FEC_ERR_BAD_ARG, "transport block %u indexes outside the batch buffers", t);
    }
    n_map += g.nof_cbs;
    aligned = aligned && (g.data_bits % 8 == 0);
    max_Z   = g.lifting_size > max_Z ? g.lifting_size : max_Z;
    any1 |= launch_flags_of(tb.base_graph, g.lifting_size);
  }
  if (n_map > e->cfg.max_batch_cbs) {
    return fail(FEC_ERR_BAD_ARG, "%u codeblocks in one call, the engine takes %u", n_map, e->cfg.max_batch_cbs);
  }
  // Batches of two kernel waves or more (many cells per slot) are pipelined like fec_run_host: groups of whole transport
  // blocks of about one wave, LLR copies in order on the copy stream, codeblock + gather kernels and the copy back of the
  // block bytes on three streams. Needs the byte-aligned gather and LLR spans that follow one another in block order.
  // Everything the LLR copies need is known after pass 0, so the first of them are put on the wire BEFORE the host segments
  // the blocks (0.17 ms for 46 config-3 blocks).
  const uint32_t wave      = FEC_PACKED_CTAS_PER_SM * (uint32_t)e->sm_count;
  bool           pipelined = !dm && aligned && n > 1 && n_map >= 2 * wave;
  for (uint32_t t = 1; t < n && pipelined; ++t) {
    pipelined = tbs[t].llr_offset >= tbs[t - 1].llr_offset + tbs[t - 1].nof_llrs;
  }
  // Whatever way this call ends, the caller's LLR buffer must not be in use by a copy any more.
  struct copy_guard {
    cudaStream_t s;
    bool         armed;
    ~copy_guard()
    {
      if (armed) {
        cudaStreamSynchronize(s);
      }
    }
  } guard{hx.copy_in, false};
  hx.tb_chunk_end.clear();
  if (pipelined) {
    const size_t pl0 = (size_t)tbs[0].llr_offset, pl1 = (size_t)tbs[n - 1].llr_offset + tbs[n - 1].nof_llrs;
    if (pl1 - pl0 > e->cfg.max_batch_llrs) {
      return fail(FEC_ERR_BAD_ARG, "%zu LLR bytes exceed the engine staging capacity", pl1 - pl0);
    }
    CU_TRY(cudaSetDevice(e->cfg.device));
    guard.armed = true;
    uint32_t c  = 0;
    for (uint32_t t0 = 0, t1 = 0; t0 < n; t0 = t1, ++c) {
      uint32_t cn = 0;
      for (t1 = t0; t1 < n && (cn < wave || t1 == t0); ++t1) {
        cn += hx.tb_geo[t1].nof_cbs;
      }
      if (hx.chunk_in.size() <= c) {
        cudaEvent_t ev = nullptr;
        CU_TRY(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        hx.chunk_in.push_back(ev);
      }
      // Only the first three groups go now (about the transfer time of the segmentation that follows): the descriptor
      // block has to reach the device before the first kernel, and one copy engine serves the queue in issue order.
      if (c < 3) {
        const size_t a = (size_t)tbs[t0].llr_offset, b = (size_t)tbs[t1 - 1].llr_offset + tbs[t1 - 1].nof_llrs;
        CU_TRY(cudaMemcpyAsync(hx.d_llr + (a - pl0), h_llr + a, b - a, cudaMemcpyHostToDevice, hx.copy_in));
        CU_TRY(cudaEventRecord(hx.chunk_in[c], hx.copy_in));
      }
      hx.tb_chunk_end.push_back(t1);
    }
  }
  // One pinned block travels to the device in one copy: jobs | codeblock map | scratch (zeros) | codeblock descriptors.
  auto           up16      = [](size_t x) { return (x + 15) & ~(size_t)15; };
  const size_t   off_map   = up16((size_t)n * sizeof(fec_tb_job));
  const size_t   off_scr   = up16(off_map + (size_t)n_map * sizeof(fec_tb_cb_ref));
  const size_t   off_descs = up16(off_scr + (size_t)n * 2 * sizeof(uint32_t));
  const size_t   meta_size = off_descs + (size_t)n_map * sizeof(fec_cb_desc);
  fec_tb_job*    h_jobs    = reinterpret_cast<fec_tb_job*>(hx.h_meta);
  fec_tb_cb_ref* h_map     = reinterpret_cast<fec_tb_cb_ref*>(hx.h_meta + off_map);
  fec_cb_desc*   h_descs   = reinterpret_cast<fec_cb_desc*>(hx.h_meta + off_descs);
  memset(hx.h_meta + off_scr, 0, (size_t)n * 2 * sizeof(uint32_t));
  // Pass 1: segmentation -- one descriptor per codeblock, the codeblock map, the jobs.
  size_t   l0 = ~(size_t)0, l1 = 0;
  uint32_t m = 0;
  for (uint32_t t = 0; t < n; ++t) {
    const fec_tb_desc&     tb = tbs[t];
    const fec_tb_geometry& g  = hx.tb_geo[t];
    fec_tb_job&            J  = h_jobs[t];
    J.out_offset              = tb.out_offset;
    J.nof_cbs                 = g.nof_cbs;
    J.data_bits               = g.data_bits;
    J.stream_bits             = g.stream_bits;
    J.tbs_bytes               = tb.tbs / 8;
    J.map_first               = m;
    J.max_iter                = tb.max_iter;
    if (g.data_bits % 8 == 0 && g.nof_cbs > 1 &&
        (hx.wc_key[0] != g.nof_cbs || hx.wc_key[1] != g.data_bits || hx.wc_key[2] != g.stream_bits)) {
      // weight of codeblock r = x^(8 * stream bytes after its payload) mod g24A, built backwards with one multiply each
      hx.wc_key[0] = g.nof_cbs;
      hx.wc_key[1] = g.data_bits;
      hx.wc_key[2] = g.stream_bits;
      hx.wc_weights.assign(g.nof_cbs, 1u);
      const uint32_t dbytes = g.data_bits / 8, last = g.stream_bits / 8 - (g.nof_cbs - 1) * dbytes;
      uint32_t       w      = fec_tb_crc24a_xpow8_host(last); // after codeblock C-2 comes the last payload
      const uint32_t step   = fec_tb_crc24a_xpow8_host(dbytes);
      for (int r = (int)g.nof_cbs - 2; r >= 0; --r) {
        hx.wc_weights[r] = w;
        w                = fec_tb_crc24a_mulmod_host(w, step);
      }
    }
    uint32_t cw_off = 0;
    for (uint32_t r = 0; r < g.nof_cbs; ++r, ++m) {
      const uint32_t slot = tb_slot_of(tb, slots, r);
      if (slot >= e->cfg.harq_slots) {
        return fail(FEC_ERR_BAD_ARG, "transport block %u codeblock %u: harq slot %u >= %u", t, r, slot, e->cfg.harq_slots);
      }
      fec_tb_cb_ref& ref = h_map[m];
      ref.slot           = slot;
      ref.tb             = t;
      ref.weight         = (g.data_bits % 8 == 0 && g.nof_cbs > 1) ? hx.wc_weights[r] : 1u;
      fec_cb_desc& d     = h_descs[m];
      tb_cb_desc(tb, g, r, cw_off, slot, d);
      // A codeblock whose CRC passed in an earlier transmission is not decoded again, but the new LLRs are still
      // combined into its soft buffer: the transport-block CRC may yet fail and put it back in play
      // (pusch_decoder_impl.cpp:333-344).
      // (new data resets the flags, pusch_decoder_impl.cpp:124-128; the reset itself is committed after validation)
      const bool already_ok = !(tb.flags & FEC_FLAG_NEW_DATA) && e->cb_ok[slot].load() != 0;
      if (already_ok) {
        d.mode = FEC_MODE_DEMATCH;
      }
      rc = fec_desc_prepare(&d);
      if (rc != FEC_OK) {
        return rc;
      }
      ref.result_index = already_ok ? -1 : (int32_t)m;
      once             = once && reads_input_once(d);
      l0               = d.llr_offset < l0 ? (size_t)d.llr_offset : l0;
      l1               = d.llr_offset + d.E > l1 ? (size_t)(d.llr_offset + d.E) : l1;
      cw_off += d.E;
    }
    if (cw_off != tb.nof_llrs) {
      return fail(FEC_ERR_BAD_ARG, "transport block %u: rate-matched lengths sum to %u, codeword has %u LLRs", t, cw_off,
                  tb.nof_llrs);
    }
  }
  if (l1 - l0 > e->cfg.max_batch_llrs) {
    return fail(FEC_ERR_BAD_ARG, "%zu LLR bytes exceed the engine staging capacity", l1 - l0);
  }
  // Every check has passed: only now does the call change engine state (codeblock CRC flags of new transmissions, the
  // "reads as zeros" marks of freed slots). A CUDA failure further down puts the marks back (run_guard).
  run_guard rg(e, &hx);
  for (uint32_t t = 0; t < n; ++t) {
    if (tbs[t].flags & FEC_FLAG_NEW_DATA) {
      for (uint32_t r = 0; r < h_jobs[t].nof_cbs; ++r) {
        e->cb_ok[h_map[h_jobs[t].map_first + r].slot].store(0);
      }
    }
  }
  for (uint32_t i = 0; i < n_map; ++i) {
    apply_fresh(e, h_descs[i], &rg.taken);
    h_descs[i].llr_offset -= l0; // the staged LLRs start at l0
  }
  t_seg = now();
  CU_TRY(cudaSetDevice(e->cfg.device));
  cudaStream_t         st      = hx.stream;
  const fec_tb_job*    d_jobs  = reinterpret_cast<const fec_tb_job*>(hx.d_meta);
  const fec_tb_cb_ref* d_map   = reinterpret_cast<const fec_tb_cb_ref*>(hx.d_meta + off_map);
  uint32_t*            d_scr   = reinterpret_cast<uint32_t*>(hx.d_meta + off_scr);
  const fec_cb_desc*   d_descs = reinterpret_cast<const fec_cb_desc*>(hx.d_meta + off_descs);
  // results come back in one copy too: codeblock results | transport block results
  const size_t   off_tbres = up16((size_t)n_map * sizeof(fec_cb_result));
  fec_cb_result* d_cbres   = reinterpret_cast<fec_cb_result*>(hx.d_rblk);
  fec_tb_result* d_tbres   = reinterpret_cast<fec_tb_result*>(hx.d_rblk + off_tbres);
  CU_TRY(cudaMemcpyAsync(hx.d_meta, hx.h_meta, meta_size, cudaMemcpyHostToDevice, st));
  if (pipelined) {
    CU_TRY(cudaEventRecord(hx.pipe_done[0], st));
    for (int k = 0; k < 3; ++k) {
      CU_TRY(cudaStreamWaitEvent(hx.pipe[k], hx.pipe_done[0], 0));
    }
    for (uint32_t c = 0, t0 = 0; c < hx.tb_chunk_end.size(); t0 = hx.tb_chunk_end[c], ++c) {
      const uint32_t t1 = hx.tb_chunk_end[c];
      const uint32_t m0 = h_jobs[t0].map_first, cn = h_jobs[t1 - 1].map_first + h_jobs[t1 - 1].nof_cbs - m0;
      cudaStream_t   ps = hx.pipe[c % 3];
      if (c >= 3) { // the first three groups' LLRs went on the wire before the segmentation
        const size_t a = (size_t)tbs[t0].llr_offset, b = (size_t)tbs[t1 - 1].llr_offset + tbs[t1 - 1].nof_llrs;
        CU_TRY(cudaMemcpyAsync(hx.d_llr + (a - l0), h_llr + a, b - a, cudaMemcpyHostToDevice, hx.copy_in));
        CU_TRY(cudaEventRecord(hx.chunk_in[c], hx.copy_in));
      }
      CU_TRY(cudaStreamWaitEvent(ps, hx.chunk_in[c], 0));
      rc = launch(e, ps, d_descs + m0, cn, hx.d_llr, e->d_msg, d_cbres + m0, max_Z, any1);
      if (rc != FEC_OK) {
        return rc;
      }
      fec_tb_gather_kernel<<<cn, 256, 0, ps>>>(d_jobs, d_map, d_cbres, e->d_msg, hx.d_out, d_tbres, d_scr, e->d_tab, m0);
      CU_TRY(cudaGetLastError());
      e->launches.fetch_add(1);
      for (uint32_t t = t0; t < t1; ++t) {
        CU_TRY(cudaMemcpyAsync(hx.h_tbout + tbs[t].out_offset, hx.d_out + tbs[t].out_offset, tbs[t].tbs / 8,
                               cudaMemcpyDeviceToHost, ps));
      }
      // The chunk's own slice of the results follows, so that its blocks can be handed over while later chunks still run.
      CU_TRY(cudaMemcpyAsync(hx.h_rblk + (size_t)m0 * sizeof(fec_cb_result), hx.d_rblk + (size_t)m0 * sizeof(fec_cb_result),
                             (size_t)cn * sizeof(fec_cb_result), cudaMemcpyDeviceToHost, ps));
      CU_TRY(cudaMemcpyAsync(hx.h_rblk + off_tbres + (size_t)t0 * sizeof(fec_tb_result),
                             hx.d_rblk + off_tbres + (size_t)t0 * sizeof(fec_tb_result), (size_t)(t1 - t0) * sizeof(fec_tb_result),
                             cudaMemcpyDeviceToHost, ps));
      if (hx.chunk_out.size() <= c) {
        cudaEvent_t ev = nullptr;
        CU_TRY(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        hx.chunk_out.push_back(ev);
      }
      CU_TRY(cudaEventRecord(hx.chunk_out[c], ps));
    }
  } else {
    const int8_t* zc_llr = nullptr;
    if (dm) {
      rc = demod_stage(e, st, *dm, l0, l1, hx.d_llr);
      if (rc != FEC_OK) {
        return rc;
      }
    } else {
      // pinned input: the kernel reads it in place (the transfer overlaps the decoding inside the one launch)
      // (only for a call that is alone on the engine, see fec_run_host)
      zc_llr = (once && e->ctx_busy.load(std::memory_order_relaxed) <= 1)
                   ? static_cast<const int8_t*>(pinned_device_ptr(h_llr + l0))
                   : nullptr;
      if (!zc_llr) {
        CU_TRY(cudaMemcpyAsync(hx.d_llr, h_llr + l0, l1 - l0, cudaMemcpyHostToDevice, st));
      }
    }
    rc = launch(e, st, d_descs, n_map, zc_llr ? zc_llr : hx.d_llr, e->d_msg, d_cbres, max_Z, any1);
    if (rc != FEC_OK) {
      return rc;
    }
    if (aligned) {
      fec_tb_gather_kernel<<<n_map, 256, 0, st>>>(d_jobs, d_map, d_cbres, e->d_msg, hx.d_out, d_tbres, d_scr, e->d_tab, 0u);
    } else {
      fec_tb_assemble_kernel<<<n, 1024, 0, st>>>(d_jobs, d_map, d_cbres, e->d_msg, hx.d_out, d_tbres);
    }
    CU_TRY(cudaGetLastError());
    e->launches.fetch_add(1);
    CU_TRY(cudaMemcpyAsync(hx.h_rblk, hx.d_rblk, off_tbres + (size_t)n * sizeof(fec_tb_result), cudaMemcpyDeviceToHost, st));
    // The transport block bytes come back unconditionally into a pinned bounce buffer (a block that is not delivered must
    // leave the caller's buffer untouched, and waiting for the results first would cost a second round trip).
    for (uint32_t t = 0; t < n; ++t) {
      CU_TRY(cudaMemcpyAsync(hx.h_tbout + tbs[t].out_offset, hx.d_out + tbs[t].out_offset, tbs[t].tbs / 8,
                             cudaMemcpyDeviceToHost, st));
    }
  }
  // One event behind everything: the pipelined groups ran on three streams.
  if (pipelined) {
    for (size_t c = 0; c < hx.tb_chunk_end.size(); ++c) {
      CU_TRY(cudaStreamWaitEvent(st, hx.chunk_out[c], 0));
    }
  }
  CU_TRY(cudaEventRecord(hx.done, st));
  rg.ok = true; // from here on tb_complete owns the marks
  guard.armed = false; // and the wait for the copies out of the caller's LLR buffer
  P.tbs.assign(tbs, tbs + n);
  P.n         = n;
  P.n_map     = n_map;
  P.off_map   = off_map;
  P.off_tbres = off_tbres;
  P.pipelined = pipelined;
  P.h_out     = h_out;
  P.taken.swap(rg.taken);
  P.t_in  = t_in;
  P.t_seg = t_seg;
  P.t_enq = now();
  e->st_tb_calls.fetch_add(1, std::memory_order_relaxed);
  e->st_tbs.fetch_add(n, std::memory_order_relaxed);
  (void)trace;
  return FEC_OK;
}

// Second half: waits for the GPU (chunk by chunk for a pipelined call, so that the host copies of the early groups overlap the
// kernels of the late ones), hands over the transport blocks whose codeblocks are all good and updates the per-codeblock
// state. A CUDA failure comes back as failed blocks (nothing delivered, every codeblock failed) AND as FEC_ERR_CUDA.
int tb_complete(fec_engine* e, host_ctx& hx, fec_tb_result* h_results, fec_cb_result* h_cb_results)
{
  static const bool trace = getenv("FEC_B200_TRACE") != nullptr;
  const auto        now   = [] { return std::chrono::steady_clock::now(); };
  tb_pending&       P     = *hx.pending;
  const uint32_t    n     = P.n;
  const fec_tb_desc* tbs  = P.tbs.data();
  uint8_t*          h_out = P.h_out;
  const fec_tb_job*    h_jobs  = reinterpret_cast<const fec_tb_job*>(hx.h_meta);
  const fec_tb_cb_ref* h_map   = reinterpret_cast<const fec_tb_cb_ref*>(hx.h_meta + P.off_map);
  const fec_cb_result* h_cbres = reinterpret_cast<const fec_cb_result*>(hx.h_rblk);
  const fec_tb_result* h_tbres = reinterpret_cast<const fec_tb_result*>(hx.h_rblk + P.off_tbres);
  auto                 t_sync  = P.t_enq;
  stats_batch          sb;
  // (the per-slot flags below are published by the mutex of the context release that follows this call: relaxed stores)
  const auto           deliver = [&](uint32_t t) {
    const fec_tb_result& R = h_tbres[t];
    const fec_tb_job&    J = h_jobs[t];
    h_results[t]           = R;
    if (R.tb_delivered) {
      memcpy(h_out + tbs[t].out_offset, hx.h_tbout + tbs[t].out_offset, tbs[t].tbs / 8);
    }
    for (uint32_t r = 0; r < J.nof_cbs; ++r) {
      const fec_tb_cb_ref& ref = h_map[J.map_first + r];
      fec_cb_result        cr;
      if (ref.result_index >= 0) {
        cr = h_cbres[ref.result_index];
        if (cr.status & FEC_RES_CRC_PASS) {
          e->cb_ok[ref.slot].store(1, std::memory_order_relaxed);
        }
        sb.add(cr);
      } else {
        cr.nof_ldpc_iterations = 0;
        cr.status              = FEC_RES_SKIPPED | FEC_RES_CRC_PASS | FEC_RES_DONE;
        cr.nof_layers          = 0;
        cr.reserved            = 0;
      }
      if (h_cb_results) {
        h_cb_results[(size_t)t * FEC_MAX_TB_CBS + r] = cr;
      }
    }
    if (R.tb_crc_ok) {
      e->st_tb_crc_ok.fetch_add(1, std::memory_order_relaxed);
      // join_and_notify: the buffer is released -> free the HARQ context entries.
      for (uint32_t r = 0; r < J.nof_cbs; ++r) {
        const uint32_t slot = h_map[J.map_first + r].slot;
        e->fresh[slot].store(1, std::memory_order_relaxed);
        e->cb_ok[slot].store(0, std::memory_order_relaxed);
      }
    } else if (R.tb_delivered) {
      // All codeblock CRCs good but the TB CRC is not: at least one false positive, start over (reset_codeblocks_crc).
      for (uint32_t r = 0; r < J.nof_cbs; ++r) {
        e->cb_ok[h_map[J.map_first + r].slot].store(0);
      }
    }
  };
  cudaSetDevice(e->cfg.device);
  cudaError_t err = cudaSuccess;
  uint32_t    t   = 0;
  if (P.pipelined) {
    for (size_t c = 0; c < hx.tb_chunk_end.size() && err == cudaSuccess; ++c) {
      err = cudaEventSynchronize(hx.chunk_out[c]);
      for (; err == cudaSuccess && t < hx.tb_chunk_end[c]; ++t) {
        deliver(t);
      }
    }
  }
  if (err == cudaSuccess) {
    err    = cudaEventSynchronize(hx.done);
    t_sync = now();
  }
  for (; err == cudaSuccess && t < n; ++t) {
    deliver(t);
  }
  sb.flush(e);
  if (err != cudaSuccess) {
    // The device faulted: whatever was not handed over yet failed; the freed-slot marks this call consumed are put back.
    cudaGetLastError();
    cudaStreamSynchronize(hx.copy_in);
    e->st_faults.fetch_add(1, std::memory_order_relaxed);
    for (; t < n; ++t) {
      fec_tb_result R = {};
      R.nof_codeblocks = (uint16_t)h_jobs[t].nof_cbs;
      h_results[t]     = R;
      for (uint32_t r = 0; h_cb_results && r < h_jobs[t].nof_cbs; ++r) {
        fec_cb_result cr = {};
        cr.nof_ldpc_iterations = (uint16_t)tbs[t].max_iter;
        cr.status              = FEC_RES_DONE;
        h_cb_results[(size_t)t * FEC_MAX_TB_CBS + r] = cr;
      }
    }
    for (uint32_t sl : P.taken) {
      e->fresh[sl].store(1);
    }
    P.taken.clear();
    return fail(FEC_ERR_CUDA, "transport-block call failed on the device: %s", cudaGetErrorString(err));
  }
  P.taken.clear();
  if (trace) {
    const auto us = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
      return std::chrono::duration<double, std::micro>(b - a).count();
    };
    fprintf(stderr, "fec_b200 trace: tb_run n=%u cbs=%u segment %.1f us, enqueue %.1f us, wait %.1f us, deliver %.1f us\n", n, P.n_map,
            us(P.t_in, P.t_seg), us(P.t_seg, P.t_enq), us(P.t_enq, t_sync), us(t_sync, now()));
  }
  return FEC_OK;
}

// Synchronous call: a free context, submit, complete. dm (device-side demodulation) uses engine-wide staging: context 0.
int tb_run(fec_engine* e, const fec_tb_desc* tbs, uint32_t n, const uint32_t* slots, const int8_t* h_llr, size_t llr_bytes,
           const demod_source* dm, uint8_t* h_out, size_t out_bytes, fec_tb_result* h_results, fec_cb_result* h_cb_results)
{
  if (!e || !h_results) {
    return fail(FEC_ERR_BAD_ARG, "null argument");
  }
  if (n == 0) {
    return FEC_OK;
  }
  ctx_lease lease(e, dm ? 0 : -1);
  int       rc = tb_submit(e, *lease.c, tbs, n, slots, h_llr, llr_bytes, dm, h_out, out_bytes);
  if (rc != FEC_OK) {
    return rc;
  }
  return tb_complete(e, *lease.c, h_results, h_cb_results);
}
} // namespace

int fec_tb_run_host(fec_engine* e, const fec_tb_desc* tbs, uint32_t n, const uint32_t* slots, const int8_t* h_llr,
                    size_t llr_bytes, uint8_t* h_out, size_t out_bytes, fec_tb_result* h_results, fec_cb_result* h_cb_results)
{
  if (!h_llr) {
    return fail(FEC_ERR_BAD_ARG, "null argument");
  }
  return tb_run(e, tbs, n, slots, h_llr, llr_bytes, nullptr, h_out, out_bytes, h_results, h_cb_results);
}

// ---- asynchronous transport-block interface ------------------------------------------------------------------------------------
namespace {
// Completion thread: takes the in-flight calls that carry a callback in submission order, waits for each on its event,
// hands the blocks over and calls back. Started by the first such call.
void tb_completion_loop(fec_engine* e)
{
  cudaSetDevice(e->cfg.device);
  for (;;) {
    host_ctx* c = nullptr;
    {
      std::unique_lock<std::mutex> lk(e->tb_mtx);
      e->tb_cv.wait(lk, [e] { return e->tb_stop || !e->tb_fifo.empty(); });
      if (e->tb_fifo.empty()) {
        return; // stop requested and nothing in flight
      }
      c = e->tb_fifo.front();
      e->tb_fifo.erase(e->tb_fifo.begin());
    }
    tb_pending& P = *c->pending;
    P.res.assign(P.n, fec_tb_result());
    const int             rc     = tb_complete(e, *c, P.res.data(), nullptr);
    const fec_tb_callback cb     = P.cb;
    void*                 user   = P.user;
    const uint64_t        ticket = P.ticket;
    std::vector<fec_tb_result> res;
    res.swap(P.res);
    P.ticket = 0;
    ctx_release(e, c); // before the callback: it may enqueue the next block
    cb(user, ticket, rc, res.data(), (uint32_t)res.size());
  }
}
} // namespace

int fec_tb_enqueue(fec_engine* e, const fec_tb_desc* tbs, uint32_t n, const uint32_t* slots, const int8_t* h_llr, size_t llr_bytes,
                   uint8_t* h_out, size_t out_bytes, fec_tb_callback callback, void* user, uint64_t* ticket)
{
  if (!e || !tbs || !h_llr || !h_out || !ticket || n == 0) {
    return fail(FEC_ERR_BAD_ARG, "null argument");
  }
  host_ctx* c = ctx_acquire(e, -1, false);
  if (c == nullptr) {
    e->st_queue_full.fetch_add(1, std::memory_order_relaxed);
    return fail(FEC_ERR_QUEUE_FULL, "all %zu host-call contexts are in flight", e->ctxs.size());
  }
  const int rc = tb_submit(e, *c, tbs, n, slots, h_llr, llr_bytes, nullptr, h_out, out_bytes);
  if (rc != FEC_OK) {
    ctx_release(e, c);
    return rc;
  }
  tb_pending& P = *c->pending;
  P.ticket      = ((uint64_t)e->tb_ticket.fetch_add(1) << 8) | c->index;
  P.cb          = callback;
  P.user        = user;
  *ticket       = P.ticket;
  if (callback != nullptr) {
    std::lock_guard<std::mutex> lk(e->tb_mtx);
    if (!e->tb_thread.joinable()) {
      e->tb_thread = std::thread(tb_completion_loop, e);
    }
    e->tb_fifo.push_back(c);
    e->tb_cv.notify_one();
  }
  return FEC_OK;
}

namespace {
int tb_ticket_ctx(fec_engine* e, uint64_t ticket, host_ctx** out)
{
  const uint32_t idx = (uint32_t)(ticket & 0xff);
  if (!e || ticket == 0 || idx >= e->ctxs.size() || e->ctxs[idx]->pending == nullptr || e->ctxs[idx]->pending->ticket != ticket ||
      e->ctxs[idx]->pending->cb != nullptr) {
    return fail(FEC_ERR_BAD_ARG, "unknown transport-block ticket (already collected, or a call with a callback)");
  }
  *out = e->ctxs[idx];
  return FEC_OK;
}
} // namespace

int fec_tb_poll(fec_engine* e, uint64_t ticket, fec_tb_result* h_results, fec_cb_result* h_cb_results)
{
  host_ctx* c  = nullptr;
  int       rc = tb_ticket_ctx(e, ticket, &c);
  if (rc != FEC_OK) {
    return rc;
  }
  if (!h_results) {
    return fail(FEC_ERR_BAD_ARG, "null argument");
  }
  cudaSetDevice(e->cfg.device);
  const cudaError_t q = cudaEventQuery(c->done);
  if (q == cudaErrorNotReady) {
    return 0;
  }
  rc                 = tb_complete(e, *c, h_results, h_cb_results); // an error state of the event is reported from in there
  c->pending->ticket = 0;
  ctx_release(e, c);
  return rc == FEC_OK ? 1 : rc;
}

int fec_tb_wait(fec_engine* e, uint64_t ticket, fec_tb_result* h_results, fec_cb_result* h_cb_results)
{
  host_ctx* c  = nullptr;
  int       rc = tb_ticket_ctx(e, ticket, &c);
  if (rc != FEC_OK) {
    return rc;
  }
  if (!h_results) {
    return fail(FEC_ERR_BAD_ARG, "null argument");
  }
  rc                 = tb_complete(e, *c, h_results, h_cb_results);
  c->pending->ticket = 0;
  ctx_release(e, c);
  return rc;
}

// ---- counters ---------------------------------------------------------------------------------------------------------------
int fec_engine_get_stats(fec_engine* e, fec_engine_stats* out)
{
  if (!e || !out) {
    return fail(FEC_ERR_BAD_ARG, "null argument");
  }
  memset(out, 0, sizeof(*out));
  out->launches            = e->launches.load();
  out->cb_launches         = e->st_batches.load();
  out->codeblocks          = e->st_codeblocks.load();
  out->results             = e->st_results.load();
  out->crc_pass            = e->st_crc_pass.load();
  out->queue_full          = e->st_queue_full.load();
  out->faults              = e->st_faults.load();
  out->generic_launches    = e->st_generic.load();
  out->tb_calls            = e->st_tb_calls.load();
  out->transport_blocks    = e->st_tbs.load();
  out->tb_crc_ok           = e->st_tb_crc_ok.load();
  out->max_queue_depth     = e->st_max_depth.load();
  out->dispatches          = e->st_dispatches.load();
  out->dispatch_batches    = e->st_dispatch_slots.load();
  for (int i = 0; i < FEC_STATS_ITER_BINS; ++i) {
    out->iterations_hist[i] = e->st_iter_hist[i].load();
  }
  return FEC_OK;
}

int fec_engine_debug_fail_launches(fec_engine* e, uint32_t n)
{
  if (!e) {
    return fail(FEC_ERR_BAD_ARG, "null engine");
  }
  e->debug_fail_launches.store(n);
  return FEC_OK;
}

// ---- downlink twin: PDSCH codeblock encoding ---------------------------------------------------------------------------------
int fec_enc_desc_prepare(fec_enc_desc* d)
{
  if (d == nullptr) {
    return fail(FEC_ERR_BAD_ARG, "null descriptor");
  }
  if (d->base_graph != 1 && d->base_graph != 2) {
    return fail(FEC_ERR_BAD_ARG, "base_graph=%u", d->base_graph);
  }
  const int zi = z_index(d->lifting_size);
  if (zi < 0) {
    return fail(FEC_ERR_BAD_ARG, "lifting_size=%u is not one of the 51 NR lifting sizes", d->lifting_size);
  }
  const uint32_t Z = d->lifting_size, Kb = (d->base_graph == 1) ? 22 : 10, nb = (d->base_graph == 1) ? 66 : 50;
  const uint32_t N = nb * Z;
  d->graph_index   = (uint16_t)((d->base_graph - 1) * 51 + zi);
  if (d->nof_filler_bits >= (Kb - 2) * Z) {
    return fail(FEC_ERR_BAD_ARG, "nof_filler_bits=%u out of range", d->nof_filler_bits);
  }
  if (d->rv > 3 || (d->qm != 1 && d->qm != 2 && d->qm != 4 && d->qm != 6 && d->qm != 8) || d->E == 0 ||
      (d->E % d->qm) != 0) {
    return fail(FEC_ERR_BAD_ARG, "rv=%u qm=%u E=%u", d->rv, d->qm, d->E);
  }
  const uint32_t Ncb = (d->Nref > 0) ? (d->Nref < N ? d->Nref : N) : N;
  if (Ncb <= (Kb - 2) * Z) {
    return fail(FEC_ERR_BAD_ARG, "Nref=%u leaves no parity in the circular buffer", d->Nref);
  }
  // k0 as the reference computes it (ldpc_rate_matcher_impl.cpp:61-63): floor(f * Ncb / N) * Z in double precision.
  static const double F1[4] = {0, 17, 33, 56};
  static const double F2[4] = {0, 13, 25, 43};
  const double        f     = (d->base_graph == 1) ? F1[d->rv] : F2[d->rv];
  d->k0                     = (uint32_t)(uint16_t)std::floor(f * (double)Ncb / (double)N) * Z;
  return FEC_OK;
}

int fec_encode_launch_device(fec_engine* e, const fec_enc_desc* d_descs, uint32_t n, const uint8_t* d_in, uint8_t* d_out,
                             uint32_t max_Z)
{
  if (!e || !d_descs || !d_in || !d_out) {
    return fail(FEC_ERR_BAD_ARG, "null argument");
  }
  if (n == 0) {
    return FEC_OK;
  }
  if (z_index(max_Z) < 0) {
    return fail(FEC_ERR_BAD_ARG, "max_Z=%u is not a lifting size", max_Z);
  }
  CU_TRY(cudaSetDevice(e->cfg.device));
  fec_encode_kernel<<<n, FEC_ENC_THREADS, 0, e->stream>>>(d_descs, d_in, d_out, e->d_tab);
  CU_TRY(cudaGetLastError());
  e->launches.fetch_add(1);
  return FEC_OK;
}

namespace {
int grow(void** p, size_t* cap, size_t need)
{
  if (need <= *cap) {
    return FEC_OK;
  }
  if (*p) {
    cudaFree(*p);
    *p   = nullptr;
    *cap = 0;
  }
  if (cudaMalloc(p, need) != cudaSuccess) {
    cudaGetLastError();
    return fail(FEC_ERR_NO_MEMORY, "cannot allocate %zu bytes of device memory", need);
  }
  *cap = need;
  return FEC_OK;
}
} // namespace

int fec_encode_run_host(fec_engine* e, fec_enc_desc* h_descs, uint32_t n, const uint8_t* h_in, size_t in_bytes,
                        uint8_t* h_out, size_t out_bytes)
{
  if (!e || !h_descs || !h_in || !h_out) {
    return fail(FEC_ERR_BAD_ARG, "null argument");
  }
  if (n == 0) {
    return FEC_OK;
  }
  uint32_t max_Z = 0;
  size_t   lo = out_bytes, hi = 0; // span of the output buffer this call writes
  for (uint32_t i = 0; i < n; ++i) {
    fec_enc_desc& d = h_descs[i];
    int           r = fec_enc_desc_prepare(&d);
    if (r != FEC_OK) {
      return r;
    }
    const uint32_t Kb      = (d.base_graph == 1) ? 22 : 10;
    const size_t   in_len  = ((size_t)Kb * d.lifting_size - d.nof_filler_bits + 7) / 8;
    const size_t   out_len = (d.flags & FEC_ENC_FLAG_UNPACKED) ? (size_t)d.E : ((size_t)d.E + 7) / 8;
    if ((size_t)d.in_offset + in_len > in_bytes || d.out_offset + out_len > out_bytes) {
      return fail(FEC_ERR_BAD_ARG, "codeblock %u: buffers too small (in %u+%zu of %zu, out %llu+%zu of %zu)", i, d.in_offset,
                  in_len, in_bytes, (unsigned long long)d.out_offset, out_len, out_bytes);
    }
    max_Z = d.lifting_size > max_Z ? d.lifting_size : max_Z;
    lo    = d.out_offset < lo ? (size_t)d.out_offset : lo;
    hi    = d.out_offset + out_len > hi ? (size_t)d.out_offset + out_len : hi;
  }
  ctx_lease lease(e, 0); // the encoder staging is engine-wide: one call at a time, on context 0's pipeline streams
  host_ctx& hx = *lease.c;
  CU_TRY(cudaSetDevice(e->cfg.device));
  int r = grow((void**)&e->d_enc_descs, &e->enc_descs_cap, (size_t)n * sizeof(fec_enc_desc));
  r     = (r == FEC_OK) ? grow((void**)&e->d_enc_in, &e->enc_in_cap, in_bytes) : r;
  r     = (r == FEC_OK) ? grow((void**)&e->d_enc_out, &e->enc_out_cap, out_bytes) : r;
  if (r != FEC_OK) {
    return r;
  }
  CU_TRY(cudaMemcpyAsync(e->d_enc_descs, h_descs, (size_t)n * sizeof(fec_enc_desc), cudaMemcpyHostToDevice, e->stream));
  // Large batches are pipelined in chunks of one kernel wave (8 CTAs per SM): segment copies in order on the copy stream,
  // kernels and the (three times larger) copies of the rate-matched output on three streams. The chunks' input and output
  // spans must follow one another; each output span runs up to the next chunk's so that together they tile [lo, hi).
  struct span_t {
    size_t i0, i1, o0, o1;
  };
  const uint32_t      wave  = 8u * (uint32_t)e->sm_count;
  uint32_t            chunk = (n >= 2 * wave) ? wave : n;
  std::vector<span_t> spans;
  for (uint32_t c0 = 0; c0 < n && chunk != n; c0 += chunk) {
    span_t         sp = {~(size_t)0, 0, ~(size_t)0, 0};
    const uint32_t c1 = (n - c0 < chunk) ? n : c0 + chunk;
    for (uint32_t i = c0; i < c1; ++i) {
      const fec_enc_desc& d       = h_descs[i];
      const size_t        in_len  = ((size_t)((d.base_graph == 1) ? 22 : 10) * d.lifting_size - d.nof_filler_bits + 7) / 8;
      const size_t        out_len = (d.flags & FEC_ENC_FLAG_UNPACKED) ? (size_t)d.E : ((size_t)d.E + 7) / 8;
      sp.i0 = d.in_offset < sp.i0 ? (size_t)d.in_offset : sp.i0;
      sp.i1 = d.in_offset + in_len > sp.i1 ? (size_t)d.in_offset + in_len : sp.i1;
      sp.o0 = d.out_offset < sp.o0 ? (size_t)d.out_offset : sp.o0;
      sp.o1 = d.out_offset + out_len > sp.o1 ? (size_t)d.out_offset + out_len : sp.o1;
    }
    if (!spans.empty() && (sp.i0 < spans.back().i1 || sp.o0 < spans.back().o1)) {
      chunk = n; // interleaved buffers: one launch
      break;
    }
    if (!spans.empty()) {
      spans.back().o1 = sp.o0;
    }
    spans.push_back(sp);
  }
  if (chunk == n) {
    CU_TRY(cudaMemcpyAsync(e->d_enc_in, h_in, in_bytes, cudaMemcpyHostToDevice, e->stream));
    CU_TRY(cudaMemsetAsync(e->d_enc_out + lo, 0, hi - lo, e->stream)); // bytes between two outputs read as zeros
    r = fec_encode_launch_device(e, e->d_enc_descs, n, e->d_enc_in, e->d_enc_out, max_Z);
    if (r != FEC_OK) {
      return r;
    }
    CU_TRY(cudaMemcpyAsync(h_out + lo, e->d_enc_out + lo, hi - lo, cudaMemcpyDeviceToHost, e->stream));
    CU_TRY(cudaStreamSynchronize(e->stream));
    return FEC_OK;
  }
  CU_TRY(cudaEventRecord(hx.pipe_done[0], e->stream));
  for (int k = 0; k < 3; ++k) {
    CU_TRY(cudaStreamWaitEvent(hx.pipe[k], hx.pipe_done[0], 0));
  }
  for (uint32_t c = 0; c < spans.size(); ++c) {
    const span_t&  sp = spans[c];
    const uint32_t c0 = c * chunk, cn = (c + 1 == spans.size()) ? n - c0 : chunk;
    cudaStream_t   ps = hx.pipe[c % 3];
    CU_TRY(cudaMemcpyAsync(e->d_enc_in + sp.i0, h_in + sp.i0, sp.i1 - sp.i0, cudaMemcpyHostToDevice, hx.copy_in));
    if (hx.chunk_in.size() <= c) {
      cudaEvent_t ev = nullptr;
      CU_TRY(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
      hx.chunk_in.push_back(ev);
    }
    CU_TRY(cudaEventRecord(hx.chunk_in[c], hx.copy_in));
    CU_TRY(cudaStreamWaitEvent(ps, hx.chunk_in[c], 0));
    CU_TRY(cudaMemsetAsync(e->d_enc_out + sp.o0, 0, sp.o1 - sp.o0, ps));
    fec_encode_kernel<<<cn, FEC_ENC_THREADS, 0, ps>>>(e->d_enc_descs + c0, e->d_enc_in, e->d_enc_out, e->d_tab);
    CU_TRY(cudaGetLastError());
    e->launches.fetch_add(1);
    CU_TRY(cudaMemcpyAsync(h_out + sp.o0, e->d_enc_out + sp.o0, sp.o1 - sp.o0, cudaMemcpyDeviceToHost, ps));
  }
  for (int k = 0; k < 3; ++k) {
    CU_TRY(cudaEventRecord(hx.pipe_done[k], hx.pipe[k]));
    CU_TRY(cudaStreamWaitEvent(e->stream, hx.pipe_done[k], 0));
  }
  CU_TRY(cudaStreamSynchronize(e->stream));
  return FEC_OK;
}

// ---- the step in front of the path: soft demodulation + descrambling ---------------------------------------------------------
namespace {
const fec_demod_tables* demod_host_tables()
{
  static const fec_demod_tables* tab = [] {
    fec_demod_tables* t = new fec_demod_tables;
    fec_demod_build_tables(t);
    return t;
  }();
  return tab;
}

const fec_demod_host_jump* demod_host_jump()
{
  static const fec_demod_host_jump* jump = [] {
    fec_demod_host_jump* j = new fec_demod_host_jump;
    fec_demod_build_jump(demod_host_tables(), j);
    return j;
  }();
  return jump;
}

int demod_tables_alloc(fec_engine* e)
{
  if (e->d_dtab != nullptr) {
    return FEC_OK;
  }
  cudaError_t err = cudaMalloc((void**)&e->d_dtab, sizeof(fec_demod_tables));
  if (err == cudaSuccess) {
    err = cudaMemcpy(e->d_dtab, demod_host_tables(), sizeof(fec_demod_tables), cudaMemcpyHostToDevice);
  }
  if (err != cudaSuccess) {
    cudaGetLastError();
    return fail(FEC_ERR_CUDA, "demodulator tables: %s", cudaGetErrorString(err));
  }
  return FEC_OK;
}

int demod_launch(fec_engine* e, cudaStream_t st, const fec_demod_desc* d_descs, uint32_t n, uint32_t max_symbols,
                 const float* d_symbols, const float* d_noise, int8_t* d_llr, uint32_t flavour)
{
  if (n == 0 || max_symbols == 0) {
    return FEC_OK;
  }
  const uint32_t tiles = (max_symbols + FEC_DEMOD_THREADS - 1) / FEC_DEMOD_THREADS;
  if (tiles > 65535u) {
    return fail(FEC_ERR_BAD_ARG, "%u symbols in one demodulation call", max_symbols);
  }
  fec_demod_kernel<<<dim3(n, tiles), FEC_DEMOD_THREADS, 0, st>>>(d_descs, reinterpret_cast<const float2*>(d_symbols), d_noise,
                                                                d_llr, e->d_dtab, flavour);
  CU_TRY(cudaGetLastError());
  e->launches.fetch_add(1);
  return FEC_OK;
}

// True when the non-empty demodulation calls tile [l0, l1) exactly: each starts where the previous one (by LLR offset) ended.
// (A sum-of-lengths test alone would accept two overlapping calls plus a gap of the same size.)
bool demod_calls_tile(const fec_demod_desc* d, uint32_t n, size_t l0, size_t l1)
{
  std::vector<std::pair<size_t, size_t>> span;
  span.reserve(n);
  for (uint32_t i = 0; i < n; ++i) {
    if (d[i].nof_symbols != 0) {
      span.emplace_back((size_t)d[i].llr_offset, (size_t)d[i].nof_symbols * d[i].qm);
    }
  }
  std::sort(span.begin(), span.end());
  size_t at = l0;
  for (const auto& sp : span) {
    if (sp.first != at) {
      return false;
    }
    at += sp.second;
  }
  return at == l1;
}

// Copies the symbols the descriptors use to the device and demodulates them into d_llr, whose first byte is LLR index l0.
int demod_stage(fec_engine* e, cudaStream_t st, const demod_source& dm, size_t l0, size_t l1, int8_t* d_llr)
{
  if (!dm.descs || !dm.h_symbols || !dm.h_noise) {
    return fail(FEC_ERR_BAD_ARG, "null argument");
  }
  if (dm.flavour > FEC_DEMOD_FLAVOUR_AVX512) {
    return fail(FEC_ERR_BAD_ARG, "flavour=%u", dm.flavour);
  }
  int rc = demod_tables_alloc(e);
  if (rc != FEC_OK) {
    return rc;
  }
  size_t   s0 = ~(size_t)0, s1 = 0, covered = 0;
  uint32_t max_symbols = 0;
  if ((size_t)dm.n > e->h_dm_descs_cap) {
    cudaFreeHost(e->h_dm_descs);
    e->h_dm_descs     = nullptr;
    e->h_dm_descs_cap = 0;
    const size_t cap  = (size_t)dm.n * 2;
    if (cudaHostAlloc((void**)&e->h_dm_descs, cap * sizeof(fec_demod_desc), cudaHostAllocDefault) != cudaSuccess) {
      cudaGetLastError();
      return fail(FEC_ERR_NO_MEMORY, "cannot allocate pinned memory for %u demodulation descriptors", dm.n);
    }
    e->h_dm_descs_cap = cap;
  }
  memcpy(e->h_dm_descs, dm.descs, (size_t)dm.n * sizeof(fec_demod_desc));
  for (uint32_t i = 0; i < dm.n; ++i) {
    fec_demod_desc& d = e->h_dm_descs[i];
    rc                = fec_demod_desc_prepare(&d);
    if (rc != FEC_OK) {
      return rc;
    }
    const size_t bytes = (size_t)d.nof_symbols * d.qm;
    if (d.sym_offset + d.nof_symbols > dm.nof_symbols || d.llr_offset < l0 || d.llr_offset + bytes > l1) {
      return fail(FEC_ERR_BAD_ARG, "demodulation %u indexes outside the symbol or LLR buffers", i);
    }
    if (d.nof_symbols == 0) {
      continue;
    }
    s0          = d.sym_offset < s0 ? (size_t)d.sym_offset : s0;
    s1          = d.sym_offset + d.nof_symbols > s1 ? (size_t)(d.sym_offset + d.nof_symbols) : s1;
    max_symbols = d.nof_symbols > max_symbols ? d.nof_symbols : max_symbols;
    covered += bytes;
  }
  if (covered != l1 - l0 || !demod_calls_tile(e->h_dm_descs, dm.n, l0, l1)) {
    return fail(FEC_ERR_BAD_ARG, "the demodulation calls (%zu LLRs) do not tile the %zu LLRs the transport blocks read", covered,
                l1 - l0);
  }
  for (uint32_t i = 0; i < dm.n; ++i) {
    e->h_dm_descs[i].sym_offset -= s0;
    e->h_dm_descs[i].llr_offset -= l0;
  }
  rc = grow((void**)&e->d_dm_descs, &e->dm_descs_cap, (size_t)dm.n * sizeof(fec_demod_desc));
  rc = (rc == FEC_OK) ? grow((void**)&e->d_dm_sym, &e->dm_sym_cap, (s1 - s0) * 2 * sizeof(float)) : rc;
  rc = (rc == FEC_OK) ? grow((void**)&e->d_dm_noise, &e->dm_noise_cap, (s1 - s0) * sizeof(float)) : rc;
  if (rc != FEC_OK) {
    return rc;
  }
  CU_TRY(cudaMemcpyAsync(e->d_dm_descs, e->h_dm_descs, (size_t)dm.n * sizeof(fec_demod_desc), cudaMemcpyHostToDevice, st));
  CU_TRY(cudaMemcpyAsync(e->d_dm_sym, dm.h_symbols + 2 * s0, (s1 - s0) * 2 * sizeof(float), cudaMemcpyHostToDevice, st));
  CU_TRY(cudaMemcpyAsync(e->d_dm_noise, dm.h_noise + s0, (s1 - s0) * sizeof(float), cudaMemcpyHostToDevice, st));
  return demod_launch(e, st, e->d_dm_descs, dm.n, max_symbols, e->d_dm_sym, e->d_dm_noise, d_llr, dm.flavour);
}
} // namespace

int fec_demod_desc_prepare(fec_demod_desc* d)
{
  if (d == nullptr) {
    return fail(FEC_ERR_BAD_ARG, "null descriptor");
  }
  if (d->qm != 1 && d->qm != 2 && d->qm != 4 && d->qm != 6 && d->qm != 8) {
    return fail(FEC_ERR_BAD_ARG, "demodulation: qm=%u", d->qm);
  }
  if ((uint64_t)d->seq_offset + (uint64_t)d->nof_symbols * d->qm + 1600u > 0xffffffffull) {
    return fail(FEC_ERR_BAD_ARG, "demodulation: sequence position beyond 2^32");
  }
  fec_demod_lfsr_states(demod_host_jump(), d->c_init, d->seq_offset, &d->x1_state, &d->x2_state);
  return FEC_OK;
}

int fec_demod_launch_device(fec_engine* e, const fec_demod_desc* d_descs, uint32_t n, uint32_t max_symbols,
                            const float* d_symbols, const float* d_noise_vars, int8_t* d_llr, uint32_t flavour)
{
  if (!e || !d_descs || !d_symbols || !d_noise_vars || !d_llr) {
    return fail(FEC_ERR_BAD_ARG, "null argument");
  }
  if (flavour > FEC_DEMOD_FLAVOUR_AVX512) {
    return fail(FEC_ERR_BAD_ARG, "flavour=%u", flavour);
  }
  CU_TRY(cudaSetDevice(e->cfg.device));
  int rc = demod_tables_alloc(e);
  if (rc != FEC_OK) {
    return rc;
  }
  return demod_launch(e, e->stream, d_descs, n, max_symbols, d_symbols, d_noise_vars, d_llr, flavour);
}

int fec_demod_run_host(fec_engine* e, const fec_demod_desc* h_descs, uint32_t n, const float* h_symbols,
                       const float* h_noise_vars, size_t nof_symbols, int8_t* h_llr, size_t llr_bytes, uint32_t flavour)
{
  if (!e || !h_descs || !h_symbols || !h_noise_vars || !h_llr) {
    return fail(FEC_ERR_BAD_ARG, "null argument");
  }
  if (n == 0) {
    return FEC_OK;
  }
  size_t l0 = ~(size_t)0, l1 = 0, covered = 0;
  for (uint32_t i = 0; i < n; ++i) {
    const size_t bytes = (size_t)h_descs[i].nof_symbols * h_descs[i].qm;
    if (h_descs[i].llr_offset + bytes > llr_bytes) {
      return fail(FEC_ERR_BAD_ARG, "demodulation %u writes outside the LLR buffer", i);
    }
    if (bytes == 0) {
      continue;
    }
    l0 = h_descs[i].llr_offset < l0 ? (size_t)h_descs[i].llr_offset : l0;
    l1 = h_descs[i].llr_offset + bytes > l1 ? (size_t)(h_descs[i].llr_offset + bytes) : l1;
    covered += bytes;
  }
  if (covered == 0) {
    return FEC_OK;
  }
  if (covered != l1 - l0 || !demod_calls_tile(h_descs, n, l0, l1)) {
    return fail(FEC_ERR_BAD_ARG, "the demodulation calls must tile one contiguous span of the LLR buffer");
  }
  ctx_lease lease(e, 0); // the demodulator staging is engine-wide: one call at a time
  CU_TRY(cudaSetDevice(e->cfg.device));
  int rc = grow((void**)&e->d_dm_llr, &e->dm_llr_cap, l1 - l0);
  if (rc != FEC_OK) {
    return rc;
  }
  const demod_source dm = {h_descs, n, h_symbols, h_noise_vars, nof_symbols, flavour};
  rc                    = demod_stage(e, e->stream, dm, l0, l1, e->d_dm_llr);
  if (rc != FEC_OK) {
    return rc;
  }
  CU_TRY(cudaMemcpyAsync(h_llr + l0, e->d_dm_llr, l1 - l0, cudaMemcpyDeviceToHost, e->stream));
  CU_TRY(cudaStreamSynchronize(e->stream));
  return FEC_OK;
}

int fec_tb_run_host_symbols(fec_engine* e, const fec_tb_desc* tbs, uint32_t n, const uint32_t* slots,
                            const fec_demod_desc* h_demod, uint32_t n_demod, const float* h_symbols,
                            const float* h_noise_vars, size_t nof_symbols, uint32_t flavour, uint8_t* h_out, size_t out_bytes,
                            fec_tb_result* h_results, fec_cb_result* h_cb_results)
{
  if (!h_demod || !h_symbols || !h_noise_vars) {
    return fail(FEC_ERR_BAD_ARG, "null argument");
  }
  const demod_source dm = {h_demod, n_demod, h_symbols, h_noise_vars, nof_symbols, flavour};
  return tb_run(e, tbs, n, slots, nullptr, ~(size_t)0 >> 1, &dm, h_out, out_bytes, h_results, h_cb_results);
}


// ---- allocation-free transport-block assembly (no counterpart in the reference) ----
// fec_tb_gather_kernel with its scratch (jobs | codeblock map | per-block {CRC accumulator, arrival counter}) in a file-scope
// device array instead of cudaMalloc'ed engine memory. Every GPU that loads this library holds the array until exit.
namespace {
constexpr size_t NOALLOC_OFF_MAP   = ((size_t)FEC_NOALLOC_MAX_TBS * sizeof(fec_tb_job) + 255) & ~(size_t)255;
constexpr size_t NOALLOC_OFF_ACC   = (NOALLOC_OFF_MAP + (size_t)FEC_NOALLOC_MAX_CBS * sizeof(fec_tb_cb_ref) + 255) & ~(size_t)255;
constexpr size_t NOALLOC_BYTES     = (NOALLOC_OFF_ACC + (size_t)FEC_NOALLOC_MAX_TBS * 2 * sizeof(uint32_t) + 255) & ~(size_t)255;
}
static __device__ __align__(256) unsigned char fec_noalloc_scratch[NOALLOC_BYTES];
namespace {
// In host code `fec_noalloc_scratch` is only a handle: the device address comes from cudaGetSymbolAddress, once per device,
// with that device current.
constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex     noalloc_mtx;                                  // guards the two arrays below and serialises the calls
unsigned char* noalloc_dev_ptr[NOALLOC_MAX_DEVICES] = {};
unsigned char  noalloc_host_stage[NOALLOC_BYTES];            // host image of the scratch, rebuilt by every call
}

int fec_tb_assemble_device_noalloc(fec_engine* e, const fec_tb_desc* tbs, uint32_t n, const uint8_t* d_cb_msgs,
                                   const fec_cb_result* d_cb_results, uint8_t* d_out, size_t out_bytes,
                                   fec_tb_result* d_tb_results)
{
  if (!e || !tbs || !d_cb_msgs || !d_cb_results || !d_out || !d_tb_results) {
    return fail(FEC_ERR_BAD_ARG, "null argument");
  }
  if (n == 0) {
    return FEC_OK;
  }
  if (n > FEC_NOALLOC_MAX_TBS) {
    return fail(FEC_ERR_BAD_ARG, "fec_tb_assemble_device_noalloc: %u transport blocks, the static scratch takes %u", n,
                FEC_NOALLOC_MAX_TBS);
  }
  if (e->cfg.device < 0 || e->cfg.device >= NOALLOC_MAX_DEVICES) {
    return fail(FEC_ERR_BAD_ARG, "fec_tb_assemble_device_noalloc: device ordinal %d", e->cfg.device);
  }
  std::lock_guard<std::mutex> lk(noalloc_mtx);
  CU_TRY(cudaSetDevice(e->cfg.device));
  if (noalloc_dev_ptr[e->cfg.device] == nullptr) {
    void* p = nullptr;
    CU_TRY(cudaGetSymbolAddress(&p, fec_noalloc_scratch));
    noalloc_dev_ptr[e->cfg.device] = static_cast<unsigned char*>(p);
  }
  unsigned char* d_scr = noalloc_dev_ptr[e->cfg.device];
  fec_tb_job*    jobs  = reinterpret_cast<fec_tb_job*>(noalloc_host_stage);
  fec_tb_cb_ref* map   = reinterpret_cast<fec_tb_cb_ref*>(noalloc_host_stage + NOALLOC_OFF_MAP);
  uint32_t       m     = 0;
  for (uint32_t t = 0; t < n; ++t) {
    fec_tb_geometry g;
    if (!fec_tb_geometry_of(tbs[t], g)) {
      return fail(FEC_ERR_BAD_ARG, "transport block %u: tbs=%u bg=%u rv=%u qm=%u layers=%u llrs=%u", t, tbs[t].tbs, tbs[t].base_graph,
                  tbs[t].rv, tbs[t].qm, tbs[t].nof_layers, tbs[t].nof_llrs);
    }
    if (g.data_bits % 8 != 0) {
      return fail(FEC_ERR_BAD_ARG, "transport block %u: payloads of %u bits are not whole bytes (use fec_tb_run_host)", t, g.data_bits);
    }
    if (m + g.nof_cbs > FEC_NOALLOC_MAX_CBS) {
      return fail(FEC_ERR_BAD_ARG, "fec_tb_assemble_device_noalloc: more than %u codeblocks in one call", FEC_NOALLOC_MAX_CBS);
    }
    if (tbs[t].out_offset + tbs[t].tbs / 8 > out_bytes) {
      return fail(FEC_ERR_BAD_ARG, "transport block %u does not fit the output buffer", t);
    }
    fec_tb_job& J = jobs[t];
    memset(&J, 0, sizeof(J));
    J.out_offset  = tbs[t].out_offset;
    J.nof_cbs     = g.nof_cbs;
    J.data_bits   = g.data_bits;
    J.stream_bits = g.stream_bits;
    J.tbs_bytes   = tbs[t].tbs / 8;
    J.map_first   = m;
    J.max_iter    = tbs[t].max_iter;
    // weight of codeblock r = x^(8 * stream bytes after its payload) mod g24A, built backwards (as tb_run does)
    const uint32_t dbytes = g.data_bits / 8;
    uint32_t       w      = 1;
    const uint32_t step   = fec_tb_crc24a_xpow8_host(dbytes);
    if (g.nof_cbs > 1) {
      w = fec_tb_crc24a_xpow8_host(g.stream_bits / 8 - (g.nof_cbs - 1) * dbytes);
    }
    for (int r = (int)g.nof_cbs - 1; r >= 0; --r) {
      fec_tb_cb_ref& ref = map[m + r];
      ref.slot           = m + r; // codeblock messages are densely packed in call order
      ref.result_index   = (int32_t)(m + r);
      ref.tb             = t;
      if (r == (int)g.nof_cbs - 1) {
        ref.weight = 1u;
      } else {
        ref.weight = w;
        w          = fec_tb_crc24a_mulmod_host(w, step);
      }
    }
    m += g.nof_cbs;
  }
  // The array keeps the previous call's contents: the accumulators and arrival counters must start at zero.
  memset(noalloc_host_stage + NOALLOC_OFF_ACC, 0, (size_t)n * 2 * sizeof(uint32_t));
  const cudaStream_t st = e->stream;
  CU_TRY(cudaMemcpyAsync(d_scr, noalloc_host_stage, (size_t)n * sizeof(fec_tb_job), cudaMemcpyHostToDevice, st));
  CU_TRY(cudaMemcpyAsync(d_scr + NOALLOC_OFF_MAP, noalloc_host_stage + NOALLOC_OFF_MAP, (size_t)m * sizeof(fec_tb_cb_ref),
                         cudaMemcpyHostToDevice, st));
  CU_TRY(cudaMemcpyAsync(d_scr + NOALLOC_OFF_ACC, noalloc_host_stage + NOALLOC_OFF_ACC, (size_t)n * 2 * sizeof(uint32_t),
                         cudaMemcpyHostToDevice, st));
  fec_tb_gather_kernel<<<m, 256, 0, st>>>(reinterpret_cast<const fec_tb_job*>(d_scr),
                                          reinterpret_cast<const fec_tb_cb_ref*>(d_scr + NOALLOC_OFF_MAP), d_cb_results, d_cb_msgs,
                                          d_out, d_tb_results, reinterpret_cast<uint32_t*>(d_scr + NOALLOC_OFF_ACC), e->d_tab, 0u);
  CU_TRY(cudaGetLastError());
  e->launches.fetch_add(1);
  return FEC_OK;
}

} // extern "C"


// ---- UL-SCH demultiplexing (fec_ulsch.cuh) --------------------------------------------------------------------------------
int fec_ulsch_demux_sizes(const fec_ulsch_cfg* cfg, fec_ulsch_sizes* out)
{
  if (!cfg || !out) {
    return fail(FEC_ERR_BAD_ARG, "null argument");
  }
  fec_ulsch_plan P;
  if (!fec_ulsch_plan_build(*cfg, P)) {
    return fail(FEC_ERR_BAD_ARG, "invalid UL-SCH multiplexing configuration");
  }
  out->nof_codeword_bits  = (uint32_t)P.flags.size() * P.bits_per_re;
  out->nof_sch_bits       = P.nof_re[0] * P.bits_per_re;
  out->nof_harq_ack_bits  = P.nof_re[1] * P.bits_per_re;
  out->nof_csi_part1_bits = P.nof_re[2] * P.bits_per_re;
  out->nof_csi_part2_bits = P.nof_re[3] * P.bits_per_re;
  return FEC_OK;
}

namespace {
// The plan of a configuration, resident in HBM. Called and used under e->ulsch_mtx (an eviction frees device memory, which
// waits for the kernels that read it).
int ulsch_plan_get(fec_engine* e, const fec_ulsch_cfg& cfg, fec_engine::ulsch_entry** out)
{
  for (auto& en : e->ulsch_cache) {
    if (memcmp(&en.cfg, &cfg, sizeof(cfg)) == 0) {
      en.stamp = ++e->ulsch_clock;
      *out     = &en;
      return FEC_OK;
    }
  }
  fec_engine::ulsch_entry en;
  en.cfg = cfg;
  if (!fec_ulsch_plan_build(cfg, en.plan)) {
    return fail(FEC_ERR_BAD_ARG, "invalid UL-SCH multiplexing configuration");
  }
  CU_TRY(cudaSetDevice(e->cfg.device));
  if (e->ulsch_cache.size() >= 16) { // drop the least recently used
    size_t lru = 0;
    for (size_t i = 1; i < e->ulsch_cache.size(); ++i) {
      lru = (e->ulsch_cache[i].stamp < e->ulsch_cache[lru].stamp) ? i : lru;
    }
    cudaFree(e->ulsch_cache[lru].d_segs);
    cudaFree(e->ulsch_cache[lru].d_flags);
    e->ulsch_cache.erase(e->ulsch_cache.begin() + lru);
  }
  const size_t fb = en.plan.flags.size(), sb = en.plan.segs.size() * sizeof(fec_ulsch_seg);
  CU_TRY(cudaMalloc((void**)&en.d_flags, fb ? fb : 1));
  CU_TRY(cudaMalloc((void**)&en.d_segs, sb ? sb : 1));
  CU_TRY(cudaMemcpy(en.d_flags, en.plan.flags.data(), fb, cudaMemcpyHostToDevice));
  CU_TRY(cudaMemcpy(en.d_segs, en.plan.segs.data(), sb, cudaMemcpyHostToDevice));
  en.stamp = ++e->ulsch_clock;
  e->ulsch_cache.push_back(std::move(en));
  *out = &e->ulsch_cache.back();
  return FEC_OK;
}

int ulsch_launch(fec_engine* e, cudaStream_t st, const fec_ulsch_cfg& cfg, const int8_t* d_cw, const uint8_t* d_seq,
                 int8_t* d_sch, int8_t* d_harq, int8_t* d_csi1, int8_t* d_csi2, fec_ulsch_sizes* sizes)
{
  std::lock_guard<std::mutex> lk(e->ulsch_mtx);
  fec_engine::ulsch_entry*    en = nullptr;
  int                         rc = ulsch_plan_get(e, cfg, &en);
  if (rc != FEC_OK) {
    return rc;
  }
  const fec_ulsch_plan& P = en->plan;
  fec_ulsch_launch      L;
  L.bits_per_re = P.bits_per_re;
  L.qm          = cfg.qm;
  L.mode[0]     = 0;
  L.mode[1]     = (cfg.nof_harq_ack_bits <= 2) ? cfg.nof_harq_ack_bits : 0;
  L.mode[2]     = (cfg.nof_csi_part1_bits <= 2) ? cfg.nof_csi_part1_bits : 0;
  L.mode[3]     = (cfg.nof_csi_part2_bits <= 2) ? cfg.nof_csi_part2_bits : 0;
  L.harq_punctures = (cfg.nof_harq_ack_bits == 1 || cfg.nof_harq_ack_bits == 2) ? 1u : 0u;
  const bool needs_seq = cfg.qm > 1 && ((P.nof_re[1] && L.mode[1]) || (P.nof_re[2] && L.mode[2]) || (P.nof_re[3] && L.mode[3]));
  if (needs_seq && !d_seq) {
    return fail(FEC_ERR_BAD_ARG, "a UCI field of one or two bits needs the scrambling sequence of the codeword");
  }
  if (!d_cw || (P.nof_re[0] && !d_sch) || (P.nof_re[1] && !d_harq) || (P.nof_re[2] && !d_csi1) || (P.nof_re[3] && !d_csi2)) {
    return fail(FEC_ERR_BAD_ARG, "null buffer for a stream that receives soft bits");
  }
  if (sizes) {
    sizes->nof_codeword_bits  = (uint32_t)P.flags.size() * P.bits_per_re;
    sizes->nof_sch_bits       = P.nof_re[0] * P.bits_per_re;
    sizes->nof_harq_ack_bits  = P.nof_re[1] * P.bits_per_re;
    sizes->nof_csi_part1_bits = P.nof_re[2] * P.bits_per_re;
    sizes->nof_csi_part2_bits = P.nof_re[3] * P.bits_per_re;
  }
  if (P.segs.empty()) {
    return FEC_OK;
  }
  CU_TRY(cudaSetDevice(e->cfg.device));
  fec_ulsch_demux_kernel<<<(unsigned)P.segs.size(), FEC_ULSCH_SEG, 0, st>>>(en->d_segs, en->d_flags, L, d_cw, d_seq, d_sch, d_harq,
                                                                           d_csi1, d_csi2);
  CU_TRY(cudaGetLastError());
  e->launches.fetch_add(1);
  return FEC_OK;
}
} // namespace

int fec_ulsch_demux_launch_device(fec_engine* e, const fec_ulsch_cfg* cfg, const int8_t* d_codeword,
                                  const uint8_t* d_scrambling_seq, int8_t* d_sch, int8_t* d_harq_ack, int8_t* d_csi_part1,
                                  int8_t* d_csi_part2)
{
  if (!e || !cfg) {
    return fail(FEC_ERR_BAD_ARG, "null argument");
  }
  return ulsch_launch(e, e->stream, *cfg, d_codeword, d_scrambling_seq, d_sch, d_harq_ack, d_csi_part1, d_csi_part2, nullptr);
}

int fec_ulsch_demux_run_host(fec_engine* e, const fec_ulsch_cfg* cfg, const int8_t* h_codeword, const uint8_t* h_scrambling_seq,
                             int8_t* h_sch, int8_t* h_harq_ack, int8_t* h_csi_part1, int8_t* h_csi_part2)
{
  if (!e || !cfg || !h_codeword) {
    return fail(FEC_ERR_BAD_ARG, "null argument");
  }
  fec_ulsch_sizes z;
  int             rc = fec_ulsch_demux_sizes(cfg, &z);
  if (rc != FEC_OK) {
    return rc;
  }
  CU_TRY(cudaSetDevice(e->cfg.device));
  // one device allocation for everything (this is the convenience path; the production path keeps all of it in HBM)
  const size_t a16 = 15;
  const size_t o_cw = 0, o_seq = (o_cw + z.nof_codeword_bits + a16) & ~a16, o_sch = (o_seq + (z.nof_codeword_bits + 7) / 8 + a16) & ~a16,
               o_h = (o_sch + z.nof_sch_bits + a16) & ~a16, o_c1 = (o_h + z.nof_harq_ack_bits + a16) & ~a16,
               o_c2 = (o_c1 + z.nof_csi_part1_bits + a16) & ~a16, total = o_c2 + z.nof_csi_part2_bits + 16;
  uint8_t* d = nullptr;
  CU_TRY(cudaMalloc((void**)&d, total));
  cudaStream_t st = e->stream;
  cudaError_t  err = cudaMemcpyAsync(d + o_cw, h_codeword, z.nof_codeword_bits, cudaMemcpyHostToDevice, st);
  if (err == cudaSuccess && h_scrambling_seq) {
    err = cudaMemcpyAsync(d + o_seq, h_scrambling_seq, (z.nof_codeword_bits + 7) / 8, cudaMemcpyHostToDevice, st);
  }
  if (err == cudaSuccess) {
    rc = ulsch_launch(e, st, *cfg, (const int8_t*)(d + o_cw), h_scrambling_seq ? d + o_seq : nullptr, (int8_t*)(d + o_sch),
                      (int8_t*)(d + o_h), (int8_t*)(d + o_c1), (int8_t*)(d + o_c2), nullptr);
  }
  const struct {
    int8_t*  h;
    size_t   off;
    uint32_t n;
  } outs[4] = {{h_sch, o_sch, z.nof_sch_bits}, {h_harq_ack, o_h, z.nof_harq_ack_bits}, {h_csi_part1, o_c1, z.nof_csi_part1_bits},
               {h_csi_part2, o_c2, z.nof_csi_part2_bits}};
  for (int k = 0; k < 4 && err == cudaSuccess && rc == FEC_OK; ++k) {
    if (outs[k].n) {
      if (!outs[k].h) {
        rc = fail(FEC_ERR_BAD_ARG, "null buffer for a stream that receives soft bits");
        break;
      }
      err = cudaMemcpyAsync(outs[k].h, d + outs[k].off, outs[k].n, cudaMemcpyDeviceToHost, st);
    }
  }
  const cudaError_t serr = cudaStreamSynchronize(st);
  cudaFree(d);
  if (rc != FEC_OK) {
    return rc;
  }
  if (err != cudaSuccess || serr != cudaSuccess) {
    return fail(FEC_ERR_CUDA, "UL-SCH demultiplexing failed on the device: %s", cudaGetErrorString(err != cudaSuccess ? err : serr));
  }
  return FEC_OK;
}

#ifdef FEC_PROFILE
// Profiling build only (profiles/scripts/phase_clocks.py): the phase clocks accumulated by fec_cb_packed_kernel.
extern "C" int fec_debug_profile_read(unsigned long long* out, int reset)
{
  cudaDeviceSynchronize();
  if (cudaMemcpyFromSymbol(out, fec_prof_acc, sizeof(fec_prof_acc)) != cudaSuccess) {
    return FEC_ERR_CUDA;
  }
  if (reset) {
    unsigned long long zero[FEC_PROF_PHASES] = {};
    zero[11] = ~0ull; // earliest CTA start (atomicMin)
    cudaMemcpyToSymbol(fec_prof_acc, zero, sizeof(zero));
  }
  return FEC_OK;
}
#endif

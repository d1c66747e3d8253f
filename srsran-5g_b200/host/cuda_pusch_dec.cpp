// srsran-5g_b200/host/cuda_pusch_dec.cpp -- see cuda_pusch_dec.h. Host-only C++17; all device work is behind the C ABI.
#include "cuda_pusch_dec.h"
#include "srsran/support/error_handling.h"
#include <cstring>
#include <thread>
#include <vector>

using namespace srsran;
using namespace srsran::b200;

namespace {

uint8_t to_fec_crc(hal::hw_dec_cb_crc_type t)
{
  // Same numbering on both sides (hw_accelerator_pusch_dec.h:36 / fec_b200.h).
  return static_cast<uint8_t>(t);
}

uint8_t to_fec_crc(crc_generator_poly poly)
{
  switch (poly) {
    case crc_generator_poly::CRC24A:
      return FEC_CRC24A;
    case crc_generator_poly::CRC24B:
      return FEC_CRC24B;
    case crc_generator_poly::CRC16:
      return FEC_CRC16;
    default:
      report_fatal_error("The CUDA LDPC decoder checks CRC16, CRC24A or CRC24B only.");
  }
}

} // namespace

fec_engine_handle::fec_engine_handle(const fec_engine_cfg& cfg)
{
  // No CPU fallback: a missing device is fatal at start-up, exactly where the bbdev factory would fail.
  report_fatal_error_if_not(fec_engine_create(&cfg, &engine) == FEC_OK, "B200 FEC engine: {}", fec_last_error());
}

fec_engine_handle::~fec_engine_handle()
{
  fec_engine_destroy(engine);
}

cuda_pusch_dec_accelerator::cuda_pusch_dec_accelerator(std::shared_ptr<fec_engine_handle> engine_) :
  engine(std::move(engine_))
{
}

cuda_pusch_dec_accelerator::~cuda_pusch_dec_accelerator()
{
  if (has_queue) {
    fec_queue_free(engine->get(), queue);
  }
}

void cuda_pusch_dec_accelerator::reserve_queue()
{
  if (!has_queue) {
    report_fatal_error_if_not(fec_queue_reserve(engine->get(), &queue) == FEC_OK, "B200 FEC engine: {}", fec_last_error());
    has_queue = true;
  }
}

void cuda_pusch_dec_accelerator::free_queue()
{
  // The reference calls this once per transport block (pusch_decoder_hw_impl.cpp:439); the engine queue itself is kept
  // for the lifetime of the instance so that its pinned staging is not re-created every slot.
  for (operation& op : ops) {
    op.enqueued = false;
    op.done     = false;
    op.rejected = false;
  }
  dirty = false;
}

void cuda_pusch_dec_accelerator::configure_operation(const hal::hw_pusch_decoder_configuration& config, unsigned cb_index)
{
  srsran_assert(cb_index < MAX_NOF_SEGMENTS, "Codeblock index {} out of range.", cb_index);
  ops[cb_index].cfg      = config;
  ops[cb_index].enqueued = false;
  ops[cb_index].done     = false;
  ops[cb_index].rejected = false;
}

bool cuda_pusch_dec_accelerator::enqueue_operation(span<const int8_t> data, span<const int8_t> /*aux_data*/, unsigned cb_index)
{
  srsran_assert(has_queue, "enqueue_operation() before reserve_queue().");
  const hal::hw_pusch_decoder_configuration& c = ops[cb_index].cfg;
  srsran_assert(data.size() == c.cw_length, "Codeblock {}: {} soft bits given, cw_length is {}.", cb_index, data.size(),
                c.cw_length);
  fec_cb_desc d     = {};
  d.harq_slot       = c.absolute_cb_id;
  d.E               = c.cw_length;
  d.Nref            = c.Nref;
  d.nof_filler_bits = c.nof_filler_bits;
  d.lifting_size    = static_cast<uint16_t>(c.lifting_size);
  d.base_graph      = static_cast<uint8_t>(c.base_graph_index);
  d.rv              = static_cast<uint8_t>(c.rv);
  d.qm              = static_cast<uint8_t>(get_bits_per_symbol(c.modulation));
  d.crc_type        = to_fec_crc(c.cb_crc_type);
  d.max_iter        = static_cast<uint8_t>(c.max_nof_ldpc_iterations);
  d.mode            = FEC_MODE_CHAIN;
  d.flags           = (c.new_data ? FEC_FLAG_NEW_DATA : 0u) | (c.use_early_stop ? FEC_FLAG_EARLY_STOP : 0u);
  const int rc      = fec_queue_enqueue(engine->get(), queue, &d, data.data(), cb_index);
  if (rc == FEC_ERR_QUEUE_FULL) {
    return false; // the consumer stops enqueuing and drains (pusch_decoder_hw_impl.cpp:253-257)
  }
  ops[cb_index].enqueued = true;
  if (rc != FEC_OK) {
    // Anything else (a parameter the engine rejects, a CUDA failure) is a failed codeblock, not a dead DU: the operation
    // counts as accepted and completes at once with CRC_pass = false (the consumer then treats it like any decoding failure).
    ops[cb_index].result                     = {};
    ops[cb_index].result.nof_ldpc_iterations = static_cast<uint16_t>(c.max_nof_ldpc_iterations);
    ops[cb_index].result.status              = FEC_RES_DONE;
    ops[cb_index].done                       = true;
    ops[cb_index].rejected                   = true;
    return true;
  }
  dirty = true;
  return true;
}

void cuda_pusch_dec_accelerator::poll()
{
  std::array<uint64_t, 64>      tags;
  std::array<fec_cb_result, 64> results;
  for (;;) {
    const int n = fec_queue_poll(engine->get(), queue, tags.data(), results.data(), tags.size());
    for (int i = 0; i < n; ++i) {
      operation& op = ops[tags[i]];
      op.result     = results[i];
      op.done       = true;
    }
    if (n < static_cast<int>(tags.size())) {
      break;
    }
  }
}

bool cuda_pusch_dec_accelerator::dequeue_operation(span<uint8_t> data, span<int8_t> /*aux_data*/, unsigned segment_index)
{
  operation& op = ops[segment_index];
  if (!op.enqueued) {
    return false;
  }
  if (!op.done) {
    if (dirty) {
      // First dequeue after a run of enqueues: launch whatever is still staged (the consumer enqueues every codeblock
      // of the transport block before it starts polling because HARQ is external).
      fec_queue_flush(engine->get(), queue);
      dirty = false;
    }
    poll();
    if (!op.done) {
      // The consumer spins on dequeue_operation (pusch_decoder_hw_impl.cpp:289-317): give the core away between two looks, so
      // that more decoder threads than cores do not starve the threads that still have codeblocks to enqueue.
      if (++idle_polls >= 64) {
        idle_polls = 0;
        std::this_thread::yield();
      }
      return false;
    }
  }
  if (!op.rejected) {
    fec_queue_fetch(engine->get(), queue, segment_index, data.data(), static_cast<uint32_t>(data.size()));
  }
  return true;
}

void cuda_pusch_dec_accelerator::read_operation_outputs(hal::hw_pusch_decoder_outputs& out,
                                                        unsigned                       cb_index,
                                                        unsigned /*absolute_cb_id*/)
{
  const operation& op     = ops[cb_index];
  out.CRC_pass            = (op.result.status & FEC_RES_CRC_PASS) != 0;
  out.nof_ldpc_iterations = op.result.nof_ldpc_iterations;
}

void cuda_pusch_dec_accelerator::free_harq_context_entry(unsigned absolute_cb_id)
{
  fec_harq_free(engine->get(), absolute_cb_id);
}

std::shared_ptr<hal::hw_accelerator_pusch_dec_factory> srsran::b200::create_cuda_pusch_dec_acc_factory(const fec_engine_cfg& cfg)
{
  return std::make_shared<cuda_pusch_dec_factory>(cfg);
}

pusch_decoder_cuda::pusch_decoder_cuda(std::shared_ptr<fec_engine_handle> engine_, unsigned max_nof_softbits, bool asynchronous_) :
  engine(std::move(engine_)), softbits_cap(max_nof_softbits), asynchronous(asynchronous_)
{
  static_assert(sizeof(log_likelihood_ratio) == 1, "LLRs are int8 on both sides of the boundary.");
  softbits = static_cast<log_likelihood_ratio*>(fec_host_alloc(max_nof_softbits));
  report_fatal_error_if_not(softbits != nullptr, "B200 FEC engine: cannot pin {} bytes for the codeword buffer.", max_nof_softbits);
}

pusch_decoder_cuda::~pusch_decoder_cuda()
{
  fec_host_free(softbits);
}

pusch_decoder_buffer& pusch_decoder_cuda::new_data(span<uint8_t>           transport_block_,
                                                   unique_rx_buffer        rm_buffer_,
                                                   pusch_decoder_notifier& notifier_,
                                                   const configuration&    cfg)
{
  transport_block = transport_block_;
  rm_buffer       = std::move(rm_buffer_);
  notifier        = &notifier_;
  current_config  = cfg;
  softbits_count  = 0;
  return *this;
}

span<log_likelihood_ratio> pusch_decoder_cuda::get_next_block_view(unsigned block_size)
{
  srsran_assert(softbits_count + block_size <= softbits_cap, "Codeword of more than {} soft bits.", softbits_cap);
  return span<log_likelihood_ratio>(softbits + softbits_count, block_size);
}

void pusch_decoder_cuda::on_new_softbits(span<const log_likelihood_ratio> in)
{
  span<log_likelihood_ratio> block = get_next_block_view(in.size());
  if (block.data() != in.data()) {
    std::memcpy(block.data(), in.data(), in.size());
  }
  softbits_count += in.size();
}

void pusch_decoder_cuda::on_end_softbits()
{
  nof_cbs = rm_buffer->get_nof_codeblocks();
  srsran_assert(nof_cbs <= MAX_NOF_SEGMENTS, "Too many codeblocks.");
  for (unsigned i = 0; i != nof_cbs; ++i) {
    slots[i] = rm_buffer->get_absolute_codeblock_id(i);
  }
  fec_tb_desc tb = {};
  tb.tbs         = static_cast<uint32_t>(transport_block.size() * 8);
  tb.nof_llrs    = softbits_count;
  tb.Nref        = current_config.Nref;
  tb.slot_list   = 0;
  tb.base_graph  = (current_config.base_graph == ldpc_base_graph_type::BG1) ? 1 : 2;
  tb.rv          = static_cast<uint8_t>(current_config.rv);
  tb.qm          = static_cast<uint8_t>(get_bits_per_symbol(current_config.mod));
  tb.nof_layers  = static_cast<uint8_t>(current_config.nof_layers);
  tb.max_iter    = static_cast<uint8_t>(current_config.nof_ldpc_iterations);
  tb.flags       = (current_config.new_data ? FEC_FLAG_NEW_DATA : 0u) | (current_config.use_early_stop ? FEC_FLAG_EARLY_STOP : 0u);
  if (asynchronous) {
    // The call returns once the work is on the GPU's queues; the engine's completion thread hands the block over and
    // notifies, as pusch_decoder_impl does from its executor (pusch_decoder_impl.cpp:372-381). When every host-call context
    // of the engine is in flight the block is decoded synchronously instead (back-pressure, never a drop).
    uint64_t  ticket = 0;
    const int rc = fec_tb_enqueue(engine->get(), &tb, 1, slots.data(), reinterpret_cast<const int8_t*>(softbits), softbits_count,
                                  transport_block.data(), transport_block.size(), &pusch_decoder_cuda::on_completion, this, &ticket);
    if (rc == FEC_OK) {
      return;
    }
    if (rc != FEC_ERR_QUEUE_FULL) {
      finish(rc, fec_tb_result{});
      return;
    }
  }
  fec_tb_result res = {};
  const int     rc  = fec_tb_run_host(engine->get(), &tb, 1, slots.data(), reinterpret_cast<const int8_t*>(softbits),
                                      softbits_count, transport_block.data(), transport_block.size(), &res, cb_results.data());
  have_cb_results   = (rc == FEC_OK);
  finish(rc, res);
}

void pusch_decoder_cuda::on_completion(void* user, uint64_t /*ticket*/, int status, const fec_tb_result* results, uint32_t n)
{
  pusch_decoder_cuda* self = static_cast<pusch_decoder_cuda*>(user);
  self->have_cb_results    = false;
  self->finish(status, (n != 0) ? results[0] : fec_tb_result{});
}

void pusch_decoder_cuda::finish(int rc, const fec_tb_result& res)
{
  pusch_decoder_result stats;
  stats.ldpc_decoder_stats.reset();
  if (rc != FEC_OK) {
    // A failure of the engine (a rejected configuration, a CUDA fault) is a failed transport block, not a dead DU
    // (pusch_decoder_hw_impl.cpp:319-332 treats a codeblock the accelerator could not decode the same way).
    stats.tb_crc_ok            = false;
    stats.nof_codeblocks_total = nof_cbs;
    for (unsigned i = 0; i != nof_cbs; ++i) {
      stats.ldpc_decoder_stats.update(current_config.nof_ldpc_iterations);
    }
    rm_buffer.unlock();
    notifier->on_sch_data(stats);
    return;
  }
  srsran_assert(res.nof_codeblocks == nof_cbs, "Wrong number of codeblocks {} (expected {}).", nof_cbs, res.nof_codeblocks);
  stats.tb_crc_ok            = res.tb_crc_ok != 0;
  stats.nof_codeblocks_total = res.nof_codeblocks;
  if (have_cb_results) {
    // Keep the pool's view of the codeblock CRCs in step with the engine's (the engine is authoritative).
    span<bool> cb_crcs = rm_buffer->get_codeblocks_crc();
    for (unsigned i = 0; i != nof_cbs; ++i) {
      const fec_cb_result& cr = cb_results[i];
      const bool           ok = (cr.status & FEC_RES_CRC_PASS) != 0;
      if (!(cr.status & FEC_RES_SKIPPED)) {
        stats.ldpc_decoder_stats.update(ok ? cr.nof_ldpc_iterations : current_config.nof_ldpc_iterations);
      }
      cb_crcs[i] = ok;
    }
  } else {
    // Asynchronous completion carries the block summary only: the same statistics from its sum / minimum / maximum (the
    // reference's sample statistics expose count, mean, min and max; one observation each of min and max and the rest at
    // the value that restores the sum reproduces all four).
    const unsigned nd = res.nof_cb_decoded;
    if (nd == 1) {
      stats.ldpc_decoder_stats.update(res.iter_sum);
    } else if (nd >= 2) {
      stats.ldpc_decoder_stats.update(res.iter_min);
      stats.ldpc_decoder_stats.update(res.iter_max);
      const unsigned rest = res.iter_sum - res.iter_min - res.iter_max;
      for (unsigned i = 2; i != nd; ++i) {
        const unsigned v = rest / (nd - 2) + ((i - 2) < rest % (nd - 2) ? 1 : 0);
        stats.ldpc_decoder_stats.update(v);
      }
    }
  }
  if (!stats.tb_crc_ok && res.tb_delivered) {
    rm_buffer->reset_codeblocks_crc();
  }
  if (stats.tb_crc_ok) {
    rm_buffer.release();
  } else {
    rm_buffer.unlock();
  }
  notifier->on_sch_data(stats);
}

std::optional<unsigned> ldpc_decoder_cuda::decode(bit_buffer&                      output,
                                                  span<const log_likelihood_ratio> input,
                                                  crc_calculator*                  crc,
                                                  const configuration&             cfg)
{
  const codeblock_metadata& m = cfg.block_conf;
  fec_cb_desc               d = {};
  d.E                         = static_cast<uint32_t>(input.size());
  d.nof_filler_bits           = m.cb_specific.nof_filler_bits;
  d.lifting_size              = static_cast<uint16_t>(m.tb_common.lifting_size);
  d.base_graph                = static_cast<uint8_t>(m.tb_common.base_graph);
  d.max_iter                  = static_cast<uint8_t>(cfg.algorithm_conf.max_iterations);
  d.mode                      = FEC_MODE_DECODE;
  d.crc_type                  = (crc != nullptr) ? to_fec_crc(crc->get_generator_poly()) : static_cast<uint8_t>(FEC_CRC16);
  d.flags                     = (crc != nullptr) ? FEC_FLAG_EARLY_STOP : FEC_FLAG_NO_CRC;
  fec_cb_result        res    = {};
  span<uint8_t>        bytes  = output.get_buffer();
  static_assert(sizeof(log_likelihood_ratio) == 1, "LLRs are int8 on both sides of the boundary.");
  const int rc = fec_run_host(engine->get(), &d, 1, reinterpret_cast<const int8_t*>(input.data()), input.size(),
                              bytes.data(), bytes.size(), &res, nullptr);
  if (rc != FEC_OK) {
    return std::nullopt; // an engine failure is a decoding failure (fec_last_error() has the reason)
  }
  if ((crc != nullptr) && (res.status & FEC_RES_CRC_PASS)) {
    return res.nof_ldpc_iterations;
  }
  return std::nullopt;
}

void ldpc_rate_dematcher_cuda::rate_dematch(span<log_likelihood_ratio>       output,
                                            span<const log_likelihood_ratio> input,
                                            bool                             new_data,
                                            const codeblock_metadata&        cfg)
{
  fec_cb_desc d     = {};
  d.harq_slot       = scratch_slot;
  d.E               = static_cast<uint32_t>(input.size());
  d.Nref            = cfg.tb_common.Nref;
  d.nof_filler_bits = cfg.cb_specific.nof_filler_bits;
  d.lifting_size    = static_cast<uint16_t>(cfg.tb_common.lifting_size);
  d.base_graph      = static_cast<uint8_t>(cfg.tb_common.base_graph);
  d.rv              = static_cast<uint8_t>(cfg.tb_common.rv);
  d.qm              = static_cast<uint8_t>(get_bits_per_symbol(cfg.tb_common.mod));
  d.crc_type        = FEC_CRC24B;
  d.max_iter        = 1;
  d.mode            = FEC_MODE_DEMATCH;
  d.flags           = new_data ? FEC_FLAG_NEW_DATA : 0u;
  fec_cb_result res = {};
  int8_t*       out = reinterpret_cast<int8_t*>(output.data());
  int           rc  = fec_harq_write(engine->get(), scratch_slot, out, static_cast<uint32_t>(output.size()));
  rc                = (rc == FEC_OK)
                          ? fec_run_host(engine->get(), &d, 1, reinterpret_cast<const int8_t*>(input.data()), input.size(), nullptr, 0,
                                         &res, nullptr)
                          : rc;
  rc                = (rc == FEC_OK) ? fec_harq_read(engine->get(), scratch_slot, out, static_cast<uint32_t>(output.size())) : rc;
  if (rc != FEC_OK) {
    // The interface has no failure channel: an engine failure leaves erasures (LLR 0 = no information), which the decoder
    // downstream turns into a failed codeblock; fec_last_error() has the reason.
    std::memset(out, 0, output.size());
  }
}

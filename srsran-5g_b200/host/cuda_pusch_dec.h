// srsran-5g_b200/host/cuda_pusch_dec.h -- reference-side binding of the B200 FEC engine.
//
// These classes derive from the reference's own interfaces (headers of the reference tree, found through the include
// path given at build time; nothing of the reference is copied here) and forward to the C ABI in include/fec_b200.h:
//
//   cuda_pusch_dec_accelerator : hal::hw_accelerator_pusch_dec
//        include/srsran/hal/phy/upper/channel_processors/pusch/hw_accelerator_pusch_dec.h:83-115 -- the object
//        pusch_decoder_hw_impl drives (lib/phy/upper/channel_processors/pusch/pusch_decoder_hw_impl.cpp:149-358)
//   cuda_pusch_dec_factory     : hal::hw_accelerator_pusch_dec_factory
//        include/srsran/hal/phy/upper/channel_processors/pusch/hw_accelerator_pusch_dec_factory.h:31-36 -- what
//        pusch_decoder_factory_hw_configuration::hw_decoder_factory expects (.../pusch/factories.h:74-83)
//   pusch_decoder_cuda         : pusch_decoder + pusch_decoder_buffer
//        include/srsran/phy/upper/channel_processors/pusch/pusch_decoder.h:54-101, pusch_decoder_buffer.h:36-63 -- whole
//        transport blocks in one engine call (segmentation, codeblock decoding, TB assembly and TB CRC on the device);
//        takes the place of pusch_decoder_impl / pusch_decoder_hw_impl (lib/phy/upper/channel_processors/pusch/)
//   ldpc_decoder_cuda          : ldpc_decoder        include/srsran/phy/upper/channel_coding/ldpc/ldpc_decoder.h:37-75
//   ldpc_rate_dematcher_cuda   : ldpc_rate_dematcher include/srsran/phy/upper/channel_coding/ldpc/ldpc_rate_dematcher.h:35-58
//
// One fec_engine (one GPU) is shared by every accelerator instance the factory creates; each instance owns one engine
// queue, is driven by one thread (the reference pins instances to threads), and never blocks: enqueue returns false when
// the staging ring is full, dequeue returns false until the operation has completed.
#pragma once

#include "fec_b200.h"
#include "srsran/hal/phy/upper/channel_processors/pusch/hw_accelerator_pusch_dec.h"
#include "srsran/hal/phy/upper/channel_processors/pusch/hw_accelerator_pusch_dec_factory.h"
#include "srsran/phy/upper/channel_coding/crc_calculator.h"
#include "srsran/phy/upper/channel_processors/pusch/pusch_decoder.h"
#include "srsran/phy/upper/channel_processors/pusch/pusch_decoder_buffer.h"
#include "srsran/phy/upper/channel_processors/pusch/pusch_decoder_notifier.h"
#include "srsran/phy/upper/channel_processors/pusch/pusch_decoder_result.h"
#include "srsran/phy/upper/unique_rx_buffer.h"
#include "srsran/phy/upper/channel_coding/ldpc/ldpc_decoder.h"
#include "srsran/phy/upper/channel_coding/ldpc/ldpc_rate_dematcher.h"
#include "srsran/ran/sch/sch_constants.h"
#include <array>
#include <memory>

namespace srsran {
namespace b200 {

/// Owns one engine (one GPU); shared by the accelerator instances and adapters created from it.
class fec_engine_handle
{
public:
  explicit fec_engine_handle(const fec_engine_cfg& cfg);
  ~fec_engine_handle();
  fec_engine_handle(const fec_engine_handle&)            = delete;
  fec_engine_handle& operator=(const fec_engine_handle&) = delete;
  fec_engine*        get() const { return engine; }

private:
  fec_engine* engine = nullptr;
};

/// hal::hw_accelerator_pusch_dec over the queue entry points of the C ABI. HARQ soft buffers are external (in HBM).
class cuda_pusch_dec_accelerator : public hal::hw_accelerator_pusch_dec
{
public:
  explicit cuda_pusch_dec_accelerator(std::shared_ptr<fec_engine_handle> engine_);
  ~cuda_pusch_dec_accelerator() override;

  // See hal::hw_accelerator for the documentation.
  bool enqueue_operation(span<const int8_t> data, span<const int8_t> aux_data = {}, unsigned cb_index = 0) override;
  bool dequeue_operation(span<uint8_t> data, span<int8_t> aux_data = {}, unsigned segment_index = 0) override;
  // See hal::hw_accelerator_pusch_dec for the documentation.
  void reserve_queue() override;
  void free_queue() override;
  void configure_operation(const hal::hw_pusch_decoder_configuration& config, unsigned cb_index = 0) override;
  void read_operation_outputs(hal::hw_pusch_decoder_outputs& out,
                              unsigned                       cb_index       = 0,
                              unsigned                       absolute_cb_id = 0) override;
  void free_harq_context_entry(unsigned absolute_cb_id) override;
  bool is_harq_external() const override { return true; }

private:
  struct operation {
    hal::hw_pusch_decoder_configuration cfg      = {};
    fec_cb_result                       result   = {};
    bool                                enqueued = false;
    bool                                done     = false;
    bool                                rejected = false; // the engine refused it: completes as a failed codeblock
  };

  void poll();

  std::shared_ptr<fec_engine_handle>       engine;
  uint32_t                                 queue     = 0;
  bool                                     has_queue = false;
  bool                                     dirty     = false; // operations enqueued since the last flush
  unsigned                                 idle_polls = 0;    // consecutive dequeue attempts that found nothing
  std::array<operation, MAX_NOF_SEGMENTS>  ops;
};

/// Factory handed to pusch_decoder_factory_hw_configuration::hw_decoder_factory.
class cuda_pusch_dec_factory : public hal::hw_accelerator_pusch_dec_factory
{
public:
  explicit cuda_pusch_dec_factory(const fec_engine_cfg& cfg) : engine(std::make_shared<fec_engine_handle>(cfg)) {}
  /// Accelerators over an engine that other adapters (transport-block decoder, encoder, demodulator) share.
  explicit cuda_pusch_dec_factory(std::shared_ptr<fec_engine_handle> engine_) : engine(std::move(engine_)) {}
  std::unique_ptr<hal::hw_accelerator_pusch_dec> create() override
  {
    return std::make_unique<cuda_pusch_dec_accelerator>(engine);
  }

private:
  std::shared_ptr<fec_engine_handle> engine;
};

/// The call that replaces create_bbdev_pusch_dec_acc_factory (lib/hal/phy/upper/channel_processors/pusch/
/// hw_accelerator_factories.cpp:28-36, a stub returning nullptr in the open-source tree).
std::shared_ptr<hal::hw_accelerator_pusch_dec_factory> create_cuda_pusch_dec_acc_factory(const fec_engine_cfg& cfg);

/// pusch_decoder over the transport-block entry point of the C ABI (fec_tb_run_host). The rx buffer only supplies the
/// absolute codeblock identifiers (the pool must be created with external_soft_bits = true): soft bits, codeblock CRC
/// flags and the decoded messages of earlier transmissions live in the engine. Decoding starts at on_end_softbits; calls
/// of different instances (threads) run concurrently on the GPU. Synchronous by default, asynchronous on request.
class pusch_decoder_cuda : public pusch_decoder, private pusch_decoder_buffer
{
public:
  /// \param max_nof_softbits capacity of the (pinned) codeword buffer, e.g. pusch_constants::get_max_codeword_size().
  /// \param asynchronous     on_end_softbits returns as soon as the work is queued on the GPU (fec_tb_enqueue) and the notifier
  ///                         is called from the engine's completion thread; the instance is busy until then.
  pusch_decoder_cuda(std::shared_ptr<fec_engine_handle> engine_, unsigned max_nof_softbits, bool asynchronous = false);
  ~pusch_decoder_cuda() override;

  // See pusch_decoder for the documentation.
  pusch_decoder_buffer& new_data(span<uint8_t>           transport_block,
                                 unique_rx_buffer        rm_buffer,
                                 pusch_decoder_notifier& notifier,
                                 const configuration&    cfg) override;
  void                  set_nof_softbits(units::bits /*nof_softbits*/) override {}

private:
  // See pusch_decoder_buffer for the documentation.
  span<log_likelihood_ratio> get_next_block_view(unsigned block_size) override;
  void                       on_new_softbits(span<const log_likelihood_ratio> softbits) override;
  void                       on_end_softbits() override;

  std::shared_ptr<fec_engine_handle> engine;
  log_likelihood_ratio*              softbits       = nullptr; // pinned
  unsigned                           softbits_cap   = 0;
  unsigned                           softbits_count = 0;
  span<uint8_t>                      transport_block;
  unique_rx_buffer                   rm_buffer;
  pusch_decoder_notifier*            notifier = nullptr;
  configuration                      current_config;
  std::array<uint32_t, MAX_NOF_SEGMENTS>      slots;
  std::array<fec_cb_result, MAX_NOF_SEGMENTS> cb_results;
  bool                                        asynchronous    = false;
  bool                                        have_cb_results = false;
  unsigned                                    nof_cbs         = 0;

  static void on_completion(void* user, uint64_t ticket, int status, const fec_tb_result* results, uint32_t n);
  void        finish(int rc, const fec_tb_result& res);
};

/// ldpc_decoder over FEC_MODE_DECODE (synchronous, one codeblock). The crc pointer only selects the polynomial.
class ldpc_decoder_cuda : public ldpc_decoder
{
public:
  explicit ldpc_decoder_cuda(std::shared_ptr<fec_engine_handle> engine_) : engine(std::move(engine_)) {}
  std::optional<unsigned> decode(bit_buffer&                      output,
                                 span<const log_likelihood_ratio> input,
                                 crc_calculator*                  crc,
                                 const configuration&             cfg) override;

private:
  std::shared_ptr<fec_engine_handle> engine;
};

/// ldpc_rate_dematcher over FEC_MODE_DEMATCH (synchronous, one codeblock, through a scratch HARQ slot).
class ldpc_rate_dematcher_cuda : public ldpc_rate_dematcher
{
public:
  ldpc_rate_dematcher_cuda(std::shared_ptr<fec_engine_handle> engine_, uint32_t scratch_slot_) :
    engine(std::move(engine_)), scratch_slot(scratch_slot_)
  {
  }
  void rate_dematch(span<log_likelihood_ratio>       output,
                    span<const log_likelihood_ratio> input,
                    bool                             new_data,
                    const codeblock_metadata&        cfg) override;

private:
  std::shared_ptr<fec_engine_handle> engine;
  uint32_t                           scratch_slot;
};

} // namespace b200
} // namespace srsran

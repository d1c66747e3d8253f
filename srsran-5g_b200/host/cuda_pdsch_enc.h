// srsran-5g_b200/host/cuda_pdsch_enc.h -- reference-side binding of the downlink twin of the B200 FEC engine.
//
//   cuda_pdsch_enc_accelerator : hal::hw_accelerator_pdsch_enc
//        include/srsran/hal/phy/upper/channel_processors/hw_accelerator_pdsch_enc.h:81-106 -- the object
//        pdsch_encoder_hw_impl drives (lib/phy/upper/channel_processors/pdsch/pdsch_encoder_hw_impl.cpp:31-186)
//   cuda_pdsch_enc_factory     : hal::hw_accelerator_pdsch_enc_factory
//        include/srsran/hal/phy/upper/channel_processors/hw_accelerator_pdsch_enc_factory.h:31-36 -- what
//        pdsch_encoder_factory_hw_configuration::hw_encoder_factory expects (.../pdsch/factories.h:57-64)
//
// CB mode (is_cb_mode_supported() = true): the reference's segmenter attaches the CRCs and hands over one packed segment per
// codeblock; the accelerator LDPC-encodes and rate-matches it. The consumer enqueues every codeblock of a transport block
// and then dequeues them in order; the adapter batches what is pending and encodes it in ONE launch at the first dequeue.
// The engine (one GPU) is shared with the PUSCH decoder accelerators.
#pragma once

#include "cuda_pusch_dec.h"
#include "fec_b200.h"
#include "srsran/hal/phy/upper/channel_processors/hw_accelerator_pdsch_enc.h"
#include "srsran/hal/phy/upper/channel_processors/hw_accelerator_pdsch_enc_factory.h"
#include <array>
#include <memory>
#include <vector>

namespace srsran {
namespace b200 {

class cuda_pdsch_enc_accelerator : public hal::hw_accelerator_pdsch_enc
{
public:
  explicit cuda_pdsch_enc_accelerator(std::shared_ptr<fec_engine_handle> engine_);
  ~cuda_pdsch_enc_accelerator() override;

  // See hal::hw_accelerator for the documentation.
  bool enqueue_operation(span<const uint8_t> data, span<const uint8_t> aux_data = {}, unsigned cb_index = 0) override;
  bool dequeue_operation(span<uint8_t> data, span<uint8_t> aux_data = {}, unsigned segment_index = 0) override;
  // See hal::hw_accelerator_pdsch_enc for the documentation.
  void     reserve_queue() override {}
  void     free_queue() override;
  void     configure_operation(const hal::hw_pdsch_encoder_configuration& config, unsigned cb_index = 0) override;
  bool     is_cb_mode_supported() const override { return true; }
  unsigned get_max_supported_buff_size() const override { return FEC_MAX_CB_IN_BYTES; }

private:
  enum class op_state { idle, configured, pending, done };
  struct operation {
    hal::hw_pdsch_encoder_configuration cfg   = {};
    op_state                            state = op_state::idle;
    uint32_t                            out_offset = 0;
  };

  void run_pending();

  std::shared_ptr<fec_engine_handle>      engine;
  std::array<operation, MAX_NOF_SEGMENTS> ops;
  std::vector<fec_enc_desc>               descs;     // pending batch
  std::vector<unsigned>                   desc_cb;   // codeblock index of each pending descriptor
  uint8_t*                                h_in  = nullptr; // pinned: packed segments of the pending batch
  uint8_t*                                h_out = nullptr; // pinned: packed rate-matched codeblocks
  size_t                                  in_used = 0, out_used = 0, in_cap = 0, out_cap = 0;
};

class cuda_pdsch_enc_factory : public hal::hw_accelerator_pdsch_enc_factory
{
public:
  explicit cuda_pdsch_enc_factory(std::shared_ptr<fec_engine_handle> engine_) : engine(std::move(engine_)) {}
  explicit cuda_pdsch_enc_factory(const fec_engine_cfg& cfg) : engine(std::make_shared<fec_engine_handle>(cfg)) {}
  std::unique_ptr<hal::hw_accelerator_pdsch_enc> create() override
  {
    return std::make_unique<cuda_pdsch_enc_accelerator>(engine);
  }

private:
  std::shared_ptr<fec_engine_handle> engine;
};

/// The call that replaces create_bbdev_pdsch_enc_acc_factory (lib/hal/phy/upper/channel_processors/hw_accelerator_factories.cpp,
/// a stub returning nullptr in the open-source tree).
std::shared_ptr<hal::hw_accelerator_pdsch_enc_factory> create_cuda_pdsch_enc_acc_factory(const fec_engine_cfg& cfg);

} // namespace b200
} // namespace srsran

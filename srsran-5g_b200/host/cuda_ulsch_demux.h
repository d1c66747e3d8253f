// srsran-5g_b200/host/cuda_ulsch_demux.h -- reference-side binding of the UL-SCH demultiplexer of the B200 FEC engine.
//
//   cuda_ulsch_demultiplex : ulsch_demultiplex
//        include/srsran/phy/upper/channel_processors/pusch/ulsch_demultiplex.h:44-101 -- the object pusch_processor_impl
//        places between the demodulator and the decoders (lib/phy/upper/channel_processors/pusch/pusch_processor_impl.cpp);
//        its pusch_codeword_buffer face is what pusch_demodulator_impl writes into
//        (include/srsran/phy/upper/channel_processors/pusch/pusch_codeword_buffer.h)
//   cuda_ulsch_demultiplex_factory : ulsch_demultiplex_factory
//        include/srsran/phy/upper/channel_processors/pusch/factories.h -- what create_pusch_processor_factory_sw takes
//
// The adapter collects the descrambled soft bits and the scrambling sequence of the codeword as the demodulator delivers
// them and sorts them on the device at on_end_codeword (fec_ulsch_demux_run_host): one host -> device -> host round trip per
// codeword, exact, not fast -- the efficient integration keeps the codeword in HBM (fec_demod_launch_device ->
// fec_ulsch_demux_launch_device -> the codeblock kernel) and only the three small UCI streams return.
// CSI Part 2, whose size the reference only learns by decoding CSI Part 1 (set_csi_part2 is called from inside the CSI Part 1
// buffer's on_end_softbits), takes a second pass with fec_ulsch_cfg.csi2_after_csi1 = 1.
#pragma once

#include "cuda_pusch_dec.h"
#include "fec_b200.h"
#include "srsran/phy/upper/channel_processors/pusch/factories.h"
#include "srsran/phy/upper/channel_processors/pusch/pusch_codeword_buffer.h"
#include "srsran/phy/upper/channel_processors/pusch/pusch_decoder_buffer.h"
#include "srsran/phy/upper/channel_processors/pusch/ulsch_demultiplex.h"
#include <memory>
#include <vector>

namespace srsran {
namespace b200 {

class cuda_ulsch_demultiplex : public ulsch_demultiplex, private pusch_codeword_buffer
{
public:
  explicit cuda_ulsch_demultiplex(std::shared_ptr<fec_engine_handle> engine_) : engine(std::move(engine_)) {}

  // See ulsch_demultiplex for the documentation.
  void set_csi_part2(pusch_decoder_buffer& csi_part2, unsigned nof_csi_part2_bits, unsigned nof_csi_part2_enc_bits) override;
  pusch_codeword_buffer& demultiplex(pusch_decoder_buffer& sch_data,
                                     pusch_decoder_buffer& harq_ack,
                                     pusch_decoder_buffer& csi_part1,
                                     const configuration&  config) override;

private:
  // See pusch_codeword_buffer for the documentation.
  span<log_likelihood_ratio> get_next_block_view(unsigned block_size) override;
  void on_new_block(span<const log_likelihood_ratio> data, const bit_buffer& scrambling_seq) override;
  void on_end_codeword() override;

  /// One pass over the collected codeword with the current configuration; fills the four stream vectors.
  void run(bool csi2_after_csi1);

  std::shared_ptr<fec_engine_handle> engine;
  fec_ulsch_cfg                      cfg          = {};
  pusch_decoder_buffer*              sch          = nullptr;
  pusch_decoder_buffer*              harq         = nullptr;
  pusch_decoder_buffer*              csi1         = nullptr;
  pusch_decoder_buffer*              csi2         = nullptr;
  bool                               csi2_at_start = false; ///< set_csi_part2 arrived before the first soft bit
  std::vector<int8_t>                codeword;              ///< descrambled soft bits, in arrival order
  std::vector<uint8_t>               seq;                   ///< packed scrambling sequence (MSB first)
  size_t                             nof_bits = 0;
  std::vector<log_likelihood_ratio>  view;                  ///< scratch handed out by get_next_block_view
  std::vector<int8_t>                out[4];
};

class cuda_ulsch_demultiplex_factory : public ulsch_demultiplex_factory
{
public:
  explicit cuda_ulsch_demultiplex_factory(std::shared_ptr<fec_engine_handle> engine_) : engine(std::move(engine_)) {}
  std::unique_ptr<ulsch_demultiplex> create() override { return std::make_unique<cuda_ulsch_demultiplex>(engine); }

private:
  std::shared_ptr<fec_engine_handle> engine;
};

} // namespace b200
} // namespace srsran

// oracle/ref_ulsch_shim.cpp -- C-ABI wrapper around the reference's UNMODIFIED ulsch_demultiplex_impl
// (lib/phy/upper/channel_processors/pusch/ulsch_demultiplex_impl.{h,cpp}), compiled where it lies by oracle/Makefile into
// oracle/_ref/libref_fec.so. TEST INFRASTRUCTURE: it pins oracle/fec_ulsch.c (the C restatement) and generates the committed
// vectors; nothing in the product loads it.
//
// The reference demultiplexer is driven exactly as pusch_demodulator_impl drives it: blocks of descrambled soft bits plus
// the scrambling sequence of each block through pusch_codeword_buffer::on_new_block, then on_end_codeword. The four
// destinations are plain collectors. set_csi_part2 is called right after demultiplex(), as the reference's own unit test does
// (ulsch_demultiplex_test.cpp:131-136), or -- csi2_after_csi1 != 0 -- from the CSI Part 1 collector's on_end_softbits, as
// pusch_processor_impl does once CSI Part 1 is decoded (pusch_processor_impl.cpp:73-101).
#include "phy/upper/channel_processors/pusch/ulsch_demultiplex_impl.h"
#include "srsran/phy/upper/channel_processors/pusch/pusch_decoder_buffer.h"
#include "srsran/srsvec/bit.h"
#include <cstdint>
#include <functional>
#include <vector>

using namespace srsran;

namespace {

class collector : public pusch_decoder_buffer
{
public:
  collector(int8_t* out_, unsigned cap_) : out(out_), cap(cap_) {}
  span<log_likelihood_ratio> get_next_block_view(unsigned block_size) override
  {
    tmp.resize(block_size);
    return tmp;
  }
  void on_new_softbits(span<const log_likelihood_ratio> sb) override
  {
    for (const log_likelihood_ratio& v : sb) {
      if (n < cap) {
        out[n] = static_cast<int8_t>(v.to_int());
      }
      ++n;
    }
  }
  void on_end_softbits() override
  {
    ended = true;
    if (at_end) {
      at_end();
    }
  }
  unsigned              n     = 0;
  bool                  ended = false;
  std::function<void()> at_end;

private:
  std::vector<log_likelihood_ratio> tmp;
  int8_t*                           out;
  unsigned                          cap;
};

} // namespace

extern "C" {

/// cfg[0..15] = {bits per symbol, nof_layers, nof_prb, start_symbol_index, nof_symbols, nof_harq_ack_rvd, dmrs type (1 / 2),
/// dmrs symbol mask (bit l = OFDM symbol l), nof_cdm_groups_without_data, nof_harq_ack_bits, nof_enc_harq_ack_bits,
/// nof_csi_part1_bits, nof_enc_csi_part1_bits, nof_csi_part2_bits, nof_enc_csi_part2_bits, csi2_after_csi1}.
/// codeword: n descrambled soft bits; scrambling: n unpacked bits; block: soft bits per on_new_block call.
/// counts[0..3] = soft bits delivered to {SCH, HARQ-ACK, CSI Part 1, CSI Part 2}; counts[4..7] = their on_end_softbits flags.
int ref_ulsch_demux(const uint32_t* cfg,
                    const int8_t*   codeword,
                    uint32_t        n,
                    const uint8_t*  scrambling,
                    uint32_t        block,
                    int8_t*         o_sch,
                    int8_t*         o_harq,
                    int8_t*         o_csi1,
                    int8_t*         o_csi2,
                    uint32_t        cap,
                    uint32_t*       counts)
{
  static const modulation_scheme mods[9] = {modulation_scheme::PI_2_BPSK,
                                            modulation_scheme::PI_2_BPSK,
                                            modulation_scheme::QPSK,
                                            modulation_scheme::QPSK,
                                            modulation_scheme::QAM16,
                                            modulation_scheme::QAM16,
                                            modulation_scheme::QAM64,
                                            modulation_scheme::QAM64,
                                            modulation_scheme::QAM256};
  ulsch_demultiplex::configuration c;
  c.modulation         = mods[cfg[0] > 8 ? 8 : cfg[0]];
  c.nof_layers         = cfg[1];
  c.nof_prb            = cfg[2];
  c.start_symbol_index = cfg[3];
  c.nof_symbols        = cfg[4];
  c.nof_harq_ack_rvd   = cfg[5];
  c.dmrs               = (cfg[6] == 1) ? dmrs_type::TYPE1 : dmrs_type::TYPE2;
  c.dmrs_symbol_mask   = symbol_slot_mask(14);
  for (unsigned l = 0; l != 14; ++l) {
    c.dmrs_symbol_mask.set(l, ((cfg[7] >> l) & 1u) != 0);
  }
  c.nof_cdm_groups_without_data = cfg[8];
  c.nof_harq_ack_bits           = cfg[9];
  c.nof_enc_harq_ack_bits       = cfg[10];
  c.nof_csi_part1_bits          = cfg[11];
  c.nof_enc_csi_part1_bits      = cfg[12];

  collector sch(o_sch, cap), harq(o_harq, cap), csi1(o_csi1, cap), csi2(o_csi2, cap);
  auto      demux = std::make_unique<ulsch_demultiplex_impl>();

  pusch_codeword_buffer& cw = demux->demultiplex(sch, harq, csi1, c);
  if (cfg[14] != 0) {
    if (cfg[15] == 0) {
      demux->set_csi_part2(csi2, cfg[13], cfg[14]);
    } else {
      csi1.at_end = [&]() { demux->set_csi_part2(csi2, cfg[13], cfg[14]); };
    }
  }

  std::vector<log_likelihood_ratio> data(block);
  for (uint32_t pos = 0; pos < n;) {
    const uint32_t len = (n - pos < block) ? (n - pos) : block;
    for (uint32_t i = 0; i != len; ++i) {
      data[i] = log_likelihood_ratio(static_cast<int>(codeword[pos + i]));
    }
    dynamic_bit_buffer seq(len);
    srsvec::bit_pack(seq, span<const uint8_t>(scrambling + pos, len));
    cw.on_new_block(span<const log_likelihood_ratio>(data.data(), len), seq);
    pos += len;
  }
  cw.on_end_codeword();

  counts[0] = sch.n;
  counts[1] = harq.n;
  counts[2] = csi1.n;
  counts[3] = csi2.n;
  counts[4] = sch.ended;
  counts[5] = harq.ended;
  counts[6] = csi1.ended;
  counts[7] = csi2.ended;
  return 0;
}

} // extern "C"

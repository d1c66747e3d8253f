// srsran-5g_b200/host/cuda_ulsch_demux.cpp -- see cuda_ulsch_demux.h.
#include "cuda_ulsch_demux.h"
#include "srsran/ran/sch/modulation_scheme.h"
#include "srsran/support/error_handling.h"

using namespace srsran;
using namespace srsran::b200;

pusch_codeword_buffer& cuda_ulsch_demultiplex::demultiplex(pusch_decoder_buffer& sch_data,
                                                           pusch_decoder_buffer& harq_ack,
                                                           pusch_decoder_buffer& csi_part1,
                                                           const configuration&  config)
{
  cfg                             = {};
  cfg.qm                          = static_cast<uint8_t>(get_bits_per_symbol(config.modulation));
  cfg.nof_layers                  = static_cast<uint8_t>(config.nof_layers);
  cfg.start_symbol_index          = static_cast<uint8_t>(config.start_symbol_index);
  cfg.nof_symbols                 = static_cast<uint8_t>(config.nof_symbols);
  cfg.dmrs_type                   = (config.dmrs == dmrs_type::TYPE1) ? 1 : 2;
  cfg.nof_cdm_groups_without_data = static_cast<uint8_t>(config.nof_cdm_groups_without_data);
  cfg.nof_prb                     = static_cast<uint16_t>(config.nof_prb);
  for (unsigned l = 0; l != config.dmrs_symbol_mask.size(); ++l) {
    if (config.dmrs_symbol_mask.test(l)) {
      cfg.dmrs_symbol_mask |= static_cast<uint16_t>(1u << l);
    }
  }
  cfg.nof_harq_ack_rvd       = config.nof_harq_ack_rvd;
  cfg.nof_harq_ack_bits      = config.nof_harq_ack_bits;
  cfg.nof_enc_harq_ack_bits  = config.nof_enc_harq_ack_bits;
  cfg.nof_csi_part1_bits     = config.nof_csi_part1_bits;
  cfg.nof_enc_csi_part1_bits = config.nof_enc_csi_part1_bits;

  sch           = &sch_data;
  harq          = (config.nof_harq_ack_bits != 0) ? &harq_ack : nullptr;
  csi1          = (config.nof_csi_part1_bits != 0) ? &csi_part1 : nullptr;
  csi2          = nullptr;
  csi2_at_start = false;
  nof_bits      = 0;
  codeword.clear();
  seq.clear();
  return *this;
}

void cuda_ulsch_demultiplex::set_csi_part2(pusch_decoder_buffer& csi_part2,
                                           unsigned              nof_csi_part2_bits,
                                           unsigned              nof_csi_part2_enc_bits)
{
  csi2                       = &csi_part2;
  cfg.nof_csi_part2_bits     = nof_csi_part2_bits;
  cfg.nof_enc_csi_part2_bits = nof_csi_part2_enc_bits;
  // Before the first soft bit: known from the start (the reference's unit test). Later: from inside the CSI Part 1 buffer's
  // on_end_softbits (pusch_processor_impl.cpp:73-101), i.e. during on_end_codeword here.
  csi2_at_start = (nof_bits == 0);
}

span<log_likelihood_ratio> cuda_ulsch_demultiplex::get_next_block_view(unsigned block_size)
{
  view.resize(block_size);
  return view;
}

void cuda_ulsch_demultiplex::on_new_block(span<const log_likelihood_ratio> data, const bit_buffer& scrambling_seq)
{
  srsran_assert(data.size() == scrambling_seq.size(), "Demodulated and descrambled data sizes must be equal.");
  const size_t n = data.size();
  codeword.resize(nof_bits + n);
  seq.resize((nof_bits + n + 7) / 8, 0);
  for (size_t i = 0; i != n; ++i) {
    codeword[nof_bits + i] = static_cast<int8_t>(data[i].to_int());
  }
  size_t i = 0;
  if (nof_bits % 8 == 0) { // whole bytes of the sequence at once
    for (; i + 8 <= n; i += 8) {
      seq[(nof_bits + i) / 8] = scrambling_seq.extract(static_cast<unsigned>(i), 8);
    }
  }
  for (; i != n; ++i) {
    if (scrambling_seq.extract(static_cast<unsigned>(i), 1)) {
      seq[(nof_bits + i) / 8] |= static_cast<uint8_t>(0x80u >> ((nof_bits + i) % 8));
    }
  }
  nof_bits += n;
}

void cuda_ulsch_demultiplex::run(bool csi2_after_csi1)
{
  cfg.csi2_after_csi1 = csi2_after_csi1 ? 1 : 0;
  fec_ulsch_sizes z   = {};
  int             rc  = fec_ulsch_demux_sizes(&cfg, &z);
  if (rc != FEC_OK || z.nof_codeword_bits != nof_bits) {
    report_fatal_error("UL-SCH demultiplexer: {} soft bits received, the configuration holds {} ({}).",
                       nof_bits,
                       z.nof_codeword_bits,
                       fec_last_error());
  }
  const uint32_t sizes[4] = {z.nof_sch_bits, z.nof_harq_ack_bits, z.nof_csi_part1_bits, z.nof_csi_part2_bits};
  for (int k = 0; k != 4; ++k) {
    out[k].assign(sizes[k] ? sizes[k] : 1, 0);
  }
  rc = fec_ulsch_demux_run_host(
      engine->get(), &cfg, codeword.data(), seq.data(), out[0].data(), out[1].data(), out[2].data(), out[3].data());
  if (rc != FEC_OK) {
    // A device fault must not take the DU down: every stream reads as erasures (all-zero soft bits), the decoders fail their
    // CRCs and the transmission is repeated.
    for (int k = 0; k != 4; ++k) {
      std::fill(out[k].begin(), out[k].end(), 0);
    }
  }
  for (int k = 0; k != 4; ++k) {
    out[k].resize(sizes[k]);
  }
}

static void deliver(pusch_decoder_buffer* dst, const std::vector<int8_t>& bits, bool end)
{
  if (dst == nullptr) {
    return;
  }
  if (!bits.empty()) {
    span<log_likelihood_ratio> v = dst->get_next_block_view(static_cast<unsigned>(bits.size()));
    for (size_t i = 0; i != bits.size(); ++i) {
      v[i] = log_likelihood_ratio(static_cast<int>(bits[i]));
    }
    dst->on_new_softbits(v);
  }
  if (end) {
    dst->on_end_softbits();
  }
}

void cuda_ulsch_demultiplex::on_end_codeword()
{
  // First pass: with CSI Part 2 if it was known from the start, without it otherwise.
  if (!csi2_at_start) {
    cfg.nof_csi_part2_bits = cfg.nof_enc_csi_part2_bits = 0;
  }
  run(false);
  deliver(harq, out[1], !out[1].empty());
  harq                              = nullptr;
  pusch_decoder_buffer* csi1_now    = csi1;
  csi1                              = nullptr;
  const std::vector<int8_t> sch_one = out[0], csi2_one = out[3];
  // Delivering the end of CSI Part 1 lets the caller decode it and call set_csi_part2 (which fills in cfg and csi2).
  const std::vector<int8_t> csi1_bits = out[2];
  deliver(csi1_now, csi1_bits, !csi1_bits.empty());
  if (!csi2_at_start && csi2 != nullptr && cfg.nof_enc_csi_part2_bits != 0) {
    run(true); // second pass: CSI Part 2 placed from the symbol in which CSI Part 1 ended
    deliver(csi2, out[3], !out[3].empty());
    deliver(sch, out[0], true);
  } else {
    deliver(csi2, csi2_one, !csi2_one.empty());
    deliver(sch, sch_one, true);
  }
  csi2 = nullptr;
  sch  = nullptr;
}

// tests/cpp/dropin_ulsch_demultiplex.cpp -- drop-in acceptance of srsran::b200::cuda_ulsch_demultiplex: the reference's own
// ulsch_demultiplex_impl (the unmodified source, linked from oracle/_ref/libref_fec.so) and the CUDA adapter are driven
// through the reference's interfaces only (ulsch_demultiplex::demultiplex / set_csi_part2, pusch_codeword_buffer::
// on_new_block / on_end_codeword, pusch_decoder_buffer collectors) on the same random configurations, with CSI Part 2
// announced either right after demultiplex() (ulsch_demultiplex_test.cpp:131-136) or from inside the CSI Part 1 buffer's
// on_end_softbits (pusch_processor_impl.cpp:73-101). Every stream must be identical. TEST INFRASTRUCTURE.
//   usage: dropin_ulsch_demultiplex [cases]
#include "cuda_ulsch_demux.h"
#include "phy/upper/channel_processors/pusch/ulsch_demultiplex_impl.h"
#include "srsran/srsvec/bit.h"
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <random>
#include <vector>

using namespace srsran;

namespace {

class collector : public pusch_decoder_buffer
{
public:
  span<log_likelihood_ratio> get_next_block_view(unsigned block_size) override
  {
    tmp.resize(block_size);
    return tmp;
  }
  void on_new_softbits(span<const log_likelihood_ratio> sb) override
  {
    for (const log_likelihood_ratio& v : sb) {
      bits.push_back(static_cast<int8_t>(v.to_int()));
    }
  }
  void on_end_softbits() override
  {
    ++ended;
    if (at_end) {
      at_end();
    }
  }
  std::vector<int8_t>   bits;
  int                   ended = 0;
  std::function<void()> at_end;

private:
  std::vector<log_likelihood_ratio> tmp;
};

struct outcome {
  collector sch, harq, csi1, csi2;
};

struct test_case {
  ulsch_demultiplex::configuration cfg;
  unsigned                         csi2_bits = 0, csi2_enc = 0;
  bool                             csi2_late = false;
  std::vector<int8_t>              codeword;
  std::vector<uint8_t>             seq;
  unsigned                         block = 4096;
};

void drive(ulsch_demultiplex& demux, const test_case& tc, outcome& o)
{
  pusch_codeword_buffer& cw = demux.demultiplex(o.sch, o.harq, o.csi1, tc.cfg);
  if (tc.csi2_enc != 0) {
    if (!tc.csi2_late) {
      demux.set_csi_part2(o.csi2, tc.csi2_bits, tc.csi2_enc);
    } else {
      o.csi1.at_end = [&demux, &o, &tc]() { demux.set_csi_part2(o.csi2, tc.csi2_bits, tc.csi2_enc); };
    }
  }
  std::vector<log_likelihood_ratio> data(tc.block);
  for (size_t pos = 0; pos < tc.codeword.size();) {
    const size_t len = std::min<size_t>(tc.block, tc.codeword.size() - pos);
    for (size_t i = 0; i != len; ++i) {
      data[i] = log_likelihood_ratio(static_cast<int>(tc.codeword[pos + i]));
    }
    dynamic_bit_buffer seq(len);
    srsvec::bit_pack(seq, span<const uint8_t>(tc.seq.data() + pos, len));
    cw.on_new_block(span<const log_likelihood_ratio>(data.data(), len), seq);
    pos += len;
  }
  cw.on_end_codeword();
}

test_case make_case(std::mt19937& rng, int kind)
{
  auto       U = [&rng](int lo, int hi) { return std::uniform_int_distribution<int>(lo, hi)(rng); };
  test_case  tc;
  const int  qms[5]  = {1, 2, 4, 6, 8};
  const modulation_scheme mods[9] = {modulation_scheme::PI_2_BPSK, modulation_scheme::PI_2_BPSK, modulation_scheme::QPSK,
                                     modulation_scheme::QPSK,      modulation_scheme::QAM16,     modulation_scheme::QAM16,
                                     modulation_scheme::QAM64,     modulation_scheme::QAM64,     modulation_scheme::QAM256};
  const int  qm      = qms[U(0, 4)];
  auto&      c       = tc.cfg;
  c.modulation       = mods[qm];
  c.nof_layers       = U(1, 4);
  c.nof_prb          = (kind % 7 == 0) ? U(100, 273) : U(1, 40);
  c.start_symbol_index = U(0, 2);
  c.nof_symbols      = U(6, 14 - c.start_symbol_index);
  c.dmrs             = U(0, 1) ? dmrs_type::TYPE1 : dmrs_type::TYPE2;
  c.nof_cdm_groups_without_data = U(1, c.dmrs == dmrs_type::TYPE1 ? 2 : 3);
  const unsigned masks[5] = {0x0004, 0x0804, 0x0884, 0x0924, 0x000c};
  unsigned       mask     = masks[U(0, 4)];
  bool           any      = false;
  for (unsigned l = c.start_symbol_index; l != c.start_symbol_index + c.nof_symbols; ++l) {
    any = any || ((mask >> l) & 1u);
  }
  if (!any) {
    mask = 1u << c.start_symbol_index;
  }
  c.dmrs_symbol_mask = symbol_slot_mask(14);
  unsigned re_uci = 0, re_tot = 0;
  const unsigned per_prb_dmrs = 12 - c.nof_cdm_groups_without_data * (c.dmrs == dmrs_type::TYPE1 ? 6 : 4);
  for (unsigned l = 0; l != 14; ++l) {
    c.dmrs_symbol_mask.set(l, (mask >> l) & 1u);
    if (l >= c.start_symbol_index && l < c.start_symbol_index + c.nof_symbols) {
      re_uci += ((mask >> l) & 1u) ? 0 : 12 * c.nof_prb;
      re_tot += (((mask >> l) & 1u) ? per_prb_dmrs : 12) * c.nof_prb;
    }
  }
  const unsigned B      = qm * c.nof_layers;
  const int      budget = std::max(1u, re_uci / 9);
  auto           enc    = [&]() { return static_cast<unsigned>(U(1, std::max(2, budget))) * B; };
  c.nof_harq_ack_rvd = c.nof_harq_ack_bits = c.nof_enc_harq_ack_bits = c.nof_csi_part1_bits = c.nof_enc_csi_part1_bits = 0;
  const int k = kind % 6;
  if (k == 1 || k == 2 || k == 4 || k == 5) {
    const int acks[6]       = {1, 2, 3, 5, 11, 20};
    c.nof_harq_ack_bits     = acks[U(0, 5)];
    c.nof_enc_harq_ack_bits = enc();
    if (c.nof_harq_ack_bits <= 2) {
      c.nof_harq_ack_rvd = c.nof_enc_harq_ack_bits + U(0, 3) * B;
    }
  }
  if (k >= 2) {
    const int n1[5]          = {1, 2, 4, 7, 20};
    c.nof_csi_part1_bits     = n1[U(0, 4)];
    c.nof_enc_csi_part1_bits = enc();
  }
  if (k >= 4) {
    const int n2[5] = {1, 2, 3, 9, 30};
    tc.csi2_bits    = n2[U(0, 4)];
    tc.csi2_enc     = enc();
    tc.csi2_late    = (k == 5);
  }
  tc.codeword.resize(static_cast<size_t>(re_tot) * B);
  tc.seq.resize(tc.codeword.size());
  for (size_t i = 0; i != tc.codeword.size(); ++i) {
    tc.codeword[i] = static_cast<int8_t>(U(-120, 120));
    tc.seq[i]      = static_cast<uint8_t>(U(0, 1));
  }
  const unsigned blocks[4] = {64, 777, 4096, 1u << 20};
  tc.block                 = blocks[U(0, 3)];
  return tc;
}

bool same(const collector& a, const collector& b, const char* what, int i)
{
  if (a.bits != b.bits || a.ended != b.ended) {
    std::printf("MISMATCH case %d stream %s: %zu vs %zu soft bits, ended %d vs %d\n", i, what, a.bits.size(), b.bits.size(),
                a.ended, b.ended);
    return false;
  }
  return true;
}

} // namespace

int main(int argc, char** argv)
{
  const int      ncases = (argc > 1) ? std::atoi(argv[1]) : 120;
  fec_engine_cfg ecfg;
  fec_engine_default_cfg(&ecfg);
  ecfg.harq_slots    = 16;
  ecfg.max_batch_cbs = 16;
  auto engine = std::make_shared<b200::fec_engine_handle>(ecfg);
  b200::cuda_ulsch_demultiplex_factory factory(engine);
  std::mt19937                         rng(38212);
  size_t uci_bits = 0, late = 0;
  for (int i = 0; i != ncases; ++i) {
    const test_case tc = make_case(rng, i);
    outcome         ref_out, cuda_out;
    ulsch_demultiplex_impl ref;
    drive(ref, tc, ref_out);
    std::unique_ptr<ulsch_demultiplex> dut = factory.create();
    drive(*dut, tc, cuda_out);
    if (!same(ref_out.sch, cuda_out.sch, "sch", i) || !same(ref_out.harq, cuda_out.harq, "harq_ack", i) ||
        !same(ref_out.csi1, cuda_out.csi1, "csi_part1", i) || !same(ref_out.csi2, cuda_out.csi2, "csi_part2", i)) {
      return 1;
    }
    uci_bits += ref_out.harq.bits.size() + ref_out.csi1.bits.size() + ref_out.csi2.bits.size();
    late += (tc.csi2_late && !ref_out.csi2.bits.empty()) ? 1 : 0;
  }
  std::printf("OK cases=%d uci_soft_bits=%zu csi2_after_csi1=%zu\n", ncases, uci_bits, late);
  return 0;
}

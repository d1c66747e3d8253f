// tests/cpp/dropin_pdsch_encoder.cpp -- TEST INFRASTRUCTURE: drop-in acceptance of the downlink adapter.
//
// Two encoders run side by side on the same transport blocks:
//   (1) the reference's own software encoder, pdsch_encoder_impl (segmenter + ldpc_encoder ("auto": the widest SIMD the host has) + ldpc_rate_matcher from
//       oracle/_ref, the class itself compiled for this test from the source where it lies) -- the expected codeword;
//   (2) the reference's UNMODIFIED hardware-accelerated consumer, pdsch_encoder_hw_impl, driving
//       srsran::b200::cuda_pdsch_enc_accelerator through the HAL (hal::hw_accelerator_pdsch_enc, CB mode).
// Both are called through pdsch_encoder::encode only. Exit code 0 = every codeword identical.
#include "cuda_pdsch_enc.h"
#include "phy/upper/channel_processors/pdsch/pdsch_encoder_hw_impl.h"
#include "phy/upper/channel_processors/pdsch/pdsch_encoder_impl.h"
#include "srsran/phy/upper/channel_coding/channel_coding_factories.h"
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

using namespace srsran;

namespace {

modulation_scheme mod_of(unsigned qm)
{
  switch (qm) {
    case 1:
      return modulation_scheme::PI_2_BPSK;
    case 2:
      return modulation_scheme::QPSK;
    case 4:
      return modulation_scheme::QAM16;
    case 6:
      return modulation_scheme::QAM64;
    default:
      return modulation_scheme::QAM256;
  }
}

struct tb_case {
  unsigned tbs, bg, qm, nl, nsym, rv, Nref;
};

} // namespace

int main(int argc, char** argv)
{
  const unsigned n_random = (argc > 1) ? std::atoi(argv[1]) : 40;

  auto sw = std::make_unique<pdsch_encoder_impl>(create_ldpc_segmenter_tx_factory_sw(
                                                     {create_crc_calculator_factory_sw("auto")})
                                                     ->create(),
                                                 create_ldpc_encoder_factory_sw("auto")->create(),
                                                 create_ldpc_rate_matcher_factory_sw()->create());

  fec_engine_cfg cfg;
  fec_engine_default_cfg(&cfg);
  cfg.harq_slots    = 16;
  cfg.max_batch_cbs = 256;
  auto factory      = b200::create_cuda_pdsch_enc_acc_factory(cfg);
  auto crcf         = create_crc_calculator_factory_sw("auto");
  pdsch_encoder_hw_impl::sch_crc crcs;
  crcs.crc16  = crcf->create(crc_generator_poly::CRC16);
  crcs.crc24A = crcf->create(crc_generator_poly::CRC24A);
  crcs.crc24B = crcf->create(crc_generator_poly::CRC24B);
  auto hw = std::make_unique<pdsch_encoder_hw_impl>(
      crcs, create_ldpc_segmenter_tx_factory_sw({create_crc_calculator_factory_sw("auto")})->create(), factory->create());

  std::mt19937          rng(2026);
  std::vector<tb_case>  cases;
  cases.push_back({1277992, 1, 8, 4, 1362816 / 8, 0, 0}); // BASELINE config 3: 152 codeblocks
  cases.push_back({1277992, 1, 8, 4, 1362816 / 8, 2, 0});
  for (unsigned i = 0; i != n_random; ++i) {
    tb_case c;
    c.bg  = 1 + rng() % 2;
    c.qm  = std::array<unsigned, 5>{1, 2, 4, 6, 8}[rng() % 5];
    c.nl  = 1 + rng() % 4;
    c.tbs = 8 * (3 + rng() % ((c.bg == 1 ? 60000 : 9000) / 8));
    const double rate = 0.2 + 0.7 * (rng() % 1000) / 1000.0;
    c.nsym            = std::max(1u, static_cast<unsigned>((c.tbs + 24) / rate / (c.qm * c.nl))) * c.nl;
    c.rv              = rng() % 4;
    c.Nref            = (rng() % 3 == 0) ? 20000 + rng() % 6000 : 0;
    cases.push_back(c);
  }

  unsigned checked = 0, total_cbs = 0;
  double   us_sw = 0, us_hw = 0;
  for (const tb_case& c : cases) {
    std::vector<uint8_t> tb(c.tbs / 8);
    for (uint8_t& b : tb) {
      b = rng() & 0xff;
    }
    pdsch_encoder::configuration ecfg;
    ecfg.base_graph     = (c.bg == 1) ? ldpc_base_graph_type::BG1 : ldpc_base_graph_type::BG2;
    ecfg.rv             = c.rv;
    ecfg.mod            = mod_of(c.qm);
    ecfg.Nref           = c.Nref;
    ecfg.nof_layers     = c.nl;
    ecfg.nof_ch_symbols = c.nsym;
    std::vector<uint8_t> cw_sw(c.nsym * c.qm, 0xee), cw_hw(c.nsym * c.qm, 0xdd);
    auto t0 = std::chrono::steady_clock::now();
    sw->encode(cw_sw, tb, ecfg);
    auto t1 = std::chrono::steady_clock::now();
    hw->encode(cw_hw, tb, ecfg);
    auto t2 = std::chrono::steady_clock::now();
    if (checked < 2) {
      us_sw += std::chrono::duration<double, std::micro>(t1 - t0).count();
      us_hw += std::chrono::duration<double, std::micro>(t2 - t1).count();
    }
    if (cw_sw != cw_hw) {
      size_t first = 0;
      while (cw_sw[first] == cw_hw[first]) {
        ++first;
      }
      std::printf("MISMATCH tbs=%u bg=%u qm=%u nl=%u nsym=%u rv=%u Nref=%u first differing bit %zu\n", c.tbs, c.bg, c.qm, c.nl,
                  c.nsym, c.rv, c.Nref, first);
      return 1;
    }
    ++checked;
  }
  {
    // Latency of the config-3 block through pdsch_encoder::encode, p50 of 30 calls each.
    const tb_case&       c = cases[0];
    std::vector<uint8_t> tb(c.tbs / 8, 0x5a), cw(c.nsym * c.qm);
    pdsch_encoder::configuration ecfg;
    ecfg.base_graph     = ldpc_base_graph_type::BG1;
    ecfg.rv             = 0;
    ecfg.mod            = mod_of(c.qm);
    ecfg.Nref           = 0;
    ecfg.nof_layers     = c.nl;
    ecfg.nof_ch_symbols = c.nsym;
    std::vector<double> a, b;
    for (unsigned i = 0; i != 30; ++i) {
      auto t0 = std::chrono::steady_clock::now();
      sw->encode(cw, tb, ecfg);
      auto t1 = std::chrono::steady_clock::now();
      hw->encode(cw, tb, ecfg);
      auto t2 = std::chrono::steady_clock::now();
      a.push_back(std::chrono::duration<double, std::micro>(t1 - t0).count());
      b.push_back(std::chrono::duration<double, std::micro>(t2 - t1).count());
    }
    std::sort(a.begin(), a.end());
    std::sort(b.begin(), b.end());
    std::printf("LATENCY config-3 transport block (152 codeblocks, p50 of 30): reference software pdsch_encoder_impl %.0f us, "
                "pdsch_encoder_hw_impl over the CUDA HAL adapter %.0f us\n", a[15], b[15]);
  }
  (void)us_sw;
  (void)us_hw;
  std::printf("OK transport_blocks=%u\n", checked);
  (void)total_cbs;
  return 0;
}

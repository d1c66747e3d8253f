// tests/cpp/dropin_pusch_decoder.cpp -- TEST INFRASTRUCTURE: drop-in acceptance of the reference-side adapters.
//
// Three decoders run side by side on the same HARQ transmissions:
//   (1) the reference's own software decoder, pusch_decoder_impl (from oracle/_ref) -- the expected results;
//   (2) srsran::b200::pusch_decoder_cuda -- whole transport blocks through fec_tb_run_host;
//   (3) the reference's UNMODIFIED hardware-accelerated consumer, pusch_decoder_hw_impl (compiled for this test from the
//       source where it lies), driving srsran::b200::cuda_pusch_dec_accelerator instances through the HAL
//       (hal::hw_accelerator_pusch_dec: reserve_queue / configure / enqueue / dequeue / read outputs / free HARQ entry).
// Everything goes through the reference's interfaces only -- rx_buffer_pool::reserve, pusch_decoder::new_data, pusch_decoder_buffer::
// on_new_softbits / on_end_softbits, pusch_decoder_notifier::on_sch_data -- over the HARQ transmissions found in the
// input file (written by tests/test_gpu_dropin.py), and compares transport block bytes, tb_crc_ok, the number of
// codeblocks and the LDPC iteration statistics of every transmission. Exit code 0 = identical.
#include "cuda_pusch_dec.h"
#include "phy/upper/channel_processors/pusch/pusch_decoder_hw_impl.h"
#include "phy/upper/channel_processors/pusch/pusch_decoder_impl.h"
#include "srsran/phy/upper/channel_coding/channel_coding_factories.h"
#include "srsran/phy/upper/rx_buffer_pool.h"
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

using namespace srsran;

namespace {

struct capture : public pusch_decoder_notifier {
  pusch_decoder_result r;
  bool                 got = false;
  void                 on_sch_data(const pusch_decoder_result& result) override
  {
    r   = result;
    got = true;
  }
};

std::unique_ptr<pusch_decoder> make_reference_decoder()
{
  auto crcf      = create_crc_calculator_factory_sw("auto");
  auto make_crcs = [&crcf](auto& c) {
    c.crc16  = crcf->create(crc_generator_poly::CRC16);
    c.crc24A = crcf->create(crc_generator_poly::CRC24A);
    c.crc24B = crcf->create(crc_generator_poly::CRC24B);
  };
  pusch_codeblock_decoder::sch_crc cb_crcs;
  make_crcs(cb_crcs);
  std::vector<std::unique_ptr<pusch_codeblock_decoder>> cbd;
  cbd.push_back(std::make_unique<pusch_codeblock_decoder>(create_ldpc_rate_dematcher_factory_sw("generic")->create(),
                                                          create_ldpc_decoder_factory_sw("generic")->create(), cb_crcs));
  auto                        pool = std::make_shared<pusch_decoder_impl::codeblock_decoder_pool>(std::move(cbd));
  pusch_decoder_impl::sch_crc crcs;
  make_crcs(crcs);
  return std::make_unique<pusch_decoder_impl>(create_ldpc_segmenter_rx_factory_sw()->create(), pool, std::move(crcs), nullptr,
                                              MAX_RB, 4);
}

std::unique_ptr<pusch_decoder> make_hal_decoder(std::shared_ptr<hal::hw_accelerator_pusch_dec_factory> factory)
{
  auto                           crcf = create_crc_calculator_factory_sw("auto");
  pusch_decoder_hw_impl::sch_crc crcs;
  crcs.crc16  = crcf->create(crc_generator_poly::CRC16);
  crcs.crc24A = crcf->create(crc_generator_poly::CRC24A);
  crcs.crc24B = crcf->create(crc_generator_poly::CRC24B);
  std::vector<std::unique_ptr<hal::hw_accelerator_pusch_dec>> accs;
  accs.push_back(factory->create());
  auto pool = std::make_shared<pusch_decoder_hw_impl::hw_decoder_pool>(std::move(accs));
  return std::make_unique<pusch_decoder_hw_impl>(create_ldpc_segmenter_rx_factory_sw()->create(), crcs, pool, nullptr);
}

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

} // namespace

int main(int argc, char** argv)
{
  if (argc < 2) {
    std::fprintf(stderr, "usage: %s <input file>\n", argv[0]);
    return 2;
  }
  FILE* f = std::fopen(argv[1], "rb");
  if (!f) {
    return 2;
  }
  constexpr unsigned POOL_CBS = 512;
  fec_engine_cfg     ecfg;
  fec_engine_default_cfg(&ecfg);
  ecfg.harq_slots    = POOL_CBS;
  ecfg.max_batch_cbs = 256;
  auto engine        = std::make_shared<b200::fec_engine_handle>(ecfg);

  rx_buffer_pool_config pc;
  pc.max_codeblock_size   = ldpc::MAX_CODEBLOCK_SIZE;
  pc.nof_buffers          = 8;
  pc.nof_codeblocks       = POOL_CBS;
  pc.expire_timeout_slots = 100000;
  pc.external_soft_bits   = false;
  auto pool_ref           = create_rx_buffer_pool(pc);
  pc.external_soft_bits   = true; // soft bits stay in HBM
  auto pool_cuda          = create_rx_buffer_pool(pc);
  auto pool_hal           = create_rx_buffer_pool(pc);

  std::unique_ptr<pusch_decoder> dec_ref  = make_reference_decoder();
  std::unique_ptr<pusch_decoder> dec_cuda = std::make_unique<b200::pusch_decoder_cuda>(engine, 4u * 1024 * 1024);
  ecfg.max_batch_cbs                      = 162; // one transport block per launch
  std::unique_ptr<pusch_decoder> dec_hal  = make_hal_decoder(b200::create_cuda_pusch_dec_acc_factory(ecfg)); // own engine

  uint32_t nof_blocks = 0;
  if (std::fread(&nof_blocks, 4, 1, f) != 1) {
    return 2;
  }
  const unsigned nof_timed = (argc > 2) ? std::atoi(argv[2]) : 0; // latency mode: repeat block 0 / transmission 0
  unsigned slot_count = 0, nof_tx_total = 0, nof_bad_tx = 0;
  for (uint32_t b = 0; b != nof_blocks; ++b) {
    uint32_t h[6]; // tbs, bg, qm, nof_layers, G, nof_tx
    if (std::fread(h, 4, 6, f) != 6) {
      return 2;
    }
    const unsigned        tbs = h[0], G = h[4];
    ldpc_base_graph_type  bg  = (h[1] == 1) ? ldpc_base_graph_type::BG1 : ldpc_base_graph_type::BG2;
    const unsigned        ncb = ldpc::compute_nof_codeblocks(units::bits(tbs), bg);
    trx_buffer_identifier id(static_cast<uint16_t>(b + 1), 0);
    std::vector<int8_t>   llr(G);
    bool                  done = false;
    for (uint32_t t = 0; t != h[5]; ++t) {
      uint32_t rv = 0;
      if (std::fread(&rv, 4, 1, f) != 1 || std::fread(llr.data(), 1, G, f) != G) {
        return 2;
      }
      if (done) {
        continue; // both decoders released the buffer after the good transmission
      }
      pusch_decoder::configuration cfg;
      cfg.base_graph          = bg;
      cfg.rv                  = rv;
      cfg.mod                 = mod_of(h[2]);
      cfg.Nref                = 0;
      cfg.nof_layers          = h[3];
      cfg.nof_ldpc_iterations = 6;
      cfg.use_early_stop      = true;
      cfg.new_data            = (t == 0);
      std::vector<uint8_t> tb_ref(tbs / 8, 0xAA), tb_cuda(tbs / 8, 0xAA), tb_hal(tbs / 8, 0xAA);
      capture              n_ref, n_cuda, n_hal;
      span<const log_likelihood_ratio> soft(reinterpret_cast<const log_likelihood_ratio*>(llr.data()), G);
      slot_point slot(0, slot_count++ % 10240);
      {
        unique_rx_buffer buf = pool_ref->get_pool().reserve(slot, id, ncb, cfg.new_data);
        if (!buf.is_valid()) {
          std::fprintf(stderr, "reference pool: reservation failed\n");
          return 3;
        }
        pusch_decoder_buffer& in = dec_ref->new_data(tb_ref, std::move(buf), n_ref, cfg);
        in.on_new_softbits(soft);
        in.on_end_softbits();
      }
      {
        unique_rx_buffer buf = pool_cuda->get_pool().reserve(slot, id, ncb, cfg.new_data);
        if (!buf.is_valid()) {
          std::fprintf(stderr, "cuda pool: reservation failed\n");
          return 3;
        }
        pusch_decoder_buffer& in = dec_cuda->new_data(tb_cuda, std::move(buf), n_cuda, cfg);
        // in two pieces, the second through the zero-copy view, as the UL-SCH demultiplexer may do
        in.on_new_softbits(soft.first(G / 2));
        span<log_likelihood_ratio> view = in.get_next_block_view(G - G / 2);
        std::memcpy(view.data(), soft.data() + G / 2, G - G / 2);
        in.on_new_softbits(view);
        in.on_end_softbits();
      }
      {
        unique_rx_buffer buf = pool_hal->get_pool().reserve(slot, id, ncb, cfg.new_data);
        if (!buf.is_valid()) {
          std::fprintf(stderr, "hal pool: reservation failed\n");
          return 3;
        }
        pusch_decoder_buffer& in = dec_hal->new_data(tb_hal, std::move(buf), n_hal, cfg);
        in.on_new_softbits(soft);
        in.on_end_softbits();
      }
      ++nof_tx_total;
      const bool hal_same = n_hal.got && n_ref.r.tb_crc_ok == n_hal.r.tb_crc_ok &&
                            n_ref.r.nof_codeblocks_total == n_hal.r.nof_codeblocks_total && tb_ref == tb_hal &&
                            ((t != 0) || (n_ref.r.ldpc_decoder_stats.get_min() == n_hal.r.ldpc_decoder_stats.get_min() &&
                                          n_ref.r.ldpc_decoder_stats.get_max() == n_hal.r.ldpc_decoder_stats.get_max() &&
                                          n_ref.r.ldpc_decoder_stats.get_mean() == n_hal.r.ldpc_decoder_stats.get_mean()));
      if (!hal_same) {
        std::fprintf(stderr, "MISMATCH (HAL path) block %u tx %u: crc %d/%d cbs %u/%u bytes %s\n", b, t, n_ref.r.tb_crc_ok,
                     n_hal.r.tb_crc_ok, n_ref.r.nof_codeblocks_total, n_hal.r.nof_codeblocks_total,
                     tb_ref == tb_hal ? "same" : "differ");
        return 1;
      }
      const bool same = n_ref.got && n_cuda.got && n_ref.r.tb_crc_ok == n_cuda.r.tb_crc_ok &&
                        n_ref.r.nof_codeblocks_total == n_cuda.r.nof_codeblocks_total && tb_ref == tb_cuda;
      // iteration statistics are comparable when every codeblock was decoded in this transmission
      const bool stats_same = (t != 0) || (n_ref.r.ldpc_decoder_stats.get_nof_observations() ==
                                               n_cuda.r.ldpc_decoder_stats.get_nof_observations() &&
                                           n_ref.r.ldpc_decoder_stats.get_min() == n_cuda.r.ldpc_decoder_stats.get_min() &&
                                           n_ref.r.ldpc_decoder_stats.get_max() == n_cuda.r.ldpc_decoder_stats.get_max() &&
                                           n_ref.r.ldpc_decoder_stats.get_mean() == n_cuda.r.ldpc_decoder_stats.get_mean());
      if (!same || !stats_same) {
        std::fprintf(stderr, "MISMATCH block %u tx %u: crc %d/%d cbs %u/%u bytes %s stats %s\n", b, t, n_ref.r.tb_crc_ok,
                     n_cuda.r.tb_crc_ok, n_ref.r.nof_codeblocks_total, n_cuda.r.nof_codeblocks_total,
                     tb_ref == tb_cuda ? "same" : "differ", stats_same ? "same" : "differ");
        return 1;
      }
      nof_bad_tx += n_ref.r.tb_crc_ok ? 0 : 1;
      done = n_ref.r.tb_crc_ok;
      if (nof_timed != 0 && b == 0 && t == 0) {
        // Latency of one transport block (new data every time) through the reference's interfaces, call to notification.
        auto timed = [&](pusch_decoder& dec, rx_buffer_pool_controller& pool, std::vector<uint8_t>& tb, const char* name,
                         unsigned reps) {
          std::vector<double> us;
          for (unsigned i = 0; i != reps; ++i) {
            capture          n;
            slot_point       sl(0, slot_count++ % 10240);
            unique_rx_buffer buf = pool.get_pool().reserve(sl, id, ncb, true);
            if (!buf.is_valid()) {
              std::fprintf(stderr, "%s: reservation failed in the timed loop\n", name);
              return;
            }
            // The UL-SCH demultiplexer writes straight into the decoder's buffer (get_next_block_view); the clock starts
            // when the soft bits are there.
            pusch_decoder_buffer&      in   = dec.new_data(tb, std::move(buf), n, cfg);
            span<log_likelihood_ratio> view = in.get_next_block_view(G);
            std::memcpy(view.data(), soft.data(), G);
            auto t0 = std::chrono::steady_clock::now();
            in.on_new_softbits(view);
            in.on_end_softbits();
            us.push_back(std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count());
            if (!n.got || !n.r.tb_crc_ok) {
              std::fprintf(stderr, "%s: timed transmission failed\n", name);
              return;
            }
          }
          std::sort(us.begin(), us.end());
          std::printf("LATENCY %s tbs=%u codeblocks=%u reps=%u p50_us=%.1f p99_us=%.1f max_us=%.1f\n", name, tbs, ncb, reps,
                      us[us.size() / 2], us[std::min<size_t>(us.size() - 1, us.size() * 99 / 100)], us.back());
        };
        timed(*dec_cuda, *pool_cuda, tb_cuda, "pusch_decoder_cuda", nof_timed);
        timed(*dec_hal, *pool_hal, tb_hal, "pusch_decoder_hw_impl+cuda_hal", nof_timed);
        timed(*dec_ref, *pool_ref, tb_ref, "pusch_decoder_impl(cpu,1thread)", std::max(3u, nof_timed / 20));
      }
    }
  }
  std::fclose(f);
  std::printf("OK blocks=%u transmissions=%u failed_transmissions=%u\n", nof_blocks, nof_tx_total, nof_bad_tx);
  return 0;
}

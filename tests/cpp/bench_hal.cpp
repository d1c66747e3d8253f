// tests/cpp/bench_hal.cpp -- MEASUREMENT HARNESS: throughput and LOADED latency through the boundary the reference consumes.
//
// T threads, each with its own decoder instance(s), decode BASELINE config-3 transport blocks (273 PRB, 4 layers, 256QAM:
// 152 codeblocks, 1 277 992 bits) back to back for a fixed time, the shape of the reference's own accelerator benchmark
// (tests/benchmarks/phy/upper/channel_processors/pusch/pusch_decoder_hwacc_benchmark.cpp:385-506). Three consumers:
//   hal       the reference's UNMODIFIED pusch_decoder_hw_impl (compiled from the source where it lies) over one
//             srsran::b200::cuda_pusch_dec_accelerator per thread (hal::hw_accelerator_pusch_dec: per-codeblock enqueue /
//             dequeue; the engine's cross-queue batcher coalesces the threads' batches);
//   tb        srsran::b200::pusch_decoder_cuda, synchronous (fec_tb_run_host on a host-call context per call);
//   tb_async  srsran::b200::pusch_decoder_cuda, asynchronous, two instances per thread (fec_tb_enqueue + completion thread).
// Everything goes through the reference's interfaces (rx_buffer_pool::reserve, pusch_decoder::new_data,
// pusch_decoder_buffer::get_next_block_view / on_new_softbits / on_end_softbits, pusch_decoder_notifier::on_sch_data).
// Throughput counts the whole loop, including the copy of the block's soft bits into the decoder's buffer; latency runs from
// on_end_softbits to the notification, measured while all T threads are loading the GPU. One JSON line per (mode, T).
#include "cuda_pusch_dec.h"
#include "phy/upper/channel_processors/pusch/pusch_decoder_hw_impl.h"
#include "srsran/phy/upper/channel_coding/channel_coding_factories.h"
#include "srsran/phy/upper/rx_buffer_pool.h"
#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <random>
#include <string>
#include <thread>
#include <vector>

using namespace srsran;
using clk = std::chrono::steady_clock;

namespace {

constexpr unsigned TBS = 1277992, G = 1362816, NCB = 152;

struct waiter : public pusch_decoder_notifier {
  std::mutex              m;
  std::condition_variable cv;
  bool                    got = false;
  pusch_decoder_result    r;
  clk::time_point         t_done;
  void                    on_sch_data(const pusch_decoder_result& result) override
  {
    std::lock_guard<std::mutex> lk(m);
    r      = result;
    t_done = clk::now();
    got    = true;
    cv.notify_one();
  }
  void wait()
  {
    std::unique_lock<std::mutex> lk(m);
    cv.wait(lk, [this] { return got; });
    got = false;
  }
};

std::unique_ptr<pusch_decoder> make_hal_decoder(hal::hw_accelerator_pusch_dec_factory& factory)
{
  auto                           crcf = create_crc_calculator_factory_sw("auto");
  pusch_decoder_hw_impl::sch_crc crcs;
  crcs.crc16  = crcf->create(crc_generator_poly::CRC16);
  crcs.crc24A = crcf->create(crc_generator_poly::CRC24A);
  crcs.crc24B = crcf->create(crc_generator_poly::CRC24B);
  std::vector<std::unique_ptr<hal::hw_accelerator_pusch_dec>> accs;
  accs.push_back(factory.create());
  auto pool = std::make_shared<pusch_decoder_hw_impl::hw_decoder_pool>(std::move(accs));
  return std::make_unique<pusch_decoder_hw_impl>(create_ldpc_segmenter_rx_factory_sw()->create(), crcs, pool, nullptr);
}

struct result_t {
  double              seconds = 0;
  uint64_t            blocks  = 0, bad = 0;
  std::vector<double> lat_us;
};

} // namespace

// One mode at one thread count. Returns false if a decoder failed.
static bool run_mode(const std::string& mode, unsigned T, double seconds, double esn0_db, unsigned max_wait_us, unsigned batch_cbs)
{
  const unsigned inst_per_thread = (mode == "tb_async") ? 2 : 1;
  const unsigned n_inst          = T * inst_per_thread;
  static const unsigned hctx_cap = getenv("BENCH_HAL_CONTEXTS") ? (unsigned)std::atoi(getenv("BENCH_HAL_CONTEXTS")) : 16u;
  fec_engine_cfg ecfg;
  fec_engine_default_cfg(&ecfg);
  ecfg.harq_slots      = n_inst * 2 * 162 + 162;
  ecfg.max_batch_cbs   = 162;
  ecfg.max_batch_llrs  = G + 4096;
  ecfg.nof_queues      = T;
  ecfg.queue_batch_cbs = batch_cbs; // default 38, a quarter of the block: the GPU starts while the consumer is still enqueuing
  ecfg.queue_batches   = 6;
  ecfg.max_wait_us     = max_wait_us;
  ecfg.host_contexts   = std::min(hctx_cap, std::max(2u, n_inst + 2));
  auto engine          = std::make_shared<b200::fec_engine_handle>(ecfg);
  b200::cuda_pusch_dec_factory factory_shared(engine);

  rx_buffer_pool_config pc;
  pc.max_codeblock_size   = ldpc::MAX_CODEBLOCK_SIZE;
  pc.nof_buffers          = n_inst * 2;
  pc.nof_codeblocks       = n_inst * 2 * 162;
  pc.expire_timeout_slots = 1000000;
  pc.external_soft_bits   = true; // soft bits stay in HBM
  auto pool               = create_rx_buffer_pool(pc);
  std::mutex pool_mtx;    // the reference's pool is driven from one thread per cell; the bench shares one

  // soft bits of one block: all-zero codeword (zero payload, zero CRCs), BPSK-per-bit AWGN, quantised like bench.py
  std::vector<int8_t> llr(G);
  {
    std::mt19937                    rng(7);
    std::normal_distribution<float> nd(0.f, 1.f);
    const float sigma2 = 0.5f * std::pow(10.f, -(float)esn0_db / 10.f), sigma = std::sqrt(sigma2);
    for (unsigned i = 0; i != G; ++i) {
      const float y = 1.f + sigma * nd(rng);
      const float v = std::max(-20.f, std::min(20.f, 2.f * y / sigma2)) * 6.f;
      llr[i]        = (int8_t)std::lround(v);
    }
  }
  pusch_decoder::configuration cfg;
  cfg.base_graph          = ldpc_base_graph_type::BG1;
  cfg.rv                  = 0;
  cfg.mod                 = modulation_scheme::QAM256;
  cfg.Nref                = 0;
  cfg.nof_layers          = 4;
  cfg.nof_ldpc_iterations = 6;
  cfg.use_early_stop      = true;
  cfg.new_data            = true;

  std::vector<result_t>    results(T);
  std::atomic<bool>        stop{false}, failed{false};
  std::atomic<unsigned>    ready{0};
  std::vector<std::thread> threads;
  std::atomic<uint32_t>    slot_count{0};
  for (unsigned k = 0; k != T; ++k) {
    threads.emplace_back([&, k] {
      std::vector<std::unique_ptr<pusch_decoder>> dec;
      for (unsigned i = 0; i != inst_per_thread; ++i) {
        if (mode == "hal") {
          dec.push_back(make_hal_decoder(factory_shared));
        } else {
          dec.push_back(std::make_unique<b200::pusch_decoder_cuda>(engine, G + 64, mode == "tb_async"));
        }
      }
      std::vector<waiter>               note(inst_per_thread);
      std::vector<std::vector<uint8_t>> tb(inst_per_thread, std::vector<uint8_t>(TBS / 8));
      std::vector<clk::time_point>      t_start(inst_per_thread);
      std::vector<bool>                 busy(inst_per_thread, false);
      result_t&                         R = results[k];
      auto collect = [&](unsigned i) {
        note[i].wait();
        busy[i] = false;
        R.lat_us.push_back(std::chrono::duration<double, std::micro>(note[i].t_done - t_start[i]).count());
        ++R.blocks;
        if (!note[i].r.tb_crc_ok) {
          ++R.bad;
        }
      };
      ready.fetch_add(1);
      while (ready.load() != T) {
        std::this_thread::yield();
      }
      const auto t0 = clk::now();
      unsigned   i  = 0;
      while (!stop.load(std::memory_order_relaxed)) {
        if (busy[i]) {
          collect(i);
        }
        trx_buffer_identifier id(static_cast<uint16_t>(1 + k * inst_per_thread + i), 0);
        unique_rx_buffer      buf;
        {
          std::lock_guard<std::mutex> lk(pool_mtx);
          buf = pool->get_pool().reserve(slot_point(0, slot_count.fetch_add(1) % 10240), id, NCB, true);
        }
        if (!buf.is_valid()) {
          failed = true;
          break;
        }
        pusch_decoder_buffer&      in   = dec[i]->new_data(tb[i], std::move(buf), note[i], cfg);
        span<log_likelihood_ratio> view = in.get_next_block_view(G);
        std::memcpy(view.data(), llr.data(), G);
        t_start[i] = clk::now();
        in.on_new_softbits(view);
        in.on_end_softbits();
        busy[i] = true;
        if (inst_per_thread == 1) {
          collect(i);
        }
        i = (i + 1) % inst_per_thread;
      }
      for (unsigned j = 0; j != inst_per_thread; ++j) {
        if (busy[j]) {
          collect(j);
        }
      }
      R.seconds = std::chrono::duration<double>(clk::now() - t0).count();
    });
  }
  std::this_thread::sleep_for(std::chrono::duration<double>(seconds));
  stop = true;
  for (auto& t : threads) {
    t.join();
  }
  uint64_t            blocks = 0, bad = 0;
  double              secs = 0;
  std::vector<double> lat;
  for (const result_t& R : results) {
    blocks += R.blocks;
    bad += R.bad;
    secs = std::max(secs, R.seconds);
    // the first blocks of a thread include one-time allocations (staging of its queue / context): not steady state
    lat.insert(lat.end(), R.lat_us.begin() + std::min<size_t>(R.lat_us.size() / 10, 5), R.lat_us.end());
  }
  std::sort(lat.begin(), lat.end());
  fec_engine_stats st;
  fec_engine_get_stats(engine->get(), &st);
  auto pct = [&](double p) { return lat.empty() ? 0.0 : lat[std::min<size_t>(lat.size() - 1, (size_t)(p * lat.size()))]; };
  std::printf("{\"mode\": \"%s\", \"threads\": %u, \"seconds\": %.3f, \"transport_blocks\": %llu, \"tb_crc_fail\": %llu, "
              "\"gbps\": %.3f, \"lat_us_p50\": %.1f, \"lat_us_p99\": %.1f, \"lat_us_max\": %.1f, \"tbs_bits\": %u, "
              "\"codeblocks_per_launch\": %.1f, \"queue_full\": %llu, \"dispatches\": %llu, \"batches_per_dispatch\": %.2f, "
              "\"max_wait_us\": %u}\n",
              mode.c_str(), T, secs, (unsigned long long)blocks, (unsigned long long)bad, blocks * (double)TBS / secs / 1e9,
              pct(0.5), pct(0.99), lat.empty() ? 0.0 : lat.back(), TBS,
              st.cb_launches ? (double)st.codeblocks / st.cb_launches : 0.0, (unsigned long long)st.queue_full,
              (unsigned long long)st.dispatches, st.dispatches ? (double)st.dispatch_batches / st.dispatches : 0.0, max_wait_us);
  std::fflush(stdout);
  return !failed.load() && bad == 0 && blocks != 0;
}

int main(int argc, char** argv)
{
  // usage: bench_hal [seconds per point] [comma-separated thread counts] [modes] [Es/N0 dB] [max_wait_us] [queue_batch_cbs]
  const double      seconds = (argc > 1) ? std::atof(argv[1]) : 2.0;
  const std::string tlist   = (argc > 2) ? argv[2] : "1,4,16,32";
  const std::string mlist   = (argc > 3) ? argv[3] : "hal,tb,tb_async";
  const double      esn0    = (argc > 4) ? std::atof(argv[4]) : 10.0;
  const unsigned    wait_us = (argc > 5) ? std::atoi(argv[5]) : 0;
  const unsigned    batch   = (argc > 6) ? std::atoi(argv[6]) : 38;
  auto split = [](const std::string& s) {
    std::vector<std::string> out;
    size_t                   a = 0;
    while (a <= s.size()) {
      const size_t b = s.find(',', a);
      out.push_back(s.substr(a, b == std::string::npos ? std::string::npos : b - a));
      if (b == std::string::npos) {
        break;
      }
      a = b + 1;
    }
    return out;
  };
  bool ok = true;
  for (const std::string& m : split(mlist)) {
    for (const std::string& t : split(tlist)) {
      ok = run_mode(m, (unsigned)std::atoi(t.c_str()), seconds, esn0, wait_us, batch) && ok;
    }
  }
  return ok ? 0 : 1;
}

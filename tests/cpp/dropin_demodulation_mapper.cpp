// tests/cpp/dropin_demodulation_mapper.cpp -- TEST INFRASTRUCTURE.
// Acceptance test of the reference-side soft-demodulator binding: the reference's own demodulation_mapper_impl (compiled
// into this binary from /root/reference with this binary's instruction-set flags, i.e. one of the three flavours) against
// cuda_demodulation_mapper created through the demodulation_mapper_factory interface, on the same inputs, bit for bit.
// Usage: dropin_demodulation_mapper [nof_random_cases]
#include "cuda_demod_mapper.h"
#include "phy/upper/channel_modulation/demodulation_mapper_impl.h"
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <limits>
#include <random>
#include <vector>

using namespace srsran;

int main(int argc, char** argv)
{
  unsigned       n_cases = argc > 1 ? std::atoi(argv[1]) : 60;
  fec_engine_cfg cfg;
  fec_engine_default_cfg(&cfg);
  cfg.harq_slots    = 16;
  cfg.max_batch_cbs = 16;
  std::shared_ptr<demodulation_mapper_factory> factory = b200::create_cuda_demodulation_mapper_factory(cfg);
  std::unique_ptr<demodulation_mapper>         cuda    = factory->create();
  demodulation_mapper_impl                     ref;

  const modulation_scheme mods[] = {modulation_scheme::PI_2_BPSK, modulation_scheme::BPSK, modulation_scheme::QPSK,
                                    modulation_scheme::QAM16,     modulation_scheme::QAM64, modulation_scheme::QAM256};
  std::mt19937            rng(2026);
  unsigned                total = 0;
  double                  us_ref = 0, us_cuda = 0;
  for (unsigned c = 0; c != n_cases; ++c) {
    modulation_scheme mod = mods[c % 6];
    unsigned          qm  = get_bits_per_symbol(mod);
    // lengths around the SIMD block sizes and the size of one OFDM symbol of a wide allocation
    const unsigned lens[] = {1, 3, 7, 8, 15, 16, 31, 32, 33, 255, 257, 3276, 13104};
    unsigned       n      = lens[(c / 6) % 13] + (c >= 78 ? rng() % 50 : 0);
    std::vector<cf_t>  sym(n);
    std::vector<float> nv(n);
    std::uniform_real_distribution<float> amp(-1.4F, 1.4F), var(0.002F, 0.4F);
    for (unsigned i = 0; i != n; ++i) {
      sym[i] = cf_t(amp(rng), amp(rng));
      nv[i]  = var(rng);
    }
    if (n >= 16) {
      nv[rng() % n]  = 0.0F;
      nv[rng() % n]  = -1.0F;
      nv[rng() % n]  = std::numeric_limits<float>::quiet_NaN();
      sym[rng() % n] = cf_t(0.0F, 0.0F);
      sym[rng() % n] = cf_t(1e-10F, 0.5F);
    }
    std::vector<log_likelihood_ratio> a(n * qm), b(n * qm);
    auto t0 = std::chrono::steady_clock::now();
    ref.demodulate_soft(a, sym, nv, mod);
    auto t1 = std::chrono::steady_clock::now();
    cuda->demodulate_soft(b, sym, nv, mod);
    auto t2 = std::chrono::steady_clock::now();
    us_ref += std::chrono::duration<double, std::micro>(t1 - t0).count();
    us_cuda += std::chrono::duration<double, std::micro>(t2 - t1).count();
    for (unsigned i = 0; i != n * qm; ++i) {
      if (a[i].to_int() != b[i].to_int()) {
        std::printf("MISMATCH case %u mod %s n %u llr %u: reference %d cuda %d\n", c, to_string(mod).c_str(), n, i,
                    a[i].to_int(), b[i].to_int());
        return 1;
      }
    }
    total += n * qm;
  }
  std::printf("flavour=%u reference_us=%.0f cuda_us=%.0f\n", b200::native_demod_flavour(), us_ref, us_cuda);
  std::printf("OK cases=%u soft_bits=%u\n", n_cases, total);
  return 0;
}

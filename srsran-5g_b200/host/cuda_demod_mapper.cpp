// srsran-5g_b200/host/cuda_demod_mapper.cpp -- see cuda_demod_mapper.h.
#include "cuda_demod_mapper.h"
#include "srsran/support/error_handling.h"
#include "srsran/support/srsran_assert.h"

using namespace srsran;
using namespace srsran::b200;

void cuda_demodulation_mapper::demodulate_soft(span<log_likelihood_ratio> llrs,
                                               span<const cf_t>           symbols,
                                               span<const float>          noise_vars,
                                               modulation_scheme          mod)
{
  // Same preconditions and wording as demodulation_mapper_impl.cpp:82-83.
  srsran_assert(symbols.size() == noise_vars.size(), "Inputs symbols and noise_vars must have the same length.");
  srsran_assert(symbols.size() * get_bits_per_symbol(mod) == llrs.size(), "Input and output lengths are incompatible.");
  if (symbols.empty()) {
    return;
  }
  fec_demod_desc d = {};
  d.nof_symbols    = static_cast<uint32_t>(symbols.size());
  d.qm             = static_cast<uint8_t>(get_bits_per_symbol(mod));
  d.flags          = FEC_DEMOD_FLAG_NO_DESCRAMBLE | ((mod == modulation_scheme::PI_2_BPSK) ? FEC_DEMOD_FLAG_PI2_BPSK : 0);
  static_assert(sizeof(cf_t) == 2 * sizeof(float) && sizeof(log_likelihood_ratio) == 1, "layout");
  int rc = fec_demod_run_host(engine->get(),
                              &d,
                              1,
                              reinterpret_cast<const float*>(symbols.data()),
                              noise_vars.data(),
                              symbols.size(),
                              reinterpret_cast<int8_t*>(llrs.data()),
                              llrs.size(),
                              flavour);
  report_fatal_error_if_not(rc == FEC_OK, "B200 FEC engine: soft demodulation failed: {}", fec_last_error());
}

std::shared_ptr<demodulation_mapper_factory> srsran::b200::create_cuda_demodulation_mapper_factory(const fec_engine_cfg& cfg,
                                                                                                   uint32_t flavour)
{
  return std::make_shared<cuda_demodulation_mapper_factory>(std::make_shared<fec_engine_handle>(cfg), flavour);
}

// srsran-5g_b200/host/cuda_demod_mapper.h -- reference-side binding of the soft demodulator of the B200 FEC engine.
//
//   cuda_demodulation_mapper         : demodulation_mapper
//        include/srsran/phy/upper/channel_modulation/demodulation_mapper.h:35-70 -- the object pusch_demodulator_impl calls
//        once per OFDM symbol (lib/phy/upper/channel_processors/pusch/pusch_demodulator_impl.cpp:355-375)
//   cuda_demodulation_mapper_factory : demodulation_mapper_factory
//        include/srsran/phy/upper/channel_modulation/channel_modulation_factories.h:41-48 -- what
//        create_pusch_demodulator_factory_sw takes as its demodulation factory
//
// The reference picks its demodulator arithmetic at compile time (AVX-512 / AVX2 / scalar) and the variants differ in the
// last place; `flavour` names the build whose results the adapter reproduces bit for bit (FEC_DEMOD_FLAVOUR_*; the default
// is what this translation unit itself is compiled for, i.e. what the reference would be in the same build).
//
// Each call is one host -> device -> host round trip: the drop-in is exact, not fast. The efficient integration hands the
// equalised symbols of the whole allocation to fec_tb_run_host_symbols and never sees the LLRs.
#pragma once

#include "cuda_pusch_dec.h"
#include "fec_b200.h"
#include "srsran/phy/upper/channel_modulation/channel_modulation_factories.h"
#include "srsran/phy/upper/channel_modulation/demodulation_mapper.h"
#include <memory>

namespace srsran {
namespace b200 {

/// The flavour of the reference this translation unit would compile to (same #if conditions as the reference sources).
constexpr uint32_t native_demod_flavour()
{
#if defined(__AVX512F__) && defined(__AVX512BW__) && defined(__AVX2__) && (!defined(__GNUC__) || (__GNUC__ > 9))
  return FEC_DEMOD_FLAVOUR_AVX512;
#elif defined(__AVX2__)
  return FEC_DEMOD_FLAVOUR_AVX2;
#else
  return FEC_DEMOD_FLAVOUR_GENERIC;
#endif
}

class cuda_demodulation_mapper : public demodulation_mapper
{
public:
  cuda_demodulation_mapper(std::shared_ptr<fec_engine_handle> engine_, uint32_t flavour_) :
    engine(std::move(engine_)), flavour(flavour_)
  {
  }

  // See demodulation_mapper for the documentation.
  void demodulate_soft(span<log_likelihood_ratio> llrs,
                       span<const cf_t>           symbols,
                       span<const float>          noise_vars,
                       modulation_scheme          mod) override;

private:
  std::shared_ptr<fec_engine_handle> engine;
  uint32_t                           flavour;
};

class cuda_demodulation_mapper_factory : public demodulation_mapper_factory
{
public:
  explicit cuda_demodulation_mapper_factory(std::shared_ptr<fec_engine_handle> engine_,
                                            uint32_t                           flavour_ = native_demod_flavour()) :
    engine(std::move(engine_)), flavour(flavour_)
  {
  }
  std::unique_ptr<demodulation_mapper> create() override
  {
    return std::make_unique<cuda_demodulation_mapper>(engine, flavour);
  }

private:
  std::shared_ptr<fec_engine_handle> engine;
  uint32_t                           flavour;
};

/// The call that replaces create_demodulation_mapper_factory (lib/phy/upper/channel_modulation/channel_modulation_factories.cpp).
std::shared_ptr<demodulation_mapper_factory> create_cuda_demodulation_mapper_factory(const fec_engine_cfg& cfg,
                                                                                     uint32_t flavour = native_demod_flavour());

} // namespace b200
} // namespace srsran

// oracle/ref_demod_shim.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// C-ABI wrapper around the UNMODIFIED reference soft demodulator and scrambling-sequence generator, compiled from the
// sources where they lie under /root/reference by oracle/Makefile. The reference selects its SIMD code at COMPILE time
// for these files (no run-time dispatch), and its CMake builds them with -march=native, so the library is built in three
// flavours (oracle/_ref/libref_demod_<flavour>.so): avx512 (AVX-512F/BW/VL + AVX2 + FMA: what -march=native gives on the
// hosts used here), avx2 (AVX2 + FMA) and generic (baseline x86-64, scalar code only). Nothing from the reference is
// copied; this file instantiates the reference classes and forwards calls.
//
// Reference entry points wrapped (file:line under /root/reference):
//   demodulation_mapper_impl::demodulate_soft   lib/phy/upper/channel_modulation/demodulation_mapper_impl.cpp:77-106
//   pseudo_random_generator_impl::init/advance/generate/apply_xor
//                                               lib/phy/upper/sequence_generators/pseudo_random_generator_impl.cpp
#include "phy/upper/channel_modulation/demodulation_mapper_impl.h"
#include "phy/upper/sequence_generators/pseudo_random_generator_impl.h"
#include "srsran/adt/bit_buffer.h"
#include <cstdint>
#include <vector>

using namespace srsran;

namespace {
modulation_scheme scheme_of(int qm, int pi2)
{
  switch (qm) {
    case 1:
      return pi2 ? modulation_scheme::PI_2_BPSK : modulation_scheme::BPSK;
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

extern "C" {

// One demodulation_mapper::demodulate_soft call: n symbols (interleaved re, im), n noise variances -> n * qm LLRs.
void ref_demodulate_soft(int qm, int pi2, unsigned n, const float* symbols, const float* noise, int8_t* llr)
{
  demodulation_mapper_impl demod;
  demod.demodulate_soft(span<log_likelihood_ratio>(reinterpret_cast<log_likelihood_ratio*>(llr), (size_t)n * qm),
                        span<const cf_t>(reinterpret_cast<const cf_t*>(symbols), n),
                        span<const float>(noise, n),
                        scheme_of(qm, pi2));
}

// The scrambling sequence as pusch_demodulator_impl uses it: init(c_init), skip `offset` bits, generate nbits (packed MSB
// first, as bit_buffer stores them).
void ref_prg_generate(unsigned c_init, unsigned offset, unsigned nbits, uint8_t* packed)
{
  pseudo_random_generator_impl prg;
  prg.init(c_init);
  prg.advance(offset);
  dynamic_bit_buffer seq(nbits);
  prg.generate(seq);
  span<const uint8_t> bytes = seq.get_buffer();
  for (unsigned i = 0; i != (nbits + 7) / 8; ++i) {
    packed[i] = bytes[i];
  }
}

// Descrambling of soft bits with the generator's own LLR overload.
void ref_prg_apply_xor_llr(unsigned c_init, unsigned offset, unsigned n, const int8_t* in, int8_t* out)
{
  pseudo_random_generator_impl prg;
  prg.init(c_init);
  prg.advance(offset);
  prg.apply_xor(span<log_likelihood_ratio>(reinterpret_cast<log_likelihood_ratio*>(out), n),
                span<const log_likelihood_ratio>(reinterpret_cast<const log_likelihood_ratio*>(in), n));
}

} // extern "C"

// oracle/ref_shim.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// Thin C-ABI wrapper around the UNMODIFIED reference classes, compiled from the sources where they lie under
// /root/reference by oracle/Makefile into oracle/_ref/libref_fec.so. Nothing from the reference is copied: this file
// only instantiates the reference's public factories and forwards calls. It is used
//   * by tests/ to pin the C restatement (oracle/fec_oracle.c) and to generate tests/golden fixtures,
//   * by bench.py's cpu_baseline / --impl reference leg (AVX-512 / AVX2 decoder timed on the host cores).
// The product path (srsran-5g_b200/csrc) never links or loads it.
//
// Reference entry points wrapped (file:line under /root/reference):
//   create_ldpc_decoder_factory_sw        lib/phy/upper/channel_coding/channel_coding_factories.cpp:102-135
//   create_ldpc_rate_dematcher_factory_sw lib/phy/upper/channel_coding/channel_coding_factories.cpp:168-203
//   create_ldpc_encoder_factory_sw / create_ldpc_rate_matcher_factory_sw / create_crc_calculator_factory_sw (same file)
//   ldpc_graph_impl getters               lib/phy/upper/channel_coding/ldpc/ldpc_graph_impl.h
//   pusch_codeblock_decoder::decode       lib/phy/upper/channel_processors/pusch/pusch_codeblock_decoder.cpp:34-71 (re-stated
//                                         in ref_cb_chain below through the same three reference objects)
//   pusch_decoder_impl + rx_buffer_pool   lib/phy/upper/channel_processors/pusch/pusch_decoder_impl.{h,cpp},
//                                         lib/phy/upper/rx_buffer_pool_impl.cpp (whole transport blocks: ref_tb_* below)
#include "ldpc_graph_impl.h"
#include "ldpc_luts_impl.h"
#include "phy/upper/channel_processors/pusch/pusch_decoder_impl.h"
#include "srsran/phy/upper/channel_processors/pusch/pusch_decoder_notifier.h"
#include "srsran/phy/upper/channel_processors/pusch/pusch_decoder_result.h"
#include "srsran/phy/upper/rx_buffer_pool.h"
#include "srsran/phy/upper/unique_rx_buffer.h"
#include "srsran/phy/upper/channel_coding/channel_coding_factories.h"
#include "srsran/phy/upper/channel_coding/ldpc/ldpc_encoder_buffer.h"
#include "srsran/ran/sch/sch_segmentation.h"
#include "srsran/ran/sch/tbs_calculator.h"
#include "srsran/srsvec/bit.h"
#include <atomic>
#include <chrono>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

using namespace srsran;

namespace {

// Per-thread set of reference objects, keyed by implementation name ("generic", "avx2", "avx512", "auto").
struct ref_objs {
  std::unique_ptr<ldpc_decoder>        dec;
  std::unique_ptr<ldpc_rate_dematcher> dem;
  std::unique_ptr<ldpc_encoder>        enc;
  std::unique_ptr<ldpc_rate_matcher>   rm;
  std::unique_ptr<crc_calculator>      crc[3]; // 24A, 24B, 16
};

const char* impl_name(int impl)
{
  switch (impl) {
    case 0:
      return "generic";
    case 1:
      return "avx2";
    case 2:
      return "avx512";
    default:
      return "auto";
  }
}

ref_objs* get_objs(int impl)
{
  thread_local std::map<int, std::unique_ptr<ref_objs>> cache;
  auto                                                 it = cache.find(impl);
  if (it != cache.end()) {
    return it->second.get();
  }
  auto        o    = std::make_unique<ref_objs>();
  const char* name = impl_name(impl);
  auto        decf = create_ldpc_decoder_factory_sw(name);
  auto        demf = create_ldpc_rate_dematcher_factory_sw(name);
  if (!decf || !demf) {
    return nullptr;
  }
  o->dec = decf->create();
  o->dem = demf->create();
  // Tx side and CRC: generic encoder is bit-identical to the SIMD ones; CRC implementations agree bit for bit.
  o->enc        = create_ldpc_encoder_factory_sw("generic")->create();
  o->rm         = create_ldpc_rate_matcher_factory_sw()->create();
  auto crcf     = create_crc_calculator_factory_sw("auto");
  o->crc[0]     = crcf->create(crc_generator_poly::CRC24A);
  o->crc[1]     = crcf->create(crc_generator_poly::CRC24B);
  o->crc[2]     = crcf->create(crc_generator_poly::CRC16);
  ref_objs* ret = o.get();
  cache[impl]   = std::move(o);
  return ret;
}

modulation_scheme mod_from_qm(unsigned qm)
{
  switch (qm) {
    case 1:
      return modulation_scheme::BPSK;
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

codeblock_metadata make_meta(unsigned bg,
                             unsigned Z,
                             unsigned rv,
                             unsigned qm,
                             unsigned Nref,
                             unsigned E,
                             unsigned fillers,
                             unsigned crc_bits)
{
  codeblock_metadata m;
  m.tb_common.base_graph        = (bg == 1) ? ldpc_base_graph_type::BG1 : ldpc_base_graph_type::BG2;
  m.tb_common.lifting_size      = static_cast<ldpc::lifting_size_t>(Z);
  m.tb_common.rv                = rv;
  m.tb_common.mod               = mod_from_qm(qm);
  m.tb_common.Nref              = Nref;
  m.tb_common.cw_length         = E;
  m.cb_specific.full_length     = ((bg == 1) ? 66 : 50) * Z;
  m.cb_specific.rm_length       = E;
  m.cb_specific.nof_filler_bits = fillers;
  m.cb_specific.cw_offset       = 0;
  m.cb_specific.nof_crc_bits    = crc_bits;
  return m;
}

// Encoder-buffer view over an unpacked codeblock (one bit per byte, fillers marked with ldpc::FILLER_BIT), so that the
// reference rate matcher can be driven with codeblocks produced by any encoder.
class span_encoder_buffer : public ldpc_encoder_buffer
{
public:
  explicit span_encoder_buffer(span<const uint8_t> bits_) : bits(bits_) {}
  unsigned get_codeblock_length() const override { return bits.size(); }
  void     write_codeblock(span<uint8_t> data, unsigned offset) const override
  {
    std::memcpy(data.data(), bits.data() + offset, data.size());
  }

private:
  span<const uint8_t> bits;
};

} // namespace

extern "C" {

/// Returns 1 if the named reference implementation is available on this CPU (0 generic, 1 avx2, 2 avx512, 3 auto).
int ref_impl_available(int impl)
{
  return get_objs(impl) != nullptr;
}

/// Dumps one lifted graph: for every base row the connected columns (ascending) and the shifts mod Z.
/// cols/shifts are [M][20]; deg is [M]. Returns M. bg in {1,2}.
int ref_graph(int bg, int Z, uint16_t* cols, uint16_t* shifts, uint16_t* deg)
{
  ldpc_graph_impl g((bg == 1) ? ldpc_base_graph_type::BG1 : ldpc_base_graph_type::BG2,
                          static_cast<ldpc::lifting_size_t>(Z));
  unsigned              M = g.get_nof_BG_check_nodes();
  for (unsigned m = 0; m != M; ++m) {
    const auto& row = g.get_adjacency_row(m);
    unsigned    d   = 0;
    for (unsigned e = 0; e != ldpc::MAX_BG_CHECK_EDGES; ++e) {
      if (row[e] == ldpc::NO_EDGE) {
        break;
      }
      cols[m * 20 + d]   = row[e];
      shifts[m * 20 + d] = g.get_lifted_node(m, row[e]);
      ++d;
    }
    deg[m] = d;
  }
  return M;
}

/// Lifting-set index of Z (TS 38.212 Table 5.3.2-1), 255 if invalid.
int ref_lifting_index(int Z)
{
  return ldpc::get_lifting_index(static_cast<ldpc::lifting_size_t>(Z));
}

/// CRC remainder over the first nbits of a packed (MSB-first) buffer. poly: 0=CRC24A, 1=CRC24B, 2=CRC16.
unsigned ref_crc(int poly, const uint8_t* bytes, unsigned nbits)
{
  ref_objs*  o = get_objs(3);
  bit_buffer bb = bit_buffer::from_bytes(span<uint8_t>(const_cast<uint8_t*>(bytes), (nbits + 7) / 8)).first(nbits);
  return o->crc[poly]->calculate(bb);
}

/// LDPC-encodes K=Kb*Z message bits (one bit per byte, fillers as 0) into N=66Z|50Z unpacked bits.
void ref_encode(int bg, int Z, const uint8_t* msg_bits, uint8_t* out_bits)
{
  ref_objs*                         o  = get_objs(0);
  unsigned                          K  = ((bg == 1) ? 22 : 10) * Z;
  unsigned                          N  = ((bg == 1) ? 66 : 50) * Z;
  dynamic_bit_buffer                in(K);
  srsvec::bit_pack(in, span<const uint8_t>(msg_bits, K));
  codeblock_metadata                m   = make_meta(bg, Z, 0, 1, 0, N, 0, 24);
  const ldpc_encoder_buffer&        buf = o->enc->encode(in, m.tb_common);
  buf.write_codeblock(span<uint8_t>(out_bits, N), 0);
}

/// Rate-matches an unpacked codeblock (fillers must be marked 254) into E unpacked bits.
void ref_rate_match(int            bg,
                    int            Z,
                    const uint8_t* cb_bits,
                    uint8_t*       out_bits,
                    unsigned       E,
                    unsigned       rv,
                    unsigned       qm,
                    unsigned       Nref,
                    unsigned       fillers)
{
  ref_objs*           o = get_objs(0);
  unsigned            N = ((bg == 1) ? 66 : 50) * Z;
  span_encoder_buffer buf(span<const uint8_t>(cb_bits, N));
  dynamic_bit_buffer  out(E);
  codeblock_metadata  m = make_meta(bg, Z, rv, qm, Nref, E, fillers, 24);
  o->rm->rate_match(out, buf, m);
  srsvec::bit_unpack(span<uint8_t>(out_bits, E), out);
}

/// ldpc_rate_dematcher::rate_dematch on a caller-owned N-byte soft buffer (updated in place).
int ref_rate_dematch(int           impl,
                     int           bg,
                     int           Z,
                     int8_t*       soft,
                     const int8_t* in,
                     unsigned      E,
                     int           new_data,
                     unsigned      rv,
                     unsigned      qm,
                     unsigned      Nref,
                     unsigned      fillers)
{
  ref_objs* o = get_objs(impl);
  if (!o) {
    return -1;
  }
  unsigned           N = ((bg == 1) ? 66 : 50) * Z;
  codeblock_metadata m = make_meta(bg, Z, rv, qm, Nref, E, fillers, 24);
  o->dem->rate_dematch(span<log_likelihood_ratio>(reinterpret_cast<log_likelihood_ratio*>(soft), N),
                       span<const log_likelihood_ratio>(reinterpret_cast<const log_likelihood_ratio*>(in), E),
                       new_data != 0,
                       m);
  return 0;
}

/// ldpc_decoder::decode. crc_poly: -1 = nullptr (no early stop), 0=CRC24A, 1=CRC24B, 2=CRC16.
/// out must hold ceil(Kb*Z/8) bytes and is passed in as-is (so "left untouched" behaviour is observable).
/// Returns the iteration count, or -1 for nullopt.
int ref_decode(int           impl,
               int           bg,
               int           Z,
               uint8_t*      out,
               const int8_t* llr,
               unsigned      n_llr,
               unsigned      fillers,
               int           crc_poly,
               unsigned      max_iter,
               float         scaling)
{
  ref_objs* o = get_objs(impl);
  if (!o) {
    return -2;
  }
  unsigned                    K = ((bg == 1) ? 22 : 10) * Z;
  bit_buffer                  bb = bit_buffer::from_bytes(span<uint8_t>(out, (K + 7) / 8)).first(K);
  ldpc_decoder::configuration cfg;
  cfg.block_conf = make_meta(bg, Z, 0, 1, 0, n_llr, fillers, (crc_poly == 2) ? 16 : 24);
  cfg.algorithm_conf.max_iterations = max_iter;
  cfg.algorithm_conf.scaling_factor = scaling;
  crc_calculator*         crc       = (crc_poly < 0) ? nullptr : o->crc[crc_poly].get();
  std::optional<unsigned> r =
      o->dec->decode(bb, span<const log_likelihood_ratio>(reinterpret_cast<const log_likelihood_ratio*>(llr), n_llr), crc, cfg);
  return r.has_value() ? static_cast<int>(*r) : -1;
}

/// One codeblock through dematch -> decode -> CRC exactly as pusch_codeblock_decoder::decode composes them.
/// Returns iterations (>=1) when the CB CRC passes, -1 otherwise.
int ref_cb_chain(int           impl,
                 int           bg,
                 int           Z,
                 int8_t*       soft,
                 const int8_t* in,
                 unsigned      E,
                 int           new_data,
                 unsigned      rv,
                 unsigned      qm,
                 unsigned      Nref,
                 unsigned      fillers,
                 int           crc_poly,
                 unsigned      max_iter,
                 int           early_stop,
                 uint8_t*      out)
{
  ref_objs* o = get_objs(impl);
  if (!o) {
    return -2;
  }
  unsigned           N = ((bg == 1) ? 66 : 50) * Z;
  unsigned           K = ((bg == 1) ? 22 : 10) * Z;
  codeblock_metadata m = make_meta(bg, Z, rv, qm, Nref, E, fillers, (crc_poly == 2) ? 16 : 24);
  span<log_likelihood_ratio> rm_buffer(reinterpret_cast<log_likelihood_ratio*>(soft), N);
  o->dem->rate_dematch(
      rm_buffer, span<const log_likelihood_ratio>(reinterpret_cast<const log_likelihood_ratio*>(in), E), new_data != 0, m);
  bit_buffer                  bb = bit_buffer::from_bytes(span<uint8_t>(out, (K + 7) / 8)).first(K);
  ldpc_decoder::configuration cfg;
  cfg.block_conf                    = m;
  cfg.algorithm_conf.max_iterations = max_iter;
  crc_calculator* crc               = o->crc[crc_poly].get();
  if (early_stop) {
    std::optional<unsigned> r = o->dec->decode(bb, rm_buffer, crc, cfg);
    return r.has_value() ? static_cast<int>(*r) : -1;
  }
  o->dec->decode(bb, rm_buffer, nullptr, cfg);
  if (crc->calculate(bb.first(K - fillers)) == 0) {
    return static_cast<int>(max_iter);
  }
  return -1;
}

/// CPU baseline: n codeblocks (same shape) through ref_cb_chain on n_threads std::threads, statically partitioned.
/// llr is [n][E]; soft is [n][N] scratch (zero-initialised by the caller); out is [n][ceil(K/8)]; iters is [n].
/// Returns elapsed seconds of the parallel region (steady_clock).
double ref_batch_chain(int           impl,
                       int           n_threads,
                       unsigned      n,
                       int           bg,
                       int           Z,
                       int8_t*       soft,
                       const int8_t* llr,
                       unsigned      E,
                       int           new_data,
                       unsigned      rv,
                       unsigned      qm,
                       unsigned      Nref,
                       unsigned      fillers,
                       int           crc_poly,
                       unsigned      max_iter,
                       int           early_stop,
                       uint8_t*      out,
                       int*          iters)
{
  unsigned                 N  = ((bg == 1) ? 66 : 50) * Z;
  unsigned                 Kb = (((bg == 1) ? 22 : 10) * Z + 7) / 8;
  std::vector<std::thread> th;
  std::atomic<int>         ready{0};
  std::atomic<bool>        go{false};
  auto                     work = [&](int t) {
    get_objs(impl); // construct decoder objects outside the timed region
    ready.fetch_add(1);
    while (!go.load(std::memory_order_acquire)) {
    }
    for (unsigned i = t; i < n; i += n_threads) {
      iters[i] = ref_cb_chain(impl,
                              bg,
                              Z,
                              soft + size_t(i) * N,
                              llr + size_t(i) * E,
                              E,
                              new_data,
                              rv,
                              qm,
                              Nref,
                              fillers,
                              crc_poly,
                              max_iter,
                              early_stop,
                              out + size_t(i) * Kb);
    }
  };
  for (int t = 0; t != n_threads; ++t) {
    th.emplace_back(work, t);
  }
  while (ready.load() != n_threads) {
  }
  auto t0 = std::chrono::steady_clock::now();
  go.store(true, std::memory_order_release);
  for (auto& t : th) {
    t.join();
  }
  auto t1 = std::chrono::steady_clock::now();
  return std::chrono::duration<double>(t1 - t0).count();
}

/// TBS for a PUSCH allocation (tbs_calculator_calculate, lib/ran/sch/tbs_calculator.cpp).
unsigned ref_tbs(unsigned nof_symb, unsigned nof_dmrs_prb, unsigned nof_oh_prb, unsigned qm, float target_code_rate_x1024, unsigned nof_layers, unsigned n_prb)
{
  tbs_calculator_configuration c;
  c.nof_symb_sh      = nof_symb;
  c.nof_dmrs_prb     = nof_dmrs_prb;
  c.nof_oh_prb       = nof_oh_prb;
  c.mcs_descr        = {mod_from_qm(qm), target_code_rate_x1024};
  c.nof_layers       = nof_layers;
  c.tb_scaling_field = 0;
  c.n_prb            = n_prb;
  return tbs_calculator_calculate(c);
}

/// Rx segmentation metadata for a TB: fills per-CB rm_length/cw_offset arrays (size >= 162), returns number of CBs.
/// scalars[0..5] = {bg, Z, nof_filler_bits, full_length, nof_crc_bits(cb), K'(info bits per CB incl. crc)}.
int ref_segment(unsigned tbs, int bg, unsigned rv, unsigned qm, unsigned Nref, unsigned nof_layers, unsigned nof_ch_symbols, unsigned* rm_len, unsigned* cw_off, unsigned* scalars)
{
  auto             seg = create_ldpc_segmenter_rx_factory_sw()->create();
  segmenter_config cfg;
  cfg.base_graph     = (bg == 1) ? ldpc_base_graph_type::BG1 : ldpc_base_graph_type::BG2;
  cfg.rv             = rv;
  cfg.mod            = mod_from_qm(qm);
  cfg.Nref           = Nref;
  cfg.nof_layers     = nof_layers;
  cfg.nof_ch_symbols = nof_ch_symbols;
  std::vector<log_likelihood_ratio>                       llrs(size_t(nof_ch_symbols) * qm);
  static_vector<described_rx_codeblock, MAX_NOF_SEGMENTS> cbs;
  seg->segment(cbs, llrs, tbs, cfg);
  for (unsigned i = 0; i != cbs.size(); ++i) {
    rm_len[i] = cbs[i].second.cb_specific.rm_length;
    cw_off[i] = cbs[i].second.cb_specific.cw_offset;
  }
  const codeblock_metadata& m = cbs[0].second;
  scalars[0]                  = (m.tb_common.base_graph == ldpc_base_graph_type::BG1) ? 1 : 2;
  scalars[1]                  = m.tb_common.lifting_size;
  scalars[2]                  = m.cb_specific.nof_filler_bits;
  scalars[3]                  = m.cb_specific.full_length;
  scalars[4]                  = m.cb_specific.nof_crc_bits;
  scalars[5]                  = 0;
  return cbs.size();
}

} // extern "C"


// ---- whole transport blocks through the reference's software PUSCH decoder and its rx buffer pool ------------------------
namespace {

struct ref_tb_notifier : public pusch_decoder_notifier {
  pusch_decoder_result result;
  bool                 notified = false;
  void                 on_sch_data(const pusch_decoder_result& r) override
  {
    result   = r;
    notified = true;
  }
};

struct ref_tb_ctx {
  std::unique_ptr<rx_buffer_pool_controller> pool;
  std::unique_ptr<pusch_decoder>             dec;
  unsigned                                   slot_count = 0;
};

} // namespace

extern "C" {

/// Creates a software PUSCH decoder (generic LDPC kernels) with an rx buffer pool of nof_buffers x nof_codeblocks.
void* ref_tb_create(unsigned nof_buffers, unsigned nof_codeblocks)
{
  auto ctx = new ref_tb_ctx();
  rx_buffer_pool_config pc;
  pc.max_codeblock_size   = ldpc::MAX_CODEBLOCK_SIZE;
  pc.nof_buffers          = nof_buffers;
  pc.nof_codeblocks       = nof_codeblocks;
  pc.expire_timeout_slots = 1000000;
  pc.external_soft_bits   = false;
  ctx->pool               = create_rx_buffer_pool(pc);
  auto crcf               = create_crc_calculator_factory_sw("auto");
  auto make_crcs          = [&crcf]() {
    pusch_decoder_impl::sch_crc c;
    c.crc16  = crcf->create(crc_generator_poly::CRC16);
    c.crc24A = crcf->create(crc_generator_poly::CRC24A);
    c.crc24B = crcf->create(crc_generator_poly::CRC24B);
    return c;
  };
  pusch_codeblock_decoder::sch_crc cb_crcs;
  cb_crcs.crc16  = crcf->create(crc_generator_poly::CRC16);
  cb_crcs.crc24A = crcf->create(crc_generator_poly::CRC24A);
  cb_crcs.crc24B = crcf->create(crc_generator_poly::CRC24B);
  std::vector<std::unique_ptr<pusch_codeblock_decoder>> cbd;
  cbd.push_back(std::make_unique<pusch_codeblock_decoder>(create_ldpc_rate_dematcher_factory_sw("generic")->create(),
                                                          create_ldpc_decoder_factory_sw("generic")->create(), cb_crcs));
  auto dpool = std::make_shared<pusch_decoder_impl::codeblock_decoder_pool>(std::move(cbd));
  ctx->dec   = std::make_unique<pusch_decoder_impl>(create_ldpc_segmenter_rx_factory_sw()->create(), dpool, make_crcs(),
                                                  nullptr, MAX_RB, 4);
  return ctx;
}

void ref_tb_destroy(void* h)
{
  delete static_cast<ref_tb_ctx*>(h);
}

/// One transmission of a transport block. buffer_id selects the rx buffer (rnti << 8 | harq). tb_out (tbs / 8 bytes) is
/// written only when the reference writes its transport_block. info[0..5] = {tb_crc_ok, nof_codeblocks, iterations min,
/// iterations max, nof iteration observations, reservation ok}. Returns 0, or -1 when the buffer could not be reserved.
int ref_tb_decode(void*         h,
                  unsigned      buffer_id,
                  unsigned      tbs,
                  int           bg,
                  unsigned      rv,
                  unsigned      qm,
                  unsigned      Nref,
                  unsigned      nof_layers,
                  int           new_data,
                  unsigned      max_iter,
                  int           early_stop,
                  const int8_t* llrs,
                  unsigned      nof_llrs,
                  uint8_t*      tb_out,
                  unsigned*     info)
{
  ref_tb_ctx*          ctx = static_cast<ref_tb_ctx*>(h);
  ldpc_base_graph_type g   = (bg == 1) ? ldpc_base_graph_type::BG1 : ldpc_base_graph_type::BG2;
  unsigned             ncb = ldpc::compute_nof_codeblocks(units::bits(tbs), g);
  slot_point           slot(0, ctx->slot_count++ % 10240);
  unique_rx_buffer     buf = ctx->pool->get_pool().reserve(
      slot, trx_buffer_identifier(static_cast<uint16_t>(buffer_id >> 8), static_cast<uint8_t>(buffer_id & 0xff)), ncb, new_data != 0);
  info[5] = buf.is_valid() ? 1 : 0;
  if (!buf.is_valid()) {
    return -1;
  }
  pusch_decoder::configuration cfg;
  cfg.base_graph          = g;
  cfg.rv                  = rv;
  cfg.mod                 = mod_from_qm(qm);
  cfg.Nref                = Nref;
  cfg.nof_layers          = nof_layers;
  cfg.nof_ldpc_iterations = max_iter;
  cfg.use_early_stop      = early_stop != 0;
  cfg.new_data            = new_data != 0;
  ref_tb_notifier       notifier;
  pusch_decoder_buffer& in = ctx->dec->new_data(span<uint8_t>(tb_out, tbs / 8), std::move(buf), notifier, cfg);
  in.on_new_softbits(span<const log_likelihood_ratio>(reinterpret_cast<const log_likelihood_ratio*>(llrs), nof_llrs));
  in.on_end_softbits();
  info[0] = notifier.result.tb_crc_ok ? 1 : 0;
  info[1] = notifier.result.nof_codeblocks_total;
  info[2] = notifier.result.ldpc_decoder_stats.get_min();
  info[3] = notifier.result.ldpc_decoder_stats.get_max();
  info[4] = notifier.result.ldpc_decoder_stats.get_nof_observations();
  return notifier.notified ? 0 : -2;
}

/// Peeks at the rx buffer of a transport block between transmissions: codeblock CRC flags and soft bits (cb_len each).
/// Returns the number of codeblocks, or -1 when the buffer does not exist (released after a good TB CRC).
int ref_tb_peek(void* h, unsigned buffer_id, unsigned nof_cbs, unsigned cb_len, uint8_t* crc_flags, int8_t* soft)
{
  ref_tb_ctx*      ctx = static_cast<ref_tb_ctx*>(h);
  slot_point       slot(0, ctx->slot_count++ % 10240);
  unique_rx_buffer buf = ctx->pool->get_pool().reserve(
      slot, trx_buffer_identifier(static_cast<uint16_t>(buffer_id >> 8), static_cast<uint8_t>(buffer_id & 0xff)), nof_cbs, false);
  if (!buf.is_valid()) {
    return -1;
  }
  span<bool> crcs = buf->get_codeblocks_crc();
  for (unsigned i = 0; i != nof_cbs; ++i) {
    crc_flags[i] = crcs[i] ? 1 : 0;
    if (soft != nullptr) {
      span<log_likelihood_ratio> sb = buf->get_codeblock_soft_bits(i, cb_len);
      std::memcpy(soft + size_t(i) * cb_len, sb.data(), cb_len);
    }
  }
  buf.unlock();
  return static_cast<int>(nof_cbs);
}

// ---- rx buffer pool alone (differential test of the engine's HARQ buffer pool) ---------------------------------------------
namespace {
struct ref_pool_ctx {
  std::unique_ptr<rx_buffer_pool_controller> pool;
  unsigned                                   numerology = 0;
  std::map<unsigned, unique_rx_buffer>       held; // reservations handed out and not yet unlocked / released
  unsigned                                   next = 0;
};
} // namespace

void* ref_pool_create(unsigned nof_buffers, unsigned nof_codeblocks, unsigned expire_timeout_slots, unsigned numerology)
{
  auto                  ctx = new ref_pool_ctx();
  rx_buffer_pool_config pc;
  pc.max_codeblock_size   = ldpc::MAX_CODEBLOCK_SIZE;
  pc.nof_buffers          = nof_buffers;
  pc.nof_codeblocks       = nof_codeblocks;
  pc.expire_timeout_slots = expire_timeout_slots;
  pc.external_soft_bits   = true;
  ctx->pool               = create_rx_buffer_pool(pc);
  ctx->numerology         = numerology;
  return ctx;
}

void ref_pool_destroy(void* h)
{
  auto ctx = static_cast<ref_pool_ctx*>(h);
  ctx->held.clear();
  delete ctx;
}

/// rx_buffer_pool::reserve. Returns a handle >= 0 for a valid unique_rx_buffer, -1 for an invalid one.
int ref_pool_reserve(void* h, unsigned slot, unsigned id, unsigned nof_codeblocks, int new_data)
{
  auto             ctx = static_cast<ref_pool_ctx*>(h);
  unique_rx_buffer buf =
      ctx->pool->get_pool().reserve(slot_point(ctx->numerology, slot), trx_buffer_identifier((uint16_t)(id >> 8), (uint8_t)(id & 0xffu)), nof_codeblocks, new_data != 0);
  if (!buf.is_valid()) {
    return -1;
  }
  unsigned handle = ctx->next++;
  ctx->held.emplace(handle, std::move(buf));
  return (int)handle;
}

int ref_pool_codeblocks(void* h, unsigned handle, unsigned* ids)
{
  auto ctx = static_cast<ref_pool_ctx*>(h);
  auto it  = ctx->held.find(handle);
  if (it == ctx->held.end()) {
    return -1;
  }
  unsigned n = it->second.get().get_nof_codeblocks();
  for (unsigned i = 0; i != n; ++i) {
    ids[i] = it->second.get().get_absolute_codeblock_id(i);
  }
  return (int)n;
}

void ref_pool_unlock(void* h, unsigned handle)
{
  static_cast<ref_pool_ctx*>(h)->held.erase(handle); // the destructor unlocks
}

void ref_pool_release(void* h, unsigned handle)
{
  auto ctx = static_cast<ref_pool_ctx*>(h);
  auto it  = ctx->held.find(handle);
  if (it != ctx->held.end()) {
    it->second.release();
    ctx->held.erase(it);
  }
}

void ref_pool_run_slot(void* h, unsigned slot)
{
  auto ctx = static_cast<ref_pool_ctx*>(h);
  ctx->pool->get_pool().run_slot(slot_point(ctx->numerology, slot));
}

} // extern "C"

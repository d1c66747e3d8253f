// srsran-5g_b200/host/cuda_pdsch_enc.cpp -- see cuda_pdsch_enc.h. Host-only C++17; all device work is behind the C ABI.
#include "cuda_pdsch_enc.h"
#include "srsran/srsvec/bit.h"
#include "srsran/support/error_handling.h"
#include <cstring>

using namespace srsran;
using namespace srsran::b200;

namespace {
constexpr size_t MAX_RM_BYTES = 1u << 20; // per codeblock; grows on demand

uint32_t round4(uint32_t x)
{
  return (x + 3u) & ~3u;
}
} // namespace

cuda_pdsch_enc_accelerator::cuda_pdsch_enc_accelerator(std::shared_ptr<fec_engine_handle> engine_) :
  engine(std::move(engine_))
{
  in_cap  = (size_t)MAX_NOF_SEGMENTS * FEC_MAX_CB_IN_BYTES;
  out_cap = (size_t)MAX_NOF_SEGMENTS * 8192;
  h_in    = static_cast<uint8_t*>(fec_host_alloc(in_cap));
  h_out   = static_cast<uint8_t*>(fec_host_alloc(out_cap));
  report_fatal_error_if_not(h_in != nullptr && h_out != nullptr, "B200 FEC engine: cannot pin the encoder staging buffers.");
  descs.reserve(MAX_NOF_SEGMENTS);
  desc_cb.reserve(MAX_NOF_SEGMENTS);
}

cuda_pdsch_enc_accelerator::~cuda_pdsch_enc_accelerator()
{
  fec_host_free(h_in);
  fec_host_free(h_out);
}

void cuda_pdsch_enc_accelerator::free_queue()
{
  // Called once per transport block (pdsch_encoder_hw_impl.cpp:185).
  for (operation& op : ops) {
    op.state = op_state::idle;
  }
  descs.clear();
  desc_cb.clear();
  in_used = out_used = 0;
}

void cuda_pdsch_enc_accelerator::configure_operation(const hal::hw_pdsch_encoder_configuration& config, unsigned cb_index)
{
  srsran_assert(cb_index < MAX_NOF_SEGMENTS, "Codeblock index {} out of range.", cb_index);
  report_fatal_error_if_not(config.cb_mode && !config.attach_cb_crc && !config.attach_tb_crc,
                            "The CUDA PDSCH encoder serves the CB mode (CRCs attached by the segmenter).");
  ops[cb_index].cfg   = config;
  ops[cb_index].state = op_state::configured;
}

bool cuda_pdsch_enc_accelerator::enqueue_operation(span<const uint8_t> data, span<const uint8_t> /*aux_data*/, unsigned cb_index)
{
  operation& op = ops[cb_index];
  srsran_assert(op.state == op_state::configured, "enqueue_operation() of codeblock {} before configure_operation().", cb_index);
  const hal::hw_pdsch_encoder_configuration& c = op.cfg;
  const uint32_t out_len = round4((c.rm_length + 7) / 8);
  if (in_used + data.size() + 4 > in_cap) {
    return false; // the consumer drains what it has enqueued and retries
  }
  if (out_used + out_len > out_cap) {
    // The staging may only be replaced while no encoded codeblock is waiting in it for its dequeue_operation().
    bool busy = !descs.empty();
    for (const operation& o : ops) {
      busy = busy || (o.state == op_state::done);
    }
    if (busy) {
      return false;
    }
    out_used = 0;
    fec_host_free(h_out);
    out_cap = std::max<size_t>(2 * out_cap, out_len);
    h_out   = static_cast<uint8_t*>(fec_host_alloc(out_cap));
    report_fatal_error_if_not(h_out != nullptr && out_len <= MAX_RM_BYTES * 8, "B200 FEC engine: encoder output staging.");
  }
  fec_enc_desc d     = {};
  d.in_offset        = static_cast<uint32_t>(in_used);
  d.out_offset       = out_used;
  d.E                = c.rm_length;
  d.Nref             = c.Nref;
  d.lifting_size     = static_cast<uint16_t>(c.lifting_size);
  d.nof_filler_bits  = static_cast<uint16_t>(c.nof_filler_bits);
  d.base_graph       = static_cast<uint8_t>(c.base_graph_index == ldpc_base_graph_type::BG1 ? 1 : 2);
  d.rv               = static_cast<uint8_t>(c.rv);
  d.qm               = static_cast<uint8_t>(get_bits_per_symbol(c.modulation));
  const unsigned K   = (d.base_graph == 1 ? 22u : 10u) * c.lifting_size;
  const unsigned seg = (K - c.nof_filler_bits + 7) / 8;
  srsran_assert(data.size() == seg, "Codeblock {}: {} bytes given, the segment has {}.", cb_index, data.size(), seg);
  std::memcpy(h_in + in_used, data.data(), data.size());
  in_used += round4(static_cast<uint32_t>(data.size()));
  op.out_offset = static_cast<uint32_t>(out_used);
  out_used += out_len;
  descs.push_back(d);
  desc_cb.push_back(cb_index);
  op.state = op_state::pending;
  return true;
}

void cuda_pdsch_enc_accelerator::run_pending()
{
  if (descs.empty()) {
    return;
  }
  const int rc = fec_encode_run_host(engine->get(), descs.data(), static_cast<uint32_t>(descs.size()), h_in, in_used, h_out,
                                     out_used);
  report_fatal_error_if_not(rc == FEC_OK, "B200 FEC engine: {}", fec_last_error());
  for (unsigned cb : desc_cb) {
    ops[cb].state = op_state::done;
  }
  descs.clear();
  desc_cb.clear();
  in_used = 0; // the output staging keeps its contents until free_queue()
}

bool cuda_pdsch_enc_accelerator::dequeue_operation(span<uint8_t> data, span<uint8_t> aux_data, unsigned segment_index)
{
  operation& op = ops[segment_index];
  if (op.state == op_state::pending) {
    run_pending();
  }
  if (op.state != op_state::done) {
    return false;
  }
  const unsigned nbytes = (op.cfg.rm_length + 7) / 8;
  const uint8_t* packed = h_out + op.out_offset;
  if (!aux_data.empty()) {
    std::memcpy(aux_data.data(), packed, std::min<size_t>(nbytes, aux_data.size()));
  }
  if (op.cfg.do_unpack && !data.empty()) {
    // One bit per byte, as hw_accelerator_pdsch_enc delivers the codeblock when do_unpack is set.
    const unsigned n  = std::min<unsigned>(op.cfg.rm_length, data.size());
    bit_buffer     bb = bit_buffer::from_bytes(span<uint8_t>(h_out + op.out_offset, nbytes));
    srsvec::bit_unpack(data.first(n), bb.first(n));
  }
  op.state = op_state::idle;
  return true;
}

std::shared_ptr<hal::hw_accelerator_pdsch_enc_factory>
srsran::b200::create_cuda_pdsch_enc_acc_factory(const fec_engine_cfg& cfg)
{
  return std::make_shared<cuda_pdsch_enc_factory>(cfg);
}

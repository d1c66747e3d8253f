// fec_harq_pool.cpp -- HARQ buffer pool of the engine: which HARQ slots of the HBM soft-buffer
// store belong to which (UE, HARQ process), for how long, and who may touch them. Host-only C++; it drives the engine through
// its public entry points (fec_harq_free, fec_tb_reset_slots) and hands out the slot lists fec_tb_run_host takes.
//
// The behaviour is the reference's rx_buffer_pool, state for state:
//   rx_buffer_pool_impl::reserve / run_slot / stop     lib/phy/upper/rx_buffer_pool_impl.cpp:36-155
//   rx_buffer_impl (available / reserved / locked, reserve, expire, release, try_lock / unlock)
//                                                      lib/phy/upper/rx_buffer_impl.h:40-258
//   rx_buffer_codeblock_pool (FIFO free list of codeblock identifiers)   lib/phy/upper/rx_buffer_codeblock_pool.h:62-103
//   unique_rx_buffer (a reservation is handed out locked)                include/srsran/phy/upper/unique_rx_buffer.h:59-78
//   slot_point comparison with wrap-around                               include/srsran/ran/slot_point.h:160-174, 208-221
// including the corners: the identifier of a released buffer lingers until the next run_slot; a buffer whose codeblock
// reservation fails is emptied but keeps its identifier; the codeblock count of a retransmission is compared with the count
// of the last reservation even after a release.
#include "../../include/fec_b200.h"
#include <deque>
#include <mutex>
#include <new>
#include <vector>

namespace {

enum class buf_state : uint8_t { available, reserved, locked };

struct pool_buffer {
  buf_state             state = buf_state::available;
  std::vector<uint32_t> cb_ids;        // engine HARQ slots (absolute codeblock identifiers)
  uint32_t              nof_crc = 0;   // size of the CRC flag list: survives release, as rx_buffer_impl::crc does
};

constexpr uint32_t INVALID_ID   = 0xffffffffu;
constexpr int64_t  NO_EXPIRATION = -1;

} // namespace

struct fec_harq_pool {
  fec_engine*              engine = nullptr;
  fec_harq_pool_cfg        cfg    = {};
  uint32_t                 nof_slots_hyper = 0; // slots per hyper system frame: 1024 frames x 10 subframes x 2^numerology
  bool                     stopped = false;
  std::vector<uint32_t>    identifiers;
  std::vector<int64_t>     expirations;
  std::vector<pool_buffer> buffers;
  std::deque<uint32_t>     free_list;
  std::mutex               mtx;

  // slot_point::operator< / operator<= on slot counts
  bool slot_lt(uint32_t a, uint32_t b) const
  {
    const int v = (int)b - (int)a;
    if (v > 0) {
      return v < (int)nof_slots_hyper / 2;
    }
    return v < -(int)nof_slots_hyper / 2;
  }
  bool slot_le(uint32_t a, uint32_t b) const { return a == b || slot_lt(a, b); }

  void free_cb(uint32_t cb)
  {
    if (engine != nullptr) {
      // The slot reads as zeros for its next owner and forgets the codeblock's CRC flag / decoded message.
      fec_harq_free(engine, cb);
      fec_tb_reset_slots(engine, cb, 1);
    }
    free_list.push_back(cb);
  }

  void free_all(pool_buffer& b)
  {
    for (uint32_t cb : b.cb_ids) {
      free_cb(cb);
    }
    b.cb_ids.clear();
  }

  // rx_buffer_impl::reserve
  int reserve_cbs(pool_buffer& b, uint32_t nof_codeblocks, bool reset_crc)
  {
    if (b.state == buf_state::locked) {
      return FEC_POOL_IN_USE;
    }
    while (b.cb_ids.size() > nof_codeblocks) {
      const uint32_t cb = b.cb_ids.back();
      b.cb_ids.pop_back();
      free_cb(cb);
    }
    while (b.cb_ids.size() < nof_codeblocks) {
      if (free_list.empty()) {
        free_all(b);
        b.state = buf_state::available;
        return FEC_POOL_NO_CODEBLOCKS;
      }
      b.cb_ids.push_back(free_list.front());
      free_list.pop_front();
    }
    b.nof_crc = nof_codeblocks;
    if (reset_crc && engine != nullptr) {
      for (uint32_t cb : b.cb_ids) {
        fec_tb_reset_slots(engine, cb, 1);
      }
    }
    b.state = buf_state::reserved;
    return FEC_POOL_OK;
  }
};

extern "C" {

int fec_harq_pool_create(fec_engine* e, const fec_harq_pool_cfg* cfg, fec_harq_pool** out)
{
  if (cfg == nullptr || out == nullptr || cfg->nof_buffers == 0 || cfg->nof_codeblocks == 0 || cfg->numerology > 4) {
    return FEC_ERR_BAD_ARG;
  }
  fec_harq_pool* p = new (std::nothrow) fec_harq_pool;
  if (p == nullptr) {
    return FEC_ERR_NO_MEMORY;
  }
  p->engine          = e;
  p->cfg             = *cfg;
  p->nof_slots_hyper = 1024u * 10u * (1u << cfg->numerology);
  p->identifiers.assign(cfg->nof_buffers, INVALID_ID);
  p->expirations.assign(cfg->nof_buffers, NO_EXPIRATION);
  p->buffers.resize(cfg->nof_buffers);
  for (uint32_t i = 0; i < cfg->nof_codeblocks; ++i) {
    p->free_list.push_back(cfg->first_slot + i);
  }
  *out = p;
  return FEC_OK;
}

void fec_harq_pool_destroy(fec_harq_pool* p)
{
  delete p;
}

int fec_harq_pool_reserve(fec_harq_pool* p, uint32_t slot, uint32_t id, uint32_t nof_codeblocks, int new_data, uint32_t* buffer)
{
  if (p == nullptr || buffer == nullptr || id == INVALID_ID || slot >= p->nof_slots_hyper || nof_codeblocks > FEC_MAX_TB_CBS) {
    return FEC_ERR_BAD_ARG;
  }
  *buffer = FEC_POOL_INVALID_BUFFER;
  std::lock_guard<std::mutex> lk(p->mtx);
  if (p->stopped) {
    return FEC_POOL_STOPPED;
  }
  const uint32_t nb = (uint32_t)p->identifiers.size();
  uint32_t       i  = 0;
  while (i < nb && p->identifiers[i] != id) {
    ++i;
  }
  if (i == nb) {
    if (!new_data) {
      return FEC_POOL_RETX_UNKNOWN;
    }
    i = 0;
    while (i < nb && p->identifiers[i] != INVALID_ID) {
      ++i;
    }
  }
  if (i == nb) {
    return FEC_POOL_NO_BUFFER;
  }
  pool_buffer& b = p->buffers[i];
  if (!new_data && nof_codeblocks != b.nof_crc) {
    return FEC_POOL_CB_MISMATCH;
  }
  const int status = p->reserve_cbs(b, nof_codeblocks, new_data != 0);
  if (status != FEC_POOL_OK) {
    return status;
  }
  p->identifiers[i] = id;
  p->expirations[i] = (int64_t)((slot + p->cfg.expire_timeout_slots) % p->nof_slots_hyper);
  // unique_rx_buffer: the reservation is handed out locked
  b.state = buf_state::locked;
  *buffer = i;
  return FEC_POOL_OK;
}

int fec_harq_pool_codeblocks(fec_harq_pool* p, uint32_t buffer, uint32_t* slots, uint32_t max)
{
  if (p == nullptr || buffer >= p->buffers.size()) {
    return FEC_ERR_BAD_ARG;
  }
  std::lock_guard<std::mutex> lk(p->mtx);
  const pool_buffer&          b = p->buffers[buffer];
  for (uint32_t k = 0; k < b.cb_ids.size() && k < max && slots != nullptr; ++k) {
    slots[k] = b.cb_ids[k];
  }
  return (int)b.cb_ids.size();
}

int fec_harq_pool_unlock(fec_harq_pool* p, uint32_t buffer)
{
  if (p == nullptr || buffer >= p->buffers.size()) {
    return FEC_ERR_BAD_ARG;
  }
  std::lock_guard<std::mutex> lk(p->mtx);
  if (p->buffers[buffer].state != buf_state::locked) {
    return FEC_ERR_BAD_ARG; // "Failed to unlock. Invalid state."
  }
  p->buffers[buffer].state = buf_state::reserved;
  return FEC_OK;
}

int fec_harq_pool_release(fec_harq_pool* p, uint32_t buffer)
{
  if (p == nullptr || buffer >= p->buffers.size()) {
    return FEC_ERR_BAD_ARG;
  }
  std::lock_guard<std::mutex> lk(p->mtx);
  pool_buffer&                b = p->buffers[buffer];
  if (b.state != buf_state::locked) {
    return FEC_ERR_BAD_ARG; // "Failed to release. Invalid state."
  }
  p->free_all(b);
  b.state = buf_state::available;
  return FEC_OK;
}

int fec_harq_pool_run_slot(fec_harq_pool* p, uint32_t slot)
{
  if (p == nullptr || slot >= p->nof_slots_hyper) {
    return FEC_ERR_BAD_ARG;
  }
  std::lock_guard<std::mutex> lk(p->mtx);
  for (uint32_t i = 0; i < p->identifiers.size(); ++i) {
    if (p->identifiers[i] == INVALID_ID) {
      continue;
    }
    pool_buffer& b = p->buffers[i];
    if (b.state == buf_state::locked) {
      p->expirations[i] = (int64_t)((slot + p->cfg.expire_timeout_slots) % p->nof_slots_hyper);
      continue;
    }
    bool is_free = false;
    if (p->expirations[i] != NO_EXPIRATION && p->slot_le((uint32_t)p->expirations[i], slot)) {
      // rx_buffer_impl::expire: only a reserved buffer can expire
      if (b.state == buf_state::reserved) {
        b.state = buf_state::available;
        p->free_all(b);
        is_free = true;
      }
    } else {
      is_free = (b.state == buf_state::available);
    }
    if (is_free) {
      p->identifiers[i] = INVALID_ID;
      p->expirations[i] = NO_EXPIRATION;
    }
  }
  return FEC_OK;
}

int fec_harq_pool_stop(fec_harq_pool* p)
{
  if (p == nullptr) {
    return FEC_ERR_BAD_ARG;
  }
  std::lock_guard<std::mutex> lk(p->mtx);
  p->stopped = true;
  return FEC_OK;
}

int fec_harq_pool_nof_free_codeblocks(fec_harq_pool* p)
{
  if (p == nullptr) {
    return FEC_ERR_BAD_ARG;
  }
  std::lock_guard<std::mutex> lk(p->mtx);
  return (int)p->free_list.size();
}

} // extern "C"

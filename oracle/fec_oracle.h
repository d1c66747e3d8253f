/* oracle/fec_oracle.h -- TEST INFRASTRUCTURE (the parity oracle), NOT PRODUCT CODE.
 *
 * Plain-C CPU restatement of the reference's PUSCH FEC decode path. Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference leg may load it; the product path (srsran-5g_b200/csrc) never does.
 *
 * Parity status: PINNED. The reference ships no golden vectors for this path (its *.tar.gz fixtures are absent from the
 * checkout), so the oracle is pinned against the reference's own code compiled in the build container
 * (oracle/_ref/libref_fec.so, see oracle/Makefile) by tests/test_oracle_vs_ref.py, and against fixtures generated from
 * that build and committed under tests/golden/ (generator: tests/golden/make_golden.py).
 */
#ifndef FEC_ORACLE_H
#define FEC_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* CRC polynomial selectors (same numbering as hal::hw_dec_cb_crc_type would NOT be: this is oracle-local). */
enum { ORC_CRC24A = 0, ORC_CRC24B = 1, ORC_CRC16 = 2 };

typedef struct {
  int      bg;       /* 1 or 2 */
  int      Z;        /* lifting size */
  int      M;        /* base check rows (46/42) */
  int      Kb;       /* 22/10 */
  int      n_full;   /* 68/52 */
  int      n_edges;  /* 316/197 */
  uint8_t  deg[46];
  uint16_t row_start[47];
  uint8_t  col[316];
  uint16_t shift[316]; /* already reduced mod Z */
} orc_graph;

/* Builds the lifted graph; returns 0, or -1 if (bg, Z) is invalid. */
int orc_graph_init(orc_graph* g, int bg, int Z);

/* LLR algebra. */
int8_t orc_sat_add(int8_t x, int8_t y);
int8_t orc_promo(int8_t c2v, int8_t v2c);

/* CRC remainder over the first nbits of an MSB-first packed buffer. */
uint32_t orc_crc(int poly, const uint8_t* bytes, uint32_t nbits);

/* ldpc_rate_dematcher_impl::rate_dematch. soft: N = 66Z|50Z bytes, updated in place. Returns 0 / -1 on bad args. */
int orc_rate_dematch(int bg, int Z, int8_t* soft, const int8_t* in, uint32_t E, int new_data, uint32_t rv, uint32_t qm,
                     uint32_t Nref, uint32_t fillers);

/* ldpc_decoder_generic::decode. crc_poly < 0 means crc == nullptr. scale_lut: 128 entries (orc_make_scale_lut).
 * out: ceil(Kb*Z/8) bytes, MSB-first. Returns iterations (>=1) or -1 for nullopt. */
int orc_decode(int bg, int Z, uint8_t* out, const int8_t* llr, uint32_t n_llr, uint32_t fillers, int crc_poly,
               uint32_t max_iter, float scaling);

/* pusch_codeblock_decoder::decode: dematch + decode + CRC. Returns iterations when the CRC passes, else -1. */
int orc_cb_chain(int bg, int Z, int8_t* soft, const int8_t* in, uint32_t E, int new_data, uint32_t rv, uint32_t qm,
                 uint32_t Nref, uint32_t fillers, int crc_poly, uint32_t max_iter, int early_stop, uint8_t* out);

/* Scaling LUT: lut[x] = x for x > 120 else (int)roundf(x * sf). */
void orc_make_scale_lut(float sf, uint8_t lut[128]);

/* ---- transmit side / channel (fec_tx.c): input generation only ---- */
/* Appends the CRC of the first nbits (unpacked bits, one per byte) right after them; returns crc length. */
int orc_crc_attach_bits(int poly, uint8_t* bits, uint32_t nbits);
/* LDPC encode: msg = Kb*Z unpacked bits (fillers as 0) -> cb = 66Z|50Z unpacked bits (first 2Z punctured). */
int orc_encode(int bg, int Z, const uint8_t* msg, uint8_t* cb);
/* Rate match (bit selection from k0 skipping fillers + row-column interleave). cb fillers are identified by position. */
int orc_rate_match(int bg, int Z, const uint8_t* cb, uint8_t* out, uint32_t E, uint32_t rv, uint32_t qm, uint32_t Nref,
                   uint32_t fillers);
/* log_likelihood_ratio::quantize(v, range). */
int8_t orc_quantize(float v, float range);
/* BPSK-per-bit AWGN channel on E coded bits: y = (1-2b) + sigma*n, llr = quantize(2y/sigma^2, 20).
 * Gaussian noise from a xorshift64* + Box-Muller generator seeded by `seed` (deterministic, portable). */
void orc_awgn_llr(const uint8_t* bits, int8_t* llr, uint32_t E, float esn0_db, uint64_t seed);
/* Deterministic random bits. */
void orc_random_bits(uint8_t* bits, uint32_t n, uint64_t seed);

#ifdef __cplusplus
}
#endif
#endif

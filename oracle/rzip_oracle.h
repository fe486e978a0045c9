/* TEST INFRASTRUCTURE — CPU restatement (oracle) of modern-rzip's rzip pre-processor.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this.  The product (libmrz_gpu.so, the mrzip host binary) never links,
 * loads or calls anything in oracle/.
 *
 * Parity: PINNED.  tests/test_oracle_vs_reference.py runs the unmodified reference binary
 * (oracle/_ref/mrzip, built from /root/reference by oracle/Makefile.ref) on seeded inputs,
 * parses its archive, and compares stream 0 / stream 1 bytes with this restatement; the
 * committed fixtures under tests/golden/ were minted the same way (tests/golden/make_golden.py).
 *
 * Restates (file:line relative to /root/reference):
 *   src/rzip.c:65-73    levels[]
 *   src/rzip.c:161-227  put_match / put_literal token encoding
 *   src/rzip.c:229-328  hash table: insert_hash, clean_one_from_hash
 *   src/rzip.c:330-424  tags and match_len
 *   src/rzip.c:426-462  find_best_match
 *   src/rzip.c:507-667  hash_search
 *   src/rzip.c:669-673  init_hash_indexes (glibc TYPE_3 random(), unseeded)
 */
#ifndef RZIP_ORACLE_H
#define RZIP_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    int64_t matches, match_bytes, literals, literal_bytes, tag_hits, tag_misses, inserts;
    /* extra observables (not printed by the reference) used to steer tests */
    int64_t victim_events, displaced, minbit_overwrites, sweeps_wrapped, final_min_mask;
    int64_t lookups, slots_scanned;
} orc_stats;

/* one emitted token before splitting into <=0xFFFF pieces */
typedef struct {
    int64_t pos;  /* literal: start; match: start */
    int64_t len;
    int64_t ofs;  /* match: source offset (pos - dist); literal: -1 */
} orc_token;

/* rzip.c:65-73 */
void orc_level(int level, int64_t *mb_used, int *initial_freq, int *max_chain_len);

/* rzip.c:669-673 restated on glibc's TYPE_3 additive-feedback generator with seed 1.
 * skip = number of random() draws already consumed in the process (512 per earlier file). */
void orc_hash_index(int64_t hash_index[256], int64_t skip);

/* rzip.c:1006-1008 */
int orc_chunk_bytes(int64_t chunk_size);

uint32_t orc_crc32(uint32_t crc, const uint8_t *p, int64_t n);

/* hash_search for one chunk.
 *  s0/s1: caller buffers (capacity s0_cap / s1_cap); lengths returned in *s0_len / *s1_len.
 *  s0 includes the terminator and the 4 CRC bytes.
 *  tokens (optional, may be NULL): un-split token list, capacity tok_cap, count in *ntok.
 *  victim_round: in/out (static in rzip.c:259).
 * returns 0, or -1 when a buffer is too small. */
int orc_rzip_chunk(const uint8_t *buf, int64_t n, const int64_t hash_index[256], int level,
                   int chunk_bytes, int64_t *victim_round,
                   uint8_t *s0, int64_t s0_cap, int64_t *s0_len,
                   uint8_t *s1, int64_t s1_cap, int64_t *s1_len,
                   orc_token *tokens, int64_t tok_cap, int64_t *ntok,
                   orc_stats *stats);

#ifdef __cplusplus
}
#endif
#endif

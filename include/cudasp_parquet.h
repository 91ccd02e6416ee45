/* cudasp_parquet.h -- Parquet files straight into the pinned staging slots of the C ABI (libcudasp_b200.so).
 *
 * SURVEY.md 8f rank 1 names it: "read Parquet/Arrow fixed-size-binary(64) + list<int64> straight into pinned staging (e.g. a
 * sibling cudasp_scan_parquet(path, ...))".  The reference has no counterpart: its users go through DuckDB's read_parquet and
 * AccumulateInput (README.md:100-110, src/cudasp_extension.cu:154-224), and its README names that table scan as the scaling
 * limit (README.md:136-140).  Measured here on 8 B200: DuckDB's own decode of the four columns takes as long as the whole
 * scan from a DuckDB table -- so this entry decodes the pages itself (csrc/parquet_ingest.cpp: thrift footer,
 * data pages V1/V2, PLAIN and dictionary encodings, RLE/bit-packed levels) with one thread per row group in flight.  A page of
 * plain 64- or 33-byte keys without NULLs is staged into the pinned slots from where it lies in the file mapping; any other
 * page is decoded first and takes the Arrow ingest's path (the same NULL rules, key formats, piece cutting and slot recycling
 * as cudasp_arrow.h).  txid and height are read only for the rows that matched.  DESIGN.md 9.7: 100 M rows on 8 B200 in
 * 0.300 s against 0.416 s through read_parquet.
 *
 * Deliberately a SUBSET of Parquet: uncompressed or Snappy-compressed pages; the four columns
 *     txid BYTE_ARRAY | FIXED_LEN_BYTE_ARRAY, height INT32, tweak_key BYTE_ARRAY | FIXED_LEN_BYTE_ARRAY, outputs LIST<INT64>
 * found by name, else the first four top-level fields in that order (the reference takes its columns by position,
 * src/cudasp_extension.cu:159-163).  Anything else (other codecs, encodings, nested shapes, encrypted files) returns
 * CUDASP_ERR_UNSUPPORTED with the feature named in cudasp_last_error(): the caller then uses cudasp_scan over read_parquet.
 */
#ifndef CUDASP_PARQUET_H
#define CUDASP_PARQUET_H

#include <stdint.h>

#include "cudasp.h"

#ifdef __cplusplus
extern "C" {
#endif

#define CUDASP_ERR_UNSUPPORTED (-5) /* a Parquet feature outside the subset above; nothing was scanned */
#define CUDASP_ERR_IO (-6)          /* the file cannot be opened / mapped / is not a Parquet file */

/* The rows that matched, in file order then row order.  txid i = txids[txid_offsets[i] .. txid_offsets[i + 1]), likewise the keys
 * (64 or 33 bytes each, as stored).  Owned by the library: cudasp_parquet_result_free. */
typedef struct cudasp_parquet_result {
  uint64_t n_matches;
  uint64_t rows_scanned;        /* rows of all files, including rows the NULL rules drop */
  const uint32_t *file_index;   /* n_matches: index into `paths` */
  const uint64_t *row_in_file;  /* n_matches */
  const int32_t *heights;       /* n_matches */
  const uint8_t *txids;
  const uint64_t *txid_offsets; /* n_matches + 1 */
  const uint8_t *tweak_keys;
  const uint64_t *key_offsets;  /* n_matches + 1 */
} cudasp_parquet_result;

/* Scans every row of the files with the query's keys.  Row rules are the reference's AccumulateInput (src/cudasp_extension.cu:
 * 195-221): a row with a NULL txid, height or tweak_key is dropped; a NULL outputs list is an empty list; NULL list children are
 * skipped; a tweak_key that is neither 64 nor 33 bytes can never match.  `threads` host threads decode row groups side by side
 * (0 = one per host core, at most one per row group); every function of the handle stays usable from other threads meanwhile. */
int cudasp_scan_parquet(cudasp_query *q, const char *const *paths, int n_paths, int threads, cudasp_parquet_result **out);
void cudasp_parquet_result_free(cudasp_parquet_result *r);

#ifdef __cplusplus
}
#endif
#endif /* CUDASP_PARQUET_H */

/* splice_oracle.c — plain-C restatement of the byte/integer part of the hot path (TEST INFRASTRUCTURE ONLY).
 * Never linked into the product library; used by tests and by bench.py's cpu_baseline leg as the checker /
 * reported CPU baseline of the splice.
 *   oracle_attach_positions : row-major positions of attachment tokens
 *       == batch_idx*S + token_range built by  src/multimeditron/model/data_loader.py:124-141 from
 *          src/multimeditron/model/prompt_tokenizers.py:255-286 (SURVEY.md §3.4 invariant)
 *   oracle_embed_rows       : out[r] = table[ids[r]]                      src/multimeditron/model/model.py:433
 *   oracle_splice_rows      : dest[rows[k]] = src[k] (same dtype)         src/multimeditron/model/model.py:441-444
 */
#include <stdint.h>
#include <string.h>

int64_t oracle_attach_positions(const int64_t* ids, int64_t n, int64_t attach, int32_t* dest, int32_t* sop) {
  int64_t k = 0;
  for (int64_t i = 0; i < n; ++i) {
    if (ids[i] == attach) {
      if (dest) dest[k] = (int32_t)i;
      if (sop) sop[i] = (int32_t)k;
      ++k;
    } else if (sop) {
      sop[i] = -1;
    }
  }
  return k;
}

int oracle_embed_rows(uint8_t* out, const uint8_t* table, int64_t vocab, const int64_t* ids, int64_t n,
                      int64_t row_bytes) {
  for (int64_t i = 0; i < n; ++i) {
    if (ids[i] < 0 || ids[i] >= vocab) return 1;
    memcpy(out + i * row_bytes, table + ids[i] * row_bytes, (size_t)row_bytes);
  }
  return 0;
}

int oracle_splice_rows(uint8_t* dest, int64_t n_dest_rows, const uint8_t* src, const int32_t* rows, int64_t n_rows,
                       int64_t row_bytes) {
  for (int64_t k = 0; k < n_rows; ++k) {
    if (rows[k] < 0 || rows[k] >= n_dest_rows) return 1;
    memcpy(dest + (int64_t)rows[k] * row_bytes, src + k * row_bytes, (size_t)row_bytes);
  }
  return 0;
}

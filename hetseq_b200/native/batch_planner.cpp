// hetseq_b200 native host library: batch planning.
//
// Replaces the reference's only native component, the Cython extension
// data/data_utils_fast.pyx:21-62 (batch_by_size_fast). Same batching rule, but sizes arrive as one
// contiguous array instead of a Python callback per sample (the reference makes N Python calls --
// 6.8 M for the full BERT corpus), and batches are returned as offsets into the index order
// (every batch is a contiguous slice of it) instead of a list of Python lists.
#include <algorithm>
#include <cstdint>

#define HS_EXPORT extern "C" __attribute__((visibility("default")))

// sizes[i] is the token count of the sample at position i of the ordered index list.
// out_offsets must hold n+1 entries; batch b covers positions [out_offsets[b], out_offsets[b+1]).
// Returns the number of batches, or -1 if the sample at *bad_pos exceeds max_tokens.
HS_EXPORT int64_t hs_batch_by_size(const int64_t* sizes, int64_t n, int64_t max_tokens, int64_t max_sentences,
                                   int64_t bsz_mult, int64_t* out_offsets, int64_t* bad_pos) {
  if (bsz_mult < 1) bsz_mult = 1;
  int64_t nb = 0;
  int64_t start = 0;       // first position of the open batch
  int64_t sample_len = 0;  // max size inside the open batch (including carried samples)
  out_offsets[0] = 0;
  for (int64_t i = 0; i < n; ++i) {
    const int64_t sz = sizes[i];
    sample_len = std::max(sample_len, sz);
    if (sample_len > max_tokens) {
      if (bad_pos) *bad_pos = i;
      return -1;
    }
    const int64_t len = i - start;  // samples already in the open batch
    const bool full = len > 0 && (len == max_sentences || (len + 1) * sample_len > max_tokens);
    if (full) {
      const int64_t mod_len = std::max(bsz_mult * (len / bsz_mult), len % bsz_mult);
      start += mod_len;
      out_offsets[++nb] = start;
      // the running max restarts over the carried samples [start, i) and the sample being added
      // (data_utils_fast.pyx:40-41,56-57: its length was appended to sample_lens before the flush)
      sample_len = 0;
      for (int64_t j = start; j <= i; ++j) sample_len = std::max(sample_len, sizes[j]);
    }
  }
  if (n - start > 0) out_offsets[++nb] = n;
  return nb;
}

// Round-robin sharding of `num_batches` batch ids over `num_shards` consumers, padded with -1 so that
// every shard sees ceil(num_batches / num_shards) entries (reference: data/iterators.py:244-275).
// `order` (may be null) is a permutation applied first (the per-epoch shuffle).
HS_EXPORT int64_t hs_shard_plan(const int64_t* order, int64_t num_batches, int64_t num_shards, int64_t shard_id,
                                int64_t* out_ids) {
  if (num_shards < 1 || shard_id < 0 || shard_id >= num_shards) return -1;
  const int64_t per = (num_batches + num_shards - 1) / num_shards;
  for (int64_t k = 0; k < per; ++k) {
    const int64_t pos = shard_id + k * num_shards;
    out_ids[k] = pos < num_batches ? (order ? order[pos] : pos) : -1;
  }
  return per;
}

// N1: host input pipeline text -> token ids.  HOST code only (no kernels): the reference turns every text
// into a [T,300] float tensor on the CPU, one word2vec lookup per word (SimpleDataset.text_to_embedding,
// simple_two_tower.py:88-110 == enhanced_two_tower.py:144-166), and ships 786 MB of floats per step over PCIe at the
// bench shape; here the host only produces the ids (2.6 MB per step) and the gather runs on the device (K-A).
//
// Semantics kept from the reference: `words = text.lower().split()[:max_len]` (truncation happens BEFORE the
// vocabulary filter), words missing from the vocabulary are skipped, the row is padded at the END (the reference pads
// with zero vectors; pad_id is the table row that reads as zeros).  A text with no known word is all padding.
// Splitting follows Python's str.split(): any run of str.isspace() characters, recognised here in UTF-8.  ASCII letters
// are lower-cased here; the Python wrapper lower-cases the whole string first, which covers non-ASCII case folding.
#include <algorithm>
#include <string>
#include <string_view>
#include <thread>
#include <unordered_map>
#include <vector>

#include "common.cuh"

struct tt_vocab {
  std::string blob;
  std::unordered_map<std::string_view, int32_t> map;
};

namespace tt {

// length in bytes of the whitespace character starting at p (0 if none): Python's str.isspace() set
static inline int space_len(const unsigned char* p, const unsigned char* end) {
  const unsigned char c = p[0];
  if (c == 0x20 || (c >= 0x09 && c <= 0x0D) || (c >= 0x1C && c <= 0x1F)) return 1;
  if (c < 0xC2) return 0;
  if (c == 0xC2 && p + 1 < end && (p[1] == 0x85 || p[1] == 0xA0)) return 2;
  if (p + 2 < end) {
    if (c == 0xE1 && p[1] == 0x9A && p[2] == 0x80) return 3;                                   // U+1680
    if (c == 0xE2 && p[1] == 0x80 && ((p[2] >= 0x80 && p[2] <= 0x8A) || p[2] == 0xA8 || p[2] == 0xA9 || p[2] == 0xAF))
      return 3;                                                                                // U+2000-200A, 2028, 2029, 202F
    if (c == 0xE2 && p[1] == 0x81 && p[2] == 0x9F) return 3;                                   // U+205F
    if (c == 0xE3 && p[1] == 0x80 && p[2] == 0x80) return 3;                                   // U+3000
  }
  return 0;
}

template <typename IdT>
static void tokenize_range(const tt_vocab* v, const char* blob, const int64_t* offs, int64_t t0, int64_t t1, int max_len,
                           int64_t pad_id, int mode, IdT* out, int32_t* lens) {
  std::string word;
  for (int64_t t = t0; t < t1; ++t) {
    const unsigned char* p = reinterpret_cast<const unsigned char*>(blob + offs[t]);
    const unsigned char* end = reinterpret_cast<const unsigned char*>(blob + offs[t + 1]);
    IdT* row = out + t * (int64_t)max_len;
    int n = 0, words = 0;
    while (p < end && n < max_len) {
      int sl;
      while (p < end && (sl = space_len(p, end)) > 0) p += sl;
      if (p >= end) break;
      word.clear();
      while (p < end && space_len(p, end) == 0) {
        unsigned char c = *p++;
        if (c >= 'A' && c <= 'Z') c += 32;
        word.push_back((char)c);
      }
      if (mode == 0 && words >= max_len) break;          // reference: only the first max_len WORDS are looked at
      ++words;
      auto it = v->map.find(std::string_view(word));
      if (it != v->map.end()) row[n++] = (IdT)it->second;
    }
    if (lens) lens[t] = n;
    for (; n < max_len; ++n) row[n] = (IdT)pad_id;
  }
}

}  // namespace tt

extern "C" {

tt_vocab* tt_vocab_create(const char* words_blob, const int64_t* word_offsets, int64_t nwords, const int32_t* ids) {
  if (!words_blob || !word_offsets || nwords < 0) { tt::set_error("vocab_create: bad args"); return nullptr; }
  tt_vocab* v = new tt_vocab();
  v->blob.assign(words_blob, (size_t)word_offsets[nwords]);
  v->map.reserve((size_t)nwords * 2);
  for (int64_t i = 0; i < nwords; ++i) {
    std::string_view key(v->blob.data() + word_offsets[i], (size_t)(word_offsets[i + 1] - word_offsets[i]));
    v->map.emplace(key, ids ? ids[i] : (int32_t)i);      // duplicate words: the first occurrence wins
  }
  return v;
}

void tt_vocab_destroy(tt_vocab* v) { delete v; }

int64_t tt_vocab_size(const tt_vocab* v) { return v ? (int64_t)v->map.size() : 0; }

int tt_tokenize_batch(const tt_vocab* v, const char* text_blob, const int64_t* text_offsets, int64_t ntexts, int32_t max_len,
                      int64_t pad_id, int32_t mode, int32_t ids_are_i64, void* out_ids, int32_t* out_lengths,
                      int32_t nthreads) {
  TT_CHECK_ARG(v && text_blob && text_offsets && out_ids && ntexts >= 0 && max_len > 0, "tokenize_batch: bad args");
  TT_CHECK_ARG(mode == 0 || mode == 1, "tokenize_batch: mode must be 0 (reference) or 1 (filter, then truncate)");
  if (ntexts == 0) return TT_OK;
  int nt = nthreads > 0 ? nthreads : (int)std::thread::hardware_concurrency();
  nt = (int)std::max<int64_t>(1, std::min<int64_t>(nt, ntexts / 64 + 1));
  auto work = [&](int64_t a, int64_t b) {
    if (ids_are_i64) tt::tokenize_range<int64_t>(v, text_blob, text_offsets, a, b, max_len, pad_id, mode, (int64_t*)out_ids, out_lengths);
    else tt::tokenize_range<int32_t>(v, text_blob, text_offsets, a, b, max_len, pad_id, mode, (int32_t*)out_ids, out_lengths);
  };
  if (nt == 1) { work(0, ntexts); return TT_OK; }
  std::vector<std::thread> pool;
  const int64_t per = (ntexts + nt - 1) / nt;
  for (int i = 0; i < nt; ++i) {
    const int64_t a = i * per, b = std::min<int64_t>(ntexts, a + per);
    if (a < b) pool.emplace_back(work, a, b);
  }
  for (auto& th : pool) th.join();
  return TT_OK;
}

}  // extern "C"

// Sorted distinct ids of a batch and the position of every id among them (torch.unique(return_inverse=True)), the way
// the GPU engine's dedupe kernel does it: through a table with one entry per vocabulary row instead of a sort of all
// B x F ids.  Only the distinct ids (a quarter of the batch for Criteo-shaped data) are sorted.
//
// scratch[V] must be all zero on entry and is all zero again on return (only touched entries are reset), so the caller
// keeps one per vocabulary size.  Returns the number of distinct ids, or -1 - i if ids[i] lies outside [0, V).
#include <algorithm>
#include <cstdint>

namespace {

// ascending sort of n distinct ids < V: std::sort for a few thousand, else an LSD radix sort with 11-bit digits
// (two passes for a Criteo-sized vocabulary) -- a batch of mostly distinct ids has tens of thousands of them
void sort_ids(int64_t* a, int64_t n, int64_t V, int64_t* tmp) {
  if (n < 4096) {
    std::sort(a, a + n);
    return;
  }
  int bits = 1;
  while ((int64_t(1) << bits) < V) ++bits;
  int64_t* src = a;
  int64_t* dst = tmp;
  for (int shift = 0; shift < bits; shift += 11) {
    int64_t count[2049] = {0};
    for (int64_t i = 0; i < n; ++i) ++count[((src[i] >> shift) & 2047) + 1];
    for (int d = 0; d < 2048; ++d) count[d + 1] += count[d];
    for (int64_t i = 0; i < n; ++i) dst[count[(src[i] >> shift) & 2047]++] = src[i];
    std::swap(src, dst);
  }
  if (src != a) std::copy(src, src + n, a);
}

}  // namespace

extern "C" int64_t dfcpu_unique_inverse(const int64_t* ids, int64_t n, int64_t V, int32_t* scratch, int64_t* uniq,
                                        int64_t* inverse) {
  int64_t u = 0;
  for (int64_t i = 0; i < n; ++i) {
    const int64_t id = ids[i];
    if (id < 0 || id >= V) {
      for (int64_t j = 0; j < u; ++j) scratch[uniq[j]] = 0;
      return -1 - i;
    }
    if (!scratch[id]) {
      scratch[id] = 1;
      uniq[u++] = id;
    }
  }
  sort_ids(uniq, u, V, inverse);          // `inverse` is not written yet: it serves as the sort's second buffer
  for (int64_t j = 0; j < u; ++j) scratch[uniq[j]] = static_cast<int32_t>(j + 1);
  for (int64_t i = 0; i < n; ++i) inverse[i] = scratch[ids[i]] - 1;
  for (int64_t j = 0; j < u; ++j) scratch[uniq[j]] = 0;
  return u;
}

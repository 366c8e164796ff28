// Cross-sample merge of per-sample context tables into one tract vector (host side of the genome_set step).
//
// Restates g_tract_vector_concatenate_tracts() (reference src/genome_set.c:250-289), the tract_id chain of
// new_g_tract_vector_from_genomic_context_list() (:208-222) and context_histograms_overlap()
// (src/context_histogram.c:87-110) on flat arrays, as they arrive from the allgather of the per-sample tables
// (tatajuba_b200/genome_set.py).  The work is a sequential chain over 10^4..10^6 table rows (every tract id depends
// on the previous group's first element), so it stays on the host like the mapper it follows; no device code here.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "tjb_gpu.h"

namespace {

// edit distance with unit costs where leading / trailing overhangs of either string are free at the END of the
// alignment: the minimum over the last row and the last column of the DP table
// (biomcmc_levenshtein_distance(..., 1, 1, skip_borders = true), submodules/biomcmc-lib/lib/lowlevel.c:112-142)
uint32_t border_free_levenshtein (const char *a, uint32_t na, const char *b, uint32_t nb, std::vector<uint32_t> &tab)
{
  const uint32_t cols = nb + 1;
  tab.assign ((size_t) (na + 1) * cols, 0u);
  for (uint32_t i = 0; i <= na; i++) tab[(size_t) i * cols] = i;
  for (uint32_t j = 0; j <= nb; j++) tab[j] = j;
  for (uint32_t i = 1; i <= na; i++) {
    const uint32_t *up = &tab[(size_t) (i - 1) * cols];
    uint32_t *row = &tab[(size_t) i * cols];
    for (uint32_t j = 1; j <= nb; j++) {
      uint32_t best = up[j - 1] + (a[i - 1] == b[j - 1] ? 0u : 1u);
      best = std::min (best, row[j - 1] + 1u);
      best = std::min (best, up[j] + 1u);
      row[j] = best;
    }
  }
  uint32_t d = tab[(size_t) na * cols + nb];
  for (uint32_t i = 0; i <= na; i++) d = std::min (d, tab[(size_t) i * cols + nb]);
  for (uint32_t j = 0; j <= nb; j++) d = std::min (d, tab[(size_t) na * cols + j]);
  return d;
}

struct Tables {
  const int32_t *location, *base, *mode_len;
  const char *names;
  int32_t stride, lev;
};

// do rows x (the later one) and y (the first of the current tract) describe the same tract?
bool rows_overlap (const Tables &t, int32_t x, int32_t y, std::vector<uint32_t> &tab, std::vector<char> &yname)
{
  const int32_t apart = std::abs (t.location[x] - t.location[y]);
  if (!apart) return true;
  const int32_t tract = std::min (t.mode_len[x], t.mode_len[y]);
  if (apart > tract - 1 || t.base[x] != t.base[y]) return false;
  // the reference measures the first name and uses that length for both (context_histogram.c:104-105)
  const char *xn = t.names + (size_t) x * t.stride;
  const uint32_t n = (uint32_t) strnlen (xn, (size_t) t.stride);
  yname.assign (n, '\0');
  const char *yn = t.names + (size_t) y * t.stride;
  memcpy (yname.data (), yn, std::min ((size_t) n, strnlen (yn, (size_t) t.stride)));
  return border_free_levenshtein (xn, n, yname.data (), n, tab) <= (uint32_t) t.lev;
}

} // namespace

extern "C" int tjb_merge_tracts (int32_t n_samples, const int32_t *n_hist, const int32_t *location, const int32_t *base,
                                 const int32_t *mode_len, const char *names, int32_t name_stride, int32_t levenshtein_distance,
                                 int32_t *out_sample, int32_t *out_index, int32_t *out_tract_id)
{
  if (n_samples < 0 || (n_samples && !n_hist) || name_stride < 1) return TJB_ERR_ARG;
  std::vector<int32_t> first (n_samples + 1, 0);
  for (int32_t s = 0; s < n_samples; s++) { if (n_hist[s] < 0) return TJB_ERR_ARG; first[s + 1] = first[s] + n_hist[s]; }
  const int32_t total = first[n_samples];
  if (!total) return 0;
  if (!location || !base || !mode_len || !names || !out_sample || !out_index || !out_tract_id) return TJB_ERR_ARG;

  // 1. percolate the samples into one vector ordered by location.  Samples are merged one after the other into the
  //    accumulated vector; on equal locations the NEW sample's row goes first, then ONE accumulated row
  //    (genome_set.c:274-281), which is not what a stable sort by (location, sample) would give.
  std::vector<int32_t> acc (total), next (total);
  int32_t n_acc = n_hist[0];
  for (int32_t j = 0; j < n_acc; j++) acc[j] = j;
  for (int32_t s = 1; s < n_samples; s++) {
    int32_t a = 0, b = first[s], o = 0;
    const int32_t b_end = first[s + 1];
    while (a < n_acc && b < b_end) {
      const int32_t la = location[acc[a]], lb = location[b];
      if (la > lb) next[o++] = b++;
      else if (la < lb) next[o++] = acc[a++];
      else { next[o++] = b++; next[o++] = acc[a++]; }
    }
    while (b < b_end) next[o++] = b++;
    while (a < n_acc) next[o++] = acc[a++];
    acc.swap (next);
    n_acc = o;
  }

  // 2. tract ids: every row is compared with the FIRST row of the current tract (genome_set.c:208-222)
  const Tables t = {location, base, mode_len, names, name_stride, levenshtein_distance};
  std::vector<uint32_t> tab;
  std::vector<char> yname;
  std::vector<int32_t> sample_of (total);
  for (int32_t s = 0; s < n_samples; s++) for (int32_t j = first[s]; j < first[s + 1]; j++) sample_of[j] = s;
  int32_t prev = 0, id = 0;
  for (int32_t i = 0; i < n_acc; i++) {
    if (i && !rows_overlap (t, acc[i], acc[prev], tab, yname)) { id++; prev = i; }
    out_sample[i] = sample_of[acc[i]];
    out_index[i] = acc[i] - first[sample_of[acc[i]]];
    out_tract_id[i] = id;
  }
  return n_acc;
}

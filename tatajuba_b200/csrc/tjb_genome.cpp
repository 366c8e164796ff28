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
#include <memory>
#include <string>
#include <thread>
#include <utility>
#include <vector>
#if defined(__linux__)
#include <sched.h>
#endif

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

  // 1. percolate the samples into one vector ordered by location.  The reference merges the samples one after the
  //    other into the accumulated vector; on equal locations the NEW sample's row goes first, then ONE accumulated row
  //    (genome_set.c:274-281), which is not what a stable sort by (location, sample) would give — and costs
  //    n_samples x total row moves (0.15 s for 96 samples of 20 K rows).  The rule only ever reorders rows of EQUAL
  //    location, and what it does to such a group does not depend on the rest: so the rows are first merged by location
  //    alone (a stable radix sort: rows of equal location stay in (sample, index) order), and then
  //    every group of equal location gets the order the sequential merges would have left it in — the reverse of the
  //    sample order when every sample has at most one row there (the usual case), the literal interleaving otherwise.
  //    Rows carry their location and sample along, so that nothing after the sort reads the input tables at random.
  //    Many samples of one genome (dense locations) are handled in SLICES of the location range by a few threads: the
  //    sort and the tie rule never look across a location boundary, and the tract-id chain (step 2) can be cut wherever
  //    two neighbouring rows are at least max (mode_len) apart: such a row cannot overlap the first row of any earlier
  //    tract (rows_overlap() needs a distance below the shorter of the two mode lengths), so it starts a tract whatever
  //    came before.
  struct Row { uint32_t loc; int32_t sample, row, pad; };
  int workers = (int) std::thread::hardware_concurrency ();
#if defined(__linux__)
  cpu_set_t set;
  if (sched_getaffinity (0, sizeof (set), &set) == 0 && CPU_COUNT (&set) > 0) workers = std::min (workers > 0 ? workers : 1 << 20, CPU_COUNT (&set));
#endif
  if (const char *e = getenv ("TJB_MERGE_THREADS")) workers = atoi (e);
  workers = std::max (1, std::min (std::min (workers, 32), total / 4096));
  auto parallel_for = [] (int n, auto &&fn) { // fn (0 .. n-1), 0 on this thread
    std::vector<std::thread> th;
    for (int r = 1; r < n; r++) th.emplace_back ([&fn, r] { fn (r); });
    fn (0);
    for (auto &t : th) t.join ();
  };
  // the two row arrays (16 bytes per row each) are kept by the calling thread from one call to the next: on fresh
  // memory, page faults were 40 % of a merge of 1.9 M rows
  struct Scratch { void *p = nullptr; size_t cap = 0; ~Scratch () { free (p); } };
  static thread_local Scratch scratch;
  const size_t need = 2 * sizeof (Row) * (size_t) total;
  if (scratch.cap < need) {
    free (scratch.p);
    scratch.p = malloc (need + need / 8);
    scratch.cap = scratch.p ? need + need / 8 : 0;
    if (!scratch.p) return TJB_ERR_NOMEM;
  }
  Row *acc = (Row*) scratch.p, *other = acc + total;
  // rows with their (biased: locations may be negative, nothing forbids it) location; every worker takes a share of
  // the rows (and is the first to touch that part of both arrays)
  struct Seen { uint32_t differ = 0, lo = 0xffffffffu, hi = 0; int32_t max_mode_len = 0; bool ascending = true; };
  std::vector<Seen> seen (workers);
  const uint32_t loc0 = (uint32_t) location[0] ^ 0x80000000u;
  parallel_for (workers, [&] (int w) {
    const int32_t j0 = (int32_t) ((int64_t) total * w / workers), j1 = (int32_t) ((int64_t) total * (w + 1) / workers);
    int32_t s = (int32_t) (std::upper_bound (first.begin (), first.end (), j0) - first.begin ()) - 1;
    Seen v;
    for (int32_t j = j0; j < j1; j++) {
      while (j >= first[s + 1]) s++;
      acc[j].loc = (uint32_t) location[j] ^ 0x80000000u; acc[j].sample = s; acc[j].row = j; acc[j].pad = 0;
      v.differ |= acc[j].loc ^ loc0;
      v.lo = std::min (v.lo, acc[j].loc); v.hi = std::max (v.hi, acc[j].loc);
      v.max_mode_len = std::max (v.max_mode_len, mode_len[j]);
      if (j > first[s] && location[j] < location[j - 1]) v.ascending = false;
      other[j].loc = 0; // first touch
    }
    seen[w] = v;
  });
  uint32_t differ = 0, lo = 0xffffffffu, hi = 0;
  int32_t max_mode_len = 0;
  bool ascending = true; // within every sample, as the reference's merge expects (genome_set.c:274)
  for (const Seen &v : seen) { differ |= v.differ; lo = std::min (lo, v.lo); hi = std::max (hi, v.hi); max_mode_len = std::max (max_mode_len, v.max_mode_len); ascending = ascending && v.ascending; }
  const bool dense = ascending && (uint64_t) (hi - lo) <= 8ull * (uint64_t) total + 1024;

  // slices: [bound[r], bound[r + 1]) in biased locations, rows [slice_row[r], slice_row[r + 1]) of the sorted vector
  const int n_slices = dense ? workers : 1;
  std::vector<uint64_t> bound (n_slices + 1);
  for (int r = 0; r <= n_slices; r++) bound[r] = (uint64_t) lo + ((uint64_t) (hi - lo) + 1) * (uint64_t) r / (uint64_t) n_slices;
  std::vector<int32_t> slice_row (n_slices + 1, 0), cut ((size_t) (n_slices + 1) * (size_t) std::max (n_samples, 1));
  if (dense) for (int r = 0; r <= n_slices; r++) { // where every sample's rows cross the slice boundaries
    int32_t rows = 0;
    for (int32_t s = 0; s < n_samples; s++) {
      const Row *b0 = acc + first[s], *b1 = acc + first[s + 1];
      const Row *c = std::partition_point (b0, b1, [&] (const Row &x) { return (uint64_t) x.loc < bound[r]; });
      cut[(size_t) r * n_samples + s] = (int32_t) (c - acc);
      rows += (int32_t) (c - b0);
    }
    slice_row[r] = rows;
  }

  // sort + tie rule of the rows [r0, r1) of the sorted vector (a whole number of location groups)
  auto tie_rule = [&] (Row *v, int32_t r0, int32_t r1) {
    std::vector<Row> grp, merged;
    for (int32_t g0 = r0; g0 < r1; ) {
      int32_t g1 = g0 + 1;
      bool twice = false; // some sample has two rows at this location
      while (g1 < r1 && v[g1].loc == v[g0].loc) { twice = twice || v[g1].sample == v[g1 - 1].sample; g1++; }
      if (!twice) std::reverse (v + g0, v + g1);
      else { // the sequential merges, on this group alone: new sample's row, then one accumulated row, while both have some
        grp.clear ();
        int32_t i = g0;
        while (i < g1) {
          int32_t e = i + 1;
          while (e < g1 && v[e].sample == v[i].sample) e++;
          merged.clear ();
          int32_t a = 0, b = i;
          const int32_t na = (int32_t) grp.size ();
          while (a < na && b < e) { merged.push_back (v[b++]); merged.push_back (grp[a++]); }
          while (b < e) merged.push_back (v[b++]);
          while (a < na) merged.push_back (grp[a++]);
          grp.swap (merged);
          i = e;
        }
        std::copy (grp.begin (), grp.end (), v + g0);
      }
      g0 = g1;
    }
  };
  // one slice of the dense case: a counting sort by location, done block of locations by block of locations with one
  // cursor per sample, so that the part of the output being written (~128 KB) stays in cache (sample after sample over
  // the whole range would touch every line of the output once per sample)
  auto sort_slice = [&] (int r) {
    const uint64_t b_lo = bound[r], range = bound[r + 1] - bound[r];
    const int32_t r0 = slice_row[r], r1 = slice_row[r + 1];
    if (r1 == r0) return;
    std::vector<uint32_t> start ((size_t) range + 1, 0u);
    std::vector<int32_t> cursor (cut.begin () + (size_t) r * n_samples, cut.begin () + (size_t) (r + 1) * n_samples);
    const int32_t *stop = cut.data () + (size_t) (r + 1) * n_samples;
    for (int32_t s = 0; s < n_samples; s++) for (int32_t j = cursor[s]; j < stop[s]; j++) start[(uint64_t) acc[j].loc - b_lo + 1]++;
    start[0] = (uint32_t) r0;
    for (size_t c = 1; c < start.size (); c++) start[c] += start[c - 1];
    const uint64_t block = std::max<uint64_t> (1, range * 8192 / (uint64_t) (r1 - r0));
    for (uint64_t b0 = 0; b0 < range; b0 += block) {
      const uint64_t limit = b_lo + std::min (range, b0 + block); // exclusive
      for (int32_t s = 0; s < n_samples; s++) {
        int32_t j = cursor[s];
        for (; j < stop[s] && (uint64_t) acc[j].loc < limit; j++) other[start[(uint64_t) acc[j].loc - b_lo]++] = acc[j];
        cursor[s] = j;
      }
    }
    tie_rule (other, r0, r1);
  };
  if (dense) {
    parallel_for (n_slices, sort_slice);
    std::swap (acc, other);
  } else {
    const int DIGIT = 8;
    const uint32_t digit_mask = (1u << DIGIT) - 1;
    std::vector<uint32_t> start ((size_t) 1 << DIGIT);
    for (int shift = 0; shift < 32 && (differ >> shift); shift += DIGIT) { // digits above the highest differing bit are equal
      std::fill (start.begin (), start.end (), 0u);
      for (int32_t j = 0; j < total; j++) start[(acc[j].loc >> shift) & digit_mask]++;
      uint32_t run = 0;
      for (auto &c : start) { const uint32_t n = c; c = run; run += n; }
      for (int32_t j = 0; j < total; j++) other[start[(acc[j].loc >> shift) & digit_mask]++] = acc[j];
      std::swap (acc, other);
    }
    tie_rule (acc, 0, total);
  }

  // 2. tract ids: every row is compared with the FIRST row of the current tract (genome_set.c:208-222).  A slice whose
  //    first row is max_mode_len or more away from the row before it starts a chain of its own; the others are chained
  //    on by the thread of the slice before them.
  const Tables t = {location, base, mode_len, names, name_stride, levenshtein_distance};
  std::vector<char> own_chain (n_slices, 0);
  std::vector<int32_t> tracts (n_slices, 0); // tracts started in the chain that begins at slice r
  for (int r = 0; r < n_slices; r++) {
    const int32_t r0 = slice_row[r];
    own_chain[r] = r == 0 || (dense && r0 > 0 && r0 < total && (int64_t) acc[r0].loc - (int64_t) acc[r0 - 1].loc >= (int64_t) std::max (max_mode_len, 1));
  }
  if (!dense) slice_row[n_slices] = total;
  auto chain = [&] (int r) {
    if (!own_chain[r]) return;
    int r_end = r + 1;
    while (r_end < n_slices && !own_chain[r_end]) r_end++;
    const int32_t i0 = slice_row[r], i1 = slice_row[r_end];
    std::vector<uint32_t> tab;
    std::vector<char> yname;
    int32_t prev = i0, id = 0;
    for (int32_t i = i0; i < i1; i++) {
      if (i > i0 && acc[i].loc != acc[prev].loc && !rows_overlap (t, acc[i].row, acc[prev].row, tab, yname)) { id++; prev = i; }
      out_sample[i] = acc[i].sample;
      out_index[i] = acc[i].row - first[acc[i].sample];
      out_tract_id[i] = id;
    }
    tracts[r] = i1 > i0 ? id + 1 : 0;
  };
  parallel_for (n_slices, chain);
  std::vector<int32_t> id_offset (n_slices, 0);
  for (int r = 0, run = 0; r < n_slices; r++) { if (own_chain[r]) { id_offset[r] = run; run += tracts[r]; } else id_offset[r] = id_offset[r - 1]; }
  auto renumber = [&] (int r) {
    if (!id_offset[r]) return;
    for (int32_t i = slice_row[r]; i < slice_row[r + 1]; i++) out_tract_id[i] += id_offset[r];
  };
  parallel_for (n_slices, renumber);
  return total;
}

// ---------------------------------------------------------------------------------------------------------
// Per-sample context histograms after the mapper: rows 11-13 of the hot-path table.
//
// Restates, on the flat arrays of the C ABI, what new_genomic_context_list() does with a hopo_counter once
// finalise_hopo_counter() has returned (reference src/context_histogram.c:237-271): the mapped elements are walked in
// location order and each one either joins the LAST histogram (same base, location within -m, flanks within -d
// substitutions, or names within -l edits) or starts a new one (:245-268, distance :25-48, Hamming per flank
// src/hopo_counter.c:61-68); then finalise_genomic_context_hist() (:274-305): the weighted tract-length histogram of
// every context (empirical_frequency.c:202-277: bins by length, most frequent first, ties by length descending) and
// genomic_context_merge_histograms_at_same_location() (:364-421), including its two quirks: the list is cut to j
// entries where j is the INDEX of the last one (:379-382: the last histogram of every sample is dropped), and the
// context lists are united by "append unless equal to the first differing one" (:411-418).  The GFF3 feature of a
// context (:331-351) is a table look-up by position that stays with the reference's annotation code.
// A sequential chain (every element is compared with the histogram the previous one went to): host code, like the
// mapper it follows.
namespace {

inline int32_t sext_bits (uint64_t v, int lo, int width)
{
  const int64_t m = 1LL << (width - 1);
  const int64_t x = (int64_t) ((v >> lo) & ((1ULL << width) - 1));
  return (int32_t) ((x ^ m) - m);
}

struct Elem { // the fields of a mapped hopo_element (src/hopo_counter.h:34-49)
  uint64_t ctx[2];
  int32_t base, length, count, mismatches, multi, neg_strand, location, loc2d[3];
};

Elem read_elem (const uint8_t *p)
{
  Elem e;
  uint64_t bits;
  int32_t tail[4];
  memcpy (e.ctx, p, 16); memcpy (&bits, p + 16, 8); memcpy (tail, p + 24, 16);
  e.base = sext_bits (bits, 0, 2); e.length = sext_bits (bits, 2, 10); e.count = sext_bits (bits, 12, 20);
  e.mismatches = sext_bits (bits, 32, 12); e.multi = sext_bits (bits, 44, 3); e.neg_strand = sext_bits (bits, 47, 2);
  e.location = tail[0]; e.loc2d[0] = tail[1]; e.loc2d[1] = tail[2]; e.loc2d[2] = tail[3];
  return e;
}

struct Hist {
  std::vector<uint64_t> ctx;                 // pairs
  int32_t base, multi, indel, neg_strand, mismatches;
  std::string name;
  int32_t integral, location, loc2d[3], mode_count, mode_length, mode_id;
  std::vector<std::pair<int32_t, int32_t>> bins; // (tract length, reads): raw while elements arrive, then the histogram
  int32_t hmin = 0, hmax = 0;
};

const char kBase[4] = {'A', 'C', 'G', 'T'};

// "LEFT.B.RIGHT" of a context, reverse complemented for a hit on the negative strand (src/hopo_counter.c:471-493).
// Four bases per table look-up: kQuad[b] = the four letters of the byte b of 2-bit codes, first base first.
struct QuadTable {
  char fwd[256][4], rev[256][4]; // rev: complemented, last base first
  QuadTable ()
  {
    for (int b = 0; b < 256; b++)
      for (int j = 0; j < 4; j++) {
        const int code = (b >> (2 * j)) & 3;
        fwd[b][j] = kBase[code];
        rev[b][3 - j] = kBase[3 - code];
      }
  }
};
const QuadTable kQuad;

// k letters of a flank, first base first, into dst
inline void flank_letters (uint64_t ctx, int k, char *dst)
{
  int i = 0;
  for (; i + 4 <= k; i += 4, ctx >>= 8) memcpy (dst + i, kQuad.fwd[ctx & 0xff], 4);
  for (; i < k; i++, ctx >>= 2) dst[i] = kBase[ctx & 3];
}
// the same flank complemented, last base first, ending at dst_end (exclusive)
inline void flank_letters_rc (uint64_t ctx, int k, char *dst_end)
{
  int i = 0;
  for (; i + 4 <= k; i += 4, ctx >>= 8) memcpy (dst_end - i - 4, kQuad.rev[ctx & 0xff], 4);
  for (; i < k; i++, ctx >>= 2) dst_end[-1 - i] = kBase[3 - (ctx & 3)];
}

std::string context_name (const uint64_t *ctx, int32_t base, int k, bool neg)
{
  std::string s ((size_t) 2 * k + 3, '.');
  if (!neg) {
    flank_letters (ctx[0], k, &s[0]);
    s[k + 1] = kBase[base & 3];
    flank_letters (ctx[1], k, &s[k + 3]);
  } else { // position n - 1 - i of the forward name, complemented
    const int n = 2 * k + 3;
    flank_letters_rc (ctx[0], k, &s[0] + n);
    s[n - 1 - (k + 1)] = kBase[(~base) & 3];
    flank_letters_rc (ctx[1], k, &s[0] + n - (k + 3));
  }
  return s;
}

// differing bases of two flanks, counted up to max_dist (src/hopo_counter.c:61-68)
int flank_distance (uint64_t a, uint64_t b, int max_dist)
{
  uint64_t d = a ^ b;
  int dist = 0;
  while (d && dist < max_dist) { if (d & 3) dist++; d >>= 2; }
  return dist;
}

const int kMaxDist = 0xffff; // CH_MAX_DIST

// src/context_histogram.c:25-48
int hist_elem_distance (const Hist &h, const Elem &e, int max_distance, int location_difference, int *idx_match)
{
  *idx_match = -1;
  if (h.base != e.base) return kMaxDist;
  if (std::abs (e.location - h.location) > location_difference) return kMaxDist;
  int this_max = 0;
  for (size_t c = 0; c < h.ctx.size () / 2; c++) {
    int distance = flank_distance (h.ctx[2 * c], e.ctx[0], 2 * max_distance);
    if (distance >= 2 * max_distance) return distance;
    distance += flank_distance (h.ctx[2 * c + 1], e.ctx[1], 2 * max_distance - distance);
    if (distance >= 2 * max_distance) return distance;
    if (distance > this_max) this_max = distance;
    if (distance == 0) { *idx_match = (int) c; return 0; }
  }
  return this_max;
}

Hist new_hist (const Elem &e, std::string &&name) // :133-167
{
  Hist h;
  h.ctx = {e.ctx[0], e.ctx[1]};
  h.base = e.base; h.multi = e.multi; h.indel = 0;
  h.neg_strand = e.neg_strand ? 1 : 0; h.mismatches = e.mismatches;
  h.name = std::move (name);
  h.integral = e.count; h.location = e.location;
  for (int i = 0; i < 3; i++) h.loc2d[i] = e.loc2d[i];
  h.mode_count = e.count; h.mode_length = e.length; h.mode_id = 0;
  h.bins.push_back ({e.length, e.count});
  return h;
}

void add_elem (Hist &h, const Elem &e, std::string &&name, int idx_match) // :182-222
{
  if (idx_match < 0) { idx_match = (int) (h.ctx.size () / 2); h.ctx.push_back (e.ctx[0]); h.ctx.push_back (e.ctx[1]); }
  if (h.mode_count < e.count) {
    h.mode_count = e.count; h.mode_length = e.length; h.mode_id = idx_match;
    h.location = e.location;
    for (int i = 0; i < 3; i++) h.loc2d[i] = e.loc2d[i];
    h.mismatches = e.mismatches; h.neg_strand = e.neg_strand ? 1 : 0;
    h.name = std::move (name);
  }
  if ((h.multi ^ e.multi) == 1) h.multi = 2;
  h.integral += e.count;
  h.bins.push_back ({e.length, e.count});
}

// (value, weight) pairs -> bins by value with summed weights, most frequent first, ties by value descending
// (empirical_frequency.c:254-277 after the sort of :219 / :245)
void make_histogram (std::vector<std::pair<int32_t, int32_t>> &v, int32_t *vmin, int32_t *vmax)
{
  if (v.size () == 1) { *vmin = *vmax = v[0].first; return; } // the common case: one tract length seen in this context
  std::sort (v.begin (), v.end ());
  std::vector<std::pair<int32_t, int32_t>> out;
  for (const auto &p : v) {
    if (!out.empty () && out.back ().first == p.first) out.back ().second += p.second;
    else out.push_back (p);
  }
  *vmin = out.front ().first; *vmax = out.back ().first;
  std::sort (out.begin (), out.end (), [] (const std::pair<int32_t, int32_t> &a, const std::pair<int32_t, int32_t> &b) {
    return a.second != b.second ? a.second > b.second : a.first > b.first; });
  v.swap (out);
}

bool hists_overlap (const Hist &c1, const Hist &c2, int lev, std::vector<uint32_t> &tab) // :87-110
{
  const int32_t apart = std::abs (c1.location - c2.location);
  if (!apart) return true;
  const int32_t tract = std::min (c1.mode_length, c2.mode_length);
  if (apart > tract - 1 || c1.base != c2.base) return false;
  const uint32_t n = (uint32_t) c1.name.size ();
  std::string other = c2.name;
  other.resize (n, '\0'); // the reference reads the second name over the first one's length
  return border_free_levenshtein (c1.name.data (), n, other.data (), n, tab) <= (uint32_t) lev;
}

void accumulate (Hist &to, Hist &from) // :387-421
{
  std::vector<std::pair<int32_t, int32_t>> both = to.bins;
  both.insert (both.end (), from.bins.begin (), from.bins.end ());
  make_histogram (both, &to.hmin, &to.hmax);
  to.bins.swap (both);
  if (to.mode_count < from.mode_count) {
    to.location = from.location; to.mode_count = from.mode_count; to.mode_length = from.mode_length;
    to.mode_id = from.mode_id; to.mismatches = from.mismatches;
    for (int i = 0; i < 3; i++) to.loc2d[i] = from.loc2d[i];
    to.ctx.swap (from.ctx);
    to.name.swap (from.name);
  }
  to.integral += from.integral;
  const size_t n1 = to.ctx.size () / 2;
  for (size_t i = 0; i < from.ctx.size () / 2; i++)
    for (size_t j = 0; j < n1; j++)
      if (to.ctx[2 * j] != from.ctx[2 * i] || to.ctx[2 * j + 1] != from.ctx[2 * i + 1]) { // :411-418, as written
        to.ctx.push_back (from.ctx[2 * i]); to.ctx.push_back (from.ctx[2 * i + 1]);
        break;
      }
}

template <typename T> T *alloc_array (size_t n) { return (T*) malloc ((n ? n : 1) * sizeof (T)); }

} // namespace

extern "C" void tjb_context_table_free (tjb_context_table *t)
{
  if (!t) return;
  void *p[] = {t->location, t->loc2d, t->base, t->neg_strand, t->mismatches, t->multi, t->indel, t->integral, t->n_context,
               t->mode_context_count, t->mode_context_length, t->mode_context_id, t->h_min, t->h_max, t->names, t->ctx_first,
               t->contexts, t->h_first, t->h_length, t->h_freq};
  for (void *q : p) free (q);
  memset (t, 0, sizeof (*t));
}

extern "C" int tjb_context_histograms (const tjb_hist_opts *opt, const void *elem, int32_t n_elem, tjb_context_table *out)
{
  if (!opt || !out || n_elem < 0 || (n_elem && !elem) || opt->kmer_size < 2 || opt->kmer_size > 32) return TJB_ERR_ARG;
  memset (out, 0, sizeof (*out));
  const int k = opt->kmer_size;
  out->name_stride = 2 * k + 4;
  if (!n_elem) return 0;
  const uint8_t *p = (const uint8_t*) elem;
  std::vector<Hist> hist;
  hist.reserve ((size_t) n_elem); // most elements open a histogram of their own
  std::vector<uint32_t> tab;
  // 1. elements in location order against the last histogram (:237-268)
  for (int32_t i = 0; i < n_elem; i++) {
    const Elem e = read_elem (p + (size_t) 40 * i);
    std::string name = context_name (e.ctx, e.base, k, e.neg_strand != 0);
    if (hist.empty ()) { hist.push_back (new_hist (e, std::move (name))); continue; }
    Hist &last = hist.back ();
    int idx_match;
    int distance = hist_elem_distance (last, e, opt->max_distance_per_flank, opt->min_tract_size, &idx_match);
    if (distance < opt->max_distance_per_flank) add_elem (last, e, std::move (name), idx_match);
    else {
      if (distance < kMaxDist) { // same base, close by: try again, now allowing indels between the names
        const uint32_t n = (uint32_t) last.name.size ();
        distance = (int) border_free_levenshtein (last.name.data (), n, name.data (), n, tab);
      }
      if (distance < opt->levenshtein_distance) { add_elem (last, e, std::move (name), idx_match); last.indel = 1; }
      else hist.push_back (new_hist (e, std::move (name)));
    }
  }
  // 2. tract-length histogram of every context (:279-286); elements of weight 0 do not take part
  for (Hist &h : hist) {
    auto &v = h.bins;
    v.erase (std::remove_if (v.begin (), v.end (), [] (const std::pair<int32_t, int32_t> &b) { return b.second == 0; }), v.end ());
    if (v.empty ()) return TJB_ERR_ARG; // the reference stops with "vector of empirical frequencies with all freqs=0"
    make_histogram (v, &h.hmin, &h.hmax);
  }
  // 3. neighbours that overlap are merged (:364-385); the list then keeps j entries, j being the index of the last one
  std::vector<Hist> merged;
  merged.reserve (hist.size ());
  merged.push_back (std::move (hist[0]));
  for (size_t i = 1; i < hist.size (); i++) {
    if (hists_overlap (merged.back (), hist[i], opt->levenshtein_distance, tab)) accumulate (merged.back (), hist[i]);
    else merged.push_back (std::move (hist[i]));
  }
  const size_t n_out = merged.size () - 1; // :379-382: n_hist = j
  // 4. flat arrays
  size_t n_ctx = 0, n_bins = 0;
  for (size_t i = 0; i < n_out; i++) { n_ctx += merged[i].ctx.size () / 2; n_bins += merged[i].bins.size (); }
  out->n_hist = (int32_t) n_out;
  out->location = alloc_array<int32_t> (n_out); out->loc2d = alloc_array<int32_t> (3 * n_out); out->base = alloc_array<int32_t> (n_out);
  out->neg_strand = alloc_array<int32_t> (n_out); out->mismatches = alloc_array<int32_t> (n_out); out->multi = alloc_array<int32_t> (n_out);
  out->indel = alloc_array<int32_t> (n_out); out->integral = alloc_array<int32_t> (n_out); out->n_context = alloc_array<int32_t> (n_out);
  out->mode_context_count = alloc_array<int32_t> (n_out); out->mode_context_length = alloc_array<int32_t> (n_out);
  out->mode_context_id = alloc_array<int32_t> (n_out); out->h_min = alloc_array<int32_t> (n_out); out->h_max = alloc_array<int32_t> (n_out);
  out->names = alloc_array<char> (n_out * (size_t) out->name_stride);
  out->ctx_first = alloc_array<int32_t> (n_out + 1); out->contexts = alloc_array<uint64_t> (2 * n_ctx);
  out->h_first = alloc_array<int32_t> (n_out + 1); out->h_length = alloc_array<int32_t> (n_bins); out->h_freq = alloc_array<int32_t> (n_bins);
  if (!out->location || !out->loc2d || !out->base || !out->neg_strand || !out->mismatches || !out->multi || !out->indel || !out->integral ||
      !out->n_context || !out->mode_context_count || !out->mode_context_length || !out->mode_context_id || !out->h_min || !out->h_max ||
      !out->names || !out->ctx_first || !out->contexts || !out->h_first || !out->h_length || !out->h_freq) {
    tjb_context_table_free (out);
    return TJB_ERR_NOMEM;
  }
  memset (out->names, 0, n_out * (size_t) out->name_stride);
  size_t c = 0, b = 0;
  for (size_t i = 0; i < n_out; i++) {
    const Hist &h = merged[i];
    out->location[i] = h.location; out->base[i] = h.base; out->neg_strand[i] = h.neg_strand; out->mismatches[i] = h.mismatches;
    for (int j = 0; j < 3; j++) out->loc2d[3 * i + j] = h.loc2d[j];
    out->multi[i] = h.multi; out->indel[i] = h.indel; out->integral[i] = h.integral; out->n_context[i] = (int32_t) (h.ctx.size () / 2);
    out->mode_context_count[i] = h.mode_count; out->mode_context_length[i] = h.mode_length; out->mode_context_id[i] = h.mode_id;
    out->h_min[i] = h.hmin; out->h_max[i] = h.hmax;
    memcpy (out->names + i * (size_t) out->name_stride, h.name.data (), std::min (h.name.size (), (size_t) out->name_stride - 1));
    out->ctx_first[i] = (int32_t) c;
    memcpy (out->contexts + 2 * c, h.ctx.data (), h.ctx.size () * sizeof (uint64_t));
    c += h.ctx.size () / 2;
    out->h_first[i] = (int32_t) b;
    for (const auto &bin : h.bins) { out->h_length[b] = bin.first; out->h_freq[b] = bin.second; b++; }
  }
  out->ctx_first[n_out] = (int32_t) c; out->h_first[n_out] = (int32_t) b;
  return (int) n_out;
}

// libtjb_gpu: context, workspace, stream orchestration and the C ABI declared in include/tjb_gpu.h.
// Device work only — there is no host fallback for any compute step in this file.
#include <chrono>
#include <sys/mman.h>
#include <sys/stat.h>
#include <climits>
#include <cstdlib>
#include <cstring>
#include <condition_variable>
#include <map>
#include <mutex>
#include <thread>
#include <string>
#include <vector>

#include "tjb_gpu.h"
#include "tjb_common.cuh"
#include "tjb_scan.cuh"
#include "tjb_bucket.cuh"
#include "tjb_sort.cuh"
#include "tjb_reduce.cuh"
#include "tjb_union.cuh"
#include "tjb_hostpack.h"

#define TJB_VERSION_STR "tatajuba_b200 0.1 (sm_100a)"
#define TJB_CHUNK_BYTES (32u << 20)      // files path: pinned chunk of decoded reads (two per decoder thread)
#define TJB_HOST_CHUNK_BYTES (32u << 20) // host-buffer path: H2D chunk overlapped with the scan of the previous one

enum KernelId { K_SCAN = 0, K_AGGREGATE, K_PLAN, K_PARTITION, K_BUCKET, K_TOTALS, K_PACK, K_COVMAX,
                // global multi-pass path, only for buckets too large for shared memory
                K_SORT_HIST, K_SORT_SCATTER, K_PSCAN, K_RLE, K_GROUP, K_UNION, K_UNPACK, K_TABLE, K_COUNT };
static const char *kKernelNames[K_COUNT] = {"tjb_scan_kernel", "aggregate_kernel", "plan_kernel", "wpartition_kernel", "wbucket_reduce_kernel", "totals_kernel",
                                            "pack_kernel", "cov_max_kernel", "rsort_hist_kernel",
                                            "rsort_scatter_kernel", "pscan_kernels", "rle_kernels", "group_kernels", "union_kernels", "unpack_kernels",
                                            "table_kernels"};

struct DevBuf {
  void *p = nullptr;
  size_t cap = 0;
};

struct PendingEvent { int kid; cudaEvent_t a, b; };

// device scalars of one sample, one allocation so that they come back in one copy
struct ResultBlock {
  unsigned long long counters[CTR_WORDS];
  u32 meta[META_WORDS];
  int cov;
  u32 pad;
};

struct tjb_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  std::string err;
  int profiling = 0; // 0 off, 1 scan kernel only, 2 every kernel
  bool t1_recorded = false;
  int result_mode = TJB_RESULT_MALLOC;
  bool result_dma = false; // results leave through a device staging array + DMA instead of mapped stores of the pack kernel
  int64_t launches = 0;
  int64_t h2d_bytes = 0; // host-buffer path: bytes sent host -> device
  size_t last_keep = 0, last_idx = 0, last_entries = 0; // sizes of the previous sample: the next one is sized ahead
  size_t want_entries = 0, want_records = 0;            // set when a sample has to be redone with a larger table / record array
  tjb_kernel_stat stats[K_COUNT];
  std::vector<PendingEvent> pending;
  std::vector<cudaEvent_t> free_events;
  int sm_count = 148;

  uint8_t *h_pinned[4] = {nullptr, nullptr, nullptr, nullptr}; // files path: two chunks per decoder thread
  cudaEvent_t h_free[4] = {nullptr, nullptr, nullptr, nullptr}; // pinned chunk i may be overwritten once this has fired
  DevBuf d_chunk[2];
  DevBuf d_long;                             // a read longer than a chunk: copied and scanned on its own
  // host-buffer path with 2-bit packed transfers (tjb_hostpack.h): pinned and device staging per chunk slot
  uint8_t *h_pk[2] = {nullptr, nullptr}; uint32_t *h_exc_pos[2] = {nullptr, nullptr}; uint8_t *h_exc_val[2] = {nullptr, nullptr};
  DevBuf d_pk[2], d_exc_pos[2], d_exc_val[2];

  // ---- the sample being counted ----
  tjb_opts opt = {0, 0, 0, 0, 0};
  bool in_sample = false;
  int scan_launches = 0;                    // of this sample
  bool records_pending = false, records_aggregated = false; // scan records not yet / already added to the table
  int64_t exported_rows = -1;               // rows of the table as exported for an exchange (-1: not exported yet)
  int rec_words = 0;
  int bin_bits = 8, bucket_shift = 10;      // partition plan: 2^bin_bits bins, bucket windows of 2^bucket_shift records
  DevBuf tab;                               // aggregation table (tjb_table.cuh): counts, then keys
  size_t tab_slots = 0;
  DevBuf recs;                              // tract observations as the scan found them (one record each)
  size_t rec_capacity = 0;
  DevBuf ovf;                               // observations the table could not take, raw records
  size_t ovf_capacity = 0;
  DevBuf wkey, wcnt, okey;                  // weighted records bucket by bucket; compact per-bucket output
  size_t list_cap = 0;
  ResultBlock *d_res = nullptr, *h_res = nullptr; // device scalars and their pinned mirror
  u32 *d_tile_ctr = nullptr;                // scan kernel: next window to hand out
  u32 *d_totals = nullptr;                  // small device scalars (global multi-pass path)
  u32 *d_bin_hist = nullptr, *d_lut = nullptr;                               // per bin (2^TJB_MAX_BIN_BITS)
  u32 *d_bkt_start = nullptr, *d_cursor = nullptr, *d_koff = nullptr, *d_ioff = nullptr; // per bucket (views of the DevBufs below)
  uint4 *d_bstat = nullptr;
  DevBuf b_bkt_start, b_cursor, b_koff, b_ioff, b_bstat;
  u32 max_buckets = 0;
  // results: pinned, device-mapped arena {elem[keep_cap] (40 B), idx_initial[idx_cap], idx_final[idx_cap]}
  char *h_arena = nullptr, *d_arena = nullptr;
  size_t arena_keep_cap = 0, arena_idx_cap = 0;
  u32 cov_slots = 0;                        // slots of the coverage table of the last pack launch
  DevBuf out_cnt, out_ii, out_if, cov_tab, out40, idx_i, idx_f, qrows;
  DevBuf hist, ptmp, heads, pos, seg, drec, dcount, keep, kpos, krec, cw, csum, gkeep, gkpos;
  DevBuf uni_a, uni_b;                      // union of query rows

  cudaEvent_t ev_t0 = nullptr, ev_t1 = nullptr;
  cudaStream_t copy_stream = nullptr; // host-buffer path: H2D copies overlap the scan of the previous chunk
  cudaEvent_t ev_ready = nullptr, ev_copied[2] = {nullptr, nullptr}, ev_scanned[2] = {nullptr, nullptr};
};

// ------------------------------------------------------------------------------------------- helpers ----
#define CU(call)                                                                                        \
  do {                                                                                                  \
    cudaError_t e_ = (call);                                                                            \
    if (e_ != cudaSuccess) {                                                                            \
      ctx->err = std::string (#call) + ": " + cudaGetErrorString (e_);                                  \
      return TJB_ERR_CUDA;                                                                              \
    }                                                                                                   \
  } while (0)
#define TRY(call)                \
  do {                           \
    int rc_ = (call);            \
    if (rc_ < 0) return rc_;     \
  } while (0)

static int ensure (tjb_ctx *ctx, DevBuf &b, size_t bytes)
{
  if (bytes <= b.cap) return 0;
  size_t want = bytes + bytes / 4 + 256;
  if (b.p) { CU (cudaStreamSynchronize (ctx->stream)); CU (cudaFree (b.p)); b.p = nullptr; b.cap = 0; }
  CU (cudaMalloc (&b.p, want));
  b.cap = want;
  return 0;
}

static cudaEvent_t get_event (tjb_ctx *ctx)
{
  if (!ctx->free_events.empty ()) { cudaEvent_t e = ctx->free_events.back (); ctx->free_events.pop_back (); return e; }
  cudaEvent_t e = nullptr;
  cudaEventCreate (&e);
  return e;
}

struct LaunchScope { // counts a launch and, when profiling, brackets it with events on the context's stream
  tjb_ctx *ctx; int kid; bool timed; cudaEvent_t a = nullptr, b = nullptr;
  LaunchScope (tjb_ctx *c, int k, int64_t alg_bytes) : ctx (c), kid (k)
  {
    timed = ctx->profiling >= 2 || (ctx->profiling == 1 && kid == K_SCAN);
    ctx->launches++;
    ctx->stats[kid].launches++;
    ctx->stats[kid].bytes_algorithmic += alg_bytes;
    if (timed) { a = get_event (ctx); b = get_event (ctx); cudaEventRecord (a, ctx->stream); }
  }
  ~LaunchScope ()
  {
    if (timed) { cudaEventRecord (b, ctx->stream); ctx->pending.push_back ({kid, a, b}); }
  }
};
#define LAUNCH(kid, alg_bytes, ...) do { LaunchScope ls_ (ctx, kid, (int64_t) (alg_bytes)); __VA_ARGS__; } while (0)

static void resolve_events (tjb_ctx *ctx) // call only after the stream has been synchronised
{
  for (auto &p : ctx->pending) {
    float ms = 0.f;
    if (cudaEventElapsedTime (&ms, p.a, p.b) == cudaSuccess) ctx->stats[p.kid].ms_total += ms;
    ctx->free_events.push_back (p.a);
    ctx->free_events.push_back (p.b);
  }
  ctx->pending.clear ();
}

struct Trace { // TJB_TRACE=1: host wall-clock of the stages of one call, on stderr
  bool on; std::chrono::steady_clock::time_point t0; std::string line;
  Trace () : on (getenv ("TJB_TRACE") != nullptr), t0 (std::chrono::steady_clock::now ()) {}
  void mark (const char *what)
  {
    if (!on) return;
    auto t = std::chrono::steady_clock::now ();
    char b[64]; snprintf (b, sizeof b, " %s=%.0fus", what, std::chrono::duration<double, std::micro> (t - t0).count ());
    line += b; t0 = t;
  }
  ~Trace () { if (on) fprintf (stderr, "[tjb]%s\n", line.c_str ()); }
};

static inline u32 nblk (size_t n, u32 per) { return (u32) ((n + per - 1) / per); }
static inline size_t pow2_at_least (size_t x) { size_t p = 1; while (p < x) p <<= 1; return p; }

// ------------------------------------------------------------------------------------ scan + sort -------
static int exclusive_scan (tjb_ctx *ctx, const u32 *in, u32 *out, size_t n, u32 *tmp, u32 *total_out)
{
  if (n <= PSCAN_TILE) {
    LAUNCH (K_PSCAN, 8 * n, (pscan_scan_kernel<<<1, PSCAN_THREADS, 0, ctx->stream>>> (in, out, nullptr, n, total_out)));
    return 0;
  }
  u32 nb = nblk (n, PSCAN_TILE);
  LAUNCH (K_PSCAN, 4 * n, (pscan_reduce_kernel<<<nb, PSCAN_THREADS, 0, ctx->stream>>> (in, tmp, n)));
  TRY (exclusive_scan (ctx, tmp, tmp, nb, tmp + nb, nullptr));
  LAUNCH (K_PSCAN, 8 * n, (pscan_scan_kernel<<<nb, PSCAN_THREADS, 0, ctx->stream>>> (in, out, tmp, n, total_out)));
  return 0;
}

static void add_digits (std::vector<DigitDesc> &v, int word, int lo, int nbits)
{
  for (int done = 0; done < nbits; done += RSORT_RADIX_BITS) {
    int nb = nbits - done < RSORT_RADIX_BITS ? nbits - done : RSORT_RADIX_BITS;
    v.push_back ({word, lo + done, nb});
  }
}
// least significant digit first: flag < length < ctx1 < ctx0 < base (see tjb_common.cuh)
static std::vector<DigitDesc> tract_digits (int W, int k)
{
  std::vector<DigitDesc> v;
  if (W == 2) { add_digits (v, 0, 0, 12 + 2 * k); add_digits (v, 1, 0, 2 * k + 1); }
  else { add_digits (v, 0, 0, 12); add_digits (v, 1, 0, 2 * k); add_digits (v, 2, 0, 2 * k); add_digits (v, 0, 12, 1); }
  return v;
}

// S = words per record (the digits only look at the key words)
template <int S>
static int radix_sort (tjb_ctx *ctx, u64 *a, u64 *b, size_t n, const std::vector<DigitDesc> &digits, u64 **sorted)
{
  u32 n_tiles = nblk (n, RSORT_TILE);
  size_t hn = (size_t) RSORT_RADIX * n_tiles;
  TRY (ensure (ctx, ctx->hist, hn * sizeof (u32)));
  TRY (ensure (ctx, ctx->ptmp, pscan_tmp_elems (hn > n ? hn : n) * sizeof (u32) + 64));
  u32 *hist = (u32*) ctx->hist.p;
  for (const DigitDesc &dd : digits) {
    LAUNCH (K_SORT_HIST, 8 * S * n, (rsort_hist_kernel<S><<<n_tiles, RSORT_THREADS, 0, ctx->stream>>> (a, n, dd, hist, n_tiles)));
    TRY (exclusive_scan (ctx, hist, hist, hn, (u32*) ctx->ptmp.p, nullptr));
    LAUNCH (K_SORT_SCATTER, 16 * S * n, (rsort_scatter_kernel<S><<<n_tiles, RSORT_THREADS, 0, ctx->stream>>> (a, b, n, dd, hist, n_tiles)));
    u64 *t = a; a = b; b = t;
  }
  *sorted = a;
  return 0;
}

// records per bucket (2^bshift, the window of the partition plan): a sample's records spread over every CTA the bucket
// kernel can have resident, between 64 (buckets of < 128 records are ordered by comparison counting) and BK_CAPH
static int choose_bucket_shift (size_t est_records, int sm_count)
{
  int bshift = 6;
  while ((1u << bshift) < BK_CAPH && (est_records >> (bshift + 1)) > (size_t) sm_count * 3) bshift++;
  return bshift;
}
// bins of the partition plan: enough that a bin holds a quarter of a bucket window on average
static int choose_bin_bits (size_t est_records, int bshift, int k)
{
  const int W = tjb_rec_words (k);
  const int max_bits = W == 2 ? (4 * k + 1 < TJB_MAX_BIN_BITS ? 4 * k + 1 : TJB_MAX_BIN_BITS)
                              : (2 * k + 1 < TJB_MAX_BIN_BITS ? 2 * k + 1 : TJB_MAX_BIN_BITS);
  int F = 4;
  while (F < TJB_MAX_BIN_BITS && (est_records >> F) > ((size_t) 1 << bshift) / 4) F++;
  return F < max_bits ? F : max_bits;
}

// Table entries to expect before anything of the sample has been seen: a run of >= m equal bases starts at a position
// with probability 3/4 * 4^-(m-1); that many observations, every one of them distinct in the worst case.
static size_t estimate_entries (size_t n_bytes_hint, int min_tract)
{
  const int m = min_tract < 2 ? 2 : (min_tract > 8 ? 8 : min_tract);
  if (!n_bytes_hint) return (size_t) 1 << 20; // size unknown
  const size_t est = (n_bytes_hint * 3 / 4) >> (2 * (m - 1));
  return est < 4096 ? 4096 : est;
}

#define TJB_MAX_TABLE_SLOTS ((size_t) 1 << 28)

// Sizes and clears the per-sample device state.  Everything that depends on how many DISTINCT observations the sample
// holds (table slots, overflow list, weighted record arrays, bins) is sized from the previous sample of this context
// (samples of one run are alike) or, for the first one, from the input size; a sample that does not fit is detected
// at the synchronisation that ends it and redone with ctx->want_entries.
static int begin_sample (tjb_ctx *ctx, const tjb_opts *opt, size_t n_bytes_hint)
{
  if (!opt || opt->kmer_size < 2 || opt->kmer_size > 32 || opt->min_tract_size < 1 || opt->min_coverage < 0) {
    ctx->err = "options out of range (kmer_size 2..32, min_tract_size >= 1, min_coverage >= 0)";
    return TJB_ERR_ARG;
  }
  CU (cudaSetDevice (ctx->device));
  ctx->opt = *opt;
  const int W = tjb_rec_words (opt->kmer_size);
  ctx->rec_words = W;
  size_t est = ctx->want_entries ? ctx->want_entries
             : (ctx->last_entries ? ctx->last_entries + ctx->last_entries / 4 + 1024 : estimate_entries (n_bytes_hint, opt->min_tract_size));
  ctx->want_entries = 0;
  // slots per expected entry: collisions make the lanes of a warp probe in turns, so a table that stays L2 resident is
  // kept sparse (measured at 58 K entries: 3x: aggregate 79 us + partition 12; 8x: 60 + 19; 16x: 57 + 22); a large one
  // is not: its memset and the partition's walk grow with the slots
  static const int tab_factor = getenv ("TJB_TABLE_FACTOR") ? atoi (getenv ("TJB_TABLE_FACTOR")) : 0; // experiments
  const size_t factor = tab_factor >= 2 && tab_factor <= 64 ? (size_t) tab_factor : (est <= ((size_t) 1 << 18) ? 8 : 3);
  size_t slots = pow2_at_least (factor * est);
  if (slots < ((size_t) 1 << 16)) slots = (size_t) 1 << 16;
  if (slots > TJB_MAX_TABLE_SLOTS) slots = TJB_MAX_TABLE_SLOTS;
  ctx->tab_slots = slots;
  ctx->ovf_capacity = est / 8 < 65536 ? 65536 : est / 8;
  ctx->list_cap = slots / 2 + ctx->ovf_capacity;
  ctx->bucket_shift = choose_bucket_shift (est, ctx->sm_count);
  { static const int forced = getenv ("TJB_BUCKET_SHIFT") ? atoi (getenv ("TJB_BUCKET_SHIFT")) : 0; // experiments
    if (forced >= 5 && forced <= 10) ctx->bucket_shift = forced; }
  ctx->bin_bits = choose_bin_bits (est, ctx->bucket_shift, opt->kmer_size);
  TRY (ensure (ctx, ctx->tab, table_bytes (slots, W)));
  TRY (ensure (ctx, ctx->ovf, ctx->ovf_capacity * W * 8));
  // the scan's records: a guess from the input size (a record per 16 bytes is ~10x what a random genome gives at -m 4);
  // records that do not fit are counted, never written, and the sample is redone with the exact size
  ctx->rec_capacity = ctx->want_records ? ctx->want_records : n_bytes_hint / 16 + 65536;
  ctx->want_records = 0;
  if (ctx->recs.cap / ((size_t) W * 8) > ctx->rec_capacity) ctx->rec_capacity = ctx->recs.cap / ((size_t) W * 8); // what is there already
  TRY (ensure (ctx, ctx->recs, ctx->rec_capacity * W * 8));
  CU (cudaMemsetAsync (ctx->tab.p, 0, table_bytes (slots, W), ctx->stream));
  {
    const u32 n_bins = 1u << ctx->bin_bits;
    const u32 g = n_bins / 256 < 1 ? 1 : (n_bins / 256 > (u32) ctx->sm_count * 4 ? (u32) ctx->sm_count * 4 : n_bins / 256);
    sample_prologue_kernel<<<g, 256, 0, ctx->stream>>> ((u32*) ctx->d_res, (int) (sizeof (ResultBlock) / 4), (int) (offsetof (ResultBlock, cov) / 4), ctx->d_tile_ctr, ctx->d_bin_hist, n_bins);
    ctx->launches++; ctx->stats[K_TABLE].launches++;
    CU (cudaGetLastError ());
  }
  ctx->scan_launches = 0;
  ctx->records_pending = ctx->records_aggregated = false;
  ctx->in_sample = true;
  ctx->exported_rows = -1;
  return 0;
}

static AggTable table_of (tjb_ctx *ctx)
{
  AggTable t;
  t.cnt = (unsigned long long*) ctx->tab.p; t.keys = (u64*) ctx->tab.p + ctx->tab_slots; t.mask = (u32) (ctx->tab_slots - 1);
  return t;
}

static int launch_scan (tjb_ctx *ctx, const uint8_t *d_seqs, size_t n)
{
  long long n_tiles = (long long) ((n + SW_TILE - 1) / SW_TILE); // one window of SW_TILE owned bytes per warp at a time
  if (!n_tiles) return 0;
  const tjb_opts *opt = &ctx->opt;
  int min_len = opt->min_tract_size < 2 ? 2 : opt->min_tract_size;
  if (ctx->records_aggregated) { ctx->err = "the sample's table has been exported: no more reads can be scanned into it"; return TJB_ERR_ARG; }
  ctx->records_pending = true;
  if (ctx->scan_launches++) CU (cudaMemsetAsync (ctx->d_tile_ctr, 0, sizeof (u32), ctx->stream)); // window tickets of this launch (the first: cleared by the prologue)
  long long want = (n_tiles + SCAN_WARPS - 1) / SCAN_WARPS;
  long long grid = want < (long long) ctx->sm_count * SCAN_CTAS_PER_SM ? want : (long long) ctx->sm_count * SCAN_CTAS_PER_SM;
  // record positions are reserved in blocks once every warp has eight or more windows to scan (from ~110 MB of reads on;
  // what a warp leaves unused of its last block, at most a block per warp, is small against the record array then:
  // bytes / 16 + 64 Ki positions)
#ifndef SCAN_REC_BLOCK3
#define SCAN_REC_BLOCK3 256u
#endif
  const u32 rec_block = ((ctx->rec_words == 3 || SCAN_BLOCKS_W2) && n_tiles >= 8 * grid * SCAN_WARPS) ? (ctx->rec_words == 3 ? SCAN_REC_BLOCK3 : 128u) : 0u; // (W = 2: exact, see flush ())
  // algorithmic bytes of this launch: every sequence byte read once (+ 8W bytes per emitted record, added at finalise)
  if (ctx->rec_words == 2)
    LAUNCH (K_SCAN, n, (tjb_scan_kernel<2><<<(u32) grid, SCAN_THREADS, scan_smem_bytes<2> (), ctx->stream>>> (d_seqs, (long long) n, opt->kmer_size, min_len, (u64*) ctx->recs.p, ctx->rec_capacity, ctx->d_res->counters, n_tiles, ctx->d_tile_ctr, rec_block)));
  else
    LAUNCH (K_SCAN, n, (tjb_scan_kernel<3><<<(u32) grid, SCAN_THREADS, scan_smem_bytes<3> (), ctx->stream>>> (d_seqs, (long long) n, opt->kmer_size, min_len, (u64*) ctx->recs.p, ctx->rec_capacity, ctx->d_res->counters, n_tiles, ctx->d_tile_ctr, rec_block)));
  CU (cudaGetLastError ());
  return 0;
}

// ONE pool of packer threads per process, shared by every context (the reference feeds several GPUs from OpenMP
// threads of one process: a pool per context would oversubscribe the host).  Chunks of different contexts take turns;
// the pool spins while at least one call is feeding a sample and is parked otherwise.  Never destroyed.
static std::mutex g_packer_mu;       // pool creation and the session count
static std::mutex g_packer_run_mu;   // one chunk at a time
static tjb_packer *g_packer = nullptr;
static int g_packer_sessions = 0;

// the pool, created on first use; nullptr when TJB_HOST_PACK=0
static tjb_packer *host_pool ()
{
  const char *e = getenv ("TJB_HOST_PACK"); // "0" disables, a number > 0 sets the thread count (first use decides)
  if (e && atoi (e) == 0) return nullptr;
  std::lock_guard<std::mutex> lk (g_packer_mu);
  if (!g_packer) g_packer = tjb_packer_create (e ? atoi (e) : 0);
  return g_packer;
}

struct PackerSession { // the packer threads spin while at least one call is feeding a sample
  bool on;
  explicit PackerSession (bool enable) : on (enable)
  {
    if (!on) return;
    std::lock_guard<std::mutex> lk (g_packer_mu);
    if (g_packer_sessions++ == 0) tjb_packer_begin (g_packer);
  }
  ~PackerSession ()
  {
    if (!on) return;
    std::lock_guard<std::mutex> lk (g_packer_mu);
    if (--g_packer_sessions == 0) tjb_packer_end (g_packer);
  }
};

// --------------------------------------------------------------------------------------- finalise -------
static int qc_mode_of (const tjb_opts *opt) { return opt->remove_biased ? QC_BOTH_STRANDS : QC_SEEN_TWICE; }

// Global multi-pass reduce of ONE bucket that does not fit in shared memory: the weighted records are gathered into
// (key, count) rows, LSD radix sorted over every key digit, equal keys merged, QC filter, context groups.  Results land
// where wbucket_reduce_kernel would have put them.
template <int W>
static int reduce_oversize_bucket (tjb_ctx *ctx, int qc_mode, int min_coverage, int k, u32 b, u32 start, u32 n)
{
  constexpr int S = W + 1;
  cudaStream_t st = ctx->stream;
  TRY (ensure (ctx, ctx->drec, (size_t) n * S * 8));
  TRY (ensure (ctx, ctx->krec, (size_t) n * S * 8));
  u64 *in = (u64*) ctx->drec.p, *tmp = (u64*) ctx->krec.p;
  u64 *wkey = (u64*) ctx->wkey.p + (size_t) start * W;
  LAUNCH (K_RLE, (16 * W + 16) * (size_t) n, (wgather_kernel<W><<<nblk (n, RED_THREADS), RED_THREADS, 0, st>>> (wkey, (const unsigned long long*) ctx->wcnt.p + start, n, in)));
  u64 *sorted = nullptr;
  TRY ((radix_sort<S> (ctx, in, tmp, n, tract_digits (W, k), &sorted)));
  TRY (ensure (ctx, ctx->heads, (size_t) n * 4)); TRY (ensure (ctx, ctx->pos, (size_t) n * 4)); TRY (ensure (ctx, ctx->seg, (size_t) n * 4));
  TRY (ensure (ctx, ctx->ptmp, pscan_tmp_elems (n) * 4 + 64));
  u32 *heads = (u32*) ctx->heads.p, *pos = (u32*) ctx->pos.p, *seg = (u32*) ctx->seg.p, *ptmp = (u32*) ctx->ptmp.p;
  u32 h_tot[4];
  LAUNCH (K_RLE, 16 * S * (size_t) n, (head_kernel<W, 0, S><<<nblk (n, RED_THREADS), RED_THREADS, 0, st>>> (sorted, n, heads)));
  TRY (exclusive_scan (ctx, heads, pos, n, ptmp, ctx->d_totals + 0));
  CU (cudaMemcpyAsync (h_tot, ctx->d_totals, 4, cudaMemcpyDeviceToHost, st));
  LAUNCH (K_RLE, 12 * (size_t) n, (seg_start_kernel<<<nblk (n, RED_THREADS), RED_THREADS, 0, st>>> (heads, pos, n, seg)));
  CU (cudaStreamSynchronize (st));
  const u32 n_dist = h_tot[0];
  // distinct records: the bucket's slice of the weighted list is free now (its rows were gathered)
  u64 *drec = wkey;
  TRY (ensure (ctx, ctx->dcount, (size_t) n_dist * 4)); TRY (ensure (ctx, ctx->keep, (size_t) n_dist * 4)); TRY (ensure (ctx, ctx->kpos, (size_t) n_dist * 4));
  LAUNCH (K_RLE, (16 * W + 12) * (size_t) n_dist, (wrle_emit_kernel<W><<<nblk (n_dist, RED_THREADS), RED_THREADS, 0, st>>> (sorted, n, seg, n_dist, qc_mode, drec, (int*) ctx->dcount.p, (u32*) ctx->keep.p)));
  TRY (exclusive_scan (ctx, (u32*) ctx->keep.p, (u32*) ctx->kpos.p, n_dist, ptmp, ctx->d_totals + 1));
  CU (cudaMemcpyAsync (h_tot, ctx->d_totals + 1, 4, cudaMemcpyDeviceToHost, st));
  CU (cudaStreamSynchronize (st));
  const u32 n_keep = h_tot[0];
  u32 n_idx = 0;
  if (n_keep) {
    u64 *krec = (u64*) ctx->okey.p + (size_t) start * W;
    int *kcount = (int*) ctx->out_cnt.p + start;
    LAUNCH (K_RLE, (16 * W + 8) * (size_t) n_keep, (compact_kernel<W><<<nblk (n_dist, RED_THREADS), RED_THREADS, 0, st>>> (drec, (int*) ctx->dcount.p, (u32*) ctx->keep.p, (u32*) ctx->kpos.p, n_dist, krec, kcount)));
    TRY (ensure (ctx, ctx->cw, (size_t) n_keep * 4)); TRY (ensure (ctx, ctx->csum, (size_t) n_keep * 4));
    TRY (ensure (ctx, ctx->gkeep, (size_t) n_keep * 4)); TRY (ensure (ctx, ctx->gkpos, (size_t) n_keep * 4));
    LAUNCH (K_GROUP, 16 * W * (size_t) n_keep, (head_kernel<W, 1, W><<<nblk (n_keep, RED_THREADS), RED_THREADS, 0, st>>> (krec, n_keep, heads)));
    TRY (exclusive_scan (ctx, heads, pos, n_keep, ptmp, ctx->d_totals + 0));
    LAUNCH (K_GROUP, 12 * (size_t) n_keep, (seg_start_kernel<<<nblk (n_keep, RED_THREADS), RED_THREADS, 0, st>>> (heads, pos, n_keep, seg)));
    LAUNCH (K_GROUP, 8 * (size_t) n_keep, (icopy_kernel<<<nblk (n_keep, RED_THREADS), RED_THREADS, 0, st>>> (kcount, n_keep, (u32*) ctx->cw.p)));
    TRY (exclusive_scan (ctx, (u32*) ctx->cw.p, (u32*) ctx->csum.p, n_keep, ptmp, ctx->d_totals + 2));
    CU (cudaMemcpyAsync (h_tot, ctx->d_totals, 4, cudaMemcpyDeviceToHost, st));
    CU (cudaStreamSynchronize (st));
    const u32 n_groups = h_tot[0];
    LAUNCH (K_GROUP, 12 * (size_t) n_groups, (group_keep_kernel<<<nblk (n_groups, RED_THREADS), RED_THREADS, 0, st>>> (seg, n_groups, n_keep, (u32*) ctx->csum.p, ctx->d_totals + 2, min_coverage, (u32*) ctx->gkeep.p)));
    TRY (exclusive_scan (ctx, (u32*) ctx->gkeep.p, (u32*) ctx->gkpos.p, n_groups, ptmp, ctx->d_totals + 3));
    CU (cudaMemcpyAsync (h_tot, ctx->d_totals + 3, 4, cudaMemcpyDeviceToHost, st));
    LAUNCH (K_GROUP, 16 * (size_t) n_groups, (group_emit_kernel<<<nblk (n_groups, RED_THREADS), RED_THREADS, 0, st>>> (seg, (u32*) ctx->gkeep.p, (u32*) ctx->gkpos.p, n_groups, n_keep, (int*) ctx->out_ii.p + start, (int*) ctx->out_if.p + start)));
    CU (cudaStreamSynchronize (st));
    n_idx = h_tot[0];
  }
  uint4 stat = make_uint4 (n_dist, n_keep, n_idx, 0);
  CU (cudaMemcpyAsync (ctx->d_bstat + b, &stat, sizeof (stat), cudaMemcpyHostToDevice, st));
  CU (cudaStreamSynchronize (st));
  return 0;
}

static DigitList digit_list (int W, int k)
{
  DigitList dl;
  std::vector<DigitDesc> v = tract_digits (W, k);
  dl.n = (int) v.size ();
  for (int i = 0; i < dl.n; i++) { dl.word[i] = (unsigned char) v[i].word; dl.shift[i] = (unsigned char) v[i].shift; dl.nbits[i] = (unsigned char) v[i].nbits; }
  return dl;
}

// the result arena: pinned host memory the device can store into (results are written by pack_kernel, no copy engine
// and no staging copy on the way); grown geometrically, contents valid until the context counts its next sample
static int ensure_arena (tjb_ctx *ctx, size_t keep, size_t idx)
{
  if (keep <= ctx->arena_keep_cap && idx <= ctx->arena_idx_cap) return 0;
  CU (cudaStreamSynchronize (ctx->stream));
  if (ctx->h_arena) CU (cudaFreeHost (ctx->h_arena));
  ctx->h_arena = ctx->d_arena = nullptr;
  const size_t kc = keep + keep / 4 + 4096 > ctx->arena_keep_cap ? keep + keep / 4 + 4096 : ctx->arena_keep_cap;
  const size_t ic = idx + idx / 4 + 4096 > ctx->arena_idx_cap ? idx + idx / 4 + 4096 : ctx->arena_idx_cap;
  ctx->arena_keep_cap = ctx->arena_idx_cap = 0;
  CU (cudaHostAlloc ((void**) &ctx->h_arena, kc * 40 + ic * 8, cudaHostAllocMapped | cudaHostAllocPortable));
  CU (cudaHostGetDevicePointer ((void**) &ctx->d_arena, ctx->h_arena, 0));
  ctx->arena_keep_cap = kc; ctx->arena_idx_cap = ic;
  return 0;
}

#define TJB_INTERNAL_RETRY 100 // the sample did not fit the table / lists it was sized for; ctx->want_entries is set

// the scan's records into the aggregation table, once (grid sized for the record capacity: the count is on the device)
template <int W>
static int aggregate_launch (tjb_ctx *ctx)
{
  if (!ctx->records_pending) return 0;
  const size_t want = nblk (ctx->rec_capacity < ((size_t) 1 << 24) ? ctx->rec_capacity : ((size_t) 1 << 24), 256 * AGG_UNROLL);
  const u32 g = (u32) (want < (size_t) ctx->sm_count * 8 ? (want ? want : 1) : (size_t) ctx->sm_count * 8);
  LAUNCH (K_AGGREGATE, 0, (aggregate_kernel<W><<<g, 256, 0, ctx->stream>>> ((const u64*) ctx->recs.p, ctx->d_res->counters, ctx->rec_capacity, table_of (ctx), (u64*) ctx->ovf.p, ctx->ovf_capacity, ctx->opt.kmer_size, ctx->d_bin_hist, ctx->bin_bits)));
  CU (cudaGetLastError ());
  ctx->records_pending = false;
  ctx->records_aggregated = true;
  return 0;
}

// plan -> partition -> bucket reduce -> totals, all asynchronous (the scalars stay on the device)
template <int W>
static int reduce_launch (tjb_ctx *ctx, int qc_mode, int min_coverage)
{
  const int k = ctx->opt.kmer_size, F = ctx->bin_bits;
  cudaStream_t st = ctx->stream;
  const unsigned long long cap = ctx->list_cap;
  TRY (ensure (ctx, ctx->wkey, (size_t) cap * W * 8)); TRY (ensure (ctx, ctx->wcnt, (size_t) cap * 8)); TRY (ensure (ctx, ctx->okey, (size_t) cap * W * 8));
  TRY (ensure (ctx, ctx->out_cnt, (size_t) cap * 4)); TRY (ensure (ctx, ctx->out_ii, (size_t) cap * 4)); TRY (ensure (ctx, ctx->out_if, (size_t) cap * 4));
  const u32 grid = (u32) ctx->sm_count;
  // per-bucket tables: a bucket starts where the running record count crosses a multiple of 2^bucket_shift
  const u32 max_buckets = (u32) (cap >> ctx->bucket_shift) + 2;
  TRY (ensure (ctx, ctx->b_bkt_start, ((size_t) max_buckets + 1) * 4));
  TRY (ensure (ctx, ctx->b_cursor, (size_t) max_buckets * BK_CURSOR_STRIDE * 4));
  TRY (ensure (ctx, ctx->b_koff, (size_t) max_buckets * 4)); TRY (ensure (ctx, ctx->b_ioff, (size_t) max_buckets * 4));
  TRY (ensure (ctx, ctx->b_bstat, (size_t) max_buckets * sizeof (uint4)));
  ctx->d_bkt_start = (u32*) ctx->b_bkt_start.p; ctx->d_cursor = (u32*) ctx->b_cursor.p;
  ctx->d_koff = (u32*) ctx->b_koff.p; ctx->d_ioff = (u32*) ctx->b_ioff.p; ctx->d_bstat = (uint4*) ctx->b_bstat.p;
  ctx->max_buckets = max_buckets;
  u32 *d_meta = ctx->d_res->meta;
  unsigned long long *d_counters = ctx->d_res->counters;
  TRY ((aggregate_launch<W> (ctx)));
  {
    const int ipt = F > 10 ? 1 << (F - 10) : 1; // bins per thread of the single planning CTA
#define PLAN_CASE(I) case I: LAUNCH (K_PLAN, 10 << F, (plan_kernel<I><<<1, PLAN_THREADS, 0, st>>> (ctx->d_bin_hist, F, ctx->bucket_shift, cap, ctx->d_bkt_start, ctx->d_lut, ctx->d_cursor, d_meta, max_buckets))); break;
    if (F > PLAN_REG_BIN_BITS)
      LAUNCH (K_PLAN, 18 << F, (plan_big_kernel<<<1, PLAN_THREADS, 0, st>>> (ctx->d_bin_hist, F, ctx->bucket_shift, cap, ctx->d_bkt_start, ctx->d_lut, ctx->d_cursor, d_meta, max_buckets)));
    else
    switch (ipt) { PLAN_CASE (1) PLAN_CASE (2) PLAN_CASE (4) PLAN_CASE (8) PLAN_CASE (16) PLAN_CASE (32) PLAN_CASE (64) }
#undef PLAN_CASE
  }
  {
    const size_t want = nblk (ctx->tab_slots, 256 * PART_UNROLL);
    const u32 g = (u32) (want < (size_t) grid * PART_GRID ? (want ? want : 1) : (size_t) grid * PART_GRID);
    LAUNCH (K_PARTITION, 8 * ctx->tab_slots, (wpartition_kernel<W><<<g, 256, 0, st>>> (table_of (ctx), (const u64*) ctx->ovf.p, d_counters, ctx->ovf_capacity, k, F, ctx->d_lut, ctx->d_cursor, (u64*) ctx->wkey.p, (unsigned long long*) ctx->wcnt.p, cap)));
  }
  LAUNCH (K_BUCKET, 0, (wbucket_reduce_kernel<W><<<grid * (W == 2 ? BK_CTAS2 : 2), BK_THREADS, wbucket_smem_bytes<W> (), st>>> ((const u64*) ctx->wkey.p, (const unsigned long long*) ctx->wcnt.p, ctx->d_bkt_start, d_meta, cap, digit_list (W, k), qc_mode, min_coverage, (u64*) ctx->okey.p, (u32*) ctx->out_cnt.p, (u32*) ctx->out_ii.p, (u32*) ctx->out_if.p, ctx->d_bstat)));
  LAUNCH (K_TOTALS, 0, (totals_kernel<<<1, PLAN_THREADS, 0, st>>> (ctx->d_bstat, d_meta, ctx->d_koff, ctx->d_ioff)));
  CU (cudaGetLastError ());
  return 0;
}

// after a synchronisation that brought ctx->h_res back: did the sample fit what it was sized for?
static bool sample_overflowed (tjb_ctx *ctx)
{
  const ResultBlock *hr = ctx->h_res;
  const unsigned long long n_ovf = hr->counters[CTR_OVERFLOW], n_ent = hr->counters[CTR_ENTRIES];
  if (hr->counters[CTR_RECORDS] > ctx->rec_capacity) { // the record array was too small: the rest of the pipeline saw a part of the sample
    ctx->want_records = (size_t) hr->counters[CTR_RECORDS] + 1024;
    return true;
  }
  if (n_ovf > ctx->ovf_capacity || n_ent > ctx->tab_slots / 2 || n_ent + n_ovf > ctx->list_cap) {
    // the table or the overflow list was too small for this sample: everything after the scan worked on a part of it
    ctx->want_entries = (size_t) (n_ent + n_ovf) * 2 + 4096;
    if (ctx->want_entries < (size_t) n_ovf * 8 + 4096) ctx->want_entries = (size_t) n_ovf * 8 + 4096; // the overflow list is an eighth of it
    return true;
  }
  ctx->last_entries = (size_t) (n_ent + n_ovf);
  return false;
}

// rare: buckets that do not fit in shared memory go through the global multi-pass path; totals again, h_res refreshed
template <int W>
static int reduce_oversize_buckets (tjb_ctx *ctx, int qc_mode, int min_coverage)
{
  cudaStream_t st = ctx->stream;
  const u32 n_buckets = ctx->h_res->meta[META_NBUCKETS];
  std::vector<u32> h_start (n_buckets + 1);
  std::vector<uint4> h_stat (n_buckets);
  CU (cudaMemcpyAsync (h_start.data (), ctx->d_bkt_start, (n_buckets + 1) * sizeof (u32), cudaMemcpyDeviceToHost, st));
  CU (cudaMemcpyAsync (h_stat.data (), ctx->d_bstat, n_buckets * sizeof (uint4), cudaMemcpyDeviceToHost, st));
  CU (cudaStreamSynchronize (st));
  for (u32 b = 0; b < n_buckets; b++)
    if (h_stat[b].w) TRY ((reduce_oversize_bucket<W> (ctx, qc_mode, min_coverage, ctx->opt.kmer_size, b, h_start[b], h_start[b + 1] - h_start[b])));
  LAUNCH (K_TOTALS, 0, (totals_kernel<<<1, PLAN_THREADS, 0, st>>> (ctx->d_bstat, ctx->d_res->meta, ctx->d_koff, ctx->d_ioff)));
  CU (cudaMemcpyAsync (ctx->h_res, ctx->d_res, sizeof (ResultBlock), cudaMemcpyDeviceToHost, st));
  CU (cudaStreamSynchronize (st));
  return 0;
}

// reduce -> pack + coverage -> ONE synchronisation (a second one only when a sample turns out larger than the arena
// it was packed into, or holds buckets that do not fit in shared memory)
template <int W>
static int finalise_t (tjb_ctx *ctx, tjb_counter_out *out, int qc_mode)
{
  const tjb_opts *opt = &ctx->opt;
  const int k = opt->kmer_size;
  cudaStream_t st = ctx->stream;
  const unsigned long long cap = ctx->list_cap;
  const u32 grid = (u32) ctx->sm_count;
  Trace tr;
  TRY ((reduce_launch<W> (ctx, qc_mode, opt->min_coverage)));
  u32 *d_meta = ctx->d_res->meta;
  // elements + group indices + coverage, into the arena (capacity: the previous sample's counts + 25 %, at least 4 Ki)
  const int cov_init = INT_MIN;
  auto launch_pack = [&] (bool again) -> int {
    const u32 max_keep = (u32) ctx->arena_keep_cap, max_idx = (u32) ctx->arena_idx_cap;
    u32 slots = 1024;
    while (slots < 4u * max_keep) slots <<= 1; // two keys per record, load factor <= 0.5
    TRY (ensure (ctx, ctx->cov_tab, (size_t) slots * 8));
    u32 *cov_keys = (u32*) ctx->cov_tab.p, *cov_vals = cov_keys + slots;
    u64 *o40 = (u64*) ctx->d_arena;
    int *oi = (int*) (ctx->d_arena + (size_t) max_keep * 40), *of = oi + max_idx;
    if (ctx->result_dma) {
      TRY (ensure (ctx, ctx->out40, (size_t) max_keep * 40)); TRY (ensure (ctx, ctx->idx_i, (size_t) max_idx * 4)); TRY (ensure (ctx, ctx->idx_f, (size_t) max_idx * 4));
      o40 = (u64*) ctx->out40.p; oi = (int*) ctx->idx_i.p; of = (int*) ctx->idx_f.p;
    }
    CU (cudaMemsetAsync (cov_keys, 0, (size_t) slots * 8, st)); // keys and sums
    if (again) CU (cudaMemcpyAsync (&ctx->d_res->cov, &cov_init, 4, cudaMemcpyHostToDevice, st)); // (the prologue set it for the first launch)
    LAUNCH (K_PACK, (8 * W + 4 + 40) * (size_t) ctx->last_keep + 16 * (size_t) ctx->last_idx,
            (pack_kernel<W><<<grid * 4, 256, 0, st>>> ((const u64*) ctx->okey.p, (u32*) ctx->out_cnt.p, (u32*) ctx->out_ii.p, (u32*) ctx->out_if.p, ctx->d_bkt_start, ctx->d_bstat, ctx->d_koff, ctx->d_ioff, d_meta, k, o40, oi, of, cov_keys, cov_vals, slots - 1, max_keep, max_idx, cap, &ctx->d_res->cov, d_meta)));
    ctx->cov_slots = slots;
    CU (cudaGetLastError ());
    if (ctx->result_dma) { // one copy of what the previous sample needed (+ 12 %); the rest follows after the counts if it was more
      const size_t sk = ctx->last_keep ? ctx->last_keep + ctx->last_keep / 8 + 256 : 0, si = ctx->last_idx ? ctx->last_idx + ctx->last_idx / 8 + 256 : 0;
      const size_t ck = sk < max_keep ? sk : max_keep, ci = si < max_idx ? si : max_idx;
      if (ck) CU (cudaMemcpyAsync (ctx->h_arena, ctx->out40.p, ck * 40, cudaMemcpyDeviceToHost, st));
      if (ci) {
        CU (cudaMemcpyAsync (ctx->h_arena + (size_t) max_keep * 40, ctx->idx_i.p, ci * 4, cudaMemcpyDeviceToHost, st));
        CU (cudaMemcpyAsync (ctx->h_arena + (size_t) max_keep * 40 + (size_t) max_idx * 4, ctx->idx_f.p, ci * 4, cudaMemcpyDeviceToHost, st));
      }
    }
    CU (cudaMemcpyAsync (ctx->h_res, ctx->d_res, sizeof (ResultBlock), cudaMemcpyDeviceToHost, st));
    return 0;
  };
  TRY (ensure_arena (ctx, ctx->last_keep + ctx->last_keep / 8 + 256, ctx->last_idx + ctx->last_idx / 8 + 256));
  TRY (launch_pack (false));
  tr.mark ("enqueue");
  // The result arrays of the default mode are malloc()ed (the reference free()s them) and fresh heap memory costs a page
  // fault per 4 KiB on first touch (measured: up to 1.2 ms for 2.8 MB, depending on what the allocator has cached).
  // The host has nothing to do until the kernels above finish, so arrays sized like the previous sample's result
  // (+12 %) are allocated and touched now; they are trimmed or replaced once the real sizes are known.
  struct Spec { void *elem = nullptr, *ii = nullptr, *fi = nullptr; size_t keep = 0, idx = 0;
                ~Spec () { free (elem); free (ii); free (fi); } } spec;
  const bool pool_copy = ctx->result_mode == TJB_RESULT_MALLOC && ctx->last_keep * 40 >= ((size_t) 512 << 10) && host_pool () != nullptr;
  PackerSession copy_session (pool_copy);
  if (ctx->result_mode == TJB_RESULT_MALLOC && ctx->last_keep && ctx->last_idx) {
    spec.keep = ctx->last_keep + ctx->last_keep / 8 + 256;
    spec.idx = ctx->last_idx + ctx->last_idx / 8 + 256;
    spec.elem = malloc (spec.keep * 40); spec.ii = malloc (spec.idx * 4); spec.fi = malloc (spec.idx * 4);
    if (spec.elem && spec.ii && spec.fi) {
      auto touch = [] (void *p, size_t n) { // one batched fault-in (Linux >= 5.14) instead of a trap per page
        const uintptr_t lo = (uintptr_t) p & ~(uintptr_t) 4095, hi = ((uintptr_t) p + n + 4095) & ~(uintptr_t) 4095;
#ifdef MADV_POPULATE_WRITE
        if (madvise ((void*) lo, hi - lo, MADV_POPULATE_WRITE) == 0) return;
#else
        if (madvise ((void*) lo, hi - lo, 23 /* MADV_POPULATE_WRITE */) == 0) return;
#endif
        volatile char *c = (volatile char*) p;
        for (size_t o = 0; o < n; o += 4096) c[o] = 0;
        c[n - 1] = 0;
      };
      touch (spec.elem, spec.keep * 40); touch (spec.ii, spec.idx * 4); touch (spec.fi, spec.idx * 4);
    } else spec.keep = spec.idx = 0;
  }
  tr.mark ("prealloc");
  ctx->t1_recorded = false;
  if (ctx->profiling >= 2) { CU (cudaEventRecord (ctx->ev_t1, st)); ctx->t1_recorded = true; } // end of the sample's device work: no second round trip for it
  CU (cudaStreamSynchronize (st)); // ---- the synchronisation of the sample
  tr.mark ("sync");
  const ResultBlock *hr = ctx->h_res;
  if (sample_overflowed (ctx)) return TJB_INTERNAL_RETRY;
  out->n_raw = (int64_t) hr->counters[CTR_NRAW];
  out->n_skipped_nonacgt = (int64_t) hr->counters[CTR_SKIP];
  if (hr->counters[CTR_NRAW] == 0) { out->status = TJB_NO_HT; return 0; }
  if (hr->counters[CTR_NRAW] > (unsigned long long) INT32_MAX) { ctx->err = "more than INT32_MAX tract observations"; return TJB_ERR_TOO_LARGE; }
  const u32 n_buckets = hr->meta[META_NBUCKETS];
  bool repack = false;
  if (hr->meta[META_NOVERSIZE]) { TRY ((reduce_oversize_buckets<W> (ctx, qc_mode, opt->min_coverage))); repack = true; }
  const u32 n_keep = hr->meta[META_NKEEP], n_idx = hr->meta[META_NIDX];
  out->n_distinct = hr->meta[META_NDIST];
  // algorithmic bytes of the launches above, now that the counts are known
  ctx->stats[K_SCAN].bytes_algorithmic += (int64_t) 8 * W * out->n_raw;          // the scan writes one record per observation
  ctx->stats[K_AGGREGATE].bytes_algorithmic += (int64_t) (16 * W + 16) * out->n_raw; // record read, key compared, count word updated
  ctx->stats[K_PARTITION].bytes_algorithmic += (int64_t) (16 * W + 16) * (int64_t) ctx->last_entries;
  ctx->stats[K_BUCKET].bytes_algorithmic += (int64_t) (8 * W + 8) * (int64_t) ctx->last_entries + (int64_t) (8 * W + 4) * n_keep + 8 * (int64_t) n_idx;
  ctx->stats[K_TOTALS].bytes_algorithmic += (int64_t) 24 * n_buckets;
  if (!n_keep) { out->status = TJB_NO_HT_AFTER_QC; return 0; }
  if (!n_idx) { out->status = TJB_NO_CTX_ABOVE_COVERAGE; return 0; }
  // ---- reference-layout elements, global group indices, coverage (:420-438) ---------------------------------
  const size_t sent_keep = ctx->last_keep ? ctx->last_keep + ctx->last_keep / 8 + 256 : 0, sent_idx = ctx->last_idx ? ctx->last_idx + ctx->last_idx / 8 + 256 : 0;
  if (repack || n_keep > ctx->arena_keep_cap || n_idx > ctx->arena_idx_cap) {
    TRY (ensure_arena (ctx, n_keep, n_idx)); // not packed completely: the arena was sized for a smaller sample
    ctx->last_keep = n_keep; ctx->last_idx = n_idx;
    TRY (launch_pack (true));
    CU (cudaStreamSynchronize (st));
  } else if (ctx->result_dma && (n_keep > sent_keep || n_idx > sent_idx)) { // packed, but only a part has been copied
    const size_t mk = ctx->arena_keep_cap, mi = ctx->arena_idx_cap;
    CU (cudaMemcpyAsync (ctx->h_arena, ctx->out40.p, (size_t) n_keep * 40, cudaMemcpyDeviceToHost, st));
    CU (cudaMemcpyAsync (ctx->h_arena + mk * 40, ctx->idx_i.p, (size_t) n_idx * 4, cudaMemcpyDeviceToHost, st));
    CU (cudaMemcpyAsync (ctx->h_arena + mk * 40 + mi * 4, ctx->idx_f.p, (size_t) n_idx * 4, cudaMemcpyDeviceToHost, st));
    CU (cudaStreamSynchronize (st));
  }
  if (hr->meta[META_COVSCAN]) { // a count wrapped its 20-bit field: the running maximum is not the answer, scan the table
    const int cov_init2 = INT_MIN;
    u32 *cov_keys = (u32*) ctx->cov_tab.p, *cov_vals = cov_keys + ctx->cov_slots;
    CU (cudaMemcpyAsync (&ctx->d_res->cov, &cov_init2, 4, cudaMemcpyHostToDevice, st));
    LAUNCH (K_COVMAX, 8 * (size_t) ctx->cov_slots, (cov_max_kernel<<<nblk (ctx->cov_slots, 256 * 8), 256, 0, st>>> (cov_keys, cov_vals, ctx->cov_slots, &ctx->d_res->cov)));
    CU (cudaMemcpyAsync (ctx->h_res, ctx->d_res, sizeof (ResultBlock), cudaMemcpyDeviceToHost, st));
    CU (cudaStreamSynchronize (st));
  }
  tr.mark ("pack");
  ctx->last_keep = n_keep; ctx->last_idx = n_idx;
  const char *a_elem = ctx->h_arena, *a_ii = ctx->h_arena + ctx->arena_keep_cap * 40, *a_fi = a_ii + ctx->arena_idx_cap * 4;
  if (ctx->result_mode == TJB_RESULT_ARENA) { // the caller reads the arena; valid until this context counts its next sample
    out->elem = (void*) a_elem; out->idx_initial = (int32_t*) a_ii; out->idx_final = (int32_t*) a_fi;
    out->arena = 1;
  } else {
    if (spec.keep >= n_keep && spec.idx >= n_idx) { // allocated ahead: give back what is not needed
      void *e = realloc (spec.elem, (size_t) n_keep * 40), *a = realloc (spec.ii, (size_t) n_idx * 4), *b = realloc (spec.fi, (size_t) n_idx * 4);
      out->elem = e ? e : spec.elem; out->idx_initial = (int32_t*) (a ? a : spec.ii); out->idx_final = (int32_t*) (b ? b : spec.fi);
      spec.elem = spec.ii = spec.fi = nullptr;
    } else {
      out->elem = malloc ((size_t) n_keep * 40);
      out->idx_initial = (int32_t*) malloc ((size_t) n_idx * 4);
      out->idx_final = (int32_t*) malloc ((size_t) n_idx * 4);
    }
    if (!out->elem || !out->idx_initial || !out->idx_final) { ctx->err = "malloc failed"; return TJB_ERR_NOMEM; }
    if (pool_copy) {
      std::lock_guard<std::mutex> lk (g_packer_run_mu);
      tjb_packer_copy (g_packer, out->elem, a_elem, (size_t) n_keep * 40);
    } else memcpy (out->elem, a_elem, (size_t) n_keep * 40);
    memcpy (out->idx_initial, a_ii, (size_t) n_idx * 4);
    memcpy (out->idx_final, a_fi, (size_t) n_idx * 4);
    tr.mark ("copy");
  }
  out->n_elem = (int32_t) n_keep;
  out->n_idx = (int32_t) n_idx;
  out->coverage = hr->cov;
  out->status = TJB_OK;
  return 0;
}

static int finalise (tjb_ctx *ctx, tjb_counter_out *out)
{
  const int qc = qc_mode_of (&ctx->opt);
  int rc = ctx->rec_words == 2 ? finalise_t<2> (ctx, out, qc) : finalise_t<3> (ctx, out, qc);
  ctx->in_sample = false;
  cudaStreamSynchronize (ctx->stream);
  resolve_events (ctx);
  if (rc == TJB_INTERNAL_RETRY) return rc;
  if (rc < 0 || out->status != TJB_OK) { // the reference leaves an "empty counter" (n_elem = 0)
    if (!out->arena) { free (out->elem); free (out->idx_initial); free (out->idx_final); }
    out->elem = nullptr; out->idx_initial = out->idx_final = nullptr;
    out->n_elem = out->n_idx = 0; out->coverage = 0; out->arena = 0;
  }
  return rc;
}

static void out_reset (tjb_counter_out *out) { memset (out, 0, sizeof (*out)); }

// device time of the call between ev_t0 and the end of the sample's device work (diagnostics, profiling level 2)
static int device_time (tjb_ctx *ctx, tjb_counter_out *out)
{
  if (ctx->profiling < 2) return 0; // level 1 times the scan kernel only (bench.py inside its timed region): nothing else is recorded
  if (!ctx->t1_recorded) CU (cudaEventRecord (ctx->ev_t1, ctx->stream)); // (paths that do not end in finalise ())
  CU (cudaEventSynchronize (ctx->ev_t1));
  ctx->t1_recorded = false;
  float ms = 0.f;
  cudaEventElapsedTime (&ms, ctx->ev_t0, ctx->ev_t1);
  out->ms_device = ms;
  return 0;
}

// ---- union of query rows over samples (tjb_union.cuh): the rows go through the aggregation table and the bucket
// reduce like the observations of a sample, with nothing filtered --------------------------------------------------
template <int W>
static int union_t (tjb_ctx *ctx, const u64 *d_rows, u32 n, int k, u64 *d_union, int64_t *n_union, int *d_index)
{
  cudaStream_t st = ctx->stream;
  tjb_opts opt = {k, 2, 0, 0, 0};
  for (int attempt = 0; attempt < 4; attempt++) {
    if (!ctx->want_entries) ctx->want_entries = (size_t) n + 1024; // every row distinct in the worst case
    const size_t keep_entries = ctx->last_entries;                 // sizes of the samples this context counts stay as they are
    TRY (begin_sample (ctx, &opt, 0));
    LAUNCH (K_UNION, (24 + 8 * W + 16) * (size_t) n, (rows_insert_kernel<W><<<nblk (n, UNI_THREADS), UNI_THREADS, 0, st>>> (d_rows, n, k, table_of (ctx), (u64*) ctx->ovf.p, ctx->ovf_capacity, ctx->d_res->counters, ctx->d_bin_hist, ctx->bin_bits)));
    TRY ((reduce_launch<W> (ctx, QC_KEEP_ALL, INT_MIN)));
    auto emit = [&] () -> int {
      LAUNCH (K_UNION, (8 * W + 24) * (size_t) n, (union_rows_kernel<W><<<(u32) ctx->sm_count * 4, UNI_THREADS, 0, st>>> ((const u64*) ctx->okey.p, ctx->d_bkt_start, ctx->d_bstat, ctx->d_koff, ctx->d_res->meta, k, d_union, n, ctx->list_cap)));
      if (d_index)
        LAUNCH (K_UNION, (24 + 4) * (size_t) n, (union_index_kernel<<<nblk (n, UNI_THREADS), UNI_THREADS, 0, st>>> (d_rows, n, d_union, ctx->d_res->meta, n, d_index)));
      CU (cudaGetLastError ());
      CU (cudaMemcpyAsync (ctx->h_res, ctx->d_res, sizeof (ResultBlock), cudaMemcpyDeviceToHost, st));
      CU (cudaStreamSynchronize (st));
      return 0;
    };
    TRY (emit ());
    ctx->in_sample = false;
    const bool redo = sample_overflowed (ctx);
    ctx->last_entries = keep_entries;
    if (redo) continue;
    if (ctx->h_res->meta[META_NOVERSIZE]) { TRY ((reduce_oversize_buckets<W> (ctx, QC_KEEP_ALL, INT_MIN))); TRY (emit ()); }
    resolve_events (ctx);
    *n_union = ctx->h_res->meta[META_NKEEP];
    return 0;
  }
  ctx->err = "union: table overflow persisted";
  return TJB_ERR_CUDA;
}

// ---- a sample's table as rows, and rows added to a table (one sample counted in pieces on several GPUs) ------------
// table entries + overflow list -> rows of W + 1 words {key words without the flag bits, forward | reverse << 32}
template <int W>
__global__ void __launch_bounds__ (256)
table_export_kernel (AggTable tab, const u64 *__restrict__ ovf, const unsigned long long *__restrict__ counters, unsigned long long ovf_capacity,
                     u64 *__restrict__ rows, unsigned long long *__restrict__ n_rows, unsigned long long max_rows)
{
  const size_t stride = (size_t) gridDim.x * 256, n_slots = (size_t) tab.mask + 1;
  const unsigned long long n_ovf = counters[CTR_OVERFLOW] < ovf_capacity ? counters[CTR_OVERFLOW] : ovf_capacity;
  for (size_t i = (size_t) blockIdx.x * 256 + threadIdx.x; i < n_slots + n_ovf; i += stride) {
    Rec<W> key;
    unsigned long long c;
    if (i < n_slots) {
      c = tab.cnt[i];
      if (!c) continue;
      key = table_key_at<W> (tab, (u32) i);
    } else {
      key = load_rec<W> (ovf, i - n_slots);
      c = flag_weight (key.w[0]);
      key.w[0] &= ~3ULL;
    }
    const unsigned long long pos = atomicAdd (n_rows, 1ULL);
    if (pos < max_rows) {
#pragma unroll
      for (int j = 0; j < W; j++) rows[(size_t) (W + 1) * pos + j] = key.w[j];
      rows[(size_t) (W + 1) * pos + W] = c;
    }
  }
}

// rows of another rank's table added to this context's table (sum of counts per strand = what one scan of all the
// chunks would have counted: + and | are associative)
template <int W>
__global__ void __launch_bounds__ (256)
table_import_kernel (const u64 *__restrict__ rows, unsigned long long n, int k, AggTable tab, u64 *__restrict__ ovf, unsigned long long ovf_capacity,
                     unsigned long long *__restrict__ counters, u32 *__restrict__ bin_hist, int bin_bits)
{
  for (size_t i = (size_t) blockIdx.x * 256 + threadIdx.x; i < n; i += (size_t) gridDim.x * 256) {
    Rec<W> key;
#pragma unroll
    for (int j = 0; j < W; j++) key.w[j] = rows[(size_t) (W + 1) * i + j];
    unsigned long long c = rows[(size_t) (W + 1) * i + W];
    const int res = table_add<W> (tab, key, c);
    if (res == TAB_NEW) { atomicAdd (&counters[CTR_ENTRIES], 1ULL); atomicAdd (&bin_hist[rec_bin<W> (key, k, bin_bits)], 1u); }
    else if (res == TAB_DEFER) {
      // The overflow list holds observations of weight 1, not weighted rows.  An entry is only busy for the few
      // instructions between its claim and its publication, so the row is simply tried again (bounded: nobody is
      // waited for); if it still cannot be placed the table is too full and the sample is flagged for a redo.
      int again = TAB_DEFER;
      for (int t = 0; t < 64 && again == TAB_DEFER; t++) { __nanosleep (200); again = table_add<W> (tab, key, c); }
      if (again == TAB_NEW) { atomicAdd (&counters[CTR_ENTRIES], 1ULL); atomicAdd (&bin_hist[rec_bin<W> (key, k, bin_bits)], 1u); }
      else if (again == TAB_DEFER) atomicAdd (&counters[CTR_OVERFLOW], ovf_capacity + 1);
    }
  }
  (void) ovf;
}

// ---- file feed: one decoder thread per FASTQ file, two pinned chunks each -----------------
struct FeedSlot { uint8_t *pin = nullptr; cudaEvent_t free_ev = nullptr; int64_t got = 0; int state = 0; /* 0 empty, 1 full */
                  const uint8_t *long_read = nullptr; /* got bytes in the reader's own buffer instead of pin */ };
struct FileFeed {
  std::mutex mu;
  std::condition_variable cv;
  FeedSlot slot[2][2];
  int64_t n_reads[2] = {0, 0}, n_bases[2] = {0, 0};
  double ms_decode[2] = {0., 0.};
  bool abort = false;
};
static void file_feed_worker (FileFeed *f, int p, tjb_reader *rd, int device)
{
  cudaSetDevice (device);
  for (int i = 0;; i ^= 1) {
    FeedSlot &s = f->slot[p][i];
    {
      std::unique_lock<std::mutex> lk (f->mu);
      f->cv.wait (lk, [&] { return s.state == 0 || f->abort; });
      if (f->abort) return;
    }
    cudaEventSynchronize (s.free_ev); // the copy engine has drained this pinned chunk (no-op before its first use)
    auto t0 = std::chrono::steady_clock::now ();
    int64_t got = tjb_reader_fill (rd, s.pin, TJB_CHUNK_BYTES, &f->n_reads[p], &f->n_bases[p]);
    const uint8_t *long_read = nullptr;
    if (got == TJB_ERR_TOO_LARGE) { // a read longer than a chunk: handed over where it lies in the reader
      const int64_t len = tjb_reader_take_long (rd, &long_read, &f->n_reads[p], &f->n_bases[p]);
      if (len > 0) got = len; else long_read = nullptr;
    }
    f->ms_decode[p] += std::chrono::duration<double, std::milli> (std::chrono::steady_clock::now () - t0).count ();
    { std::lock_guard<std::mutex> lk (f->mu); s.got = got; s.long_read = long_read; s.state = 1; }
    f->cv.notify_all ();
    if (got <= 0) return;
    if (long_read) { // the reader's buffer must stay as it is until the consumer has copied it
      std::unique_lock<std::mutex> lk (f->mu);
      f->cv.wait (lk, [&] { return s.state == 0 || f->abort; });
      if (f->abort) return;
    }
  }
}

// sequence bytes to expect from a FASTQ/FASTA file, from its size: half of a FASTQ is qualities, gzip -1 shrinks
// such a file ~4x.  Only sizes the aggregation table of a context's first sample (a wrong guess costs a redo).
static size_t sequence_bytes_hint (const char *const *files, int n_files)
{
  size_t total = 0;
  for (int f = 0; f < n_files; f++) {
    struct stat sb;
    if (stat (files[f], &sb) != 0 || sb.st_size <= 0) return 0;
    const size_t nl = strlen (files[f]);
    const bool gz = nl > 3 && strcmp (files[f] + nl - 3, ".gz") == 0;
    total += (size_t) sb.st_size * (gz ? 4 : 1) / 2;
  }
  return total;
}

// ------------------------------------------------------------------------------------------- C ABI ------
extern "C" {

const char *tjb_version (void) { return TJB_VERSION_STR; }

const char *tjb_strerror (int code)
{
  switch (code) {
    case TJB_OK: return "ok";
    case TJB_NO_HT: return "no homopolymer tract found (sample excluded)";
    case TJB_NO_HT_AFTER_QC: return "no tract left after strand-bias / singleton filter (sample excluded)";
    case TJB_NO_CTX_ABOVE_COVERAGE: return "no context reaches the minimum coverage (sample excluded)";
    case TJB_ERR_CUDA: return "CUDA error (no usable sm_100 device? there is no CPU fallback)";
    case TJB_ERR_ARG: return "invalid argument";
    case TJB_ERR_IO: return "I/O error";
    case TJB_ERR_NOMEM: return "out of memory";
    case TJB_ERR_TOO_LARGE: return "input too large";
    case TJB_ERR_NOALLOC_LIMIT: return "batch larger than the fixed scratch of tjb_tract_rows_noalloc (TJB_NOALLOC_MAX_BYTES = 4 MiB of reads)";
    default: return "unknown code";
  }
}

const char *tjb_ctx_last_error (const tjb_ctx *ctx) { return ctx ? ctx->err.c_str () : "null context"; }

int tjb_ctx_create (int device, tjb_ctx **pctx)
{
  if (!pctx) return TJB_ERR_ARG;
  *pctx = nullptr;
  int n_dev = 0;
  if (cudaGetDeviceCount (&n_dev) != cudaSuccess || device < 0 || device >= n_dev) return TJB_ERR_CUDA;
  tjb_ctx *ctx = new tjb_ctx ();
  ctx->device = device;
  for (int i = 0; i < K_COUNT; i++) { memset (&ctx->stats[i], 0, sizeof (tjb_kernel_stat)); strncpy (ctx->stats[i].name, kKernelNames[i], 47); }
  auto fail = [&] (cudaError_t e) { (void) e; delete ctx; return TJB_ERR_CUDA; };
  cudaError_t e;
  if ((e = cudaSetDevice (device)) != cudaSuccess) return fail (e);
  cudaDeviceProp prop;
  if ((e = cudaGetDeviceProperties (&prop, device)) != cudaSuccess) return fail (e);
  if (prop.major < 10) { delete ctx; return TJB_ERR_CUDA; } // kernels are built for sm_100a only
  ctx->sm_count = prop.multiProcessorCount;
  if ((e = cudaStreamCreateWithFlags (&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess) return fail (e);
  if ((e = cudaMalloc ((void**) &ctx->d_res, sizeof (ResultBlock))) != cudaSuccess) return fail (e);
  if ((e = cudaMallocHost ((void**) &ctx->h_res, sizeof (ResultBlock))) != cudaSuccess) return fail (e);
  if ((e = cudaMalloc (&ctx->d_totals, 16 * sizeof (u32))) != cudaSuccess) return fail (e);
  if ((e = cudaMalloc (&ctx->d_tile_ctr, sizeof (u32))) != cudaSuccess) return fail (e);
  if ((e = cudaStreamCreateWithFlags (&ctx->copy_stream, cudaStreamNonBlocking)) != cudaSuccess) return fail (e);
  if ((e = cudaMalloc (&ctx->d_bin_hist, TJB_MAX_BINS * sizeof (u32))) != cudaSuccess) return fail (e);
  if ((e = cudaMalloc (&ctx->d_lut, TJB_MAX_BINS * sizeof (u32))) != cudaSuccess) return fail (e);
  if ((e = cudaFuncSetAttribute (tjb_scan_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) scan_smem_bytes<2> ())) != cudaSuccess) return fail (e);
  if ((e = cudaFuncSetAttribute (tjb_scan_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) scan_smem_bytes<3> ())) != cudaSuccess) return fail (e);
  if ((e = cudaFuncSetAttribute (wbucket_reduce_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) wbucket_smem_bytes<2> ())) != cudaSuccess) return fail (e);
  if ((e = cudaFuncSetAttribute (wbucket_reduce_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) wbucket_smem_bytes<3> ())) != cudaSuccess) return fail (e);
  cudaEventCreate (&ctx->ev_t0); cudaEventCreate (&ctx->ev_t1);
  cudaEventCreateWithFlags (&ctx->ev_ready, cudaEventDisableTiming);
  for (int i = 0; i < 2; i++) { cudaEventCreateWithFlags (&ctx->ev_copied[i], cudaEventDisableTiming); cudaEventCreateWithFlags (&ctx->ev_scanned[i], cudaEventDisableTiming); }
  const char *rp = getenv ("TJB_RESULT_PATH"); // "dma": results through a device staging array and the copy engine
  ctx->result_dma = rp && strcmp (rp, "dma") == 0;
  *pctx = ctx;
  return 0;
}

void tjb_ctx_destroy (tjb_ctx *ctx)
{
  if (!ctx) return;
  cudaSetDevice (ctx->device);
  if (ctx->stream) cudaStreamSynchronize (ctx->stream);
  resolve_events (ctx);
  for (cudaEvent_t e : ctx->free_events) cudaEventDestroy (e);
  for (int i = 0; i < 2; i++) {
    if (ctx->h_pk[i]) cudaFreeHost (ctx->h_pk[i]);
    if (ctx->h_exc_pos[i]) cudaFreeHost (ctx->h_exc_pos[i]);
    if (ctx->h_exc_val[i]) cudaFreeHost (ctx->h_exc_val[i]);
  }
  DevBuf *bufs[] = {&ctx->d_pk[0], &ctx->d_pk[1], &ctx->d_exc_pos[0], &ctx->d_exc_pos[1], &ctx->d_exc_val[0], &ctx->d_exc_val[1],
                    &ctx->d_chunk[0], &ctx->d_chunk[1], &ctx->d_long, &ctx->tab, &ctx->recs, &ctx->ovf, &ctx->wkey, &ctx->wcnt, &ctx->okey,
                    &ctx->hist, &ctx->ptmp, &ctx->heads, &ctx->pos, &ctx->seg, &ctx->drec, &ctx->dcount, &ctx->keep, &ctx->kpos, &ctx->krec,
                    &ctx->cw, &ctx->csum, &ctx->gkeep, &ctx->gkpos, &ctx->out40, &ctx->idx_i, &ctx->idx_f, &ctx->qrows, &ctx->out_cnt, &ctx->out_ii,
                    &ctx->out_if, &ctx->cov_tab, &ctx->uni_a, &ctx->uni_b, &ctx->b_bkt_start, &ctx->b_cursor, &ctx->b_koff, &ctx->b_ioff, &ctx->b_bstat};
  for (DevBuf *b : bufs) if (b->p) cudaFree (b->p);
  for (int i = 0; i < 4; i++) { if (ctx->h_pinned[i]) cudaFreeHost (ctx->h_pinned[i]); if (ctx->h_free[i]) cudaEventDestroy (ctx->h_free[i]); }
  void *small[] = {ctx->d_res, ctx->d_totals, ctx->d_tile_ctr, ctx->d_bin_hist, ctx->d_lut};
  for (void *q : small) if (q) cudaFree (q);
  if (ctx->h_res) cudaFreeHost (ctx->h_res);
  if (ctx->h_arena) cudaFreeHost (ctx->h_arena);
  if (ctx->copy_stream) { cudaStreamSynchronize (ctx->copy_stream); cudaStreamDestroy (ctx->copy_stream); }
  if (ctx->ev_ready) cudaEventDestroy (ctx->ev_ready);
  for (int i = 0; i < 2; i++) { if (ctx->ev_copied[i]) cudaEventDestroy (ctx->ev_copied[i]); if (ctx->ev_scanned[i]) cudaEventDestroy (ctx->ev_scanned[i]); }
  if (ctx->ev_t0) cudaEventDestroy (ctx->ev_t0);
  if (ctx->ev_t1) cudaEventDestroy (ctx->ev_t1);
  if (ctx->stream) cudaStreamDestroy (ctx->stream);
  delete ctx;
}

void *tjb_ctx_stream (tjb_ctx *ctx) { return ctx ? (void*) ctx->stream : nullptr; }

int tjb_ctx_set_profiling (tjb_ctx *ctx, int on) { if (!ctx) return TJB_ERR_ARG; ctx->profiling = on; return 0; }

int tjb_ctx_set_result_mode (tjb_ctx *ctx, int mode)
{
  if (!ctx || (mode != TJB_RESULT_MALLOC && mode != TJB_RESULT_ARENA)) return TJB_ERR_ARG;
  ctx->result_mode = mode;
  return 0;
}

int tjb_ctx_kernel_stats (tjb_ctx *ctx, tjb_kernel_stat *stats, int max_stats)
{
  if (!ctx) return TJB_ERR_ARG;
  int n = 0;
  for (int i = 0; i < K_COUNT; i++) if (ctx->stats[i].launches) { if (stats && n < max_stats) stats[n] = ctx->stats[i]; n++; }
  return n;
}

void tjb_ctx_reset_kernel_stats (tjb_ctx *ctx)
{
  if (!ctx) return;
  for (int i = 0; i < K_COUNT; i++) { ctx->stats[i].launches = 0; ctx->stats[i].ms_total = 0; ctx->stats[i].bytes_algorithmic = 0; }
  ctx->launches = 0;
  ctx->h2d_bytes = 0;
}

int64_t tjb_ctx_launch_count (const tjb_ctx *ctx) { return ctx ? ctx->launches : 0; }
int64_t tjb_ctx_htod_bytes (const tjb_ctx *ctx) { return ctx ? ctx->h2d_bytes : 0; }

#define TJB_MAX_ATTEMPTS 5 // a sample is redone with a larger table when it holds more distinct observations than it was sized for

int tjb_count_sample_device (tjb_ctx *ctx, const void *d_seqs, size_t n_bytes, const tjb_opts *opt, tjb_counter_out *out)
{
  if (!ctx || !out || (n_bytes && !d_seqs)) return TJB_ERR_ARG;
  if (((uintptr_t) d_seqs) & 15) { ctx->err = "device buffer must be 16-byte aligned"; return TJB_ERR_ARG; }
  int rc = 0;
  Trace tr;
  for (int attempt = 0; attempt < TJB_MAX_ATTEMPTS; attempt++) {
    out_reset (out);
    TRY (begin_sample (ctx, opt, n_bytes));
    if (ctx->profiling >= 2) CU (cudaEventRecord (ctx->ev_t0, ctx->stream));
    TRY (launch_scan (ctx, (const uint8_t*) d_seqs, n_bytes));
    tr.mark ("device:scan_enqueue");
    rc = finalise (ctx, out);
    tr.mark ("finalise");
    if (rc != TJB_INTERNAL_RETRY) break;
  }
  if (rc == TJB_INTERNAL_RETRY) { ctx->err = "aggregation table overflow persisted"; return TJB_ERR_CUDA; }
  if (rc < 0) return rc;
  TRY (device_time (ctx, out));
  tr.mark ("post");
  return out->status;
}

// A read longer than a chunk (the reference accepts any length, src/kseq.h:172-212; e.g. an assembly given as a sample)
// is copied into a device buffer of its own and scanned there; rare, so everything around it is synchronous.
static int scan_long_read (tjb_ctx *ctx, const uint8_t *h_seq, size_t len)
{
  CU (cudaStreamSynchronize (ctx->stream));      // the buffer's previous user, if any, is done
  CU (cudaStreamSynchronize (ctx->copy_stream));
  TRY (ensure (ctx, ctx->d_long, len + 32));
  CU (cudaMemsetAsync ((uint8_t*) ctx->d_long.p + len, '\n', 16, ctx->stream)); // the separator behind it
  CU (cudaMemcpyAsync (ctx->d_long.p, h_seq, len, cudaMemcpyHostToDevice, ctx->stream));
  ctx->h2d_bytes += (int64_t) len;
  TRY (launch_scan (ctx, (const uint8_t*) ctx->d_long.p, len + 1));
  CU (cudaStreamSynchronize (ctx->stream));      // the host memory may change hands after this call
  return 0;
}

// cut [p, p+len) at the last '\n' so that a chunk always holds whole reads
static size_t whole_reads_prefix (const uint8_t *p, size_t len, size_t remaining)
{
  if (len >= remaining) return remaining;
  const void *nl = memrchr (p, '\n', len);
  return nl ? (size_t) ((const uint8_t*) nl - p) + 1 : 0;
}

// ---- 2-bit packed chunks (tjb_hostpack.h) back to the bytes the scan reads -------------------------------------------
// one thread per packed word = 16 bases = one 16-byte store; PRMT looks the letters up ("ACTG" by gray code)
__global__ void __launch_bounds__ (256) unpack_kernel (const u32 *__restrict__ packed, size_t n_words, uint4 *__restrict__ out)
{
  for (size_t i = (size_t) blockIdx.x * 256 + threadIdx.x; i < n_words; i += (size_t) gridDim.x * 256) {
    const u32 p = packed[i];
    u32 w[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
      u32 y = (p >> (8 * j)) & 0xffu;      // four 2-bit codes -> four nibbles
      y = (y | (y << 4)) & 0x0F0Fu;
      y = (y | (y << 2)) & 0x3333u;
      w[j] = __byte_perm (SCAN_GRAY_LUT, 0u, y);
    }
    out[i] = make_uint4 (w[0], w[1], w[2], w[3]);
  }
}
// every byte that is not one of "ACGT" (read separators, N, lower case, ...) travels as (position, byte)
__global__ void __launch_bounds__ (256) unpack_exc_kernel (const u32 *__restrict__ pos, const uint8_t *__restrict__ val, u32 n, uint8_t *__restrict__ out)
{
  const u32 i = blockIdx.x * 256 + threadIdx.x;
  if (i < n) out[pos[i]] = val[i];
}

#define TJB_PACK_MIN_BYTES ((size_t) 8 << 20)          // below this the packer threads are not worth waking
#define TJB_PACK_EXC_CAP (TJB_HOST_CHUNK_BYTES / 16)   // exceptions per chunk before it is sent unpacked (reads of >= ~20 bases)
static bool packed_transfer_enabled (tjb_ctx *ctx, size_t n_bytes)
{
  if (n_bytes < TJB_PACK_MIN_BYTES) return false;
  tjb_packer *pool = host_pool ();
  // sixteen packer threads turn 80-100 GB/s of reads into 2-bit codes (tools/pack_bench.py), fewer than ten do not beat
  // the copy engine's ~55 GB/s of plain bytes (ranks that share a host split its cores): then the bytes are sent as they are
  if (!pool || tjb_packer_threads (pool) < 9) return false;
  for (int i = 0; i < 2; i++) {
    if (!ctx->h_pk[i] && cudaMallocHost ((void**) &ctx->h_pk[i], TJB_HOST_CHUNK_BYTES / 4 + 64) != cudaSuccess) return false;
    if (!ctx->h_exc_pos[i] && cudaMallocHost ((void**) &ctx->h_exc_pos[i], (size_t) TJB_PACK_EXC_CAP * 4) != cudaSuccess) return false;
    if (!ctx->h_exc_val[i] && cudaMallocHost ((void**) &ctx->h_exc_val[i], TJB_PACK_EXC_CAP) != cudaSuccess) return false;
    if (ensure (ctx, ctx->d_pk[i], TJB_HOST_CHUNK_BYTES / 4 + 64) || ensure (ctx, ctx->d_exc_pos[i], (size_t) TJB_PACK_EXC_CAP * 4) ||
        ensure (ctx, ctx->d_exc_val[i], TJB_PACK_EXC_CAP)) return false;
  }
  return true;
}

// can the copy engine read this host range directly (pinned memory)?  Pageable memory is staged by the driver, chunk
// by chunk and synchronously, so sending such chunks unpacked does not overlap with anything.
static bool host_range_is_pinned (const void *p)
{
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes (&at, p) != cudaSuccess) { cudaGetLastError (); return false; }
  return at.type == cudaMemoryTypeHost;
}

// Host buffer: chunks are copied on the context's copy stream while the previous chunk is scanned on the compute
// stream (two device chunk buffers).  Large inputs cross PCIe 2-bit packed (host threads pack chunk i+1 while chunk i is
// copied, unpacked and scanned).  TJB_PLAIN_EVERY=n sends every n-th chunk of a pinned buffer as it is, so that the copy
// engine and the packer threads work on different chunks at the same time: measured slower (both draw on the host's
// memory bandwidth), off by default.
int tjb_count_sample_host (tjb_ctx *ctx, const uint8_t *seqs, size_t n_bytes, const tjb_opts *opt, tjb_counter_out *out)
{
  if (!ctx || !out || (n_bytes && !seqs)) return TJB_ERR_ARG;
  const size_t max_chunk = TJB_HOST_CHUNK_BYTES;
  int rc = 0;
  CU (cudaSetDevice (ctx->device));
  const bool packed = packed_transfer_enabled (ctx, n_bytes);
  static const int mix_env = getenv ("TJB_PLAIN_EVERY") ? atoi (getenv ("TJB_PLAIN_EVERY")) : 0; // n > 1: every n-th chunk unpacked (measured: no gain, off)
  const int plain_every = packed && mix_env > 1 && host_range_is_pinned (seqs) ? mix_env : 0;
  PackerSession session (packed);
  for (int attempt = 0; attempt < TJB_MAX_ATTEMPTS; attempt++) {
    out_reset (out);
    TRY (begin_sample (ctx, opt, n_bytes));
    if (ctx->profiling >= 2) CU (cudaEventRecord (ctx->ev_t0, ctx->stream));
    for (int i = 0; i < 2; i++) TRY (ensure (ctx, ctx->d_chunk[i], (max_chunk < n_bytes ? max_chunk : n_bytes) + 16));
    CU (cudaEventRecord (ctx->ev_ready, ctx->stream));            // workspace set up / previous sample done
    CU (cudaStreamWaitEvent (ctx->copy_stream, ctx->ev_ready, 0));
    size_t off = 0;
    int which = 0, chunk_no = 0;
    bool used[2] = {false, false};
    while (off < n_bytes) {
      size_t len = whole_reads_prefix (seqs + off, max_chunk < n_bytes - off ? max_chunk : n_bytes - off, n_bytes - off);
      if (!len) { // one read longer than a chunk: on its own, as it is
        const size_t rest = n_bytes - off;
        const void *nl = memchr (seqs + off + max_chunk, '\n', rest - max_chunk);
        const size_t seq_len = nl ? (size_t) ((const uint8_t*) nl - (seqs + off)) : rest;
        TRY (scan_long_read (ctx, seqs + off, seq_len));
        off += nl ? seq_len + 1 : seq_len;
        continue;
      }
      long n_exc = -1;
      const bool plain_turn = plain_every && (chunk_no % plain_every) == plain_every - 1;
      if (packed && !plain_turn) {
        if (used[which]) CU (cudaEventSynchronize (ctx->ev_copied[which])); // the slot's pinned staging has been read
        std::lock_guard<std::mutex> lk (g_packer_run_mu);
        n_exc = tjb_packer_run (g_packer, seqs + off, len, ctx->h_pk[which], ctx->h_exc_pos[which], ctx->h_exc_val[which], TJB_PACK_EXC_CAP);
      }
      if (used[which]) CU (cudaStreamWaitEvent (ctx->copy_stream, ctx->ev_scanned[which], 0)); // buffer free again
      if (n_exc >= 0) {
        const size_t n_words = (len + 15) / 16;
        CU (cudaMemcpyAsync (ctx->d_pk[which].p, ctx->h_pk[which], (len + 3) / 4, cudaMemcpyHostToDevice, ctx->copy_stream));
        if (n_exc) {
          CU (cudaMemcpyAsync (ctx->d_exc_pos[which].p, ctx->h_exc_pos[which], (size_t) n_exc * 4, cudaMemcpyHostToDevice, ctx->copy_stream));
          CU (cudaMemcpyAsync (ctx->d_exc_val[which].p, ctx->h_exc_val[which], (size_t) n_exc, cudaMemcpyHostToDevice, ctx->copy_stream));
        }
        ctx->h2d_bytes += (int64_t) ((len + 3) / 4 + (size_t) n_exc * 5);
        const u32 g = (u32) (nblk (n_words, 256) < (size_t) ctx->sm_count * 8 ? nblk (n_words, 256) : (size_t) ctx->sm_count * 8);
        unpack_kernel<<<g, 256, 0, ctx->copy_stream>>> ((const u32*) ctx->d_pk[which].p, n_words, (uint4*) ctx->d_chunk[which].p);
        ctx->launches++; ctx->stats[K_UNPACK].launches++; ctx->stats[K_UNPACK].bytes_algorithmic += (int64_t) (len + (len + 3) / 4);
        if (n_exc) {
          unpack_exc_kernel<<<(u32) nblk ((size_t) n_exc, 256), 256, 0, ctx->copy_stream>>> ((const u32*) ctx->d_exc_pos[which].p, (const uint8_t*) ctx->d_exc_val[which].p, (u32) n_exc, (uint8_t*) ctx->d_chunk[which].p);
          ctx->launches++; ctx->stats[K_UNPACK].launches++; ctx->stats[K_UNPACK].bytes_algorithmic += (int64_t) n_exc * 6;
        }
        CU (cudaGetLastError ());
      } else {
        CU (cudaMemcpyAsync (ctx->d_chunk[which].p, seqs + off, len, cudaMemcpyHostToDevice, ctx->copy_stream));
        ctx->h2d_bytes += (int64_t) len;
      }
      CU (cudaEventRecord (ctx->ev_copied[which], ctx->copy_stream));
      CU (cudaStreamWaitEvent (ctx->stream, ctx->ev_copied[which], 0));
      TRY (launch_scan (ctx, (const uint8_t*) ctx->d_chunk[which].p, len));
      CU (cudaEventRecord (ctx->ev_scanned[which], ctx->stream));
      used[which] = true;
      off += len;
      which ^= 1;
      chunk_no++;
    }
    rc = finalise (ctx, out);
    if (rc != TJB_INTERNAL_RETRY) break;
  }
  if (rc == TJB_INTERNAL_RETRY) { ctx->err = "aggregation table overflow persisted"; return TJB_ERR_CUDA; }
  if (rc < 0) return rc;
  TRY (device_time (ctx, out));
  return out->status;
}

static int count_files_once (tjb_ctx *ctx, const char *const *files, int n_files, const tjb_opts *opt, tjb_counter_out *out)
{
  out_reset (out);
  TRY (begin_sample (ctx, opt, sequence_bytes_hint (files, n_files)));
  for (int i = 0; i < 2 * n_files; i++) {
    if (!ctx->h_pinned[i]) CU (cudaMallocHost ((void**) &ctx->h_pinned[i], TJB_CHUNK_BYTES));
    if (!ctx->h_free[i]) CU (cudaEventCreateWithFlags (&ctx->h_free[i], cudaEventDisableTiming));
  }
  for (int i = 0; i < 2; i++) TRY (ensure (ctx, ctx->d_chunk[i], TJB_CHUNK_BYTES));
  if (ctx->profiling >= 2) CU (cudaEventRecord (ctx->ev_t0, ctx->stream));
  CU (cudaEventRecord (ctx->ev_ready, ctx->stream));
  CU (cudaStreamWaitEvent (ctx->copy_stream, ctx->ev_ready, 0));
  // One decoder thread per file (R1 and R2 inflate and parse concurrently; the order of the reads does not matter to
  // the counter), two pinned chunks each; this thread copies (copy stream) and scans (compute stream) whatever chunk
  // is ready, the copy of a chunk running under the scan of the one before.
  FileFeed feed;
  tjb_reader *rd[2] = {nullptr, nullptr};
  for (int f = 0; f < n_files; f++) {
    rd[f] = tjb_reader_open (files[f]);
    if (!rd[f]) { for (int g = 0; g < f; g++) tjb_reader_close (rd[g]); ctx->err = std::string ("cannot open ") + files[f]; return TJB_ERR_IO; }
    for (int i = 0; i < 2; i++) { feed.slot[f][i].pin = ctx->h_pinned[2 * f + i]; feed.slot[f][i].free_ev = ctx->h_free[2 * f + i]; }
  }
  std::thread workers[2];
  for (int f = 0; f < n_files; f++) workers[f] = std::thread (file_feed_worker, &feed, f, rd[f], ctx->device);
  int rc = 0, alive = n_files, turn[2] = {0, 0}, which = 0;
  bool finished[2] = {false, false}, used[2] = {false, false};
  while (alive > 0 && rc >= 0) {
    int p = -1;
    {
      std::unique_lock<std::mutex> lk (feed.mu);
      feed.cv.wait (lk, [&] { for (int q = 0; q < n_files; q++) if (!finished[q] && feed.slot[q][turn[q]].state == 1) { p = q; return true; } return false; });
    }
    FeedSlot &sl = feed.slot[p][turn[p]];
    if (sl.got <= 0) { // end of this file (0) or a decode error (< 0)
      finished[p] = true; alive--;
      if (sl.got < 0) { rc = (int) sl.got; ctx->err = std::string ("decode failed: ") + files[p]; }
      continue;
    }
    if (sl.long_read) {
      rc = scan_long_read (ctx, sl.long_read, (size_t) sl.got);
      if (rc >= 0 && cudaEventRecord (sl.free_ev, ctx->copy_stream) != cudaSuccess) rc = TJB_ERR_CUDA; // (the pinned slot was not used)
      { std::lock_guard<std::mutex> lk (feed.mu); sl.state = 0; sl.long_read = nullptr; }
      feed.cv.notify_all ();
      turn[p] ^= 1;
      continue;
    }
    cudaError_t e = cudaSuccess;
    if (used[which]) e = cudaStreamWaitEvent (ctx->copy_stream, ctx->ev_scanned[which], 0); // device chunk free again
    if (e == cudaSuccess) e = cudaMemcpyAsync (ctx->d_chunk[which].p, sl.pin, (size_t) sl.got, cudaMemcpyHostToDevice, ctx->copy_stream);
    if (e == cudaSuccess) e = cudaEventRecord (sl.free_ev, ctx->copy_stream); // the decoder waits for it before refilling
    if (e == cudaSuccess) e = cudaEventRecord (ctx->ev_copied[which], ctx->copy_stream);
    if (e == cudaSuccess) e = cudaStreamWaitEvent (ctx->stream, ctx->ev_copied[which], 0);
    if (e != cudaSuccess) { ctx->err = std::string ("chunk copy: ") + cudaGetErrorString (e); rc = TJB_ERR_CUDA; }
    if (rc >= 0) rc = launch_scan (ctx, (const uint8_t*) ctx->d_chunk[which].p, (size_t) sl.got);
    if (rc >= 0 && cudaEventRecord (ctx->ev_scanned[which], ctx->stream) != cudaSuccess) rc = TJB_ERR_CUDA;
    used[which] = true;
    { std::lock_guard<std::mutex> lk (feed.mu); sl.state = 0; }
    feed.cv.notify_all ();
    turn[p] ^= 1;
    which ^= 1;
  }
  { std::lock_guard<std::mutex> lk (feed.mu); feed.abort = true; }
  feed.cv.notify_all ();
  for (int f = 0; f < n_files; f++) { workers[f].join (); tjb_reader_close (rd[f]); }
  if (rc < 0) return rc;
  for (int f = 0; f < n_files; f++) { out->n_reads += feed.n_reads[f]; out->n_bases += feed.n_bases[f]; out->ms_decode += feed.ms_decode[f]; }
  return finalise (ctx, out);
}

int tjb_count_sample_files (tjb_ctx *ctx, const char *const *files, int n_files, const tjb_opts *opt, tjb_counter_out *out)
{
  if (!ctx || !out || !files || n_files < 1 || n_files > 2) return TJB_ERR_ARG;
  int rc = 0;
  for (int attempt = 0; attempt < TJB_MAX_ATTEMPTS; attempt++) { // a redo reads the files again
    rc = count_files_once (ctx, files, n_files, opt, out);
    if (rc != TJB_INTERNAL_RETRY) break;
  }
  if (rc == TJB_INTERNAL_RETRY) { ctx->err = "aggregation table overflow persisted"; return TJB_ERR_CUDA; }
  if (rc < 0) return rc;
  TRY (device_time (ctx, out));
  size_t nl = strlen (files[0]);
  out->name = (char*) malloc (nl + 1);
  if (out->name) memcpy (out->name, files[0], nl + 1);
  return out->status;
}

int tjb_count_sample (const char *const *files, int n_files, const tjb_opts *opt, int device, tjb_counter_out *out)
{
  struct Cache {
    std::map<int, tjb_ctx*> m;
    ~Cache () { for (auto &kv : m) tjb_ctx_destroy (kv.second); }
  };
  static thread_local Cache cache;
  tjb_ctx *ctx = nullptr;
  auto it = cache.m.find (device);
  if (it == cache.m.end ()) {
    int rc = tjb_ctx_create (device, &ctx);
    if (rc < 0) return rc;
    cache.m[device] = ctx;
  } else ctx = it->second;
  return tjb_count_sample_files (ctx, files, n_files, opt, out);
}

int tjb_device_count (void)
{
  int n = 0;
  if (cudaGetDeviceCount (&n) != cudaSuccess) { cudaGetLastError (); return 0; }
  return n;
}

// ---- one sample in pieces: read chunks counted on several GPUs, their tables exchanged and added ------
int tjb_sample_begin (tjb_ctx *ctx, const tjb_opts *opt, size_t n_bytes_hint)
{
  if (!ctx) return TJB_ERR_ARG;
  TRY (begin_sample (ctx, opt, n_bytes_hint));
  if (ctx->profiling >= 2) CU (cudaEventRecord (ctx->ev_t0, ctx->stream));
  return 0;
}

int64_t tjb_ctx_table_capacity (const tjb_ctx *ctx) { return ctx ? (int64_t) (ctx->tab_slots / 2) : 0; }

int tjb_ctx_reserve_entries (tjb_ctx *ctx, int64_t n_entries)
{
  if (!ctx || n_entries < 0) return TJB_ERR_ARG;
  ctx->want_entries = (size_t) n_entries + (size_t) n_entries / 8 + 4096;
  return 0;
}

int tjb_sample_scan_device (tjb_ctx *ctx, const void *d_seqs, size_t n_bytes)
{
  if (!ctx || !ctx->in_sample || (n_bytes && !d_seqs)) return TJB_ERR_ARG;
  if (((uintptr_t) d_seqs) & 15) { ctx->err = "device buffer must be 16-byte aligned"; return TJB_ERR_ARG; }
  return launch_scan (ctx, (const uint8_t*) d_seqs, n_bytes);
}

int tjb_sample_export (tjb_ctx *ctx, void *d_rows, int64_t max_rows, int64_t *n_rows, int32_t *row_words)
{
  if (!ctx || !ctx->in_sample || !n_rows || max_rows < 0) return TJB_ERR_ARG;
  const int W = ctx->rec_words;
  if (row_words) *row_words = W + 1;
  if (ctx->exported_rows < 0) { // table -> rows, once per sample (kept in the context until the sample ends)
    if (W == 2) TRY ((aggregate_launch<2> (ctx))); else TRY ((aggregate_launch<3> (ctx)));
    const size_t cap_rows = ctx->list_cap;
    TRY (ensure (ctx, ctx->uni_a, cap_rows * (W + 1) * 8 + 16));
    unsigned long long *d_n = (unsigned long long*) ((char*) ctx->uni_a.p + cap_rows * (W + 1) * 8);
    CU (cudaMemsetAsync (d_n, 0, 8, ctx->stream));
    const u32 g = (u32) ctx->sm_count * 8;
    if (W == 2) LAUNCH (K_TABLE, 8 * ctx->tab_slots, (table_export_kernel<2><<<g, 256, 0, ctx->stream>>> (table_of (ctx), (const u64*) ctx->ovf.p, ctx->d_res->counters, ctx->ovf_capacity, (u64*) ctx->uni_a.p, d_n, cap_rows)));
    else LAUNCH (K_TABLE, 8 * ctx->tab_slots, (table_export_kernel<3><<<g, 256, 0, ctx->stream>>> (table_of (ctx), (const u64*) ctx->ovf.p, ctx->d_res->counters, ctx->ovf_capacity, (u64*) ctx->uni_a.p, d_n, cap_rows)));
    CU (cudaGetLastError ());
    unsigned long long h_n = 0;
    CU (cudaMemcpyAsync (&h_n, d_n, 8, cudaMemcpyDeviceToHost, ctx->stream));
    CU (cudaMemcpyAsync (ctx->h_res, ctx->d_res, sizeof (ResultBlock), cudaMemcpyDeviceToHost, ctx->stream));
    CU (cudaStreamSynchronize (ctx->stream));
    if (ctx->h_res->counters[CTR_RECORDS] > ctx->rec_capacity) ctx->want_records = (size_t) ctx->h_res->counters[CTR_RECORDS] + 1024;
    if (h_n > cap_rows || ctx->h_res->counters[CTR_OVERFLOW] > ctx->ovf_capacity || ctx->h_res->counters[CTR_RECORDS] > ctx->rec_capacity) {
      ctx->want_entries = (size_t) h_n * 2 + 4096; // the caller starts the sample again (tjb_sample_begin sizes from this)
      ctx->err = "aggregation table too small for this chunk: begin the sample again";
      return TJB_ERR_TOO_LARGE;
    }
    ctx->exported_rows = (int64_t) h_n;
  }
  *n_rows = ctx->exported_rows;
  if (d_rows) {
    const int64_t n = ctx->exported_rows < max_rows ? ctx->exported_rows : max_rows;
    if (n) CU (cudaMemcpyAsync (d_rows, ctx->uni_a.p, (size_t) n * (W + 1) * 8, cudaMemcpyDeviceToDevice, ctx->stream));
    CU (cudaStreamSynchronize (ctx->stream));
  }
  return 0;
}

int tjb_sample_import (tjb_ctx *ctx, const void *d_rows, int64_t n_rows)
{
  if (!ctx || !ctx->in_sample || n_rows < 0 || (n_rows && !d_rows)) return TJB_ERR_ARG;
  if (!n_rows) return 0;
  const u32 g = (u32) (nblk ((size_t) n_rows, 256) < (size_t) ctx->sm_count * 8 ? nblk ((size_t) n_rows, 256) : (size_t) ctx->sm_count * 8);
  if (ctx->rec_words == 2) LAUNCH (K_TABLE, 40 * (size_t) n_rows, (table_import_kernel<2><<<g, 256, 0, ctx->stream>>> ((const u64*) d_rows, (unsigned long long) n_rows, ctx->opt.kmer_size, table_of (ctx), (u64*) ctx->ovf.p, ctx->ovf_capacity, ctx->d_res->counters, ctx->d_bin_hist, ctx->bin_bits)));
  else LAUNCH (K_TABLE, 48 * (size_t) n_rows, (table_import_kernel<3><<<g, 256, 0, ctx->stream>>> ((const u64*) d_rows, (unsigned long long) n_rows, ctx->opt.kmer_size, table_of (ctx), (u64*) ctx->ovf.p, ctx->ovf_capacity, ctx->d_res->counters, ctx->d_bin_hist, ctx->bin_bits)));
  CU (cudaGetLastError ());
  return 0;
}

int tjb_sample_finalise (tjb_ctx *ctx, tjb_counter_out *out)
{
  if (!ctx || !ctx->in_sample || !out) return TJB_ERR_ARG;
  out_reset (out);
  int rc = finalise (ctx, out);
  if (rc == TJB_INTERNAL_RETRY) { ctx->err = "aggregation table too small for this sample: begin the sample again"; return TJB_ERR_TOO_LARGE; }
  if (rc < 0) return rc;
  TRY (device_time (ctx, out));
  return out->status;
}

// ---- cross-sample exchange: query rows and their union (tjb_union.cuh) ------------------------------------------
int tjb_query_rows (const tjb_counter_out *in, uint64_t *rows)
{
  if (!in || (in->n_idx && (!rows || !in->elem || !in->idx_initial))) return TJB_ERR_ARG;
  const uint64_t *e = (const uint64_t*) in->elem;
  for (int32_t g = 0; g < in->n_idx; g++) { // first = longest record of the group (hopo_counter.c:513-520)
    const uint64_t *r = e + 5 * (size_t) in->idx_initial[g];
    rows[3 * (size_t) g] = r[0];
    rows[3 * (size_t) g + 1] = r[1];
    rows[3 * (size_t) g + 2] = (r[2] & 1) | (((r[2] >> 2) & 0x3ff) << 8); // base:2 length:10 bit-fields
  }
  return in->n_idx;
}

int tjb_query_rows_device (tjb_ctx *ctx, void *d_rows, int64_t max_rows)
{
  if (!ctx || max_rows < 0 || (max_rows && !d_rows)) return TJB_ERR_ARG;
  if (ctx->in_sample) { ctx->err = "query rows are taken from the last counted sample"; return TJB_ERR_ARG; }
  const int64_t n_idx = (int64_t) ctx->last_idx;
  if (ctx->h_res->meta[META_NIDX] != (u32) n_idx) return 0; // the last sample was dropped: no rows
  if (n_idx > max_rows) { ctx->err = "query row buffer too small"; return TJB_ERR_ARG; }
  if (!n_idx) return 0;
  const u32 g = (u32) ctx->sm_count * 4;
  const int k = ctx->opt.kmer_size;
  if (ctx->rec_words == 2) LAUNCH (K_UNION, 40 * (size_t) n_idx, (query_rows_kernel<2><<<g, UNI_THREADS, 0, ctx->stream>>> ((const u64*) ctx->okey.p, (const u32*) ctx->out_ii.p, ctx->d_bkt_start, ctx->d_bstat, ctx->d_ioff, ctx->d_res->meta, k, (u64*) d_rows, (u32) max_rows, ctx->list_cap)));
  else LAUNCH (K_UNION, 48 * (size_t) n_idx, (query_rows_kernel<3><<<g, UNI_THREADS, 0, ctx->stream>>> ((const u64*) ctx->okey.p, (const u32*) ctx->out_ii.p, ctx->d_bkt_start, ctx->d_bstat, ctx->d_ioff, ctx->d_res->meta, k, (u64*) d_rows, (u32) max_rows, ctx->list_cap)));
  CU (cudaGetLastError ());
  CU (cudaStreamSynchronize (ctx->stream));
  return (int) n_idx;
}

int tjb_union_queries_device (tjb_ctx *ctx, const void *d_rows, int64_t n_rows, int kmer_size, void *d_union, int64_t *n_union, void *d_index)
{
  if (!ctx || !n_union || n_rows < 0 || (n_rows && (!d_rows || !d_union))) return TJB_ERR_ARG;
  if (kmer_size < 2 || kmer_size > 32) { ctx->err = "kmer_size out of range (2..32)"; return TJB_ERR_ARG; }
  if (n_rows > (int64_t) INT32_MAX) { ctx->err = "more than INT32_MAX query rows"; return TJB_ERR_TOO_LARGE; }
  *n_union = 0;
  if (!n_rows) return 0;
  CU (cudaSetDevice (ctx->device));
  const int W = tjb_rec_words (kmer_size);
  return W == 2 ? union_t<2> (ctx, (const u64*) d_rows, (u32) n_rows, kmer_size, (u64*) d_union, n_union, (int*) d_index)
                : union_t<3> (ctx, (const u64*) d_rows, (u32) n_rows, kmer_size, (u64*) d_union, n_union, (int*) d_index);
}

void tjb_counter_out_free (tjb_counter_out *out)
{
  if (!out) return;
  if (!out->arena) { free (out->elem); free (out->idx_initial); free (out->idx_final); }
  free (out->name);
  out->elem = nullptr; out->idx_initial = out->idx_final = nullptr; out->name = nullptr;
  out->arena = 0;
}

} // extern "C"

// ---- tjb_tract_rows_noalloc: scan + count of one small batch of reads without any allocation ------------------------
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The scan's records, the
// aggregation table, its overflow list and the counters live in ONE file-scope device array of a fixed size (every GPU
// that loads the library holds it until exit); the caller passes the reads and the output rows.  Nothing here calls
// cudaMalloc, cudaFree, cudaHostAlloc or a library routine that allocates; everything is enqueued on the caller's stream
// and nothing is synchronised.  Every call shares the array: it is cleared at the start of a call (stream ordered), and
// two calls that overlap on the GPU (different streams, or several host threads) race.
#define NA_SLOTS   ((size_t) 1 << 17)   // table slots
#define NA_REC_CAP ((size_t) 1 << 18)   // tract observations per call
#define NA_OVF_CAP ((size_t) 1 << 14)
#define NA_BIN_BITS 8                   // the aggregate kernel bins new entries for the partition plan: kept, unused here
#define NA_OFF_COUNTERS 0               // CTR_WORDS x u64
#define NA_OFF_NROWS    64              // u64
#define NA_OFF_TILE     128             // u32
#define NA_OFF_BINS     256             // 2^NA_BIN_BITS x u32
#define NA_OFF_TABLE    (NA_OFF_BINS + (4u << NA_BIN_BITS))
#define NA_OFF_OVF      (NA_OFF_TABLE + NA_SLOTS * 40)          // sized for W = 3 (40 bytes per slot, 24 per record)
#define NA_OFF_RECS     (NA_OFF_OVF + NA_OVF_CAP * 24)
#define NA_SCRATCH_BYTES (NA_OFF_RECS + NA_REC_CAP * 24)        // 11.9 MB
static_assert (CTR_WORDS * 8 <= NA_OFF_NROWS && NA_OFF_TABLE % 256 == 0 && NA_OFF_OVF % 256 == 0 && NA_OFF_RECS % 256 == 0, "scratch layout");
static __device__ __align__ (256) unsigned char g_noalloc_scratch[NA_SCRATCH_BYTES];

// status[0] = rows written, or -1 when the batch did not fit the fixed scratch (more than NA_REC_CAP observations, an
// overfull table) or the caller's rows array; status[1] = tract observations; status[2] = skipped non-ACGT runs
__global__ void noalloc_status_kernel (const unsigned long long *__restrict__ counters, const unsigned long long *__restrict__ n_rows,
                                       unsigned long long max_rows, long long *__restrict__ status)
{
  const bool fits = counters[CTR_RECORDS] <= NA_REC_CAP && counters[CTR_OVERFLOW] <= NA_OVF_CAP && *n_rows <= max_rows;
  status[0] = fits ? (long long) *n_rows : -1LL;
  status[1] = (long long) counters[CTR_RECORDS];
  status[2] = (long long) counters[CTR_SKIP];
}

#define NA_MAX_DEVICES 64
static std::mutex g_noalloc_mu;
static unsigned char *g_noalloc_ptr[NA_MAX_DEVICES]; // device address of the array on each device, looked up once
static int g_noalloc_sms[NA_MAX_DEVICES];

// once per device, with that device current: the DEVICE address of the array (its name is only a host-side handle)
static int noalloc_setup (int device, unsigned char **ptr, int *sms)
{
  std::lock_guard<std::mutex> lk (g_noalloc_mu);
  if (!g_noalloc_ptr[device]) {
    unsigned char *p = nullptr;
    if (cudaGetSymbolAddress ((void**) &p, g_noalloc_scratch) != cudaSuccess || !p) { cudaGetLastError (); return TJB_ERR_CUDA; }
    int n_sm = 0;
    if (cudaDeviceGetAttribute (&n_sm, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || n_sm < 1) { cudaGetLastError (); return TJB_ERR_CUDA; }
    if (cudaFuncSetAttribute (tjb_scan_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) scan_smem_bytes<2> ()) != cudaSuccess ||
        cudaFuncSetAttribute (tjb_scan_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) scan_smem_bytes<3> ()) != cudaSuccess) { cudaGetLastError (); return TJB_ERR_CUDA; }
    g_noalloc_sms[device] = n_sm;
    g_noalloc_ptr[device] = p;
  }
  *ptr = g_noalloc_ptr[device]; *sms = g_noalloc_sms[device];
  return 0;
}

template <int W>
static int noalloc_enqueue (unsigned char *base, int sms, const uint8_t *d_seqs, size_t n, const tjb_opts *opt, u64 *d_rows, unsigned long long max_rows,
                            long long *d_status, cudaStream_t st)
{
  unsigned long long *counters = (unsigned long long*) (base + NA_OFF_COUNTERS), *n_rows = (unsigned long long*) (base + NA_OFF_NROWS);
  u32 *tile_ctr = (u32*) (base + NA_OFF_TILE), *bins = (u32*) (base + NA_OFF_BINS);
  u64 *ovf = (u64*) (base + NA_OFF_OVF), *recs = (u64*) (base + NA_OFF_RECS);
  AggTable tab;
  tab.cnt = (unsigned long long*) (base + NA_OFF_TABLE); tab.keys = (u64*) (base + NA_OFF_TABLE) + NA_SLOTS; tab.mask = (u32) (NA_SLOTS - 1);
  // what must start at zero: counters, row count, window tickets, bins, the table (keys and counts)
  if (cudaMemsetAsync (base, 0, NA_OFF_TABLE + table_bytes (NA_SLOTS, W), st) != cudaSuccess) return TJB_ERR_CUDA;
  const long long n_tiles = (long long) ((n + SW_TILE - 1) / SW_TILE);
  if (n_tiles) {
    const int min_len = opt->min_tract_size < 2 ? 2 : opt->min_tract_size;
    const long long want = (n_tiles + SCAN_WARPS - 1) / SCAN_WARPS;
    const long long grid = want < (long long) sms * SCAN_CTAS_PER_SM ? want : (long long) sms * SCAN_CTAS_PER_SM;
    tjb_scan_kernel<W><<<(u32) grid, SCAN_THREADS, scan_smem_bytes<W> (), st>>> (d_seqs, (long long) n, opt->kmer_size, min_len, recs, NA_REC_CAP, counters, n_tiles, tile_ctr, 0u);
    const u32 g_agg = (u32) (nblk (NA_REC_CAP, 256 * AGG_UNROLL) < (size_t) sms * 8 ? nblk (NA_REC_CAP, 256 * AGG_UNROLL) : (size_t) sms * 8);
    aggregate_kernel<W><<<g_agg, 256, 0, st>>> (recs, counters, NA_REC_CAP, tab, ovf, NA_OVF_CAP, opt->kmer_size, bins, NA_BIN_BITS);
    table_export_kernel<W><<<(u32) sms * 4, 256, 0, st>>> (tab, ovf, counters, NA_OVF_CAP, d_rows, n_rows, max_rows);
  }
  noalloc_status_kernel<<<1, 1, 0, st>>> (counters, n_rows, max_rows, d_status);
  return cudaGetLastError () == cudaSuccess ? 0 : TJB_ERR_CUDA;
}

extern "C" {

size_t tjb_noalloc_max_bytes (void) { return TJB_NOALLOC_MAX_BYTES; }

int tjb_tract_rows_noalloc (int device, const void *d_seqs, size_t n_bytes, const tjb_opts *opt, void *d_rows, int64_t max_rows,
                            int64_t *d_status, void *stream)
{
  if (device < 0 || device >= NA_MAX_DEVICES || !opt || !d_status || max_rows < 0 || (n_bytes && !d_seqs) || (max_rows && !d_rows)) return TJB_ERR_ARG;
  if (opt->kmer_size < 2 || opt->kmer_size > 32 || opt->min_tract_size < 1) return TJB_ERR_ARG;
  if (n_bytes > TJB_NOALLOC_MAX_BYTES) return TJB_ERR_NOALLOC_LIMIT; // before any CUDA call: a clear error, nothing enqueued
  int current = -1;
  if (cudaGetDevice (&current) != cudaSuccess) { cudaGetLastError (); return TJB_ERR_CUDA; }
  if (current != device && cudaSetDevice (device) != cudaSuccess) { cudaGetLastError (); return TJB_ERR_CUDA; }
  unsigned char *base = nullptr;
  int sms = 0;
  int rc = noalloc_setup (device, &base, &sms);
  if (rc == 0)
    rc = tjb_rec_words (opt->kmer_size) == 2
       ? noalloc_enqueue<2> (base, sms, (const uint8_t*) d_seqs, n_bytes, opt, (u64*) d_rows, (unsigned long long) max_rows, (long long*) d_status, (cudaStream_t) stream)
       : noalloc_enqueue<3> (base, sms, (const uint8_t*) d_seqs, n_bytes, opt, (u64*) d_rows, (unsigned long long) max_rows, (long long*) d_status, (cudaStream_t) stream);
  if (current != device && current >= 0) cudaSetDevice (current);
  return rc;
}

} // extern "C"

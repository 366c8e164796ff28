// libtjb_gpu: context, workspace, stream orchestration and the C ABI declared in include/tjb_gpu.h.
// Device work only — there is no host fallback for any compute step in this file.
#include <chrono>
#include <sys/mman.h>
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

enum KernelId { K_SCAN = 0, K_PLAN, K_PARTITION, K_BUCKET, K_TOTALS, K_PACK, K_COVMAX, K_RECHIST,
                // global multi-pass path, only for buckets too large for shared memory
                K_SORT_HIST, K_SORT_SCATTER, K_PSCAN, K_RLE, K_GROUP, K_UNION, K_UNPACK, K_COUNT };
static const char *kKernelNames[K_COUNT] = {"tjb_scan_kernel", "plan_kernel", "partition_kernel", "bucket_reduce_kernel", "totals_kernel",
                                            "pack_kernel", "cov_max_kernel", "rec_hist_kernel", "rsort_hist_kernel",
                                            "rsort_scatter_kernel", "pscan_kernels", "rle_kernels", "group_kernels", "union_kernels", "unpack_kernels"};

struct DevBuf {
  void *p = nullptr;
  size_t cap = 0;
};

struct PendingEvent { int kid; cudaEvent_t a, b; };

struct tjb_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  std::string err;
  int profiling = 0; // 0 off, 1 scan kernel only, 2 every kernel
  int64_t launches = 0;
  int64_t h2d_bytes = 0; // host-buffer path: bytes sent host -> device
  size_t last_keep = 0, last_idx = 0; // sizes of the previous sample's result: the next one is allocated ahead (finalise_t)
  tjb_kernel_stat stats[K_COUNT];
  std::vector<PendingEvent> pending;
  std::vector<cudaEvent_t> free_events;
  int sm_count = 148;

  uint8_t *h_pinned[4] = {nullptr, nullptr, nullptr, nullptr}; // files path: two chunks per decoder thread
  cudaEvent_t h_free[4] = {nullptr, nullptr, nullptr, nullptr}; // pinned chunk i may be overwritten once this has fired
  DevBuf d_chunk[2];
  // host-buffer path with 2-bit packed transfers (tjb_hostpack.h): pinned and device staging per chunk slot
  uint8_t *h_pk[2] = {nullptr, nullptr}; uint32_t *h_exc_pos[2] = {nullptr, nullptr}; uint8_t *h_exc_val[2] = {nullptr, nullptr};
  DevBuf d_pk[2], d_exc_pos[2], d_exc_val[2];

  DevBuf rec_a, rec_b; // raw tract observations + sort ping-pong
  size_t rec_capacity = 0; // records rec_a can hold
  int rec_words = 0;
  unsigned long long n_raw_known = 0, n_skip_known = 0;
  const uint8_t *pend_ptr = nullptr; size_t pend_len = 0; bool pend = false; // scan launched, overflow not checked yet
  unsigned long long *d_counters = nullptr; // [0] = n_raw, [1] = skipped non-ACGTU runs
  u32 *d_tile_ctr = nullptr;                // scan kernel: next window to hand out
  u32 *d_totals = nullptr;                  // small device scalars (global multi-pass path)
  int *d_cov = nullptr;
  // bucketed reduce: bin histogram (filled by the scan), splitters, per-bucket results
  int bin_bits = PLAN_REG_BIN_BITS;
  bool hist_dirty = false;                  // a chunk was re-scanned after an overflow: rebuild the histogram from the records
  u32 *d_bin_hist = nullptr, *d_lut = nullptr, *d_meta = nullptr;            // per bin (2^TJB_MAX_BIN_BITS)
  u32 *d_bkt_start = nullptr, *d_cursor = nullptr, *d_koff = nullptr, *d_ioff = nullptr; // per bucket (views of the DevBufs below)
  uint4 *d_bstat = nullptr;
  DevBuf b_bkt_start, b_cursor, b_koff, b_ioff, b_bstat;
  u32 max_buckets = 0;
  unsigned long long *h_sync = nullptr;     // pinned: [0..1] counters, [2..] meta words
  char *h_out = nullptr; size_t h_out_cap = 0; // pinned staging of the results
  DevBuf out_cnt, out_ii, out_if, cov_tab;
  DevBuf hist, ptmp, heads, pos, seg, drec, dcount, keep, kpos, krec, kcount, cw, csum, gkeep, gkpos, out40, idx_i, idx_f;

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

template <int W>
static int radix_sort (tjb_ctx *ctx, u64 *a, u64 *b, size_t n, const std::vector<DigitDesc> &digits, u64 **sorted)
{
  u32 n_tiles = nblk (n, RSORT_TILE);
  size_t hn = (size_t) RSORT_RADIX * n_tiles;
  TRY (ensure (ctx, ctx->hist, hn * sizeof (u32)));
  TRY (ensure (ctx, ctx->ptmp, pscan_tmp_elems (hn > n ? hn : n) * sizeof (u32) + 64));
  u32 *hist = (u32*) ctx->hist.p;
  for (const DigitDesc &dd : digits) {
    LAUNCH (K_SORT_HIST, 8 * W * n, (rsort_hist_kernel<W><<<n_tiles, RSORT_THREADS, 0, ctx->stream>>> (a, n, dd, hist, n_tiles)));
    TRY (exclusive_scan (ctx, hist, hist, hn, (u32*) ctx->ptmp.p, nullptr));
    LAUNCH (K_SORT_SCATTER, 16 * W * n, (rsort_scatter_kernel<W><<<n_tiles, RSORT_THREADS, 0, ctx->stream>>> (a, b, n, dd, hist, n_tiles)));
    u64 *t = a; a = b; b = t;
  }
  *sorted = a;
  return 0;
}

static int grow_records (tjb_ctx *ctx, size_t want_records)
{
  if (want_records <= ctx->rec_capacity) return 0;
  size_t bytes = want_records * ctx->rec_words * sizeof (u64);
  void *np = nullptr;
  CU (cudaMalloc (&np, bytes));
  if (ctx->rec_a.p && ctx->n_raw_known)
    CU (cudaMemcpyAsync (np, ctx->rec_a.p, (size_t) ctx->n_raw_known * ctx->rec_words * sizeof (u64), cudaMemcpyDeviceToDevice, ctx->stream));
  CU (cudaStreamSynchronize (ctx->stream));
  if (ctx->rec_a.p) CU (cudaFree (ctx->rec_a.p));
  ctx->rec_a.p = np; ctx->rec_a.cap = bytes;
  ctx->rec_capacity = want_records;
  return 0;
}

// bins of the partition plan: enough that a bin holds ~BK_CAP/8 records on average, at most what the context has.
// Expected records: a run of >= m equal bases starts at a position with probability 3/4 * 4^-(m-1).
static int choose_bin_bits (size_t n_bytes_hint, int k, int min_tract)
{
  const int W = tjb_rec_words (k);
  const int max_bits = W == 2 ? (4 * k + 1 < TJB_MAX_BIN_BITS ? 4 * k + 1 : TJB_MAX_BIN_BITS)
                              : (2 * k + 1 < TJB_MAX_BIN_BITS ? 2 * k + 1 : TJB_MAX_BIN_BITS);
  const int m = min_tract < 2 ? 2 : (min_tract > 8 ? 8 : min_tract);
  int F = 4;
  if (n_bytes_hint == 0) F = PLAN_REG_BIN_BITS; // size unknown (streamed files)
  else { size_t est = (n_bytes_hint * 3 / 4) >> (2 * (m - 1)); while (F < TJB_MAX_BIN_BITS && (est >> F) > BK_CAP / 8) F++; }
  return F < max_bits ? F : max_bits;
}

static int begin_sample (tjb_ctx *ctx, const tjb_opts *opt, size_t n_bytes_hint)
{
  if (!opt || opt->kmer_size < 2 || opt->kmer_size > 32 || opt->min_tract_size < 1 || opt->min_coverage < 0) {
    ctx->err = "options out of range (kmer_size 2..32, min_tract_size >= 1, min_coverage >= 0)";
    return TJB_ERR_ARG;
  }
  CU (cudaSetDevice (ctx->device));
  int W = tjb_rec_words (opt->kmer_size);
  if (W != ctx->rec_words) { ctx->rec_words = W; ctx->rec_capacity = ctx->rec_a.cap / (W * sizeof (u64)); }
  ctx->n_raw_known = ctx->n_skip_known = 0;
  ctx->pend = false;
  ctx->hist_dirty = false;
  ctx->bin_bits = choose_bin_bits (n_bytes_hint, opt->kmer_size, opt->min_tract_size);
  CU (cudaMemsetAsync (ctx->d_counters, 0, 2 * sizeof (unsigned long long), ctx->stream));
  CU (cudaMemsetAsync (ctx->d_bin_hist, 0, sizeof (u32) << ctx->bin_bits, ctx->stream));
  return 0;
}

static int launch_scan (tjb_ctx *ctx, const uint8_t *d_seqs, size_t n, const tjb_opts *opt)
{
  long long n_tiles = (long long) ((n + SW_TILE - 1) / SW_TILE); // one window of SW_TILE owned bytes per warp at a time
  if (!n_tiles) return 0;
  int min_len = opt->min_tract_size < 2 ? 2 : opt->min_tract_size;
  CU (cudaMemsetAsync (ctx->d_tile_ctr, 0, sizeof (u32), ctx->stream)); // window tickets of this launch
  long long want = (n_tiles + SCAN_WARPS - 1) / SCAN_WARPS;
  long long grid = want < (long long) ctx->sm_count * SCAN_CTAS_PER_SM ? want : (long long) ctx->sm_count * SCAN_CTAS_PER_SM;
  // algorithmic bytes of this launch: every sequence byte read once (+ 8W bytes per emitted record, added at finalise)
  if (ctx->rec_words == 2)
    LAUNCH (K_SCAN, n, (tjb_scan_kernel<2><<<(u32) grid, SCAN_THREADS, scan_smem_bytes<2> (), ctx->stream>>> (d_seqs, (long long) n, opt->kmer_size, min_len, (u64*) ctx->rec_a.p, ctx->rec_capacity, ctx->d_counters, n_tiles, ctx->d_bin_hist, ctx->bin_bits, ctx->d_tile_ctr)));
  else
    LAUNCH (K_SCAN, n, (tjb_scan_kernel<3><<<(u32) grid, SCAN_THREADS, scan_smem_bytes<3> (), ctx->stream>>> (d_seqs, (long long) n, opt->kmer_size, min_len, (u64*) ctx->rec_a.p, ctx->rec_capacity, ctx->d_counters, n_tiles, ctx->d_bin_hist, ctx->bin_bits, ctx->d_tile_ctr)));
  CU (cudaGetLastError ());
  return 0;
}

// Launches the scan of one device-resident chunk (asynchronous).  The record array is grown beforehand to an estimate;
// overflowing records are counted but never written, so the overflow is detected at the next synchronisation
// (scan_resolve(), or the first synchronisation of finalise for a single-chunk sample) and the chunk re-scanned.
static int scan_launch (tjb_ctx *ctx, const uint8_t *d_seqs, size_t n, const tjb_opts *opt)
{
  size_t guess = (size_t) ctx->n_raw_known + n / 16 + 65536;
  TRY (grow_records (ctx, guess));
  TRY (launch_scan (ctx, d_seqs, n, opt));
  ctx->pend_ptr = d_seqs; ctx->pend_len = n; ctx->pend = true;
  return 0;
}
// re-scan of the pending chunk after an overflow: cnt0 = the (too large) record count the first attempt reported
static int scan_redo (tjb_ctx *ctx, const tjb_opts *opt, unsigned long long cnt0)
{
  unsigned long long back[2] = {ctx->n_raw_known, ctx->n_skip_known};
  TRY (grow_records (ctx, (size_t) cnt0 + 1024));
  CU (cudaMemcpyAsync (ctx->d_counters, back, sizeof (back), cudaMemcpyHostToDevice, ctx->stream));
  CU (cudaStreamSynchronize (ctx->stream)); // `back` lives on this stack frame
  ctx->hist_dirty = true;                   // the first attempt already counted this chunk's bins
  return launch_scan (ctx, ctx->pend_ptr, ctx->pend_len, opt);
}
static int scan_resolve (tjb_ctx *ctx, const tjb_opts *opt)
{
  if (!ctx->pend) return 0;
  unsigned long long cnt[2];
  CU (cudaMemcpyAsync (cnt, ctx->d_counters, sizeof (cnt), cudaMemcpyDeviceToHost, ctx->stream));
  CU (cudaStreamSynchronize (ctx->stream));
  if (cnt[0] > ctx->rec_capacity) {
    TRY (scan_redo (ctx, opt, cnt[0]));
    CU (cudaMemcpyAsync (cnt, ctx->d_counters, sizeof (cnt), cudaMemcpyDeviceToHost, ctx->stream));
    CU (cudaStreamSynchronize (ctx->stream));
  }
  ctx->pend = false;
  ctx->n_raw_known = cnt[0];
  ctx->n_skip_known = cnt[1];
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
// Global multi-pass reduce of ONE bucket that does not fit in shared memory: LSD radix sort over every digit,
// run-length count, QC filter, context groups.  Results land where bucket_reduce_kernel would have put them.
template <int W>
static int reduce_oversize_bucket (tjb_ctx *ctx, const tjb_opts *opt, u32 b, u32 start, u32 n)
{
  const int k = opt->kmer_size;
  cudaStream_t st = ctx->stream;
  TRY (ensure (ctx, ctx->drec, (size_t) n * W * 8)); // scratch of n records
  u64 *in = (u64*) ctx->rec_b.p + (size_t) start * W, *tmp = (u64*) ctx->drec.p;
  u64 *sorted = nullptr;
  TRY ((radix_sort<W> (ctx, in, tmp, n, tract_digits (W, k), &sorted)));
  TRY (ensure (ctx, ctx->heads, (size_t) n * 4)); TRY (ensure (ctx, ctx->pos, (size_t) n * 4)); TRY (ensure (ctx, ctx->seg, (size_t) n * 4));
  TRY (ensure (ctx, ctx->ptmp, pscan_tmp_elems (n) * 4 + 64));
  u32 *heads = (u32*) ctx->heads.p, *pos = (u32*) ctx->pos.p, *seg = (u32*) ctx->seg.p, *ptmp = (u32*) ctx->ptmp.p;
  u32 h_tot[4];
  LAUNCH (K_RLE, 16 * W * (size_t) n, (head_kernel<W, 0><<<nblk (n, RED_THREADS), RED_THREADS, 0, st>>> (sorted, n, heads)));
  TRY (exclusive_scan (ctx, heads, pos, n, ptmp, ctx->d_totals + 0));
  CU (cudaMemcpyAsync (h_tot, ctx->d_totals, 4, cudaMemcpyDeviceToHost, st));
  LAUNCH (K_RLE, 12 * (size_t) n, (seg_start_kernel<<<nblk (n, RED_THREADS), RED_THREADS, 0, st>>> (heads, pos, n, seg)));
  CU (cudaStreamSynchronize (st));
  const u32 n_dist = h_tot[0];
  // distinct records: into whichever of {in, tmp} does not hold the sorted array
  u64 *drec = (sorted == in) ? tmp : in;
  TRY (ensure (ctx, ctx->dcount, (size_t) n_dist * 4)); TRY (ensure (ctx, ctx->keep, (size_t) n_dist * 4)); TRY (ensure (ctx, ctx->kpos, (size_t) n_dist * 4));
  LAUNCH (K_RLE, (16 * W + 12) * (size_t) n_dist, (rle_emit_kernel<W><<<nblk (n_dist, RED_THREADS), RED_THREADS, 0, st>>> (sorted, n, seg, n_dist, opt->remove_biased, drec, (int*) ctx->dcount.p, (u32*) ctx->keep.p)));
  TRY (exclusive_scan (ctx, (u32*) ctx->keep.p, (u32*) ctx->kpos.p, n_dist, ptmp, ctx->d_totals + 1));
  CU (cudaMemcpyAsync (h_tot, ctx->d_totals + 1, 4, cudaMemcpyDeviceToHost, st));
  CU (cudaStreamSynchronize (st));
  const u32 n_keep = h_tot[0];
  u32 n_idx = 0;
  if (n_keep) {
    // kept records go straight to the bucket's output slot (rec_a doubles as the compact output array)
    u64 *krec = (u64*) ctx->rec_a.p + (size_t) start * W;
    int *kcount = (int*) ctx->out_cnt.p + start;
    LAUNCH (K_RLE, (16 * W + 8) * (size_t) n_keep, (compact_kernel<W><<<nblk (n_dist, RED_THREADS), RED_THREADS, 0, st>>> (drec, (int*) ctx->dcount.p, (u32*) ctx->keep.p, (u32*) ctx->kpos.p, n_dist, krec, kcount)));
    TRY (ensure (ctx, ctx->cw, (size_t) n_keep * 4)); TRY (ensure (ctx, ctx->csum, (size_t) n_keep * 4));
    TRY (ensure (ctx, ctx->gkeep, (size_t) n_keep * 4)); TRY (ensure (ctx, ctx->gkpos, (size_t) n_keep * 4));
    LAUNCH (K_GROUP, 16 * W * (size_t) n_keep, (head_kernel<W, 1><<<nblk (n_keep, RED_THREADS), RED_THREADS, 0, st>>> (krec, n_keep, heads)));
    TRY (exclusive_scan (ctx, heads, pos, n_keep, ptmp, ctx->d_totals + 0));
    LAUNCH (K_GROUP, 12 * (size_t) n_keep, (seg_start_kernel<<<nblk (n_keep, RED_THREADS), RED_THREADS, 0, st>>> (heads, pos, n_keep, seg)));
    LAUNCH (K_GROUP, 8 * (size_t) n_keep, (icopy_kernel<<<nblk (n_keep, RED_THREADS), RED_THREADS, 0, st>>> (kcount, n_keep, (u32*) ctx->cw.p)));
    TRY (exclusive_scan (ctx, (u32*) ctx->cw.p, (u32*) ctx->csum.p, n_keep, ptmp, ctx->d_totals + 2));
    CU (cudaMemcpyAsync (h_tot, ctx->d_totals, 4, cudaMemcpyDeviceToHost, st));
    CU (cudaStreamSynchronize (st));
    const u32 n_groups = h_tot[0];
    LAUNCH (K_GROUP, 12 * (size_t) n_groups, (group_keep_kernel<<<nblk (n_groups, RED_THREADS), RED_THREADS, 0, st>>> (seg, n_groups, n_keep, (u32*) ctx->csum.p, ctx->d_totals + 2, opt->min_coverage, (u32*) ctx->gkeep.p)));
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

template <int W>
static int finalise_t (tjb_ctx *ctx, const tjb_opts *opt, tjb_counter_out *out, bool *overflow)
{
  const int k = opt->kmer_size, F = ctx->bin_bits;
  cudaStream_t st = ctx->stream;
  const unsigned long long cap = ctx->rec_capacity;
  *overflow = false;
  Trace tr;
  if (!cap) { // nothing was ever scanned (empty input)
    out->status = TJB_NO_HT;
    return 0;
  }
  // workspace that scales with the record capacity; rec_a is reused as the compact output array once partitioned
  TRY (ensure (ctx, ctx->rec_b, (size_t) cap * W * 8));
  TRY (ensure (ctx, ctx->out_cnt, (size_t) cap * 4)); TRY (ensure (ctx, ctx->out_ii, (size_t) cap * 4)); TRY (ensure (ctx, ctx->out_if, (size_t) cap * 4));
  u64 *rec_a = (u64*) ctx->rec_a.p, *rec_b = (u64*) ctx->rec_b.p;
  const u32 grid = (u32) ctx->sm_count;
  // per-bucket tables: a bucket starts where the running record count crosses a multiple of BK_CAPH
  const u32 max_buckets = (u32) (cap / BK_CAPH) + 2;
  TRY (ensure (ctx, ctx->b_bkt_start, ((size_t) max_buckets + 1) * 4));
  TRY (ensure (ctx, ctx->b_cursor, (size_t) max_buckets * BK_CURSOR_STRIDE * 4));
  TRY (ensure (ctx, ctx->b_koff, (size_t) max_buckets * 4)); TRY (ensure (ctx, ctx->b_ioff, (size_t) max_buckets * 4));
  TRY (ensure (ctx, ctx->b_bstat, (size_t) max_buckets * sizeof (uint4)));
  ctx->d_bkt_start = (u32*) ctx->b_bkt_start.p; ctx->d_cursor = (u32*) ctx->b_cursor.p;
  ctx->d_koff = (u32*) ctx->b_koff.p; ctx->d_ioff = (u32*) ctx->b_ioff.p; ctx->d_bstat = (uint4*) ctx->b_bstat.p;
  ctx->max_buckets = max_buckets;

  if (ctx->hist_dirty) {
    CU (cudaMemsetAsync (ctx->d_bin_hist, 0, sizeof (u32) << F, st));
    LAUNCH (K_RECHIST, 0, (rec_hist_kernel<W><<<grid * 8, 256, 0, st>>> (rec_a, ctx->d_counters, cap, k, F, ctx->d_bin_hist)));
    ctx->hist_dirty = false;
  }
  {
    const int ipt = F > 10 ? 1 << (F - 10) : 1; // bins per thread of the single planning CTA
#define PLAN_CASE(I) case I: LAUNCH (K_PLAN, 10 << F, (plan_kernel<I><<<1, PLAN_THREADS, 0, st>>> (ctx->d_bin_hist, F, ctx->d_counters, cap, ctx->d_bkt_start, ctx->d_lut, ctx->d_cursor, ctx->d_meta, max_buckets))); break;
    if (F > PLAN_REG_BIN_BITS)
      LAUNCH (K_PLAN, 18 << F, (plan_big_kernel<<<1, PLAN_THREADS, 0, st>>> (ctx->d_bin_hist, F, ctx->d_counters, cap, ctx->d_bkt_start, ctx->d_lut, ctx->d_cursor, ctx->d_meta, max_buckets)));
    else
    switch (ipt) { PLAN_CASE (1) PLAN_CASE (2) PLAN_CASE (4) PLAN_CASE (8) PLAN_CASE (16) PLAN_CASE (32) PLAN_CASE (64) }
#undef PLAN_CASE
  }
  LAUNCH (K_PARTITION, 0, (partition_kernel<W><<<grid * PART_GRID, 256, 0, st>>> (rec_a, rec_b, ctx->d_counters, cap, k, F, ctx->d_lut, ctx->d_cursor)));
  LAUNCH (K_BUCKET, 0, (bucket_reduce_kernel<W><<<grid * (W == 2 ? BK_CTAS2 : 2), BK_THREADS, bucket_smem_bytes<W> (), st>>> (rec_b, ctx->d_bkt_start, ctx->d_meta, cap, digit_list (W, k), opt->remove_biased, opt->min_coverage, rec_a, (u32*) ctx->out_cnt.p, (u32*) ctx->out_ii.p, (u32*) ctx->out_if.p, ctx->d_bstat)));
  LAUNCH (K_TOTALS, 0, (totals_kernel<<<1, PLAN_THREADS, 0, st>>> (ctx->d_bstat, ctx->d_meta, ctx->d_koff, ctx->d_ioff)));
  CU (cudaGetLastError ());
  // elements + group indices + coverage, into arrays for max_keep records / max_idx groups
  const int cov_init = INT_MIN;
  auto launch_pack = [&] (u32 max_keep, u32 max_idx) -> int {
    u32 slots = 1024;
    while (slots < 4u * max_keep) slots <<= 1; // two keys per record, load factor <= 0.5
    TRY (ensure (ctx, ctx->cov_tab, (size_t) slots * 8));
    TRY (ensure (ctx, ctx->out40, (size_t) max_keep * 24)); TRY (ensure (ctx, ctx->idx_i, (size_t) max_idx * 4)); TRY (ensure (ctx, ctx->idx_f, (size_t) max_idx * 4));
    u32 *cov_keys = (u32*) ctx->cov_tab.p, *cov_vals = cov_keys + slots;
    CU (cudaMemsetAsync (cov_keys, 0xff, (size_t) slots * 4, st));
    CU (cudaMemsetAsync (cov_vals, 0, (size_t) slots * 4, st));
    CU (cudaMemcpyAsync (ctx->d_cov, &cov_init, 4, cudaMemcpyHostToDevice, st));
    LAUNCH (K_PACK, (8 * W + 4 + 24) * (size_t) max_keep + 16 * (size_t) max_idx,
            (pack_kernel<W><<<grid * 4, 256, 0, st>>> ((const u64*) ctx->rec_a.p, (u32*) ctx->out_cnt.p, (u32*) ctx->out_ii.p, (u32*) ctx->out_if.p, ctx->d_bkt_start, ctx->d_bstat, ctx->d_koff, ctx->d_ioff, ctx->d_meta, k, (u64*) ctx->out40.p, (int*) ctx->idx_i.p, (int*) ctx->idx_f.p, cov_keys, cov_vals, slots - 1, max_keep, max_idx, cap)));
    LAUNCH (K_COVMAX, 8 * (size_t) slots, (cov_max_kernel<<<nblk (slots, 256 * 8), 256, 0, st>>> (cov_keys, cov_vals, slots, ctx->d_cov)));
    CU (cudaGetLastError ());
    return 0;
  };
  // ---- synchronisation 1: counts --------------------------------------------------------------------------
  unsigned long long *hs = ctx->h_sync;
  u32 *hm = (u32*) (hs + 2);
  CU (cudaMemcpyAsync (hs, ctx->d_counters, 2 * sizeof (unsigned long long), cudaMemcpyDeviceToHost, st));
  CU (cudaMemcpyAsync (hm, ctx->d_meta, META_WORDS * sizeof (u32), cudaMemcpyDeviceToHost, st));
  cudaEvent_t ev_counts = get_event (ctx);
  CU (cudaEventRecord (ev_counts, st));
  // The counts are only known after this round trip, but samples of one run are alike: launch the last kernels now,
  // sized from the previous sample (+12 %), so that they run while the host handles the counts; they are launched
  // again with the exact sizes in the rare case that the guess was too small.
  bool spec_packed = false;
  u32 spec_max_keep = 0, spec_max_idx = 0;
  if (ctx->last_keep && ctx->last_idx) {
    spec_max_keep = (u32) (ctx->last_keep + ctx->last_keep / 8 + 256);
    spec_max_idx = (u32) (ctx->last_idx + ctx->last_idx / 8 + 256);
    TRY (launch_pack (spec_max_keep, spec_max_idx));
    spec_packed = true;
  }
  tr.mark ("enqueue1");
  // the copy-out below is bound by ONE host thread copying out of the pinned staging buffer (~20 GB/s): when the
  // previous result was large, wake the host pool now (its wake-up hides under the kernels) and let it share the copy
  const bool pool_copy = ctx->last_keep * 40 >= ((size_t) 512 << 10) && host_pool () != nullptr;
  PackerSession copy_session (pool_copy);
  // The result arrays are malloc()ed (the reference free()s them) and fresh heap memory costs a page fault per 4 KiB
  // on first touch (measured: up to 1.2 ms for the 2.8 MB of config 2, depending on what the allocator has cached).
  // The host has nothing to do until the kernels above finish (~0.4 ms), so arrays sized like the previous sample's
  // result (+12 %) are allocated and touched now; they are trimmed or replaced once the real sizes are known.
  struct Spec { void *elem = nullptr, *ii = nullptr, *fi = nullptr; size_t keep = 0, idx = 0;
                ~Spec () { free (elem); free (ii); free (fi); } } spec;
  if (ctx->last_keep && ctx->last_idx) {
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
  CU (cudaEventSynchronize (ev_counts));
  ctx->free_events.push_back (ev_counts);
  tr.mark ("sync1");
  if (hs[0] > cap) { *overflow = true; return 0; } // the record array overflowed: everything above worked on garbage
  out->n_raw = (int64_t) hs[0];
  out->n_skipped_nonacgt = (int64_t) hs[1];
  if (hs[0] == 0) { out->status = TJB_NO_HT; return 0; }
  if (hs[0] > (unsigned long long) INT32_MAX) { ctx->err = "more than INT32_MAX tract observations"; return TJB_ERR_TOO_LARGE; }
  const u32 n = (u32) hs[0], n_buckets = hm[META_NBUCKETS];
  if (hm[META_NOVERSIZE]) { // rare: buckets that do not fit in shared memory go through the global multi-pass path
    std::vector<u32> h_start (n_buckets + 1);
    std::vector<uint4> h_stat (n_buckets);
    CU (cudaMemcpyAsync (h_start.data (), ctx->d_bkt_start, (n_buckets + 1) * sizeof (u32), cudaMemcpyDeviceToHost, st));
    CU (cudaMemcpyAsync (h_stat.data (), ctx->d_bstat, n_buckets * sizeof (uint4), cudaMemcpyDeviceToHost, st));
    CU (cudaStreamSynchronize (st));
    for (u32 b = 0; b < n_buckets; b++)
      if (h_stat[b].w) TRY ((reduce_oversize_bucket<W> (ctx, opt, b, h_start[b], h_start[b + 1] - h_start[b])));
    rec_a = (u64*) ctx->rec_a.p;
    LAUNCH (K_TOTALS, 0, (totals_kernel<<<1, PLAN_THREADS, 0, st>>> (ctx->d_bstat, ctx->d_meta, ctx->d_koff, ctx->d_ioff)));
    CU (cudaMemcpyAsync (hm, ctx->d_meta, META_WORDS * sizeof (u32), cudaMemcpyDeviceToHost, st));
    CU (cudaStreamSynchronize (st));
  }
  const u32 n_keep = hm[META_NKEEP], n_idx = hm[META_NIDX];
  out->n_distinct = hm[META_NDIST];
  // algorithmic bytes of the launches above, now that the counts are known
  ctx->stats[K_PARTITION].bytes_algorithmic += (int64_t) 16 * W * n;
  ctx->stats[K_BUCKET].bytes_algorithmic += (int64_t) 8 * W * n + (int64_t) (8 * W + 4) * n_keep + 8 * (int64_t) n_idx;
  ctx->stats[K_TOTALS].bytes_algorithmic += (int64_t) 24 * n_buckets;
  if (!n_keep) { out->status = TJB_NO_HT_AFTER_QC; return 0; }
  if (!n_idx) { out->status = TJB_NO_CTX_ABOVE_COVERAGE; return 0; }

  // ---- reference-layout elements, global group indices, coverage (:420-438) ---------------------------------
  if (!(spec_packed && n_keep <= spec_max_keep && n_idx <= spec_max_idx && !hm[META_NOVERSIZE]))
    TRY (launch_pack (n_keep, n_idx)); // not launched ahead, or the previous sample's sizes were too small
  tr.mark ("enqueue2");
  // copy out into malloc'ed host memory (the reference free()s these).  Measured on B200 for 2.8 MB: the driver's
  // pageable path 250 us; pinned staging in 4 chunks with the host copy of chunk i under the DMA of chunk i+1 220 us
  // (the single-threaded host copy into the caller's memory is what is left).
  ctx->last_keep = n_keep; ctx->last_idx = n_idx;
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
  int cov = 0;
  {
    // pinned staging in chunks: the DMA of chunk i+1 overlaps the host copy of chunk i into the caller's memory
    // an element crosses PCIe as its 24 varying bytes; the host appends the constant 16 while it copies
    const size_t eb = (size_t) n_keep * 24, ib = (size_t) n_idx * 4, total = eb + 2 * ib + 16;
    if (total > ctx->h_out_cap) {
      if (ctx->h_out) CU (cudaFreeHost (ctx->h_out));
      ctx->h_out = nullptr; ctx->h_out_cap = 0;
      CU (cudaMallocHost ((void**) &ctx->h_out, total + total / 4));
      ctx->h_out_cap = total + total / 4;
    }
    static const int nch_env = getenv ("TJB_OUT_CHUNKS") ? atoi (getenv ("TJB_OUT_CHUNKS")) : 0;
    const int NCH = nch_env >= 1 && nch_env <= 8 ? nch_env : 4;
    cudaEvent_t ev[8];
    size_t lo[9];
    for (int c = 0; c <= NCH; c++) lo[c] = (size_t) n_keep * c / NCH; // in elements
    lo[NCH] = n_keep;
    // the index arrays travel with the first chunk, so that their host copy runs under the DMA of the later chunks
    for (int c = 0; c < NCH; c++) {
      if (c == 0) {
        CU (cudaMemcpyAsync (ctx->h_out + eb, ctx->idx_i.p, ib, cudaMemcpyDeviceToHost, st));
        CU (cudaMemcpyAsync (ctx->h_out + eb + ib, ctx->idx_f.p, ib, cudaMemcpyDeviceToHost, st));
        CU (cudaMemcpyAsync (ctx->h_out + eb + 2 * ib, ctx->d_cov, 4, cudaMemcpyDeviceToHost, st));
      }
      CU (cudaMemcpyAsync (ctx->h_out + 24 * lo[c], (const char*) ctx->out40.p + 24 * lo[c], 24 * (lo[c + 1] - lo[c]), cudaMemcpyDeviceToHost, st));
      ev[c] = get_event (ctx);
      CU (cudaEventRecord (ev[c], st));
    }
    for (int c = 0; c < NCH; c++) {
      CU (cudaEventSynchronize (ev[c]));
      if (pool_copy) {
        std::lock_guard<std::mutex> lk (g_packer_run_mu);
        tjb_packer_expand (g_packer, (char*) out->elem + 40 * lo[c], ctx->h_out + 24 * lo[c], lo[c + 1] - lo[c]);
      } else
        tjb_expand_rows ((char*) out->elem + 40 * lo[c], ctx->h_out + 24 * lo[c], lo[c + 1] - lo[c]);
      if (c == 0) {
        memcpy (out->idx_initial, ctx->h_out + eb, ib);
        memcpy (out->idx_final, ctx->h_out + eb + ib, ib);
        memcpy (&cov, ctx->h_out + eb + 2 * ib, 4);
      }
      ctx->free_events.push_back (ev[c]);
    }
  }
  tr.mark ("d2h_enqueue");
  CU (cudaStreamSynchronize (st)); // ---- synchronisation 2
  tr.mark ("sync2");
  out->n_elem = (int32_t) n_keep;
  out->n_idx = (int32_t) n_idx;
  out->coverage = cov;
  out->status = TJB_OK;
  return 0;
}

#define TJB_INTERNAL_RETRY 100 // the record array overflowed; ctx->h_sync[0] holds the number of records needed

static int finalise (tjb_ctx *ctx, const tjb_opts *opt, tjb_counter_out *out)
{
  bool overflow = false;
  int rc = ctx->rec_words == 2 ? finalise_t<2> (ctx, opt, out, &overflow) : finalise_t<3> (ctx, opt, out, &overflow);
  ctx->pend = false;
  if (rc == 0 && overflow) return TJB_INTERNAL_RETRY;
  if (rc == 0) {
    // scan wrote 8W bytes per record on top of the sequence bytes it read
    ctx->stats[K_SCAN].bytes_algorithmic += (int64_t) out->n_raw * 8 * ctx->rec_words;
  }
  cudaStreamSynchronize (ctx->stream);
  resolve_events (ctx);
  if (rc < 0 || out->status != TJB_OK) { // the reference leaves an "empty counter" (n_elem = 0)
    free (out->elem); free (out->idx_initial); free (out->idx_final);
    out->elem = nullptr; out->idx_initial = out->idx_final = nullptr;
    out->n_elem = out->n_idx = 0; out->coverage = 0;
  }
  return rc;
}

static void out_reset (tjb_counter_out *out) { memset (out, 0, sizeof (*out)); }

// ---- union of query rows over samples (tjb_union.cuh): pack, LSD sort, heads, emit, index -----------------------
template <int W>
static int union_t (tjb_ctx *ctx, const u64 *d_rows, u32 n, int k, u64 *d_union, int64_t *n_union, int *d_index)
{
  cudaStream_t st = ctx->stream;
  TRY (grow_records (ctx, n));
  TRY (ensure (ctx, ctx->rec_b, (size_t) n * W * 8));
  TRY (ensure (ctx, ctx->heads, (size_t) n * 4)); TRY (ensure (ctx, ctx->pos, (size_t) n * 4));
  TRY (ensure (ctx, ctx->ptmp, pscan_tmp_elems (n) * 4 + 64));
  u64 *a = (u64*) ctx->rec_a.p, *b = (u64*) ctx->rec_b.p, *sorted = nullptr;
  u32 *heads = (u32*) ctx->heads.p, *pos = (u32*) ctx->pos.p;
  LAUNCH (K_UNION, (24 + 8 * W) * (size_t) n, (rows_to_rec_kernel<W><<<nblk (n, UNI_THREADS), UNI_THREADS, 0, st>>> (d_rows, n, k, a)));
  TRY ((radix_sort<W> (ctx, a, b, n, tract_digits (W, k), &sorted)));
  LAUNCH (K_UNION, 16 * W * (size_t) n, (head_kernel<W, 0><<<nblk (n, RED_THREADS), RED_THREADS, 0, st>>> (sorted, n, heads)));
  TRY (exclusive_scan (ctx, heads, pos, n, (u32*) ctx->ptmp.p, ctx->d_totals + 0));
  LAUNCH (K_UNION, (8 * W + 32) * (size_t) n, (union_emit_kernel<W><<<nblk (n, UNI_THREADS), UNI_THREADS, 0, st>>> (sorted, heads, pos, n, ctx->d_totals + 0, k, d_union)));
  if (d_index)
    LAUNCH (K_UNION, (24 + 4) * (size_t) n, (union_index_kernel<W><<<nblk (n, UNI_THREADS), UNI_THREADS, 0, st>>> (d_rows, n, k, sorted, heads, pos, ctx->d_totals + 0, d_index)));
  CU (cudaGetLastError ());
  u32 h_tot = 0;
  CU (cudaMemcpyAsync (&h_tot, ctx->d_totals, 4, cudaMemcpyDeviceToHost, st));
  CU (cudaStreamSynchronize (st));
  resolve_events (ctx);
  *n_union = h_tot;
  return 0;
}

// ---- file feed: one decoder thread per FASTQ file, two pinned chunks each -----------------
struct FeedSlot { uint8_t *pin = nullptr; cudaEvent_t free_ev = nullptr; int64_t got = 0; int state = 0; /* 0 empty, 1 full */ };
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
    const int64_t got = tjb_reader_fill (rd, s.pin, TJB_CHUNK_BYTES, &f->n_reads[p], &f->n_bases[p]);
    f->ms_decode[p] += std::chrono::duration<double, std::milli> (std::chrono::steady_clock::now () - t0).count ();
    { std::lock_guard<std::mutex> lk (f->mu); s.got = got; s.state = 1; }
    f->cv.notify_all ();
    if (got <= 0) return;
  }
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
  if ((e = cudaMalloc (&ctx->d_counters, 2 * sizeof (unsigned long long))) != cudaSuccess) return fail (e);
  if ((e = cudaMalloc (&ctx->d_totals, 16 * sizeof (u32))) != cudaSuccess) return fail (e);
  if ((e = cudaMalloc (&ctx->d_tile_ctr, sizeof (u32))) != cudaSuccess) return fail (e);
  if ((e = cudaMalloc (&ctx->d_cov, sizeof (int))) != cudaSuccess) return fail (e);
  if ((e = cudaStreamCreateWithFlags (&ctx->copy_stream, cudaStreamNonBlocking)) != cudaSuccess) return fail (e);
  if ((e = cudaMalloc (&ctx->d_bin_hist, TJB_MAX_BINS * sizeof (u32))) != cudaSuccess) return fail (e);
  if ((e = cudaMalloc (&ctx->d_lut, TJB_MAX_BINS * sizeof (u32))) != cudaSuccess) return fail (e);
  if ((e = cudaMalloc (&ctx->d_meta, META_WORDS * sizeof (u32))) != cudaSuccess) return fail (e);
  if ((e = cudaMallocHost ((void**) &ctx->h_sync, 16 * sizeof (unsigned long long))) != cudaSuccess) return fail (e);
  if ((e = cudaFuncSetAttribute (tjb_scan_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) scan_smem_bytes<2> ())) != cudaSuccess) return fail (e);
  if ((e = cudaFuncSetAttribute (tjb_scan_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) scan_smem_bytes<3> ())) != cudaSuccess) return fail (e);
  if ((e = cudaFuncSetAttribute (bucket_reduce_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) bucket_smem_bytes<2> ())) != cudaSuccess) return fail (e);
  if ((e = cudaFuncSetAttribute (bucket_reduce_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) bucket_smem_bytes<3> ())) != cudaSuccess) return fail (e);
  cudaEventCreate (&ctx->ev_t0); cudaEventCreate (&ctx->ev_t1);
  cudaEventCreateWithFlags (&ctx->ev_ready, cudaEventDisableTiming);
  for (int i = 0; i < 2; i++) { cudaEventCreateWithFlags (&ctx->ev_copied[i], cudaEventDisableTiming); cudaEventCreateWithFlags (&ctx->ev_scanned[i], cudaEventDisableTiming); }
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
                    &ctx->d_chunk[0], &ctx->d_chunk[1], &ctx->rec_a, &ctx->rec_b, &ctx->hist, &ctx->ptmp, &ctx->heads, &ctx->pos, &ctx->seg,
                    &ctx->drec, &ctx->dcount, &ctx->keep, &ctx->kpos, &ctx->krec, &ctx->kcount, &ctx->cw, &ctx->csum, &ctx->gkeep,
                    &ctx->gkpos, &ctx->out40, &ctx->idx_i, &ctx->idx_f, &ctx->out_cnt, &ctx->out_ii, &ctx->out_if, &ctx->cov_tab};
  for (DevBuf *b : bufs) if (b->p) cudaFree (b->p);
  for (int i = 0; i < 4; i++) { if (ctx->h_pinned[i]) cudaFreeHost (ctx->h_pinned[i]); if (ctx->h_free[i]) cudaEventDestroy (ctx->h_free[i]); }
  if (ctx->d_counters) cudaFree (ctx->d_counters);
  if (ctx->d_totals) cudaFree (ctx->d_totals);
  if (ctx->d_tile_ctr) cudaFree (ctx->d_tile_ctr);
  if (ctx->d_cov) cudaFree (ctx->d_cov);
  void *small[] = {ctx->d_bin_hist, ctx->d_lut, ctx->d_meta, ctx->b_bkt_start.p, ctx->b_cursor.p, ctx->b_koff.p, ctx->b_ioff.p, ctx->b_bstat.p};
  for (void *q : small) if (q) cudaFree (q);
  if (ctx->h_sync) cudaFreeHost (ctx->h_sync);
  if (ctx->h_out) cudaFreeHost (ctx->h_out);
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

int tjb_count_sample_device (tjb_ctx *ctx, const void *d_seqs, size_t n_bytes, const tjb_opts *opt, tjb_counter_out *out)
{
  if (!ctx || !out || (n_bytes && !d_seqs)) return TJB_ERR_ARG;
  if (((uintptr_t) d_seqs) & 15) { ctx->err = "device buffer must be 16-byte aligned"; return TJB_ERR_ARG; }
  size_t want = n_bytes / 16 + 65536; // records; the scan counts what does not fit, then the sample is redone
  int rc = 0;
  Trace tr;
  for (int attempt = 0; attempt < 3; attempt++) {
    out_reset (out);
    TRY (begin_sample (ctx, opt, n_bytes));
    CU (cudaEventRecord (ctx->ev_t0, ctx->stream));
    TRY (grow_records (ctx, want));
    TRY (launch_scan (ctx, (const uint8_t*) d_seqs, n_bytes, opt));
    tr.mark ("device:scan_enqueue");
    rc = finalise (ctx, opt, out);
    tr.mark ("finalise");
    if (rc != TJB_INTERNAL_RETRY) break;
    want = (size_t) ctx->h_sync[0] + 1024;
  }
  if (rc == TJB_INTERNAL_RETRY) { ctx->err = "record array overflow persisted"; return TJB_ERR_CUDA; }
  if (rc < 0) return rc;
  CU (cudaEventRecord (ctx->ev_t1, ctx->stream));
  CU (cudaEventSynchronize (ctx->ev_t1));
  float ms = 0.f;
  cudaEventElapsedTime (&ms, ctx->ev_t0, ctx->ev_t1);
  out->ms_device = ms;
  tr.mark ("post");
  return out->status;
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
  if (!pool || tjb_packer_threads (pool) < 4) return false; // PCIe beats a few cores
  for (int i = 0; i < 2; i++) {
    if (!ctx->h_pk[i] && cudaMallocHost ((void**) &ctx->h_pk[i], TJB_HOST_CHUNK_BYTES / 4 + 64) != cudaSuccess) return false;
    if (!ctx->h_exc_pos[i] && cudaMallocHost ((void**) &ctx->h_exc_pos[i], (size_t) TJB_PACK_EXC_CAP * 4) != cudaSuccess) return false;
    if (!ctx->h_exc_val[i] && cudaMallocHost ((void**) &ctx->h_exc_val[i], TJB_PACK_EXC_CAP) != cudaSuccess) return false;
    if (ensure (ctx, ctx->d_pk[i], TJB_HOST_CHUNK_BYTES / 4 + 64) || ensure (ctx, ctx->d_exc_pos[i], (size_t) TJB_PACK_EXC_CAP * 4) ||
        ensure (ctx, ctx->d_exc_val[i], TJB_PACK_EXC_CAP)) return false;
  }
  return true;
}

// Host buffer: chunks are copied on the context's copy stream while the previous chunk is scanned on the compute
// stream (two device chunk buffers); one optimistic pass, redone with the exact record capacity after an overflow.
// Large inputs cross PCIe 2-bit packed (host threads pack chunk i+1 while chunk i is copied, unpacked and scanned).
int tjb_count_sample_host (tjb_ctx *ctx, const uint8_t *seqs, size_t n_bytes, const tjb_opts *opt, tjb_counter_out *out)
{
  if (!ctx || !out || (n_bytes && !seqs)) return TJB_ERR_ARG;
  const size_t max_chunk = TJB_HOST_CHUNK_BYTES;
  size_t want = n_bytes / 16 + 65536;
  int rc = 0;
  CU (cudaSetDevice (ctx->device));
  const bool packed = packed_transfer_enabled (ctx, n_bytes);
  PackerSession session (packed);
  for (int attempt = 0; attempt < 3; attempt++) {
    out_reset (out);
    TRY (begin_sample (ctx, opt, n_bytes));
    CU (cudaEventRecord (ctx->ev_t0, ctx->stream));
    TRY (grow_records (ctx, want));
    for (int i = 0; i < 2; i++) TRY (ensure (ctx, ctx->d_chunk[i], (max_chunk < n_bytes ? max_chunk : n_bytes) + 16));
    CU (cudaEventRecord (ctx->ev_ready, ctx->stream));            // workspace set up / previous sample done
    CU (cudaStreamWaitEvent (ctx->copy_stream, ctx->ev_ready, 0));
    size_t off = 0;
    int which = 0;
    bool used[2] = {false, false};
    while (off < n_bytes) {
      size_t len = whole_reads_prefix (seqs + off, max_chunk < n_bytes - off ? max_chunk : n_bytes - off, n_bytes - off);
      if (!len) { ctx->err = "a single read exceeds the chunk size"; return TJB_ERR_TOO_LARGE; }
      long n_exc = -1;
      if (packed) {
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
      TRY (launch_scan (ctx, (const uint8_t*) ctx->d_chunk[which].p, len, opt));
      CU (cudaEventRecord (ctx->ev_scanned[which], ctx->stream));
      used[which] = true;
      off += len;
      which ^= 1;
    }
    rc = finalise (ctx, opt, out);
    if (rc != TJB_INTERNAL_RETRY) break;
    want = (size_t) ctx->h_sync[0] + 1024;
  }
  if (rc == TJB_INTERNAL_RETRY) { ctx->err = "record array overflow persisted"; return TJB_ERR_CUDA; }
  if (rc < 0) return rc;
  CU (cudaEventRecord (ctx->ev_t1, ctx->stream));
  CU (cudaEventSynchronize (ctx->ev_t1));
  float ms = 0.f;
  cudaEventElapsedTime (&ms, ctx->ev_t0, ctx->ev_t1);
  out->ms_device = ms;
  return out->status;
}

int tjb_count_sample_files (tjb_ctx *ctx, const char *const *files, int n_files, const tjb_opts *opt, tjb_counter_out *out)
{
  if (!ctx || !out || !files || n_files < 1 || n_files > 2) return TJB_ERR_ARG;
  out_reset (out);
  TRY (begin_sample (ctx, opt, 0));
  for (int i = 0; i < 2 * n_files; i++) {
    if (!ctx->h_pinned[i]) CU (cudaMallocHost ((void**) &ctx->h_pinned[i], TJB_CHUNK_BYTES));
    if (!ctx->h_free[i]) CU (cudaEventCreateWithFlags (&ctx->h_free[i], cudaEventDisableTiming));
  }
  for (int i = 0; i < 2; i++) TRY (ensure (ctx, ctx->d_chunk[i], TJB_CHUNK_BYTES));
  CU (cudaEventRecord (ctx->ev_t0, ctx->stream));
  // One decoder thread per file (R1 and R2 inflate and parse concurrently; the order of the reads does not matter to
  // the counter), two pinned chunks each; this thread copies and scans whatever chunk is ready.
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
  bool finished[2] = {false, false};
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
    rc = scan_resolve (ctx, opt); // previous chunk: record count known, overflow handled
    if (rc >= 0) {
      cudaError_t e = cudaMemcpyAsync (ctx->d_chunk[which].p, sl.pin, (size_t) sl.got, cudaMemcpyHostToDevice, ctx->stream);
      if (e == cudaSuccess) e = cudaEventRecord (sl.free_ev, ctx->stream); // the decoder waits for it before refilling
      if (e != cudaSuccess) { ctx->err = std::string ("chunk copy: ") + cudaGetErrorString (e); rc = TJB_ERR_CUDA; }
    }
    if (rc >= 0) rc = scan_launch (ctx, (const uint8_t*) ctx->d_chunk[which].p, (size_t) sl.got, opt);
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
  TRY (scan_resolve (ctx, opt));
  rc = finalise (ctx, opt, out);
  if (rc < 0) return rc;
  CU (cudaEventRecord (ctx->ev_t1, ctx->stream));
  CU (cudaEventSynchronize (ctx->ev_t1));
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

int tjb_union_queries_device (tjb_ctx *ctx, const void *d_rows, int64_t n_rows, int kmer_size, void *d_union, int64_t *n_union, void *d_index)
{
  if (!ctx || !n_union || n_rows < 0 || (n_rows && (!d_rows || !d_union))) return TJB_ERR_ARG;
  if (kmer_size < 2 || kmer_size > 32) { ctx->err = "kmer_size out of range (2..32)"; return TJB_ERR_ARG; }
  if (n_rows > (int64_t) INT32_MAX) { ctx->err = "more than INT32_MAX query rows"; return TJB_ERR_TOO_LARGE; }
  *n_union = 0;
  if (!n_rows) return 0;
  CU (cudaSetDevice (ctx->device));
  const int W = tjb_rec_words (kmer_size);
  if (W != ctx->rec_words) { ctx->rec_words = W; ctx->rec_capacity = ctx->rec_a.cap / (W * sizeof (u64)); }
  ctx->n_raw_known = 0; // nothing of a previous sample needs to survive a reallocation
  return W == 2 ? union_t<2> (ctx, (const u64*) d_rows, (u32) n_rows, kmer_size, (u64*) d_union, n_union, (int*) d_index)
                : union_t<3> (ctx, (const u64*) d_rows, (u32) n_rows, kmer_size, (u64*) d_union, n_union, (int*) d_index);
}

void tjb_counter_out_free (tjb_counter_out *out)
{
  if (!out) return;
  free (out->elem); free (out->idx_initial); free (out->idx_final); free (out->name);
  out->elem = nullptr; out->idx_initial = out->idx_final = nullptr; out->name = nullptr;
}

} // extern "C"

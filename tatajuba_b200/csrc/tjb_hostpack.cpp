// Host-side 2-bit packing of sequence bytes (see tjb_hostpack.h).  Plain C++ threads; the inner loop is AVX2 when the
// CPU has it (checked at run time), scalar otherwise.
#include "tjb_hostpack.h"

#include <atomic>
#include <sched.h>
#include <stdio.h>
#include <condition_variable>
#include <mutex>
#include <string.h>
#include <thread>
#include <vector>
#if defined(__x86_64__)
#include <immintrin.h>
#define TJB_PAUSE() _mm_pause ()
#else
#define TJB_PAUSE() do { } while (0)
#endif

// spin-wait step: a few hundred pauses, then give the core away (several ranks may share the host cores)
static inline void spin_step (unsigned &spins)
{
  if (++spins < 512) TJB_PAUSE (); else { spins = 0; std::this_thread::yield (); }
}

static inline bool is_acgt (uint8_t b) { return b == 'A' || b == 'C' || b == 'G' || b == 'T'; }

long tjb_pack_scalar (const uint8_t *src, size_t n, uint8_t *packed, uint32_t *exc_pos, uint8_t *exc_val, size_t cap, uint32_t pos0)
{
  size_t ne = 0;
  for (size_t i = 0; i < n; i += 4) {
    unsigned v = 0;
    for (size_t j = 0; j < 4 && i + j < n; j++) {
      const uint8_t b = src[i + j];
      v |= ((b >> 1) & 3u) << (2 * j);
      if (!is_acgt (b)) {
        if (ne >= cap) return -1;
        exc_pos[ne] = pos0 + (uint32_t) (i + j);
        exc_val[ne] = b;
        ne++;
      }
    }
    packed[i >> 2] = (uint8_t) v;
  }
  return (long) ne;
}

#if defined(__x86_64__)
// n is a multiple of 32 here
__attribute__ ((target ("avx2"))) static long pack_avx2 (const uint8_t *src, size_t n, uint8_t *packed, uint32_t *exc_pos, uint8_t *exc_val,
                                                          size_t cap, uint32_t pos0)
{
  const __m256i three = _mm256_set1_epi8 (3);
  const __m256i lut = _mm256_setr_epi8 ('A', 'C', 'T', 'G', 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 'A', 'C', 'T', 'G', 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0);
  const __m256i w14 = _mm256_set1_epi16 (0x0401);     // bytes (1, 4): c0 + 4 c1 per 16-bit lane
  const __m256i w116 = _mm256_set1_epi32 (0x00100001); // words (1, 16): + 16 (c2 + 4 c3) per 32-bit lane
  const __m256i pick = _mm256_setr_epi8 (0, 4, 8, 12, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, 0, 4, 8, 12, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1);
  size_t ne = 0;
  for (size_t i = 0; i < n; i += 32) {
    const __m256i v = _mm256_loadu_si256 ((const __m256i*) (src + i));
    const __m256i c = _mm256_and_si256 (_mm256_srli_epi16 (v, 1), three);
    const __m256i ok = _mm256_cmpeq_epi8 (_mm256_shuffle_epi8 (lut, c), v);
    const __m256i t = _mm256_madd_epi16 (_mm256_maddubs_epi16 (c, w14), w116); // one packed byte per 32-bit lane
    const __m256i q = _mm256_shuffle_epi8 (t, pick);
    uint32_t lo = (uint32_t) _mm256_cvtsi256_si32 (q), hi = (uint32_t) _mm256_extract_epi32 (q, 4);
    memcpy (packed + (i >> 2), &lo, 4);
    memcpy (packed + (i >> 2) + 4, &hi, 4);
    uint32_t bad = ~(uint32_t) _mm256_movemask_epi8 (ok);
    while (bad) {
      const int j = __builtin_ctz (bad);
      bad &= bad - 1;
      if (ne >= cap) return -1;
      exc_pos[ne] = pos0 + (uint32_t) (i + j);
      exc_val[ne] = src[i + j];
      ne++;
    }
  }
  return (long) ne;
}
#endif

static long pack_range (const uint8_t *src, size_t n, uint8_t *packed, uint32_t *exc_pos, uint8_t *exc_val, size_t cap, uint32_t pos0)
{
  size_t body = 0;
  long ne = 0;
#if defined(__x86_64__)
  static const bool have_avx2 = __builtin_cpu_supports ("avx2");
  if (have_avx2) {
    body = n & ~(size_t) 31;
    ne = pack_avx2 (src, body, packed, exc_pos, exc_val, cap, pos0);
    if (ne < 0) return -1;
  }
#endif
  const long nt = tjb_pack_scalar (src + body, n - body, packed + (body >> 2), exc_pos + ne, exc_val + ne, cap - (size_t) ne, pos0 + (uint32_t) body);
  return nt < 0 ? -1 : ne + nt;
}

#define TJB_COPY_BLOCK ((size_t) 32 << 10) // bytes per unit of a copy job
#define TJB_PACK_BLOCK ((size_t) 64 << 10) // bytes per unit of work (a multiple of 32): shares are drawn, not dealt

struct tjb_packer {
  int T = 1;
  std::vector<std::thread> th;
  std::mutex mu;
  std::condition_variable cv;
  bool active = false, quit = false;      // guarded by mu
  std::atomic<bool> spinning{false};
  std::atomic<uint64_t> gen{0};           // jobs issued so far; every worker takes part in every job exactly once
  std::atomic<uint64_t> done{0};
  // the job
  const uint8_t *src = nullptr; size_t n = 0, cap = 0, n_blocks = 0;
  uint8_t *packed = nullptr; uint32_t *exc_pos = nullptr; uint8_t *exc_val = nullptr;
  uint8_t *copy_dst = nullptr;            // non-null: the job is a copy src -> copy_dst (tjb_packer_copy / _expand)
  bool expand = false;                    // ... of n 24-byte rows into 40-byte elements
  std::atomic<size_t> next_block{0}, exc_total{0};
  std::atomic<bool> overflow{false};
  std::vector<std::vector<uint32_t>> lpos; // exceptions of the block a participant is working on
  std::vector<std::vector<uint8_t>> lval;
};

// participants 0 .. T-1 are the workers, T is the calling thread.  Blocks are drawn from a counter, so a participant
// that lost its core for a while costs one block, not a share; a block's exceptions are appended to the chunk's list
// at an offset reserved with one atomic (their order does not matter: the device scatters them by position).
// destination is written once and not read back here: streaming stores skip the read-for-ownership of its lines
static void copy_block (uint8_t *d, const uint8_t *s, size_t n)
{
#if defined(__x86_64__)
  size_t head = (16 - ((uintptr_t) d & 15)) & 15;
  if (head > n) head = n;
  memcpy (d, s, head);
  d += head; s += head; n -= head;
  const size_t blocks = n / 64;
  for (size_t i = 0; i < blocks; i++, d += 64, s += 64) {
    const __m128i a = _mm_loadu_si128 ((const __m128i*) s), b = _mm_loadu_si128 ((const __m128i*) (s + 16));
    const __m128i c = _mm_loadu_si128 ((const __m128i*) (s + 32)), e = _mm_loadu_si128 ((const __m128i*) (s + 48));
    _mm_stream_si128 ((__m128i*) d, a); _mm_stream_si128 ((__m128i*) (d + 16), b);
    _mm_stream_si128 ((__m128i*) (d + 32), c); _mm_stream_si128 ((__m128i*) (d + 48), e);
  }
  _mm_sfence ();
  memcpy (d, s, n - blocks * 64);
#else
  memcpy (d, s, n);
#endif
}

// 24-byte rows {ctx0, ctx1, bit-fields} -> 40-byte hopo_element: the trailing four int32 are all -1 at this point
static void expand_block (uint8_t *dst, const uint8_t *src, size_t n_rows)
{
  size_t i = 0;
#if defined(__x86_64__)
  auto one = [&] (size_t r) {
    long long v[3];
    memcpy (v, src + 24 * r, 24);
    long long *d = (long long*) (dst + 40 * r); // malloc()ed base, multiples of 40: 8-byte aligned
    _mm_stream_si64 (d, v[0]); _mm_stream_si64 (d + 1, v[1]); _mm_stream_si64 (d + 2, v[2]);
    _mm_stream_si64 (d + 3, -1LL); _mm_stream_si64 (d + 4, -1LL);
  };
  if (n_rows && ((uintptr_t) dst & 15)) one (i++); // elements are 40 bytes: every other one starts 16-byte aligned
  const __m128i ones = _mm_set1_epi32 (-1);
  for (; i + 2 <= n_rows; i += 2) { // two elements = 48 bytes in, five 16-byte streaming stores out
    const __m128i a = _mm_loadu_si128 ((const __m128i*) (src + 24 * i)), b = _mm_loadu_si128 ((const __m128i*) (src + 24 * i + 16));
    const __m128i c = _mm_loadu_si128 ((const __m128i*) (src + 24 * i + 32));
    __m128i *d = (__m128i*) (dst + 40 * i);
    _mm_stream_si128 (d, a);                              // ctx0, ctx1
    _mm_stream_si128 (d + 1, _mm_unpacklo_epi64 (b, ones)); // bits, -1
    _mm_stream_si128 (d + 2, _mm_unpackhi_epi64 (ones, b)); // -1, ctx0'
    _mm_stream_si128 (d + 3, c);                          // ctx1', bits'
    _mm_stream_si128 (d + 4, ones);                       // -1, -1
  }
  if (i < n_rows) one (i++);
  _mm_sfence ();
#else
  for (; i < n_rows; i++) {
    long long v[3];
    memcpy (v, src + 24 * i, 24);
    long long *d = (long long*) (dst + 40 * i);
    d[0] = v[0]; d[1] = v[1]; d[2] = v[2]; d[3] = -1LL; d[4] = -1LL;
  }
#endif
}
void tjb_expand_rows (void *dst, const void *src, size_t n_rows) { expand_block ((uint8_t*) dst, (const uint8_t*) src, n_rows); }
#define TJB_EXPAND_BLOCK ((size_t) 1024)  // rows per unit of an expand job

static void packer_take_part (tjb_packer *p, int who)
{
  if (p->copy_dst && p->expand) {
    for (;;) {
      const size_t b = p->next_block.fetch_add (1, std::memory_order_relaxed);
      if (b >= p->n_blocks) return;
      const size_t lo = b * TJB_EXPAND_BLOCK, hi = lo + TJB_EXPAND_BLOCK < p->n ? lo + TJB_EXPAND_BLOCK : p->n;
      expand_block (p->copy_dst + 40 * lo, p->src + 24 * lo, hi - lo);
    }
  }
  if (p->copy_dst) {
    for (;;) {
      const size_t b = p->next_block.fetch_add (1, std::memory_order_relaxed);
      if (b >= p->n_blocks) return;
      const size_t lo = b * TJB_COPY_BLOCK, hi = lo + TJB_COPY_BLOCK < p->n ? lo + TJB_COPY_BLOCK : p->n;
      copy_block (p->copy_dst + lo, p->src + lo, hi - lo);
    }
  }
  uint32_t *lp = p->lpos[who].data ();
  uint8_t *lv = p->lval[who].data ();
  for (;;) {
    const size_t b = p->next_block.fetch_add (1, std::memory_order_relaxed);
    if (b >= p->n_blocks) break;
    const size_t lo = b * TJB_PACK_BLOCK, hi = lo + TJB_PACK_BLOCK < p->n ? lo + TJB_PACK_BLOCK : p->n;
    const long cnt = pack_range (p->src + lo, hi - lo, p->packed + (lo >> 2), lp, lv, TJB_PACK_BLOCK, (uint32_t) lo);
    if (cnt > 0) {
      const size_t off = p->exc_total.fetch_add ((size_t) cnt, std::memory_order_relaxed);
      if (off + (size_t) cnt <= p->cap) {
        memcpy (p->exc_pos + off, lp, (size_t) cnt * sizeof (uint32_t));
        memcpy (p->exc_val + off, lv, (size_t) cnt);
      } else p->overflow.store (true, std::memory_order_relaxed);
    }
  }
}

static void packer_worker (tjb_packer *p, int tid)
{
  uint64_t seen = 0;
  unsigned spins = 0;
  for (;;) {
    {
      std::unique_lock<std::mutex> lk (p->mu);
      p->cv.wait (lk, [&] { return p->quit || p->active; });
      if (p->quit) return;
    }
    while (p->spinning.load (std::memory_order_acquire)) {
      const uint64_t g = p->gen.load (std::memory_order_acquire);
      if (g == seen) { spin_step (spins); continue; }
      seen++; // jobs are issued one at a time: g == seen + 1
      packer_take_part (p, tid);
      p->done.fetch_add (1, std::memory_order_acq_rel);
    }
    // end () cleared `active` under the mutex before `spinning`: the wait at the top parks this worker again
  }
}

// hardware threads this process may actually use: the affinity mask, capped by a cgroup v2 CPU quota if there is one
// (spinning workers beyond that are worse than none)
static int usable_cpus ()
{
  int n = (int) std::thread::hardware_concurrency ();
#if defined(__linux__)
  cpu_set_t set;
  if (sched_getaffinity (0, sizeof (set), &set) == 0) { const int c = CPU_COUNT (&set); if (c > 0 && (n <= 0 || c < n)) n = c; }
  if (FILE *f = fopen ("/sys/fs/cgroup/cpu.max", "r")) {
    long long quota = 0, period = 0;
    if (fscanf (f, "%lld %lld", &quota, &period) == 2 && quota > 0 && period > 0) {
      const int q = (int) ((quota + period - 1) / period);
      if (q < n) n = q;
    }
    fclose (f);
  }
#endif
  return n < 1 ? 1 : n;
}

tjb_packer *tjb_packer_create (int n_threads)
{
  if (n_threads <= 0) { // leave two hardware threads to the caller and the driver: the workers spin
    n_threads = usable_cpus () - 2;
    if (n_threads > 32) n_threads = 32;
    if (n_threads < 1) n_threads = 1;
  }
  tjb_packer *p = new tjb_packer;
  p->T = n_threads;
  p->lpos.resize (n_threads + 1); p->lval.resize (n_threads + 1);
  for (int t = 0; t <= n_threads; t++) { p->lpos[t].resize (TJB_PACK_BLOCK); p->lval[t].resize (TJB_PACK_BLOCK); }
  for (int t = 0; t < n_threads; t++) p->th.emplace_back (packer_worker, p, t);
  return p;
}

void tjb_packer_destroy (tjb_packer *p)
{
  if (!p) return;
  { std::lock_guard<std::mutex> lk (p->mu); p->quit = true; p->active = false; }
  p->spinning.store (false, std::memory_order_release);
  p->cv.notify_all ();
  for (auto &t : p->th) t.join ();
  delete p;
}

int tjb_packer_threads (const tjb_packer *p) { return p ? p->T : 0; }

void tjb_packer_begin (tjb_packer *p)
{
  { std::lock_guard<std::mutex> lk (p->mu); p->active = true; p->spinning.store (true, std::memory_order_release); }
  p->cv.notify_all ();
}

void tjb_packer_end (tjb_packer *p)
{
  { std::lock_guard<std::mutex> lk (p->mu); p->active = false; p->spinning.store (false, std::memory_order_release); }
  p->cv.notify_all ();
}

long tjb_packer_run (tjb_packer *p, const uint8_t *src, size_t n, uint8_t *packed, uint32_t *exc_pos, uint8_t *exc_val, size_t cap)
{
  if (!n) return 0;
  p->src = src; p->n = n; p->packed = packed; p->exc_pos = exc_pos; p->exc_val = exc_val; p->cap = cap;
  p->copy_dst = nullptr;
  p->n_blocks = (n + TJB_PACK_BLOCK - 1) / TJB_PACK_BLOCK;
  p->next_block.store (0, std::memory_order_relaxed);
  p->exc_total.store (0, std::memory_order_relaxed);
  p->overflow.store (false, std::memory_order_relaxed);
  const uint64_t g = p->gen.fetch_add (1, std::memory_order_acq_rel) + 1;
  packer_take_part (p, p->T); // the caller packs too instead of only waiting
  unsigned spins = 0;
  while (p->done.load (std::memory_order_acquire) < (uint64_t) p->T * g) spin_step (spins);
  const size_t total = p->exc_total.load (std::memory_order_relaxed);
  return (p->overflow.load (std::memory_order_relaxed) || total > cap) ? -1 : (long) total;
}

void tjb_packer_copy (tjb_packer *p, void *dst, const void *src, size_t n)
{
  if (!n) return;
  p->src = (const uint8_t*) src; p->n = n; p->copy_dst = (uint8_t*) dst; p->expand = false;
  p->n_blocks = (n + TJB_COPY_BLOCK - 1) / TJB_COPY_BLOCK;
  p->next_block.store (0, std::memory_order_relaxed);
  const uint64_t g = p->gen.fetch_add (1, std::memory_order_acq_rel) + 1;
  packer_take_part (p, p->T);
  unsigned spins = 0;
  while (p->done.load (std::memory_order_acquire) < (uint64_t) p->T * g) spin_step (spins);
}

void tjb_packer_expand (tjb_packer *p, void *dst, const void *src, size_t n_rows)
{
  if (!n_rows) return;
  p->src = (const uint8_t*) src; p->n = n_rows; p->copy_dst = (uint8_t*) dst; p->expand = true;
  p->n_blocks = (n_rows + TJB_EXPAND_BLOCK - 1) / TJB_EXPAND_BLOCK;
  p->next_block.store (0, std::memory_order_relaxed);
  const uint64_t g = p->gen.fetch_add (1, std::memory_order_acq_rel) + 1;
  packer_take_part (p, p->T);
  unsigned spins = 0;
  while (p->done.load (std::memory_order_acquire) < (uint64_t) p->T * g) spin_step (spins);
}

// test hook (tests/test_hostpack.py): one chunk through a fresh pool of `threads` workers
extern "C" long tjb_hostpack_selftest (const uint8_t *src, size_t n, int threads, uint8_t *packed, uint32_t *exc_pos, uint8_t *exc_val, size_t cap)
{
  if (threads == 0) return tjb_pack_scalar (src, n, packed, exc_pos, exc_val, cap, 0);
  tjb_packer *p = tjb_packer_create (threads);
  tjb_packer_begin (p);
  long r = -2;
  for (int rep = 0; rep < 3; rep++) r = tjb_packer_run (p, src, n, packed, exc_pos, exc_val, cap); // the pool is reusable
  tjb_packer_end (p);
  tjb_packer_begin (p); // and can be parked and woken again
  const long r2 = tjb_packer_run (p, src, n, packed, exc_pos, exc_val, cap);
  // the same pool copies: packed -> a scratch buffer and back must leave it unchanged
  const size_t np = (n + 3) / 4;
  std::vector<uint8_t> tmp (np + 1, 0xAA);
  tjb_packer_begin (p);
  tjb_packer_copy (p, tmp.data (), packed, np);
  bool same = memcmp (tmp.data (), packed, np) == 0 && tmp[np] == 0xAA;
  { // and expands 24-byte rows: the packed bytes serve as rows
    const size_t rows = np / 24;
    std::vector<long long> e (5 * rows + 1, 0x5555);
    tjb_packer_expand (p, e.data (), packed, rows);
    for (size_t i = 0; i < rows && same; i++)
      same = memcmp (&e[5 * i], packed + 24 * i, 24) == 0 && e[5 * i + 3] == -1LL && e[5 * i + 4] == -1LL;
    same = same && e[5 * rows] == 0x5555;
  }
  tjb_packer_end (p);
  tjb_packer_destroy (p);
  return !same ? -4 : r == r2 ? r : -3;
}
extern "C" void tjb_hostpack_expand_hook (void *dst, const void *src, size_t n_rows) { tjb_expand_rows (dst, src, n_rows); }

// measurement hook (tools/pack_bench.py): seconds of the best of `reps` passes over src[0, n) in chunks of `chunk` bytes
// through a pool of `threads` workers, like tjb_count_sample_host() drives it (without the copies)
#include <chrono>
extern "C" double tjb_hostpack_bench (const uint8_t *src, size_t n, int threads, int reps, size_t chunk)
{
  tjb_packer *p = tjb_packer_create (threads);
  std::vector<uint8_t> packed (chunk / 4 + 64), val (chunk / 16);
  std::vector<uint32_t> pos (chunk / 16);
  double best = 1e30;
  tjb_packer_begin (p);
  for (int r = 0; r < reps; r++) {
    const auto t0 = std::chrono::steady_clock::now ();
    for (size_t off = 0; off < n; off += chunk)
      tjb_packer_run (p, src + off, n - off < chunk ? n - off : chunk, packed.data (), pos.data (), val.data (), pos.size ());
    const double s = std::chrono::duration<double> (std::chrono::steady_clock::now () - t0).count ();
    if (s < best) best = s;
  }
  tjb_packer_end (p);
  tjb_packer_destroy (p);
  return best;
}

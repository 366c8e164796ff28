/* Host-side FASTQ/FASTA(.gz) decoder feeding the GPU path.
 *
 * Stands in for the reference's reader loop  `while (bmc2_kseq_read (seq) >= 0) update_hopo_counter_from_seq (...)`
 * (src/hopo_counter.c:153) and follows the record rules of its klib-style parser (src/kseq.h:172-212):
 *   - records start at the next '>' or '@'; the rest of that line is the header;
 *   - sequence lines are concatenated (empty lines skipped, one trailing '\r' per line dropped) until a line
 *     starts with '>', '@' or '+';
 *   - after a '+' line, quality lines are concatenated until at least as long as the sequence;
 *   - a record whose quality is missing or of a different length, and everything after it, is discarded
 *     (the parser returns -2 and the reference's loop stops on any negative value).
 * Instead of handing one read at a time to a scalar scan, reads are packed as "SEQ\n" into large chunks
 * (pinned, double buffered by the caller) so that the device scans them with '\n' as the read separator.
 * This is our own implementation (block-wise memchr over the decoded pieces), not the klib macros.  Regular .gz files
 * are inflated by tjb_inflate.c (own decoder on a producer thread); everything else (plain text, pipes) is read through
 * zlib's gzread() like the reference does.  TJB_ZLIB=1 in the environment sends .gz files through zlib as well.
 */
#include <stdlib.h>
#include <string.h>
#include <zlib.h>
#include "tjb_gpu.h"

#define TJB_IOBUF (1u << 20)

struct tjb_reader {
  gzFile fp;             /* zlib path ... */
  tjb_gzstream *gz;      /* ... or our own inflate: io then points at the stream's current piece */
  uint8_t *io_own;       /* buffer of the zlib path */
  const uint8_t *io;     /* decoded bytes being parsed */
  size_t io_pos, io_end;
  int io_eof;
  int pending_header;    /* a header char ('>' / '@') was consumed while reading sequence lines */
  int finished;          /* EOF or truncated record seen: nothing more to deliver */
  uint8_t *seq;          /* current record */
  size_t seq_len, seq_cap;
  uint8_t *qual;         /* its quality string (only its length matters, but the '\r' rule needs the bytes) */
  size_t qual_len, qual_cap;
  int seq_ready;         /* a complete record is waiting in seq[] (did not fit in the previous chunk) */
};

static int refill (tjb_reader *r)
{
  if (r->io_eof) return 0;
  if (r->gz) {
    const int64_t n = tjb_gz_next (r->gz, &r->io);
    r->io_pos = 0;
    /* corrupt or truncated data: what could be decoded has been delivered, the file ends here (gzread() < 0) */
    if (n <= 0) { r->io_end = 0; r->io_eof = 1; return 0; }
    r->io_end = (size_t) n;
    return 1;
  }
  int got = gzread (r->fp, r->io_own, TJB_IOBUF);
  r->io = r->io_own;
  r->io_pos = 0;
  if (got <= 0) { r->io_end = 0; r->io_eof = 1; return 0; }
  r->io_end = (size_t) got;
  if ((unsigned) got < TJB_IOBUF && gzeof (r->fp)) r->io_eof = 1;
  return 1;
}

static inline int next_char (tjb_reader *r)
{
  if (r->io_pos >= r->io_end && !refill (r)) return -1;
  return r->io[r->io_pos++];
}

static int buf_reserve (uint8_t **buf, size_t len, size_t *cap_io, size_t extra)
{
  if (len + extra <= *cap_io) return 1;
  size_t cap = *cap_io ? *cap_io : 4096;
  while (cap < len + extra) cap *= 2;
  uint8_t *p = (uint8_t*) realloc (*buf, cap);
  if (!p) return 0;
  *buf = p; *cap_io = cap;
  return 1;
}

/* consume the rest of the current line (up to and including '\n', or to EOF); when buf is given the bytes are
 * appended to it.  Returns 1 normally, -1 when out of memory. */
static int take_line (tjb_reader *r, uint8_t **buf, size_t *len, size_t *cap)
{
  for (;;) {
    if (r->io_pos >= r->io_end && !refill (r)) break;
    const uint8_t *p = r->io + r->io_pos;
    size_t avail = r->io_end - r->io_pos;
    const uint8_t *nl = (const uint8_t*) memchr (p, '\n', avail);
    size_t n = nl ? (size_t) (nl - p) : avail;
    if (buf) {
      if (!buf_reserve (buf, *len, cap, n + 1)) return -1;
      memcpy (*buf + *len, p, n);
      *len += n;
    }
    r->io_pos += n + (nl ? 1 : 0);
    if (nl) break;
  }
  return 1;
}

tjb_reader *tjb_reader_open (const char *path)
{
  tjb_reader *r = (tjb_reader*) calloc (1, sizeof (tjb_reader));
  if (!r) return NULL;
  const char *force_zlib = getenv ("TJB_ZLIB");
  if (!(force_zlib && force_zlib[0] == '1')) r->gz = tjb_gz_open (path); /* NULL unless a regular gzip file */
  if (r->gz) return r;
  r->fp = gzopen (path, "rb");
  if (!r->fp) { free (r); return NULL; }
  gzbuffer (r->fp, 1u << 20);
  r->io_own = (uint8_t*) malloc (TJB_IOBUF);
  if (!r->io_own) { gzclose (r->fp); free (r); return NULL; }
  return r;
}

void tjb_reader_close (tjb_reader *r)
{
  if (!r) return;
  if (r->fp) gzclose (r->fp);
  tjb_gz_close (r->gz);
  free (r->io_own); free (r->seq); free (r->qual); free (r);
}

/* parse one record into seq[]; returns 1 = record ready, 0 = no more records, -1 = out of memory */
static int next_record (tjb_reader *r)
{
  int c;
  if (r->finished) return 0;
  if (!r->pending_header) { /* jump to the next header line */
    while ((c = next_char (r)) != -1 && c != '>' && c != '@');
    if (c == -1) { r->finished = 1; return 0; }
  }
  r->pending_header = 0;
  r->seq_len = 0;
  /* header line; EOF straight after the header char ends the file (the reference's getuntil() < 0) */
  if (r->io_pos >= r->io_end && !refill (r)) { r->finished = 1; return 0; }
  take_line (r, NULL, NULL, NULL);
  /* sequence lines */
  for (;;) {
    c = next_char (r);
    if (c == -1 || c == '>' || c == '+' || c == '@') break;
    if (c == '\n') continue;
    if (!buf_reserve (&r->seq, r->seq_len, &r->seq_cap, 2)) return -1;
    r->seq[r->seq_len++] = (uint8_t) c;
    if (take_line (r, &r->seq, &r->seq_len, &r->seq_cap) < 0) return -1;
    if (r->seq_len > 1 && r->seq[r->seq_len - 1] == '\r') r->seq_len--; /* src/kseq.h:124 */
  }
  if (c == '>' || c == '@') r->pending_header = 1;
  if (c != '+') { if (c == -1) r->finished = 1; return 1; } /* FASTA record (or last record before EOF) */
  /* '+' line */
  while ((c = next_char (r)) != -1 && c != '\n');
  if (c == -1) { r->finished = 1; return 0; } /* no quality string: record discarded, file ends (-2) */
  /* quality lines until at least seq_len bytes; same per-line '\r' rule on the accumulated string */
  r->qual_len = 0;
  for (;;) {
    if (r->io_pos >= r->io_end && !refill (r)) break; /* getuntil() < 0 */
    if (take_line (r, &r->qual, &r->qual_len, &r->qual_cap) < 0) return -1;
    if (r->qual_len > 1 && r->qual[r->qual_len - 1] == '\r') r->qual_len--;
    if (r->qual_len >= r->seq_len) break;
  }
  if (r->qual_len != r->seq_len) { r->finished = 1; return 0; } /* truncated / mismatching quality (-2) */
  return 1;
}

/* The common case without the per-line machinery: a four-line FASTQ record that lies completely in the decoded piece
 * (header, one sequence line, '+' line, one quality line of the same length, no '\r').  Copies the sequence straight to
 * dst and returns its length + 1; 0 = not that shape (next_record() decides), -1 = does not fit into dst any more. */
static inline int64_t fast_record (tjb_reader *r, uint8_t *dst, size_t room)
{
  const uint8_t *p = r->io + r->io_pos, *const end = r->io + r->io_end;
  if (p >= end || *p != '@') return 0;
  const uint8_t *nl = (const uint8_t*) memchr (p, '\n', (size_t) (end - p));
  if (!nl) return 0;
  const uint8_t *seq = nl + 1;
  nl = (const uint8_t*) memchr (seq, '\n', (size_t) (end - seq));
  if (!nl) return 0;
  const size_t len = (size_t) (nl - seq);
  if (!len || seq[len - 1] == '\r' || seq[0] == '>' || seq[0] == '@' || seq[0] == '+') return 0;
  if (end - nl < 3 || nl[1] != '+') return 0;
  nl = (const uint8_t*) memchr (nl + 1, '\n', (size_t) (end - nl - 1));
  if (!nl) return 0;
  const uint8_t *qual = nl + 1;
  if ((size_t) (end - qual) < len + 1 || qual[len] != '\n' || memchr (qual, '\n', len)) return 0;
  if (len > 1 && qual[len - 1] == '\r') return 0;
  if (len + 1 > room) return -1;
  memcpy (dst, seq, len);
  dst[len] = '\n';
  r->io_pos = (size_t) (qual + len + 1 - r->io);
  return (int64_t) len + 1;
}

int64_t tjb_reader_fill (tjb_reader *r, uint8_t *dst, size_t cap, int64_t *n_reads, int64_t *n_bases)
{
  size_t w = 0;
  for (;;) {
    if (!r->seq_ready && !r->pending_header && !r->finished) {
      const int64_t got = fast_record (r, dst + w, cap - w);
      if (got > 0) {
        w += (size_t) got;
        if (n_reads) (*n_reads)++;
        if (n_bases) (*n_bases) += got - 1;
        continue;
      }
      if (got < 0 && w) break; /* chunk full: the record is parsed again by the next call */
    }
    if (!r->seq_ready) {
      int rc = next_record (r);
      if (rc < 0) return TJB_ERR_NOMEM;
      if (rc == 0) break;
      r->seq_ready = 1;
    }
    if (r->seq_len + 1 > cap) {                          /* a single read larger than a whole chunk: */
      if (w) break;                                      /* what the chunk holds goes first, */
      return TJB_ERR_TOO_LARGE;                          /* then the caller takes it with tjb_reader_take_long() */
    }
    if (w + r->seq_len + 1 > cap) break;                 /* keep it for the next chunk */
    memcpy (dst + w, r->seq, r->seq_len);
    w += r->seq_len;
    dst[w++] = '\n';
    if (n_reads) (*n_reads)++;
    if (n_bases) (*n_bases) += (int64_t) r->seq_len;
    r->seq_ready = 0;
  }
  return (int64_t) w;
}

/* After tjb_reader_fill() returned TJB_ERR_TOO_LARGE: the read that does not fit into a chunk, as it stands in the
 * reader's own buffer (no separator behind it; valid until the next call on this reader).  The reference accepts reads of
 * any length (src/kseq.h grows its string), e.g. an assembled genome given as a sample.  Returns its length, 0 when no
 * read is waiting. */
int64_t tjb_reader_take_long (tjb_reader *r, const uint8_t **seq, int64_t *n_reads, int64_t *n_bases)
{
  if (!r || !seq || !r->seq_ready) return 0;
  *seq = r->seq;
  r->seq_ready = 0;
  if (n_reads) (*n_reads)++;
  if (n_bases) (*n_bases) += (int64_t) r->seq_len;
  return (int64_t) r->seq_len;
}

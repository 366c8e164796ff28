#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "tjb_gpu.h"
int main (int argc, char **argv)
{
  for (int a = 1; a < argc; a++) {
    for (size_t cap = 64; cap <= 4096; cap *= 4) {
      tjb_reader *r = tjb_reader_open (argv[a]);
      if (!r) { printf ("open failed %s\n", argv[a]); return 1; }
      uint8_t *buf = malloc (cap);
      int64_t nr = 0, nb = 0, total = 0, longs = 0;
      for (;;) {
        int64_t got = tjb_reader_fill (r, buf, cap, &nr, &nb);
        if (got == TJB_ERR_TOO_LARGE) {
          const uint8_t *q; int64_t n = tjb_reader_take_long (r, &q, &nr, &nb);
          if (n <= 0) { printf ("take_long gave %lld\n", (long long) n); return 1; }
          unsigned long s = 0; for (int64_t i = 0; i < n; i++) s += q[i];
          total += n + 1; longs++; (void) s;
          continue;
        }
        if (got < 0) { printf ("error %lld\n", (long long) got); return 1; }
        if (!got) break;
        total += got;
      }
      printf ("%s cap %zu: %lld bytes, %lld reads, %lld bases, %lld long\n", argv[a], cap, (long long) total, (long long) nr, (long long) nb, (long long) longs);
      free (buf);
      tjb_reader_close (r);
    }
  }
  return 0;
}

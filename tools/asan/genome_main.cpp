#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <algorithm>
#include "tjb_gpu.h"
int main (int argc, char **argv)
{
  // histograms on the golden elements
  FILE *f = fopen ("elem.bin", "rb"); fseek (f, 0, SEEK_END); long sz = ftell (f); fseek (f, 0, SEEK_SET);
  std::vector<unsigned char> e (sz); if (fread (e.data (), 1, sz, f) != (size_t) sz) return 1; fclose (f);
  tjb_hist_opts o = {atoi (argv[1]), atoi (argv[2]), atoi (argv[3]), atoi (argv[4])};
  for (int n : {0, 1, 2, 7, (int) (sz / 40)}) {
    tjb_context_table t;
    int rc = tjb_context_histograms (&o, e.data (), n, &t);
    printf ("hist n=%d -> %d\n", n, rc);
    tjb_context_table_free (&t);
  }
  // merges of growing and shrinking sizes (the thread-local row arrays are reused and regrown)
  srand (3);
  for (int total_samples : {1, 3, 40, 96, 5}) {
    std::vector<int32_t> n_hist, loc, base, mlen; std::vector<char> names;
    const int stride = 12;
    for (int s = 0; s < total_samples; s++) {
      int n = rand () % 3000;
      std::vector<int32_t> l (n); for (auto &x : l) x = (rand () % 4000) * (rand () % 3 ? 8 : 1) - 100;
      std::sort (l.begin (), l.end ());
      n_hist.push_back (n);
      for (int i = 0; i < n; i++) { loc.push_back (l[i]); base.push_back (rand () % 2); mlen.push_back (2 + rand () % 12);
        char nm[stride]; for (int j = 0; j < stride - 1; j++) nm[j] = "ACGT"[rand () % 4]; nm[stride - 1] = 0; names.insert (names.end (), nm, nm + stride); }
    }
    size_t tot = loc.size ();
    std::vector<int32_t> os (tot), oi (tot), ot (tot);
    int rc = tjb_merge_tracts (total_samples, n_hist.data (), loc.data (), base.data (), mlen.data (), names.data (), stride, 2, os.data (), oi.data (), ot.data ());
    printf ("merge samples=%d rows=%zu -> %d, tracts %d\n", total_samples, tot, rc, tot ? ot[tot - 1] + 1 : 0);
  }
  return 0;
}

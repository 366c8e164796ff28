#!/bin/bash
# AddressSanitizer + UBSan over the host C / C++ of the library (no GPU): the FASTQ / FASTA(.gz) reader with reads
# longer than the chunk at several chunk sizes, tjb_context_histograms on the golden elements, tjb_merge_tracts on random
# tables of changing size (its row arrays are kept between calls).  usage: bash tools/asan/run.sh   (from the repo root)
set -e
W=$(mktemp -d); R=$PWD
python - "$W" <<'PY'
import gzip, random, sys, numpy as np
w = sys.argv[1]; random.seed(1)
seq = lambda n: ''.join(random.choice('ACGTN') for _ in range(n))
recs = [seq(n) for n in (10, 500, 70, 3000, 1, 64, 63, 65, 9000, 20)]
fq = ''.join(f"@r{i}\n{s}\n+\n{'I' * len(s)}\n" for i, s in enumerate(recs))
open(f"{w}/a.fastq", "w").write(fq); gzip.open(f"{w}/a.fastq.gz", "wt").write(fq)
open(f"{w}/a.fasta", "w").write(''.join(f">r{i}\n" + '\n'.join(s[j:j + 60] for j in range(0, len(s), 60)) + "\n" for i, s in enumerate(recs)))
open(f"{w}/trunc.fastq", "w").write("@r0\nACGT\n+\nIII")
z = np.load("tests/golden/hist/context_histograms.npz")
open(f"{w}/elem.bin", "wb").write(z["elem"].tobytes())
open(f"{w}/opts", "w").write(f"{int(z['k'])} {int(z['m'])} {int(z['d'])} {int(z['lev'])}")
PY
F="-g -O1 -fsanitize=address,undefined -fno-omit-frame-pointer -pthread -I$R/include"
gcc -std=gnu11 $F tools/asan/reader_main.c tatajuba_b200/csrc/tjb_fastq.c tatajuba_b200/csrc/tjb_inflate.c -lz -o $W/rd
g++ -std=c++17 $F tools/asan/genome_main.cpp tatajuba_b200/csrc/tjb_genome.cpp -o $W/gn
(cd $W && ./rd a.fastq a.fastq.gz a.fasta trunc.fastq && ./gn $(cat opts))
rm -rf $W

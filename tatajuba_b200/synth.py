"""Seeded synthetic inputs for tests and bench: a random microbial genome with injected
homopolymer-length polymorphism, paired-end reads packed as "SEQ\\n" (the layout csrc/tjb_fastq.c produces),
and FASTQ writers.  numpy only; measurement harness, not part of the compute path."""
import gzip

import numpy as np

_ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)
_COMP = np.zeros(256, dtype=np.uint8)
for _a, _b in zip(b"ACGTNacgtn", b"TGCANtgcan"):
    _COMP[_a] = _b


def random_genome(length, seed=1):
    rng = np.random.default_rng(seed)
    return _ACGT[rng.integers(0, 4, size=length, dtype=np.uint8)]


def inject_tract_polymorphism(genome, fraction=0.02, seed=100, min_len=4):
    """+-1 base length change on a fraction of the homopolymer tracts with len >= min_len"""
    rng = np.random.default_rng(seed)
    change = np.flatnonzero(np.diff(genome)) + 1
    starts = np.concatenate(([0], change))
    lens = np.diff(np.concatenate((starts, [genome.size])))
    vals = genome[starts]
    cand = np.flatnonzero(lens >= min_len)
    pick = cand[rng.random(cand.size) < fraction]
    lens = lens.copy()
    lens[pick] += rng.choice(np.array([-1, 1]), size=pick.size)
    return np.repeat(vals, lens)


def revcomp(a):
    return _COMP[a[..., ::-1]]


def paired_reads(genome, coverage, read_len=150, insert=(250, 500), seed=7, error_rate=0.0, max_pairs=None):
    """-> uint8 array shaped (2*n_pairs, read_len+1): rows are R1/R2 alternating, last column '\\n'"""
    rng = np.random.default_rng(seed)
    G = genome.size
    n_pairs = int(G * coverage / (2 * read_len))
    if max_pairs is not None:
        n_pairs = min(n_pairs, max_pairs)
    ins = rng.integers(insert[0], insert[1] + 1, size=n_pairs)
    start = (rng.random(n_pairs) * (G - ins)).astype(np.int64)
    flip = rng.random(n_pairs) < 0.5
    win = np.lib.stride_tricks.sliding_window_view(genome, read_len)
    left = win[start]                       # forward strand, fragment start
    right = win[start + ins - read_len]     # forward strand, fragment end
    r1 = np.where(flip[:, None], revcomp(right), left)
    r2 = np.where(flip[:, None], left, revcomp(right))
    out = np.empty((2 * n_pairs, read_len + 1), dtype=np.uint8)
    out[0::2, :read_len] = r1
    out[1::2, :read_len] = r2
    out[:, read_len] = ord("\n")
    if error_rate > 0:  # substitutions at random positions (a position drawn twice is substituted twice: harmless)
        n_err = int(rng.binomial(2 * n_pairs * read_len, error_rate))
        row = rng.integers(0, 2 * n_pairs, size=n_err)
        col = rng.integers(0, read_len, size=n_err)
        out[row, col] = _ACGT[rng.integers(0, 4, size=n_err)]
    return out


def sample_buffer(genome_len, coverage, seed=1, sample=0, fraction=0.02, read_len=150, error_rate=0.0,
                  max_pairs=None):
    """packed "SEQ\\n" buffer (1-D uint8) of one synthetic sample; R1 and R2 interleaved"""
    g = inject_tract_polymorphism(random_genome(genome_len, seed), fraction, seed=100 + sample)
    return paired_reads(g, coverage, read_len, seed=1000 + sample, error_rate=error_rate,
                        max_pairs=max_pairs).reshape(-1)


def write_fastq(path, rows, prefix="r", gz=None):
    """rows: (n, L+1) uint8 with trailing '\\n' column (as from paired_reads) -> 4-line FASTQ"""
    gz = path.endswith(".gz") if gz is None else gz
    L = rows.shape[1] - 1
    qual = b"I" * L
    op = (lambda p: gzip.open(p, "wb", compresslevel=1)) if gz else (lambda p: open(p, "wb"))
    with op(path) as f:
        parts = []
        for i, r in enumerate(rows):
            parts.append(b"@%s%d\n%s+\n%s\n" % (prefix.encode(), i, r.tobytes(), qual))
            if len(parts) >= 8192:
                f.write(b"".join(parts))
                parts = []
        f.write(b"".join(parts))


def adversarial_reads(k=25, seed=3, wrap_lengths=False):
    """hand-made edge cases: tracts at both read edges, runs of N, lower case, reads shorter
    than 2k+m, a tract >= 512, empty read, U bases, N inside flanks.  Returns a list of bytes (no newlines)."""
    rng = np.random.default_rng(seed)

    def rnd(n):
        return _ACGT[rng.integers(0, 4, size=n)].tobytes()

    def flank(n):  # random flank without homopolymer runs >= 2 so that only the planted tract is found
        out = bytearray()
        while len(out) < n:
            c = b"ACGT"[rng.integers(0, 4)]
            if not out or out[-1] != c:
                out.append(c)
        return bytes(out)
    reads = []
    for base in b"ACGTacgtUu":
        b = bytes([base])
        for ln in (2, 3, 4, 5, 9):
            reads.append(flank(k) + b * ln + flank(k))          # exactly fits: s == k and e == L-k-1
            reads.append(flank(k - 1) + b * ln + flank(k))      # one short on the left
            reads.append(flank(k) + b * ln + flank(k - 1))      # one short on the right
            reads.append(flank(k + 7) + b * ln + flank(k + 3))
    reads.append(flank(k) + b"A" * 511 + flank(k))
    if wrap_lengths:  # >= 512 wraps the reference's signed 10-bit length field; the reference itself then dies in
        # generate_tract_as_string (malloc of a negative size, hopo_counter.c:450-451) if such a record leads a group
        reads.append(flank(k) + b"A" * 512 + flank(k))
        reads.append(flank(k) + b"T" * 700 + flank(k))
    reads.append(b"")                                            # empty read
    reads.append(b"A")
    reads.append(b"A" * 5)
    reads.append(b"A" * (2 * k + 4))                             # one run covering the whole read
    reads.append(rnd(2 * k + 3))
    fl = bytearray(flank(k)); fl[min(3, k - 1)] = ord("N"); fl[k - 1] = ord("n")
    reads.append(bytes(fl) + b"C" * 6 + flank(k))                # N inside the left flank -> packs as 0
    reads.append(flank(k) + b"G" * 6 + bytes(fl))                # N inside the right flank, reverse strand
    reads.append(flank(k) + b"AAAA" + b"CCCC" + b"GGGG" + b"TTTT" + flank(k))  # adjacent tracts as flanks
    for _ in range(40):
        reads.append(rnd(int(rng.integers(2 * k - 5, 2 * k + 120))))
    return reads


def nrun_reads(k=25, seed=17, n_reads=400, p_run=0.03, lead_tract=True):
    """reads with qualifying runs of bytes that are none of ACGTU ("NNNN", "nnn", "----", "RRRRR") placed AFTER a real
    tract of the same read: the reference then re-emits that tract's context, base and flag with the run's length
    (src/hopo_counter.c:233-248; well defined, unlike a run that is the FIRST tract of its read, which emits
    uninitialised stack).  lead_tract=True starts every read with flank + tract + flank so that no run is first;
    False leaves first-of-read runs in (only for the restatement / CUDA comparison: they are fenced)."""
    rng = np.random.default_rng(seed)
    odd = [b"N", b"n", b"-", b"R", b"*", b"."]
    fl = (b"GATC" * 8)[:max(k, 2)]

    def body(n):
        out = bytearray()
        while len(out) < n:
            u = rng.random()
            if u < p_run:
                out += odd[int(rng.integers(0, len(odd)))] * int(rng.integers(2, 12))
            elif u < 2.5 * p_run:
                out += bytes([b"ACGTacgtU"[int(rng.integers(0, 9))]]) * int(rng.integers(2, 12))
            else:
                out.append(b"ACGT"[int(rng.integers(0, 4))])
        return bytes(out)
    reads = []
    for i in range(n_reads):
        lead = b""
        if lead_tract:
            base = [c for c in b"ACGT" if c != fl[-1] and c != fl[0]][i % 2]  # must not extend into a flank
            lead = fl + bytes([base]) * int(rng.integers(6, 10)) + fl
        reads.append(lead + body(int(rng.integers(2 * k, 2 * k + 400))) + fl)
    return reads


def write_reference(fasta_path, gff_path, genome, contig="chr1", gene_every=1200, gene_len=900, extra_contig=3000):
    """reference genome as FASTA (60 columns) and a GFF3 with one gene + CDS every `gene_every` bases (SURVEY.md §8(d)
    config C1; reader rules: biomcmc-lib gff3_format.c:197-279), followed by a second, short contig of random bases
    (more than one reference sequence, like a draft assembly)."""
    g = genome.tobytes() if hasattr(genome, "tobytes") else bytes(genome)
    contigs = [(contig, g)]
    if extra_contig:
        contigs.append((contig + "_b", random_genome(extra_contig, seed=len(g) % 9973 + 17).tobytes()))
    with open(fasta_path, "wb") as f:
        for name, seq in contigs:
            f.write(b">" + name.encode() + b"\n")
            for i in range(0, len(seq), 60):
                f.write(seq[i:i + 60] + b"\n")
    with open(gff_path, "w") as f:
        f.write("##gff-version 3\n")
        for name, seq in contigs:
            f.write(f"##sequence-region {name} 1 {len(seq)}\n")
        n = 0
        for name, seq in contigs:
            f.write(f"{name}\tsynth\tregion\t1\t{len(seq)}\t.\t+\t.\tID={name}:1..{len(seq)};Name={name}\n")
            for start in range(101, len(seq) - gene_len, gene_every):
                n += 1
                strand = "+" if n % 2 else "-"
                f.write(f"{name}\tsynth\tgene\t{start}\t{start + gene_len - 1}\t.\t{strand}\t.\tID=gene-{n:05d};Name=g{n}\n")
                f.write(f"{name}\tsynth\tCDS\t{start}\t{start + gene_len - 1}\t.\t{strand}\t0\tID=cds-{n:05d};Parent=gene-{n:05d};"
                        f"product=hypothetical protein {n}\n")

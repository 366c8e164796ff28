#!/usr/bin/env python
"""bench.py — FASTQ bases/s of the hot path (HT scan -> context histograms) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--legs a,b,...]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Headline workload (BASELINE.json configs[1]): one E. coli-size sample per GPU — 5 Mbp random genome with injected
tract-length polymorphism, 100x, 2x150 bp paired end = 500 Mbases (3.33 M reads), default -k 25 -m 4 -i 5.
A step = one full pass of the replaced span (scan + aggregate + sort + QC + groups + coverage + results on the host)
over that sample.  Every timed number is gated on parity: before any timing the unmodified reference (oracle/_ref)
counts the SAME full sample and its state at the mapper call must equal ours, device-resident and from host memory.

  value   reads resident in HBM when the timed region starts, one context (one sample at a time), results in the pinned
          arena; `config.two_contexts` adds the throughput with the steps dealt to two contexts of the GPU (the
          reference's own loop over samples is a parallel one, src/genome_set.c:66-68)
  e2e     the same through the reference-facing C ABI call with the decoded reads in pinned HOST memory: host -> device
          copies inside, results in malloc()ed host memory as the binding needs them
  extra legs (rank 0, in the same JSON line): the sample with 0.1 % and 1 % substitution errors, BASELINE configs[4]
          (k=32 m=3 300x), configs[3] (96 samples, strong scaling over the ranks, exchange + union + merge inside the
          timed region), the cross-sample exchange on its own, FASTQ files (host decode inside) beside the reference.
--impl reference: the unmodified reference (oracle/_ref) on the host cores, the same samples, one thread per sample
like its OpenMP loop (N samples on min(N, cores) threads at --gpus N).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

GENOME, COVERAGE, READ_LEN = 5_000_000, 100, 150
K, M, MIN_COV = 25, 4, 5
METRIC = "FASTQ bases/sec for HT scan+context histograms"
WORKLOAD = "configs[1]: one sample per GPU, 5 Mbp genome, 100x, 2x150 bp PE, k=25 m=4 i=5"
C4_SAMPLES, C4_GENOME, C4_COVERAGE, C4_FRACTION = 96, 1_700_000, 50, 0.05
ALL_LEGS = ["errors", "c5", "c4", "exchange", "files"]


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """samples SM clocks / throttle reasons during the timed region (NVML in-process; nvidia-smi as a fallback)"""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False
        # NVML in-process (a query costs microseconds); spawning nvidia-smi enumerates every GPU of the box under the
        # driver's lock and was seen to stretch the steps of a 2-rank run by 25 %.
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.mx = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
            pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)  # the first query of a process is the slow one
            self.nvml = (pynvml, h)
        except Exception:
            pass

    def run(self):
        if self.nvml:
            nv, h = self.nvml
            bits = [nv.nvmlClocksThrottleReasonHwSlowdown, nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                    nv.nvmlClocksThrottleReasonSwThermalSlowdown, nv.nvmlClocksThrottleReasonSwPowerCap]
            try:
                while True:  # at least one sample, taken as the timed region starts
                    sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                    self.rows.append([str(sm), str(self.mx)] + ["Active" if r & b else "Not Active" for b in bits])
                    if self.stop_flag:
                        break
                    time.sleep(0.005)
                return
            except Exception as e:
                self.nvml = None
                sys.stderr.write(f"[bench] NVML sampling failed ({e}); falling back to nvidia-smi\n")
        while not self.stop_flag:
            try:
                o = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i",
                                    str(self.index)], capture_output=True, text=True, timeout=5).stdout.strip()
                if o:
                    self.rows.append([x.strip() for x in o.split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        self.stop_flag = True
        self.join(timeout=6)
        sm = sorted(int(r[0]) for r in self.rows if r[0].isdigit())
        mx = max([int(r[1]) for r in self.rows if r[1].isdigit()] or [0])
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i] == "Active" for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": reasons,
                "samples": len(self.rows), "source": "nvml" if self.nvml else "nvidia-smi"}


def make_sample(sample, error_rate=0.0, genome=GENOME, coverage=COVERAGE, fraction=0.02):
    from tatajuba_b200 import synth
    return synth.sample_buffer(genome, coverage, seed=1, sample=sample, read_len=READ_LEN, error_rate=error_rate,
                               fraction=fraction)


class quiet_stderr:
    """the reference prints per-sample banners on stderr"""

    def __enter__(self):
        sys.stderr.flush()
        self.devnull, self.saved = os.open(os.devnull, os.O_WRONLY), os.dup(2)
        os.dup2(self.devnull, 2)

    def __exit__(self, *a):
        os.dup2(self.saved, 2)
        os.close(self.devnull)
        os.close(self.saved)


def reference_run(buf, k=K, m=M, cov=MIN_COV, rb=True):
    """the unmodified reference on one sample, one thread -> (boundary state, seconds of the replaced span)"""
    from oracle import pyoracle
    ref = pyoracle.Reference()
    b = ref.run_buffer(buf, k, m, cov, rb)
    return b, b.secs_scan + b.secs_finalise


def assert_parity(hc, want, what):
    from oracle import pyoracle
    got = pyoracle.Boundary(hc.status, hc.n_raw, hc.elem, hc.idx_initial, hc.idx_final, hc.coverage)
    pyoracle.assert_same_boundary(got, want, what)


def n_bases_of(buf):
    return buf.size // (READ_LEN + 1) * READ_LEN


# --------------------------------------------------------------------------------------- reference arm ----
def run_reference_arm(args, rank):
    """--impl reference: the reference's CPU implementation on the host cores; rank 0 only.  One step = every one of
    the N samples of the workload (N = --gpus) once, one thread per sample on min(N, cores) threads."""
    if rank != 0:
        return
    from oracle import pyoracle
    kind = "reference" if pyoracle.Reference.available() else "port"
    n = args.gpus
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    threads = min(n, cores)
    with ThreadPoolExecutor(threads) as ex:
        bufs = list(ex.map(make_sample, range(n)))
    bases = sum(n_bases_of(b) for b in bufs)
    workers = [pyoracle.Reference() if kind == "reference" else pyoracle.Restatement() for _ in range(n)]

    def one(i):
        return workers[i].run_buffer(bufs[i], K, M, MIN_COV, True)
    times = []
    with quiet_stderr(), ThreadPoolExecutor(threads) as ex:  # ctypes releases the GIL inside the C call
        for i in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            list(ex.map(one, range(n)))
            dt = time.perf_counter() - t0
            if i >= args.warmup:
                times.append(dt)
    dt = sum(times) / len(times)
    v = bases / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": "bases/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8/u64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "same_config": True, "samples_per_step": n, "bases_per_step": bases,
                   "span": "update_hopo_counter_from_seq per read + finalise_hopo_counter up to the mapper call, reads "
                           "decoded in host memory (like the GPU arm's e2e)"},
        "cpu_baseline": {"value": v, "unit": "bases/s", "cores": threads, "kind": kind, "host_cores": cores,
                         "sample": f"the full workload: {n} sample(s) of {bases // n / 1e6:.0f} Mbases, one thread per sample "
                                   "(the reference's OpenMP loop is over samples, src/genome_set.c:66-68)"},
        "gpu_launches": 0,
        "e2e": {"value": v, "unit": "bases/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    if "files" in args.legs and kind == "reference":
        line["from_files"] = files_leg_reference(bufs[0])
    print(json.dumps(line))


FILES_READS = 400_000  # 60 Mbases: writing the full sample as .gz from Python would take minutes


def write_files(d, buf):
    from tatajuba_b200 import synth
    rows = buf[: FILES_READS * (READ_LEN + 1)].reshape(FILES_READS, READ_LEN + 1)
    out = {}
    for ext in (".fastq", ".fastq.gz"):
        r1, r2 = os.path.join(d, "s_R1" + ext), os.path.join(d, "s_R2" + ext)
        synth.write_fastq(r1, rows[0::2])
        synth.write_fastq(r2, rows[1::2])
        out["gz" if ext.endswith(".gz") else "plain"] = [r1, r2]
    return out


def files_leg_reference(buf):
    from oracle import pyoracle
    ref = pyoracle.Reference()
    leg = {"sample": f"first {FILES_READS} reads of the workload as paired FASTQ (R1/R2), plain and gzip -1; kseq decode inside",
           "threads": 1}
    with tempfile.TemporaryDirectory() as d, quiet_stderr():
        for kind, files in write_files(d, buf).items():
            best = None
            for _ in range(2):
                b = ref.run_files(files, K, M, MIN_COV, True)
                dt = b.secs_scan + b.secs_finalise
                best = dt if best is None else min(best, dt)
            leg[kind] = {"value": FILES_READS * READ_LEN / best, "unit": "bases/s", "ms": best * 1e3}
    return leg


# --------------------------------------------------------------------------------------------- our arm ----
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--legs", default=",".join(ALL_LEGS), help="extra legs to run (comma separated; 'none' for the headline only)")
    ap.add_argument("--contexts", type=int, default=1, help="contexts per GPU the steps of the headline are dealt to (the C4 leg always uses 2)")
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the parity gate / CPU baseline (kernel experiments only)")
    args = ap.parse_args()
    args.legs = [x for x in args.legs.split(",") if x and x != "none"]
    rank = int(os.environ.get("RANK", 0))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    if args.impl == "reference":
        run_reference_arm(args, rank)
        return

    import torch
    import torch.distributed as dist
    from tatajuba_b200 import Context, TatajubaOptions, genome_set
    assert torch.cuda.is_available(), "bench.py needs a GPU (there is no CPU fallback)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def all_ranks_ok(ok):
        if world == 1:
            return ok
        t = torch.tensor([1 if ok else 0], dtype=torch.int64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        return bool(t.item())

    opt = TatajubaOptions(K, M, MIN_COV, True, paired_end=True)
    ctxs = [Context(local_rank) for _ in range(max(1, args.contexts))]
    ctx = ctxs[0]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    by_rank = {}

    def timed(fn, tag=None):
        """device time of fn() between two events on an otherwise idle stream, bracketed by barriers; max over ranks"""
        barrier()
        e0.record()
        out = fn()
        torch.cuda.synchronize()
        e1.record()
        barrier()
        mine = e0.elapsed_time(e1)
        if tag and world > 1:  # every rank's own time, for the record (the reported one is the maximum)
            t = torch.tensor([mine], dtype=torch.float64, device="cuda")
            allr = [torch.zeros_like(t) for _ in range(world)]
            dist.all_gather(allr, t)
            by_rank[tag] = [round(float(x.item()), 4) for x in allr]
        elif tag:
            by_rank[tag] = [round(mine, 4)]
        return max_over_ranks(mine), out

    def count_steps(steps, count_one, contexts):
        return genome_set.count_samples(contexts, steps, count_one)

    def kernel_table(c, run, steps=5):
        """per-kernel device times (events around every launch): a separate, untimed pass"""
        c.set_profiling(2)
        c.reset_kernel_stats()
        for _ in range(steps):
            run()
        st = c.kernel_stats()
        c.set_profiling(0)
        return {s["name"]: {"launches": s["launches"] // steps, "us": round(s["ms_total"] / steps * 1e3, 1)} for s in st}

    # ---- the sample, and the parity gate at full size (outside every timed region) ---------------------------------
    buf = make_sample(rank)
    n_bytes, n_bases, n_reads = buf.size, n_bases_of(buf), buf.size // (READ_LEN + 1)
    h_pinned = torch.from_numpy(buf).pin_memory()
    d_buf = h_pinned.cuda()
    torch.cuda.synchronize()
    parity, cpu_base = "skipped (--no-cpu-baseline)", None
    if not args.no_cpu_baseline:
        from oracle import pyoracle
        kind = "reference" if pyoracle.Reference.available() else "port"
        with quiet_stderr():
            if kind == "reference":
                want, secs = reference_run(buf)
            else:
                t0 = time.perf_counter()
                want = pyoracle.Restatement().run_buffer(buf, K, M, MIN_COV, True)
                secs = time.perf_counter() - t0
        assert_parity(ctx.count_device(d_buf.data_ptr(), n_bytes, opt), want, "parity gate, device-resident")
        assert_parity(ctx.count_host(h_pinned, opt), want, "parity gate, host buffer")
        ctx.set_result_mode(True)
        assert_parity(ctx.count_device(d_buf.data_ptr(), n_bytes, opt), want, "parity gate, arena results")
        ctx.set_result_mode(False)
        parity = f"{kind}, full sample (state at the mapper call: elements, group indices, coverage; device, host and arena paths)"
        cpu_base = {"value": n_bases / secs, "unit": "bases/s", "cores": 1, "kind": kind, "seconds": secs,
                    "sample": f"the full sample ({n_bases / 1e6:.0f} Mbases) decoded in host memory, one thread (the reference "
                              "parallelises over samples only); the same run is the parity gate"}

    # ---- value: device-resident, steps dealt to the contexts of this GPU, results in the arena -----------------------
    barrier()  # ranks leave the parity gate at different times: wait for them BEFORE the warm-up, not with idle GPUs after it
    for c in ctxs:
        c.set_result_mode(True)
        for _ in range(args.warmup):
            hc = c.count_device(d_buf.data_ptr(), n_bytes, opt)
        assert hc.status == 0
    stats0 = dict(n_raw=hc.n_raw, n_distinct=hc.n_distinct, n_elem=hc.n_elem, n_idx=hc.n_idx, coverage=hc.coverage)
    for c in ctxs:
        c.set_profiling(1)  # the scan kernel is timed live, inside the timed region (events on the context's stream)
        c.reset_kernel_stats()
    sampler = ClockSampler(local_rank)
    if rank == 0 and not os.environ.get("BENCH_NO_SAMPLER"):
        sampler.start()
    step_wall = []

    def one_step(c, s):
        t0 = time.perf_counter()
        r = c.count_device(d_buf.data_ptr(), n_bytes, opt)
        step_wall.append(time.perf_counter() - t0)
        return r
    ms, _ = timed(lambda: count_steps(args.steps, one_step, ctxs), tag="value_ms_by_rank")
    if os.environ.get("BENCH_DEBUG"):
        sys.stderr.write(f"[bench] rank {rank}: {ms:.3f} ms for {args.steps} steps; host wall per step (us): "
                         + " ".join(f"{x * 1e6:.0f}" for x in step_wall) + "\n")
    clocks = sampler.summary() if rank == 0 else None
    scan_ms = scan_launches = scan_bytes = 0
    launches = 0
    for c in ctxs:
        for s in c.kernel_stats():
            if s["name"] == "tjb_scan_kernel":
                scan_ms += s["ms_total"]; scan_launches += s["launches"]; scan_bytes += s["bytes_algorithmic"]
        launches += c.launch_count()
        c.set_profiling(0)
    ms_single = ms
    two = None
    if len(ctxs) == 1:  # the same steps dealt to two contexts (sample i's reduce and copy-out under sample i+1's scan)
        pair = [ctx, Context(local_rank)]
        pair[1].set_result_mode(True)
        for _ in range(args.warmup):
            pair[1].count_device(d_buf.data_ptr(), n_bytes, opt)
        ms2, _ = timed(lambda: count_steps(args.steps, lambda c, s: c.count_device(d_buf.data_ptr(), n_bytes, opt), pair))
        two = {"ms_per_step": ms2 / args.steps, "value": world * n_bases * args.steps / (ms2 * 1e-3), "unit": "bases/s"}
        pair[1].close()
    else:
        ms_single, _ = timed(lambda: [ctx.count_device(d_buf.data_ptr(), n_bytes, opt) for _ in range(args.steps)])
    kernels = kernel_table(ctx, lambda: ctx.count_device(d_buf.data_ptr(), n_bytes, opt))

    # ---- e2e: pinned host buffer -> C ABI -> malloc()ed host results --------------------------------------------------
    for c in ctxs:
        c.set_result_mode(False)
        for _ in range(2):
            hh = c.count_host(h_pinned, opt)
        c.reset_kernel_stats()
    ms_e2e, _ = timed(lambda: count_steps(args.steps, lambda c, s: c.count_host(h_pinned, opt), ctxs))
    h2d = sum(c.h2d_bytes() for c in ctxs) // args.steps
    d2h = 40 * hh.n_elem + 8 * hh.n_idx + 88
    ms_e2e_single, _ = timed(lambda: [ctx.count_host(h_pinned, opt) for _ in range(max(3, args.steps // 2))])
    ms_e2e_single /= max(3, args.steps // 2) / args.steps
    packing = os.environ.get("TJB_HOST_PACK")
    os.environ["TJB_HOST_PACK"] = "0"
    ms_plain, _ = timed(lambda: [ctx.count_host(h_pinned, opt) for _ in range(3)])
    if packing is None:
        del os.environ["TJB_HOST_PACK"]
    else:
        os.environ["TJB_HOST_PACK"] = packing
    transfer = {"mode": "2-bit packed + exceptions" if h2d < n_bytes else "plain", "host_bytes_per_step": n_bytes,
                "one_context": {"ms_per_step": ms_e2e_single / args.steps},
                "plain": {"ms_per_step": ms_plain / 3, "value": world * n_bases / (ms_plain / 3 * 1e-3), "unit": "bases/s"}}

    extra = {}
    # ---- the sample with substitution errors (errors create singleton contexts: the opposite of the 64:1 collapse) ----
    if "errors" in args.legs and world == 1:
        extra["errors"] = {}
        for err in (0.001, 0.01):
            extra["errors"][str(err)] = single_sample_leg(ctxs, opt, make_sample(0, error_rate=err), args, timed, count_steps, kernel_table)
    if "c5" in args.legs and world == 1:
        from tatajuba_b200 import synth
        c5 = synth.sample_buffer(GENOME, 300, seed=3)
        extra["c5"] = single_sample_leg(ctxs, TatajubaOptions(32, 3, 5, True, paired_end=True), c5, args, timed, count_steps,
                                        kernel_table, k=32, m=3, steps=max(3, args.steps // 2))
        extra["c5"]["workload"] = "configs[4]: k=32 (3-word records), m=3, 300x of a 5 Mbp genome = 1.5 Gbases"
        del c5
    if "c5" in args.legs and world > 1:
        try:  # an extra leg: what goes wrong in it on every rank alike is reported in the line, not allowed to take the headline along
            extra["c5_sharded"] = c5_sharded_leg(ctx, rank, world, timed, all_ranks_ok)
        except Exception as e:  # noqa: BLE001
            extra["c5_sharded"] = {"error": f"{type(e).__name__}: {e}"[:300]}
            print(f"[bench] c5_sharded leg failed on rank {rank}: {e}", file=sys.stderr)
    if "c4" in args.legs:
        extra["c4"] = c4_leg(ctxs, rank, world, timed, all_ranks_ok, args)
    if "exchange" in args.legs:
        for c in ctxs:
            c.set_result_mode(True)
        hh = ctx.count_device(d_buf.data_ptr(), n_bytes, opt)
        rows = genome_set.query_rows_device(ctx, hh).clone()

        def exchange():
            all_rows, sizes = genome_set.allgather_tables_device({rank: rows}, world)
            return genome_set.union_queries_device(ctx, all_rows, K), sizes
        exchange()
        ms_x = min(timed(exchange)[0] for _ in range(3))
        (uni, _idx), sizes = exchange()
        extra["exchange"] = {"ms": ms_x, "samples": world, "rows": int(sum(sizes)), "n_union": int(uni.shape[0]),
                             "collective": "none (1 rank)" if world == 1 else "nccl all_gather_into_tensor x2 (counts, rows), device to device",
                             "path": "query rows written on the device, gathered GPU to GPU, union through the aggregation table"}
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    if "files" in args.legs and world == 1:
        leg = {"sample": f"first {FILES_READS} reads of the workload as paired FASTQ (R1/R2), plain and gzip -1; host decode inside",
               "decoder_threads": 2}
        ctx.set_result_mode(False)
        with tempfile.TemporaryDirectory() as d:
            for kind, files in write_files(d, buf).items():
                best = None
                for _ in range(3):
                    t0 = time.perf_counter()
                    hf = ctx.count_files(files, opt)
                    dt = time.perf_counter() - t0
                    best = dt if best is None else min(best, dt)
                leg[kind] = {"value": hf.n_bases / best, "unit": "bases/s", "ms": best * 1e3}
        from oracle import pyoracle
        if pyoracle.Reference.available() and not args.no_cpu_baseline:
            leg["reference"] = files_leg_reference(buf)
        extra["from_files"] = leg

    peak, peak_src = load_peaks()
    scan_ms_l = scan_ms / max(scan_launches, 1)
    scan_bytes_l = scan_bytes / max(scan_launches, 1)
    achieved = scan_bytes_l / (scan_ms_l * 1e-3) / 1e9 if scan_ms_l else 0.0
    kernel_us = sum(v["us"] for v in kernels.values())
    line = {
        "metric": METRIC, "value": world * n_bases * args.steps / (ms * 1e-3), "unit": "bases/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8/u64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "bases_per_gpu": n_bases, "reads_per_gpu": n_reads, "input_bytes": n_bytes,
                   "l2": "input (503 MB) larger than L2 (126 MB); no flush needed", "parity": parity,
                   "contexts_per_gpu": len(ctxs), "one_context_ms_per_step": ms_single / args.steps, "two_contexts": two,
                   "value_ms_by_rank": by_rank.get("value_ms_by_rank"), **stats0},
        "e2e": {"value": world * n_bases * args.steps / (ms_e2e * 1e-3), "unit": "bases/s", "ms_per_step": ms_e2e / args.steps,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "results": "malloc()ed host arrays (reference layout)",
                "transfer": transfer},
        "gpu_launches": launches,
        "roofline": {"kernel": "tjb_scan_kernel", "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "peak_source": peak_src, "traffic": None,
                     "ms_per_launch": scan_ms_l, "algorithmic_bytes_per_launch": scan_bytes_l,
                     "share_of_kernel_time": kernels.get("tjb_scan_kernel", {}).get("us", 0) / kernel_us if kernel_us else None,
                     "pipeline": {"algorithmic_bytes_per_step": int(n_bytes + 96 * stats0["n_raw"]),
                                  "note": "SURVEY.md 8(d): (1 + 96 rho) N bytes per sample over one context's step time",
                                  "frac": (n_bytes + 96 * stats0["n_raw"]) / (ms_single / args.steps * 1e-3) / 1e9 / peak}},
        "kernels": kernels,
        "clocks": clocks,
    }
    line.update(extra)
    tp = os.path.join(ROOT, "profiles", "scan_traffic.json")  # dram bytes of one scan launch from an ncu --set full capture
    if os.path.exists(tp):
        tr = json.load(open(tp))
        line["roofline"]["traffic"] = tr["dram_bytes_read"] + tr["dram_bytes_write"]
        line["roofline"]["traffic_source"] = tr.get("source")
    if cpu_base:
        line["cpu_baseline"] = cpu_base
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def single_sample_leg(ctxs, opt, buf, args, timed, count_steps, kernel_table, k=K, m=M, steps=None):
    """one sample, device-resident: parity against the reference at full size, steps over the contexts, kernel table"""
    import torch
    steps = steps or args.steps
    d = torch.from_numpy(buf).cuda()
    n = buf.size
    leg = {"bases": n_bases_of(buf)}
    for c in ctxs:
        c.set_result_mode(True)
    if not args.no_cpu_baseline:
        from oracle import pyoracle
        if pyoracle.Reference.available():
            with quiet_stderr():
                want, secs = reference_run(buf, k, m)
            assert_parity(ctxs[0].count_device(d.data_ptr(), n, opt), want, f"parity, k={k} m={m}")
            leg["parity"] = "reference, full sample"
            leg["reference_bases_per_s"] = leg["bases"] / secs
    for c in ctxs:
        for _ in range(2):
            hc = c.count_device(d.data_ptr(), n, opt)
    ms, _ = timed(lambda: count_steps(steps, lambda c, s: c.count_device(d.data_ptr(), n, opt), ctxs))
    ms1, _ = timed(lambda: [ctxs[0].count_device(d.data_ptr(), n, opt) for _ in range(steps)])
    leg.update({"ms_per_step": ms / steps, "value": leg["bases"] * steps / (ms * 1e-3), "unit": "bases/s",
                "one_context_ms_per_step": ms1 / steps, "n_raw": hc.n_raw, "n_distinct": hc.n_distinct, "n_elem": hc.n_elem,
                "kernels": kernel_table(ctxs[0], lambda: ctxs[0].count_device(d.data_ptr(), n, opt), steps=3)})
    return leg


def c5_sharded_leg(ctx, rank, world, timed, all_ranks_ok):
    """BASELINE configs[4] (k=32, m=3, 300x of 5 Mbp = 1.5 Gbases) as ONE sample split by read over the ranks (SURVEY.md
    8e "by read chunk within a sample"): every rank scans its chunk from HBM into its own table, the tables cross NVLink
    as rows (one NCCL allgather), every rank adds the others' rows and finalises the complete table.  Strong scaling:
    the sample is the same at every N.  Parity: every rank's result against its own one-GPU count of the whole sample
    (which the N = 1 `c5` leg compares with the compiled reference)."""
    import numpy as np
    import torch
    from oracle import pyoracle
    from tatajuba_b200 import TatajubaOptions, genome_set, synth
    opt5 = TatajubaOptions(32, 3, 5, True, paired_end=True)
    buf = synth.sample_buffer(GENOME, 300, seed=3)
    n_bases = n_bases_of(buf)
    rows = buf.reshape(-1, READ_LEN + 1)
    lo, hi = rows.shape[0] * rank // world, rows.shape[0] * (rank + 1) // world
    chunk = torch.from_numpy(np.ascontiguousarray(rows[lo:hi]).reshape(-1)).cuda()
    whole = torch.from_numpy(buf).cuda()
    ctx.set_result_mode(True)
    one = ctx.count_device(whole.data_ptr(), whole.numel(), opt5)
    want = pyoracle.Boundary(one.status, one.n_raw, one.elem.copy(), one.idx_initial.copy(), one.idx_final.copy(), one.coverage)
    ms_one = min(timed(lambda: ctx.count_device(whole.data_ptr(), whole.numel(), opt5))[0] for _ in range(3))
    del whole

    def run():
        return genome_set.count_sample_sharded(ctx, chunk.data_ptr(), chunk.numel(), opt5, total_bytes_hint=buf.size)
    got = run()
    ok = True
    try:
        pyoracle.assert_same_boundary(pyoracle.Boundary(got.status, got.n_raw, got.elem, got.idx_initial, got.idx_final, got.coverage),
                                      want, f"rank {rank}: configs[4] in {world} read chunks against one GPU")
    except AssertionError as e:
        ok = False
        print(f"[bench] {e}", file=sys.stderr)
    if not all_ranks_ok(ok):  # every rank learns it and leaves the leg together
        raise RuntimeError("configs[4] sharded by read chunk differs from the one-GPU result")
    run()
    ms = min(timed(run, "c5_sharded" if i == 0 else None)[0] for i in range(3))
    return {"workload": "configs[4] as ONE sample split by read over the ranks: k=32, m=3, 300x of a 5 Mbp genome = 1.5 Gbases",
            "scaling": "strong", "ranks": world, "ms_per_sample": ms, "value": n_bases / (ms * 1e-3), "unit": "bases/s",
            "one_gpu_ms_per_sample": ms_one, "speedup_over_one_gpu": ms_one / ms,
            "exchange": "table rows, one nccl all_gather_into_tensor, device to device; QC / groups / coverage on the complete table on every rank",
            "parity": "every rank against its own one-GPU count of the whole sample (that count against the reference: the c5 leg at N = 1)",
            "n_raw": int(got.n_raw), "n_elem": int(got.n_elem)}


def c4_leg(ctxs, rank, world, timed, all_ranks_ok, args):
    """BASELINE configs[3]: 96 samples of a 1.7 Mbp population (5 % of the tracts differ between samples), 50x PE.  Strong
    scaling: the samples are dealt to the ranks, every rank deals ITS samples dynamically to its contexts (reads in
    pinned host memory: H2D inside), then the query rows are exchanged (one allgather, GPU to GPU) and every rank
    computes the union of the mapper queries; rank 0 runs the cross-sample merge (tjb_merge_tracts) on a stand-in
    post-mapper table.  Two timed parts, added up: first sample -> union, and the merge; building the stand-in tables
    (in place of the mapper) happens between them and is not timed."""
    import torch
    from tatajuba_b200 import Context, TatajubaOptions, genome_set
    opt = TatajubaOptions(K, M, MIN_COV, True, paired_end=True)
    if len(ctxs) < 2:  # two contexts per GPU: sample i's reduce, copy-out and host work under sample i+1's copy-in and scan
        ctxs = list(ctxs) + [Context(ctxs[0].device)]
    mine = genome_set.samples_of_rank(C4_SAMPLES, rank, world)
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    threads = max(1, cores // max(1, int(os.environ.get("LOCAL_WORLD_SIZE", world))))
    with ThreadPoolExecutor(threads) as ex:
        bufs = list(ex.map(lambda s: make_sample(s, genome=C4_GENOME, coverage=C4_COVERAGE, fraction=C4_FRACTION), mine))
    pinned = [torch.from_numpy(b).pin_memory() for b in bufs]
    bases_total = C4_SAMPLES * n_bases_of(bufs[0])
    parity = "skipped"
    for c in ctxs:
        c.set_result_mode(False)
    if not args.no_cpu_baseline:
        from oracle import pyoracle
        if pyoracle.Reference.available():
            refs = [pyoracle.Reference() for _ in range(threads)]
            with quiet_stderr(), ThreadPoolExecutor(threads) as ex:
                wants = list(ex.map(lambda i: refs[i % threads].run_buffer(bufs[i], K, M, MIN_COV, True), range(len(bufs))))
            res = genome_set.count_samples(ctxs, len(bufs), lambda c, i: c.count_host(pinned[i], opt))
            for i, hc in enumerate(res):
                assert_parity(hc, wants[i], f"c4 sample {mine[i]}")
            parity = f"reference, every one of the {C4_SAMPLES} samples at full size"
    ok = all_ranks_ok(True)
    assert ok
    for c in ctxs:
        c.set_result_mode(True)
    ctx = ctxs[0]

    resident = [p.cuda() for p in pinned]  # 8.2 GB at one GPU
    torch.cuda.synchronize()

    stages = {}

    def front(from_host=True):
        """count of every sample -> query rows on the device -> allgather -> union of the mapper queries"""
        rows = {}
        t0 = time.perf_counter()

        def done(c, i, hc):  # on the context's thread, before it counts its next sample
            rows[mine[i]] = genome_set.query_rows_device(c, hc).clone()
            return hc.n_idx
        one = (lambda c, i: c.count_host(pinned[i], opt)) if from_host else (lambda c, i: c.count_device(resident[i].data_ptr(), resident[i].numel(), opt))
        genome_set.count_samples(ctxs, len(bufs), one, on_done=done)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        all_rows, sizes = genome_set.allgather_tables_device(rows, C4_SAMPLES)
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        uni, idx = genome_set.union_queries_device(ctx, all_rows, K)
        t3 = time.perf_counter()
        stages[from_host] = {"count_samples": (t1 - t0) * 1e3, "allgather": (t2 - t1) * 1e3, "union": (t3 - t2) * 1e3}
        return uni, idx, sizes

    def by_rank(ms):  # outside the timed parts: every rank's own counting time (a rank that is done early waits in the allgather)
        if world == 1:
            return [round(ms, 2)]
        import torch.distributed as dist
        mine_ms = torch.tensor([ms], dtype=torch.float64, device="cuda")
        allr = [torch.zeros_like(mine_ms) for _ in range(world)]
        dist.all_gather(allr, mine_ms)
        return [round(float(x.item()), 2) for x in allr]

    def standin_tables(idx, sizes):
        """NOT timed (it stands in for the mapper, which is out of scope): post-mapper tables for rank 0's merge, one
        histogram per union row a sample has, located by the row's rank in the union"""
        if rank != 0:
            return None
        tables, off, idx_h = [], 0, idx.cpu().numpy()
        for n in sizes:
            loc = (np.sort(idx_h[off:off + n]) * 8).astype(np.int32)  # tracts 8 bp apart: neighbours are different tracts
            off += n
            tables.append({"location": loc, "base": np.zeros(n, np.int32), "mode_context_length": np.full(n, 4, np.int32),
                           "name": np.zeros((n, 1), np.uint8)})
        return tables

    def back(tables):
        """the cross-sample merge of the tract ids (tjb_merge_tracts: host code, rank 0, all host threads)"""
        return len(genome_set.merge_tracts(tables, 2)[0]) if rank == 0 else 0

    best_by_mode, front_by_mode, back_by_mode = {}, {}, {}
    out = None
    for from_host in (False, True):
        front(from_host)
        best = None
        for _ in range(3):
            ms_f, (uni, idx, sizes) = timed(lambda: front(from_host))
            stages[from_host]["count_samples_by_rank"] = by_rank(stages[from_host]["count_samples"])
            tables = standin_tables(idx, sizes)
            ms_b, n_concat = timed(lambda: back(tables))
            if best is None or ms_f + ms_b < best:
                best, front_by_mode[from_host], back_by_mode[from_host] = ms_f + ms_b, ms_f, ms_b
                stages[from_host]["merge_tracts_rank0"] = ms_b
                kept = dict(stages[from_host])
            out = (int(uni.shape[0]), int(sum(sizes)), n_concat)
        stages[from_host] = kept
        best_by_mode[from_host] = best
    best = best_by_mode[False]
    return {"workload": f"configs[3]: {C4_SAMPLES} samples x {C4_GENOME / 1e6:.1f} Mbp x {C4_COVERAGE}x PE, tract polymorphism on {C4_FRACTION:.0%} of the tracts",
            "scaling": "strong (the 96 samples are dealt to the ranks)", "ms": best, "value": bases_total / (best * 1e-3), "unit": "bases/s",
            "samples_per_rank": len(mine), "contexts_per_gpu": len(ctxs), "bases_total": bases_total, "parity": parity,
            "inside": "count of every sample (reads resident in HBM), query rows on the device, allgather, union of mapper queries, merge of tract ids (rank 0, host code); "
                      "the stand-in post-mapper tables are built between the two timed parts, outside them",
            "device_span": {"ms": front_by_mode[False], "value": bases_total / (front_by_mode[False] * 1e-3), "unit": "bases/s",
                            "what": "count + query rows + allgather + union: the part that runs on the GPUs and shards over them"},
            "merge_rank0_ms": back_by_mode[False],
            "stages_ms": {k: (round(v, 2) if not isinstance(v, list) else v) for k, v in stages[False].items()},
            "from_host": {"ms": best_by_mode[True], "stages_ms": {k: (round(v, 2) if not isinstance(v, list) else v) for k, v in stages[True].items()}, "value": bases_total / (best_by_mode[True] * 1e-3), "unit": "bases/s",
                          "note": "the same with the reads of every sample in pinned host memory: H2D inside"},
            "n_union": out[0], "rows": out[1], "merged_rows": out[2]}


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""bench.py — FASTQ bases/s of the hot path (HT scan -> sort -> context histograms) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json configs[1]): one E. coli-size sample per GPU — 5 Mbp random genome with injected
tract-length polymorphism, 100x, 2x150 bp paired end = 500 Mbases (3.33 M reads), default -k 25 -m 4 -i 5.
A step = one full pass of the replaced span (scan + sort + reduce + QC + groups + coverage + copy-out) over that
sample.  `value`: reads resident in HBM when the timed region starts.  `e2e`: the same through the C ABI with the
decoded reads in pinned HOST memory (H2D inside the call, results copied back to malloc'ed host memory).
Weak scaling: every rank processes its own sample; no data-path collective on this part of the path.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

GENOME, COVERAGE, READ_LEN = 5_000_000, 100, 150
K, M, MIN_COV = 25, 4, 5
CPU_SAMPLE_READS = 1_000_000
METRIC = "FASTQ bases/sec for HT scan+context histograms"


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """samples nvidia-smi clocks / throttle reasons during the timed region"""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False
        # NVML in-process (a query costs microseconds); spawning nvidia-smi enumerates every GPU of the box under the
        # driver's lock and was seen to stretch the steps of a 2-rank run by 25 %.  Initialised here, before the timed
        # region, so that the first sample is taken as soon as the thread starts.
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
                    time.sleep(0.02)  # a query takes the driver's lock for ~1 ms on an 8-GPU box: keep them rare
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


def make_sample(rank, n_reads=None):
    from tatajuba_b200 import synth
    buf = synth.sample_buffer(GENOME, COVERAGE, seed=1, sample=rank, read_len=READ_LEN,
                              max_pairs=None if n_reads is None else n_reads // 2)
    return buf


def cpu_baseline(buf, n_reads):
    """the reference's own code (oracle/_ref) — or the plain-C port if it did not travel — on a bounded sample"""
    from oracle import pyoracle
    sample = buf[: n_reads * (READ_LEN + 1)]
    bases = n_reads * READ_LEN
    if pyoracle.Reference.available():
        ref, kind = pyoracle.Reference(), "reference"
        devnull, saved = os.open(os.devnull, os.O_WRONLY), os.dup(2)
        os.dup2(devnull, 2)  # the reference prints per-sample banners on stderr
        try:
            t0 = time.perf_counter()
            ref.run_buffer(sample.tobytes(), K, M, MIN_COV, True)
            dt = time.perf_counter() - t0
        finally:
            os.dup2(saved, 2)
            os.close(devnull)
    else:
        o, kind = pyoracle.Restatement(), "port"
        t0 = time.perf_counter()
        o.run_buffer(sample, K, M, MIN_COV, True)
        dt = time.perf_counter() - t0
    return {"value": bases / dt, "unit": "bases/s", "cores": 1, "kind": kind, "seconds": dt,
            "sample": f"first {n_reads} reads ({bases / 1e6:.0f} Mbases) of the workload, decoded in host memory; "
                      "update_hopo_counter_from_seq per read + finalise_hopo_counter up to the mapper call; the "
                      "reference runs one sample on one thread (OpenMP is over samples only)"}


def cpu_all_cores(buf, n_reads):
    """the reference's own parallelism is one OpenMP thread per SAMPLE (genome_set.c:67): with as many samples as host
    cores in flight, every thread runs the same bounded sample; aggregate bases/s (information beside the 1-thread
    number, which is what the reference does with the single sample of this workload)"""
    from concurrent.futures import ThreadPoolExecutor
    from oracle import pyoracle
    if not pyoracle.Reference.available():
        return None
    cores = os.cpu_count() or 1
    sample = buf[: n_reads * (READ_LEN + 1)].tobytes()
    refs = [pyoracle.Reference() for _ in range(cores)]
    devnull, saved = os.open(os.devnull, os.O_WRONLY), os.dup(2)
    os.dup2(devnull, 2)
    try:
        with ThreadPoolExecutor(cores) as ex:  # ctypes releases the GIL inside the C call
            t0 = time.perf_counter()
            list(ex.map(lambda r: r.run_buffer(sample, K, M, MIN_COV, True), refs))
            dt = time.perf_counter() - t0
    finally:
        os.dup2(saved, 2)
        os.close(devnull)
    return {"value": cores * n_reads * READ_LEN / dt, "unit": "bases/s", "cores": cores, "seconds": dt,
            "sample": f"{cores} concurrent copies of the bounded sample, one per thread"}


def run_reference_arm(args, rank):
    """--impl reference: the reference's CPU implementation on the host cores; rank 0 only"""
    if rank != 0:
        return
    n_reads = CPU_SAMPLE_READS
    buf = make_sample(0, n_reads)
    times = []
    base = None
    for i in range(args.warmup + args.steps):
        base = cpu_baseline(buf, n_reads)
        if i >= args.warmup:
            times.append(base["seconds"])
    dt = sum(times) / len(times)
    v = n_reads * READ_LEN / dt
    base["value"] = v
    try:
        base["all_cores"] = cpu_all_cores(buf, n_reads)
    except Exception as e:  # information only
        base["all_cores"] = {"error": str(e)}
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "bases/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8/u64", "data": "synthetic",
        "config": {"workload": "configs[1]: 5 Mbp genome, 100x, 2x150 bp PE, k=25 m=4 i=5 (bounded sample per step)",
                   "reads_per_step": n_reads},
        "cpu_baseline": base, "gpu_launches": 0,
        "e2e": {"value": v, "unit": "bases/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", 0))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    if args.impl == "reference":
        run_reference_arm(args, rank)
        return

    # The result arrays are malloc()ed by the library and free()d by this script every step.  Keep freed memory in the
    # heap (no trim, no mmap per array): otherwise whether a step pays ~0.5 ms of page faults for fresh pages depends
    # on what else touched the allocator (seen: the first timed steps after an NCCL barrier), which is not the path.
    try:
        import ctypes
        libc = ctypes.CDLL("libc.so.6")
        libc.mallopt(-1, 1 << 30)  # M_TRIM_THRESHOLD
        libc.mallopt(-3, 1 << 30)  # M_MMAP_THRESHOLD
    except Exception:
        pass
    import torch
    import torch.distributed as dist
    from tatajuba_b200 import Context, TatajubaOptions
    assert torch.cuda.is_available(), "bench.py needs a GPU (there is no CPU fallback)"
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        if os.environ.get("BENCH_DEBUG"):
            sys.stderr.write(f"[bench] rank {rank}: {x:.3f} ms\n")
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    opt = TatajubaOptions(K, M, MIN_COV, True, paired_end=True)
    buf = make_sample(rank)
    n_bytes = buf.size
    n_reads = n_bytes // (READ_LEN + 1)
    n_bases = n_reads * READ_LEN
    ctx = Context(local_rank)
    stream = torch.cuda.ExternalStream(ctx.stream(), device=local_rank)
    h_pinned = torch.from_numpy(buf).pin_memory()
    d_buf = h_pinned.cuda()
    torch.cuda.synchronize()

    # ---- device-resident arm (value) + roofline of the dominant kernel --------------------------------
    for _ in range(args.warmup):
        hc = ctx.count_device(d_buf.data_ptr(), n_bytes, opt)
    assert hc.status == 0
    ctx.set_profiling(2)
    ctx.reset_kernel_stats()
    sampler = ClockSampler(local_rank)
    if rank == 0 and not os.environ.get("BENCH_NO_SAMPLER"):  # one poller per job: 8 nvidia-smi loops would steal host cores from the ranks' launch threads
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for _ in range(args.steps):
        hc = ctx.count_device(d_buf.data_ptr(), n_bytes, opt)
    e1.record(stream)
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1))
    clocks = sampler.summary() if rank == 0 else None
    stats = ctx.kernel_stats()
    launches = ctx.launch_count()
    ctx.set_profiling(0)

    # ---- end-to-end arm: pinned host buffer -> C ABI -> host results ----------------------------------
    # (large host buffers cross PCIe 2-bit packed + exceptions, packed by host threads inside the timed region and
    # rebuilt on the device; TJB_HOST_PACK=0 = the bytes as they are, timed next to it over 3 steps)
    def e2e_arm(steps):
        for _ in range(2):
            ctx.count_host(h_pinned, opt)
        barrier()
        ctx.reset_kernel_stats()
        e0.record(stream)
        for _ in range(steps):
            h = ctx.count_host(h_pinned, opt)
        e1.record(stream)
        barrier()
        return max_over_ranks(e0.elapsed_time(e1)) / steps, ctx.h2d_bytes() // steps, h
    packing = os.environ.get("TJB_HOST_PACK")
    os.environ["TJB_HOST_PACK"] = "0"
    ms_plain, h2d_plain, _ = e2e_arm(3)
    if packing is None:
        del os.environ["TJB_HOST_PACK"]
    else:
        os.environ["TJB_HOST_PACK"] = packing
    ms_e2e_step, h2d, hh = e2e_arm(args.steps)
    ms_e2e = ms_e2e_step * args.steps
    d2h = 24 * hh.n_elem + 8 * hh.n_idx + 64  # an element crosses PCIe without its 16 constant bytes
    transfer = {"mode": "2-bit packed + exceptions" if h2d < n_bytes else "plain", "host_bytes_per_step": n_bytes,
                "host_threads": os.cpu_count() if packing is None else packing,
                "plain": {"ms_per_step": ms_plain, "h2d_bytes_per_step": h2d_plain,
                          "value": world * n_bases / (ms_plain * 1e-3), "unit": "bases/s"}}

    # ---- cross-sample exchange: query rows -> allgather over the ranks -> union on the GPU ---------
    from tatajuba_b200 import genome_set
    exch = []
    for _ in range(3):
        barrier()
        t0 = time.perf_counter()
        rows = genome_set.query_rows(hh)
        tables = genome_set.allgather_tables({rank: rows}, world)
        union, _index = genome_set.union_queries(ctx, tables, K)
        torch.cuda.synchronize()
        exch.append((time.perf_counter() - t0) * 1e3)
    ms_exchange = max_over_ranks(min(exch))
    exchange = {"ms": ms_exchange, "samples": world, "rows": int(sum(len(t) for t in tables)), "n_union": int(len(union)),
                "collective": "none (1 rank)" if world == 1 else "nccl all_gather_into_tensor x2",
                "note": "outside the timed steps: once per run, after every sample has been counted"}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    # ---- from FASTQ files (host decode inside): bounded sample, 1-GPU runs only (information, SURVEY.md 8d iii) -----
    files_leg = None
    if world == 1:
        import tempfile
        from tatajuba_b200 import synth
        n_fr = 400_000  # reads (60 Mbases): writing the full sample as .gz from Python would take minutes
        rows = buf[: n_fr * (READ_LEN + 1)].reshape(n_fr, READ_LEN + 1)
        files_leg = {"sample": f"first {n_fr} reads of the workload written as paired FASTQ (R1/R2), plain and gzip -1",
                     "decoder_threads": 2}
        with tempfile.TemporaryDirectory() as d:
            for ext in (".fastq", ".fastq.gz"):
                r1, r2 = os.path.join(d, "s_R1" + ext), os.path.join(d, "s_R2" + ext)
                synth.write_fastq(r1, rows[0::2])
                synth.write_fastq(r2, rows[1::2])
                best = None
                for _ in range(3):
                    t0 = time.perf_counter()
                    hf = ctx.count_files([r1, r2], opt)
                    dt = time.perf_counter() - t0
                    best = dt if best is None else min(best, dt)
                files_leg["gz" if ext.endswith(".gz") else "plain"] = {"value": hf.n_bases / best, "unit": "bases/s",
                                                                        "ms": best * 1e3, "bases": int(hf.n_bases)}
    peak, peak_src = load_peaks()
    scan = next(s for s in stats if s["name"] == "tjb_scan_kernel")
    scan_ms = scan["ms_total"] / scan["launches"]
    scan_bytes = scan["bytes_algorithmic"] / scan["launches"]
    achieved = scan_bytes / (scan_ms * 1e-3) / 1e9
    kernel_ms = sum(s["ms_total"] for s in stats)
    line = {
        "metric": METRIC, "value": world * n_bases * args.steps / (ms * 1e-3), "unit": "bases/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8/u64", "data": "synthetic",
        "config": {"workload": "configs[1]: one sample per GPU, 5 Mbp genome, 100x, 2x150 bp PE, k=25 m=4 i=5",
                   "bases_per_gpu": n_bases, "reads_per_gpu": n_reads, "input_bytes": n_bytes,
                   "l2": "input (503 MB) larger than L2 (126 MB); no flush needed",
                   "malloc": "M_TRIM_THRESHOLD / M_MMAP_THRESHOLD = 1 GiB: freed result arrays stay in the heap", "n_raw": hc.n_raw,
                   "n_distinct": hc.n_distinct, "n_elem": hc.n_elem, "n_idx": hc.n_idx, "coverage": hc.coverage},
        "e2e": {"value": world * n_bases * args.steps / (ms_e2e * 1e-3), "unit": "bases/s",
                "ms_per_step": ms_e2e / args.steps, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "transfer": transfer},
        "gpu_launches": launches,
        "roofline": {"kernel": "tjb_scan_kernel", "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "peak_source": peak_src, "traffic": None,
                     "ms_per_launch": scan_ms, "algorithmic_bytes_per_launch": scan_bytes,
                     "share_of_kernel_time": scan["ms_total"] / kernel_ms if kernel_ms else None},
        "kernels": {s["name"]: {"launches": s["launches"], "ms_total": round(s["ms_total"], 4)} for s in stats},
        "clocks": clocks,
        "exchange": exchange,
        "from_files": files_leg,
    }
    tp = os.path.join(ROOT, "profiles", "scan_traffic.json")  # dram bytes of one scan launch from an ncu --set full capture
    if os.path.exists(tp):
        tr = json.load(open(tp))
        line["roofline"]["traffic"] = tr["dram_bytes_read"] + tr["dram_bytes_write"]
        line["roofline"]["traffic_source"] = tr.get("source")
    if not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(buf, min(CPU_SAMPLE_READS, n_reads))
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

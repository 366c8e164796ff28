#!/usr/bin/env python
"""prints the numbers of a bench.py JSON line that are looked at after every run"""
import json
import sys

d = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
kern = lambda k: " ".join(f"{n.replace('_kernel', '')}={v['us']}" for n, v in k.items())  # noqa: E731
print(f"step {d['ms_per_step']:.3f} ms  value {d['value'] / 1e9:.0f} G  two ctx {d['config'].get('two_contexts')}")
print(f"e2e {d['e2e']['ms_per_step']:.2f} ms {d['e2e']['value'] / 1e9:.1f} G  plain {d['e2e']['transfer']['plain']['ms_per_step']:.2f}  h2d {d['e2e']['h2d_bytes_per_step']}")
r = d["roofline"]
print(f"roofline frac {r['frac']:.3f} scan {r['ms_per_launch'] * 1e3:.1f} us  pipeline frac {r['pipeline']['frac']:.3f}  launches {d['gpu_launches']}  clocks {d['clocks']}")
print("  ", kern(d["kernels"]))
for e, v in d.get("errors", {}).items():
    print(f"err {e}: {v['ms_per_step']:.3f} ms (1 ctx {v['one_context_ms_per_step']:.3f})  n_dist {v['n_distinct']}  ", kern(v["kernels"]))
if "c5" in d:
    v = d["c5"]
    print(f"c5: {v['ms_per_step']:.3f} ms {v['value'] / 1e9:.0f} G  ref {v.get('reference_bases_per_s', 0) / 1e9:.3f} G  ", kern(v["kernels"]))
if "c5_sharded" in d and "error" in d["c5_sharded"]:
    print("c5 sharded: FAILED", d["c5_sharded"]["error"])
elif "c5_sharded" in d:
    v = d["c5_sharded"]
    print(f"c5 sharded over {v['ranks']} ranks: {v['ms_per_sample']:.3f} ms {v['value'] / 1e9:.0f} G  one GPU {v['one_gpu_ms_per_sample']:.3f} ms  speed-up {v['speedup_over_one_gpu']:.2f}")
if "c4" in d:
    v = d["c4"]
    print(f"c4: {v['ms']:.1f} ms {v['value'] / 1e9:.1f} G  from host {v['from_host']['ms']:.1f} ms {v['from_host']['value'] / 1e9:.1f} G  union {v['n_union']} of {v['rows']}  parity {v['parity'][:30]}")
if "exchange" in d:
    print("exchange", d["exchange"]["ms"], "ms", d["exchange"]["rows"], "rows")
if "from_files" in d:
    f = d["from_files"]
    print("files plain %.2f G gz %.2f G | reference plain %.3f gz %.3f" % (f["plain"]["value"] / 1e9, f["gz"]["value"] / 1e9,
          f.get("reference", {}).get("plain", {}).get("value", 0) / 1e9, f.get("reference", {}).get("gz", {}).get("value", 0) / 1e9))
if "cpu_baseline" in d:
    print("cpu baseline %.3f G (%s, %d core)" % (d["cpu_baseline"]["value"] / 1e9, d["cpu_baseline"]["kind"], d["cpu_baseline"]["cores"]))

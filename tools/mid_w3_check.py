import sys, time
sys.path.insert(0, '.')
import torch
from oracle import pyoracle
from tatajuba_b200 import Context, TatajubaOptions, synth
ctx = Context(0)
for genome, cov, k, m in [(1_000_000, 150, 32, 3), (1_000_000, 120, 27, 2)]:
    buf = synth.sample_buffer(genome, cov, seed=7, error_rate=0.002)
    opt = TatajubaOptions(k, m, 5, True)
    want = pyoracle.Restatement().run_buffer(buf, k, m, 5, True)
    d = torch.from_numpy(buf).cuda()
    for rep in range(2):
        hc = ctx.count_device(d.data_ptr(), d.numel(), opt)
        b = pyoracle.Boundary(hc.status, hc.n_raw, hc.elem, hc.idx_initial, hc.idx_final, hc.coverage, hc.n_skipped_nonacgt)
        pyoracle.assert_same_boundary(b, want, f"{buf.size} bytes k={k} m={m}")
    print("ok", buf.size, k, m, hc.n_raw, hc.n_elem, flush=True)

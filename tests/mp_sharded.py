"""One sample split by read chunk over the ranks (torchrun, NCCL): every rank scans its chunk, the aggregation tables
are exchanged and added, every rank finalises: bit-identical to counting the whole sample on one GPU and to the
compiled reference.  Launched by tests/test_sharded.py; prints SHARDED_OK on success."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pyoracle  # noqa: E402
from tatajuba_b200 import Context, TatajubaOptions, genome_set, synth  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = Context(local)
    for k, m, rb, err, cov in [(25, 4, True, 0.0, 60), (32, 3, True, 0.003, 80), (25, 4, False, 0.01, 40)]:
        buf = synth.sample_buffer(500_000, cov, seed=91, error_rate=err)
        opt = TatajubaOptions(k, m, 5, rb)
        rows = buf.reshape(-1, 151)
        lo, hi = rows.shape[0] * rank // world, rows.shape[0] * (rank + 1) // world
        chunk = torch.from_numpy(np.ascontiguousarray(rows[lo:hi]).reshape(-1)).cuda()
        hc = genome_set.count_sample_sharded(ctx, chunk.data_ptr(), chunk.numel(), opt, total_bytes_hint=buf.size)
        whole = torch.from_numpy(buf).cuda()
        one = ctx.count_device(whole.data_ptr(), whole.numel(), opt)
        b = lambda h: pyoracle.Boundary(h.status, h.n_raw, h.elem, h.idx_initial, h.idx_final, h.coverage)  # noqa: E731
        pyoracle.assert_same_boundary(b(hc), b(one), f"rank {rank}: sharded vs one GPU, k={k}")
        if pyoracle.Reference.available() and rank == 0:
            pyoracle.assert_same_boundary(b(hc), pyoracle.Reference().run_buffer(buf, k, m, 5, rb), f"sharded vs reference, k={k}")
    dist.barrier()
    if rank == 0:
        print("SHARDED_OK")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

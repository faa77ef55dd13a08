"""A/B of the transposed local-to-local driver on N ranks (torchrun): slab exchange on the peer regions
(hmb200_peer_exchange_sum) against NCCL point-to-point calls (exchange_sum_slabs_), same operator, same inputs.
usage: python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 scripts/l2l_transposed_ab.py [N] [steps]"""
import json
import os
import sys
import types

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from htool_python_b200 import B200DistributedOperator  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 300000
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 30
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl")
    args = types.SimpleNamespace(n=n, leaf=64, eps=1e-3, eta=10.0, kernel="laplace", wavenumber=0.0)
    tree, H, flat, _ = bench.assemble(args, world, rank, lambda m: None)
    ops = {"nccl_p2p": B200DistributedOperator(flat, tree.partition, fused_allgather=False, build_transpose=1, no_matmat=1),
           "peer_regions": B200DistributedOperator(flat, tree.partition, build_transpose=1, no_matmat=1)}
    assert ops["peer_regions"].fused and not ops["nccl_p2p"].fused
    s = ops["nccl_p2p"].sizes[rank]
    x = torch.from_numpy(np.random.default_rng(rank).random(s)).cuda()
    out, res = {}, {}
    for name, op in ops.items():
        y = torch.zeros(s, dtype=torch.float64, device="cuda")
        for _ in range(5):
            op.matvec_local_to_local_(x, y, "T", 1.0, 0.0)
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            op.matvec_local_to_local_(x, y, "T", 1.0, 0.0)
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / steps], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out[name] = round(float(t.item()), 4)
        res[name] = y.clone()
    same = torch.tensor([int(torch.equal(res["nccl_p2p"], res["peer_regions"]))], device="cuda")
    dist.all_reduce(same, op=dist.ReduceOp.MIN)
    status = ops["peer_regions"].peer_status()
    if rank == 0:
        print(json.dumps({"what": "matvec_local_to_local_('T') ms per call, max over ranks", "n": n, "ranks": world, "steps": steps,
                          "ms": out, "bit_identical": bool(same.item()), "peer_status": status}), flush=True)
    for op in ops.values():
        op.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

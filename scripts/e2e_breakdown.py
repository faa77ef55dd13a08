"""Where the time of B200DistributedOperator.matvec(x, out='shared') goes (run under torchrun on >= 2 GPUs):
whole call / library call only (one cached CUDA graph + stream sync) / device time of that graph / host flag barrier."""
import os, sys, time, types
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch, torch.distributed as dist
import bench
from htool_python_b200 import B200DistributedOperator, _lib

world, rank, lr = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
os.environ["OMP_NUM_THREADS"] = str(max(1, (os.cpu_count() or 8) // world))
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
args = types.SimpleNamespace(n=int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000, eps=1e-3, eta=10.0, leaf=64, kernel="laplace", wavenumber=0.0)
tree, H, flat, _ = bench.assemble(args, world, rank, lambda m: None)
op = B200DistributedOperator(flat, tree.partition, device=lr)
n = args.n
x = torch.rand(n, dtype=torch.float64).pin_memory().numpy()
for _ in range(5):
    op.matvec(x, out="shared")
def timeit(fn, reps=100):
    dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    t = torch.tensor([(time.perf_counter() - t0) / reps * 1e3], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)
full = timeit(lambda: op.matvec(x, out="shared"))
c0, c1 = rank * n // world, (rank + 1) * n // world
src = torch.from_numpy(x)
fn = _lib.lib().hmb200_matvec_user_chunked_d
def lib_only():
    k = op._shm_calls; op._shm_calls += 1
    res = op._shm_y[k & 1]
    _lib.check(fn(op.engine._h, src.data_ptr() + 8 * c0, c0, c1 - c0, res.ctypes.data + 8 * c0, c0, c1 - c0))
libt = timeit(lib_only)
xc = torch.rand(n, dtype=torch.float64, device="cuda")
dev = timeit(lambda: op.matvec_cluster_fused_(xc))
if rank == 0:
    print(f"e2e-breakdown: P={world} N={n}: matvec(x, out='shared') {full:.4f} ms | library call only (graph + sync) {libt:.4f} ms | device-resident fused product {dev:.4f} ms")
op.close()
dist.destroy_process_group()

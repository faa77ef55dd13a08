#!/usr/bin/env python
"""bench.py -- H-matvecs/sec of the H-matrix product engine (BASELINE.json metric), 1..8 B200 vs the host CPU.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--config 1|2|3|4] [--n POINTS] ...
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

`--config c` selects BASELINE.json configs[c] (default 1 = the metric's own: Laplace N=1M, double, single RHS):
    1  Laplace 1/r, N=1M, single-RHS matvec            2  Helmholtz exp(ikr)/r, complex<double>, N=500k, matvec
    3  Laplace N=1M, 64 right-hand sides (DMMA path)    4  Laplace N=2M + the DDM solver's unpreconditioned GMRES

A "step" is one product y = A x (Y = A X with --rhs / config 3) of the whole operator:
  value  : device-resident (x, y replicated in cluster numbering in HBM; at N > 1 every rank multiplies its block-row slab and
           the all-gather is fused into the kernel epilogue), CUDA events on the launching stream, barrier + synchronize on
           both sides, max over ranks;
  e2e    : the same product through the reference-facing API (`B200HMatrix.matvec` / `B200DistributedOperator.matvec`, `@` for
           blocks) with page-locked HOST vectors in user numbering: H2D, both permutations, product, D2H inside the region;
  sustained: the same step repeated for >= 400 steps (about a second) with clocks sampled: the regime a GMRES solve runs in;
  roofline: the dominant kernel's ALGORITHMIC bytes (flops for the multi-RHS path) / its CUDA-event time against
           MEASURED_PEAKS.json hbm_gbs (against the FP64 tensor rate measured in this run, hmb200_fp64_rate);
  cpu_baseline / parity: the reference's own OpenMP product (oracle/_ref, compiled from the reference's sources) on the same
           htool::HMatrix object(s) on this box's host cores.  At N > 1 EVERY rank checks its slab of the replicated result
           against the reference product of its own slab, the ranks compare a 64-bit hash of the full vector
           (`replicas_bit_identical`), and the CPU baseline is the emulated-ranks figure 1 / max_p t(slab_p).
The H-matrix is assembled by the REFERENCE's builder (HMatrixTreeBuilder::build, OpenMP, native generator) on the host
before the timed region; strong scaling: the operator is split into the reference's block-row slabs.
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

UNIT = "matvecs/s"
WORKLOADS = {"laplace": "Laplace 1/(4 pi r), double,", "helmholtz": "Helmholtz exp(ikr)/(4 pi r), complex<double>,",
             "laplace_reg": "regularised Laplace 1/(1e-5 + 4 pi r) (the reference's GeneratorTestDoubleSymmetric), double,",
             "invdist_reg": "1/(1e-1 + |x-y|) (the generator of the reference's example/use_hmatrix.py, example/define_generators.py:15-17), double,"}
GEOMETRIES = {"sphere": "points on the unit sphere surface (seed 0)", "cylinder": "points on a cylinder of radius 1 and height 4 (theta, z uniform, seed 0)"}
CONFIGS = {0: dict(kernel="invdist_reg", n=20_000, rhs=1, geometry="cylinder", symmetry="S"), 1: dict(kernel="laplace", n=1_000_000, rhs=1), 2: dict(kernel="helmholtz", n=500_000, rhs=1),
           3: dict(kernel="laplace", n=1_000_000, rhs=64), 4: dict(kernel="laplace", n=2_000_000, rhs=1, gmres=True)}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", type=int, default=1, choices=[0, 1, 2, 3, 4], help="BASELINE.json configs[c]; sets --kernel/--n/--rhs/--gmres "
                    "(0: the reference's CPU-runnable use_hmatrix.py case, N=20k on a cylinder, symmetric 'S','L'; sweep --leaf 10/64/128)")
    ap.add_argument("--symmetry", default=None, choices=["N", "S"], help="symmetry given to the reference's HMatrixTreeBuilder ('S' stores the lower "
                    "triangle, UPLO 'L'); default: what the config's reference example uses")
    ap.add_argument("--n", type=int, default=None)
    ap.add_argument("--eps", type=float, default=1e-3)
    ap.add_argument("--eta", type=float, default=10.0)
    ap.add_argument("--leaf", type=int, default=64)
    ap.add_argument("--kernel", default=None, choices=["laplace", "helmholtz", "laplace_reg", "invdist_reg"])
    ap.add_argument("--wavenumber", type=float, default=0.0, help="helmholtz: k (default: 10 points per wavelength)")
    ap.add_argument("--rhs", type=int, default=None, help="right-hand sides per step (64 = BASELINE.json configs[3], multi-RHS tensor-core path)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--also-rhs", type=int, default=64, help="single-RHS runs on the Laplace / Helmholtz configs: also time and check ONE more workload on the "
                    "operator already assembled -- blocks of this many right-hand sides (BASELINE.json configs[3]: device value, e2e through matmat, "
                    "per-rank parity against the reference) -- reported under 'also' (0: skip)")
    ap.add_argument("--sustained-steps", type=int, default=400, help="steps of the sustained region reported next to the burst value (0: skip)")
    ap.add_argument("--gmres", action="store_true", help="also solve A x = b with the device-resident GMRES (BASELINE.json configs[4]; adds a 'gmres' object)")
    ap.add_argument("--gmres-block", type=int, default=0, help="with --gmres: also solve a block of this many right-hand sides in lockstep "
                    "(HPDDM's GMRES with mu > 1 on the multi-RHS product) and column by column, tol 1e-3")
    ap.add_argument("--gmres-args", default="-hpddm_krylov_method gmres -hpddm_schwarz_method none -hpddm_variant right -hpddm_tol 1e-6 -hpddm_gmres_restart 40 -hpddm_max_it 200")
    ap.add_argument("--no-gmres-regularised", action="store_true", help="with --gmres: skip the second solve on the reference's regularised test kernel")
    ap.add_argument("--transposed", action="store_true", help="also time y = A^T x (and A^H x for complex) on the same operator: adds a 'transposed' object "
                    "(packs the transposed streams too: twice the device memory)")
    ap.add_argument("--transposed-exchange", choices=["nccl", "peer"], default="nccl", help="with --transposed at N > 1: NCCL all-reduce of the full-length "
                    "contributions (default, measured faster) or the rank-ordered reduce-scatter + all-gather on the peer regions")
    ap.add_argument("--engine-opts", default="", help="k=v,k=v overrides of hmb200_options (tuning runs)")
    a = ap.parse_args()
    preset = CONFIGS[a.config]
    a.geometry = preset.get("geometry", "sphere")
    if a.symmetry is None:
        a.symmetry = preset.get("symmetry", "N")
    if a.symmetry != "N" and a.gpus > 1:
        raise SystemExit("bench.py: --symmetry S is a single-GPU case (the block-row slabs of the distributed product are assembled 'N')")
    if a.kernel is None:
        a.kernel = preset["kernel"]
    if a.n is None:
        a.n = int(os.environ.get("HMB200_BENCH_N", preset["n"]))
    if a.rhs is None:
        a.rhs = preset["rhs"]
    if preset.get("gmres"):
        a.gmres = True
    return a


def config_index(args, mu):
    if args.kernel == "helmholtz":
        return 2
    if args.kernel == "invdist_reg":
        return 0
    if mu > 1:
        return 3
    return 4 if (args.gmres and args.n >= 2_000_000) else 1


def metric_name(args, mu):
    kind = {"laplace": "Laplace 1/r, double", "helmholtz": "Helmholtz exp(ikr)/r, complex<double>", "laplace_reg": "regularised Laplace, double", "invdist_reg": "1/(0.1+|x-y|) on a cylinder, double"}[args.kernel]
    n = f"{args.n // 1_000_000}M" if args.n % 1_000_000 == 0 else (f"{args.n // 1000}k" if args.n % 1000 == 0 else str(args.n))
    return f"H-matvecs/sec (N={n} {kind}, {'single RHS' if mu == 1 else str(mu) + ' RHS per product'})"


def workload_name(args, n, mu):
    return (f"{WORKLOADS[args.kernel]} H-matrix{' (symmetric, lower triangle stored)' if getattr(args, 'symmetry', 'N') == 'S' else ''}, N={n} {GEOMETRIES[getattr(args, 'geometry', 'sphere')]}, eps={args.eps}, eta={args.eta}, "
            f"leaf {args.leaf}, {'single-RHS matvec' if mu == 1 else str(mu) + '-RHS matmat (FP64 tensor-core path)'} (BASELINE.json configs[{config_index(args, mu)}])")


def config_of(args, mu):
    """the workload description -- IDENTICAL on both arms (engine-only facts go under `impl_config`)"""
    if args.kernel == "invdist_reg" and args.n <= 100_000:  # configs[0]: ~70 MB of factors would sit in the 126 MB L2 from step to step
        return {"workload": workload_name(args, args.n, mu), "rhs_per_step": mu,
                "l2": "operator smaller than L2: a 512 MB buffer is overwritten before every timed step (outside the per-step events), steps timed one by one; the e2e loop is not flushed"}
    return {"workload": workload_name(args, args.n, mu), "rhs_per_step": mu,
            "l2": "inputs larger than L2 (factor stream >> 126 MB, read once per step)"}


def sphere_points(n, seed=0):
    """N points uniformly on the unit sphere surface (normalised Gaussian, seed 0) -- SURVEY.md 8(d) C2."""
    g = np.random.default_rng(seed).standard_normal((3, n))
    g /= np.linalg.norm(g, axis=0)
    return np.asfortranarray(g)


def cylinder_points(n, seed=0):
    """N points on a cylinder: x = cos(theta), y = sin(theta), z ~ U(0, 4) (the shape of the reference's
    example/create_geometry.py without its clustering of the points)."""
    rng = np.random.default_rng(seed)
    theta, z = rng.random(n) * 2 * np.pi, rng.random(n) * 4.0
    return np.asfortranarray(np.stack([np.cos(theta), np.sin(theta), z]))


def host_threads(world):
    try:
        cores = len(os.sched_getaffinity(0))
    except Exception:
        cores = os.cpu_count() or 1
    return max(1, cores // max(world, 1)), cores


def assemble(args, world, rank, log, kernel=None):
    """Cluster tree + this rank's block-row slab, assembled by the reference's own builders (bridge)."""
    from htool_python_b200 import FlatHMatrix, bridge

    t0 = time.time()
    pts = cylinder_points(args.n) if getattr(args, "geometry", "sphere") == "cylinder" else sphere_points(args.n)
    sym = getattr(args, "symmetry", "N")
    tree = bridge.create_cluster_tree(pts, 2, world, args.leaf)
    t1 = time.time()
    slab = rank if world > 1 else -1
    # keep every admissible block below 46340 x 46340: the reference's ACA stop test `q*(n1+n2) > n1*n2` is int arithmetic
    # (lib/htool/include/htool/hmatrix/lrmat/sympartialACA.hpp:102); above that it overflows and the block is built DENSE
    depth = 0
    while args.n / 2 ** depth > 46340:
        depth += 1
    depth += 1 if depth else 0  # the builder's depth test is off by one against size = n / 2^depth (measured)
    kernel = kernel or getattr(args, "kernel", "laplace")
    k = getattr(args, "wavenumber", 0.0) or 2 * np.pi * np.sqrt(args.n / (4 * np.pi)) / 10.0  # ~10 points per wavelength
    H = bridge.build_hmatrix(kernel, tree, epsilon=args.eps, eta=args.eta, symmetry=sym, UPLO="L" if sym != "N" else "N", target_partition_number=slab, wavenumber=k,
                             partition_number_for_symmetry=slab, minimal_target_depth=depth, minimal_source_depth=depth)
    t2 = time.time()
    flat = FlatHMatrix.from_htool(H)
    big = (flat.rank < 0) & (flat.nrows.astype(np.int64) * flat.ncols > 1 << 22)
    if big.any():
        raise RuntimeError(f"the reference built {int(big.sum())} dense blocks above 4M coefficients (ACA int overflow?)")
    log(f"rank {rank}: tree {t1 - t0:.1f}s, assembly {t2 - t1:.1f}s, leaves {flat.n_leaves}, W {flat.stored_coefficients:.4e}")
    return tree, H, flat, dict(tree_s=round(t1 - t0, 2), assembly_s=round(t2 - t1, 2), minimal_block_depth=depth)


class ClockSampler(threading.Thread):
    """samples SM clock + throttle reasons of one GPU through NVML while a timed region runs; `mark()` opens a new window"""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.sm, self.reasons, self.sm_max = index, False, [], [], None
        self.t = []
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.sm_max = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap", nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown", nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonHwPowerBrakeSlowdown: "hw_power_brake", nv.nvmlClocksThrottleReasonSyncBoost: "sync_boost"}
        while not self.stop_flag:
            try:
                sm = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.sm.append(sm)
                self.reasons.append({name for bit, name in names.items() if r & bit})
                self.t.append(time.perf_counter())
            except Exception:
                pass
            time.sleep(0.004)

    def window(self, t0, t1):
        """clocks object of the samples taken in [t0, t1] (all samples when the region was shorter than the sampling period)"""
        if self.nv is None or not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": self.sm_max, "reasons": []}
        idx = [i for i, t in enumerate(self.t) if t0 <= t <= t1] or list(range(len(self.sm)))
        reasons = set()
        for i in idx:
            reasons |= self.reasons[i]
        return {"sm_mhz": float(np.median([self.sm[i] for i in idx])), "sm_max_mhz": self.sm_max, "reasons": sorted(reasons), "samples": len(idx)}

    def stop(self):
        self.stop_flag = True


def w_split(flat):
    """algorithmic coefficients streamed by phase 1 (V factors) and phase 2 (U factors + dense leaves); DESIGN.md"""
    r = np.maximum(flat.rank, 0).astype(np.float64)
    lr = flat.rank >= 0
    w1 = float(np.sum(r[lr] * flat.ncols[lr]))
    w2 = float(np.sum(r[lr] * flat.nrows[lr]) + np.sum(flat.nrows[~lr].astype(np.float64) * flat.ncols[~lr]))
    return w1, w2


def fp64_rates(device):
    """measured DMMA / DFMA issue rates of this GPU (TFLOP/s), hmb200_fp64_rate"""
    import ctypes

    from htool_python_b200 import _lib

    a, b = ctypes.c_double(0), ctypes.c_double(0)
    _lib.check(_lib.lib().hmb200_fp64_rate(int(device), ctypes.byref(a), ctypes.byref(b)))
    return round(a.value, 2), round(b.value, 2)


def run_b200(args):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    thr, cores = host_threads(world)
    if os.environ.get("OMP_NUM_THREADS") in (None, "", "1"):  # torchrun pins it to 1; the host-side assembly wants its share of cores
        os.environ["OMP_NUM_THREADS"] = str(thr)
    import torch
    import torch.distributed as dist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the engine has no CPU fallback")
    bound = []
    if world > 1 and not os.environ.get("HMB200_BENCH_NO_BIND"):  # like `mpirun --bind-to` for the reference's ranks
        from htool_python_b200.distributed import bind_to_gpu_cpus

        bound = bind_to_gpu_cpus(local_rank, int(os.environ.get("LOCAL_WORLD_SIZE", world)))
        if bound:
            thr = max(1, min(thr, len(bound)))
            os.environ["OMP_NUM_THREADS"] = str(thr)
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def log(msg):
        print(f"[bench] {msg}", file=sys.stderr, flush=True)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        t = torch.tensor([float(v)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(v):
        t = torch.tensor([float(v)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    from htool_python_b200 import B200DistributedOperator, B200HMatrix

    tree, H, flat, build_t = assemble(args, world, rank, log)
    opts = {k: int(v) for k, v in (kv.split("=") for kv in args.engine_opts.split(",") if kv)}
    if args.transposed:
        opts.setdefault("build_transpose", 1)
        if args.rhs <= 1:
            opts.setdefault("no_matmat", 1)
    t0 = time.time()
    if world > 1:
        op = B200DistributedOperator(flat, tree.partition, device=local_rank, fused_allgather=False if os.environ.get("HMB200_NO_FUSED") else None, **opts)
        engine = op.engine
    else:
        op = B200HMatrix(flat, device=local_rank, **opts)
        engine = op.engine
    info = engine.info
    log(f"rank {rank}: engine created in {time.time() - t0:.1f}s (pack {info['pack_seconds']:.1f}s upload {info['upload_seconds']:.1f}s) "
        f"stream1 {info['stream1_bytes'] / 1e9:.2f} GB stream2 {info['stream2_bytes'] / 1e9:.2f} GB bins {info['n_bins']}")

    n = args.n
    rng = np.random.default_rng(1)
    x_user = rng.random(n)  # U(0,1), seed 1
    if flat.dtype == "z":
        x_user = x_user + 1j * rng.random(n)
    tdt = torch.complex128 if flat.dtype == "z" else torch.float64
    perm_s, perm_t = flat.source_permutation, flat.target_permutation
    mu = max(1, args.rhs)
    if mu > 1:  # X: n x mu, Fortran-ordered for the user API, row-major in cluster numbering for the operator interface
        x_user = np.asfortranarray(np.stack([np.roll(x_user, 7 * j) * (1.0 + 0.01 * j) for j in range(mu)], axis=1))
    x_cluster_host = np.ascontiguousarray(x_user[perm_s])
    x_cluster = torch.from_numpy(x_cluster_host).to(dev)
    y_cluster = torch.zeros((n, mu) if mu > 1 else n, dtype=tdt, device=dev)
    fused = world > 1 and op.fused
    last = {}

    if mu > 1 and world > 1:
        def step():
            last["y"] = op.matmat_cluster_(x_cluster.view(-1), y_cluster, mu)
    elif mu > 1:
        def step():
            engine.add_matrix_product_row_major("N", 1.0, x_cluster, 0.0, y_cluster, mu)
            last["y"] = y_cluster
    elif fused:
        def step():  # slab product whose phase-2 epilogue stores the rows into every rank's replicated vector (no NCCL call)
            last["y"] = op.matvec_cluster_fused_(x_cluster)
    elif world > 1:
        def step():
            last["y"] = op.matvec_cluster_(x_cluster, y_cluster)
    else:
        def step():
            engine.add_vector_product("N", 1.0, x_cluster, 0.0, y_cluster)
            last["y"] = y_cluster
    if mu > 1 and world > 1:  # page-locked Fortran-ordered block; every rank moves 1/P of the rows of X and Y over PCIe
        x_blk = torch.empty(n * mu, dtype=tdt).pin_memory().numpy().reshape((n, mu), order="F")
        x_blk[...] = x_user
        apply_user = lambda: op.matmat(x_blk, out="shared")  # noqa: E731
    elif mu > 1:  # page-locked Fortran-ordered blocks: the library DMAs straight from / into them
        x_blk = torch.empty(n * mu, dtype=tdt).pin_memory().numpy().reshape((n, mu), order="F")
        x_blk[...] = x_user
        y_blk = torch.empty(n * mu, dtype=tdt).pin_memory().numpy().reshape((n, mu), order="F")
        apply_user = lambda: op.matmat(x_blk, out=y_blk)  # noqa: E731
    else:  # page-locked host buffers (torch pin_memory): the library DMAs straight from / into them
        x_pin = torch.from_numpy(np.ascontiguousarray(x_user)).pin_memory().numpy()
        if world > 1:
            apply_user = lambda: op.matvec(x_pin, out="shared")  # noqa: E731
        else:
            y_pin = torch.zeros(n, dtype=tdt).pin_memory().numpy()
            apply_user = lambda: op.matvec(x_pin, out=y_pin)  # noqa: E731

    flush_buf = torch.empty(512 << 20, dtype=torch.uint8, device=dev) if config_of(args, mu)["l2"].startswith("operator smaller") else None

    def timed_region(steps):
        """`steps` products between two events on the launching stream; returns (ms per step as the max over ranks, host window)"""
        barrier()
        if flush_buf is not None:  # small operator: evict it from L2 before every step, one event pair per step
            t_a = time.perf_counter()
            pairs = []
            for _ in range(steps):
                flush_buf.fill_(1)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                step()
                e1.record()
                pairs.append((e0, e1))
            barrier()
            return max_over_ranks(sum(a.elapsed_time(b) for a, b in pairs)) / steps, (t_a, time.perf_counter())
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t_a = time.perf_counter()
        ev0.record()
        for _ in range(steps):
            step()
        ev1.record()
        barrier()
        return max_over_ranks(ev0.elapsed_time(ev1)) / steps, (t_a, time.perf_counter())

    # ---- device-resident timed region ------------------------------------------------------------------------
    # NVML is initialised and the sampling thread started BEFORE the region (nvmlInit takes a rank-dependent time: inside the
    # barrier-to-event window it skews the ranks and the skew lands in the first step's wait); only rank 0 samples its GPU
    sampler = ClockSampler(local_rank)
    if rank == 0 and not os.environ.get("HMB200_BENCH_NO_SAMPLER"):
        sampler.start()
    for _ in range(args.warmup):
        step()
    barrier()
    launches0 = engine.launch_count
    ms_step, win = timed_region(args.steps)
    clocks = sampler.window(*win)
    launches = engine.launch_count - launches0
    # per-kernel durations: the SAME steps once more with CUDA events recorded around every kernel on the launching stream
    # (the timed region above replays CUDA graphs, which carry no per-kernel events)
    engine.enable_timing(True)
    for _ in range(args.steps):
        if flush_buf is not None:
            flush_buf.fill_(1)
        step()
    barrier()
    kms = engine.kernel_ms_avg()
    engine.enable_timing(False)
    if fused and op.peer_status() != 0:
        raise SystemExit(f"bench.py: rank {rank}: a peer did not publish its epoch (status {op.peer_status()})")
    y_dev_host = last["y"].detach().cpu().numpy().copy()  # replicated result of the device-resident path (cluster numbering)

    # ---- sustained region: the same step for about a second or more, clocks sampled --------------------------------------
    sustained = None
    if args.sustained_steps > 0:
        s_steps = max(args.sustained_steps, args.steps)
        ms_sus, win_s = timed_region(s_steps)
        sustained = {"steps": s_steps, "ms_per_step": round(ms_sus, 4), "value": round(mu * 1e3 / ms_sus, 3), "unit": UNIT,
                     "region_s": round(ms_sus * s_steps * 1e-3, 3), "clocks": sampler.window(*win_s)}

    # ---- end to end through the public API with host buffers ----------------------------------------------------
    for _ in range(max(4, args.warmup)):  # the chunked host product captures its graph at the second sight of (x, result buffer): 2 buffers x 2
        y_user = apply_user()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        y_user = apply_user()
    torch.cuda.synchronize()
    e2e_step = max_over_ranks(time.perf_counter() - t0) / args.steps
    y_user = np.array(y_user, copy=True)
    sampler.stop()
    shared_result = world > 1 and ((mu == 1 and getattr(op, "_shm", None) is not None) or (mu > 1 and getattr(op, "_shmb", None) is not None))

    # ---- transposed / conjugate-transposed products (trans != 'N' of the operator interface) -----------------------------------
    tr_obj = None
    if args.transposed and mu == 1:
        if world > 1:
            op.transposed_exchange = args.transposed_exchange
        tr_obj = run_transposed(args, op, engine, H, flat, x_cluster_host, world, rank, dev, thr, barrier, max_over_ranks)

    # ---- GMRES with the Krylov basis resident on the device (BASELINE.json configs[4]) --------------------------------
    gm = None
    if args.gmres and mu == 1:
        gm = run_gmres(args, op, flat, tree, H, world, rank, dev, thr, log, barrier)

    # ---- roofline of the dominant kernel (max over ranks of its time; this rank's bytes) ---------------------------
    w1, w2 = w_split(flat)
    es = flat.elem_bytes
    alg = {"phase1": es * (w1 + flat.n_cols), "phase2": es * (w2 + flat.n_rows)}
    dom = "phase2" if kms["phase2"] >= kms["phase1"] else "phase1"
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6553.6))
    dmma_tf, dfma_tf = fp64_rates(local_rank)
    achieved = alg[dom] / (kms[dom] * 1e-3) / 1e9 if kms[dom] > 0 else 0.0
    w_all = sum_over_ranks(flat.stored_coefficients)
    agg_bytes = es * (w_all + mu * (flat.n_cols * world + n))  # every rank reads the whole x; y is written once
    roofline = {"bound": "hbm", "kernel": f"hmatrix_stream_kernel<{'double' if es == 8 else 'double2'},{2 if dom == 'phase2' else 1}>", "achieved": round(achieved, 1),
                "peak": peak, "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback (B200_PROFILING.md)", "unit": "GB/s",
                "frac": round(achieved / peak, 4), "traffic": None, "algorithmic_bytes_per_launch": alg[dom],
                "kernel_ms": {k: round(v, 4) for k, v in kms.items()},
                "whole_product_GBps": round(flat.algorithmic_bytes(1) / (kms["total"] * 1e-3) / 1e9, 1) if kms["total"] > 0 else None}
    if mu > 1:  # multi-RHS: bounded by the FP64 tensor (DMMA) pipe; kernel_ms are per pass of 64 (32 complex) columns
        passes = -(-mu // (64 if es == 8 else 32))
        flops = (2 if es == 8 else 8) * mu * flat.stored_coefficients
        tf = flops / (kms["total"] * passes * 1e-3) / 1e12 if kms["total"] > 0 else 0.0
        roofline = {"bound": "tensor", "kernel": "hmatrix_mm_kernel<T,phase 1|2> panel GEMMs (mma.sync.m8n8k4.f64, SASS DMMA)", "achieved": round(tf, 2), "peak": dmma_tf,
                    "peak_source": "FP64 tensor-core (DMMA) issue rate measured in this run on this GPU (hmb200_fp64_rate; MEASURED_PEAKS.json has no FP64 figure)",
                    "unit": "TFLOP/s", "frac": round(tf / dmma_tf, 4) if dmma_tf else None, "traffic": None, "algorithmic_flops_per_step": flops, "passes": passes,
                    "kernel_ms_per_pass": {k: round(v, 4) for k, v in kms.items()}}
    try:  # traffic of the dominant kernel from the committed ncu capture of this exact workload (null otherwise)
        tr = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json"))).get(f"{args.kernel}_{n}_{world}" + (f"_rhs{mu}" if mu > 1 else ""))
        if tr and args.eps == 1e-3 and args.eta == 10.0 and args.leaf == 64:
            roofline["traffic"] = tr["phase2" if (mu > 1 or dom == "phase2") else "phase1"]
            roofline["traffic_source"] = tr["source"]
    except Exception:
        pass

    # ---- parity + CPU baseline: every rank against the reference's product of ITS OWN slab ---------------------------------
    cpu = parity = parity_e2e = None
    replicas_identical = None
    if not args.no_cpu_baseline:
        o, s = (op.offs[rank], op.sizes[rank]) if world > 1 else (0, n)
        t_slab, y_ref_slab = cpu_reference_slab(H, flat, x_cluster_host, mu, thr, budget_s=15.0, log=log, sync=barrier)
        nrm = float(np.linalg.norm(y_ref_slab))
        parity = max_over_ranks(np.linalg.norm(y_dev_host.reshape(n, -1)[o:o + s] - y_ref_slab.reshape(s, -1)) / nrm)
        y_e2e_cluster = y_user[perm_t[o:o + s]]  # root_cluster_to_global undone on this rank's slab
        parity_e2e = max_over_ranks(np.linalg.norm(y_e2e_cluster.reshape(s, -1) - y_ref_slab.reshape(s, -1)) / nrm)
        t_max = max_over_ranks(t_slab)
        cpu = {"value": round(mu / t_max, 4), "unit": UNIT, "cores": thr * world, "kind": "reference",
               "ms_per_product": round(t_max * 1e3, 2),
               "sample": (f"full products ({mu} RHS each) of the same N={n} operator (best of up to 10 within 15 s), reference OpenMP path "
                          "openmp_internal_add_hmatrix_{vector,matrix}_product compiled from the reference sources (oracle/_ref)" +
                          ("" if world == 1 else f"; EMULATED RANKS: each of the {world} processes multiplies its own block-row slab concurrently with "
                           f"{thr} threads, value = 1 / max_p t(slab_p), no MPI_Allgatherv included (no MPI on the box)"))}
        log(f"parity vs reference CPU product (max over ranks): device path {parity:.3e}, e2e path {parity_e2e:.3e}")
    if world > 1:  # 64-bit hash of the full replicated vector of every rank: all replicas must hold the same bits
        hs = [None] * world
        dist.all_gather_object(hs, (hashlib.blake2b(y_dev_host.tobytes(), digest_size=8).hexdigest(), hashlib.blake2b(y_user.tobytes(), digest_size=8).hexdigest()))
        replicas_identical = len(set(hs)) == 1

    # ---- one more workload on the same operator: blocks of right-hand sides (BASELINE.json configs[3]) ---------------------------------
    also = None
    if mu == 1 and args.also_rhs > 1 and not args.no_cpu_baseline and args.kernel in ("laplace", "helmholtz") and info["mm_stream_bytes"] > 0:
        also = {f"rhs{args.also_rhs}": run_also_rhs(args, args.also_rhs, op, engine, H, flat, x_user, world, rank, dev, thr, dmma_tf, w_all, log, barrier, max_over_ranks)}

    if rank == 0:
        h2d = es * n * mu * (1 if (world == 1 or (mu == 1 and fused) or shared_result) else world)
        d2h = es * n * mu * (1 if (world == 1 or shared_result) else world)
        line = {
            "metric": metric_name(args, mu), "value": round(mu * 1e3 / ms_step, 3), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": round(ms_step, 4), "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64" if flat.dtype == "d" else "c128",
            "data": "synthetic",
            "config": config_of(args, mu),
            "impl_config": {"partition": f"block-row slabs x{world}",
                            "allgather": ("none (1 GPU)" if world == 1 else "fused into the phase-2 epilogue: peer stores over NVLink + epoch flags (CUDA IPC)" if fused else "NCCL all-gather"),
                            "launch": "CUDA graph replay of the product (captured by libhmb200.so)" if mu == 1 else "eager",
                            "stored_coefficients_rank0": flat.stored_coefficients, "stored_coefficients_all_ranks": w_all, "leaves_rank0": flat.n_leaves,
                            "assembly": "reference HMatrixTreeBuilder (OpenMP, native generator), outside the timed region", **build_t,
                            "cpu_binding_rank0": (f"{len(bound)} cores next to the GPU (NVML ideal CPU set, split between the ranks)" if bound else "none"),
                            "e2e_api": ("B200HMatrix.matvec(x, out) on page-locked host numpy buffers, user numbering" if world == 1 else
                                        "B200DistributedOperator.matvec(x, out='shared'): page-locked host x (every rank uploads its 1/P chunk, NVLink exchange); the "
                                        "replicated result lands in ONE node-shared page-locked host vector (every rank downloads its 1/P chunk, host flags)")
                            if mu == 1 else ("B200HMatrix.matmat(X, out) on page-locked Fortran-ordered host blocks, user numbering" if world == 1 else
                                             "B200DistributedOperator.matmat(X, out='shared'): page-locked Fortran-ordered X (every rank uploads 1/P of its rows), "
                                             "result in ONE node-shared page-locked block (every rank downloads 1/P of the rows)")},
            "clocks": clocks,
            "e2e": {"value": round(mu / e2e_step, 3), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": round(e2e_step * 1e3, 4)},
            "gpu_launches": int(launches),
            "roofline": roofline,
            # the whole job against the bytes ALL ranks stream (the slabs hold more coefficients than the one-piece operator)
            "aggregate_algorithmic_GBps": round(agg_bytes / (ms_step * 1e-3) / 1e9, 1),
            "fp64_dmma_tflops": dmma_tf, "fp64_dfma_tflops": dfma_tf,
            "sustained": sustained,
            "cpu_baseline": cpu,
            "parity_rel_err_vs_reference": parity,
            "parity_rel_err_vs_reference_e2e": parity_e2e,
            "replicas_bit_identical": replicas_identical,
        }
        if gm is not None:
            line["gmres"] = gm
        if tr_obj is not None:
            line["transposed"] = tr_obj
        if also is not None:
            line["also"] = also
        print(json.dumps(line), flush=True)
    barrier()
    op.close()
    if world > 1:
        dist.destroy_process_group()


def run_transposed(args, op, engine, H, flat, x_cluster_host, world, rank, dev, threads, barrier, max_over_ranks):
    """y = op(A) x for op = transpose (and conjugate transpose for complex operators), device-resident, cluster numbering: one
    engine call at N = 1; at N > 1 the `trans != 'N'` branch of the global-to-global driver (every rank applies op(A_p) to ITS slab
    of x, full-length results summed over the ranks: add_distributed_operator_vector_product_global_to_global.hpp:63-84).  Parity:
    every rank's contribution against the reference's openmp_internal_add_hmatrix_vector_product(trans) of its own slab."""
    import torch

    import helpers_ref

    n = args.n
    tdt = torch.complex128 if flat.dtype == "z" else torch.float64
    o, s = (op.offs[rank], op.sizes[rank]) if world > 1 else (0, n)
    x_dev = torch.from_numpy(x_cluster_host).to(dev)
    out = {}
    for tr in ["T"] + (["C"] if flat.dtype == "z" else []):
        y = torch.zeros(n, dtype=tdt, device=dev)
        if world > 1:
            step = lambda: op.matvec_cluster_transposed_(x_dev, y, tr)  # noqa: E731
        else:
            step = lambda: engine.add_vector_product(tr, 1.0, x_dev, 0.0, y)  # noqa: E731
        for _ in range(args.warmup):
            step()
        barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        for _ in range(args.steps):
            step()
        ev1.record()
        barrier()
        ms = max_over_ranks(ev0.elapsed_time(ev1)) / args.steps
        rec = {"value": round(1e3 / ms, 3), "unit": UNIT, "ms_per_step": round(ms, 4)}
        if not args.no_cpu_baseline:  # this rank's contribution op(A_p) x_p against the reference on the same slab
            part = torch.zeros(n, dtype=tdt, device=dev)
            engine.add_vector_product(tr, 1.0, x_dev[o:o + s].contiguous(), 0.0, part)
            ro = helpers_ref.ref_ops()
            ro.omp_set_num_threads(threads)
            ref = np.zeros(n, dtype=flat.np_dtype)
            getattr(ro, "internal_vector_product_addr_" + flat.dtype)(H.address, tr, 1.0, np.ascontiguousarray(x_cluster_host[o:o + s]), 0.0, ref, True)
            rec["parity_rel_err_vs_reference"] = max_over_ranks(np.linalg.norm(part.cpu().numpy() - ref) / np.linalg.norm(ref))
        out[tr] = rec
    if world == 1:
        out["driver"] = "engine call (1 GPU)"
    elif getattr(op, "fused", False) and getattr(op, "transposed_exchange", "nccl") == "peer":
        out["driver"] = ("op(A_p) x_p on every rank into its own peer-mapped buffer; rank-ordered slab sums pulled over NVLink "
                         "(hmb200_peer_exchange_sum) + peer all-gather; no NCCL call")
        out["peer_status"] = int(op.peer_status())
    else:
        out["driver"] = "op(A_p) x_p on every rank + NCCL all-reduce of the full-length results"
    return out


def gmres_once(solver, dop, b, gargs, barrier):
    import torch

    x = np.zeros_like(b)
    barrier()
    t0 = time.perf_counter()
    it = solver.solve(x, b, gargs)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    res = float(np.linalg.norm(dop.matvec(x) - b) / np.linalg.norm(b))
    return x, int(it), dt, res


def run_gmres(args, op, flat, tree, H, world, rank, dev, threads, log, barrier):
    """b = A x_ref, then GMRES from x = 0 with HPDDM's options; device solver vs the reference's HPDDM GMRES
    (oracle/_ref/libhpddm_ref.so) driving the reference's OpenMP product on the host cores.  Two operators: the benchmark's own
    first-kind kernel (which does not reach 1e-6 within max_it: labelled) and the reference's regularised test kernel
    (lib/htool/include/htool/testing/generator_test.hpp:180-187), on which iterations-to-1e-6 is a real comparison."""
    from htool_python_b200 import B200DistributedOperator, B200Solver

    n = args.n
    max_it = int(args.gmres_args.split("-hpddm_max_it")[1].split()[0]) if "-hpddm_max_it" in args.gmres_args else 100
    dop = op if world > 1 else B200DistributedOperator.from_engine(op.engine, flat)
    solver = B200Solver(dop)
    x_ref = np.random.default_rng(3).random(n)
    b = np.array(dop.matvec(x_ref), copy=True)
    solver.solve(np.zeros(n), b, args.gmres_args + " -hpddm_max_it 3")  # warm-up (allocations, collectives)
    x, it, dt, res = gmres_once(solver, dop, b, args.gmres_args, barrier)
    out = {"operator": args.kernel, "iterations": it, "operator_calls": int(solver.last_info["operator_calls"]), "solve_s": round(dt, 4), "true_residual": res,
           "reached_max_it": bool(it >= max_it), "args": args.gmres_args, "n": n, "n_gpus": world, "ms_per_iteration": round(1e3 * dt / max(it, 1), 3)}
    log(f"gmres [{args.kernel}]: {it} iterations, {dt:.3f} s, true residual {res:.2e}")
    ref_here = world == 1 and not args.no_cpu_baseline
    if ref_here:
        import helpers_ref

        ro = helpers_ref.ref_ops()
        ro.omp_set_num_threads(threads)
        fn = getattr(ro, "vector_product_user_addr_" + flat.dtype)
        t0 = time.perf_counter()
        x_cpu, it_cpu, calls_cpu = helpers_ref.hpddm_solve(lambda v: fn(H.address, np.ascontiguousarray(v), True), b, args=args.gmres_args)
        dt_cpu = time.perf_counter() - t0
        out.update({"reference_iterations": int(it_cpu), "reference_operator_calls": int(calls_cpu), "reference_solve_s": round(dt_cpu, 3),
                    "reference_cores": threads, "solution_rel_diff_vs_reference": float(np.linalg.norm(x - x_cpu) / np.linalg.norm(x_cpu)),
                    "reference": "HPDDM::IterativeMethod::GMRES (lib/hpddm, compiled from the reference sources) + reference OpenMP product"})
        log(f"gmres reference (CPU, {threads} threads): {it_cpu} iterations, {dt_cpu:.2f} s")
    # a tolerance this first-kind operator reaches within max_it, so that the iteration COUNTS are a real comparison
    loose = args.gmres_args + " -hpddm_tol 1e-3"
    _, it3, dt3, _ = gmres_once(solver, dop, b, loose, barrier)
    out["tol_1e-3"] = {"iterations": it3, "solve_s": round(dt3, 4)}
    if ref_here:
        t0 = time.perf_counter()
        _, it3_cpu, _ = helpers_ref.hpddm_solve(lambda v: fn(H.address, np.ascontiguousarray(v), True), b, args=loose)
        out["tol_1e-3"].update({"reference_iterations": int(it3_cpu), "reference_solve_s": round(time.perf_counter() - t0, 3)})
    log(f"gmres tol 1e-3: {out['tol_1e-3']}")
    if args.gmres_block > 1:
        mu = args.gmres_block
        B = np.asfortranarray(dop @ np.asfortranarray(np.random.default_rng(4).random((n, mu))))
        X = np.zeros_like(B, order="F")
        solver.solve(X, B, loose + " -hpddm_max_it 3")  # warm-up (block buffers)
        blk = {"mu": mu, "args": loose}
        for mode in ("lockstep", "columns"):
            X[:] = 0
            barrier()
            t0 = time.perf_counter()
            itb = solver.solve(X, B, loose + " -hpddm_block_rhs " + mode)
            __import__("torch").cuda.synchronize()
            dtb = time.perf_counter() - t0
            R = (dop @ X) - B
            blk[mode] = {"iterations": int(itb), "operator_calls": int(solver.get_information()["nb_mat_vec_prod"]), "solve_s": round(dtb, 4),
                         "max_true_residual": float(max(np.linalg.norm(R[:, c]) / np.linalg.norm(B[:, c]) for c in range(mu)))}
            if mode == "lockstep":
                X_lock = X.copy()
        blk["max_rel_diff_lockstep_vs_columns"] = float(max(np.linalg.norm(X_lock[:, c] - X[:, c]) / np.linalg.norm(X[:, c]) for c in range(mu)))
        solver.set_hpddm_args("-hpddm_block_rhs lockstep")
        out["block"] = blk
        log(f"gmres block: {blk}")
    if not args.no_gmres_regularised and flat.dtype == "d" and args.kernel != "laplace_reg":
        out["regularised"] = run_gmres_regularised(args, world, rank, dev, threads, log, barrier)
    return out


def run_gmres_regularised(args, world, rank, dev, threads, log, barrier):
    """iterations-to-1e-6 on an operator where 1e-6 is reachable: 1/(1e-5 + 4 pi r) on the same points.  The device solver runs
    on all ranks; the reference count comes from HPDDM + the reference's OpenMP product of the WHOLE operator on rank 0's
    host cores (assembled there in one piece when N > 1), cached in profiles/ because it does not depend on the GPU run."""
    import torch

    from htool_python_b200 import B200DistributedOperator, B200HMatrix, B200Solver

    n = args.n
    tree, H, flat, bt = assemble(args, world, rank, log, kernel="laplace_reg")
    if world > 1:
        dop = B200DistributedOperator(flat, tree.partition, device=dev.index)
        own = dop
    else:
        own = B200HMatrix(flat, device=dev.index)
        dop = B200DistributedOperator.from_engine(own.engine, flat)
    solver = B200Solver(dop)
    x_ref = np.random.default_rng(3).random(n)
    b = np.array(dop.matvec(x_ref), copy=True)
    solver.solve(np.zeros(n), b, args.gmres_args + " -hpddm_max_it 3")
    x, it, dt, res = gmres_once(solver, dop, b, args.gmres_args, barrier)
    out = {"operator": "laplace_reg (lib/htool/include/htool/testing/generator_test.hpp:180-187)", "iterations": it,
           "operator_calls": int(solver.last_info["operator_calls"]), "solve_s": round(dt, 4), "true_residual": res, "reached_max_it": False if res < 1e-5 else None,
           "ms_per_iteration": round(1e3 * dt / max(it, 1), 3), "rel_err_vs_x_ref": float(np.linalg.norm(x - x_ref) / np.linalg.norm(x_ref)), "n": n, "n_gpus": world}
    log(f"gmres [laplace_reg]: {it} iterations, {dt:.3f} s, true residual {res:.2e}")
    own.close()
    del solver, dop, own, H, flat
    torch.cuda.empty_cache()
    cache = os.path.join(ROOT, "profiles", f"gmres_reference_laplace_reg_N{n}.json")
    ref = None
    if os.path.exists(cache):
        try:
            ref = json.load(open(cache))
            if ref.get("args") != args.gmres_args or ref.get("eps") != args.eps:
                ref = None
        except Exception:
            ref = None
    if ref is None and not args.no_cpu_baseline and rank == 0:
        import helpers_ref

        _, cores = host_threads(1)
        ro = helpers_ref.ref_ops()
        ro.omp_set_num_threads(cores)
        os.environ["OMP_NUM_THREADS"] = str(cores)
        _, Hw, flatw, _ = assemble(args, 1, 0, log, kernel="laplace_reg") if world > 1 else assemble(args, 1, 0, log, kernel="laplace_reg")
        fn = getattr(ro, "vector_product_user_addr_" + flatw.dtype)
        bw = fn(Hw.address, np.ascontiguousarray(x_ref), True)  # the reference's own right-hand side of the one-piece operator
        t0 = time.perf_counter()
        x_cpu, it_cpu, calls_cpu = helpers_ref.hpddm_solve(lambda v: fn(Hw.address, np.ascontiguousarray(v), True), bw, args=args.gmres_args)
        dt_cpu = time.perf_counter() - t0
        ref = {"iterations": int(it_cpu), "operator_calls": int(calls_cpu), "solve_s": round(dt_cpu, 3), "cores": cores, "args": args.gmres_args, "eps": args.eps, "n": n,
               "x_norm": float(np.linalg.norm(x_cpu)), "rel_err_vs_x_ref": float(np.linalg.norm(x_cpu - x_ref) / np.linalg.norm(x_ref)),
               "true_residual": float(np.linalg.norm(fn(Hw.address, np.ascontiguousarray(x_cpu), True) - bw) / np.linalg.norm(bw)),
               "rhs_rel_diff_device_vs_reference": float(np.linalg.norm(b - bw) / np.linalg.norm(bw)),
               "solution_rel_diff_device_vs_reference": float(np.linalg.norm(x - x_cpu) / np.linalg.norm(x_cpu)),
               "what": "HPDDM::IterativeMethod::GMRES (lib/hpddm, compiled from the reference sources) + the reference's OpenMP product of the one-piece operator, host cores"}
        log(f"gmres reference [laplace_reg] (CPU, {cores} threads): {it_cpu} iterations, {dt_cpu:.2f} s")
        try:
            os.makedirs(os.path.dirname(cache), exist_ok=True)
            json.dump(ref, open(cache, "w"), indent=1)
            os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
            json.dump(ref, open(os.path.join(ROOT, "gpurun_out", os.path.basename(cache)), "w"), indent=1)
        except Exception:
            pass
        del Hw, flatw
    barrier()
    if ref is not None:
        out["reference"] = ref
        out["iterations_equal"] = bool(ref["iterations"] == it)
        # both iterates satisfy ||A x - b|| <= tol ||b||, so they differ by at most about 2 tol cond(A); cond(A) < 3 for this kernel
        # (spectrum within [1e5, 1e5 + N/(4 pi)])
        out["solution_rel_diff_bound"] = "2 * tol * cond(A) <= 6e-6"
    return out


def run_also_rhs(args, mu, op, engine, H, flat, x_vec, world, rank, dev, thr, dmma_tf, w_all, log, barrier, max_over_ranks):
    """Y = A X for blocks of `mu` right-hand sides on the operator the single-RHS run has assembled (BASELINE.json configs[3] when mu = 64
    on configs[1]'s operator): device-resident row-major product timed with CUDA events (max over ranks), the same through the public
    `matmat` with page-locked host blocks, and on EVERY rank its slab of both results against the reference's
    openmp_internal_add_hmatrix_matrix_product_row_major of its own slab.  Fewer steps than the headline (the block product is 4-10x longer)."""
    import torch
    import torch.distributed as dist

    n, es = args.n, flat.elem_bytes
    tdt = torch.complex128 if flat.dtype == "z" else torch.float64
    steps, warm = max(3, min(10, args.steps)), 3
    X_user = np.asfortranarray(np.stack([np.roll(x_vec, 7 * j) * (1.0 + 0.01 * j) for j in range(mu)], axis=1))  # as `--rhs mu` builds it
    Xc_host = np.ascontiguousarray(X_user[flat.source_permutation])
    Xc = torch.from_numpy(Xc_host).to(dev)
    Yc = torch.zeros((n, mu), dtype=tdt, device=dev)
    last = {}
    if world > 1:
        def step():
            last["y"] = op.matmat_cluster_(Xc.view(-1), Yc, mu)
    else:
        def step():
            engine.add_matrix_product_row_major("N", 1.0, Xc, 0.0, Yc, mu)
            last["y"] = Yc
    for _ in range(warm):
        step()
    barrier()
    l0 = engine.launch_count
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(steps):
        step()
    ev1.record()
    barrier()
    ms = max_over_ranks(ev0.elapsed_time(ev1)) / steps
    launches = engine.launch_count - l0
    y_dev = last["y"].detach().reshape(n, mu).cpu().numpy().copy()
    # end to end: page-locked Fortran-ordered host blocks through the public matmat (N > 1: every rank moves 1/P of the rows)
    x_blk = torch.empty(n * mu, dtype=tdt).pin_memory().numpy().reshape((n, mu), order="F")
    x_blk[...] = X_user
    if world > 1:
        apply_user = lambda: op.matmat(x_blk, out="shared")  # noqa: E731
    else:
        y_blk = torch.empty(n * mu, dtype=tdt).pin_memory().numpy().reshape((n, mu), order="F")
        apply_user = lambda: op.matmat(x_blk, out=y_blk)  # noqa: E731
    for _ in range(2):
        apply_user()
    barrier()
    e_steps = max(3, steps // 2)
    t0 = time.perf_counter()
    for _ in range(e_steps):
        Y_user = apply_user()
    torch.cuda.synchronize()
    e2e_step = max_over_ranks(time.perf_counter() - t0) / e_steps
    Y_user = np.array(Y_user, copy=True)
    shared = world > 1 and getattr(op, "_shmb", None) is not None
    # parity on this rank's slab, CPU baseline = the reference's row-major block product of the slab
    o, sz = (op.offs[rank], op.sizes[rank]) if world > 1 else (0, n)
    t_slab, ref = cpu_reference_slab(H, flat, Xc_host, mu, thr, budget_s=4.0, log=log, sync=barrier)
    nrm = float(np.linalg.norm(ref))
    parity = max_over_ranks(np.linalg.norm(y_dev[o:o + sz] - ref.reshape(sz, mu)) / nrm)
    parity_e2e = max_over_ranks(np.linalg.norm(Y_user[flat.target_permutation[o:o + sz]] - ref.reshape(sz, mu)) / nrm)
    t_max = max_over_ranks(t_slab)
    identical = None
    if world > 1:
        hs = [None] * world
        dist.all_gather_object(hs, (hashlib.blake2b(y_dev.tobytes(), digest_size=8).hexdigest(), hashlib.blake2b(Y_user.tobytes(), digest_size=8).hexdigest()))
        identical = len(set(hs)) == 1
    flops = (2 if flat.dtype == "d" else 8) * mu * w_all
    tf = flops / (ms * 1e-3) / 1e12
    log(f"also {mu} RHS: {ms:.3f} ms per block ({mu * 1e3 / ms:.0f} RHS/s), e2e {e2e_step * 1e3:.2f} ms, parity {parity:.2e} / e2e {parity_e2e:.2e}")
    return {"metric": metric_name(args, mu), "config": config_of(args, mu), "value": round(mu * 1e3 / ms, 3), "unit": UNIT, "steps": steps, "warmup": warm,
            "ms_per_step": round(ms, 4), "gpu_launches": int(launches),
            "e2e": {"value": round(mu / e2e_step, 3), "unit": UNIT, "steps": e_steps, "ms_per_step": round(e2e_step * 1e3, 4),
                    "h2d_bytes_per_step": es * n * mu * (1 if (world == 1 or shared) else world), "d2h_bytes_per_step": es * n * mu * (1 if (world == 1 or shared) else world),
                    "api": "matmat(X, out) on page-locked Fortran-ordered host blocks, user numbering" + (" (node-shared result block, 1/P of the rows per rank over PCIe)" if shared else "")},
            "roofline": {"bound": "tensor", "kernel": "hmatrix_mm_kernel<T,phase 1|2> panel GEMMs (mma.sync.m8n8k4.f64, SASS DMMA)", "achieved": round(tf, 2),
                         "peak": dmma_tf * world if dmma_tf else None, "peak_source": "hmb200_fp64_rate (DMMA issue rate measured in this run) x GPUs", "unit": "TFLOP/s",
                         "frac": round(tf / (dmma_tf * world), 4) if dmma_tf else None, "algorithmic_flops_per_step": flops},
            "cpu_baseline": {"value": round(mu / t_max, 4), "unit": UNIT, "cores": thr * world, "kind": "reference", "ms_per_product": round(t_max * 1e3, 2),
                             "sample": f"full {mu}-RHS row-major block products of the same operator (best of >= 2 within 4 s), reference OpenMP path" +
                                       ("" if world == 1 else f"; emulated ranks: each of the {world} processes multiplies its own slab, value = mu / max_p t(slab_p)")},
            "parity_rel_err_vs_reference": parity, "parity_rel_err_vs_reference_e2e": parity_e2e, "replicas_bit_identical": identical}


def cpu_reference_slab(H, flat, x_cluster_host, mu, threads, budget_s, log, sync=None):
    """times the REFERENCE's OpenMP product of this process's H (the whole operator, or this rank's block-row slab) in cluster
    numbering -- openmp_internal_add_hmatrix_vector_product / ..._matrix_product_row_major, what RestrictedGlobalToLocalHMatrix
    calls -- and returns (best seconds, result)"""
    import helpers_ref

    ro = helpers_ref.ref_ops()
    ro.omp_set_num_threads(threads)
    x = np.ascontiguousarray(x_cluster_host)
    y = np.zeros((flat.n_rows, mu) if mu > 1 else flat.n_rows, dtype=flat.np_dtype)
    if mu > 1:
        fn = getattr(ro, "internal_matrix_product_row_major_addr_" + flat.dtype)
        call = lambda: fn(H.address, "N", 1.0, x, 0.0, y, mu, True)  # noqa: E731
    else:
        fn = getattr(ro, "internal_vector_product_addr_" + flat.dtype)
        call = lambda: fn(H.address, "N", 1.0, x, 0.0, y, True)  # noqa: E731
    call()  # warm-up
    if sync:
        sync()  # emulated ranks: all processes multiply at the same time
    times = []
    t_all = time.perf_counter()
    while len(times) < 10 and (time.perf_counter() - t_all < budget_s or len(times) < 2):
        t0 = time.perf_counter()
        call()
        times.append(time.perf_counter() - t0)
    best = min(times)
    log(f"cpu reference product: {len(times)} runs, best {best * 1e3:.1f} ms, {threads} threads")
    return best, y


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    thr, cores = host_threads(1)
    os.environ.setdefault("OMP_NUM_THREADS", str(thr))

    def log(msg):
        print(f"[bench-ref] {msg}", file=sys.stderr, flush=True)

    import helpers_ref

    ro = helpers_ref.ref_ops()
    ro.omp_set_num_threads(thr)
    tree, H, flat, build_t = assemble(args, 1, 0, log)
    rng = np.random.default_rng(1)
    x_user = rng.random(args.n)
    if flat.dtype == "z":
        x_user = x_user + 1j * rng.random(args.n)
    mu = max(1, args.rhs)
    if mu > 1:
        x_user = np.asfortranarray(np.stack([np.roll(x_user, 7 * j) * (1.0 + 0.01 * j) for j in range(mu)], axis=1))
    fn = getattr(ro, ("matrix_product_user_addr_" if mu > 1 else "vector_product_user_addr_") + flat.dtype)
    for _ in range(max(1, min(args.warmup, 3))):
        fn(H.address, x_user, True)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        fn(H.address, x_user, True)
    dt = (time.perf_counter() - t0) / args.steps
    v = round(mu / dt, 4)
    line = {"impl": "reference", "metric": metric_name(args, mu), "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": round(dt * 1e3, 3), "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64" if flat.dtype == "d" else "c128",
            "data": "synthetic", "config": config_of(args, mu), "impl_config": {**build_t, "threads": thr},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": thr, "kind": "reference",
                             "sample": f"each step = one full product ({mu} RHS), reference add_hmatrix_{'matrix' if mu > 1 else 'vector'}_product(exec_compat::par) from oracle/_ref, "
                                       "one process (no MPI on the box: the reference arm does not scale with --gpus)"},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    a = parse()
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)

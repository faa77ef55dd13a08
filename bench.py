#!/usr/bin/env python
"""bench.py -- H-matvecs/sec of the N=1M Laplace 1/r H-matrix (BASELINE.json metric), 1..8 B200 vs the host CPU.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--n POINTS]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one H-matrix-vector product y = A x of the whole operator:
  value  : device-resident (x, y replicated in cluster numbering in HBM; slab product + NCCL all-gather for N > 1),
           timed with CUDA events on the launching stream, barrier + synchronize on both sides, max over ranks;
  e2e    : the same product through the reference-facing API (`B200HMatrix * x` / `B200DistributedOperator * x`) with
           HOST numpy vectors in user numbering: pinned H2D, both permutations, product, D2H inside the timed region;
  roofline: the dominant kernel's ALGORITHMIC bytes / its CUDA-event time vs MEASURED_PEAKS.json hbm_gbs;
  cpu_baseline: the reference's own OpenMP product (oracle/_ref, compiled from the reference's sources) on the same
           H-matrix object on this box's host cores.
The H-matrix is assembled by the REFERENCE's builder (HMatrixTreeBuilder::build, OpenMP, native generator) on the host
before the timed region; strong scaling: the N=1M operator is split into the reference's block-row slabs.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "H-matvecs/sec (N=1M Laplace 1/r, double, single RHS)"
UNIT = "matvecs/s"
WORKLOADS = {"laplace": "Laplace 1/(4 pi r), double,", "helmholtz": "Helmholtz exp(ikr)/(4 pi r), complex<double>,"}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--n", type=int, default=int(os.environ.get("HMB200_BENCH_N", 1_000_000)))
    ap.add_argument("--eps", type=float, default=1e-3)
    ap.add_argument("--eta", type=float, default=10.0)
    ap.add_argument("--leaf", type=int, default=64)
    ap.add_argument("--kernel", default="laplace", choices=["laplace", "helmholtz"],
                    help="laplace = BASELINE.json configs[1] (the metric); helmholtz = configs[2] (complex<double>, off-metric check)")
    ap.add_argument("--wavenumber", type=float, default=0.0, help="helmholtz: k (default: 10 points per wavelength)")
    ap.add_argument("--rhs", type=int, default=1, help="right-hand sides per step (64 = BASELINE.json configs[3], DMMA multi-RHS path; off-metric)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--gmres", action="store_true", help="also solve A x = b with the device-resident GMRES (BASELINE.json configs[4]; adds a 'gmres' object)")
    ap.add_argument("--gmres-block", type=int, default=0, help="with --gmres: also solve a block of this many right-hand sides in lockstep "
                    "(HPDDM's GMRES with mu > 1 on the multi-RHS product) and column by column, tol 1e-3")
    ap.add_argument("--gmres-args", default="-hpddm_krylov_method gmres -hpddm_schwarz_method none -hpddm_variant right -hpddm_tol 1e-6 -hpddm_gmres_restart 40 -hpddm_max_it 200")
    ap.add_argument("--engine-opts", default="", help="k=v,k=v overrides of hmb200_options (tuning runs)")
    return ap.parse_args()


def workload_name(args, n, mu):
    cfg = (1 if mu == 1 else 3) if args.kernel == "laplace" else 2
    return (f"{WORKLOADS[args.kernel]} H-matrix, N={n} points on the unit sphere surface (seed 0), eps={args.eps}, eta={args.eta}, "
            f"leaf {args.leaf}, {'single-RHS matvec' if mu == 1 else str(mu) + '-RHS matmat (DMMA path)'} (BASELINE.json configs[{cfg}])")


def sphere_points(n, seed=0):
    """N points uniformly on the unit sphere surface (normalised Gaussian, seed 0) -- SURVEY.md 8(d) C2."""
    g = np.random.default_rng(seed).standard_normal((3, n))
    g /= np.linalg.norm(g, axis=0)
    return np.asfortranarray(g)


def host_threads(world):
    try:
        cores = len(os.sched_getaffinity(0))
    except Exception:
        cores = os.cpu_count() or 1
    return max(1, cores // max(world, 1)), cores


def assemble(args, world, rank, log):
    """Cluster tree + this rank's block-row slab, assembled by the reference's own builders (bridge)."""
    from htool_python_b200 import FlatHMatrix, bridge

    t0 = time.time()
    pts = sphere_points(args.n)
    tree = bridge.create_cluster_tree(pts, 2, world, args.leaf)
    t1 = time.time()
    slab = rank if world > 1 else -1
    # keep every admissible block below 46340 x 46340: the reference's ACA stop test `q*(n1+n2) > n1*n2` is int arithmetic
    # (lib/htool/include/htool/hmatrix/lrmat/sympartialACA.hpp:102); above that it overflows and the block is built DENSE
    depth = 0
    while args.n / 2 ** depth > 46340:
        depth += 1
    depth += 1 if depth else 0  # the builder's depth test is off by one against size = n / 2^depth (measured)
    kernel = getattr(args, "kernel", "laplace")
    k = getattr(args, "wavenumber", 0.0) or 2 * np.pi * np.sqrt(args.n / (4 * np.pi)) / 10.0  # ~10 points per wavelength
    H = bridge.build_hmatrix(kernel, tree, epsilon=args.eps, eta=args.eta, target_partition_number=slab, wavenumber=k,
                             partition_number_for_symmetry=slab, minimal_target_depth=depth, minimal_source_depth=depth)
    t2 = time.time()
    flat = FlatHMatrix.from_htool(H)
    big = (flat.rank < 0) & (flat.nrows.astype(np.int64) * flat.ncols > 1 << 22)
    if big.any():
        raise RuntimeError(f"the reference built {int(big.sum())} dense blocks above 4M coefficients (ACA int overflow?)")
    log(f"rank {rank}: tree {t1 - t0:.1f}s, assembly {t2 - t1:.1f}s, leaves {flat.n_leaves}, W {flat.stored_coefficients:.4e}")
    return tree, H, flat, dict(tree_s=round(t1 - t0, 2), assembly_s=round(t2 - t1, 2), minimal_block_depth=depth)


class ClockSampler(threading.Thread):
    """samples SM clock + throttle reasons of one GPU through NVML while the timed region runs"""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.sm, self.reasons, self.sm_max = index, False, [], set(), None
        self.t, self.window = [], None  # sample times; (t0, t1) of the timed region
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
                self.sm.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                self.t.append(time.perf_counter())
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.004)

    def result(self):
        self.stop_flag = True
        if self.nv is None or not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": self.sm_max, "reasons": []}
        sm = self.sm
        if self.window:  # samples inside the timed region; a region shorter than the sampling period falls back to all samples
            inside = [v for v, t in zip(self.sm, self.t) if self.window[0] <= t <= self.window[1]]  # (the thread runs from the warm-up on)
            sm = inside or sm
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": self.sm_max, "reasons": sorted(self.reasons), "samples": len(sm)}


def w_split(flat):
    """algorithmic coefficients streamed by phase 1 (V factors) and phase 2 (U factors + dense leaves); DESIGN.md"""
    r = np.maximum(flat.rank, 0).astype(np.float64)
    lr = flat.rank >= 0
    w1 = float(np.sum(r[lr] * flat.ncols[lr]))
    w2 = float(np.sum(r[lr] * flat.nrows[lr]) + np.sum(flat.nrows[~lr].astype(np.float64) * flat.ncols[~lr]))
    return w1, w2


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

    from htool_python_b200 import B200DistributedOperator, B200HMatrix

    tree, H, flat, build_t = assemble(args, world, rank, log)
    opts = {k: int(v) for k, v in (kv.split("=") for kv in args.engine_opts.split(",") if kv)}
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
    perm_s = flat.source_permutation
    mu = max(1, args.rhs)
    if mu > 1:  # X: n x mu, Fortran-ordered for the user API, row-major in cluster numbering for the operator interface
        x_user = np.asfortranarray(np.stack([np.roll(x_user, 7 * j) * (1.0 + 0.01 * j) for j in range(mu)], axis=1))
        x_cluster = torch.from_numpy(np.ascontiguousarray(x_user[perm_s])).to(dev)
        y_cluster = torch.zeros(n, mu, dtype=tdt, device=dev)
    else:
        x_cluster = torch.from_numpy(np.ascontiguousarray(x_user[perm_s])).to(dev)
        y_cluster = torch.zeros(n, dtype=tdt, device=dev)

    if mu > 1 and world > 1:
        def step():
            op.matmat_cluster_(x_cluster.view(-1), y_cluster, mu)
    elif mu > 1:
        def step():
            engine.add_matrix_product_row_major("N", 1.0, x_cluster, 0.0, y_cluster, mu)
    elif world > 1 and op.fused:
        def step():  # slab product whose phase-2 epilogue stores the rows into every rank's replicated vector (no NCCL call)
            op.matvec_cluster_fused_(x_cluster)
    elif world > 1:
        def step():
            op.matvec_cluster_(x_cluster, y_cluster)
    else:
        def step():
            engine.add_vector_product("N", 1.0, x_cluster, 0.0, y_cluster)
    if mu > 1:
        apply_user = lambda: op @ x_user  # noqa: E731
    else:  # page-locked host buffers (torch pin_memory): the library DMAs straight from / into them
        x_pin = torch.from_numpy(np.ascontiguousarray(x_user)).pin_memory().numpy()
        y_pin = torch.zeros(n, dtype=tdt).pin_memory().numpy()
        apply_user = lambda: op.matvec(x_pin, out=y_pin)  # noqa: E731

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
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_region = time.perf_counter()
    ev0.record()
    for _ in range(args.steps):
        step()
    ev1.record()
    barrier()
    sampler.window = (t_region, time.perf_counter())
    clocks = sampler.result()
    ms_total = ev0.elapsed_time(ev1)
    launches = engine.launch_count - launches0
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item()) / args.steps
    # per-kernel durations: the SAME steps once more with CUDA events recorded around every kernel on the launching stream
    # (the timed region above replays CUDA graphs, which carry no per-kernel events)
    engine.enable_timing(True)
    for _ in range(args.steps):
        step()
    barrier()
    kms = engine.kernel_ms_avg()
    engine.enable_timing(False)
    if world > 1 and op.fused and op.peer_status() != 0:
        raise SystemExit(f"bench.py: rank {rank}: a peer did not publish its epoch (status {op.peer_status()})")

    # ---- end to end through the public API with host buffers ----------------------------------------------------
    for _ in range(2):
        y_user = apply_user()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        y_user = apply_user()
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_step = float(t.item()) / args.steps

    # ---- GMRES with the Krylov basis resident on the device (off-metric extra) -----------------------------------------
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
    achieved = alg[dom] / (kms[dom] * 1e-3) / 1e9 if kms[dom] > 0 else 0.0
    roofline = {"bound": "hbm", "kernel": f"hmatrix_stream_kernel<{'double' if es == 8 else 'double2'},{2 if dom == 'phase2' else 1}>", "achieved": round(achieved, 1),
                "peak": peak, "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback (B200_PROFILING.md)", "unit": "GB/s",
                "frac": round(achieved / peak, 4), "traffic": None, "algorithmic_bytes_per_launch": alg[dom],
                "kernel_ms": {k: round(v, 4) for k, v in kms.items()},
                "whole_product_GBps": round(flat.algorithmic_bytes(1) / (kms["total"] * 1e-3) / 1e9, 1) if kms["total"] > 0 else None}
    if mu > 1:  # multi-RHS: bounded by the FP64 tensor (DMMA) pipe; kernel_ms are per pass of 64 (32 complex) columns
        passes = -(-mu // (64 if es == 8 else 32))
        flops = (2 if es == 8 else 8) * mu * flat.stored_coefficients
        tf = flops / (kms["total"] * passes * 1e-3) / 1e12 if kms["total"] > 0 else 0.0
        roofline = {"bound": "tensor", "kernel": "hmatrix_mm_kernel<T,1|2> (mma.sync.m8n8k4.f64)", "achieved": round(tf, 2), "peak": 37.0,
                    "peak_source": "measured DMMA issue rate on this B200 (profiles/r01_dmma_rate.log); MEASURED_PEAKS.json has no FP64 figure",
                    "unit": "TFLOP/s", "frac": round(tf / 37.0, 4), "traffic": None, "algorithmic_flops_per_step": flops, "passes": passes,
                    "kernel_ms_per_pass": {k: round(v, 4) for k, v in kms.items()}}

    try:  # traffic of the dominant kernel from the committed ncu capture of this exact workload (null otherwise)
        tr = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json"))).get(f"{args.kernel}_{n}_{world}" + (f"_rhs{mu}" if mu > 1 else ""))
        if tr and args.eps == 1e-3 and args.eta == 10.0 and args.leaf == 64:
            roofline["traffic"] = tr["phase2" if (mu > 1 or dom == "phase2") else "phase1"]
            roofline["traffic_source"] = tr["source"]
    except Exception:
        pass

    # ---- CPU baseline + parity (rank 0, N=1 only) ---------------------------------------------------------------------
    cpu = None
    parity = None
    if world == 1 and not args.no_cpu_baseline:
        cpu, y_ref = cpu_reference_product(H, flat, x_user, thr, budget_s=15.0, log=log)
        parity = float(np.linalg.norm(y_user - y_ref) / np.linalg.norm(y_ref))
        log(f"parity vs reference CPU product: {parity:.3e}")

    if rank == 0:
        line = {
            "metric": METRIC, "value": round(mu * 1e3 / ms_step, 3), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": round(ms_step, 4), "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64" if flat.dtype == "d" else "c128",
            "data": "synthetic",
            "config": {"workload": workload_name(args, n, mu),
                       "rhs_per_step": mu, "partition": f"block-row slabs x{world}",
                       "allgather": ("none (1 GPU)" if world == 1 else "fused into the phase-2 epilogue: peer stores over NVLink + epoch flags (CUDA IPC)" if (mu == 1 and op.fused) else "NCCL all-gather"),
                       "launch": "CUDA graph replay of the product (captured by libhmb200.so)" if mu == 1 else "eager", "l2": "inputs larger than L2 (factor stream >> 126 MB, read once per step)",
                       "stored_coefficients_rank0": flat.stored_coefficients, "leaves_rank0": flat.n_leaves,
                       "assembly": "reference HMatrixTreeBuilder (OpenMP, native generator), outside the timed region", **build_t},
            "clocks": clocks,
            # bytes summed over the ranks: at N > 1 the single-RHS `matvec` uploads 1/N of x per rank (chunks exchanged over
            # NVLink); every rank downloads the full y (the reference API returns the replicated result on every rank)
            "e2e": {"value": round(mu / e2e_step, 3), "unit": UNIT,
                    "h2d_bytes_per_step": es * n * mu * (1 if (world == 1 or (mu == 1 and op.fused)) else world), "d2h_bytes_per_step": es * n * mu * world,
                    "ms_per_step": round(e2e_step * 1e3, 4)},
            "gpu_launches": int(launches),
            "roofline": roofline,
            "cpu_baseline": cpu,
            "parity_rel_err_vs_reference": parity,
        }
        line["config"]["e2e_api"] = "B200HMatrix.matvec(x, out) / B200DistributedOperator.matvec(x, out) on page-locked host numpy buffers, user numbering"
        if gm is not None:
            line["gmres"] = gm
        print(json.dumps(line), flush=True)
    barrier()
    op.close()
    if world > 1:
        dist.destroy_process_group()


def run_gmres(args, op, flat, tree, H, world, rank, dev, threads, log, barrier):
    """b = A x_ref, then GMRES from x = 0 with HPDDM's options; device solver vs (rank 0, N=1) the reference's HPDDM GMRES
    (oracle/_ref/libhpddm_ref.so) driving the reference's OpenMP product on the host cores."""
    import torch

    from htool_python_b200 import B200DistributedOperator, B200Solver

    n = args.n
    dop = op if world > 1 else B200DistributedOperator.from_engine(op.engine, flat)
    solver = B200Solver(dop)
    x_ref = np.random.default_rng(3).random(n)
    b = dop.matvec(x_ref)
    x = np.zeros(n)
    solver.solve(x, b, args.gmres_args + " -hpddm_max_it 3")  # warm-up (allocations, NCCL)
    x[:] = 0
    barrier()
    t0 = time.perf_counter()
    it = solver.solve(x, b, args.gmres_args)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    res = float(np.linalg.norm(dop.matvec(x) - b) / np.linalg.norm(b))
    out = {"iterations": int(it), "operator_calls": int(solver.last_info["operator_calls"]), "solve_s": round(dt, 4), "true_residual": res,
           "args": args.gmres_args, "n": n, "n_gpus": world, "ms_per_iteration": round(1e3 * dt / max(it, 1), 3)}
    log(f"gmres: {it} iterations, {dt:.3f} s, true residual {res:.2e}")
    if world == 1 and not args.no_cpu_baseline:
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
    x[:] = 0
    barrier()
    t0 = time.perf_counter()
    it3 = solver.solve(x, b, loose)
    torch.cuda.synchronize()
    out["tol_1e-3"] = {"iterations": int(it3), "solve_s": round(time.perf_counter() - t0, 4)}
    if world == 1 and not args.no_cpu_baseline:
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
            torch.cuda.synchronize()
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
    return out


def cpu_reference_product(H, flat, x_user, threads, budget_s, log):
    """times the REFERENCE's OpenMP product (add_hmatrix_vector_product(exec_compat::par), user numbering) on the host"""
    import helpers_ref

    ro = helpers_ref.ref_ops()
    ro.omp_set_num_threads(threads)
    mu = x_user.shape[1] if x_user.ndim == 2 else 1
    # multi-RHS: add_hmatrix_matrix_product(exec_compat::par, ...) on the Fortran-ordered block (hmatrix.hpp:119-138 path)
    fn = getattr(ro, ("matrix_product_user_addr_" if mu > 1 else "vector_product_user_addr_") + flat.dtype)
    y = fn(H.address, x_user, True)  # warm-up
    times = []
    t_all = time.perf_counter()
    while len(times) < 10 and (time.perf_counter() - t_all < budget_s or len(times) < 2):
        t0 = time.perf_counter()
        y = fn(H.address, x_user, True)
        times.append(time.perf_counter() - t0)
    best = min(times)
    log(f"cpu reference product: {len(times)} runs, best {best * 1e3:.1f} ms, {threads} threads")
    return {"value": round(mu / best, 4), "unit": UNIT, "cores": threads, "kind": "reference",
            "sample": f"{len(times)} full products ({mu} RHS each) of the same N={flat.n_cols} operator (best), reference OpenMP path "
                      "add_hmatrix_{vector,matrix}_product(exec_compat::par) compiled from the reference sources (oracle/_ref)",
            "ms_per_product": round(best * 1e3, 2)}, y


def run_reference(args):
    world = int(os.environ.get("WORLD_SIZE", "1"))
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
    x_user = np.random.default_rng(1).random(args.n)
    if flat.dtype == "z":
        x_user = x_user + 1j * np.random.default_rng(2).random(args.n)
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
    v = round(max(1, args.rhs) / dt, 4)
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": round(dt * 1e3, 3), "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args, args.n, mu), "rhs_per_step": mu, **build_t},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": thr, "kind": "reference",
                             "sample": "each step = one full matvec, reference add_hmatrix_vector_product(exec_compat::par) from oracle/_ref"},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    a = parse()
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)

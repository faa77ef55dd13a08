"""First GPU contact: probe the box, run smoke + a few parity cases, time a mid-size product."""
import os, sys, time, subprocess, json
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from htool_python_b200 import B200HMatrix, FlatHMatrix, bridge, Engine
from oracle import flat_oracle as fo

def sh(c):
    try: return subprocess.run(c, shell=True, capture_output=True, text=True, timeout=60).stdout.strip()
    except Exception as e: return f"ERR {e}"
print("nproc", os.cpu_count(), "| mem:", sh("free -g | sed -n 2p"), "| mpirun:", sh("which mpirun mpiexec"), flush=True)
print(sh("nvidia-smi --query-gpu=name,memory.total,clocks.max.sm,clocks.sm,power.limit --format=csv"), flush=True)
print("cgroup cpu:", sh("cat /sys/fs/cgroup/cpu.max 2>/dev/null"), "| cgroup mem:", sh("cat /sys/fs/cgroup/memory.max 2>/dev/null"), flush=True)

import __graft_entry__ as ge
ge.smoke()

rng = np.random.default_rng(1)
def sphere(N, seed=0):
    g = np.random.default_rng(seed).standard_normal((3, N)); g /= np.linalg.norm(g, axis=0); return np.asfortranarray(g)

# parity cases
for kernel, sym, uplo, N, leaf in [("laplace_reg","N","N",5000,64),("laplace_reg","S","L",5000,64),("complex_reg","N","N",3000,32),("complex_herm","H","U",3000,32),("invdist_reg","N","N",3000,10)]:
    pts = sphere(N, 3)
    tree = bridge.create_cluster_tree(pts, 2, 1, leaf)
    H = bridge.build_hmatrix(kernel, tree, epsilon=1e-4, eta=10.0, symmetry=sym, UPLO=uplo)
    flat = FlatHMatrix.from_htool(H)
    dt = flat.np_dtype
    x = (rng.random(N) + (1j*rng.random(N) if flat.dtype=="z" else 0)).astype(dt)
    y0 = (rng.random(N) + (1j*rng.random(N) if flat.dtype=="z" else 0)).astype(dt)
    alpha, beta = (1.5, -0.5) if flat.dtype=="d" else (1.5-0.3j, 0.2+0.7j)
    e = Engine(flat)
    y = y0.copy(); e.add_vector_product("N", alpha, x, beta, y)
    yo = y0.copy(); fo.add_vector_product(flat, "N", alpha, x, beta, yo)
    yu = e.matvec_user(x); yuo = fo.matvec_user(flat, x)
    print(f"parity {kernel} {sym}{uplo} N={N} leaf={leaf}: cluster {np.linalg.norm(y-yo)/np.linalg.norm(yo):.2e} user {np.linalg.norm(yu-yuo)/np.linalg.norm(yuo):.2e}", flush=True)
    e.close()

# timing
for N in (100000, 400000):
    t=time.time(); pts = sphere(N); tree = bridge.create_cluster_tree(pts, 2, 1, 64); t_tree=time.time()-t
    t=time.time(); H = bridge.build_hmatrix("laplace", tree, epsilon=1e-3, eta=10.0); t_build=time.time()-t
    flat = FlatHMatrix.from_htool(H)
    info_ref = H.get_local_information()
    t=time.time(); e = Engine(flat); t_create=time.time()-t
    inf = e.info
    print(f"N={N}: tree {t_tree:.1f}s build {t_build:.1f}s create {t_create:.1f}s (pack {inf['pack_seconds']:.1f}s upload {inf['upload_seconds']:.1f}s) leaves={flat.n_leaves} W={flat.stored_coefficients:.3e} CR={info_ref['Compression_ratio']} rank_mean={info_ref['Rank_mean']}", flush=True)
    print("   info", {k:v for k,v in inf.items()}, flush=True)
    x = torch.rand(N, dtype=torch.float64, device="cuda"); y = torch.zeros(N, dtype=torch.float64, device="cuda")
    e.enable_timing(True)
    for it in range(6):
        e.add_vector_product("N", 1.0, x, 0.0, y); torch.cuda.synchronize()
        ms = e.last_kernel_ms()
        ab = flat.algorithmic_bytes(1)
        print(f"   it{it}: {ms}  alg GB/s total {ab/ms['total']/1e6:.0f}  p1 {inf['stream1_bytes']/max(ms['phase1'],1e-9)/1e6:.0f} p2 {inf['stream2_bytes']/max(ms['phase2'],1e-9)/1e6:.0f}", flush=True)
    # correctness at size vs oracle (CPU seconds)
    xh = x.cpu().numpy(); yo = np.zeros(N); t=time.time(); fo.add_vector_product(flat, "N", 1.0, xh, 0.0, yo); t_or=time.time()-t
    print(f"   rel err vs oracle {np.linalg.norm(y.cpu().numpy()-yo)/np.linalg.norm(yo):.2e} (oracle {t_or:.1f}s)", flush=True)
    e.close(); del H, flat

"""`B200DistributedOperator`: the block-row product of the reference's `DistributedOperator`, one process per GPU.

Mirrors (paths under /root/reference/lib/htool/include/htool/distributed_operator):
  * the partition of the target cluster tree into contiguous row slabs, one per rank
    (implementations/partition_from_cluster.hpp:11-42, utility.hpp:56: rank p holds cluster_on_partition(p) x source root);
  * `__mul__` / `__matmul__` = add_distributed_operator_{vector,matrix}_product_global_to_global
    (linalg/add_distributed_operator_vector_product_global_to_global.hpp:18-118,
     linalg/add_distributed_operator_matrix_product_global_to_global.hpp:138-280):
        x_user --global_to_root_cluster--> x_cluster (replicated)
        slab_p = H_p x_cluster                       (this rank's engine, hand-written sm_100a kernels)
        all-gather of the slabs                      (reference: MPI_Allgatherv :76; here NCCL over NVLink)
        y_cluster --root_cluster_to_global--> y_user
  * the solver entry (linalg/add_distributed_operator_vector_product_local_to_local.hpp:19-89): the gathered,
    replicated cluster-numbered vector is exactly the input of the next product, so `matvec_cluster_` keeps it on
    the device and GMRES never leaves the GPU.

The collective runs on `torch.distributed` (backend "nccl" on GPUs; "gloo" works for the CPU tests of the host logic with
a stand-in slab product).  Slabs are disjoint and contiguous in cluster numbering, so the all-gather is in place.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import numpy as np

from . import _lib


def partition_displacements(partition: Sequence[Tuple[int, int]]):
    """(offsets, sizes) of the rank slabs; checks they tile [0, n) contiguously (partition_from_cluster.hpp:22-25)."""
    offs = [int(o) for o, _ in partition]
    sizes = [int(s) for _, s in partition]
    cur = 0
    for o, s in zip(offs, sizes):
        if o != cur or s < 0:
            raise ValueError("partition slabs must be contiguous and ordered by rank")
        cur += s
    return offs, sizes


def allgather_slabs_(buf, offs: List[int], sizes: List[int], mu: int, group=None):
    """In-place all-gather of row slabs of the replicated row-major (n x mu) tensor `buf`: afterwards every rank holds
    every slab (MPI_Allgatherv of add_distributed_operator_vector_product_global_to_global.hpp:76)."""
    import torch.distributed as dist

    if not dist.is_initialized():
        return buf
    world = dist.get_world_size(group)
    if world == 1:
        return buf
    rank = dist.get_rank(group)
    flat = buf.view(-1)
    if len(set(sizes)) == 1 and flat.numel() == world * sizes[0] * mu:
        mine = flat[offs[rank] * mu:(offs[rank] + sizes[rank]) * mu]
        dist.all_gather_into_tensor(flat, mine, group=group)  # in place: slab p already sits at p*count
    else:  # unequal slabs: one broadcast per rank (P <= 8)
        for p in range(world):
            if sizes[p]:
                dist.broadcast(flat[offs[p] * mu:(offs[p] + sizes[p]) * mu], src=dist.get_global_rank(group, p) if group is not None else p, group=group)
    return buf


def exchange_sum_slabs_(full, offs: List[int], sizes: List[int], out_local, beta=0.0, group=None):
    """out_local = beta * out_local + sum over the ranks r (ascending) of full_r[offs[me] : offs[me] + sizes[me]], where
    `full_r` is rank r's full-length contribution: the exchange step of the TRANSPOSED local-to-local product
    (add_distributed_operator_vector_product_local_to_local.hpp:60-87: MPI_Alltoallv of the slabs, scal by beta, then one axpy
    per rank in rank order).  Point-to-point sends / receives (both NCCL and gloo have them), summed in rank order so that the
    result does not depend on arrival order."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    mine = full[offs[rank]:offs[rank] + sizes[rank]]
    if beta == 0:
        out_local.zero_()
    else:
        out_local.mul_(beta)
    if world == 1:
        out_local.add_(mine)
        return out_local
    rbuf = torch.empty((world, sizes[rank]), dtype=full.dtype, device=full.device)
    peer = lambda q: dist.get_global_rank(group, q) if group is not None else q  # noqa: E731
    ops, keep = [], []  # `keep`: the send buffers stay referenced until every request has completed
    for q in range(world):
        if q == rank:
            continue
        if sizes[q]:
            keep.append(full[offs[q]:offs[q] + sizes[q]].contiguous())
            ops.append(dist.P2POp(dist.isend, keep[-1], peer(q), group))
        if sizes[rank]:
            ops.append(dist.P2POp(dist.irecv, rbuf[q], peer(q), group))
    for req in (dist.batch_isend_irecv(ops) if ops else []):
        req.wait()
    del keep
    for q in range(world):
        out_local.add_(mine if q == rank else rbuf[q])
    return out_local


def _nvml_handle_of_cuda_device(pynvml, index: int):
    """NVML handle of CUDA device `index`.  NVML numbers ALL physical GPUs in PCI order and knows nothing of CUDA_VISIBLE_DEVICES /
    CUDA_DEVICE_ORDER, so the CUDA index is only an NVML index on a full, unreordered allocation: match by UUID, and fall back to
    the index only when no device mask is set.  Raises when the device cannot be identified (the caller then pins nothing, which
    is better than pinning to the cores of another GPU's NUMA node)."""
    import os

    import torch

    try:
        uuid = str(torch.cuda.get_device_properties(index).uuid)
        uuid = uuid if uuid.startswith(("GPU-", "MIG-")) else "GPU-" + uuid
        try:
            return pynvml.nvmlDeviceGetHandleByUUID(uuid)
        except TypeError:
            return pynvml.nvmlDeviceGetHandleByUUID(uuid.encode())
    except Exception:
        if os.environ.get("CUDA_VISIBLE_DEVICES") or os.environ.get("CUDA_DEVICE_ORDER"):
            raise
        return pynvml.nvmlDeviceGetHandleByIndex(index)


def bind_to_gpu_cpus(local_rank: int, local_world: int) -> list:
    """Pins this process to its share of the host cores next to GPU `local_rank` (NVML's ideal CPU set of the device, split
    between the ranks whose GPUs report the same set) -- what `mpirun --bind-to` does for the reference's ranks.  One process
    per GPU with host-synchronous calls is sensitive to scheduler jitter: every collective waits for the slowest rank.
    Returns the CPU list (empty when NVML or sched_setaffinity is unavailable; nothing is changed then)."""
    import os

    try:
        import pynvml

        pynvml.nvmlInit()
        allowed = sorted(os.sched_getaffinity(0))
        words = (max(allowed) // 64) + 1
        sets = []
        for g in range(local_world):
            mask = pynvml.nvmlDeviceGetCpuAffinity(_nvml_handle_of_cuda_device(pynvml, g), words)
            sets.append(tuple(c for c in allowed if (mask[c // 64] >> (c % 64)) & 1))
        mine = sets[local_rank]
        if not mine:
            return []
        sharers = [g for g in range(local_world) if sets[g] == mine]
        k, m = sharers.index(local_rank), len(sharers)
        per = max(1, len(mine) // m)
        cpus = list(mine[k * per:(k + 1) * per]) or list(mine)
        os.sched_setaffinity(0, cpus)
        return cpus
    except Exception:
        return []


class _RawDeviceArray:
    """`__cuda_array_interface__` view of device memory owned by libhmb200.so (wrapped by torch.as_tensor without a copy)"""

    def __init__(self, ptr, n, is_complex):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": "<c16" if is_complex else "<f8", "data": (int(ptr), False), "version": 2, "strides": None}


def _device_tensor(ptr, n, is_complex, device):
    import torch

    with torch.cuda.device(device):
        return torch.as_tensor(_RawDeviceArray(ptr, n, is_complex), device=device)


class B200DistributedOperator:
    """Drop-in for the product surface of `Htool.DistributedOperator` (`shape`, `*`, `@`;
    src/htool/distributed_operator/distributed_operator.hpp:16-65) backed by one `Engine` per rank.

    slab_flat : FlatHMatrix of THIS rank's row slab (n_rows = local size, n_cols = global source size), as assembled by
                the reference with target_partition_number = rank
    partition : [(offset, size)] of every rank's slab in target cluster numbering
    """

    @classmethod
    def from_engine(cls, engine, flat, group=None):
        """single-process operator over an EXISTING engine (e.g. `B200HMatrix(...).engine`): nothing is packed or uploaded again"""
        return cls(flat, [(0, flat.n_rows)], group=group, device=engine.device, engine=engine)

    def __init__(self, slab_flat, partition, group=None, device: Optional[int] = None, engine=None, fused_allgather=None, **options):
        import torch
        import torch.distributed as dist

        from .engine import Engine

        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.offs, self.sizes = partition_displacements(partition)
        if len(self.offs) != self.world:
            raise ValueError(f"partition has {len(self.offs)} slabs but the process group has {self.world} ranks")
        if slab_flat.n_rows != self.sizes[self.rank] or slab_flat.target_offset != self.offs[self.rank]:
            raise ValueError("slab_flat is not this rank's slab of the partition")
        self.n_target = sum(self.sizes)
        self.n_source = slab_flat.n_cols
        self._owns_engine = engine is None
        self.engine = engine if engine is not None else Engine(slab_flat, device=device, **options)
        self.device = torch.device("cuda", self.engine.device)
        self.is_complex = self.engine.is_complex
        self.np_dtype = self.engine.np_dtype
        self.torch_dtype = torch.complex128 if self.is_complex else torch.float64
        self.es = 16 if self.is_complex else 8
        if slab_flat.target_permutation is None or len(slab_flat.target_permutation) != self.n_target:
            raise ValueError("slab_flat must carry the full target / source permutations")
        self.perm_t = torch.from_numpy(np.ascontiguousarray(slab_flat.target_permutation, np.int32)).to(self.device)
        self.perm_s = torch.from_numpy(np.ascontiguousarray(slab_flat.source_permutation, np.int32)).to(self.device)
        self._x = {}  # mu -> (x_cluster, y_cluster) replicated device buffers
        self._pin = {}
        self.fused = False
        self._peer_failed = 0
        self.split_input_upload = True  # `matvec` on several GPUs: each rank uploads 1/P of x, NVLink exchange (needs identical x on all ranks)
        self._ybuf = None
        self._t_slab = None  # slab-sized device temporary of the transposed global product
        # `matvec_cluster_transposed_` at world > 1: "nccl" = one all-reduce of the full-length contributions (measured faster: 1.42
        # vs 1.53 ms at N=1M on 2 GPUs, profiles/r02_bench_N1M_transposed.json); "peer" = rank-ordered reduce-scatter + all-gather
        # on the peer regions (no NCCL call, the same bits on every rank and run)
        self.transposed_exchange = "nccl"
        if fused_allgather is None:
            fused_allgather = self.world > 1
        if fused_allgather:
            self._setup_fused_allgather()

    def _setup_fused_allgather(self):
        """peer-mapped replicated output buffers (CUDA IPC over NVLink) for `matvec_cluster_fused_` (include/hmb200.h,
        hmb200_peer_*).  Falls back to the NCCL all-gather with a warning when IPC mapping is not available."""
        import ctypes
        import warnings

        import torch
        import torch.distributed as dist

        lib = _lib.lib()
        handle = (ctypes.c_ubyte * 64)()
        ok = lib.hmb200_peer_create(self.engine._h, self.world, self.rank, self.n_target, self.offs[self.rank], handle) == 0
        err = "" if ok else _lib.last_error()
        handles = [bytes(handle) if ok else None]
        if self.world > 1:
            handles = [None] * self.world
            dist.all_gather_object(handles, bytes(handle) if ok else None, group=self.group)
        if ok and all(hd is not None for hd in handles):
            blob = (ctypes.c_ubyte * (64 * self.world)).from_buffer_copy(b"".join(handles))
            ok = lib.hmb200_peer_attach(self.engine._h, blob) == 0
            err = "" if ok else _lib.last_error()
        else:
            ok = False
        flags = [ok]
        if self.world > 1:
            flags = [None] * self.world
            dist.all_gather_object(flags, ok, group=self.group)
        if not all(flags):
            warnings.warn(f"hmb200: fused all-gather unavailable ({err or 'a peer could not map the shared region'}); using NCCL")
            return
        bufs = []
        for i in range(2):
            ptr = ctypes.c_void_p()
            _lib.check(lib.hmb200_peer_buffer(self.engine._h, i, ctypes.byref(ptr)))
            bufs.append(_device_tensor(ptr.value, self.n_target, self.is_complex, self.device))
        self._ybuf = {b.data_ptr(): b for b in bufs}
        self.fused = True
        if self.world > 1 and self.n_source == self.n_target:
            self._setup_shared_result()

    def _setup_shared_result(self):
        """One POSIX-shared, page-locked host buffer (two alternating result vectors + one flag line per rank) for the
        replicated result of `matvec` on the ranks of one node: every rank downloads only ITS 1/P chunk of the user-numbered
        result into it and a host-side flag tells the others, so a product moves n (not P*n) elements over PCIe.  The
        reference's contract (src/htool/distributed_operator/distributed_operator.hpp:23-37: every rank returns the full
        vector) is kept: the buffer is the same physical memory in every process."""
        import warnings
        from multiprocessing import resource_tracker, shared_memory

        import torch
        import torch.distributed as dist

        self._shm = None
        buf_bytes = (self.n_target * self.es + 4095) // 4096 * 4096
        nbytes = 4096 + 2 * buf_bytes
        name, shm, base = [None], None, None
        if self.rank == 0:
            try:
                shm = shared_memory.SharedMemory(create=True, size=nbytes)
                np.frombuffer(shm.buf, dtype=np.uint8, count=4096)[:] = 0
                name[0] = shm.name
            except Exception as e:  # no /dev/shm: keep the per-rank full download
                warnings.warn(f"hmb200: shared result buffer unavailable ({e}); every rank downloads the full vector")
        dist.broadcast_object_list(name, src=dist.get_global_rank(self.group, 0) if self.group is not None else 0, group=self.group)
        ok = name[0] is not None
        if ok and self.rank != 0:
            try:
                shm = shared_memory.SharedMemory(name=name[0])
                try:  # the attaching processes must not unlink the segment at exit (3.12 registers them with the tracker)
                    resource_tracker.unregister(shm._name, "shared_memory")
                except Exception:
                    pass
            except Exception:
                ok = False
        if ok:
            base = np.frombuffer(shm.buf, dtype=np.uint8)
            with torch.cuda.device(self.device):
                ok = int(torch.cuda.cudart().cudaHostRegister(base.ctypes.data, nbytes, 0)) == 0
        oks = [None] * self.world
        dist.all_gather_object(oks, bool(ok), group=self.group)
        if not all(oks):  # a rank could not attach or page-lock: nobody keeps the segment (rank 0 removes the name)
            if ok:
                torch.cuda.cudart().cudaHostUnregister(base.ctypes.data)
            base = None  # a live view would make close() raise BufferError
            if shm is not None:
                try:
                    shm.close()
                    if self.rank == 0:
                        shm.unlink()
                except Exception as e:
                    warnings.warn(f"hmb200: could not release the shared result segment ({e})")
            return
        self._shm, self._shm_base, self._shm_calls = shm, base, 0
        self._shm_flags = np.frombuffer(shm.buf, dtype=np.int64, count=512)[::8][:self.world]  # one 64-byte line per rank
        self._shm_y = [np.frombuffer(shm.buf, dtype=self.np_dtype, count=self.n_target, offset=4096 + b * buf_bytes) for b in range(2)]
        self._shm_y_t = [torch.from_numpy(a) for a in self._shm_y]
        inv = torch.empty(self.n_target, dtype=torch.int32, device=self.device)
        inv[self.perm_t.long()] = torch.arange(self.n_target, dtype=torch.int32, device=self.device)
        self._inv_perm_t = inv  # y_user[i] = y_cluster[inv[i]]
        self.shared_result_timeout_s = 120.0
        # the whole chunked product as one call / one cached CUDA graph (hmb200_matvec_user_chunked_*)
        ps = np.ascontiguousarray(self.perm_s.cpu().numpy(), np.int32)
        it = np.ascontiguousarray(inv.cpu().numpy(), np.int32)
        _lib.check(_lib.lib().hmb200_set_global_permutations(self.engine._h, ps.ctypes.data, len(ps), it.ctypes.data, len(it)))
        self._pin_x_chunk = torch.empty(self.n_source, dtype=self.torch_dtype).pin_memory()

    @property
    def shape(self):
        return (self.n_target, self.n_source)

    # ---- device-resident products in cluster numbering -------------------------------------------------------
    def matvec_cluster_(self, x_cluster, y_cluster):
        """y_cluster (replicated, length n_target) = A x_cluster; both CUDA tensors; slab product + all-gather."""
        if self.fused:
            y_cluster.copy_(self.matvec_cluster_fused_(x_cluster))
            return y_cluster
        o, s = self.offs[self.rank], self.sizes[self.rank]
        self.engine.add_vector_product("N", 1.0, x_cluster, 0.0, y_cluster[o:o + s])
        return allgather_slabs_(y_cluster, self.offs, self.sizes, 1, self.group)

    def matvec_cluster_fused_(self, x_cluster):
        """returns the replicated y = A x (cluster numbering) as a view of one of the operator's two peer-mapped buffers: the
        slab product stores its rows into every rank's buffer from the phase-2 epilogue (no NCCL call).  Collective; the
        returned tensor stays valid until the product after the next one."""
        import ctypes

        if not self.fused:
            raise RuntimeError("fused all-gather is not set up on this operator")
        if self._peer_failed:
            self.check_peers()
        eng = self.engine
        px, _ = eng._ptr(x_cluster, self.n_source, False)
        out = ctypes.c_void_p()
        one = (ctypes.c_double * 2)(1.0, 0.0)
        fn = _lib.lib().hmb200_vector_product_allgather_z if self.is_complex else _lib.lib().hmb200_vector_product_allgather_d
        _lib.check(fn(eng._h, one, px, ctypes.byref(out)))
        return self._ybuf[out.value]

    def peer_allgather_(self, local_slab):
        """replicated vector (view of one of the operator's two peer-mapped buffers) whose slabs are the ranks' `local_slab`
        tensors: a copy kernel stores the slab into every rank's buffer over NVLink, epoch flags, wait (hmb200_peer_allgather).
        Collective; the MPI_Allgatherv of add_distributed_operator_vector_product_local_to_local.hpp:44."""
        import ctypes

        if not self.fused:
            raise RuntimeError("fused all-gather is not set up on this operator")
        ptr, _ = self.engine._ptr(local_slab, self.sizes[self.rank], False)
        out = ctypes.c_void_p()
        _lib.check(_lib.lib().hmb200_peer_allgather(self.engine._h, ptr, ctypes.byref(out)))
        return self._ybuf[out.value]

    def _peer_allgather_range(self, part, offset, count):
        import ctypes

        out = ctypes.c_void_p()
        self.engine.set_stream(__import__("torch").cuda.current_stream(self.device).cuda_stream)
        _lib.check(_lib.lib().hmb200_peer_allgather_range(self.engine._h, part.data_ptr(), int(offset), int(count), ctypes.byref(out)))
        return self._ybuf[out.value]

    PEER_ALLREDUCE_MAX = 512  # doubles (include/hmb200.h)

    def peer_allreduce_(self, t):
        """in-place sum over the ranks of a small float64 / complex128 CUDA tensor, added in rank order on every rank
        (hmb200_peer_allreduce_f64); collective"""
        import torch

        if not self.fused:
            raise RuntimeError("fused all-gather is not set up on this operator")
        n = t.numel() * (2 if t.is_complex() else 1)
        if not t.is_contiguous() or t.dtype not in (torch.float64, torch.complex128) or n > self.PEER_ALLREDUCE_MAX:
            raise ValueError("contiguous float64 / complex128 tensor of at most 512 doubles expected")
        self.engine.set_stream(torch.cuda.current_stream(self.device).cuda_stream)
        _lib.check(_lib.lib().hmb200_peer_allreduce_f64(self.engine._h, t.data_ptr(), n))
        return t

    def peer_exchange_sum_(self, x_local, y_local, trans: str = "T", alpha=1.0, beta=0.0):
        """y_local = beta y_local + sum over the ranks (ascending) of the slab of alpha op(A_q) x_q, op = 'T' / 'C': every rank's
        full-length contribution stays in its own peer-mapped buffer and the slabs are pulled over NVLink and added in rank
        order by one kernel (hmb200_peer_exchange_sum) -- the MPI_Alltoallv + axpy loop of
        add_distributed_operator_vector_product_local_to_local.hpp:60-87 without NCCL.  Collective."""
        import ctypes

        import torch

        if not self.fused:
            raise RuntimeError("fused all-gather is not set up on this operator")
        if self._peer_failed:
            self.check_peers()
        eng, s = self.engine, self.sizes[self.rank]
        eng.set_stream(torch.cuda.current_stream(self.device).cuda_stream)
        px, _ = eng._ptr(x_local, s, False)
        py, _ = eng._ptr(y_local, s, True)
        a = (ctypes.c_double * 2)(complex(alpha).real, complex(alpha).imag)
        b = (ctypes.c_double * 2)(complex(beta).real, complex(beta).imag)
        _lib.check(_lib.lib().hmb200_peer_exchange_sum(eng._h, trans.encode(), a, px, b, py))
        return y_local

    def peer_status(self, sync: bool = True) -> int:
        """0, or 1 + the rank whose epoch did not arrive within a wait kernel's bound (101 + rank: an all-reduce).  The word
        lives in mapped host memory; with sync=False it covers the waits that have already completed on the stream."""
        import ctypes

        if not self.fused:
            return 0
        if sync:
            import torch

            torch.cuda.current_stream(self.device).synchronize()
        err = ctypes.c_int(0)
        _lib.check(_lib.lib().hmb200_peer_status(self.engine._h, ctypes.byref(err)))
        return err.value

    def check_peers(self):
        """to be called right after a host-side synchronisation: raises if a device-side wait of the peer collectives gave up
        (a late or dead rank), instead of letting a partly gathered vector through.  The epochs are then out of step, so the
        operator refuses further collectives."""
        if not self.fused:
            return
        st = self.peer_status(sync=False)
        if st or self._peer_failed:
            self._peer_failed = self._peer_failed or st
            what = f"all-reduce, rank {self._peer_failed - 101}" if self._peer_failed > 100 else f"rank {self._peer_failed - 1}"
            raise RuntimeError(f"hmb200: a peer did not arrive within the wait bound ({what}; HMB200_PEER_TIMEOUT_S); "
                               "the replicated vectors are incomplete and this operator must be rebuilt")

    def matvec_cluster_transposed_(self, x_cluster, y_cluster, trans: str = "T"):
        """y_cluster (replicated, length n_source) = op(A) x_cluster with op = transpose ('T') or conjugate transpose ('C'):
        every rank applies op(A_p) to ITS slab of x and the full-length results are summed over the ranks -- the `trans != 'N'`
        branch of add_distributed_operator_vector_product_global_to_global.hpp:63-84 (MPI_Allreduce :78).  The engine must have
        been created with build_transpose=1."""
        import torch.distributed as dist

        if trans not in ("T", "C"):
            raise ValueError("trans must be 'T' or 'C'")
        o, s = self.offs[self.rank], self.sizes[self.rank]
        if self.world > 1 and self.fused and self.transposed_exchange == "peer" and self.n_target == self.n_source:
            # reduce-scatter (rank-ordered sums pulled over NVLink) + all-gather, both on the peer regions: no NCCL call, and
            # every rank holds the same bits
            import torch

            if self._t_slab is None:
                self._t_slab = torch.empty(s, dtype=self.torch_dtype, device=self.device)
            self.peer_exchange_sum_(x_cluster[o:o + s], self._t_slab, trans, 1.0, 0.0)
            y_cluster.copy_(self.peer_allgather_(self._t_slab))
            return y_cluster
        self.engine.add_vector_product(trans, 1.0, x_cluster[o:o + s], 0.0, y_cluster)
        if self.world > 1:
            import torch

            dist.all_reduce(torch.view_as_real(y_cluster) if y_cluster.is_complex() else y_cluster, group=self.group)
        return y_cluster

    def matvec_local_to_local_(self, x_local, y_local, trans: str = "N", alpha=1.0, beta=0.0):
        """y_local = beta y_local + alpha op(A) x on LOCAL slabs (device tensors, cluster numbering): the solver-side driver
        internal_add_distributed_operator_vector_product_local_to_local (…vector_product_local_to_local.hpp:19-89).
        'N': all-gather of the input slabs (:44), then the rank's global-to-local product.  'T' / 'C': op(A_p) applied to the
        rank's slab gives a full-length contribution; the slabs are exchanged and summed in rank order (:60-87).  Square
        operators with one partition for both sides (what the solver uses); 'T'/'C' need build_transpose=1."""
        import torch

        if self.n_target != self.n_source:
            raise RuntimeError("the local-to-local driver needs a square operator")
        o, s = self.offs[self.rank], self.sizes[self.rank]
        if trans == "N":
            if self.fused:
                xg = self.peer_allgather_(x_local)
            else:
                xg = self._buffers(1)[0][1]
                xg[o:o + s].copy_(x_local)
                allgather_slabs_(xg, self.offs, self.sizes, 1, self.group)
            self.engine.add_vector_product("N", alpha, xg, beta, y_local)
            return y_local
        if trans not in ("T", "C"):
            raise ValueError("trans must be 'N', 'T' or 'C'")
        if self.fused:
            return self.peer_exchange_sum_(x_local, y_local, trans, alpha, beta)
        full = self._buffers(1)[0][2]
        self.engine.add_vector_product(trans, alpha, x_local, 0.0, full)
        return exchange_sum_slabs_(full, self.offs, self.sizes, y_local, beta, self.group if torch.distributed.is_initialized() else None)

    def _setup_fused_matmat(self, mu: int):
        """replicated (n_target x mu) blocks in a second peer-mapped region (hmb200_peer_mm_create / _attach); collective"""
        import ctypes

        import torch.distributed as dist

        lib = _lib.lib()
        if getattr(self, "_mm_mu", 0):
            # another width: no owner may free its blocks while a peer still stores into them or has them open, so every rank
            # drains its stream and closes its imports first, and the ranks meet before anyone re-creates (the exchange of the
            # status is that barrier).  Views handed out for the old width (matmat_cluster_fused_) are dead from here on.
            self._mm_ybuf, self._mm_mu = {}, 0
            ok = lib.hmb200_peer_mm_detach(self.engine._h) == 0
            err = "" if ok else _lib.last_error()
            oks = [None] * self.world
            dist.all_gather_object(oks, ok, group=self.group)
            if not all(oks):
                raise RuntimeError(f"hmb200: the replicated multi-RHS blocks of the previous width could not be released ({err or 'a peer failed'})")
        handle = (ctypes.c_ubyte * 64)()
        ok = lib.hmb200_peer_mm_create(self.engine._h, int(mu), handle) == 0
        err = "" if ok else _lib.last_error()
        handles = [None] * self.world
        dist.all_gather_object(handles, bytes(handle) if ok else None, group=self.group)
        if ok and all(hd is not None for hd in handles):
            blob = (ctypes.c_ubyte * (64 * self.world)).from_buffer_copy(b"".join(handles))
            ok = lib.hmb200_peer_mm_attach(self.engine._h, blob) == 0
            err = "" if ok else _lib.last_error()
        else:
            ok = False
        flags = [None] * self.world
        dist.all_gather_object(flags, ok, group=self.group)
        if not all(flags):
            raise RuntimeError(f"hmb200: replicated multi-RHS blocks could not be shared between the ranks ({err or 'a peer failed'})")
        bufs = []
        for i in range(2):
            ptr = ctypes.c_void_p()
            _lib.check(lib.hmb200_peer_mm_buffer(self.engine._h, i, ctypes.byref(ptr)))
            bufs.append(_device_tensor(ptr.value, self.n_target * mu, self.is_complex, self.device))
        self._mm_ybuf = {b.data_ptr(): b for b in bufs}
        self._mm_mu = int(mu)

    def matmat_cluster_fused_(self, X_cluster, mu: int):
        """returns the replicated Y = A X (row-major n_target x mu, cluster numbering) as a view of one of the operator's two
        peer-mapped blocks: the multi-RHS slab product stores every bin's rows into the block of every rank from its epilogue
        (no NCCL call).  Collective; the returned tensor stays valid until the product after the next -- or until a product with
        another `mu`, which releases and re-creates the blocks on every rank."""
        import ctypes

        if not self.fused:
            raise RuntimeError("fused all-gather is not set up on this operator")
        if self._peer_failed:
            self.check_peers()
        if getattr(self, "_mm_mu", 0) != int(mu):
            self._setup_fused_matmat(mu)
        px, _ = self.engine._ptr(X_cluster, self.n_source * mu, False)
        out = ctypes.c_void_p()
        one = (ctypes.c_double * 2)(1.0, 0.0)
        fn = _lib.lib().hmb200_matrix_product_allgather_z if self.is_complex else _lib.lib().hmb200_matrix_product_allgather_d
        _lib.check(fn(self.engine._h, one, px, int(mu), ctypes.byref(out)))
        return self._mm_ybuf[out.value]

    def matmat_cluster_(self, X_cluster, Y_cluster, mu: int):
        """row-major (n x mu) blocks (add_distributed_operator_matrix_product_row_major_global_to_global.hpp:18-84): slab product +
        all-gather of the rows, fused into the kernel when the peer regions are set up.  Returns the replicated block (a view of
        the operator's own buffer on the fused path, `Y_cluster` otherwise)."""
        if self.fused and self.world > 1 and self.engine.info["mm_stream_bytes"] > 0:
            return self.matmat_cluster_fused_(X_cluster, mu)
        o, s = self.offs[self.rank], self.sizes[self.rank]
        self.engine.add_matrix_product_row_major("N", 1.0, X_cluster, 0.0, Y_cluster.view(-1)[o * mu:(o + s) * mu], mu)
        return allgather_slabs_(Y_cluster, self.offs, self.sizes, mu, self.group)

    def internal_sub_vector_product_global_to_local(self, x, offset: int):
        """y_local = A_p[:, offset:offset+len(x)] x : the window product GenEO's coarse-space assembly uses
        (src/htool/distributed_operator/distributed_operator.hpp:67-78 ->
        internal_add_distributed_operator_vector_sub_product_global_to_local); x in source CLUSTER numbering, host array;
        returns this rank's slab (length local target size) like the reference."""
        x = np.ascontiguousarray(x, dtype=self.np_dtype)
        if x.ndim != 1:
            raise RuntimeError("Wrong dimension for DistributedOperator-vector product")
        out = np.zeros(self.sizes[self.rank], dtype=self.np_dtype)
        self.engine.add_sub_matrix_product_to_local(x, out, 1, int(offset), x.shape[0])
        return out

    # ---- user-numbering products on host arrays (what the reference's Python API does) ------------------------
    def _permute(self, gather: bool, perm, n, src, dst, mu):
        import torch

        fn = _lib.lib().hmb200_permute_gather if gather else _lib.lib().hmb200_permute_scatter
        _lib.check(fn(perm.data_ptr(), 0, n, src.data_ptr(), dst.data_ptr(), self.es, mu, _lib.DEVICE,
                      torch.cuda.current_stream(self.device).cuda_stream))

    def _buffers(self, mu):
        import torch

        if mu not in self._x:
            mk = lambda n: torch.empty(n * mu, dtype=self.torch_dtype, device=self.device)  # noqa: E731
            self._x[mu] = (mk(self.n_source), mk(self.n_source), mk(self.n_target), mk(self.n_target))
            self._pin[mu] = (torch.empty(self.n_source * mu, dtype=self.torch_dtype).pin_memory(),
                             torch.empty(self.n_target * mu, dtype=self.torch_dtype).pin_memory())
        return self._x[mu], self._pin[mu]

    def __mul__(self, x):
        return self.matvec(x)

    def matvec(self, x, out=None):
        """`A * x` (user numbering, host arrays); `out`, when given, receives the result in place.  Page-locked numpy buffers
        (`torch.Tensor.pin_memory().numpy()`) are DMA'd directly, pageable ones go through the operator's pinned staging.
        `out="shared"` returns the operator's own result buffer without a copy: on several ranks of one node that is ONE
        page-locked POSIX-shared vector every rank sees (each rank uploads 1/P of x and downloads 1/P of y), valid until the
        call after the next."""
        import torch

        x = np.asarray(x)
        if x.ndim != 1:
            raise RuntimeError("Wrong dimension for HMatrix-vector product")  # distributed_operator.hpp:25-27
        if x.shape[0] != self.n_source:
            raise RuntimeError("Wrong size for HMatrix-vector product")
        if isinstance(out, str):
            bad = out != "shared"
        else:
            bad = out is not None and (not isinstance(out, np.ndarray) or out.dtype != self.np_dtype or out.shape != (self.n_target,) or not out.flags.c_contiguous)
        if bad:
            raise ValueError("out must be a contiguous numpy array of the operator's dtype and target size, or the string 'shared'")
        if self.fused and self.world > 1 and getattr(self, "_shm", None) is not None and self.split_input_upload:
            return self._matvec_chunked(x, out)
        (xu, xc, yc, yu), (pin_x, pin_y) = self._buffers(1)
        with torch.cuda.device(self.device):
            src = self._as_pinned_tensor(x)
            if self.fused and self.world > 1 and self.split_input_upload and self.n_source == self.n_target:
                # the input is replicated on the hosts (the contract of the global-to-global product): every rank uploads
                # only ITS 1/P chunk of the user-numbered vector and the chunks are exchanged over NVLink
                c0, c1 = self.rank * self.n_source // self.world, (self.rank + 1) * self.n_source // self.world
                if src is None:
                    pin_x.numpy()[c0:c1] = x[c0:c1].astype(self.np_dtype, copy=False)
                    src = pin_x
                xu[c0:c1].copy_(src[c0:c1], non_blocking=True)
                xu = self._peer_allgather_range(xu[c0:c1], c0, c1 - c0)
            else:
                if src is None:
                    pin_x.numpy()[:] = x.astype(self.np_dtype, copy=False)
                    src = pin_x
                xu.copy_(src, non_blocking=True)
            self._permute(True, self.perm_s, self.n_source, xu, xc, 1)     # global_to_root_cluster (:108)
            yc = self.matvec_cluster_fused_(xc) if self.fused else self.matvec_cluster_(xc, yc)
            if self.fused and self.world > 1 and getattr(self, "_shm", None) is not None:
                return self._finish_shared(yc, yu, out)
            self._permute(False, self.perm_t, self.n_target, yc, yu, 1)    # root_cluster_to_global (:117)
            dst = self._as_pinned_tensor(out) if isinstance(out, np.ndarray) else None
            (dst if dst is not None else pin_y).copy_(yu, non_blocking=True)
            torch.cuda.current_stream(self.device).synchronize()
        self.check_peers()
        if out is None or isinstance(out, str):
            return pin_y.numpy().copy() if out is None else pin_y.numpy()
        if dst is None:
            out[:] = pin_y.numpy()
        return out

    def _matvec_chunked(self, x, out):
        """the distributed user-numbering product as ONE library call per rank (one cached CUDA graph: H2D of this rank's 1/P
        chunk of x, all-gather of the chunks over NVLink, permutation, slab product with the fused all-gather, permutation of this
        rank's 1/P chunk of y, D2H into the node-shared page-locked result vector), then the host-side flags."""
        k = self._shm_calls
        self._shm_calls += 1
        b = k & 1
        n, es = self.n_source, self.es
        c0, c1 = self.rank * n // self.world, (self.rank + 1) * n // self.world
        src = self._as_pinned_tensor(x)
        if src is None:  # pageable input: stage this rank's chunk
            self._pin_x_chunk.numpy()[c0:c1] = x[c0:c1].astype(self.np_dtype, copy=False)
            src = self._pin_x_chunk
        res = self._shm_y[b]
        fn = _lib.lib().hmb200_matvec_user_chunked_z if self.is_complex else _lib.lib().hmb200_matvec_user_chunked_d
        self.engine.set_stream(__import__("torch").cuda.current_stream(self.device).cuda_stream)
        st = fn(self.engine._h, src.data_ptr() + es * c0, c0, c1 - c0, res.ctypes.data + es * c0, c0, c1 - c0)
        if st != 0:
            self._peer_failed = self._peer_failed or self.peer_status(sync=False) or 1
            _lib.check(st)
        self._publish_and_wait(self._shm_flags, k + 1)
        if isinstance(out, str):
            return res
        if out is None:
            return res.copy()
        np.copyto(out, res)
        return out

    def _finish_shared(self, yc, yu, out):
        """root_cluster_to_global of THIS rank's 1/P chunk of the user-numbered result, one D2H copy of that chunk into the
        node-shared page-locked buffer, host-side flag, wait for the other ranks' flags.  Two result buffers alternate: a
        rank can be at most one call ahead of the others, so a returned vector stays intact until the call after the next."""
        import torch

        k = self._shm_calls
        self._shm_calls += 1
        b = k & 1
        c0, c1 = self.rank * self.n_target // self.world, (self.rank + 1) * self.n_target // self.world
        _lib.check(_lib.lib().hmb200_permute_gather(self._inv_perm_t.data_ptr() + 4 * c0, 0, c1 - c0, yc.data_ptr(), yu.data_ptr() + self.es * c0,
                                                    self.es, 1, _lib.DEVICE, torch.cuda.current_stream(self.device).cuda_stream))
        self._shm_y_t[b][c0:c1].copy_(yu[c0:c1], non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        self.check_peers()
        self._publish_and_wait(self._shm_flags, k + 1)
        res = self._shm_y[b]
        if isinstance(out, str):  # out="shared": the node-shared buffer itself (valid until the call after the next)
            return res
        if out is None:
            return res.copy()
        np.copyto(out, res)
        return out

    def _publish_and_wait(self, fl, value):
        """this rank's part of the shared result is in place: publish `value` in its flag word (release store) and wait until every
        rank has done the same (acquire loads with back-off, in C with the GIL released: hmb200_host_flag_publish / _flags_wait)"""
        import ctypes

        lib = _lib.lib()
        stride = fl.strides[0] // 8
        lib.hmb200_host_flag_publish(fl.ctypes.data + fl.strides[0] * self.rank, int(value))
        laggard = ctypes.c_int(-1)
        st = lib.hmb200_host_flags_wait(fl.ctypes.data, stride, len(fl), int(value), float(self.shared_result_timeout_s), ctypes.byref(laggard))
        if st != 0:
            raise RuntimeError(f"hmb200: rank {laggard.value} did not deliver its part of the result within {self.shared_result_timeout_s} s")

    def _as_pinned_tensor(self, a):
        """torch view of a page-locked, contiguous numpy array of the operator's dtype (None otherwise)"""
        import torch

        if not isinstance(a, np.ndarray) or a.dtype != self.np_dtype or not a.flags.c_contiguous:
            return None
        try:
            t = torch.from_numpy(a)
            return t if t.is_pinned() else None
        except Exception:
            return None

    def _setup_shared_block(self, mu: int):
        """node-shared page-locked result blocks (two, n_target x mu column-major) for `matmat`; collective.  Returns False (and
        keeps the per-rank full transfer) when the shared memory cannot be set up on every rank."""
        from multiprocessing import resource_tracker, shared_memory

        import torch
        import torch.distributed as dist

        old = getattr(self, "_shmb", None)
        if old is not None and old["mu"] == mu:
            return True
        self._release_shared_block()
        blk = (self.n_target * mu * self.es + 4095) // 4096 * 4096
        nbytes = 4096 + 2 * blk
        name, shm, base = [None], None, None
        if self.rank == 0:
            try:
                shm = shared_memory.SharedMemory(create=True, size=nbytes)
                np.frombuffer(shm.buf, dtype=np.uint8, count=4096)[:] = 0
                name[0] = shm.name
            except Exception:
                shm = None
        dist.broadcast_object_list(name, src=dist.get_global_rank(self.group, 0) if self.group is not None else 0, group=self.group)
        ok = name[0] is not None
        if ok and self.rank != 0:
            try:
                shm = shared_memory.SharedMemory(name=name[0])
                try:
                    resource_tracker.unregister(shm._name, "shared_memory")
                except Exception:
                    pass
            except Exception:
                ok = False
        if ok:
            base = np.frombuffer(shm.buf, dtype=np.uint8)
            with torch.cuda.device(self.device):
                ok = int(torch.cuda.cudart().cudaHostRegister(base.ctypes.data, nbytes, 0)) == 0
        oks = [None] * self.world
        dist.all_gather_object(oks, bool(ok), group=self.group)
        if not all(oks):
            if ok:
                torch.cuda.cudart().cudaHostUnregister(base.ctypes.data)
            base = None  # a live view would make close() raise BufferError
            if shm is not None:
                shm.close()
                if self.rank == 0:
                    shm.unlink()
            return False
        self._shmb = {"mu": mu, "shm": shm, "base": base, "calls": 0,
                      "flags": np.frombuffer(shm.buf, dtype=np.int64, count=512)[::8][:self.world],
                      "Y": [np.frombuffer(shm.buf, dtype=self.np_dtype, count=self.n_target * mu, offset=4096 + b * blk).reshape((self.n_target, mu), order="F") for b in range(2)],
                      "pin_x": None}
        return True

    def _release_shared_block(self):
        blk = getattr(self, "_shmb", None)
        if blk is None:
            return
        import torch

        try:
            torch.cuda.current_stream(self.device).synchronize()
            torch.cuda.cudart().cudaHostUnregister(blk["base"].ctypes.data)
        except Exception:
            pass
        shm = blk["shm"]
        blk["flags"] = blk["Y"] = blk["base"] = None
        self._shmb = None
        try:
            shm.close()
            if self.rank == 0:
                shm.unlink()
        except Exception:
            pass

    def matmat(self, X, out=None):
        """`A @ X` for a Fortran-ordered (n_source, mu) host block in user numbering.  On several ranks of one node (peer regions
        set up, square operator) every rank uploads its 1/P of the ROWS of X and downloads its 1/P of the rows of Y
        (hmb200_matmat_user_chunked_*); the result lands in a node-shared page-locked block that `out="shared"` returns without
        a copy (valid until the call after the next, or until a call with another number of columns: that releases the shared block
        and the peer-mapped device blocks on every rank, so copy what must outlive it).  Page-locked X is DMA'd directly, pageable X
        goes through a pinned chunk."""
        import torch

        X = np.asarray(X)
        if X.ndim != 2:
            raise RuntimeError("Wrong dimension for HMatrix-matrix product")
        if X.shape[0] != self.n_source:
            raise RuntimeError("Wrong size for HMatrix-matrix product")
        mu = int(X.shape[1])
        chunked = (self.fused and self.world > 1 and self.n_source == self.n_target and getattr(self, "_shm", None) is not None
                   and self.engine.info["mm_stream_bytes"] > 0 and mu > 1)
        if chunked:
            if getattr(self, "_mm_mu", 0) != mu:
                self._setup_fused_matmat(mu)
            chunked = self._setup_shared_block(mu)
        if not chunked:
            Y = self.__matmul__(X)
            if isinstance(out, np.ndarray):
                np.copyto(out, Y)
                return out
            return Y
        if self._peer_failed:
            self.check_peers()
        blk = self._shmb
        k = blk["calls"]
        blk["calls"] += 1
        n, es = self.n_source, self.es
        c0, c1 = self.rank * n // self.world, (self.rank + 1) * n // self.world
        fn = _lib.lib().hmb200_matmat_user_chunked_z if self.is_complex else _lib.lib().hmb200_matmat_user_chunked_d
        self.engine.set_stream(torch.cuda.current_stream(self.device).cuda_stream)
        Y = blk["Y"][k & 1]
        if X.dtype == self.np_dtype and X.flags.f_contiguous and self._is_pinned(X):
            xptr, ldx = X.ctypes.data, n  # page-locked block: strided DMA straight from it
        else:
            # pageable X: stage this rank's rows of every column in a pinned (c1 - c0) x mu block.  The library addresses element
            # (i, j) as base + i + j*ld and reads rows [c0, c1), so pass base = stage - c0 elements and ld = c1 - c0
            if blk["pin_x"] is None or blk["pin_x"].numel() < (c1 - c0) * mu:
                blk["pin_x"] = torch.empty((c1 - c0) * mu, dtype=self.torch_dtype).pin_memory()
            stage = blk["pin_x"].numpy()[:(c1 - c0) * mu].reshape((c1 - c0, mu), order="F")
            stage[...] = X[c0:c1, :]
            xptr, ldx = stage.ctypes.data - es * c0, c1 - c0
        st = fn(self.engine._h, xptr, ldx, c0, c1 - c0, Y.ctypes.data, n, c0, c1 - c0, mu)
        if st != 0:
            self._peer_failed = self._peer_failed or self.peer_status(sync=False) or 1
            _lib.check(st)
        self._publish_and_wait(blk["flags"], k + 1)
        if isinstance(out, str):
            return Y
        if out is None:
            return np.array(Y, order="F")
        np.copyto(out, Y)
        return out

    def _is_pinned(self, a) -> bool:
        import torch

        try:
            return bool(torch.from_numpy(a.reshape(-1, order="F") if a.ndim == 2 else a).is_pinned())
        except Exception:
            return False

    def __matmul__(self, X):
        """X: (n_source, mu) Fortran-ordered like the reference (distributed_operator.hpp:39-65); result F-ordered."""
        import torch

        X = np.asarray(X)
        if X.ndim != 2:
            raise RuntimeError("Wrong dimension for HMatrix-matrix product")
        if X.shape[0] != self.n_source:
            raise RuntimeError("Wrong size for HMatrix-matrix product")
        mu = X.shape[1]
        (xu, xc, yc, yu), (pin_x, pin_y) = self._buffers(mu)
        with torch.cuda.device(self.device):
            # host side of the reference does per-column permutation + transposes to row-major and back (:176-187, :202-204);
            # here both are one layout kernel each (hmb200_block_to_rows / hmb200_rows_to_block, permutation fused)
            pin_x.numpy()[:] = np.asfortranarray(X, dtype=self.np_dtype).reshape(-1, order="F")
            xu.copy_(pin_x, non_blocking=True)
            st = torch.cuda.current_stream(self.device).cuda_stream
            lib = _lib.lib()
            _lib.check(lib.hmb200_block_to_rows(self.perm_s.data_ptr(), self.n_source, mu, xu.data_ptr(), self.n_source, xc.data_ptr(), self.es, st))
            yr = self.matmat_cluster_(xc, yc.view(self.n_target, mu), mu)
            _lib.check(lib.hmb200_rows_to_block(self.perm_t.data_ptr(), self.n_target, mu, yr.data_ptr(), yu.data_ptr(), self.n_target, self.es, st))
            pin_y.copy_(yu, non_blocking=True)
            torch.cuda.current_stream(self.device).synchronize()
        self.check_peers()
        return np.array(pin_y.numpy().reshape((self.n_target, mu), order="F"), order="F")

    def close(self):
        self._release_shared_block()
        shm = getattr(self, "_shm", None)
        if shm is not None:
            import torch

            try:
                torch.cuda.current_stream(self.device).synchronize()
                torch.cuda.cudart().cudaHostUnregister(self._shm_base.ctypes.data)
            except Exception:
                pass
            self._shm_flags = self._shm_y = self._shm_y_t = self._shm_base = None
            self._shm = None
            try:
                shm.close()
                if self.rank == 0:
                    shm.unlink()
            except Exception:
                pass
        if self._owns_engine:
            self.engine.close()

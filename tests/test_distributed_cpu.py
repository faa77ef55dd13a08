"""Host-side logic of the block-row DistributedOperator path on CPU: world_size-2 `gloo` process group.

The slab product itself is played by the ORACLE here (tests only); what is checked is everything around it that the GPU
path shares: the reference's row-slab assembly (target_partition_number), slab flattening (offsets relative to the slab
root), partition displacements, the in-place all-gather of unequal / equal slabs and the two global permutations --
restating add_distributed_operator_vector_product_global_to_global.hpp:18-118.
"""
import os
import socket

import numpy as np
import pytest

import helpers
from helpers import TOL, rel_err

pytestmark = pytest.mark.skipif(not helpers.bridge_available(), reason="bridge not built")


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, kernel, n, mu, q):
    import torch
    import torch.distributed as dist

    from htool_python_b200 import FlatHMatrix, bridge
    from htool_python_b200.distributed import allgather_slabs_, partition_displacements
    from oracle import flat_oracle as fo

    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(5)
        g = rng.standard_normal((3, n))
        pts = np.asfortranarray(g / np.linalg.norm(g, axis=0))
        tree = bridge.create_cluster_tree(pts, 2, world, 32)
        offs, sizes = partition_displacements(tree.partition)
        assert sum(sizes) == n and len(offs) == world
        Hp = bridge.build_hmatrix(kernel, tree, epsilon=1e-4, eta=10.0, target_partition_number=rank, partition_number_for_symmetry=rank, wavenumber=3.0)
        slab = FlatHMatrix.from_htool(Hp)
        assert slab.n_rows == sizes[rank] and slab.target_offset == offs[rank] and slab.n_cols == n
        dt = slab.np_dtype
        X = helpers.rand_vec(rng, n, dt, mu)
        perm = slab.source_permutation
        Xc = np.ascontiguousarray(X[perm])                              # global_to_root_cluster (:108)
        Yc = torch.zeros(n * mu, dtype=torch.from_numpy(np.zeros(1, dt)).dtype)
        mine = np.zeros((sizes[rank], mu), dtype=dt)
        fo.add_matrix_product_row_major(slab, "N", 1.0, Xc, 0.0, mine, mu)  # stand-in for the engine's slab product
        Yc.view(n, mu)[offs[rank]:offs[rank] + sizes[rank]] = torch.from_numpy(mine)
        allgather_slabs_(Yc, offs, sizes, mu)                           # MPI_Allgatherv (:76)
        Y = np.empty((n, mu), dtype=dt)
        Y[slab.target_permutation] = Yc.view(n, mu).numpy()             # root_cluster_to_global (:117)
        # transposed local-to-local driver (…vector_product_local_to_local.hpp:60-87): op(A_p) x_p is a full-length
        # contribution; slabs exchanged point-to-point and summed in rank order on top of beta * out
        from htool_python_b200.distributed import exchange_sum_slabs_

        for tr in ["T"] + (["C"] if dt == np.complex128 else []):
            x_loc = helpers.rand_vec(np.random.default_rng(100 + rank), sizes[rank], dt)
            contrib = np.zeros(n, dt)
            fo.add_vector_product(slab, tr, 1.5, x_loc, 0.0, contrib)
            out0 = helpers.rand_vec(np.random.default_rng(200 + rank), sizes[rank], dt)
            beta = 0.5
            out_t = torch.from_numpy(out0.copy())
            exchange_sum_slabs_(torch.from_numpy(contrib), offs, sizes, out_t, beta)
            everyone = [None] * world
            dist.all_gather_object(everyone, contrib)
            expect = beta * out0
            for r in range(world):
                expect = expect + everyone[r][offs[rank]:offs[rank] + sizes[rank]]
            assert np.array_equal(out_t.numpy(), expect), tr
            out_z = torch.from_numpy(np.full(sizes[rank], np.nan).astype(dt))
            exchange_sum_slabs_(torch.from_numpy(contrib), offs, sizes, out_z, 0.0)   # beta = 0 must not read `out`
            assert np.array_equal(out_z.numpy(), expect - beta * out0 + 0) or rel_err(out_z.numpy(), expect - beta * out0) < 1e-13
        if rank == 0:
            # the same operator assembled in one piece by the reference, product by the oracle in user numbering
            tree1 = bridge.create_cluster_tree(pts, 2, 1, 32)
            H = bridge.build_hmatrix(kernel, tree1, epsilon=1e-4, eta=10.0, wavenumber=3.0)
            full = FlatHMatrix.from_htool(H)
            Y_ref = np.stack([fo.matvec_user(full, np.ascontiguousarray(X[:, j])) for j in range(mu)], axis=1)
            q.put(("ok", rel_err(Y, Y_ref), [int(s) for s in sizes]))
    except Exception as e:  # pragma: no cover
        import traceback

        q.put(("error", traceback.format_exc(), None))
        raise
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("kernel,n,mu", [("laplace_reg", 3000, 1), ("complex_reg", 2001, 3)])
def test_block_row_product_two_ranks_gloo(kernel, n, mu):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, kernel, n, mu, q)) for r in range(2)]
    for p in procs:
        p.start()
    status, err, sizes = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
    assert status == "ok", err
    # two independently compressed operators (slabs vs one piece): agreement to the compression accuracy, not 1e-12
    assert err < 5e-4, err
    assert all(p.exitcode == 0 for p in procs)


def test_partition_displacements_rejects_gaps():
    from htool_python_b200.distributed import partition_displacements

    assert partition_displacements([(0, 3), (3, 4)]) == ([0, 3], [3, 4])
    with pytest.raises(ValueError):
        partition_displacements([(0, 3), (4, 4)])


def _gmres_worker(rank, world, port, n, q):
    """Host logic of the distributed solver on two `gloo` ranks: row-slab bookkeeping, input all-gather, all-reduced
    dot products; the slab product and the vector algebra are numpy stand-ins (tests/numpy_space.py)."""
    import torch.distributed as dist

    from htool_python_b200 import FlatHMatrix, bridge
    from htool_python_b200.distributed import partition_displacements
    from htool_python_b200.solver import B200Solver
    from numpy_space import NumpyKrylovSpace
    from oracle import flat_oracle as fo
    from oracle import gmres_oracle as go

    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(11)
        g = rng.standard_normal((3, n))
        pts = np.asfortranarray(g / np.linalg.norm(g, axis=0))
        tree = bridge.create_cluster_tree(pts, 2, world, 32)
        offs, sizes = partition_displacements(tree.partition)
        Hp = bridge.build_hmatrix("laplace_reg", tree, epsilon=1e-6, eta=10.0, target_partition_number=rank, partition_number_for_symmetry=rank)
        slab = FlatHMatrix.from_htool(Hp)
        b = rng.random(n)
        args = "-hpddm_krylov_method gmres -hpddm_tol 1e-6 -hpddm_gmres_restart 25 -hpddm_max_it 150 -hpddm_schwarz_method none -hpddm_variant right"
        solver = B200Solver(None, space=NumpyKrylovSpace(slab, offs, sizes, rank, world))
        x = np.zeros(n)
        it = solver.solve(x, b, args)
        # every rank must hold the same solution and iteration count; compare with the one-process oracle GMRES whose
        # operator is the concatenation of the same slabs
        # a block of 3 right-hand sides in lockstep (one row-major block product + one all-reduce of all columns per step)
        B = np.asfortranarray(rng.random((n, 3)))
        XB = np.zeros_like(B, order="F")
        itB = solver.solve(XB, B, args)
        callsB = solver.last_info["operator_calls"]
        slabs = [None] * world
        dist.all_gather_object(slabs, (it, x, itB, XB))
        if rank == 0:
            ps, pt = slab.source_permutation, slab.target_permutation
            flats = [slab] + [FlatHMatrix.from_htool(bridge.build_hmatrix("laplace_reg", tree, epsilon=1e-6, eta=10.0, target_partition_number=p, partition_number_for_symmetry=p)) for p in range(1, world)]

            def matvec(v):
                vc = np.ascontiguousarray(v[ps])
                yc = np.zeros(n)
                for p, f in enumerate(flats):
                    fo.add_vector_product(f, "N", 1.0, vc, 0.0, yc[offs[p]:offs[p] + sizes[p]])
                y = np.empty(n)
                y[pt] = yc
                return y

            x_ref, it_ref, info = go.gmres(matvec, b, **go.parse_hpddm_args(args))
            same = all(s[0] == it and np.array_equal(s[1], x) for s in slabs)
            XB_ref, itB_ref, infoB = go.gmres_block(lambda Y: np.stack([matvec(Y[:, c]) for c in range(Y.shape[1])], axis=1), B, **go.parse_hpddm_args(args))
            sameB = all(s[2] == itB and np.array_equal(s[3], XB) for s in slabs)
            q.put(("ok", (it, it_ref, same and sameB, rel_err(x, x_ref), np.linalg.norm(matvec(x) - b) / np.linalg.norm(b),
                          (itB, callsB), (itB_ref, infoB["operator_calls"]), rel_err(XB, XB_ref))))
    except Exception:  # pragma: no cover
        import traceback

        q.put(("error", traceback.format_exc()))
        raise
    finally:
        dist.destroy_process_group()


def test_distributed_gmres_two_ranks_gloo():
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gmres_worker, args=(r, 2, port, 1500, q)) for r in range(2)]
    for p in procs:
        p.start()
    status, res = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
    assert status == "ok", res
    it, it_ref, same, err, resid, block, block_ref, err_block = res
    assert it == it_ref and 0 < it < 150, (it, it_ref)
    assert same
    assert err < 1e-7 and resid < 2e-6, (err, resid)
    assert block == block_ref and err_block < 1e-7, (block, block_ref, err_block)


def test_peer_wait_failure_raises_and_poisons_the_operator():
    """a device-side wait that gave up (error word != 0) must surface at the next host synchronisation point and make the
    operator refuse further collectives, instead of letting a partly gathered vector through"""
    import pytest

    from htool_python_b200.distributed import B200DistributedOperator, bind_to_gpu_cpus

    op = object.__new__(B200DistributedOperator)
    op.fused, op._peer_failed = True, 0
    status = {"v": 0}
    op.peer_status = lambda sync=True: status["v"]
    op.check_peers()  # healthy: nothing happens
    status["v"] = 3   # rank 2 never published its epoch
    with pytest.raises(RuntimeError, match="rank 2"):
        op.check_peers()
    status["v"] = 0   # the epochs stay out of step: still refused
    with pytest.raises(RuntimeError, match="must be rebuilt"):
        op.check_peers()
    with pytest.raises(RuntimeError):
        op.matvec_cluster_fused_(None)
    status["v"] = 104  # an all-reduce wait
    op2 = object.__new__(B200DistributedOperator)
    op2.fused, op2._peer_failed = True, 0
    op2.peer_status = lambda sync=True: status["v"]
    with pytest.raises(RuntimeError, match="all-reduce, rank 3"):
        op2.check_peers()
    assert isinstance(bind_to_gpu_cpus(0, 1), list)  # no NVML here: nothing is changed, empty list

/*
 * hmb200.h -- C ABI of the B200-native H-matrix product engine (libhmb200.so).
 *
 * Drop-in boundary for ONE path of htool-ddm/htool_python: y += A*x / Y += A*X on an assembled htool::HMatrix and the
 * block-row DistributedOperator product built on it.  Plain pointers and sizes only; no torch / pybind types.
 * Every entry point cites the reference interface it replaces (paths relative to /root/reference; `core/` is
 * lib/htool/include/htool/).  INTEGRATION.md shows the reference-side bindings (pybind11 operator subclass / ctypes).
 *
 * Conventions (identical to the reference's operator interface,
 * core/distributed_operator/interfaces/virtual_global_to_local_operator.hpp:8-35):
 *   out = beta*out + alpha*op(A)*in, vectors in CLUSTER numbering, multi-RHS blocks row-major (row i, rhs j at i*mu+j).
 *   Scalars are passed by pointer (double for _d, two doubles re,im for _z) so the ABI has no complex type.
 *   `mem` says where in/out live: HMB200_HOST (copied through pinned staging, synchronous on return, like the
 *   reference which owns host buffers) or HMB200_DEVICE (device pointers on `stream`, asynchronous).
 *   All functions return 0 on success, non-zero on error; hmb200_last_error() returns the message (thread-local).
 *   There is no CPU fallback: if no CUDA device is usable, create fails.
 */
#ifndef HMB200_H
#define HMB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HMB200_HOST 0
#define HMB200_DEVICE 1

/* leaf flags */
#define HMB200_LEAF_MIRROR 1u                       /* leaf is also in leaves_for_symmetry (core/hmatrix/hmatrix.hpp:262-266) */
#define HMB200_LEAF_DENSE_KIND(f) (((f) >> 1) & 7u) /* 0 general, 1 'S'/'L', 2 'S'/'U', 3 'H'/'L', 4 'H'/'U' */

/*
 * Flat leaf table of an assembled htool::HMatrix block (what get_leaves_from enumerates, core/hmatrix/hmatrix.hpp:247-274,
 * addressed as in core/hmatrix/linalg/add_hmatrix_vector_product.hpp:90-103).  All pointers are HOST pointers that stay
 * owned by the caller (they point straight into the reference's Matrix / LowRankMatrix buffers); they are only read
 * during hmb200_create_*.
 */
typedef struct hmb200_flat {
    int32_t n_rows, n_cols;      /* size of the root block (target x source cluster) */
    int64_t n_leaves;
    const int32_t *row_off;      /* leaf target offset - root target offset */
    const int32_t *col_off;      /* leaf source offset - root source offset */
    const int32_t *nrows;        /* leaf target size */
    const int32_t *ncols;        /* leaf source size */
    const int32_t *rank;         /* -1 dense (core/hmatrix/hmatrix.hpp:142-144), >= 0 low-rank */
    const uint8_t *flags;        /* HMB200_LEAF_* */
    const void *const *a;        /* dense: nrows x ncols col-major (ld nrows); low-rank: U nrows x rank col-major */
    const void *const *b;        /* low-rank: V rank x ncols col-major (ld rank) (core/hmatrix/lrmat/lrmat.hpp:47-50) */
    int32_t symmetry_for_leaves; /* 'N', 'S' or 'H' (HMatrix::get_symmetry_for_leaves) */
    int32_t n_row_cuts;          /* optional: target leaf-cluster boundaries (0 ... n_rows), may be 0 / NULL */
    const int32_t *row_cuts;
} hmb200_flat;

/* tuning knobs; zero / NULL means default */
typedef struct hmb200_options {
    int32_t stage_bytes;     /* bytes of one staged packet in shared memory (default 49152) */
    int32_t stage_z_bytes;   /* bytes of the per-stage operand (x / t) area (default 8192) */
    int32_t n_stages;        /* depth of the TMA ring (default 3) */
    int32_t consumer_warps;  /* consumer warps per CTA (default 16) */
    int32_t piece1_bytes;    /* target bytes of a phase-1 (V) piece (default 8192) */
    int32_t piece2_bytes;    /* max bytes of a phase-2 (U / dense) piece (default 8192) */
    int32_t bin_rows;        /* max rows of a target bin, <= 64 (default 64) */
    int32_t unit_stages;     /* phase-1 stages per work unit (default 32) */
    int32_t ctas;            /* persistent CTAs (default: number of SMs) */
    int32_t build_transpose; /* also pack the transposed operator so trans='T'/'C' products work (default 0) */
    int32_t no_matmat;       /* 1: do not pack the two panel streams of the multi-RHS kernels (halves the device memory; the
                                multi-RHS entry points then fall back to one streamed product per right-hand side) */
    int32_t mm_task_cols;    /* multi-RHS phase 1: source columns per work unit (default 1024); wider clusters are split and their
                                partial T rows summed */
    int32_t mm1_kg;          /* multi-RHS phase 1: k-groups (8 columns real / 4 complex) per packet of a 128-row panel (even; default 6) */
    int32_t mm2_kg;          /* multi-RHS phase 2: k-groups per packet of a 64-row panel (even; default 8) */
    int32_t reserved0;
    int32_t reserved[1];
} hmb200_options;

typedef struct hmb200_handle hmb200_handle;

/* information about a created engine */
typedef struct hmb200_info {
    int32_t n_rows, n_cols;
    int32_t is_complex;
    int32_t n_bins;
    int64_t n_leaves, n_pieces1, n_pieces2, n_stages1, n_stages2, n_reduce_segments;
    int64_t stream1_bytes, stream2_bytes; /* packed bytes streamed per product by phase 1 / phase 2 */
    int64_t table_bytes;                  /* stage / operand-copy / unit tables */
    int64_t workspace_bytes;              /* x | t | partial workspace */
    double stored_coefficients;           /* W = sum r(m+n) + sum m n, mirrored leaves counted once,
                                             symmetric dense leaves m(n+1)/2 (core/hmatrix/hmatrix_output.hpp:153-208) */
    double pack_seconds, upload_seconds;
    int64_t mm_stream_bytes, mm_tiles, mm_stages, mm_rows; /* multi-RHS panel streams (both phases): bytes, phase-1 units, packets; workspace rows of the multi-RHS path */
} hmb200_info;

/* ---- life cycle ------------------------------------------------------------------------------------------------
 * Replaces: the per-call leaf enumeration of the reference (get_leaves_from is re-run by every product,
 * core/hmatrix/linalg/add_hmatrix_vector_product.hpp:72) by a one-time flatten + pack + upload. */
hmb200_handle *hmb200_create_d(const hmb200_flat *flat, int device, void *stream, const hmb200_options *opt);
hmb200_handle *hmb200_create_z(const hmb200_flat *flat, int device, void *stream, const hmb200_options *opt);
void hmb200_destroy(hmb200_handle *h);
int hmb200_get_info(const hmb200_handle *h, hmb200_info *info);
int hmb200_set_stream(hmb200_handle *h, void *stream);
/* device-resident single-RHS products are replayed from CUDA graphs (one per (op, in, out, alpha, beta), captured the
 * second time a combination is seen); on by default, 0 turns it off */
int hmb200_set_graphs(hmb200_handle *h, int on);

/* ---- products (cluster numbering) ------------------------------------------------------------------------------
 * Replaces VirtualGlobalToLocalOperator::add_vector_product
 *   (core/distributed_operator/interfaces/virtual_global_to_local_operator.hpp:16) as implemented by
 *   RestrictedGlobalToLocalHMatrix -> openmp_internal_add_hmatrix_vector_product
 *   (core/distributed_operator/implementations/global_to_local_operators/hmatrix.hpp:27-29,
 *    core/hmatrix/linalg/add_hmatrix_vector_product.hpp:106-170). */
int hmb200_add_vector_product_d(hmb200_handle *h, char trans, const double *alpha, const double *in, const double *beta, double *out, int mem);
int hmb200_add_vector_product_z(hmb200_handle *h, char trans, const double *alpha, const double *in, const double *beta, double *out, int mem);

/* Replaces VirtualGlobalToLocalOperator::add_matrix_product_row_major (same header :25) ->
 *   openmp_internal_add_hmatrix_matrix_product_row_major (core/hmatrix/linalg/add_hmatrix_matrix_product_row_major.hpp:111-178). */
int hmb200_add_matrix_product_row_major_d(hmb200_handle *h, char trans, const double *alpha, const double *in, const double *beta, double *out, int mu, int mem);
int hmb200_add_matrix_product_row_major_z(hmb200_handle *h, char trans, const double *alpha, const double *in, const double *beta, double *out, int mu, int mem);

/* Replaces VirtualGlobalToLocalOperator::add_sub_matrix_product_to_local (same header :33) as implemented by
 *   RestrictedGlobalToLocalOperator / LocalToLocalHMatrix (core/distributed_operator/implementations/global_to_local_operators/
 *   restricted_operator.hpp:173-194, .../local_to_local_operators/hmatrix.hpp:34-51): out += A[:, offset:offset+size] * in,
 *   `in` = size x mu row-major rows of the source window, `offset` in the numbering of the source tree (the operator's own
 *   source cluster starts at the source offset given to hmb200_set_cluster_offsets, default 0).  The window is clipped to
 *   the operator's source range exactly like the reference and extended by zero on the device. */
int hmb200_set_cluster_offsets(hmb200_handle *h, int target_offset, int source_offset);
int hmb200_add_sub_matrix_product_to_local_d(hmb200_handle *h, const double *in, double *out, int mu, int offset, int size, int mem);
int hmb200_add_sub_matrix_product_to_local_z(hmb200_handle *h, const double *in, double *out, int mu, int offset, int size, int mem);

/* ---- block-row product with the all-gather fused into the kernel (one process per GPU, GPUs of one node) -------------
 * Replaces the slab product + MPI_Allgatherv of add_distributed_operator_vector_product_global_to_global
 *   (core/distributed_operator/linalg/add_distributed_operator_vector_product_global_to_global.hpp:63-76): the phase-2
 *   epilogue stores every output row of this rank's slab into the replicated vector of EVERY rank through peer-mapped
 *   memory (CUDA IPC over NVLink), the last CTA publishes an epoch flag to all ranks, and a one-warp kernel waits for the
 *   epochs of all ranks.  No NCCL call, no host synchronisation; two replicated buffers alternate.
 *   1. every rank: hmb200_peer_create(h, n_ranks, rank, n_global, slab_offset, handle[64]) -- allocates its shared region
 *   2. exchange the 64-byte handles (any transport), then hmb200_peer_attach(h, handles: n_ranks x 64 bytes in rank order)
 *   3. collectively, the same number of times on every rank: hmb200_vector_product_allgather_{d,z}(h, alpha, in, &y):
 *      y (device pointer into this rank's region, n_global elements, cluster numbering) = alpha * A * in on all ranks;
 *      asynchronous on the handle's stream.  hmb200_peer_status reads the error word (non-zero: 1 + the rank that did not
 *      arrive within the bound of a product / gather wait, 101 + rank for an all-reduce).  The word lives in page-locked
 *      mapped host memory, so the call costs nothing and covers every wait that has COMPLETED on the stream: call it after
 *      any host-side synchronisation.  After an error the epochs are out of step: destroy the operator. */
int hmb200_peer_create(hmb200_handle *h, int n_ranks, int rank, int64_t n_global, int slab_offset, unsigned char *ipc_handle_out);
int hmb200_peer_attach(hmb200_handle *h, const unsigned char *ipc_handles);
int hmb200_peer_buffer(hmb200_handle *h, int index, void **ptr);
int hmb200_peer_status(hmb200_handle *h, int *err);
/* the same for row-major blocks of mu right-hand sides (MPI_Allgatherv of
 * core/distributed_operator/linalg/add_distributed_operator_matrix_product_row_major_global_to_global.hpp:75): a second shared region
 * per rank holding two replicated (n_global x mu) blocks.  hmb200_peer_mm_create after hmb200_peer_create (again for another
 * mu: the old blocks are released), exchange the handles, hmb200_peer_mm_attach; then, collectively,
 * hmb200_matrix_product_allgather_{d,z}(h, alpha, in (device, n_cols x mu row-major), mu, &Y): Y (device pointer into this
 * rank's region, n_global x mu row-major, cluster numbering) = alpha * A * in on all ranks -- the bin epilogue of the multi-RHS
 * kernel stores every tile row to all ranks, same epoch flags; counts as one collective of the sequence above.
 * Changing mu is a two-phase collective: (1) every rank calls hmb200_peer_mm_detach (drains its stream, closes the blocks it
 * imported from its peers); (2) the ranks synchronise (any barrier); (3) every rank calls hmb200_peer_mm_create for the new mu --
 * only then does an owner free its old blocks -- and the handles are exchanged and attached as before.  hmb200_peer_mm_create
 * refuses to run while this rank still has blocks of its peers attached.  Pointers returned for the old mu are dead after (1). */
int hmb200_peer_mm_create(hmb200_handle *h, int mu, unsigned char *ipc_handle_out);
int hmb200_peer_mm_detach(hmb200_handle *h);
int hmb200_peer_mm_attach(hmb200_handle *h, const unsigned char *ipc_handles);
int hmb200_peer_mm_buffer(hmb200_handle *h, int index, void **ptr);
int hmb200_matrix_product_allgather_d(hmb200_handle *h, const double *alpha, const double *in, int mu, void **y_replicated);
int hmb200_matrix_product_allgather_z(hmb200_handle *h, const double *alpha, const double *in, int mu, void **y_replicated);
/* y = A x in USER numbering across the ranks of one node with n / P elements per rank and direction over PCIe (the replicated-input
 * / replicated-result contract of the reference's DistributedOperator.__mul__, src/htool/distributed_operator/distributed_operator.hpp:23-37,
 * add_distributed_operator_vector_product_global_to_global.hpp:96-118).  After hmb200_peer_attach and
 * hmb200_set_global_permutations(h, source permutation (n_cols, cluster slot -> user index), its size, INVERSE target permutation
 * (user index -> cluster slot, n_global), its size), collectively:
 *   hmb200_matvec_user_chunked_{d,z}(h, x_chunk, c0, count, y_chunk, r0, rcount)
 * uploads this rank's chunk x_user[c0 : c0+count] (page-locked host pointer), all-gathers the chunks over the peer regions,
 * permutes, multiplies (fused all-gather of the rows), and downloads y_user[r0 : r0+rcount] into y_chunk (page-locked host
 * pointer, e.g. this rank's window of one node-shared vector).  One cached CUDA graph; synchronous on return; counts as two
 * collectives of the sequence above. */
int hmb200_set_global_permutations(hmb200_handle *h, const int32_t *source_perm, int64_t n_source, const int32_t *inverse_target_perm, int64_t n_target);
int hmb200_matvec_user_chunked_d(hmb200_handle *h, const double *x_chunk, int64_t c0, int64_t count, double *y_chunk, int64_t r0, int64_t rcount);
int hmb200_matvec_user_chunked_z(hmb200_handle *h, const double *x_chunk, int64_t c0, int64_t count, double *y_chunk, int64_t r0, int64_t rcount);
/* the same for column-major blocks of mu right-hand sides in user numbering (DistributedOperator.__matmul__,
 * core/distributed_operator/linalg/add_distributed_operator_matrix_product_global_to_global.hpp:138-280): X / Y are the BASES of
 * the page-locked host blocks (element (i, j) at i + j*ld; only the rows named below are touched), every rank uploads rows [c0, c0+count) of X and
 * downloads rows [r0, r0+rcount) of Y (all columns).  Needs hmb200_peer_mm_create / _attach for this mu.  Synchronous; counts as
 * two collectives. */
int hmb200_matmat_user_chunked_d(hmb200_handle *h, const double *X, int64_t ldx, int64_t c0, int64_t count, double *Y, int64_t ldy, int64_t r0, int64_t rcount, int mu);
int hmb200_matmat_user_chunked_z(hmb200_handle *h, const double *X, int64_t ldx, int64_t c0, int64_t count, double *Y, int64_t ldy, int64_t r0, int64_t rcount, int mu);
/* the solver's two collectives on the same peer regions (no NCCL / MPI on the iteration path):
 *   hmb200_peer_allgather     local slab (n_rows elements, device) -> replicated vector on every rank
 *                             (MPI_Allgatherv of ..._vector_product_local_to_local.hpp:44); counts as one collective of step 3
 *   hmb200_peer_allreduce_f64 in-place sum over the ranks of n <= 512 device doubles, added in rank order on every rank
 *                             (the MPI_Allreduce of HPDDM's Arnoldi step, lib/hpddm/include/HPDDM_iterative.hpp:549,740) */
int hmb200_peer_allgather(hmb200_handle *h, const void *local_slab, void **y_replicated);
/* the same with an arbitrary disjoint range [elem_offset, elem_offset+count) per rank (e.g. equal chunks of a user-numbered
 * input that every rank holds on the host: each rank uploads 1/P of it and the chunks are exchanged over NVLink) */
int hmb200_peer_allgather_range(hmb200_handle *h, const void *local_part, int64_t elem_offset, int64_t count, void **y_replicated);
int hmb200_peer_allreduce_f64(hmb200_handle *h, double *vals, int n);
/* the exchange of the TRANSPOSED local-to-local product on the same regions (trans = 'T' or 'C', needs build_transpose=1 and a
 * square operator): y_local = beta y_local + sum over the ranks q, ascending, of [alpha op(A_q) x_q](slab of this rank).  Every
 * rank leaves its full-length contribution in its own replicated buffer and pulls its slab of all contributions over NVLink --
 * the MPI_Alltoallv + scal + one axpy per rank of ..._vector_product_local_to_local.hpp:60-87.  x_local / y_local: device, n_rows
 * elements; alpha, beta as in hmb200_vector_product_*; collective, counts as one product of step 3. */
int hmb200_peer_exchange_sum(hmb200_handle *h, char trans, const double *alpha, const void *x_local, const double *beta, void *y_local);
int hmb200_vector_product_allgather_d(hmb200_handle *h, const double *alpha, const double *in, void **y_replicated);
int hmb200_vector_product_allgather_z(hmb200_handle *h, const double *alpha, const double *in, void **y_replicated);

/* ---- permutations (bit-exact) -----------------------------------------------------------------------------------
 * gather : out[i] = in[perm[off+i]-off]   user_to_cluster / global_to_root_cluster (core/clustering/cluster_node.hpp:110-119,162-175)
 * scatter: out[perm[off+i]-off] = in[i]   cluster_to_user / root_cluster_to_global (core/clustering/cluster_node.hpp:99-109,149-161)
 * `perm` follows `mem` like in/out; `elem_bytes` is 8 (double) or 16 (complex<double>); `mu` columns of length n stored
 * one after the other (column-major block, as add_hmatrix_matrix_product permutes per column,
 * core/hmatrix/linalg/add_hmatrix_matrix_product.hpp:190-204). */
int hmb200_permute_gather(const int32_t *perm, int off, int n, const void *in, void *out, int elem_bytes, int mu, int mem, void *stream);
int hmb200_permute_scatter(const int32_t *perm, int off, int n, const void *in, void *out, int elem_bytes, int mu, int mem, void *stream);

/* ---- user-numbering convenience (what HMatrix.__mul__ / __matmul__ do, src/htool/hmatrix/hmatrix.hpp:101-138):
 * y = A x with both permutations fused into the device pipeline.  x, y: user numbering; X, Y column-major n x mu
 * (the Fortran-ordered arrays of the Python API).  perms are device-resident after hmb200_set_permutations. */
int hmb200_set_permutations(hmb200_handle *h, const int32_t *target_perm_local, const int32_t *source_perm_local);
int hmb200_matvec_user_d(hmb200_handle *h, const double *x, double *y, int mem);
int hmb200_matvec_user_z(hmb200_handle *h, const double *x, double *y, int mem);
int hmb200_matmat_user_d(hmb200_handle *h, const double *X, double *Y, int mu, int mem);
int hmb200_matmat_user_z(hmb200_handle *h, const double *X, double *Y, int mu, int mem);

/* ---- Krylov vector kernels (device pointers only, asynchronous on `stream`) ---------------------------------------
 * The BLAS-1/2 of the GMRES the reference's solver runs (DDM::solve -> HPDDM::IterativeMethod::GMRES,
 * core/solvers/ddm.hpp:193, lib/hpddm/include/HPDDM_GMRES.hpp:30-196) on basis vectors that stay in HBM:
 *   dots    h[j] = <V_j, w>, j < k                 (Blas::gemv 'C' of orthogonalization, lib/hpddm/include/HPDDM_iterative.hpp:541-548)
 *   combine w += sign * sum_j h[j] V_j, *norm2 = ||w||^2   (Blas::gemv 'N' :551-552 and the dot of Arnoldi :738-739; with sign = +1
 *           the basis combination of addSol :316-326)
 *   axpby   out = alpha*x + beta*y, *norm2 = ||out||^2     (residual b - A x, HPDDM_GMRES.hpp:62-77)
 *   scale   w /= sqrt(*norm2)                               (HPDDM_iterative.hpp:743-745)
 * V holds basis vector j at V + j*ldv (elements); h, norm2 are DEVICE scalars (all-reduced between calls by the caller
 * when the rows are distributed); norm2 may be NULL.  `scratch` must hold hmb200_krylov_scratch_elems(n, k) elements.
 * Two-stage fixed-order reductions: results are bit-reproducible run to run. */
int64_t hmb200_krylov_scratch_elems(int64_t n, int k);
int hmb200_krylov_dots(const void *V, int64_t ldv, int k, const void *w, int64_t n, void *h, void *scratch, int elem_bytes, void *stream);
int hmb200_krylov_combine(const void *V, int64_t ldv, int k, const void *h, double sign, void *w, int64_t n, double *norm2, void *scratch, int elem_bytes, void *stream);
int hmb200_krylov_axpby(const double *alpha, const void *x, const double *beta, const void *y, void *out, int64_t n, double *norm2, void *scratch, int elem_bytes, void *stream);
int hmb200_krylov_scale(void *w, int64_t n, const double *norm2, int elem_bytes, void *stream);
/* hmb200_krylov_dots for callers that cannot bring a workspace.  (No counterpart in the reference: its solver runs on host
 * vectors and never needs one.)  Some users of this library are not able to allocate device memory at the point where they
 * need the dot products: an operator or orthogonalisation callback that a foreign solver calls with raw pointers only, code
 * that was handed a stream it must not stall (cudaMalloc / cudaFree synchronise the whole device), a process whose HBM is
 * fully planned in advance, or a C / Fortran / ctypes host with no device allocator.  This variant takes the same V, ldv, k,
 * w, n, h as hmb200_krylov_dots and writes the same h (same kernels, same summation order), but keeps its partial sums in an
 * array that is built into the library's device code, so neither the caller nor the library allocates anything.
 *   Limit: ceil(n / 2048) * k * elem_bytes <= HMB200_STATIC_SCRATCH_BYTES (4 MiB: e.g. n = 1M rows with k <= 1072 real or
 *     536 complex basis vectors).  Anything larger is refused with a non-zero status and a message; use hmb200_krylov_dots.
 *   Cost: each GPU that loads the library's kernels keeps those 4 MiB until the process ends.
 *   First call per device: looks up the array's address on the CURRENT device (the one V, w, h live on must be current).
 *     Make that first call during set-up, before any stream capture.
 *   Sharing: one array per GPU for all calls; it is not cleared between calls and does not need to be (every partial sum that
 *     is read was written by the same call).  Calls that may overlap on one GPU (two streams, two threads) corrupt each
 *     other's partial sums: issue them one after the other on one stream, or order them with events. */
#define HMB200_STATIC_SCRATCH_BYTES (4 * 1024 * 1024)
int hmb200_krylov_dots_static(const void *V, int64_t ldv, int k, const void *w, int64_t n, void *h, int elem_bytes, void *stream);

/* timing of the kernels of the last product on the handle's stream (CUDA events), in milliseconds:
 * ms[0] phase 1 (t = V x), ms[1] partial reduce, ms[2] phase 2 (y += U t, y += D x), ms[3] whole product */
int hmb200_last_kernel_ms(hmb200_handle *h, float ms[4]);
/* the same, averaged over the (up to 256) products issued since hmb200_enable_timing(h, 1); events are recorded on the
 * launching stream without synchronising, and collected here */
int hmb200_kernel_ms_avg(hmb200_handle *h, float ms[4], int *n_products);
int hmb200_enable_timing(hmb200_handle *h, int on);
/* number of kernel launches issued by this handle since creation */
int64_t hmb200_launch_count(const hmb200_handle *h);
/* Layout change between the reference's COLUMN-major blocks in user numbering (row i, rhs j at i + j*col_stride) and the row-major
 * blocks in cluster numbering of the operator interface (cluster row i, rhs j at i*mu + j), fused with the permutation (perm:
 * device pointer, cluster slot -> user index, or NULL for none): the per-column user_to_cluster / cluster_to_user plus the
 * transposes of core/hmatrix/linalg/add_hmatrix_matrix_product.hpp:29-31,190-204 and of HPDDMOperator::GMV
 * (core/wrappers/wrapper_hpddm.hpp:108-116,129-135).  Device pointers, asynchronous on `stream`. */
int hmb200_block_to_rows(const int32_t *perm, int n, int mu, const void *src_colmajor, long long col_stride, void *dst_rowmajor, int elem_bytes, void *stream);
int hmb200_rows_to_block(const int32_t *perm, int n, int mu, const void *src_rowmajor, void *dst_colmajor, long long col_stride, int elem_bytes, void *stream);
/* Diagnostics: measured FP64 issue rates of `device` in TFLOP/s -- tensor-core DMMA (mma.sync.m8n8k4.f64) and plain DFMA, 8
 * warps per SM with 8 independent accumulators each; the roofline denominators of the multi-RHS path (no reference
 * counterpart: bench.py reports them next to the achieved rate). */
int hmb200_fp64_rate(int device, double *dmma_tflops, double *dfma_tflops);

/* Host-side completion flags of the chunked user-numbering products (no reference counterpart: the reference's ranks meet in
 * MPI_Allgatherv).  Every rank owns one int64 word of a node-shared segment (word q at flags + q*stride) and, after its part of
 * the result has landed in the shared block, publishes the call counter there with release semantics; _wait returns 0 once
 * all n words are >= value (acquire loads; spins briefly, then yields, then sleeps in 50 us steps), 2 after timeout_s seconds
 * with *laggard = the first rank still behind.  Host code only: usable without a GPU. */
int hmb200_host_flag_publish(int64_t *flag, int64_t value);
int hmb200_host_flags_wait(const int64_t *flags, int64_t stride, int n, int64_t value, double timeout_s, int *laggard);

const char *hmb200_last_error(void);
const char *hmb200_version(void);

#ifdef __cplusplus
}
#endif
#endif /* HMB200_H */

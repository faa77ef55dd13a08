// hmb200.cu -- C ABI of the engine (include/hmb200.h): create (plan + pack + upload), products, permutations.
#include "hmb200.h"
#include "internal.h"
#include "planner.h"
#include "../kernels/stream_kernels.cuh"
#include "../kernels/mm_kernels.cuh"
#include "../kernels/peer_kernels.cuh"

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <algorithm>
#include <iterator>
#include <vector>

using namespace hmb200;

static thread_local std::string g_last_error;
void hmb200i_set_error(const char *msg) { g_last_error = msg ? msg : ""; }

// ABI layout pinned (mirrored by htool_python_b200/_lib.py and checked in tests/test_capi_symbols.py)
static_assert(sizeof(hmb200_flat) == 96 && sizeof(hmb200_options) == 64 && sizeof(hmb200_info) == 152, "hmb200 ABI layout changed");
static_assert(sizeof(PacketHeader) == 16 && sizeof(PieceDesc) == 16 && sizeof(StageEntry) == 32 && sizeof(ZCopy) == 8 && sizeof(WorkUnit) == 8 && sizeof(ReduceSeg) == 16, "packet layout changed");

#define CUDA_CHECK(expr)                                                                                               \
    do {                                                                                                               \
        cudaError_t _e = (expr);                                                                                       \
        if (_e != cudaSuccess)                                                                                         \
            throw std::runtime_error(std::string(#expr) + " failed: " + cudaGetErrorString(_e));                       \
    } while (0)

namespace {

struct PackedOp {
    bool valid = false;
    char op    = 'N';
    int32_t n_in = 0, n_out = 0;
    // index 0: phase 1, 1: phase 2 (single RHS); 2, 3: the multi-RHS panel streams of phase 1 / phase 2 (plan.h "mm streams")
    uint8_t *stream[4]    = {nullptr, nullptr, nullptr, nullptr};
    StageEntry *stages[4] = {nullptr, nullptr, nullptr, nullptr};
    ZCopy *zcopies[4]     = {nullptr, nullptr, nullptr, nullptr};
    WorkUnit *units[4]    = {nullptr, nullptr, nullptr, nullptr};
    uint32_t n_units[4]   = {0, 0, 0, 0};
    uint32_t n_zcopies[4] = {0, 0, 0, 0};
    ReduceSeg *reduce     = nullptr;
    uint32_t n_reduce     = 0;
    ReduceSeg *reduce_mm  = nullptr;
    uint32_t *reduce_mm_rows = nullptr;
    uint32_t n_reduce_mm  = 0;
    uint64_t xz_elems     = 0;
    uint64_t mm_rows      = 0;
    uint32_t mm_zero_row  = 0;
    uint32_t mm_packet_bytes[2] = {0, 0}; // largest packet of the two panel streams
    bool has_mm           = false;
    // bookkeeping
    uint64_t stream_bytes[4] = {0, 0, 0, 0};
    uint64_t n_stages[4] = {0, 0, 0, 0}, n_pieces[4] = {0, 0, 0, 0};
    uint64_t table_bytes = 0;
    int n_bins = 0;
};

} // namespace

struct hmb200_handle {
    int device          = 0;
    cudaStream_t stream = nullptr;
    int es              = 8;
    Options opt;
    int n_sms = 0;
    size_t smem_optin = 0;
    int32_t n_rows = 0, n_cols = 0;
    PackedOp ops[3]; // N, T, C
    void *xz = nullptr;
    uint64_t xz_elems = 0;
    void *ybuf = nullptr; // device staging for host-resident outputs / user-numbering paths
    uint64_t ybuf_elems = 0;
    void *ybuf2 = nullptr; // cluster-numbered row-major result of the user-numbering multi-RHS product
    uint64_t ybuf2_elems = 0;
    void *ws_mm = nullptr; // multi-RHS workspace rows [x | t | partials] at stride ld (grown on demand)
    uint64_t ws_mm_elems = 0;
    void *xbuf = nullptr; // device staging for multi-RHS inputs
    uint64_t xbuf_elems = 0;
    void *xsub = nullptr; // zero-extended source window of add_sub_matrix_product_to_local
    uint64_t xsub_elems = 0;
    int target_offset = 0, source_offset = 0; // offsets of the root block in the target / source tree numbering
    // CUDA graphs of the device-resident product: one instantiated graph per (op, in, out, alpha, beta), captured on a private
    // stream the second time a key is seen and replayed on the caller's stream (removes ~20 us of launch latency per product)
    struct GraphKey {
        char op; const void *in; void *out; double a[2], b[2]; int fused;
        bool operator<(const GraphKey &o) const { return std::memcmp(this, &o, sizeof(GraphKey)) < 0; }
    };
    struct GraphEntry { cudaGraphExec_t exec = nullptr; int64_t launches = 0; int seen = 0; uint64_t last_use = 0; };
    std::map<GraphKey, GraphEntry> graphs;
    uint64_t graph_tick = 0;
    static constexpr size_t kMaxGraphKeys = 256;
    cudaStream_t cap_stream = nullptr;
    bool use_graphs = true;
    bool use_pdl = false; // programmatic dependent launches inside a product: opt-in with HMB200_PDL=1 (measured: no gain on top of the CUDA-graph replay, profiles/r02_tried_and_dropped.md)
    // fused all-gather: one IPC-shared region per rank = [replicated y, buffer 0 | buffer 1 | control block]; the control
    // block holds flags[2][8] (written by the peers), and -- local use only -- epoch[2], the CTA counter and the error word
    struct Peer {
        int n_ranks = 0, rank = 0, slab_offset = 0;
        int64_t n_global = 0;
        size_t buf_bytes = 0, region_bytes = 0;
        uint8_t *region[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr}; // [rank] own allocation, others IPC-mapped
        bool attached = false;
        // replicated row-major blocks (n_global x mm_mu) of the fused multi-RHS product: a second IPC region per rank, two blocks
        uint8_t *mm_region[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
        size_t mm_buf_bytes = 0;
        int mm_mu = 0;
        bool mm_attached = false;
        uint64_t products = 0; // collectives on the replicated buffers so far (fused products + gathers): selects the buffer
        uint64_t reduces = 0;  // all-reduces so far: selects the slot parity
        long long wait_cycles = 240000000000ll; // bound of the device-side waits (~2 minutes; HMB200_PEER_TIMEOUT_S overrides)
        // error word of the bounded waits: page-locked MAPPED host memory, written by the kernels through err_dev and read by
        // the host without any CUDA call, so every host-side synchronisation point can check it for free
        int *err_host = nullptr, *err_dev = nullptr;
    } peer;
    void *pinned = nullptr;
    size_t pinned_bytes = 0;
    uint32_t *counters = nullptr; // [2]
    int32_t *perm_t = nullptr, *perm_s = nullptr;
    int64_t gperm_n_target = 0;                    // n_target of the first hmb200_set_global_permutations call (sizes ginv_t, gy)
    int32_t *gperm_s = nullptr, *ginv_t = nullptr; // distributed user-numbering product: full source permutation, inverse target permutation
    void *gx = nullptr, *gy = nullptr;             // ... and its chunk staging (n_global elements each, never reallocated: cached graphs hold them)
    // per-kernel timing: a ring of CUDA-event quads recorded on the launching stream, collected lazily (no sync in the
    // product path); kTimingSlots products are remembered
    static constexpr int kTimingSlots = 256;
    std::vector<cudaEvent_t> ev; // 4 per slot
    int64_t timed_products = 0;  // products recorded since timing was enabled
    bool timing = false;
    float last_ms[4] = {0, 0, 0, 0};
    int64_t launches = 0;
    hmb200_info info{};
};

namespace {

template <typename V>
V *upload_vec(const std::vector<V> &v, cudaStream_t s) {
    V *d = nullptr;
    size_t bytes = std::max<size_t>(v.size() * sizeof(V), 64);
    CUDA_CHECK(cudaMalloc(&d, bytes));
    if (!v.empty()) CUDA_CHECK(cudaMemcpyAsync(d, v.data(), v.size() * sizeof(V), cudaMemcpyHostToDevice, s));
    return d;
}

// host staging copy between the caller's (pageable) vectors and the pinned buffer: multi-threaded above 1 MiB, so that
// the HOST entry points are not bound by one core's memcpy rate (8 MB vectors at N=1M)
void staged_copy(void *dst, const void *src, size_t bytes) {
    constexpr size_t kChunk = 256 << 10;
    if (bytes < (1u << 20)) {
        std::memcpy(dst, src, bytes);
        return;
    }
    const int64_t n = (int64_t)((bytes + kChunk - 1) / kChunk);
#pragma omp parallel for schedule(static) num_threads(8)
    for (int64_t i = 0; i < n; i++) {
        const size_t o = (size_t)i * kChunk;
        std::memcpy((uint8_t *)dst + o, (const uint8_t *)src + o, std::min(kChunk, bytes - o));
    }
}

// a caller buffer that is already page-locked (cudaHostAlloc / cudaHostRegister / torch pin_memory) is DMA'd directly
bool host_is_pinned(const void *p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost;
}
// host source of an H2D copy: the caller's buffer itself when page-locked, else the pinned slot after a staging copy
const void *stage_in(const void *user, void *slot, size_t bytes) {
    if (host_is_pinned(user)) return user;
    staged_copy(slot, user, bytes);
    return slot;
}

void ensure_pinned(hmb200_handle *h, size_t bytes) {
    if (h->pinned_bytes >= bytes) return;
    if (h->pinned) CUDA_CHECK(cudaFreeHost(h->pinned));
    h->pinned = nullptr;
    CUDA_CHECK(cudaHostAlloc(&h->pinned, bytes, cudaHostAllocDefault));
    h->pinned_bytes = bytes;
}
void ensure_dev(void **p, uint64_t *have, uint64_t need_elems, int es) {
    if (*have >= need_elems) return;
    if (*p) CUDA_CHECK(cudaFree(*p));
    *p = nullptr;
    CUDA_CHECK(cudaMalloc(p, (size_t)(need_elems + 2) * es));
    *have = need_elems;
}

// packs one phase's stream window by window through two pinned buffers and uploads it
void upload_stream(hmb200_handle *h, const hmb200_flat &flat, const Plan &plan, int phase, uint8_t **dptr, double &pack_s, double &up_s) {
    struct View { const std::vector<StageEntry> &stages; uint64_t stream_bytes; };
    const View ph = phase == 1 ? View{plan.phase1.stages, plan.phase1.stream_bytes} : phase == 2 ? View{plan.phase2.stages, plan.phase2.stream_bytes}
                  : phase == 3 ? View{plan.mm1.stages, plan.mm1.stream_bytes} : View{plan.mm2.stages, plan.mm2.stream_bytes};
    CUDA_CHECK(cudaMalloc(dptr, (size_t)std::max<uint64_t>(ph.stream_bytes, 128)));
    if (ph.stages.empty()) return;
    const uint64_t window = std::min<uint64_t>(std::max<uint64_t>(ph.stream_bytes, 1 << 20), 256ull << 20);
    uint8_t *buf[2] = {nullptr, nullptr};
    cudaEvent_t done[2];
    for (int i = 0; i < 2; i++) {
        CUDA_CHECK(cudaHostAlloc((void **)&buf[i], (size_t)window + 256, cudaHostAllocDefault));
        CUDA_CHECK(cudaEventCreateWithFlags(&done[i], cudaEventDisableTiming));
    }
    size_t s0 = 0;
    int k = 0;
    const size_t ns = ph.stages.size();
    while (s0 < ns) {
        const uint64_t base = ph.stages[s0].gmem_off;
        size_t s1 = s0 + 1;
        while (s1 < ns && ph.stages[s1].gmem_off + ph.stages[s1].bytes - base <= window) s1++;
        const uint64_t end = (s1 < ns) ? ph.stages[s1].gmem_off : ph.stream_bytes;
        if (end - base > window + 256) throw std::runtime_error("internal: packet larger than the upload window");
        CUDA_CHECK(cudaEventSynchronize(done[k]));
        auto t0 = std::chrono::steady_clock::now();
        pack_stages(flat, plan, phase, s0, s1, buf[k]);
        auto t1 = std::chrono::steady_clock::now();
        CUDA_CHECK(cudaMemcpyAsync(*dptr + base, buf[k], (size_t)(end - base), cudaMemcpyHostToDevice, h->stream));
        CUDA_CHECK(cudaEventRecord(done[k], h->stream));
        pack_s += std::chrono::duration<double>(t1 - t0).count();
        s0 = s1;
        k ^= 1;
    }
    auto t2 = std::chrono::steady_clock::now();
    CUDA_CHECK(cudaStreamSynchronize(h->stream));
    up_s += std::chrono::duration<double>(std::chrono::steady_clock::now() - t2).count();
    for (int i = 0; i < 2; i++) {
        CUDA_CHECK(cudaFreeHost(buf[i]));
        CUDA_CHECK(cudaEventDestroy(done[i]));
    }
}

void build_op(hmb200_handle *h, const hmb200_flat &flat, char op, PackedOp &po) {
    Plan plan;
    build_plan(flat, h->es, op, h->opt, plan);
    po.op    = op;
    po.n_in  = plan.n_in;
    po.n_out = plan.n_out;
    double pack_s = 0, up_s = 0;
    po.has_mm = h->opt.build_matmat != 0;
    for (int ph = 0; ph < 2; ph++) {
        const PhasePlan &pp = ph == 0 ? plan.phase1 : plan.phase2;
        upload_stream(h, flat, plan, ph + 1, &po.stream[ph], pack_s, up_s);
        po.stages[ph]       = upload_vec(pp.stages, h->stream);
        po.zcopies[ph]      = upload_vec(pp.zcopies, h->stream);
        po.units[ph]        = upload_vec(pp.units, h->stream);
        po.n_units[ph]      = (uint32_t)pp.units.size();
        po.n_zcopies[ph]    = (uint32_t)pp.zcopies.size();
        po.stream_bytes[ph] = pp.stream_bytes;
        po.n_stages[ph]     = pp.stages.size();
        po.table_bytes += pp.stages.size() * sizeof(StageEntry) + pp.zcopies.size() * sizeof(ZCopy) + pp.units.size() * sizeof(WorkUnit);
    }
    for (int ph = 2; ph < (po.has_mm ? 4 : 2); ph++) {
        const MmPhasePlan &pp = ph == 2 ? plan.mm1 : plan.mm2;
        upload_stream(h, flat, plan, ph + 1, &po.stream[ph], pack_s, up_s);
        po.stages[ph]       = upload_vec(pp.stages, h->stream);
        po.zcopies[ph]      = upload_vec(pp.zcopies, h->stream);
        po.units[ph]        = upload_vec(pp.units, h->stream);
        po.n_units[ph]      = (uint32_t)pp.units.size();
        po.n_zcopies[ph]    = (uint32_t)pp.zcopies.size();
        po.stream_bytes[ph] = pp.stream_bytes;
        po.n_stages[ph]     = pp.stages.size();
        po.n_pieces[ph]     = pp.pieces.size();
        po.mm_packet_bytes[ph - 2] = pp.max_packet_bytes;
        po.table_bytes += pp.stages.size() * sizeof(StageEntry) + pp.zcopies.size() * sizeof(ZCopy) + pp.units.size() * sizeof(WorkUnit);
    }
    po.n_pieces[0] = plan.pieces1.size();
    po.n_pieces[1] = plan.pieces2.size();
    po.reduce_mm   = upload_vec(plan.reduce_mm, h->stream);
    po.reduce_mm_rows = upload_vec(plan.reduce_mm_rows, h->stream);
    po.n_reduce_mm = (uint32_t)plan.reduce_mm.size();
    po.mm_rows     = plan.mm_rows;
    po.mm_zero_row = plan.mm_zero_row;
    po.reduce      = upload_vec(plan.reduce, h->stream);
    po.n_reduce    = (uint32_t)plan.reduce.size();
    po.table_bytes += plan.reduce.size() * sizeof(ReduceSeg);
    po.xz_elems = plan.xz_elems;
    po.n_bins   = (int)plan.bin_start.size() - 1;
    CUDA_CHECK(cudaStreamSynchronize(h->stream));
    po.valid = true;
    if (op == 'N') {
        h->info.stored_coefficients = plan.stored_coefficients;
        h->info.n_bins              = po.n_bins;
    }
    h->info.pack_seconds += pack_s;
    h->info.upload_seconds += up_s;
}

void free_op(PackedOp &po) {
    cudaFree(po.reduce_mm);
    cudaFree(po.reduce_mm_rows);
    for (int i = 0; i < 4; i++) {
        cudaFree(po.stream[i]);
        cudaFree(po.stages[i]);
        cudaFree(po.zcopies[i]);
        cudaFree(po.units[i]);
    }
    cudaFree(po.reduce);
    po = PackedOp();
}

template <typename T>
void set_kernel_attrs(size_t smem1, size_t smem2) {
    CUDA_CHECK(cudaFuncSetAttribute(hmatrix_stream_kernel<T, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem1));
    CUDA_CHECK(cudaFuncSetAttribute(hmatrix_stream_kernel<T, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
}

// workspace row stride for a pass of nc right-hand sides: bank-conflict-free B-fragment loads (mm_kernels.cuh)
inline int mm_ld(int nc, int es) {
    const int nc8 = (nc + 7) & ~7;
    if (es == 16) return nc8 + 2;
    return nc8 + ((nc8 % 16 == 0) ? 4 : 12);
}
inline int mm_max_ld(int es) { return es == 16 ? 34 : 68; }
inline int mm_pass_cols(int es) { return es == 8 ? 64 : 32; }

// shared memory of the multi-RHS kernels for a pass at row stride ld: ring depth that fits the opt-in limit (>= 2 or no launch)
struct MmLaunchCfg { int stage_bytes, z_rows, n_stages; size_t smem; };
inline MmLaunchCfg mm_launch_cfg(int phase, int es, uint32_t max_packet_bytes, int kg_per_packet, int ld, size_t smem_optin) {
    MmLaunchCfg c{};
    c.stage_bytes = (int)((std::max<uint32_t>(max_packet_bytes, 16) + 127) / 128 * 128);
    c.z_rows      = kg_per_packet * (es == 8 ? 8 : 4);
    int ns        = 6;
    while (ns > 1 && mm_smem_bytes(phase, es, c.stage_bytes, c.z_rows, ns, ld) > smem_optin) ns--;
    c.n_stages = ns;
    c.smem     = mm_smem_bytes(phase, es, c.stage_bytes, c.z_rows, ns, ld);
    return c;
}
inline bool mm_fits(const hmb200_handle *h, size_t smem_optin) {
    const Options &o = h->opt;
    const int ld = mm_max_ld(h->es), kgc = h->es == 8 ? 8 : 4;
    const uint32_t p1 = 16 + (uint32_t)o.mm1_kg * (kMmRows1 / 8) * 512 + kMmRows1 * 4, p2 = 16 + (uint32_t)o.mm2_kg * (kMmRows2 / 8) * 512;
    (void)kgc;
    return mm_launch_cfg(1, h->es, p1, o.mm1_kg, ld, smem_optin).n_stages >= 2 && mm_launch_cfg(2, h->es, p2, o.mm2_kg, ld, smem_optin).n_stages >= 2;
}

hmb200_handle *create_impl(const hmb200_flat *flat, int device, void *stream, const hmb200_options *opt, int es) {
    if (!flat) throw std::runtime_error("flat is NULL");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        throw std::runtime_error("no CUDA device available (the engine has no CPU fallback)");
    CUDA_CHECK(cudaSetDevice(device));
    cudaDeviceProp prop;
    CUDA_CHECK(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) throw std::runtime_error("libhmb200.so holds sm_100a code only (B200); this device is sm_" + std::to_string(prop.major) + std::to_string(prop.minor));
    hmb200_handle *h = new hmb200_handle();
    try {
        h->device = device;
        h->stream = (cudaStream_t)stream;
        h->es     = es;
        h->opt    = resolve_options(opt);
        h->n_sms  = prop.multiProcessorCount;
        h->n_rows = flat->n_rows;
        h->n_cols = flat->n_cols;
        // fit the ring into the opt-in shared memory of this device
        while (h->opt.n_stages > 2 && stream_smem_bytes(2, es, h->opt.stage_bytes, h->opt.stage_z_bytes, h->opt.n_stages, h->opt.consumer_warps) > (size_t)prop.sharedMemPerBlockOptin)
            h->opt.n_stages--;
        h->info = hmb200_info{};
        h->smem_optin = (size_t)prop.sharedMemPerBlockOptin;
        if (const char *e = std::getenv("HMB200_PDL")) h->use_pdl = std::atoi(e) != 0;
        if (h->opt.build_matmat && !mm_fits(h, h->smem_optin))
            h->opt.build_matmat = 0; // packet sizes chosen by the caller leave no room for two stages of the multi-RHS kernels
        build_op(h, *flat, 'N', h->ops[0]);
        if (h->opt.build_transpose) {
            const char sym = (char)flat->symmetry_for_leaves;
            if (sym != 'H') build_op(h, *flat, 'T', h->ops[1]);
            if (es == 16 && sym != 'S') build_op(h, *flat, 'C', h->ops[2]);
        }
        for (auto &po : h->ops)
            if (po.valid) h->xz_elems = std::max(h->xz_elems, po.xz_elems);
        CUDA_CHECK(cudaMalloc(&h->xz, (size_t)(h->xz_elems + 4) * es));
        CUDA_CHECK(cudaMemsetAsync(h->xz, 0, (size_t)(h->xz_elems + 4) * es, h->stream));
        CUDA_CHECK(cudaMalloc(&h->counters, 64));
        size_t s1 = stream_smem_bytes(1, es, h->opt.stage_bytes, h->opt.stage_z_bytes, h->opt.n_stages, h->opt.consumer_warps);
        size_t s2 = stream_smem_bytes(2, es, h->opt.stage_bytes, h->opt.stage_z_bytes, h->opt.n_stages, h->opt.consumer_warps);
        if (es == 8) set_kernel_attrs<double>(s1, s2); else set_kernel_attrs<double2>(s1, s2);
        CUDA_CHECK(cudaStreamSynchronize(h->stream));
        // info
        const PackedOp &n = h->ops[0];
        h->info.n_rows = h->n_rows; h->info.n_cols = h->n_cols; h->info.is_complex = es == 16;
        h->info.n_leaves = flat->n_leaves;
        h->info.n_pieces1 = n.n_pieces[0]; h->info.n_pieces2 = n.n_pieces[1];
        h->info.n_stages1 = n.n_stages[0]; h->info.n_stages2 = n.n_stages[1];
        h->info.n_reduce_segments = n.n_reduce;
        h->info.stream1_bytes = n.stream_bytes[0]; h->info.stream2_bytes = n.stream_bytes[1];
        h->info.table_bytes = n.table_bytes;
        h->info.workspace_bytes = (int64_t)h->xz_elems * es;
        h->info.mm_stream_bytes = n.stream_bytes[2] + n.stream_bytes[3]; h->info.mm_tiles = n.n_units[2]; h->info.mm_stages = n.n_stages[2] + n.n_stages[3]; h->info.mm_rows = n.mm_rows;
    } catch (...) {
        hmb200_destroy(h);
        throw;
    }
    return h;
}

} // namespace

// ------------------------------------------------------------------------------------------------------------
// product
// ------------------------------------------------------------------------------------------------------------
namespace {

template <typename T>
__global__ void strided_gather_kernel(const T *__restrict__ in, int n, int mu, int j, T *__restrict__ out) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < (size_t)n; i += (size_t)gridDim.x * blockDim.x)
        out[i] = in[i * mu + j];
}
template <typename T>
__global__ void strided_scatter_kernel(const T *__restrict__ in, int n, int mu, int j, T *__restrict__ out) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < (size_t)n; i += (size_t)gridDim.x * blockDim.x)
        out[i * mu + j] = in[i];
}

PackedOp &select_op(hmb200_handle *h, char trans) {
    int idx = trans == 'N' ? 0 : (trans == 'T' ? 1 : (trans == 'C' ? (h->es == 8 ? 1 : 2) : -1));
    if (idx < 0) throw std::runtime_error("trans must be 'N', 'T' or 'C'");
    if (!h->ops[idx].valid)
        throw std::runtime_error(std::string("operator for trans='") + trans + "' was not packed (create with build_transpose=1, or unsupported with this symmetry)");
    return h->ops[idx];
}

// control block of rank q's shared region (after the two replicated buffers)
inline unsigned long long *peer_ctrl_flags(const hmb200_handle::Peer &pr, int q, int buf) {
    return reinterpret_cast<unsigned long long *>(pr.region[q] + 2 * pr.buf_bytes) + 8 * buf;
}
inline unsigned long long *peer_ctrl_epoch(const hmb200_handle::Peer &pr, int buf) {
    return reinterpret_cast<unsigned long long *>(pr.region[pr.rank] + 2 * pr.buf_bytes + 128) + buf;
}
inline uint32_t *peer_ctrl_done(const hmb200_handle::Peer &pr) { return reinterpret_cast<uint32_t *>(pr.region[pr.rank] + 2 * pr.buf_bytes + 160); }
// all-reduce area of rank q (after the 256-byte control block): flags[2][8], epoch[2] (local use), then slots[2][8][kPeerReduceMax]
inline unsigned long long *peer_red_flags(const hmb200_handle::Peer &pr, int q, int parity) {
    return reinterpret_cast<unsigned long long *>(pr.region[q] + 2 * pr.buf_bytes + 256) + 8 * parity;
}
inline unsigned long long *peer_red_epoch(const hmb200_handle::Peer &pr, int parity) {
    return reinterpret_cast<unsigned long long *>(pr.region[pr.rank] + 2 * pr.buf_bytes + 256 + 128) + parity;
}
inline double *peer_red_slots(const hmb200_handle::Peer &pr, int q, int parity) {
    return reinterpret_cast<double *>(pr.region[q] + 2 * pr.buf_bytes + 512) + (size_t)parity * kMaxPeers * kPeerReduceMax;
}
inline int *peer_ctrl_err(const hmb200_handle::Peer &pr) { return pr.err_dev; }

// kernel launch as a programmatic dependent of the previous kernel in the stream (PDL); plain launch when `pdl` is false
template <typename... KArgs, typename... Args>
void launch_kernel(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, bool pdl, Args &&...args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = pdl ? 1 : 0;
    CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...));
}

// runs out = beta*out + alpha*op(A)*xz[0:n_in] on the handle's stream; x must already be in xz; y is a device pointer;
// peer_buf >= 0: the phase-2 epilogue also stores the rows into buffer `peer_buf` of every rank and publishes the epoch
template <typename T>
void run_product(hmb200_handle *h, PackedOp &po, const double *alpha, const double *beta, void *y, const int32_t *out_perm, int peer_buf = -1) {
    const Options &o = h->opt;
    const int threads = (o.consumer_warps + 1) * 32;
    const int ctas    = o.ctas > 0 ? o.ctas : h->n_sms;
    StreamParams p{};
    p.xz = h->xz; p.y = y; p.out_perm = out_perm;
    p.alpha_re = alpha[0]; p.alpha_im = h->es == 16 ? alpha[1] : 0.0;
    p.beta_re  = beta[0];  p.beta_im  = h->es == 16 ? beta[1] : 0.0;
    p.stage_bytes = o.stage_bytes; p.z_bytes = o.stage_z_bytes; p.n_stages = o.n_stages; p.consumer_warps = o.consumer_warps;
    CUDA_CHECK(cudaMemsetAsync(h->counters, 0, 64, h->stream));
    cudaEvent_t *ev = nullptr;
    if (h->timing) {
        if (h->ev.empty()) {
            h->ev.resize(4 * hmb200_handle::kTimingSlots, nullptr);
            for (auto &e : h->ev) CUDA_CHECK(cudaEventCreate(&e));
        }
        ev = h->ev.data() + 4 * (h->timed_products % hmb200_handle::kTimingSlots);
        h->timed_products++;
        CUDA_CHECK(cudaEventRecord(ev[0], h->stream));
    }
    if (po.n_units[0] > 0) {
        p.stream = po.stream[0]; p.stages = po.stages[0]; p.zcopies = po.zcopies[0]; p.units = po.units[0];
        p.n_units = po.n_units[0]; p.n_zcopies = po.n_zcopies[0]; p.counter = h->counters;
        size_t smem = stream_smem_bytes(1, h->es, o.stage_bytes, o.stage_z_bytes, o.n_stages, o.consumer_warps);
        hmatrix_stream_kernel<T, 1><<<(unsigned)std::min<uint32_t>((uint32_t)ctas, po.n_units[0]), threads, smem, h->stream>>>(p);
        h->launches++;
    }
    if (ev) CUDA_CHECK(cudaEventRecord(ev[1], h->stream));
    // the kernels after phase 1 are programmatic dependents of their predecessors (PDL): their CTAs are scheduled as SMs free
    // up and start streaming packets before the predecessor's last CTA has finished; per-kernel events would serialise them
    const bool pdl = h->use_pdl && !ev;
    bool prev_kernel = po.n_units[0] > 0;
    if (po.n_reduce > 0) {
        unsigned blocks = (po.n_reduce + 7) / 8;
        launch_kernel(reduce_partials_kernel<T>, dim3(blocks), dim3(256), 0, h->stream, pdl && prev_kernel, (const ReduceSeg *)po.reduce, po.n_reduce, (T *)h->xz);
        h->launches++;
        prev_kernel = true;
    }
    if (ev) CUDA_CHECK(cudaEventRecord(ev[2], h->stream));
    if (po.n_units[1] > 0) {
        p.stream = po.stream[1]; p.stages = po.stages[1]; p.zcopies = po.zcopies[1]; p.units = po.units[1];
        p.n_units = po.n_units[1]; p.n_zcopies = po.n_zcopies[1]; p.counter = h->counters + 1;
        if (peer_buf >= 0) { // replicate the slab into every rank's buffer `peer_buf` and publish the epoch
            const hmb200_handle::Peer &pr = h->peer;
            const size_t slab_b = (size_t)pr.slab_offset * sizeof(T);
            for (int q = 0; q < pr.n_ranks; q++) {
                if (q != pr.rank) p.peer_y[p.n_peers++] = pr.region[q] + (size_t)peer_buf * pr.buf_bytes + slab_b;
                p.peer_flags[p.n_flags++] = peer_ctrl_flags(pr, q, peer_buf) + pr.rank;
            }
            p.epoch = peer_ctrl_epoch(pr, peer_buf);
            p.done  = peer_ctrl_done(pr);
        }
        size_t smem = stream_smem_bytes(2, h->es, o.stage_bytes, o.stage_z_bytes, o.n_stages, o.consumer_warps);
        p.pdl = (pdl && prev_kernel) ? 1 : 0;
        launch_kernel(hmatrix_stream_kernel<T, 2>, dim3((unsigned)std::min<uint32_t>((uint32_t)ctas, po.n_units[1])), dim3(threads), smem, h->stream, p.pdl != 0, p);
        h->launches++;
    }
    if (ev) CUDA_CHECK(cudaEventRecord(ev[3], h->stream));
    CUDA_CHECK(cudaGetLastError());
}

// averages the per-kernel times of the last min(timed_products, kTimingSlots) products into ms[4]; returns the count
int collect_timing(hmb200_handle *h, float ms[4], bool last_only) {
    for (int i = 0; i < 4; i++) ms[i] = 0;
    const int64_t have = std::min<int64_t>(h->timed_products, hmb200_handle::kTimingSlots);
    if (have == 0 || h->ev.empty()) return 0;
    const int64_t first = last_only ? h->timed_products - 1 : h->timed_products - have;
    double acc[4] = {0, 0, 0, 0};
    int n = 0;
    for (int64_t k = first; k < h->timed_products; k++, n++) {
        cudaEvent_t *ev = h->ev.data() + 4 * (k % hmb200_handle::kTimingSlots);
        CUDA_CHECK(cudaEventSynchronize(ev[3]));
        float t;
        for (int i = 0; i < 3; i++) { CUDA_CHECK(cudaEventElapsedTime(&t, ev[i], ev[i + 1])); acc[i] += t; }
        CUDA_CHECK(cudaEventElapsedTime(&t, ev[0], ev[3]));
        acc[3] += t;
    }
    for (int i = 0; i < 4; i++) ms[i] = (float)(acc[i] / n);
    return n;
}

void clear_graphs(hmb200_handle *h) {
    for (auto &kv : h->graphs)
        if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
    h->graphs.clear();
}

// the cache is full: forget the keys seen once (one-off pointers, no graph yet), then the least recently used half of the rest.
// Replaying callers (a solver cycling through a few buffers) keep their graphs however many one-off pointers pass by.
void trim_graphs(hmb200_handle *h) {
    for (auto it = h->graphs.begin(); it != h->graphs.end();)
        it = it->second.exec ? std::next(it) : h->graphs.erase(it);
    if (h->graphs.size() <= hmb200_handle::kMaxGraphKeys / 2) return;
    std::vector<uint64_t> ticks;
    ticks.reserve(h->graphs.size());
    for (auto &kv : h->graphs) ticks.push_back(kv.second.last_use);
    std::nth_element(ticks.begin(), ticks.begin() + ticks.size() / 2, ticks.end());
    const uint64_t keep_from = ticks[ticks.size() / 2];
    for (auto it = h->graphs.begin(); it != h->graphs.end();) {
        if (it->second.last_use < keep_from) {
            cudaGraphExecDestroy(it->second.exec); // stream-ordered after the launches already enqueued
            it = h->graphs.erase(it);
        } else ++it;
    }
}

// issues `issue()` (a stream-ordered sequence on h->stream without host synchronisation) either eagerly or -- from the second
// time `key` is seen -- as a replay of a CUDA graph captured on a private stream
template <typename F>
void run_cached(hmb200_handle *h, const hmb200_handle::GraphKey &key, F &&issue) {
    if (!h->use_graphs || h->timing) {
        issue();
        return;
    }
    {   // the caller is capturing h->stream into a graph of its own (e.g. torch.cuda.graph): record the plain sequence there
        cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(h->stream, &st) != cudaSuccess) { cudaGetLastError(); st = cudaStreamCaptureStatusActive; }
        if (st != cudaStreamCaptureStatusNone) {
            issue();
            return;
        }
    }
    if (h->graphs.size() > hmb200_handle::kMaxGraphKeys && !h->graphs.count(key)) trim_graphs(h);
    auto &ge = h->graphs[key];
    ge.last_use = ++h->graph_tick;
    // first use of a key runs eagerly (one-off pointers never pay for a capture).  Fused device-resident products always recur
    // (the operator's own two buffers); the chunked host product ('U') is keyed on the CALLER's host pointers, which may well be
    // one-off, so it waits for the second sight like everything else.
    if (!ge.exec && ++ge.seen < 2 && (key.fused == 0 || key.op == 'U')) {
        issue();
        return;
    }
    if (!ge.exec) {
        if (!h->cap_stream) CUDA_CHECK(cudaStreamCreateWithFlags(&h->cap_stream, cudaStreamNonBlocking));
        cudaStream_t user = h->stream;
        const int64_t l0 = h->launches;
        cudaGraph_t g = nullptr;
        CUDA_CHECK(cudaStreamBeginCapture(h->cap_stream, cudaStreamCaptureModeThreadLocal));
        h->stream = h->cap_stream;
        try {
            issue();
        } catch (...) {
            h->stream = user;
            cudaStreamEndCapture(h->cap_stream, &g);
            if (g) cudaGraphDestroy(g);
            h->graphs.erase(key);
            throw;
        }
        h->stream = user;
        CUDA_CHECK(cudaStreamEndCapture(h->cap_stream, &g));
        ge.launches = h->launches - l0;
        h->launches = l0;
        cudaError_t ie = cudaGraphInstantiate(&ge.exec, g, 0);
        cudaGraphDestroy(g);
        if (ie != cudaSuccess) {
            h->graphs.erase(key);
            throw std::runtime_error(std::string("cudaGraphInstantiate failed: ") + cudaGetErrorString(ie));
        }
    }
    CUDA_CHECK(cudaGraphLaunch(ge.exec, h->stream));
    h->launches += ge.launches;
}

template <typename T>
int vector_product(hmb200_handle *h, char trans, const double *alpha, const void *in, const double *beta, void *out, int mem) {
    CUDA_CHECK(cudaSetDevice(h->device));
    PackedOp &po = select_op(h, trans);
    const size_t in_b = (size_t)po.n_in * sizeof(T), out_b = (size_t)po.n_out * sizeof(T);
    const bool beta_zero = beta[0] == 0.0 && (h->es == 8 || beta[1] == 0.0);
    if (mem == HMB200_DEVICE) {
        hmb200_handle::GraphKey key;
        std::memset(&key, 0, sizeof(key));
        key.op = po.op; key.in = in; key.out = out;
        key.a[0] = alpha[0]; key.a[1] = h->es == 16 ? alpha[1] : 0.0; key.b[0] = beta[0]; key.b[1] = h->es == 16 ? beta[1] : 0.0;
        run_cached(h, key, [&] {
            CUDA_CHECK(cudaMemcpyAsync(h->xz, in, in_b, cudaMemcpyDeviceToDevice, h->stream));
            run_product<T>(h, po, alpha, beta, out, nullptr);
        });
    } else {
        ensure_pinned(h, in_b + out_b + 64);
        ensure_dev(&h->ybuf, &h->ybuf_elems, (uint64_t)po.n_out, h->es);
        uint8_t *pin = (uint8_t *)h->pinned, *pout = pin + ((in_b + 63) & ~(size_t)63);
        CUDA_CHECK(cudaMemcpyAsync(h->xz, stage_in(in, pin, in_b), in_b, cudaMemcpyHostToDevice, h->stream));
        void *hout = host_is_pinned(out) ? out : (void *)pout;
        if (!beta_zero) {
            if (hout != out) staged_copy(pout, out, out_b);
            CUDA_CHECK(cudaMemcpyAsync(h->ybuf, hout, out_b, cudaMemcpyHostToDevice, h->stream));
        }
        run_product<T>(h, po, alpha, beta, h->ybuf, nullptr);
        CUDA_CHECK(cudaMemcpyAsync(hout, h->ybuf, out_b, cudaMemcpyDeviceToHost, h->stream));
        CUDA_CHECK(cudaStreamSynchronize(h->stream));
        if (hout != out) staged_copy(out, pout, out_b);
    }
    return 0;
}

// ---- multi-RHS on the DMMA panel kernels (kernels/mm_kernels.cuh) --------------------------------------------------------------
// warp tiling by the number of n-tiles (8 columns each) of the pass; the eight consumer warps form WM x WN, each MT x NT tiles:
//   n-tiles   phase 1 (16 m-tiles)          phase 2 (8 m-tiles)
//   1         8 x 1 warps, 2 x 1 tiles      8 x 1 warps, 1 x 1 tiles
//   2         8 x 1 warps, 2 x 2 tiles      8 x 1 warps, 1 x 2 tiles        (measured faster than 4 x 2 warps at 16 columns)
//   3, 4      4 x 2 warps, 4 x 2 tiles      4 x 2 warps, 2 x 2 tiles
//   5 .. 8    2 x 4 warps, 8 x 2 tiles      4 x 2 warps, 2 x 4 tiles        (real only; complex passes hold <= 4 n-tiles)
// complex, 3-4 n-tiles, phase 1: 2 x 4 warps, 8 x 1 tiles
template <typename T, int PHASE>
void launch_mm(int nt, unsigned grid, size_t smem, cudaStream_t st, const MmParams &p) {
    auto go = [&](auto kern) {
        CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<grid, kMmThreads, smem, st>>>(p);
    };
    if constexpr (PHASE == 1) {
        if (nt <= 1) go(hmatrix_mm_kernel<T, 1, 1, 2, 1>);
        else if (nt == 2) go(hmatrix_mm_kernel<T, 1, 1, 2, 2>);
        else if (nt <= 4) {
            if constexpr (sizeof(T) == 8) go(hmatrix_mm_kernel<T, 1, 2, 4, 2>);
            else go(hmatrix_mm_kernel<T, 1, 4, 8, 1>);
        } else if constexpr (sizeof(T) == 8) go(hmatrix_mm_kernel<T, 1, 4, 8, 2>);
        else throw std::runtime_error("internal: complex passes hold at most 32 columns");
    } else {
        if (nt <= 1) go(hmatrix_mm_kernel<T, 2, 1, 1, 1>);
        else if (nt == 2) go(hmatrix_mm_kernel<T, 2, 1, 1, 2>);
        else if (nt <= 4) go(hmatrix_mm_kernel<T, 2, 2, 2, 2>);
        else if constexpr (sizeof(T) == 8) go(hmatrix_mm_kernel<T, 2, 2, 2, 4>);
        else throw std::runtime_error("internal: complex passes hold at most 32 columns");
    }
}

// Y(:, c0:c0+nc) = beta*Y + alpha*op(A)*X(:, c0:c0+nc) for every pass of <= 64 (32 complex) columns.
//   x: device; row-major n_in x mu (perm_in == nullptr), or COLUMN-major user block gathered through perm_in
//      row_perm != nullptr (row-major x only): row i of the product's input is row row_perm[i] of x (user -> cluster numbering)
//   y: device, row-major n_out x mu (stride mu)
//   peer_buf >= 0: the bin epilogue also stores the rows into block `peer_buf` of every rank (row stride mu) and the last pass
//   publishes the epoch (fused all-gather of the row slabs)
template <typename T>
void run_mm_product(hmb200_handle *h, PackedOp &po, const double *alpha, const double *beta, const T *x, const int32_t *perm_in, T *y, int mu, int peer_buf = -1,
                    const int32_t *row_perm = nullptr) {
    const Options &o = h->opt;
    const int PC = mm_pass_cols(h->es);
    const int ctas = o.ctas > 0 ? o.ctas : h->n_sms;
    ensure_dev(&h->ws_mm, &h->ws_mm_elems, (po.mm_rows + 4) * (uint64_t)mm_max_ld(h->es), h->es);
    T *ws = (T *)h->ws_mm;
    for (int c0 = 0; c0 < mu; c0 += PC) {
        const int nc = std::min(PC, mu - c0), nc8 = (nc + 7) & ~7;
        const int ld = mm_ld(nc, h->es);
        const int nt = nc8 / 8;
        // the zero rows (operand of K padding) at this pass's stride; x rows of the pass
        CUDA_CHECK(cudaMemsetAsync(ws + (size_t)po.mm_zero_row * ld, 0, (size_t)kMmZeroRows * ld * sizeof(T), h->stream));
        {
            const size_t total = (size_t)po.n_in * nc;
            const unsigned g = (unsigned)std::min<size_t>((total + 255) / 256 + 1, 148 * 16);
            if (perm_in) mm_load_rows_kernel<T><<<g, 256, 0, h->stream>>>(x + (size_t)c0 * po.n_in, 0, po.n_in, nc, ws, ld, perm_in, po.n_in);
            else {
                const bool vec_ok = ((uintptr_t)(x + c0) % 16 == 0) && (((size_t)mu * sizeof(T)) % 16 == 0);
                const unsigned gb = (unsigned)std::min<size_t>(((size_t)po.n_in + 7) / 8, (size_t)h->n_sms * 32);
                mm_load_rows_fast_kernel<T><<<std::max(gb, 1u), dim3(32, 8), 0, h->stream>>>(x + c0, mu, po.n_in, nc, ws, ld, vec_ok ? 1 : 0, row_perm);
            }
            h->launches++;
        }
        CUDA_CHECK(cudaMemsetAsync(h->counters, 0, 64, h->stream));
        cudaEvent_t *ev = nullptr;
        if (h->timing) {
            if (h->ev.empty()) {
                h->ev.resize(4 * hmb200_handle::kTimingSlots, nullptr);
                for (auto &e : h->ev) CUDA_CHECK(cudaEventCreate(&e));
            }
            ev = h->ev.data() + 4 * (h->timed_products % hmb200_handle::kTimingSlots);
            h->timed_products++;
            CUDA_CHECK(cudaEventRecord(ev[0], h->stream));
        }
        MmParams p{};
        p.ws = ws; p.ld = ld; p.ncols = nc; p.y = y + c0; p.ldy = mu;
        p.alpha_re = alpha[0]; p.alpha_im = h->es == 16 ? alpha[1] : 0.0;
        p.beta_re  = beta[0];  p.beta_im  = h->es == 16 ? beta[1] : 0.0;
        if (po.n_units[2] > 0) {
            const MmLaunchCfg c = mm_launch_cfg(1, h->es, po.mm_packet_bytes[0], o.mm1_kg, ld, h->smem_optin);
            p.stream = po.stream[2]; p.stages = po.stages[2]; p.zcopies = po.zcopies[2]; p.units = po.units[2];
            p.n_units = po.n_units[2]; p.counter = h->counters; p.stage_bytes = c.stage_bytes; p.z_rows = c.z_rows; p.n_stages = c.n_stages;
            launch_mm<T, 1>(nt, (unsigned)std::min<uint32_t>((uint32_t)ctas, po.n_units[2]), c.smem, h->stream, p);
            h->launches++;
        }
        if (ev) CUDA_CHECK(cudaEventRecord(ev[1], h->stream));
        if (po.n_reduce_mm > 0) {
            unsigned blocks = std::min<unsigned>((po.n_reduce_mm + 7) / 8, 148 * 32);
            reduce_rows_kernel<T><<<blocks, 256, 0, h->stream>>>(po.reduce_mm, po.reduce_mm_rows, po.n_reduce_mm, ws, ld, nc8);
            h->launches++;
        }
        if (ev) CUDA_CHECK(cudaEventRecord(ev[2], h->stream));
        if (po.n_units[3] > 0) {
            const MmLaunchCfg c = mm_launch_cfg(2, h->es, po.mm_packet_bytes[1], o.mm2_kg, ld, h->smem_optin);
            p.stream = po.stream[3]; p.stages = po.stages[3]; p.zcopies = po.zcopies[3]; p.units = po.units[3];
            p.n_units = po.n_units[3]; p.counter = h->counters + 1; p.stage_bytes = c.stage_bytes; p.z_rows = c.z_rows; p.n_stages = c.n_stages;
            if (peer_buf >= 0) { // replicate the slab's rows into every rank's block; the epoch is published by the LAST pass only
                const hmb200_handle::Peer &pr = h->peer;
                const size_t slab_b = ((size_t)pr.slab_offset * mu + c0) * sizeof(T);
                for (int q = 0; q < pr.n_ranks; q++) {
                    if (q != pr.rank) p.peer_y[p.n_peers++] = pr.mm_region[q] + (size_t)peer_buf * pr.mm_buf_bytes + slab_b;
                    if (c0 + PC >= mu) p.peer_flags[p.n_flags++] = peer_ctrl_flags(pr, q, peer_buf) + pr.rank;
                }
                p.epoch = peer_ctrl_epoch(pr, peer_buf);
                p.done  = peer_ctrl_done(pr);
            }
            launch_mm<T, 2>(nt, (unsigned)std::min<uint32_t>((uint32_t)ctas, po.n_units[3]), c.smem, h->stream, p);
            h->launches++;
        }
        if (ev) CUDA_CHECK(cudaEventRecord(ev[3], h->stream));
        CUDA_CHECK(cudaGetLastError());
    }
}

// multi-RHS, row-major blocks (VirtualGlobalToLocalOperator::add_matrix_product_row_major).  With the multi-RHS stream
// packed (default) the factors are streamed once per pass of 64 (32 complex) right-hand sides through the DMMA kernels;
// otherwise (no_matmat) one streamed single-RHS product per column.
template <typename T>
int matrix_product(hmb200_handle *h, char trans, const double *alpha, const void *in, const double *beta, void *out, int mu, int mem) {
    CUDA_CHECK(cudaSetDevice(h->device));
    if (mu <= 0) throw std::runtime_error("mu must be positive");
    PackedOp &po = select_op(h, trans);
    const size_t in_b = (size_t)po.n_in * mu * sizeof(T), out_b = (size_t)po.n_out * mu * sizeof(T);
    const bool beta_zero = beta[0] == 0.0 && (h->es == 8 || beta[1] == 0.0);
    const T *din = (const T *)in;
    T *dout      = (T *)out;
    uint8_t *pout = nullptr;
    ensure_dev(&h->ybuf, &h->ybuf_elems, (uint64_t)po.n_out * (mem == HMB200_HOST ? (mu + 1) : 1), h->es);
    if (mem == HMB200_HOST) {
        ensure_pinned(h, in_b + out_b + 64);
        ensure_dev(&h->xbuf, &h->xbuf_elems, (uint64_t)po.n_in * mu, h->es);
        uint8_t *pin = (uint8_t *)h->pinned;
        pout         = pin + ((in_b + 63) & ~(size_t)63);
        CUDA_CHECK(cudaMemcpyAsync(h->xbuf, stage_in(in, pin, in_b), in_b, cudaMemcpyHostToDevice, h->stream));
        dout = (T *)h->ybuf + po.n_out; // [column scratch | row-major block]
        if (host_is_pinned(out)) pout = (uint8_t *)out;
        if (!beta_zero) {
            if ((void *)pout != out) staged_copy(pout, out, out_b);
            CUDA_CHECK(cudaMemcpyAsync(dout, pout, out_b, cudaMemcpyHostToDevice, h->stream));
        }
        din = (const T *)h->xbuf;
    }
    if (po.has_mm && mu > 1) {
        run_mm_product<T>(h, po, alpha, beta, din, nullptr, dout, mu);
    } else {
        T *ycol = (T *)h->ybuf;
        const int tb = 256;
        const unsigned gin = (unsigned)std::min<size_t>(((size_t)po.n_in + tb - 1) / tb + 1, 4096), gout = (unsigned)std::min<size_t>(((size_t)po.n_out + tb - 1) / tb + 1, 4096);
        for (int j = 0; j < mu; j++) {
            strided_gather_kernel<T><<<gin, tb, 0, h->stream>>>(din, po.n_in, mu, j, (T *)h->xz);
            if (!beta_zero) strided_gather_kernel<T><<<gout, tb, 0, h->stream>>>(dout, po.n_out, mu, j, ycol);
            run_product<T>(h, po, alpha, beta, ycol, nullptr);
            strided_scatter_kernel<T><<<gout, tb, 0, h->stream>>>(ycol, po.n_out, mu, j, dout);
            h->launches += beta_zero ? 2 : 3;
        }
    }
    CUDA_CHECK(cudaGetLastError());
    if (mem == HMB200_HOST) {
        CUDA_CHECK(cudaMemcpyAsync(pout, dout, out_b, cudaMemcpyDeviceToHost, h->stream));
        CUDA_CHECK(cudaStreamSynchronize(h->stream));
        if ((void *)pout != out) staged_copy(out, pout, out_b);
    }
    return 0;
}

// out += A[:, offset:offset+size] * in (row-major size x mu window of the source, `offset` in source-tree numbering):
// clip to the operator's source range and extend by zero, as the reference does on the host
// (core/distributed_operator/implementations/local_to_local_operators/hmatrix.hpp:34-51)
template <typename T>
int sub_matrix_product(hmb200_handle *h, const void *in, void *out, int mu, int offset, int size, int mem) {
    CUDA_CHECK(cudaSetDevice(h->device));
    if (mu <= 0) throw std::runtime_error("mu must be positive");
    if (size < 0) throw std::runtime_error("negative window size");
    PackedOp &po = h->ops[0];
    const int src_off = h->source_offset, src_end = src_off + po.n_in;
    const int w0 = std::max(offset, src_off), w1 = std::min(src_end, offset + size);
    if (w1 - w0 <= 0) return 0; // empty intersection: out is left untouched (beta = 1)
    const double one[2] = {1.0, 0.0};
    const size_t win_b = (size_t)(w1 - w0) * mu * sizeof(T), out_b = (size_t)po.n_out * mu * sizeof(T);
    const T *win = (const T *)in + (size_t)(w0 - offset) * mu;
    ensure_dev(&h->xsub, &h->xsub_elems, (uint64_t)po.n_in * mu, h->es);
    T *xs = (T *)h->xsub;
    const bool whole = (w0 == src_off && w1 == src_end);
    if (!whole) CUDA_CHECK(cudaMemsetAsync(xs, 0, (size_t)po.n_in * mu * sizeof(T), h->stream));
    if (mem == HMB200_DEVICE) {
        CUDA_CHECK(cudaMemcpyAsync(xs + (size_t)(w0 - src_off) * mu, win, win_b, cudaMemcpyDeviceToDevice, h->stream));
        return matrix_product<T>(h, 'N', one, xs, one, out, mu, HMB200_DEVICE);
    }
    ensure_pinned(h, win_b + out_b + 64);
    ensure_dev(&h->ybuf2, &h->ybuf2_elems, (uint64_t)po.n_out * mu, h->es);
    uint8_t *pin = (uint8_t *)h->pinned, *pout = pin + ((win_b + 63) & ~(size_t)63);
    CUDA_CHECK(cudaMemcpyAsync(xs + (size_t)(w0 - src_off) * mu, stage_in(win, pin, win_b), win_b, cudaMemcpyHostToDevice, h->stream));
    void *hout = host_is_pinned(out) ? out : (void *)pout;
    if (hout != out) staged_copy(pout, out, out_b);
    CUDA_CHECK(cudaMemcpyAsync(h->ybuf2, hout, out_b, cudaMemcpyHostToDevice, h->stream));
    matrix_product<T>(h, 'N', one, xs, one, h->ybuf2, mu, HMB200_DEVICE);
    CUDA_CHECK(cudaMemcpyAsync(hout, h->ybuf2, out_b, cudaMemcpyDeviceToHost, h->stream));
    CUDA_CHECK(cudaStreamSynchronize(h->stream));
    if (hout != out) staged_copy(out, pout, out_b);
    return 0;
}

// y_user = A x_user (alpha = 1, beta = 0) with the two permutations on the device:
// gather (user_to_cluster) -> product -> scatter fused into the phase-2 epilogue (cluster_to_user)
template <typename T>
int matvec_user(hmb200_handle *h, const void *x, void *y, int mem, int mu) {
    CUDA_CHECK(cudaSetDevice(h->device));
    if (!h->perm_t || !h->perm_s) throw std::runtime_error("hmb200_set_permutations has not been called");
    PackedOp &po = h->ops[0];
    const size_t in_b = (size_t)po.n_in * mu * sizeof(T), out_b = (size_t)po.n_out * mu * sizeof(T);
    const double one[2] = {1.0, 0.0}, zero[2] = {0.0, 0.0};
    const T *dx = (const T *)x;
    T *dy       = (T *)y;
    uint8_t *pout = nullptr;
    if (mem == HMB200_HOST) {
        ensure_pinned(h, in_b + out_b + 64);
        ensure_dev(&h->xbuf, &h->xbuf_elems, (uint64_t)po.n_in * mu, h->es);
        ensure_dev(&h->ybuf, &h->ybuf_elems, (uint64_t)po.n_out * mu, h->es);
        uint8_t *pin = (uint8_t *)h->pinned;
        pout         = pin + ((in_b + 63) & ~(size_t)63);
        CUDA_CHECK(cudaMemcpyAsync(h->xbuf, stage_in(x, pin, in_b), in_b, cudaMemcpyHostToDevice, h->stream));
        if (host_is_pinned(y)) pout = (uint8_t *)y;
        dx = (const T *)h->xbuf;
        dy = (T *)h->ybuf;
    }
    const int tb = 256;
    const unsigned gin = (unsigned)std::min<size_t>(((size_t)po.n_in + tb - 1) / tb + 1, 4096);
    if (po.has_mm && mu > 1) {
        // column-major user blocks: the permutations and the row-major <-> column-major layout change are fused into the
        // workspace load and the final store (reference: per-column permutation + 3 transposes, add_hmatrix_matrix_product.hpp:20-90,190-204)
        ensure_dev(&h->ybuf2, &h->ybuf2_elems, (uint64_t)po.n_out * mu, h->es);
        T *yc = (T *)h->ybuf2;
        run_mm_product<T>(h, po, one, zero, dx, h->perm_s, yc, mu);
        const size_t total = (size_t)po.n_out * mu;
        mm_store_user_kernel<T><<<(unsigned)std::min<size_t>((total + 255) / 256 + 1, 148 * 16), 256, 0, h->stream>>>(yc, po.n_out, mu, dy, h->perm_t, po.n_out);
        h->launches++;
    } else
    for (int j = 0; j < mu; j++) {
        permute_gather_kernel<T><<<gin, tb, 0, h->stream>>>(h->perm_s, 0, po.n_in, dx + (size_t)j * po.n_in, (T *)h->xz, 1);
        h->launches++;
        run_product<T>(h, po, one, zero, dy + (size_t)j * po.n_out, h->perm_t);
    }
    if (mem == HMB200_HOST) {
        CUDA_CHECK(cudaMemcpyAsync(pout, dy, out_b, cudaMemcpyDeviceToHost, h->stream));
        CUDA_CHECK(cudaStreamSynchronize(h->stream));
        if ((void *)pout != y) staged_copy(y, pout, out_b);
    }
    return 0;
}

// y_replicated = alpha * A_slab * in on EVERY rank: the slab product whose phase-2 epilogue stores each output row into the
// replicated vector of all ranks (peer memory over NVLink), then a wait for the epochs of all ranks.  Buffers alternate so
// that a rank running one product ahead never overwrites a vector its peers may still be reading.
template <typename T>
int fused_product(hmb200_handle *h, const double *alpha, const void *in, void **y_out) {
    CUDA_CHECK(cudaSetDevice(h->device));
    hmb200_handle::Peer &pr = h->peer;
    if (!pr.attached) throw std::runtime_error("hmb200_peer_attach has not been called");
    PackedOp &po = h->ops[0];
    const int buf = (int)(pr.products & 1);
    pr.products++;
    uint8_t *ybuf = pr.region[pr.rank] + (size_t)buf * pr.buf_bytes;
    const double zero[2] = {0.0, 0.0};
    const size_t in_b = (size_t)po.n_in * sizeof(T);
    hmb200_handle::GraphKey key;
    std::memset(&key, 0, sizeof(key));
    key.op = 'N'; key.in = in; key.out = ybuf; key.fused = 1 + buf;
    key.a[0] = alpha[0]; key.a[1] = h->es == 16 ? alpha[1] : 0.0;
    run_cached(h, key, [&] {
        CUDA_CHECK(cudaMemcpyAsync(h->xz, in, in_b, cudaMemcpyDeviceToDevice, h->stream));
        run_product<T>(h, po, alpha, zero, ybuf + (size_t)pr.slab_offset * sizeof(T), nullptr, buf);
        launch_kernel(peer_wait_kernel, dim3(1), dim3(32), 0, h->stream, h->use_pdl && !h->timing && po.n_units[1] > 0,
                      (const unsigned long long *)peer_ctrl_flags(pr, pr.rank, buf), pr.n_ranks, (const unsigned long long *)peer_ctrl_epoch(pr, buf), peer_ctrl_err(pr), pr.wait_cycles);
        h->launches++;
        CUDA_CHECK(cudaGetLastError());
    });
    if (y_out) *y_out = ybuf;
    return 0;
}

// all-gather of the local row slabs into the replicated buffer of every rank (the solver's operator input), then the wait
template <typename T>
int peer_allgather(hmb200_handle *h, const void *local_slab, void **y_out, int64_t elem_offset = -1, int64_t count = -1) {
    CUDA_CHECK(cudaSetDevice(h->device));
    hmb200_handle::Peer &pr = h->peer;
    if (!pr.attached) throw std::runtime_error("hmb200_peer_attach has not been called");
    if (elem_offset < 0) { elem_offset = pr.slab_offset; count = h->n_rows; } // default: this rank's row slab
    if (count < 0 || elem_offset + count > pr.n_global) throw std::runtime_error("range does not fit the replicated vector");
    const int buf = (int)(pr.products & 1);
    pr.products++;
    PeerGatherParams p{};
    p.src = local_slab; p.n_local = count; p.n_targets = pr.n_ranks;
    for (int q = 0; q < pr.n_ranks; q++) {
        p.dst[q]   = pr.region[q] + (size_t)buf * pr.buf_bytes + (size_t)elem_offset * sizeof(T);
        p.flags[q] = peer_ctrl_flags(pr, q, buf) + pr.rank;
    }
    p.epoch = peer_ctrl_epoch(pr, buf);
    p.done  = peer_ctrl_done(pr);
    const unsigned grid = (unsigned)std::max<int64_t>(1, std::min<int64_t>((count + 2047) / 2048, 64));
    peer_gather_kernel<T><<<grid, 256, 0, h->stream>>>(p);
    peer_wait_kernel<<<1, 32, 0, h->stream>>>(peer_ctrl_flags(pr, pr.rank, buf), pr.n_ranks, peer_ctrl_epoch(pr, buf), peer_ctrl_err(pr), pr.wait_cycles);
    h->launches += 2;
    CUDA_CHECK(cudaGetLastError());
    if (y_out) *y_out = pr.region[pr.rank] + (size_t)buf * pr.buf_bytes;
    return 0;
}

// y_local = beta * y_local + sum over the ranks q (ascending) of [alpha * op(A_q) x_q] restricted to this rank's slab: the transposed
// local-to-local product.  The rank's full-length contribution goes into ITS OWN replicated buffer, the epoch is published, and
// after the wait a kernel pulls the rank's slab of every contribution over NVLink.  Same buffers, epochs and alternation as the
// fused all-gather, so the two can be mixed freely: a rank overwrites a buffer only after the wait of the product in between,
// i.e. after every peer has finished reading it.
template <typename T>
int peer_exchange_sum(hmb200_handle *h, char trans, const double *alpha, const void *x_local, const double *beta, void *y_local) {
    CUDA_CHECK(cudaSetDevice(h->device));
    hmb200_handle::Peer &pr = h->peer;
    if (!pr.attached) throw std::runtime_error("hmb200_peer_attach has not been called");
    if (trans != 'T' && trans != 'C') throw std::runtime_error("trans must be 'T' or 'C'");
    PackedOp &po = select_op(h, trans);
    if (po.n_out != pr.n_global || po.n_in != h->n_rows) throw std::runtime_error("the transposed exchange needs a square operator whose columns are the replicated vector");
    const int buf = (int)(pr.products & 1);
    pr.products++;
    uint8_t *ybuf = pr.region[pr.rank] + (size_t)buf * pr.buf_bytes;
    const double zero[2] = {0.0, 0.0};
    vector_product<T>(h, trans, alpha, x_local, zero, ybuf, HMB200_DEVICE);
    PeerGatherParams g{};
    g.n_targets = pr.n_ranks;
    for (int q = 0; q < pr.n_ranks; q++) g.flags[q] = peer_ctrl_flags(pr, q, buf) + pr.rank;
    g.epoch = peer_ctrl_epoch(pr, buf);
    peer_publish_kernel<<<1, 32, 0, h->stream>>>(g);
    peer_wait_kernel<<<1, 32, 0, h->stream>>>(peer_ctrl_flags(pr, pr.rank, buf), pr.n_ranks, peer_ctrl_epoch(pr, buf), peer_ctrl_err(pr), pr.wait_cycles);
    PeerSumParams sp{};
    sp.n_ranks = pr.n_ranks;
    sp.n = h->n_rows;
    for (int q = 0; q < pr.n_ranks; q++) sp.src[q] = pr.region[q] + (size_t)buf * pr.buf_bytes + (size_t)pr.slab_offset * sizeof(T);
    sp.out = y_local;
    sp.beta[0] = beta[0]; sp.beta[1] = h->es == 16 ? beta[1] : 0.0;
    const unsigned grid = (unsigned)std::max<int64_t>(1, std::min<int64_t>(((int64_t)h->n_rows + 255) / 256, 148 * 4));
    peer_sum_slabs_kernel<T><<<grid, 256, 0, h->stream>>>(sp);
    h->launches += 3;
    CUDA_CHECK(cudaGetLastError());
    return 0;
}

// Y_replicated (n_global x mu, row-major) = alpha * A_slab * in on EVERY rank: the multi-RHS slab product whose bin epilogue stores
// each tile's rows into the replicated block of all ranks, then the wait for the epochs of all ranks (the MPI_Allgatherv of
// add_distributed_operator_matrix_product_row_major_global_to_global.hpp:75 inside the kernel)
template <typename T>
int fused_mm_product(hmb200_handle *h, const double *alpha, const void *in, int mu, void **y_out) {
    CUDA_CHECK(cudaSetDevice(h->device));
    hmb200_handle::Peer &pr = h->peer;
    if (!pr.attached || !pr.mm_attached) throw std::runtime_error("hmb200_peer_attach / hmb200_peer_mm_attach have not been called");
    if (mu != pr.mm_mu) throw std::runtime_error("the replicated blocks were created for another number of right-hand sides");
    PackedOp &po = h->ops[0];
    if (!po.has_mm) throw std::runtime_error("the multi-RHS streams were not packed (no_matmat)");
    const int buf = (int)(pr.products & 1);
    pr.products++;
    T *yb = reinterpret_cast<T *>(pr.mm_region[pr.rank] + (size_t)buf * pr.mm_buf_bytes);
    const double zero[2] = {0.0, 0.0};
    run_mm_product<T>(h, po, alpha, zero, (const T *)in, nullptr, yb + (size_t)pr.slab_offset * mu, mu, buf);
    peer_wait_kernel<<<1, 32, 0, h->stream>>>(peer_ctrl_flags(pr, pr.rank, buf), pr.n_ranks, peer_ctrl_epoch(pr, buf), peer_ctrl_err(pr), pr.wait_cycles);
    h->launches++;
    CUDA_CHECK(cudaGetLastError());
    if (y_out) *y_out = yb;
    return 0;
}

// y_user[r0 : r0 + rcnt] of y = A x on every rank, from this rank's chunk x_user[c0 : c0 + cnt] of the replicated input, as ONE
// cached CUDA graph: H2D of the chunk, all-gather of the chunks over the peer regions, global_to_root_cluster, the slab product
// with the fused all-gather of its rows, root_cluster_to_global of the rank's chunk of the result, D2H.  The distributed
// global-to-global product (add_distributed_operator_vector_product_global_to_global.hpp:96-118) with n / P elements per rank and
// direction over PCIe.  Host pointers must be page-locked; synchronous on return.
template <typename T>
int user_matvec_chunked(hmb200_handle *h, const void *x_host, int64_t c0, int64_t cnt, void *y_host, int64_t r0, int64_t rcnt) {
    CUDA_CHECK(cudaSetDevice(h->device));
    hmb200_handle::Peer &pr = h->peer;
    if (!pr.attached) throw std::runtime_error("hmb200_peer_attach has not been called");
    if (!h->gperm_s || !h->ginv_t) throw std::runtime_error("hmb200_set_global_permutations has not been called");
    PackedOp &po = h->ops[0];
    if ((int64_t)po.n_in != pr.n_global) throw std::runtime_error("the chunked user-numbering product needs a square operator");
    if (c0 < 0 || cnt < 0 || c0 + cnt > pr.n_global || r0 < 0 || rcnt < 0 || r0 + rcnt > pr.n_global) throw std::runtime_error("chunk outside the vector");
    if ((cnt && !host_is_pinned(x_host)) || (rcnt && !host_is_pinned(y_host))) throw std::runtime_error("the host chunks must be page-locked (cudaHostAlloc / cudaHostRegister)");
    const int buf_x = (int)(pr.products & 1);
    pr.products++;
    const int buf_y = (int)(pr.products & 1);
    pr.products++;
    hmb200_handle::GraphKey key;
    std::memset(&key, 0, sizeof(key));
    key.op = 'U'; key.in = x_host; key.out = y_host; key.fused = 16 + buf_x;
    key.a[0] = (double)c0; key.a[1] = (double)cnt; key.b[0] = (double)r0; key.b[1] = (double)rcnt;
    const double one[2] = {1.0, 0.0}, zero[2] = {0.0, 0.0};
    run_cached(h, key, [&] {
        if (cnt) CUDA_CHECK(cudaMemcpyAsync(h->gx, x_host, (size_t)cnt * sizeof(T), cudaMemcpyHostToDevice, h->stream));
        {   // all-gather of the user-numbered chunks into buffer buf_x of every rank
            PeerGatherParams p{};
            p.src = h->gx; p.n_local = cnt; p.n_targets = pr.n_ranks;
            for (int q = 0; q < pr.n_ranks; q++) {
                p.dst[q]   = pr.region[q] + (size_t)buf_x * pr.buf_bytes + (size_t)c0 * sizeof(T);
                p.flags[q] = peer_ctrl_flags(pr, q, buf_x) + pr.rank;
            }
            p.epoch = peer_ctrl_epoch(pr, buf_x);
            p.done  = peer_ctrl_done(pr);
            const unsigned grid = (unsigned)std::max<int64_t>(1, std::min<int64_t>((cnt + 2047) / 2048, 64));
            peer_gather_kernel<T><<<grid, 256, 0, h->stream>>>(p);
            peer_wait_kernel<<<1, 32, 0, h->stream>>>(peer_ctrl_flags(pr, pr.rank, buf_x), pr.n_ranks, peer_ctrl_epoch(pr, buf_x), peer_ctrl_err(pr), pr.wait_cycles);
        }
        const int tb = 256;
        const T *xu = reinterpret_cast<const T *>(pr.region[pr.rank] + (size_t)buf_x * pr.buf_bytes);
        permute_gather_kernel<T><<<(unsigned)std::min<size_t>(((size_t)po.n_in + tb - 1) / tb + 1, 4096), tb, 0, h->stream>>>(h->gperm_s, 0, po.n_in, xu, (T *)h->xz, 1);
        uint8_t *yb = pr.region[pr.rank] + (size_t)buf_y * pr.buf_bytes;
        run_product<T>(h, po, one, zero, yb + (size_t)pr.slab_offset * sizeof(T), nullptr, buf_y);
        peer_wait_kernel<<<1, 32, 0, h->stream>>>(peer_ctrl_flags(pr, pr.rank, buf_y), pr.n_ranks, peer_ctrl_epoch(pr, buf_y), peer_ctrl_err(pr), pr.wait_cycles);
        if (rcnt) {
            permute_gather_kernel<T><<<(unsigned)std::min<size_t>(((size_t)rcnt + tb - 1) / tb + 1, 4096), tb, 0, h->stream>>>(h->ginv_t + r0, 0, (int)rcnt, (const T *)yb, (T *)h->gy, 1);
            CUDA_CHECK(cudaMemcpyAsync(y_host, h->gy, (size_t)rcnt * sizeof(T), cudaMemcpyDeviceToHost, h->stream));
        }
        h->launches += 5;
        CUDA_CHECK(cudaGetLastError());
    });
    CUDA_CHECK(cudaStreamSynchronize(h->stream));
    const int err = *(volatile int *)pr.err_host;
    if (err) throw std::runtime_error("a peer did not arrive within the wait bound (status " + std::to_string(err) + ")");
    return 0;
}

// rows [r0, r0 + rcnt) of Y_user = A X_user on every rank, from this rank's rows [c0, c0 + cnt) of the replicated column-major
// user block X: strided H2D of the chunk, layout change to rows, all-gather of the chunks over the multi-RHS peer blocks, row
// gather to cluster numbering fused into the workspace load, slab product with the fused all-gather of its rows, gather of the
// rank's chunk of the user-numbered result, strided D2H.  The distributed matrix product of
// add_distributed_operator_matrix_product_global_to_global.hpp:138-280 with n mu / P elements per rank and direction over PCIe.
template <typename T>
int user_matmat_chunked(hmb200_handle *h, const void *x_host, int64_t ldx, int64_t c0, int64_t cnt, void *y_host, int64_t ldy, int64_t r0, int64_t rcnt, int mu) {
    CUDA_CHECK(cudaSetDevice(h->device));
    hmb200_handle::Peer &pr = h->peer;
    if (!pr.attached || !pr.mm_attached) throw std::runtime_error("hmb200_peer_attach / hmb200_peer_mm_attach have not been called");
    if (mu != pr.mm_mu) throw std::runtime_error("the replicated blocks were created for another number of right-hand sides");
    if (!h->gperm_s || !h->ginv_t) throw std::runtime_error("hmb200_set_global_permutations has not been called");
    PackedOp &po = h->ops[0];
    if (!po.has_mm) throw std::runtime_error("the multi-RHS streams were not packed (no_matmat)");
    if ((int64_t)po.n_in != pr.n_global) throw std::runtime_error("the chunked user-numbering product needs a square operator");
    if (c0 < 0 || cnt < 0 || c0 + cnt > pr.n_global || r0 < 0 || rcnt < 0 || r0 + rcnt > pr.n_global || ldx < cnt || ldy < rcnt) throw std::runtime_error("chunk outside the block");
    if ((cnt && !host_is_pinned((const T *)x_host + c0)) || (rcnt && !host_is_pinned((T *)y_host + r0))) throw std::runtime_error("the host blocks must be page-locked (cudaHostAlloc / cudaHostRegister)");
    const int buf_x = (int)(pr.products & 1);
    pr.products++;
    const int buf_y = (int)(pr.products & 1);
    pr.products++;
    const uint64_t chunk = (uint64_t)std::max<int64_t>(std::max(cnt, rcnt), 1) * mu;
    ensure_dev(&h->xbuf, &h->xbuf_elems, 2 * chunk, h->es); // [column-major chunk | the same as rows]
    ensure_dev(&h->ybuf2, &h->ybuf2_elems, chunk, h->es);
    T *xcol = (T *)h->xbuf, *xrow = xcol + chunk, *ycol = (T *)h->ybuf2;
    if (cnt) {
        CUDA_CHECK(cudaMemcpy2DAsync(xcol, (size_t)cnt * sizeof(T), (const T *)x_host + c0, (size_t)ldx * sizeof(T), (size_t)cnt * sizeof(T), (size_t)mu, cudaMemcpyHostToDevice, h->stream));
        const unsigned g = (unsigned)std::min<size_t>(((size_t)cnt * mu + 255) / 256 + 1, 148 * 16);
        mm_load_rows_kernel<T><<<g, 256, 0, h->stream>>>(xcol, 0, (int)cnt, mu, xrow, mu, nullptr, cnt); // column-major chunk -> rows
    }
    {   // all-gather of the row chunks (user numbering) into block buf_x of every rank
        PeerGatherParams p{};
        p.src = xrow; p.n_local = cnt * mu; p.n_targets = pr.n_ranks;
        for (int q = 0; q < pr.n_ranks; q++) {
            p.dst[q]   = pr.mm_region[q] + (size_t)buf_x * pr.mm_buf_bytes + (size_t)c0 * mu * sizeof(T);
            p.flags[q] = peer_ctrl_flags(pr, q, buf_x) + pr.rank;
        }
        p.epoch = peer_ctrl_epoch(pr, buf_x);
        p.done  = peer_ctrl_done(pr);
        const unsigned grid = (unsigned)std::max<int64_t>(1, std::min<int64_t>((cnt * mu + 2047) / 2048, 4 * h->n_sms));
        peer_gather_kernel<T><<<grid, 256, 0, h->stream>>>(p);
        peer_wait_kernel<<<1, 32, 0, h->stream>>>(peer_ctrl_flags(pr, pr.rank, buf_x), pr.n_ranks, peer_ctrl_epoch(pr, buf_x), peer_ctrl_err(pr), pr.wait_cycles);
    }
    const T *xu = reinterpret_cast<const T *>(pr.mm_region[pr.rank] + (size_t)buf_x * pr.mm_buf_bytes);
    T *yb       = reinterpret_cast<T *>(pr.mm_region[pr.rank] + (size_t)buf_y * pr.mm_buf_bytes);
    const double one[2] = {1.0, 0.0}, zero[2] = {0.0, 0.0};
    run_mm_product<T>(h, po, one, zero, xu, nullptr, yb + (size_t)pr.slab_offset * mu, mu, buf_y, h->gperm_s);
    peer_wait_kernel<<<1, 32, 0, h->stream>>>(peer_ctrl_flags(pr, pr.rank, buf_y), pr.n_ranks, peer_ctrl_epoch(pr, buf_y), peer_ctrl_err(pr), pr.wait_cycles);
    if (rcnt) {
        const unsigned g = (unsigned)std::min<size_t>(((size_t)rcnt * mu + 255) / 256 + 1, 148 * 16);
        mm_gather_user_chunk_kernel<T><<<g, 256, 0, h->stream>>>(yb, mu, h->ginv_t, r0, (int)rcnt, ycol);
        CUDA_CHECK(cudaMemcpy2DAsync((T *)y_host + r0, (size_t)ldy * sizeof(T), ycol, (size_t)rcnt * sizeof(T), (size_t)rcnt * sizeof(T), (size_t)mu, cudaMemcpyDeviceToHost, h->stream));
    }
    h->launches += 6;
    CUDA_CHECK(cudaGetLastError());
    CUDA_CHECK(cudaStreamSynchronize(h->stream));
    const int err = *(volatile int *)pr.err_host;
    if (err) throw std::runtime_error("a peer did not arrive within the wait bound (status " + std::to_string(err) + ")");
    return 0;
}

template <typename F>
int guarded(F &&f) {
    try {
        return f();
    } catch (const std::exception &e) {
        g_last_error = e.what();
        return 1;
    }
}

} // namespace

// ------------------------------------------------------------------------------------------------------------
// extern "C"
// ------------------------------------------------------------------------------------------------------------
extern "C" {

hmb200_handle *hmb200_create_d(const hmb200_flat *flat, int device, void *stream, const hmb200_options *opt) {
    try {
        return create_impl(flat, device, stream, opt, 8);
    } catch (const std::exception &e) {
        g_last_error = e.what();
        return nullptr;
    }
}
hmb200_handle *hmb200_create_z(const hmb200_flat *flat, int device, void *stream, const hmb200_options *opt) {
    try {
        return create_impl(flat, device, stream, opt, 16);
    } catch (const std::exception &e) {
        g_last_error = e.what();
        return nullptr;
    }
}

void hmb200_destroy(hmb200_handle *h) {
    if (!h) return;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream); else cudaDeviceSynchronize();
    clear_graphs(h);
    if (h->cap_stream) cudaStreamDestroy(h->cap_stream);
    for (int q = 0; q < h->peer.n_ranks; q++) {
        if (!h->peer.region[q]) continue;
        if (q == h->peer.rank) cudaFree(h->peer.region[q]); else cudaIpcCloseMemHandle(h->peer.region[q]);
    }
    for (int q = 0; q < h->peer.n_ranks; q++) {
        if (!h->peer.mm_region[q]) continue;
        if (q == h->peer.rank) cudaFree(h->peer.mm_region[q]); else cudaIpcCloseMemHandle(h->peer.mm_region[q]);
    }
    if (h->peer.err_host) cudaFreeHost(h->peer.err_host);
    for (auto &po : h->ops) free_op(po);
    cudaFree(h->xz);
    cudaFree(h->ybuf);
    cudaFree(h->xbuf);
    cudaFree(h->xsub);
    cudaFree(h->ws_mm);
    cudaFree(h->ybuf2);
    cudaFree(h->counters);
    cudaFree(h->perm_t);
    cudaFree(h->perm_s);
    cudaFree(h->gperm_s);
    cudaFree(h->ginv_t);
    cudaFree(h->gx);
    cudaFree(h->gy);
    if (h->pinned) cudaFreeHost(h->pinned);
    for (auto &ev : h->ev)
        if (ev) cudaEventDestroy(ev);
    h->ev.clear();
    delete h;
}

int hmb200_get_info(const hmb200_handle *h, hmb200_info *info) {
    if (!h || !info) { g_last_error = "null argument"; return 1; }
    *info = h->info;
    return 0;
}
int hmb200_set_stream(hmb200_handle *h, void *stream) {
    if (!h) { g_last_error = "null handle"; return 1; }
    h->stream = (cudaStream_t)stream;
    return 0;
}

int hmb200_add_vector_product_d(hmb200_handle *h, char trans, const double *alpha, const double *in, const double *beta, double *out, int mem) {
    return guarded([&] { if (!h || h->es != 8) throw std::runtime_error("handle is not a real (double) engine"); return vector_product<double>(h, trans, alpha, in, beta, out, mem); });
}
int hmb200_add_vector_product_z(hmb200_handle *h, char trans, const double *alpha, const double *in, const double *beta, double *out, int mem) {
    return guarded([&] { if (!h || h->es != 16) throw std::runtime_error("handle is not a complex engine"); return vector_product<double2>(h, trans, alpha, in, beta, out, mem); });
}
int hmb200_add_matrix_product_row_major_d(hmb200_handle *h, char trans, const double *alpha, const double *in, const double *beta, double *out, int mu, int mem) {
    return guarded([&] { if (!h || h->es != 8) throw std::runtime_error("handle is not a real (double) engine"); return matrix_product<double>(h, trans, alpha, in, beta, out, mu, mem); });
}
int hmb200_add_matrix_product_row_major_z(hmb200_handle *h, char trans, const double *alpha, const double *in, const double *beta, double *out, int mu, int mem) {
    return guarded([&] { if (!h || h->es != 16) throw std::runtime_error("handle is not a complex engine"); return matrix_product<double2>(h, trans, alpha, in, beta, out, mu, mem); });
}
int hmb200_peer_create(hmb200_handle *h, int n_ranks, int rank, int64_t n_global, int slab_offset, unsigned char *ipc_handle_out) {
    return guarded([&] {
        if (!h || !ipc_handle_out) throw std::runtime_error("null argument");
        if (n_ranks < 1 || n_ranks > kMaxPeers || rank < 0 || rank >= n_ranks) throw std::runtime_error("between 1 and 8 ranks are supported");
        if (slab_offset < 0 || (int64_t)slab_offset + h->n_rows > n_global) throw std::runtime_error("slab does not fit the replicated vector");
        CUDA_CHECK(cudaSetDevice(h->device));
        hmb200_handle::Peer &pr = h->peer;
        if (pr.region[pr.rank]) throw std::runtime_error("peer region already created");
        pr.n_ranks = n_ranks; pr.rank = rank; pr.n_global = n_global; pr.slab_offset = slab_offset;
        {   // a peer that never arrives must not hang the GPU for ever: the waits give up after this many SM clocks and set the
            // error word (hmb200_peer_status); ranks may legitimately be seconds apart (host-side work), hence the generous default
            int khz = 0;
            cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, h->device);
            const char *ev = std::getenv("HMB200_PEER_TIMEOUT_S");
            const double secs = ev && std::atof(ev) > 0 ? std::atof(ev) : 120.0;
            pr.wait_cycles = (long long)(secs * 1e3 * (khz > 0 ? khz : 1965000));
        }
        pr.buf_bytes    = (((size_t)n_global * h->es) + 255) & ~(size_t)255;
        pr.region_bytes = 2 * pr.buf_bytes + 512 + sizeof(double) * 2 * kMaxPeers * kPeerReduceMax;
        void *r = nullptr;
        CUDA_CHECK(cudaMalloc(&r, pr.region_bytes));
        CUDA_CHECK(cudaMemset(r, 0, pr.region_bytes));
        pr.region[rank] = (uint8_t *)r;
        CUDA_CHECK(cudaHostAlloc((void **)&pr.err_host, 64, cudaHostAllocMapped));
        std::memset(pr.err_host, 0, 64);
        CUDA_CHECK(cudaHostGetDevicePointer((void **)&pr.err_dev, pr.err_host, 0));
        cudaIpcMemHandle_t ih;
        CUDA_CHECK(cudaIpcGetMemHandle(&ih, r));
        static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
        std::memcpy(ipc_handle_out, &ih, 64);
        return 0;
    });
}
int hmb200_peer_attach(hmb200_handle *h, const unsigned char *ipc_handles) {
    return guarded([&] {
        if (!h) throw std::runtime_error("null handle");
        hmb200_handle::Peer &pr = h->peer;
        if (!pr.region[pr.rank]) throw std::runtime_error("hmb200_peer_create has not been called");
        if (pr.n_ranks > 1 && !ipc_handles) throw std::runtime_error("null argument");
        CUDA_CHECK(cudaSetDevice(h->device));
        for (int q = 0; q < pr.n_ranks; q++) {
            if (q == pr.rank || pr.region[q]) continue;
            cudaIpcMemHandle_t ih;
            std::memcpy(&ih, ipc_handles + 64 * (size_t)q, 64);
            void *m = nullptr;
            CUDA_CHECK(cudaIpcOpenMemHandle(&m, ih, cudaIpcMemLazyEnablePeerAccess));
            pr.region[q] = (uint8_t *)m;
        }
        pr.attached = true;
        return 0;
    });
}
int hmb200_peer_mm_create(hmb200_handle *h, int mu, unsigned char *ipc_handle_out) {
    return guarded([&] {
        if (!h || !ipc_handle_out) throw std::runtime_error("null argument");
        hmb200_handle::Peer &pr = h->peer;
        if (!pr.region[pr.rank]) throw std::runtime_error("hmb200_peer_create has not been called");
        if (mu < 1) throw std::runtime_error("mu must be positive");
        CUDA_CHECK(cudaSetDevice(h->device));
        CUDA_CHECK(cudaStreamSynchronize(h->stream));
        // A previous set of blocks (another mu) is released first.  Freeing an exported region that a peer still has open is
        // undefined, and a peer may still be storing into it: the owner frees only after EVERY rank has run
        // hmb200_peer_mm_detach (stream drained, imports closed) and the ranks have met at a barrier -- this rank's own
        // imports being closed is the part of that protocol that can be checked here.
        for (int q = 0; q < pr.n_ranks; q++)
            if (q != pr.rank && pr.mm_region[q])
                throw std::runtime_error("multi-RHS peer blocks of another width are still attached: call hmb200_peer_mm_detach on every rank, then synchronise the ranks, before hmb200_peer_mm_create");
        if (pr.mm_region[pr.rank]) {
            CUDA_CHECK(cudaFree(pr.mm_region[pr.rank]));
            pr.mm_region[pr.rank] = nullptr;
        }
        pr.mm_attached  = false;
        pr.mm_mu        = mu;
        pr.mm_buf_bytes = (((size_t)pr.n_global * mu * h->es) + 255) & ~(size_t)255;
        void *r = nullptr;
        CUDA_CHECK(cudaMalloc(&r, 2 * pr.mm_buf_bytes));
        CUDA_CHECK(cudaMemset(r, 0, 2 * pr.mm_buf_bytes));
        pr.mm_region[pr.rank] = (uint8_t *)r;
        cudaIpcMemHandle_t ih;
        CUDA_CHECK(cudaIpcGetMemHandle(&ih, r));
        std::memcpy(ipc_handle_out, &ih, 64);
        return 0;
    });
}
int hmb200_peer_mm_detach(hmb200_handle *h) {
    return guarded([&] {
        if (!h) throw std::runtime_error("null handle");
        hmb200_handle::Peer &pr = h->peer;
        CUDA_CHECK(cudaSetDevice(h->device));
        CUDA_CHECK(cudaStreamSynchronize(h->stream)); // this rank's stores into the peers' blocks have landed
        pr.mm_attached = false;
        for (int q = 0; q < pr.n_ranks; q++) {
            if (q == pr.rank || !pr.mm_region[q]) continue;
            uint8_t *m      = pr.mm_region[q];
            pr.mm_region[q] = nullptr;
            CUDA_CHECK(cudaIpcCloseMemHandle(m));
        }
        return 0;
    });
}
int hmb200_peer_mm_attach(hmb200_handle *h, const unsigned char *ipc_handles) {
    return guarded([&] {
        if (!h) throw std::runtime_error("null handle");
        hmb200_handle::Peer &pr = h->peer;
        if (!pr.mm_region[pr.rank]) throw std::runtime_error("hmb200_peer_mm_create has not been called");
        if (pr.n_ranks > 1 && !ipc_handles) throw std::runtime_error("null argument");
        CUDA_CHECK(cudaSetDevice(h->device));
        for (int q = 0; q < pr.n_ranks; q++) {
            if (q == pr.rank || pr.mm_region[q]) continue;
            cudaIpcMemHandle_t ih;
            std::memcpy(&ih, ipc_handles + 64 * (size_t)q, 64);
            void *m = nullptr;
            CUDA_CHECK(cudaIpcOpenMemHandle(&m, ih, cudaIpcMemLazyEnablePeerAccess));
            pr.mm_region[q] = (uint8_t *)m;
        }
        pr.mm_attached = true;
        return 0;
    });
}
int hmb200_peer_mm_buffer(hmb200_handle *h, int index, void **ptr) {
    if (!h || !ptr || index < 0 || index > 1 || !h->peer.mm_region[h->peer.rank]) { g_last_error = "no multi-RHS peer region / bad index"; return 1; }
    *ptr = h->peer.mm_region[h->peer.rank] + (size_t)index * h->peer.mm_buf_bytes;
    return 0;
}
int hmb200_matrix_product_allgather_d(hmb200_handle *h, const double *alpha, const double *in, int mu, void **y_replicated) {
    return guarded([&] { if (!h || h->es != 8) throw std::runtime_error("handle is not a real (double) engine"); return fused_mm_product<double>(h, alpha, in, mu, y_replicated); });
}
int hmb200_matrix_product_allgather_z(hmb200_handle *h, const double *alpha, const double *in, int mu, void **y_replicated) {
    return guarded([&] { if (!h || h->es != 16) throw std::runtime_error("handle is not a complex engine"); return fused_mm_product<double2>(h, alpha, in, mu, y_replicated); });
}
int hmb200_set_global_permutations(hmb200_handle *h, const int32_t *source_perm, int64_t n_source, const int32_t *inverse_target_perm, int64_t n_target) {
    return guarded([&] {
        if (!h || !source_perm || !inverse_target_perm) throw std::runtime_error("null argument");
        if (n_source != h->n_cols || n_target < h->n_rows) throw std::runtime_error("permutation sizes do not match the operator");
        // the buffers below are sized by the first call: a later call must bring the same sizes (a larger n_target would
        // overrun ginv_t / gy)
        if (h->ginv_t && n_target != h->gperm_n_target) throw std::runtime_error("hmb200_set_global_permutations was called before with another target size");
        h->gperm_n_target = n_target;
        CUDA_CHECK(cudaSetDevice(h->device));
        if (!h->gperm_s) CUDA_CHECK(cudaMalloc(&h->gperm_s, sizeof(int32_t) * (size_t)std::max<int64_t>(n_source, 1)));
        if (!h->ginv_t) CUDA_CHECK(cudaMalloc(&h->ginv_t, sizeof(int32_t) * (size_t)std::max<int64_t>(n_target, 1)));
        if (!h->gx) CUDA_CHECK(cudaMalloc(&h->gx, (size_t)std::max<int64_t>(n_source, 1) * h->es));
        if (!h->gy) CUDA_CHECK(cudaMalloc(&h->gy, (size_t)std::max<int64_t>(n_target, 1) * h->es));
        CUDA_CHECK(cudaMemcpyAsync(h->gperm_s, source_perm, sizeof(int32_t) * (size_t)n_source, cudaMemcpyHostToDevice, h->stream));
        CUDA_CHECK(cudaMemcpyAsync(h->ginv_t, inverse_target_perm, sizeof(int32_t) * (size_t)n_target, cudaMemcpyHostToDevice, h->stream));
        CUDA_CHECK(cudaStreamSynchronize(h->stream));
        return 0;
    });
}
int hmb200_matvec_user_chunked_d(hmb200_handle *h, const double *x_chunk, int64_t c0, int64_t count, double *y_chunk, int64_t r0, int64_t rcount) {
    return guarded([&] { if (!h || h->es != 8) throw std::runtime_error("handle is not a real (double) engine"); return user_matvec_chunked<double>(h, x_chunk, c0, count, y_chunk, r0, rcount); });
}
int hmb200_matvec_user_chunked_z(hmb200_handle *h, const double *x_chunk, int64_t c0, int64_t count, double *y_chunk, int64_t r0, int64_t rcount) {
    return guarded([&] { if (!h || h->es != 16) throw std::runtime_error("handle is not a complex engine"); return user_matvec_chunked<double2>(h, x_chunk, c0, count, y_chunk, r0, rcount); });
}
int hmb200_matmat_user_chunked_d(hmb200_handle *h, const double *X, int64_t ldx, int64_t c0, int64_t count, double *Y, int64_t ldy, int64_t r0, int64_t rcount, int mu) {
    return guarded([&] { if (!h || h->es != 8) throw std::runtime_error("handle is not a real (double) engine"); return user_matmat_chunked<double>(h, X, ldx, c0, count, Y, ldy, r0, rcount, mu); });
}
int hmb200_matmat_user_chunked_z(hmb200_handle *h, const double *X, int64_t ldx, int64_t c0, int64_t count, double *Y, int64_t ldy, int64_t r0, int64_t rcount, int mu) {
    return guarded([&] { if (!h || h->es != 16) throw std::runtime_error("handle is not a complex engine"); return user_matmat_chunked<double2>(h, X, ldx, c0, count, Y, ldy, r0, rcount, mu); });
}
int hmb200_peer_buffer(hmb200_handle *h, int index, void **ptr) {
    if (!h || !ptr || index < 0 || index > 1 || !h->peer.region[h->peer.rank]) { g_last_error = "no peer region / bad index"; return 1; }
    *ptr = h->peer.region[h->peer.rank] + (size_t)index * h->peer.buf_bytes;
    return 0;
}
int hmb200_peer_status(hmb200_handle *h, int *err) {
    return guarded([&] {
        if (!h || !err || !h->peer.region[h->peer.rank]) throw std::runtime_error("no peer region");
        *err = *(volatile int *)h->peer.err_host; // covers the waits of all work that has completed on the stream
        return 0;
    });
}
int hmb200_peer_allgather(hmb200_handle *h, const void *local_slab, void **y_replicated) {
    return guarded([&] {
        if (!h || !local_slab) throw std::runtime_error("null argument");
        return h->es == 8 ? peer_allgather<double>(h, local_slab, y_replicated) : peer_allgather<double2>(h, local_slab, y_replicated);
    });
}
int hmb200_peer_allgather_range(hmb200_handle *h, const void *local_part, int64_t elem_offset, int64_t count, void **y_replicated) {
    return guarded([&] {
        if (!h || (!local_part && count > 0) || elem_offset < 0) throw std::runtime_error("null argument / negative offset");
        return h->es == 8 ? peer_allgather<double>(h, local_part, y_replicated, elem_offset, count) : peer_allgather<double2>(h, local_part, y_replicated, elem_offset, count);
    });
}
int hmb200_peer_exchange_sum(hmb200_handle *h, char trans, const double *alpha, const void *x_local, const double *beta, void *y_local) {
    return guarded([&] {
        if (!h || !alpha || !beta || (h->n_rows > 0 && (!x_local || !y_local))) throw std::runtime_error("null argument");
        return h->es == 8 ? peer_exchange_sum<double>(h, trans, alpha, x_local, beta, y_local) : peer_exchange_sum<double2>(h, trans, alpha, x_local, beta, y_local);
    });
}
int hmb200_peer_allreduce_f64(hmb200_handle *h, double *vals, int n) {
    return guarded([&] {
        if (!h || !vals) throw std::runtime_error("null argument");
        hmb200_handle::Peer &pr = h->peer;
        if (!pr.attached) throw std::runtime_error("hmb200_peer_attach has not been called");
        if (n < 0 || n > kPeerReduceMax) throw std::runtime_error("at most 512 doubles per all-reduce");
        if (n == 0) return 0;
        CUDA_CHECK(cudaSetDevice(h->device));
        const int parity = (int)(pr.reduces & 1);
        pr.reduces++;
        PeerReduceParams p{};
        p.vals = vals; p.n = n; p.n_ranks = pr.n_ranks; p.rank = pr.rank;
        for (int q = 0; q < pr.n_ranks; q++) {
            p.slots[q] = peer_red_slots(pr, q, parity);
            p.flags[q] = peer_red_flags(pr, q, parity) + pr.rank;
        }
        p.my_flags = peer_red_flags(pr, pr.rank, parity);
        p.epoch = peer_red_epoch(pr, parity);
        p.err = peer_ctrl_err(pr);
        p.max_cycles = pr.wait_cycles;
        peer_allreduce_kernel<<<1, 256, 0, h->stream>>>(p);
        h->launches++;
        CUDA_CHECK(cudaGetLastError());
        return 0;
    });
}
int hmb200_vector_product_allgather_d(hmb200_handle *h, const double *alpha, const double *in, void **y_replicated) {
    return guarded([&] { if (!h || h->es != 8) throw std::runtime_error("handle is not a real (double) engine"); return fused_product<double>(h, alpha, in, y_replicated); });
}
int hmb200_vector_product_allgather_z(hmb200_handle *h, const double *alpha, const double *in, void **y_replicated) {
    return guarded([&] { if (!h || h->es != 16) throw std::runtime_error("handle is not a complex engine"); return fused_product<double2>(h, alpha, in, y_replicated); });
}
int hmb200_set_graphs(hmb200_handle *h, int on) {
    if (!h) { g_last_error = "null handle"; return 1; }
    h->use_graphs = on != 0;
    if (!on) clear_graphs(h);
    return 0;
}
int hmb200_set_cluster_offsets(hmb200_handle *h, int target_offset, int source_offset) {
    if (!h) { g_last_error = "null handle"; return 1; }
    h->target_offset = target_offset;
    h->source_offset = source_offset;
    return 0;
}
int hmb200_add_sub_matrix_product_to_local_d(hmb200_handle *h, const double *in, double *out, int mu, int offset, int size, int mem) {
    return guarded([&] { if (!h || h->es != 8) throw std::runtime_error("handle is not a real (double) engine"); return sub_matrix_product<double>(h, in, out, mu, offset, size, mem); });
}
int hmb200_add_sub_matrix_product_to_local_z(hmb200_handle *h, const double *in, double *out, int mu, int offset, int size, int mem) {
    return guarded([&] { if (!h || h->es != 16) throw std::runtime_error("handle is not a complex engine"); return sub_matrix_product<double2>(h, in, out, mu, offset, size, mem); });
}

static int permute_impl(bool gather, const int32_t *perm, int off, int n, const void *in, void *out, int elem_bytes, int mu, int mem, void *stream) {
    return guarded([&] {
        if (elem_bytes != 8 && elem_bytes != 16) throw std::runtime_error("elem_bytes must be 8 or 16");
        if (n < 0 || mu < 0) throw std::runtime_error("negative size");
        if (n == 0 || mu == 0) return 0;
        cudaStream_t s = (cudaStream_t)stream;
        const size_t bytes = (size_t)n * mu * elem_bytes;
        const int32_t *dperm = perm;
        const void *din = in;
        void *dout = out;
        int32_t *tp = nullptr; void *ti = nullptr, *to = nullptr;
        if (mem == HMB200_HOST) { // staging from the stream-ordered pool: after the first call no allocation reaches the driver
            CUDA_CHECK(cudaMallocAsync((void **)&tp, sizeof(int32_t) * (size_t)(off + n), s));
            CUDA_CHECK(cudaMallocAsync(&ti, bytes, s));
            CUDA_CHECK(cudaMallocAsync(&to, bytes, s));
            CUDA_CHECK(cudaMemcpyAsync(tp, perm, sizeof(int32_t) * (size_t)(off + n), cudaMemcpyHostToDevice, s));
            CUDA_CHECK(cudaMemcpyAsync(ti, in, bytes, cudaMemcpyHostToDevice, s));
            dperm = tp; din = ti; dout = to;
        }
        const int tb = 256;
        const unsigned g = (unsigned)std::min<size_t>(((size_t)n * mu + tb - 1) / tb, 148 * 16);
        if (elem_bytes == 8) {
            if (gather) permute_gather_kernel<double><<<g, tb, 0, s>>>(dperm, off, n, (const double *)din, (double *)dout, mu);
            else permute_scatter_kernel<double><<<g, tb, 0, s>>>(dperm, off, n, (const double *)din, (double *)dout, mu);
        } else {
            if (gather) permute_gather_kernel<double2><<<g, tb, 0, s>>>(dperm, off, n, (const double2 *)din, (double2 *)dout, mu);
            else permute_scatter_kernel<double2><<<g, tb, 0, s>>>(dperm, off, n, (const double2 *)din, (double2 *)dout, mu);
        }
        CUDA_CHECK(cudaGetLastError());
        if (mem == HMB200_HOST) {
            CUDA_CHECK(cudaMemcpyAsync(out, to, bytes, cudaMemcpyDeviceToHost, s));
            cudaFreeAsync(tp, s); cudaFreeAsync(ti, s); cudaFreeAsync(to, s);
            CUDA_CHECK(cudaStreamSynchronize(s));
        }
        return 0;
    });
}
int hmb200_permute_gather(const int32_t *perm, int off, int n, const void *in, void *out, int elem_bytes, int mu, int mem, void *stream) {
    return permute_impl(true, perm, off, n, in, out, elem_bytes, mu, mem, stream);
}
int hmb200_permute_scatter(const int32_t *perm, int off, int n, const void *in, void *out, int elem_bytes, int mu, int mem, void *stream) {
    return permute_impl(false, perm, off, n, in, out, elem_bytes, mu, mem, stream);
}

// column-major (user) block <-> row-major block, optionally through a cluster permutation; device pointers
static int block_layout(bool to_rows, const int32_t *perm, int n, int mu, const void *src, void *dst, long long col_stride, int elem_bytes, void *stream) {
    return guarded([&] {
        if (elem_bytes != 8 && elem_bytes != 16) throw std::runtime_error("elem_bytes must be 8 or 16");
        if (n < 0 || mu < 0 || col_stride < n) throw std::runtime_error("bad block shape");
        if (n == 0 || mu == 0) return 0;
        cudaStream_t s = (cudaStream_t)stream;
        const unsigned g = (unsigned)std::min<size_t>(((size_t)n * mu + 255) / 256 + 1, 148 * 16);
        if (elem_bytes == 8) {
            if (to_rows) mm_load_rows_kernel<double><<<g, 256, 0, s>>>((const double *)src, 0, n, mu, (double *)dst, mu, perm, col_stride);
            else mm_store_user_kernel<double><<<g, 256, 0, s>>>((const double *)src, n, mu, (double *)dst, perm, col_stride);
        } else {
            if (to_rows) mm_load_rows_kernel<double2><<<g, 256, 0, s>>>((const double2 *)src, 0, n, mu, (double2 *)dst, mu, perm, col_stride);
            else mm_store_user_kernel<double2><<<g, 256, 0, s>>>((const double2 *)src, n, mu, (double2 *)dst, perm, col_stride);
        }
        CUDA_CHECK(cudaGetLastError());
        return 0;
    });
}
int hmb200_block_to_rows(const int32_t *perm, int n, int mu, const void *src_colmajor, long long col_stride, void *dst_rowmajor, int elem_bytes, void *stream) {
    return block_layout(true, perm, n, mu, src_colmajor, dst_rowmajor, col_stride, elem_bytes, stream);
}
int hmb200_rows_to_block(const int32_t *perm, int n, int mu, const void *src_rowmajor, void *dst_colmajor, long long col_stride, int elem_bytes, void *stream) {
    return block_layout(false, perm, n, mu, src_rowmajor, dst_colmajor, col_stride, elem_bytes, stream);
}

int hmb200_set_permutations(hmb200_handle *h, const int32_t *target_perm_local, const int32_t *source_perm_local) {
    return guarded([&] {
        if (!h || !target_perm_local || !source_perm_local) throw std::runtime_error("null argument");
        CUDA_CHECK(cudaSetDevice(h->device));
        if (!h->perm_t) CUDA_CHECK(cudaMalloc(&h->perm_t, sizeof(int32_t) * (size_t)std::max(h->n_rows, 1)));
        if (!h->perm_s) CUDA_CHECK(cudaMalloc(&h->perm_s, sizeof(int32_t) * (size_t)std::max(h->n_cols, 1)));
        CUDA_CHECK(cudaMemcpyAsync(h->perm_t, target_perm_local, sizeof(int32_t) * (size_t)h->n_rows, cudaMemcpyHostToDevice, h->stream));
        CUDA_CHECK(cudaMemcpyAsync(h->perm_s, source_perm_local, sizeof(int32_t) * (size_t)h->n_cols, cudaMemcpyHostToDevice, h->stream));
        CUDA_CHECK(cudaStreamSynchronize(h->stream));
        return 0;
    });
}
int hmb200_matvec_user_d(hmb200_handle *h, const double *x, double *y, int mem) {
    return guarded([&] { if (!h || h->es != 8) throw std::runtime_error("handle is not a real (double) engine"); return matvec_user<double>(h, x, y, mem, 1); });
}
int hmb200_matvec_user_z(hmb200_handle *h, const double *x, double *y, int mem) {
    return guarded([&] { if (!h || h->es != 16) throw std::runtime_error("handle is not a complex engine"); return matvec_user<double2>(h, x, y, mem, 1); });
}
int hmb200_matmat_user_d(hmb200_handle *h, const double *X, double *Y, int mu, int mem) {
    return guarded([&] { if (!h || h->es != 8) throw std::runtime_error("handle is not a real (double) engine"); return matvec_user<double>(h, X, Y, mem, mu); });
}
int hmb200_matmat_user_z(hmb200_handle *h, const double *X, double *Y, int mu, int mem) {
    return guarded([&] { if (!h || h->es != 16) throw std::runtime_error("handle is not a complex engine"); return matvec_user<double2>(h, X, Y, mem, mu); });
}

int hmb200_last_kernel_ms(hmb200_handle *h, float ms[4]) {
    return guarded([&] { if (!h || !ms) throw std::runtime_error("null argument"); collect_timing(h, ms, true); return 0; });
}
int hmb200_kernel_ms_avg(hmb200_handle *h, float ms[4], int *n_products) {
    return guarded([&] {
        if (!h || !ms) throw std::runtime_error("null argument");
        int n = collect_timing(h, ms, false);
        if (n_products) *n_products = n;
        return 0;
    });
}
int hmb200_enable_timing(hmb200_handle *h, int on) {
    if (!h) { g_last_error = "null handle"; return 1; }
    if (on && !h->timing) h->timed_products = 0; // (re)start the averaging window
    h->timing = on != 0;
    return 0;
}
int64_t hmb200_launch_count(const hmb200_handle *h) { return h ? h->launches : 0; }

const char *hmb200_last_error(void) { return g_last_error.c_str(); }
const char *hmb200_version(void) { return "hmb200 0.1 (sm_100a)"; }

} // extern "C"

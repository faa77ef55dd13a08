// host_flags.cpp -- completion flags between the processes of one node (host code only, no CUDA).
//
// The chunked user-numbering products (hmb200_matvec_user_chunked_*, hmb200_matmat_user_chunked_*) leave 1/P of the result in a
// node-shared page-locked buffer per rank; a rank may read the whole result once every rank has published the call counter in
// its own flag word of the same segment.  The flag is the only ordering between a rank's D2H data and its readers in other
// processes, so it is stored with release and read with acquire semantics (plain loads and stores happen to work on x86 only),
// and the wait backs off instead of burning a core that the other ranks of the node need.
#include "hmb200.h"
#include "internal.h"

#include <chrono>
#include <sched.h>
#include <string>
#include <time.h>

#if defined(__x86_64__) || defined(__i386__)
#include <immintrin.h>
static inline void cpu_relax() { _mm_pause(); }
#elif defined(__aarch64__)
static inline void cpu_relax() { __asm__ __volatile__("yield"); }
#else
static inline void cpu_relax() {}
#endif

extern "C" {

int hmb200_host_flag_publish(int64_t *flag, int64_t value) {
    if (!flag) { hmb200i_set_error("null flag"); return 1; }
    __atomic_store_n(flag, value, __ATOMIC_RELEASE);
    return 0;
}

int hmb200_host_flags_wait(const int64_t *flags, int64_t stride, int n, int64_t value, double timeout_s, int *laggard) {
    if (!flags || n < 0 || stride < 1) { hmb200i_set_error("bad flag array"); return 1; }
    const auto t0 = std::chrono::steady_clock::now();
    int q = 0; // flags only grow: ranks that have been seen at `value` need not be looked at again
    for (unsigned spins = 0;; spins++) {
        while (q < n && __atomic_load_n(flags + (int64_t)q * stride, __ATOMIC_ACQUIRE) >= value) q++;
        if (q == n) return 0;
        if (spins < 2000) { cpu_relax(); continue; }      // a few microseconds: the usual skew between the ranks
        if ((spins & 63) == 0) {
            const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
            if (dt > timeout_s) {
                if (laggard) *laggard = q;
                hmb200i_set_error(("rank " + std::to_string(q) + " did not publish its part of the result within " + std::to_string(timeout_s) + " s").c_str());
                return 2;
            }
            if (dt > 2e-3) { // a late peer (e.g. an exception on its way): stop competing for the cores
                const timespec ts{0, 50000};
                nanosleep(&ts, nullptr);
                continue;
            }
        }
        sched_yield();
    }
}
}

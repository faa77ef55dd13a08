// oracle/hpddm_ref.cpp -- TEST INFRASTRUCTURE ONLY (checker, never shipped, never on the product path).
//
// Thin C shim around the REFERENCE's own Krylov solver: HPDDM::IterativeMethod::solve (the call DDM::solve makes,
// lib/htool/include/htool/solvers/ddm.hpp:193) compiled from the vendored HPDDM headers where they lie
// (lib/hpddm/include, GMRES = HPDDM_GMRES.hpp:30-196) against oracle/mpi_stub/mpi.h, our one-rank stand-in for <mpi.h>
// (no MPI in this image; HPDDM's own HPDDM_MPI=0 mode does not compile in this version).  No reference code is copied: this file only defines an operator that forwards GMV to a C callback, exactly
// like HPDDM's own C interface does (lib/hpddm/interface/hpddm_c.cpp:39-54).
//
// It gives the real HPDDM iteration counts / iterates for one process (no preconditioner = htool's
// "-hpddm_schwarz_method none" with one subdomain), against which oracle/gmres_oracle.py and the device solver are pinned.
#define HPDDM_MPI 1
#define HPDDM_NUMBERING 'F'
#include <HPDDM.hpp>

#include <complex>
#include <cstring>

namespace {
typedef int (*matvec_cb)(void *ctx, const void *in, void *out, int mu);

template <class K>
struct CallbackOperator : public HPDDM::EmptyOperator<K> {
    matvec_cb mv_;
    void *ctx_;
    CallbackOperator(int n, matvec_cb mv, void *ctx) : HPDDM::EmptyOperator<K>(n), mv_(mv), ctx_(ctx) {}
    int GMV(const K *const in, K *const out, const int &mu = 1) const { return mv_(ctx_, in, out, mu); }
    // no preconditioner: identity (what Prcndtnr::NO does for htool's HPDDMOperator, ddm.hpp:136-138)
    template <bool>
    int apply(const K *const in, K *const out, const unsigned short &mu = 1, K * = nullptr, const unsigned short & = 0) const {
        std::memcpy(out, in, sizeof(K) * (size_t)mu * this->getDof());
        return 0;
    }
};

template <class K>
int solve(int n, matvec_cb mv, void *ctx, const K *b, K *x, int mu, const char *args) {
    HPDDM::Option &opt = *HPDDM::Option::get();
    std::string a(args ? args : "");
    opt.parse(a, false);
    CallbackOperator<K> A(n, mv, ctx);
    return HPDDM::IterativeMethod::solve(A, b, x, mu, 0);
}
} // namespace

extern "C" {
__attribute__((visibility("default"))) int hpddm_ref_solve_d(int n, matvec_cb mv, void *ctx, const double *b, double *x, int mu, const char *args) {
    return solve<double>(n, mv, ctx, b, x, mu, args);
}
__attribute__((visibility("default"))) int hpddm_ref_solve_z(int n, matvec_cb mv, void *ctx, const double *b, double *x, int mu, const char *args) {
    return solve<std::complex<double>>(n, mv, ctx, reinterpret_cast<const std::complex<double> *>(b), reinterpret_cast<std::complex<double> *>(x), mu, args);
}
}

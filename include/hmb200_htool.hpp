// hmb200_htool.hpp -- the reference-side binding of libhmb200.so: what a maintainer of htool / htool_python adds to run the
// B200 engine behind the reference's own operator interface.  Header-only; needs the reference's headers on the include path.
//
//   hmb200::FlatTable<T>                 leaf table of an assembled htool::HMatrix (what get_leaves_from enumerates,
//                                        lib/htool/include/htool/hmatrix/hmatrix.hpp:247-274) with pointers straight into the
//                                        reference's U / V / dense buffers -> hmb200_flat (include/hmb200.h)
//   hmb200::RestrictedGlobalToLocalB200<T>   drop-in for RestrictedGlobalToLocalHMatrix
//                                        (distributed_operator/implementations/global_to_local_operators/hmatrix.hpp:15-35): overrides the
//                                        three pure / virtual hooks of RestrictedGlobalToLocalOperator
//                                        (.../restricted_operator.hpp:38,47,49) with calls into the C ABI, host pointers, synchronous.
//
// Register it like the stock operator: DistributedOperator::add_global_to_local_operator(&op)
// (distributed_operator/distributed_operator.hpp:51-53).  Compiled and exercised by tests/native/reference_binding.cpp under the
// reference's own drivers (global-to-global vector / matrix products, the sub-matrix product, DDM::solve -> HPDDMOperator::GMV).
#pragma once
#include "hmb200.h"

#include <htool/distributed_operator/implementations/global_to_local_operators/restricted_operator.hpp>
#include <htool/hmatrix/hmatrix.hpp>
#include <htool/misc/logger.hpp>

#include <complex>
#include <stdexcept>
#include <unordered_set>
#include <vector>

namespace hmb200 {

template <typename T>
struct FlatTable {
    std::vector<int32_t> row_off, col_off, nrows, ncols, rank;
    std::vector<uint8_t> flags;
    std::vector<const void *> a, b;
    hmb200_flat flat{};

    explicit FlatTable(const htool::HMatrix<T, double> &H) {
        auto leaves = htool::get_leaves_from(H); // (leaves, leaves_for_symmetry)
        std::unordered_set<const htool::HMatrix<T, double> *> mirrored(leaves.second.begin(), leaves.second.end());
        const char sym = H.get_symmetry_for_leaves();
        const int t0 = H.get_target_cluster().get_offset(), s0 = H.get_source_cluster().get_offset();
        for (const auto *leaf : leaves.first) {
            const auto &t = leaf->get_target_cluster();
            const auto &s = leaf->get_source_cluster();
            row_off.push_back(t.get_offset() - t0);
            col_off.push_back(s.get_offset() - s0);
            nrows.push_back(t.get_size());
            ncols.push_back(s.get_size());
            uint8_t f = (sym != 'N' && mirrored.count(leaf)) ? HMB200_LEAF_MIRROR : 0;
            if (leaf->is_dense()) {
                rank.push_back(-1);
                a.push_back(leaf->get_dense_data()->data());
                b.push_back(nullptr);
                const char ls = leaf->get_symmetry(), lu = leaf->get_UPLO();
                if (ls == 'S') f |= (uint8_t)((lu == 'L' ? 1 : 2) << 1);
                if (ls == 'H') f |= (uint8_t)((lu == 'L' ? 3 : 4) << 1);
            } else if (leaf->is_low_rank()) {
                const auto *lr = leaf->get_low_rank_data();
                rank.push_back(lr->rank_of());
                a.push_back(lr->get_U().data());
                b.push_back(lr->get_V().data());
            } else {
                throw std::runtime_error("hmb200: leaf is neither dense nor low-rank");
            }
            flags.push_back(f);
        }
        flat.n_rows = H.get_target_cluster().get_size();
        flat.n_cols = H.get_source_cluster().get_size();
        flat.n_leaves = (int64_t)rank.size();
        flat.row_off = row_off.data(); flat.col_off = col_off.data(); flat.nrows = nrows.data(); flat.ncols = ncols.data();
        flat.rank = rank.data(); flat.flags = flags.data(); flat.a = a.data(); flat.b = b.data();
        flat.symmetry_for_leaves = sym;
        flat.n_row_cuts = 0; flat.row_cuts = nullptr;
    }
};

template <typename T> struct Abi;
template <> struct Abi<double> {
    static hmb200_handle *create(const hmb200_flat *f, int dev, const hmb200_options *o) { return hmb200_create_d(f, dev, nullptr, o); }
    static int vec(hmb200_handle *h, char tr, const double *al, const double *in, const double *be, double *out) { return hmb200_add_vector_product_d(h, tr, al, in, be, out, HMB200_HOST); }
    static int mat(hmb200_handle *h, char tr, const double *al, const double *in, const double *be, double *out, int mu) { return hmb200_add_matrix_product_row_major_d(h, tr, al, in, be, out, mu, HMB200_HOST); }
    static int sub(hmb200_handle *h, const double *in, double *out, int mu, int off, int size) { return hmb200_add_sub_matrix_product_to_local_d(h, in, out, mu, off, size, HMB200_HOST); }
};
template <> struct Abi<std::complex<double>> {
    static hmb200_handle *create(const hmb200_flat *f, int dev, const hmb200_options *o) { return hmb200_create_z(f, dev, nullptr, o); }
    static int vec(hmb200_handle *h, char tr, const double *al, const double *in, const double *be, double *out) { return hmb200_add_vector_product_z(h, tr, al, in, be, out, HMB200_HOST); }
    static int mat(hmb200_handle *h, char tr, const double *al, const double *in, const double *be, double *out, int mu) { return hmb200_add_matrix_product_row_major_z(h, tr, al, in, be, out, mu, HMB200_HOST); }
    static int sub(hmb200_handle *h, const double *in, double *out, int mu, int off, int size) { return hmb200_add_sub_matrix_product_to_local_z(h, in, out, mu, off, size, HMB200_HOST); }
};

template <typename T>
class RestrictedGlobalToLocalB200 final : public htool::RestrictedGlobalToLocalOperator<T> {
    hmb200_handle *m_h = nullptr;
    static const double *dp(const T *p) { return reinterpret_cast<const double *>(p); }
    static double *dp(T *p) { return reinterpret_cast<double *>(p); }
    static void report(int status) { // the reference logs and carries on (add_hmatrix_vector_product.hpp:58-61); so does the binding
        if (status) htool::Logger::get_instance().log(htool::LogLevel::ERROR, hmb200_last_error());
    }

  public:
    // `hmatrix`: the rank's block (cluster_on_partition(rank) x source root), already assembled by the reference.  Its
    // coefficients are copied to the GPU here, once; the htool::HMatrix itself is not kept.
    RestrictedGlobalToLocalB200(const htool::HMatrix<T, double> &hmatrix, const htool::LocalRenumbering &target, const htool::LocalRenumbering &source, int device, bool with_transpose = true)
        : htool::RestrictedGlobalToLocalOperator<T>(target, source, false, false) {
        FlatTable<T> table(hmatrix);
        hmb200_options opt{};
        opt.build_transpose = with_transpose ? 1 : 0;
        m_h = Abi<T>::create(&table.flat, device, &opt);
        if (!m_h) throw std::runtime_error(hmb200_last_error());
        if (hmb200_set_cluster_offsets(m_h, hmatrix.get_target_cluster().get_offset(), hmatrix.get_source_cluster().get_offset())) throw std::runtime_error(hmb200_last_error());
    }
    RestrictedGlobalToLocalB200(const RestrictedGlobalToLocalB200 &)            = delete;
    RestrictedGlobalToLocalB200 &operator=(const RestrictedGlobalToLocalB200 &) = delete;
    ~RestrictedGlobalToLocalB200() { hmb200_destroy(m_h); }

    // cluster numbering, out = beta*out + alpha*op(A)*in, in place, host pointers, synchronous: the contract of
    // RestrictedGlobalToLocalHMatrix::local_add_vector_product (global_to_local_operators/hmatrix.hpp:27-29)
    void local_add_vector_product(char trans, T alpha, const T *in, T beta, T *out) const override {
        report(Abi<T>::vec(m_h, trans, dp(&alpha), dp(in), dp(&beta), dp(out)));
    }
    void local_add_matrix_product_row_major(char trans, T alpha, const T *in, T beta, T *out, int mu) const override {
        report(Abi<T>::mat(m_h, trans, dp(&alpha), dp(in), dp(&beta), dp(out), mu));
    }
    // out += A[:, offset:offset+size] in : the base class has already clipped the window to the source range
    // (restricted_operator.hpp:173-194); the zero extension happens on the device instead of in a host temporary (:49-56)
    void local_add_sub_matrix_product_to_local(const T *const in, T *const out, int mu, int offset, int size) const override {
        report(Abi<T>::sub(m_h, dp(in), dp(out), mu, offset, size));
    }
    hmb200_handle *handle() const { return m_h; }
};

} // namespace hmb200

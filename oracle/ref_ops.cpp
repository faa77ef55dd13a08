// oracle/ref_ops.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// Thin pybind11 shim that calls the REFERENCE'S OWN CPU products (compiled from the headers where they lie under
// /root/reference/lib/htool/include by oracle/Makefile; no arithmetic of ours in here) so that
//   * tests/ can compare the CUDA engine and the flat C restatement against the real reference on the identical
//     htool::HMatrix object, in cluster numbering (not bound by the reference's Python module) and with OpenMP;
//   * bench.py's cpu_baseline / --impl reference legs can time the reference's OpenMP product on the host cores.
//
// Entry points and the reference functions they forward to (lib/htool/include/htool/...):
//   internal_vector_product            hmatrix/linalg/add_hmatrix_vector_product.hpp:56-104 (seq) / :106-170 (openmp)
//   vector_product_user                hmatrix/linalg/add_hmatrix_vector_product.hpp:172-197
//   internal_matrix_product_row_major  hmatrix/linalg/add_hmatrix_matrix_product_row_major.hpp:57-109 / :111-178
//   matrix_product_user                hmatrix/linalg/add_hmatrix_matrix_product.hpp:170-205
//
// The H-matrix is passed either as the reference's Python object (Htool.HMatrix / Htool.ComplexHMatrix) or as the raw
// address of an htool::HMatrix owned by htool_python_b200._htool_bridge.NativeHMatrix (`.address`, `.dtype`).
#include <pybind11/complex.h>
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>

#include <htool/hmatrix/hmatrix.hpp>
#include <htool/hmatrix/linalg/add_hmatrix_matrix_product.hpp>
#include <htool/hmatrix/linalg/add_hmatrix_matrix_product_row_major.hpp>
#include <htool/hmatrix/linalg/add_hmatrix_vector_product.hpp>
#include <htool/matrix/matrix_view.hpp>

#include <complex>
#include <cstdint>
#if defined(_OPENMP)
#    include <omp.h>
#endif

namespace py = pybind11;
using namespace htool;
using zdouble = std::complex<double>;

template <typename T>
using arr = py::array_t<T, py::array::c_style>;

template <typename T>
static void internal_vec(const HMatrix<T, double> &H, char trans, T alpha, arr<T> in, T beta, arr<T> out, bool parallel) {
    const T *x = in.data();
    T *y       = out.mutable_data();
    py::gil_scoped_release release;
    if (parallel)
        openmp_internal_add_hmatrix_vector_product(trans, alpha, H, x, beta, y);
    else
        sequential_internal_add_hmatrix_vector_product(trans, alpha, H, x, beta, y);
}

template <typename T>
static void internal_mat(const HMatrix<T, double> &H, char trans, T alpha, arr<T> in, T beta, arr<T> out, int mu, bool parallel) {
    const T *x = in.data();
    T *y       = out.mutable_data();
    py::gil_scoped_release release;
    if (parallel)
        openmp_internal_add_hmatrix_matrix_product_row_major(trans, 'N', alpha, H, x, beta, y, mu);
    else
        sequential_internal_add_hmatrix_matrix_product_row_major(trans, 'N', alpha, H, x, beta, y, mu);
}

template <typename T>
static arr<T> user_vec(const HMatrix<T, double> &H, arr<T> in, bool parallel) {
    arr<T> out(H.nb_rows());
    std::fill_n(out.mutable_data(), H.nb_rows(), T(0));
    const T *x = in.data();
    T *y       = out.mutable_data();
    {
        py::gil_scoped_release release;
        if (parallel)
            add_hmatrix_vector_product(exec_compat::par, 'N', T(1), H, x, T(0), y);
        else
            add_hmatrix_vector_product(exec_compat::seq, 'N', T(1), H, x, T(0), y);
    }
    return out;
}

template <typename T>
static py::array_t<T, py::array::f_style> user_mat(const HMatrix<T, double> &H, py::array_t<T, py::array::f_style> in, bool parallel) {
    py::array_t<T, py::array::f_style> out({(py::ssize_t)H.nb_rows(), in.shape(1)});
    std::fill_n(out.mutable_data(), out.size(), T(0));
    MatrixView<const T> B(in.shape(0), in.shape(1), in.data());
    MatrixView<T> C(H.nb_rows(), in.shape(1), out.mutable_data());
    {
        py::gil_scoped_release release;
        if (parallel)
            add_hmatrix_matrix_product(exec_compat::par, 'N', 'N', T(1), H, B, T(0), C);
        else
            add_hmatrix_matrix_product(exec_compat::seq, 'N', 'N', T(1), H, B, T(0), C);
    }
    return out;
}

template <typename T>
static const HMatrix<T, double> &from_address(uintptr_t a) { return *reinterpret_cast<const HMatrix<T, double> *>(a); }

template <typename T>
static void declare(py::module &m, const char *sfx) {
    std::string s(sfx);
    // by reference-module object
    m.def("internal_vector_product", [](const HMatrix<T, double> &H, char trans, T alpha, arr<T> in, T beta, arr<T> out, bool parallel) { internal_vec<T>(H, trans, alpha, in, beta, out, parallel); });
    m.def("internal_matrix_product_row_major", [](const HMatrix<T, double> &H, char trans, T alpha, arr<T> in, T beta, arr<T> out, int mu, bool parallel) { internal_mat<T>(H, trans, alpha, in, beta, out, mu, parallel); });
    m.def("vector_product_user", [](const HMatrix<T, double> &H, arr<T> in, bool parallel) { return user_vec<T>(H, in, parallel); });
    m.def("matrix_product_user", [](const HMatrix<T, double> &H, py::array_t<T, py::array::f_style> in, bool parallel) { return user_mat<T>(H, in, parallel); });
    // by raw address (NativeHMatrix.address)
    m.def(("internal_vector_product_addr_" + s).c_str(), [](uintptr_t a, char trans, T alpha, arr<T> in, T beta, arr<T> out, bool parallel) { internal_vec<T>(from_address<T>(a), trans, alpha, in, beta, out, parallel); });
    m.def(("internal_matrix_product_row_major_addr_" + s).c_str(), [](uintptr_t a, char trans, T alpha, arr<T> in, T beta, arr<T> out, int mu, bool parallel) { internal_mat<T>(from_address<T>(a), trans, alpha, in, beta, out, mu, parallel); });
    m.def(("vector_product_user_addr_" + s).c_str(), [](uintptr_t a, arr<T> in, bool parallel) { return user_vec<T>(from_address<T>(a), in, parallel); });
    m.def(("matrix_product_user_addr_" + s).c_str(), [](uintptr_t a, py::array_t<T, py::array::f_style> in, bool parallel) { return user_mat<T>(from_address<T>(a), in, parallel); });
}

PYBIND11_MODULE(htool_ref_ops, m) {
    m.doc() = "reference CPU products (oracle; test infrastructure only)";
    declare<double>(m, "d");
    declare<zdouble>(m, "z");
    m.def("omp_max_threads", []() {
#if defined(_OPENMP)
        return omp_get_max_threads();
#else
        return 1;
#endif
    });
    m.def("omp_set_num_threads", [](int n) {
#if defined(_OPENMP)
        omp_set_num_threads(n);
#else
        (void)n;
#endif
    });
}

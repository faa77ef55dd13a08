// tests/native/reference_binding.cpp -- TEST INFRASTRUCTURE: the reference-side binding (include/hmb200_htool.hpp) compiled for
// real against the reference's headers, and driven by the REFERENCE'S OWN code paths on one rank (oracle/mpi_stub/mpi.h stands
// in for <mpi.h>: there is no MPI in the image):
//   * htool::DistributedOperator with the GPU operator registered through add_global_to_local_operator
//     (distributed_operator/distributed_operator.hpp:51-53), next to a second DistributedOperator holding the stock
//     RestrictedGlobalToLocalHMatrix on the SAME htool::HMatrix;
//   * add_distributed_operator_vector_product_global_to_global / ..._matrix_product_row_major_global_to_global (user numbering,
//     N / T / C), VirtualGlobalToLocalOperator::add_sub_matrix_product_to_local with clipped windows;
//   * DDM::solve (solvers/ddm.hpp:127-229) -> HPDDM::IterativeMethod::solve -> HPDDMOperator::GMV (wrappers/wrapper_hpddm.hpp:102-145)
//     -> ..._local_to_local -> the operator, with mu = 1 and mu > 1.
// Returns the relative differences GPU operator vs stock operator; tests/test_gpu_reference_binding.py asserts on them.
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <htool/distributed_operator/distributed_operator.hpp>
#include <htool/distributed_operator/implementations/global_to_local_operators/hmatrix.hpp>
#include <htool/distributed_operator/implementations/partition_from_cluster.hpp>
#include <htool/distributed_operator/linalg/add_distributed_operator_matrix_product_row_major_global_to_global.hpp>
#include <htool/distributed_operator/linalg/add_distributed_operator_vector_product_global_to_global.hpp>
#include <htool/solvers/ddm.hpp>
#include <htool/solvers/utility.hpp>

#include <hmb200_htool.hpp>

#include <cmath>
#include <complex>
#include <map>
#include <random>
#include <string>

namespace py = pybind11;
using namespace htool;

namespace {

template <typename T> T rnd(std::mt19937_64 &g);
template <> double rnd<double>(std::mt19937_64 &g) { return std::uniform_real_distribution<double>(-1, 1)(g); }
template <> std::complex<double> rnd<std::complex<double>>(std::mt19937_64 &g) { std::uniform_real_distribution<double> d(-1, 1); double a = d(g), b = d(g); return {a, b}; }

template <typename T>
double rel_diff(const std::vector<T> &a, const std::vector<T> &b) {
    double num = 0, den = 0;
    for (size_t i = 0; i < a.size(); i++) { num += std::norm(a[i] - b[i]); den += std::norm(b[i]); }
    return den > 0 ? std::sqrt(num / den) : std::sqrt(num);
}

template <typename T>
std::map<std::string, double> check(uintptr_t address, int device, int mu, const std::string &hpddm_args, bool with_transpose, bool with_solver) {
    HMatrix<T, double> &H = *reinterpret_cast<HMatrix<T, double> *>(address);
    const Cluster<double> &tc = H.get_target_cluster(), &sc = H.get_source_cluster();
    if (!tc.is_root() || !sc.is_root()) throw std::runtime_error("one-rank check: the H-matrix must be the root block");
    const int nt = tc.get_size(), ns = sc.get_size();
    std::map<std::string, double> res;
    std::mt19937_64 gen(1234);

    PartitionFromCluster<T, double> tp(tc), sp(sc);
    DistributedOperator<T> d_ref(tp, sp, MPI_COMM_WORLD), d_gpu(tp, sp, MPI_COMM_WORLD);
    RestrictedGlobalToLocalHMatrix<T, double> op_ref(H, LocalRenumbering(tc), LocalRenumbering(sc), false, false);
    hmb200::RestrictedGlobalToLocalB200<T> op_gpu(H, LocalRenumbering(tc), LocalRenumbering(sc), device, with_transpose);
    d_ref.add_global_to_local_operator(&op_ref);
    d_gpu.add_global_to_local_operator(&op_gpu); // the reference's own registration call

    const T alpha = T(1.25), beta = T(-0.75);
    // ---- global-to-global vector product, user numbering (the product behind DistributedOperator.__mul__) ----
    // 'T' on a Hermitian-stored / 'C' on a symmetric-stored block are refused by the reference itself (add_hmatrix_vector_product.hpp:112-115)
    const char sfl = H.get_symmetry_for_leaves();
    const std::string ops = !with_transpose ? "N" : (std::is_same<T, double>::value ? "NT" : (sfl == 'S' ? "NT" : (sfl == 'H' ? "NC" : "NTC")));
    for (char trans : ops) {
        const int ni = trans == 'N' ? ns : nt, no = trans == 'N' ? nt : ns;
        std::vector<T> x(ni), y0(no), y1, y2;
        for (auto &v : x) v = rnd<T>(gen);
        for (auto &v : y0) v = rnd<T>(gen);
        y1 = y0; y2 = y0;
        add_distributed_operator_vector_product_global_to_global(trans, alpha, d_ref, x.data(), beta, y1.data(), (T *)nullptr); // work == nullptr: the drivers allocate their own buffers
        add_distributed_operator_vector_product_global_to_global(trans, alpha, d_gpu, x.data(), beta, y2.data(), (T *)nullptr);
        res[std::string("vector_global_to_global_") + trans] = rel_diff(y2, y1);
    }
    // ---- global-to-global matrix product, row-major blocks stored as (mu x n) Matrix (..._row_major_global_to_global.hpp:18) ----
    {
        Matrix<T> X(mu, ns), Y1(mu, nt), Y2(mu, nt);
        for (int i = 0; i < mu * ns; i++) X.data()[i] = rnd<T>(gen);
        for (int i = 0; i < mu * nt; i++) Y1.data()[i] = Y2.data()[i] = rnd<T>(gen);
        internal_add_distributed_operator_matrix_product_row_major_global_to_global('N', alpha, d_ref, X, beta, Y1, (T *)nullptr);
        internal_add_distributed_operator_matrix_product_row_major_global_to_global('N', alpha, d_gpu, X, beta, Y2, (T *)nullptr);
        std::vector<T> a(Y2.data(), Y2.data() + (size_t)mu * nt), b(Y1.data(), Y1.data() + (size_t)mu * nt);
        res["matrix_row_major_global_to_global_N"] = rel_diff(a, b);
    }
    // ---- add_sub_matrix_product_to_local through the virtual interface: whole range, interior, clipped left / right, outside ----
    {
        const int windows[5][2] = {{0, ns}, {ns / 3, ns / 4}, {-17, 200}, {ns - 100, 300}, {ns + 5, 40}};
        double worst = 0;
        for (int m : {1, mu})
            for (auto &w : windows) {
                const int off = w[0], size = w[1];
                std::vector<T> in((size_t)size * m), y1((size_t)nt * m), y2;
                for (auto &v : in) v = rnd<T>(gen);
                for (auto &v : y1) v = rnd<T>(gen);
                y2 = y1;
                if (m > 1 && off < 0) continue; // the reference's own clipping of a multi-RHS window left of the range mis-steps the pointer (restricted_operator.hpp:184: no * mu)
                static_cast<const VirtualGlobalToLocalOperator<T> &>(op_ref).add_sub_matrix_product_to_local(in.data(), y1.data(), m, off, size);
                static_cast<const VirtualGlobalToLocalOperator<T> &>(op_gpu).add_sub_matrix_product_to_local(in.data(), y2.data(), m, off, size);
                worst = std::max(worst, rel_diff(y2, y1));
            }
        res["sub_matrix_product_to_local"] = worst;
    }
    // ---- the reference's solver: DDM::solve -> HPDDM GMRES -> HPDDMOperator::GMV -> local-to-local driver -> operator ----
    if (with_solver) {
        HPDDM::Option &opt = *HPDDM::Option::get();
        std::string args = hpddm_args;
        opt.parse(args, false);
        for (int m : {1, mu}) {
            std::vector<T> xs((size_t)ns * m), rhs((size_t)nt * m), x1((size_t)ns * m, T(0)), x2((size_t)ns * m, T(0));
            for (auto &v : xs) v = rnd<T>(gen);
            for (int c = 0; c < m; c++) {
                std::fill_n(rhs.data() + (size_t)c * nt, nt, T(0));
                add_distributed_operator_vector_product_global_to_global('N', T(1), d_ref, xs.data() + (size_t)c * ns, T(0), rhs.data() + (size_t)c * nt, (T *)nullptr);
            }
            DDMSolverBuilder<T, double> b_ref(d_ref, H), b_gpu(d_gpu, H); // block-Jacobi builder, never factorised (-hpddm_schwarz_method none)
            b_ref.solver.solve(rhs.data(), x1.data(), m);
            b_gpu.solver.solve(rhs.data(), x2.data(), m);
            const std::string sfx = m == 1 ? "_mu1" : "_mu";
            res["ddm_iterations_reference" + sfx] = std::stod(b_ref.solver.get_information("Nb_it"));
            res["ddm_iterations_gpu" + sfx]       = std::stod(b_gpu.solver.get_information("Nb_it"));
            res["ddm_solution_rel_diff" + sfx]    = rel_diff(x2, x1);
            res["ddm_error_vs_exact_gpu" + sfx]   = rel_diff(x2, xs);
        }
    }
    return res;
}

} // namespace

PYBIND11_MODULE(htool_b200_binding, m) {
    m.doc() = "TEST INFRASTRUCTURE: hmb200::RestrictedGlobalToLocalB200 under the reference's own DistributedOperator / DDM drivers";
    m.def("check_d", &check<double>, py::arg("address"), py::arg("device"), py::arg("mu"), py::arg("hpddm_args"), py::arg("with_transpose") = true, py::arg("with_solver") = true);
    m.def("check_z", &check<std::complex<double>>, py::arg("address"), py::arg("device"), py::arg("mu"), py::arg("hpddm_args"), py::arg("with_transpose") = true, py::arg("with_solver") = true);
}

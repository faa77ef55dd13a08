// htool_python_b200/csrc/bridge/htool_bridge.cpp
//
// The flattening layer: a thin pybind11 shim compiled against the reference's header-only core
// (/root/reference/lib/htool/include, included by path, nothing copied).  It does two things:
//
//  1. flat_view(H): walks an ALREADY ASSEMBLED htool::HMatrix -- either the reference's own Python object
//     (Htool.HMatrix / Htool.ComplexHMatrix, registered globally by the reference module,
//     src/htool/hmatrix/hmatrix.hpp:30) or a NativeHMatrix built below -- and returns the leaf table
//     (offsets, sizes, ranks, symmetry flags and the raw addresses of the U/V/dense buffers) that
//     hmb200_create_{d,z} (include/hmb200.h) packs and uploads.  The Python HMatrix exposes no leaf data
//     (src/htool/hmatrix/hmatrix.hpp:31-138), so this must be C++.
//     Leaf enumeration = the reference's get_leaves_from (hmatrix/hmatrix.hpp:247-274); addressing as in
//     hmatrix/linalg/add_hmatrix_vector_product.hpp:90-103.
//
//  2. build_hmatrix(...): assembles an H-matrix for a synthetic kernel with a native C++ generator by calling the
//     reference's OWN ClusterTreeBuilder / HMatrixTreeBuilder::build(exec_compat::par, ...)
//     (clustering/tree_builder/tree_builder.hpp:52, hmatrix/tree_builder/tree_builder.hpp:204-215).  Going through
//     the Python VirtualGenerator trampoline costs one GIL-bound callback per ACA row/column and is impractical
//     at N=1M; the compressed matrix is still assembled by reference code.
//
// No product arithmetic lives here: the CPU products of the reference are only reachable through oracle/ref_ops.cpp.
#include <pybind11/complex.h>
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <htool/clustering/cluster_node.hpp>
#include <htool/clustering/tree_builder/tree_builder.hpp>
#include <htool/hmatrix/hmatrix.hpp>
#include <htool/hmatrix/hmatrix_output.hpp>
#include <htool/hmatrix/interfaces/virtual_generator.hpp>
#include <htool/hmatrix/tree_builder/tree_builder.hpp>

#include <cmath>
#include <complex>
#include <cstdint>
#include <memory>
#include <stack>
#include <string>
#include <unordered_set>
#include <vector>

namespace py = pybind11;
using namespace htool;
using zdouble = std::complex<double>;

// ------------------------------------------------------------------------------------------------------------------
// native generators (user numbering, column-major M x N block, VirtualGenerator::copy_submatrix contract:
// hmatrix/interfaces/virtual_generator.hpp:21-33)
// ------------------------------------------------------------------------------------------------------------------
enum KernelId : int {
    K_LAPLACE      = 0, // 1/(4 pi r), 0 where r == 0
    K_LAPLACE_REG  = 1, // 1/(1e-5 + 4 pi r)          (testing/generator_test.hpp:180-187)
    K_INVDIST_REG  = 2, // 1/(0.1 + r)                (example/define_generators.py:15-17)
    K_HELMHOLTZ    = 3, // exp(i k r)/(4 pi r), 0 where r == 0
    K_COMPLEX_REG  = 4, // (1+i)/(1e-5 + 4 pi r)      (testing/generator_test.hpp:189-196)
    K_COMPLEX_HERM = 5, // (1 + sign(xt-xs) i)/(1e-5 + 4 pi r)   (testing/generator_test.hpp:198-205)
};

template <typename T>
struct KernelEval;
template <>
struct KernelEval<double> {
    static inline double eval(int id, double r, double, double) {
        switch (id) {
        case K_LAPLACE:
            return r > 0 ? 1. / (4 * M_PI * r) : 0.;
        case K_LAPLACE_REG:
            return 1. / (1e-5 + 4 * M_PI * r);
        default:
            return 1. / (0.1 + r);
        }
    }
};
template <>
struct KernelEval<zdouble> {
    static inline zdouble eval(int id, double r, double k, double dx0) {
        switch (id) {
        case K_HELMHOLTZ:
            return r > 0 ? zdouble(std::cos(k * r), std::sin(k * r)) / (4 * M_PI * r) : zdouble(0);
        case K_COMPLEX_HERM: {
            double s = dx0 > 0 ? 1. : (dx0 < 0 ? -1. : 0.);
            return zdouble(1, s) / (1e-5 + 4 * M_PI * r);
        }
        default:
            return zdouble(1, 1) / (1e-5 + 4 * M_PI * r);
        }
    }
};

template <typename T>
class NativeKernelGenerator : public VirtualGenerator<T> {
    int m_id, m_dim;
    double m_k;
    std::shared_ptr<std::vector<double>> m_tp, m_sp; // (dim, N) Fortran order == point-major xyzxyz...
  public:
    NativeKernelGenerator(int id, int dim, double k, std::shared_ptr<std::vector<double>> tp, std::shared_ptr<std::vector<double>> sp) : m_id(id), m_dim(dim), m_k(k), m_tp(tp), m_sp(sp) {}
    void copy_submatrix(int M, int N, const int *rows, const int *cols, T *ptr) const override {
        const double *tp = m_tp->data(), *sp = m_sp->data();
        for (int j = 0; j < N; j++) {
            const double *y = sp + (size_t)m_dim * cols[j];
            for (int i = 0; i < M; i++) {
                const double *x = tp + (size_t)m_dim * rows[i];
                double r2       = 0;
                for (int d = 0; d < m_dim; d++)
                    r2 += (x[d] - y[d]) * (x[d] - y[d]);
                ptr[i + (size_t)M * j] = KernelEval<T>::eval(m_id, std::sqrt(r2), m_k, x[0] - y[0]);
            }
        }
    }
};

// ------------------------------------------------------------------------------------------------------------------
// native cluster tree + H-matrix holders
// ------------------------------------------------------------------------------------------------------------------
struct NativeClusterTree {
    std::shared_ptr<std::vector<double>> points; // point-major copy of the coordinates
    int dim = 0, n = 0;
    std::shared_ptr<Cluster<double>> root;
};

template <typename T>
struct NativeHMatrix {
    std::shared_ptr<NativeClusterTree> target, source;
    std::shared_ptr<NativeKernelGenerator<T>> generator;
    std::unique_ptr<HMatrix<T, double>> hmatrix;
    const HMatrix<T, double> &get() const { return *hmatrix; }
};

static std::shared_ptr<NativeClusterTree> make_cluster_tree(py::array_t<double, py::array::f_style | py::array::forcecast> coordinates, int number_of_children, int size_of_partition, int maximal_leaf_size) {
    if (coordinates.ndim() != 2)
        throw std::runtime_error("coordinates must be a (dim, N) Fortran-ordered array");
    auto t    = std::make_shared<NativeClusterTree>();
    t->dim    = (int)coordinates.shape(0);
    t->n      = (int)coordinates.shape(1);
    t->points = std::make_shared<std::vector<double>>(coordinates.data(), coordinates.data() + (size_t)t->dim * t->n);
    ClusterTreeBuilder<double> builder;
    builder.set_maximal_leaf_size(maximal_leaf_size);
    {
        py::gil_scoped_release release;
        t->root = std::make_shared<Cluster<double>>(builder.create_cluster_tree(t->n, t->dim, t->points->data(), number_of_children, size_of_partition));
    }
    return t;
}

template <typename T>
static std::shared_ptr<NativeHMatrix<T>> build_hmatrix(int kernel_id, double wavenumber, std::shared_ptr<NativeClusterTree> target, std::shared_ptr<NativeClusterTree> source, double epsilon, double eta, char symmetry, char UPLO, int target_partition_number, int partition_number_for_symmetry, bool parallel, int minimal_target_depth, int minimal_source_depth) {
    auto h       = std::make_shared<NativeHMatrix<T>>();
    h->target    = target;
    h->source    = source;
    h->generator = std::make_shared<NativeKernelGenerator<T>>(kernel_id, target->dim, wavenumber, target->points, source->points);
    HMatrixTreeBuilder<T, double> tree_builder(epsilon, eta, symmetry, UPLO);
    // the reference's own setters (src/htool/hmatrix/hmatrix_tree_builder.hpp binds them too): admissible blocks are only
    // accepted below these depths.  Needed at N >= ~90k per side: sympartialACA's `q*(n1+n2) > n1*n2` is `int` arithmetic
    // (hmatrix/lrmat/sympartialACA.hpp:102) and overflows for blocks above 46340 x 46340, which turns them into DENSE blocks.
    tree_builder.set_minimal_target_depth(minimal_target_depth);
    tree_builder.set_minimal_source_depth(minimal_source_depth);
    {
        py::gil_scoped_release release;
        if (parallel)
            h->hmatrix = std::make_unique<HMatrix<T, double>>(tree_builder.build(exec_compat::par, *h->generator, *target->root, *source->root, target_partition_number, partition_number_for_symmetry));
        else
            h->hmatrix = std::make_unique<HMatrix<T, double>>(tree_builder.build(exec_compat::seq, *h->generator, *target->root, *source->root, target_partition_number, partition_number_for_symmetry));
    }
    return h;
}

// ------------------------------------------------------------------------------------------------------------------
// flat view
// ------------------------------------------------------------------------------------------------------------------
static void collect_leaf_cluster_cuts(const Cluster<double> &c, int base, std::vector<int32_t> &cuts) {
    std::stack<const Cluster<double> *> st;
    st.push(&c);
    std::vector<std::pair<int, int>> leaves;
    while (!st.empty()) {
        const Cluster<double> *cur = st.top();
        st.pop();
        if (cur->is_leaf())
            leaves.emplace_back(cur->get_offset() - base, cur->get_size());
        for (const auto &child : cur->get_children())
            st.push(child.get());
    }
    std::sort(leaves.begin(), leaves.end());
    cuts.clear();
    for (auto &l : leaves)
        cuts.push_back(l.first);
    cuts.push_back(c.get_size());
}

template <typename V>
static py::array_t<V> to_numpy(const std::vector<V> &v) {
    py::array_t<V> a((py::ssize_t)v.size());
    if (!v.empty())
        std::memcpy(a.mutable_data(), v.data(), sizeof(V) * v.size());
    return a;
}

template <typename T>
static py::dict flat_view(const HMatrix<T, double> &H) {
    auto leaves_pair               = get_leaves_from(H);
    const auto &leaves             = leaves_pair.first;
    const auto &leaves_for_symmetry = leaves_pair.second;
    std::unordered_set<const HMatrix<T, double> *> mirrored(leaves_for_symmetry.begin(), leaves_for_symmetry.end());
    const char sym_for_leaves = H.get_symmetry_for_leaves();

    const int t_off = H.get_target_cluster().get_offset();
    const int s_off = H.get_source_cluster().get_offset();
    size_t nl       = leaves.size();
    std::vector<int32_t> row_off(nl), col_off(nl), nrows(nl), ncols(nl), rank(nl);
    std::vector<uint8_t> flags(nl);
    std::vector<uint64_t> a_ptr(nl), b_ptr(nl);
    double stored = 0;
    for (size_t l = 0; l < nl; l++) {
        const auto *leaf = leaves[l];
        row_off[l]       = leaf->get_target_cluster().get_offset() - t_off;
        col_off[l]       = leaf->get_source_cluster().get_offset() - s_off;
        nrows[l]         = leaf->get_target_cluster().get_size();
        ncols[l]         = leaf->get_source_cluster().get_size();
        uint8_t f        = 0;
        if (sym_for_leaves != 'N' && mirrored.count(leaf))
            f |= 1;
        if (leaf->is_dense()) {
            rank[l]  = -1;
            a_ptr[l] = reinterpret_cast<uint64_t>(leaf->get_dense_data()->data());
            b_ptr[l] = 0;
            char s = leaf->get_symmetry(), u = leaf->get_UPLO();
            if (s == 'S')
                f |= (u == 'L' ? 1 : 2) << 1;
            else if (s == 'H')
                f |= (u == 'L' ? 3 : 4) << 1;
            stored += (s != 'N') ? nrows[l] * (ncols[l] + 1.) / 2. : double(nrows[l]) * ncols[l]; // hmatrix_output.hpp:172-178
        } else if (leaf->is_low_rank()) {
            const auto *lr = leaf->get_low_rank_data();
            rank[l]        = lr->rank_of();
            a_ptr[l]       = reinterpret_cast<uint64_t>(lr->get_U().data());
            b_ptr[l]       = reinterpret_cast<uint64_t>(lr->get_V().data());
            stored += double(rank[l]) * (nrows[l] + ncols[l]); // hmatrix_output.hpp:166
        } else {
            throw std::runtime_error("flat_view: leaf is neither dense nor low-rank");
        }
        flags[l] = f;
    }

    std::vector<int32_t> row_cuts, col_cuts;
    collect_leaf_cluster_cuts(H.get_target_cluster(), t_off, row_cuts);
    collect_leaf_cluster_cuts(H.get_source_cluster(), s_off, col_cuts);

    py::dict d;
    d["dtype"]               = std::is_same<T, double>::value ? "d" : "z";
    d["n_rows"]              = H.get_target_cluster().get_size();
    d["n_cols"]              = H.get_source_cluster().get_size();
    d["target_offset"]       = t_off;
    d["source_offset"]       = s_off;
    d["symmetry_for_leaves"] = std::string(1, sym_for_leaves);
    d["symmetry"]            = std::string(1, H.get_symmetry());
    d["UPLO"]                = std::string(1, H.get_UPLO());
    d["row_off"]             = to_numpy(row_off);
    d["col_off"]             = to_numpy(col_off);
    d["nrows"]               = to_numpy(nrows);
    d["ncols"]               = to_numpy(ncols);
    d["rank"]                = to_numpy(rank);
    d["flags"]               = to_numpy(flags);
    d["a_ptr"]               = to_numpy(a_ptr);
    d["b_ptr"]               = to_numpy(b_ptr);
    d["row_cuts"]            = to_numpy(row_cuts);
    d["col_cuts"]            = to_numpy(col_cuts);
    d["stored_coefficients"] = stored;
    // full permutations of the two cluster trees (cluster slot -> user index, clustering/cluster_tree_data.hpp:21)
    const auto &tp           = H.get_target_cluster().get_permutation();
    const auto &sp           = H.get_source_cluster().get_permutation();
    d["target_permutation"]  = py::array_t<int32_t>((py::ssize_t)tp.size(), tp.data());
    d["source_permutation"]  = py::array_t<int32_t>((py::ssize_t)sp.size(), sp.data());
    d["target_global_size"]  = (int)tp.size(); // (get_root_cluster() dangles once the root Cluster has been moved)
    d["source_global_size"]  = (int)sp.size();
    return d;
}

template <typename T>
static std::map<std::string, std::string> info_of(const HMatrix<T, double> &H) { return get_hmatrix_information(H); }

template <typename T>
static void declare_native(py::module &m, const char *name) {
    using N = NativeHMatrix<T>;
    py::class_<N, std::shared_ptr<N>>(m, name)
        .def_property_readonly("shape", [](const N &s) { return std::make_pair(s.get().nb_rows(), s.get().nb_cols()); })
        .def_property_readonly("address", [](const N &s) { return reinterpret_cast<uintptr_t>(s.hmatrix.get()); }, "address of the underlying htool::HMatrix (for oracle/ref_ops only)")
        .def_property_readonly("dtype", [](const N &) { return std::is_same<T, double>::value ? "d" : "z"; })
        .def("flat_view", [](const N &s) { return flat_view(s.get()); })
        .def("get_local_information", [](const N &s) { return info_of(s.get()); })
        .def("get_tree_parameters", [](const N &s) { return get_tree_parameters(s.get()); });
}

PYBIND11_MODULE(_htool_bridge, m) {
    m.doc() = "Flattening layer between the reference's htool::HMatrix and the B200 engine";

    py::class_<NativeClusterTree, std::shared_ptr<NativeClusterTree>>(m, "NativeClusterTree")
        .def_property_readonly("size", [](const NativeClusterTree &t) { return t.n; })
        .def_property_readonly("permutation", [](const NativeClusterTree &t) {
            const auto &p = t.root->get_permutation();
            return py::array_t<int32_t>((py::ssize_t)p.size(), p.data());
        })
        .def_property_readonly("partition", [](const NativeClusterTree &t) {
            std::vector<std::pair<int, int>> out;
            for (const auto *c : t.root->get_clusters_on_partition())
                out.emplace_back(c->get_offset(), c->get_size());
            return out;
        })
        .def_property_readonly("maximal_depth", [](const NativeClusterTree &t) { return t.root->get_maximal_depth(); })
        .def_property_readonly("minimal_depth", [](const NativeClusterTree &t) { return t.root->get_minimal_depth(); });

    declare_native<double>(m, "NativeHMatrix");
    declare_native<zdouble>(m, "NativeComplexHMatrix");

    m.def("create_cluster_tree", &make_cluster_tree, py::arg("coordinates"), py::arg("number_of_children") = 2, py::arg("size_of_partition") = 1, py::arg("maximal_leaf_size") = 64);
    m.def("build_hmatrix", &build_hmatrix<double>, py::arg("kernel_id"), py::arg("wavenumber"), py::arg("target"), py::arg("source"), py::arg("epsilon"), py::arg("eta"), py::arg("symmetry") = 'N', py::arg("UPLO") = 'N', py::arg("target_partition_number") = -1, py::arg("partition_number_for_symmetry") = -1, py::arg("parallel") = true, py::arg("minimal_target_depth") = 0, py::arg("minimal_source_depth") = 0);
    m.def("build_complex_hmatrix", &build_hmatrix<zdouble>, py::arg("kernel_id"), py::arg("wavenumber"), py::arg("target"), py::arg("source"), py::arg("epsilon"), py::arg("eta"), py::arg("symmetry") = 'N', py::arg("UPLO") = 'N', py::arg("target_partition_number") = -1, py::arg("partition_number_for_symmetry") = -1, py::arg("parallel") = true, py::arg("minimal_target_depth") = 0, py::arg("minimal_source_depth") = 0);

    // the reference's own Python objects (only resolvable once `import Htool` has registered the types)
    m.def("flat_view", [](const HMatrix<double, double> &H) { return flat_view(H); });
    m.def("flat_view", [](const HMatrix<zdouble, double> &H) { return flat_view(H); });
    m.def("flat_view", [](const NativeHMatrix<double> &H) { return flat_view(H.get()); });
    m.def("flat_view", [](const NativeHMatrix<zdouble> &H) { return flat_view(H.get()); });

    m.attr("KERNELS") = py::dict(py::arg("laplace") = (int)K_LAPLACE, py::arg("laplace_reg") = (int)K_LAPLACE_REG, py::arg("invdist_reg") = (int)K_INVDIST_REG, py::arg("helmholtz") = (int)K_HELMHOLTZ, py::arg("complex_reg") = (int)K_COMPLEX_REG, py::arg("complex_herm") = (int)K_COMPLEX_HERM);
}

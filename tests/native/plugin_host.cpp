// tests/native/plugin_host.cpp -- TEST INFRASTRUCTURE: hosts the REFERENCE's own Python plugin classes in a module without MPI.
//
// The stock module only registers LocalRenumbering / RestrictedGlobalToLocalOperator / VirtualLocalToLocalOperator under HAVE_MPI
// (src/htool/main.cpp:97-111) although their binding headers need no MPI.  This file includes those headers from where they lie
// under $(REF)/src/htool (nothing is copied) and registers the same classes under the same names, so that the Python plugin of
// htool_python_b200/plugin.py is driven by the reference's real trampolines (zero-copy views of C++ buffers, in-place contract,
// src/htool/local_operator/local_operator.hpp:14-28).  `drive_*` call an operator through the C++ virtual interface, the way
// DistributedOperator does (core/distributed_operator/distributed_operator.hpp:47-53 raw operator pointers).
#include <pybind11/complex.h>
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include "clustering/cluster_node.hpp"
#include "local_operator/local_operator.hpp"
#include "local_operator/local_renumbering.hpp"
#include "local_operator/virtual_local_to_local_operator.hpp"

template <typename T, typename Op>
void declare_drivers(py::module &m, const char *suffix) {
    using Arr = py::array_t<T, py::array::c_style>;
    m.def((std::string("drive_add_vector_product") + suffix).c_str(), [](const Op &op, char trans, T alpha, Arr in, T beta, Arr out) { op.add_vector_product(trans, alpha, in.data(), beta, out.mutable_data()); }, py::arg("op"), py::arg("trans"), py::arg("alpha"), py::arg("in").noconvert(), py::arg("beta"), py::arg("out").noconvert());
    m.def((std::string("drive_add_matrix_product_row_major") + suffix).c_str(), [](const Op &op, char trans, T alpha, Arr in, T beta, Arr out, int mu) { op.add_matrix_product_row_major(trans, alpha, in.data(), beta, out.mutable_data(), mu); }, py::arg("op"), py::arg("trans"), py::arg("alpha"), py::arg("in").noconvert(), py::arg("beta"), py::arg("out").noconvert(), py::arg("mu"));
    m.def((std::string("drive_add_sub_matrix_product_to_local") + suffix).c_str(), [](const Op &op, Arr in, Arr out, int mu, int offset, int size) { op.add_sub_matrix_product_to_local(in.data(), out.mutable_data(), mu, offset, size); }, py::arg("op"), py::arg("in").noconvert(), py::arg("out").noconvert(), py::arg("mu"), py::arg("offset"), py::arg("size"));
}

PYBIND11_MODULE(htool_plugin_host, m) {
    // htool::Cluster is NOT registered here (the stock module does that; two registrations in one interpreter collide): the
    // LocalRenumbering(cluster) constructor works once `Htool` is imported, LocalRenumbering(offset, size, permutation) always
    declare_local_renumbering<double>(m, "LocalRenumbering");
    declare_global_to_local_operator<double>(m, "");
    declare_global_to_local_operator<std::complex<double>>(m, "Complex");
    declare_virtual_local_to_local_operator<double>(m, "");
    declare_virtual_local_to_local_operator<std::complex<double>>(m, "Complex");
    declare_drivers<double, htool::VirtualGlobalToLocalOperator<double>>(m, "_g2l_d");
    declare_drivers<std::complex<double>, htool::VirtualGlobalToLocalOperator<std::complex<double>>>(m, "_g2l_z");
    declare_drivers<double, htool::VirtualLocalToLocalOperator<double>>(m, "_l2l_d");
    declare_drivers<std::complex<double>, htool::VirtualLocalToLocalOperator<std::complex<double>>>(m, "_l2l_z");
}

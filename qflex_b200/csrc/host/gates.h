// Gate tensors (complex64, row-major, last index fastest).
// Conventions (reference src/read_circuit.cpp:27-165):
//   one-qubit gate  : indices (in, out), 2x2
//   two-qubit gate  : split into two rank-3 tensors (in, virtual, out), virtual
//                     dimension 2 for cz/cx (balanced SVD) and 4 for fsim.
#ifndef QFLEX_HOST_GATES_H_
#define QFLEX_HOST_GATES_H_

#include <complex>
#include <string>
#include <tuple>
#include <vector>

namespace qflex {

using s_type = std::complex<float>;

// delta_0, delta_1, h, hz_1_2, t, x_1_2, y_1_2, rz(theta/pi), plus the halves
// cz_q1/cz_q2/cx_q1/cx_q2.  Throws "Invalid gate name provided: ...".
std::vector<s_type> gate_array(const std::string& gate_name,
                               const std::vector<double>& params);

// cz / cx / fsim(theta/pi, phi/pi) -> (tensor on q1, tensor on q2, dims).
std::tuple<std::vector<s_type>, std::vector<s_type>, std::vector<std::size_t>>
gate_arrays(const std::string& gate_name, const std::vector<double>& params);

}  // namespace qflex

#endif  // QFLEX_HOST_GATES_H_

// Gate tables.  Values and float-precision evaluation order follow the
// reference (src/read_circuit.cpp:27-165) so that site tensors agree to the
// last bit: constants are rounded double->float exactly once, rz and fsim are
// evaluated in complex<float>.
#include "gates.h"

#include <cmath>

#include "errors.h"

namespace qflex {
namespace {

constexpr float kR = static_cast<float>(M_SQRT1_2);

// balanced-SVD halves of CZ and CX (in, virtual, out)
constexpr double kCzA = 0.3446133714352872, kCzB = 1.1381806476131544;
constexpr double kCzC = 1.0484937059720079, kCzD = 0.5611368023131075;
constexpr double kCxA = 0.8408964152537143;  // 2^(-1/4)
constexpr double kCxB = 0.5946035575013604;  // 2^(-3/4)

std::vector<s_type> reals(std::initializer_list<double> v) {
  std::vector<s_type> out;
  for (double x : v) out.emplace_back(static_cast<float>(x), 0.f);
  return out;
}

// fSim(theta, phi) split across the two qubits; both halves are the same
// 2x4x2 array (reference src/read_circuit.cpp:105-139).
std::vector<s_type> fsim_half(float theta, float phi) {
  const s_type half(0.5f);
  const s_type e = std::exp(s_type(0.0, -1 * phi / 2));
  const s_type c(std::cos(theta), 0.0);
  const s_type c1 = half * (e + c);
  const s_type c2 = half * (e - c);
  const s_type c3 = s_type(0.0, -1 / 2.) * s_type(std::sin(theta), 0.0);
  const s_type s1 = std::sqrt(c1), s2 = std::sqrt(c2), s3 = std::sqrt(c3);
  const s_type s4 = s3;
  const s_type d = std::exp(s_type(0.0, phi / 4.));
  const s_type z(0.f, 0.f), i(0.f, 1.f), mi(0.f, -1.f);
  return {d * s1, z, d * s2,        z, z,  s3, z,       i * s4,
          z, std::conj(d) * s1, z, -std::conj(d) * s2, s3, z, mi * s4, z};
}

}  // namespace

std::vector<s_type> gate_array(const std::string& name,
                               const std::vector<double>& params) {
  if (name == "rz") {
    const double angle = M_PI * params.at(0);
    const s_type phase = std::exp(s_type(0., angle / 2));
    return {std::conj(phase), 0., 0., phase};
  }
  if (name == "delta_0") return reals({1, 0});
  if (name == "delta_1") return reals({0, 1});
  if (name == "h") return {kR, kR, kR, -kR};
  if (name == "hz_1_2") return {{0.5f, 0.5f}, {kR, 0.f}, {0.f, -kR}, {0.5f, 0.5f}};
  if (name == "t") return {1.f, 0.f, 0.f, {kR, kR}};
  if (name == "x_1_2")
    return {{0.5f, 0.5f}, {0.5f, -0.5f}, {0.5f, -0.5f}, {0.5f, 0.5f}};
  if (name == "y_1_2")
    return {{0.5f, 0.5f}, {0.5f, 0.5f}, {-0.5f, -0.5f}, {0.5f, 0.5f}};
  if (name == "cz_q1") return reals({-kCzA, 0, kCzB, 0, 0, -kCzB, 0, -kCzA});
  if (name == "cz_q2") return reals({-kCzC, 0, kCzD, 0, 0, kCzD, 0, kCzC});
  if (name == "cx_q1") return reals({kCxA, 0, kCxA, 0, 0, -kCxA, 0, kCxA});
  if (name == "cx_q2")
    return reals({kCxB, -kCxB, kCxB, kCxB, -kCxB, kCxB, kCxB, kCxB});
  if (name == "cz") return reals({1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, -1});
  if (name == "cx") return reals({1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 0, 1, 0, 0, 1, 0});
  throw ERROR_MSG("Invalid gate name provided: ", name);
}

std::tuple<std::vector<s_type>, std::vector<s_type>, std::vector<std::size_t>>
gate_arrays(const std::string& name, const std::vector<double>& params) {
  if (name == "cz")
    return {gate_array("cz_q1", params), gate_array("cz_q2", params), {2, 2, 2}};
  if (name == "cx")
    return {gate_array("cx_q1", params), gate_array("cx_q2", params), {2, 2, 2}};
  if (name == "fsim") {
    const double theta = M_PI * params.at(0), phi = M_PI * params.at(1);
    const auto half = fsim_half(static_cast<float>(theta), static_cast<float>(phi));
    return {half, half, {2, 4, 2}};
  }
  throw ERROR_MSG("Invalid gate name provided: ", name);
}

}  // namespace qflex

// Forwarding header.  The reference's read_circuit.h declares the network builder and
// flatten_grid_of_tensors on host Tensors (src/read_circuit.h:73-111); callers above
// EvaluateCircuit (src/pybind_main.h:14) only need it for the types it drags in.  Here
// the network is built and flattened by the planner (plan.h) on the device.
#ifndef QFLEX_HOST_FWD_READ_CIRCUIT_H_
#define QFLEX_HOST_FWD_READ_CIRCUIT_H_
#include <complex>
#include <fstream>
#include <iostream>
#include <sstream>
#include <string>
#include <vector>

#include "circuit.h"
#include "contraction_utils.h"
#include "global.h"
#include "tensor.h"
#endif

// Forwarding header: reference front-ends (src/main.cpp:3, src/pybind_main.h) include
// "circuit.h" for QflexCircuit/QflexGate (src/circuit.h:12-36); here they live in input.h.
#ifndef QFLEX_HOST_FWD_CIRCUIT_H_
#define QFLEX_HOST_FWD_CIRCUIT_H_
#include <fstream>
#include <sstream>

#include "input.h"
#endif

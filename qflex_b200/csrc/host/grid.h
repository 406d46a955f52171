// Forwarding header for QflexGrid (reference src/grid.h:13-20) -> input.h.
#ifndef QFLEX_HOST_FWD_GRID_H_
#define QFLEX_HOST_FWD_GRID_H_
#include <fstream>

#include "input.h"
#endif

// Forwarding header for QflexOrdering (reference src/ordering.h:13-19) -> input.h.
#ifndef QFLEX_HOST_FWD_ORDERING_H_
#define QFLEX_HOST_FWD_ORDERING_H_
#include <fstream>

#include "input.h"
#endif

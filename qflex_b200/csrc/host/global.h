// Forwarding header: qflex::global::{verbose, memory_limit, track_memory_seconds}
// (reference src/global.h:9-15) are declared in errors.h.
#ifndef QFLEX_HOST_FWD_GLOBAL_H_
#define QFLEX_HOST_FWD_GLOBAL_H_
#include "errors.h"
#endif

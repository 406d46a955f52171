// Forwarding header: qflex::utils::{concat, readable_memory_string,
// from_readable_memory_string, get_date_time} (reference src/utils.h:14-60) are
// declared in errors.h next to ERROR_MSG / WARN_MSG.
#ifndef QFLEX_HOST_FWD_UTILS_H_
#define QFLEX_HOST_FWD_UTILS_H_
#include <iostream>
#include <map>
#include <string>

#include "errors.h"
#endif

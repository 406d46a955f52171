// Forwarding header.  The reference's host `Tensor` (src/tensor.h:34-344) has no
// counterpart object here: tensors are spans of the device arena described by
// SymTensor (plan.h) and moved by the C ABI of include/qflex_b200.h.  Front-ends only
// use the scalar type.
#ifndef QFLEX_HOST_FWD_TENSOR_H_
#define QFLEX_HOST_FWD_TENSOR_H_
#include <complex>

#include "gates.h"
#include "plan.h"
#endif

// kb_todo.cu -- entry points declared in include/katib_b200.h whose kernels are not written yet would live here and
// fail loudly (KB_ERR_UNSUPPORTED).  Every declared entry point is implemented now; nothing computes on the CPU.
#include "kb_internal.cuh"

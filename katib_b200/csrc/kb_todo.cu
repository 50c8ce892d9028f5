// kb_todo.cu -- entry points declared in include/katib_b200.h whose kernels are not written yet.
// They fail loudly (KB_ERR_UNSUPPORTED); nothing here computes on the CPU.
#include "kb_internal.cuh"
#define KB_TODO(name) return kb_fail(ctx, KB_ERR_UNSUPPORTED, name ": not implemented yet")
extern "C" int kb_cmaes_create(kb_ctx *ctx, int, const double *, double, int, const double *, const double *) { KB_TODO("kb_cmaes_create"); }
extern "C" int kb_cmaes_ask(kb_ctx *ctx, int, const double *, uint64_t, double *) { KB_TODO("kb_cmaes_ask"); }
extern "C" int kb_cmaes_tell(kb_ctx *ctx, const double *, const double *) { KB_TODO("kb_cmaes_tell"); }
extern "C" int kb_cmaes_get_state(kb_ctx *ctx, double *, double *, double *, double *, double *, double *, double *, int *) { KB_TODO("kb_cmaes_get_state"); }
extern "C" int kb_cmaes_set_eig(kb_ctx *ctx, const double *, const double *) { KB_TODO("kb_cmaes_set_eig"); }
extern "C" int kb_eigh(kb_ctx *ctx, int, const double *, double *, double *) { KB_TODO("kb_eigh"); }
int kb_cmaes_free(kb_ctx *) { return KB_OK; }

// Single translation unit for libsc_b200.so: the kernels are templates shared by the entry points,
// and one TU keeps exactly one instance of each (and of the per-kernel attribute caches).
#include "sae_step.cu"
#include "sae_step_h.cu"
#include "infer.cu"
#include "fista.cu"
#include "noalloc.cu"
#include "gather.cu"

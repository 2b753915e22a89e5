// kernels.cu — dispatch from (variant, durations, what) to the kernel instantiations (kernels_variant.cu) and the
// shared-memory sizing they share.
#include <cuda_runtime.h>

#include "kernels.h"

namespace tb {

// (lanes, rt, streams, mode, curvature) tuples the Makefile builds (LANE_OBJS there); mode 0 lock-step, 1 ring, 2 split
// The streams / ring variants (eval_kernel.cuh) and the producer / consumer split (eval_kernel_split.cuh) measured slower
// than lock-step on every workload:
// they are built only with `make EXPERIMENTS=1` (-DTB_EXPERIMENTS)
#ifdef TB_EXPERIMENTS
#define TB_EXPERIMENT_VARIANTS(X) X(1, 0, 1, 1, 0) X(1, 0, 2, 0, 0) X(1, 0, 2, 1, 0) X(1, 1, 2, 1, 0) X(2, 0, 1, 2, 0) X(2, 1, 1, 2, 0)
#else
#define TB_EXPERIMENT_VARIANTS(X)
#endif
#define TB_VARIANTS(X) X(1, 0, 1, 0, 0) X(1, 1, 1, 0, 0) X(2, 0, 1, 0, 0) X(2, 1, 1, 0, 0) X(4, 0, 1, 0, 0) X(4, 1, 1, 0, 0) \
                       X(1, 0, 1, 0, 1) X(1, 1, 1, 0, 1) X(2, 0, 1, 0, 1) X(2, 1, 1, 0, 1) TB_EXPERIMENT_VARIANTS(X)

#define TB_DECL(L, R, S, G, C)                                                                                             \
  cudaError_t LaunchV##L##_##R##_##S##_##G##_##C(const DeviceProgram&, const ProblemConst*, int, const double*, double*, double*, \
                                                 int*, int, int, int, cudaStream_t);                                        \
  int OccupancyV##L##_##R##_##S##_##G##_##C(const DeviceProgram&);
TB_VARIANTS(TB_DECL)
#undef TB_DECL

size_t EvalSharedBytes(const DeviceProgram& P, int lanes, int streams, int mode) {
  return static_cast<size_t>(P.blob.bytes) + sizeof(double) * streams * static_cast<size_t>(StreamDoubles(P, lanes, mode)) + 16;
}

bool EvalVariantBuilt(const DeviceProgram& P, const EvalVariant& v) {
  const int rt = P.blob.rt != 0, curv = P.curvature_dims != 0;
#define TB_HAS(L, R, S, G, C) if (v.lanes == L && rt == R && v.streams == S && v.mode() == G && curv == C) return true;
  TB_VARIANTS(TB_HAS)
#undef TB_HAS
  return false;
}

cudaError_t LaunchEval(const DeviceProgram& P, const ProblemConst* pcs, int what, const double* x, double* g,
                       double* vals, int* status, int n_problems, const EvalVariant& v, int max_ctas, cudaStream_t stream) {
  if (n_problems <= 0) return cudaSuccess;
  const int rt = P.blob.rt != 0, curv = P.curvature_dims != 0;
#define TB_GO(L, R, S, G, C)                                                                  \
  if (v.lanes == L && rt == R && v.streams == S && v.mode() == G && curv == C)       \
    return LaunchV##L##_##R##_##S##_##G##_##C(P, pcs, what, x, g, vals, status, n_problems, max_ctas, v.skew, stream);
  TB_VARIANTS(TB_GO)
#undef TB_GO
  return cudaErrorInvalidValue;
}

int EvalOccupancy(const DeviceProgram& P, const EvalVariant& v) {
  const int rt = P.blob.rt != 0, curv = P.curvature_dims != 0;
#define TB_OCC(L, R, S, G, C) \
  if (v.lanes == L && rt == R && v.streams == S && v.mode() == G && curv == C) return OccupancyV##L##_##R##_##S##_##G##_##C(P);
  TB_VARIANTS(TB_OCC)
#undef TB_OCC
  return 0;
}

}  // namespace tb

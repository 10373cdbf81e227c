// Instantiations of the streaming kernel for ONE nr_comp (compile with -DPPLS_INST_R=r).
// (R, M, T): T*R even and <= 32; M bounded by the register budget (regs_estimate <= 255).
#include "em_stream_kernel.cuh"

#ifndef PPLS_INST_R
#error "compile with -DPPLS_INST_R=<1..8>"
#endif

namespace ppls {

#define PPLS_CAT_(a, b) a##b
#define PPLS_CAT(a, b) PPLS_CAT_(a, b)

template <int R, int M, int T, bool KEEP>
static stream_kernel_t kernel_if() {
  if constexpr (T * R <= 32 && (T * R) % 2 == 0 && regs_estimate(R, M, T, KEEP) <= 255 + 12) return em_stream_kernel<R, M, T, KEEP>;
  else return nullptr;
}

stream_kernel_t PPLS_CAT(lookup_kernel_r, PPLS_INST_R)(int M, int T, int keep) {
  constexpr int R = PPLS_INST_R;
#define PPLS_K(m, t) if (M == m && T == t) return keep ? kernel_if<R, m, t, true>() : kernel_if<R, m, t, false>();
  PPLS_K(1, 2) PPLS_K(2, 2) PPLS_K(3, 2) PPLS_K(4, 2) PPLS_K(5, 2) PPLS_K(6, 2)
  PPLS_K(1, 4) PPLS_K(2, 4) PPLS_K(3, 4) PPLS_K(4, 4) PPLS_K(5, 4) PPLS_K(6, 4)
  PPLS_K(1, 8) PPLS_K(2, 8) PPLS_K(3, 8) PPLS_K(4, 8) PPLS_K(5, 8) PPLS_K(6, 8)
#undef PPLS_K
  return nullptr;
}

}  // namespace ppls

// tk_inst.cu — instantiates tk_kernel for ONE (arithmetic, lane layout, extras) combination,
// chosen on the command line:  -DH2D_INST_ARITH=0|1|2  -DH2D_INST_LAY=0|1|2  -DH2D_INST_XTR=0|1.
// heat2d_b200/build.py compiles this file once per combination, in parallel (all K = 1..8 of
// one combination take 20-60 s of ptxas time; a single translation unit took minutes).
// Experiment hook: -DH2D_TK_KMIN=8 builds only the deepest kernels.
#include "tk_kernel.cuh"

#if !defined(H2D_INST_ARITH) || !defined(H2D_INST_LAY) || !defined(H2D_INST_XTR)
#error "define H2D_INST_ARITH, H2D_INST_LAY and H2D_INST_XTR"
#endif
#ifndef H2D_TK_KMIN
#define H2D_TK_KMIN 1
#endif

namespace h2d {

// the 8-column layout exists for sweeps of at most TK_V8_KMAX levels
constexpr int INST_KMAX = H2D_INST_LAY == LAY_V8 ? TK_V8_KMAX : TK_KMAX;

template <int K, bool IMM>
static tk_fn_t one() {
  if (K < H2D_TK_KMIN || K > INST_KMAX) return nullptr;
  constexpr int KK = K < H2D_TK_KMIN ? H2D_TK_KMIN : (K > INST_KMAX ? INST_KMAX : K);
  return tk_kernel<KK, H2D_INST_ARITH, H2D_INST_LAY, IMM, (H2D_INST_XTR != 0)>;
}

template <bool IMM>
static tk_fn_t pick_k(int k) {
  switch (k) {
    case 1: return one<1, IMM>();
    case 2: return one<2, IMM>();
    case 3: return one<3, IMM>();
    case 4: return one<4, IMM>();
    case 5: return one<5, IMM>();
    case 6: return one<6, IMM>();
    case 7: return one<7, IMM>();
    case 8: return one<8, IMM>();
    default: return nullptr;
  }
}

template <>
tk_fn_t tk_get<H2D_INST_ARITH, H2D_INST_LAY, (H2D_INST_XTR != 0)>(int k, bool imm) {
#if H2D_INST_ARITH == 0
  return imm ? pick_k<true>(k) : pick_k<false>(k);
#else
  (void)imm;   // strict modes take their coefficients as doubles: no immediate-operand variant
  return pick_k<false>(k);
#endif
}

}  // namespace h2d

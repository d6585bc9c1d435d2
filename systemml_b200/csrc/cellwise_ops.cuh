// cellwise_ops.cuh -- the value functions of the cellwise engine (binary and unary), Java semantics.
// Included by cellwise.cu and ALSO embedded as text into the library (build.py writes cellwise_ops_embed.inc)
// so that fused.cu can hand the very same definitions to NVRTC when it compiles a fused cellwise chain:
// one source of truth for what "+", "%%", "min" ... mean.  Needs only the SYSML_OP_* / SYSML_UN_* macros of
// include/sysml_b200.h and the CUDA math functions -- no other header.
//
// Reference: runtime/functionobjects/{Plus,Minus,Multiply,Divide,Power,Modulus,IntegerDivide,Minus1Multiply,
// MinusNz,LessThan...,And,Or,Builtin}.java; opcode numbering = LibMatrixCUDA.getBinaryOp (LibMatrixCUDA.java:1551-1575).
#ifndef SYSML_CELLWISE_OPS_CUH
#define SYSML_CELLWISE_OPS_CUH

namespace sysml_ops {

// Java Math.max/min: NaN propagates, -0.0 < +0.0 (functionobjects/Builtin.java:356-361)
__device__ __forceinline__ double jmax(double a, double b) {
  if (a != a) return a;
  if (b != b) return b;
  if (a == 0.0 && b == 0.0) return __longlong_as_double(__double_as_longlong(a) & __double_as_longlong(b));
  return (a >= b) ? a : b;
}
__device__ __forceinline__ double jmin(double a, double b) {
  if (a != a) return a;
  if (b != b) return b;
  if (a == 0.0 && b == 0.0) return __longlong_as_double(__double_as_longlong(a) | __double_as_longlong(b));
  return (a <= b) ? a : b;
}
// IntegerDivide.executeIntDiv (functionobjects/IntegerDivide.java:62-73): floor(x/y + 2^-53) via long cast
__device__ __forceinline__ double intdiv(double x, double y) {
  const double r = __ddiv_rn(x, y);
  if (isnan(r) || isinf(r)) return r;
  return (double)(long long)floor(__dadd_rn(r, 0x1.0p-53));
}

template <int OP>
__device__ __forceinline__ double binop(double x, double y) {
  switch (OP) {
    case SYSML_OP_PLUS: return __dadd_rn(x, y);
    case SYSML_OP_MINUS: return __dsub_rn(x, y);
    case SYSML_OP_MULTIPLY: return __dmul_rn(x, y);
    case SYSML_OP_DIVIDE: return __ddiv_rn(x, y);
    case SYSML_OP_POWER: return pow(x, y);
    case SYSML_OP_LESS: return (x < y) ? 1.0 : 0.0;
    case SYSML_OP_LESSEQUAL: return (x <= y) ? 1.0 : 0.0;
    case SYSML_OP_GREATER: return (x > y) ? 1.0 : 0.0;
    case SYSML_OP_GREATEREQUAL: return (x >= y) ? 1.0 : 0.0;
    case SYSML_OP_EQUAL: return (x == y) ? 1.0 : 0.0;
    case SYSML_OP_NOTEQUAL: return (x != y) ? 1.0 : 0.0;
    case SYSML_OP_MIN: return jmin(x, y);
    case SYSML_OP_MAX: return jmax(x, y);
    case SYSML_OP_AND: return ((x != 0.0) && (y != 0.0)) ? 1.0 : 0.0;  // functionobjects/And.java:46-48
    case SYSML_OP_OR: return ((x != 0.0) || (y != 0.0)) ? 1.0 : 0.0;   // functionobjects/Or.java:46-48
    case SYSML_OP_MINUS1MULTIPLY: return __dsub_rn(1.0, __dmul_rn(x, y));  // no FMA: Java rounds twice
    case SYSML_OP_MINUSNZ: return (x != 0.0) ? __dsub_rn(x, y) : 0.0;
    case SYSML_OP_MODULUS:  // functionobjects/Modulus.java:49-53
      if (y == 0.0) return __longlong_as_double(0x7ff8000000000000LL);
      return __dsub_rn(x, __dmul_rn(intdiv(x, y), y));
    case SYSML_OP_INTDIV: return intdiv(x, y);
  }
  return 0.0;
}

template <int OP>
__device__ __forceinline__ double unop(double x) {
  switch (OP) {
    case SYSML_UN_EXP: return exp(x);
    case SYSML_UN_SQRT: return sqrt(x);
    case SYSML_UN_ROUND: {
      // Builtin ROUND = (double) Math.round(x) (functionobjects/Builtin.java:314): the closest long, ties towards +infinity,
      // computed exactly (floor(x + 0.5) with a ROUNDED sum is wrong for 0.49999999999999994 and for odd integers
      // above 2^52); NaN -> 0; beyond the long range the value saturates at +-2^63; a long has no negative zero
      if (x != x) return 0.0;
      if (fabs(x) >= 4503599627370496.0) return (x >= 9223372036854775808.0) ? 9223372036854775808.0 : ((x <= -9223372036854775808.0) ? -9223372036854775808.0 : x);
      const double r = floor(x);  // exact below 2^52, and so is x - r
      return __dadd_rn((__dsub_rn(x, r) >= 0.5) ? __dadd_rn(r, 1.0) : r, 0.0);
    }
    case SYSML_UN_ABS: return fabs(x);
    case SYSML_UN_LOG: return log(x);
    case SYSML_UN_FLOOR: return floor(x);
    case SYSML_UN_CEIL: return ceil(x);
    case SYSML_UN_SIGN: return (x > 0.0) ? 1.0 : ((x < 0.0) ? -1.0 : x);  // Math.signum: zeros of either sign and NaN pass through
    case SYSML_UN_SIGMOID: return 1.0 / (1.0 + exp(-x));
    case SYSML_UN_SIN: return sin(x);
    case SYSML_UN_COS: return cos(x);
    case SYSML_UN_TAN: return tan(x);
    case SYSML_UN_TANH: return tanh(x);
  }
  return 0.0;
}

}  // namespace sysml_ops

#endif  // SYSML_CELLWISE_OPS_CUH

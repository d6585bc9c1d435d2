// cellwise_rt.cuh -- the value functions of cellwise_ops.cuh behind a RUN-TIME opcode (a switch over the templated
// definitions; one source of truth), and a small interpreter for the postfix programs of include/sysml_b200.h
// (SYSML_FUSED_TOKEN) evaluated once per ROW inside a streaming kernel -- where a per-row cost of a few dozen
// instructions does not show next to the row's kilobytes.  The fused cellwise kernels do NOT use this: they are compiled
// per program (fused.cu).
#ifndef SYSML_CELLWISE_RT_CUH
#define SYSML_CELLWISE_RT_CUH

#include "cellwise_ops.cuh"

namespace sysml_rt {

#define SYSML_RT_ALL_BINOPS(X) \
  X(0) X(1) X(2) X(3) X(4) X(5) X(6) X(7) X(8) X(9) X(10) X(11) X(12) X(13) X(14) X(15) X(16) X(17) X(18)
#define SYSML_RT_ALL_UNOPS(X) X(0) X(1) X(2) X(3) X(4) X(5) X(6) X(7) X(8) X(9) X(10) X(11) X(12)

static __device__ __noinline__ double binop_rt(int op, double x, double y) {
  switch (op) {
#define X(o) \
  case o: return sysml_ops::binop<o>(x, y);
    SYSML_RT_ALL_BINOPS(X)
#undef X
  }
  return 0.0;
}
static __device__ __noinline__ double unop_rt(int op, double x) {
  switch (op) {
#define X(o) \
  case o: return sysml_ops::unop<o>(x);
    SYSML_RT_ALL_UNOPS(X)
#undef X
  }
  return 0.0;
}

// a validated postfix program (host side: validate below), at most 24 tokens, stack depth at most 6
struct RowProgram {
  int n_tokens;
  int n_results;  // values left on the stack: 1 or 2
  int tok[24];
  double k[4];
};

// in0..in2: the row's inputs; r0 (and r1 when the program leaves two values; result 0 is the deeper one).
// One thread runs this per row on the critical path of its tile, so it is written for latency: the stack is six
// registers that shift (no local memory), the operators are inlined switches, the tokens come from the kernel's
// constant bank.
static __device__ __forceinline__ void row_eval(const RowProgram& f, double in0, double in1, double in2, double& r0, double& r1) {
  double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0, s4 = 0.0, s5 = 0.0;  // s0 = top of stack
  for (int t = 0; t < f.n_tokens; t++) {
    const int tk = f.tok[t];
    const int kind = (tk >> 24) & 0xff, arg = tk & 0xffffff;
    if (kind == SYSML_FUSED_BINARY) {
      double v;
      switch (arg) {
#define X(o) \
  case o: v = sysml_ops::binop<o>(s1, s0); break;
        SYSML_RT_ALL_BINOPS(X)
#undef X
        default: v = 0.0;
      }
      s0 = v; s1 = s2; s2 = s3; s3 = s4; s4 = s5;
    } else if (kind == SYSML_FUSED_UNARY) {
      switch (arg) {
#define X(o) \
  case o: s0 = sysml_ops::unop<o>(s0); break;
        SYSML_RT_ALL_UNOPS(X)
#undef X
      }
    } else if (kind == SYSML_FUSED_NIP) {
      s1 = s2; s2 = s3; s3 = s4; s4 = s5;
    } else {
      double v;
      if (kind == SYSML_FUSED_IN) v = arg == 0 ? in0 : (arg == 1 ? in1 : in2);
      else if (kind == SYSML_FUSED_CONST) v = f.k[arg];
      else v = arg == 0 ? s0 : (arg == 1 ? s1 : (arg == 2 ? s2 : (arg == 3 ? s3 : (arg == 4 ? s4 : s5))));  // PICK
      s5 = s4; s4 = s3; s3 = s2; s2 = s1; s1 = s0; s0 = v;
    }
  }
  if (f.n_results > 1) {
    r0 = s1;
    r1 = s0;
  } else {
    r0 = s0;
    r1 = 0.0;
  }
}

// host: checks a program against the interpreter's limits; returns nullptr or what is wrong with it
static inline const char* validate(const int32_t* prog, int n_tokens, int n_inputs, int n_consts, int n_results) {
  if (n_tokens < 1 || n_tokens > 24) return "between 1 and 24 tokens";
  if (n_consts < 0 || n_consts > 4) return "at most 4 constants";
  if (n_results < 1 || n_results > 2) return "one or two results";
  int sp = 0;
  for (int t = 0; t < n_tokens; t++) {
    const int kind = (prog[t] >> 24) & 0xff, arg = prog[t] & 0xffffff;
    switch (kind) {
      case SYSML_FUSED_IN:
        if (arg >= n_inputs) return "input index out of range";
        sp++;
        break;
      case SYSML_FUSED_CONST:
        if (arg >= n_consts) return "constant index out of range";
        sp++;
        break;
      case SYSML_FUSED_BINARY:
        if (arg > SYSML_OP_INTDIV) return "unknown binary opcode";
        if (sp < 2) return "stack underflow at a binary operator";
        sp--;
        break;
      case SYSML_FUSED_UNARY:
        if (arg > SYSML_UN_TANH) return "unknown unary opcode";
        if (sp < 1) return "stack underflow at a unary operator";
        break;
      case SYSML_FUSED_PICK:
        if (arg >= sp) return "pick reaches below the stack";
        sp++;
        break;
      case SYSML_FUSED_NIP:
        if (sp < 2) return "stack underflow at nip";
        sp--;
        break;
      default: return "unknown token kind";
    }
    if (sp > 6) return "stack deeper than 6";
  }
  if (sp != n_results) return "program must leave exactly one value per result on the stack";
  return nullptr;
}

}  // namespace sysml_rt

#endif  // SYSML_CELLWISE_RT_CUH

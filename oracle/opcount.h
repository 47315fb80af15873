/* CPU ORACLE — TEST INFRASTRUCTURE ONLY.
 *
 * Op-count build of mjstep_oracle.c ("binding rule": the oracle carries a
 * compile-time counter of the fp64 operations it actually executes).  The oracle source
 * is compiled unchanged as C++ with this header force-included; the last line renames
 * `double` to a counting wrapper of the same size and layout, so every +, -, *, /, sqrt
 * and transcendental the oracle performs on its way through a step bumps a thread-local
 * counter.  Recipe: oracle/Makefile target _build/liboracle_ops.so; driver:
 * tools/count_ops.py -> profiles/r1_oracle_opcount.json -> bench.py roofline.
 */
#ifndef BP_ORACLE_OPCOUNT_H_
#define BP_ORACLE_OPCOUNT_H_
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <type_traits>

enum { OP_ADD, OP_MUL, OP_DIV, OP_SQRT, OP_TRANS, OP_CMP, OP_NCLASS };
static __thread unsigned long long o_ops[OP_NCLASS];

struct cdouble {
  double v;
  cdouble() = default;
  template <class T, class = typename std::enable_if<std::is_arithmetic<T>::value>::type>
  cdouble(T x) : v((double)x) {}
  explicit operator double() const { return v; }
  explicit operator int() const { return (int)v; }
  explicit operator long() const { return (long)v; }
  explicit operator bool() const { return v != 0; }
  cdouble operator-() const { return cdouble(-v); }
  cdouble& operator+=(cdouble o) { o_ops[OP_ADD]++; v += o.v; return *this; }
  cdouble& operator-=(cdouble o) { o_ops[OP_ADD]++; v -= o.v; return *this; }
  cdouble& operator*=(cdouble o) { o_ops[OP_MUL]++; v *= o.v; return *this; }
  cdouble& operator/=(cdouble o) { o_ops[OP_DIV]++; v /= o.v; return *this; }
};
static_assert(sizeof(cdouble) == sizeof(double), "layout");

#define BP_BIN(op, cls)                                                                   \
  static inline cdouble operator op(cdouble a, cdouble b) { o_ops[cls]++; return cdouble(a.v op b.v); } \
  template <class T, class = typename std::enable_if<std::is_arithmetic<T>::value>::type>  \
  static inline cdouble operator op(cdouble a, T b) { o_ops[cls]++; return cdouble(a.v op (double)b); } \
  template <class T, class = typename std::enable_if<std::is_arithmetic<T>::value>::type>  \
  static inline cdouble operator op(T a, cdouble b) { o_ops[cls]++; return cdouble((double)a op b.v); }
BP_BIN(+, OP_ADD) BP_BIN(-, OP_ADD) BP_BIN(*, OP_MUL) BP_BIN(/, OP_DIV)
#undef BP_BIN
#define BP_CMP(op)                                                                        \
  static inline bool operator op(cdouble a, cdouble b) { o_ops[OP_CMP]++; return a.v op b.v; } \
  template <class T, class = typename std::enable_if<std::is_arithmetic<T>::value>::type>  \
  static inline bool operator op(cdouble a, T b) { o_ops[OP_CMP]++; return a.v op (double)b; } \
  template <class T, class = typename std::enable_if<std::is_arithmetic<T>::value>::type>  \
  static inline bool operator op(T a, cdouble b) { o_ops[OP_CMP]++; return (double)a op b.v; }
BP_CMP(<) BP_CMP(>) BP_CMP(<=) BP_CMP(>=) BP_CMP(==) BP_CMP(!=)
#undef BP_CMP

#define BP_FN1(name, cls) \
  static inline cdouble name(cdouble a) { o_ops[cls]++; return cdouble(::name(a.v)); }
BP_FN1(sqrt, OP_SQRT) BP_FN1(sin, OP_TRANS) BP_FN1(cos, OP_TRANS) BP_FN1(tan, OP_TRANS)
BP_FN1(exp, OP_TRANS) BP_FN1(log, OP_TRANS) BP_FN1(log1p, OP_TRANS) BP_FN1(acos, OP_TRANS)
BP_FN1(asin, OP_TRANS) BP_FN1(atan, OP_TRANS) BP_FN1(tanh, OP_TRANS) BP_FN1(asinh, OP_TRANS)
#undef BP_FN1
static inline cdouble fabs(cdouble a) { return cdouble(::fabs(a.v)); }
static inline cdouble floor(cdouble a) { return cdouble(::floor(a.v)); }
static inline cdouble ceil(cdouble a) { return cdouble(::ceil(a.v)); }
static inline cdouble atan2(cdouble a, cdouble b) { o_ops[OP_TRANS]++; return cdouble(::atan2(a.v, b.v)); }
static inline cdouble pow(cdouble a, cdouble b) { o_ops[OP_TRANS]++; return cdouble(::pow(a.v, b.v)); }
static inline cdouble fmax(cdouble a, cdouble b) { o_ops[OP_CMP]++; return cdouble(::fmax(a.v, b.v)); }
static inline cdouble fmin(cdouble a, cdouble b) { o_ops[OP_CMP]++; return cdouble(::fmin(a.v, b.v)); }
#undef isnan
#undef isinf
static inline bool isnan(cdouble a) { return a.v != a.v; }
static inline bool isinf(cdouble a) { return std::isinf(a.v); }

/* counters of the calling thread: out[OP_NCLASS] */
extern "C" void o_opcount_read(unsigned long long* out) {
  for (int k = 0; k < OP_NCLASS; k++) out[k] = o_ops[k];
}
extern "C" void o_opcount_reset(void) { memset(o_ops, 0, sizeof(o_ops)); }

#define double cdouble
#endif

// Forward-mode dual numbers with a short vector of tangent directions.
//
// The adjoint build of a model library (-DTCLB_AD) compiles the SAME generated node code with real_t = Dual: one pass
// over a node then yields the derivatives of everything the node writes (fields and objective contribution) with
// respect to TCLB_AD_DIRS of the values it read.  The adjoint kernel runs the node code once per chunk of inputs and
// scatters  sum_i d(out_i)/d(in_k) * adj(out_i)  to where in_k was read from: a discrete adjoint of the stage, exact to
// rounding, for ANY model the generator can lower.  The reference gets its adjoint from Tapenade source transformation
// (tools/makeAD, src/ADTools.cu); nothing of that tool chain exists here.
#pragma once
#include <cmath>
#include <type_traits>

#ifndef TCLB_AD_DIRS
#define TCLB_AD_DIRS 4
#endif
#ifndef TCLB_HD
#ifdef __CUDACC__
#define TCLB_HD __host__ __device__ inline
#else
#define TCLB_HD inline
#endif
#endif

template <class T, int N>
struct Dual {
    T v;
    T d[N];
    TCLB_HD Dual() {}
    TCLB_HD Dual(T x) : v(x) { for (int i = 0; i < N; i++) d[i] = 0; }
    TCLB_HD Dual(int x) : v((T)x) { for (int i = 0; i < N; i++) d[i] = 0; }
    TCLB_HD Dual(unsigned x) : v((T)x) { for (int i = 0; i < N; i++) d[i] = 0; }
    TCLB_HD Dual(long long x) : v((T)x) { for (int i = 0; i < N; i++) d[i] = 0; }
    template <class U, class = typename std::enable_if<std::is_floating_point<U>::value && !std::is_same<U, T>::value>::type>
    TCLB_HD Dual(U x) : v((T)x) { for (int i = 0; i < N; i++) d[i] = 0; }
    TCLB_HD explicit operator T() const { return v; }
    TCLB_HD explicit operator int() const { return (int)v; }
    TCLB_HD explicit operator bool() const { return v != 0; }
    TCLB_HD Dual &operator+=(const Dual &o) { v += o.v; for (int i = 0; i < N; i++) d[i] += o.d[i]; return *this; }
    TCLB_HD Dual &operator-=(const Dual &o) { v -= o.v; for (int i = 0; i < N; i++) d[i] -= o.d[i]; return *this; }
    TCLB_HD Dual &operator*=(const Dual &o) { for (int i = 0; i < N; i++) d[i] = d[i] * o.v + v * o.d[i]; v *= o.v; return *this; }
    TCLB_HD Dual &operator/=(const Dual &o) {
        const T r = (T)1 / o.v; const T q = v * r;
        for (int i = 0; i < N; i++) d[i] = (d[i] - q * o.d[i]) * r;
        v = q; return *this;
    }
    TCLB_HD Dual operator-() const { Dual r; r.v = -v; for (int i = 0; i < N; i++) r.d[i] = -d[i]; return r; }
    TCLB_HD Dual operator+() const { return *this; }
};

#define TCLB_DUAL_T template <class T, int N> TCLB_HD
#define TCLB_DUAL_BIN(op, opeq)                                                                             \
    TCLB_DUAL_T Dual<T, N> operator op(const Dual<T, N> &a, const Dual<T, N> &b) { Dual<T, N> r = a; r opeq b; return r; } \
    template <class T, int N, class U, class = typename std::enable_if<std::is_arithmetic<U>::value>::type>  \
    TCLB_HD Dual<T, N> operator op(const Dual<T, N> &a, U b) { Dual<T, N> r = a; r opeq Dual<T, N>((T)b); return r; } \
    template <class T, int N, class U, class = typename std::enable_if<std::is_arithmetic<U>::value>::type>  \
    TCLB_HD Dual<T, N> operator op(U a, const Dual<T, N> &b) { Dual<T, N> r((T)a); r opeq b; return r; }
TCLB_DUAL_BIN(+, +=)
TCLB_DUAL_BIN(-, -=)
TCLB_DUAL_BIN(*, *=)
TCLB_DUAL_BIN(/, /=)
#undef TCLB_DUAL_BIN

#define TCLB_DUAL_CMP(op)                                                                                   \
    TCLB_DUAL_T bool operator op(const Dual<T, N> &a, const Dual<T, N> &b) { return a.v op b.v; }            \
    template <class T, int N, class U, class = typename std::enable_if<std::is_arithmetic<U>::value>::type>  \
    TCLB_HD bool operator op(const Dual<T, N> &a, U b) { return a.v op (T)b; }                                \
    template <class T, int N, class U, class = typename std::enable_if<std::is_arithmetic<U>::value>::type>  \
    TCLB_HD bool operator op(U a, const Dual<T, N> &b) { return (T)a op b.v; }
TCLB_DUAL_CMP(<)
TCLB_DUAL_CMP(>)
TCLB_DUAL_CMP(<=)
TCLB_DUAL_CMP(>=)
TCLB_DUAL_CMP(==)
TCLB_DUAL_CMP(!=)
#undef TCLB_DUAL_CMP

// f(a) with derivative df: chain rule on every direction
TCLB_DUAL_T Dual<T, N> tclb_dual_chain(const Dual<T, N> &a, T f, T df) {
    Dual<T, N> r; r.v = f;
    for (int i = 0; i < N; i++) r.d[i] = df * a.d[i];
    return r;
}
TCLB_DUAL_T Dual<T, N> sqrt(const Dual<T, N> &a) { const T s = std::sqrt(a.v); return tclb_dual_chain(a, s, s > 0 ? (T)0.5 / s : (T)0); }
TCLB_DUAL_T Dual<T, N> exp(const Dual<T, N> &a) { const T e = std::exp(a.v); return tclb_dual_chain(a, e, e); }
TCLB_DUAL_T Dual<T, N> log(const Dual<T, N> &a) { return tclb_dual_chain(a, (T)std::log(a.v), (T)1 / a.v); }
TCLB_DUAL_T Dual<T, N> sin(const Dual<T, N> &a) { return tclb_dual_chain(a, (T)std::sin(a.v), (T)std::cos(a.v)); }
TCLB_DUAL_T Dual<T, N> cos(const Dual<T, N> &a) { return tclb_dual_chain(a, (T)std::cos(a.v), (T)-std::sin(a.v)); }
TCLB_DUAL_T Dual<T, N> tan(const Dual<T, N> &a) { const T t = std::tan(a.v); return tclb_dual_chain(a, t, (T)1 + t * t); }
TCLB_DUAL_T Dual<T, N> tanh(const Dual<T, N> &a) { const T t = std::tanh(a.v); return tclb_dual_chain(a, t, (T)1 - t * t); }
TCLB_DUAL_T Dual<T, N> atan(const Dual<T, N> &a) { return tclb_dual_chain(a, (T)std::atan(a.v), (T)1 / ((T)1 + a.v * a.v)); }
TCLB_DUAL_T Dual<T, N> sinh(const Dual<T, N> &a) { return tclb_dual_chain(a, (T)std::sinh(a.v), (T)std::cosh(a.v)); }
TCLB_DUAL_T Dual<T, N> cosh(const Dual<T, N> &a) { return tclb_dual_chain(a, (T)std::cosh(a.v), (T)std::sinh(a.v)); }
TCLB_DUAL_T Dual<T, N> asin(const Dual<T, N> &a) { return tclb_dual_chain(a, (T)std::asin(a.v), (T)1 / std::sqrt((T)1 - a.v * a.v)); }
TCLB_DUAL_T Dual<T, N> acos(const Dual<T, N> &a) { return tclb_dual_chain(a, (T)std::acos(a.v), (T)-1 / std::sqrt((T)1 - a.v * a.v)); }
TCLB_DUAL_T Dual<T, N> cbrt(const Dual<T, N> &a) { const T c = std::cbrt(a.v); return tclb_dual_chain(a, c, c != 0 ? (T)1 / ((T)3 * c * c) : (T)0); }
TCLB_DUAL_T Dual<T, N> log10(const Dual<T, N> &a) { return tclb_dual_chain(a, (T)std::log10(a.v), (T)0.4342944819032518 / a.v); }
TCLB_DUAL_T Dual<T, N> log2(const Dual<T, N> &a) { return tclb_dual_chain(a, (T)std::log2(a.v), (T)1.4426950408889634 / a.v); }
TCLB_DUAL_T Dual<T, N> erf(const Dual<T, N> &a) { return tclb_dual_chain(a, (T)std::erf(a.v), (T)1.1283791670955126 * (T)std::exp(-a.v * a.v)); }
TCLB_DUAL_T Dual<T, N> atan2(const Dual<T, N> &y, const Dual<T, N> &x) {
    Dual<T, N> r; r.v = std::atan2(y.v, x.v);
    const T q = x.v * x.v + y.v * y.v, iq = q > 0 ? (T)1 / q : (T)0;
    for (int i = 0; i < N; i++) r.d[i] = (x.v * y.d[i] - y.v * x.d[i]) * iq;
    return r;
}
template <class T, int N, class U, class = typename std::enable_if<std::is_arithmetic<U>::value>::type>
TCLB_HD Dual<T, N> atan2(const Dual<T, N> &y, U x) { return atan2(y, Dual<T, N>((T)x)); }
template <class T, int N, class U, class = typename std::enable_if<std::is_arithmetic<U>::value>::type>
TCLB_HD Dual<T, N> atan2(U y, const Dual<T, N> &x) { return atan2(Dual<T, N>((T)y), x); }
TCLB_DUAL_T Dual<T, N> fabs(const Dual<T, N> &a) { return a.v < 0 ? -a : a; }
TCLB_DUAL_T Dual<T, N> abs(const Dual<T, N> &a) { return a.v < 0 ? -a : a; }
TCLB_DUAL_T Dual<T, N> floor(const Dual<T, N> &a) { return Dual<T, N>((T)std::floor(a.v)); }
TCLB_DUAL_T Dual<T, N> ceil(const Dual<T, N> &a) { return Dual<T, N>((T)std::ceil(a.v)); }
TCLB_DUAL_T Dual<T, N> pow(const Dual<T, N> &a, const Dual<T, N> &b) {
    const T p = std::pow(a.v, b.v);
    Dual<T, N> r; r.v = p;
    const T da = b.v != 0 ? b.v * std::pow(a.v, b.v - (T)1) : (T)0;
    bool bvar = false;
    for (int i = 0; i < N; i++) bvar = bvar || b.d[i] != 0;
    const T db = bvar && a.v > 0 ? p * std::log(a.v) : (T)0;
    for (int i = 0; i < N; i++) r.d[i] = da * a.d[i] + db * b.d[i];
    return r;
}
template <class T, int N, class U, class = typename std::enable_if<std::is_arithmetic<U>::value>::type>
TCLB_HD Dual<T, N> pow(const Dual<T, N> &a, U b) { return tclb_dual_chain(a, (T)std::pow(a.v, (T)b), b != 0 ? (T)b * (T)std::pow(a.v, (T)b - (T)1) : (T)0); }
template <class T, int N, class U, class = typename std::enable_if<std::is_arithmetic<U>::value>::type>
TCLB_HD Dual<T, N> pow(U a, const Dual<T, N> &b) { const T p = std::pow((T)a, b.v); return tclb_dual_chain(b, p, a > 0 ? p * (T)std::log((T)a) : (T)0); }
TCLB_DUAL_T Dual<T, N> fmax(const Dual<T, N> &a, const Dual<T, N> &b) { return a.v >= b.v ? a : b; }
TCLB_DUAL_T Dual<T, N> fmin(const Dual<T, N> &a, const Dual<T, N> &b) { return a.v <= b.v ? a : b; }
TCLB_DUAL_T Dual<T, N> max(const Dual<T, N> &a, const Dual<T, N> &b) { return a.v >= b.v ? a : b; }
TCLB_DUAL_T Dual<T, N> min(const Dual<T, N> &a, const Dual<T, N> &b) { return a.v <= b.v ? a : b; }
template <class T, int N, class U, class = typename std::enable_if<std::is_arithmetic<U>::value>::type>
TCLB_HD Dual<T, N> fmax(const Dual<T, N> &a, U b) { return a.v >= (T)b ? a : Dual<T, N>((T)b); }
template <class T, int N, class U, class = typename std::enable_if<std::is_arithmetic<U>::value>::type>
TCLB_HD Dual<T, N> fmin(const Dual<T, N> &a, U b) { return a.v <= (T)b ? a : Dual<T, N>((T)b); }
template <class T, int N, class U, class = typename std::enable_if<std::is_arithmetic<U>::value>::type>
TCLB_HD Dual<T, N> max(const Dual<T, N> &a, U b) { return a.v >= (T)b ? a : Dual<T, N>((T)b); }
template <class T, int N, class U, class = typename std::enable_if<std::is_arithmetic<U>::value>::type>
TCLB_HD Dual<T, N> min(const Dual<T, N> &a, U b) { return a.v <= (T)b ? a : Dual<T, N>((T)b); }
TCLB_DUAL_T bool isnan(const Dual<T, N> &a) { return a.v != a.v; }
TCLB_DUAL_T bool isfinite(const Dual<T, N> &a) { return (a.v - a.v) == 0; }
#undef TCLB_DUAL_T

// value part of a scalar, whatever the build (the plain overloads live in tclb_node_rt.h)
template <class T, int N> TCLB_HD T tclb_value(const Dual<T, N> &x) { return x.v; }

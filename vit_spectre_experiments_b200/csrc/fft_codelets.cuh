// Small-N FFT codelets for the spectral token mixer (sm_100a).
//
// Every transform is fully unrolled at compile time with twiddles folded to
// literal constants (constexpr Taylor trig in double, rounded once to fp32).
// The codelets are generic over a "lane vector" type V:
//   V = float   one channel per thread (also what the host self-test compiles)
//   V = float2  two adjacent channels per thread, arithmetic issued as
//               Blackwell packed FP32 (FADD2 / FMUL2 / FFMA2)
// so one thread owns whole length-N transforms of 1 or 2 channels and no
// cross-lane traffic is needed.
//
// Conventions (un-normalised; the kernels fold the reference's norm="ortho"
// 1/(H*W) -- spectre_vit/models/spectre/spectre.py:136 -- into the filter):
//   CFFT<N,D>  : X[k] = sum_n x[n] exp(D * 2*pi*i*k*n/N),  D=-1 forward, +1 inverse
//   RFFT<N>    : real x[N] -> X[0..N/2], forward sign, Im X[0] (= Im X[N/2]) set to 0
//   IRFFT<N>   : X[0..N/2] -> real x[N] = sum over the Hermitian extension, and
//                -- like torch.fft.irfft / pocketfft / cuFFT C2R -- IGNORES
//                Im X[0] and (N even) Im X[N/2]   
// N even: radix-2 split; N odd: direct DFT using the x[j] +- x[N-j] symmetry.
#pragma once

#include <type_traits>

#if defined(__CUDACC__)
#include <cuda_runtime.h>
#define SB_HD __host__ __device__ __forceinline__
#else
#include <cmath>
#define SB_HD inline
#endif

// 1: issue pair arithmetic as packed FADD2/FMUL2/FFMA2; 0: two scalar ops.
// Measured on B200 (profiles/r01_pipes.txt): scalar FFMA sustains 4.0 warp-instr/clk/SM,
// FFMA2 1.58 (= 3.2 scalar-equivalent), FADD2 1.6, FMUL2 1.97 -- packed math halves
// issue slots but not FP-pipe time, so it only pays where issue is the limiter.
#ifndef SB200_PACKED_MATH
#define SB200_PACKED_MATH 0
#endif

namespace sb200 {

// ----------------------------------------------------------------- constexpr trig
constexpr double kPi = 3.14159265358979323846264338327950288;

constexpr double cx_sin_small(double a) {  // |a| <= pi/4
  double a2 = a * a, term = a, sum = a;
  for (int i = 1; i < 14; ++i) {
    term *= -a2 / ((2.0 * i) * (2.0 * i + 1.0));
    sum += term;
  }
  return sum;
}
constexpr double cx_cos_small(double a) {
  double a2 = a * a, term = 1.0, sum = 1.0;
  for (int i = 1; i < 14; ++i) {
    term *= -a2 / ((2.0 * i - 1.0) * (2.0 * i));
    sum += term;
  }
  return sum;
}
// cos / sin of 2*pi*k/n with exact quadrant reduction in integers.
constexpr double cos2pi(int k, int n) {
  k = ((k % n) + n) % n;
  int q = (8 * k + n) / (2 * n);  // nearest quadrant
  double a = (kPi / 2.0) * static_cast<double>(4 * k - q * n) / static_cast<double>(n);
  q &= 3;
  return q == 0 ? cx_cos_small(a) : q == 1 ? -cx_sin_small(a) : q == 2 ? -cx_cos_small(a) : cx_sin_small(a);
}
constexpr double sin2pi(int k, int n) {
  k = ((k % n) + n) % n;
  int q = (8 * k + n) / (2 * n);
  double a = (kPi / 2.0) * static_cast<double>(4 * k - q * n) / static_cast<double>(n);
  q &= 3;
  return q == 0 ? cx_sin_small(a) : q == 1 ? cx_cos_small(a) : q == 2 ? -cx_sin_small(a) : -cx_cos_small(a);
}

// ----------------------------------------------------------------- compile-time loop
template <int I, int N, class F>
SB_HD void static_for(F&& f) {
  if constexpr (I < N) {
    f(std::integral_constant<int, I>{});
    static_for<I + 1, N>(f);
  }
}

// ----------------------------------------------------------------- lane-vector ops
template <class V>
struct VOps;

template <>
struct VOps<float> {
  static SB_HD float zero() { return 0.f; }
  static SB_HD float add(float a, float b) { return a + b; }
  static SB_HD float sub(float a, float b) { return a - b; }
  static SB_HD float neg(float a) { return -a; }
  static SB_HD float muls(float a, float s) { return a * s; }
  static SB_HD float fmas(float a, float s, float c) { return fmaf(a, s, c); }  // a*s + c
  static SB_HD float mul(float a, float b) { return a * b; }
  static SB_HD float fma(float a, float b, float c) { return fmaf(a, b, c); }
  static SB_HD float fnma(float a, float b, float c) { return fmaf(-a, b, c); }  // c - a*b
};

#if defined(__CUDACC__)
template <>
struct VOps<float2> {
  static SB_HD float2 zero() { return make_float2(0.f, 0.f); }
#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ >= 1000) && SB200_PACKED_MATH
  static SB_HD float2 add(float2 a, float2 b) { return __fadd2_rn(a, b); }
  static SB_HD float2 sub(float2 a, float2 b) { return __ffma2_rn(b, make_float2(-1.f, -1.f), a); }
  static SB_HD float2 muls(float2 a, float s) { return __fmul2_rn(a, make_float2(s, s)); }
  static SB_HD float2 fmas(float2 a, float s, float2 c) { return __ffma2_rn(a, make_float2(s, s), c); }
  static SB_HD float2 mul(float2 a, float2 b) { return __fmul2_rn(a, b); }
  static SB_HD float2 fma(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }
  static SB_HD float2 fnma(float2 a, float2 b, float2 c) { return __ffma2_rn(make_float2(-a.x, -a.y), b, c); }
#else
  static SB_HD float2 add(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
  static SB_HD float2 sub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
  static SB_HD float2 muls(float2 a, float s) { return make_float2(a.x * s, a.y * s); }
  static SB_HD float2 fmas(float2 a, float s, float2 c) { return make_float2(fmaf(a.x, s, c.x), fmaf(a.y, s, c.y)); }
  static SB_HD float2 mul(float2 a, float2 b) { return make_float2(a.x * b.x, a.y * b.y); }
  static SB_HD float2 fma(float2 a, float2 b, float2 c) { return make_float2(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y)); }
  static SB_HD float2 fnma(float2 a, float2 b, float2 c) { return make_float2(fmaf(-a.x, b.x, c.x), fmaf(-a.y, b.y, c.y)); }
#endif
  static SB_HD float2 neg(float2 a) { return make_float2(-a.x, -a.y); }
};
#endif

template <class V>
struct Cx {
  V re, im;
};

template <class V>
SB_HD Cx<V> cadd(Cx<V> a, Cx<V> b) {
  return {VOps<V>::add(a.re, b.re), VOps<V>::add(a.im, b.im)};
}
template <class V>
SB_HD Cx<V> csub(Cx<V> a, Cx<V> b) {
  return {VOps<V>::sub(a.re, b.re), VOps<V>::sub(a.im, b.im)};
}
// a * b (complex, both lane vectors)
template <class V>
SB_HD Cx<V> cmul(Cx<V> a, Cx<V> b) {
  using O = VOps<V>;
  return {O::fnma(a.im, b.im, O::mul(a.re, b.re)), O::fma(a.im, b.re, O::mul(a.re, b.im))};
}
// a * conj(b)
template <class V>
SB_HD Cx<V> cmul_conj(Cx<V> a, Cx<V> b) {
  using O = VOps<V>;
  return {O::fma(a.im, b.im, O::mul(a.re, b.re)), O::fnma(a.re, b.im, O::mul(a.im, b.re))};
}
// acc += conj(a) * b
template <class V>
SB_HD void cmac_conj(Cx<V>& acc, Cx<V> a, Cx<V> b) {
  using O = VOps<V>;
  acc.re = O::fma(a.im, b.im, O::fma(a.re, b.re, acc.re));
  acc.im = O::fnma(a.im, b.re, O::fma(a.re, b.im, acc.im));
}

// ----------------------------------------------------------------- complex FFT
template <int N, int D, class V>
struct CFFT {
  static_assert(D == 1 || D == -1, "direction");
  static SB_HD void run(Cx<V>* x) {
    using O = VOps<V>;
    if constexpr (N == 1) {
      return;
    } else if constexpr (N == 2) {
      Cx<V> a = x[0], b = x[1];
      x[0] = cadd(a, b);
      x[1] = csub(a, b);
    } else if constexpr (N % 4 == 2) {
      // Good-Thomas prime-factor split N = 2 * M (M odd): no twiddles.
      constexpr int M = N / 2;
      Cx<V> s[M], d[M];
      static_for<0, M>([&](auto nc) {
        constexpr int n2 = decltype(nc)::value;
        s[n2] = cadd(x[(2 * n2) % N], x[(M + 2 * n2) % N]);
        d[n2] = csub(x[(2 * n2) % N], x[(M + 2 * n2) % N]);
      });
      CFFT<M, D, V>::run(s);
      CFFT<M, D, V>::run(d);
      static_for<0, N>([&](auto kc) {
        constexpr int k = decltype(kc)::value;
        if constexpr (k % 2 == 0) x[k] = s[k % M];
        else x[k] = d[k % M];
      });
    } else if constexpr (N % 2 == 0) {
      constexpr int M = N / 2;
      Cx<V> e[M], o[M];
      static_for<0, M>([&](auto kc) {
        constexpr int k = decltype(kc)::value;
        e[k] = x[2 * k];
        o[k] = x[2 * k + 1];
      });
      CFFT<M, D, V>::run(e);
      CFFT<M, D, V>::run(o);
      static_for<0, M>([&](auto kc) {
        constexpr int k = decltype(kc)::value;
        if constexpr (k == 0) {
          x[0] = cadd(e[0], o[0]);
          x[M] = csub(e[0], o[0]);
        } else if constexpr (4 * k == N) {
          // w = D*i :  t = (-D*o.im, D*o.re)
          if constexpr (D == 1) {
            x[k] = {O::sub(e[k].re, o[k].im), O::add(e[k].im, o[k].re)};
            x[k + M] = {O::add(e[k].re, o[k].im), O::sub(e[k].im, o[k].re)};
          } else {
            x[k] = {O::add(e[k].re, o[k].im), O::sub(e[k].im, o[k].re)};
            x[k + M] = {O::sub(e[k].re, o[k].im), O::add(e[k].im, o[k].re)};
          }
        } else {
          constexpr float c = static_cast<float>(cos2pi(k, N));
          constexpr float s = static_cast<float>(D * sin2pi(k, N));
          // t = o * (c + i s)
          V tr = O::fmas(o[k].im, -s, O::muls(o[k].re, c));
          V ti = O::fmas(o[k].im, c, O::muls(o[k].re, s));
          x[k] = {O::add(e[k].re, tr), O::add(e[k].im, ti)};
          x[k + M] = {O::sub(e[k].re, tr), O::sub(e[k].im, ti)};
        }
      });
    } else {
      constexpr int M = (N - 1) / 2;
      Cx<V> a[M + 1], b[M + 1];
      const Cx<V> x0 = x[0];
      Cx<V> dc = x0;
      static_for<1, M + 1>([&](auto jc) {
        constexpr int j = decltype(jc)::value;
        a[j] = cadd(x[j], x[N - j]);
        b[j] = csub(x[j], x[N - j]);
        dc = cadd(dc, a[j]);
      });
      x[0] = dc;
      static_for<1, M + 1>([&](auto kc) {
        constexpr int k = decltype(kc)::value;
        Cx<V> p = x0, q;
        static_for<1, M + 1>([&](auto jc) {
          constexpr int j = decltype(jc)::value;
          constexpr float cj = static_cast<float>(cos2pi(j * k, N));
          constexpr float sj = static_cast<float>(sin2pi(j * k, N));
          p.re = O::fmas(a[j].re, cj, p.re);
          p.im = O::fmas(a[j].im, cj, p.im);
          if constexpr (j == 1) {
            q.re = O::muls(b[j].re, sj);
            q.im = O::muls(b[j].im, sj);
          } else {
            q.re = O::fmas(b[j].re, sj, q.re);
            q.im = O::fmas(b[j].im, sj, q.im);
          }
        });
        // X[k] = p + i*D*q ; X[N-k] = p - i*D*q
        if constexpr (D == 1) {
          x[k] = {O::sub(p.re, q.im), O::add(p.im, q.re)};
          x[N - k] = {O::add(p.re, q.im), O::sub(p.im, q.re)};
        } else {
          x[k] = {O::add(p.re, q.im), O::sub(p.im, q.re)};
          x[N - k] = {O::sub(p.re, q.im), O::add(p.im, q.re)};
        }
      });
    }
  }
};

// ----------------------------------------------------------------- real -> half-complex
template <int N, class V>
struct RFFT {
  static SB_HD void run(const V* x, Cx<V>* X) {
    using O = VOps<V>;
    if constexpr (N == 1) {
      X[0] = {x[0], O::zero()};
    } else if constexpr (N == 2) {
      X[0] = {O::add(x[0], x[1]), O::zero()};
      X[1] = {O::sub(x[0], x[1]), O::zero()};
    } else if constexpr (N % 4 == 2) {
      // prime-factor split N = 2 * M (M odd): two real M-point DFTs, no twiddles
      constexpr int M = N / 2, Mh = M / 2;
      V s[M], d[M];
      static_for<0, M>([&](auto nc) {
        constexpr int n2 = decltype(nc)::value;
        s[n2] = O::add(x[(2 * n2) % N], x[(M + 2 * n2) % N]);
        d[n2] = O::sub(x[(2 * n2) % N], x[(M + 2 * n2) % N]);
      });
      Cx<V> S[Mh + 1], Dd[Mh + 1];
      RFFT<M, V>::run(s, S);
      RFFT<M, V>::run(d, Dd);
      static_for<0, M + 1>([&](auto kc) {
        constexpr int k = decltype(kc)::value;
        constexpr int k2 = k % M;
        if constexpr (k % 2 == 0) {
          if constexpr (k2 <= Mh) X[k] = S[k2];
          else X[k] = {S[M - k2].re, O::neg(S[M - k2].im)};
        } else {
          if constexpr (k2 <= Mh) X[k] = Dd[k2];
          else X[k] = {Dd[M - k2].re, O::neg(Dd[M - k2].im)};
        }
      });
    } else if constexpr (N % 2 == 0) {
      constexpr int M = N / 2, Mh = M / 2;
      V xe[M], xo[M];
      static_for<0, M>([&](auto kc) {
        constexpr int k = decltype(kc)::value;
        xe[k] = x[2 * k];
        xo[k] = x[2 * k + 1];
      });
      Cx<V> Ev[Mh + 1], Ov[Mh + 1];
      RFFT<M, V>::run(xe, Ev);
      RFFT<M, V>::run(xo, Ov);
      X[0] = {O::add(Ev[0].re, Ov[0].re), O::zero()};
      X[M] = {O::sub(Ev[0].re, Ov[0].re), O::zero()};
      static_for<1, (M - 1) / 2 + 1>([&](auto kc) {
        constexpr int k = decltype(kc)::value;
        constexpr float c = static_cast<float>(cos2pi(k, N));
        constexpr float s = static_cast<float>(sin2pi(k, N));
        // t = O[k] * (c - i s)
        V tr = O::fmas(Ov[k].im, s, O::muls(Ov[k].re, c));
        V ti = O::fmas(Ov[k].re, -s, O::muls(Ov[k].im, c));
        X[k] = {O::add(Ev[k].re, tr), O::add(Ev[k].im, ti)};
        X[M - k] = {O::sub(Ev[k].re, tr), O::sub(ti, Ev[k].im)};
      });
      if constexpr (M % 2 == 0) {
        X[Mh] = {Ev[Mh].re, O::neg(Ov[Mh].re)};
      }
    } else {
      constexpr int M = (N - 1) / 2;
      V a[M + 1], b[M + 1];
      V dc = x[0];
      static_for<1, M + 1>([&](auto jc) {
        constexpr int j = decltype(jc)::value;
        a[j] = O::add(x[j], x[N - j]);
        b[j] = O::sub(x[j], x[N - j]);
        dc = O::add(dc, a[j]);
      });
      X[0] = {dc, O::zero()};
      static_for<1, M + 1>([&](auto kc) {
        constexpr int k = decltype(kc)::value;
        V p = x[0], q;
        static_for<1, M + 1>([&](auto jc) {
          constexpr int j = decltype(jc)::value;
          constexpr float cj = static_cast<float>(cos2pi(j * k, N));
          constexpr float nsj = static_cast<float>(-sin2pi(j * k, N));
          p = O::fmas(a[j], cj, p);
          if constexpr (j == 1) q = O::muls(b[j], nsj);
          else q = O::fmas(b[j], nsj, q);
        });
        X[k] = {p, q};
      });
    }
  }
};

// ----------------------------------------------------------------- half-complex -> real
template <int N, class V>
struct IRFFT {
  static SB_HD void run(const Cx<V>* X, V* x) {
    using O = VOps<V>;
    if constexpr (N == 1) {
      x[0] = X[0].re;
    } else if constexpr (N == 2) {
      x[0] = O::add(X[0].re, X[1].re);
      x[1] = O::sub(X[0].re, X[1].re);
    } else if constexpr (N % 4 == 2) {
      // prime-factor split (transpose of the RFFT one)
      constexpr int M = N / 2, Mh = M / 2;
      Cx<V> S[Mh + 1], Dd[Mh + 1];
      static_for<0, Mh + 1>([&](auto kc) {
        constexpr int k2 = decltype(kc)::value;
        // even-k bin with k % M == k2, and odd-k bin with k % M == k2 (k in [0, N))
        constexpr int ke = (k2 % 2 == 0) ? k2 : k2 + M;
        constexpr int ko = (k2 % 2 == 1) ? k2 : k2 + M;
        if constexpr (ke <= M) S[k2] = X[ke];
        else S[k2] = {X[N - ke].re, O::neg(X[N - ke].im)};
        if constexpr (ko <= M) Dd[k2] = X[ko];
        else Dd[k2] = {X[N - ko].re, O::neg(X[N - ko].im)};
      });
      V s[M], d[M];
      IRFFT<M, V>::run(S, s);
      IRFFT<M, V>::run(Dd, d);
      static_for<0, M>([&](auto nc) {
        constexpr int n2 = decltype(nc)::value;
        x[(2 * n2) % N] = O::add(s[n2], d[n2]);
        x[(M + 2 * n2) % N] = O::sub(s[n2], d[n2]);
      });
    } else if constexpr (N % 2 == 0) {
      constexpr int M = N / 2, Mh = M / 2;
      Cx<V> Ev[Mh + 1], Ov[Mh + 1];
      Ev[0] = {O::add(X[0].re, X[M].re), O::zero()};
      Ov[0] = {O::sub(X[0].re, X[M].re), O::zero()};
      static_for<1, (M - 1) / 2 + 1>([&](auto kc) {
        constexpr int k = decltype(kc)::value;
        constexpr float c = static_cast<float>(cos2pi(k, N));
        constexpr float s = static_cast<float>(sin2pi(k, N));
        // sum = X[k] + conj(X[M-k]) ; d = X[k] - conj(X[M-k]) ; O[k] = d * (c + i s)
        Ev[k] = {O::add(X[k].re, X[M - k].re), O::sub(X[k].im, X[M - k].im)};
        V dr = O::sub(X[k].re, X[M - k].re), di = O::add(X[k].im, X[M - k].im);
        Ov[k] = {O::fmas(di, -s, O::muls(dr, c)), O::fmas(di, c, O::muls(dr, s))};
      });
      if constexpr (M % 2 == 0) {
        Ev[Mh] = {O::muls(X[Mh].re, 2.f), O::zero()};
        Ov[Mh] = {O::muls(X[Mh].im, -2.f), O::zero()};
      }
      V xe[M], xo[M];
      IRFFT<M, V>::run(Ev, xe);
      IRFFT<M, V>::run(Ov, xo);
      static_for<0, M>([&](auto kc) {
        constexpr int k = decltype(kc)::value;
        x[2 * k] = xe[k];
        x[2 * k + 1] = xo[k];
      });
    } else {
      constexpr int M = (N - 1) / 2;
      V dc = X[0].re;
      static_for<1, M + 1>([&](auto kc) {
        constexpr int k = decltype(kc)::value;
        dc = O::fmas(X[k].re, 2.f, dc);
      });
      x[0] = dc;
      static_for<1, M + 1>([&](auto nc) {
        constexpr int n = decltype(nc)::value;
        V p = X[0].re, q;
        static_for<1, M + 1>([&](auto kc) {
          constexpr int k = decltype(kc)::value;
          constexpr float ck = static_cast<float>(2.0 * cos2pi(k * n, N));
          constexpr float sk = static_cast<float>(2.0 * sin2pi(k * n, N));
          p = O::fmas(X[k].re, ck, p);
          if constexpr (k == 1) q = O::muls(X[k].im, sk);
          else q = O::fmas(X[k].im, sk, q);
        });
        x[n] = O::sub(p, q);
        x[N - n] = O::add(p, q);
      });
    }
  }
};

}  // namespace sb200

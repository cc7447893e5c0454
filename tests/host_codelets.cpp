// Host build of the device FFT codelets (V = float) so the arithmetic can be
// checked against numpy on a CPU-only machine.  Built by tests/test_codelets_host.py.
#include "fft_codelets.cuh"
using namespace sb200;

template <int N>
static void cfft_t(int dir, float* re, float* im) {
  Cx<float> x[N];
  for (int i = 0; i < N; ++i) x[i] = {re[i], im[i]};
  if (dir < 0) CFFT<N, -1, float>::run(x); else CFFT<N, 1, float>::run(x);
  for (int i = 0; i < N; ++i) { re[i] = x[i].re; im[i] = x[i].im; }
}
template <int N>
static void rfft_t(const float* x, float* re, float* im) {
  Cx<float> X[N / 2 + 1];
  RFFT<N, float>::run(x, X);
  for (int i = 0; i <= N / 2; ++i) { re[i] = X[i].re; im[i] = X[i].im; }
}
template <int N>
static void irfft_t(const float* re, const float* im, float* x) {
  Cx<float> X[N / 2 + 1];
  for (int i = 0; i <= N / 2; ++i) X[i] = {re[i], im[i]};
  IRFFT<N, float>::run(X, x);
}

#define SIZES(F) F(2) F(3) F(4) F(5) F(6) F(7) F(8) F(9) F(10) F(12) F(14) F(16)
extern "C" {
int codelet_cfft(int n, int dir, float* re, float* im) {
  switch (n) {
#define F(N) case N: cfft_t<N>(dir, re, im); return 0;
    SIZES(F)
#undef F
  }
  return -1;
}
int codelet_rfft(int n, const float* x, float* re, float* im) {
  switch (n) {
#define F(N) case N: rfft_t<N>(x, re, im); return 0;
    SIZES(F)
#undef F
  }
  return -1;
}
int codelet_irfft(int n, const float* re, const float* im, float* x) {
  switch (n) {
#define F(N) case N: irfft_t<N>(re, im, x); return 0;
    SIZES(F)
#undef F
  }
  return -1;
}
}

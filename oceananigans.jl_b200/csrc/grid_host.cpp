// grid_host.cpp — host-side set-up arithmetic that must round exactly like the reference:
//   regular spacings Δ = FT(BigFloat(L)/N)                    src/Grids/grid_generation.jl:98-135
//   stretched z spacings incl. halos                          src/Grids/grid_generation.jl:34-95
//   reconstruction coefficients (exact rational → FT, last = 1 - Σ)   src/Advection/reconstruction_coefficients.jl:100-115
//   Poisson eigenvalues (Float64)                             src/Solvers/poisson_eigenvalues.jl:8-31
// Compiled by g++ (uses __float128 / exact integer rationals); no CUDA here.
#include "grid_host.h"
#include <cmath>
#include <cstdint>
#include <numeric>
#include <stdexcept>

namespace nhhost {

// ---- exact rationals (small) ------------------------------------------------------------------------------
struct Frac {
  __int128 n, d;
  Frac(__int128 nn = 0, __int128 dd = 1) : n(nn), d(dd) { norm(); }
  static __int128 gcd(__int128 a, __int128 b) { if (a < 0) a = -a; if (b < 0) b = -b; while (b) { __int128 t = a % b; a = b; b = t; } return a ? a : 1; }
  void norm() { if (d < 0) { n = -n; d = -d; } __int128 g = gcd(n, d); n /= g; d /= g; }
  Frac operator+(const Frac& o) const { return Frac(n * o.d + o.n * d, d * o.d); }
  Frac operator*(const Frac& o) const { return Frac(n * o.n, d * o.d); }
  Frac operator/(const Frac& o) const { return Frac(n * o.d, d * o.n); }
};

// correctly rounded p/q for small integers: IEEE division of two exactly representable values
template <typename FT> static FT round_frac(const Frac& f) { return (FT)((FT)(double)f.n / (FT)(double)f.d); }

// stencil_coefficients(i=50, r, xr=1:100, xi=1:100; order) on a uniform grid — reconstruction_coefficients.jl:100-115
template <typename FT>
static void stencil_coefficients(int r, int order, FT* out) {
  Frac exact[8];
  for (int j = 0; j < order; ++j) {
    Frac acc(0);
    for (int m = j + 1; m <= order; ++m) {
      Frac num(0);
      for (int l = 0; l <= order; ++l) {
        if (l == m) continue;
        Frac p(1);
        for (int q = 0; q <= order; ++q)
          if (q != m && q != l) p = p * Frac(r - q + 1);
        num = num + p;
      }
      Frac den(1);
      for (int l = 0; l <= order; ++l)
        if (l != m) den = den * Frac(m - l);
      acc = acc + num / den;
    }
    exact[j] = acc;
  }
  FT s = 0;
  for (int j = 0; j < order - 1; ++j) {
    out[j] = round_frac<FT>(exact[j]);
    s = j == 0 ? out[0] : (FT)(s + out[j]);
  }
  out[order - 1] = (FT)(FT(1) - s);
}

template <typename FT>
void make_scheme_tables(Tables<FT>& T) {
  for (auto& row : T.sym) for (auto& v : row) v = 0;
  stencil_coefficients<FT>(0, 2, T.sym[1]);          // Centered(2):  r = B-1 = 0, order 2B = 2
  stencil_coefficients<FT>(1, 4, T.sym[2]);          // Centered(4)
  for (int B = 1; B <= 3; ++B) {
    for (int q = 0; q < 5; ++q) T.upl[B][q] = T.upr[B][q] = 0;
    if (B == 1) { T.upl[1][0] = T.upr[1][0] = FT(1); continue; }
    stencil_coefficients<FT>(B - 2, 2 * B - 1, T.upl[B]);
    stencil_coefficients<FT>(B - 1, 2 * B - 1, T.upr[B]);
  }
  for (int B = 2; B <= 3; ++B)
    for (int r = 0; r < B; ++r) stencil_coefficients<FT>(r, B, T.wc[B][r]);
  T.cstar[2][0] = round_frac<FT>(Frac(2, 3)); T.cstar[2][1] = round_frac<FT>(Frac(1, 3)); T.cstar[2][2] = 0;
  T.cstar[3][0] = round_frac<FT>(Frac(3, 10)); T.cstar[3][1] = round_frac<FT>(Frac(3, 5)); T.cstar[3][2] = round_frac<FT>(Frac(1, 10));
  T.eps = (FT)1e-8f;                                   // ε = 1f-8 (weno_interpolants.jl:70)
}
template void make_scheme_tables<float>(Tables<float>&);
template void make_scheme_tables<double>(Tables<double>&);

// Δ = FT(BigFloat(c2 - c1)/N) with c1, c2 already converted to FT; L = FT(c2 - c1)
template <typename FT>
void regular_spacing(double c1d, double c2d, int N, FT& delta, FT& L) {
  FT c1 = (FT)c1d, c2 = (FT)c2d;
  if (!(c1 < c2)) throw std::invalid_argument("coordinate interval must be increasing");
  __float128 Lq = (__float128)c2 - (__float128)c1;
  L = (FT)Lq;
  delta = (FT)(Lq / (__float128)N);
}
template void regular_spacing<float>(double, double, int, float&, float&);
template void regular_spacing<double>(double, double, int, double&, double&);

// stretched coordinate: faces (N+1) → Δᶜ[k] for k = 1-H … , Δᶠ[k] for k = -H …  (grid_generation.jl:34-95)
// dzc gets TF-1 entries with first index 1-H; dzf gets TC+1 entries with first index -H.
template <typename FT>
void stretched_spacings(const double* faces, int N, int H, bool bounded, std::vector<FT>& dzc, std::vector<FT>& dzf, FT& L) {
  std::vector<FT> Fi(N + 1);
  for (int i = 0; i <= N; ++i) Fi[i] = (FT)faces[i];
  for (int i = 0; i < N; ++i)
    if (!(Fi[i + 1] >= Fi[i])) throw std::invalid_argument("The values of z must increase as the index increases!");
  L = (FT)(Fi[N] - Fi[0]);
  std::vector<FT> dlo(H), dhi(H);
  if (bounded) {
    for (int i = 0; i < H; ++i) { dlo[i] = (FT)(Fi[1] - Fi[0]); dhi[i] = (FT)(Fi[N] - Fi[N - 1]); }
  } else {
    for (int i = 1; i <= H; ++i) dlo[i - 1] = (FT)(Fi[N - H + i] - Fi[N - H + i - 1]);
    for (int i = 1; i <= H; ++i) dhi[H - i] = (FT)(Fi[i] - Fi[i - 1]);
  }
  auto seqsum = [](const std::vector<FT>& v, int from) { FT s = 0; for (size_t q = from; q < v.size(); ++q) s = (FT)(s + v[q]); return s; };
  std::vector<FT> F;
  for (int i = 0; i < H; ++i) F.push_back((FT)(Fi[0] - seqsum(dlo, i)));
  for (int i = 0; i <= N; ++i) F.push_back(Fi[i]);
  {
    std::vector<FT> hi(H);
    for (int i = 0; i < H; ++i) hi[i] = (FT)(Fi[N] + seqsum(dhi, i));
    for (int i = H - 1; i >= 0; --i) F.push_back(hi[i]);
  }
  int TC = N + 2 * H;
  std::vector<FT> C(TC);
  for (int i = 0; i < TC; ++i) C[i] = (FT)((FT)(F[i + 1] + F[i]) / FT(2));
  std::vector<FT> d_raw(TC > 0 ? TC - 1 : 0);
  for (int i = 1; i < TC; ++i) d_raw[i - 1] = (FT)(C[i] - C[i - 1]);
  int TF = N + 2 * H + (bounded ? 1 : 0);
  dzc.resize(TF - 1);
  for (int i = 0; i < TF - 1; ++i) dzc[i] = (FT)(F[i + 1] - F[i]);
  dzf.clear();
  FT first = d_raw.empty() ? (FT)(F[1] - F[0]) : d_raw[0];
  dzf.push_back(first); dzf.push_back(first);
  for (auto v : d_raw) dzf.push_back(v);
}
template void stretched_spacings<float>(const double*, int, int, bool, std::vector<float>&, std::vector<float>&, float&);
template void stretched_spacings<double>(const double*, int, int, bool, std::vector<double>&, std::vector<double>&, double&);

// poisson_eigenvalues — Float64 arithmetic on (i-1)·π; h = L/N evaluated in FT by the caller
std::vector<double> poisson_eigenvalues(int N, double h, int topo) {
  std::vector<double> lam(N, 0.0);
  if (topo == 2) return lam;
  const double pi = 3.141592653589793;
  for (int i = 0; i < N; ++i) {
    double s = topo == 0 ? std::sin(i * pi / N) : std::sin(i * pi / (2 * N));
    double x = 2 * s / h;
    lam[i] = x * x;
  }
  return lam;
}

// exp(-2πi m/N), m < N, and exp(-iπk/(2N)), k < N: evaluated in long double with octant reduction
void twiddles(int N, std::vector<double>& tw, std::vector<double>& tw4) {
  tw.resize(2 * (size_t)N); tw4.resize(2 * (size_t)N);
  const long double pi = 3.14159265358979323846264338327950288L;
  for (int m = 0; m < N; ++m) {
    // angle = 2π m/N = (π/4)·(8m/N): reduce to an octant for accuracy
    long long num = 8LL * m; long long oct = num / N; long long rem = num % N;
    long double a = (pi / 4) * ((long double)rem / N);
    long double c = cosl(a), s = sinl(a), cc, ss;
    const long double r2 = 0.70710678118654752440084436210484904L;
    switch (oct) {
      case 0: cc = c; ss = s; break;
      case 1: cc = r2 * (c - s); ss = r2 * (c + s); break;
      case 2: cc = -s; ss = c; break;
      case 3: cc = -r2 * (c + s); ss = r2 * (c - s); break;
      case 4: cc = -c; ss = -s; break;
      case 5: cc = -r2 * (c - s); ss = -r2 * (c + s); break;
      case 6: cc = s; ss = -c; break;
      default: cc = r2 * (c + s); ss = -r2 * (c - s); break;
    }
    if (rem == 0) {   // exact multiples of π/4
      static const long double C8[8] = {1, r2, 0, -r2, -1, -r2, 0, r2}, S8[8] = {0, r2, 1, r2, 0, -r2, -1, -r2};
      cc = C8[oct]; ss = S8[oct];
    }
    tw[2 * m] = (double)cc; tw[2 * m + 1] = (double)(-ss);
    long double b = pi * ((long double)m / (2.0L * N));
    tw4[2 * m] = (double)cosl(b); tw4[2 * m + 1] = (double)(-sinl(b));
  }
}

// ---- correctly rounded cube root ------------------------------------------------------------------------------------
// x = m·2^(E-P) with m a P-bit integer; E = 3q + rem; g = m·2^(rem-P) ∈ [1/2, 4).  r = ⌊cbrt(g)·2^S⌋ with S = P when
// g ≥ 1 (cbrt(g) ∈ [1, 1.59)) and P+1 otherwise (cbrt(g) ∈ [0.79, 1)) has one bit more than the result's mantissa, so
// the rounded mantissa is (r+1)>>1 (no ties: an odd (P+1)-bit integer cubed is never m·2^s).  r is found by bisection on
// r³ ≤ X = m·2^(rem-P+3S) with 192-bit integers.
namespace {
struct U192 { unsigned long long hi, mid, lo; };
inline bool le(const U192& a, const U192& b) {
  if (a.hi != b.hi) return a.hi < b.hi;
  if (a.mid != b.mid) return a.mid < b.mid;
  return a.lo <= b.lo;
}
inline U192 cube(unsigned long long r) {          // r < 2^55
  typedef unsigned __int128 u128;
  u128 r2 = (u128)r * r;
  unsigned long long a = (unsigned long long)(r2 >> 64), b = (unsigned long long)r2;
  u128 br = (u128)b * r, ar = (u128)a * r + (br >> 64);
  return U192{(unsigned long long)(ar >> 64), (unsigned long long)ar, (unsigned long long)br};
}
inline U192 shifted(unsigned long long m, int sh) {   // m·2^sh, m < 2^53, sh < 128
  typedef unsigned __int128 u128;
  if (sh < 64) { u128 v = (u128)m << sh; return U192{0, (unsigned long long)(v >> 64), (unsigned long long)v}; }
  u128 v = (u128)m << (sh - 64);
  return U192{(unsigned long long)(v >> 64), (unsigned long long)v, 0};
}
}  // namespace

template <typename FT>
FT cbrt_correctly_rounded(FT x) {
  if (!(x > FT(0)) || !std::isfinite((double)x)) return std::cbrt(x);
  const int P = sizeof(FT) == 8 ? 53 : 24;
  int E;
  double f = std::frexp((double)x, &E);                       // x = f·2^E, f ∈ [1/2, 1)
  unsigned long long m = (unsigned long long)std::ldexp(f, P);   // exact: x has P significant bits
  int q = E >= 0 ? E / 3 : -((-E + 2) / 3);                  // floor division
  int rem = E - 3 * q;
  const int S = ((m << rem) >= (1ull << P)) ? P : P + 1;
  const U192 X = shifted(m, rem - P + 3 * S);
  unsigned long long lo = 0, hi = 1ull << 55;                 // invariant: lo³ ≤ X < hi³
  while (hi - lo > 1) {
    unsigned long long mid = lo + (hi - lo) / 2;
    if (le(cube(mid), X)) lo = mid; else hi = mid;
  }
  unsigned long long y = (lo + 1) >> 1;
  return (FT)std::ldexp((double)y, q - (S - 1));
}
template float cbrt_correctly_rounded<float>(float);
template double cbrt_correctly_rounded<double>(double);

}  // namespace nhhost

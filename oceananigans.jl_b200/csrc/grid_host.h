// grid_host.h — host set-up arithmetic (see grid_host.cpp)
#pragma once
#include <vector>
namespace nhhost {
template <typename FT>
struct Tables {           // must mirror SchemeTables<FT> in common.cuh field for field
  FT sym[3][4];
  FT upl[4][5];
  FT upr[4][5];
  FT wc[4][3][3];
  FT cstar[4][3];
  FT eps;
};
template <typename FT> void make_scheme_tables(Tables<FT>& T);
template <typename FT> void regular_spacing(double c1, double c2, int N, FT& delta, FT& L);
template <typename FT> void stretched_spacings(const double* faces, int N, int H, bool bounded, std::vector<FT>& dzc,
                                               std::vector<FT>& dzf, FT& L);
// correctly rounded cube root of a positive finite number (exact integer arithmetic; implementation-independent, so the
// NumPy oracle's cbrt_correctly_rounded gives the same bits).  The reference's Base.cbrt is accurate to < 1 ulp.
template <typename FT> FT cbrt_correctly_rounded(FT x);
std::vector<double> poisson_eigenvalues(int N, double h, int topo);
void twiddles(int N, std::vector<double>& tw, std::vector<double>& tw4);
}

"""Host-side check of tiled_adv_flux (csrc/tendency_generic.cuh, TG_FASTFLUX) against the generic flux functions:
   python tools/check_fast_flux_host.py
Extracts the device functions from the CUDA headers, compiles them as plain C++ (g++, -ffp-contract=off) with the WENO
reconstruction replaced by an argument-sensitive stand-in, and compares the two flux evaluations bit for bit at every flux
point 0 … N+1 of a (12, 10, 9) grid, triply periodic and with walls in x and z, for every (field, direction) pair.
Development aid for the round-2 item "enable TG_FASTFLUX" (DESIGN.md §7b); no GPU needed."""
import os
import subprocess
import sys
import tempfile

R = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oceananigans.jl_b200", "csrc")


def host(s):
    return (s.replace("__host__ __device__ __forceinline__", "inline").replace("__device__ __forceinline__", "inline"))


def main():
    rec, gen, com = (open(os.path.join(R, f)).read() for f in ("recon.cuh", "tendency_generic.cuh", "common.cuh"))
    i = com.index("template <typename FT>\nstruct GridDev")
    j = com.index("template <typename FT>\nstruct TendParams")
    common = host(com[i:com.index("};", j) + 2])
    r1 = host(rec[rec.index("// A getter is a functor"):rec.index("// ---- IEEE-exact Float64 division")])
    c = rec.index("// _biased_interpolate:")
    r2 = host(rec[c:rec.index("\n}\n", c) + 3])
    stubs = """
template <typename FT> inline FT weno5_from_q(const SchemeTables<FT>&, FT q0, FT q1, FT q2, FT q3, FT q4) { return q0 + FT(2)*q1 + FT(3)*q2*q2 + FT(5)*q3 + FT(7)*q4*q0; }
template <typename FT> inline FT weno3_from_q(const SchemeTables<FT>&, FT q0, FT q1, FT q2) { return q0 + FT(2)*q1 + FT(3)*q2*q2; }
"""
    g1 = host(gen[gen.index("template <typename FT>\nstruct FieldGet"):gen.index("// ---- Coriolis")])
    h = gen.index("template <typename FT, int a, int b>\n__device__ __forceinline__ FT tiled_adv_flux")
    g2 = host(gen[h:gen.index("template <typename FT, int a>    // a = 0,1,2")]).replace("TG_FASTFLUX &&", "true &&")
    hdr = ("#include <cstdio>\n#include <cstdlib>\n#include <cstring>\n#include <vector>\n#include <cmath>\n#define NH_MAXTR 8\n"
           "enum {NH_PERIODIC=0,NH_BOUNDED=1,NH_FLAT=2};\nenum {NH_ADV_NONE=0,NH_ADV_CENTERED=1,NH_ADV_UPWIND=2,NH_ADV_WENO=3};\n"
           + common + r1 + stubs + r2 + g1 + g2)
    main_cpp = r"""
#include "hdr.h"
typedef double FT;
template <int a, int b> int check(const TendParams<FT>& P, const FT* c, const char* tag) {
  const GridDev<FT>& g = P.g; int bad = 0, tot = 0;
  for (int k = 0; k <= g.N[2] + 1; ++k) for (int j = 0; j <= g.N[1] + 1; ++j) for (int i = 0; i <= g.N[0] + 1; ++i) {
    constexpr int av = a == 3 ? 0 : a;
    FT ref = a == 3 ? tracer_flux<FT, b>(P, c, i, j, k) : momentum_flux<FT, av, b>(P, i, j, k);
    FT got = tiled_adv_flux<FT, a, b>(P, c, i, j, k);
    ++tot;
    if (memcmp(&ref, &got, sizeof(FT)) != 0) ++bad;
  }
  printf("%s field %d direction %d: %d of %d flux points differ\n", tag, a, b, bad, tot);
  return bad;
}
int main() {
  int bad = 0;
  for (int walls = 0; walls < 2; ++walls) {
    TendParams<FT> P; memset(&P, 0, sizeof(P));
    GridDev<FT>& g = P.g;
    int N[3] = {12, 10, 9};
    for (int e = 0; e < 3; ++e) { g.N[e] = N[e]; g.H[e] = 3; g.topo[e] = NH_PERIODIC; }
    if (walls) { g.topo[2] = NH_BOUNDED; g.topo[0] = NH_BOUNDED; }
    g.ox = 16; g.oy = 3; g.oz = 3; g.px = 48; g.py = N[1] + 6; g.pz = N[2] + 6; g.sy = g.px; g.sz = (long long)g.px * g.py;
    g.dx = 0.1; g.dy = 0.13;
    std::vector<FT> dzc(N[2] + 8), dzf(N[2] + 8);
    for (size_t q = 0; q < dzc.size(); ++q) { dzc[q] = 0.07 + 0.003 * q; dzf[q] = 0.071 + 0.0029 * q; }
    g.dzc = dzc.data(); g.dzf = dzf.data(); g.zoff = 3;
    size_t vol = (size_t)g.px * g.py * g.pz;
    std::vector<FT> U[3], c(vol);
    srand(1);
    for (int e = 0; e < 3; ++e) { U[e].resize(vol); for (auto& x : U[e]) x = rand() / (double)RAND_MAX - 0.5; P.U[e] = U[e].data(); }
    for (auto& x : c) x = rand() / (double)RAND_MAX;
    for (int e = 0; e < 3; ++e) { P.adv[e].kind = NH_ADV_WENO; P.adv[e].B = 3; }
    FT* t = &P.tab.sym[0][0];
    for (size_t q = 0; q < sizeof(P.tab) / sizeof(FT); ++q) t[q] = 0.1 + 0.01 * q;
    const char* tag = walls ? "(B,P,B)" : "(P,P,P)";
    bad += check<0,0>(P,c.data(),tag) + check<0,1>(P,c.data(),tag) + check<0,2>(P,c.data(),tag);
    bad += check<1,0>(P,c.data(),tag) + check<1,1>(P,c.data(),tag) + check<1,2>(P,c.data(),tag);
    bad += check<2,0>(P,c.data(),tag) + check<2,1>(P,c.data(),tag) + check<2,2>(P,c.data(),tag);
    bad += check<3,0>(P,c.data(),tag) + check<3,1>(P,c.data(),tag) + check<3,2>(P,c.data(),tag);
  }
  return bad != 0;
}
"""
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, "hdr.h"), "w").write(hdr)
        open(os.path.join(d, "t.cpp"), "w").write(main_cpp)
        subprocess.check_call(["g++", "-O0", "-std=c++17", "-ffp-contract=off", "t.cpp", "-o", "t"], cwd=d)
        return subprocess.call([os.path.join(d, "t")])


if __name__ == "__main__":
    sys.exit(main())

// Drop-in proof for the black-box interface bbfmm::kernel (bbfmm.h:451-592).  This file uses nothing but the reference's own
// API and is compiled TWICE by oracle/Makefile: against /root/reference/bbfmm/bbfmm.h (CPU, target bbfmm_ref) and against
// include/compat/bbfmm/bbfmm.h + libaniso_b200.so (B200, target bbfmm_dropin).  Both binaries read points and charges from
// argv[1] ("n np rank maxLevel" then n lines "x y q") and print the potentials, one per line, for a kernel that is NOT one
// of the aniso families: K(a, b) = log|a - b| + 0.25 (a.x - b.x)  (unsymmetric), first through runCache, then - after the
// re-initialisation the reference's callers do - through runFast with the charges doubled.
#include <cstdio>

#include "bbfmm.h"

using namespace bbfmm;

int main(int argc, char** argv) {
    if (argc < 2) return 2;
    FILE* f = fopen(argv[1], "r");
    if (!f) return 3;
    int n, np, rank, maxLevel;
    if (fscanf(f, "%d %d %d %d", &n, &np, &rank, &maxLevel) != 4) return 4;
    vector<point> pts((size_t)n);
    Vector charge(n);
    for (int i = 0; i < n; ++i) {
        double x, y, q;
        if (fscanf(f, "%lf %lf %lf", &x, &y, &q) != 3) return 5;
        pts[i].x = x;
        pts[i].y = y;
        charge(i) = q;
    }
    fclose(f);
    kernel k;
    k.eval = [](point& a, point& b) {
        const double dx = a.x - b.x, dy = a.y - b.y, r2 = dx * dx + dy * dy;
        return (r2 == 0.0 ? 0.0 : 0.5 * log(r2)) + 0.25 * dx;
    };
    k.initialize(np, pts, pts, charge, n, n, rank, maxLevel);
    Vector pot;
    k.runCache(pot);
    for (int i = 0; i < n; ++i) printf("%.17g\n", pot(i));
    for (int i = 0; i < n; ++i) charge(i) *= 2.0;
    k.initialize(np, pts, pts, charge, n, n, rank, maxLevel);
    Vector pot2;
    k.runFast(pot2);
    for (int i = 0; i < n; ++i) printf("%.17g\n", pot2(i));
    return 0;
}

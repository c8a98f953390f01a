// Host-side property check of the rank-major multipole layout of the sharded upward pass (DevTree::pack_*, mc_node in
// aniso_b200/csrc/device.cuh).  Compiled with nvcc, runs on the CPU (tests/test_host_cpu.py).  For every rank count P, first
// M2L level L0 and depth:
//   1. (level, index) -> position is a bijection onto [0, P * pack_tot)
//   2. the nodes rank q owns (the q-th of P equal Morton ranges of every level) fill exactly [q * pack_tot, (q+1) * pack_tot)
//   3. inside a rank's block the levels follow each other, Morton order preserved (so the 4^d descendants of a sub-tree root
//      at one level stay consecutive: what subtree_m2m and up_mid_kernel rely on)
//   4. without packing (pack_logP < 0) the position is the level-major node number
#include <cstdio>
#include <vector>

#include "device.cuh"

using namespace an;

int main() {
    int checked = 0;
    for (int logP = 1; logP <= 3; ++logP)
        for (int L0 = 2; L0 <= 3; ++L0)
            for (int deepest = L0; deepest <= 8; ++deepest) {
                const int P = 1 << logP;
                if ((1 << (2 * L0)) % P) continue;
                DevTree T;
                T.nlevels = deepest + 1;
                T.pack_logP = logP;
                T.pack_L0 = L0;
                T.pack_tot = ((((1 << (2 * T.nlevels)) - 1) / 3) - (((1 << (2 * L0)) - 1) / 3)) >> logP;
                std::vector<int> owner((size_t)P * T.pack_tot, -1);
                for (int L = L0; L <= deepest; ++L) {
                    const int n = 1 << (2 * L), per = n / P;
                    int prev = -1, prev_rank = -1;
                    for (int i = 0; i < n; ++i) {
                        const int pos = mc_node(T, L, i), q = i / per;
                        if (pos < 0 || pos >= P * T.pack_tot) { printf("FAIL range P %d L0 %d deepest %d L %d i %d pos %d\n", P, L0, deepest, L, i, pos); return 1; }
                        if (owner[pos] != -1) { printf("FAIL collision P %d L %d i %d pos %d\n", P, L, i, pos); return 1; }
                        owner[pos] = q;
                        if (pos / T.pack_tot != q) { printf("FAIL block P %d L %d i %d pos %d rank %d\n", P, L, i, pos, q); return 1; }
                        if (q == prev_rank && pos != prev + 1) { printf("FAIL order P %d L %d i %d\n", P, L, i); return 1; }
                        prev = pos;
                        prev_rank = q;
                    }
                }
                for (size_t k = 0; k < owner.size(); ++k)
                    if (owner[k] < 0) { printf("FAIL hole P %d L0 %d deepest %d at %zu\n", P, L0, deepest, k); return 1; }
                // levels follow each other inside a block: the first node of level L + 1 of rank 0 comes right after its level L
                for (int L = L0; L < deepest; ++L)
                    if (mc_node(T, L + 1, 0) != mc_node(T, L, (1 << (2 * L)) / P - 1) + 1) { printf("FAIL level order P %d L %d\n", P, L); return 1; }
                ++checked;
            }
    DevTree U;  // unpacked: level-major numbering
    for (int L = 0; L <= 8; ++L)
        for (int i = 0; i < (1 << (2 * L)); i += 37)
            if (mc_node(U, L, i) != ((1 << (2 * L)) - 1) / 3 + i) { printf("FAIL unpacked L %d i %d\n", L, i); return 1; }
    printf("mc_node ok: %d layouts\n", checked);
    return 0;
}

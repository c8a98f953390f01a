// Host-side check of the symmetric block storage the plan builder hands to the streaming kernels (BlockCSR::symmetric,
// symmetrize() in aniso_b200/csrc/plan.cpp).  The reference stores and applies every (target, source) block of the V / X and
// U / W lists separately (bbfmm.h:863-946); the product keeps one block per pair and sends the transposed product to a scratch
// row of the partner.  Compiled with g++ together with plan.cpp and run on the CPU by tests/test_host_cpu.py.  Per operator:
//   1. every stored entry is a list entry of its group's node; nothing is stored twice
//   2. every list entry (t <- s) is covered exactly once: stored itself, or as the mirror of the stored entry (s <- t)
//   3. an entry has a mirror row iff the partner entry exists among the groups of this plan (sharded plans hold only the
//      owned leaves and their ancestors); the row lies in the partner's scratch range behind its own first row, and no
//      scratch row is written by two entries; scratch_rows = groups + mirrors
//   4. block offsets are the running sum of R * R; nnz = stored blocks * R * R
//   5. without symmetric storage (ragged P2P tiles, generic plans) the entries are the lists themselves
#include <cstdio>
#include <cstdlib>
#include <map>
#include <set>
#include <utility>
#include <vector>

#include "plan.hpp"

using namespace an;

static int fails = 0;
#define EXPECT(c, ...)                       \
    do {                                     \
        if (!(c)) {                          \
            if (fails++ < 20) {              \
                printf("FAIL %s: ", #c);     \
                printf(__VA_ARGS__);         \
                printf("\n");                \
            }                                \
        }                                    \
    } while (0)

static std::set<int> list_of(const HostPlan& p, int id, bool m2l) {
    const TreeNode& n = p.nodes[id];
    std::set<int> s;
    for (const std::vector<int>* L : {m2l ? &n.V : &n.U, m2l ? &n.X : &n.W})
        for (int e : *L)
            if (!p.nodes[e].empty) s.insert(e);
    return s;
}

static void check(const HostPlan& p, const BlockCSR& b, bool m2l, const char* what) {
    const int groups = (int)b.nrow.size(), ne = (int)b.col_off.size();
    std::map<int, int> group_of;
    for (int g = 0; g < groups; ++g) {
        EXPECT(!group_of.count(b.group_node[g]), "%s node %d has two groups", what, b.group_node[g]);
        group_of[b.group_node[g]] = g;
    }
    std::set<std::pair<int, int>> stored, covered;
    std::set<int> scratch_used;
    long mirrors = 0;
    int64_t off = 0;
    for (int g = 0; g < groups; ++g) {
        const int t = b.group_node[g];
        const std::set<int> L = list_of(p, t, m2l);
        for (int e = b.ent_ptr[g]; e < b.ent_ptr[g + 1]; ++e) {
            const int s = b.src_node[e];
            EXPECT(b.tgt_node[e] == t, "%s entry %d target %d in group of %d", what, e, b.tgt_node[e], t);
            EXPECT(L.count(s), "%s stored (%d <- %d) is not a list entry", what, t, s);
            EXPECT(stored.insert({t, s}).second, "%s (%d <- %d) stored twice", what, t, s);
            EXPECT(covered.insert({t, s}).second, "%s (%d <- %d) covered twice", what, t, s);
            EXPECT(b.blk_off[e] == off, "%s entry %d block offset", what, e);
            off += (int64_t)b.nrow[g] * b.ncol[e];
            if (!b.symmetric) continue;
            auto gs = group_of.find(s);
            const bool partner = s != t && gs != group_of.end() && list_of(p, s, m2l).count(t);
            EXPECT((b.mirror_row[e] >= 0) == partner, "%s (%d <- %d) mirror row %d, partner %d", what, t, s, b.mirror_row[e], (int)partner);
            if (b.mirror_row[e] >= 0 && partner) {
                ++mirrors;
                EXPECT(b.mirror_row[e] > b.srow_ptr[gs->second] && b.mirror_row[e] < b.srow_ptr[gs->second + 1],
                       "%s (%d <- %d) mirror row %d outside (%d, %d)", what, t, s, b.mirror_row[e], b.srow_ptr[gs->second], b.srow_ptr[gs->second + 1]);
                EXPECT(scratch_used.insert(b.mirror_row[e]).second, "%s scratch row %d written twice", what, b.mirror_row[e]);
                EXPECT(covered.insert({s, t}).second, "%s (%d <- %d) covered twice (as a mirror)", what, s, t);
            }
        }
    }
    EXPECT(b.nnz == off, "%s nnz", what);
    long expected = 0;
    for (int g = 0; g < groups; ++g) {
        const int t = b.group_node[g];
        for (int s : list_of(p, t, m2l)) {
            ++expected;
            EXPECT(covered.count({t, s}), "%s list entry (%d <- %d) is not covered", what, t, s);
        }
    }
    EXPECT((long)covered.size() == expected, "%s covered %zu expected %ld", what, covered.size(), expected);
    if (b.symmetric) {
        EXPECT(b.scratch_rows == groups + mirrors, "%s scratch rows %d groups %d mirrors %ld", what, b.scratch_rows, groups, mirrors);
        EXPECT((long)scratch_used.size() == mirrors, "%s scratch rows used", what);
        EXPECT(ne + mirrors == expected, "%s stored %d + mirrors %ld != list entries %ld", what, ne, mirrors, expected);
        for (int g = 0; g < groups; ++g) EXPECT(b.srow_ptr[g + 1] > b.srow_ptr[g], "%s group %d has no own scratch row", what, g);
    } else
        EXPECT(ne == expected, "%s stored %d list entries %ld", what, ne, expected);
}

int main() {
    struct Cfg {
        int sz, deg, np, ml;
    };
    // complete trees (16, 32: every P2P tile uniform -> both operators symmetric), a grid that is no power of two and a capped
    // depth (ragged leaves: P2P stays unsymmetric, M2L symmetric), a higher degree, and np 8
    const Cfg cfgs[] = {{16, 1, 4, 20}, {32, 1, 4, 20}, {24, 1, 4, 20}, {20, 3, 4, 20}, {64, 1, 4, 2}, {32, 1, 8, 20}, {12, 2, 4, 20}};
    int plans = 0, symmetric_ops = 0;
    for (const Cfg& c : cfgs) {
        HostPlan p;
        build_plan(p, c.sz, c.deg, 2, 0.5, 10, c.np, c.ml);
        check(p, p.m2l, true, "m2l");
        check(p, p.p2p, false, "p2p");
        symmetric_ops += p.m2l.symmetric + p.p2p.symmetric;
        // every non-empty node with a non-empty V / X list has a group, every non-empty leaf has a P2P group
        size_t want = 0;
        for (size_t id = 1; id < p.nodes.size(); ++id) want += !p.nodes[id].empty && !list_of(p, (int)id, true).empty();
        EXPECT(p.m2l.nrow.size() == want, "m2l groups %zu want %zu", p.m2l.nrow.size(), want);
        EXPECT(p.p2p.nrow.size() == p.leaves.size(), "p2p groups");
        ++plans;
        for (int nranks : {2, 3}) {  // sharded plans: the owned leaves and their ancestors only
            std::vector<int> lb, le;
            partition_leaves(p, nranks, lb, le);
            for (int r = 0; r < nranks; ++r) {
                flatten_lists(p, lb[r], le[r]);
                check(p, p.m2l, true, "m2l (sharded)");
                check(p, p.p2p, false, "p2p (sharded)");
                EXPECT((int)p.p2p.nrow.size() == le[r] - lb[r], "sharded p2p groups");
                ++plans;
            }
        }
    }
    EXPECT(symmetric_ops >= 8, "only %d operators used symmetric storage", symmetric_ops);
    if (fails) {
        printf("%d failures\n", fails);
        return 1;
    }
    printf("plan symmetric storage ok: %d plans, %d symmetric operators\n", plans, symmetric_ops);
    return 0;
}

// Host-only check of rpt_b200/csrc/bvh_sah.hpp (built by tests/test_host_logic.py with nvcc, runs without a GPU):
// every primitive sits in exactly one leaf, every child box contains its primitives' boxes (padded), leaves hold
// <= leaf_max items, and the depth stays at or below 36 (two such levels + a sentinel fit kStackSize = 80).
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

#include "../../rpt_b200/csrc/bvh_sah.hpp"

using rptx::BvhNode;
using rptx::PrimBox;

static int fail(const char* what, int a = 0, int b = 0) { std::printf("FAIL %s %d %d\n", what, a, b); return 1; }

static int check(const std::vector<PrimBox>& boxes, float pad, int leaf_max, const char* name) {
    std::vector<BvhNode> nodes;
    std::vector<uint32_t> perm;
    int root = 0;
    rptx_host::build_sah_host(boxes, pad, leaf_max, nodes, perm, &root);
    const uint32_t n = (uint32_t)boxes.size();
    if (perm.size() != n) return fail("perm size");
    std::vector<int> seen(n, 0), covered(n, 0);
    for (uint32_t p : perm) { if (p >= n) return fail("perm range"); seen[p]++; }
    for (uint32_t i = 0; i < n; i++) if (seen[i] != 1) return fail("perm not a permutation", (int)i, seen[i]);
    if (nodes.size() + 1 > n && n > (uint32_t)leaf_max) return fail("too many nodes", (int)nodes.size(), (int)n);
    struct Item { int ref; int depth; float mn[3], mx[3]; bool boxed; };
    std::vector<Item> stack;
    Item r{};
    r.ref = root; r.depth = 0; r.boxed = false;
    if (n) stack.push_back(r);
    int max_depth = 0;
    while (!stack.empty()) {
        const Item it = stack.back();
        stack.pop_back();
        if (it.depth > max_depth) max_depth = it.depth;
        if (it.ref < 0) {
            const uint32_t code = (uint32_t)(~it.ref), first = code >> 3, count = (code & 7u) + 1u;
            if ((int)count > leaf_max) return fail("leaf too large", (int)count, leaf_max);
            if (first + count > n) return fail("leaf range");
            for (uint32_t i = first; i < first + count; i++) {
                covered[i]++;
                const PrimBox& b = boxes[perm[i]];
                if (it.boxed)
                    for (int k = 0; k < 3; k++)
                        if (!(it.mn[k] <= b.mn[k] - pad * 0.999f && it.mx[k] >= b.mx[k] + pad * 0.999f)) return fail("leaf box does not contain a primitive", (int)i, k);
            }
            continue;
        }
        if ((size_t)it.ref >= nodes.size()) return fail("node index");
        const BvhNode& nd = nodes[it.ref];
        if (it.boxed)
            for (int k = 0; k < 3; k++)
                if (!(it.mn[k] <= std::fmin(nd.lmin[k], nd.rmin[k]) + 2 * pad + 1e-30f && it.mx[k] >= std::fmax(nd.lmax[k], nd.rmax[k]) - 2 * pad - 1e-30f))
                    return fail("child boxes leave the parent box", it.ref, k);
        Item l{}, rr{};
        l.ref = nd.left; rr.ref = nd.right; l.depth = rr.depth = it.depth + 1; l.boxed = rr.boxed = true;
        for (int k = 0; k < 3; k++) { l.mn[k] = nd.lmin[k]; l.mx[k] = nd.lmax[k]; rr.mn[k] = nd.rmin[k]; rr.mx[k] = nd.rmax[k]; }
        stack.push_back(l); stack.push_back(rr);
    }
    for (uint32_t i = 0; i < n; i++) if (covered[i] != 1) return fail("position not in exactly one leaf", (int)i, covered[i]);
    if (max_depth > 36) return fail("too deep for the traversal stack", max_depth);
    std::printf("ok %s n=%u nodes=%zu depth=%d\n", name, n, nodes.size(), max_depth);
    return 0;
}

int main() {
    std::mt19937 rng(7);
    std::uniform_real_distribution<float> u(0.f, 1.f);
    int rc = 0;
    for (uint32_t n : {0u, 1u, 3u, 4u, 5u, 9u, 100u, 6320u}) {
        std::vector<PrimBox> b(n);
        for (auto& x : b) for (int k = 0; k < 3; k++) { const float c = u(rng) * 10.f - 5.f, e = u(rng) * 0.2f; x.mn[k] = c - e; x.mx[k] = c + e; }
        rc |= check(b, 1e-6f, 4, "uniform");
        rc |= check(b, 1e-6f, 1, "uniform leaf1");
    }
    {   // all centroids identical, and a geometric progression along one axis (the SAH wants 1 | n-1 splits)
        std::vector<PrimBox> same(1000), prog(2000);
        for (auto& x : same) for (int k = 0; k < 3; k++) { x.mn[k] = -1.f; x.mx[k] = 1.f; }
        float pos = 1e-20f;
        for (auto& x : prog) { for (int k = 0; k < 3; k++) { x.mn[k] = 0.f; x.mx[k] = 1e-3f; } x.mn[0] = pos; x.mx[0] = pos * 1.5f; pos *= 1.05f; }
        rc |= check(same, 0.f, 4, "coincident");
        rc |= check(prog, 0.f, 1, "progression");
    }
    return rc;
}

// Binned-SAH BVH2 over f32 boxes for SMALL inputs (a teapot's 6 K triangles, a scene's ~1 K instances), in the
// node format of bvh_build.cuh.  A top-down SAH build of a few thousand primitives is ~2 ms of sequential work:
// it runs on the host at commit time, the result is uploaded like any other scene table.  Larger inputs (and the
// photon maps, 10^7-10^8 items) keep the GPU LBVH of bvh_build.cu.  Against that LBVH the SAH tree saves node
// visits of k_extend on the mesh scenes (profiles/r1_v2_summary.md).
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <numeric>
#include <vector>

#include "bvh_build.cuh"
#include "common.cuh"

namespace rptx_host {

constexpr uint32_t kSahMaxPrims = 1u << 16;
constexpr int kSahBins = 16;
// SAH splits down to this depth, halving below: depth <= 20 + 16 per tree, <= 74 pending entries for TLAS + sentinel + BLAS
// (trace.cuh kStackSize = 80).  Real meshes stay far below (teapot: 13).
constexpr uint32_t kSahFreeDepth = 20;

struct SahBox {
    float mn[3], mx[3];
    void reset() { for (int k = 0; k < 3; k++) { mn[k] = INFINITY; mx[k] = -INFINITY; } }
    void grow(const rptx::PrimBox& b) { for (int k = 0; k < 3; k++) { mn[k] = std::fmin(mn[k], b.mn[k]); mx[k] = std::fmax(mx[k], b.mx[k]); } }
    void grow(const SahBox& b) { for (int k = 0; k < 3; k++) { mn[k] = std::fmin(mn[k], b.mn[k]); mx[k] = std::fmax(mx[k], b.mx[k]); } }
    double half_area() const {
        const double dx = (double)mx[0] - mn[0], dy = (double)mx[1] - mn[1], dz = (double)mx[2] - mn[2];
        return (dx < 0 || dy < 0 || dz < 0) ? 0.0 : dx * dy + dy * dz + dz * dx;
    }
};

inline float sah_down(double x) { float f = (float)x; if ((double)f > x) f = std::nextafterf(f, -INFINITY); return f; }
inline float sah_up(double x) { float f = (float)x; if ((double)f < x) f = std::nextafterf(f, INFINITY); return f; }

// nodes: at most n - 1 entries, node 0 is the root; perm[i] = input index of the primitive at leaf-order position i.
// Leaves hold <= leaf_max (<= 8) consecutive positions.  Child boxes are the exact unions padded outward by `pad`.
inline void build_sah_host(const std::vector<rptx::PrimBox>& boxes, float pad, int leaf_max, std::vector<rptx::BvhNode>& nodes,
                           std::vector<uint32_t>& perm, int* root_ref) {
    using rptx::BvhNode;
    const uint32_t n = (uint32_t)boxes.size();
    perm.resize(n);
    std::iota(perm.begin(), perm.end(), 0u);
    nodes.clear();
    if (n <= (uint32_t)leaf_max) {
        *root_ref = n ? rptx::make_leaf_ref(0u, n) : 0;
        return;
    }
    std::vector<float> cen(3 * (size_t)n);
    for (uint32_t i = 0; i < n; i++)
        for (int k = 0; k < 3; k++) cen[3 * (size_t)i + k] = 0.5f * boxes[i].mn[k] + 0.5f * boxes[i].mx[k];
    struct Task { uint32_t node, first, count, depth; };  // depth: the traversal stack holds kStackSize entries
    std::vector<Task> stack;
    nodes.reserve(n);
    nodes.push_back(BvhNode{});
    stack.push_back({0u, 0u, n, 0u});
    *root_ref = 0;
    auto range_box = [&](uint32_t first, uint32_t count) {
        SahBox b;
        b.reset();
        for (uint32_t i = first; i < first + count; i++) b.grow(boxes[perm[i]]);
        return b;
    };
    while (!stack.empty()) {
        const Task t = stack.back();
        stack.pop_back();
        // centroid bounds
        float cmn[3] = {INFINITY, INFINITY, INFINITY}, cmx[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (uint32_t i = t.first; i < t.first + t.count; i++)
            for (int k = 0; k < 3; k++) {
                const float c = cen[3 * (size_t)perm[i] + k];
                cmn[k] = std::fmin(cmn[k], c); cmx[k] = std::fmax(cmx[k], c);
            }
        // best binned split over the three axes
        double best_cost = INFINITY;
        int best_axis = -1, best_split = 0;
        for (int a = 0; a < 3; a++) {
            const float ext = cmx[a] - cmn[a];
            if (!(ext > 0.0f)) continue;
            const float scale = (float)kSahBins / ext;
            SahBox bb[kSahBins];
            uint32_t bc[kSahBins] = {0};
            for (int b = 0; b < kSahBins; b++) bb[b].reset();
            for (uint32_t i = t.first; i < t.first + t.count; i++) {
                const uint32_t p = perm[i];
                int b = (int)((cen[3 * (size_t)p + a] - cmn[a]) * scale);
                b = std::min(std::max(b, 0), kSahBins - 1);
                bc[b]++;
                bb[b].grow(boxes[p]);
            }
            double right_area[kSahBins];
            uint32_t right_cnt[kSahBins];
            SahBox acc;
            acc.reset();
            uint32_t cnt = 0;
            for (int b = kSahBins - 1; b > 0; b--) {
                acc.grow(bb[b]); cnt += bc[b];
                right_area[b] = acc.half_area(); right_cnt[b] = cnt;
            }
            acc.reset();
            cnt = 0;
            for (int s = 1; s < kSahBins; s++) {  // left = bins [0, s), right = bins [s, kSahBins)
                acc.grow(bb[s - 1]); cnt += bc[s - 1];
                if (cnt == 0 || right_cnt[s] == 0) continue;
                const double cost = acc.half_area() * cnt + right_area[s] * right_cnt[s];
                if (cost < best_cost) { best_cost = cost; best_axis = a; best_split = s; }
            }
        }
        uint32_t nl;
        if (best_axis >= 0) {
            const int a = best_axis;
            const float scale = (float)kSahBins / (cmx[a] - cmn[a]);
            auto mid = std::partition(perm.begin() + t.first, perm.begin() + t.first + t.count, [&](uint32_t p) {
                int b = (int)((cen[3 * (size_t)p + a] - cmn[a]) * scale);
                b = std::min(std::max(b, 0), kSahBins - 1);
                return b < best_split;
            });
            nl = (uint32_t)(mid - (perm.begin() + t.first));
        } else {
            nl = t.count / 2;  // all centroids coincide: any split is as good
        }
        if (nl == 0 || nl == t.count || t.depth >= kSahFreeDepth) nl = t.count / 2;  // bounded depth: halving from here on
        const uint32_t nr = t.count - nl;
        const SahBox lb = range_box(t.first, nl), rb = range_box(t.first + nl, nr);
        int refs[2];
        const uint32_t firsts[2] = {t.first, t.first + nl}, counts[2] = {nl, nr};
        for (int c = 0; c < 2; c++) {
            if (counts[c] <= (uint32_t)leaf_max) {
                refs[c] = rptx::make_leaf_ref(firsts[c], counts[c]);
            } else {
                refs[c] = (int)nodes.size();
                nodes.push_back(BvhNode{});
                stack.push_back({(uint32_t)refs[c], firsts[c], counts[c], t.depth + 1});
            }
        }
        BvhNode nd;
        for (int k = 0; k < 3; k++) {
            nd.lmin[k] = sah_down((double)lb.mn[k] - (double)pad); nd.lmax[k] = sah_up((double)lb.mx[k] + (double)pad);
            nd.rmin[k] = sah_down((double)rb.mn[k] - (double)pad); nd.rmax[k] = sah_up((double)rb.mx[k] + (double)pad);
        }
        nd.left = refs[0]; nd.right = refs[1]; nd.pad0 = 0; nd.pad1 = 0;
        nodes[t.node] = nd;
    }
}

}  // namespace rptx_host

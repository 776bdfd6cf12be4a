// Collapse of the binary radix tree into an 8-wide BVH with quantised child boxes
// (80-byte nodes, layout of Ylitie, Karras, Laine 2017).  Runs on the device, one
// tree level per launch: each thread owns one wide node, pulls up to 8 children out
// of the binary tree by repeatedly opening the child with the largest surface area,
// assigns them to octant-ordered slots, allocates its inner children / primitive
// ranges with atomics and quantises the child boxes conservatively.
#include "lj_bvh_build.h"
#include "lj_alloc.h"

#include <cooperative_groups.h>
#include <cfloat>
#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { snprintf(err, 256, "%s: %s", #x, cudaGetErrorString(e_)); return -1; } } while (0)

namespace {

constexpr int MAX_LEAF_CAP = 3;   // primitives per leaf slot (unary count in 3 bits)

struct Work { int node2; int node8; };

// number of primitives / first sorted slot under binary-tree child id (inner >= 0, leaf = ~slot)
__device__ __forceinline__ int child_count(int c, const int *count) { return c < 0 ? 1 : count[c]; }

__device__ __forceinline__ void child_box(int c, const float4 *node_lo, const float4 *node_hi, const int *sorted_prim,
                                          const float4 *prim_lo, const float4 *prim_hi, float *lo, float *hi) {
    float4 l, h;
    if (c < 0) { int p = sorted_prim[~c]; l = prim_lo[p]; h = prim_hi[p]; } else { l = node_lo[c]; h = node_hi[c]; }
    lo[0] = l.x; lo[1] = l.y; lo[2] = l.z; hi[0] = h.x; hi[1] = h.y; hi[2] = h.z;
}

// One work item = one 8-wide node: gathers up to 8 children of binary node w.node2, writes node w.node8, queues the inner children.
__device__ void collapse_item(const Work w, Work *items_out, int *n_out,
                              const int *left, const int *right, const float4 *node_lo, const float4 *node_hi,
                              const int *count, const int *sorted_prim,
                              const float4 *prim_lo, const float4 *prim_hi, const float4 *tri_verts,
                              Bvh8Node *nodes8, int *node_counter, float4 *tri_verts8, int *prim_counter, int n_prims_total, int MAX_LEAF, int open_leaves) {
    int child[8]; int nchild;
    float plo[3], phi[3];
    if (w.node2 < 0) {            // degenerate tree: a single primitive
        child[0] = w.node2; nchild = 1;
        child_box(w.node2, node_lo, node_hi, sorted_prim, prim_lo, prim_hi, plo, phi);
    } else {
        child[0] = left[w.node2]; child[1] = right[w.node2]; nchild = 2;
        float4 l = node_lo[w.node2], h = node_hi[w.node2];
        plo[0] = l.x; plo[1] = l.y; plo[2] = l.z; phi[0] = h.x; phi[1] = h.y; phi[2] = h.z;
    }
    // open the largest expandable child until 8 children
    while (nchild < 8) {
        int best = -1; float best_area = -1.f;
        for (int i = 0; i < nchild; i++) {
            int c = child[i];
            if (c < 0 || child_count(c, count) <= MAX_LEAF) continue;
            float4 l = node_lo[c], h = node_hi[c];
            float dx = h.x - l.x, dy = h.y - l.y, dz = h.z - l.z;
            float a = dx * dy + dy * dz + dz * dx;
            if (a > best_area) { best_area = a; best = i; }
        }
        if (best < 0) break;
        int c = child[best];
        child[best] = left[c];
        child[nchild++] = right[c];
    }
    // Slots left over (the bottom of the tree: every child already holds <= MAX_LEAF primitives): split multi-primitive
    // leaf children as well, largest box first, so that each primitive (or pair) gets its own, tighter child box.  Same
    // number of nodes, fewer primitive tests.
    while (open_leaves && nchild < 8) {
        int best = -1; float best_area = -1.f;
        for (int i = 0; i < nchild; i++) {
            int c = child[i];
            if (c < 0) continue;
            float4 l = node_lo[c], h = node_hi[c];
            float dx = h.x - l.x, dy = h.y - l.y, dz = h.z - l.z;
            float a = dx * dy + dy * dz + dz * dx;
            if (a > best_area) { best_area = a; best = i; }
        }
        if (best < 0) break;
        int c = child[best];
        child[best] = left[c];
        child[nchild++] = right[c];
    }
    // greedy assignment of children to octant slots: slot bit a set <=> child lies on the + side of axis a
    float cx = 0.5f * (plo[0] + phi[0]), cy = 0.5f * (plo[1] + phi[1]), cz = 0.5f * (plo[2] + phi[2]);
    float off[8][3];
    float clo[8][3], chi[8][3];
    for (int i = 0; i < nchild; i++) {
        child_box(child[i], node_lo, node_hi, sorted_prim, prim_lo, prim_hi, clo[i], chi[i]);
        off[i][0] = 0.5f * (clo[i][0] + chi[i][0]) - cx;
        off[i][1] = 0.5f * (clo[i][1] + chi[i][1]) - cy;
        off[i][2] = 0.5f * (clo[i][2] + chi[i][2]) - cz;
    }
    int slot_child[8];
    for (int s = 0; s < 8; s++) slot_child[s] = -1;
    unsigned child_done = 0;
    for (int round = 0; round < nchild; round++) {
        float best = -FLT_MAX; int bc = -1, bs = -1;
        for (int i = 0; i < nchild; i++) {
            if (child_done & (1u << i)) continue;
            for (int s = 0; s < 8; s++) {
                if (slot_child[s] >= 0) continue;
                float cost = ((s & 1) ? off[i][0] : -off[i][0]) + ((s & 2) ? off[i][1] : -off[i][1]) + ((s & 4) ? off[i][2] : -off[i][2]);
                if (cost > best) { best = cost; bc = i; bs = s; }
            }
        }
        slot_child[bs] = bc; child_done |= 1u << bc;
    }
    // classify, allocate
    int n_inner = 0, n_leaf_prims = 0;
    unsigned imask = 0;
    for (int s = 0; s < 8; s++) {
        int i = slot_child[s];
        if (i < 0) continue;
        int c = child[i];
        int cnt = child_count(c, count);
        if (c >= 0 && cnt > MAX_LEAF) { imask |= 1u << s; n_inner++; } else n_leaf_prims += cnt;
    }
    int child_base = n_inner ? atomicAdd(node_counter, n_inner) : 0;
    int prim_base = n_leaf_prims ? atomicAdd(prim_counter, n_leaf_prims) : 0;
    int out_base = n_inner ? atomicAdd(n_out, n_inner) : 0;

    Bvh8Node nd;
    // quantisation frame: origin = node lo, per-axis power-of-two scale so that 255 steps cover the extent
    int e[3];
    for (int a = 0; a < 3; a++) {
        float ext = phi[a] - plo[a];
        int ea;
        if (!(ext > 0.f)) ea = -100;
        else {
            ea = (int)ceilf(log2f(ext / 255.f));
            // make sure 255 * 2^e really reaches the far side in float arithmetic
            while (plo[a] + 255.f * exp2f((float)ea) < phi[a]) ea++;
        }
        if (ea < -126) ea = -126;
        e[a] = ea;
    }
    nd.px = plo[0]; nd.py = plo[1]; nd.pz = plo[2];
    nd.ex = (uint8_t)(e[0] + 127); nd.ey = (uint8_t)(e[1] + 127); nd.ez = (uint8_t)(e[2] + 127);
    nd.imask = (uint8_t)imask; nd.child_base = child_base; nd.prim_base = prim_base;
    int inner_rank = 0, prim_off = 0;
    for (int s = 0; s < 8; s++) {
        int i = slot_child[s];
        uint8_t q[6] = {0, 0, 0, 0, 0, 0};
        if (i < 0) { nd.meta[s] = 0; }
        else {
            int c = child[i];
            int cnt = child_count(c, count);
            if (imask & (1u << s)) {
                nd.meta[s] = (uint8_t)((1u << 5) | (24 + s));
                Work o; o.node2 = c; o.node8 = child_base + inner_rank;
                items_out[out_base + inner_rank] = o;
                inner_rank++;
            } else {
                unsigned unary = (1u << cnt) - 1u;
                nd.meta[s] = (uint8_t)((unary << 5) | prim_off);
                // enumerate the (at most MAX_LEAF) leaves of this small subtree
                int stk[MAX_LEAF_CAP + 1]; int sp = 0, k = 0;
                stk[sp++] = c;
                while (sp) {
                    int x = stk[--sp];
                    if (x >= 0) { stk[sp++] = right[x]; stk[sp++] = left[x]; continue; }
                    int src = ~x, dst = prim_base + prim_off + k++;
                    tri_verts8[3 * dst + 0] = tri_verts[3 * src + 0];
                    tri_verts8[3 * dst + 1] = tri_verts[3 * src + 1];
                    tri_verts8[3 * dst + 2] = tri_verts[3 * src + 2];
                }
                prim_off += cnt;
            }
            for (int a = 0; a < 3; a++) {
                float scale = exp2f((float)e[a]);
                float p = plo[a];
                int ql = (int)floorf((clo[i][a] - p) / scale);
                int qh = (int)ceilf((chi[i][a] - p) / scale);
                ql = max(0, min(255, ql)); qh = max(0, min(255, qh));
                while (ql > 0 && p + (float)ql * scale > clo[i][a]) ql--;
                while (qh < 255 && p + (float)qh * scale < chi[i][a]) qh++;
                q[a] = (uint8_t)ql; q[3 + a] = (uint8_t)qh;
            }
        }
        nd.qlox[s] = q[0]; nd.qloy[s] = q[1]; nd.qloz[s] = q[2];
        nd.qhix[s] = q[3]; nd.qhiy[s] = q[4]; nd.qhiz[s] = q[5];
    }
    nodes8[w.node8] = nd;
}

struct CollapseArgs {
    const int *left, *right; const float4 *node_lo, *node_hi; const int *count, *sorted_prim;
    const float4 *prim_lo, *prim_hi, *tri_verts;
    Bvh8Node *nodes8; float4 *tri_verts8; int n_prims_total, max_leaf, open_leaves;
};
// counters: [0] 8-wide nodes allocated, [1] primitive slots, [2..4] queue sizes of three consecutive levels (rotating), [5] levels

// one level per launch (LJ_BVH8_COOP=0): the host reads the size of the next level back after each
__global__ void k_collapse_level(const Work *items, int n_items, Work *items_out, int *n_out, int *counters, CollapseArgs a) {
    int wi = blockIdx.x * blockDim.x + threadIdx.x;
    if (wi >= n_items) return;
    collapse_item(items[wi], items_out, n_out, a.left, a.right, a.node_lo, a.node_hi, a.count, a.sorted_prim, a.prim_lo, a.prim_hi, a.tri_verts,
                  a.nodes8, counters, a.tri_verts8, counters + 1, a.n_prims_total, a.max_leaf, a.open_leaves);
}
__global__ void k_collapse_root(Work *q0, const int *d_root2, int n) {
    Work root; root.node2 = (n >= 2) ? *d_root2 : ~0; root.node8 = 0;
    q0[0] = root;
}

// every level in ONE cooperative launch (the default): a grid barrier between levels instead of a launch, a memset and a
// blocking read-back.  The queue size of level L is counters[2 + L % 3]; level L fills the counter of L + 1 and zeroes
// the one of L + 2 (last read before the previous barrier).
__global__ void __launch_bounds__(64) k_collapse_all(Work *q0, Work *q1, int *counters, const int *d_root2, CollapseArgs a) {
    namespace cg = cooperative_groups;
    cg::grid_group grid = cg::this_grid();
    const int gtid = blockIdx.x * blockDim.x + threadIdx.x, gsize = gridDim.x * blockDim.x;
    int level = 0, n_items = 1;
    while (n_items > 0 && level <= 128) {
        const Work *items = (level & 1) ? q1 : q0;
        Work *items_out = (level & 1) ? q0 : q1;
        int *n_out = counters + 2 + (level + 1) % 3;
        if (gtid == 0) counters[2 + (level + 2) % 3] = 0;
        for (int wi = gtid; wi < n_items; wi += gsize) {
            Work w;
            if (level == 0) { w.node2 = (a.n_prims_total >= 2) ? *d_root2 : ~0; w.node8 = 0; } else w = items[wi];
            collapse_item(w, items_out, n_out, a.left, a.right, a.node_lo, a.node_hi, a.count, a.sorted_prim, a.prim_lo, a.prim_hi, a.tri_verts,
                          a.nodes8, counters, a.tri_verts8, counters + 1, a.n_prims_total, a.max_leaf, a.open_leaves);
        }
        grid.sync();
        n_items = *(volatile int *)n_out;
        level++;
    }
    if (gtid == 0) counters[5] = level;
}

} // namespace

// d_root2 (device): root id of the binary tree (ignored when n < 2).  d_counters_out (device, 6 ints): the counters described
// above, left for the caller to read back together with its own results: [0] nodes, [1] slots, [5] levels (> 128 = did not terminate).
int lj_bvh8_collapse(int n, const int *d_root2, const float4 *tri_verts, const int *left, const int *right, const int *count,
                     const float4 *node_lo, const float4 *node_hi, const int *sorted_prim,
                     const float4 *prim_lo, const float4 *prim_hi, BvhBuildOutput &out, int *counters, char *err) {
    if (n <= 0) return 0;
    Work *q0 = nullptr, *q1 = nullptr;
    CK(lj_malloc(&q0, sizeof(Work) * (size_t)n)); CK(lj_malloc(&q1, sizeof(Work) * (size_t)n));
    CK(lj_malloc(&out.bvh8, sizeof(Bvh8Node) * (size_t)n));
    CK(lj_malloc(&out.tri_verts8, sizeof(float4) * 3 * (size_t)n));
    int h_counters[6] = {1 /* nodes allocated: root */, 0 /* prims */, 1, 0, 0, 0};
    CK(cudaMemcpyAsync(counters, h_counters, sizeof h_counters, cudaMemcpyHostToDevice, 0));
    CollapseArgs a;
    a.left = left; a.right = right; a.node_lo = node_lo; a.node_hi = node_hi; a.count = count; a.sorted_prim = sorted_prim;
    a.prim_lo = prim_lo; a.prim_hi = prim_hi; a.tri_verts = tri_verts; a.nodes8 = out.bvh8; a.tri_verts8 = out.tri_verts8;
    a.n_prims_total = n;
    a.max_leaf = MAX_LEAF_CAP;        // LJ_MAX_LEAF: experiment knob (1..3)
    if (const char *e = getenv("LJ_MAX_LEAF")) { int v = atoi(e); if (v >= 1 && v <= MAX_LEAF_CAP) a.max_leaf = v; }
    a.open_leaves = 1;                // LJ_OPEN_LEAVES=0: experiment knob
    if (const char *e = getenv("LJ_OPEN_LEAVES")) a.open_leaves = atoi(e) != 0;
    bool coop = true;
    if (const char *e = getenv("LJ_BVH8_COOP")) coop = atoi(e) != 0;
    if (coop) {
        int per_sm = 0, dev = 0, sm_count = 0;
        CK(cudaGetDevice(&dev));
        CK(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_collapse_all, 64, 0));
        int grid = per_sm * sm_count;
        if (grid < 1) { snprintf(err, 256, "bvh8 collapse: no resident block for the cooperative launch"); return -1; }
        const int want = (n / 4 + 63) / 64 + 1;         // the widest level holds fewer than n / 4 nodes
        if (grid > want) grid = want;
        void *args[] = {&q0, &q1, &counters, &d_root2, &a};
        CK(cudaLaunchCooperativeKernel((const void *)k_collapse_all, dim3(grid), dim3(64), args, 0, 0));
    } else {
        k_collapse_root<<<1, 1>>>(q0, d_root2, n);
        int n_items = 1, levels = 0;
        while (n_items > 0) {
            int *n_out = counters + 2 + (levels + 1) % 3;
            CK(cudaMemsetAsync(n_out, 0, sizeof(int), 0));
            k_collapse_level<<<(n_items + 63) / 64, 64>>>(q0, n_items, q1, n_out, counters, a);
            CK(cudaGetLastError());
            CK(cudaMemcpy(&n_items, n_out, sizeof(int), cudaMemcpyDeviceToHost));
            Work *t = q0; q0 = q1; q1 = t;
            if (++levels > 128) break;
        }
        CK(cudaMemcpyAsync(counters + 5, &levels, sizeof(int), cudaMemcpyHostToDevice, 0));
        CK(cudaStreamSynchronize(0));       // `levels` is a stack variable
    }
    lj_free(q0); lj_free(q1);
    return 0;
}

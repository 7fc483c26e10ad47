// Exact kNN of ONE query by ONE warp over the bucket hierarchy (device code shared by knn_warp_kernel,
// pcp_knn.cu, and the overflow drain of the tiled kernel, pcp_knn_tile.cu).
// Replaces kdtree_knn_search (kdtree.py:141-172) and KNNResultSet.add_point (result_set.py:35-53).
#pragma once

#include "pcp_device.cuh"
#include "pcp_internal.cuh"

namespace pcp {
namespace warpq {

constexpr unsigned kDone = 0xffffffffu;   // "child visited / absent" marker (> every float bit pattern in use)

__device__ __forceinline__ unsigned long long shfl64(unsigned long long v, int src) {
    return __shfl_sync(kFullMask, v, src);
}

// Ascending bitonic sort of one 64-bit key per lane across the warp.
__device__ __forceinline__ unsigned long long warp_sort32(unsigned long long key, int lane) {
#pragma unroll
    for (int size = 2; size <= 32; size <<= 1) {
#pragma unroll
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            const unsigned long long other = __shfl_xor_sync(kFullMask, key, stride);
            const bool up = ((lane & size) == 0) || size == 32;      // final merge is ascending everywhere
            const bool lower = (lane & stride) == 0;
            const bool take_min = (up == lower);
            const unsigned long long mn = key < other ? key : other;
            const unsigned long long mx = key < other ? other : key;
            key = take_min ? mn : mx;
        }
    }
    return key;
}

// Ascending sort of a BITONIC sequence held one key per lane (5 compare-exchange stages).
__device__ __forceinline__ unsigned long long warp_merge32(unsigned long long key, int lane) {
#pragma unroll
    for (int stride = 16; stride > 0; stride >>= 1) {
        const unsigned long long other = __shfl_xor_sync(kFullMask, key, stride);
        const bool lower = (lane & stride) == 0;
        const unsigned long long mn = key < other ? key : other;
        const unsigned long long mx = key < other ? other : key;
        key = lower ? mn : mx;
    }
    return key;
}

constexpr int kMergeThreshold = 10;   // survivors per bucket above which sort+merge beats serial insertion

struct QueryState {
    float qx, qy, qz;
    float gate;                 // squared-distance gate: d2_gate of the current k-th distance (+inf until k found)
    unsigned long long key;     // this lane's slot of the sorted 32-entry list
    bool first;                 // no bucket scanned yet
    unsigned n_test, n_scan, n_ins, n_open;
};

// p: this lane's point of the bucket (lane j = point j), loaded by the caller
__device__ __forceinline__ void scan_bucket(QueryState &st, const float4 p, int lane, int k) {
    ++st.n_scan;
    const float d2 = ref_d2(st.qx, st.qy, st.qz, p.x, p.y, p.z);
    if (st.first) {
        // pad points (NaN coordinates) sort to the end as empty keys
        const unsigned long long nk = (d2 == d2) ? pack_key(__fsqrt_rn(d2), __float_as_uint(p.w)) : kEmptyKey;
        st.key = warp_sort32(nk, lane);
        st.first = false;
    } else {
        const bool pass = d2 <= st.gate;                                 // NaN never passes
        unsigned m = __ballot_sync(kFullMask, pass);
        const unsigned long long mine = pack_key(__fsqrt_rn(d2), __float_as_uint(p.w));
        if (__popc(m) > kMergeThreshold) {
            // many survivors: sort them and bitonic-merge with the list (cost independent of the count)
            const unsigned long long sorted_new = warp_sort32(pass ? mine : kEmptyKey, lane);
            const unsigned long long rev = shfl64(sorted_new, 31 - lane);
            st.key = warp_merge32(st.key < rev ? st.key : rev, lane);    // 32 smallest of the 64, bitonic -> sorted
            st.n_ins += kMergeThreshold;
        } else {
            while (m) {
                const int j = __ffs(m) - 1;
                m &= m - 1;
                const unsigned long long nk = shfl64(mine, j);
                const bool gt = st.key > nk;                             // a suffix of the lanes
                const unsigned bal = __ballot_sync(kFullMask, gt);
                const unsigned long long up = __shfl_up_sync(kFullMask, st.key, 1);
                const int pos = __ffs(bal) - 1;                          // -1: nk is larger than all 32 -> dropped
                st.key = gt ? (lane == pos ? nk : up) : st.key;
                ++st.n_ins;
            }
        }
    }
    st.gate = fminf(st.gate, d2_gate(key_dist(shfl64(st.key, k - 1))));   // (a seeded gate stays until beaten)
}

// Search: `st` holds the query (qx, qy, qz) on entry and the sorted result list (st.key, lane i = i-th best) on
// return.  lbs / cds: [kMaxLevels][32] words of per-warp shared memory, nodes: [kMaxLevels].
// gate0: a squared-distance gate already known to hold the k nearest neighbours and everything tied with the
// k-th (+inf when nothing is known): boxes and points beyond it are never looked at.
__device__ __forceinline__ void search(const IndexView &ix, QueryState &st, int lane, int k, unsigned (*lbs)[32],
                                       unsigned (*cds)[32], unsigned *nodes, float gate0) {
    st.gate = gate0;
    st.key = kEmptyKey;
    st.first = true;
    st.n_test = st.n_scan = st.n_ins = st.n_open = 0;
    const int top = ix.n_levels - 1;
    if (top == 0) {
        scan_bucket(st, __ldg(ix.pts + lane), lane, k);
    } else if (top > 0) {
        // children of `node` (a level-`lv` node) live on level lv-1
        auto open_node = [&](int lv, unsigned node) {
            ++st.n_open;
            if (lane == 0) nodes[lv] = node;
            const long long c = (long long)node * kFan + lane;
            unsigned bits = kDone, cbits = kDone;
            if (c < ix.level_cnt[lv - 1]) {
                const long long b = ix.level_off[lv - 1] + c;
                const float4 lo = __ldg(ix.box_lo + b), hi = __ldg(ix.box_hi + b);
                bits = __float_as_uint(box_lb2(st.qx, st.qy, st.qz, lo, hi));
                const float cx = 0.5f * (lo.x + hi.x) - st.qx, cy = 0.5f * (lo.y + hi.y) - st.qy,
                            cz = 0.5f * (lo.z + hi.z) - st.qz;
                cbits = __float_as_uint(cx * cx + cy * cy + cz * cz);   // ordering hint only, never used to prune
            }
            lbs[lv][lane] = bits;
            cds[lv][lane] = cbits;
            __syncwarp();
        };
        // nearest pending child of the open node on `level` (ties: nearest box centre), or -1; `bound` = its
        // lower bound.  Does not mark it visited.
        auto select = [&](int level, unsigned &bound) -> int {
            const unsigned mine = lbs[level][lane];
            bound = __reduce_min_sync(kFullMask, mine);
            if (bound == kDone) return -1;
            unsigned tie = __ballot_sync(kFullMask, mine == bound);
            if (tie & (tie - 1)) {   // several children at the same bound (typically 0: the query is inside them)
                const unsigned cd = (mine == bound) ? cds[level][lane] : kDone;
                const unsigned cbest = __reduce_min_sync(kFullMask, cd);
                tie = __ballot_sync(kFullMask, cd == cbest);
            }
            return __ffs(tie) - 1;
        };
        auto mark = [&](int level, int sel) {
            if (lane == sel) lbs[level][lane] = kDone;
            __syncwarp();
        };
        int level = top;
        open_node(top, 0u);
        while (true) {
            unsigned bound;
            int sel = select(level, bound);
            // exact prune: every point below a child is at least `bound` away (in d2 arithmetic)
            if (sel < 0 || !(__uint_as_float(bound) <= st.gate)) {
                if (level == top) break;
                ++level;
                continue;
            }
            mark(level, sel);
            ++st.n_test;
            if (level > 1) {
                const unsigned child = nodes[level] * kFan + sel;
                --level;
                open_node(level, child);
                continue;
            }
            // level 1: buckets.  The visiting order depends only on the bounds, so the NEXT bucket's points are
            // requested before the current one is ranked (a query is one dependent chain of L2 round trips
            // otherwise); if the tightened gate then excludes it, so it does every bucket after it.
            const float4 *base = ix.pts + (long long)nodes[1] * kFan * kLeaf + lane;
            float4 cur = __ldg(base + sel * kLeaf);
            while (true) {
                unsigned nb;
                const int nxt = select(1, nb);
                const bool more = nxt >= 0 && __uint_as_float(nb) <= st.gate;
                float4 pf = cur;
                if (more) pf = __ldg(base + nxt * kLeaf);
                scan_bucket(st, cur, lane, k);
                if (!more || !(__uint_as_float(nb) <= st.gate)) break;
                mark(1, nxt);
                ++st.n_test;
                cur = pf;
            }
        }
    }

}

}  // namespace warpq
}  // namespace pcp

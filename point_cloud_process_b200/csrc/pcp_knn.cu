// Exact kNN over the bucket hierarchy -- one warp per query.
// Replaces kdtree_knn_search (kdtree.py:141-172), octree_knn_search (octree.py:262-306) and
// KNNResultSet.add_point (result_set.py:35-53).
//
// The warp owns ONE query.  Its result list is the warp itself: lane i holds the i-th best packed
// (distance, index) key, so the list is a 32-entry sorted register file distributed over the lanes
// and every k <= 32 is served by the same code.  The warp walks the 32-ary AABB hierarchy
// nearest-first: on every node lane j evaluates the exact lower bound of child j (two coalesced
// float4 loads), a redux.min picks the nearest unvisited child, and the walk of a node stops as
// soon as that bound exceeds the current k-th distance gate.  A bucket is scanned with one
// coalesced 512-byte load (lane j = point j), the distance is the reference's unfused float32
// expression, and the few candidates inside the gate are inserted with ballot + shfl_up.  The first
// bucket is bitonic-sorted straight into the list.
#include <cfloat>
#include <cstdlib>
#include <cstring>

#include "pcp_device.cuh"
#include "pcp_internal.cuh"
#include "pcp_knn_warp.cuh"

namespace pcp {

namespace {

constexpr int kWarpsPerCta = 8;

// LIST = false: warp w of the grid takes query w (SELF: sorted slot w).
// LIST = true:  persistent warps drain the overflow list of the tiled kernel (pcp_knn_tile.cu) WHILE that kernel may
//               still be running (launched as its programmatic dependent): list[1 + e] = group id + 1 once entry e is
//               published (0 before), sched[1] = groups the tiled kernel has finished (handed-over ones included).
//               Warp item w -> query lane (w % 32) of entry w / 32; a warp waits for its entry until it appears or
//               until all n_groups groups are accounted for without it.
template <bool SELF, bool LIST>
__global__ void __launch_bounds__(kWarpsPerCta * 32)
knn_warp_kernel(const IndexView ix, const float *__restrict__ queries, long long nq, long long stride, int k,
                int *__restrict__ out_idx, float *__restrict__ out_dist, int *__restrict__ out_cnt,
                unsigned long long *__restrict__ ctr, const unsigned *list, const unsigned *sched, unsigned n_groups,
                const unsigned *__restrict__ perm, const float *hints, unsigned hint_groups) {
    // per-level traversal state: lane j keeps the pending lower bound of child j of the open node
    __shared__ unsigned s_lb[kWarpsPerCta][kMaxLevels][32];
    __shared__ unsigned s_cd[kWarpsPerCta][kMaxLevels][32];   // squared distance to the child's box centre
    __shared__ unsigned s_node[kWarpsPerCta][kMaxLevels];
    __shared__ unsigned s_entry;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long n_work = LIST ? (long long)n_groups * 32 : (SELF ? ix.n_leaf * kLeaf : nq);
    for (long long item = blockIdx.x * (long long)kWarpsPerCta + warp; item < n_work;
         item += LIST ? gridDim.x * (long long)kWarpsPerCta : n_work) {
    long long wid = item;
    if (LIST) {
        // the 8 items of a CTA belong to one entry: ONE thread polls for it (thousands of warps polling the words
        // the tiled kernel's scheduler lives on would starve that kernel's own atomics), the others wait at the barrier
        __syncthreads();                         // (everyone has read s_entry of the previous round)
        if (threadIdx.x == 0) {
            const unsigned *entry = list + 1 + (item >> 5);
            unsigned g1 = 0, spins = 0;
            for (;;) {
                asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(g1) : "l"(entry) : "memory");
                if (g1) break;
                if ((spins++ & 3u) == 0u) {
                    unsigned done;
                    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(done) : "l"(sched + 1) : "memory");
                    if (done >= n_groups) {      // every publication precedes its group's count: look once more
                        __threadfence();
                        asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(g1) : "l"(entry) : "memory");
                        break;
                    }
                }
                __nanosleep(2000);
            }
            __threadfence();                     // acquire: the entry's gates are read below
            s_entry = g1;
        }
        __syncthreads();
        const unsigned g1 = s_entry;
        if (!g1) break;                          // no such entry (whole CTA): neither will any later one exist
        wid = (long long)(g1 - 1u) * 32 + (item & 31);                       // slot (SELF) / sorted position
    }
    warpq::QueryState st;
    long long qid = wid;
    if (SELF) {
        const float4 p = __ldg(ix.pts + wid);
        st.qx = p.x; st.qy = p.y; st.qz = p.z;
        qid = (long long)(int)__float_as_uint(p.w);
        if (qid < 0) continue;                     // pad slot of the last bucket (whole warp skips)
    } else {
        if (LIST) {
            if (wid >= nq) continue;
            qid = perm ? (long long)perm[wid] : wid;
        }
        const float *q = queries + qid * stride;
        st.qx = __ldg(q); st.qy = __ldg(q + 1); st.qz = __ldg(q + 2);
    }
    // LIST: the gate the tiled kernel had reached for this query when it gave the group up (first hint_groups entries)
    float gate0 = __int_as_float(0x7f800000);
    if (LIST && (item >> 5) < (long long)hint_groups) gate0 = __ldcg(hints + item);
    warpq::search(ix, st, lane, k, s_lb[warp], s_cd[warp], s_node[warp], gate0);

    if (ctr && lane == 0) {
        atomicAdd(ctr + 0, (unsigned long long)st.n_scan * kLeaf);
        if (LIST) atomicAdd(ctr + 7, (unsigned long long)st.n_scan * kLeaf);   // share of [0] spent on overflow groups
        if (!LIST) {   // (overflow queries of the tiled kernel: [1..5] keep describing the tiled kernel alone)
            atomicAdd(ctr + 1, (unsigned long long)st.n_open * kFan);
            atomicAdd(ctr + 2, (unsigned long long)st.n_open);
            atomicAdd(ctr + 3, (unsigned long long)st.n_ins);
            atomicAdd(ctr + 4, 1ull);
            atomicAdd(ctr + 5, (unsigned long long)st.n_scan);
        }
    }
    // lane i < k writes slot i: one coalesced row per query
    const unsigned id = key_index(st.key);
    const bool filled = id != 0xffffffffu;
    if (lane < k) {
        // result_set.py:19-22: unfilled slots read (1e10, 0)
        out_idx[qid * k + lane] = filled ? (int)id : 0;
        out_dist[qid * k + lane] = filled ? key_dist(st.key) : PCP_KNN_EMPTY_DIST;
    }
    if (out_cnt) {
        const int cnt = __popc(__ballot_sync(kFullMask, filled && lane < k));
        if (lane == 0) out_cnt[qid] = cnt;
    }
    __syncwarp();
    }
    // completion of this grid must imply completion of the tiled kernel (whose last groups may still be writing results)
    if (LIST) asm volatile("griddepcontrol.wait;" ::: "memory");
}

__global__ void fill_empty_kernel(long long nq, int k, int *idx, float *dist, int *cnt) {
    long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i < nq * k) { idx[i] = 0; dist[i] = PCP_KNN_EMPTY_DIST; }
    if (cnt && i < nq) cnt[i] = 0;
}

}  // namespace

// Overflow groups of the tiled kernel, one warp per query, persistent grid.  overlap: launched with programmatic
// stream serialisation -- the grid may start once every CTA of the tiled kernel in front of it on the stream has
// executed griddepcontrol.launch_dependents (all of them are resident by then, so this grid can only take what
// they leave or free) and consumes the list as it is published.  PCP_KNN_OVERLAP=0 keeps the two kernels serial.
int knn_warp_list_device(pcp_index *ix, bool self, const float *d_q, int64_t nq, int64_t stride, const uint32_t *d_perm,
                         const uint32_t *d_list, const uint32_t *d_sched, uint32_t n_groups, const float *d_hints,
                         uint32_t hint_groups, int k, int32_t *d_idx, float *d_dist, int32_t *d_count, bool overlap) {
    pcp_ctx *ctx = ix->ctx;
    const IndexView v = make_view(ix);
    static const bool allow = !(getenv("PCP_KNN_OVERLAP") && atoi(getenv("PCP_KNN_OVERLAP")) == 0);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)ctx->sm_count * 8);
    cfg.blockDim = dim3(kWarpsPerCta * 32);
    cfg.stream = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = overlap && allow ? 1 : 0;
    unsigned long long *ctr = ctx->ctr();
    const long long n = self ? ix->n : nq, st = self ? 3 : stride;
    const float *q = self ? nullptr : d_q;
    const unsigned *perm = self ? nullptr : d_perm;
    if (self)
        PCP_CUDA(cudaLaunchKernelEx(&cfg, knn_warp_kernel<true, true>, v, q, n, st, k, d_idx, d_dist, d_count, ctr,
                                    (const unsigned *)d_list, (const unsigned *)d_sched, n_groups, perm, d_hints, hint_groups));
    else
        PCP_CUDA(cudaLaunchKernelEx(&cfg, knn_warp_kernel<false, true>, v, q, n, st, k, d_idx, d_dist, d_count, ctr,
                                    (const unsigned *)d_list, (const unsigned *)d_sched, n_groups, perm, d_hints, hint_groups));
    ctx->launches += 1;
    PCP_CUDA(cudaGetLastError());
    return PCP_OK;
}

int knn_kernel_choice() {
    static const int choice = [] {
        const char *e = getenv("PCP_KNN_KERNEL");
        if (!e) return 0;
        return strcmp(e, "warp") == 0 ? 1 : strcmp(e, "tile") == 0 ? 2 : 0;
    }();
    return choice;
}

// Below this many queries the warp-per-query kernel wins: it needs no query sort and exposes 32x more warps.
constexpr int64_t kTileMinQueries = 8192;

int knn_device(pcp_index *ix, const float *d_q, int64_t nq, int64_t stride, const uint32_t *d_perm, int k,
               int32_t *d_idx, float *d_dist, int32_t *d_count, bool side_ok) {
    (void)d_perm;
    if (nq == 0) return PCP_OK;
    pcp_ctx *ctx = ix->ctx;
    const int choice = knn_kernel_choice();
    if (ix->n > 0 && (choice == 2 || (choice == 0 && nq >= kTileMinQueries))) {
        // group queries by Morton order in the index's frame, 32 per warp
        // The regrouping needs the index's bounding box only.  When the query data is certainly complete --
        // uploaded by the library on the side stream, or the very array the index was built from -- it runs on
        // the side stream, beside the rest of an index build that may still be queued on the main stream.
        if (side_ok || (d_q == ix->built_from && stride == ix->built_stride && nq == ix->n)) {
            PCP_TRY(ctx->q_perm.ensure(sizeof(uint32_t) * (size_t)nq));
            cudaStream_t side = ctx->copy_stream;
            if (ctx->capturing) {
                // inside a graph capture: fork the side stream from the build's bounding-box event when this capture
                // holds the build (first search after it), else from the main stream as it stands; join below
                if (!ctx->cap_has_bbox) PCP_CUDA(cudaEventRecord(ctx->ev_cap_bbox, ctx->stream));
                ctx->cap_has_bbox = false;
                PCP_CUDA(cudaStreamWaitEvent(side, ctx->ev_cap_bbox, 0));
            } else {
                PCP_CUDA(cudaStreamWaitEvent(side, ix->b.ev_bbox, 0));          // boxes of THIS index final
                PCP_CUDA(cudaStreamWaitEvent(side, ctx->ev_search_end, 0));     // last reader of q_perm done
            }
            PCP_TRY(sort_queries_on(ix, d_q, nq, stride, ctx->q_perm.as<uint32_t>(), side, ctx->q_keys, ctx->cub_tmp_side));
            cudaEvent_t ready = ctx->capturing ? ctx->ev_cap_q : ctx->ev_q_ready;
            PCP_CUDA(cudaEventRecord(ready, side));
            PCP_CUDA(cudaStreamWaitEvent(ctx->stream, ready, 0));
            PCP_TRY(knn_tile_device(ix, d_q, nq, stride, ctx->q_perm.as<uint32_t>(), k, d_idx, d_dist, d_count));
            if (!ctx->capturing) PCP_CUDA(cudaEventRecord(ctx->ev_search_end, ctx->stream));
            return PCP_OK;
        }
        PCP_TRY(ctx->scratch[3].ensure(sizeof(uint32_t) * (size_t)nq));
        PCP_TRY(sort_queries(ix, d_q, nq, stride, ctx->scratch[3].as<uint32_t>()));
        return knn_tile_device(ix, d_q, nq, stride, ctx->scratch[3].as<uint32_t>(), k, d_idx, d_dist, d_count);
    }
    if (ix->n == 0) {   // empty cloud: every KNNResultSet stays as constructed
        long long tot = std::max<long long>(nq * k, nq);
        fill_empty_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, ctx->stream>>>(nq, k, d_idx, d_dist, d_count);
        ctx->launches += 1;
        PCP_CUDA(cudaGetLastError());
        return PCP_OK;
    }
    const IndexView v = make_view(ix);
    const unsigned grid = (unsigned)((nq + kWarpsPerCta - 1) / kWarpsPerCta);
    knn_warp_kernel<false, false><<<grid, kWarpsPerCta * 32, 0, ctx->stream>>>(v, d_q, nq, stride, k, d_idx, d_dist,
                                                                               d_count, ctx->ctr(), nullptr, nullptr, 0u, nullptr, nullptr, 0u);
    ctx->launches += 1;
    PCP_CUDA(cudaGetLastError());
    return PCP_OK;
}

int knn_self_device(pcp_index *ix, int k, int32_t *d_idx, float *d_dist, int32_t *d_count) {
    if (ix->n == 0) return PCP_OK;
    pcp_ctx *ctx = ix->ctx;
    const int choice = knn_kernel_choice();
    if (choice == 2 || (choice == 0 && ix->n >= kTileMinQueries)) return knn_tile_self_device(ix, k, d_idx, d_dist, d_count);
    const IndexView v = make_view(ix);
    const long long n_work = v.n_leaf * kLeaf;
    const unsigned grid = (unsigned)((n_work + kWarpsPerCta - 1) / kWarpsPerCta);
    knn_warp_kernel<true, false><<<grid, kWarpsPerCta * 32, 0, ctx->stream>>>(v, nullptr, ix->n, 3, k, d_idx, d_dist,
                                                                              d_count, ctx->ctr(), nullptr, nullptr, 0u, nullptr, nullptr, 0u);
    ctx->launches += 1;
    PCP_CUDA(cudaGetLastError());
    return PCP_OK;
}

}  // namespace pcp

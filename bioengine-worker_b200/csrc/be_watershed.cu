// Seed detection and marker flooding — the back half of the fibsem post-processing
// (ref: apps/fibsem-mito-analysis/analysis_deployment.py:171-175):
//
//     coords = feature.peak_local_max(dist, min_distance=8, labels=closed)
//     labels = segmentation.watershed(-dist, measure.label(m), mask=closed)
//
//   be_peak_local_max_i32   skimage's peak_local_max on the squared distance (an integer, same order and same ties as
//                           the fp64 distance): bounding box of the border-stripped mask, separable (2d+1)^2 window
//                           maximum clipped to that box, "every pixel is a maximum" special case, then the greedy
//                           thinning to Chebyshev distance >= d.  The greedy pass walks the candidates by (value
//                           descending, raster order); here a candidate is accepted once every stronger candidate in
//                           its window is rejected and rejected once one of them is accepted — decisions are final, so
//                           they can be taken in any order: two grid-wide rounds settle everything except long plateaus
//                           (ridges of equal distance), which one block finishes round by round.
//   be_watershed_i32        skimage's flood (connectivity 1, no compactness, no lines): a priority queue ordered by
//                           (height, push order); a popped pixel gives its label to its unlabelled 4-neighbours (up,
//                           left, right, down) as it pushes them.  Connected regions of the mask flood independently, so
//                           one block owns one region and keeps that region's queue as an array in push order.  Per step
//                           it takes ALL entries of the lowest height at once — their pushes are ordered by (position of
//                           the pusher, neighbour number), which is the order the sequential flood would give — unless
//                           one of them uncovers a lower pixel: then only the entries up to and including that one are
//                           taken, because the sequential flood would run down into the new basin before it returns.
//                           A contested pixel goes to the entry around it that was pushed first (every queued pixel carries
//                           its push number), which each entry works out for itself from the records of the pixels two
//                           steps away.  Markers that tie in height leave in raster order (first in, first out); skimage's
//                           binary heap orders such complete ties by its array layout, which no parallel flood can
//                           reproduce — the oracle implements both and the tests say where they differ.
#include "be_common.cuh"

namespace {

constexpr int kThreads = 256;

// ---------------------------------------------------------------------------------------------------------------
// peak_local_max
// ---------------------------------------------------------------------------------------------------------------

struct PlmState {
    int ymin, ymax, xmin, xmax;  // bounding box of the mask pixels at least `d` away from every border (ymax < ymin: none)
    int dmin;                    // image.min() on the squared distance
    int not_max;                 // mask pixels inside the box that are below their window maximum
    int n_cand;                  // candidates written to the list
    int undecided;
};

__global__ void plm_init_kernel(PlmState* st) {
    st->ymin = st->xmin = INT32_MAX;
    st->ymax = st->xmax = -1;
    st->dmin = INT32_MAX;
    st->not_max = 0;
    st->n_cand = 0;
    st->undecided = 0;
}

__global__ void __launch_bounds__(kThreads) plm_roi_kernel(const int32_t* __restrict__ d2, const uint8_t* __restrict__ mask,
                                                           int Y, int X, int d, PlmState* st) {
    __shared__ int s[5];
    if (threadIdx.x == 0) {
        s[0] = s[2] = INT32_MAX;
        s[1] = s[3] = -1;
        s[4] = INT32_MAX;
    }
    __syncthreads();
    const int64_t n = (int64_t)Y * X;
    int ymin = INT32_MAX, ymax = -1, xmin = INT32_MAX, xmax = -1, dmin = INT32_MAX;
    for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kThreads) {
        dmin = min(dmin, d2[i]);
        if (!mask[i]) continue;
        const int y = (int)(i / X), x = (int)(i - (int64_t)y * X);
        if (y < d || y >= Y - d || x < d || x >= X - d) continue;
        ymin = min(ymin, y);
        ymax = max(ymax, y);
        xmin = min(xmin, x);
        xmax = max(xmax, x);
    }
    atomicMin(&s[0], ymin);
    atomicMax(&s[1], ymax);
    atomicMin(&s[2], xmin);
    atomicMax(&s[3], xmax);
    atomicMin(&s[4], dmin);
    __syncthreads();
    if (threadIdx.x == 0) {
        atomicMin(&st->ymin, s[0]);
        atomicMax(&st->ymax, s[1]);
        atomicMin(&st->xmin, s[2]);
        atomicMax(&st->xmax, s[3]);
        atomicMin(&st->dmin, s[4]);
    }
}

// value of a pixel inside the box as the reference sees it: pixels outside the label hold the lowest number
__device__ __forceinline__ int plm_value(const int32_t* d2, const uint8_t* mask, int64_t i) { return mask[i] ? d2[i] : INT32_MIN; }

__global__ void __launch_bounds__(kThreads) plm_rowmax_kernel(const int32_t* __restrict__ d2, const uint8_t* __restrict__ mask,
                                                              int32_t* __restrict__ rowmax, int Y, int X, int d,
                                                              const PlmState* __restrict__ st) {
    const int ymin = st->ymin, ymax = st->ymax, xmin = st->xmin, xmax = st->xmax;
    if (ymax < ymin) return;
    const int bw = xmax - xmin + 1;
    const int64_t n = (int64_t)(ymax - ymin + 1) * bw;
    for (int64_t j = (int64_t)blockIdx.x * kThreads + threadIdx.x; j < n; j += (int64_t)gridDim.x * kThreads) {
        const int y = ymin + (int)(j / bw), x = xmin + (int)(j % bw);
        const int a = max(x - d, xmin), b = min(x + d, xmax);
        int m = INT32_MIN;
        for (int xx = a; xx <= b; ++xx) m = max(m, plm_value(d2, mask, (int64_t)y * X + xx));
        rowmax[(int64_t)y * X + x] = m;
    }
}

// status: 0 no candidate, 1 undecided, 2 accepted, 3 rejected
__global__ void __launch_bounds__(kThreads) plm_peak_kernel(const int32_t* __restrict__ d2, const uint8_t* __restrict__ mask,
                                                            const int32_t* __restrict__ rowmax, uint8_t* __restrict__ status,
                                                            int Y, int X, int d, PlmState* st) {
    const int ymin = st->ymin, ymax = st->ymax, xmin = st->xmin, xmax = st->xmax, dmin = st->dmin;
    if (ymax < ymin) return;
    const int bw = xmax - xmin + 1, bh = ymax - ymin + 1;
    const bool plain = d == 0 || (bw == 1 && bh == 1);  // footprint or box of one pixel: image > threshold
    const int64_t n = (int64_t)bh * bw;
    int not_max = 0;
    for (int64_t j = (int64_t)blockIdx.x * kThreads + threadIdx.x; j < n; j += (int64_t)gridDim.x * kThreads) {
        const int y = ymin + (int)(j / bw), x = xmin + (int)(j % bw);
        const int64_t i = (int64_t)y * X + x;
        if (!mask[i]) continue;
        const int v = d2[i];
        bool is_max = true;
        if (!plain) {
            const int a = max(y - d, ymin), b = min(y + d, ymax);
            int m = INT32_MIN;
            for (int yy = a; yy <= b; ++yy) m = max(m, rowmax[(int64_t)yy * X + x]);
            is_max = v == m;
            not_max += !is_max;
        }
        status[i] = (is_max && v > dmin) ? 1 : 0;
    }
    if (!plain) {
        not_max = __reduce_add_sync(0xffffffffu, not_max);
        if ((threadIdx.x & 31) == 0 && not_max) atomicAdd(&st->not_max, not_max);
    }
}

// "trivial image" branch of skimage's _get_peak_mask: when every pixel of the label is a window maximum there are no
// peaks, except the pixels a binary opening (3x3 cross, outside of the box = 0) removes from the label.
__global__ void __launch_bounds__(kThreads) plm_trivial_kernel(const int32_t* __restrict__ d2, const uint8_t* __restrict__ mask,
                                                               uint8_t* __restrict__ status, int Y, int X, int d,
                                                               const PlmState* __restrict__ st) {
    const int ymin = st->ymin, ymax = st->ymax, xmin = st->xmin, xmax = st->xmax, dmin = st->dmin;
    if (ymax < ymin || st->not_max != 0) return;
    const int bw = xmax - xmin + 1, bh = ymax - ymin + 1;
    if (d == 0 || (bw == 1 && bh == 1)) return;
    auto in_label = [&](int y, int x) {
        return y >= ymin && y <= ymax && x >= xmin && x <= xmax && mask[(int64_t)y * X + x] != 0;
    };
    auto eroded = [&](int y, int x) {
        return in_label(y, x) && in_label(y - 1, x) && in_label(y + 1, x) && in_label(y, x - 1) && in_label(y, x + 1);
    };
    const int64_t n = (int64_t)bh * bw;
    for (int64_t j = (int64_t)blockIdx.x * kThreads + threadIdx.x; j < n; j += (int64_t)gridDim.x * kThreads) {
        const int y = ymin + (int)(j / bw), x = xmin + (int)(j % bw);
        const int64_t i = (int64_t)y * X + x;
        if (!mask[i]) continue;
        const bool opened = eroded(y, x) || eroded(y - 1, x) || eroded(y + 1, x) || eroded(y, x - 1) || eroded(y, x + 1);
        status[i] = (!opened && d2[i] > dmin) ? 1 : 0;
    }
}

__global__ void __launch_bounds__(kThreads) plm_list_kernel(const uint8_t* __restrict__ status, int32_t* __restrict__ list,
                                                            int X, PlmState* st) {
    const int ymin = st->ymin, ymax = st->ymax, xmin = st->xmin, xmax = st->xmax;
    if (ymax < ymin) return;
    const int bw = xmax - xmin + 1;
    const int64_t n = (int64_t)(ymax - ymin + 1) * bw;
    for (int64_t j = (int64_t)blockIdx.x * kThreads + threadIdx.x; j < n; j += (int64_t)gridDim.x * kThreads) {
        const int y = ymin + (int)(j / bw), x = xmin + (int)(j % bw);
        const int64_t i = (int64_t)y * X + x;
        if (status[i] == 1) list[atomicAdd(&st->n_cand, 1)] = (int32_t)i;
    }
}

// One look at candidate i by a whole warp (its lanes share the (2d-1)^2 window): 2 accept, 3 reject, 1 still waiting for a
// stronger neighbour.  "Stronger" = larger value, or equal value and earlier in raster order — the order of the
// reference's stable sort.  One thread walking the 225 positions of the d = 8 window was 80 us per round for a few
// thousand candidates; a warp per candidate needs 8 positions per lane.
__device__ __forceinline__ int plm_decide_warp(const int32_t* __restrict__ d2, const volatile uint8_t* status, int Y, int X,
                                               int d, int i) {
    const int lane = threadIdx.x & 31;
    const int y = i / X, x = i - y * X;
    const int v = d2[i];
    const int side = 2 * d - 1;
    bool rejected = false, blocked = false;
    for (int p = lane; p < side * side; p += 32) {
        const int yy = y - d + 1 + p / side, xx = x - d + 1 + p % side;
        if (yy < 0 || yy >= Y || xx < 0 || xx >= X) continue;
        const int j = yy * X + xx;
        const int s = status[j];
        if (s == 0 || s == 3 || j == i) continue;
        const int w = d2[j];
        if (w > v || (w == v && j < i)) {
            rejected |= s == 2;
            blocked = true;
        }
    }
    if (__any_sync(0xffffffffu, rejected)) return 3;
    return __any_sync(0xffffffffu, blocked) ? 1 : 2;
}

__global__ void __launch_bounds__(kThreads) plm_round_kernel(const int32_t* __restrict__ d2, uint8_t* status,
                                                             const int32_t* __restrict__ list, int Y, int X, int d,
                                                             const PlmState* __restrict__ st) {
    const int n = st->n_cand;
    const int warps = gridDim.x * (kThreads / 32);
    for (int k = blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5); k < n; k += warps) {   // a warp per candidate
        const int i = list[k];
        if (status[i] != 1) continue;
        const int s = plm_decide_warp(d2, status, Y, X, d, i);
        if (s != 1 && (threadIdx.x & 31) == 0) status[i] = (uint8_t)s;
    }
}

// what the grid-wide rounds left (plateaus): one block, rounds separated by block barriers
__global__ void __launch_bounds__(1024) plm_finish_kernel(const int32_t* __restrict__ d2, uint8_t* status,
                                                          int32_t* __restrict__ list, int Y, int X, int d, PlmState* st) {
    __shared__ int s_n, s_left;
    const int n = st->n_cand;
    // compact the undecided ones to the front of the list
    if (threadIdx.x == 0) s_n = 0;
    __syncthreads();
    for (int base = 0; base < n; base += 1024) {
        const int k = base + threadIdx.x;
        const int i = k < n ? list[k] : -1;
        const bool open = i >= 0 && status[i] == 1;
        __syncthreads();  // everyone has read its slot before slots are overwritten (writes go to positions <= k)
        if (open) list[atomicAdd(&s_n, 1)] = i;
        __syncthreads();
    }
    int m = s_n;
    while (m > 0) {
        if (threadIdx.x == 0) s_left = 0;
        __syncthreads();
        for (int k = threadIdx.x >> 5; k < m; k += 32) {   // a warp per candidate
            const int i = list[k];
            if (((volatile uint8_t*)status)[i] != 1) continue;
            const int s = plm_decide_warp(d2, status, Y, X, d, i);
            if ((threadIdx.x & 31) == 0) {
                if (s != 1) {
                    status[i] = (uint8_t)s;
                    __threadfence_block();
                } else {
                    atomicAdd(&s_left, 1);
                }
            }
        }
        __syncthreads();
        m = s_left ? m : 0;
        __syncthreads();
    }
}

__global__ void __launch_bounds__(kThreads) plm_out_kernel(const uint8_t* __restrict__ status, uint8_t* __restrict__ peaks,
                                                           int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kThreads)
        peaks[i] = status[i] == 2;
}

// ---------------------------------------------------------------------------------------------------------------
// flood
// ---------------------------------------------------------------------------------------------------------------

constexpr int kPixBits = 27;  // queue entry = height key (32 bits) << 32 | pixel index (27 bits)
constexpr uint32_t kPixMask = (1u << kPixBits) - 1u;

__device__ __forceinline__ uint32_t height_key(int32_t h, int negate) {
    return (uint32_t)(negate ? -h : h) ^ 0x80000000u;  // unsigned order = signed order of the height
}

__global__ void __launch_bounds__(kThreads) ws_area_kernel(const int32_t* __restrict__ comp, const int32_t* __restrict__ markers,
                                                           int32_t* __restrict__ out, int64_t n, int32_t* __restrict__ area,
                                                           int32_t* __restrict__ n_seed, const int32_t* __restrict__ n_comp_p,
                                                           int64_t cap) {
    const int lane = threadIdx.x & 31;
    const int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    const bool in = i < n;
    const int32_t n_comp = (int32_t)min((int64_t)*n_comp_p, cap - 1);
    int32_t c = in ? comp[i] : 0;
    if (c > n_comp) c = 0;   // a region id beyond the stated count is treated as outside the mask (no table is that long)
    if (in) {
        const int32_t m = markers[i];
        out[i] = m;
        if (m != 0 && c > 0) atomicAdd(&n_seed[c], 1);
    }
    const int32_t cp = __shfl_up_sync(0xffffffffu, c, 1);
    const bool fg = c > 0;
    const bool start = fg && (lane == 0 || cp != c);
    const unsigned brks = __ballot_sync(0xffffffffu, !fg || start);
    if (!start) return;
    const unsigned stop = brks & (lane == 31 ? 0u : (0xffffffffu << (lane + 1)));
    atomicAdd(&area[c], (stop ? (__ffs(stop) - 1) : 32) - lane);  // one atomic per run of equal region ids
}

// exclusive prefix sum of area[1..n_comp] → offset[c] (one block; regions are few compared with pixels)
__global__ void __launch_bounds__(1024) ws_offsets_kernel(const int32_t* __restrict__ area, const int32_t* __restrict__ n_comp_p,
                                                          int64_t cap, int32_t* __restrict__ offset, int32_t* __restrict__ ticket) {
    __shared__ int s_warp[32];
    __shared__ int s_carry;
    const int n_comp = (int)min((int64_t)*n_comp_p, cap - 1);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) {
        s_carry = 0;
        *ticket = 0;
    }
    if (threadIdx.x < 8) ((unsigned long long*)(ticket + 2))[threadIdx.x] = 0;
    __syncthreads();
    for (int base = 1; base <= n_comp; base += 1024) {
        const int c = base + threadIdx.x;
        const int a = c <= n_comp ? area[c] : 0;
        int v = a;
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= o) v += t;
        }
        if (lane == 31) s_warp[warp] = v;
        __syncthreads();
        if (warp == 0) {
            int w = s_warp[lane];
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += t;
            }
            s_warp[lane] = w;
        }
        __syncthreads();
        const int excl = s_carry + (warp ? s_warp[warp - 1] : 0) + v - a;
        if (c <= n_comp) offset[c] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = excl + a;
        __syncthreads();
    }
}

__global__ void __launch_bounds__(kThreads) ws_seed_kernel(const int32_t* __restrict__ comp, const int32_t* __restrict__ markers,
                                                           const int32_t* __restrict__ height, int negate, int64_t n,
                                                           const int32_t* __restrict__ offset, int32_t* __restrict__ fill,
                                                           uint64_t* __restrict__ queue, const int32_t* __restrict__ n_comp_p,
                                                           int64_t cap) {
    const int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    if (i >= n) return;
    const int32_t c = comp[i];
    if (c <= 0 || c > (int32_t)min((int64_t)*n_comp_p, cap - 1) || markers[i] == 0) return;
    const int pos = atomicAdd(&fill[c], 1);
    queue[offset[c] + pos] = ((uint64_t)height_key(height[i], negate) << 32) | (uint32_t)i;
}

struct FloodArgs {
    const int32_t* comp;
    const int32_t* height;
    const int32_t* n_comp;
    const int32_t* offset;
    const int32_t* n_seed;
    int32_t* out;
    uint64_t* qa;
    uint64_t* qb;
    unsigned long long* rec;   // per pixel: height key << 32 | gone << 31 | push number (all ones: never queued)
    int32_t* ticket;
    int64_t cap;
    int Y, X, negate;
};

constexpr int kFloodThreads = 512;   // threads of a flood block: the passes over a long queue are per-thread loops

__device__ __forceinline__ unsigned long long block_scan_incl(unsigned long long v, unsigned long long* s_warp) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
    }
    if (lane == 31) s_warp[warp] = v;
    __syncthreads();
    unsigned long long add = 0;
    for (int w = 0; w < warp; ++w) add += s_warp[w];
    return v + add;
}

__global__ void __launch_bounds__(kFloodThreads, 2) ws_flood_kernel(FloodArgs a) {
    __shared__ int s_c, s_istar;
    __shared__ unsigned long long s_warp[kFloodThreads / 32], s_tot;
    __shared__ uint32_t s_hnext;
    const int tid = threadIdx.x;
    const int X = a.X, Y = a.Y;
    const int n_comp = (int)min((int64_t)*a.n_comp, a.cap - 1);
    const int dq[4] = {-X, -1, 1, X};
    for (;;) {
        __syncthreads();
        if (tid == 0) s_c = atomicAdd(a.ticket, 1) + 1;
        __syncthreads();
        const int c = s_c;
        if (c > n_comp) return;
        int m = a.n_seed[c];
        if (m == 0) continue;
        uint64_t* cur = a.qa + a.offset[c];
        uint64_t* nxt = a.qb + a.offset[c];
        // seeds arrived in any order: put them in raster order (their pixel index), find the lowest height.  Every pixel
        // that enters the queue gets a record (height key, push number): the push number is its place in the queue order
        // for good, whatever position it holds in the array
        if (tid == 0) s_hnext = 0xffffffffu;
        __syncthreads();
        for (int i = tid; i < m; i += kFloodThreads) {
            const uint64_t e = cur[i];
            const uint32_t p = (uint32_t)e & kPixMask;
            int rank = 0;
            for (int j = 0; j < m; ++j) rank += ((uint32_t)cur[j] & kPixMask) < p;
            nxt[rank] = e;
            a.rec[p] = (e & 0xffffffff00000000ull) | (unsigned)rank;
            atomicMin(&s_hnext, (uint32_t)(e >> 32));
        }
        __syncthreads();
        {
            uint64_t* t = cur;
            cur = nxt;
            nxt = t;
        }
        uint32_t step = 0, pushed = (uint32_t)m;   // push numbers handed out so far
        unsigned long long sum_m = 0, max_m = 0;
        while (m > 0) {
            sum_m += (unsigned long long)m;
            max_m = max(max_m, (unsigned long long)m);
            const uint32_t hmin = s_hnext;
            if (tid == 0) s_istar = INT32_MAX;
            __syncthreads();
            // A: every entry of the lowest height looks at its four neighbours and at THEIR other neighbours.  A free
            // neighbour q goes to the entry with the lowest push number among the lowest-height entries around q — the one
            // the sequential flood pops first — so an entry knows what it wins without asking anybody: all its loads (4 x
            // {region, label, height} + its own and 8 second-ring records) are issued together, one round trip to L2.
            // what the push phase (C) needs of this thread's first such entry — its label and its neighbours' heights — is
            // loaded here as well and kept in registers: most steps take a handful of entries, one per thread at most, and
            // C then costs no second round trip
            int st_i = -1, st_label = 0, st_h[4] = {0, 0, 0, 0};
            for (int i = tid; i < m; i += kFloodThreads) {
                const uint64_t e = cur[i];
                if ((uint32_t)(e >> 32) != hmin) continue;
                const int p = (int)((uint32_t)e & kPixMask);
                const int y = p / X, x = p - y * X;
                const bool ok[4] = {y > 0, x > 0, x + 1 < X, y + 1 < Y};
                int rc[4], ro[4], rh[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const int q = ok[k] ? p + dq[k] : p;
                    rc[k] = a.comp[q];
                    ro[k] = __ldcg(a.out + q);
                    rh[k] = a.height[q];
                }
                if (st_i < 0) {
                    st_i = i;
                    st_label = __ldcg(a.out + p);
#pragma unroll
                    for (int k = 0; k < 4; ++k) st_h[k] = rh[k];
                }
                // second ring: U2, UL, UR, L2, R2, DL, DR, D2
                const bool ok2[8] = {y > 1, ok[0] && ok[1], ok[0] && ok[2], x > 1, x + 2 < X, ok[3] && ok[1], ok[3] && ok[2], y + 2 < Y};
                const int d2[8] = {-2 * X, -X - 1, -X + 1, -2, 2, X - 1, X + 1, 2 * X};
                unsigned long long r2[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) r2[j] = ok2[j] ? __ldcg(a.rec + (p + d2[j])) : ~0ull;
                const uint32_t my = (uint32_t)__ldcg(a.rec + p) & 0x7fffffffu;
                // does a waiting lowest-height entry with an earlier push number sit on that pixel?
                auto beats = [&](unsigned long long r) {
                    return (uint32_t)(r >> 32) == hmin && !((uint32_t)r & 0x80000000u) && ((uint32_t)r & 0x7fffffffu) < my;
                };
                const bool lose[4] = {beats(r2[0]) || beats(r2[1]) || beats(r2[2]),    // up:    U2, UL, UR
                                      beats(r2[1]) || beats(r2[3]) || beats(r2[5]),    // left:  UL, L2, DL
                                      beats(r2[2]) || beats(r2[4]) || beats(r2[6]),    // right: UR, R2, DR
                                      beats(r2[5]) || beats(r2[6]) || beats(r2[7])};   // down:  DL, DR, D2
                unsigned won = 0x10u;   // bit 4: "is a lowest-height entry"; bits 0..3: neighbour k is mine
                bool lower = false;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    if (!ok[k] || rc[k] != c || ro[k] != 0) continue;
                    if (!lose[k]) won |= 1u << k;
                    lower |= height_key(rh[k], a.negate) < hmin;
                }
                cur[i] = e | ((uint64_t)won << kPixBits);
                if (lower) atomicMin(&s_istar, i);
            }
            __syncthreads();
            const int istar = s_istar;
            if (tid == 0) s_hnext = 0xffffffffu;
            // B: survivors keep their order at the front; the pixels won follow, ordered by (winner's position, neighbour).
            // A thread owns a contiguous stretch of the queue, so one block-wide scan per step places everything; a
            // leaving entry gets the position of its first push written over its height.
            const int L = (m + kFloodThreads - 1) / kFloodThreads;
            const int lo = min(tid * L, m), hi = min(lo + L, m);
            int kept = 0, push = 0;
            for (int i = lo; i < hi; ++i) {
                const unsigned flags = ((uint32_t)cur[i]) >> kPixBits;
                if (flags && i <= istar) push += __popc(flags & 0xfu);
                else ++kept;
            }
            const unsigned long long packed = (unsigned long long)kept | ((unsigned long long)push << 32);
            const unsigned long long incl = block_scan_incl(packed, s_warp);
            if (tid == kFloodThreads - 1) s_tot = incl;
            __syncthreads();
            const int kept_total = (int)(uint32_t)s_tot, push_total = (int)(s_tot >> 32);
            int kept_pos = (int)(uint32_t)(incl - packed), push_pos = kept_total + (int)((incl - packed) >> 32);
            uint32_t hnext = 0xffffffffu;
            for (int i = lo; i < hi; ++i) {
                const uint64_t e = cur[i];
                const unsigned flags = ((uint32_t)e) >> kPixBits;
                if (flags && i <= istar) {
                    cur[i] = ((uint64_t)(uint32_t)push_pos << 32) | (uint32_t)e;
                    push_pos += __popc(flags & 0xfu);
                } else {
                    nxt[kept_pos++] = e & ~((uint64_t)0x1fu << kPixBits);
                    hnext = min(hnext, (uint32_t)(e >> 32));
                }
            }
            __syncthreads();
            // C: the leaving entries are marked as gone, label and push what they won
            for (int i = tid; i < m && i <= istar; i += kFloodThreads) {
                const uint64_t e = cur[i];
                const unsigned flags = ((uint32_t)e) >> kPixBits;
                if (!flags) continue;
                const int p = (int)((uint32_t)e & kPixMask);
                a.rec[p] = ((unsigned long long)hmin << 32) | 0x80000000ull;
                if (!(flags & 0xfu)) continue;
                int pos = (int)(e >> 32);
                int32_t label = st_label;
                int rh[4] = {st_h[0], st_h[1], st_h[2], st_h[3]};
                if (i != st_i) {   // a thread's second, third ... entry of one step: loaded now
                    label = __ldcg(a.out + p);
#pragma unroll
                    for (int k = 0; k < 4; ++k) rh[k] = a.height[(flags & (1u << k)) ? p + dq[k] : p];
                }
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (flags & (1u << k)) {
                        const int q = p + dq[k];
                        const uint32_t hk = height_key(rh[k], a.negate);
                        a.out[q] = label;
                        a.rec[q] = ((unsigned long long)hk << 32) | (pushed + (uint32_t)(pos - kept_total));
                        nxt[pos++] = ((uint64_t)hk << 32) | (uint32_t)q;
                        hnext = min(hnext, hk);
                    }
            }
            hnext = __reduce_min_sync(0xffffffffu, hnext);
            if ((tid & 31) == 0) atomicMin(&s_hnext, hnext);
            __syncthreads();
            m = kept_total + push_total;
            pushed += (uint32_t)push_total;
            uint64_t* t = cur;
            cur = nxt;
            nxt = t;
            ++step;
        }
        if (tid == 0) {  // diagnostics next to the ticket: steps, entries walked, longest queue, most steps of one region
            unsigned long long* stats = (unsigned long long*)(a.ticket + 2);
            atomicAdd(&stats[0], (unsigned long long)step);
            atomicAdd(&stats[1], sum_m);
            atomicMax(&stats[2], max_m);
            atomicMax(&stats[3], (unsigned long long)step);
        }
    }
}

}  // namespace

extern "C" size_t be_peak_local_max_workspace_bytes(int64_t Y, int64_t X) {
    if (Y <= 0 || X <= 0) return 256;
    const size_t n = (size_t)Y * X;
    return 256 + be_align_up(4 * n, 256) + be_align_up(n, 256) + be_align_up(4 * n, 256);
}

extern "C" int be_peak_local_max_i32(const int32_t* d2, const uint8_t* mask, uint8_t* peaks, int64_t Y, int64_t X,
                                     int min_distance, void* ws, be_stream_t stream) {
    BE_REQUIRE(d2 && mask && peaks && ws, BE_ERR_INVALID_ARG, "be_peak_local_max_i32: null pointer");
    BE_REQUIRE(Y > 0 && X > 0, BE_ERR_INVALID_ARG, "be_peak_local_max_i32: empty input");
    BE_REQUIRE(Y * X < ((int64_t)1 << 31), BE_ERR_TOO_LARGE, "be_peak_local_max_i32: more than 2^31 pixels");
    BE_REQUIRE(min_distance >= 0 && min_distance <= 64, BE_ERR_INVALID_ARG, "be_peak_local_max_i32: min_distance %d not in 0..64",
               min_distance);
    cudaStream_t s = (cudaStream_t)stream;
    const int64_t n = Y * X;
    char* w = (char*)ws;
    PlmState* st = (PlmState*)w;
    w += 256;
    int32_t* rowmax = (int32_t*)w;
    w += be_align_up(4 * (size_t)n, 256);
    uint8_t* status = (uint8_t*)w;
    w += be_align_up((size_t)n, 256);
    int32_t* list = (int32_t*)w;
    const unsigned grid = (unsigned)min((n + kThreads - 1) / kThreads, (int64_t)BE_SM_COUNT * 8);
    const int d = min_distance;
    BE_CUDA(cudaMemsetAsync(status, 0, (size_t)n, s));
    plm_init_kernel<<<1, 1, 0, s>>>(st);
    BE_LAUNCH_CHECK();
    plm_roi_kernel<<<grid, kThreads, 0, s>>>(d2, mask, (int)Y, (int)X, d, st);
    BE_LAUNCH_CHECK();
    plm_rowmax_kernel<<<grid, kThreads, 0, s>>>(d2, mask, rowmax, (int)Y, (int)X, d, st);
    BE_LAUNCH_CHECK();
    plm_peak_kernel<<<grid, kThreads, 0, s>>>(d2, mask, rowmax, status, (int)Y, (int)X, d, st);
    BE_LAUNCH_CHECK();
    plm_trivial_kernel<<<grid, kThreads, 0, s>>>(d2, mask, status, (int)Y, (int)X, d, st);
    BE_LAUNCH_CHECK();
    plm_list_kernel<<<grid, kThreads, 0, s>>>(status, list, (int)X, st);
    BE_LAUNCH_CHECK();
    for (int r = 0; r < 2; ++r) {
        plm_round_kernel<<<grid, kThreads, 0, s>>>(d2, status, list, (int)Y, (int)X, d, st);
        BE_LAUNCH_CHECK();
    }
    plm_finish_kernel<<<1, 1024, 0, s>>>(d2, status, list, (int)Y, (int)X, d, st);
    BE_LAUNCH_CHECK();
    plm_out_kernel<<<grid, kThreads, 0, s>>>(status, peaks, n);
    BE_LAUNCH_CHECK();
    return BE_OK;
}

static int64_t ws_region_cap(int64_t n) { return (n + 1) / 2 + 2; }  // most 4-connected regions n pixels can hold, + id 0

extern "C" size_t be_watershed_workspace_bytes(int64_t Y, int64_t X) {
    if (Y <= 0 || X <= 0) return 256;
    const size_t n = (size_t)Y * X;
    const size_t cap = (size_t)ws_region_cap((int64_t)n);
    return 256 + 4 * be_align_up(4 * cap, 256) + 3 * be_align_up(8 * n, 256);
}

extern "C" int be_watershed_i32(const int32_t* height, int negate, const int32_t* markers, const int32_t* regions,
                                const int32_t* n_regions, int32_t* out, int64_t Y, int64_t X, void* ws, be_stream_t stream) {
    BE_REQUIRE(height && markers && regions && n_regions && out && ws, BE_ERR_INVALID_ARG, "be_watershed_i32: null pointer");
    BE_REQUIRE(Y > 0 && X > 0, BE_ERR_INVALID_ARG, "be_watershed_i32: empty input");
    BE_REQUIRE(Y * X <= (int64_t)kPixMask, BE_ERR_TOO_LARGE, "be_watershed_i32: more than 2^27 pixels");
    cudaStream_t s = (cudaStream_t)stream;
    const int64_t n = Y * X, cap = ws_region_cap(n);
    char* w = (char*)ws;
    int32_t* ticket = (int32_t*)w;
    w += 256;
    const size_t cap_bytes = be_align_up(4 * (size_t)cap, 256), q_bytes = be_align_up(8 * (size_t)n, 256);
    int32_t* area = (int32_t*)w;
    int32_t* n_seed = (int32_t*)(w + cap_bytes);
    int32_t* fill = (int32_t*)(w + 2 * cap_bytes);
    int32_t* offset = (int32_t*)(w + 3 * cap_bytes);
    w += 4 * cap_bytes;
    uint64_t* qa = (uint64_t*)w;
    uint64_t* qb = (uint64_t*)(w + q_bytes);
    unsigned long long* rec = (unsigned long long*)(w + 2 * q_bytes);
    BE_CUDA(cudaMemsetAsync(area, 0, 3 * cap_bytes, s));                 // area, n_seed, fill
    BE_CUDA(cudaMemsetAsync(rec, 0xff, 8 * (size_t)n, s));
    const unsigned gpx = (unsigned)((n + kThreads - 1) / kThreads);
    ws_area_kernel<<<gpx, kThreads, 0, s>>>(regions, markers, out, n, area, n_seed, n_regions, cap);
    BE_LAUNCH_CHECK();
    ws_offsets_kernel<<<1, 1024, 0, s>>>(area, n_regions, cap, offset, ticket);
    BE_LAUNCH_CHECK();
    ws_seed_kernel<<<gpx, kThreads, 0, s>>>(regions, markers, height, negate, n, offset, fill, qa, n_regions, cap);
    BE_LAUNCH_CHECK();
    FloodArgs a{regions, height, n_regions, offset, n_seed, out, qa, qb, rec, ticket, cap, (int)Y, (int)X, negate};
    ws_flood_kernel<<<BE_SM_COUNT * 2, kFloodThreads, 0, s>>>(a);
    BE_LAUNCH_CHECK();
    return BE_OK;
}

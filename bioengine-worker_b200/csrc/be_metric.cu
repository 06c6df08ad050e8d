// Screening metric and per-label statistics (integer work, HBM-bound).
//
//   be_label_overlap_i32   sparse (true, pred) pair histogram in an open-addressing table; one atomic per
//                          run of equal pairs (warp ballot run detection), areas/max labels from the table
//   be_match_counts        IoU >= thr pair counting with "second partner" detection (exact tp for thr >= 0.5)
//   be_label_stats_i32     exact integer raw moments + bounding boxes, one set of atomics per x-run
#include "be_common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr unsigned long long kEmpty = ~0ull;

__device__ __forceinline__ unsigned long long mix64(unsigned long long k) {
    k ^= k >> 33;
    k *= 0xff51afd7ed558ccdull;
    k ^= k >> 33;
    k *= 0xc4ceb9fe1a85ec53ull;
    k ^= k >> 33;
    return k;
}

__device__ __forceinline__ void table_add(unsigned long long* keys, unsigned* counts, int64_t capacity,
                                          unsigned long long key, unsigned cnt, int32_t* status) {
    const unsigned long long maskc = (unsigned long long)capacity - 1ull;
    unsigned long long slot = mix64(key) & maskc;
    for (int64_t probe = 0; probe < capacity; ++probe) {
        unsigned long long old = keys[slot];
        if (old != key) {
            if (old != kEmpty) {
                slot = (slot + 1) & maskc;
                continue;
            }
            old = atomicCAS(&keys[slot], kEmpty, key);
            if (old != kEmpty && old != key) {
                slot = (slot + 1) & maskc;
                continue;
            }
        }
        atomicAdd(&counts[slot], cnt);
        return;
    }
    *status = 1;  // table full
}

// ---- label-pair overlap ----
//
// One CTA owns a 128 x 128 px tile of the image pair.  Each lane owns 4 consecutive pixels of a row; a lane whose 4 pairs
// are equal joins a warp-level run with its neighbours, mixed lanes split into sub-runs.  Runs are accumulated in a
// shared-memory hash table first: a component is a compact blob, so a tile sees each (true, pred) pair many times (once
// per row it crosses) but holds only a few hundred distinct pairs, and the global atomics shrink from one per run to one
// per (tile, pair).  Background-on-background runs are only counted.  At the end of the tile every entry adds its count to
// the two per-label areas, and entries with both labels > 0 also go to the image's global pair table (only those pairs
// can match; the areas carry what the IoU needs from the background row and column).

constexpr int kOvTile = 128;        // tile edge in pixels
constexpr int kOvSlots = 2048;      // shared-memory table slots (24 KB)
constexpr int kOvProbes = 48;       // give up on the shared table after this many probes: straight to global

// How much of an image's tables a call uses: the host passes the allocated sizes; if per-image label bounds are known on
// the device (e.g. the component counts the labelling kernel just wrote) the pair table is cut to a power of two that
// fits them, so clearing and scanning cost what the image needs, not what the worst image would.
struct OvSizes {
    int64_t capacity;   // pair-table slots in use (power of two, <= allocated)
    int64_t n_true, n_pred;   // area entries in use (<= cap_true / cap_pred)
};
__device__ __forceinline__ OvSizes ov_sizes(const int32_t* __restrict__ bounds, int64_t img, int64_t capacity,
                                            int64_t cap_true, int64_t cap_pred) {
    OvSizes z{capacity, cap_true, cap_pred};
    if (bounds) {
        const int64_t bt = min((int64_t)max(bounds[img * 2 + 0], 0) + 1, cap_true);
        const int64_t bp = min((int64_t)max(bounds[img * 2 + 1], 0) + 1, cap_pred);
        z.n_true = bt;
        z.n_pred = bp;
        const unsigned long long want = 4ull * (unsigned long long)(bt + bp) + 64ull;
        int64_t c = 1ll << (64 - __clzll((long long)(want - 1)));
        z.capacity = min(c, capacity);
    }
    return z;
}

__global__ void __launch_bounds__(kThreads) overlap_init_kernel(unsigned long long* __restrict__ keys,
                                                                unsigned* __restrict__ counts, int64_t capacity,
                                                                int32_t* __restrict__ area_true, int64_t cap_true,
                                                                int32_t* __restrict__ area_pred, int64_t cap_pred,
                                                                int32_t* __restrict__ max_lbl, int32_t* __restrict__ status,
                                                                const int32_t* __restrict__ bounds) {
    const int64_t img = blockIdx.y;
    be_pdl_wait();   // the label counts in `bounds` may come from the kernel just before
    const OvSizes z = ov_sizes(bounds, img, capacity, cap_true, cap_pred);
    const int64_t stride = (int64_t)gridDim.x * kThreads;
    const int64_t i0 = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    for (int64_t i = i0; i < z.capacity; i += stride) {
        keys[img * capacity + i] = kEmpty;
        counts[img * capacity + i] = 0;
    }
    for (int64_t i = i0; i < z.n_true; i += stride) area_true[img * cap_true + i] = 0;
    for (int64_t i = i0; i < z.n_pred; i += stride) area_pred[img * cap_pred + i] = 0;
    if (i0 == 0) {
        max_lbl[img * 2] = max_lbl[img * 2 + 1] = 0;
        status[img] = 0;
    }
}

struct OvGlobal {
    unsigned long long* K;
    unsigned* Cn;
    int64_t capacity;
    int32_t *area_true, *area_pred;
    int64_t n_true, n_pred;
    int32_t* st;
};

// one (pair, count) into the global structures; (t, 0) / (0, p) only touch the areas (background areas come per CTA)
__device__ __forceinline__ void ov_global_add(const OvGlobal& G, unsigned long long key, unsigned cnt, unsigned* bg_true,
                                              unsigned* bg_pred) {
    const int64_t t = (int64_t)(key >> 32), p = (int64_t)(key & 0xffffffffull);
    if (t >= G.n_true || p >= G.n_pred) {
        *G.st = 2;
        return;
    }
    if (t) atomicAdd(&G.area_true[t], (int)cnt); else atomicAdd(bg_true, cnt);
    if (p) atomicAdd(&G.area_pred[p], (int)cnt); else atomicAdd(bg_pred, cnt);
    if (t && p) table_add(G.K, G.Cn, G.capacity, key, cnt, G.st);
}

// the CTA's shared table: a cheap 32-bit hash of the two labels, the first probe is the claim itself (no look first)
__device__ __forceinline__ void ov_add(unsigned long long* sk, unsigned* sc, const OvGlobal& G, unsigned long long key,
                                       unsigned cnt, unsigned* bg_true, unsigned* bg_pred) {
    const unsigned h = ((unsigned)(key >> 32) * 0x9E3779B1u) ^ ((unsigned)key * 0x85EBCA6Bu);
    unsigned slot = h >> (32 - 11);
    static_assert(kOvSlots == (1 << 11), "slot bits");
#pragma unroll 1
    for (int probe = 0; probe < kOvProbes; ++probe) {
        const unsigned long long old = atomicCAS(&sk[slot], kEmpty, key);
        if (old == kEmpty || old == key) {
            atomicAdd(&sc[slot], cnt);
            return;
        }
        slot = (slot + 1) & (kOvSlots - 1);
    }
    ov_global_add(G, key, cnt, bg_true, bg_pred);   // crowded tile (far more than ~1500 distinct pairs): bypass
}

__global__ void __launch_bounds__(kThreads) overlap_kernel(const int32_t* __restrict__ lab_true,
                                                           const int32_t* __restrict__ lab_pred, int64_t n_px, int64_t X,
                                                           int tiles_x, unsigned long long* __restrict__ keys,
                                                           unsigned* __restrict__ counts, int64_t capacity,
                                                           int32_t* __restrict__ area_true, int64_t cap_true,
                                                           int32_t* __restrict__ area_pred, int64_t cap_pred,
                                                           int32_t* __restrict__ max_lbl, int32_t* __restrict__ status,
                                                           const int32_t* __restrict__ bounds) {
    __shared__ unsigned long long sk[kOvSlots];
    __shared__ unsigned sc[kOvSlots];
    __shared__ unsigned s_bg[3];   // pixels of (0,0), of (t,0) [pred background], of (0,p) [true background]
    __shared__ int s_max[2];
    const int64_t img = blockIdx.y;
    const int32_t* t = lab_true + img * n_px;
    const int32_t* p = lab_pred + img * n_px;
    be_pdl_wait();   // the cleared tables
    const OvSizes z = ov_sizes(bounds, img, capacity, cap_true, cap_pred);
    OvGlobal G{keys + img * capacity, counts + img * capacity, z.capacity, area_true + img * cap_true,
               area_pred + img * cap_pred, z.n_true, z.n_pred, status + img};
    for (int i = threadIdx.x; i < kOvSlots; i += kThreads) {
        sk[i] = kEmpty;
        sc[i] = 0;
    }
    if (threadIdx.x < 3) s_bg[threadIdx.x] = 0;
    if (threadIdx.x < 2) s_max[threadIdx.x] = 0;
    __syncthreads();

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t Y = n_px / X;
    const int tx = blockIdx.x % tiles_x, ty = blockIdx.x / tiles_x;
    const int64_t x = (int64_t)tx * kOvTile + lane * 4;
    const int64_t y_end = min((int64_t)(ty + 1) * kOvTile, Y);
    const bool vec_ok = ((((uintptr_t)t | (uintptr_t)p) & 15) == 0) && ((X & 3) == 0);
    unsigned bg00 = 0;
    int mt = 0, mp = 0;
    for (int64_t y = (int64_t)ty * kOvTile + warp; y < y_end; y += kThreads / 32) {
        unsigned long long k[4];
        int cnt = 0;
        if (x < X) {
            const int64_t b = y * X + x;
            int32_t tv[4], pv[4];
            if (vec_ok && x + 4 <= X) {
                const int4 a = __ldg(reinterpret_cast<const int4*>(t + b));
                const int4 c = __ldg(reinterpret_cast<const int4*>(p + b));
                tv[0] = a.x; tv[1] = a.y; tv[2] = a.z; tv[3] = a.w;
                pv[0] = c.x; pv[1] = c.y; pv[2] = c.z; pv[3] = c.w;
                cnt = 4;
            } else {
                for (int j = 0; j < 4 && x + j < X; ++j) {
                    tv[j] = t[b + j];
                    pv[j] = p[b + j];
                    cnt = j + 1;
                }
            }
            for (int j = 0; j < cnt; ++j) {
                if ((tv[j] | pv[j]) < 0) {
                    *G.st = 2;
                    tv[j] = pv[j] = 0;
                }
                mt = max(mt, tv[j]);
                mp = max(mp, pv[j]);
                k[j] = ((unsigned long long)(unsigned)tv[j] << 32) | (unsigned)pv[j];
            }
        }
        const bool uniform = (cnt == 4) && k[0] == k[1] && k[1] == k[2] && k[2] == k[3];
        const unsigned long long k0 = uniform ? k[0] : kEmpty;
        const unsigned long long prev = __shfl_up_sync(0xffffffffu, k0, 1);
        const bool start = uniform && (lane == 0 || prev != k0);
        const unsigned starts = __ballot_sync(0xffffffffu, start);
        const unsigned unif = __ballot_sync(0xffffffffu, uniform);
        if (start) {
            // run = consecutive uniform lanes up to the next start or the first non-uniform lane
            const unsigned stop = (starts | ~unif) & (lane == 31 ? 0u : (0xffffffffu << (lane + 1)));
            const int end = stop ? (__ffs(stop) - 1) : 32;
            const unsigned c = 4u * (unsigned)(end - lane);
            if (k0 == 0ull) bg00 += c; else ov_add(sk, sc, G, k0, c, &s_bg[2], &s_bg[1]);
        } else if (!uniform && cnt > 0) {
            unsigned long long cur = k[0];
            unsigned c = 1;
            for (int j = 1; j < cnt; ++j) {
                if (k[j] == cur) {
                    ++c;
                } else {
                    if (cur == 0ull) bg00 += c; else ov_add(sk, sc, G, cur, c, &s_bg[2], &s_bg[1]);
                    cur = k[j];
                    c = 1;
                }
            }
            if (cur == 0ull) bg00 += c; else ov_add(sk, sc, G, cur, c, &s_bg[2], &s_bg[1]);
        }
    }
    bg00 = __reduce_add_sync(0xffffffffu, bg00);
    mt = __reduce_max_sync(0xffffffffu, mt);
    mp = __reduce_max_sync(0xffffffffu, mp);
    if (lane == 0) {
        if (bg00) atomicAdd(&s_bg[0], bg00);
        if (mt) atomicMax(&s_max[0], mt);
        if (mp) atomicMax(&s_max[1], mp);
    }
    __syncthreads();
    // the tile's pairs leave for the global structures: one area update per label and pair, one table update per pair
    for (int i = threadIdx.x; i < kOvSlots; i += kThreads) {
        const unsigned long long key = sk[i];
        if (key != kEmpty) ov_global_add(G, key, sc[i], &s_bg[2], &s_bg[1]);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        // background row / column of the overlap matrix: area_true[0] = (0,0) + (0,p), area_pred[0] = (0,0) + (t,0)
        const unsigned a0 = s_bg[0] + s_bg[2], b0 = s_bg[0] + s_bg[1];
        if (a0) atomicAdd(&G.area_true[0], (int)a0);
        if (b0) atomicAdd(&G.area_pred[0], (int)b0);
        if (s_max[0]) atomicMax(&max_lbl[img * 2 + 0], s_max[0]);
        if (s_max[1]) atomicMax(&max_lbl[img * 2 + 1], s_max[1]);
    }
}

__global__ void __launch_bounds__(kThreads) match_kernel(const unsigned long long* __restrict__ keys,
                                                         const unsigned* __restrict__ counts, int64_t capacity,
                                                         const int32_t* __restrict__ area_true, int64_t cap_true,
                                                         const int32_t* __restrict__ area_pred, int64_t cap_pred,
                                                         const double* __restrict__ thr, int n_thr,
                                                         unsigned long long* __restrict__ out,
                                                         unsigned* __restrict__ bits, int64_t words_true,
                                                         int64_t words_pred, const int32_t* __restrict__ bounds) {
    const int64_t img = blockIdx.y;
    const int64_t used = ov_sizes(bounds, img, capacity, cap_true, cap_pred).capacity;
    for (int64_t s = (int64_t)blockIdx.x * kThreads + threadIdx.x; s < used; s += (int64_t)gridDim.x * kThreads) {
        const unsigned long long k = keys[img * capacity + s];
        if (k == kEmpty) continue;
        const int64_t t = (int64_t)(k >> 32), p = (int64_t)(k & 0xffffffffull);
        if (t == 0 || p == 0 || t >= cap_true || p >= cap_pred) continue;
        const unsigned c = counts[img * capacity + s];
        const long long uni = (long long)area_true[img * cap_true + t] + (long long)area_pred[img * cap_pred + p] - (long long)c;
        // the reference's comparison: fp64 quotient against the fp64 threshold
        const double iou = __ddiv_rn((double)c, (double)uni);
        for (int j = 0; j < n_thr; ++j) {
            if (iou >= thr[j]) {
                unsigned long long* o = out + (img * n_thr + j) * 2;
                atomicAdd(&o[0], 1ull);
                unsigned* bt = bits + (img * n_thr + j) * (words_true + words_pred);
                unsigned* bp = bt + words_true;
                const unsigned mt = 1u << (t & 31), mp = 1u << (p & 31);
                unsigned extra = 0;
                if (atomicOr(&bt[t >> 5], mt) & mt) ++extra;
                if (atomicOr(&bp[p >> 5], mp) & mp) ++extra;
                if (extra) atomicAdd(&o[1], (unsigned long long)extra);
            }
        }
    }
}

// ---- screening table rows ----
//
// One thread per (image, column): turns a batch's {pairs, extra} counts and max labels into rows of the screening
// table [n_images][n_thr + 2] = {tp per threshold, n_true, n_pred} at the batch's image indices, and folds the batch's
// status words into two sticky flags, so a shard runs without the host looking at anything until the table is merged.
__global__ void screen_rows_kernel(const unsigned long long* __restrict__ counts, const int32_t* __restrict__ max_lbl,
                                   const int32_t* __restrict__ status, int64_t N, int n_thr,
                                   const double* __restrict__ thr, const int64_t* __restrict__ rows,
                                   long long* __restrict__ table, int32_t* __restrict__ flags) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int cols = n_thr + 2;
    if (i >= N * cols) return;
    const int64_t img = i / cols;
    const int k = (int)(i - img * cols);
    long long v;
    if (k < n_thr) {
        const unsigned long long pairs = counts[(img * n_thr + k) * 2], extra = counts[(img * n_thr + k) * 2 + 1];
        v = (long long)(pairs - extra);
        if (extra && thr[k] < 0.5) atomicOr(&flags[1], 1);   // below 0.5 the assignment is not implied by the counts
    } else {
        v = max_lbl[img * 2 + (k - n_thr)];
        if (k == n_thr && status[img]) atomicMax(&flags[0], status[img]);
    }
    table[rows[img] * cols + k] = v;
}

// ---- label statistics ----

__device__ __forceinline__ unsigned long long series_sum(unsigned long long a, unsigned long long b) {
    return (a + b) * (b - a + 1) / 2;  // a + (a+1) + ... + b
}
__device__ __forceinline__ unsigned long long squares_below(unsigned long long n) {
    return n ? (n - 1) * n * (2 * n - 1) / 6 : 0;  // 0^2 + ... + (n-1)^2
}

__global__ void __launch_bounds__(kThreads) label_stats_kernel(const int32_t* __restrict__ labels, int64_t Z, int64_t Y,
                                                               int64_t X, int64_t cap,
                                                               unsigned long long* __restrict__ sums,
                                                               int32_t* __restrict__ bbox, int32_t* __restrict__ status) {
    const int64_t img = blockIdx.y;
    const int64_t vol = Z * Y * X;
    const int32_t* lab = labels + img * vol;
    const int lane = threadIdx.x & 31;
    const int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    // whole warps stay alive so the shuffles below are full-mask
    const bool in = i < vol;
    const int32_t l = in ? lab[i] : 0;
    const int64_t rowi = in ? i / X : -1;
    const int x = in ? (int)(i - rowi * X) : 0;
    const int32_t lp = __shfl_up_sync(0xffffffffu, l, 1);
    const int64_t rp = __shfl_up_sync(0xffffffffu, rowi, 1);
    const bool fg = in && l > 0;
    const bool start = fg && (lane == 0 || lp != l || rp != rowi);
    const bool brk = !fg || start;  // anything that ends the run of the lane before it
    const unsigned brks = __ballot_sync(0xffffffffu, brk);
    if (!start) return;
    if (l >= cap) {
        status[img] = 2;
        return;
    }
    const unsigned stop = brks & (lane == 31 ? 0u : (0xffffffffu << (lane + 1)));
    const int end = stop ? (__ffs(stop) - 1) : 32;
    const unsigned long long len = (unsigned long long)(end - lane);
    const unsigned long long xa = (unsigned long long)x, xb = xa + len - 1;
    const unsigned long long y = (unsigned long long)(rowi % Y), z = (unsigned long long)(rowi / Y);
    const unsigned long long sx = series_sum(xa, xb);
    const unsigned long long sxx = squares_below(xb + 1) - squares_below(xa);
    unsigned long long* S = sums + (img * cap + l) * 10;
    atomicAdd(&S[0], len);
    if (z) atomicAdd(&S[1], z * len);
    atomicAdd(&S[2], y * len);
    atomicAdd(&S[3], sx);
    if (z) atomicAdd(&S[4], z * z * len);
    atomicAdd(&S[5], y * y * len);
    atomicAdd(&S[6], sxx);
    if (z) atomicAdd(&S[7], z * y * len);
    if (z) atomicAdd(&S[8], z * sx);
    atomicAdd(&S[9], y * sx);
    int32_t* B = bbox + (img * cap + l) * 6;
    atomicMin(&B[0], (int)z);
    atomicMin(&B[1], (int)y);
    atomicMin(&B[2], (int)xa);
    atomicMax(&B[3], (int)z);
    atomicMax(&B[4], (int)y);
    atomicMax(&B[5], (int)xb);
}

__global__ void bbox_init_kernel(int32_t* bbox, int64_t n_labels_total) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_labels_total) return;
    int32_t* B = bbox + i * 6;
    B[0] = B[1] = B[2] = 0x7fffffff;
    B[3] = B[4] = B[5] = -1;
}

// ---- pixel-level binary metrics ----

__device__ __forceinline__ float atomic_min_f32(float* addr, float v) {
    // ordered-int trick: valid for any non-NaN floats
    return (v >= 0.f) ? __int_as_float(atomicMin((int*)addr, __float_as_int(v)))
                      : __uint_as_float(atomicMax((unsigned*)addr, __float_as_uint(v)));
}
__device__ __forceinline__ float atomic_max_f32(float* addr, float v) {
    return (v >= 0.f) ? __int_as_float(atomicMax((int*)addr, __float_as_int(v)))
                      : __uint_as_float(atomicMin((unsigned*)addr, __float_as_uint(v)));
}

__global__ void __launch_bounds__(kThreads) minmax_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                                          int64_t n, float* __restrict__ out /* amin amax bmin bmax */) {
    float lo_a = __int_as_float(0x7f800000), hi_a = __int_as_float(0xff800000), lo_b = lo_a, hi_b = hi_a;
    for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kThreads) {
        const float x = a[i], y = b[i];
        lo_a = fminf(lo_a, x);
        hi_a = fmaxf(hi_a, x);
        lo_b = fminf(lo_b, y);
        hi_b = fmaxf(hi_b, y);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lo_a = fminf(lo_a, __shfl_xor_sync(0xffffffffu, lo_a, o));
        hi_a = fmaxf(hi_a, __shfl_xor_sync(0xffffffffu, hi_a, o));
        lo_b = fminf(lo_b, __shfl_xor_sync(0xffffffffu, lo_b, o));
        hi_b = fmaxf(hi_b, __shfl_xor_sync(0xffffffffu, hi_b, o));
    }
    if ((threadIdx.x & 31) == 0) {
        atomic_min_f32(&out[0], lo_a);
        atomic_max_f32(&out[1], hi_a);
        atomic_min_f32(&out[2], lo_b);
        atomic_max_f32(&out[3], hi_b);
    }
}

__global__ void minmax_init_kernel(float* out) {
    out[0] = out[2] = __int_as_float(0x7f800000);
    out[1] = out[3] = __int_as_float(0xff800000);
}

// thresholds chosen on the device exactly like the reference: 0.5 if the value range lies inside [0, 1], else 0.0
__global__ void __launch_bounds__(kThreads) confusion_kernel(const float* __restrict__ pred, const float* __restrict__ target,
                                                             int64_t n, const float* __restrict__ mm, int adaptive,
                                                             float thr_pred, float thr_target,
                                                             unsigned long long* __restrict__ counts /* tp fp fn tn */) {
    if (adaptive) {
        thr_pred = (mm[0] >= 0.f && mm[1] <= 1.f) ? 0.5f : 0.0f;
        thr_target = (mm[2] >= 0.f && mm[3] <= 1.f) ? 0.5f : 0.0f;
    }
    unsigned tp = 0, fp = 0, fn = 0, tn = 0;
    for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kThreads) {
        const bool p = pred[i] > thr_pred, t = target[i] > thr_target;
        tp += p & t;
        fp += p & !t;
        fn += !p & t;
        tn += !p & !t;
    }
    tp = __reduce_add_sync(0xffffffffu, tp);
    fp = __reduce_add_sync(0xffffffffu, fp);
    fn = __reduce_add_sync(0xffffffffu, fn);
    tn = __reduce_add_sync(0xffffffffu, tn);
    if ((threadIdx.x & 31) == 0) {
        if (tp) atomicAdd(&counts[0], (unsigned long long)tp);
        if (fp) atomicAdd(&counts[1], (unsigned long long)fp);
        if (fn) atomicAdd(&counts[2], (unsigned long long)fn);
        if (tn) atomicAdd(&counts[3], (unsigned long long)tn);
    }
}

inline bool is_pow2(int64_t v) { return v > 0 && (v & (v - 1)) == 0; }

}  // namespace

extern "C" int be_label_overlap_i32(const int32_t* lab_true, const int32_t* lab_pred, int64_t N, int64_t n_px,
                                    int64_t row_len, uint64_t* keys, uint32_t* counts, int64_t capacity,
                                    int32_t* area_true, int64_t cap_true, int32_t* area_pred, int64_t cap_pred,
                                    int32_t* max_lbl, int32_t* status, const int32_t* label_bounds, be_stream_t stream) {
    BE_REQUIRE(lab_true && lab_pred && keys && counts && area_true && area_pred && max_lbl && status, BE_ERR_INVALID_ARG,
               "be_label_overlap_i32: null pointer");
    BE_REQUIRE(N > 0 && n_px > 0, BE_ERR_INVALID_ARG, "be_label_overlap_i32: empty input");
    BE_REQUIRE(is_pow2(capacity), BE_ERR_INVALID_ARG, "be_label_overlap_i32: capacity %lld is not a power of two",
               (long long)capacity);
    BE_REQUIRE(cap_true > 0 && cap_pred > 0, BE_ERR_INVALID_ARG, "be_label_overlap_i32: label capacities must be positive");
    BE_REQUIRE(N <= 65535, BE_ERR_TOO_LARGE, "be_label_overlap_i32: N %lld > 65535", (long long)N);
    BE_REQUIRE(n_px < ((int64_t)1 << 31), BE_ERR_TOO_LARGE, "be_label_overlap_i32: n_px >= 2^31");
    BE_REQUIRE(row_len >= 0 && (row_len == 0 || n_px % row_len == 0), BE_ERR_INVALID_ARG,
               "be_label_overlap_i32: row_len %lld does not divide n_px %lld", (long long)row_len, (long long)n_px);
    cudaStream_t s = (cudaStream_t)stream;
    // row_len 0 = no row structure known: rows of 128 px (the last one ragged), i.e. flat chunks of 16 K pixels per CTA
    int64_t X = row_len > 0 ? row_len : kOvTile;
    int64_t Y = (n_px + X - 1) / X;
    if (row_len == 0 && n_px % X) {   // ragged tail: fall back to one long row per image
        X = n_px;
        Y = 1;
    }
    const int64_t tiles_x = (X + kOvTile - 1) / kOvTile, tiles_y = (Y + kOvTile - 1) / kOvTile;
    BE_REQUIRE(tiles_x * tiles_y < ((int64_t)1 << 31), BE_ERR_TOO_LARGE, "be_label_overlap_i32: too many tiles");
    {
        int64_t most = capacity > cap_pred ? capacity : cap_pred;
        if (cap_true > most) most = cap_true;
        int64_t gx = (most + kThreads * 8 - 1) / (kThreads * 8);
        const int64_t cap_blocks = max((int64_t)1, (int64_t)BE_SM_COUNT * 8 / N);
        if (gx > cap_blocks) gx = cap_blocks;
        BE_CUDA(be_launch_pdl(overlap_init_kernel, dim3((unsigned)gx, (unsigned)N), dim3(kThreads), 0, s,
                              (unsigned long long*)keys, counts, capacity, area_true, cap_true, area_pred, cap_pred, max_lbl,
                              status, label_bounds));
    }
    BE_CUDA(be_launch_pdl(overlap_kernel, dim3((unsigned)(tiles_x * tiles_y), (unsigned)N), dim3(kThreads), 0, s, lab_true,
                          lab_pred, n_px, X, (int)tiles_x, (unsigned long long*)keys, counts, capacity, area_true, cap_true,
                          area_pred, cap_pred, max_lbl, status, label_bounds));
    return BE_OK;
}

extern "C" size_t be_match_workspace_bytes(int64_t N, int64_t cap_true, int64_t cap_pred, int n_thr) {
    if (N <= 0 || cap_true <= 0 || cap_pred <= 0 || n_thr <= 0) return 256;
    const int64_t words = (cap_true + 31) / 32 + (cap_pred + 31) / 32;
    return be_align_up(sizeof(uint32_t) * N * n_thr * words, 256);
}

extern "C" int be_match_counts(const uint64_t* keys, const uint32_t* counts, int64_t N, int64_t capacity,
                               const int32_t* area_true, int64_t cap_true, const int32_t* area_pred, int64_t cap_pred,
                               const double* thr, int n_thr, int64_t* counts_out, void* ws, const int32_t* label_bounds,
                               be_stream_t stream) {
    BE_REQUIRE(keys && counts && area_true && area_pred && thr && counts_out && ws, BE_ERR_INVALID_ARG,
               "be_match_counts: null pointer");
    BE_REQUIRE(N > 0 && is_pow2(capacity) && cap_true > 0 && cap_pred > 0, BE_ERR_INVALID_ARG,
               "be_match_counts: bad sizes");
    BE_REQUIRE(n_thr >= 1 && n_thr <= 16, BE_ERR_INVALID_ARG, "be_match_counts: n_thr %d not in 1..16", n_thr);
    BE_REQUIRE(N <= 65535, BE_ERR_TOO_LARGE, "be_match_counts: N %lld > 65535", (long long)N);
    cudaStream_t s = (cudaStream_t)stream;
    const int64_t wt = (cap_true + 31) / 32, wp = (cap_pred + 31) / 32;
    BE_CUDA(cudaMemsetAsync(ws, 0, sizeof(uint32_t) * N * n_thr * (wt + wp), s));
    BE_CUDA(cudaMemsetAsync(counts_out, 0, sizeof(int64_t) * N * n_thr * 2, s));
    int64_t gx = (capacity + kThreads - 1) / kThreads;
    const int64_t most = max((int64_t)1, (int64_t)BE_SM_COUNT * 16 / N);   // the used part of a table is usually far shorter
    if (gx > most) gx = most;
    match_kernel<<<dim3((unsigned)gx, (unsigned)N), kThreads, 0, s>>>(
        (const unsigned long long*)keys, counts, capacity, area_true, cap_true, area_pred, cap_pred, thr, n_thr,
        (unsigned long long*)counts_out, (unsigned*)ws, wt, wp, label_bounds);
    BE_LAUNCH_CHECK();
    return BE_OK;
}

extern "C" int be_screen_rows_i64(const int64_t* counts, const int32_t* max_lbl, const int32_t* status, int64_t N,
                                  int n_thr, const double* thr, const int64_t* rows, int64_t* table, int32_t* flags,
                                  be_stream_t stream) {
    BE_REQUIRE(counts && max_lbl && status && thr && rows && table && flags, BE_ERR_INVALID_ARG,
               "be_screen_rows_i64: null pointer");
    BE_REQUIRE(N > 0 && n_thr >= 1 && n_thr <= 16, BE_ERR_INVALID_ARG, "be_screen_rows_i64: bad sizes");
    const int64_t n = N * (n_thr + 2);
    screen_rows_kernel<<<(unsigned)((n + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
        (const unsigned long long*)counts, max_lbl, status, N, n_thr, thr, rows, (long long*)table, flags);
    BE_LAUNCH_CHECK();
    return BE_OK;
}

extern "C" int be_label_stats_i32(const int32_t* labels, int64_t N, int64_t Z, int64_t Y, int64_t X, int64_t cap,
                                  int64_t* sums, int32_t* bbox, int32_t* status, be_stream_t stream) {
    BE_REQUIRE(labels && sums && bbox && status, BE_ERR_INVALID_ARG, "be_label_stats_i32: null pointer");
    BE_REQUIRE(N > 0 && Z > 0 && Y > 0 && X > 0 && cap > 0, BE_ERR_INVALID_ARG, "be_label_stats_i32: empty input");
    BE_REQUIRE(N <= 65535, BE_ERR_TOO_LARGE, "be_label_stats_i32: N %lld > 65535", (long long)N);
    const int64_t vol = Z * Y * X;
    BE_REQUIRE(vol < ((int64_t)1 << 31), BE_ERR_TOO_LARGE, "be_label_stats_i32: image >= 2^31 voxels");
    cudaStream_t s = (cudaStream_t)stream;
    BE_CUDA(cudaMemsetAsync(sums, 0, sizeof(int64_t) * N * cap * 10, s));
    BE_CUDA(cudaMemsetAsync(status, 0, sizeof(int32_t) * N, s));
    bbox_init_kernel<<<(unsigned)((N * cap + 255) / 256), 256, 0, s>>>(bbox, N * cap);
    BE_LAUNCH_CHECK();
    label_stats_kernel<<<dim3((unsigned)((vol + kThreads - 1) / kThreads), (unsigned)N), kThreads, 0, s>>>(
        labels, Z, Y, X, cap, (unsigned long long*)sums, bbox, status);
    BE_LAUNCH_CHECK();
    return BE_OK;
}

extern "C" int be_binary_confusion_f32(const float* pred, const float* target, int64_t n, int adaptive, float thr_pred,
                                       float thr_target, float* minmax, int64_t* counts, be_stream_t stream) {
    BE_REQUIRE(pred && target && minmax && counts, BE_ERR_INVALID_ARG, "be_binary_confusion_f32: null pointer");
    BE_REQUIRE(n > 0, BE_ERR_INVALID_ARG, "be_binary_confusion_f32: empty input");
    cudaStream_t s = (cudaStream_t)stream;
    int64_t blocks = (n + kThreads * 8 - 1) / (kThreads * 8);
    if (blocks > BE_SM_COUNT * 16) blocks = BE_SM_COUNT * 16;
    minmax_init_kernel<<<1, 1, 0, s>>>(minmax);
    BE_LAUNCH_CHECK();
    minmax_kernel<<<(unsigned)blocks, kThreads, 0, s>>>(pred, target, n, minmax);
    BE_LAUNCH_CHECK();
    BE_CUDA(cudaMemsetAsync(counts, 0, sizeof(int64_t) * 4, s));
    confusion_kernel<<<(unsigned)blocks, kThreads, 0, s>>>(pred, target, n, minmax, adaptive, thr_pred, thr_target,
                                                          (unsigned long long*)counts);
    BE_LAUNCH_CHECK();
    return BE_OK;
}

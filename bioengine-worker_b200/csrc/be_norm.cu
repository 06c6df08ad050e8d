// Pre-/post-processing statistics and elementwise kernels (HBM-bound; no tensor cores by design).
//
//   be_meanvar_f32     one pass, shifted fp64 sums, deterministic two-stage reduction
//   be_quantile_f32    exact order statistics by 3-level (11/11/10 bit) radix select + numpy's linear lerp
//   be_normalize_f32   (x - sub) / div in fp64, optional clip
//   be_elementwise_f32 clip / scale_linear / sigmoid
//   be_binarize_f32_u8 x > thr
#include "be_common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kMaxBlocksPerSeg = BE_SM_COUNT * 8;
constexpr int kBins = 2048;
constexpr int kMaxRanks = 4;  // lo / lo+1 for up to two quantiles

// How a segment is cut into per-block chunks (multiple of 4 elements so chunks keep float4 alignment).
struct SegSplit {
    int bps;
    int64_t chunk;
};
inline SegSplit split_segment(int64_t n_seg, int64_t seg_len) {
    // Few fat blocks: a block's fixed cost (zeroing and flushing a shared histogram, the block reduction) is paid
    // once per ~10^4 elements, and the whole grid is about 8 CTAs per SM across all segments.
    const int64_t min_per_block = (int64_t)kThreads * 16;
    int64_t bps = kMaxBlocksPerSeg / (n_seg < 1 ? 1 : n_seg);
    if (bps < 1) bps = 1;
    const int64_t most = (seg_len + min_per_block - 1) / min_per_block;
    if (bps > most) bps = most;
    if (bps < 1) bps = 1;
    int64_t chunk = (seg_len + bps - 1) / bps;
    chunk = (chunk + 3) / 4 * 4;
    bps = (seg_len + chunk - 1) / chunk;
    if (bps < 1) bps = 1;
    return {(int)bps, chunk};
}

// Apply f to every element of p[0..n) with the whole block; float4 body, scalar head/tail.  Four independent 16-byte
// loads are issued before any of them is consumed: 64 B in flight per thread is what keeps HBM busy at ~2k threads/SM.
template <class F>
__device__ __forceinline__ void block_foreach(const float* __restrict__ p, int64_t n, F f) {
    int64_t head = ((16 - ((uintptr_t)p & 15)) & 15) >> 2;
    if (head > n) head = n;
    if ((int64_t)threadIdx.x < head) f(p[threadIdx.x]);
    const float4* p4 = reinterpret_cast<const float4*>(p + head);
    const int64_t n4 = (n - head) >> 2;
    int64_t i = threadIdx.x;
    for (; i + 3 * kThreads < n4; i += 4 * kThreads) {
        const float4 a = ld_stream_f4(p4 + i);
        const float4 b = ld_stream_f4(p4 + i + kThreads);
        const float4 c = ld_stream_f4(p4 + i + 2 * kThreads);
        const float4 d = ld_stream_f4(p4 + i + 3 * kThreads);
        f(a.x); f(a.y); f(a.z); f(a.w);
        f(b.x); f(b.y); f(b.z); f(b.w);
        f(c.x); f(c.y); f(c.z); f(c.w);
        f(d.x); f(d.y); f(d.z); f(d.w);
    }
    for (; i < n4; i += kThreads) {
        const float4 v = ld_stream_f4(p4 + i);
        f(v.x); f(v.y); f(v.z); f(v.w);
    }
    const int64_t done = head + (n4 << 2);
    if (done + (int64_t)threadIdx.x < n) f(p[done + threadIdx.x]);
}

// Returns true in every thread of exactly one block per segment: the last one to arrive.
__device__ __forceinline__ bool last_block_of_segment(unsigned* counter, int bps) {
    __shared__ bool s_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = (atomicAdd(counter, 1u) == (unsigned)(bps - 1));
    __syncthreads();
    if (s_last) __threadfence();
    return s_last;
}

// ---------------------------------------------------------------------------------------------------
// mean / variance
// ---------------------------------------------------------------------------------------------------

__global__ void __launch_bounds__(kThreads) meanvar_kernel(const float* __restrict__ x, int64_t seg_len, int bps,
                                                           int64_t chunk, double eps, double* __restrict__ partials,
                                                           unsigned* __restrict__ counters, double* __restrict__ stats,
                                                           double* __restrict__ norm) {
    const int64_t seg = blockIdx.y;
    const float* xs = x + seg * seg_len;
    const double shift = (double)xs[0];  // sums of (x - x[0]) stay well conditioned for any offset
    const int64_t begin = (int64_t)blockIdx.x * chunk;
    const int64_t end = min(begin + chunk, seg_len);

    double s1 = 0.0, s2 = 0.0;
    block_foreach(xs + begin, end - begin, [&](float v) {
        const double d = (double)v - shift;
        s1 += d;
        s2 = fma(d, d, s2);
    });

    __shared__ double sh1[kThreads / 32], sh2[kThreads / 32];
    s1 = warp_sum(s1);
    s2 = warp_sum(s2);
    if ((threadIdx.x & 31) == 0) {
        sh1[threadIdx.x >> 5] = s1;
        sh2[threadIdx.x >> 5] = s2;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0, b = 0.0;
        for (int w = 0; w < kThreads / 32; ++w) {
            a += sh1[w];
            b += sh2[w];
        }
        partials[(seg * bps + blockIdx.x) * 2 + 0] = a;
        partials[(seg * bps + blockIdx.x) * 2 + 1] = b;
    }
    if (!last_block_of_segment(&counters[seg], bps)) return;

    // fixed-order sum of the per-block partials => bitwise reproducible statistics
    double a = 0.0, b = 0.0;
    for (int i = threadIdx.x; i < bps; i += kThreads) {
        a += __ldcg(&partials[(seg * bps + i) * 2 + 0]);
        b += __ldcg(&partials[(seg * bps + i) * 2 + 1]);
    }
    a = warp_sum(a);
    b = warp_sum(b);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) {
        sh1[threadIdx.x >> 5] = a;
        sh2[threadIdx.x >> 5] = b;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        a = 0.0;
        b = 0.0;
        for (int w = 0; w < kThreads / 32; ++w) {
            a += sh1[w];
            b += sh2[w];
        }
        const double n = (double)seg_len;
        const double m = a / n;
        double var = b / n - m * m;
        if (var < 0.0) var = 0.0;
        const double mean = shift + m;
        stats[seg * 2 + 0] = mean;
        stats[seg * 2 + 1] = var;
        if (norm) {
            norm[seg * 2 + 0] = mean;
            norm[seg * 2 + 1] = sqrt(var) + eps;
        }
        counters[seg] = 0;
    }
}

// ---------------------------------------------------------------------------------------------------
// radix select
// ---------------------------------------------------------------------------------------------------

__device__ __forceinline__ uint32_t f32_to_key(float v) {
    const uint32_t u = __float_as_uint(v);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key_to_f32(uint32_t k) {
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

struct QState {
    uint32_t prefix[kMaxRanks];  // key bits fixed so far (right aligned)
    uint32_t krem[kMaxRanks];    // rank of the target among the elements sharing `prefix`
    int32_t group[kMaxRanks];    // histogram slot of each rank (ranks with equal prefix share one)
    uint32_t gprefix[kMaxRanks];
    int32_t n_groups;
    int32_t n_ranks;
    uint32_t n_valid;
    uint32_t pad;
    double gamma[kMaxRanks / 2];
};

struct QWorkspace {
    unsigned* counters;  // [3][n_seg]
    unsigned* nan_count; // [n_seg]
    unsigned* hist0;     // [n_seg][kBins]
    unsigned* hist12;    // [n_seg][kMaxRanks][kBins]
    QState* state;       // [n_seg]
    size_t zero_bytes;   // leading part that must be zero on entry
    size_t total_bytes;
};

inline QWorkspace carve_quantile_ws(void* ws, int64_t n_seg) {
    QWorkspace w;
    char* p = (char*)ws;
    size_t off = 0;
    w.counters = (unsigned*)(p + off);
    off += be_align_up(sizeof(unsigned) * 3 * n_seg, 256);
    w.nan_count = (unsigned*)(p + off);
    off += be_align_up(sizeof(unsigned) * n_seg, 256);
    w.hist0 = (unsigned*)(p + off);
    off += be_align_up(sizeof(unsigned) * kBins * n_seg, 256);
    w.hist12 = (unsigned*)(p + off);
    off += be_align_up(sizeof(unsigned) * kBins * kMaxRanks * n_seg, 256);
    w.zero_bytes = off;
    w.state = (QState*)(p + off);
    off += be_align_up(sizeof(QState) * n_seg, 256);
    w.total_bytes = off;
    return w;
}

// Block-cooperative: find the bin holding the k-th (0-based) element of a histogram in global memory.
// Every thread gets (bin, k - #elements in lower bins).  kBins / kThreads = 8 bins per thread.
__device__ void find_bin(const unsigned* __restrict__ hist, uint32_t k, uint32_t* bin_out, uint32_t* krem_out) {
    __shared__ uint32_t s_warp[kThreads / 32];
    __shared__ uint32_t s_res[2];
    constexpr int per = kBins / kThreads;
    uint32_t c[per];
    uint32_t local = 0;
#pragma unroll
    for (int i = 0; i < per; ++i) {
        c[i] = __ldcg(&hist[threadIdx.x * per + i]);
        local += c[i];
    }
    // exclusive scan of `local` over the block
    uint32_t incl = local;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if ((threadIdx.x & 31) >= o) incl += t;
    }
    __syncthreads();  // protects s_warp / s_res from the previous call
    if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = incl;
    __syncthreads();
    uint32_t base = 0;
    for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) base += s_warp[w];
    uint32_t excl = base + incl - local;
    if (k >= excl && k < excl + local) {
        uint32_t r = k - excl;
#pragma unroll
        for (int i = 0; i < per; ++i) {
            if (r < c[i]) {
                s_res[0] = threadIdx.x * per + i;
                s_res[1] = r;
                break;
            }
            r -= c[i];
        }
    }
    __syncthreads();
    *bin_out = s_res[0];
    *krem_out = s_res[1];
}

// thread 0 only: ranks with the same prefix share one histogram slot on the next level
__device__ void regroup(QState* st) {
    int ng = 0;
    for (int j = 0; j < st->n_ranks; ++j) {
        int g = -1;
        for (int i = 0; i < ng; ++i)
            if (st->gprefix[i] == st->prefix[j]) g = i;
        if (g < 0) {
            g = ng++;
            st->gprefix[g] = st->prefix[j];
        }
        st->group[j] = g;
    }
    st->n_groups = ng;
}

// Per-thread run combining: consecutive elements of a thread's stream that fall in the same bin become one
// shared-memory atomic.  Natural images put neighbouring pixels in the same coarse bin, which is exactly the
// case where un-combined atomics would serialise on one address.
struct RunCounter {
    uint32_t bin = 0xffffffffu;
    uint32_t cnt = 0;
    __device__ __forceinline__ void add(unsigned* sh, uint32_t b) {
        if (b == bin) {
            ++cnt;
        } else {
            if (cnt) atomicAdd(&sh[bin], cnt);
            bin = b;
            cnt = 1;
        }
    }
    __device__ __forceinline__ void flush(unsigned* sh) {
        if (cnt) atomicAdd(&sh[bin], cnt);
        cnt = 0;
    }
};

__global__ void __launch_bounds__(kThreads) qsel_level0_kernel(const float* __restrict__ x, int64_t seg_len, int bps,
                                                               int64_t chunk, const double* __restrict__ q, int n_q,
                                                               QWorkspace w, int64_t n_seg) {
    __shared__ unsigned sh[kBins];
    __shared__ unsigned s_nan;
    const int64_t seg = blockIdx.y;
    const float* xs = x + seg * seg_len;
    for (int i = threadIdx.x; i < kBins; i += kThreads) sh[i] = 0;
    if (threadIdx.x == 0) s_nan = 0;
    __syncthreads();

    const int64_t begin = (int64_t)blockIdx.x * chunk;
    const int64_t end = min(begin + chunk, seg_len);
    RunCounter rc;
    uint32_t nans = 0;
    block_foreach(xs + begin, end - begin, [&](float v) {
        if (v != v) {
            ++nans;
        } else {
            rc.add(sh, f32_to_key(v) >> 21);
        }
    });
    rc.flush(sh);
    if (nans) atomicAdd(&s_nan, nans);
    __syncthreads();
    unsigned* gh = w.hist0 + seg * kBins;
    for (int i = threadIdx.x; i < kBins; i += kThreads)
        if (sh[i]) atomicAdd(&gh[i], sh[i]);
    if (threadIdx.x == 0 && s_nan) atomicAdd(&w.nan_count[seg], s_nan);

    if (!last_block_of_segment(&w.counters[seg], bps)) return;

    QState* st = &w.state[seg];
    __shared__ uint32_t s_rank[kMaxRanks];
    if (threadIdx.x == 0) {
        const uint32_t n_valid = (uint32_t)seg_len - __ldcg(&w.nan_count[seg]);
        st->n_valid = n_valid;
        st->n_ranks = 2 * n_q;
        for (int i = 0; i < n_q; ++i) {
            // numpy: virtual index q*(n-1) in fp64, neighbours floor / floor+1 clipped to n-1
            double h = (n_valid > 0) ? q[i] * (double)(n_valid - 1) : 0.0;
            if (h < 0.0) h = 0.0;
            double lo = floor(h);
            uint32_t ilo = (uint32_t)lo;
            if (n_valid > 0 && ilo > n_valid - 1) ilo = n_valid - 1;
            uint32_t ihi = (n_valid > 0 && ilo + 1 < n_valid) ? ilo + 1 : ilo;
            s_rank[2 * i] = ilo;
            s_rank[2 * i + 1] = ihi;
            st->gamma[i] = h - lo;
        }
    }
    __syncthreads();
    const int n_ranks = 2 * n_q;
    const uint32_t n_valid = (uint32_t)seg_len - __ldcg(&w.nan_count[seg]);
    if (n_valid > 0) {
        for (int j = 0; j < n_ranks; ++j) {
            uint32_t bin, krem;
            find_bin(gh, s_rank[j], &bin, &krem);
            if (threadIdx.x == 0) {
                st->prefix[j] = bin;
                st->krem[j] = krem;
            }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        regroup(st);
        __threadfence();
    }
}

template <int LEVEL>
__global__ void __launch_bounds__(kThreads) qsel_level12_kernel(const float* __restrict__ x, int64_t seg_len, int bps,
                                                                int64_t chunk, int n_q, double eps, QWorkspace w,
                                                                int64_t n_seg, double* __restrict__ quant,
                                                                double* __restrict__ norm) {
    constexpr int kBits = (LEVEL == 1) ? 11 : 10;
    constexpr int kShift = (LEVEL == 1) ? 10 : 0;        // digit position in the key
    constexpr int kHiShift = (LEVEL == 1) ? 21 : 10;     // bits above the digit form the prefix
    constexpr uint32_t kMask = (1u << kBits) - 1u;
    __shared__ unsigned sh[kMaxRanks][kBins];
    const int64_t seg = blockIdx.y;
    const float* xs = x + seg * seg_len;
    QState* st = &w.state[seg];
    const int n_groups = st->n_groups;
    const uint32_t n_valid = st->n_valid;
    uint32_t gp[kMaxRanks];
#pragma unroll
    for (int g = 0; g < kMaxRanks; ++g) gp[g] = (g < n_groups) ? st->gprefix[g] : 0xffffffffu;

    for (int i = threadIdx.x; i < n_groups * kBins; i += kThreads) (&sh[0][0])[i] = 0;
    __syncthreads();

    const int64_t begin = (int64_t)blockIdx.x * chunk;
    const int64_t end = min(begin + chunk, seg_len);
    RunCounter rc[kMaxRanks];
    // A bucket of keys is an interval of floats, so almost every element is rejected by two float compares against the
    // hull of the (<= 4, usually 1-2) wanted buckets; NaNs fail both.  The inclusive float test lets both zeros through
    // at a bucket edge; the exact key test below sorts them out.
    float fmin = __int_as_float(0x7f800000), fmax = __int_as_float(0xff800000);
#pragma unroll
    for (int g = 0; g < kMaxRanks; ++g) {
        if (g < n_groups) {
            fmin = fminf(fmin, key_to_f32(gp[g] << kHiShift));
            fmax = fmaxf(fmax, key_to_f32((gp[g] << kHiShift) | ((1u << kHiShift) - 1u)));
        }
    }
    if (n_valid > 0) {
        block_foreach(xs + begin, end - begin, [&](float v) {
            if (v >= fmin && v <= fmax) {
                const uint32_t key = f32_to_key(v);
                const uint32_t hi = key >> kHiShift;
                const uint32_t digit = (key >> kShift) & kMask;
#pragma unroll
                for (int g = 0; g < kMaxRanks; ++g)
                    if (hi == gp[g]) rc[g].add(sh[g], digit);  // gp of unused groups can never match (prefix < 2^22)
            }
        });
    }
#pragma unroll
    for (int g = 0; g < kMaxRanks; ++g) rc[g].flush(sh[g]);
    __syncthreads();
    unsigned* gh = w.hist12 + seg * (kMaxRanks * kBins);
    for (int i = threadIdx.x; i < n_groups * kBins; i += kThreads) {
        const unsigned c = (&sh[0][0])[i];
        if (c) atomicAdd(&gh[i], c);
    }

    if (!last_block_of_segment(&w.counters[LEVEL * n_seg + seg], bps)) return;

    const int n_ranks = st->n_ranks;
    if (n_valid > 0) {
        for (int j = 0; j < n_ranks; ++j) {
            uint32_t bin, krem;
            find_bin(gh + st->group[j] * kBins, st->krem[j], &bin, &krem);
            __syncthreads();
            if (threadIdx.x == 0) {
                st->prefix[j] = (st->prefix[j] << kBits) | bin;
                st->krem[j] = krem;
            }
        }
    }
    __syncthreads();
    if (LEVEL == 1) {
        // the same histogram storage serves level 2
        for (int i = threadIdx.x; i < kMaxRanks * kBins; i += kThreads) gh[i] = 0;
        if (threadIdx.x == 0) regroup(st);
    } else if (threadIdx.x == 0) {
        for (int i = 0; i < n_q; ++i) {
            double r;
            if (n_valid == 0) {
                r = __longlong_as_double(0x7ff8000000000000LL);
            } else {
                // numpy _lerp: fp32 difference, fp64 product and sum, switch of anchor at gamma >= 0.5
                const float a = key_to_f32(st->prefix[2 * i]);
                const float b = key_to_f32(st->prefix[2 * i + 1]);
                const double d = (double)__fsub_rn(b, a);
                const double t = st->gamma[i];
                r = (t >= 0.5) ? __dsub_rn((double)b, __dmul_rn(d, __dsub_rn(1.0, t)))
                               : __dadd_rn((double)a, __dmul_rn(d, t));
            }
            quant[seg * n_q + i] = r;
        }
        if (norm) {
            const double lo = quant[seg * n_q + 0], hi = quant[seg * n_q + 1];
            norm[seg * 2 + 0] = lo;
            norm[seg * 2 + 1] = __dadd_rn(__dsub_rn(hi, lo), eps);
        }
    }
}

// ---------------------------------------------------------------------------------------------------
// elementwise
// ---------------------------------------------------------------------------------------------------

__device__ __forceinline__ float normalize_one(float v, double sub, double div, double rdiv, int clip, float lo,
                                               float hi) {
    float r = be_div_to_f32(__dsub_rn((double)v, sub), div, rdiv);
    if (clip && r == r) r = fminf(fmaxf(r, lo), hi);  // NaN passes through, as in np.clip
    return r;
}

__global__ void __launch_bounds__(kThreads) normalize_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                             int64_t seg_len, const double* __restrict__ norm, int clip,
                                                             float lo, float hi) {
    const int64_t seg = blockIdx.y;
    const double sub = norm[seg * 2], div = norm[seg * 2 + 1];
    const double rdiv = be_fast_rcp(div);
    const float* xs = x + seg * seg_len;
    float* ys = y + seg * seg_len;
    const bool vec = ((((uintptr_t)xs | (uintptr_t)ys) & 15) == 0);
    const int64_t stride = (int64_t)gridDim.x * kThreads;
    if (vec) {
        const int64_t n4 = seg_len >> 2;
        for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n4; i += stride) {
            float4 v = ld_stream_f4(reinterpret_cast<const float4*>(xs) + i);
            v.x = normalize_one(v.x, sub, div, rdiv, clip, lo, hi);
            v.y = normalize_one(v.y, sub, div, rdiv, clip, lo, hi);
            v.z = normalize_one(v.z, sub, div, rdiv, clip, lo, hi);
            v.w = normalize_one(v.w, sub, div, rdiv, clip, lo, hi);
            st_stream_f4(reinterpret_cast<float4*>(ys) + i, v);
        }
        for (int64_t i = (n4 << 2) + (int64_t)blockIdx.x * kThreads + threadIdx.x; i < seg_len; i += stride)
            ys[i] = normalize_one(xs[i], sub, div, rdiv, clip, lo, hi);
    } else {
        for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < seg_len; i += stride)
            ys[i] = normalize_one(xs[i], sub, div, rdiv, clip, lo, hi);
    }
}

template <int OP>
__device__ __forceinline__ float ew_one(float v, float a, float b) {
    if (OP == BE_OP_CLIP) return (v != v) ? v : fminf(fmaxf(v, a), b);  // NaN passes through, as in np.clip
    if (OP == BE_OP_SCALE_LINEAR) return __fadd_rn(__fmul_rn(v, a), b);
    return be_sigmoid_f32(v);
}

template <int OP>
__global__ void __launch_bounds__(kThreads) elementwise_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                               int64_t n, float a, float b) {
    const int64_t stride = (int64_t)gridDim.x * kThreads;
    const int64_t tid = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    if (((((uintptr_t)x | (uintptr_t)y) & 15) == 0)) {
        const int64_t n4 = n >> 2;
        for (int64_t i = tid; i < n4; i += stride) {
            float4 v = ld_stream_f4(reinterpret_cast<const float4*>(x) + i);
            v.x = ew_one<OP>(v.x, a, b);
            v.y = ew_one<OP>(v.y, a, b);
            v.z = ew_one<OP>(v.z, a, b);
            v.w = ew_one<OP>(v.w, a, b);
            st_stream_f4(reinterpret_cast<float4*>(y) + i, v);
        }
        for (int64_t i = (n4 << 2) + tid; i < n; i += stride) y[i] = ew_one<OP>(x[i], a, b);
    } else {
        for (int64_t i = tid; i < n; i += stride) y[i] = ew_one<OP>(x[i], a, b);
    }
}

__global__ void __launch_bounds__(kThreads) binarize_kernel(const float* __restrict__ x, uint8_t* __restrict__ y,
                                                            int64_t n, float thr) {
    const int64_t stride = (int64_t)gridDim.x * kThreads;
    const int64_t tid = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    if (((((uintptr_t)x) & 15) == 0) && ((((uintptr_t)y) & 3) == 0)) {
        const int64_t n4 = n >> 2;
        for (int64_t i = tid; i < n4; i += stride) {
            const float4 v = ld_stream_f4(reinterpret_cast<const float4*>(x) + i);
            uchar4 m;
            m.x = v.x > thr;
            m.y = v.y > thr;
            m.z = v.z > thr;
            m.w = v.w > thr;
            reinterpret_cast<uchar4*>(y)[i] = m;
        }
        for (int64_t i = (n4 << 2) + tid; i < n; i += stride) y[i] = x[i] > thr;
    } else {
        for (int64_t i = tid; i < n; i += stride) y[i] = x[i] > thr;
    }
}

inline int stream_grid(int64_t n_items_per_thread_units) {
    int64_t blocks = (n_items_per_thread_units + kThreads - 1) / kThreads;
    const int64_t cap = (int64_t)BE_SM_COUNT * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

}  // namespace

// ---------------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------------

extern "C" size_t be_meanvar_workspace_bytes(int64_t n_seg, int64_t seg_len) {
    if (n_seg <= 0 || seg_len <= 0) return 256;
    const SegSplit sp = split_segment(n_seg, seg_len);
    return be_align_up(sizeof(unsigned) * n_seg, 256) + be_align_up(sizeof(double) * 2 * sp.bps * n_seg, 256);
}

extern "C" int be_meanvar_f32(const float* x, int64_t n_seg, int64_t seg_len, double eps, double* stats, double* norm,
                              void* ws, be_stream_t stream) {
    BE_REQUIRE(x && stats && ws, BE_ERR_INVALID_ARG, "be_meanvar_f32: null pointer");
    BE_REQUIRE(n_seg > 0 && seg_len > 0, BE_ERR_INVALID_ARG, "be_meanvar_f32: empty input (n_seg=%lld seg_len=%lld)",
               (long long)n_seg, (long long)seg_len);
    BE_REQUIRE(n_seg <= 65535, BE_ERR_TOO_LARGE, "be_meanvar_f32: n_seg %lld > 65535", (long long)n_seg);
    cudaStream_t s = (cudaStream_t)stream;
    const SegSplit sp = split_segment(n_seg, seg_len);
    unsigned* counters = (unsigned*)ws;
    double* partials = (double*)((char*)ws + be_align_up(sizeof(unsigned) * n_seg, 256));
    BE_CUDA(cudaMemsetAsync(counters, 0, sizeof(unsigned) * n_seg, s));
    meanvar_kernel<<<dim3(sp.bps, (unsigned)n_seg), kThreads, 0, s>>>(x, seg_len, sp.bps, sp.chunk, eps, partials,
                                                                      counters, stats, norm);
    BE_LAUNCH_CHECK();
    return BE_OK;
}

extern "C" size_t be_quantile_workspace_bytes(int64_t n_seg, int64_t seg_len) {
    (void)seg_len;
    if (n_seg <= 0) return 256;
    return carve_quantile_ws(nullptr, n_seg).total_bytes;
}

extern "C" int be_quantile_f32(const float* x, int64_t n_seg, int64_t seg_len, const double* q, int n_q, double eps,
                               double* quant, double* norm, void* ws, be_stream_t stream) {
    BE_REQUIRE(x && q && quant && ws, BE_ERR_INVALID_ARG, "be_quantile_f32: null pointer");
    BE_REQUIRE(n_seg > 0 && seg_len > 0, BE_ERR_INVALID_ARG, "be_quantile_f32: empty input (n_seg=%lld seg_len=%lld)",
               (long long)n_seg, (long long)seg_len);
    BE_REQUIRE(n_q >= 1 && 2 * n_q <= kMaxRanks, BE_ERR_INVALID_ARG, "be_quantile_f32: n_q must be 1 or 2, got %d", n_q);
    BE_REQUIRE(!norm || n_q == 2, BE_ERR_INVALID_ARG, "be_quantile_f32: norm output needs n_q == 2");
    BE_REQUIRE(seg_len < (int64_t)1 << 31, BE_ERR_TOO_LARGE, "be_quantile_f32: seg_len %lld >= 2^31", (long long)seg_len);
    BE_REQUIRE(n_seg <= 65535, BE_ERR_TOO_LARGE, "be_quantile_f32: n_seg %lld > 65535", (long long)n_seg);
    cudaStream_t s = (cudaStream_t)stream;
    const SegSplit sp = split_segment(n_seg, seg_len);
    QWorkspace w = carve_quantile_ws(ws, n_seg);
    BE_CUDA(cudaMemsetAsync(ws, 0, w.zero_bytes, s));
    const dim3 grid(sp.bps, (unsigned)n_seg);
    qsel_level0_kernel<<<grid, kThreads, 0, s>>>(x, seg_len, sp.bps, sp.chunk, q, n_q, w, n_seg);
    BE_LAUNCH_CHECK();
    qsel_level12_kernel<1><<<grid, kThreads, 0, s>>>(x, seg_len, sp.bps, sp.chunk, n_q, eps, w, n_seg, quant, norm);
    BE_LAUNCH_CHECK();
    qsel_level12_kernel<2><<<grid, kThreads, 0, s>>>(x, seg_len, sp.bps, sp.chunk, n_q, eps, w, n_seg, quant, norm);
    BE_LAUNCH_CHECK();
    return BE_OK;
}

// ---- quantiles for callers that must not allocate -------------------------------------------------------------------
//
// Some hosts of this library cannot give it a workspace: a microscope acquisition loop that is not allowed to touch
// the CUDA allocator between frames (cudaMalloc / cudaFree serialise the whole device and can stall a live stream), or
// an embedding runtime that has handed every byte of device memory to its own pool and passes plugins nothing but the
// input and output pointers.  For them the scratch of the quantile kernels lives in the library image itself: one
// fixed array per GPU, present from the moment the library is loaded on that GPU until the process exits.
//
// `g_quantile_scratch` below is that array.  On the host the name is only a handle for the symbol — it is not a
// device address and must never be handed to a kernel; the address is looked up once per device with
// cudaGetSymbolAddress while that device is current, and cached.
constexpr int kNoAllocMaxSegments = 8;
constexpr size_t kNoAllocScratchBytes = 344064;  // >= carve_quantile_ws(., 8).total_bytes (checked at run time)
static __device__ __align__(256) unsigned char g_quantile_scratch[kNoAllocScratchBytes];

constexpr int kMaxDevices = 64;
static void* g_quantile_scratch_addr[kMaxDevices];  // per-device address of g_quantile_scratch, filled on first use

extern "C" int be_quantile_f32_noalloc(const float* x, int64_t n_seg, int64_t seg_len, const double* q, int n_q,
                                       double eps, double* quant, double* norm, be_stream_t stream) {
    BE_REQUIRE(n_seg >= 1 && n_seg <= kNoAllocMaxSegments, BE_ERR_TOO_LARGE,
               "be_quantile_f32_noalloc: n_seg %lld outside 1..%d (the built-in scratch is sized for %d segments; use "
               "be_quantile_f32 with a caller workspace for more)",
               (long long)n_seg, kNoAllocMaxSegments, kNoAllocMaxSegments);
    BE_REQUIRE(carve_quantile_ws(nullptr, n_seg).total_bytes <= kNoAllocScratchBytes, BE_ERR_TOO_LARGE,
               "be_quantile_f32_noalloc: built-in scratch too small (library bug)");
    int dev = -1;
    BE_CUDA(cudaGetDevice(&dev));
    BE_REQUIRE(dev >= 0 && dev < kMaxDevices, BE_ERR_TOO_LARGE, "be_quantile_f32_noalloc: device ordinal %d >= %d", dev,
               kMaxDevices);
    void* scratch = g_quantile_scratch_addr[dev];
    if (!scratch) {
        // first call on this device: resolve the symbol's device address (no allocation; the array is part of the
        // loaded module).  Two threads racing here both store the same value.
        BE_CUDA(cudaGetSymbolAddress(&scratch, g_quantile_scratch));
        BE_REQUIRE(scratch != nullptr, BE_ERR_CUDA, "be_quantile_f32_noalloc: cudaGetSymbolAddress returned null");
        g_quantile_scratch_addr[dev] = scratch;
    }
    // be_quantile_f32 zeroes the counters and histograms it needs at the start of every call, so whatever the previous
    // call left in the shared array is harmless.
    return be_quantile_f32(x, n_seg, seg_len, q, n_q, eps, quant, norm, scratch, stream);
}

extern "C" int be_normalize_f32(const float* x, float* y, int64_t n_seg, int64_t seg_len, const double* norm, int clip,
                                float clip_lo, float clip_hi, be_stream_t stream) {
    BE_REQUIRE(x && y && norm, BE_ERR_INVALID_ARG, "be_normalize_f32: null pointer");
    BE_REQUIRE(n_seg > 0 && seg_len > 0, BE_ERR_INVALID_ARG, "be_normalize_f32: empty input");
    BE_REQUIRE(n_seg <= 65535, BE_ERR_TOO_LARGE, "be_normalize_f32: n_seg %lld > 65535", (long long)n_seg);
    int gx = stream_grid((seg_len + 3) / 4);
    if (n_seg > 1) gx = (int)min((int64_t)gx, max((int64_t)1, (int64_t)BE_SM_COUNT * 16 / n_seg));
    normalize_kernel<<<dim3(gx, (unsigned)n_seg), kThreads, 0, (cudaStream_t)stream>>>(x, y, seg_len, norm, clip,
                                                                                       clip_lo, clip_hi);
    BE_LAUNCH_CHECK();
    return BE_OK;
}

extern "C" int be_elementwise_f32(const float* x, float* y, int64_t n, int op, float a, float b, be_stream_t stream) {
    BE_REQUIRE(x && y, BE_ERR_INVALID_ARG, "be_elementwise_f32: null pointer");
    BE_REQUIRE(n > 0, BE_ERR_INVALID_ARG, "be_elementwise_f32: empty input");
    const int g = stream_grid((n + 3) / 4);
    cudaStream_t s = (cudaStream_t)stream;
    switch (op) {
        case BE_OP_CLIP: elementwise_kernel<BE_OP_CLIP><<<g, kThreads, 0, s>>>(x, y, n, a, b); break;
        case BE_OP_SCALE_LINEAR: elementwise_kernel<BE_OP_SCALE_LINEAR><<<g, kThreads, 0, s>>>(x, y, n, a, b); break;
        case BE_OP_SIGMOID: elementwise_kernel<BE_OP_SIGMOID><<<g, kThreads, 0, s>>>(x, y, n, a, b); break;
        default: BE_REQUIRE(false, BE_ERR_INVALID_ARG, "be_elementwise_f32: unknown op %d", op);
    }
    BE_LAUNCH_CHECK();
    return BE_OK;
}

extern "C" int be_binarize_f32_u8(const float* x, uint8_t* y, int64_t n, float thr, be_stream_t stream) {
    BE_REQUIRE(x && y, BE_ERR_INVALID_ARG, "be_binarize_f32_u8: null pointer");
    BE_REQUIRE(n > 0, BE_ERR_INVALID_ARG, "be_binarize_f32_u8: empty input");
    binarize_kernel<<<stream_grid((n + 3) / 4), kThreads, 0, (cudaStream_t)stream>>>(x, y, n, thr);
    BE_LAUNCH_CHECK();
    return BE_OK;
}

// Pre-/post-processing statistics and elementwise kernels (HBM-bound; no tensor cores by design).
//
//   be_meanvar_f32     one pass, shifted fp64 sums, deterministic two-stage reduction
//   be_quantile_f32    exact order statistics + numpy's linear lerp: sampled-pivot single pass for large segments
//                      (qsel_sample / partition kernels, then one thread-block-cluster launch for the selection), 3-level (11/11/10 bit) radix select otherwise
//   be_normalize_f32   (x - sub) / div in fp64, optional clip
//   be_elementwise_f32 clip / scale_linear / sigmoid
//   be_binarize_f32_u8 x > thr
#include <cooperative_groups.h>
#include <cuda_fp16.h>
#include <stdlib.h>

#include <type_traits>

#include "be_common.cuh"

namespace cg = cooperative_groups;

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
inline SegSplit split_segment(int64_t n_seg, int64_t seg_len, int64_t max_blocks = kMaxBlocksPerSeg) {
    // Few fat blocks: a block's fixed cost (zeroing and flushing a shared histogram, the block reduction) is paid
    // once per ~10^4 elements, and the whole grid is about 8 CTAs per SM across all segments.
    const int64_t min_per_block = (int64_t)kThreads * 16;
    int64_t bps = max_blocks / (n_seg < 1 ? 1 : n_seg);
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
    if (threadIdx.x == 0) s_res[0] = s_res[1] = 0;  // k beyond the histogram: a defined answer (bin 0)
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
    // at a bucket edge; the exact key test below sorts them out.  The buckets that hold -inf / +inf also span NaN
    // encodings (keys below key(-inf), above key(+inf)); their bounds are clamped to the infinities' keys first, so the
    // hull stays a float interval that contains the infinities (a NaN bound would be dropped by fminf / fmaxf and the
    // hull would come out empty).
    constexpr uint32_t kKeyNegInf = 0x007fffffu, kKeyPosInf = 0xff800000u;
    float fmin = __int_as_float(0x7f800000), fmax = __int_as_float(0xff800000);
#pragma unroll
    for (int g = 0; g < kMaxRanks; ++g) {
        if (g < n_groups) {
            const uint32_t klo = gp[g] << kHiShift, khi = klo | ((1u << kHiShift) - 1u);
            fmin = fminf(fmin, key_to_f32(klo < kKeyNegInf ? kKeyNegInf : klo));
            fmax = fmaxf(fmax, key_to_f32(khi > kKeyPosInf ? kKeyPosInf : khi));
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

// y = x * gain[seg] + offset[seg] in fp32, the per-segment constants read from device memory ({offset, gain} pairs, fp64,
// as be_meanvar_f32 writes its `norm` output for a reference tensor: {ref_mean, ref_std + eps})
__global__ void __launch_bounds__(kThreads) scale_linear_seg_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                                    int64_t seg_len, const double* __restrict__ params) {
    const int64_t seg = blockIdx.y;
    const float offset = (float)params[seg * 2], gain = (float)params[seg * 2 + 1];
    const float* xs = x + seg * seg_len;
    float* ys = y + seg * seg_len;
    for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < seg_len; i += (int64_t)gridDim.x * kThreads)
        ys[i] = __fadd_rn(__fmul_rn(xs[i], gain), offset);
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

// ---------------------------------------------------------------------------------------------------
// sampled-pivot select: one streaming pass over a large segment instead of three
// ---------------------------------------------------------------------------------------------------
//
// For every (segment, quantile) TASK a 2048-element sample is sorted and two pivots flo <= fhi are taken 4 sigma + 4
// (of the sample-rank binomial) either side of the wanted rank.  One pass over the segment then counts the elements
// below flo, equal to flo, equal to fhi, and appends the (about 2-3 %) strictly in between to a candidate list.  The
// exact order statistics are found by the same 11/11/10-bit radix select, run over the candidate list only, with the
// tie counts entering the histograms as "virtual" elements at key(flo) / key(fhi).  If the wanted rank falls outside
// the bracket or the list overflows (both essentially impossible, but data-dependent), the task's source is switched
// to the whole segment and the very same radix kernels do the full three passes: the result is exact either way.

constexpr int kSample = 2048;       // one compare-exchange pair per thread and stage
constexpr int kSampleLarge = 8192;  // segments of 2^26 elements and more: brackets half as wide for ~15 us more sorting
constexpr int64_t kSampleLargeLen = (int64_t)1 << 26;
constexpr int kSampleThreads = 1024;
constexpr int64_t kSampleMinLen = (int64_t)1 << 18;  // below this the plain radix select is used
constexpr int kStage = 1024;                         // per-block staging of candidates in shared memory

struct QTask {
    const float* src;     // candidate list, or the segment itself after a fallback
    uint32_t len;
    uint32_t rank[2];     // ranks (lo, hi) inside the multiset {virtual ties} + src
    uint32_t vkey[2];     // key(flo), key(fhi)
    uint32_t vcnt[2];     // how many elements equal flo / fhi
    uint32_t total;       // multiset size (0 => all-NaN segment => NaN result)
    double gamma;
    // partition phase
    float flo, fhi;
    uint32_t below, eq_lo, eq_hi, n_mid;
    uint32_t kmax, nkmin;  // max key and max ~key over the candidate list (zero-initialised, raised with atomicMax)
    uint32_t klo, range;   // selection: digits are cut from (key - klo), which lies in [0, range]
};

struct SWorkspace {
    unsigned* counters;   // [n_seg] completion tickets of the partition pass
    unsigned* nan_count;  // [n_seg]
    unsigned* seg_done;   // [n_seg]
    QTask* tasks;         // [n_tasks]   (zeroed: counters inside)
    size_t zero_bytes;    // everything up to here is zeroed at the start of a call
    float* cand;          // [n_tasks][cap]
    int64_t cap;
    size_t total_bytes;
};

inline int64_t sample_cap(int64_t seg_len) {
    int64_t c = seg_len / 16;
    if (c < 16384) c = 16384;
    return (c + 63) / 64 * 64;
}

inline int64_t sample_segments_per_pass(int64_t n_seg, int64_t seg_len) {
    const int64_t per_seg = sample_cap(seg_len) * 2 * (int64_t)sizeof(float);
    int64_t s = ((int64_t)256 << 20) / per_seg;
    if (s < 1) s = 1;
    if (s > 4096) s = 4096;
    return n_seg < s ? n_seg : s;
}

inline SWorkspace carve_sample_ws(void* ws, int64_t n_seg, int n_q, int64_t seg_len) {
    SWorkspace w;
    const int64_t nt = n_seg * n_q;
    char* p = (char*)ws;
    size_t off = 0;
    w.counters = (unsigned*)(p + off);
    off += be_align_up(sizeof(unsigned) * n_seg, 256);
    w.nan_count = (unsigned*)(p + off);
    off += be_align_up(sizeof(unsigned) * n_seg, 256);
    w.seg_done = (unsigned*)(p + off);
    off += be_align_up(sizeof(unsigned) * n_seg, 256);
    w.tasks = (QTask*)(p + off);
    off += be_align_up(sizeof(QTask) * nt, 256);
    w.zero_bytes = off;
    w.cap = sample_cap(seg_len);
    w.cand = (float*)(p + off);
    off += be_align_up(sizeof(float) * w.cap * nt, 256);
    w.total_bytes = off;
    return w;
}

__device__ __forceinline__ uint32_t hash32(uint32_t x) {
    x ^= x >> 16;
    x *= 0x7feb352du;
    x ^= x >> 15;
    x *= 0x846ca68bu;
    x ^= x >> 16;
    return x;
}

// one block per segment: sample, bitonic sort in shared memory, pivots for each quantile
template <int kSample>
__global__ void __launch_bounds__(kSampleThreads) qsel_sample_kernel(const float* __restrict__ x, int64_t seg_len,
                                                                     const double* __restrict__ q, int n_q,
                                                                     SWorkspace w) {
    __shared__ uint32_t key[kSample];
    __shared__ unsigned s_valid;
    be_pdl_launch_dependents();   // the partition pass may be placed on the other SMs now; it waits for the pivots
    const int64_t seg = blockIdx.x;
    const float* xs = x + seg * seg_len;
    const uint32_t stride = (uint32_t)(seg_len / kSample);  // >= 64
    if (threadIdx.x == 0) s_valid = 0;
    __syncthreads();
    unsigned valid = 0;
    for (int i = threadIdx.x; i < kSample; i += kSampleThreads) {
        const int64_t pos = (int64_t)i * stride + hash32((uint32_t)i * 0x9e3779b9u + (uint32_t)seg) % stride;
        const float v = __ldg(xs + pos);
        const bool ok = (v == v);
        key[i] = ok ? f32_to_key(v) : 0xffffffffu;  // NaNs sort last
        valid += ok;
    }
    valid = __reduce_add_sync(0xffffffffu, valid);
    if ((threadIdx.x & 31) == 0) atomicAdd(&s_valid, valid);
    __syncthreads();
    for (int k = 2; k <= kSample; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int p = threadIdx.x; p < kSample / 2; p += kSampleThreads) {
                const int i = ((p & ~(j - 1)) << 1) | (p & (j - 1));  // lower element of the p-th pair at distance j
                const int l = i | j;
                const uint32_t a = key[i], b = key[l];
                if ((a > b) == ((i & k) == 0)) {
                    key[i] = b;
                    key[l] = a;
                }
            }
            __syncthreads();
        }
    }
    if ((int)threadIdx.x < n_q) {
        const int i = threadIdx.x;
        const int mv = (int)s_valid;
        QTask* t = &w.tasks[seg * n_q + i];
        float flo = __int_as_float(0xff800000), fhi = __int_as_float(0x7f800000);
        if (mv > 0) {
            const double qi = fmin(fmax(q[i], 0.0), 1.0);
            const double sp = qi * (double)(mv - 1);
            const int d = (int)(4.0 * sqrt((double)mv * qi * (1.0 - qi))) + 4;   // misses about once in 1e5 calls
            const int lo = (int)floor(sp) - d, hi = (int)ceil(sp) + 1 + d;
            if (lo >= 0) flo = key_to_f32(key[lo]);
            if (hi < mv) fhi = key_to_f32(key[hi]);
        }
        t->flo = flo;
        t->fhi = fhi;
    }
}

// the single streaming pass: classify every element against the (<= 2) brackets of its segment
// The pivots of a segment without a sort: the block reads the 2048 sample positions and picks the 2 NQ wanted order
// statistics with an 8-bit radix select — rounds of {shared-memory histogram per wanted rank, one warp per rank walks its
// 256 bins} — instead of the 66 compare-exchange stages of a bitonic sort (sample kernel 12.9 -> ~8 us).  (Doing this in
// every block of the partition pass, to save the launch, was no faster: 592 blocks fetching the same 2048 sectors.)
// Digits are cut from (key - lowest sample key), starting at the top bit of the sample's key range: an image's values
// share their leading exponent bits, and digits taken from the raw key would send the whole sample to one or two bins
// (2048 atomics on one shared-memory word).  Same sample positions and ranks as qsel_sample_kernel, hence the same
// pivots as its sorted array gives.
template <int NQ, int kThreads>
__device__ __forceinline__ void qsel_own_pivots(const float* __restrict__ xs, int64_t seg_len, int64_t seg,
                                                const double* __restrict__ q, float* flo, float* fhi) {
    constexpr int kSel = 2 * NQ, kPer = kSample / kThreads;
    __shared__ unsigned s_hist[kSel][256];
    __shared__ unsigned s_pref[kSel];
    __shared__ int s_rank[kSel];
    __shared__ unsigned s_valid, s_min, s_nmax;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) {
        s_valid = 0;
        s_min = 0xffffffffu;
        s_nmax = 0xffffffffu;
    }
    __syncthreads();
    const uint32_t stride = (uint32_t)(seg_len / kSample);
    uint32_t key[kPer];
    unsigned okmask = 0, kmin = 0xffffffffu, nkmax = 0xffffffffu;
#pragma unroll
    for (int j = 0; j < kPer; ++j) {
        const int i = tid + j * kThreads;
        const int64_t pos = (int64_t)i * stride + hash32((uint32_t)i * 0x9e3779b9u + (uint32_t)seg) % stride;
        const float v = __ldg(xs + pos);
        key[j] = f32_to_key(v);
        if (v == v) {   // NaNs rank last and are never wanted: they stay out of the histograms
            okmask |= 1u << j;
            kmin = min(kmin, key[j]);
            nkmax = min(nkmax, ~key[j]);
        }
    }
    const unsigned valid = __reduce_add_sync(0xffffffffu, (unsigned)__popc(okmask));
    kmin = __reduce_min_sync(0xffffffffu, kmin);
    nkmax = __reduce_min_sync(0xffffffffu, nkmax);
    if (lane == 0) {
        atomicAdd(&s_valid, valid);
        atomicMin(&s_min, kmin);
        atomicMin(&s_nmax, nkmax);
    }
    __syncthreads();
    const int mv = (int)s_valid;
    const unsigned base = s_min, range = mv > 0 ? ~s_nmax - s_min : 0u;
    if (tid < kSel) {
        const int i = tid >> 1, side = tid & 1;
        int rank = -1;
        if (mv > 0) {
            const double qi = fmin(fmax(q[i], 0.0), 1.0);
            const double sp = qi * (double)(mv - 1);
            const int d = (int)(4.0 * sqrt((double)mv * qi * (1.0 - qi))) + 4;   // misses about once in 1e5 calls
            const int lo = (int)floor(sp) - d, hi = (int)ceil(sp) + 1 + d;
            if (side == 0 && lo >= 0) rank = lo;
            if (side == 1 && hi < mv) rank = hi;
        }
        s_rank[tid] = rank;
        s_pref[tid] = 0;
    }
#pragma unroll
    for (int j = 0; j < kPer; ++j) key[j] -= base;
    const int bits = 32 - __clz(range | 1u);
    int prev = 32;   // bits [prev, 32) of (key - base) are settled
    for (int shift = max(bits - 8, 0);; shift = max(shift - 8, 0)) {
        for (int b = tid; b < kSel * 256; b += kThreads) (&s_hist[0][0])[b] = 0;
        __syncthreads();
#pragma unroll
        for (int sel = 0; sel < kSel; ++sel) {
            if (s_rank[sel] < 0) continue;
            const unsigned pref = s_pref[sel];
#pragma unroll
            for (int j = 0; j < kPer; ++j)
                if (((okmask >> j) & 1u) && (prev >= 32 || (key[j] >> prev) == (pref >> prev)))
                    atomicAdd(&s_hist[sel][(key[j] >> shift) & 255u], 1u);
        }
        __syncthreads();
        if (warp < kSel && s_rank[warp] >= 0) {
            unsigned c[8], tot = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                c[j] = s_hist[warp][lane * 8 + j];
                tot += c[j];
            }
            unsigned incl = tot;
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned t = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += t;
            }
            unsigned cum = incl - tot;
            const unsigned r = (unsigned)s_rank[warp];
            if (r >= cum && r < cum + tot) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    if (r < cum + c[j]) {
                        s_pref[warp] |= (unsigned)(lane * 8 + j) << shift;   // bits shared with the last digit agree
                        s_rank[warp] = (int)(r - cum);
                        break;
                    }
                    cum += c[j];
                }
            }
        }
        __syncthreads();
        prev = shift;
        if (shift == 0) break;
    }
#pragma unroll
    for (int i = 0; i < NQ; ++i) {
        flo[i] = s_rank[2 * i] >= 0 ? key_to_f32(s_pref[2 * i] + base) : __int_as_float(0xff800000);
        fhi[i] = s_rank[2 * i + 1] >= 0 ? key_to_f32(s_pref[2 * i + 1] + base) : __int_as_float(0x7f800000);
    }
}

// one block per segment: the pivots of its NQ quantiles into the task records
template <int NQ>
__global__ void __launch_bounds__(kSampleThreads) qsel_pivot_kernel(const float* __restrict__ x, int64_t seg_len,
                                                                    const double* __restrict__ q, SWorkspace w) {
    be_pdl_wait();                // whatever wrote x, and the last reader of the task records
    be_pdl_launch_dependents();   // the partition pass may be placed now; it waits for the pivots
    const int64_t seg = blockIdx.x;
    // this segment's counters and task records start from zero (no memset between this kernel and the one before it)
    if (threadIdx.x == 0) w.counters[seg] = w.nan_count[seg] = w.seg_done[seg] = 0u;
    {
        uint32_t* t32 = reinterpret_cast<uint32_t*>(w.tasks + seg * NQ);
        for (int i = threadIdx.x; i < (int)(NQ * sizeof(QTask) / sizeof(uint32_t)); i += kSampleThreads) t32[i] = 0u;
    }
    __syncthreads();
    float flo[NQ], fhi[NQ];
    qsel_own_pivots<NQ, kSampleThreads>(x + seg * seg_len, seg_len, seg, q, flo, fhi);
    if ((int)threadIdx.x < NQ) {
        QTask* t = &w.tasks[seg * NQ + threadIdx.x];
        t->flo = flo[threadIdx.x];
        t->fhi = fhi[threadIdx.x];
    }
}

template <int NQ>
__global__ void __launch_bounds__(kThreads, 4) qsel_partition_kernel(const float* __restrict__ x, int64_t seg_len, int bps,
                                                                  int64_t chunk, const double* __restrict__ q,
                                                                  SWorkspace w, int64_t n_seg) {
    __shared__ float s_buf[NQ][kStage];
    __shared__ unsigned s_n[NQ], s_cnt[NQ][3], s_nan, s_base[NQ], s_kmax[NQ], s_nkmin[NQ];
    const int64_t seg = blockIdx.y;
    const float* xs = x + seg * seg_len;
    QTask* tasks = w.tasks + seg * NQ;
    float flo[NQ], fhi[NQ];
    be_pdl_wait();   // the pivots
#pragma unroll
    for (int i = 0; i < NQ; ++i) {
        flo[i] = tasks[i].flo;
        fhi[i] = tasks[i].fhi;
    }
    if (threadIdx.x < NQ) {
        s_n[threadIdx.x] = 0;
        s_cnt[threadIdx.x][0] = s_cnt[threadIdx.x][1] = s_cnt[threadIdx.x][2] = 0;
        s_kmax[threadIdx.x] = s_nkmin[threadIdx.x] = 0;
    }
    if (threadIdx.x == 0) s_nan = 0;
    __syncthreads();

    unsigned below[NQ], eqlo[NQ], eqhi[NQ], nans = 0;
#pragma unroll
    for (int i = 0; i < NQ; ++i) below[i] = eqlo[i] = eqhi[i] = 0;

    // Two brackets that do not overlap (the usual case: a low and a high percentile) leave ~96 % of the elements strictly
    // between them, and such an element only counts as "below the upper bracket": two compares and one add.  Everything
    // else — below or inside the lower bracket, inside or above the upper one, NaN — goes through the rare path, which
    // then does all the counting for it.
    const bool fastmid = (NQ == 2) && (fhi[0] < flo[NQ - 1]);

    // Work is dealt out in spans of 4096 elements (16 per thread), span j going to block j mod bps: a block's spans are
    // spread over the whole segment.  With one contiguous chunk per block the blocks that happened to hold a bright
    // structure (most of an image's top percentile sits in a few places) had several times the rare-path work of the
    // others, and a kernel that is one wave of blocks ends with its slowest block.
    constexpr int64_t kSpan = (int64_t)kThreads * 16;
    constexpr int kSpansPerRound = 8;   // staged candidates move to the global lists every 8 spans (32 K elements)
    unsigned kx[NQ], nk[NQ];   // key range of the candidates this thread moved
#pragma unroll
    for (int i = 0; i < NQ; ++i) kx[i] = nk[i] = 0;
    const int head = (int)(((16 - ((uintptr_t)xs & 15)) & 15) >> 2);   // elements before the first 16-byte boundary
    const int64_t n_body = (seg_len > head) ? seg_len - head : 0;
    const int64_t n_spans = (n_body + kSpan - 1) / kSpan;

    auto stream = [&](auto fast_tag) {
        constexpr bool FAST = decltype(fast_tag)::value;
        // rare path, one element: ties with the pivots are counted, the rest of the bracket is appended
        auto take_inside = [&](float v) {
            if (v != v) {   // NaNs are flagged by the common path (they fail its comparisons) and counted here
                ++nans;
                return;
            }
            if (FAST) {
                if (v <= fhi[0]) {
                    ++below[NQ - 1];
                    if (v < flo[0]) {
                        ++below[0];
                        return;
                    }
                } else if (v > fhi[NQ - 1]) {
                    return;
                }
            }
#pragma unroll
            for (int i = 0; i < NQ; ++i) {
                if (v >= flo[i] && v <= fhi[i]) {
                    if (v == flo[i]) {
                        ++eqlo[i];
                    } else if (v == fhi[i]) {
                        ++eqhi[i];
                    } else {
                        const unsigned pos = atomicAdd(&s_n[i], 1u);
                        if (pos < kStage) {
                            s_buf[i][pos] = v;
                        } else {  // staging full: straight to the global list (still counted if that is full too)
                            const unsigned g = atomicAdd(&tasks[i].n_mid, 1u);
                            if (g < w.cap) w.cand[(seg * NQ + i) * w.cap + g] = v;
                            atomicMax(&s_kmax[i], f32_to_key(v));
                            atomicMax(&s_nkmin[i], ~f32_to_key(v));
                        }
                    }
                }
            }
        };
        // common path, one element; returns whether the element needs the rare path
        auto classify = [&](float v) -> bool {
            if (FAST) {
                const bool mid = (v > fhi[0]) & (v < flo[NQ - 1]);
                below[NQ - 1] += mid;
                return !mid;
            }
            bool flagged = false;
#pragma unroll
            for (int i = 0; i < NQ; ++i) {
                const bool lt = v < flo[i];
                below[i] += lt;
                flagged |= !(lt | (v > fhi[i]));   // inside the bracket — or NaN, which the rare path sorts out
            }
            return flagged;
        };
        // the staged candidates move to the global lists (one reservation per block and bracket)
        auto flush = [&]() {
            __syncthreads();   // staging is complete
            if (threadIdx.x < NQ) {
                const unsigned n_st = min(s_n[threadIdx.x], (unsigned)kStage);
                s_base[threadIdx.x] = n_st ? atomicAdd(&tasks[threadIdx.x].n_mid, n_st) : 0u;
            }
            __syncthreads();
#pragma unroll
            for (int i = 0; i < NQ; ++i) {
                const unsigned n_st = min(s_n[i], (unsigned)kStage), base = s_base[i];
                float* dst = w.cand + (seg * NQ + i) * w.cap;
                for (unsigned k = threadIdx.x; k < n_st; k += kThreads) {
                    const float v = s_buf[i][k];
                    if (base + k < w.cap) dst[base + k] = v;
                    kx[i] = max(kx[i], f32_to_key(v));
                    nk[i] = max(nk[i], ~f32_to_key(v));
                }
            }
            __syncthreads();   // staging read out: free again
            if (threadIdx.x < NQ) s_n[threadIdx.x] = 0;
            __syncthreads();
        };

        if (blockIdx.x == 0 && (int)threadIdx.x < head && (int64_t)threadIdx.x < seg_len)
            if (classify(xs[threadIdx.x])) take_inside(xs[threadIdx.x]);
        const float* body = xs + head;
        int in_round = 0;
        for (int64_t span = blockIdx.x; span < n_spans; span += bps) {
            const float* p = body + span * kSpan;
            const int64_t n = min(kSpan, n_body - span * kSpan);
            // 16 elements per thread: all of them take the cheap path first, and the few per cent that need more are
            // handled afterwards from a 16-bit flag word, so the divergent code runs once per 16 elements
            const float4* p4 = reinterpret_cast<const float4*>(p);
            const int n4 = (int)(n >> 2);
            float4 v[4];
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if ((int)threadIdx.x + k * kThreads < n4) v[k] = __ldg(p4 + threadIdx.x + k * kThreads);   // allocating: the rare path re-reads from L1
            unsigned flags = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if ((int)threadIdx.x + k * kThreads < n4) {
                    flags |= (unsigned)classify(v[k].x) << (4 * k + 0);
                    flags |= (unsigned)classify(v[k].y) << (4 * k + 1);
                    flags |= (unsigned)classify(v[k].z) << (4 * k + 2);
                    flags |= (unsigned)classify(v[k].w) << (4 * k + 3);
                }
            }
            while (flags) {  // (measured: faster than 16 statically indexed copies of the rare path)
                const int b = __ffs(flags) - 1;
                flags &= flags - 1;
                // the element is read again (an L1 hit) rather than picked out of v[]: indexing the register array
                // dynamically would put all of v[] in local memory, with four 16-byte stores on the common path
                take_inside(__ldg(p + (((int)threadIdx.x + (b >> 2) * kThreads) << 2) + (b & 3)));
            }
            for (int i = (n4 << 2) + threadIdx.x; i < n; i += kThreads)   // the segment's last, ragged span
                if (classify(p[i])) take_inside(p[i]);
            if (++in_round == kSpansPerRound) {
                flush();
                in_round = 0;
            }
        }
        flush();
    };
    if (fastmid)
        stream(std::true_type{});
    else
        stream(std::false_type{});
#pragma unroll
    for (int i = 0; i < NQ; ++i) {
        const unsigned b = __reduce_add_sync(0xffffffffu, below[i]);
        const unsigned l = __reduce_add_sync(0xffffffffu, eqlo[i]);
        const unsigned h = __reduce_add_sync(0xffffffffu, eqhi[i]);
        if ((threadIdx.x & 31) == 0) {
            if (b) atomicAdd(&s_cnt[i][0], b);
            if (l) atomicAdd(&s_cnt[i][1], l);
            if (h) atomicAdd(&s_cnt[i][2], h);
        }
    }
    nans = __reduce_add_sync(0xffffffffu, nans);
    if ((threadIdx.x & 31) == 0 && nans) atomicAdd(&s_nan, nans);
    __syncthreads();
    if (threadIdx.x < NQ) {
        const int i = threadIdx.x;
        if (s_cnt[i][0]) atomicAdd(&tasks[i].below, s_cnt[i][0]);
        if (s_cnt[i][1]) atomicAdd(&tasks[i].eq_lo, s_cnt[i][1]);
        if (s_cnt[i][2]) atomicAdd(&tasks[i].eq_hi, s_cnt[i][2]);
    }
    if (threadIdx.x == 0 && s_nan) atomicAdd(&w.nan_count[seg], s_nan);
    // the key range of the candidates (what the selection cuts its digits from)
#pragma unroll
    for (int i = 0; i < NQ; ++i) {
        if (__any_sync(0xffffffffu, kx[i] | nk[i])) {
            const unsigned a = __reduce_max_sync(0xffffffffu, kx[i]), b = __reduce_max_sync(0xffffffffu, nk[i]);
            if ((threadIdx.x & 31) == 0) {
                atomicMax(&s_kmax[i], a);
                atomicMax(&s_nkmin[i], b);
            }
        }
    }
    __syncthreads();
    if (threadIdx.x < NQ && (s_kmax[threadIdx.x] | s_nkmin[threadIdx.x])) {
        atomicMax(&tasks[threadIdx.x].kmax, s_kmax[threadIdx.x]);
        atomicMax(&tasks[threadIdx.x].nkmin, s_nkmin[threadIdx.x]);
    }

    if (!last_block_of_segment(&w.counters[seg], bps)) return;

    // all counts of this segment are final: turn each bracket into a selection task (fields are read into registers
    // up front and the task is written back once, so the serial tail is a handful of memory round trips)
    if (threadIdx.x < NQ) {
        const int i = threadIdx.x;
        QTask* t = &tasks[i];
        const uint32_t n_nan = __ldcg(&w.nan_count[seg]);
        const uint32_t below_c = __ldcg(&t->below), eq_lo = __ldcg(&t->eq_lo), eq_hi = __ldcg(&t->eq_hi);
        const uint32_t n_mid = __ldcg(&t->n_mid);
        const uint32_t kmax = __ldcg(&t->kmax), kmin = ~__ldcg(&t->nkmin);
        const float tflo = flo[i], tfhi = fhi[i];
        const double qi = q[i];
        const uint32_t n_valid = (uint32_t)seg_len - n_nan;
        // numpy: virtual index q*(n-1) in fp64, neighbours floor / floor+1 clipped to n-1
        double h = (n_valid > 0) ? qi * (double)(n_valid - 1) : 0.0;
        if (h < 0.0) h = 0.0;
        const double lo = floor(h);
        uint32_t ilo = (uint32_t)lo;
        if (n_valid > 0 && ilo > n_valid - 1) ilo = n_valid - 1;
        const uint32_t ihi = (n_valid > 0 && ilo + 1 < n_valid) ? ilo + 1 : ilo;
        const uint32_t inside = eq_lo + n_mid + eq_hi;
        const bool ok = n_mid <= (uint32_t)w.cap && ilo >= below_c && ihi < below_c + inside;
        QTask r;
        r.gamma = h - lo;
        r.flo = tflo;
        r.fhi = tfhi;
        r.below = below_c;
        r.eq_lo = eq_lo;
        r.eq_hi = eq_hi;
        r.n_mid = n_mid;
        r.kmax = r.nkmin = 0;
        if (ok || n_valid == 0) {
            r.src = w.cand + (seg * NQ + i) * w.cap;
            r.len = n_valid ? n_mid : 0;
            r.rank[0] = ilo - below_c;
            r.rank[1] = ihi - below_c;
            r.vkey[0] = f32_to_key(tflo);
            r.vkey[1] = f32_to_key(tfhi);
            r.vcnt[0] = eq_lo;
            r.vcnt[1] = eq_hi;
            r.total = n_valid ? inside : 0;
            r.klo = (n_mid && kmax >= kmin) ? kmin : 0u;
            r.range = (n_mid && kmax >= kmin) ? kmax - kmin : 0u;
        } else {  // bracket missed or list overflowed: select on the whole segment instead
            r.klo = 0;
            r.range = 0xffffffffu;
            r.src = xs;
            r.len = (uint32_t)seg_len;
            r.rank[0] = ilo;
            r.rank[1] = ihi;
            r.vkey[0] = r.vkey[1] = 0;
            r.vcnt[0] = r.vcnt[1] = 0;
            r.total = n_valid;
        }
        *t = r;
        __threadfence();
    }
}

// ---- the selection stage as ONE launch: a thread-block cluster per task ----
//
// The radix levels only ever look at the ~1-2 % of a segment the partition pass kept, so relaunching a grid per level
// costs more in launch and completion-ticket latency than the passes themselves.  Here a cluster of kClusterCtas CTAs
// owns a task and the levels are separated by cluster barriers:
//   * digits are taken from (key - key(flo)), top bits of the bracket's actual width first, so the first level already
//     spreads the candidates over all 2048 bins (plain key digits would put a narrow bracket into one or two bins);
//   * every CTA histograms its slice of the list in shared memory; the eight histograms are summed through distributed
//     shared memory (each CTA sums 1/8 of the bins and stores the sums into the leader's table), the leader adds the
//     virtual ties, finds the bins of the two ranks and publishes the new prefixes;
//   * after the first level each CTA keeps only the elements of the selected bins (~1/2048 of its slice) in a small
//     shared-memory list, so the remaining levels cost next to nothing.  A CTA whose selected elements do not fit the
//     list (heavy ties, or a fallback to a whole segment) streams its slice again instead.
constexpr int kClusterMax = 16;   // CTAs per cluster: 16 where the device grants it (non-portable size), else 8
constexpr int kSelThreads = 1024;
constexpr int kSmallCap = 3072;

struct ClusterShared {
    unsigned sh[2][kBins];    // this CTA's histograms (one per group)
    unsigned tot[2][kBins];   // leader: the cluster's histograms
    float small[kSmallCap];   // elements of the bins selected at level 0
    unsigned n_small;
    QState st;
    QTask t;
    uint32_t fixed_key[2];    // ranks that fall on a tie with a pivot need no selection: their key is known
    int fixed[2];
    uint32_t s_warp[kSelThreads / 32];
    uint32_t s_res[2];
};

// block-cooperative (kSelThreads): the bin holding the k-th (0-based) element of a shared-memory histogram
__device__ void find_bin_sm(ClusterShared& S, const unsigned* hist, uint32_t k, uint32_t* bin_out, uint32_t* krem_out) {
    constexpr int per = kBins / kSelThreads;
    uint32_t c[per];
    uint32_t local = 0;
#pragma unroll
    for (int i = 0; i < per; ++i) {
        c[i] = hist[threadIdx.x * per + i];
        local += c[i];
    }
    uint32_t incl = local;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if ((threadIdx.x & 31) >= o) incl += t;
    }
    __syncthreads();  // protects s_warp / s_res from the previous call
    if ((threadIdx.x & 31) == 31) S.s_warp[threadIdx.x >> 5] = incl;
    if (threadIdx.x == 0) S.s_res[0] = S.s_res[1] = 0;   // k beyond the histogram (an empty list): bin 0
    __syncthreads();
    // exclusive prefix of the warp totals: every warp scans the 32 totals itself (one shuffle scan, no serial loop)
    uint32_t wt = S.s_warp[threadIdx.x & 31], winc = wt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, winc, o);
        if ((threadIdx.x & 31) >= o) winc += t;
    }
    const uint32_t base = __shfl_sync(0xffffffffu, winc - wt, threadIdx.x >> 5);
    const uint32_t excl = base + incl - local;
    if (k >= excl && k < excl + local) {
        uint32_t r = k - excl;
#pragma unroll
        for (int i = 0; i < per; ++i) {
            if (r < c[i]) {
                S.s_res[0] = threadIdx.x * per + i;
                S.s_res[1] = r;
                break;
            }
            r -= c[i];
        }
    }
    __syncthreads();
    *bin_out = S.s_res[0];
    *krem_out = S.s_res[1];
}

// how the bracket's key range is cut into digits: level l takes `bits[l]` bits at `shift[l]` of (key - klo)
struct DigitPlan {
    uint32_t klo;
    int shift[3], bits[3];
};

__device__ __forceinline__ DigitPlan make_plan(const QTask& t) {
    DigitPlan p;
    p.klo = t.klo;
    int left = 32 - __clz(t.range);
    const int want[3] = {11, 11, 10};
#pragma unroll
    for (int l = 0; l < 3; ++l) {
        p.bits[l] = min(want[l], left);
        left -= p.bits[l];
        p.shift[l] = left;
    }
    return p;   // 32 bits are always used up by 11 + 11 + 10
}

// apply f to the elements of this CTA's slice: the small list if it is valid, the source otherwise
template <class F>
__device__ __forceinline__ void slice_foreach(const float* xs, int64_t n, F f) {
    int64_t head = ((16 - ((uintptr_t)xs & 15)) & 15) >> 2;
    if (head > n) head = n;
    if ((int64_t)threadIdx.x < head) f(xs[threadIdx.x]);
    const float4* p4 = reinterpret_cast<const float4*>(xs + head);
    const int64_t n4 = (n - head) >> 2;
    int64_t i = threadIdx.x;
    for (; i + 3 * kSelThreads < n4; i += 4 * kSelThreads) {   // four loads in flight: a slice is only a few rounds long
        const float4 a = ld_stream_f4(p4 + i);
        const float4 b = ld_stream_f4(p4 + i + kSelThreads);
        const float4 c = ld_stream_f4(p4 + i + 2 * kSelThreads);
        const float4 e = ld_stream_f4(p4 + i + 3 * kSelThreads);
        f(a.x); f(a.y); f(a.z); f(a.w);
        f(b.x); f(b.y); f(b.z); f(b.w);
        f(c.x); f(c.y); f(c.z); f(c.w);
        f(e.x); f(e.y); f(e.z); f(e.w);
    }
    for (; i + kSelThreads < n4; i += 2 * kSelThreads) {
        const float4 a = ld_stream_f4(p4 + i);
        const float4 b = ld_stream_f4(p4 + i + kSelThreads);
        f(a.x); f(a.y); f(a.z); f(a.w);
        f(b.x); f(b.y); f(b.z); f(b.w);
    }
    for (; i < n4; i += kSelThreads) {
        const float4 v = ld_stream_f4(p4 + i);
        f(v.x); f(v.y); f(v.z); f(v.w);
    }
    const int64_t done = head + (n4 << 2);
    if (done + (int64_t)threadIdx.x < n) f(xs[done + threadIdx.x]);
}

template <int LEVEL>
__device__ __forceinline__ void cluster_level(cg::cluster_group& cluster, ClusterShared& S, const DigitPlan& plan,
                                              const float* xs, int64_t n, bool use_small) {
    constexpr int kGroups = (LEVEL == 0) ? 1 : 2;
    const unsigned rank = cluster.block_rank();
    const int shift = plan.shift[LEVEL], hi_shift = (LEVEL == 0) ? 0 : plan.shift[LEVEL - 1];
    const uint32_t mask = (1u << plan.bits[LEVEL]) - 1u;
    const uint32_t klo = plan.klo;
    const int n_groups = (LEVEL == 0) ? 1 : S.st.n_groups;
    uint32_t gp[kGroups];
#pragma unroll
    for (int g = 0; g < kGroups; ++g) gp[g] = (LEVEL == 0) ? 0u : ((g < n_groups) ? S.st.gprefix[g] : 0xffffffffu);
    for (int i = threadIdx.x; i < n_groups * kBins; i += kSelThreads) {
        (&S.sh[0][0])[i] = 0;
        if (rank == 0) (&S.tot[0][0])[i] = 0;
    }
    __syncthreads();
    RunCounter rc[kGroups];
    auto count = [&](float v) {
        if (v == v) {  // a fallback source may hold NaN
            const uint32_t d = f32_to_key(v) - klo;
            const uint32_t digit = (d >> shift) & mask;
            if (LEVEL == 0) {
                rc[0].add(S.sh[0], digit);
            } else {
                const uint32_t hi = (hi_shift >= 32) ? 0u : (d >> hi_shift);
#pragma unroll
                for (int g = 0; g < kGroups; ++g)
                    if (hi == gp[g]) rc[g].add(S.sh[g], digit);
            }
        }
    };
    if (use_small) {
        for (unsigned i = threadIdx.x; i < S.n_small; i += kSelThreads) count(S.small[i]);
    } else {
        slice_foreach(xs, n, count);
    }
#pragma unroll
    for (int g = 0; g < kGroups; ++g) rc[g].flush(S.sh[g]);
    cluster.sync();   // every CTA's histogram is complete, the leader's table is zeroed
    {
        // push this CTA's non-empty bins into the leader's table: remote shared-memory atomics, nothing waits on them
        unsigned* lead_tot = cluster.map_shared_rank(&S.tot[0][0], 0);
        for (int i = threadIdx.x; i < n_groups * kBins; i += kSelThreads) {
            const unsigned c = (&S.sh[0][0])[i];
            if (c) atomicAdd(lead_tot + i, c);
        }
    }
    cluster.sync();   // the leader's table is complete
    if (rank == 0) {
        {
            for (int j = 0; j < 2; ++j) {
                uint32_t bin, krem;
                find_bin_sm(S, &S.tot[0][0] + S.st.group[j] * kBins, S.st.krem[j], &bin, &krem);
                __syncthreads();
                if (threadIdx.x == 0) {
                    S.st.prefix[j] = (S.st.prefix[j] << plan.bits[LEVEL]) | bin;
                    S.st.krem[j] = krem;
                }
            }
        }
        __syncthreads();
        if (LEVEL < 2 && threadIdx.x == 0) {
            regroup(&S.st);
            if (LEVEL == 0) {
                // few enough elements left in the selected bins: they are gathered into the leader's list and the leader
                // finishes alone, without cluster barriers
                const uint32_t b0 = S.st.gprefix[0], b1 = (S.st.n_groups > 1) ? S.st.gprefix[1] : b0;
                const unsigned left = S.tot[0][b0] + ((b1 != b0) ? S.tot[0][b1] : 0u);
                S.st.pad = left <= (unsigned)kSmallCap;
            }
        }
    }
    cluster.sync();   // the leader's state is final
    if (LEVEL < 2 && rank != 0) {
        if (threadIdx.x == 0) S.st = *cluster.map_shared_rank(&S.st, 0);
        __syncthreads();
    }
}

// a level on the leader's gathered list: block-local, no cluster barrier
template <int LEVEL>
__device__ __forceinline__ void leader_level(ClusterShared& S, const DigitPlan& plan) {
    const int shift = plan.shift[LEVEL], hi_shift = plan.shift[LEVEL - 1];
    const uint32_t mask = (1u << plan.bits[LEVEL]) - 1u, klo = plan.klo;
    const int n_groups = S.st.n_groups;
    for (int i = threadIdx.x; i < n_groups * kBins; i += kSelThreads) (&S.tot[0][0])[i] = 0;
    __syncthreads();
    for (unsigned i = threadIdx.x; i < S.n_small; i += kSelThreads) {
        const uint32_t d = f32_to_key(S.small[i]) - klo;
        const uint32_t hi = d >> hi_shift;
        for (int g = 0; g < n_groups; ++g)
            if (hi == S.st.gprefix[g]) atomicAdd(&S.tot[g][(d >> shift) & mask], 1u);
    }
    __syncthreads();
    for (int j = 0; j < 2; ++j) {
        uint32_t bin, krem;
        find_bin_sm(S, &S.tot[0][0] + S.st.group[j] * kBins, S.st.krem[j], &bin, &krem);
        __syncthreads();
        if (threadIdx.x == 0) {
            S.st.prefix[j] = (S.st.prefix[j] << plan.bits[LEVEL]) | bin;
            S.st.krem[j] = krem;
        }
    }
    __syncthreads();
    if (LEVEL < 2 && threadIdx.x == 0) regroup(&S.st);
    __syncthreads();
}

__global__ void __launch_bounds__(kSelThreads)
    qsel_cluster_kernel(int n_q, double eps, SWorkspace w, double* __restrict__ quant, double* __restrict__ norm) {
    __shared__ ClusterShared S;
    cg::cluster_group cluster = cg::this_cluster();
    const unsigned r = cluster.block_rank();
    const int64_t task = blockIdx.y;
    be_pdl_wait();   // the partition pass
    if (threadIdx.x == 0) {
        S.t = w.tasks[task];
        S.st.n_ranks = 2;
        S.st.n_valid = S.t.total;
        S.st.gamma[0] = S.t.gamma;
        S.st.n_groups = 1;
        S.st.gprefix[0] = 0;
        S.st.group[0] = S.st.group[1] = 0;
        S.st.prefix[0] = S.st.prefix[1] = 0;
        // the multiset is {vcnt[0] x flo} + list + {vcnt[1] x fhi}: a rank inside a run of ties is settled right away
        for (int j = 0; j < 2; ++j) {
            const uint32_t k = S.t.rank[j];
            S.fixed[j] = 1;
            if (k < S.t.vcnt[0]) {
                S.fixed_key[j] = S.t.vkey[0];
            } else if (k - S.t.vcnt[0] < S.t.len) {
                S.fixed[j] = 0;
                S.st.krem[j] = k - S.t.vcnt[0];
            } else {
                S.fixed_key[j] = S.t.vkey[1];
            }
        }
        // a settled rank shadows the other one through the levels (same histogram, result ignored)
        if (S.fixed[0] && !S.fixed[1]) S.st.krem[0] = S.st.krem[1];
        if (S.fixed[1] && !S.fixed[0]) S.st.krem[1] = S.st.krem[0];
        S.n_small = 0;
    }
    __syncthreads();
    const DigitPlan plan = make_plan(S.t);
    // this CTA's slice of the source
    const int64_t len = S.t.len;
    const int64_t n_cta = cluster.num_blocks();
    int64_t per = (len + n_cta - 1) / n_cta;
    per = (per + 3) & ~(int64_t)3;
    const int64_t begin = min(len, (int64_t)r * per), n = min(len, begin + per) - begin;
    const float* xs = S.t.src + begin;

    cluster_level<0>(cluster, S, plan, xs, n, false);

    const uint32_t b0 = S.st.gprefix[0], b1 = (S.st.n_groups > 1) ? S.st.gprefix[1] : b0;
    const int shift0 = plan.shift[0];
    const uint32_t klo0 = plan.klo, mask0 = (1u << plan.bits[0]) - 1u;
    if (S.st.pad) {
        // the usual case: the elements of the selected level-0 bins (~1/2048 of the list) go to the leader's list
        // through distributed shared memory, and the leader runs the remaining levels on its own
        const unsigned mine = S.sh[0][b0] + ((b1 != b0) ? S.sh[0][b1] : 0u);   // this CTA's level-0 counts
        if (mine > 0) {
            float* lead_small = cluster.map_shared_rank(&S.small[0], 0);
            unsigned* lead_n = cluster.map_shared_rank(&S.n_small, 0);
            slice_foreach(xs, n, [&](float v) {
                if (v == v) {
                    const uint32_t digit = ((f32_to_key(v) - klo0) >> shift0) & mask0;
                    if (digit == b0 || digit == b1) lead_small[atomicAdd(lead_n, 1u)] = v;
                }
            });
        }
        cluster.sync();   // the leader's list is complete
        if (r != 0) return;
        leader_level<1>(S, plan);
        leader_level<2>(S, plan);
    } else {
        // heavy ties or a fallback to a whole segment: every CTA keeps its share if that fits, else streams again
        const unsigned mine = S.sh[0][b0] + ((b1 != b0) ? S.sh[0][b1] : 0u);
        const bool use_small = mine <= (unsigned)kSmallCap;
        if (use_small && mine > 0) {
            slice_foreach(xs, n, [&](float v) {
                if (v == v) {
                    const uint32_t digit = ((f32_to_key(v) - klo0) >> shift0) & mask0;
                    if (digit == b0 || digit == b1) S.small[atomicAdd(&S.n_small, 1u)] = v;
                }
            });
        }
        __syncthreads();
        cluster_level<1>(cluster, S, plan, xs, n, use_small);
        cluster_level<2>(cluster, S, plan, xs, n, use_small);
    }
    if (r != 0 || threadIdx.x != 0) return;
    double res;
    if (S.t.total == 0) {
        res = __longlong_as_double(0x7ff8000000000000LL);
    } else {
        // numpy _lerp: fp32 difference, fp64 product and sum, switch of anchor at gamma >= 0.5
        const float a = key_to_f32(S.fixed[0] ? S.fixed_key[0] : plan.klo + S.st.prefix[0]);
        const float b = key_to_f32(S.fixed[1] ? S.fixed_key[1] : plan.klo + S.st.prefix[1]);
        const double d = (double)__fsub_rn(b, a);
        const double g = S.st.gamma[0];
        res = (g >= 0.5) ? __dsub_rn((double)b, __dmul_rn(d, __dsub_rn(1.0, g))) : __dadd_rn((double)a, __dmul_rn(d, g));
    }
    quant[task] = res;
    if (norm) {  // the task of a segment that finishes second writes {q0, q1 - q0 + eps}
        __threadfence();
        const int64_t seg = task / n_q;
        if (atomicAdd(&w.seg_done[seg], 1u) == (unsigned)(n_q - 1)) {
            __threadfence();
            const double lo = __ldcg(&quant[seg * n_q + 0]), hi = __ldcg(&quant[seg * n_q + 1]);
            norm[seg * 2 + 0] = lo;
            norm[seg * 2 + 1] = __dadd_rn(__dsub_rn(hi, lo), eps);
        }
    }
}

// ---- ensure_dtype(float32) on the device ----
template <typename T>
__device__ __forceinline__ float to_f32(T v) { return (float)v; }
template <>
__device__ __forceinline__ float to_f32<int32_t>(int32_t v) { return __int2float_rn(v); }
template <>
__device__ __forceinline__ float to_f32<uint32_t>(uint32_t v) { return __uint2float_rn(v); }
template <>
__device__ __forceinline__ float to_f32<double>(double v) { return __double2float_rn(v); }
template <>
__device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }

// 16 input bytes per thread and iteration (one 16-byte load), 16-byte stores
template <typename T>
__global__ void __launch_bounds__(kThreads) cast_f32_kernel(const T* __restrict__ x, float* __restrict__ y, int64_t n) {
    constexpr int kPer = 16 / (int)sizeof(T);
    const int64_t stride = (int64_t)gridDim.x * kThreads;
    const int64_t nv = n / kPer;
    if (((((uintptr_t)x) | ((uintptr_t)y)) & 15) == 0) {
        for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < nv; i += stride) {
            const int4 raw = ld_stream_i4(reinterpret_cast<const int4*>(x) + i);
            const T* e = reinterpret_cast<const T*>(&raw);
            float out[kPer];
#pragma unroll
            for (int j = 0; j < kPer; ++j) out[j] = to_f32<T>(e[j]);
            if (kPer >= 4) {
#pragma unroll
                for (int j = 0; j < kPer; j += 4)
                    st_stream_f4(reinterpret_cast<float4*>(y + i * kPer + j), make_float4(out[j], out[j + 1], out[j + 2], out[j + 3]));
            } else {
#pragma unroll
                for (int j = 0; j < kPer; ++j) y[i * kPer + j] = out[j];
            }
        }
        for (int64_t i = nv * kPer + (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += stride) y[i] = to_f32<T>(x[i]);
    } else {
        for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += stride) y[i] = to_f32<T>(x[i]);
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

// plain three-pass radix select on whole segments (small segments, and the library-resident-scratch entry point)
static int quantile_radix(const float* x, int64_t n_seg, int64_t seg_len, const double* q, int n_q, double eps,
                          double* quant, double* norm, void* ws, cudaStream_t s) {
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

// Largest cluster the selection kernel can be launched with on the current device: 16 CTAs if the device grants the
// non-portable size and can keep at least two such clusters resident, else 8; 0 if neither works.  Asked once per device.
static int select_cluster_size() {
    static int cached[64];   // 0 = not asked yet, -1 = no cluster possible
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 0;
    if (cached[dev]) return cached[dev] > 0 ? cached[dev] : 0;
    int best = -1;
    // BE_B200_CLUSTER_MAX=8 (test hook): take the portable cluster size even where 16 is available
    const char* cap = getenv("BE_B200_CLUSTER_MAX");
    const int first = (cap && atoi(cap) == 8) ? 8 : kClusterMax;
    for (int size = first; size >= 8 && best < 0; size >>= 1) {
        if (size > 8 &&
            cudaFuncSetAttribute(qsel_cluster_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) {
            cudaGetLastError();
            continue;
        }
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)size, 2);
        cfg.blockDim = dim3(kSelThreads);
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = (unsigned)size;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        int n = 0;
        if (cudaOccupancyMaxActiveClusters(&n, qsel_cluster_kernel, &cfg) == cudaSuccess && n >= (size > 8 ? 2 : 1))
            best = size;
        else
            cudaGetLastError();
    }
    cached[dev] = best;
    return best > 0 ? best : 0;
}

// sampled-pivot select: one pass over the data + radix select on ~2 % of it, `per_pass` segments at a time
static int quantile_sampled(const float* x, int64_t n_seg, int64_t seg_len, const double* q, int n_q, double eps,
                            double* quant, double* norm, void* ws, cudaStream_t s) {
    const int64_t per_pass = sample_segments_per_pass(n_seg, seg_len);
    for (int64_t s0 = 0; s0 < n_seg; s0 += per_pass) {
        const int64_t ns = (n_seg - s0 < per_pass) ? (n_seg - s0) : per_pass;
        const int64_t nt = ns * n_q;
        // the partition kernel keeps 4 blocks of 60 registers per SM resident and a block runs for most of the kernel:
        // exactly one wave; a block's spans are interleaved with the other blocks' (see the kernel)
        const SegSplit sp = split_segment(ns, seg_len, (int64_t)BE_SM_COUNT * 4);
        const SWorkspace w = carve_sample_ws(ws, ns, n_q, seg_len);
        const float* xs = x + s0 * seg_len;
        double* qs = quant + s0 * n_q;
        double* nr = norm ? norm + s0 * 2 : nullptr;
        const dim3 grid(sp.bps, (unsigned)ns);
        if (seg_len >= kSampleLargeLen) {   // volumes: an 8192-element sample, sorted by a kernel of its own
            BE_CUDA(cudaMemsetAsync(ws, 0, w.zero_bytes, s));
            qsel_sample_kernel<kSampleLarge><<<(unsigned)ns, kSampleThreads, 0, s>>>(xs, seg_len, q, n_q, w);
            BE_LAUNCH_CHECK();
            if (n_q == 1)
                BE_CUDA(be_launch_pdl(qsel_partition_kernel<1>, grid, dim3(kThreads), 0, s, xs, seg_len, sp.bps, sp.chunk, q, w, ns));
            else
                BE_CUDA(be_launch_pdl(qsel_partition_kernel<2>, grid, dim3(kThreads), 0, s, xs, seg_len, sp.bps, sp.chunk, q, w, ns));
        } else if (n_q == 1) {              // images: pivots by a radix select on the 2048-element sample
            BE_CUDA(be_launch_pdl(qsel_pivot_kernel<1>, dim3((unsigned)ns), dim3(kSampleThreads), 0, s, xs, seg_len, q, w));
            BE_CUDA(be_launch_pdl(qsel_partition_kernel<1>, grid, dim3(kThreads), 0, s, xs, seg_len, sp.bps, sp.chunk, q, w, ns));
        } else {
            BE_CUDA(be_launch_pdl(qsel_pivot_kernel<2>, dim3((unsigned)ns), dim3(kSampleThreads), 0, s, xs, seg_len, q, w));
            BE_CUDA(be_launch_pdl(qsel_partition_kernel<2>, grid, dim3(kThreads), 0, s, xs, seg_len, sp.bps, sp.chunk, q, w, ns));
        }
        {
            const int n_cta = select_cluster_size();
            BE_REQUIRE(n_cta > 0, BE_ERR_CUDA, "be_quantile_f32: the device cannot run a thread-block cluster of 8");
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3((unsigned)n_cta, (unsigned)nt);
            cfg.blockDim = dim3(kSelThreads);
            cfg.dynamicSmemBytes = 0;
            cfg.stream = s;
            cudaLaunchAttribute attr[2];
            attr[0].id = cudaLaunchAttributeClusterDimension;
            attr[0].val.clusterDim.x = (unsigned)n_cta;
            attr[0].val.clusterDim.y = 1;
            attr[0].val.clusterDim.z = 1;
            attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[1].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = attr;
            cfg.numAttrs = be_pdl_enabled() ? 2 : 1;
            BE_CUDA(cudaLaunchKernelEx(&cfg, qsel_cluster_kernel, n_q, eps, w, qs, nr));
        }
    }
    return BE_OK;
}

extern "C" size_t be_quantile_workspace_bytes(int64_t n_seg, int64_t seg_len) {
    if (n_seg <= 0) return 256;
    if (seg_len >= kSampleMinLen)
        return carve_sample_ws(nullptr, sample_segments_per_pass(n_seg, seg_len), 2, seg_len).total_bytes;
    return carve_quantile_ws(nullptr, n_seg).total_bytes;
}

static int quantile_check(const char* who, const float* x, int64_t n_seg, int64_t seg_len, const double* q, int n_q,
                          const double* quant, const double* norm, const void* ws) {
    BE_REQUIRE(x && q && quant && ws, BE_ERR_INVALID_ARG, "%s: null pointer", who);
    BE_REQUIRE(n_seg > 0 && seg_len > 0, BE_ERR_INVALID_ARG, "%s: empty input (n_seg=%lld seg_len=%lld)", who,
               (long long)n_seg, (long long)seg_len);
    BE_REQUIRE(n_q >= 1 && 2 * n_q <= kMaxRanks, BE_ERR_INVALID_ARG, "%s: n_q must be 1 or 2, got %d", who, n_q);
    BE_REQUIRE(!norm || n_q == 2, BE_ERR_INVALID_ARG, "%s: norm output needs n_q == 2", who);
    BE_REQUIRE(seg_len < (int64_t)1 << 31, BE_ERR_TOO_LARGE, "%s: seg_len %lld >= 2^31", who, (long long)seg_len);
    BE_REQUIRE(n_seg <= 65535, BE_ERR_TOO_LARGE, "%s: n_seg %lld > 65535", who, (long long)n_seg);
    return BE_OK;
}

extern "C" int be_quantile_f32(const float* x, int64_t n_seg, int64_t seg_len, const double* q, int n_q, double eps,
                               double* quant, double* norm, void* ws, be_stream_t stream) {
    const int rc = quantile_check("be_quantile_f32", x, n_seg, seg_len, q, n_q, quant, norm, ws);
    if (rc) return rc;
    if (seg_len >= kSampleMinLen)
        return quantile_sampled(x, n_seg, seg_len, q, n_q, eps, quant, norm, ws, (cudaStream_t)stream);
    return quantile_radix(x, n_seg, seg_len, q, n_q, eps, quant, norm, ws, (cudaStream_t)stream);
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
    const int rc = quantile_check("be_quantile_f32_noalloc", x, n_seg, seg_len, q, n_q, quant, norm, scratch);
    if (rc) return rc;
    // the radix select zeroes the counters and histograms it needs at the start of every call, so whatever the
    // previous call left in the shared array is harmless (the sampled path of be_quantile_f32 would need candidate
    // lists proportional to the input and is not used here)
    return quantile_radix(x, n_seg, seg_len, q, n_q, eps, quant, norm, scratch, (cudaStream_t)stream);
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

extern "C" int be_scale_linear_seg_f32(const float* x, float* y, int64_t n_seg, int64_t seg_len, const double* offset_gain,
                                       be_stream_t stream) {
    BE_REQUIRE(x && y && offset_gain, BE_ERR_INVALID_ARG, "be_scale_linear_seg_f32: null pointer");
    BE_REQUIRE(n_seg > 0 && seg_len > 0, BE_ERR_INVALID_ARG, "be_scale_linear_seg_f32: empty input");
    BE_REQUIRE(n_seg <= 65535, BE_ERR_TOO_LARGE, "be_scale_linear_seg_f32: n_seg %lld > 65535", (long long)n_seg);
    int gx = stream_grid(seg_len);
    if (n_seg > 1) gx = (int)min((int64_t)gx, max((int64_t)1, (int64_t)BE_SM_COUNT * 16 / n_seg));
    scale_linear_seg_kernel<<<dim3(gx, (unsigned)n_seg), kThreads, 0, (cudaStream_t)stream>>>(x, y, seg_len, offset_gain);
    BE_LAUNCH_CHECK();
    return BE_OK;
}

extern "C" int be_cast_f32(const void* x, int dtype, float* y, int64_t n, be_stream_t stream) {
    BE_REQUIRE(x && y, BE_ERR_INVALID_ARG, "be_cast_f32: null pointer");
    BE_REQUIRE(n > 0, BE_ERR_INVALID_ARG, "be_cast_f32: empty input");
    cudaStream_t s = (cudaStream_t)stream;
    const int g = stream_grid((n + 7) / 8);
    switch (dtype) {
        case BE_DTYPE_U8: cast_f32_kernel<uint8_t><<<g, kThreads, 0, s>>>((const uint8_t*)x, y, n); break;
        case BE_DTYPE_I8: cast_f32_kernel<int8_t><<<g, kThreads, 0, s>>>((const int8_t*)x, y, n); break;
        case BE_DTYPE_U16: cast_f32_kernel<uint16_t><<<g, kThreads, 0, s>>>((const uint16_t*)x, y, n); break;
        case BE_DTYPE_I16: cast_f32_kernel<int16_t><<<g, kThreads, 0, s>>>((const int16_t*)x, y, n); break;
        case BE_DTYPE_I32: cast_f32_kernel<int32_t><<<g, kThreads, 0, s>>>((const int32_t*)x, y, n); break;
        case BE_DTYPE_U32: cast_f32_kernel<uint32_t><<<g, kThreads, 0, s>>>((const uint32_t*)x, y, n); break;
        case BE_DTYPE_F16: cast_f32_kernel<__half><<<g, kThreads, 0, s>>>((const __half*)x, y, n); break;
        case BE_DTYPE_F64: cast_f32_kernel<double><<<g, kThreads, 0, s>>>((const double*)x, y, n); break;
        default: BE_REQUIRE(false, BE_ERR_INVALID_ARG, "be_cast_f32: unknown dtype code %d", dtype);
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

// Shared helpers for the be_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "be_b200.h"

#define BE_SM_COUNT 148  // B200: 2 dies x 74 SMs; grids for streaming kernels are sized in multiples of this

void be_set_error(const char* fmt, ...);

#define BE_REQUIRE(cond, code, ...) \
    do {                            \
        if (!(cond)) {              \
            be_set_error(__VA_ARGS__); \
            return (code);          \
        }                           \
    } while (0)

#define BE_CUDA(call)                                                                        \
    do {                                                                                     \
        cudaError_t e_ = (call);                                                             \
        if (e_ != cudaSuccess) {                                                             \
            be_set_error("%s failed at %s:%d: %s", #call, __FILE__, __LINE__, cudaGetErrorString(e_)); \
            return BE_ERR_CUDA;                                                              \
        }                                                                                    \
    } while (0)

#define BE_LAUNCH_CHECK()                                                                     \
    do {                                                                                      \
        cudaError_t e_ = cudaGetLastError();                                                  \
        if (e_ != cudaSuccess) {                                                              \
            be_set_error("kernel launch failed at %s:%d: %s", __FILE__, __LINE__, cudaGetErrorString(e_)); \
            return BE_ERR_CUDA;                                                               \
        }                                                                                     \
    } while (0)

static inline size_t be_align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// Programmatic dependent launch.  The kernels of the hot path are launched with programmatic stream serialization: the
// next kernel's blocks are placed as the previous kernel drains, so launch latency and block start-up stay off the critical
// path (nine launches per image).  be_pdl_wait() returns once the previous kernel in the stream has completed and its
// writes are visible; a kernel calls it before it touches anything another kernel may have written.  Launched without the
// attribute (or after a memset) it returns at once.
__device__ __forceinline__ void be_pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void be_pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// BE_B200_NO_PDL=1 (read once) launches everything with plain stream order: a debugging aid
static inline bool be_pdl_enabled() {
    static int on = -1;
    if (on < 0) {
        const char* e = getenv("BE_B200_NO_PDL");
        on = (e && e[0] == '1') ? 0 : 1;
    }
    return on != 0;
}

template <typename Kernel, typename... Args>
static inline cudaError_t be_launch_pdl(Kernel kernel, dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args... args) {
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cfg.attrs = attr;
    cfg.numAttrs = be_pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, args...);
}

// Streaming loads/stores: data touched once goes around L1 so it does not evict the small tables.
__device__ __forceinline__ float4 ld_stream_f4(const float4* p) {
    float4 r;  // not volatile: the compiler may (and should) hoist several of these ahead of their uses
    asm("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ int4 ld_stream_i4(const int4* p) {
    int4 r;
    asm("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ void st_stream_f4(float4* p, float4 v) {
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
                 "f"(v.w));
}

// numpy's float32 `exp` ufunc, operation for operation (AVX2+FMA3 / AVX512F loops of numpy >= 1.17, the only ones
// x86 hosts of the reference dispatch to): Cody-Waite reduction by round(x*log2e), a degree-5 / degree-2 rational
// in fused multiply-adds, one division, scale by 2^k.  It is faithfully — not correctly — rounded (about 39 % of
// inputs differ from the correctly rounded result by 1 ulp), so a sigmoid → binarize mask can only match the
// reference's bit for bit if the same roundings are made.  Verified against np.exp on 10^7 inputs
// (tests/test_numpy_exp_emulation.py restates this function in numpy).
__device__ __forceinline__ float be_numpy_expf(float x) {
    if (x != x) return x;
    if (x >= 88.72283935546875f) return __int_as_float(0x7f800000);
    if (x <= -103.97208404541015625f) return 0.0f;
    const float magic = 12582912.0f;  // 0x1.8p23: add/subtract rounds to nearest integer
    float k = __fmul_rn(x, 1.442695040888963407359924681001892137f);
    k = __fsub_rn(__fadd_rn(k, magic), magic);
    float r = __fmaf_rn(k, -6.93145752e-1f, x);
    r = __fmaf_rn(k, -1.42860677e-6f, r);
    float num = __fmaf_rn(5.082762527590693718096e-04f, r, 6.757896990527504603057e-03f);
    num = __fmaf_rn(num, r, 5.114512081637298353406e-02f);
    num = __fmaf_rn(num, r, 2.473615434895520810817e-01f);
    num = __fmaf_rn(num, r, 7.257664613233124478488e-01f);
    num = __fmaf_rn(num, r, 9.999999999980870924916e-01f);
    float den = __fmaf_rn(2.159509375685829852307e-02f, r, -2.742335390411667452936e-01f);
    den = __fmaf_rn(den, r, 1.0f);
    return ldexpf(__fdiv_rn(num, den), (int)k);
}

// bioimageio `sigmoid` on an fp32 tensor: 1 / (1 + np.exp(-x)), every step rounded to fp32 as numpy does.
__device__ __forceinline__ float be_sigmoid_f32(float x) {
    return __fdiv_rn(1.0f, __fadd_rn(1.0f, be_numpy_expf(-x)));
}

// Post-processing fused into the stitchers.  `fast` (mask-only sigmoid): the decision sigmoid(v) > thr is v > logit(thr)
// for every v farther than a safety margin from logit(thr) — [lo, hi] is that margin band, computed by the host; only
// values inside the band (a vanishing fraction) take the exact numpy-fp32 sigmoid.
struct Epilogue {
    int mode;   // BE_EPI_*
    int fast;   // mask only: 1 = sigmoid decided by the band test around logit(thr), 2 = no sigmoid, band around thr itself
    float thr, lo, hi;
    double lo2, hi2;  // the same band, a little wider, for the division-free test acc <> wacc * bound (row-walking blend)
};

// Host side of the band test: sigmoid'(x) = s(1-s) >= 0.0099 for thr in [0.01, 0.99], so 1e-4 away from logit(thr) the
// true sigmoid is >= 9.9e-7 away from thr, while the fp32 evaluation is good to a few ulp (<= 2.4e-7 near 1).
static inline Epilogue make_epilogue(int mode, float thr, bool mask_only) {
    Epilogue e{mode, 0, thr, 0.f, 0.f, 0.0, 0.0};
    if (mode == BE_EPI_SIGMOID && mask_only && thr >= 0.01f && thr <= 0.99f) {
        const double L = log((double)thr / (1.0 - (double)thr));
        e.fast = 1;
        e.lo = (float)(L - 1e-4);
        e.hi = (float)(L + 1e-4);
        // acc > wacc * hi2 (wacc > 0) puts the exact quotient above hi * (1 + 2^-21) >= the next float after hi, and
        // rounding is monotone, so float(acc / wacc) > hi; the absolute 2^-100 covers hi == 0.  Same for lo2.
        e.hi2 = (double)e.hi + fabs((double)e.hi) * 0x1p-20 + 0x1p-100;
        e.lo2 = (double)e.lo - fabs((double)e.lo) * 0x1p-20 - 0x1p-100;
    } else if (mode == BE_EPI_NONE && mask_only && thr == thr && fabs((double)thr) < 1e30) {
        // float(q) > thr: a quotient above thr * (1 + 2^-22) rounds to at least the next float after thr, one below
        // thr * (1 - 2^-22) to at most thr (rounding is monotone); the absolute term covers thr == 0 and subnormals
        e.fast = 2;
        e.lo = e.hi = thr;
        e.hi2 = (double)thr + fabs((double)thr) * 0x1p-22 + 0x1p-140;
        e.lo2 = (double)thr - fabs((double)thr) * 0x1p-22 - 0x1p-140;
    }
    return e;
}

// value > thr after the optional sigmoid: the mask the stitchers and the labelling kernel derive from a logit
__device__ __forceinline__ bool be_epilogue_mask(float v, const Epilogue& e) {
    if (e.fast == 1) return (v > e.hi) || (!(v < e.lo) && be_sigmoid_f32(v) > e.thr);
    if (e.mode == BE_EPI_SIGMOID) v = be_sigmoid_f32(v);
    return v > e.thr;
}

// (float)(n / d) with the fp64 quotient correctly rounded first — what numpy does when an fp64 expression is cast to
// fp32 — without paying for an IEEE fp64 division per element: r ~ 1/d to a few ulp (be_fast_rcp), then be_div_to_f32.
__device__ __forceinline__ double be_fast_rcp(double d) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
    r = fma(fma(-d, r, 1.0), r, r);
    r = fma(fma(-d, r, 1.0), r, r);
    return r;
}
// kept out of line on purpose: inlined, the compiler computes the full division speculatively next to the fast path
static __device__ __noinline__ float be_div_to_f32_exact(double n, double d) { return (float)__ddiv_rn(n, d); }

// exact floor(p / S) for 0 <= p < 2^24, 0 < S < 2^24, with rS = 1.0f / S: float estimate + one-step correction
__device__ __forceinline__ int be_fast_div(int p, int S, float rS) {
    int q = __float2int_rz(__int2float_rz(p) * rS);
    q -= (q * S > p);
    q += ((q + 1) * S <= p);
    return q;
}

__device__ __forceinline__ float be_div_to_f32(double n, double d, double r) {
    // q = n * r is within 2^-50 of the exact quotient Q, i.e. within 8 units of q's last mantissa bit.  fp32 keeps 24 of
    // the 53 mantissa bits: fp32(q) can differ from fp32(fp64(Q)) only if the 29 dropped bits of q sit next to the
    // rounding boundary 0x10000000.  Anything within 512 units of it (~4e-6 of inputs) takes the real division, and so
    // does every result outside the fp32 normal range except an exact zero (q == 0 means Q rounds to the same zero),
    // which covers subnormals, overflow, infinities and NaN.  One conversion and integer tests, where converting
    // q(1 + 2^-44) and q(1 - 2^-44) and comparing would cost two conversions and two fp64 FMAs.
    const double q = n * r;
    const float f = (float)q;
    const unsigned dist = ((unsigned)__double2loint(q) & 0x1fffffffu) - 0x0ffffe00u;
    const unsigned e = __float_as_uint(f) & 0x7f800000u;
    if (dist <= 0x400u || e == 0x7f800000u || (e == 0u && q != 0.0)) return be_div_to_f32_exact(n, d);
    return f;
}

// ---- TMA bulk copies + mbarriers: one warp fetches contiguous runs of global memory into a ring of shared-memory
//      stages; `full` barriers count the bytes in, `empty` barriers the readers out ----
static __device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
static __device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
static __device__ __forceinline__ void mbar_arrive(unsigned bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
static __device__ __forceinline__ void bulk_g2s(unsigned dst, const void* src, unsigned bytes, unsigned bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
// one box of a 2-D tensor map (inner coordinate x, outer coordinate y) into shared memory: a single instruction for a
// whole stage of rows, where the per-row bulk copies above are bound by the copy engine's instruction rate
static __device__ __forceinline__ void tensor_g2s_2d(unsigned dst, const void* tmap, int x, int y, unsigned bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"(tmap), "r"(x), "r"(y), "r"(bar)
                 : "memory");
}
// Waits with a suspend-time hint: the warp sleeps in hardware until the phase completes (or the hint runs out), so a copy
// warp that is a whole ring ahead does not burn issue slots polling — with plain try_wait the polling loop was a fifth
// of all instructions the blend kernel issued.
static __device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
    unsigned ok = 0;
    for (int spin = 0; !ok; ++spin) {
        asm volatile(
            "{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3; selp.u32 %0, 1, 0, p; }"
            : "=r"(ok)
            : "r"(bar), "r"(parity), "r"(100000u)
            : "memory");
        if (spin > (1 << 18)) __trap();  // a byte count that never arrives is a bug, not something to wait out
    }
}

static __device__ __forceinline__ bool mbar_try_wait(unsigned bar, unsigned parity) {
    unsigned ok;
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                 : "=r"(ok)
                 : "r"(bar), "r"(parity)
                 : "memory");
    return ok != 0;
}
static __device__ __forceinline__ float4 lds128(unsigned addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

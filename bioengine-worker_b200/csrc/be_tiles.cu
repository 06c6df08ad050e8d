// Halo-tile gather and the two stitchers.  All three are pure data movement (HBM-bound): one thread makes
// four consecutive x outputs and stores them as one 16-byte (fp32) / 4-byte (uint8) transaction.
//
//   be_tile_gather_f32   image (+ fused normalise) -> (N*T, C, Bz, By, Bx) tiles, table-driven padding
//   be_stitch_crop_f32   bioimageio inner-region reassembly, table-driven, optional sigmoid / mask epilogue
//   be_stitch_blend_f32  fibsem Gaussian-weighted overlap blend as a gather, fp64 accumulation
#include "be_common.cuh"

namespace {

constexpr int kThreads = 256;

struct Dims {
    int64_t N, C, Z, Y, X;
    int Tz, Ty, Tx, Bz, By, Bx;
};

template <bool CLIP>
__device__ __forceinline__ float norm_apply(float v, double sub, double div, double rdiv, float lo, float hi) {
    // np.clip on the fp64 quotient then a cast equals a clip of the cast value: the bounds are fp32 numbers
    float r = be_div_to_f32(__dsub_rn((double)v, sub), div, rdiv);
    if (CLIP && r == r) r = fminf(fmaxf(r, lo), hi);
    return r;
}

// ---------------------------------------------------------------------------------------------------
// gather
// ---------------------------------------------------------------------------------------------------

// grid: x = tile (n, kz, ky, kx), y = (c, iz), z = a band of tile rows.  A thread owns one 4-px column unit and walks
// the rows of its band, so the tile / plane decomposition (integer divides) is paid once per thread, not once per
// 16 bytes, and several independent row loads are in flight per thread.
template <bool NORM, bool CLIP>
__global__ void __launch_bounds__(kThreads) tile_gather_kernel(const float* __restrict__ img, float* __restrict__ tiles,
                                                               Dims d, const int32_t* __restrict__ zsrc,
                                                               const int32_t* __restrict__ ysrc,
                                                               const int32_t* __restrict__ xsrc,
                                                               const double* __restrict__ norm, int norm_mode,
                                                               float clip_lo, float clip_hi, int units_per_row,
                                                               int rows_per_band) {
    const int ucols = min(units_per_row, kThreads);     // column units the block covers per sweep
    const int u_in = threadIdx.x % ucols;
    const int r_in = threadIdx.x / ucols;               // first row offset inside the band
    const int r_step = kThreads / ucols;                // rows the block covers per sweep
    if (r_in >= r_step) return;
    unsigned t = blockIdx.x;
    const int kx = t % d.Tx;
    t /= d.Tx;
    const int ky = t % d.Ty;
    t /= d.Ty;
    const int kz = t % d.Tz;
    const int64_t n = t / d.Tz;
    const int c = blockIdx.y / d.Bz, iz = blockIdx.y - c * d.Bz;
    const int iy0 = blockIdx.z * rows_per_band, iy1 = min(iy0 + rows_per_band, d.By);

    const int64_t sz = zsrc[kz * d.Bz + iz];
    const float* plane = img + ((n * d.C + c) * d.Z + sz) * d.Y * d.X;
    const int32_t* ys = ysrc + ky * d.By;
    float* dst_tile = tiles + (((int64_t)blockIdx.x * d.C + c) * d.Bz + iz) * (int64_t)d.By * d.Bx;

    double sub = 0.0, div = 1.0, rdiv = 1.0;
    if (NORM) {
        const int64_t s = (norm_mode == 1) ? (n * d.C + c) : n;
        sub = norm[2 * s];
        div = norm[2 * s + 1];
        rdiv = be_fast_rcp(div);
    }

    for (int u = u_in; u < units_per_row; u += ucols) {
    float* dst_plane = dst_tile + u * 4;
    const int rem = d.Bx - u * 4;
    const int32_t* xs = xsrc + kx * d.Bx + u * 4;
    if (rem >= 4 && (d.Bx & 3) == 0) {
        const int4 xi = *reinterpret_cast<const int4*>(xs);
        const bool contiguous = xi.y == xi.x + 1 && xi.z == xi.x + 2 && xi.w == xi.x + 3 && ((d.X & 3) == 0) &&
                                ((xi.x & 3) == 0) && ((((uintptr_t)img) & 15) == 0);
#pragma unroll 4
        for (int iy = iy0 + r_in; iy < iy1; iy += r_step) {
            const float* src = plane + (int64_t)ys[iy] * d.X;
            float4 v;
            if (contiguous) {
                v = __ldg(reinterpret_cast<const float4*>(src + xi.x));
            } else {
                v.x = __ldg(src + xi.x);
                v.y = __ldg(src + xi.y);
                v.z = __ldg(src + xi.z);
                v.w = __ldg(src + xi.w);
            }
            if (NORM) {
                v.x = norm_apply<CLIP>(v.x, sub, div, rdiv, clip_lo, clip_hi);
                v.y = norm_apply<CLIP>(v.y, sub, div, rdiv, clip_lo, clip_hi);
                v.z = norm_apply<CLIP>(v.z, sub, div, rdiv, clip_lo, clip_hi);
                v.w = norm_apply<CLIP>(v.w, sub, div, rdiv, clip_lo, clip_hi);
            }
            st_stream_f4(reinterpret_cast<float4*>(dst_plane + (int64_t)iy * d.Bx), v);
        }
    } else {
        for (int iy = iy0 + r_in; iy < iy1; iy += r_step) {
            const float* src = plane + (int64_t)ys[iy] * d.X;
            for (int j = 0; j < rem && j < 4; ++j) {
                float v = __ldg(src + xs[j]);
                if (NORM) v = norm_apply<CLIP>(v, sub, div, rdiv, clip_lo, clip_hi);
                dst_plane[(int64_t)iy * d.Bx + j] = v;
            }
        }
    }
    }
}

// ---------------------------------------------------------------------------------------------------
// stitch: shared epilogue
// ---------------------------------------------------------------------------------------------------

// Post-processing fused into the stitchers.  `fast` (mask-only sigmoid): the decision sigmoid(v) > thr is v > logit(thr)
// for every v farther than a safety margin from logit(thr) — [lo, hi] is that margin band, computed by the host; only
// values inside the band (a vanishing fraction) take the exact numpy-fp32 sigmoid.
struct Epilogue {
    int mode;   // BE_EPI_*
    int fast;   // 1: mask only, sigmoid decided by the band test
    float thr, lo, hi;
};

__device__ __forceinline__ void finish4(float* __restrict__ out, uint8_t* __restrict__ mask, int64_t o, int cnt,
                                        float v[4], const Epilogue& e, bool aligned) {
    bool m[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        if (e.fast) {
            m[j] = (v[j] > e.hi) || (!(v[j] < e.lo) && be_sigmoid_f32(v[j]) > e.thr);
        } else {
            if (e.mode == BE_EPI_SIGMOID) v[j] = be_sigmoid_f32(v[j]);
            m[j] = v[j] > e.thr;
        }
    }
    if (cnt == 4 && aligned) {
        if (out) st_stream_f4(reinterpret_cast<float4*>(out + o), make_float4(v[0], v[1], v[2], v[3]));
        if (mask) *reinterpret_cast<uchar4*>(mask + o) = make_uchar4(m[0], m[1], m[2], m[3]);
    } else {
        for (int j = 0; j < cnt; ++j) {
            if (out) out[o + j] = v[j];
            if (mask) mask[o + j] = m[j];
        }
    }
}

// grid: x = 4-px units along the row, y = output row y, z = (n, c, z) planes (grid-strided)
struct RowIndex {
    int u, y, z, c;
    int64_t n, row;
};

__device__ __forceinline__ bool row_index(const Dims& d, int units_per_row, unsigned plane, RowIndex* r) {
    r->u = blockIdx.x * kThreads + threadIdx.x;
    if (r->u >= units_per_row) return false;
    r->y = blockIdx.y;
    r->z = plane % (unsigned)d.Z;
    const unsigned nc = plane / (unsigned)d.Z;
    r->c = nc % (unsigned)d.C;
    r->n = nc / (unsigned)d.C;
    r->row = (int64_t)plane * d.Y + r->y;
    return true;
}

// ---------------------------------------------------------------------------------------------------
// crop stitch
// ---------------------------------------------------------------------------------------------------

__global__ void __launch_bounds__(kThreads) stitch_crop_kernel(const float* __restrict__ tiles, float* __restrict__ out,
                                                               uint8_t* __restrict__ mask, Dims d,
                                                               const int32_t* __restrict__ zsel,
                                                               const int32_t* __restrict__ ysel,
                                                               const int32_t* __restrict__ xsel, Epilogue e,
                                                               int units_per_row, unsigned n_planes) {
    for (unsigned plane = blockIdx.z; plane < n_planes; plane += gridDim.z) {
        RowIndex r;
        if (!row_index(d, units_per_row, plane, &r)) return;
        const int vz = zsel[r.z], vy = ysel[r.y];
        const int kz = vz / d.Bz, iz = vz - kz * d.Bz;
        const int ky = vy / d.By, iy = vy - ky * d.By;
        const int64_t tvol = (int64_t)d.Bz * d.By * d.Bx;
        const int64_t trow = ((r.n * d.Tz + kz) * d.Ty + ky) * d.Tx;  // + kx
        const int64_t in_tile = ((int64_t)r.c * d.Bz + iz) * d.By * d.Bx + (int64_t)iy * d.Bx;

        const int x0 = r.u * 4;
        const int cnt = min(4, (int)d.X - x0);
        float v[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (j < cnt) {
                const int vx = xsel[x0 + j];
                const int kx = vx / d.Bx, ix = vx - kx * d.Bx;
                v[j] = __ldg(tiles + (trow + kx) * d.C * tvol + in_tile + ix);
            }
        }
        finish4(out, mask, r.row * d.X + x0, cnt, v, e, (d.X & 3) == 0);
    }
}

// ---------------------------------------------------------------------------------------------------
// blend stitch
// ---------------------------------------------------------------------------------------------------

struct Strides {
    int Sz, Sy, Sx;
    float rSz, rSy, rSx;   // 1.0f / stride, for be_fast_div
    int64_t tile_stride;   // elements between consecutive tiles = C * Bz * By * Bx
};

// tiles k covering output coordinate p: k*S <= p < k*S + B, 0 <= k < T
__device__ __forceinline__ void cover(int p, int B, int S, float rS, int T, int* lo, int* hi) {
    *hi = min(be_fast_div(p, S, rS), T - 1);
    const int t = p - B + 1;
    *lo = (t <= 0) ? 0 : be_fast_div(t + S - 1, S, rS);
}

__global__ void __launch_bounds__(kThreads) stitch_blend_kernel(const float* __restrict__ tiles, float* __restrict__ out,
                                                                uint8_t* __restrict__ mask, Dims d, Strides s,
                                                                const double* __restrict__ wz,
                                                                const double* __restrict__ wy,
                                                                const double* __restrict__ wx, Epilogue e,
                                                                int units_per_row, unsigned n_planes) {
    const int u = min((int)(blockIdx.x * kThreads + threadIdx.x), units_per_row - 1);
    const bool live = blockIdx.x * kThreads + threadIdx.x < (unsigned)units_per_row;  // spare threads shadow the last unit
    const int y = blockIdx.y;
    const int64_t tile_stride = s.tile_stride;
    const int x0 = u * 4;
    const int cnt = min(4, (int)d.X - x0);
    // tiles touching this 4-px unit: from the first tile covering x0 to the last one covering its last pixel (one
    // division pair; a unit is 4 px wide, so the last pixel can at most reach the next tile)
    int kx_first, kx_last;
    cover(x0, d.Bx, s.Sx, s.rSx, d.Tx, &kx_first, &kx_last);
    kx_last += (kx_last + 1 < d.Tx && (kx_last + 1) * s.Sx <= x0 + cnt - 1) ? 1 : 0;
    // the row's tile range is the same for the whole block: one thread works it out
    __shared__ int s_ky[2];
    if (threadIdx.x == 0) cover(y, d.By, s.Sy, s.rSy, d.Ty, &s_ky[0], &s_ky[1]);
    __syncthreads();
    const int ky_lo = s_ky[0], ky_hi = s_ky[1];

    for (unsigned plane = blockIdx.z; plane < n_planes; plane += gridDim.z) {
        int z = 0, c = 0;
        int64_t n = 0;
        if (n_planes > 1) {  // single 2-D image: nothing to decompose
            z = plane % (unsigned)d.Z;
            const unsigned nc = plane / (unsigned)d.Z;
            c = nc % (unsigned)d.C;
            n = nc / (unsigned)d.C;
        }
        int kz_lo, kz_hi;
        cover(z, d.Bz, s.Sz, s.rSz, d.Tz, &kz_lo, &kz_hi);

        double acc[4] = {0.0, 0.0, 0.0, 0.0}, wacc[4] = {0.0, 0.0, 0.0, 0.0};
        // ascending (kz, ky, kx): the order a sequential walk over the reference's tile_coords list adds them
        for (int kz = kz_lo; kz <= kz_hi; ++kz) {
            const int iz = z - kz * s.Sz;
            const double wzv = wz[iz];
            for (int ky = ky_lo; ky <= ky_hi; ++ky) {
                const int iy = y - ky * s.Sy;
                const double wzy = __dmul_rn(wzv, wy[iy]);
                const float* rowbase = tiles + (((n * d.Tz + kz) * d.Ty + ky) * d.Tx) * tile_stride +
                                       (((int64_t)c * d.Bz + iz) * d.By + iy) * d.Bx;
                const float* src = rowbase + kx_first * tile_stride + (x0 - kx_first * s.Sx);
                for (int kx = kx_first; kx <= kx_last; ++kx, src += tile_stride - s.Sx) {
                    const int ix0 = x0 - kx * s.Sx;  // in-tile offset of the unit's first pixel
                    if (cnt == 4 && ix0 >= 0 && ix0 + 4 <= d.Bx && ((((uintptr_t)src) & 15) == 0) && ((ix0 & 1) == 0)) {
                        // the whole unit lies inside this tile: one 16-byte tile load, two 16-byte window loads
                        const float4 p = __ldg(reinterpret_cast<const float4*>(src));
                        const double2 wa = __ldg(reinterpret_cast<const double2*>(wx + ix0));
                        const double2 wb = __ldg(reinterpret_cast<const double2*>(wx + ix0 + 2));
                        const float w0 = (float)__dmul_rn(wzy, wa.x), w1 = (float)__dmul_rn(wzy, wa.y);
                        const float w2 = (float)__dmul_rn(wzy, wb.x), w3 = (float)__dmul_rn(wzy, wb.y);
                        acc[0] = __dadd_rn(acc[0], (double)__fmul_rn(p.x, w0));
                        acc[1] = __dadd_rn(acc[1], (double)__fmul_rn(p.y, w1));
                        acc[2] = __dadd_rn(acc[2], (double)__fmul_rn(p.z, w2));
                        acc[3] = __dadd_rn(acc[3], (double)__fmul_rn(p.w, w3));
                        wacc[0] = __dadd_rn(wacc[0], (double)w0);
                        wacc[1] = __dadd_rn(wacc[1], (double)w1);
                        wacc[2] = __dadd_rn(wacc[2], (double)w2);
                        wacc[3] = __dadd_rn(wacc[3], (double)w3);
                    } else {
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const int ix = ix0 + j;
                            if (j < cnt && ix >= 0 && ix < d.Bx) {
                                const float w = (float)__dmul_rn(wzy, wx[ix]);
                                acc[j] = __dadd_rn(acc[j], (double)__fmul_rn(__ldg(src + j), w));
                                wacc[j] = __dadd_rn(wacc[j], (double)w);
                            }
                        }
                    }
                }
            }
        }
        float v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] = be_div_to_f32(acc[j], wacc[j], be_fast_rcp(wacc[j]));
        if (live) finish4(out, mask, ((int64_t)plane * d.Y + y) * d.X + x0, cnt, v, e, (d.X & 3) == 0);
    }
}

// Host side of the band test: sigmoid'(x) = s(1-s) >= 0.0099 for thr in [0.01, 0.99], so 1e-4 away from logit(thr) the
// true sigmoid is >= 9.9e-7 away from thr, while the fp32 evaluation is good to a few ulp (<= 2.4e-7 near 1).
Epilogue make_epilogue(int mode, float thr, bool mask_only) {
    Epilogue e{mode, 0, thr, 0.f, 0.f};
    if (mode == BE_EPI_SIGMOID && mask_only && thr >= 0.01f && thr <= 0.99f) {
        const double L = log((double)thr / (1.0 - (double)thr));
        e.fast = 1;
        e.lo = (float)(L - 1e-4);
        e.hi = (float)(L + 1e-4);
    }
    return e;
}

int check_dims(const char* who, int64_t N, int64_t C, int64_t Z, int64_t Y, int64_t X, int Tz, int Ty, int Tx, int Bz,
               int By, int Bx) {
    BE_REQUIRE(N > 0 && C > 0 && Z > 0 && Y > 0 && X > 0, BE_ERR_INVALID_ARG, "%s: empty image (%lld,%lld,%lld,%lld,%lld)",
               who, (long long)N, (long long)C, (long long)Z, (long long)Y, (long long)X);
    BE_REQUIRE(Tz > 0 && Ty > 0 && Tx > 0 && Bz > 0 && By > 0 && Bx > 0, BE_ERR_INVALID_ARG,
               "%s: tile grid / block shape must be positive", who);
    BE_REQUIRE(Z < (1 << 30) && Y < (1 << 30) && X < (1 << 30), BE_ERR_TOO_LARGE, "%s: axis longer than 2^30", who);
    return BE_OK;
}

int stitch_grid(const char* who, const Dims& d, int upr, dim3* grid) {
    BE_REQUIRE(d.Y <= 65535, BE_ERR_TOO_LARGE, "%s: Y %lld > 65535", who, (long long)d.Y);
    const int64_t planes = d.N * d.C * d.Z;
    BE_REQUIRE(planes < ((int64_t)1 << 31), BE_ERR_TOO_LARGE, "%s: too many planes", who);
    *grid = dim3((unsigned)((upr + kThreads - 1) / kThreads), (unsigned)d.Y, (unsigned)(planes > 65535 ? 65535 : planes));
    return BE_OK;
}

}  // namespace

extern "C" int be_tile_gather_f32(const float* img, float* tiles, int64_t N, int64_t C, int64_t Z, int64_t Y, int64_t X,
                                  int Tz, int Ty, int Tx, int Bz, int By, int Bx, const int32_t* zsrc,
                                  const int32_t* ysrc, const int32_t* xsrc, const double* norm, int norm_mode, int clip,
                                  float clip_lo, float clip_hi, be_stream_t stream) {
    BE_REQUIRE(img && tiles && zsrc && ysrc && xsrc, BE_ERR_INVALID_ARG, "be_tile_gather_f32: null pointer");
    int rc = check_dims("be_tile_gather_f32", N, C, Z, Y, X, Tz, Ty, Tx, Bz, By, Bx);
    if (rc) return rc;
    BE_REQUIRE(norm_mode >= 0 && norm_mode <= 2, BE_ERR_INVALID_ARG, "be_tile_gather_f32: norm_mode %d", norm_mode);
    BE_REQUIRE(norm_mode == 0 || norm, BE_ERR_INVALID_ARG, "be_tile_gather_f32: norm_mode %d needs norm", norm_mode);
    const Dims d{N, C, Z, Y, X, Tz, Ty, Tx, Bz, By, Bx};
    const int upr = (Bx + 3) / 4;
    const int64_t n_tiles = N * Tz * Ty * Tx;
    BE_REQUIRE(n_tiles < ((int64_t)1 << 31) - 1, BE_ERR_TOO_LARGE, "be_tile_gather_f32: too many tiles");
    BE_REQUIRE(C * Bz <= 65535, BE_ERR_TOO_LARGE, "be_tile_gather_f32: C*Bz %lld > 65535", (long long)(C * Bz));
    // bands of tile rows: enough blocks for ~8 CTAs per SM, at least 4 sweeps of work per thread where possible
    const int r_step = kThreads / (upr < kThreads ? upr : kThreads);
    const int64_t planes = n_tiles * C * Bz;
    int bands = (int)((int64_t)BE_SM_COUNT * 8 / (planes < 1 ? 1 : planes)) + 1;
    const int max_bands = (By + 4 * r_step - 1) / (4 * r_step);
    if (bands > max_bands) bands = max_bands;
    if (bands < 1) bands = 1;
    const int rows_per_band = (By + bands - 1) / bands;
    bands = (By + rows_per_band - 1) / rows_per_band;
    const dim3 grid((unsigned)n_tiles, (unsigned)(C * Bz), (unsigned)bands);
    cudaStream_t s = (cudaStream_t)stream;
    if (norm_mode && clip)
        tile_gather_kernel<true, true><<<grid, kThreads, 0, s>>>(img, tiles, d, zsrc, ysrc, xsrc, norm, norm_mode, clip_lo,
                                                                 clip_hi, upr, rows_per_band);
    else if (norm_mode)
        tile_gather_kernel<true, false><<<grid, kThreads, 0, s>>>(img, tiles, d, zsrc, ysrc, xsrc, norm, norm_mode, 0.f, 0.f,
                                                                  upr, rows_per_band);
    else
        tile_gather_kernel<false, false><<<grid, kThreads, 0, s>>>(img, tiles, d, zsrc, ysrc, xsrc, norm, 0, 0.f, 0.f, upr,
                                                                   rows_per_band);
    BE_LAUNCH_CHECK();
    return BE_OK;
}

extern "C" int be_stitch_crop_f32(const float* tiles, float* out, uint8_t* mask, int64_t N, int64_t C, int64_t Z,
                                  int64_t Y, int64_t X, int Tz, int Ty, int Tx, int Bz, int By, int Bx,
                                  const int32_t* zsel, const int32_t* ysel, const int32_t* xsel, int epilogue, float thr,
                                  be_stream_t stream) {
    BE_REQUIRE(tiles && (out || mask) && zsel && ysel && xsel, BE_ERR_INVALID_ARG, "be_stitch_crop_f32: null pointer");
    int rc = check_dims("be_stitch_crop_f32", N, C, Z, Y, X, Tz, Ty, Tx, Bz, By, Bx);
    if (rc) return rc;
    BE_REQUIRE(epilogue == BE_EPI_NONE || epilogue == BE_EPI_SIGMOID, BE_ERR_INVALID_ARG,
               "be_stitch_crop_f32: epilogue %d", epilogue);
    const Dims d{N, C, Z, Y, X, Tz, Ty, Tx, Bz, By, Bx};
    const int upr = (int)((X + 3) / 4);
    dim3 grid;
    rc = stitch_grid("be_stitch_crop_f32", d, upr, &grid);
    if (rc) return rc;
    stitch_crop_kernel<<<grid, kThreads, 0, (cudaStream_t)stream>>>(tiles, out, mask, d, zsel, ysel, xsel,
                                                                    make_epilogue(epilogue, thr, out == nullptr), upr,
                                                                    (unsigned)(N * C * Z));
    BE_LAUNCH_CHECK();
    return BE_OK;
}

extern "C" int be_stitch_blend_f32(const float* tiles, float* out, uint8_t* mask, int64_t N, int64_t C, int64_t Z,
                                   int64_t Y, int64_t X, int Tz, int Ty, int Tx, int Bz, int By, int Bx, int Sz, int Sy,
                                   int Sx, const double* wz, const double* wy, const double* wx, int epilogue, float thr,
                                   be_stream_t stream) {
    BE_REQUIRE(tiles && (out || mask) && wz && wy && wx, BE_ERR_INVALID_ARG, "be_stitch_blend_f32: null pointer");
    int rc = check_dims("be_stitch_blend_f32", N, C, Z, Y, X, Tz, Ty, Tx, Bz, By, Bx);
    if (rc) return rc;
    BE_REQUIRE(epilogue == BE_EPI_NONE || epilogue == BE_EPI_SIGMOID, BE_ERR_INVALID_ARG,
               "be_stitch_blend_f32: epilogue %d", epilogue);
    BE_REQUIRE(Sz > 0 && Sy > 0 && Sx > 0 && Sz <= Bz && Sy <= By && Sx <= Bx, BE_ERR_INVALID_ARG,
               "be_stitch_blend_f32: stride must be in 1..block (got %d,%d,%d)", Sz, Sy, Sx);
    BE_REQUIRE((int64_t)(Tz - 1) * Sz < Z && (int64_t)Tz * Sz >= Z && (int64_t)(Ty - 1) * Sy < Y && (int64_t)Ty * Sy >= Y &&
                   (int64_t)(Tx - 1) * Sx < X && (int64_t)Tx * Sx >= X,
               BE_ERR_INVALID_ARG, "be_stitch_blend_f32: tile grid must be ceil(dim/stride) per axis");
    const Dims d{N, C, Z, Y, X, Tz, Ty, Tx, Bz, By, Bx};
    BE_REQUIRE(Z < (1 << 24) && Y < (1 << 24) && X < (1 << 24), BE_ERR_TOO_LARGE, "be_stitch_blend_f32: axis longer than 2^24");
    const Strides st{Sz, Sy, Sx, 1.0f / (float)Sz, 1.0f / (float)Sy, 1.0f / (float)Sx, (int64_t)C * Bz * By * Bx};
    const int upr = (int)((X + 3) / 4);
    dim3 grid;
    rc = stitch_grid("be_stitch_blend_f32", d, upr, &grid);
    if (rc) return rc;
    stitch_blend_kernel<<<grid, kThreads, 0, (cudaStream_t)stream>>>(tiles, out, mask, d, st, wz, wy, wx,
                                                                     make_epilogue(epilogue, thr, out == nullptr), upr,
                                                                     (unsigned)(N * C * Z));
    BE_LAUNCH_CHECK();
    return BE_OK;
}

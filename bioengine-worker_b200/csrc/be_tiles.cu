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

__device__ __forceinline__ float norm_apply(float v, double sub, double div, int clip, float lo, float hi) {
    double r = __ddiv_rn(__dsub_rn((double)v, sub), div);
    if (clip && r == r) r = fmin(fmax(r, (double)lo), (double)hi);
    return (float)r;
}

// ---------------------------------------------------------------------------------------------------
// gather
// ---------------------------------------------------------------------------------------------------

template <bool NORM>
__global__ void __launch_bounds__(kThreads) tile_gather_kernel(const float* __restrict__ img, float* __restrict__ tiles,
                                                               Dims d, const int32_t* __restrict__ zsrc,
                                                               const int32_t* __restrict__ ysrc,
                                                               const int32_t* __restrict__ xsrc,
                                                               const double* __restrict__ norm, int norm_mode, int clip,
                                                               float clip_lo, float clip_hi, int units_per_row,
                                                               int64_t n_units) {
    const int64_t gid = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    if (gid >= n_units) return;
    const int u = (int)(gid % units_per_row);
    int64_t r = gid / units_per_row;  // row index over (n, kz, ky, kx, c, iz, iy)
    const int iy = (int)(r % d.By);
    r /= d.By;
    const int iz = (int)(r % d.Bz);
    r /= d.Bz;
    const int c = (int)(r % d.C);
    r /= d.C;
    const int kx = (int)(r % d.Tx);
    r /= d.Tx;
    const int ky = (int)(r % d.Ty);
    r /= d.Ty;
    const int kz = (int)(r % d.Tz);
    const int64_t n = r / d.Tz;

    const int64_t sz = zsrc[kz * d.Bz + iz];
    const int64_t sy = ysrc[ky * d.By + iy];
    const float* src = img + (((n * d.C + c) * d.Z + sz) * d.Y + sy) * d.X;
    const int32_t* xs = xsrc + kx * d.Bx + u * 4;
    float* dst = tiles + (gid / units_per_row) * d.Bx + u * 4;

    double sub = 0.0, div = 1.0;
    if (NORM) {
        const int64_t s = (norm_mode == 1) ? (n * d.C + c) : n;
        sub = norm[2 * s];
        div = norm[2 * s + 1];
    }

    const int rem = d.Bx - u * 4;
    if (rem >= 4 && (d.Bx & 3) == 0) {
        const int4 xi = *reinterpret_cast<const int4*>(xs);
        float4 v;
        const float* p0 = src + xi.x;
        if (xi.y == xi.x + 1 && xi.z == xi.x + 2 && xi.w == xi.x + 3 && (((uintptr_t)p0) & 15) == 0) {
            v = __ldg(reinterpret_cast<const float4*>(p0));
        } else {
            v.x = __ldg(src + xi.x);
            v.y = __ldg(src + xi.y);
            v.z = __ldg(src + xi.z);
            v.w = __ldg(src + xi.w);
        }
        if (NORM) {
            v.x = norm_apply(v.x, sub, div, clip, clip_lo, clip_hi);
            v.y = norm_apply(v.y, sub, div, clip, clip_lo, clip_hi);
            v.z = norm_apply(v.z, sub, div, clip, clip_lo, clip_hi);
            v.w = norm_apply(v.w, sub, div, clip, clip_lo, clip_hi);
        }
        st_stream_f4(reinterpret_cast<float4*>(dst), v);
    } else {
        for (int j = 0; j < rem && j < 4; ++j) {
            float v = __ldg(src + xs[j]);
            if (NORM) v = norm_apply(v, sub, div, clip, clip_lo, clip_hi);
            dst[j] = v;
        }
    }
}

// ---------------------------------------------------------------------------------------------------
// stitch: shared epilogue
// ---------------------------------------------------------------------------------------------------

__device__ __forceinline__ void store4(float* __restrict__ out, uint8_t* __restrict__ mask, int64_t o, int cnt,
                                       const float v[4], float thr, bool aligned) {
    if (cnt == 4 && aligned) {
        if (out) st_stream_f4(reinterpret_cast<float4*>(out + o), make_float4(v[0], v[1], v[2], v[3]));
        if (mask) {
            uchar4 m;
            m.x = v[0] > thr;
            m.y = v[1] > thr;
            m.z = v[2] > thr;
            m.w = v[3] > thr;
            *reinterpret_cast<uchar4*>(mask + o) = m;
        }
    } else {
        for (int j = 0; j < cnt; ++j) {
            if (out) out[o + j] = v[j];
            if (mask) mask[o + j] = v[j] > thr;
        }
    }
}

// ---------------------------------------------------------------------------------------------------
// crop stitch
// ---------------------------------------------------------------------------------------------------

__global__ void __launch_bounds__(kThreads) stitch_crop_kernel(const float* __restrict__ tiles, float* __restrict__ out,
                                                               uint8_t* __restrict__ mask, Dims d,
                                                               const int32_t* __restrict__ zsel,
                                                               const int32_t* __restrict__ ysel,
                                                               const int32_t* __restrict__ xsel, int epilogue, float thr,
                                                               int units_per_row, int64_t n_units) {
    const int64_t gid = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    if (gid >= n_units) return;
    const int u = (int)(gid % units_per_row);
    int64_t r = gid / units_per_row;  // (n, c, z, y)
    const int64_t row = r;
    const int y = (int)(r % d.Y);
    r /= d.Y;
    const int z = (int)(r % d.Z);
    r /= d.Z;
    const int c = (int)(r % d.C);
    const int64_t n = r / d.C;

    const int vz = zsel[z], vy = ysel[y];
    const int kz = vz / d.Bz, iz = vz - kz * d.Bz;
    const int ky = vy / d.By, iy = vy - ky * d.By;
    const int64_t tvol = (int64_t)d.Bz * d.By * d.Bx;
    const int64_t trow = ((n * d.Tz + kz) * d.Ty + ky) * d.Tx;  // + kx
    const int64_t in_tile = ((int64_t)c * d.Bz + iz) * d.By * d.Bx + (int64_t)iy * d.Bx;

    const int x0 = u * 4;
    const int cnt = min(4, (int)d.X - x0);
    float v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        if (j < cnt) {
            const int vx = xsel[x0 + j];
            const int kx = vx / d.Bx, ix = vx - kx * d.Bx;
            float t = __ldg(tiles + (trow + kx) * d.C * tvol + in_tile + ix);
            if (epilogue == BE_EPI_SIGMOID) t = be_sigmoid_f32(t);
            v[j] = t;
        }
    }
    store4(out, mask, row * d.X + x0, cnt, v, thr, (d.X & 3) == 0);
}

// ---------------------------------------------------------------------------------------------------
// blend stitch
// ---------------------------------------------------------------------------------------------------

struct Strides {
    int Sz, Sy, Sx;
};

__device__ __forceinline__ void cover(int p, int B, int S, int T, int* lo, int* hi) {
    *hi = min(p / S, T - 1);
    const int t = p - B + 1;
    *lo = (t <= 0) ? 0 : (t + S - 1) / S;
}

__global__ void __launch_bounds__(kThreads) stitch_blend_kernel(const float* __restrict__ tiles, float* __restrict__ out,
                                                                uint8_t* __restrict__ mask, Dims d, Strides s,
                                                                const double* __restrict__ wz,
                                                                const double* __restrict__ wy,
                                                                const double* __restrict__ wx, int epilogue, float thr,
                                                                int units_per_row, int64_t n_units) {
    const int64_t gid = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    if (gid >= n_units) return;
    const int u = (int)(gid % units_per_row);
    int64_t r = gid / units_per_row;
    const int64_t row = r;
    const int y = (int)(r % d.Y);
    r /= d.Y;
    const int z = (int)(r % d.Z);
    r /= d.Z;
    const int c = (int)(r % d.C);
    const int64_t n = r / d.C;

    const int64_t tvol = (int64_t)d.Bz * d.By * d.Bx;
    const int x0 = u * 4;
    const int cnt = min(4, (int)d.X - x0);
    int kz_lo, kz_hi, ky_lo, ky_hi;
    cover(z, d.Bz, s.Sz, d.Tz, &kz_lo, &kz_hi);
    cover(y, d.By, s.Sy, d.Ty, &ky_lo, &ky_hi);
    int kx_lo[4], kx_hi[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) cover(min(x0 + j, (int)d.X - 1), d.Bx, s.Sx, d.Tx, &kx_lo[j], &kx_hi[j]);

    double acc[4] = {0.0, 0.0, 0.0, 0.0}, wacc[4] = {0.0, 0.0, 0.0, 0.0};
    // ascending (kz, ky, kx): the order a sequential walk over the reference's tile_coords list adds them
    for (int kz = kz_lo; kz <= kz_hi; ++kz) {
        const int iz = z - kz * s.Sz;
        const double wzv = wz[iz];
        for (int ky = ky_lo; ky <= ky_hi; ++ky) {
            const int iy = y - ky * s.Sy;
            const double wzy = __dmul_rn(wzv, wy[iy]);
            const int64_t trow = ((n * d.Tz + kz) * d.Ty + ky) * d.Tx;
            const int64_t in_tile = (((int64_t)c * d.Bz + iz) * d.By + iy) * d.Bx;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if (j < cnt) {
                    for (int kx = kx_lo[j]; kx <= kx_hi[j]; ++kx) {
                        const int ix = x0 + j - kx * s.Sx;
                        const float w = (float)__dmul_rn(wzy, wx[ix]);
                        const float p = __ldg(tiles + (trow + kx) * d.C * tvol + in_tile + ix);
                        acc[j] = __dadd_rn(acc[j], (double)__fmul_rn(p, w));
                        wacc[j] = __dadd_rn(wacc[j], (double)w);
                    }
                }
            }
        }
    }
    float v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        float t = (float)__ddiv_rn(acc[j], wacc[j]);
        if (epilogue == BE_EPI_SIGMOID) t = be_sigmoid_f32(t);
        v[j] = t;
    }
    store4(out, mask, row * d.X + x0, cnt, v, thr, (d.X & 3) == 0);
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

inline unsigned grid_for(int64_t n_units) { return (unsigned)((n_units + kThreads - 1) / kThreads); }

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
    const int64_t rows = N * Tz * Ty * Tx * C * Bz * By;
    const int64_t n_units = rows * upr;
    BE_REQUIRE(n_units / kThreads < ((int64_t)1 << 31) - 1, BE_ERR_TOO_LARGE, "be_tile_gather_f32: too many tiles");
    cudaStream_t s = (cudaStream_t)stream;
    if (norm_mode)
        tile_gather_kernel<true><<<grid_for(n_units), kThreads, 0, s>>>(img, tiles, d, zsrc, ysrc, xsrc, norm, norm_mode,
                                                                        clip, clip_lo, clip_hi, upr, n_units);
    else
        tile_gather_kernel<false><<<grid_for(n_units), kThreads, 0, s>>>(img, tiles, d, zsrc, ysrc, xsrc, norm, 0, 0, 0.f,
                                                                         0.f, upr, n_units);
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
    const int64_t n_units = N * C * Z * Y * upr;
    BE_REQUIRE(n_units / kThreads < ((int64_t)1 << 31) - 1, BE_ERR_TOO_LARGE, "be_stitch_crop_f32: output too large");
    stitch_crop_kernel<<<grid_for(n_units), kThreads, 0, (cudaStream_t)stream>>>(tiles, out, mask, d, zsel, ysel, xsel,
                                                                                 epilogue, thr, upr, n_units);
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
    const Strides st{Sz, Sy, Sx};
    const int upr = (int)((X + 3) / 4);
    const int64_t n_units = N * C * Z * Y * upr;
    BE_REQUIRE(n_units / kThreads < ((int64_t)1 << 31) - 1, BE_ERR_TOO_LARGE, "be_stitch_blend_f32: output too large");
    stitch_blend_kernel<<<grid_for(n_units), kThreads, 0, (cudaStream_t)stream>>>(tiles, out, mask, d, st, wz, wy, wx,
                                                                                  epilogue, thr, upr, n_units);
    BE_LAUNCH_CHECK();
    return BE_OK;
}

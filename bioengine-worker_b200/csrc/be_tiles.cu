// Halo-tile gather and the two stitchers.  All three move data (HBM-bound by design): one thread makes four consecutive
// x outputs and stores them as one 16-byte (fp32) / 4-byte (uint8) transaction.
//
//   be_tile_gather_f32   image (+ fused normalise) -> (N*T, C, Bz, By, Bx) tiles, table-driven padding
//   be_stitch_crop_f32   bioimageio inner-region reassembly, table-driven, optional sigmoid / mask epilogue
//   be_stitch_blend_f32  fibsem Gaussian-weighted overlap blend as a gather, fp64 accumulation
//
// The gather and the blend each have two kernels: a general one (any geometry, per-thread loads) and, for geometries
// whose tile rows are aligned 16-byte runs, one wave of persistent blocks in which a copy warp streams the rows into a
// ring of shared-memory stages with TMA bulk copies (cp.async.bulk + mbarrier) while eight arithmetic warps work from
// shared memory.
#include <stdlib.h>

#include <cuda.h>   // CUtensorMap and the encode entry point's signature; the entry point itself is fetched at run time

#include "be_common.cuh"

namespace {

constexpr int kThreads = 256;

struct Dims {
    int64_t N, C, Z, Y, X;
    int Tz, Ty, Tx, Bz, By, Bx;
};

// ---- TMA bulk copies + mbarriers (shared by the gather and the blend): one warp fetches contiguous runs of global
//      memory into a ring of shared-memory stages; `full` barriers count the bytes in, `empty` barriers the readers out ----
constexpr int kCopyThreads = kThreads + 32;   // eight arithmetic warps + the copy warp

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
__device__ __forceinline__ void gather_direct(const float* __restrict__ img, float* __restrict__ tiles, const Dims& d,
                                              const int32_t* __restrict__ zsrc, const int32_t* __restrict__ ysrc,
                                              const int32_t* __restrict__ xsrc, const double* __restrict__ norm,
                                              int norm_mode, float clip_lo, float clip_hi, int units_per_row,
                                              int rows_per_band, unsigned tile, unsigned cz, unsigned band) {
    const int ucols = min(units_per_row, kThreads);     // column units the block covers per sweep
    const int u_in = threadIdx.x % ucols;
    const int r_in = threadIdx.x / ucols;               // first row offset inside the band
    const int r_step = kThreads / ucols;                // rows the block covers per sweep
    if (r_in >= r_step) return;
    unsigned t = tile;
    const int kx = t % d.Tx;
    t /= d.Tx;
    const int ky = t % d.Ty;
    t /= d.Ty;
    const int kz = t % d.Tz;
    const int64_t n = t / d.Tz;
    const int c = cz / d.Bz, iz = cz - c * d.Bz;
    const int iy0 = band * rows_per_band, iy1 = min(iy0 + rows_per_band, d.By);

    const int64_t sz = zsrc[kz * d.Bz + iz];
    const float* plane = img + ((n * d.C + c) * d.Z + sz) * d.Y * d.X;
    const int32_t* ys = ysrc + ky * d.By;
    float* dst_tile = tiles + (((int64_t)tile * d.C + c) * d.Bz + iz) * (int64_t)d.By * d.Bx;

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

template <bool NORM, bool CLIP>
__global__ void __launch_bounds__(kThreads) tile_gather_kernel(const float* __restrict__ img, float* __restrict__ tiles,
                                                               Dims d, const int32_t* __restrict__ zsrc,
                                                               const int32_t* __restrict__ ysrc,
                                                               const int32_t* __restrict__ xsrc,
                                                               const double* __restrict__ norm, int norm_mode,
                                                               float clip_lo, float clip_hi, int units_per_row,
                                                               int rows_per_band) {
    gather_direct<NORM, CLIP>(img, tiles, d, zsrc, ysrc, xsrc, norm, norm_mode, clip_lo, clip_hi, units_per_row,
                              rows_per_band, blockIdx.x, blockIdx.y, blockIdx.z);
}

// The same gather fed by TMA bulk copies, for the usual geometry (Bx a multiple of 4 whose unit count divides 256, X a
// multiple of 4, 16-byte aligned buffers).  A tile row whose x sources are one contiguous, aligned run of the image row
// (every tile that does not touch the left / right border) is a single cp.async.bulk.
//
// Persistent blocks: the grid is one wave (3 blocks per SM) and every block walks work items (tile, plane, band of
// rows) with a grid stride.  The copy warp runs ahead of the arithmetic warps across item boundaries, so the ring of
// kGatherStages x 16 KB stages never drains: with one block per item (the direct kernel) every block pays table
// look-ups and a first round of loads before it moves a byte, and at ~90 rows per block that start-up was a third of
// the kernel.  Border tiles (reflected sources) take the direct path inside the same launch.
constexpr int kGatherStages = 4;
constexpr int kGatherSweeps = 4;                                   // sweeps of the 256 threads per stage
constexpr int kGatherStageBytes = kGatherSweeps * kThreads * 16;   // 16 KB: 16 rows of a 256-px tile
constexpr int kGatherMaxTx = 2048;

struct GatherItem {
    unsigned tile, cz, band;
    int kx, ky, kz, c, iz, iy0, iy1;
    int64_t n;
};

__device__ __forceinline__ GatherItem gather_item(unsigned item, const Dims& d, int cz_count, int bands, int rows_per_band) {
    GatherItem it;
    it.band = item % bands;
    const unsigned rest = item / bands;
    it.cz = rest % cz_count;
    it.tile = rest / cz_count;
    unsigned t = it.tile;
    it.kx = t % d.Tx;
    t /= d.Tx;
    it.ky = t % d.Ty;
    t /= d.Ty;
    it.kz = t % d.Tz;
    it.n = t / d.Tz;
    it.c = it.cz / d.Bz;
    it.iz = it.cz - it.c * d.Bz;
    it.iy0 = it.band * rows_per_band;
    it.iy1 = min(it.iy0 + rows_per_band, d.By);
    return it;
}

template <bool NORM, bool CLIP>
__global__ void __launch_bounds__(kCopyThreads, 3) tile_gather_tma_kernel(const float* __restrict__ img, float* __restrict__ tiles,
                                                                       Dims d, const int32_t* __restrict__ zsrc,
                                                                       const int32_t* __restrict__ ysrc,
                                                                       const int32_t* __restrict__ xsrc,
                                                                       const double* __restrict__ norm, int norm_mode,
                                                                       float clip_lo, float clip_hi, int units_per_row,
                                                                       int rows_per_band, int bands, unsigned n_items,
                                                                       const __grid_constant__ CUtensorMap tmap, int use_tmap) {
    be_pdl_wait();   // the statistics kernel that wrote `norm`
    extern __shared__ __align__(128) unsigned char s_stage[];   // kGatherStages * kGatherStageBytes
    __shared__ __align__(8) unsigned long long s_bar[2 * kGatherStages];
    __shared__ unsigned s_ok[kGatherMaxTx / 32];   // bit kx: the x sources of tile column kx are one aligned run
    const int tid = threadIdx.x;
    for (int i = tid; i < kGatherMaxTx / 32; i += kCopyThreads) s_ok[i] = 0xffffffffu;
    const unsigned smem0 = (unsigned)__cvta_generic_to_shared(s_stage);
    const unsigned bar0 = (unsigned)__cvta_generic_to_shared(s_bar);
    if (tid == 0) {
#pragma unroll
        for (int i = 0; i < kGatherStages; ++i) {
            mbar_init(bar0 + 8 * i, 1);
            mbar_init(bar0 + 8 * (kGatherStages + i), kThreads / 32);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    for (int i = tid; i < d.Tx * units_per_row; i += kCopyThreads) {
        const int kx = i / units_per_row, u = i - kx * units_per_row;
        const int x_first = xsrc[kx * d.Bx];
        const int4 xi = *reinterpret_cast<const int4*>(xsrc + kx * d.Bx + 4 * u);
        const bool ok = (x_first & 3) == 0 && xi.x == x_first + 4 * u && xi.y == xi.x + 1 && xi.z == xi.x + 2 && xi.w == xi.x + 3;
        if (!ok) atomicAnd(&s_ok[kx >> 5], ~(1u << (kx & 31)));
    }
    __syncthreads();

    const int r_sweep = kThreads / units_per_row;       // rows per sweep; units_per_row divides kThreads here
    const int r_stage = r_sweep * kGatherSweeps;        // rows per stage
    const int cz_count = (int)d.C * d.Bz;
    const unsigned row_bytes = (unsigned)d.Bx * 4;

    if (tid >= kThreads) {   // the copy warp: lane r fetches rows r, r + 32, ... of every step
        const int lane = tid & 31;
        unsigned st = 0, par = 0;
        for (unsigned item = blockIdx.x; item < n_items; item += gridDim.x) {
            const GatherItem it = gather_item(item, d, cz_count, bands, rows_per_band);
            if (!((s_ok[it.kx >> 5] >> (it.kx & 31)) & 1u)) continue;
            const int64_t sz = zsrc[it.kz * d.Bz + it.iz];
            const int x_first = xsrc[it.kx * d.Bx];
            const int64_t plane_row = ((it.n * d.C + it.c) * d.Z + sz) * d.Y;   // row of the image seen as one (rows, X) matrix
            const float* plane = img + plane_row * d.X + x_first;
            const int32_t* ys = ysrc + it.ky * d.By;
            const int n_steps = (it.iy1 - it.iy0 + r_stage - 1) / r_stage;
            for (int g = 0; g < n_steps; ++g) {
                const int row0 = it.iy0 + g * r_stage;
                const int rows = min(r_stage, it.iy1 - row0);
                // the source rows first: these loads do not depend on the stage being free
                int64_t off[2];
#pragma unroll
                for (int k = 0; k < 2; ++k) off[k] = (lane + 32 * k < rows) ? (int64_t)ys[row0 + lane + 32 * k] * d.X : 0;
                // a full stage of consecutive image rows (every tile away from the bottom border) is ONE box of the tensor
                // map: one instruction instead of one bulk copy per 1-KB row
                bool box = use_tmap && rows == r_stage && r_stage <= 64;
                if (box) {
                    const int64_t first = (int64_t)ys[row0] * d.X;
                    bool run = true;
#pragma unroll
                    for (int k = 0; k < 2; ++k)
                        if (lane + 32 * k < rows) run &= off[k] == first + (int64_t)(lane + 32 * k) * d.X;
                    box = __all_sync(0xffffffffu, run);
                }
                mbar_wait(bar0 + 8 * (kGatherStages + st), par ^ 1);
                if (lane == 0) {
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic reads before async writes
                    mbar_expect_tx(bar0 + 8 * st, rows * row_bytes);
                    if (box)
                        tensor_g2s_2d(smem0 + st * kGatherStageBytes, &tmap, x_first, (int)(plane_row + ys[row0]), bar0 + 8 * st);
                }
                __syncwarp();
                if (box) {
                    if (++st == kGatherStages) {
                        st = 0;
                        par ^= 1;
                    }
                    continue;
                }
#pragma unroll
                for (int k = 0; k < 2; ++k)
                    if (lane + 32 * k < rows)
                        bulk_g2s(smem0 + st * kGatherStageBytes + (lane + 32 * k) * row_bytes, plane + off[k], row_bytes,
                                 bar0 + 8 * st);
                for (int r = lane + 64; r < rows; r += 32)   // tiles narrower than 64 px: more than 64 rows per stage
                    bulk_g2s(smem0 + st * kGatherStageBytes + r * row_bytes, plane + (int64_t)ys[row0 + r] * d.X,
                             row_bytes, bar0 + 8 * st);
                if (++st == kGatherStages) {
                    st = 0;
                    par ^= 1;
                }
            }
        }
        return;
    }

    // thread -> (row of the sweep, 4-px unit): its 16 bytes sit at (sweep * kThreads + tid) * 16 of the stage
    const int r_in = tid / units_per_row, u_in = tid - r_in * units_per_row;
    const int64_t dsweep = (int64_t)r_sweep * d.Bx;
    unsigned st = 0, par = 0, sa = smem0 + tid * 16;
    double sub = 0.0, div = 1.0, rdiv = 1.0;
    int64_t have = -1;   // segment whose normalisation constants are loaded
    for (unsigned item = blockIdx.x; item < n_items; item += gridDim.x) {
        const GatherItem it = gather_item(item, d, cz_count, bands, rows_per_band);
        if (!((s_ok[it.kx >> 5] >> (it.kx & 31)) & 1u)) {
            gather_direct<NORM, CLIP>(img, tiles, d, zsrc, ysrc, xsrc, norm, norm_mode, clip_lo, clip_hi, units_per_row,
                                      rows_per_band, it.tile, it.cz, it.band);
            continue;
        }
        if (NORM) {
            const int64_t s = (norm_mode == 1) ? (it.n * d.C + it.c) : it.n;
            if (s != have) {
                sub = norm[2 * s];
                div = norm[2 * s + 1];
                rdiv = be_fast_rcp(div);
                have = s;
            }
        }
        float* dst = tiles + ((((int64_t)it.tile * d.C + it.c) * d.Bz + it.iz) * (int64_t)d.By + it.iy0 + r_in) * d.Bx + u_in * 4;
        int row = it.iy0 + r_in;
        const int n_steps = (it.iy1 - it.iy0 + r_stage - 1) / r_stage;
        for (int g = 0; g < n_steps; ++g) {
            if (!mbar_try_wait(bar0 + 8 * st, par)) mbar_wait(bar0 + 8 * st, par);
            float4 v[kGatherSweeps];
#pragma unroll
            for (int k = 0; k < kGatherSweeps; ++k) v[k] = lds128(sa + k * kThreads * 16);
            __syncwarp();
            if ((tid & 31) == 0) mbar_arrive(bar0 + 8 * (kGatherStages + st));   // this warp has read its part of the stage
#pragma unroll
            for (int k = 0; k < kGatherSweeps; ++k, row += r_sweep, dst += dsweep) {
                if (row < it.iy1) {
                    if (NORM) {
                        v[k].x = norm_apply<CLIP>(v[k].x, sub, div, rdiv, clip_lo, clip_hi);
                        v[k].y = norm_apply<CLIP>(v[k].y, sub, div, rdiv, clip_lo, clip_hi);
                        v[k].z = norm_apply<CLIP>(v[k].z, sub, div, rdiv, clip_lo, clip_hi);
                        v[k].w = norm_apply<CLIP>(v[k].w, sub, div, rdiv, clip_lo, clip_hi);
                    }
                    st_stream_f4(reinterpret_cast<float4*>(dst), v[k]);
                }
            }
            sa += kGatherStageBytes;
            if (++st == kGatherStages) {
                st = 0;
                par ^= 1;
                sa -= kGatherStages * kGatherStageBytes;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------
// stitch: shared epilogue
// ---------------------------------------------------------------------------------------------------

// Packed mask: a row is ceil(X / 64) 64-bit words, lowest bit = first pixel; written as 32-bit pieces.  Called by every
// lane of a warp whose lanes own consecutive 4-px units of one row starting at a multiple of 128 px (x0 = the lane's
// first pixel, also for lanes past the end of the row, which pass m4 = 0 so that the last word's tail is clean).
__device__ __forceinline__ void store_mask_bits(uint32_t* __restrict__ row32, int x0, unsigned m4, int pieces_per_row) {
    unsigned c = m4 << (4 * (threadIdx.x & 7));
    c |= __shfl_xor_sync(0xffffffffu, c, 1);
    c |= __shfl_xor_sync(0xffffffffu, c, 2);
    c |= __shfl_xor_sync(0xffffffffu, c, 4);
    if ((threadIdx.x & 7) == 0 && (x0 >> 5) < pieces_per_row) row32[x0 >> 5] = c;
}

// returns the four mask bits
__device__ __forceinline__ unsigned finish4(float* __restrict__ out, uint8_t* __restrict__ mask, int64_t o, int cnt,
                                            float v[4], const Epilogue& e, bool aligned) {
    bool m[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        if (e.fast == 1) {
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
    unsigned m4 = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) m4 |= (j < cnt && m[j]) ? (1u << j) : 0u;
    return m4;
}

// grid: x = 4-px units along the row, y = output row y, z = (n, c, z) planes (grid-strided)
struct RowIndex {
    int u, y, z, c;
    int64_t n, row;
};

__device__ __forceinline__ bool row_index(const Dims& d, int units_per_row, unsigned plane, RowIndex* r) {
    r->u = blockIdx.x * kThreads + threadIdx.x;
    r->y = blockIdx.y;
    r->z = plane % (unsigned)d.Z;
    const unsigned nc = plane / (unsigned)d.Z;
    r->c = nc % (unsigned)d.C;
    r->n = nc / (unsigned)d.C;
    r->row = (int64_t)plane * d.Y + r->y;
    return r->u < units_per_row;
}

// ---------------------------------------------------------------------------------------------------
// crop stitch
// ---------------------------------------------------------------------------------------------------

__global__ void __launch_bounds__(kThreads) stitch_crop_kernel(const float* __restrict__ tiles, float* __restrict__ out,
                                                               uint8_t* __restrict__ mask, uint32_t* __restrict__ bits32,
                                                               Dims d, const int32_t* __restrict__ zsel,
                                                               const int32_t* __restrict__ ysel,
                                                               const int32_t* __restrict__ xsel, Epilogue e,
                                                               int units_per_row, unsigned n_planes) {
    const int pieces = 2 * (int)((d.X + 63) / 64);
    for (unsigned plane = blockIdx.z; plane < n_planes; plane += gridDim.z) {
        RowIndex r;
        const bool live = row_index(d, units_per_row, plane, &r);
        if (!live) {   // past the end of the row: only the packed mask wants this lane (with a clean tail)
            if (bits32) store_mask_bits(bits32 + r.row * pieces, r.u * 4, 0u, pieces);
            continue;
        }
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
        const unsigned m4 = finish4(out, mask, r.row * d.X + x0, cnt, v, e, (d.X & 3) == 0);
        if (bits32) store_mask_bits(bits32 + r.row * pieces, x0, m4, pieces);
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
                                                                uint8_t* __restrict__ mask, uint32_t* __restrict__ bits32,
                                                                Dims d, Strides s,
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
        unsigned m4 = 0;
        if (live) m4 = finish4(out, mask, ((int64_t)plane * d.Y + y) * d.X + x0, cnt, v, e, (d.X & 3) == 0);
        if (bits32) {
            const int pieces = 2 * (int)((d.X + 63) / 64);
            store_mask_bits(bits32 + ((int64_t)plane * d.Y + y) * pieces, (int)(blockIdx.x * kThreads + threadIdx.x) * 4, m4,
                            pieces);
        }
    }
}

// In-band pixels of the division-free test: the exact quotient and numpy's fp32 sigmoid.  Out of line on purpose: a
// vanishing fraction of pixels gets here and the inlined code would cost every thread registers.
static __device__ __noinline__ bool blend_decide_exact(double acc, double wacc, float lo, float hi, float thr) {
    const float v = be_div_to_f32(acc, wacc, be_fast_rcp(wacc));
    return (v > hi) || (!(v < lo) && be_sigmoid_f32(v) > thr);
}
static __device__ __noinline__ bool blend_decide_exact_plain(double acc, double wacc, float thr) {
    return be_div_to_f32(acc, wacc, be_fast_rcp(wacc)) > thr;
}

// Row-walking variant for the usual geometry (Bx <= 2*Sx, By <= 2*Sy; Sx, Bx, X multiples of 4; 16-byte aligned
// buffers): every 4-px unit is covered by the same one or two tiles along x and every tile access is one aligned
// 16-byte read.  A block owns an x range of bx*4 pixels and a band of rows; a thread owns one 4-px unit and walks the
// rows, so its x cover, eight window values and shared-memory offsets live in registers for the whole band.
//
// Data movement: the tile rows a step needs are contiguous runs of the tile buffer, so warp 0 fetches them with TMA bulk
// copies (cp.async.bulk, one per tile row segment, clipped to the block's x range) into a kStages-deep ring of
// shared-memory stages, two steps ahead of the arithmetic; an mbarrier per stage counts the bytes in.  The other
// warps issue no loads at all.  A step is one (row group, kz) pair — the kz tiles of a row are consecutive steps
// adding into the same accumulators, in the reference's ascending (kz, ky, kx) order.
//
// FAST (mask-only sigmoid): a pixel is decided by |acc - wacc*mid| > wacc*half (one DFMA + DMUL instead of the exact
// division) unless it falls in the safety band around logit(thr); only those take the division and numpy's sigmoid.
// Block: bx = 1 << log2_bx threads along x, r_step = kThreads / bx rows per step; grid (x blocks, row bands, planes).
// measured on the 4096^2 / 256-px-tile workload: 3..5 stages within 2 %; 3 blocks of 72 registers per SM beat both
// 2 x 96 (too few warps to hide the fp64 / conversion latencies) and 4 x 56 (spills)
constexpr int kStages = 4;

// four pixels of one tile row into the accumulators: w = fp32(wzy * wx), acc += fp64(fp32(p * w)), wacc += fp64(w)
__device__ __forceinline__ double2 lds_d2(unsigned addr) {
    double2 v;
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(addr));
    return v;
}
__device__ __forceinline__ void blend_add4(const float4 p, const double wzy, unsigned w_addr, unsigned w_half,
                                           double (&acc)[4], double (&wacc)[4]) {
    const float pv[4] = {p.x, p.y, p.z, p.w};
    const double2 w01 = lds_d2(w_addr), w23 = lds_d2(w_addr + w_half);
    const double wx4[4] = {w01.x, w01.y, w23.x, w23.y};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float w = (float)__dmul_rn(wzy, wx4[j]);
        acc[j] = __dadd_rn(acc[j], (double)__fmul_rn(pv[j], w));
        wacc[j] = __dadd_rn(wacc[j], (double)w);
    }
}

// launch-invariant numbers of the row-walking kernel, worked out by the host so that they are constant-bank operands
struct RowsConst {
    int64_t dky, dkz;        // element steps in the tile buffer: next tile along y / z (same in-tile position + stride)
    int64_t dout;            // element step in the output: next row of this thread
    double mid, half;        // division-free band test: |acc - wacc*mid| > wacc*half  <=>  acc/wacc outside [lo2, hi2]
    int log2_bx, units_per_row;
    unsigned n_bands, n_items;     // work items = (plane, band); band b covers rows [b*Y/n_bands, (b+1)*Y/n_bands)
    int slot_bytes, stage_bytes;   // shared memory: one (row, ky slot) = nseg tile rows; one stage = r_step rows x 2 slots
    unsigned n_planes;
};

// tiles covering row y along y: first tile, how many (1 or 2), first row at which this changes
struct RowCover {
    int ky_lo, nky, change;
    __device__ __forceinline__ void set(int y, const Dims& d, const Strides& s) {
        int hi;
        cover(y, d.By, s.Sy, s.rSy, d.Ty, &ky_lo, &hi);
        nky = hi - ky_lo + 1;
        const int up = (hi + 1 < d.Ty) ? (hi + 1) * s.Sy : 0x7fffffff;
        change = min(up, ky_lo * s.Sy + d.By);
    }
};

constexpr int kRowsThreads = kCopyThreads;

template <bool FAST>
__global__ void __launch_bounds__(kRowsThreads, 3) stitch_blend_rows_kernel(const float* __restrict__ tiles, float* __restrict__ out,
                                                                        uint8_t* __restrict__ mask, uint32_t* __restrict__ bits32,
                                                                        const Dims d,
                                                                        const Strides s, const double* __restrict__ wz,
                                                                        const double* __restrict__ wy,
                                                                        const double* __restrict__ wx, const Epilogue e,
                                                                        const RowsConst k) {
    extern __shared__ __align__(128) unsigned char s_dyn[];
    __shared__ __align__(8) unsigned long long s_bar[2 * kStages];   // full[kStages], empty[kStages]
    const int tid = threadIdx.x;
    const int log2_bx = k.log2_bx;
    const int tx = tid & ((1 << log2_bx) - 1), ty = tid >> log2_bx;
    const int r_step = kThreads >> log2_bx;
    const int u = (blockIdx.x << log2_bx) + tx;
    const bool x_live = u < k.units_per_row;
    const int x0 = min(u, k.units_per_row - 1) * 4;
    const int X0 = (blockIdx.x << log2_bx) * 4, X1 = min((int)d.X, X0 + (4 << log2_bx));   // the block's x range
    // tiles overlapping the block's x range: the segments of a stage
    int seg_first, seg_last;
    cover(X0, d.Bx, s.Sx, s.rSx, d.Tx, &seg_first, &seg_last);
    seg_last = min(d.Tx - 1, be_fast_div(X1 - 1, s.Sx, s.rSx));
    const int seg_bytes = d.Bx * 4;                      // a segment slot holds a whole tile row
    const int slot_bytes = k.slot_bytes;                 // one (row, ky slot): nseg segments
    const int stage_bytes = k.stage_bytes;               // r_step rows x 2 ky slots
    const unsigned smem0 = (unsigned)__cvta_generic_to_shared(s_dyn);
    const unsigned bar0 = (unsigned)__cvta_generic_to_shared(s_bar);

    int kxa, kxb;
    cover(x0, d.Bx, s.Sx, s.rSx, d.Tx, &kxa, &kxb);
    const bool two = kxb > kxa;
    const int ixa = x0 - kxa * s.Sx, ixb = x0 - kxb * s.Sx;
    // the x window lives in shared memory behind the stages, split so that a warp's reads are contiguous:
    // A[u] = (wx[4u], wx[4u+1]), B[u] = (wx[4u+2], wx[4u+3]); a thread reads entries ixa/4 and ixb/4
    const unsigned wxs = smem0 + kStages * stage_bytes;
    for (int i = tid; i < d.Bx / 4; i += kRowsThreads) {
        double2* A = reinterpret_cast<double2*>(s_dyn + kStages * stage_bytes);
        double2* B = A + d.Bx / 4;
        A[i] = make_double2(wx[4 * i], wx[4 * i + 1]);
        B[i] = make_double2(wx[4 * i + 2], wx[4 * i + 3]);
    }
    const unsigned wa = wxs + ixa * 4, wb = wxs + ixb * 4, wB = d.Bx * 4;   // byte addresses of A[ix/4]; B is wB further
    // this thread's two 16-byte reads inside a (row, ky slot): byte offsets from the start of a stage
    const int lda = (ty * 2) * slot_bytes + (kxa - seg_first) * seg_bytes + ixa * 4;
    const int ab = (kxb - kxa) * seg_bytes + (ixb - ixa) * 4;   // from the read of the first x tile to the second

    // producer lane i of warp 0 owns segment i: the part of tile row (kx = seg_first + i) inside [X0, X1)
    const bool producer = tid >= kThreads;
    const int lane = tid & 31;
    int seg_src = 0, seg_dst = 0, seg_len = 0;   // element offset in the tile row, byte offset in the slot, bytes
    if (producer) {
        const int kx = seg_first + lane;
        if (kx <= seg_last) {
            const int xs = max(kx * s.Sx, X0), xe = min(kx * s.Sx + d.Bx, X1);
            seg_src = xs - kx * s.Sx;
            seg_dst = lane * seg_bytes + seg_src * 4;
            seg_len = (xe - xs) * 4;
        }
    }
    int row_bytes = seg_len;   // bytes one (row, ky slot) brings in
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) row_bytes += __shfl_xor_sync(0xffffffffu, row_bytes, o);

    if (tid == 0) {
#pragma unroll
        for (int i = 0; i < kStages; ++i) {
            mbar_init(bar0 + 8 * i, 1);                           // full: the copy warp's expect_tx arrival + the bytes
            mbar_init(bar0 + 8 * (kStages + i), kThreads / 32);   // empty: one arrival per arithmetic warp
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    unsigned use = 0;   // steps this warp has been through (stage = use % kStages, parity = (use / kStages) & 1)

    // persistent blocks: the grid is one wave and a block walks work items (plane, band of rows) with a grid stride, so
    // the window table, the barriers and the x cover are set up once and the ring of stages never drains between bands
    for (unsigned item = blockIdx.y; item < k.n_items; item += gridDim.y) {
        const unsigned plane = item / k.n_bands, band = item - plane * k.n_bands;
        const int band_begin = (int)(((int64_t)band * d.Y) / k.n_bands);
        const int y_end = (int)(((int64_t)(band + 1) * d.Y) / k.n_bands);
        const int n_groups = (y_end - band_begin + r_step - 1) / r_step;
        int z = 0, c = 0;
        int64_t n = 0;
        if (k.n_planes > 1) {
            z = plane % (unsigned)d.Z;
            const unsigned nc = plane / (unsigned)d.Z;
            c = nc % (unsigned)d.C;
            n = nc / (unsigned)d.C;
        }
        int kz_lo = 0, kz_hi = 0;
        if (d.Z > 1) cover(z, d.Bz, s.Sz, s.rSz, d.Tz, &kz_lo, &kz_hi);
        const int nkz = kz_hi - kz_lo + 1;
        const int iz0 = z - kz_lo * s.Sz;
        const int total = n_groups * nkz;
        // tile (n, kz_lo, ky = 0, kx = seg_first + lane), channel c, plane iz0, row 0, first element this lane copies
        const float* lane_base = tiles + ((n * d.Tz + kz_lo) * d.Ty * d.Tx + seg_first + lane) * s.tile_stride +
                                 ((int64_t)c * d.Bz + iz0) * d.By * d.Bx + seg_src;

        if (producer) {
            // the copy warp: fetch step after step into the ring, as far ahead as free stages allow
            for (int ts = 0; ts < total; ++ts, ++use) {
                const int g = ts / nkz, kk = ts - g * nkz;
                const unsigned st = use % kStages;
                const unsigned full = bar0 + 8 * st, empty = bar0 + 8 * (kStages + st);
                int slots = 0;
                for (int r = 0; r < r_step; ++r) {
                    const int y = band_begin + g * r_step + r;
                    if (y < y_end) {
                        RowCover rc;
                        rc.set(y, d, s);
                        slots += rc.nky;
                    }
                }
                mbar_wait(empty, ((use / kStages) & 1) ^ 1);   // all arithmetic warps are done with this stage
                if (lane == 0) {
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic reads before async writes
                    mbar_expect_tx(full, (unsigned)(slots * row_bytes));
                }
                __syncwarp();
                if (seg_len > 0) {
                    for (int r = 0; r < r_step; ++r) {
                        const int y = band_begin + g * r_step + r;
                        if (y >= y_end) break;
                        RowCover rc;
                        rc.set(y, d, s);
                        const float* src = lane_base + kk * k.dkz + (int64_t)rc.ky_lo * d.Tx * s.tile_stride +
                                           (int64_t)(y - rc.ky_lo * s.Sy) * d.Bx;
                        const unsigned dst = smem0 + st * stage_bytes + (r * 2) * slot_bytes + seg_dst;
                        bulk_g2s(dst, src, (unsigned)seg_len, full);
                        if (rc.nky > 1) bulk_g2s(dst + slot_bytes, src + k.dky, (unsigned)seg_len, full);
                    }
                }
            }
            continue;
        }
        // arithmetic warps.  Loop-carried state is kept as plain registers that only ever get an add: the stage cursor
        // (index, parity, shared-memory address of this thread's first read), the row cursor (y, in-tile row, window
        // pointer, output pointer), the kz cursor.
        int y = band_begin + ty;
        RowCover rc;
        rc.set(min(y, y_end - 1), d, s);
        int iy = y - rc.ky_lo * s.Sy, kk = 0;
        bool live = x_live && y < y_end;
        uint8_t* mrow = mask ? mask + ((int64_t)plane * d.Y + y) * d.X + x0 : nullptr;
        float* orow = out ? out + ((int64_t)plane * d.Y + y) * d.X + x0 : nullptr;
        double acc[4] = {0.0, 0.0, 0.0, 0.0}, wacc[4] = {0.0, 0.0, 0.0, 0.0};
        double wzv = wz[iz0];
        // y window values of the current row, fetched a step ahead of their use
        double wy0 = wy[max(iy, 0)], wy1 = (rc.nky > 1) ? wy[iy - s.Sy] : 0.0;
        unsigned st = use % kStages, par = (use / kStages) & 1;
        unsigned sa = smem0 + st * stage_bytes + lda;   // this thread's read of the first x tile, ky slot 0

        for (int t = 0; t < total; ++t) {
            if (!mbar_try_wait(bar0 + 8 * st, par)) mbar_wait(bar0 + 8 * st, par);
            if (live) {
                const double wzy0 = __dmul_rn(wzv, wy0);
                blend_add4(lds128(sa), wzy0, wa, wB, acc, wacc);
                if (two) blend_add4(lds128(sa + ab), wzy0, wb, wB, acc, wacc);
                if (rc.nky > 1) {
                    const double wzy1 = __dmul_rn(wzv, wy1);
                    blend_add4(lds128(sa + slot_bytes), wzy1, wa, wB, acc, wacc);
                    if (two) blend_add4(lds128(sa + slot_bytes + ab), wzy1, wb, wB, acc, wacc);
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(bar0 + 8 * (kStages + st));   // this warp has read its part of the stage
            ++use;
            sa += stage_bytes;
            if (++st == kStages) {
                st = 0;
                par ^= 1;
                sa -= kStages * stage_bytes;
            }
            if (++kk < nkz) {
                wzv = wz[iz0 - kk * s.Sz];
                continue;
            }
            // last kz of this row group: finish the four pixels, move to the next group
            unsigned m4 = 0;
            if (live) {
                if (FAST) {
                    // m = acc/wacc above the band <=> dlt > 0 (sign bit clear; dlt == 0 is inside the band).  The test
                    // trusts wacc only when it is a positive normal number (high word > 0; a NaN passes that but
                    // fails the comparison), everything else takes the exact path
                    unsigned bits = 0, slow = 0;
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const double dlt = fma(wacc[j], -k.mid, acc[j]);
                        const bool clear = (fabs(dlt) > __dmul_rn(wacc[j], k.half)) & (__double2hiint(wacc[j]) > 0);
                        bits |= ((unsigned)(~__double2hiint(dlt)) >> 31) << (8 * j);
                        slow |= clear ? 0u : (1u << j);
                    }
                    if (slow) {
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            if (slow & (1u << j)) {
                                const bool m = (e.fast == 2) ? blend_decide_exact_plain(acc[j], wacc[j], e.thr)
                                                             : blend_decide_exact(acc[j], wacc[j], e.lo, e.hi, e.thr);
                                bits = (bits & ~(0xffu << (8 * j))) | ((m ? 1u : 0u) << (8 * j));
                            }
                    }
                    if (mrow) *reinterpret_cast<unsigned*>(mrow) = bits;
                    m4 = (bits & 1u) | ((bits >> 7) & 2u) | ((bits >> 14) & 4u) | ((bits >> 21) & 8u);
                } else {
                    float v[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) v[j] = be_div_to_f32(acc[j], wacc[j], be_fast_rcp(wacc[j]));
                    m4 = finish4(orow, mrow, 0, 4, v, e, true);
                }
            }
            if (bits32 && y < y_end)   // warp-uniform: a warp's lanes share their row
                store_mask_bits(bits32 + ((int64_t)plane * d.Y + y) * (2 * (int)((d.X + 63) / 64)),
                                ((blockIdx.x << log2_bx) + tx) * 4, m4, 2 * (int)((d.X + 63) / 64));
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[j] = wacc[j] = 0.0;
            kk = 0;
            if (nkz > 1) wzv = wz[iz0];
            y += r_step;
            iy += r_step;
            if (FAST) {
                if (mrow) mrow += k.dout;
            } else {
                if (mrow) mrow += k.dout;
                if (orow) orow += k.dout;
            }
            live = x_live && y < y_end;   // a band need not be a whole number of row groups
            if (y >= rc.change && y < y_end) {
                rc.set(y, d, s);
                iy = y - rc.ky_lo * s.Sy;
            }
            if (y < y_end) {
                wy0 = wy[iy];
                if (rc.nky > 1) wy1 = wy[iy - s.Sy];
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------
// 2-D, mask only: fp32 interval filter
// ---------------------------------------------------------------------------------------------------
//
// When only the mask of a 2-D blend is wanted (sigmoid -> binarize, or a plain threshold on the blended value) the exact
// quotient matters for the few pixels next to the threshold only.  This kernel walks rows like the one above, but
//   * the window is what the reference holds in memory, the fp32 table w[iy][ix] = fp32(wy[iy] * wx[ix])
//     (be_blend_window_f32 writes it once per geometry; 256 KB for 256-px tiles, resident in L2 and mostly in L1);
//   * a pixel's sums are taken in fp32: A = sum fp32(p * w) (the products are the reference's own fp32 numbers),
//     W = sum w, S = sum |p * w|.  With at most four terms |A - A_exact| <= 3u S and |W - W_exact| <= 3u W (u = 2^-24), so
//     d = A - W * mid misses the exact A_exact - W_exact * mid by well under 2^-20 (S + W |mid| + |d|), and a pixel with
//     |d| > W * half + 2^-20 (S + W |mid| + |d|) lies outside the band [mid - half, mid + half] on the side sign(d)
//     shows — the same band the fp64 test of the kernel above uses.  No fp64 instruction and no conversion is issued;
//   * the pixels that fail the test (about one in 10^4) are redone in the reference's arithmetic (fp64 sums of the fp32
//     products in ascending tile order, exact quotient, numpy's sigmoid) from the values the thread still holds;
//   * every tile element is used by exactly one thread, so nothing is staged and no thread waits for another: a thread
//     loads its (up to four) 16-byte pieces of a row and their weights straight into registers, and the warps of an SM
//     cover each other's load latency.  (The ring of TMA bulk copies the kernel above feeds from is bound by the copy
//     engine once the arithmetic is this light: ~0.8 KB per copy, about 6 B / cycle / SM = 1.8 TB/s; see profiles/README.md.)
// The mask leaves bit-packed (one 32-bit piece per eight lanes) and / or as bytes.
struct Mask2dConst {
    int64_t dky;             // element step in the tile buffer: next tile along y (same in-tile column, row - Sy)
    float mid, abs_mid, half_up;
    int log2_bx, units_per_row;
    unsigned n_bands, n_items;
    int pieces;              // 32-bit pieces per row of the packed mask
};

__global__ void __launch_bounds__(kThreads) blend_window_kernel(const double* __restrict__ wy, const double* __restrict__ wx,
                                                                double wz0, int By, int Bx, float* __restrict__ table) {
    const int i = blockIdx.x * kThreads + threadIdx.x;
    if (i >= By * Bx) return;
    const int iy = i / Bx, ix = i - iy * Bx;
    table[i] = (float)__dmul_rn(__dmul_rn(wz0, wy[iy]), wx[ix]);
}

// one pixel again, in the reference's arithmetic: fp64 sums of the fp32 products in ascending (ky, kx) order, exact quotient,
// numpy's sigmoid.  Out of line: about one pixel in 10^4 gets here, and it reloads its few operands itself so that the
// caller keeps nothing alive for it.
static __device__ __noinline__ bool mask2d_exact(const float* __restrict__ tiles, const float* __restrict__ wtab, unsigned pa,
                                                 unsigned ab, unsigned dky, int wa, int wab, int dwy, int two, int nky,
                                                 int fast, float thr, float lo, float hi) {
    double acc = 0.0, wacc = 0.0;
    for (int q = 0; q < nky; ++q) {
        for (int h = 0; h <= two; ++h) {
            const float pv = __ldg(tiles + (pa + (q ? dky : 0u) + (h ? ab : 0u)));
            const float wv = __ldg(wtab + (wa - (q ? dwy : 0) + (h ? wab : 0)));
            acc = __dadd_rn(acc, (double)__fmul_rn(pv, wv));
            wacc = __dadd_rn(wacc, (double)wv);
        }
    }
    const float v = be_div_to_f32(acc, wacc, be_fast_rcp(wacc));
    if (fast == 2) return v > thr;
    return (v > hi) || (!(v < lo) && be_sigmoid_f32(v) > thr);
}

__global__ void __launch_bounds__(kThreads, 4) stitch_blend_mask2d_kernel(const float* __restrict__ tiles,
                                                                          uint8_t* __restrict__ mask,
                                                                          uint32_t* __restrict__ bits32,
                                                                          const float* __restrict__ wtab, const Dims d,
                                                                          const Strides s, const Epilogue e,
                                                                          const Mask2dConst k) {
    be_pdl_wait();   // whatever wrote the tiles
    const int tid = threadIdx.x;
    const int log2_bx = k.log2_bx;
    const int tx = tid & ((1 << log2_bx) - 1), ty = tid >> log2_bx;
    const int r_step = kThreads >> log2_bx;
    const int u_raw = (blockIdx.x << log2_bx) + tx;
    const bool x_live = u_raw < k.units_per_row;
    const int x0 = min(u_raw, k.units_per_row - 1) * 4;
    int kxa, kxb;
    cover(x0, d.Bx, s.Sx, s.rSx, d.Tx, &kxa, &kxb);
    const int two = kxb > kxa;
    const int ixa = x0 - kxa * s.Sx;
    // element offsets (the host checked that the tile buffer has fewer than 2^31 elements): second x tile, second y tile
    const unsigned ab = (unsigned)((kxb - kxa) * s.tile_stride) - (unsigned)((kxb - kxa) * s.Sx);
    const unsigned dky = (unsigned)k.dky;
    const int wab = -(kxb - kxa) * s.Sx;   // in the window row: ixb - ixa
    const int dwy = s.Sy * d.Bx;           // the window row of the second y tile lies Sy rows up

    for (unsigned item = blockIdx.y; item < k.n_items; item += gridDim.y) {
        const unsigned plane = item / k.n_bands, band = item - plane * k.n_bands;
        const int band_begin = (int)(((int64_t)band * d.Y) / k.n_bands);
        const int y_end = (int)(((int64_t)(band + 1) * d.Y) / k.n_bands);
        const unsigned c = plane % (unsigned)d.C;
        const unsigned n = plane / (unsigned)d.C;
        // tile (n, ky = 0, kx = kxa), channel c, row 0, this thread's first pixel
        const unsigned base_a = (unsigned)((n * d.Ty * d.Tx + kxa) * s.tile_stride + (int64_t)c * d.By * d.Bx + ixa);
        RowCover rc;
        rc.change = -1;
        for (int y = band_begin + ty; y < y_end; y += r_step) {   // warp-uniform: a warp's lanes share their row
            if (y >= rc.change) rc.set(y, d, s);
            const int nky = rc.nky;
            const int iy = y - rc.ky_lo * s.Sy;
            unsigned m4 = 0;
            if (x_live) {
                const unsigned pa = base_a + (unsigned)(rc.ky_lo * d.Tx) * (unsigned)s.tile_stride + (unsigned)(iy * d.Bx);
                const int wa = iy * d.Bx + ixa;
                float A[4] = {0.f, 0.f, 0.f, 0.f}, W[4] = {0.f, 0.f, 0.f, 0.f}, S[4] = {0.f, 0.f, 0.f, 0.f};
                auto add4 = [&](const float4& pp, const float4& ww) {
                    const float pv[4] = {pp.x, pp.y, pp.z, pp.w}, wv[4] = {ww.x, ww.y, ww.z, ww.w};
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const float a = __fmul_rn(pv[j], wv[j]);
                        A[j] = __fadd_rn(A[j], a);
                        S[j] = __fadd_rn(S[j], fabsf(a));
                        W[j] = __fadd_rn(W[j], wv[j]);
                    }
                };
                // all pieces of the row are requested before the first is used: it is the number of loads in flight per SM
                // that sets the pace (a pure-load kernel with this address pattern needs 30 us at eight blocks per SM and 43
                // at three; benchmarks/readprobe.cu)
                {
                    const float4 p0 = ld_stream_f4(reinterpret_cast<const float4*>(tiles + pa));
                    float4 p1, p2, p3;
                    if (two) p1 = ld_stream_f4(reinterpret_cast<const float4*>(tiles + (pa + ab)));
                    if (nky > 1) {
                        p2 = ld_stream_f4(reinterpret_cast<const float4*>(tiles + (pa + dky)));
                        if (two) p3 = ld_stream_f4(reinterpret_cast<const float4*>(tiles + (pa + dky + ab)));
                    }
                    add4(p0, __ldg(reinterpret_cast<const float4*>(wtab + wa)));
                    if (two) add4(p1, __ldg(reinterpret_cast<const float4*>(wtab + (wa + wab))));
                    if (nky > 1) {
                        add4(p2, __ldg(reinterpret_cast<const float4*>(wtab + (wa - dwy))));
                        if (two) add4(p3, __ldg(reinterpret_cast<const float4*>(wtab + (wa - dwy + wab))));
                    }
                }
                unsigned slow = 0;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float dlt = __fmaf_rn(W[j], -k.mid, A[j]);
                    const float e1 = __fadd_rn(__fmaf_rn(W[j], k.abs_mid, S[j]), fabsf(dlt));
                    const float bound = __fmaf_rn(e1, 0x1p-20f, __fmul_rn(W[j], k.half_up));
                    const bool clear = (fabsf(dlt) > bound) & (W[j] > 0x1p-60f);   // NaN / inf sums fail the comparison
                    m4 |= (dlt > 0.f ? 1u : 0u) << j;
                    slow |= clear ? 0u : (1u << j);
                }
                while (slow) {
                    const int j = __ffs(slow) - 1;
                    slow &= slow - 1;
                    const bool m = mask2d_exact(tiles, wtab, pa + j, ab, dky, wa + j, wab, dwy, two, nky, e.fast, e.thr, e.lo, e.hi);
                    m4 = (m4 & ~(1u << j)) | ((m ? 1u : 0u) << j);
                }
                if (mask) {
                    const unsigned bytes = (m4 & 1u) | ((m4 & 2u) << 7) | ((m4 & 4u) << 14) | ((m4 & 8u) << 21);
                    *reinterpret_cast<unsigned*>(mask + ((int64_t)plane * d.Y + y) * d.X + x0) = bytes;
                }
            }
            if (bits32) store_mask_bits(bits32 + ((int64_t)plane * d.Y + y) * k.pieces, u_raw * 4, m4, k.pieces);
        }
    }
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

// cuTensorMapEncodeTiled through the runtime's driver entry-point query: no link against libcuda
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn tensor_map_encoder() {
    static EncodeTiledFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        const char* off = getenv("BE_B200_NO_TENSORMAP");
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (!(off && off[0] == '1') && cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
        else
            cudaGetLastError();
    }
    return fn;
}

extern "C" int be_tile_gather_f32(const float* img, float* tiles, int64_t N, int64_t C, int64_t Z, int64_t Y, int64_t X,
                                  int Tz, int Ty, int Tx, int Bz, int By, int Bx, const int32_t* zsrc,
                                  const int32_t* ysrc, const int32_t* xsrc, const double* norm, int norm_mode, int clip,
                                  float clip_lo, float clip_hi, be_stream_t stream) {
    BE_REQUIRE(img && tiles && zsrc && ysrc && xsrc, BE_ERR_INVALID_ARG, "be_tile_gather_f32: null pointer");
    BE_REQUIRE(((((uintptr_t)xsrc) | ((uintptr_t)tiles)) & 15) == 0, BE_ERR_INVALID_ARG,
               "be_tile_gather_f32: xsrc and tiles must be 16-byte aligned (the kernels read the x table and write the tiles "
               "with 16-byte accesses)");
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
    const bool tma = (Bx & 3) == 0 && (X & 3) == 0 && upr <= kThreads && kThreads % upr == 0 && Tx <= kGatherMaxTx &&
                     ((((uintptr_t)img) | ((uintptr_t)tiles)) & 15) == 0;
    if (tma) {
        // work items of two stages of rows each; one wave of persistent blocks walks them
        const int r_stage = kThreads / upr * kGatherSweeps;
        // the image as one (N*C*Z*Y, X) matrix behind a tensor map with a box of one stage (r_stage rows x Bx pixels)
        CUtensorMap tmap;
        memset(&tmap, 0, sizeof(tmap));
        int use_tmap = 0;
        if (Bx <= 256 && r_stage <= 64 && N * C * Z * Y < ((int64_t)1 << 31) && X < ((int64_t)1 << 31)) {
            if (EncodeTiledFn encode = tensor_map_encoder()) {
                const cuuint64_t dims[2] = {(cuuint64_t)X, (cuuint64_t)(N * C * Z * Y)};
                const cuuint64_t strides[1] = {(cuuint64_t)X * sizeof(float)};
                const cuuint32_t box[2] = {(cuuint32_t)Bx, (cuuint32_t)r_stage};
                const cuuint32_t estr[2] = {1, 1};
                use_tmap = encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(img), dims, strides, box, estr,
                                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
            }
        }
        const int rpb = 2 * r_stage;
        const int nb = (By + rpb - 1) / rpb;
        const int64_t n_items = n_tiles * C * Bz * nb;
        BE_REQUIRE(n_items < ((int64_t)1 << 31), BE_ERR_TOO_LARGE, "be_tile_gather_f32: too many work items");
        const size_t smem = (size_t)kGatherStages * kGatherStageBytes;
        const unsigned blocks = (unsigned)(n_items < (int64_t)BE_SM_COUNT * 3 ? n_items : (int64_t)BE_SM_COUNT * 3);
        if (norm_mode && clip) {
            BE_CUDA(cudaFuncSetAttribute(tile_gather_tma_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            BE_CUDA(be_launch_pdl(tile_gather_tma_kernel<true, true>, dim3(blocks), dim3(kCopyThreads), smem, s, img, tiles, d, zsrc, ysrc, xsrc, norm, norm_mode,
                                                                                 clip_lo, clip_hi, upr, rpb, nb, (unsigned)n_items, tmap, use_tmap));
        } else if (norm_mode) {
            BE_CUDA(cudaFuncSetAttribute(tile_gather_tma_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            BE_CUDA(be_launch_pdl(tile_gather_tma_kernel<true, false>, dim3(blocks), dim3(kCopyThreads), smem, s, img, tiles, d, zsrc, ysrc, xsrc, norm, norm_mode,
                                                                                  0.f, 0.f, upr, rpb, nb, (unsigned)n_items, tmap, use_tmap));
        } else {
            BE_CUDA(cudaFuncSetAttribute(tile_gather_tma_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            BE_CUDA(be_launch_pdl(tile_gather_tma_kernel<false, false>, dim3(blocks), dim3(kCopyThreads), smem, s, img, tiles, d, zsrc, ysrc, xsrc, norm, 0, 0.f,
                                                                                   0.f, upr, rpb, nb, (unsigned)n_items, tmap, use_tmap));
        }
    } else if (norm_mode && clip)
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

extern "C" int be_stitch_crop_f32(const float* tiles, float* out, uint8_t* mask, uint64_t* bits, int64_t N, int64_t C, int64_t Z,
                                  int64_t Y, int64_t X, int Tz, int Ty, int Tx, int Bz, int By, int Bx,
                                  const int32_t* zsel, const int32_t* ysel, const int32_t* xsel, int epilogue, float thr,
                                  be_stream_t stream) {
    BE_REQUIRE(tiles && (out || mask || bits) && zsel && ysel && xsel, BE_ERR_INVALID_ARG, "be_stitch_crop_f32: null pointer");
    BE_REQUIRE((((uintptr_t)bits) & 7) == 0, BE_ERR_INVALID_ARG, "be_stitch_crop_f32: bits must be 8-byte aligned");
    int rc = check_dims("be_stitch_crop_f32", N, C, Z, Y, X, Tz, Ty, Tx, Bz, By, Bx);
    if (rc) return rc;
    BE_REQUIRE(epilogue == BE_EPI_NONE || epilogue == BE_EPI_SIGMOID, BE_ERR_INVALID_ARG,
               "be_stitch_crop_f32: epilogue %d", epilogue);
    const Dims d{N, C, Z, Y, X, Tz, Ty, Tx, Bz, By, Bx};
    const int upr = (int)((X + 3) / 4);
    dim3 grid;
    rc = stitch_grid("be_stitch_crop_f32", d, upr, &grid);
    if (rc) return rc;
    if (bits) grid.x = (unsigned)(((X + 63) / 64 * 16 + kThreads - 1) / kThreads);   // lanes for the tail of the last word too
    stitch_crop_kernel<<<grid, kThreads, 0, (cudaStream_t)stream>>>(tiles, out, mask, (uint32_t*)bits, d, zsel, ysel, xsel,
                                                                    make_epilogue(epilogue, thr, out == nullptr), upr,
                                                                    (unsigned)(N * C * Z));
    BE_LAUNCH_CHECK();
    return BE_OK;
}

extern "C" int be_blend_window_f32(const double* wz, const double* wy, const double* wx, int By, int Bx, float* window,
                                   be_stream_t stream) {
    BE_REQUIRE(wy && wx && window, BE_ERR_INVALID_ARG, "be_blend_window_f32: null pointer");
    BE_REQUIRE(By > 0 && Bx > 0 && (int64_t)By * Bx < ((int64_t)1 << 30), BE_ERR_INVALID_ARG, "be_blend_window_f32: bad block shape");
    double wz0 = 1.0;
    if (wz) BE_CUDA(cudaMemcpyAsync(&wz0, wz, sizeof(double), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    if (wz) BE_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    blend_window_kernel<<<(unsigned)(((int64_t)By * Bx + kThreads - 1) / kThreads), kThreads, 0, (cudaStream_t)stream>>>(
        wy, wx, wz0, By, Bx, window);
    BE_LAUNCH_CHECK();
    return BE_OK;
}

extern "C" int be_stitch_blend_f32(const float* tiles, float* out, uint8_t* mask, uint64_t* bits, int64_t N, int64_t C,
                                   int64_t Z, int64_t Y, int64_t X, int Tz, int Ty, int Tx, int Bz, int By, int Bx, int Sz, int Sy,
                                   int Sx, const double* wz, const double* wy, const double* wx, const float* window,
                                   int epilogue, float thr, be_stream_t stream) {
    BE_REQUIRE(tiles && (out || mask || bits) && wz && wy && wx, BE_ERR_INVALID_ARG, "be_stitch_blend_f32: null pointer");
    BE_REQUIRE((((uintptr_t)window) & 15) == 0, BE_ERR_INVALID_ARG, "be_stitch_blend_f32: window must be 16-byte aligned");
    BE_REQUIRE((((uintptr_t)bits) & 7) == 0, BE_ERR_INVALID_ARG, "be_stitch_blend_f32: bits must be 8-byte aligned");
    BE_REQUIRE((((uintptr_t)wx) & 15) == 0, BE_ERR_INVALID_ARG,
               "be_stitch_blend_f32: wx must be 16-byte aligned (the kernels read the x window two doubles at a time)");
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
    const Epilogue e = make_epilogue(epilogue, thr, out == nullptr);
    const bool aligned16 = ((((uintptr_t)tiles) | ((uintptr_t)out)) & 15) == 0 && (((uintptr_t)mask) & 3) == 0 &&
                           (((uintptr_t)wx) & 7) == 0;
    // geometry of the row-walking kernel: bx threads along x, r_step rows per step, nseg tile rows per (row, ky slot)
    int log2_bx = 5;
    while ((1 << log2_bx) < upr && (1 << log2_bx) < kThreads) ++log2_bx;
    const int r_step = kThreads >> log2_bx;
    const unsigned xblocks = (unsigned)((upr + (1 << log2_bx) - 1) >> log2_bx);
    int nseg = 0;
    for (unsigned b = 0; b < xblocks; ++b) {
        const int64_t X0 = (int64_t)b * (4 << log2_bx), X1 = (X0 + (4 << log2_bx) < X) ? X0 + (4 << log2_bx) : X;
        const int64_t first = (X0 - Bx + 1 <= 0) ? 0 : (X0 - Bx + 1 + Sx - 1) / Sx;
        const int64_t last = ((X1 - 1) / Sx < Tx - 1) ? (X1 - 1) / Sx : Tx - 1;
        if ((int)(last - first + 1) > nseg) nseg = (int)(last - first + 1);
    }
    const size_t smem = (size_t)kStages * r_step * 2 * nseg * Bx * 4 + (size_t)Bx * 8;
    const bool rows_geometry = Bx <= 2 * Sx && By <= 2 * Sy && (Sx & 3) == 0 && (Bx & 3) == 0 && (X & 3) == 0 && aligned16 &&
                               nseg <= 32;
    if (rows_geometry && e.fast && window && Z == 1 && Tz == 1 && Bz == 1 &&
        N * Ty * Tx * st.tile_stride < ((int64_t)1 << 31)) {
        // 2-D, mask only: the fp32 interval filter (see stitch_blend_mask2d_kernel); one wave of blocks walking (plane, band) items
        const int64_t planes = N * C;
        BE_REQUIRE(planes < ((int64_t)1 << 31), BE_ERR_TOO_LARGE, "be_stitch_blend_f32: too many planes");
        int64_t G = (int64_t)BE_SM_COUNT * 8 / xblocks;
        if (G < 1) G = 1;
        int64_t n_bands = (Y + 8 * r_step - 1) / (8 * r_step);
        const int64_t per_block = n_bands * planes / G;
        if (per_block < 16) {
            const int64_t kk = per_block < 1 ? 1 : per_block + 1;
            n_bands = (kk * G + planes - 1) / planes;
        }
        const int64_t max_bands = (Y + r_step - 1) / r_step;
        if (n_bands > max_bands) n_bands = max_bands;
        if (n_bands < 1) n_bands = 1;
        const int64_t n_items = n_bands * planes;
        BE_REQUIRE(n_items < ((int64_t)1 << 31), BE_ERR_TOO_LARGE, "be_stitch_blend_f32: too many work items");
        if (G > n_items) G = n_items;
        if (G > 65535) G = 65535;
        Mask2dConst k;
        k.dky = (int64_t)Tx * st.tile_stride - (int64_t)Sy * Bx;
        const double mid = 0.5 * (e.hi2 + e.lo2), half = 0.5 * (e.hi2 - e.lo2);
        k.mid = (float)mid;
        k.abs_mid = fabsf(k.mid);
        k.half_up = nextafterf((float)(half * (1.0 + 0x1p-19)), INFINITY);
        k.log2_bx = log2_bx;
        k.units_per_row = upr;
        k.n_bands = (unsigned)n_bands;
        k.n_items = (unsigned)n_items;
        k.pieces = 2 * (int)((X + 63) / 64);
        // four blocks of 63 registers per SM, all (up to four) loads of a row in flight at once: measured 42 us for the
        // 4096^2 / 256-px workload against 44 (two loads at a time), 51 (three blocks) and 57 (six blocks, spilling)
        BE_CUDA(be_launch_pdl(stitch_blend_mask2d_kernel, dim3(xblocks, (unsigned)G, 1), dim3(kThreads), 0, (cudaStream_t)stream,
                              tiles, mask, (uint32_t*)bits, window, d, st, e, k));
        BE_LAUNCH_CHECK();
        return BE_OK;
    }
    if (rows_geometry && smem <= 72 * 1024) {
        // one wave of persistent blocks (3 per SM) walks (plane, band) items; bands of about 8 steps, their number
        // rounded so that every block gets the same count when there are only a few items per block
        const int64_t planes = N * C * Z;
        BE_REQUIRE(planes < ((int64_t)1 << 31), BE_ERR_TOO_LARGE, "be_stitch_blend_f32: too many planes");
        int64_t G = (int64_t)BE_SM_COUNT * 3 / xblocks;
        if (G < 1) G = 1;
        int64_t n_bands = (Y + 8 * r_step - 1) / (8 * r_step);
        const int64_t per_block = n_bands * planes / G;
        if (per_block < 16) {
            const int64_t kk = per_block < 1 ? 1 : per_block + 1;
            n_bands = (kk * G + planes - 1) / planes;
        }
        const int64_t max_bands = (Y + r_step - 1) / r_step;
        if (n_bands > max_bands) n_bands = max_bands;
        if (n_bands < 1) n_bands = 1;
        const int64_t n_items = n_bands * planes;
        BE_REQUIRE(n_items < ((int64_t)1 << 31), BE_ERR_TOO_LARGE, "be_stitch_blend_f32: too many work items");
        if (G > n_items) G = n_items;
        if (G > 65535) G = 65535;
        const dim3 g2(xblocks, (unsigned)G, 1);
        RowsConst k;
        k.dky = (int64_t)Tx * st.tile_stride - (int64_t)Sy * Bx;
        k.dkz = (int64_t)Ty * Tx * st.tile_stride - (int64_t)Sz * By * Bx;
        k.dout = (int64_t)r_step * X;
        k.mid = 0.5 * (e.hi2 + e.lo2);
        k.half = 0.5 * (e.hi2 - e.lo2) * (1.0 + 0x1p-30);
        k.log2_bx = log2_bx;
        k.units_per_row = upr;
        k.n_bands = (unsigned)n_bands;
        k.n_items = (unsigned)n_items;
        k.slot_bytes = nseg * Bx * 4;
        k.stage_bytes = r_step * 2 * k.slot_bytes;
        k.n_planes = (unsigned)(N * C * Z);
        if (e.fast)
            BE_CUDA(cudaFuncSetAttribute(stitch_blend_rows_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 72 * 1024));
        else
            BE_CUDA(cudaFuncSetAttribute(stitch_blend_rows_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 72 * 1024));
        if (e.fast)
            stitch_blend_rows_kernel<true><<<g2, kRowsThreads, smem, (cudaStream_t)stream>>>(tiles, out, mask, (uint32_t*)bits, d, st, wz, wy, wx, e, k);
        else
            stitch_blend_rows_kernel<false><<<g2, kRowsThreads, smem, (cudaStream_t)stream>>>(tiles, out, mask, (uint32_t*)bits, d, st, wz, wy, wx, e, k);
    } else {
        if (bits) grid.x = (unsigned)(((X + 63) / 64 * 16 + kThreads - 1) / kThreads);
        stitch_blend_kernel<<<grid, kThreads, 0, (cudaStream_t)stream>>>(tiles, out, mask, (uint32_t*)bits, d, st, wz, wy, wx, e, upr,
                                                                         (unsigned)(N * C * Z));
    }
    BE_LAUNCH_CHECK();
    return BE_OK;
}

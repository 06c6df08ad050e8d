// Connected-component labelling with raster-order numbering (scipy.ndimage.label / skimage.measure.label).
//
// Pipeline (all int/bit work, HBM-bound):
//   1. ccl_tile_kernel     one CTA per 64x64 (2-D) / 64x8x8 (3-D) tile: rows become 64-bit masks by warp
//                          ballot, a pixel's first parent is the start of its run (pure bit arithmetic, so
//                          x-links cost nothing), rows are linked with a shared-memory union-find where only
//                          one lane per touching run pair issues the union; writes parent = global index of
//                          the tile-local root (minimum linear index) to scratch, -1 for background.
//   2. ccl_border_kernel   unions across tile faces in global memory (atomicMin union-find).
//   3. ccl_count / scan / assign   roots (parent == self) are numbered in raster order; a root's scratch
//                          entry becomes -(id) - 1.
//   4. ccl_resolve_kernel  every pixel walks to its root and writes id (background decodes to 0).
#include "be_common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kTX = 64;          // tile width = one 64-bit row mask
constexpr int kRows = 64;        // rows per tile (TY * TZ)
constexpr int kChunk = 2048;     // pixels per block in the numbering passes

struct Geom {
    int64_t Z, Y, X;
    int ty_log2;  // TY = 1 << ty_log2 (6 in 2-D, 3 in 3-D); TZ = kRows >> ty_log2
    int tiles_x, tiles_y, tiles_z;
    int conn;
};

__device__ __forceinline__ int run_start(unsigned long long m, int x) {
    const unsigned long long zeros_below = ~m & ((1ull << x) - 1ull);
    return zeros_below ? 64 - __clzll((long long)zeros_below) : 0;
}

__device__ __forceinline__ unsigned sm_find(volatile unsigned* parent, unsigned i) {
    unsigned p;
    while ((p = parent[i]) != i) i = p;
    return i;
}

__device__ __forceinline__ void sm_unite(unsigned* parent, unsigned a, unsigned b) {
    while (true) {
        a = sm_find(parent, a);
        b = sm_find(parent, b);
        if (a == b) return;
        const unsigned hi = max(a, b), lo = min(a, b);
        const unsigned old = atomicMin(&parent[hi], lo);
        if (old == hi) return;
        a = old;
        b = lo;
    }
}

// Link row `m` (base index bm) with an already-seen neighbour row `u` (base bu).  `full` also links the two
// x-diagonals.  Per pair of touching runs exactly one lane issues the union.
__device__ __forceinline__ void link_rows(unsigned* parent, unsigned long long m, unsigned long long u, unsigned bm,
                                          unsigned bu, bool full, int lane) {
    if (!u) return;
    const unsigned long long both_left = (m << 1) & (u << 1);
    const unsigned long long need_c = m & u & ~both_left;
    const unsigned long long need_l = full ? (m & ~u & (u << 1) & ~(m << 1)) : 0ull;
    const unsigned long long need_r = full ? (m & ~u & (u >> 1) & ~(m >> 1)) : 0ull;
    if (!(need_c | need_l | need_r)) return;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int x = lane + 32 * h;
        const unsigned long long bit = 1ull << x;
        if (need_c & bit) sm_unite(parent, bm + run_start(m, x), bu + run_start(u, x));
        if (need_l & bit) sm_unite(parent, bm + run_start(m, x), bu + run_start(u, x - 1));
        if (need_r & bit) sm_unite(parent, bm + run_start(m, x), bu + run_start(u, x + 1));
    }
}

__global__ void __launch_bounds__(kThreads) ccl_tile_kernel(const uint8_t* __restrict__ mask, int32_t* __restrict__ L,
                                                            Geom g) {
    __shared__ unsigned long long rmask[kRows];
    __shared__ unsigned parent[kRows * kTX];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ty_mask = (1 << g.ty_log2) - 1;
    const int bid = blockIdx.x;
    const int tx = bid % g.tiles_x;
    const int ty = (bid / g.tiles_x) % g.tiles_y;
    const int tz = bid / (g.tiles_x * g.tiles_y);
    const int64_t x0 = (int64_t)tx * kTX, y0 = (int64_t)ty << g.ty_log2, z0 = (int64_t)tz * (kRows >> g.ty_log2);
    const int64_t vol = g.Z * g.Y * g.X;
    const uint8_t* im = mask + (int64_t)blockIdx.y * vol;
    int32_t* Li = L + (int64_t)blockIdx.y * vol;

    for (int r = warp; r < kRows; r += kThreads / 32) {
        const int64_t z = z0 + (r >> g.ty_log2), y = y0 + (r & ty_mask);
        const bool row_ok = (z < g.Z) && (y < g.Y);
        const uint8_t* p = im + (z * g.Y + y) * g.X + x0;
        const bool b0 = row_ok && (x0 + lane < g.X) && p[lane] != 0;
        const bool b1 = row_ok && (x0 + 32 + lane < g.X) && p[32 + lane] != 0;
        const unsigned lo = __ballot_sync(0xffffffffu, b0), hi = __ballot_sync(0xffffffffu, b1);
        if (lane == 0) rmask[r] = ((unsigned long long)hi << 32) | lo;
    }
    for (int i = threadIdx.x; i < kRows * kTX; i += kThreads) parent[i] = i;
    __syncthreads();

    const int TY = 1 << g.ty_log2;
    for (int r = warp; r < kRows; r += kThreads / 32) {
        const unsigned long long m = rmask[r];
        if (!m) continue;
        const int lz = r >> g.ty_log2, ly = r & ty_mask;
        const unsigned bm = r * kTX;
        if (ly > 0) link_rows(parent, m, rmask[r - 1], bm, (r - 1) * kTX, g.conn >= 2, lane);
        if (lz > 0) {
            link_rows(parent, m, rmask[r - TY], bm, (r - TY) * kTX, g.conn >= 2, lane);
            if (g.conn >= 2) {
                if (ly > 0) link_rows(parent, m, rmask[r - TY - 1], bm, (r - TY - 1) * kTX, g.conn >= 3, lane);
                if (ly < TY - 1) link_rows(parent, m, rmask[r - TY + 1], bm, (r - TY + 1) * kTX, g.conn >= 3, lane);
            }
        }
    }
    __syncthreads();

    for (int r = warp; r < kRows; r += kThreads / 32) {
        const int64_t z = z0 + (r >> g.ty_log2), y = y0 + (r & ty_mask);
        if (z >= g.Z || y >= g.Y) continue;
        const unsigned long long m = rmask[r];
        int32_t* row = Li + (z * g.Y + y) * g.X + x0;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int x = lane + 32 * h;
            if (x0 + x >= g.X) continue;
            int32_t v = -1;
            if ((m >> x) & 1ull) {
                const unsigned start = r * kTX + run_start(m, x);
                const unsigned root = sm_find(parent, start);
                const int rr = root >> 6, rx = root & 63;
                v = (int32_t)(((z0 + (rr >> g.ty_log2)) * g.Y + (y0 + (rr & ty_mask))) * g.X + x0 + rx);
            }
            row[x] = v;
        }
    }
}

__device__ __forceinline__ int32_t gl_find(volatile int32_t* L, int32_t i) {
    int32_t p;
    while ((p = L[i]) != i) i = p;
    return i;
}

__device__ __forceinline__ void gl_unite(int32_t* L, int32_t a, int32_t b) {
    while (true) {
        a = gl_find(L, a);
        b = gl_find(L, b);
        if (a == b) return;
        const int32_t hi = max(a, b), lo = min(a, b);
        const int32_t old = atomicMin(&L[hi], lo);
        if (old == hi) return;
        a = old;
        b = lo;
    }
}

// One thread per voxel; only voxels on a tile face do any memory access.
__global__ void __launch_bounds__(kThreads) ccl_border_kernel(const uint8_t* __restrict__ mask, int32_t* __restrict__ L,
                                                              Geom g) {
    const int64_t vol = g.Z * g.Y * g.X;
    const int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    if (i >= vol) return;
    const int x = (int)(i % g.X);
    const int64_t t = i / g.X;
    const int y = (int)(t % g.Y);
    const int z = (int)(t / g.Y);
    const int TY = 1 << g.ty_log2, TZ = kRows >> g.ty_log2;
    const int lx = x & (kTX - 1), ly = y & (TY - 1), lz = z & (TZ - 1);
    const bool f_xlo = lx == 0 && x > 0, f_xhi = lx == kTX - 1 && x + 1 < g.X;
    const bool f_ylo = ly == 0 && y > 0, f_yhi = ly == TY - 1 && y + 1 < g.Y;
    const bool f_zlo = lz == 0 && z > 0;
    if (!(f_xlo | f_xhi | f_ylo | f_yhi | f_zlo)) return;
    const uint8_t* im = mask + (int64_t)blockIdx.y * vol;
    int32_t* Li = L + (int64_t)blockIdx.y * vol;
    if (!im[i]) return;
    // the 13 raster-order predecessors
    for (int dz = -1; dz <= 0; ++dz) {
        if (dz < 0 && z == 0) continue;
        for (int dy = -1; dy <= 1; ++dy) {
            if (dz == 0 && dy > 0) break;
            const int qy = y + dy;
            if (qy < 0 || qy >= g.Y) continue;
            for (int dx = -1; dx <= 1; ++dx) {
                if (dz == 0 && dy == 0 && dx >= 0) break;
                const int qx = x + dx;
                if (qx < 0 || qx >= g.X) continue;
                if ((dz != 0) + (dy != 0) + (dx != 0) > g.conn) continue;
                const bool crosses = (dx < 0 && lx == 0) || (dx > 0 && lx == kTX - 1) || (dy < 0 && ly == 0) ||
                                     (dy > 0 && ly == TY - 1) || (dz < 0 && lz == 0);
                if (!crosses) continue;
                const int64_t q = ((int64_t)(z + dz) * g.Y + qy) * g.X + qx;
                if (im[q]) gl_unite(Li, (int32_t)i, (int32_t)q);
            }
        }
    }
}

// ---- raster-order numbering of the roots ----

__global__ void __launch_bounds__(kThreads) ccl_count_kernel(const int32_t* __restrict__ L, int64_t vol,
                                                             int blocks_per_image, int32_t* __restrict__ block_count) {
    const int64_t base = (int64_t)blockIdx.x * kChunk;
    const int32_t* Li = L + (int64_t)blockIdx.y * vol;
    int c = 0;
#pragma unroll
    for (int k = 0; k < kChunk / kThreads; ++k) {
        const int64_t i = base + k * kThreads + threadIdx.x;
        if (i < vol) c += (Li[i] == (int32_t)i);
    }
    c = __reduce_add_sync(0xffffffffu, c);
    __shared__ int s[kThreads / 32];
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int w = 0; w < kThreads / 32; ++w) t += s[w];
        block_count[(int64_t)blockIdx.y * blocks_per_image + blockIdx.x] = t;
    }
}

// one block per image: in-place exclusive scan of its block counts, total -> n_labels
__global__ void __launch_bounds__(1024) ccl_scan_kernel(int32_t* __restrict__ block_count, int blocks_per_image,
                                                        int32_t* __restrict__ n_labels) {
    __shared__ int s_warp[32];
    __shared__ int s_carry;
    int32_t* bc = block_count + (int64_t)blockIdx.x * blocks_per_image;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    for (int base = 0; base < blocks_per_image; base += 1024) {
        const int i = base + threadIdx.x;
        const int v = (i < blocks_per_image) ? bc[i] : 0;
        int incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if ((threadIdx.x & 31) >= o) incl += t;
        }
        if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = incl;
        __syncthreads();
        int wbase = 0;
        for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) wbase += s_warp[w];
        const int carry = s_carry;
        if (i < blocks_per_image) bc[i] = carry + wbase + incl - v;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = carry + wbase + incl;
        __syncthreads();
    }
    if (threadIdx.x == 0) n_labels[blockIdx.x] = s_carry;
}

__global__ void __launch_bounds__(kThreads) ccl_assign_kernel(int32_t* __restrict__ L, int64_t vol, int blocks_per_image,
                                                              const int32_t* __restrict__ block_offset) {
    constexpr int per = kChunk / kThreads;
    const int64_t base = (int64_t)blockIdx.x * kChunk + (int64_t)threadIdx.x * per;
    int32_t* Li = L + (int64_t)blockIdx.y * vol;
    bool is_root[per];
    int c = 0;
#pragma unroll
    for (int k = 0; k < per; ++k) {
        const int64_t i = base + k;
        is_root[k] = (i < vol) && (Li[i] == (int32_t)i);
        c += is_root[k];
    }
    int incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if ((threadIdx.x & 31) >= o) incl += t;
    }
    __shared__ int s_warp[kThreads / 32];
    if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = incl;
    __syncthreads();
    int id = block_offset[(int64_t)blockIdx.y * blocks_per_image + blockIdx.x] + incl - c;
    for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) id += s_warp[w];
#pragma unroll
    for (int k = 0; k < per; ++k)
        if (is_root[k]) Li[base + k] = -(++id) - 1;
}

__device__ __forceinline__ int32_t resolve_one(const int32_t* __restrict__ Li, int32_t v) {
    while (v >= 0) v = Li[v];
    return -v - 1;
}

__global__ void __launch_bounds__(kThreads) ccl_resolve_kernel(const int32_t* __restrict__ L, int32_t* __restrict__ out,
                                                               int64_t vol) {
    const int32_t* Li = L + (int64_t)blockIdx.y * vol;
    int32_t* oi = out + (int64_t)blockIdx.y * vol;
    const int64_t i4 = ((int64_t)blockIdx.x * kThreads + threadIdx.x) * 4;
    if (i4 >= vol) return;
    if (i4 + 4 <= vol && (((uintptr_t)(Li + i4) | (uintptr_t)(oi + i4)) & 15) == 0) {
        int4 v = *reinterpret_cast<const int4*>(Li + i4);
        v.x = resolve_one(Li, v.x);
        v.y = resolve_one(Li, v.y);
        v.z = resolve_one(Li, v.z);
        v.w = resolve_one(Li, v.w);
        *reinterpret_cast<int4*>(oi + i4) = v;
    } else {
        for (int64_t i = i4; i < min(i4 + 4, vol); ++i) oi[i] = resolve_one(Li, Li[i]);
    }
}

inline int blocks_per_image(int64_t vol) { return (int)((vol + kChunk - 1) / kChunk); }

}  // namespace

extern "C" size_t be_ccl_workspace_bytes(int64_t N, int64_t Z, int64_t Y, int64_t X) {
    if (N <= 0 || Z <= 0 || Y <= 0 || X <= 0) return 256;
    const int64_t vol = Z * Y * X;
    return be_align_up(sizeof(int32_t) * N * vol, 256) + be_align_up(sizeof(int32_t) * N * blocks_per_image(vol), 256);
}

extern "C" int be_ccl_u8_i32(const uint8_t* mask, int32_t* labels, int64_t N, int64_t Z, int64_t Y, int64_t X,
                             int connectivity, int32_t* n_labels, void* ws, be_stream_t stream) {
    BE_REQUIRE(mask && labels && n_labels && ws, BE_ERR_INVALID_ARG, "be_ccl_u8_i32: null pointer");
    BE_REQUIRE(N > 0 && Z > 0 && Y > 0 && X > 0, BE_ERR_INVALID_ARG, "be_ccl_u8_i32: empty input (%lld,%lld,%lld,%lld)",
               (long long)N, (long long)Z, (long long)Y, (long long)X);
    const int ndim = (Z > 1) ? 3 : 2;
    BE_REQUIRE(connectivity >= 1 && connectivity <= ndim, BE_ERR_INVALID_ARG,
               "be_ccl_u8_i32: connectivity %d not in 1..%d", connectivity, ndim);
    const int64_t vol = Z * Y * X;
    BE_REQUIRE(vol < ((int64_t)1 << 31) - 2, BE_ERR_TOO_LARGE, "be_ccl_u8_i32: %lld voxels per image >= 2^31", (long long)vol);
    BE_REQUIRE(N <= 65535, BE_ERR_TOO_LARGE, "be_ccl_u8_i32: N %lld > 65535", (long long)N);
    Geom g;
    g.Z = Z;
    g.Y = Y;
    g.X = X;
    g.ty_log2 = (ndim == 2) ? 6 : 3;
    const int TY = 1 << g.ty_log2, TZ = kRows >> g.ty_log2;
    g.tiles_x = (int)((X + kTX - 1) / kTX);
    g.tiles_y = (int)((Y + TY - 1) / TY);
    g.tiles_z = (int)((Z + TZ - 1) / TZ);
    g.conn = connectivity;
    const int64_t n_tiles = (int64_t)g.tiles_x * g.tiles_y * g.tiles_z;
    BE_REQUIRE(n_tiles < ((int64_t)1 << 31) - 1, BE_ERR_TOO_LARGE, "be_ccl_u8_i32: too many tiles");
    cudaStream_t s = (cudaStream_t)stream;
    int32_t* L = (int32_t*)ws;
    int32_t* block_count = (int32_t*)((char*)ws + be_align_up(sizeof(int32_t) * N * vol, 256));
    const int bpi = blocks_per_image(vol);

    ccl_tile_kernel<<<dim3((unsigned)n_tiles, (unsigned)N), kThreads, 0, s>>>(mask, L, g);
    BE_LAUNCH_CHECK();
    ccl_border_kernel<<<dim3((unsigned)((vol + kThreads - 1) / kThreads), (unsigned)N), kThreads, 0, s>>>(mask, L, g);
    BE_LAUNCH_CHECK();
    ccl_count_kernel<<<dim3(bpi, (unsigned)N), kThreads, 0, s>>>(L, vol, bpi, block_count);
    BE_LAUNCH_CHECK();
    ccl_scan_kernel<<<(unsigned)N, 1024, 0, s>>>(block_count, bpi, n_labels);
    BE_LAUNCH_CHECK();
    ccl_assign_kernel<<<dim3(bpi, (unsigned)N), kThreads, 0, s>>>(L, vol, bpi, block_count);
    BE_LAUNCH_CHECK();
    ccl_resolve_kernel<<<dim3((unsigned)((vol + kThreads * 4 - 1) / (kThreads * 4)), (unsigned)N), kThreads, 0, s>>>(
        L, labels, vol);
    BE_LAUNCH_CHECK();
    return BE_OK;
}

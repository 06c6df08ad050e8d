// Connected-component labelling with raster-order numbering (scipy.ndimage.label / skimage.measure.label).
//
// Four launches, chained by programmatic dependent launch (all int/bit work, HBM-bound):
//   1. ccl_tile_kernel<IN>  one CTA per 64x64 (2-D) / 64x8x8 (3-D) tile.  Rows become 64-bit masks — from a uint8 mask
//                           (16-byte loads + byte->bit multiply trick, or warp ballots at ragged edges), from a
//                           bit-packed mask (one word per row: what the blend stitcher can emit), or straight from fp32
//                           logits with the sigmoid / threshold decided here.  The union-find nodes are the
//                           RUNS of a row (<= 32 per row, 2048 per tile, parents in shared memory): x-links cost
//                           nothing, and rows are linked with bit arithmetic that leaves exactly one lane per touching
//                           run pair to issue the union (atomicMin, path halving).  After a per-run flatten every run
//                           takes the label of its root (global index of the component's first pixel in the tile) and
//                           every foreground pixel gets that as its parent, branch-free, 16 bytes per lane; units
//                           without a foreground pixel are not written at all.  Tile roots are flagged in a
//                           1-bit-per-pixel bitmap, the mask itself is kept bit-packed, and the tile's four edge
//                           masks are kept for step 2.
//   2. ccl_edge_kernel      (2-D) four warps per tile link its top and left edge to the neighbouring tiles from the edge
//                           masks alone, one union per touching run pair, a lane's (up to 3) unions issued as a batch;
//                           ccl_face3d_kernel (3-D) does the same per 64-voxel row word of the bit-packed mask, four
//                           threads per word (one per predecessor row).  Unions are atomicMin on the global parent array
//                           with path halving.
//   3. ccl_number_kernel    surviving roots are numbered in raster order from the root bitmap (2 MB instead of a
//                           second pass over all parents) with a single-pass decoupled look-back scan (blocks take their
//                           256 words by ticket; 128 predecessors per step); a root's entry becomes -(id) - 1.
//   4. ccl_resolve_kernel   driven by the packed mask: a foreground pixel walks to its root (<= 2-3 hops) and writes
//                           the id, background is written as 0 without touching the parent array.
// (Tried and dropped, with numbers in profiles/README.md: border unions inside the tile launch — by the last neighbour to
// finish, or by ticketed edge blocks behind the tiles — numbering + resolving in one ticketed launch, and flattening the
// merged tile roots while numbering so that the resolve needs no walk.  The fences and flags hold a CTA's slot for a
// memory round trip per tile, blocks that spin for a number crowd out the blocks that produce it, and the flattening
// costs the numbering more than it saves the resolve; four plain launches, chained so that launch latency is hidden,
// are faster.)
#include "be_common.cuh"

namespace {

typedef unsigned long long u64;

constexpr int kThreads = 256;
constexpr int kTX = 64;            // tile width = one 64-bit row mask
constexpr int kRows = 64;          // rows per tile (TY * TZ)
constexpr int kMaxRuns = kRows * 32;
constexpr int kWordsPerBlock = 256;  // bitmap words per block in the numbering passes

struct Geom {
    int64_t Z, Y, X;
    int ty_log2;  // TY = 1 << ty_log2 (6 in 2-D, 3 in 3-D); TZ = kRows >> ty_log2
    int tiles_x, tiles_y, tiles_z;
    int conn;
    unsigned* scan_zero;   // tickets + look-back states of the numbering: cleared by the tile kernel's blocks, a slice each
    int scan_words;
};

struct TileEdges {
    u64 top, bottom, left, right;
};

// index (within its row) of the run containing set bit x: number of run starts at positions <= x, minus one
__device__ __forceinline__ int run_index(u64 starts, int x) { return __popcll(starts & ((2ull << x) - 1ull)) - 1; }

__device__ __forceinline__ unsigned sm_find(volatile unsigned* parent, unsigned i) {
    while (true) {
        const unsigned p = parent[i];
        if (p == i) return i;
        const unsigned gp = parent[p];
        if (gp != p) parent[i] = gp;  // path halving: any ancestor is a valid parent
        i = gp;
    }
}

// read-only walk, for the flatten phase: there the forest is final and each thread overwrites its own entry with the
// root, which a concurrent path-halving store could replace by a mere ancestor
__device__ __forceinline__ unsigned sm_find_ro(const volatile unsigned* parent, unsigned i) {
    unsigned p;
    while ((p = parent[i]) != i) i = p;
    return i;
}

// returns the surviving (smaller) root
__device__ __forceinline__ unsigned sm_unite(unsigned* parent, unsigned a, unsigned b) {
    while (true) {
        a = sm_find(parent, a);
        b = sm_find(parent, b);
        if (a == b) return a;
        const unsigned hi = max(a, b), lo = min(a, b);
        const unsigned old = atomicMin(&parent[hi], lo);
        if (old == hi) return lo;
        a = old;
        b = lo;
    }
}

// Which lanes must issue a union between row m and an already-seen neighbour row u.  One lane per pair of touching
// runs: the straight link is dropped when the left neighbours are linked too, a diagonal link is dropped when the
// straight neighbour or the run's own continuation makes it redundant.
struct LinkMasks {
    u64 c, l, r;
};
__device__ __forceinline__ LinkMasks link_masks(u64 m, u64 u, bool diag) {
    LinkMasks k;
    k.c = m & u & ~((m << 1) & (u << 1));
    k.l = diag ? (m & ~u & (u << 1) & ~(m << 1)) : 0ull;
    k.r = diag ? (m & ~u & (u >> 1) & ~(m >> 1)) : 0ull;
    return k;
}

__device__ __forceinline__ unsigned nibble_of(unsigned w) {
    // one bit per non-zero byte, lowest address -> lowest bit
    return (((__vcmpne4(w, 0u) & 0x01010101u) * 0x01020408u) >> 24) & 0xfu;
}

// Link run `id` = [x, x+len) of one row with every run of neighbour row (mask u, starts su, id base bu) it touches.
// Touched runs are consecutive in index, so only the first needs a popcount.
__device__ __forceinline__ void link_run(unsigned* parent, unsigned id, int x, int len, u64 u, u64 su, unsigned bu,
                                         bool diag) {
    int lo = x, hi = x + len - 1;
    if (diag) {
        lo = max(lo - 1, 0);
        hi = min(hi + 1, 63);
    }
    const u64 window = ((2ull << hi) - 1ull) & ~((1ull << lo) - 1ull);
    const u64 t = u & window;
    if (!t) return;
    const int f = __ffsll((long long)t) - 1;
    unsigned j = bu + run_index(su, f);
    sm_unite(parent, id, j);
    u64 more = su & window & ~((2ull << f) - 1ull);  // further touched runs start strictly after f inside the window
    while (more) {
        more &= more - 1;
        sm_unite(parent, id, ++j);
    }
}

// what the tile kernel reads its rows from
constexpr int kInU8 = 0, kInBits = 1, kInF32 = 2;

template <int IN>
__global__ void __launch_bounds__(kThreads, 8) ccl_tile_kernel(const void* __restrict__ input, Epilogue epi,
                                                            int32_t* __restrict__ L, u64* __restrict__ rootbits,
                                                            u64* __restrict__ bitvol, TileEdges* __restrict__ edges,
                                                            Geom g) {
    __shared__ u64 rmask[kRows];
    __shared__ u64 rstart[kRows];
    __shared__ __align__(16) unsigned parent[kMaxRuns];   // union-find over run ids (row * 32 + index in row)
    __shared__ int32_t glab[kMaxRuns];                    // root run -> global index of its first pixel
    __shared__ unsigned char xstart[kMaxRuns];            // first pixel (x within the tile) of every run
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ty_mask = (1 << g.ty_log2) - 1;
    // grid: x = tile column, y = tile row, z = image * tiles_z + tile plane
    const int tx = blockIdx.x, ty = blockIdx.y;
    const int img = blockIdx.z / g.tiles_z, tz = blockIdx.z - img * g.tiles_z;
    const int bid = (tz * g.tiles_y + ty) * g.tiles_x + tx;
    const int64_t x0 = (int64_t)tx * kTX, y0 = (int64_t)ty << g.ty_log2, z0 = (int64_t)tz * (kRows >> g.ty_log2);
    const int64_t vol = g.Z * g.Y * g.X;
    int32_t* Li = L + (int64_t)img * vol;
    __shared__ u64 rroot[kRows];
    if (threadIdx.x < kRows) rroot[threadIdx.x] = 0;
    be_pdl_wait();   // the kernel that wrote the mask (and, before it, the last user of the numbering's scan states)
    {
        const int n_blocks = gridDim.x * gridDim.y * gridDim.z;
        const int lin = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        const int per = (g.scan_words + n_blocks - 1) / n_blocks;
        for (int i = lin * per + threadIdx.x; i < min((lin + 1) * per, g.scan_words); i += kThreads) g.scan_zero[i] = 0u;
    }

    // ---- A: rows -> bit masks ----
    if (IN == kInBits) {
        // one word per row, laid out like bitvol: [image][z*Y + y][tile column]
        if (threadIdx.x < kRows) {
            const int r = threadIdx.x;
            const int64_t z = z0 + (r >> g.ty_log2), y = y0 + (r & ty_mask);
            u64 v = 0;
            if (z < g.Z && y < g.Y) {
                v = __ldg(reinterpret_cast<const u64*>(input) + ((int64_t)img * g.Z * g.Y + z * g.Y + y) * g.tiles_x + tx);
                if (x0 + kTX > g.X) v &= (1ull << (g.X - x0)) - 1ull;   // bits past the row end do not count
            }
            rmask[r] = v;
            rstart[r] = v & ~(v << 1);
        }
    } else if (IN == kInF32) {
        // 16 lanes x 4 logits per row, two rows per warp instruction; the mask bit is decided here (sigmoid / threshold)
        const float* im = reinterpret_cast<const float*>(input) + (int64_t)img * vol;
        const int lane16 = lane & 15, sub = lane >> 4;
        const bool vec_ok = ((g.X & 3) == 0) && ((((uintptr_t)im) & 15) == 0);
#pragma unroll
        for (int it = 0; it < 4; ++it) {
            const int r = warp * 8 + it * 2 + sub;
            const int64_t z = z0 + (r >> g.ty_log2), y = y0 + (r & ty_mask);
            const int64_t x = x0 + lane16 * 4;
            unsigned m4 = 0;
            if (z < g.Z && y < g.Y && x < g.X) {
                const float* src = im + (z * g.Y + y) * g.X + x;
                if (vec_ok && x + 4 <= g.X) {
                    const float4 q = ld_stream_f4(reinterpret_cast<const float4*>(src));
                    m4 = (unsigned)be_epilogue_mask(q.x, epi) | ((unsigned)be_epilogue_mask(q.y, epi) << 1) |
                         ((unsigned)be_epilogue_mask(q.z, epi) << 2) | ((unsigned)be_epilogue_mask(q.w, epi) << 3);
                } else {
                    for (int j = 0; j < 4 && x + j < g.X; ++j) m4 |= (unsigned)be_epilogue_mask(src[j], epi) << j;
                }
            }
            u64 v = (u64)m4 << (4 * lane16);
#pragma unroll
            for (int o = 1; o < 16; o <<= 1) v |= __shfl_xor_sync(0xffffffffu, v, o);
            if (lane16 == 0) {
                rmask[r] = v;
                rstart[r] = v & ~(v << 1);
            }
        }
    } else if ((g.X & 15) == 0 && x0 + kTX <= g.X &&
               (((uintptr_t)(reinterpret_cast<const uint8_t*>(input) + (int64_t)img * vol)) & 15) == 0) {
        const uint8_t* im = reinterpret_cast<const uint8_t*>(input) + (int64_t)img * vol;
        // 4 lanes x 16 bytes per row, 8 rows per warp: one load instruction per warp
        const int r = warp * 8 + (lane >> 2), part = lane & 3;
        const int64_t z = z0 + (r >> g.ty_log2), y = y0 + (r & ty_mask);
        u64 v = 0;
        if (z < g.Z && y < g.Y) {
            const uint4 q = __ldg(reinterpret_cast<const uint4*>(im + (z * g.Y + y) * g.X + x0) + part);
            const unsigned bits = nibble_of(q.x) | (nibble_of(q.y) << 4) | (nibble_of(q.z) << 8) | (nibble_of(q.w) << 12);
            v = (u64)bits << (16 * part);
        }
        v |= __shfl_xor_sync(0xffffffffu, v, 1);
        v |= __shfl_xor_sync(0xffffffffu, v, 2);
        if (part == 0) {
            rmask[r] = v;
            rstart[r] = v & ~(v << 1);
        }
    } else {
        const uint8_t* im = reinterpret_cast<const uint8_t*>(input) + (int64_t)img * vol;
        for (int r = warp; r < kRows; r += kThreads / 32) {
            const int64_t z = z0 + (r >> g.ty_log2), y = y0 + (r & ty_mask);
            const bool row_ok = (z < g.Z) && (y < g.Y);
            const uint8_t* p = im + (z * g.Y + y) * g.X + x0;
            const bool b0 = row_ok && (x0 + lane < g.X) && p[lane] != 0;
            const bool b1 = row_ok && (x0 + 32 + lane < g.X) && p[32 + lane] != 0;
            const unsigned lo = __ballot_sync(0xffffffffu, b0), hi = __ballot_sync(0xffffffffu, b1);
            if (lane == 0) {
                const u64 v = ((u64)hi << 32) | lo;
                rmask[r] = v;
                rstart[r] = v & ~(v << 1);
            }
        }
    }
    __syncthreads();
    // union-find nodes: only the runs that exist (a row's share of the 2048 slots is mostly unused)
    {
        const int r = threadIdx.x >> 2, n_runs = __popcll(rstart[r]);
        for (int k = threadIdx.x & 3; k < n_runs; k += 4) parent[r * 32 + k] = r * 32 + k;
    }
    __syncthreads();

    // ---- B: every run is linked to the runs it touches in the rows before it.  Four threads share a row and take
    //         every fourth run: the loop over a row's runs is sequential per thread, but all 256 threads (64 rows)
    //         run it side by side, so the instruction stream is spent on runs, not on idle lanes. ----
    {
        const int TY = 1 << g.ty_log2;
        const int r = threadIdx.x >> 2, part = threadIdx.x & 3;
        const u64 m = rmask[r];
        u64 s = rstart[r];
        const int lz = r >> g.ty_log2, ly = r & ty_mask;
        const bool d1 = g.conn >= 2, d2 = g.conn >= 3;
        int k = part;
        for (int i = 0; i < part && s; ++i) s &= s - 1;  // skip to this thread's first run
        while (s) {
            const int x = __ffsll((long long)s) - 1;
            const unsigned id = r * 32 + k;
            xstart[id] = (unsigned char)x;
            const u64 rest = ~(m >> x);  // first zero above the run start ends the run
            const int len = rest ? __ffsll((long long)rest) - 1 : 64;
            if (ly > 0) link_run(parent, id, x, len, rmask[r - 1], rstart[r - 1], (r - 1) * 32, d1);
            if (lz > 0) {
                link_run(parent, id, x, len, rmask[r - TY], rstart[r - TY], (r - TY) * 32, d1);
                if (g.conn >= 2) {
                    if (ly > 0) link_run(parent, id, x, len, rmask[r - TY - 1], rstart[r - TY - 1], (r - TY - 1) * 32, d2);
                    if (ly < TY - 1)
                        link_run(parent, id, x, len, rmask[r - TY + 1], rstart[r - TY + 1], (r - TY + 1) * 32, d2);
                }
            }
            // on to run k + 4 of this row
            s &= s - 1;
            s &= s - 1;
            s &= s - 1;
            s &= s - 1;
            k += 4;
        }
    }
    __syncthreads();

    // ---- C: flatten per run (same walk: 4 threads per row); a root is the component's first run in the tile and gets
    //         the global index of its first pixel ----
    {
        const int r = threadIdx.x >> 2, part = threadIdx.x & 3;
        const int n_runs = __popcll(rstart[r]);
        const int64_t row_base = ((z0 + (r >> g.ty_log2)) * g.Y + (y0 + (r & ty_mask))) * g.X + x0;
        for (int k = part; k < n_runs; k += 4) {
            const unsigned id = r * 32 + k;
            const unsigned root = sm_find_ro(parent, id);
            parent[id] = root;
            if (root == id) glab[id] = (int32_t)(row_base + xstart[id]);
        }
    }
    __syncthreads();
    // ---- C2 / D: every run writes its root's label over its pixels (the per-pixel parents: they exist only where the
    //          mask has pixels, nobody reads the rest); a root run flags its first pixel in the root bitmap.  Done per run,
    //          not per 4-pixel unit: a speckled mask has fewer runs than units with a set pixel, and a run needs no
    //          bit-twiddling to find out whose label it carries ----
    {
        const int r = threadIdx.x >> 2, part = threadIdx.x & 3;
        const u64 m = rmask[r];
        const int n_runs = __popcll(rstart[r]);
        const int64_t row_base = ((z0 + (r >> g.ty_log2)) * g.Y + (y0 + (r & ty_mask))) * g.X + x0;
        int32_t* row = Li + row_base;
        const bool vec_ok = ((g.X & 3) == 0) && ((((uintptr_t)Li) & 15) == 0);   // x0 is a multiple of 64
        u64 roots = 0;
        for (int k = part; k < n_runs; k += 4) {
            const unsigned id = r * 32 + k;
            const unsigned root = parent[id];
            const int x = xstart[id];
            if (root == id) roots |= 1ull << x;
            const int32_t lab = glab[root];
            const u64 rest = ~(m >> x);
            const int len = rest ? __ffsll((long long)rest) - 1 : 64;
            if (len < 8 || !vec_ok) {
                for (int j = 0; j < len; ++j) row[x + j] = lab;
            } else {   // long runs (blob-like masks): 16-byte stores between an unaligned head and tail
                int j = x;
                const int end = x + len;
                for (; j & 3; ++j) row[j] = lab;
                const int4 v = make_int4(lab, lab, lab, lab);
                for (; j + 4 <= end; j += 4) *reinterpret_cast<int4*>(row + j) = v;
                for (; j < end; ++j) row[j] = lab;
            }
        }
        if (roots) atomicOr(&rroot[r], roots);
    }
    __syncthreads();

    // ---- bitmaps out: the root bitmap and, unless the input can be used in place, the packed mask ----
    {
        const int64_t wpr = g.tiles_x;
        if (lane < 8) {
            const int r = warp * 8 + lane;
            const int64_t z = z0 + (r >> g.ty_log2), y = y0 + (r & ty_mask);
            if (z < g.Z && y < g.Y) {
                const int64_t wi = ((int64_t)img * g.Z * g.Y + z * g.Y + y) * wpr + tx;
                rootbits[wi] = rroot[r];
                if (IN != kInBits || (g.X & (kTX - 1))) bitvol[wi] = rmask[r];  // the mask, 1 bit per voxel (3-D face links, the resolve pass); a bit-packed input with whole-word rows is used in place, one with ragged rows is copied with the tails cleaned
            }
        }
    }
    if (warp == 0 && edges) {
        const u64 a = rmask[lane], b = rmask[lane + 32];
        const unsigned l_lo = __ballot_sync(0xffffffffu, a & 1ull), l_hi = __ballot_sync(0xffffffffu, b & 1ull);
        const unsigned r_lo = __ballot_sync(0xffffffffu, (a >> 63) & 1ull), r_hi = __ballot_sync(0xffffffffu, (b >> 63) & 1ull);
        if (lane == 0) {
            TileEdges e;
            e.top = rmask[0];
            e.bottom = rmask[kRows - 1];
            e.left = ((u64)l_hi << 32) | l_lo;
            e.right = ((u64)r_hi << 32) | r_lo;
            edges[(int64_t)img * (g.tiles_x * g.tiles_y * g.tiles_z) + bid] = e;
        }
    }
    be_pdl_launch_dependents();
}

// ---- global union-find on the parent array ----

__device__ __forceinline__ int32_t gl_find(volatile int32_t* L, int32_t i) {
    while (true) {
        const int32_t p = L[i];
        if (p == i) return i;
        const int32_t gp = L[p];
        if (gp != p) L[i] = gp;
        i = gp;
    }
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

// 2-D: top edge against the tile above, left edge against the tile to the left, two corners.  A lane's unions are
// independent, so they run as a batch — all parent loads issued before any is used, then all atomics — and cost about
// three memory round trips in total instead of five per union.  pm / pu give the pixel index of bit i on the two sides
// of an edge: base + i * step.
constexpr int kEdgePairs = 3;

struct EdgeBatch {
    int32_t a[kEdgePairs], b[kEdgePairs];
    unsigned valid;
};

// the (up to three) unions of bit i of edge mask m against the neighbour's edge mask u
__device__ __forceinline__ void edge_pairs(EdgeBatch& q, u64 m, u64 u, int64_t base_m, int64_t base_u, int64_t step,
                                           bool diag, int i) {
    if (!m || !u) return;
    const LinkMasks k = link_masks(m, u, diag);
    const u64 bit = 1ull << i;
    q.a[0] = q.a[1] = q.a[2] = (int32_t)(base_m + i * step);
    q.b[0] = (int32_t)(base_u + i * step);
    q.b[1] = (int32_t)(base_u + (i - 1) * step);
    q.b[2] = (int32_t)(base_u + (i + 1) * step);
    q.valid = ((k.c & bit) ? 1u : 0u) | ((k.l & bit) ? 2u : 0u) | ((k.r & bit) ? 4u : 0u);
}

__device__ __forceinline__ void unite_batch(int32_t* Li, EdgeBatch& q) {
    if (!q.valid) return;
    // level 1: an edge pixel's parent is its tile root (or something even closer to the final root)
#pragma unroll
    for (int k = 0; k < kEdgePairs; ++k)
        if ((q.valid >> k) & 1u) {
            q.a[k] = __ldcg(Li + q.a[k]);
            q.b[k] = __ldcg(Li + q.b[k]);
        }
    // both are roots as far as this lane knows: one atomicMin per pair, all issued before any result is looked at.  If
    // the larger one had been linked elsewhere in the meantime the atomic says so (old != hi) and the pair goes through
    // the general walk with that link preserved — so there is no need to re-read the roots first
    unsigned fast = 0;
    int32_t old[kEdgePairs];
#pragma unroll
    for (int k = 0; k < kEdgePairs; ++k)
        if ((q.valid >> k) & 1u) {
            if (q.a[k] != q.b[k]) {
                const int32_t hi = max(q.a[k], q.b[k]), lo = min(q.a[k], q.b[k]);
                old[k] = atomicMin(&Li[hi], lo);
                q.a[k] = hi;
                q.b[k] = lo;
                fast |= 1u << k;
            }
        }
#pragma unroll
    for (int k = 0; k < kEdgePairs; ++k) {
        if ((fast >> k) & 1u) {
            if (old[k] != q.a[k]) gl_unite(Li, old[k], q.b[k]);   // hi was no longer a root: keep its old link alive
        }
    }
}

// Four warps per tile: (top | left edge) x (bits 0..31 | 32..63); a lane owns one edge pixel and its at most three unions.
// A lane's unions are a chain of dependent memory round trips and the kernel ends when the longest chain does, so the
// work is spread over as many lanes as there are edge pixels.
__global__ void __launch_bounds__(kThreads) ccl_edge_kernel(int32_t* __restrict__ L, const TileEdges* __restrict__ edges,
                                                            Geom g, int tiles_per_image) {
    be_pdl_wait();   // the tile kernel's parents and edge masks
    be_pdl_launch_dependents();   // the numbering kernel may be placed now: it reads the root bitmap (complete) and waits
    const int wg = blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5);
    const int t = wg >> 2, part = wg & 3;
    if (t >= tiles_per_image) return;
    const int lane = threadIdx.x & 31;
    const int i = lane + 32 * (part & 1);
    const bool left = part >> 1;
    const int tx = t % g.tiles_x, ty = t / g.tiles_x;
    const TileEdges* E = edges + (int64_t)blockIdx.y * tiles_per_image;
    int32_t* Li = L + (int64_t)blockIdx.y * g.Y * g.X;
    const TileEdges me = E[t];
    const int64_t x0 = (int64_t)tx * kTX, y0 = (int64_t)ty * kRows;
    const bool diag = g.conn >= 2;
    EdgeBatch q;
    q.valid = 0;
#pragma unroll
    for (int k = 0; k < kEdgePairs; ++k) q.a[k] = q.b[k] = 0;
    if (!left && ty > 0) edge_pairs(q, me.top, E[t - g.tiles_x].bottom, y0 * g.X + x0, (y0 - 1) * g.X + x0, 1, diag, i);
    if (left && tx > 0) edge_pairs(q, me.left, E[t - 1].right, y0 * g.X + x0, y0 * g.X + x0 - 1, g.X, diag, i);
    unite_batch(Li, q);
    if (diag && ty > 0 && part == 0 && lane == 0) {
        if (tx > 0 && (me.top & 1ull) && (E[t - g.tiles_x - 1].bottom >> 63))
            gl_unite(Li, (int32_t)(y0 * g.X + x0), (int32_t)((y0 - 1) * g.X + x0 - 1));
        if (tx + 1 < g.tiles_x && (me.top >> 63) && (E[t - g.tiles_x + 1].bottom & 1ull))
            gl_unite(Li, (int32_t)(y0 * g.X + x0 + 63), (int32_t)((y0 - 1) * g.X + x0 + 64));
    }
}

// 3-D: links across tile faces from the packed mask (1 bit per voxel, one 64-bit word per row segment).  One thread
// per word: for each raster-order predecessor row it applies the same one-union-per-touching-run-pair bit rules as the
// tile kernel, but only where the two rows sit in different tiles; links through the x faces are the single bit pairs
// 63|0 between neighbouring words.
__device__ __forceinline__ void link_words(int32_t* Li, u64 m, u64 u, int64_t pm, int64_t pu, bool diag) {
    if (!u) return;
    const LinkMasks k = link_masks(m, u, diag);
    u64 any = k.c | k.l | k.r;
    while (any) {
        const int i = __ffsll((long long)any) - 1;
        const u64 bit = 1ull << i;
        any &= any - 1;
        if (k.c & bit) gl_unite(Li, (int32_t)(pm + i), (int32_t)(pu + i));
        if (k.l & bit) gl_unite(Li, (int32_t)(pm + i), (int32_t)(pu + i - 1));
        if (k.r & bit) gl_unite(Li, (int32_t)(pm + i), (int32_t)(pu + i + 1));
    }
}

__global__ void __launch_bounds__(kThreads) ccl_face3d_kernel(int32_t* __restrict__ L, const u64* __restrict__ bitvol,
                                                              Geom g, int64_t words_per_image) {
    // four threads per word, one per predecessor row: a thread's unions are a chain of dependent memory round trips, and
    // the kernel ends when the longest chain does
    const int64_t gt = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    const int64_t w = gt >> 2;
    const int kk = (int)(gt & 3);
    if (w >= words_per_image) return;
    const int wpr = g.tiles_x;
    const u64* B = bitvol + (int64_t)blockIdx.y * words_per_image;
    be_pdl_wait();   // the tile kernel's parents and packed mask
    const u64 m = B[w];
    if (!m) return;
    int32_t* Li = L + (int64_t)blockIdx.y * g.Z * g.Y * g.X;
    const int tx = (int)(w % wpr);
    const int64_t row = w / wpr;
    const int y = (int)(row % g.Y), z = (int)(row / g.Y);
    const int TY = 1 << g.ty_log2, TZ = kRows >> g.ty_log2;
    const int ly = y & (TY - 1), lz = z & (TZ - 1);
    const int64_t x0 = (int64_t)tx * kTX;
    const int64_t pm = ((int64_t)z * g.Y + y) * g.X + x0;
    // same row, x face
    if (kk == 0 && tx > 0 && (m & 1ull) && (B[w - 1] >> 63)) gl_unite(Li, (int32_t)pm, (int32_t)(pm - 1));
    // predecessor rows (dz, dy): (0,-1) (-1,-1) (-1,0) (-1,+1)
    {
        const int k = kk;
        const int dz = (k == 0) ? 0 : -1, dy = (k == 0 || k == 1) ? -1 : (k == 2 ? 0 : 1);
        const int steps = (dz != 0) + (dy != 0);
        if (steps > g.conn) return;
        const int zz = z + dz, yy = y + dy;
        if (zz < 0 || yy < 0 || yy >= g.Y) return;
        const bool diag = steps + 1 <= g.conn;
        const bool other_tile = (dz < 0 && lz == 0) || (dy < 0 && ly == 0) || (dy > 0 && ly == TY - 1);
        if (!other_tile && !diag) return;
        const int64_t wu = ((int64_t)zz * g.Y + yy) * wpr + tx;
        const int64_t pu = ((int64_t)zz * g.Y + yy) * g.X + x0;
        const u64 u = B[wu];
        if (other_tile) link_words(Li, m, u, pm, pu, diag);
        if (diag) {
            // corner bits through the x faces; redundant when the straight neighbour in row u is set (it is linked to
            // this row by the straight rule and to its x neighbour by the same-row rule)
            if (tx > 0 && (m & 1ull) && !(u & 1ull) && (B[wu - 1] >> 63)) gl_unite(Li, (int32_t)pm, (int32_t)(pu - 1));
            if (tx + 1 < wpr && (m >> 63) && !(u >> 63) && (B[wu + 1] & 1ull))
                gl_unite(Li, (int32_t)(pm + 63), (int32_t)(pu + 64));
        }
    }
}

// ---- raster-order numbering of the surviving roots + resolve, one launch (see the file header) ----
constexpr unsigned long long kFlagAggregate = 1ull << 62, kFlagPrefix = 2ull << 62, kFlagMask = 3ull << 62;

__device__ __forceinline__ int32_t resolve_one(const int32_t* __restrict__ Li, int32_t v) {
    // roots hold -(id) - 1; a pixel is normally one or two hops from its root
    if (v >= 0) {
        v = __ldg(Li + v);
        if (v >= 0) {
            v = __ldg(Li + v);
            while (v >= 0) v = __ldg(Li + v);
        }
    }
    return -v - 1;
}

__global__ void __launch_bounds__(kThreads) ccl_number_kernel(int32_t* __restrict__ L, const u64* __restrict__ rootbits,
                                                              int64_t vol, int64_t X, int wpr, int64_t words_per_image,
                                                              int blocks_per_image, unsigned* __restrict__ tickets,
                                                              unsigned long long* __restrict__ block_state,
                                                              int32_t* __restrict__ n_labels) {
    __shared__ int s_warp[kThreads / 32];
    __shared__ unsigned s_ticket;
    __shared__ long long s_excl;
    const int img = blockIdx.y;
    // blocks take their part of the bitmap by ticket, i.e. in starting order: a block only ever waits for blocks with
    // smaller tickets, which are running already
    if (threadIdx.x == 0) s_ticket = atomicAdd(&tickets[img], 1u);
    __syncthreads();
    const int blk = (int)s_ticket;
    int32_t* Li = L + (int64_t)img * vol;
    const int64_t w = (int64_t)blk * kWordsPerBlock + threadIdx.x;
    // the root bitmap was written by the tile kernel, which had completed before the border-union kernel let this one
    // start: it is read while that kernel still runs
    u64 bits = (w < words_per_image) ? __ldcg(rootbits + (int64_t)img * words_per_image + w) : 0ull;
    be_pdl_wait();   // the border unions
    be_pdl_launch_dependents();   // the resolve kernel may be placed: it fetches its mask words and waits for the numbers
    u64 keep = 0;
    int64_t base = 0;
    if (w < words_per_image) {
        base = (w / wpr) * X + (w % wpr) * kTX;
        // still a root after the border unions?  Four candidates per round, their loads issued together
        while (bits) {
            int b[4];
            int32_t pv[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                b[i] = bits ? __ffsll((long long)bits) - 1 : -1;
                bits &= bits - 1;   // (0 & anything stays 0)
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) pv[i] = (b[i] >= 0) ? __ldcg(Li + base + b[i]) : -1;
#pragma unroll
            for (int i = 0; i < 4; ++i)
                if (b[i] >= 0 && pv[i] == (int32_t)(base + b[i])) keep |= 1ull << b[i];
        }
    }
    const int c = __popcll(keep);
    int incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if ((threadIdx.x & 31) >= o) incl += t;
    }
    if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = incl;
    __syncthreads();
    int before = incl - c, total = 0;
    for (int i = 0; i < kThreads / 32; ++i) {
        if (i < (int)(threadIdx.x >> 5)) before += s_warp[i];
        total += s_warp[i];
    }
    // Exclusive prefix over the blocks before this one: decoupled look-back by warp 0, 128 predecessors per step (four
    // per lane, nearest first), polled with plain loads — thousands of pollers spinning with atomics on the same few KB
    // is slow, and with all blocks resident at once nobody finds a published prefix early, so the step count matters.
    unsigned long long* state = block_state + (int64_t)img * blocks_per_image;
    volatile unsigned long long* vstate = state;
    if (threadIdx.x < 32) {
        const int lane = threadIdx.x;
        if (lane == 0 && blk > 0) atomicExch(&state[blk], kFlagAggregate | (unsigned long long)total);
        long long excl = 0;
        for (int end = blk; end > 0; end -= 128) {
            long long v = 0;
            int best = -1;   // nearest entry of this lane that carries an inclusive prefix
            unsigned long long sj[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int j = end - 1 - lane - 32 * i;
                sj[i] = kFlagAggregate;   // entries before block 0 contribute nothing
                if (j >= 0) {
                    do {
                        sj[i] = vstate[j];
                    } while ((sj[i] & kFlagMask) == 0);
                    if ((sj[i] & kFlagMask) == kFlagPrefix) best = max(best, j);
                }
            }
            const int nearest = __reduce_max_sync(0xffffffffu, best);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int j = end - 1 - lane - 32 * i;
                if (j >= 0 && j >= nearest) v += (long long)(sj[i] & ~kFlagMask);
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            excl += v;
            if (nearest >= 0) break;
        }
        if (lane == 0) {
            atomicExch(&state[blk], kFlagPrefix | (unsigned long long)(excl + total));
            s_excl = excl;
            if (blk == blocks_per_image - 1) n_labels[img] = (int32_t)(excl + total);
        }
    }
    __syncthreads();
    int id = (int)s_excl + before;
    while (keep) {
        const int b = __ffsll((long long)keep) - 1;
        keep &= keep - 1;
        __stcg(Li + base + b, -(++id) - 1);
    }
}

// grid: x = row (z * Y + y), y = blocks of 8-px units along the row, z = image.  The packed mask says which pixels have a
// parent entry at all: all-background units (most of a typical mask) are written without touching the parent array,
// and a unit's two 16-byte parent loads are issued together.
__global__ void __launch_bounds__(kThreads) ccl_resolve_kernel(const int32_t* __restrict__ L, const u64* __restrict__ bitvol,
                                                               int32_t* __restrict__ out, int64_t vol, int X, int wpr,
                                                               int64_t words_per_image, int units_per_row) {
    const int u = blockIdx.y * kThreads + threadIdx.x;
    if (u >= units_per_row) return;
    const int64_t row = blockIdx.x;
    const int x = u * 8;
    // the packed mask is the caller's or the tile kernel's (complete long ago): fetched while the numbering still runs
    const u64 word = __ldg(bitvol + (int64_t)blockIdx.z * words_per_image + row * wpr + (x >> 6));
    be_pdl_wait();   // the numbering (and, through it, everything before)
    const unsigned bits = (unsigned)(word >> (x & 63)) & 0xffu;
    const int32_t* Li = L + (int64_t)blockIdx.z * vol;
    const int64_t base = row * X + x;
    int32_t* o = out + (int64_t)blockIdx.z * vol + base;
    const bool vec = x + 8 <= X && (((uintptr_t)o | (uintptr_t)(Li + base)) & 15) == 0;
    int32_t p[8], v[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (bits) {
        if (vec) {
            int4 a = make_int4(-1, -1, -1, -1), b = a;
            if (bits & 0x0fu) a = __ldcg(reinterpret_cast<const int4*>(Li + base));
            if (bits & 0xf0u) b = __ldcg(reinterpret_cast<const int4*>(Li + base) + 1);
            p[0] = a.x, p[1] = a.y, p[2] = a.z, p[3] = a.w, p[4] = b.x, p[5] = b.y, p[6] = b.z, p[7] = b.w;
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) p[j] = ((bits >> j) & 1u) ? __ldcg(Li + base + j) : -1;
        }
        int32_t prev_p = 0x7fffffff, prev_v = 0;   // neighbouring pixels usually share their parent: look it up once
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            if ((bits >> j) & 1u) {
                if (p[j] != prev_p) {
                    prev_p = p[j];
                    prev_v = resolve_one(Li, p[j]);
                }
                v[j] = prev_v;
            }
        }
    }
    if (vec) {
        __stcs(reinterpret_cast<int4*>(o), make_int4(v[0], v[1], v[2], v[3]));
        __stcs(reinterpret_cast<int4*>(o) + 1, make_int4(v[4], v[5], v[6], v[7]));
    } else {
#pragma unroll
        for (int j = 0; j < 8; ++j)
            if (x + j < X) o[j] = v[j];
    }
}

struct CclWorkspace {
    int32_t* L;
    u64* rootbits;
    u64* bitvol;        // the mask, 1 bit per voxel
    TileEdges* edges;
    unsigned* tickets;                 // [N]
    unsigned long long* block_state;   // [N][blocks_per_image]
    size_t scan_bytes;                 // tickets + block_state: zeroed by every call
    int64_t words_per_image;
    int blocks_per_image;
    size_t total_bytes;
};

CclWorkspace carve_ccl_ws(void* ws, int64_t N, int64_t Z, int64_t Y, int64_t X) {
    CclWorkspace w;
    const int64_t vol = Z * Y * X;
    const int ndim = Z > 1 ? 3 : 2;
    const int TY = ndim == 2 ? 64 : 8, TZ = kRows / TY;
    const int64_t tiles_x = (X + kTX - 1) / kTX, tiles_y = (Y + TY - 1) / TY, tiles_z = (Z + TZ - 1) / TZ;
    w.words_per_image = Z * Y * tiles_x;
    w.blocks_per_image = (int)((w.words_per_image + kWordsPerBlock - 1) / kWordsPerBlock);
    char* p = (char*)ws;
    size_t off = 0;
    w.L = (int32_t*)(p + off);
    off += be_align_up(sizeof(int32_t) * N * vol, 256);
    w.rootbits = (u64*)(p + off);
    off += be_align_up(sizeof(u64) * N * w.words_per_image, 256);
    w.bitvol = (u64*)(p + off);
    off += be_align_up(sizeof(u64) * N * w.words_per_image, 256);
    w.edges = (TileEdges*)(p + off);
    off += be_align_up(sizeof(TileEdges) * N * tiles_x * tiles_y * tiles_z, 256);
    w.tickets = (unsigned*)(p + off);
    const size_t scan_begin = off;
    off += be_align_up(sizeof(unsigned) * N, 256);
    w.block_state = (unsigned long long*)(p + off);
    off += be_align_up(sizeof(unsigned long long) * N * w.blocks_per_image, 256);
    w.scan_bytes = off - scan_begin;
    w.total_bytes = off;
    return w;
}

}  // namespace

extern "C" size_t be_ccl_workspace_bytes(int64_t N, int64_t Z, int64_t Y, int64_t X) {
    if (N <= 0 || Z <= 0 || Y <= 0 || X <= 0) return 256;
    return carve_ccl_ws(nullptr, N, Z, Y, X).total_bytes;
}

static int ccl_run(const char* who, int in_kind, const void* input, Epilogue epi, int32_t* labels, int64_t N, int64_t Z,
                   int64_t Y, int64_t X, int connectivity, int32_t* n_labels, void* ws, be_stream_t stream) {
    BE_REQUIRE(input && labels && n_labels && ws, BE_ERR_INVALID_ARG, "%s: null pointer", who);
    BE_REQUIRE(N > 0 && Z > 0 && Y > 0 && X > 0, BE_ERR_INVALID_ARG, "%s: empty input (%lld,%lld,%lld,%lld)", who,
               (long long)N, (long long)Z, (long long)Y, (long long)X);
    const int ndim = (Z > 1) ? 3 : 2;
    BE_REQUIRE(connectivity >= 1 && connectivity <= ndim, BE_ERR_INVALID_ARG, "%s: connectivity %d not in 1..%d", who,
               connectivity, ndim);
    const int64_t vol = Z * Y * X;
    BE_REQUIRE(vol < ((int64_t)1 << 31) - 2, BE_ERR_TOO_LARGE, "%s: %lld voxels per image >= 2^31", who, (long long)vol);
    BE_REQUIRE(N <= 65535, BE_ERR_TOO_LARGE, "%s: N %lld > 65535", who, (long long)N);
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
    BE_REQUIRE(n_tiles < ((int64_t)1 << 31) - 1, BE_ERR_TOO_LARGE, "%s: too many tiles", who);
    cudaStream_t s = (cudaStream_t)stream;
    const CclWorkspace w = carve_ccl_ws(ws, N, Z, Y, X);

    g.scan_zero = w.tickets;
    g.scan_words = (int)(w.scan_bytes / sizeof(unsigned));
    BE_REQUIRE(g.tiles_y <= 65535 && (int64_t)g.tiles_z * N <= 65535, BE_ERR_TOO_LARGE,
               "%s: tile grid (%d, %d, %lld) exceeds the launch limits", who, g.tiles_x, g.tiles_y, (long long)g.tiles_z * N);
    const dim3 grid((unsigned)g.tiles_x, (unsigned)g.tiles_y, (unsigned)(g.tiles_z * N));
    TileEdges* edges = ndim == 2 ? w.edges : nullptr;
    // a bit-packed input whose rows are whole words is used in place of the workspace copy of the mask
    const u64* bitvol = (in_kind == kInBits && (X % kTX) == 0) ? (const u64*)input : w.bitvol;
    // no memset in this call: the tile kernel is chained to whatever ran before it like the kernels after it
    if (in_kind == kInBits)
        BE_CUDA(be_launch_pdl(ccl_tile_kernel<kInBits>, grid, dim3(kThreads), 0, s, input, epi, w.L, w.rootbits, w.bitvol, edges, g));
    else if (in_kind == kInF32)
        BE_CUDA(be_launch_pdl(ccl_tile_kernel<kInF32>, grid, dim3(kThreads), 0, s, input, epi, w.L, w.rootbits, w.bitvol, edges, g));
    else
        BE_CUDA(be_launch_pdl(ccl_tile_kernel<kInU8>, grid, dim3(kThreads), 0, s, input, epi, w.L, w.rootbits, w.bitvol, edges, g));
    // the following kernels are chained to their predecessor with programmatic stream serialization (see pdl_wait)
    cudaLaunchAttribute pdl[1];
    pdl[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    pdl[0].val.programmaticStreamSerializationAllowed = 1;
    cudaLaunchConfig_t cfg = {};
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = s;
    cfg.attrs = pdl;
    cfg.numAttrs = be_pdl_enabled() ? 1 : 0;
    if (ndim == 2) {
        if (n_tiles > 1) {
            cfg.gridDim = dim3((unsigned)((4 * n_tiles + kThreads / 32 - 1) / (kThreads / 32)), (unsigned)N);   // four warps per tile
            BE_CUDA(cudaLaunchKernelEx(&cfg, ccl_edge_kernel, w.L, (const TileEdges*)w.edges, g, (int)n_tiles));
        }
    } else {
        cfg.gridDim = dim3((unsigned)((4 * w.words_per_image + kThreads - 1) / kThreads), (unsigned)N);
        BE_CUDA(cudaLaunchKernelEx(&cfg, ccl_face3d_kernel, w.L, bitvol, g, w.words_per_image));
    }
    cfg.gridDim = dim3((unsigned)w.blocks_per_image, (unsigned)N);
    BE_CUDA(cudaLaunchKernelEx(&cfg, ccl_number_kernel, w.L, (const u64*)w.rootbits, vol, X, g.tiles_x, w.words_per_image,
                               w.blocks_per_image, w.tickets, w.block_state, n_labels));
    const int upr = (int)((X + 7) / 8);
    BE_REQUIRE((upr + kThreads - 1) / kThreads <= 65535 && N <= 65535, BE_ERR_TOO_LARGE, "%s: X %lld too long", who, (long long)X);
    cfg.gridDim = dim3((unsigned)(Z * Y), (unsigned)((upr + kThreads - 1) / kThreads), (unsigned)N);
    BE_CUDA(cudaLaunchKernelEx(&cfg, ccl_resolve_kernel, (const int32_t*)w.L, bitvol, labels, vol, (int)X, g.tiles_x,
                               w.words_per_image, upr));
    return BE_OK;
}

extern "C" int be_ccl_u8_i32(const uint8_t* mask, int32_t* labels, int64_t N, int64_t Z, int64_t Y, int64_t X,
                             int connectivity, int32_t* n_labels, void* ws, be_stream_t stream) {
    return ccl_run("be_ccl_u8_i32", kInU8, mask, make_epilogue(BE_EPI_NONE, 0.5f, false), labels, N, Z, Y, X, connectivity,
                   n_labels, ws, stream);
}

extern "C" int be_ccl_bits_i32(const uint64_t* bits, int32_t* labels, int64_t N, int64_t Z, int64_t Y, int64_t X,
                               int connectivity, int32_t* n_labels, void* ws, be_stream_t stream) {
    BE_REQUIRE((((uintptr_t)bits) & 7) == 0, BE_ERR_INVALID_ARG, "be_ccl_bits_i32: bits must be 8-byte aligned");
    return ccl_run("be_ccl_bits_i32", kInBits, bits, make_epilogue(BE_EPI_NONE, 0.5f, false), labels, N, Z, Y, X,
                   connectivity, n_labels, ws, stream);
}

extern "C" int be_ccl_f32_i32(const float* x, int epilogue, float thr, int32_t* labels, int64_t N, int64_t Z, int64_t Y,
                              int64_t X, int connectivity, int32_t* n_labels, void* ws, be_stream_t stream) {
    BE_REQUIRE(epilogue == BE_EPI_NONE || epilogue == BE_EPI_SIGMOID, BE_ERR_INVALID_ARG, "be_ccl_f32_i32: epilogue %d",
               epilogue);
    return ccl_run("be_ccl_f32_i32", kInF32, x, make_epilogue(epilogue, thr, true), labels, N, Z, Y, X, connectivity,
                   n_labels, ws, stream);
}

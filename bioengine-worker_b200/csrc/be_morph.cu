// Mask clean-up and distance transform — the deterministic part of the fibsem post-processing
// (ref: apps/fibsem-mito-analysis/analysis_deployment.py:166-169): remove_small_objects, binary_closing(disk), exact EDT.
//
//   be_label_area_i32 / be_keep_large_u8   component sizes and the size filter (with be_ccl_u8_i32 at connectivity 1)
//   be_binary_closing_u8                   dilation then erosion with a row-span structuring element, on bit-packed rows:
//                                          a 32-pixel word is OR-ed / AND-ed with shifted copies of itself and its two
//                                          neighbours, so a 9x9 disk costs ~70 bit operations per 32 pixels
//   be_edt_u8_f64                          exact Euclidean distance to the nearest zero pixel (Meijster): column scan
//                                          for the vertical distance, then the lower envelope of parabolas per row
#include "be_common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kMaxRadius = 15;

struct Spans {
    int radius;
    int hw[2 * kMaxRadius + 1];  // half width of the structuring element's row dy = -radius..radius (-1: empty row)
};

// ---- component sizes ----

__global__ void __launch_bounds__(kThreads) label_area_kernel(const int32_t* __restrict__ labels, int64_t n_px, int64_t cap,
                                                              int32_t* __restrict__ area, int32_t* __restrict__ status) {
    const int64_t img = blockIdx.y;
    const int32_t* lab = labels + img * n_px;
    const int lane = threadIdx.x & 31;
    const int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    const bool in = i < n_px;
    const int32_t l = in ? lab[i] : 0;
    const int32_t lp = __shfl_up_sync(0xffffffffu, l, 1);
    const bool fg = in && l > 0;
    const bool start = fg && (lane == 0 || lp != l);
    const unsigned brks = __ballot_sync(0xffffffffu, !fg || start);
    if (!start) return;
    if (l >= cap) {
        status[img] = 2;
        return;
    }
    const unsigned stop = brks & (lane == 31 ? 0u : (0xffffffffu << (lane + 1)));
    const int end = stop ? (__ffs(stop) - 1) : 32;
    atomicAdd(&area[img * cap + l], end - lane);  // one atomic per run of equal labels
}

__global__ void __launch_bounds__(kThreads) keep_large_kernel(const int32_t* __restrict__ labels,
                                                              const int32_t* __restrict__ area, int64_t n_px, int64_t cap,
                                                              int32_t min_size, uint8_t* __restrict__ out) {
    const int64_t img = blockIdx.y;
    const int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    if (i >= n_px) return;
    const int32_t l = labels[img * n_px + i];
    out[img * n_px + i] = (l > 0 && l < cap && area[img * cap + l] >= min_size) ? 1 : 0;
}

// ---- bit-packed morphology ----

__global__ void __launch_bounds__(kThreads) pack_kernel(const uint8_t* __restrict__ mask, unsigned* __restrict__ bits,
                                                        int64_t rows, int64_t X, int wpr) {
    const int64_t w = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    if (w >= rows * wpr) return;
    const int64_t row = w / wpr;
    const int wx = (int)(w - row * wpr);
    const uint8_t* p = mask + row * X + (int64_t)wx * 32;
    const int n = (int)min((int64_t)32, X - (int64_t)wx * 32);
    unsigned b = 0;
    for (int i = 0; i < n; ++i) b |= (unsigned)(p[i] != 0) << i;
    bits[w] = b;
}

__global__ void __launch_bounds__(kThreads) unpack_kernel(const unsigned* __restrict__ bits, uint8_t* __restrict__ mask,
                                                          int64_t rows, int64_t X, int wpr) {
    const int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    if (i >= rows * X) return;
    const int64_t row = i / X;
    const int x = (int)(i - row * X);
    mask[i] = (bits[row * wpr + (x >> 5)] >> (x & 31)) & 1u;
}

// one thread = one 32-pixel word of one output row; pixels outside the image count as 0 for both operations
// (scipy's border_value = 0), so erosion eats into objects that touch the border
template <bool DILATE>
__global__ void __launch_bounds__(kThreads) morph_kernel(const unsigned* __restrict__ in, unsigned* __restrict__ out,
                                                         int64_t planes, int Y, int64_t X, int wpr, Spans sp) {
    const int64_t w = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    if (w >= planes * Y * wpr) return;
    const int wx = (int)(w % wpr);
    const int64_t row = w / wpr;
    const int y = (int)(row % Y);
    const unsigned* plane = in + (row - y) * wpr;
    unsigned acc = DILATE ? 0u : 0xffffffffu;
    for (int dy = -sp.radius; dy <= sp.radius; ++dy) {
        const int h = sp.hw[dy + sp.radius];
        if (h < 0) continue;
        const int yy = y + dy;
        unsigned prev = 0, cur = 0, next = 0;
        if (yy >= 0 && yy < Y) {
            const unsigned* r = plane + (int64_t)yy * wpr;
            cur = r[wx];
            if (wx > 0) prev = r[wx - 1];
            if (wx + 1 < wpr) next = r[wx + 1];
        }
        unsigned v = cur;
        for (int k = 1; k <= h; ++k) {
            const unsigned left = (cur << k) | (prev >> (32 - k));    // pixel x - k
            const unsigned right = (cur >> k) | (next << (32 - k));   // pixel x + k
            v = DILATE ? (v | left | right) : (v & left & right);
        }
        acc = DILATE ? (acc | v) : (acc & v);
    }
    const int64_t rem = X - (int64_t)wx * 32;
    if (rem < 32) acc &= (1u << rem) - 1u;  // keep the padding bits of the last word zero
    out[w] = acc;
}

// ---- exact Euclidean distance transform (Meijster, Roerdink, Hesselink 2000) ----

constexpr int kInf = 1 << 29;

// phase 1: g(y, x) = distance to the nearest zero pixel in column x (kInf if the column has none).  A column is cut into
// chunks of 64 rows so that a 2048-px-wide image keeps 65 536 threads busy instead of 2048: first every (column, chunk) notes
// the first and last zero row of its chunk, then it finds the nearest zero above and below among its neighbours' notes
// (usually the adjacent chunk) and walks its 64 rows down and up with the mask bits it kept from the first walk.
constexpr int kEdtChunk = 64;

__global__ void __launch_bounds__(kThreads) edt_col_marks_kernel(const uint8_t* __restrict__ mask, int2* __restrict__ marks,
                                                                 int Y, int64_t X, int n_chunks) {
    const int64_t x = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    if (x >= X) return;
    const int chunk = blockIdx.y;
    const uint8_t* m = mask + (int64_t)blockIdx.z * Y * X;
    const int y0 = chunk * kEdtChunk, y1 = min(y0 + kEdtChunk, Y);
    int first0 = -1, last0 = -1;
    for (int y = y0; y < y1; ++y) {
        if (!m[(int64_t)y * X + x]) {
            if (first0 < 0) first0 = y;
            last0 = y;
        }
    }
    marks[((int64_t)blockIdx.z * n_chunks + chunk) * X + x] = make_int2(first0, last0);
}

__global__ void __launch_bounds__(kThreads) edt_col_dist_kernel(const uint8_t* __restrict__ mask, const int2* __restrict__ marks,
                                                                int32_t* __restrict__ g, int Y, int64_t X, int n_chunks) {
    const int64_t x = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    if (x >= X) return;
    const int chunk = blockIdx.y;
    const uint8_t* m = mask + (int64_t)blockIdx.z * Y * X;
    int32_t* gp = g + (int64_t)blockIdx.z * Y * X;
    const int2* mk = marks + (int64_t)blockIdx.z * n_chunks * X + x;
    int above = -1, below = -1;   // nearest zero rows outside this chunk
    for (int c = chunk - 1; c >= 0 && above < 0; --c) above = mk[(int64_t)c * X].y;
    for (int c = chunk + 1; c < n_chunks && below < 0; ++c) below = mk[(int64_t)c * X].x;
    const int y0 = chunk * kEdtChunk, y1 = min(y0 + kEdtChunk, Y);
    unsigned long long fg = 0;   // the chunk's mask bits, kept for the walk back up
    int last = above;
    for (int y = y0; y < y1; ++y) {
        const bool on = m[(int64_t)y * X + x] != 0;
        fg |= (unsigned long long)on << (y - y0);
        if (!on) last = y;
        gp[(int64_t)y * X + x] = last < 0 ? kInf : y - last;
    }
    int next = below;
    for (int y = y1 - 1; y >= y0; --y) {
        if (!((fg >> (y - y0)) & 1ull)) next = y;
        const int d = next < 0 ? kInf : next - y;
        const int64_t i = (int64_t)y * X + x;
        if (d < gp[i]) gp[i] = d;
    }
}

// phase 2: one block per row.  A background pixel (g = 0) separates the row: for a pixel between two of them no parabola
// from beyond either can be the lowest.  So every run of foreground pixels is a problem of its own — one thread builds the
// lower envelope of the parabolas (x - i)^2 + g(i)^2 of its run and of the background pixel on its left in its own stretch
// of shared memory, then walks the run back to front; the background pixel on its right is kept out of the stack (its
// slot belongs to the next run) and enters as a min with (end - x)^2.
// OUT = double: the distance; OUT = int32_t: its square (exact, saturated) — what the seed detection and the flood order by
template <typename OUT>
__global__ void __launch_bounds__(kThreads) edt_rows_kernel(const int32_t* __restrict__ g, OUT* __restrict__ dist, int Y,
                                                            int X) {
    extern __shared__ int smem[];
    int* s = smem;           // parabola apexes of the envelopes
    int* t = smem + X;       // t[q]: first x where s[q] is the lowest
    int* gs = smem + 2 * X;  // this row of g
    const int64_t row = blockIdx.x;
    const int32_t* gr = g + row * X;
    int big = 0;
    for (int x = threadIdx.x; x < X; x += kThreads) {
        const int v = gr[x];
        gs[x] = v;
        big |= v >= 16384;
    }
    // with every column distance of the row below 2^14 (and X <= 2^14) all squares and their sums fit 32 bits: the envelope
    // then costs 32-bit multiplies and one 32-bit division per parabola instead of 64-bit ones (a row whose columns hold no
    // zero at all carries kInf and takes the wide path)
    big = __syncthreads_or(big);
    auto put = [&](int x, long long d2) {
        if constexpr (sizeof(OUT) == 8) dist[row * X + x] = sqrt((double)d2);
        else dist[row * X + x] = (OUT)min(d2, (long long)INT32_MAX);
    };
    auto rows_of = [&](auto tag) {
        using T = decltype(tag);
        for (int x0 = threadIdx.x; x0 < X; x0 += kThreads) {
            if (gs[x0] == 0) {
                put(x0, 0);
                continue;
            }
            if (x0 > 0 && gs[x0 - 1] != 0) continue;  // not the first pixel of a run
            int end = x0 + 1;
            while (end < X && gs[end] != 0) ++end;     // [x0, end) is the run; gs[end] == 0 if end < X
            const int a = x0 > 0 ? x0 - 1 : 0;         // first candidate: the background pixel on the left, if there is one
            auto f = [&](T x, T i) {
                const T gi = gs[i];
                return (x - i) * (x - i) + gi * gi;
            };
            int q = a;  // stack lives in s[a..], t[a..]
            s[a] = a;
            t[a] = a;
            for (int u = a + 1; u < end; ++u) {
                while (q >= a && f(t[q], s[q]) > f(t[q], u)) --q;
                if (q < a) {
                    q = a;
                    s[a] = u;
                } else {
                    const T i = s[q], gi = gs[i], gu = gs[u];
                    const T w = 1 + (((T)u * u - i * i + gu * gu - gi * gi) / (2 * ((T)u - i)));
                    if (w < end) {
                        ++q;
                        s[q] = u;
                        t[q] = (int)w;
                    }
                }
            }
            for (int x = end - 1; x >= x0; --x) {
                while (q > a && t[q] > x) --q;
                T d2 = f(x, s[q]);
                if (end < X) d2 = min(d2, (T)(end - x) * (T)(end - x));
                put(x, (long long)d2);
            }
        }
    };
    if (big)
        rows_of((long long)0);
    else
        rows_of((int)0);
}

}  // namespace

extern "C" int be_label_area_i32(const int32_t* labels, int64_t N, int64_t n_px, int64_t cap, int32_t* area,
                                 int32_t* status, be_stream_t stream) {
    BE_REQUIRE(labels && area && status, BE_ERR_INVALID_ARG, "be_label_area_i32: null pointer");
    BE_REQUIRE(N > 0 && n_px > 0 && cap > 0, BE_ERR_INVALID_ARG, "be_label_area_i32: empty input");
    BE_REQUIRE(N <= 65535, BE_ERR_TOO_LARGE, "be_label_area_i32: N %lld > 65535", (long long)N);
    cudaStream_t s = (cudaStream_t)stream;
    BE_CUDA(cudaMemsetAsync(area, 0, sizeof(int32_t) * N * cap, s));
    BE_CUDA(cudaMemsetAsync(status, 0, sizeof(int32_t) * N, s));
    label_area_kernel<<<dim3((unsigned)((n_px + kThreads - 1) / kThreads), (unsigned)N), kThreads, 0, s>>>(labels, n_px, cap,
                                                                                                           area, status);
    BE_LAUNCH_CHECK();
    return BE_OK;
}

extern "C" int be_keep_large_u8(const int32_t* labels, const int32_t* area, int64_t N, int64_t n_px, int64_t cap,
                                int32_t min_size, uint8_t* out, be_stream_t stream) {
    BE_REQUIRE(labels && area && out, BE_ERR_INVALID_ARG, "be_keep_large_u8: null pointer");
    BE_REQUIRE(N > 0 && n_px > 0 && cap > 0, BE_ERR_INVALID_ARG, "be_keep_large_u8: empty input");
    BE_REQUIRE(N <= 65535, BE_ERR_TOO_LARGE, "be_keep_large_u8: N %lld > 65535", (long long)N);
    keep_large_kernel<<<dim3((unsigned)((n_px + kThreads - 1) / kThreads), (unsigned)N), kThreads, 0, (cudaStream_t)stream>>>(
        labels, area, n_px, cap, min_size, out);
    BE_LAUNCH_CHECK();
    return BE_OK;
}

extern "C" size_t be_binary_closing_workspace_bytes(int64_t N, int64_t Y, int64_t X) {
    if (N <= 0 || Y <= 0 || X <= 0) return 256;
    const int64_t wpr = (X + 31) / 32;
    return 2 * be_align_up(sizeof(unsigned) * N * Y * wpr, 256);
}

extern "C" int be_binary_closing_u8(const uint8_t* mask, uint8_t* out, int64_t N, int64_t Y, int64_t X,
                                    const int32_t* half_widths, int radius, void* ws, be_stream_t stream) {
    BE_REQUIRE(mask && out && half_widths && ws, BE_ERR_INVALID_ARG, "be_binary_closing_u8: null pointer");
    BE_REQUIRE(N > 0 && Y > 0 && X > 0, BE_ERR_INVALID_ARG, "be_binary_closing_u8: empty input");
    BE_REQUIRE(radius >= 0 && radius <= kMaxRadius, BE_ERR_INVALID_ARG, "be_binary_closing_u8: radius %d not in 0..%d", radius,
               kMaxRadius);
    BE_REQUIRE(Y < (1 << 30) && X < ((int64_t)1 << 36), BE_ERR_TOO_LARGE, "be_binary_closing_u8: image too large");
    Spans sp;
    sp.radius = radius;
    for (int i = 0; i < 2 * radius + 1; ++i) {
        BE_REQUIRE(half_widths[i] >= -1 && half_widths[i] <= 31, BE_ERR_INVALID_ARG,
                   "be_binary_closing_u8: half width %d of row %d not in -1..31", half_widths[i], i - radius);
        BE_REQUIRE(half_widths[i] == half_widths[2 * radius - i], BE_ERR_INVALID_ARG,
                   "be_binary_closing_u8: the structuring element must be symmetric (rows %d and %d differ)", i - radius,
                   radius - i);
        sp.hw[i] = half_widths[i];
    }
    cudaStream_t s = (cudaStream_t)stream;
    const int wpr = (int)((X + 31) / 32);
    const int64_t rows = N * Y, words = rows * wpr;
    unsigned* a = (unsigned*)ws;
    unsigned* b = (unsigned*)((char*)ws + be_align_up(sizeof(unsigned) * words, 256));
    const unsigned gw = (unsigned)((words + kThreads - 1) / kThreads);
    pack_kernel<<<gw, kThreads, 0, s>>>(mask, a, rows, X, wpr);
    BE_LAUNCH_CHECK();
    morph_kernel<true><<<gw, kThreads, 0, s>>>(a, b, N, (int)Y, X, wpr, sp);
    BE_LAUNCH_CHECK();
    morph_kernel<false><<<gw, kThreads, 0, s>>>(b, a, N, (int)Y, X, wpr, sp);
    BE_LAUNCH_CHECK();
    unpack_kernel<<<(unsigned)((rows * X + kThreads - 1) / kThreads), kThreads, 0, s>>>(a, out, rows, X, wpr);
    BE_LAUNCH_CHECK();
    return BE_OK;
}

extern "C" size_t be_edt_workspace_bytes(int64_t N, int64_t Y, int64_t X) {
    if (N <= 0 || Y <= 0 || X <= 0) return 256;
    const int64_t n_chunks = (Y + kEdtChunk - 1) / kEdtChunk;
    return be_align_up(sizeof(int32_t) * N * Y * X, 256) + be_align_up(sizeof(int2) * N * n_chunks * X, 256);
}

template <typename OUT>
static int edt_run(const char* who, const uint8_t* mask, OUT* dist, int64_t N, int64_t Y, int64_t X, void* ws, be_stream_t stream) {
    BE_REQUIRE(mask && dist && ws, BE_ERR_INVALID_ARG, "%s: null pointer", who);
    BE_REQUIRE(N > 0 && Y > 0 && X > 0, BE_ERR_INVALID_ARG, "%s: empty input", who);
    BE_REQUIRE(N <= 65535, BE_ERR_TOO_LARGE, "%s: N %lld > 65535", who, (long long)N);
    BE_REQUIRE(X <= 16384 && Y < (1 << 28), BE_ERR_TOO_LARGE, "%s: X %lld > 16384 (row envelope lives in shared memory)", who,
               (long long)X);
    BE_REQUIRE(N * Y < ((int64_t)1 << 31), BE_ERR_TOO_LARGE, "%s: too many rows", who);
    cudaStream_t s = (cudaStream_t)stream;
    int32_t* g = (int32_t*)ws;
    int2* marks = (int2*)((char*)ws + be_align_up(sizeof(int32_t) * N * Y * X, 256));
    const int n_chunks = (int)((Y + kEdtChunk - 1) / kEdtChunk);
    BE_REQUIRE(n_chunks <= 65535, BE_ERR_TOO_LARGE, "%s: Y %lld too large", who, (long long)Y);
    const dim3 cgrid((unsigned)((X + kThreads - 1) / kThreads), (unsigned)n_chunks, (unsigned)N);
    edt_col_marks_kernel<<<cgrid, kThreads, 0, s>>>(mask, marks, (int)Y, X, n_chunks);
    BE_LAUNCH_CHECK();
    edt_col_dist_kernel<<<cgrid, kThreads, 0, s>>>(mask, marks, g, (int)Y, X, n_chunks);
    BE_LAUNCH_CHECK();
    const size_t smem = sizeof(int) * 3 * X;
    BE_CUDA(cudaFuncSetAttribute(edt_rows_kernel<OUT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    edt_rows_kernel<OUT><<<(unsigned)(N * Y), kThreads, smem, s>>>(g, dist, (int)Y, (int)X);
    BE_LAUNCH_CHECK();
    return BE_OK;
}

extern "C" int be_edt_u8_f64(const uint8_t* mask, double* dist, int64_t N, int64_t Y, int64_t X, void* ws,
                             be_stream_t stream) {
    return edt_run<double>("be_edt_u8_f64", mask, dist, N, Y, X, ws, stream);
}

extern "C" int be_edt_sq_u8_i32(const uint8_t* mask, int32_t* dist2, int64_t N, int64_t Y, int64_t X, void* ws,
                                be_stream_t stream) {
    return edt_run<int32_t>("be_edt_sq_u8_i32", mask, dist2, N, Y, X, ws, stream);
}

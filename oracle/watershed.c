/* CPU oracle — TEST INFRASTRUCTURE ONLY (see oracle/oracle.py).
 *
 * Marker-controlled flooding as `skimage.segmentation.watershed(image, markers, mask=mask)` runs it with its defaults
 * (connectivity 1, compactness 0, no watershed line) — the call at ref: apps/fibsem-mito-analysis/
 * analysis_deployment.py:175.  scikit-image is a third-party dependency of the reference that is neither vendored nor
 * installable here (PARITY UNPINNED); this restates its published algorithm (skimage/segmentation/_watershed.py,
 * _watershed_cy.pyx, heap_general.pxi):
 *
 *   - every marker pixel enters a min-heap in raster order with (value = image[p], age = 0);
 *   - the smallest (value, age) leaves the heap; each of its 4 neighbours (order: up, left, right, down — offsets sorted
 *     by distance with a stable sort) that is inside the mask and still unlabelled takes its label AT THAT MOMENT and
 *     enters the heap with age = ++counter;
 *   - the heap is a plain binary heap with strict comparisons on (value, age), so elements that tie in both — possible
 *     only among markers — leave in an order that depends on the array layout.
 *
 * tie_mode 0 keeps that heap (sift-up from the tail on push; on pop the tail moves to the root and sifts down, left
 * child preferred on ties).  tie_mode 1 breaks complete ties by push order (first in, first out), which makes the
 * result independent of the heap layout; this is the order a bucket queue gives and what the CUDA flood reproduces.
 * The two differ only where two markers of one connected region hold exactly the same value.
 */
#include <stdint.h>
#include <stdlib.h>

typedef struct {
    double value;
    int64_t age;
    int64_t seq;
    int64_t index;
} item_t;

static int before(const item_t* a, const item_t* b, int tie_mode) {
    if (a->value != b->value) return a->value < b->value;
    if (a->age != b->age) return a->age < b->age;
    return tie_mode ? a->seq < b->seq : 0;
}

typedef struct {
    item_t* a;
    int64_t n, cap;
    int tie_mode;
} heap_t;

static int heap_push(heap_t* h, const item_t* e) {
    if (h->n == h->cap) {
        int64_t cap = h->cap ? 2 * h->cap : 1024;
        item_t* p = (item_t*)realloc(h->a, sizeof(item_t) * (size_t)cap);
        if (!p) return -1;
        h->a = p;
        h->cap = cap;
    }
    int64_t c = h->n++;
    h->a[c] = *e;
    while (c > 0) {
        const int64_t p = (c + 1) / 2 - 1;
        if (!before(&h->a[c], &h->a[p], h->tie_mode)) break;
        const item_t t = h->a[c];
        h->a[c] = h->a[p];
        h->a[p] = t;
        c = p;
    }
    return 0;
}

static void heap_pop(heap_t* h, item_t* out) {
    *out = h->a[0];
    h->n -= 1;
    if (h->n == 0) return;
    h->a[0] = h->a[h->n];
    int64_t i = 0;
    for (;;) {
        const int64_t l = 2 * i + 1, r = 2 * i + 2;
        int64_t s = i;
        if (l >= h->n) break;
        if (before(&h->a[l], &h->a[i], h->tie_mode)) s = l;
        if (r < h->n && before(&h->a[r], &h->a[s], h->tie_mode)) s = r;
        if (s == i) break;
        const item_t t = h->a[i];
        h->a[i] = h->a[s];
        h->a[s] = t;
        i = s;
    }
}

/* image, markers, mask, out: Y*X row-major.  out starts as a copy of markers (skimage floods its padded copy of them).
 * Returns 0, or -1 when memory runs out. */
int oracle_watershed(const double* image, const int32_t* markers, const uint8_t* mask, int32_t* out, int64_t Y, int64_t X,
                     int tie_mode) {
    heap_t h = {0, 0, 0, tie_mode};
    int64_t age = 1, seq = 0;
    for (int64_t i = 0; i < Y * X; ++i) out[i] = markers[i];
    for (int64_t i = 0; i < Y * X; ++i) {
        if (!markers[i]) continue;
        const item_t e = {image[i], 0, seq++, i};
        if (heap_push(&h, &e)) return -1;
    }
    while (h.n > 0) {
        item_t e;
        heap_pop(&h, &e);
        const int64_t y = e.index / X, x = e.index - y * X;
        const int64_t ny[4] = {y - 1, y, y, y + 1};
        const int64_t nx[4] = {x, x - 1, x + 1, x};
        for (int k = 0; k < 4; ++k) {
            if (ny[k] < 0 || ny[k] >= Y || nx[k] < 0 || nx[k] >= X) continue; /* skimage pads with mask = 0 */
            const int64_t q = ny[k] * X + nx[k];
            if (!mask[q] || out[q]) continue;
            age += 1;
            out[q] = out[e.index];
            const item_t n = {image[q], age, seq++, q};
            if (heap_push(&h, &n)) {
                free(h.a);
                return -1;
            }
        }
    }
    free(h.a);
    return 0;
}

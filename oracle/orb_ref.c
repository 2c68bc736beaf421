/*
 * oracle/orb_ref.c -- CPU ORACLE for the ORB-SLAM2 front-end hot path.
 *
 * TEST INFRASTRUCTURE, NOT PRODUCT.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this library.  The
 * product (liborbb200.so) never links, loads or calls anything in oracle/.
 *
 * What it is: a plain-C restatement of the reference's algorithm for
 *   ORBextractor::operator()            /root/reference/src/ORBextractor.cc:1261-1339
 *   ORBmatcher::DescriptorDistance      /root/reference/src/ORBmatcher.cc:1733-1749
 *   the best/second-best search loops   /root/reference/src/ORBmatcher.cc:214-244 (and siblings)
 *   Frame::ComputeStereoMatches         /root/reference/src/Frame.cc:577-751
 *   Frame::AssignFeaturesToGrid / GetFeaturesInArea   src/Frame.cc:261-279, 386-442
 * each function below cites the lines it follows.
 *
 * The reference delegates four primitives to OpenCV (un-vendored, version
 * unpinned by the reference: CMakeLists.txt:32-38).  They are restated here from
 * OpenCV's published 8-bit algorithms and PINNED against the OpenCV actually on
 * this image (cv2 4.13.0) by tests/test_oracle_pin.py + oracle/orb_oracle.py:
 *   cv::resize INTER_LINEAR 8U, cv::FAST 9_16 + NMS, cv::GaussianBlur 7x7 s=2 8U,
 *   cv::fastAtan2.
 * Parity status: pinned to cv2 4.13.0 outputs (upstream ships no tests/golden
 * vectors of its own: "parity unpinned by upstream tests").
 *
 * Known non-determinism of the reference that the oracle fixes:
 *   DistributeOctTree sorts (nKeys, node POINTER) (ORBextractor.cc:832); ties on
 *   nKeys are broken by heap address.  The oracle breaks ties by node creation
 *   sequence (a bump allocator's address order).
 *
 * Build: gcc -O3 -march=x86-64-v3 -ffp-contract=off -fPIC -shared (oracle/Makefile).
 * -ffp-contract=off: the float expressions must not be fused into FMAs.
 */
#include <float.h>
#include <limits.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define REF_MAX_LEVELS 32
#define EDGE_THRESHOLD 19
#define HALF_PATCH 15
#define PATCH_SIZE 31

typedef struct {
    float x, y, size, angle, response;
    int octave, class_id;
} ref_keypoint;

typedef struct {
    int x, y, resp;
} ref_cand; /* pre-octree candidate, coordinates relative to the 16-px border */

typedef struct ref_extractor {
    int nfeatures, nlevels, iniTh, minTh;
    double scaleFactor; /* the reference member is double (ORBextractor.h:117) */
    float sf[REF_MAX_LEVELS], isf[REF_MAX_LEVELS], sig2[REF_MAX_LEVELS], isig2[REF_MAX_LEVELS];
    int nfeat[REF_MAX_LEVELS];
    int umax[HALF_PATCH + 1];
    /* per-call state */
    int w[REF_MAX_LEVELS], h[REF_MAX_LEVELS];
    uint8_t *img[REF_MAX_LEVELS];  /* un-blurred pyramid, tight rows (stride = w) */
    uint8_t *blur[REF_MAX_LEVELS]; /* blurred copy */
    ref_cand *cand[REF_MAX_LEVELS];
    int ncand[REF_MAX_LEVELS];
    ref_keypoint *lvl_kp[REF_MAX_LEVELS];
    int lvl_nkp[REF_MAX_LEVELS];
} ref_extractor;

static const signed char k_pattern[256][4] = {
#include "orb_pattern.inc"
};

/* cvRound: round-half-to-even under the default rounding mode */
static inline int cv_round_f(float v) { return (int)lrintf(v); }
static inline int cv_round_d(double v) { return (int)lrint(v); }

/* ------------------------------------------------------------------------- */
/* E1  ORBextractor::ORBextractor            src/ORBextractor.cc:429-534     */
/* ------------------------------------------------------------------------- */
ref_extractor *ref_extractor_create(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh)
{
    if (nlevels < 1 || nlevels > REF_MAX_LEVELS) return NULL;
    ref_extractor *e = (ref_extractor *)calloc(1, sizeof(*e));
    e->nfeatures = nfeatures;
    e->nlevels = nlevels;
    e->iniTh = iniTh;
    e->minTh = minTh;
    e->scaleFactor = scaleFactor;
    e->sf[0] = 1.0f;
    e->sig2[0] = 1.0f;
    for (int i = 1; i < nlevels; i++) {
        e->sf[i] = (float)(e->sf[i - 1] * e->scaleFactor); /* float*double -> double -> float (:450) */
        e->sig2[i] = e->sf[i] * e->sf[i];
    }
    for (int i = 0; i < nlevels; i++) {
        e->isf[i] = 1.0f / e->sf[i];
        e->isig2[i] = 1.0f / e->sig2[i];
    }
    float factor = (float)(1.0f / e->scaleFactor);
    float nDesired = nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nlevels));
    int sum = 0;
    for (int l = 0; l < nlevels - 1; l++) {
        e->nfeat[l] = cv_round_f(nDesired);
        sum += e->nfeat[l];
        nDesired *= factor;
    }
    e->nfeat[nlevels - 1] = nfeatures - sum > 0 ? nfeatures - sum : 0;

    /* umax: row half-widths of the radius-15 disc (:510-533) */
    int v, v0;
    int vmax = (int)floor(HALF_PATCH * sqrt(2.f) / 2 + 1);
    int vmin = (int)ceil(HALF_PATCH * sqrt(2.f) / 2);
    const double hp2 = HALF_PATCH * HALF_PATCH;
    for (v = 0; v <= vmax; ++v) e->umax[v] = cv_round_d(sqrt(hp2 - v * v));
    for (v = HALF_PATCH, v0 = 0; v >= vmin; --v) {
        while (e->umax[v0] == e->umax[v0 + 1]) ++v0;
        e->umax[v] = v0;
        ++v0;
    }
    return e;
}

static void ref_free_call_state(ref_extractor *e)
{
    for (int l = 0; l < REF_MAX_LEVELS; l++) {
        free(e->img[l]);
        free(e->blur[l]);
        free(e->cand[l]);
        free(e->lvl_kp[l]);
        e->img[l] = e->blur[l] = NULL;
        e->cand[l] = NULL;
        e->lvl_kp[l] = NULL;
        e->ncand[l] = e->lvl_nkp[l] = 0;
    }
}

void ref_extractor_destroy(ref_extractor *e)
{
    if (!e) return;
    ref_free_call_state(e);
    free(e);
}

void ref_level_tables(const ref_extractor *e, float *sf, float *isf, float *sig2, float *isig2, int *nfeat, int *umax)
{
    for (int l = 0; l < e->nlevels; l++) {
        if (sf) sf[l] = e->sf[l];
        if (isf) isf[l] = e->isf[l];
        if (sig2) sig2[l] = e->sig2[l];
        if (isig2) isig2[l] = e->isig2[l];
        if (nfeat) nfeat[l] = e->nfeat[l];
    }
    if (umax)
        for (int v = 0; v <= HALF_PATCH; v++) umax[v] = e->umax[v];
}

/* ------------------------------------------------------------------------- */
/* cv::resize(src, dst, sz, 0, 0, INTER_LINEAR), 8UC1   (call site :1366)     */
/* OpenCV fixed-point linear path: 11-bit coefficients (INTER_RESIZE_COEF_SCALE */
/* = 2048), horizontal pass in int32, vertical pass with the >>4, >>16, +2>>2  */
/* cast (VResizeLinear<uchar,int,short,FixedPtCast<int,uchar,22>>).            */
/* ------------------------------------------------------------------------- */
static short sat_short(int v) { return (short)(v < -32768 ? -32768 : (v > 32767 ? 32767 : v)); }

void ref_resize_linear_u8(const uint8_t *src, int sw, int sh, int sstride, uint8_t *dst, int dw, int dh, int dstride)
{
    int *xofs = (int *)malloc(sizeof(int) * dw);
    short *ialpha = (short *)malloc(sizeof(short) * 2 * dw);
    int *H0 = (int *)malloc(sizeof(int) * dw), *H1 = (int *)malloc(sizeof(int) * dw);
    double scale_x = 1. / ((double)dw / sw), scale_y = 1. / ((double)dh / sh);
    for (int dx = 0; dx < dw; dx++) {
        float fx = (float)((dx + 0.5) * scale_x - 0.5);
        int sx = (int)floorf(fx);
        fx -= sx;
        if (sx < 0) { fx = 0; sx = 0; }
        if (sx >= sw - 1) { fx = 0; sx = sw - 1; }
        xofs[dx] = sx;
        ialpha[2 * dx] = sat_short(cv_round_f((1.f - fx) * 2048.f));
        ialpha[2 * dx + 1] = sat_short(cv_round_f(fx * 2048.f));
    }
    int prev0 = -1, prev1 = -1;
    for (int dy = 0; dy < dh; dy++) {
        float fy = (float)((dy + 0.5) * scale_y - 0.5);
        int sy = (int)floorf(fy);
        fy -= sy;
        short b0 = sat_short(cv_round_f((1.f - fy) * 2048.f));
        short b1 = sat_short(cv_round_f(fy * 2048.f));
        int y0 = sy < 0 ? 0 : (sy > sh - 1 ? sh - 1 : sy);
        int y1 = sy + 1 < 0 ? 0 : (sy + 1 > sh - 1 ? sh - 1 : sy + 1);
        /* horizontal pass of the two source rows (re-use row buffers when rows repeat) */
        if (y0 == prev1) { int *t = H0; H0 = H1; H1 = t; prev0 = prev1; prev1 = -1; }
        if (y0 != prev0) {
            const uint8_t *S = src + (size_t)y0 * sstride;
            for (int dx = 0; dx < dw; dx++) {
                int sx = xofs[dx], sx1 = sx + 1 < sw ? sx + 1 : sx;
                H0[dx] = S[sx] * ialpha[2 * dx] + S[sx1] * ialpha[2 * dx + 1];
            }
            prev0 = y0;
        }
        if (y1 != prev1) {
            const uint8_t *S = src + (size_t)y1 * sstride;
            for (int dx = 0; dx < dw; dx++) {
                int sx = xofs[dx], sx1 = sx + 1 < sw ? sx + 1 : sx;
                H1[dx] = S[sx] * ialpha[2 * dx] + S[sx1] * ialpha[2 * dx + 1];
            }
            prev1 = y1;
        }
        uint8_t *D = dst + (size_t)dy * dstride;
        for (int dx = 0; dx < dw; dx++) {
            int v = (((b0 * (H0[dx] >> 4)) >> 16) + ((b1 * (H1[dx] >> 4)) >> 16) + 2) >> 2;
            D[dx] = (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v));
        }
    }
    free(xofs); free(ialpha); free(H0); free(H1);
}

/* ------------------------------------------------------------------------- */
/* cv::GaussianBlur(m, m, Size(7,7), 2, 2, BORDER_REFLECT_101), 8UC1 (:1314)   */
/* OpenCV 4.x fixed-point path: Q8 kernel {18,34,48,56,48,34,18} (sum 256),    */
/* horizontal sums kept in 16 bits, vertical in 32, out = (acc + 2^15) >> 16.  */
/* ------------------------------------------------------------------------- */
static inline int reflect101(int p, int n)
{
    if (n == 1) return 0;
    while (p < 0 || p >= n) {
        if (p < 0) p = -p;
        else p = 2 * (n - 1) - p;
    }
    return p;
}

void ref_gaussian7_u8(const uint8_t *src, int w, int h, int sstride, uint8_t *dst, int dstride)
{
    static const int K[7] = {18, 34, 48, 56, 48, 34, 18};
    uint16_t *tmp = (uint16_t *)malloc(sizeof(uint16_t) * (size_t)w * h);
    for (int y = 0; y < h; y++) {
        const uint8_t *S = src + (size_t)y * sstride;
        uint16_t *T = tmp + (size_t)y * w;
        for (int x = 0; x < w; x++) {
            int acc = 0;
            if (x >= 3 && x < w - 3)
                for (int k = 0; k < 7; k++) acc += K[k] * S[x + k - 3];
            else
                for (int k = 0; k < 7; k++) acc += K[k] * S[reflect101(x + k - 3, w)];
            T[x] = (uint16_t)acc;
        }
    }
    for (int y = 0; y < h; y++) {
        const uint16_t *R[7];
        for (int k = 0; k < 7; k++) R[k] = tmp + (size_t)reflect101(y + k - 3, h) * w;
        uint8_t *D = dst + (size_t)y * dstride;
        for (int x = 0; x < w; x++) {
            uint32_t acc = 0;
            for (int k = 0; k < 7; k++) acc += (uint32_t)K[k] * R[k][x];
            D[x] = (uint8_t)((acc + 32768u) >> 16);
        }
    }
    free(tmp);
}

/* ------------------------------------------------------------------------- */
/* cv::FAST(roi, keys, th, true) -- FAST-9/16 + 3x3 NMS     (call sites :988,993) */
/* Score = largest t at which the pixel is still a corner (cornerScore<16>);   */
/* NMS keeps a corner iff its score is strictly greater than its 8 neighbours, */
/* where non-corners-at-th and pixels outside the ROI interior [3,n-3) score 0. */
/* Output order: row-major.                                                    */
/* ------------------------------------------------------------------------- */
static const int k_ring[16][2] = {{0, 3}, {1, 3}, {2, 2}, {3, 1}, {3, 0}, {3, -1}, {2, -2}, {1, -3},
                                  {0, -3}, {-1, -3}, {-2, -2}, {-3, -1}, {-3, 0}, {-3, 1}, {-2, 2}, {-1, 3}};

/* max over the 16 contiguous 9-arcs of min(d) and of min(-d); returns M (corner at th <=> M > th) */
static int fast_arc_strength(const uint8_t *p, int stride, int th)
{
    int v = p[0], d[25];
    for (int k = 0; k < 16; k++) d[k] = v - p[k_ring[k][1] * stride + k_ring[k][0]];
    /* quick reject: every 9-arc holds k or k+8 for each k */
    int br = 1, dk = 1;
    for (int k = 0; k < 8 && (br | dk); k++) {
        br &= (d[k] > th) | (d[k + 8] > th);
        dk &= (d[k] < -th) | (d[k + 8] < -th);
    }
    if (!(br | dk)) return 0;
    for (int k = 16; k < 25; k++) d[k] = d[k - 16];
    int best = 0;
    for (int k = 0; k < 16; k++) {
        int mn = d[k], mx = d[k];
        for (int j = 1; j < 9; j++) {
            if (d[k + j] < mn) mn = d[k + j];
            if (d[k + j] > mx) mx = d[k + j];
        }
        if (mn > best) best = mn;
        if (-mx > best) best = -mx;
    }
    return best;
}

/* returns count; out[] receives (x, y, score) relative to the ROI origin */
int ref_fast9_nms(const uint8_t *roi, int stride, int w, int h, int th, ref_cand *out, int cap)
{
    if (w < 7 || h < 7) return 0;
    uint8_t *sc = (uint8_t *)calloc((size_t)w * h, 1);
    for (int y = 3; y < h - 3; y++)
        for (int x = 3; x < w - 3; x++) {
            int M = fast_arc_strength(roi + (size_t)y * stride + x, stride, th);
            if (M > th) sc[y * w + x] = (uint8_t)(M - 1);
        }
    int n = 0;
    for (int y = 3; y < h - 3; y++)
        for (int x = 3; x < w - 3; x++) {
            int s = sc[y * w + x];
            if (!s) continue;
            const uint8_t *c = sc + y * w + x;
            if (s > c[-1] && s > c[1] && s > c[-w - 1] && s > c[-w] && s > c[-w + 1] && s > c[w - 1] && s > c[w] &&
                s > c[w + 1]) {
                if (n < cap) { out[n].x = x; out[n].y = y; out[n].resp = s; }
                n++;
            }
        }
    free(sc);
    return n;
}

/* ------------------------------------------------------------------------- */
/* E5  ExtractorNode::DivideNode + ORBextractor::DistributeOctTree             */
/*     src/ORBextractor.cc:551-609, 617-915                                    */
/* std::list is restated as a doubly linked list over a bump pool; the pool    */
/* index is the node's creation sequence (the documented pointer tie-break).   */
/* ------------------------------------------------------------------------- */
typedef struct onode {
    int x0, x1, y0, y1; /* UL.x, UR.x, UL.y, BR.y */
    int *keys, nkeys;   /* indices into the candidate array, original relative order */
    int nomore;
    int prev, next; /* list links (pool indices), -1 = none */
} onode;

typedef struct {
    onode *pool;
    int npool, cap;
    int head, tail, size;
} olist;

static int ol_new(olist *L)
{
    if (L->npool == L->cap) {
        L->cap = L->cap ? L->cap * 2 : 256;
        L->pool = (onode *)realloc(L->pool, sizeof(onode) * L->cap);
    }
    onode *n = &L->pool[L->npool];
    memset(n, 0, sizeof(*n));
    n->prev = n->next = -1;
    return L->npool++;
}
static void ol_push_front(olist *L, int i)
{
    L->pool[i].prev = -1;
    L->pool[i].next = L->head;
    if (L->head >= 0) L->pool[L->head].prev = i; else L->tail = i;
    L->head = i;
    L->size++;
}
static void ol_push_back(olist *L, int i)
{
    L->pool[i].next = -1;
    L->pool[i].prev = L->tail;
    if (L->tail >= 0) L->pool[L->tail].next = i; else L->head = i;
    L->tail = i;
    L->size++;
}
static int ol_erase(olist *L, int i) /* returns next */
{
    onode *n = &L->pool[i];
    int nx = n->next;
    if (n->prev >= 0) L->pool[n->prev].next = n->next; else L->head = n->next;
    if (n->next >= 0) L->pool[n->next].prev = n->prev; else L->tail = n->prev;
    L->size--;
    free(n->keys);
    n->keys = NULL;
    return nx;
}

/* DivideNode (:551-609): children indices returned in c[4] (n1..n4), all allocated */
static void divide_node(olist *L, int pi, const ref_cand *cand, int c[4])
{
    for (int k = 0; k < 4; k++) c[k] = ol_new(L); /* may realloc: take parent pointer afterwards */
    onode *p = &L->pool[pi];
    const int halfX = (int)ceilf((float)(p->x1 - p->x0) / 2);
    const int halfY = (int)ceilf((float)(p->y1 - p->y0) / 2);
    onode *n1 = &L->pool[c[0]], *n2 = &L->pool[c[1]], *n3 = &L->pool[c[2]], *n4 = &L->pool[c[3]];
    n1->x0 = p->x0; n1->x1 = p->x0 + halfX; n1->y0 = p->y0; n1->y1 = p->y0 + halfY;
    n2->x0 = p->x0 + halfX; n2->x1 = p->x1; n2->y0 = p->y0; n2->y1 = p->y0 + halfY;
    n3->x0 = p->x0; n3->x1 = p->x0 + halfX; n3->y0 = p->y0 + halfY; n3->y1 = p->y1;
    n4->x0 = p->x0 + halfX; n4->x1 = p->x1; n4->y0 = p->y0 + halfY; n4->y1 = p->y1;
    for (int k = 0; k < 4; k++) L->pool[c[k]].keys = (int *)malloc(sizeof(int) * (p->nkeys ? p->nkeys : 1));
    for (int i = 0; i < p->nkeys; i++) {
        const ref_cand *kp = &cand[p->keys[i]];
        onode *d;
        if ((float)kp->x < (float)n1->x1) d = ((float)kp->y < (float)n1->y1) ? n1 : n3;
        else d = ((float)kp->y < (float)n1->y1) ? n2 : n4;
        d->keys[d->nkeys++] = p->keys[i];
    }
    for (int k = 0; k < 4; k++)
        if (L->pool[c[k]].nkeys == 1) L->pool[c[k]].nomore = 1;
}

typedef struct { int size, seq; } size_seq;
static int cmp_size_seq(const void *a, const void *b)
{
    const size_seq *p = (const size_seq *)a, *q = (const size_seq *)b;
    if (p->size != q->size) return p->size < q->size ? -1 : 1;
    return p->seq < q->seq ? -1 : (p->seq > q->seq);
}

/* cand coordinates are relative to (minX, minY); returns the number of selected keys,
 * out_idx[] = candidate indices in the reference's output order */
int ref_distribute_octree(const ref_cand *cand, int ncand, int minX, int maxX, int minY, int maxY, int N,
                          int *out_idx, int cap)
{
    olist L;
    memset(&L, 0, sizeof(L));
    L.head = L.tail = -1;
    int nIni = (int)roundf((float)(maxX - minX) / (maxY - minY));
    if (nIni < 1) nIni = 1; /* the reference divides by zero here (portrait images); documented deviation */
    const float hX = (float)(maxX - minX) / nIni;
    int *ini = (int *)malloc(sizeof(int) * nIni);
    for (int i = 0; i < nIni; i++) {
        int ni = ol_new(&L);
        onode *n = &L.pool[ni];
        n->x0 = (int)(hX * (float)i);
        n->x1 = (int)(hX * (float)(i + 1));
        n->y0 = 0;
        n->y1 = maxY - minY;
        n->keys = (int *)malloc(sizeof(int) * (ncand ? ncand : 1));
        ol_push_back(&L, ni);
        ini[i] = ni;
    }
    for (int i = 0; i < ncand; i++) {
        int r = (int)((float)cand[i].x / hX);
        if (r >= nIni) r = nIni - 1; /* out-of-range write in the reference; cannot happen for x < maxX-minX */
        onode *n = &L.pool[ini[r]];
        n->keys[n->nkeys++] = i;
    }
    for (int it = L.head; it >= 0;) {
        onode *n = &L.pool[it];
        if (n->nkeys == 1) { n->nomore = 1; it = n->next; }
        else if (n->nkeys == 0) it = ol_erase(&L, it);
        else it = n->next;
    }
    int finish = 0;
    size_seq *vs = NULL, *vprev = NULL;
    int nvs = 0, capvs = 0;
#define VS_PUSH(sz, id)                                                        \
    do {                                                                       \
        if (nvs == capvs) { capvs = capvs ? capvs * 2 : 256; vs = (size_seq *)realloc(vs, sizeof(size_seq) * capvs); } \
        vs[nvs].size = (sz); vs[nvs].seq = (id); nvs++;                        \
    } while (0)
    while (!finish) {
        int prevSize = L.size;
        int nToExpand = 0;
        nvs = 0;
        for (int it = L.head; it >= 0;) {
            if (L.pool[it].nomore) { it = L.pool[it].next; continue; }
            int c[4];
            divide_node(&L, it, cand, c);
            for (int k = 0; k < 4; k++) {
                if (L.pool[c[k]].nkeys > 0) {
                    ol_push_front(&L, c[k]);
                    if (L.pool[c[k]].nkeys > 1) { nToExpand++; VS_PUSH(L.pool[c[k]].nkeys, c[k]); }
                } else { free(L.pool[c[k]].keys); L.pool[c[k]].keys = NULL; }
            }
            it = ol_erase(&L, it);
        }
        if (L.size >= N || L.size == prevSize) {
            finish = 1;
        } else if (L.size + nToExpand * 3 > N) {
            while (!finish) {
                prevSize = L.size;
                int nprev = nvs;
                vprev = (size_seq *)realloc(vprev, sizeof(size_seq) * (nprev ? nprev : 1));
                memcpy(vprev, vs, sizeof(size_seq) * nprev);
                nvs = 0;
                qsort(vprev, nprev, sizeof(size_seq), cmp_size_seq);
                for (int j = nprev - 1; j >= 0; j--) {
                    int c[4];
                    divide_node(&L, vprev[j].seq, cand, c);
                    for (int k = 0; k < 4; k++) {
                        if (L.pool[c[k]].nkeys > 0) {
                            ol_push_front(&L, c[k]);
                            if (L.pool[c[k]].nkeys > 1) VS_PUSH(L.pool[c[k]].nkeys, c[k]);
                        } else { free(L.pool[c[k]].keys); L.pool[c[k]].keys = NULL; }
                    }
                    ol_erase(&L, vprev[j].seq);
                    if (L.size >= N) break;
                }
                if (L.size >= N || L.size == prevSize) finish = 1;
            }
        }
    }
#undef VS_PUSH
    int nout = 0;
    for (int it = L.head; it >= 0; it = L.pool[it].next) {
        onode *n = &L.pool[it];
        int best = n->keys[0];
        int maxr = cand[best].resp;
        for (int k = 1; k < n->nkeys; k++)
            if (cand[n->keys[k]].resp > maxr) { best = n->keys[k]; maxr = cand[best].resp; }
        if (nout < cap) out_idx[nout] = best;
        nout++;
    }
    for (int i = 0; i < L.npool; i++) free(L.pool[i].keys);
    free(L.pool); free(ini); free(vs); free(vprev);
    return nout;
}

/* ------------------------------------------------------------------------- */
/* cv::fastAtan2(y, x) (OpenCV 4.x scalar path; call site :110)                */
/* degrees in [0,360); 7th-order odd polynomial on the octant, no FMA.         */
/* ------------------------------------------------------------------------- */
float ref_fast_atan2(float y, float x)
{
    const float scale = (float)(180.0 / 3.14159265358979323846);
    const float p1 = 0.9997878412794807f * scale, p3 = -0.3258083974640975f * scale;
    const float p5 = 0.1555786518463281f * scale, p7 = -0.04432655554792128f * scale;
    float ax = fabsf(x), ay = fabsf(y), a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + (float)DBL_EPSILON);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    } else {
        c = ax / (ay + (float)DBL_EPSILON);
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

/* E6  IC_Angle  src/ORBextractor.cc:83-111 (un-blurred level image, integer centre) */
float ref_ic_angle(const uint8_t *img, int stride, int x, int y, const int *umax)
{
    int m01 = 0, m10 = 0;
    const uint8_t *c = img + (size_t)y * stride + x;
    for (int u = -HALF_PATCH; u <= HALF_PATCH; ++u) m10 += u * c[u];
    for (int v = 1; v <= HALF_PATCH; ++v) {
        int vsum = 0, d = umax[v];
        for (int u = -d; u <= d; ++u) {
            int vp = c[u + v * stride], vm = c[u - v * stride];
            vsum += (vp - vm);
            m10 += u * (vp + vm);
        }
        m01 += v * vsum;
    }
    return ref_fast_atan2((float)m01, (float)m10);
}

/* E8  computeOrbDescriptor  src/ORBextractor.cc:116-157 (blurred level image) */
void ref_orb_descriptor(const uint8_t *img, int stride, int x, int y, float angle_deg, uint8_t *desc)
{
    const float factorPI = (float)(3.14159265358979323846 / 180.f);
    float angle = angle_deg * factorPI;
    float a = cosf(angle), b = sinf(angle);
    const uint8_t *c = img + (size_t)y * stride + x;
    for (int i = 0; i < 32; i++) {
        int val = 0;
        for (int k = 0; k < 8; k++) {
            const signed char *p = k_pattern[i * 8 + k];
            int t0 = c[cv_round_f(p[0] * b + p[1] * a) * stride + cv_round_f(p[0] * a - p[1] * b)];
            int t1 = c[cv_round_f(p[2] * b + p[3] * a) * stride + cv_round_f(p[2] * a - p[3] * b)];
            val |= (t0 < t1) << k;
        }
        desc[i] = (uint8_t)val;
    }
}

/* ------------------------------------------------------------------------- */
/* E2  ComputePyramid  src/ORBextractor.cc:1345-1394                           */
/* The 19-px REFLECT_101 border the reference adds is never read by a later    */
/* stage (FAST ROIs, IC_Angle and descriptor samples stay inside the level;    */
/* resize reads the ROI only; blur works on a clone) so it is not materialised. */
/* ------------------------------------------------------------------------- */
static void compute_pyramid(ref_extractor *e, const uint8_t *image, int w, int h, int stride)
{
    for (int l = 0; l < e->nlevels; l++) {
        float scale = e->isf[l];
        int lw = cv_round_f((float)w * scale), lh = cv_round_f((float)h * scale);
        e->w[l] = lw; e->h[l] = lh;
        e->img[l] = (uint8_t *)malloc((size_t)lw * lh + 64);
        if (l == 0)
            for (int y = 0; y < h; y++) memcpy(e->img[0] + (size_t)y * w, image + (size_t)y * stride, w);
        else
            ref_resize_linear_u8(e->img[l - 1], e->w[l - 1], e->h[l - 1], e->w[l - 1], e->img[l], lw, lh, lw);
    }
}

/* E3/E4  ComputeKeyPointsOctTree cell loop  src/ORBextractor.cc:925-1009 */
static void detect_level(ref_extractor *e, int level)
{
    const int W = e->w[level], H = e->h[level];
    const int minBX = EDGE_THRESHOLD - 3, minBY = minBX;
    const int maxBX = W - EDGE_THRESHOLD + 3, maxBY = H - EDGE_THRESHOLD + 3;
    int cap = 4096, n = 0;
    ref_cand *out = (ref_cand *)malloc(sizeof(ref_cand) * cap);
    const float width = (float)(maxBX - minBX), height = (float)(maxBY - minBY);
    const int nCols = (int)(width / 30.f), nRows = (int)(height / 30.f);
    if (nCols < 1 || nRows < 1) { e->cand[level] = out; e->ncand[level] = 0; return; }
    const int wCell = (int)ceilf(width / nCols), hCell = (int)ceilf(height / nRows);
    ref_cand cell[64 * 64];
    for (int i = 0; i < nRows; i++) {
        const float iniY = (float)(minBY + i * hCell);
        float maxY = iniY + hCell + 6;
        if (iniY >= maxBY - 3) continue;
        if (maxY > maxBY) maxY = (float)maxBY;
        for (int j = 0; j < nCols; j++) {
            const float iniX = (float)(minBX + j * wCell);
            float maxX = iniX + wCell + 6;
            if (iniX >= maxBX - 6) continue;
            if (maxX > maxBX) maxX = (float)maxBX;
            const uint8_t *roi = e->img[level] + (size_t)(int)iniY * W + (int)iniX;
            int rw = (int)maxX - (int)iniX, rh = (int)maxY - (int)iniY;
            int nc = ref_fast9_nms(roi, W, rw, rh, e->iniTh, cell, 64 * 64);
            if (nc == 0) nc = ref_fast9_nms(roi, W, rw, rh, e->minTh, cell, 64 * 64);
            for (int k = 0; k < nc; k++) {
                if (n == cap) { cap *= 2; out = (ref_cand *)realloc(out, sizeof(ref_cand) * cap); }
                out[n].x = cell[k].x + j * wCell;
                out[n].y = cell[k].y + i * hCell;
                out[n].resp = cell[k].resp;
                n++;
            }
        }
    }
    e->cand[level] = out;
    e->ncand[level] = n;
}

/* E9  ORBextractor::operator()  src/ORBextractor.cc:1261-1339; returns nkeypoints (or -1 if cap too small) */
int ref_extract(ref_extractor *e, const uint8_t *image, int w, int h, int stride, ref_keypoint *kps, uint8_t *desc,
                int cap)
{
    ref_free_call_state(e);
    if (!image || w <= 0 || h <= 0) return 0;
    compute_pyramid(e, image, w, h, stride);
    for (int l = 0; l < e->nlevels; l++) {
        detect_level(e, l);
        const int W = e->w[l], H = e->h[l];
        const int minB = EDGE_THRESHOLD - 3;
        int outcap = e->ncand[l] + 8;
        int *sel = (int *)malloc(sizeof(int) * outcap);
        int ns = ref_distribute_octree(e->cand[l], e->ncand[l], minB, W - EDGE_THRESHOLD + 3, minB,
                                       H - EDGE_THRESHOLD + 3, e->nfeat[l], sel, outcap);
        const int scaledPatch = (int)(PATCH_SIZE * e->sf[l]);
        e->lvl_kp[l] = (ref_keypoint *)malloc(sizeof(ref_keypoint) * (ns ? ns : 1));
        e->lvl_nkp[l] = ns;
        for (int i = 0; i < ns; i++) {
            ref_keypoint *k = &e->lvl_kp[l][i];
            const ref_cand *c = &e->cand[l][sel[i]];
            k->x = (float)(c->x + minB);
            k->y = (float)(c->y + minB);
            k->response = (float)c->resp;
            k->octave = l;
            k->size = (float)scaledPatch;
            k->class_id = -1;
            k->angle = -1.f;
        }
        free(sel);
    }
    for (int l = 0; l < e->nlevels; l++) /* computeOrientation :1050-1051 */
        for (int i = 0; i < e->lvl_nkp[l]; i++) {
            ref_keypoint *k = &e->lvl_kp[l][i];
            k->angle = ref_ic_angle(e->img[l], e->w[l], cv_round_f(k->x), cv_round_f(k->y), e->umax);
        }
    int total = 0;
    for (int l = 0; l < e->nlevels; l++) total += e->lvl_nkp[l];
    if (total > cap) return -1;
    int off = 0;
    for (int l = 0; l < e->nlevels; l++) {
        int n = e->lvl_nkp[l];
        if (n == 0) continue;
        e->blur[l] = (uint8_t *)malloc((size_t)e->w[l] * e->h[l] + 64);
        ref_gaussian7_u8(e->img[l], e->w[l], e->h[l], e->w[l], e->blur[l], e->w[l]);
        for (int i = 0; i < n; i++) {
            ref_keypoint *k = &e->lvl_kp[l][i];
            ref_orb_descriptor(e->blur[l], e->w[l], cv_round_f(k->x), cv_round_f(k->y), k->angle,
                               desc + (size_t)(off + i) * 32);
        }
        if (l != 0) {
            float scale = e->sf[l];
            for (int i = 0; i < n; i++) { e->lvl_kp[l][i].x *= scale; e->lvl_kp[l][i].y *= scale; }
        }
        memcpy(kps + off, e->lvl_kp[l], sizeof(ref_keypoint) * n);
        off += n;
    }
    return total;
}

/* stage accessors (tests compare the CUDA path stage by stage) */
int ref_level_size(const ref_extractor *e, int l, int *w, int *h) { *w = e->w[l]; *h = e->h[l]; return 0; }
const uint8_t *ref_level_image(const ref_extractor *e, int l) { return e->img[l]; }
const uint8_t *ref_level_blurred(const ref_extractor *e, int l) { return e->blur[l]; }
int ref_level_candidates(const ref_extractor *e, int l, ref_cand *out, int cap)
{
    int n = e->ncand[l] < cap ? e->ncand[l] : cap;
    if (out) memcpy(out, e->cand[l], sizeof(ref_cand) * n);
    return e->ncand[l];
}
int ref_level_nkeypoints(const ref_extractor *e, int l) { return e->lvl_nkp[l]; }

/* ========================================================================= */
/*                                MATCHER                                     */
/* ========================================================================= */
typedef struct { int best_dist, best_idx, second_dist, second_idx; } ref_top2;

/* M1  ORBmatcher::DescriptorDistance  src/ORBmatcher.cc:1733-1749 (SWAR popcount on 8 int32 words) */
int ref_descriptor_distance(const uint8_t *a, const uint8_t *b)
{
    const int32_t *pa = (const int32_t *)a, *pb = (const int32_t *)b;
    int dist = 0;
    for (int i = 0; i < 8; i++, pa++, pb++) {
        unsigned int v = (unsigned int)(*pa ^ *pb);
        v = v - ((v >> 1) & 0x55555555);
        v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
        dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
    }
    return dist;
}

/* same distance with the hardware popcount (labelled separately in the CPU baseline) */
static inline int dist_popcnt(const uint8_t *a, const uint8_t *b)
{
    const uint64_t *pa = (const uint64_t *)a, *pb = (const uint64_t *)b;
    return __builtin_popcountll(pa[0] ^ pb[0]) + __builtin_popcountll(pa[1] ^ pb[1]) +
           __builtin_popcountll(pa[2] ^ pb[2]) + __builtin_popcountll(pa[3] ^ pb[3]);
}

/* the running best / second-best update every Search* shares (ORBmatcher.cc:229-239):
 * strict '<' so the first-seen candidate wins ties; second-best counts duplicates */
static inline void top2_update(ref_top2 *t, int dist, int idx)
{
    if (dist < t->best_dist) {
        t->second_dist = t->best_dist; t->second_idx = t->best_idx;
        t->best_dist = dist; t->best_idx = idx;
    } else if (dist < t->second_dist) {
        t->second_dist = dist; t->second_idx = idx;
    }
}

/* X1  brute-force top-2 of every query against a flat database (loop of ORBmatcher.cc:214-240).
 * use_popcnt=0 -> the reference's bit-hack distance, 1 -> __builtin_popcountll */
void ref_bf_top2(const uint8_t *q, int nq, const uint8_t *db, int ndb, ref_top2 *out, int use_popcnt)
{
    for (int i = 0; i < nq; i++) {
        ref_top2 t = {256, -1, 256, -1};
        const uint8_t *qi = q + (size_t)i * 32;
        if (use_popcnt)
            for (int j = 0; j < ndb; j++) top2_update(&t, dist_popcnt(qi, db + (size_t)j * 32), j);
        else
            for (int j = 0; j < ndb; j++) top2_update(&t, ref_descriptor_distance(qi, db + (size_t)j * 32), j);
        out[i] = t;
    }
}

/* M2-M9 core: per-query candidate lists in CSR form, visited in list order */
void ref_match_candidates(const uint8_t *q, int nq, const uint8_t *t, const int *offs, const int *idx, ref_top2 *out)
{
    for (int i = 0; i < nq; i++) {
        ref_top2 r = {256, -1, 256, -1};
        for (int c = offs[i]; c < offs[i + 1]; c++)
            top2_update(&r, ref_descriptor_distance(q + (size_t)i * 32, t + (size_t)idx[c] * 32), idx[c]);
        out[i] = r;
    }
}

/* ------------------------------------------------------------------------- */
/* M8  the per-query loop of ORBmatcher::SearchForTriangulation                 */
/*     src/ORBmatcher.cc:757-815 with CheckDistEpipolarLine :147-164            */
/* candidates in list order; `dist > bestDist` continues, so a later candidate  */
/* replaces an earlier one at equal distance; vbMatched2 is never set (:725).   */
/* ------------------------------------------------------------------------- */
void ref_search_triangulation(const uint8_t *q, const float *qxy, const uint8_t *qstereo, int nq, const uint8_t *t, const ref_keypoint *tkp,
                              const uint8_t *tstereo, const int *offs, const int *idx, const float *F12, float ex, float ey,
                              const float *sf, const float *sig2, int th_low, int *best_idx2, int *best_dist)
{
    for (int i = 0; i < nq; i++) {
        const float x1 = qxy[2 * i], y1 = qxy[2 * i + 1];
        const float a = x1 * F12[0] + y1 * F12[3] + F12[6];
        const float b = x1 * F12[1] + y1 * F12[4] + F12[7];
        const float c = x1 * F12[2] + y1 * F12[5] + F12[8];
        int bestDist = th_low, bestIdx2 = -1;
        for (int cc = offs[i]; cc < offs[i + 1]; cc++) {
            const int i2 = idx[cc];
            const int dist = ref_descriptor_distance(q + (size_t)i * 32, t + (size_t)i2 * 32);
            if (dist > th_low || dist > bestDist) continue;
            const ref_keypoint *kp2 = tkp + i2;
            if (!qstereo[i] && !tstereo[i2]) {
                const float distex = ex - kp2->x, distey = ey - kp2->y;
                if (distex * distex + distey * distey < 100 * sf[kp2->octave]) continue;
            }
            const float num = a * kp2->x + b * kp2->y + c;
            const float den = a * a + b * b;
            if (den == 0) continue;
            const float dsqr = num * num / den;
            if (dsqr < 3.84 * sig2[kp2->octave]) { bestIdx2 = i2; bestDist = dist; }
        }
        best_idx2[i] = bestIdx2;
        best_dist[i] = bestIdx2 >= 0 ? bestDist : -1;
    }
}

/* accept test of the BoW searches (ORBmatcher.cc:242-244): best <= TH_LOW and best < ratio*second */
int ref_ratio_accept(int best, int second, int th, float ratio)
{
    return best <= th && (float)best < ratio * (float)second;
}

/* MapPoint::ComputeDistinctiveDescriptors  src/MapPoint.cc:266-340, for a batch of map points (group g = the descriptors
 * gathered at :288-294).  Distances[N][N] with a zero diagonal (:306-316), each row sorted (:329), median =
 * vDists[0.5*(N-1)] (:331), strict < keeps the first row (:333-337).  Empty group: early return (:281,296) -> -1. */
static int cmp_int(const void *a, const void *b) { return *(const int *)a - *(const int *)b; }
void ref_distinctive_descriptors(const uint8_t *desc, const int *offsets, int ngroups, int *best_idx, int *best_median)
{
    for (int g = 0; g < ngroups; g++) {
        const uint8_t *d = desc + (size_t)offsets[g] * 32;
        const int N = offsets[g + 1] - offsets[g];
        if (N <= 0) { best_idx[g] = -1; best_median[g] = 0; continue; }
        float *D = (float *)malloc(sizeof(float) * (size_t)N * N); /* float like the reference's Distances[N][N] */
        int *v = (int *)malloc(sizeof(int) * (size_t)N);
        for (int i = 0; i < N; i++) {
            D[(size_t)i * N + i] = 0;
            for (int j = i + 1; j < N; j++) {
                const int dij = ref_descriptor_distance(d + 32 * (size_t)i, d + 32 * (size_t)j);
                D[(size_t)i * N + j] = (float)dij;
                D[(size_t)j * N + i] = (float)dij;
            }
        }
        int BestMedian = 0x7fffffff, BestIdx = 0;
        for (int i = 0; i < N; i++) {
            for (int j = 0; j < N; j++) v[j] = (int)D[(size_t)i * N + j]; /* vector<int> vDists(Distances[i], ...) */
            qsort(v, (size_t)N, sizeof(int), cmp_int);
            const int median = v[(int)(0.5 * (N - 1))];
            if (median < BestMedian) { BestMedian = median; BestIdx = i; }
        }
        best_idx[g] = BestIdx;
        best_median[g] = BestMedian;
        free(D);
        free(v);
    }
}

/* F3  Frame::AssignFeaturesToGrid + PosInGrid  src/Frame.cc:261-279, 444-456.
 * CSR over cells, cell id = ix*rows + iy (mGrid[ix][iy]), insertion (keypoint-index) order inside a cell */
void ref_assign_grid(const ref_keypoint *kp, int n, float minX, float minY, float invW, float invH, int cols, int rows,
                     int *cell_start /* cols*rows+1 */, int *cell_items /* n */)
{
    int nc = cols * rows;
    int *cnt = (int *)calloc(nc + 1, sizeof(int));
    int *cell = (int *)malloc(sizeof(int) * (n ? n : 1));
    for (int i = 0; i < n; i++) {
        int px = (int)roundf((kp[i].x - minX) * invW), py = (int)roundf((kp[i].y - minY) * invH);
        cell[i] = (px < 0 || px >= cols || py < 0 || py >= rows) ? -1 : px * rows + py;
        if (cell[i] >= 0) cnt[cell[i]]++;
    }
    cell_start[0] = 0;
    for (int c = 0; c < nc; c++) cell_start[c + 1] = cell_start[c] + cnt[c];
    memset(cnt, 0, sizeof(int) * nc);
    for (int i = 0; i < n; i++)
        if (cell[i] >= 0) cell_items[cell_start[cell[i]] + cnt[cell[i]]++] = i;
    free(cnt); free(cell);
}

/* F2  Frame::GetFeaturesInArea  src/Frame.cc:386-442; returns count, out[] in the reference's order */
int ref_features_in_area(const ref_keypoint *kp, const int *cell_start, const int *cell_items, float minX, float minY,
                         float invW, float invH, int cols, int rows, float x, float y, float r, int minLevel,
                         int maxLevel, int *out, int cap)
{
    int n = 0;
    int c0 = (int)floorf((x - minX - r) * invW); if (c0 < 0) c0 = 0;
    if (c0 >= cols) return 0;
    int c1 = (int)ceilf((x - minX + r) * invW); if (c1 > cols - 1) c1 = cols - 1;
    if (c1 < 0) return 0;
    int r0 = (int)floorf((y - minY - r) * invH); if (r0 < 0) r0 = 0;
    if (r0 >= rows) return 0;
    int r1 = (int)ceilf((y - minY + r) * invH); if (r1 > rows - 1) r1 = rows - 1;
    if (r1 < 0) return 0;
    const int check = (minLevel > 0) || (maxLevel >= 0);
    for (int ix = c0; ix <= c1; ix++)
        for (int iy = r0; iy <= r1; iy++)
            for (int j = cell_start[ix * rows + iy]; j < cell_start[ix * rows + iy + 1]; j++) {
                const ref_keypoint *k = &kp[cell_items[j]];
                if (check) {
                    if (k->octave < minLevel) continue;
                    if (maxLevel >= 0 && k->octave > maxLevel) continue;
                }
                if (fabsf(k->x - x) < r && fabsf(k->y - y) < r) {
                    if (n < cap) out[n] = cell_items[j];
                    n++;
                }
            }
    return n;
}

/* F1  Frame::ComputeStereoMatches  src/Frame.cc:577-751.
 * pyramids: un-blurred levels, tight rows (stride = lw[l]).  Returns the number of valid depths. */
typedef struct { int dist, idx; } dist_idx;
static int cmp_dist_idx(const void *a, const void *b)
{
    const dist_idx *p = (const dist_idx *)a, *q = (const dist_idx *)b;
    if (p->dist != q->dist) return p->dist < q->dist ? -1 : 1;
    return p->idx < q->idx ? -1 : (p->idx > q->idx);
}

int ref_stereo_match(const ref_keypoint *kL, const uint8_t *dL, int nL, const ref_keypoint *kR, const uint8_t *dR,
                     int nR, const uint8_t *const *pyrL, const uint8_t *const *pyrR, const int *lw, const int *lh,
                     const float *sf, const float *isf, float bf, float b, float *uRight, float *depth,
                     int *best_r /* optional [nL]: Hamming arg-min or -1 */)
{
    const int TH_HIGH = 100, TH_LOW = 50, thOrbDist = (TH_HIGH + TH_LOW) / 2;
    const int nRows = lh[0];
    for (int i = 0; i < nL; i++) { uRight[i] = -1.f; depth[i] = -1.f; if (best_r) best_r[i] = -1; }
    /* row bands (:594-604) as CSR */
    int *rs = (int *)calloc(nRows + 2, sizeof(int));
    for (int iR = 0; iR < nR; iR++) {
        const float r = 2.0f * sf[kR[iR].octave];
        int maxr = (int)ceilf(kR[iR].y + r), minr = (int)floorf(kR[iR].y - r);
        for (int yi = minr; yi <= maxr; yi++) if (yi >= 0 && yi < nRows) rs[yi + 1]++;
    }
    for (int y = 0; y < nRows; y++) rs[y + 1] += rs[y];
    int total = rs[nRows];
    int *items = (int *)malloc(sizeof(int) * (total ? total : 1));
    int *fill = (int *)calloc(nRows, sizeof(int));
    for (int iR = 0; iR < nR; iR++) {
        const float r = 2.0f * sf[kR[iR].octave];
        int maxr = (int)ceilf(kR[iR].y + r), minr = (int)floorf(kR[iR].y - r);
        for (int yi = minr; yi <= maxr; yi++) if (yi >= 0 && yi < nRows) items[rs[yi] + fill[yi]++] = iR;
    }
    const float minZ = b, minD = 0, maxD = bf / minZ;
    dist_idx *vd = (dist_idx *)malloc(sizeof(dist_idx) * (nL ? nL : 1));
    int nvd = 0;
    for (int iL = 0; iL < nL; iL++) {
        const int levelL = kL[iL].octave;
        const float vL = kL[iL].y, uL = kL[iL].x;
        int row = (int)vL;
        if (row < 0 || row >= nRows) continue;
        if (rs[row + 1] == rs[row]) continue;
        const float minU = uL - maxD, maxU = uL - minD;
        if (maxU < 0) continue;
        int bestDist = TH_HIGH, bestIdxR = 0;
        for (int c = rs[row]; c < rs[row + 1]; c++) {
            const int iR = items[c];
            if (kR[iR].octave < levelL - 1 || kR[iR].octave > levelL + 1) continue;
            const float uR = kR[iR].x;
            if (uR >= minU && uR <= maxU) {
                int dist = ref_descriptor_distance(dL + (size_t)iL * 32, dR + (size_t)iR * 32);
                if (dist < bestDist) { bestDist = dist; bestIdxR = iR; }
            }
        }
        if (bestDist < thOrbDist) {
            if (best_r) best_r[iL] = bestIdxR;
            const float uR0 = kR[bestIdxR].x;
            const float scaleFactor = isf[levelL];
            const float scaleduL = roundf(uL * scaleFactor), scaledvL = roundf(vL * scaleFactor);
            const float scaleduR0 = roundf(uR0 * scaleFactor);
            const int w = 5, L = 5;
            const int W = lw[levelL], H = lh[levelL];
            const uint8_t *IL = pyrL[levelL], *IR = pyrR[levelL];
            const int cy = (int)scaledvL, cxl = (int)scaleduL, cxr = (int)scaleduR0;
            const float iniu = scaleduR0 + L - w, endu = scaleduR0 + L + w + 1;
            if (iniu < 0 || endu >= W) continue;
            /* guard the accesses the reference leaves unchecked (cv::Mat ranges would throw) */
            if (cy - w < 0 || cy + w >= H || cxl - w < 0 || cxl + w >= W || cxr - L - w < 0 || cxr + L + w >= W) continue;
            int best = INT_MAX, bestinc = 0;
            float vDists[11];
            const int lc = IL[cy * W + cxl];
            for (int inc = -L; inc <= L; inc++) {
                const int rc = IR[cy * W + cxr + inc];
                float sum = 0.f;
                for (int dy = -w; dy <= w; dy++)
                    for (int dx = -w; dx <= w; dx++) {
                        float l = (float)IL[(cy + dy) * W + cxl + dx] - (float)lc;
                        float r = (float)IR[(cy + dy) * W + cxr + inc + dx] - (float)rc;
                        sum += fabsf(l - r);
                    }
                if (sum < (float)best) { best = (int)sum; bestinc = inc; }
                vDists[L + inc] = sum;
            }
            if (bestinc == -L || bestinc == L) continue;
            const float d1 = vDists[L + bestinc - 1], d2 = vDists[L + bestinc], d3 = vDists[L + bestinc + 1];
            const float deltaR = (d1 - d3) / (2.0f * (d1 + d3 - 2.0f * d2));
            if (deltaR < -1 || deltaR > 1) continue;
            float bestuR = sf[levelL] * ((float)scaleduR0 + (float)bestinc + deltaR);
            float disparity = uL - bestuR;
            if (disparity >= minD && disparity < maxD) {
                if (disparity <= 0) { disparity = 0.01f; bestuR = (float)((double)uL - 0.01); } /* double literal (:725) */
                depth[iL] = bf / disparity;
                uRight[iL] = bestuR;
                vd[nvd].dist = best; vd[nvd].idx = iL; nvd++;
            }
        }
    }
    int nvalid = nvd;
    if (nvd > 0) {
        qsort(vd, nvd, sizeof(dist_idx), cmp_dist_idx);
        const float median = (float)vd[nvd / 2].dist;
        const float thDist = 1.5f * 1.4f * median;
        for (int i = nvd - 1; i >= 0; i--) {
            if ((float)vd[i].dist < thDist) break;
            uRight[vd[i].idx] = -1; depth[vd[i].idx] = -1; nvalid--;
        }
    }
    free(rs); free(items); free(fill); free(vd);
    return nvalid;
}

/* ========================================================================= */
/*            EuRoC prologue: stereo rectification (test infrastructure)      */
/* ========================================================================= */
/* cv::remap(src, dst, map1, map2, INTER_LINEAR) with CV_32FC1 maps, BORDER_CONSTANT 0 -- what Examples/Stereo/stereo_euroc.cc:
 * 131-132 applies to both images of every frame.  OpenCV's fixed-point path: sx = cvRound(map1 * 32), sy likewise (INTER_BITS 5);
 * integer part >> 5 saturated to short, fraction index (sy & 31) * 32 + (sx & 31) into a table of four 15-bit weights
 * w = cvRound(float((1-fy)(1-fx)) * 32768) ... (initInterTab2D; the four always sum to 32768 for the linear kernel);
 * dst = (sum w_i p_i + 16384) >> 15.  Pinned byte for byte against cv2.remap in tests/test_oracle_pin.py. */
static short g_remap_tab[1024][4];
static int g_remap_tab_ready = 0;
const short *ref_remap_table(void)
{
    if (!g_remap_tab_ready) {
        float t1[32][2];
        for (int i = 0; i < 32; i++) {
            const float x = (float)i * (1.0f / 32);
            t1[i][0] = 1.0f - x;
            t1[i][1] = x;
        }
        for (int i = 0; i < 32; i++)
            for (int j = 0; j < 32; j++) {
                const float f[4] = {t1[i][0] * t1[j][0], t1[i][0] * t1[j][1], t1[i][1] * t1[j][0], t1[i][1] * t1[j][1]};
                for (int k = 0; k < 4; k++) g_remap_tab[i * 32 + j][k] = sat_short(cv_round_f(f[k] * 32768.f));
            }
        g_remap_tab_ready = 1;
    }
    return &g_remap_tab[0][0];
}

static inline int remap_px(const uint8_t *src, int sw, int sh, int sstride, int x, int y)
{
    return (x >= 0 && x < sw && y >= 0 && y < sh) ? src[(size_t)y * sstride + x] : 0;
}

void ref_remap_linear_u8(const uint8_t *src, int sw, int sh, int sstride, const float *mapx, const float *mapy, int dw, int dh,
                         uint8_t *dst, int dstride)
{
    const short(*tab)[4] = (const short(*)[4])ref_remap_table();
    for (int y = 0; y < dh; y++)
        for (int x = 0; x < dw; x++) {
            const int sx = cv_round_f(mapx[(size_t)y * dw + x] * 32.f), sy = cv_round_f(mapy[(size_t)y * dw + x] * 32.f);
            int ix = sx >> 5, iy = sy >> 5;
            ix = ix < -32768 ? -32768 : ix > 32767 ? 32767 : ix;
            iy = iy < -32768 ? -32768 : iy > 32767 ? 32767 : iy;
            const short *w = tab[(sy & 31) * 32 + (sx & 31)];
            const int acc = remap_px(src, sw, sh, sstride, ix, iy) * w[0] + remap_px(src, sw, sh, sstride, ix + 1, iy) * w[1] +
                            remap_px(src, sw, sh, sstride, ix, iy + 1) * w[2] + remap_px(src, sw, sh, sstride, ix + 1, iy + 1) * w[3];
            const int v = (acc + 16384) >> 15;
            dst[(size_t)y * dstride + x] = (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v);
        }
}

/* cv::initUndistortRectifyMap(K, D, R, P(3x3), size, CV_32F, map1, map2) (stereo_euroc.cc:96-97), D = (k1, k2, p1, p2, k3):
 * for every destination pixel the ray through the rectified camera is rotated back (inverse of P*R), distorted and projected
 * with K.  Double arithmetic, one narrowing per map entry.  (OpenCV vectorises the row loop; entries may differ from it in the
 * last bit of the double before the narrowing -- the test allows one float ulp on a handful of entries.) */
void ref_init_undistort_rectify_map(const double *K, const double *D, const double *R, const double *P, int w, int h, float *mapx,
                                    float *mapy)
{
    double M[9], ir[9];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) M[3 * i + j] = P[3 * i] * R[j] + P[3 * i + 1] * R[3 + j] + P[3 * i + 2] * R[6 + j];
    const double det = M[0] * (M[4] * M[8] - M[5] * M[7]) - M[1] * (M[3] * M[8] - M[5] * M[6]) + M[2] * (M[3] * M[7] - M[4] * M[6]);
    const double id = 1.0 / det;
    ir[0] = (M[4] * M[8] - M[5] * M[7]) * id; ir[1] = (M[2] * M[7] - M[1] * M[8]) * id; ir[2] = (M[1] * M[5] - M[2] * M[4]) * id;
    ir[3] = (M[5] * M[6] - M[3] * M[8]) * id; ir[4] = (M[0] * M[8] - M[2] * M[6]) * id; ir[5] = (M[2] * M[3] - M[0] * M[5]) * id;
    ir[6] = (M[3] * M[7] - M[4] * M[6]) * id; ir[7] = (M[1] * M[6] - M[0] * M[7]) * id; ir[8] = (M[0] * M[4] - M[1] * M[3]) * id;
    const double fx = K[0], fy = K[4], u0 = K[2], v0 = K[5];
    const double k1 = D[0], k2 = D[1], p1 = D[2], p2 = D[3], k3 = D[4];
    for (int i = 0; i < h; i++) {
        double _x = i * ir[1] + ir[2], _y = i * ir[4] + ir[5], _w = i * ir[7] + ir[8];
        for (int j = 0; j < w; j++, _x += ir[0], _y += ir[3], _w += ir[6]) {
            const double ww = 1. / _w, x = _x * ww, y = _y * ww;
            const double x2 = x * x, y2 = y * y, r2 = x2 + y2, _2xy = 2 * x * y;
            const double kr = (1 + ((k3 * r2 + k2) * r2 + k1) * r2);
            const double xd = x * kr + p1 * _2xy + p2 * (r2 + 2 * x2), yd = y * kr + p1 * (r2 + 2 * y2) + p2 * _2xy;
            mapx[(size_t)i * w + j] = (float)(fx * xd + u0);
            mapy[(size_t)i * w + j] = (float)(fy * yd + v0);
        }
    }
}

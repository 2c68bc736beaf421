/*
 * oracle/orbslam_ref.cpp -- TEST INFRASTRUCTURE: a C-callable wrapper around the REFERENCE's OWN front end.
 *
 * oracle/Makefile compiles /root/reference/src/{ORBextractor,ORBmatcher,Frame,KeyFrame,MapPoint,Map,KeyFrameDatabase}.cc
 * and Thirdparty/DBoW2, unmodified and where they lie, together with this file into oracle/_ref/liborbslam_ref.so
 * (OpenCV stand-in: oracle/shim/, oracle/cv_shim_impl.cpp).  This file contains no algorithm: it constructs the
 * reference's classes, calls their members and copies the results into plain arrays, so that tests/ can pin the
 * restatement (oracle/orb_ref.c) and the CUDA path to the reference's own text.
 *
 *   ORBextractor: constructor tables, operator(), mvImagePyramid, DistributeOctTree        (ORBextractor.cc:430-1394)
 *   Frame: the stereo / monocular constructors (ExtractORB on two threads, UndistortKeyPoints, ComputeStereoMatches,
 *          AssignFeaturesToGrid), GetFeaturesInArea                                         (Frame.cc:65-125, 185-240, 261-751)
 *   MapPoint::ComputeDistinctiveDescriptors over real KeyFrame / MapPoint objects           (MapPoint.cc:266-340)
 *   probes of the OpenCV stand-in's CV_32F algebra (pinned to cv2 by tests/test_ref_pin.py)
 * Protected / private members are reached by deriving from the class or by `#define private public` in THIS
 * translation unit only (access specifiers do not change the object layout); the reference sources are untouched.
 */
/* everything that is not one of the reference's own class headers first, so that the access hack below touches only those */
#include <opencv2/opencv.hpp>
#include <algorithm>
#include <fstream>
#include <list>
#include <map>
#include <mutex>
#include <numeric>
#include <set>
#include <thread>
#include "ORBVocabulary.h"
#include "Thirdparty/DBoW2/DBoW2/BowVector.h"
#include "Thirdparty/DBoW2/DBoW2/FeatureVector.h"
#define private public
#define protected public
#include "Frame.h"
#include "KeyFrame.h"
#include "KeyFrameDatabase.h"
#include "Map.h"
#include "MapPoint.h"
#include "ORBextractor.h"
#include "ORBmatcher.h"
#undef private
#undef protected
#include "Converter.h"

#include <sys/mman.h>
#include <atomic>
#include <new>

using namespace ORB_SLAM2;

/*
 * Allocation order switch.  DistributeOctTree sorts pair<int, ExtractorNode*> (ORBextractor.cc:832): nodes with the
 * same key count are ordered by HEAP ADDRESS, so the reference's output depends on the allocator's reuse pattern and
 * is not a function of the image.  The restatement and the CUDA path break such ties by node creation sequence.  To
 * pin them to the reference's own text, this library can serve small allocations (std::list nodes) from a monotonic
 * arena that never reuses memory: address order == creation order, which is exactly what the reference computes under
 * a bump allocator.  oref_set_monotonic(0) restores plain malloc so that tests can also report how often glibc's
 * reuse pattern changes the outcome.  The operators below replace the global ones for code inside this library only
 * (-Wl,-Bsymbolic); arena blocks are never freed (a few hundred KB per extraction).
 */
static std::atomic<int> g_monotonic(0);
static char *g_arena = 0;
static const size_t kArenaCap = (size_t)8 << 30; /* virtual, MAP_NORESERVE */
static std::atomic<size_t> g_arena_off(0);
static void *oref_alloc(size_t n)
{
    if (g_monotonic.load(std::memory_order_relaxed) && n <= 256 && g_arena) {
        const size_t sz = (n + 15) & ~(size_t)15;
        const size_t off = g_arena_off.fetch_add(sz);
        if (off + sz <= kArenaCap) return g_arena + off;
    }
    void *p = malloc(n ? n : 1);
    if (!p) throw std::bad_alloc();
    return p;
}
static void oref_free(void *p)
{
    if (g_arena && (char *)p >= g_arena && (char *)p < g_arena + kArenaCap) return;
    free(p);
}
void *operator new(size_t n) { return oref_alloc(n); }
void *operator new[](size_t n) { return oref_alloc(n); }
void operator delete(void *p) noexcept { oref_free(p); }
void operator delete[](void *p) noexcept { oref_free(p); }
void operator delete(void *p, size_t) noexcept { oref_free(p); }
void operator delete[](void *p, size_t) noexcept { oref_free(p); }
extern "C" void oref_set_monotonic(int on)
{
    if (on && !g_arena) {
        void *m = mmap(0, kArenaCap, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
        if (m != MAP_FAILED) g_arena = (char *)m;
    }
    g_monotonic.store(on && g_arena ? 1 : 0);
}

/* src/Converter.cc:27-36 (see oracle/shim/Converter.h for why the reference file itself is not compiled) */
std::vector<cv::Mat> Converter::toDescriptorVector(const cv::Mat &Descriptors)
{
    std::vector<cv::Mat> vDesc;
    vDesc.reserve(Descriptors.rows);
    for (int j = 0; j < Descriptors.rows; j++) vDesc.push_back(Descriptors.row(j));
    return vDesc;
}

typedef struct {
    float x, y, size, angle, response;
    int octave, class_id;
} oref_keypoint; /* == cv::KeyPoint, 28 bytes */

static void put_kp(const cv::KeyPoint &k, oref_keypoint *o)
{
    o->x = k.pt.x; o->y = k.pt.y; o->size = k.size; o->angle = k.angle; o->response = k.response; o->octave = k.octave; o->class_id = k.class_id;
}

extern "C" {

/* ---- ORBextractor ------------------------------------------------------------------------------------------------- */
void *oref_extractor_create(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh)
{
    return new ORBextractor(nfeatures, scaleFactor, nlevels, iniTh, minTh);
}
void oref_extractor_destroy(void *h) { delete (ORBextractor *)h; }

void oref_extractor_tables(void *h, float *sf, float *isf, float *sig2, float *isig2, int *nfeat, int *umax)
{
    ORBextractor *e = (ORBextractor *)h;
    const std::vector<float> a = e->GetScaleFactors(), b = e->GetInverseScaleFactors(), c = e->GetScaleSigmaSquares(), d = e->GetInverseScaleSigmaSquares();
    for (int l = 0; l < e->GetLevels(); l++) { sf[l] = a[l]; isf[l] = b[l]; sig2[l] = c[l]; isig2[l] = d[l]; nfeat[l] = e->mnFeaturesPerLevel[l]; }
    for (int i = 0; i < 16; i++) umax[i] = e->umax[i];
}

/* ORBextractor::operator() (ORBextractor.cc:1261-1339); returns the keypoint count (the first `cap` are written) */
int oref_extract(void *h, const unsigned char *img, int w, int hgt, int stride, oref_keypoint *kps, unsigned char *desc, int cap)
{
    ORBextractor *e = (ORBextractor *)h;
    cv::Mat image(hgt, w, CV_8UC1, (void *)img, (size_t)stride);
    std::vector<cv::KeyPoint> keys;
    cv::Mat descriptors;
    (*e)(image, cv::Mat(), keys, descriptors);
    const int n = (int)keys.size();
    for (int i = 0; i < n && i < cap; i++) {
        put_kp(keys[i], kps + i);
        memcpy(desc + (size_t)i * 32, descriptors.ptr(i), 32);
    }
    return n;
}

/* mvImagePyramid[level] after the last operator() call */
int oref_pyramid_level(void *h, int level, int *w, int *hgt, unsigned char *out)
{
    ORBextractor *e = (ORBextractor *)h;
    if (level < 0 || level >= (int)e->mvImagePyramid.size() || e->mvImagePyramid[level].empty()) return -1;
    const cv::Mat &m = e->mvImagePyramid[level];
    *w = m.cols; *hgt = m.rows;
    if (out)
        for (int r = 0; r < m.rows; r++) memcpy(out + (size_t)r * m.cols, m.ptr(r), (size_t)m.cols);
    return 0;
}

/* ORBextractor::DistributeOctTree (ORBextractor.cc:617-915) on a caller-supplied candidate list.
   Output: index of each selected candidate in the input list (recovered through class_id), in result order. */
int oref_distribute_octree(void *h, const int *x, const int *y, const int *resp, int n, int minX, int maxX, int minY, int maxY,
                           int nFeatures, int level, int *out_idx, int cap)
{
    ORBextractor *e = (ORBextractor *)h;
    std::vector<cv::KeyPoint> v(n);
    for (int i = 0; i < n; i++) { v[i] = cv::KeyPoint((float)x[i], (float)y[i], 7.f, -1, (float)resp[i]); v[i].class_id = i; }
    const std::vector<cv::KeyPoint> r = e->DistributeOctTree(v, minX, maxX, minY, maxY, nFeatures, level);
    for (int i = 0; i < (int)r.size() && i < cap; i++) out_idx[i] = r[i].class_id;
    return (int)r.size();
}

/* ---- Frame ---------------------------------------------------------------------------------------------------------- */
struct OrefFrame {
    ORBextractor *eL, *eR;
    Frame *F;
    bool placed;
};

/* Frame::Frame(imLeft, imRight, ...) (Frame.cc:65-125) or, with imR == NULL, the monocular constructor (:185-240).
   cam = fx, fy, cx, cy; dist = k1, k2, p1, p2[, k3] (ndist = 4 | 5).  Image bounds / grid statics are recomputed. */
void *oref_frame_create(const unsigned char *imL, const unsigned char *imR, int w, int hgt, int nfeatures, float scaleFactor, int nlevels,
                        int iniTh, int minTh, const float *cam, const float *dist, int ndist, float bf, float thDepth)
{
    OrefFrame *o = new OrefFrame;
    o->eL = new ORBextractor(nfeatures, scaleFactor, nlevels, iniTh, minTh);
    o->eR = imR ? new ORBextractor(nfeatures, scaleFactor, nlevels, iniTh, minTh) : 0;
    cv::Mat K = cv::Mat::eye(3, 3, CV_32F);
    K.at<float>(0, 0) = cam[0]; K.at<float>(1, 1) = cam[1]; K.at<float>(0, 2) = cam[2]; K.at<float>(1, 2) = cam[3];
    cv::Mat D(ndist, 1, CV_32F);
    for (int i = 0; i < ndist; i++) D.at<float>(i) = dist[i];
    cv::Mat L(hgt, w, CV_8UC1, (void *)imL, (size_t)w);
    Frame::mbInitialComputations = true;
    /* The stereo constructor calls ComputeStereoMatches() (Frame.cc:97) BEFORE it assigns mb = mbf/fx (:118), and
       ComputeStereoMatches reads mb (minZ, :609): an uninitialised read in the reference.  In the running system the Frame is a
       stack temporary of Tracking::GrabImageStereo whose slot still holds the previous frame's mb, i.e. mbf/fx, from the second
       frame on; that intended value is what the restatement and the CUDA path use.  To pin them, the object is constructed in
       storage whose mb bytes already hold mbf/fx (the constructor's init list does not touch mb). */
    void *buf = ::operator new(sizeof(Frame));
    memset(buf, 0, sizeof(Frame));
    const size_t mbOff = (size_t)((char *)&((Frame *)buf)->mb - (char *)buf);
    const float mb = bf / cam[0];
    memcpy((char *)buf + mbOff, &mb, sizeof(float));
    if (imR) {
        cv::Mat R(hgt, w, CV_8UC1, (void *)imR, (size_t)w);
        o->F = new (buf) Frame(L, R, 0.0, o->eL, o->eR, (ORBVocabulary *)0, K, D, bf, thDepth);
    } else
        o->F = new (buf) Frame(L, 0.0, o->eL, (ORBVocabulary *)0, K, D, bf, thDepth);
    o->placed = true;
    return o;
}
void oref_frame_destroy(void *h)
{
    OrefFrame *o = (OrefFrame *)h;
    if (o->placed) { o->F->~Frame(); ::operator delete((void *)o->F); } else delete o->F;
    delete o->eL; delete o->eR; delete o;
}
int oref_frame_n(void *h, int right) { OrefFrame *o = (OrefFrame *)h; return right ? (int)o->F->mvKeysRight.size() : o->F->N; }
/* which: 0 = mvKeys, 1 = mvKeysRight, 2 = mvKeysUn */
void oref_frame_keys(void *h, int which, oref_keypoint *kps, unsigned char *desc)
{
    Frame *F = ((OrefFrame *)h)->F;
    const std::vector<cv::KeyPoint> &v = which == 0 ? F->mvKeys : which == 1 ? F->mvKeysRight : F->mvKeysUn;
    const cv::Mat &d = which == 1 ? F->mDescriptorsRight : F->mDescriptors;
    for (size_t i = 0; i < v.size(); i++) {
        put_kp(v[i], kps + i);
        if (desc) memcpy(desc + i * 32, d.ptr((int)i), 32);
    }
}
void oref_frame_stereo(void *h, float *uRight, float *depth)
{
    Frame *F = ((OrefFrame *)h)->F;
    for (int i = 0; i < F->N; i++) { uRight[i] = F->mvuRight[i]; depth[i] = F->mvDepth[i]; }
}
/* bounds = mnMinX, mnMaxX, mnMinY, mnMaxY, mfGridElementWidthInv, mfGridElementHeightInv (Frame.cc:517-550, :100-103) */
void oref_frame_bounds(float *b)
{
    b[0] = Frame::mnMinX; b[1] = Frame::mnMaxX; b[2] = Frame::mnMinY; b[3] = Frame::mnMaxY;
    b[4] = Frame::mfGridElementWidthInv; b[5] = Frame::mfGridElementHeightInv;
}
/* mGrid flattened cell-major [col][row] (Frame.cc:261-279): start[FRAME_GRID_COLS*FRAME_GRID_ROWS + 1], items[N] */
void oref_frame_grid(void *h, int *start, int *items)
{
    Frame *F = ((OrefFrame *)h)->F;
    int n = 0;
    for (int i = 0; i < FRAME_GRID_COLS; i++)
        for (int j = 0; j < FRAME_GRID_ROWS; j++) {
            start[i * FRAME_GRID_ROWS + j] = n;
            for (size_t k = 0; k < F->mGrid[i][j].size(); k++) items[n++] = (int)F->mGrid[i][j][k];
        }
    start[FRAME_GRID_COLS * FRAME_GRID_ROWS] = n;
}
/* Frame::GetFeaturesInArea (Frame.cc:386-442) */
int oref_frame_features_in_area(void *h, float x, float y, float r, int minLevel, int maxLevel, int *out, int cap)
{
    Frame *F = ((OrefFrame *)h)->F;
    const std::vector<size_t> v = F->GetFeaturesInArea(x, y, r, minLevel, maxLevel);
    for (size_t i = 0; i < v.size() && (int)i < cap; i++) out[i] = (int)v[i];
    return (int)v.size();
}
int oref_frame_pyramid_level(void *h, int right, int level, int *w, int *hgt, unsigned char *out)
{
    OrefFrame *o = (OrefFrame *)h;
    return oref_pyramid_level(right ? o->eR : o->eL, level, w, hgt, out);
}

/* ---- timing sessions (bench.py --impl reference / cpu_baseline kind "reference") -------------------------------------------
   What Tracking::GrabImageStereo does per frame pair with its two persistent extractors (src/Tracking.cc:124-127, 177-213): the
   stereo Frame constructor = ExtractORB of both images on two threads + UndistortKeyPoints + ComputeStereoMatches + grid. */
struct OrefSession {
    ORBextractor *eL, *eR;
    cv::Mat K, D;
    float bf, thDepth;
    void *buf;
};
void *oref_session_create(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh, const float *cam, float bf, float thDepth)
{
    OrefSession *s = new OrefSession;
    s->eL = new ORBextractor(nfeatures, scaleFactor, nlevels, iniTh, minTh);
    s->eR = new ORBextractor(nfeatures, scaleFactor, nlevels, iniTh, minTh);
    s->K = cv::Mat::eye(3, 3, CV_32F);
    s->K.at<float>(0, 0) = cam[0]; s->K.at<float>(1, 1) = cam[1]; s->K.at<float>(0, 2) = cam[2]; s->K.at<float>(1, 2) = cam[3];
    s->D = cv::Mat::zeros(4, 1, CV_32F);
    s->bf = bf; s->thDepth = thDepth;
    s->buf = ::operator new(sizeof(Frame));
    return s;
}
void oref_session_destroy(void *h)
{
    OrefSession *s = (OrefSession *)h;
    delete s->eL; delete s->eR;
    ::operator delete(s->buf);
    delete s;
}
/* one stereo pair (imR != NULL) or one monocular frame; returns N, *nStereo = keypoints with a stereo match */
int oref_session_frame(void *h, const unsigned char *imL, const unsigned char *imR, int w, int hgt, int *nStereo)
{
    OrefSession *s = (OrefSession *)h;
    cv::Mat L(hgt, w, CV_8UC1, (void *)imL, (size_t)w);
    memset(s->buf, 0, sizeof(Frame));
    const float mb = s->bf / s->K.at<float>(0, 0); /* see oref_frame_create: mb is read before the constructor assigns it */
    memcpy((char *)s->buf + (size_t)((char *)&((Frame *)s->buf)->mb - (char *)s->buf), &mb, sizeof(float));
    Frame *F;
    if (imR) {
        cv::Mat R(hgt, w, CV_8UC1, (void *)imR, (size_t)w);
        F = new (s->buf) Frame(L, R, 0.0, s->eL, s->eR, (ORBVocabulary *)0, s->K, s->D, s->bf, s->thDepth);
    } else
        F = new (s->buf) Frame(L, 0.0, s->eL, (ORBVocabulary *)0, s->K, s->D, s->bf, s->thDepth);
    const int n = F->N;
    int ns = 0;
    for (size_t i = 0; i < F->mvuRight.size(); i++) ns += F->mvuRight[i] >= 0;
    if (nStereo) *nStereo = ns;
    F->~Frame();
    return n;
}

/* A Frame built from caller-supplied keypoints / descriptors (no extraction): UndistortKeyPoints is the identity
   (k1 = 0 path is not even entered: mvKeysUn is set directly), AssignFeaturesToGrid is the reference's (Frame.cc:261-279)
   with the bounds supplied by the caller the way ComputeImageBounds leaves them for an undistorted image. */
void *oref_frame_from_keys(const oref_keypoint *kps, const unsigned char *desc, int n, int w, int hgt, float scaleFactor, int nlevels)
{
    OrefFrame *o = new OrefFrame;
    o->placed = false;
    o->eL = new ORBextractor(1000, scaleFactor, nlevels, 20, 7);
    o->eR = 0;
    Frame *F = new Frame();
    o->F = F;
    F->mpORBextractorLeft = o->eL;
    F->N = n;
    F->mvKeys.resize(n);
    for (int i = 0; i < n; i++) {
        F->mvKeys[i] = cv::KeyPoint(kps[i].x, kps[i].y, kps[i].size, kps[i].angle, kps[i].response, kps[i].octave, kps[i].class_id);
    }
    F->mvKeysUn = F->mvKeys;
    F->mDescriptors = cv::Mat(n, 32, CV_8UC1);
    if (n) memcpy(F->mDescriptors.data, desc, (size_t)n * 32);
    F->mnScaleLevels = o->eL->GetLevels();
    F->mfScaleFactor = o->eL->GetScaleFactor();
    F->mfLogScaleFactor = log(F->mfScaleFactor);
    F->mvScaleFactors = o->eL->GetScaleFactors();
    F->mvInvScaleFactors = o->eL->GetInverseScaleFactors();
    F->mvLevelSigma2 = o->eL->GetScaleSigmaSquares();
    F->mvInvLevelSigma2 = o->eL->GetInverseScaleSigmaSquares();
    Frame::mnMinX = 0.0f; Frame::mnMaxX = (float)w; Frame::mnMinY = 0.0f; Frame::mnMaxY = (float)hgt;
    Frame::mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(Frame::mnMaxX - Frame::mnMinX);
    Frame::mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(Frame::mnMaxY - Frame::mnMinY);
    Frame::mbInitialComputations = false;
    F->mvuRight.assign(n, -1.f);
    F->mvDepth.assign(n, -1.f);
    F->mvpMapPoints.assign(n, (MapPoint *)0);
    F->mvbOutlier.assign(n, false);
    F->AssignFeaturesToGrid();
    return o;
}

/* ---- MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:266-340) ----------------------------------------------------
   Every group becomes one MapPoint observed by one KeyFrame per descriptor.  mObservations is a std::map keyed by
   KeyFrame POINTER, so the observation order is the address order: the keyframes of a group are placement-constructed
   in one block, in descriptor order, which makes "first row with the least median" the lowest descriptor index.
   Output: the chosen descriptor (32 bytes) per group; `has[g]` = 0 where the reference returns early (no observations). */
int oref_distinctive_descriptors(const unsigned char *desc, const int *offsets, int ngroups, unsigned char *out, int *has)
{
    Map map;
    Frame F;
    F.N = 1;
    F.mvKeys.assign(1, cv::KeyPoint(10.f, 10.f, 31.f, 0.f, 1.f, 0, -1));
    F.mvKeysUn = F.mvKeys;
    F.mvuRight.assign(1, -1.f);
    F.mvDepth.assign(1, -1.f);
    F.mvpMapPoints.assign(1, (MapPoint *)0);
    F.mvbOutlier.assign(1, false);
    F.mnScaleLevels = 1;
    F.mfScaleFactor = 1.2f; F.mfLogScaleFactor = log(1.2f);
    F.mvScaleFactors.assign(1, 1.f); F.mvInvScaleFactors.assign(1, 1.f); F.mvLevelSigma2.assign(1, 1.f); F.mvInvLevelSigma2.assign(1, 1.f);
    F.mK = cv::Mat::eye(3, 3, CV_32F);
    F.SetPose(cv::Mat::eye(4, 4, CV_32F));
    for (int g = 0; g < ngroups; g++) {
        const int n = offsets[g + 1] - offsets[g];
        KeyFrame *block = n ? (KeyFrame *)::operator new(sizeof(KeyFrame) * (size_t)n) : 0;
        cv::Mat pos = cv::Mat::zeros(3, 1, CV_32F);
        MapPoint *mp = 0;
        for (int i = 0; i < n; i++) {
            F.mDescriptors = cv::Mat(1, 32, CV_8UC1);
            memcpy(F.mDescriptors.data, desc + (size_t)(offsets[g] + i) * 32, 32);
            KeyFrame *kf = new (block + i) KeyFrame(F, &map, (KeyFrameDatabase *)0);
            if (!mp) mp = new MapPoint(pos, kf, &map);
            mp->AddObservation(kf, 0);
        }
        if (!mp) { /* a point nobody observes: the reference returns before touching mDescriptor (MapPoint.cc:301-302) */
            has[g] = 0;
            memset(out + (size_t)g * 32, 0, 32);
            continue;
        }
        mp->ComputeDistinctiveDescriptors();
        const cv::Mat d = mp->GetDescriptor();
        has[g] = d.empty() ? 0 : 1;
        if (!d.empty()) memcpy(out + (size_t)g * 32, d.ptr(), 32);
        delete mp;
        for (int i = 0; i < n; i++) block[i].~KeyFrame();
        ::operator delete(block);
    }
    return 0;
}

/* ---- ORBmatcher::DescriptorDistance (ORBmatcher.cc:1733-1749) --------------------------------------------------------- */
int oref_descriptor_distance(const unsigned char *a, const unsigned char *b)
{
    cv::Mat A(1, 32, CV_8UC1, (void *)a), B(1, 32, CV_8UC1, (void *)b);
    return ORBmatcher::DescriptorDistance(A, B);
}

/* ---- probes of the OpenCV stand-in (pinned to cv2 by tests/test_ref_pin.py) -------------------------------------------- */
void oref_probe_gemm(const float *a, int ar, int ac, const float *b, int br, int bc, double alpha, const float *c, int cr, int cc,
                     double beta, int flags, float *d)
{
    cv::Mat A(ar, ac, CV_32F, (void *)a), B(br, bc, CV_32F, (void *)b), C, D;
    if (c) C = cv::Mat(cr, cc, CV_32F, (void *)c);
    cv::gemm(A, B, alpha, C, beta, D, flags);
    memcpy(d, D.data, sizeof(float) * D.total());
}
double oref_probe_norm(const float *a, int n) { cv::Mat A(n, 1, CV_32F, (void *)a); return cv::norm(A); }
double oref_probe_norm_l1_diff(const float *a, const float *b, int r, int c)
{
    cv::Mat A(r, c, CV_32F, (void *)a), B(r, c, CV_32F, (void *)b);
    return cv::norm(A, B, cv::NORM_L1);
}
/* the expression forms the reference writes: which = 0: R*p + t   1: -R.t()*t   2: M/s   3: s*M   4: (1.0/s)*R.t()   5: -R*t */
void oref_probe_expr(int which, const float *R, const float *p, const float *t, float s, float *out)
{
    cv::Mat mR(3, 3, CV_32F, (void *)R), mp(3, 1, CV_32F, (void *)p), mt(3, 1, CV_32F, (void *)t), r;
    switch (which) {
    case 0: r = mR * mp + mt; break;
    case 1: r = -mR.t() * mt; break;
    case 2: r = mR / s; break;
    case 3: r = s * mR; break;
    case 4: r = (1.0 / s) * mR.t(); break;
    default: r = -mR * mt; break;
    }
    memcpy(out, r.data, sizeof(float) * r.total());
}
void oref_probe_undistort(const float *pts, int n, const float *cam, const float *dist, int ndist, float *out)
{
    cv::Mat K = cv::Mat::eye(3, 3, CV_32F);
    K.at<float>(0, 0) = cam[0]; K.at<float>(1, 1) = cam[1]; K.at<float>(0, 2) = cam[2]; K.at<float>(1, 2) = cam[3];
    cv::Mat D(ndist, 1, CV_32F, (void *)dist);
    cv::Mat m(n, 2, CV_32F);
    memcpy(m.data, pts, sizeof(float) * 2 * (size_t)n);
    m = m.reshape(2);
    cv::undistortPoints(m, m, K, D, cv::Mat(), K);
    m = m.reshape(1);
    memcpy(out, m.data, sizeof(float) * 2 * (size_t)n);
}
} /* extern "C" */

// ref_match.cpp -- TEST DRIVER: the facade ORBmatcher (include/orbb200/ORBmatcher.h, device distances through liborbb200.so)
// against the REFERENCE's own ORBmatcher.cc, on the reference's own Frame / KeyFrame / MapPoint / Map classes.
//
// oracle/_ref/liborbslam_ref.so holds the unmodified /root/reference/src/{ORBmatcher,Frame,KeyFrame,MapPoint,Map,...}.cc
// (oracle/Makefile).  This driver includes the reference's headers, so `ORB_SLAM2::ORBmatcher` below IS the compiled
// reference class; the facade class of the same name is placed in namespace orbb200_facade (-DORBB200_NAMESPACE) and its
// Search* templates bind to the reference's real ORB_SLAM2::Frame / KeyFrame / MapPoint -- the exact situation of a
// drop-in.  OpenCV is the stand-in of oracle/shim (the facade headers take their ORBB200_USE_OPENCV branch).
//
// Every scenario is ONE function templated on the matcher type: it builds a world from a seed (same seed -> identical
// world, different pointers), runs the search and serialises every side effect (match vectors, Frame::mvpMapPoints,
// KeyFrame map-point slots, MapPoint observation counts / bad flags / replacements / descriptors, returned counts) into
// integers with pointers mapped to creation indices.  parity <=> run<reference>(seed) == run<facade>(seed).
//   ref_match.bin            -> one "ok ..." / "FAIL ..." line per scenario, exit code = number of failures
#include <opencv2/opencv.hpp>
#include <algorithm>
#include <cstdio>
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
// scenario set-up writes members the reference keeps protected (MapPoint::mDescriptor, mfMaxDistance, ...); the access
// change is confined to this translation unit and does not alter the object layout
#define private public
#define protected public
#include "Frame.h"
#include "KeyFrame.h"
#include "KeyFrameDatabase.h"
#include "Map.h"
#include "MapPoint.h"
#include "ORBmatcher.h"
#undef private
#undef protected

#define ORBB200_USE_OPENCV
#define ORBB200_NAMESPACE orbb200_facade
#include "orbb200/ORBmatcher.h"

using ORB_SLAM2::Frame;
using ORB_SLAM2::KeyFrame;
using ORB_SLAM2::KeyFrameDatabase;
using ORB_SLAM2::Map;
using ORB_SLAM2::MapPoint;
typedef ORB_SLAM2::ORBmatcher RefMatcher;
typedef orbb200_facade::ORBmatcher FacMatcher;
typedef std::vector<long> Trace;

// ---- deterministic world -------------------------------------------------------------------------------------------------
struct Rng {
    unsigned long long s;
    explicit Rng(unsigned long long seed) : s(seed * 0x9E3779B97F4A7C15ull + 88172645463325252ull) { for (int i = 0; i < 4; i++) next(); }
    unsigned next() { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return (unsigned)(s >> 11); }
    float f() { return (next() & 0xFFFFFF) / (float)0x1000000; }
};

static const int W = 640, H = 480, NLEV = 8;
static const float FX = 718.856f, FY = 718.856f, CX = 320.f, CY = 240.f, BF = 386.1448f;

struct World {
    Rng r;
    Map map;
    ORB_SLAM2::ORBVocabulary voc;
    KeyFrameDatabase *db;
    std::vector<Frame *> frames;
    std::vector<KeyFrame *> kfs;
    std::vector<MapPoint *> mps;
    std::map<MapPoint *, long> idx;
    // MapPoint::mObservations is a std::map keyed by KeyFrame POINTER and ComputeDistinctiveDescriptors (called by Replace) keeps the
    // first of several equally good rows in that order: keyframes are placement-constructed in one block so that their address
    // order is their creation order in both twin worlds
    enum { MAX_KFS = 8 };
    void *kfBlock;
    explicit World(unsigned long long seed) : r(seed), kfBlock(::operator new(sizeof(KeyFrame) * MAX_KFS))
    {
        db = new KeyFrameDatabase(voc);
        Frame::fx = FX; Frame::fy = FY; Frame::cx = CX; Frame::cy = CY; Frame::invfx = 1.0f / FX; Frame::invfy = 1.0f / FY;
        Frame::mnMinX = 0; Frame::mnMaxX = W; Frame::mnMinY = 0; Frame::mnMaxY = H;
        Frame::mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / (Frame::mnMaxX - Frame::mnMinX);
        Frame::mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / (Frame::mnMaxY - Frame::mnMinY);
        Frame::mbInitialComputations = false;
    }
    ~World()
    {
        for (MapPoint *p : mps) delete p;
        for (KeyFrame *k : kfs) k->~KeyFrame();
        ::operator delete(kfBlock);
        for (Frame *f : frames) delete f;
        delete db;
    }
    long id(MapPoint *p) const { if (!p) return -1; std::map<MapPoint *, long>::const_iterator it = idx.find(p); return it == idx.end() ? -2 : it->second; }

    static cv::Mat eye4() { return cv::Mat::eye(4, 4, CV_32F); }
    static cv::Mat pose(float angY, float tx, float ty, float tz, float s = 1.f)
    {
        cv::Mat T = cv::Mat::eye(4, 4, CV_32F);
        const float c = std::cos(angY), sn = std::sin(angY);
        const float R[9] = {c, 0, sn, 0, 1, 0, -sn, 0, c};
        const float t[3] = {tx, ty, tz};
        for (int i = 0; i < 3; i++) { for (int j = 0; j < 3; j++) T.at<float>(i, j) = s * R[3 * i + j]; T.at<float>(i, 3) = s * t[i]; }
        return T;
    }
    // a frame of N random features (positions, octaves, angles, descriptors, some with a stereo coordinate)
    Frame *frame(int N, const cv::Mat &Tcw = cv::Mat())
    {
        Frame *F = new Frame();
        frames.push_back(F);
        F->N = N;
        F->mvKeys.resize(N); F->mvuRight.assign(N, -1.f); F->mvDepth.assign(N, -1.f);
        F->mvpMapPoints.assign(N, (MapPoint *)0); F->mvbOutlier.assign(N, false);
        F->mDescriptors = cv::Mat(N, 32, CV_8UC1);
        F->mnScaleLevels = NLEV; F->mfScaleFactor = 1.2f; F->mfLogScaleFactor = log(1.2f);
        F->mvScaleFactors.assign(NLEV, 1.f); F->mvInvScaleFactors.assign(NLEV, 1.f); F->mvLevelSigma2.assign(NLEV, 1.f); F->mvInvLevelSigma2.assign(NLEV, 1.f);
        for (int l = 1; l < NLEV; l++) { F->mvScaleFactors[l] = F->mvScaleFactors[l - 1] * 1.2f; F->mvLevelSigma2[l] = F->mvScaleFactors[l] * F->mvScaleFactors[l]; }
        for (int l = 0; l < NLEV; l++) { F->mvInvScaleFactors[l] = 1.0f / F->mvScaleFactors[l]; F->mvInvLevelSigma2[l] = 1.0f / F->mvLevelSigma2[l]; }
        F->mbf = BF; F->mb = BF / FX; F->mThDepth = 35.f * F->mb;
        F->mK = cv::Mat::eye(3, 3, CV_32F);
        F->mK.at<float>(0, 0) = FX; F->mK.at<float>(1, 1) = FY; F->mK.at<float>(0, 2) = CX; F->mK.at<float>(1, 2) = CY;
        F->mDistCoef = cv::Mat::zeros(4, 1, CV_32F);
        for (int i = 0; i < N; i++) {
            F->mvKeys[i] = cv::KeyPoint(r.f() * W, r.f() * H, 31, r.f() * 360, 50, r.next() % NLEV);
            for (int k = 0; k < 32; k++) F->mDescriptors.ptr(i)[k] = (unsigned char)r.next();
            if (r.next() % 3 == 0) F->mvuRight[i] = F->mvKeys[i].pt.x - r.f() * 40;
        }
        F->mvKeysUn = F->mvKeys;
        F->SetPose(Tcw.empty() ? eye4() : Tcw);
        return F;
    }
    void regrid(Frame *F)
    {
        for (int i = 0; i < FRAME_GRID_COLS; i++)
            for (int j = 0; j < FRAME_GRID_ROWS; j++) F->mGrid[i][j].clear();
        F->mvKeysUn = F->mvKeys;
        F->AssignFeaturesToGrid(); // Frame.cc:261-279
    }
    KeyFrame *keyframe(Frame *F)
    {
        regrid(F);
        if (kfs.size() >= MAX_KFS) abort();
        KeyFrame *K = new ((KeyFrame *)kfBlock + kfs.size()) KeyFrame(*F, &map, db); // KeyFrame.cc:31-56: copies keys, descriptors, grid, pose
        kfs.push_back(K);
        return K;
    }
    MapPoint *point(KeyFrame *ref, const float *pw = 0)
    {
        cv::Mat pos = cv::Mat::zeros(3, 1, CV_32F);
        if (pw) for (int i = 0; i < 3; i++) pos.at<float>(i) = pw[i];
        MapPoint *p = new MapPoint(pos, ref, &map);
        idx[p] = (long)mps.size();
        mps.push_back(p);
        p->mDescriptor = cv::Mat(1, 32, CV_8UC1);
        for (int k = 0; k < 32; k++) p->mDescriptor.ptr()[k] = (unsigned char)r.next();
        p->mNormalVector = cv::Mat::zeros(3, 1, CV_32F);
        p->mfMaxDistance = 1e9f; p->mfMinDistance = 0.f;
        p->nObs = 1;
        return p;
    }
    void noisy(const unsigned char *src, unsigned char *dst, int flips)
    {
        memcpy(dst, src, 32);
        for (int f = 0; f < flips; f++) { const unsigned b = r.next() % 256; dst[b >> 3] ^= (unsigned char)(1u << (b & 7)); }
    }
    // world point whose camera coordinates under Tcw (rotation R, translation t, no scale) are pc: R^T (pc - t)
    static void world_of(const cv::Mat &T, const float *pc, float *pw)
    {
        for (int i = 0; i < 3; i++) {
            double s = 0;
            for (int k = 0; k < 3; k++) s += (double)T.at<float>(k, i) * ((double)pc[k] - (double)T.at<float>(k, 3));
            pw[i] = (float)s;
        }
    }
    // make p a point that projects near feature src of a frame-like target under the rigid pose T (camera of that target),
    // with a scale band compatible with the feature's octave
    template <class TT> void place_near(MapPoint *p, TT *K, int src, const cv::Mat &T, float jitter, int maxFlips = 70)
    {
        const float z = 3.f + 20.f * r.f();
        const float u = K->mvKeysUn[src].pt.x + (r.f() - 0.5f) * jitter, v = K->mvKeysUn[src].pt.y + (r.f() - 0.5f) * jitter;
        const float pc[3] = {(u - CX) * z / FX, (v - CY) * z / FY, r.next() % 40 ? z : -z};
        float pw[3], O[3], PO[3];
        world_of(T, pc, pw);
        const float zero[3] = {0, 0, 0};
        world_of(T, zero, O);
        for (int i = 0; i < 3; i++) PO[i] = pw[i] - O[i];
        const float nn = std::sqrt(PO[0] * PO[0] + PO[1] * PO[1] + PO[2] * PO[2]);
        for (int i = 0; i < 3; i++) { p->mWorldPos.at<float>(i) = pw[i]; p->mNormalVector.at<float>(i) = (r.next() % 15 ? 1.f : -1.f) * PO[i] / nn; }
        p->mfMaxDistance = nn * std::pow(1.2f, (float)K->mvKeysUn[src].octave) * (0.9f + 0.2f * r.f());
        p->mfMinDistance = p->mfMaxDistance / std::pow(1.2f, 7.f);
        noisy(K->mDescriptors.ptr(src), p->mDescriptor.ptr(), r.next() % maxFlips);
        p->nObs = 1 + r.next() % 5;
        p->mbBad = r.next() % 30 == 0;
    }
    // serialisers
    void put_frame(Trace &t, Frame *F) { for (int i = 0; i < F->N; i++) t.push_back(id(F->mvpMapPoints[i])); }
    void put_kf(Trace &t, KeyFrame *K) { const std::vector<MapPoint *> v = K->GetMapPointMatches(); for (size_t i = 0; i < v.size(); i++) t.push_back(id(v[i])); }
    void put_points(Trace &t)
    {
        for (size_t i = 0; i < mps.size(); i++) {
            MapPoint *p = mps[i];
            t.push_back(p->mbBad); t.push_back(p->nObs); t.push_back(id(p->mpReplaced)); t.push_back((long)p->mObservations.size());
            long h = 0;
            if (!p->mDescriptor.empty()) for (int k = 0; k < 32; k++) h = h * 131 + p->mDescriptor.ptr()[k];
            t.push_back(h);
            t.push_back(p->mnFound); t.push_back(p->mnVisible);
        }
    }
};

// ---- scenarios (each is the call pattern of one caller in Tracking / LocalMapping / LoopClosing) ----------------------------
// M2  SearchByProjection(Frame, MapPoints, th)                       ORBmatcher.cc:48-136, caller Tracking.cc:1463
template <class M> static Trace sc_local_map(unsigned long long seed, float th)
{
    World w(seed);
    Frame *F = w.frame(1500);
    w.regrid(F);
    KeyFrame *K0 = w.keyframe(w.frame(10));
    std::vector<MapPoint *> vp;
    for (int i = 0; i < 1200; i++) {
        MapPoint *m = w.point(K0);
        const int src = w.r.next() % F->N;
        m->mTrackProjX = F->mvKeys[src].pt.x + (w.r.f() - 0.5f) * 4;
        m->mTrackProjY = F->mvKeys[src].pt.y + (w.r.f() - 0.5f) * 4;
        m->mTrackProjXR = F->mvuRight[src] > 0 ? F->mvuRight[src] + (w.r.f() - 0.5f) * 6 : -1;
        m->mnTrackScaleLevel = std::min(7, F->mvKeys[src].octave + (w.r.next() % 4 == 0 ? 1 : 0));
        m->mTrackViewCos = w.r.next() % 2 ? 0.9995f : 0.9f;
        m->mbTrackInView = w.r.next() % 10 != 0;
        m->mbBad = w.r.next() % 25 == 0;
        m->nObs = w.r.next() % 6 == 0 ? 0 : 2;
        if (w.r.next() % 5) w.noisy(F->mDescriptors.ptr(src), m->mDescriptor.ptr(), w.r.next() % 70);
        vp.push_back(m);
    }
    for (size_t i = 0; i + 1 < vp.size(); i += 7) { // contested targets
        vp[i + 1]->mDescriptor = vp[i]->mDescriptor.clone(); vp[i + 1]->mTrackProjX = vp[i]->mTrackProjX; vp[i + 1]->mTrackProjY = vp[i]->mTrackProjY;
    }
    for (int j = 0; j < F->N; j += 9) F->mvpMapPoints[j] = vp[(j * 7) % vp.size()]; // features that already hold a point (with / without observations)
    M matcher(0.8f);
    Trace t;
    t.push_back(matcher.SearchByProjection(*F, vp, th));
    w.put_frame(t, F);
    return t;
}

// M3  SearchByProjection(CurrentFrame, LastFrame, th, bMono)          ORBmatcher.cc:1396-1553, caller Tracking.cc:963, 968
template <class M> static Trace sc_last_frame(unsigned long long seed, int variant, bool bMono)
{
    World w(seed);
    Frame *Last = w.frame(1300);
    const cv::Mat Tc = World::pose(0.01f, 0.05f, 0.f, variant == 1 ? -1.0f : (variant == 2 ? 1.0f : 0.1f));
    Frame *Cur = w.frame(1500, Tc);
    KeyFrame *K0 = w.keyframe(w.frame(10));
    for (int i = 0; i < Last->N; i++) {
        const float z = 4.f + 30.f * w.r.f();
        const float pw[3] = {(Last->mvKeys[i].pt.x - CX) * z / FX, (Last->mvKeys[i].pt.y - CY) * z / FY, z};
        MapPoint *m = w.point(K0, pw);
        memcpy(m->mDescriptor.ptr(), Last->mDescriptors.ptr(i), 32);
        m->nObs = w.r.next() % 5 == 0 ? 0 : 3;
        if (w.r.next() % 8) Last->mvpMapPoints[i] = m;
        Last->mvbOutlier[i] = w.r.next() % 12 == 0;
        const float *T = Tc.ptr<float>();
        const float xc = T[0] * pw[0] + T[1] * pw[1] + T[2] * z + T[3], yc = T[4] * pw[0] + T[5] * pw[1] + T[6] * z + T[7];
        const float zc = T[8] * pw[0] + T[9] * pw[1] + T[10] * z + T[11];
        const int j = i % Cur->N;
        if (zc > 0.1f && w.r.next() % 4) {
            Cur->mvKeys[j].pt.x = FX * xc / zc + CX + (w.r.f() - 0.5f) * 6;
            Cur->mvKeys[j].pt.y = FY * yc / zc + CY + (w.r.f() - 0.5f) * 6;
            Cur->mvKeys[j].octave = std::max(0, std::min(7, Last->mvKeys[i].octave + (int)(w.r.next() % 3) - 1));
            Cur->mvKeys[j].angle = std::fmod(Last->mvKeys[i].angle + (w.r.next() % 8 ? 5.f : 170.f * w.r.f()), 360.f);
            w.noisy(Last->mDescriptors.ptr(i), Cur->mDescriptors.ptr(j), w.r.next() % 80);
            if (Cur->mvuRight[j] > 0) Cur->mvuRight[j] = Cur->mvKeys[j].pt.x - BF / zc + (w.r.f() - 0.5f) * 10;
        }
    }
    w.regrid(Last);
    w.regrid(Cur);
    M matcher(0.9f, true);
    Trace t;
    t.push_back(matcher.SearchByProjection(*Cur, *Last, 15.f, bMono));
    w.put_frame(t, Cur);
    return t;
}

// M4  SearchByProjection(CurrentFrame, KeyFrame, sAlreadyFound, th, ORBdist)   ORBmatcher.cc:1555-1682, caller Tracking.cc:1735, 1758
template <class M> static Trace sc_reloc(unsigned long long seed, float th, int ORBdist)
{
    World w(seed);
    const cv::Mat Tc = World::pose(0.02f, 0.08f, -0.02f, 0.15f);
    Frame *C = w.frame(1500, Tc);
    w.regrid(C);
    KeyFrame *K = w.keyframe(w.frame(1200));
    std::set<MapPoint *> found;
    for (int i = 0; i < K->N; i++) {
        MapPoint *m = w.point(K);
        w.place_near(m, C, w.r.next() % C->N, Tc, 4.f);
        if (w.r.next() % 6) K->AddMapPoint(m, i);
        if (w.r.next() % 12 == 0) found.insert(m);
    }
    for (int j = 0; j < C->N; j++) if (w.r.next() % 4 == 0) C->mvpMapPoints[j] = w.mps[w.r.next() % w.mps.size()]; // already claimed
    M matcher(0.9f, true);
    Trace t;
    t.push_back(matcher.SearchByProjection(*C, K, found, th, ORBdist));
    w.put_frame(t, C);
    return t;
}

// M5 + Fuse(Scw)  SearchByProjection(KeyFrame, Scw, points, matched, th) / Fuse(KeyFrame, Scw, points, th, replace)
//                 ORBmatcher.cc:305-428, 1040-1164; callers LoopClosing.cc:378, 727
template <class M> static Trace sc_sim3_projection(unsigned long long seed, bool fuse)
{
    World w(seed);
    KeyFrame *K = w.keyframe(w.frame(1500));
    const float s = 1.07f;
    const cv::Mat Scw = World::pose(-0.03f, 0.3f, 0.05f, -0.2f, s);
    const cv::Mat Tcw = World::pose(-0.03f, 0.3f, 0.05f, -0.2f, 1.f); // the rigid part: Rcw = sR/s, tcw = st/s
    const int NP = 1400;
    std::vector<MapPoint *> vp;
    for (int i = 0; i < NP; i++) { MapPoint *m = w.point(K); w.place_near(m, K, w.r.next() % K->N, Tcw, 3.f); vp.push_back(m); }
    for (int i = 5; i < NP; i += 40) vp[i] = vp[i - 3];
    Trace t;
    if (!fuse) {
        std::vector<MapPoint *> matched(K->N, (MapPoint *)0);
        for (int j = 0; j < K->N; j++) if (w.r.next() % 5 == 0) matched[j] = vp[w.r.next() % NP];
        M matcher(0.75f, true);
        t.push_back(matcher.SearchByProjection(K, Scw, vp, matched, 10));
        for (size_t j = 0; j < matched.size(); j++) t.push_back(w.id(matched[j]));
    } else {
        for (int j = 0; j < K->N; j++)
            if (w.r.next() % 3 == 0) { MapPoint *m = w.point(K); m->nObs = 1 + w.r.next() % 5; m->mbBad = w.r.next() % 20 == 0; K->AddMapPoint(m, j); m->mObservations[K] = j; }
        for (int j = 0; j < K->N; j += 17) if (K->GetMapPoint(j)) vp.push_back(K->GetMapPoint(j)); // already in the keyframe
        std::vector<MapPoint *> rep(vp.size(), (MapPoint *)0);
        M matcher(0.8f);
        t.push_back(matcher.Fuse(K, Scw, vp, 4.f, rep));
        for (size_t j = 0; j < rep.size(); j++) t.push_back(w.id(rep[j]));
        w.put_kf(t, K);
        w.put_points(t);
    }
    return t;
}

// M6  SearchByBoW(KeyFrame, Frame, matches)                           ORBmatcher.cc:171-303, callers Tracking.cc:907, 1682
template <class M> static Trace sc_bow_kf_frame(unsigned long long seed)
{
    World w(seed);
    Frame *F = w.frame(1400);
    Frame *tmp = w.frame(1300);
    for (int i = 0; i < tmp->N; i++) {
        const int src = w.r.next() % F->N;
        tmp->mvKeys[i] = F->mvKeys[src];
        tmp->mvKeys[i].angle = std::fmod(F->mvKeys[src].angle + (w.r.next() % 8 ? 12.f : 200.f * w.r.f()), 360.f);
        w.noisy(F->mDescriptors.ptr(src), tmp->mDescriptors.ptr(i), w.r.next() % 60);
        const unsigned node = (unsigned)(src % 97) * 3; // same vocabulary node as its source most of the time
        tmp->mFeatVec.addFeature(w.r.next() % 9 ? node : node + 1, i);
    }
    for (int i = 0; i < F->N; i++) F->mFeatVec.addFeature((unsigned)(i % 97) * 3, i);
    w.regrid(F);
    KeyFrame *K = w.keyframe(tmp);
    for (int i = 0; i < K->N; i++)
        if (w.r.next() % 6) { MapPoint *m = w.point(K); m->mbBad = w.r.next() % 30 == 0; K->AddMapPoint(m, i); }
    std::vector<MapPoint *> got;
    M matcher(0.7f, true);
    Trace t;
    t.push_back(matcher.SearchByBoW(K, *F, got));
    t.push_back((long)got.size());
    for (size_t i = 0; i < got.size(); i++) t.push_back(w.id(got[i]));
    return t;
}

// M6b SearchByBoW(KeyFrame, KeyFrame, matches12)                      ORBmatcher.cc:558-702, caller LoopClosing.cc:280
template <class M> static Trace sc_bow_kf_kf(unsigned long long seed)
{
    World w(seed);
    Frame *f1 = w.frame(1300), *f2 = w.frame(1300);
    for (int i = 0; i < f1->N; i++) {
        f2->mvKeys[i] = f1->mvKeys[i];
        f2->mvKeys[i].angle = std::fmod(f1->mvKeys[i].angle + (w.r.next() % 8 ? 6.f : 180.f * w.r.f()), 360.f);
        w.noisy(f1->mDescriptors.ptr(i), f2->mDescriptors.ptr(i), w.r.next() % 64);
        const unsigned node = (unsigned)(i % 61) * 2;
        f1->mFeatVec.addFeature(node, i);
        f2->mFeatVec.addFeature(w.r.next() % 10 ? node : node + 2, i);
    }
    for (int i = 0; i + 61 < f1->N; i += 7) memcpy(f2->mDescriptors.ptr(i + 61), f2->mDescriptors.ptr(i), 32); // equal distances inside one node
    KeyFrame *K1 = w.keyframe(f1), *K2 = w.keyframe(f2);
    for (int i = 0; i < K1->N; i++) {
        if (w.r.next() % 5) { MapPoint *m = w.point(K1); m->mbBad = w.r.next() % 25 == 0; K1->AddMapPoint(m, i); }
        if (w.r.next() % 5) { MapPoint *m = w.point(K2); m->mbBad = w.r.next() % 25 == 0; K2->AddMapPoint(m, i); }
    }
    std::vector<MapPoint *> got;
    M matcher(0.75f, true);
    Trace t;
    t.push_back(matcher.SearchByBoW(K1, K2, got));
    t.push_back((long)got.size());
    for (size_t i = 0; i < got.size(); i++) t.push_back(w.id(got[i]));
    return t;
}

// M7  SearchForInitialization(F1, F2, prevMatched, matches12, window)  ORBmatcher.cc:430-556, caller Tracking.cc:642
template <class M> static Trace sc_initialization(unsigned long long seed)
{
    World w(seed);
    Frame *F1 = w.frame(1600), *F2 = w.frame(1600);
    for (int i = 0; i < F1->N; i++) {
        F1->mvKeys[i].octave = w.r.next() % 3 ? 0 : (int)(w.r.next() % 4);
        const int j = (i * 7) % F2->N;
        if (w.r.next() % 5) {
            F2->mvKeys[j].pt.x = F1->mvKeys[i].pt.x + (w.r.f() - 0.5f) * 60;
            F2->mvKeys[j].pt.y = F1->mvKeys[i].pt.y + (w.r.f() - 0.5f) * 60;
            F2->mvKeys[j].octave = w.r.next() % 6 ? 0 : 1;
            F2->mvKeys[j].angle = std::fmod(F1->mvKeys[i].angle + (w.r.next() % 8 ? 3.f : 200.f * w.r.f()), 360.f);
            w.noisy(F1->mDescriptors.ptr(i), F2->mDescriptors.ptr(j), w.r.next() % 50);
        }
    }
    for (int i = 0; i + 3 < F1->N; i += 11) { memcpy(F1->mDescriptors.ptr(i + 3), F1->mDescriptors.ptr(i), 32); F1->mvKeys[i + 3].pt = F1->mvKeys[i].pt; F1->mvKeys[i + 3].octave = F1->mvKeys[i].octave; }
    for (int j = 0; j < F2->N; j++) { // keep F2 inside the image so that PosInGrid accepts every feature
        F2->mvKeys[j].pt.x = std::min(std::max(F2->mvKeys[j].pt.x, 0.f), W - 1.f);
        F2->mvKeys[j].pt.y = std::min(std::max(F2->mvKeys[j].pt.y, 0.f), H - 1.f);
    }
    w.regrid(F1);
    w.regrid(F2);
    std::vector<cv::Point2f> prev(F1->N);
    for (int i = 0; i < F1->N; i++) prev[i] = F1->mvKeysUn[i].pt;
    std::vector<int> m12;
    M matcher(0.9f, true);
    Trace t;
    t.push_back(matcher.SearchForInitialization(*F1, *F2, prev, m12, 100));
    for (size_t i = 0; i < m12.size(); i++) t.push_back(m12[i]);
    for (size_t i = 0; i < prev.size(); i++) { t.push_back((long)(prev[i].x * 64)); t.push_back((long)(prev[i].y * 64)); }
    return t;
}

// M8  SearchForTriangulation(KF1, KF2, F12, pairs, bOnlyStereo)        ORBmatcher.cc:704-871, caller LocalMapping.cc:295
template <class M> static Trace sc_triangulation(unsigned long long seed, bool onlyStereo, bool checkOri)
{
    World w(seed);
    Frame *f1 = w.frame(1400);
    Frame *f2 = w.frame(1400, World::pose(0.f, -0.5f, 0.f, 0.3f));
    for (int i = 0; i < f1->N; i++) {
        f2->mvKeys[i] = f1->mvKeys[i];
        f2->mvKeys[i].pt.x = std::min(std::max(f1->mvKeys[i].pt.x + (w.r.f() - 0.5f) * 80, 0.f), W - 1.f);
        f2->mvKeys[i].pt.y = std::min(std::max(f1->mvKeys[i].pt.y + (w.r.next() % 4 ? (w.r.f() - 0.5f) * 2.f : (w.r.f() - 0.5f) * 30.f), 0.f), H - 1.f);
        f2->mvKeys[i].angle = std::fmod(f1->mvKeys[i].angle + (w.r.next() % 8 ? 4.f : 150.f * w.r.f()), 360.f);
        w.noisy(f1->mDescriptors.ptr(i), f2->mDescriptors.ptr(i), w.r.next() % 70);
        f1->mvuRight[i] = w.r.next() % 3 == 0 ? 10.f : -1.f;
        f2->mvuRight[i] = w.r.next() % 3 == 0 ? 10.f : -1.f;
        const unsigned node = (unsigned)(i % 53) * 2;
        f1->mFeatVec.addFeature(node, i);
        f2->mFeatVec.addFeature(w.r.next() % 12 ? node : node + 2, i);
    }
    for (int i = 0; i + 53 < f1->N; i += 9) { memcpy(f2->mDescriptors.ptr(i + 53), f2->mDescriptors.ptr(i), 32); f2->mvKeys[i + 53].pt.y = f2->mvKeys[i].pt.y; }
    KeyFrame *K1 = w.keyframe(f1), *K2 = w.keyframe(f2);
    for (int i = 0; i < K1->N; i++) {
        if (w.r.next() % 10 == 0) K1->AddMapPoint(w.point(K1), i);
        if (w.r.next() % 10 == 0) K2->AddMapPoint(w.point(K2), i);
    }
    cv::Mat F12 = cv::Mat::zeros(3, 3, CV_32F); // epipolar lines = image rows (pure x translation would give this; exercised as a gate)
    F12.at<float>(1, 2) = -1.f; F12.at<float>(2, 1) = 1.f;
    std::vector<std::pair<size_t, size_t> > pairs;
    M matcher(0.6f, checkOri);
    Trace t;
    t.push_back(matcher.SearchForTriangulation(K1, K2, F12, pairs, onlyStereo));
    for (size_t i = 0; i < pairs.size(); i++) { t.push_back((long)pairs[i].first); t.push_back((long)pairs[i].second); }
    return t;
}

// M9  Fuse(KeyFrame, MapPoints, th)                                   ORBmatcher.cc:873-1038, callers LocalMapping.cc:619, 639
template <class M> static Trace sc_fuse(unsigned long long seed)
{
    World w(seed);
    const cv::Mat Tk = World::pose(0.015f, 0.1f, -0.04f, 0.2f);
    KeyFrame *K = w.keyframe(w.frame(1500, Tk));
    KeyFrame *Other = w.keyframe(w.frame(1500)); // the neighbour keyframe whose points are fused into K
    const int NP = 1300;
    std::vector<MapPoint *> vp;
    for (int i = 0; i < NP; i++) {
        MapPoint *m = w.point(Other);
        w.place_near(m, K, w.r.next() % K->N, Tk, 3.f);
        const size_t oi = (size_t)(w.r.next() % Other->N);
        if (!Other->GetMapPoint(oi)) { Other->AddMapPoint(m, oi); m->mObservations[Other] = oi; } // observed elsewhere: Replace() walks this
        vp.push_back(w.r.next() % 50 ? m : (MapPoint *)0);
    }
    for (int i = 5; i < NP; i += 40) vp[i] = vp[i - 3]; // duplicates: IsInKeyFrame after the first fuse
    for (int j = 0; j < K->N; j++)
        if (w.r.next() % 3 == 0) { // keyframe features that already hold a point (-> Replace either way)
            MapPoint *m = w.point(K);
            memcpy(m->mDescriptor.ptr(), K->mDescriptors.ptr(j), 32);
            m->nObs = 1 + w.r.next() % 5; m->mbBad = w.r.next() % 20 == 0;
            K->AddMapPoint(m, j); m->mObservations[K] = j;
        }
    for (int j = 0; j < K->N; j += 17) if (K->GetMapPoint(j)) vp.push_back(K->GetMapPoint(j)); // points that get replaced appear later in the list
    M matcher;
    Trace t;
    t.push_back(matcher.Fuse(K, vp, 3.0f));
    w.put_kf(t, K);
    w.put_kf(t, Other);
    w.put_points(t);
    return t;
}

// SearchBySim3(KF1, KF2, matches12, s12, R12, t12, th)                ORBmatcher.cc:1166-1394, caller LoopClosing.cc:319
template <class M> static Trace sc_sim3(unsigned long long seed)
{
    World w(seed);
    const float s12 = 1.04f;
    const cv::Mat T12 = World::pose(0.025f, 0.25f, -0.03f, 0.1f);
    const cv::Mat T2w = World::pose(-0.024f, -0.24f, 0.03f, -0.09f); // KF2's own pose: close to, not equal to, the inverse Sim3
    Frame *f1 = w.frame(1400), *f2 = w.frame(1400, T2w);
    cv::Mat R12 = T12.rowRange(0, 3).colRange(0, 3).clone(), t12 = T12.rowRange(0, 3).col(3).clone();
    cv::Mat sR21 = (1.0 / s12) * R12.t();
    cv::Mat t21 = -sR21 * t12;
    struct P { float pc1[3], pc2[3]; };
    std::vector<P> pts(f1->N);
    for (int i = 0; i < f1->N; i++) {
        f2->mvKeys[i] = f1->mvKeys[i];
        const float z = 4.f + 20.f * w.r.f();
        P &p = pts[i];
        p.pc1[0] = (f1->mvKeys[i].pt.x - CX) * z / FX; p.pc1[1] = (f1->mvKeys[i].pt.y - CY) * z / FY; p.pc1[2] = z;
        for (int rr = 0; rr < 3; rr++) p.pc2[rr] = sR21.at<float>(rr, 0) * p.pc1[0] + sR21.at<float>(rr, 1) * p.pc1[1] + sR21.at<float>(rr, 2) * p.pc1[2] + t21.at<float>(rr);
        const float u2 = FX * p.pc2[0] / p.pc2[2] + CX, v2 = FY * p.pc2[1] / p.pc2[2] + CY;
        if (w.r.next() % 6 && u2 > 1 && u2 < 639 && v2 > 1 && v2 < 479) { f2->mvKeys[i].pt.x = u2 + (w.r.f() - 0.5f); f2->mvKeys[i].pt.y = v2 + (w.r.f() - 0.5f); }
        w.noisy(f1->mDescriptors.ptr(i), f2->mDescriptors.ptr(i), w.r.next() % 80);
    }
    KeyFrame *K1 = w.keyframe(f1), *K2 = w.keyframe(f2);
    for (int i = 0; i < K1->N; i++) {
        const P &p = pts[i];
        MapPoint *a = w.point(K1, p.pc1); // world == camera 1
        float pw2[3];
        World::world_of(T2w, p.pc2, pw2);
        MapPoint *b = w.point(K2, pw2);
        const float n2 = std::sqrt(p.pc2[0] * p.pc2[0] + p.pc2[1] * p.pc2[1] + p.pc2[2] * p.pc2[2]), n1 = std::sqrt(p.pc1[0] * p.pc1[0] + p.pc1[1] * p.pc1[1] + p.pc1[2] * p.pc1[2]);
        a->mfMaxDistance = n2 * std::pow(1.2f, (float)K2->mvKeysUn[i].octave) * (0.93f + 0.14f * w.r.f());
        b->mfMaxDistance = n1 * std::pow(1.2f, (float)K1->mvKeysUn[i].octave) * (0.93f + 0.14f * w.r.f());
        a->mfMinDistance = a->mfMaxDistance / std::pow(1.2f, 7.f); b->mfMinDistance = b->mfMaxDistance / std::pow(1.2f, 7.f);
        memcpy(a->mDescriptor.ptr(), K1->mDescriptors.ptr(i), 32); memcpy(b->mDescriptor.ptr(), K2->mDescriptors.ptr(i), 32);
        a->mbBad = w.r.next() % 30 == 0; b->mbBad = w.r.next() % 30 == 0;
        if (w.r.next() % 8) K1->AddMapPoint(a, i);
        if (w.r.next() % 8) { K2->AddMapPoint(b, i); b->mObservations[K2] = i; }
    }
    std::vector<MapPoint *> m12(K1->N, (MapPoint *)0);
    for (int i = 0; i < K1->N; i += 9) if (K2->GetMapPoint(i)) m12[i] = K2->GetMapPoint(i); // matches already known (from SearchByBoW)
    M matcher(0.75f, true);
    Trace t;
    t.push_back(matcher.SearchBySim3(K1, K2, m12, s12, R12, t12, 7.5f));
    for (size_t i = 0; i < m12.size(); i++) t.push_back(w.id(m12[i]));
    return t;
}

// M1  DescriptorDistance                                               ORBmatcher.cc:1733-1749
template <class M> static Trace sc_distance(unsigned long long seed)
{
    Rng r(seed);
    cv::Mat a(1, 32, CV_8UC1), b(1, 32, CV_8UC1);
    Trace t;
    for (int it = 0; it < 2000; it++) {
        for (int k = 0; k < 32; k++) { a.ptr()[k] = (unsigned char)r.next(); b.ptr()[k] = it % 5 ? (unsigned char)r.next() : a.ptr()[k]; }
        t.push_back(M::DescriptorDistance(a, b));
    }
    return t;
}

static int g_fails = 0;
static bool g_refOnly = false; // --ref-only: run just the compiled reference (no GPU needed): checks that every scenario is strong
static void report(const char *name, const Trace &ref, const Trace &fac, long minCount)
{
    size_t firstDiff = 0;
    const bool same = ref == fac;
    if (!same) while (firstDiff < ref.size() && firstDiff < fac.size() && ref[firstDiff] == fac[firstDiff]) firstDiff++;
    const bool enough = !ref.empty() && ref[0] >= minCount;
    printf("%s %s: reference %ld, facade %ld", same && enough ? "ok" : "FAIL", name, ref.empty() ? -1 : ref[0], fac.empty() ? -1 : fac[0]);
    if (!same) printf(" (traces differ at %zu of %zu / %zu)", firstDiff, ref.size(), fac.size());
    if (!enough) printf(" (scenario too weak: needs >= %ld)", minCount);
    printf("\n");
    g_fails += !(same && enough);
}
#define RUN(name, minCount, call_ref, call_fac)                                  \
    do {                                                                         \
        const Trace r_ = call_ref;                                               \
        if (g_refOnly) report(name, r_, r_, minCount); else report(name, r_, call_fac, minCount); \
    } while (0)


int main(int argc, char **argv)
{
    int nseeds = 2;
    for (int i = 1; i < argc; i++) { if (!strcmp(argv[i], "--ref-only")) g_refOnly = true; else nseeds = atoi(argv[i]); }
    char nm[160];
    for (int sd = 1; sd <= nseeds; sd++) {
        const unsigned long long s = (unsigned long long)sd;
        snprintf(nm, sizeof nm, "SearchByProjection(frame, local map) seed %d", sd);
        RUN(nm, 300, sc_local_map<RefMatcher>(s, 3.f), sc_local_map<FacMatcher>(s, 3.f));
        snprintf(nm, sizeof nm, "SearchByProjection(frame, local map, th=1) seed %d", sd);
        RUN(nm, 100, sc_local_map<RefMatcher>(s + 100, 1.f), sc_local_map<FacMatcher>(s + 100, 1.f));
        for (int v = 0; v < 3; v++) {
            snprintf(nm, sizeof nm, "SearchByProjection(current, last; %s) seed %d", v == 0 ? "sideways" : v == 1 ? "forward" : "backward", sd);
            RUN(nm, 150, sc_last_frame<RefMatcher>(s, v, false), sc_last_frame<FacMatcher>(s, v, false));
        }
        snprintf(nm, sizeof nm, "SearchByProjection(current, last; mono) seed %d", sd);
        RUN(nm, 150, sc_last_frame<RefMatcher>(s, 1, true), sc_last_frame<FacMatcher>(s, 1, true));
        snprintf(nm, sizeof nm, "SearchByProjection(frame, keyframe; th 10, dist 100) seed %d", sd);
        RUN(nm, 20, sc_reloc<RefMatcher>(s, 10.f, 100), sc_reloc<FacMatcher>(s, 10.f, 100));
        snprintf(nm, sizeof nm, "SearchByProjection(frame, keyframe; th 3, dist 40) seed %d", sd);
        RUN(nm, 20, sc_reloc<RefMatcher>(s + 50, 3.f, 40), sc_reloc<FacMatcher>(s + 50, 3.f, 40));
        snprintf(nm, sizeof nm, "SearchByProjection(keyframe, Scw) seed %d", sd);
        RUN(nm, 200, sc_sim3_projection<RefMatcher>(s, false), sc_sim3_projection<FacMatcher>(s, false));
        snprintf(nm, sizeof nm, "Fuse(keyframe, Scw) seed %d", sd);
        RUN(nm, 300, sc_sim3_projection<RefMatcher>(s, true), sc_sim3_projection<FacMatcher>(s, true));
        snprintf(nm, sizeof nm, "SearchByBoW(keyframe, frame) seed %d", sd);
        RUN(nm, 200, sc_bow_kf_frame<RefMatcher>(s), sc_bow_kf_frame<FacMatcher>(s));
        snprintf(nm, sizeof nm, "SearchByBoW(keyframe, keyframe) seed %d", sd);
        RUN(nm, 200, sc_bow_kf_kf<RefMatcher>(s), sc_bow_kf_kf<FacMatcher>(s));
        snprintf(nm, sizeof nm, "SearchForInitialization seed %d", sd);
        RUN(nm, 150, sc_initialization<RefMatcher>(s), sc_initialization<FacMatcher>(s));
        snprintf(nm, sizeof nm, "SearchForTriangulation(all, no orientation check) seed %d", sd);
        RUN(nm, 150, sc_triangulation<RefMatcher>(s, false, false), sc_triangulation<FacMatcher>(s, false, false));
        snprintf(nm, sizeof nm, "SearchForTriangulation(stereo only, orientation check) seed %d", sd);
        RUN(nm, 20, sc_triangulation<RefMatcher>(s, true, true), sc_triangulation<FacMatcher>(s, true, true));
        snprintf(nm, sizeof nm, "Fuse(keyframe, map points) seed %d", sd);
        RUN(nm, 300, sc_fuse<RefMatcher>(s), sc_fuse<FacMatcher>(s));
        snprintf(nm, sizeof nm, "SearchBySim3 seed %d", sd);
        RUN(nm, 200, sc_sim3<RefMatcher>(s), sc_sim3<FacMatcher>(s));
        snprintf(nm, sizeof nm, "DescriptorDistance seed %d", sd);
        RUN(nm, 0, sc_distance<RefMatcher>(s), sc_distance<FacMatcher>(s));
    }
    printf("%s: %d scenario(s) failed\n", g_fails ? "FAIL" : "ok", g_fails);
    return g_fails ? 1 : 0;
}

// facade_match.cpp -- binds the templated Search* methods of include/orbb200/ORBmatcher.h to small stand-ins of
// ORB_SLAM2::Frame / KeyFrame / MapPoint (same member names as the reference classes) and checks them against the
// sequential reference loops evaluated with the CPU oracle's DescriptorDistance (oracle/liborb_ref.so: checker only).
//   facade_match  -> prints "ok ..." lines, exit code 0 on parity
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <set>
#include <type_traits>
#include <vector>

#include "orbb200/ORBmatcher.h"

// oracle/orb_ref.c (checker only)
extern "C" int ref_descriptor_distance(const unsigned char *a, const unsigned char *b);
extern "C" void ref_assign_grid(const void *kp, int n, float minX, float minY, float invW, float invH, int cols, int rows, int *cell_start, int *cell_items);
extern "C" int ref_features_in_area(const void *kp, const int *cell_start, const int *cell_items, float minX, float minY, float invW, float invH,
                                    int cols, int rows, float x, float y, float r, int minLevel, int maxLevel, int *out, int cap);

static unsigned long long g_s = 88172645463325252ull;
static unsigned rnd() { g_s ^= g_s << 13; g_s ^= g_s >> 7; g_s ^= g_s << 17; return (unsigned)(g_s >> 11); }
static float rndf() { return (rnd() & 0xFFFFFF) / (float)0x1000000; }

struct KeyFrame;
struct MapPoint {
    // Fuse-related state (src/MapPoint.cc): observations, replacement, distance invariance, normal, scale prediction
    float maxDist = 1e9f, minDist = 0.f;
    cv::Mat normal;
    MapPoint *replacedBy = nullptr;
    std::vector<std::pair<KeyFrame *, size_t>> observations;
    bool IsInKeyFrame(KeyFrame *kf) { for (auto &o : observations) if (o.first == kf) return true; return false; }
    float GetMaxDistanceInvariance() { return 1.2f * maxDist; }
    float GetMinDistanceInvariance() { return 0.8f * minDist; }
    cv::Mat GetNormal() { return normal.clone(); }
    int GetIndexInKeyFrame(KeyFrame *kf) { for (auto &o : observations) if (o.first == kf) return (int)o.second; return -1; }
    template <class T> int PredictScale(const float &currentDist, T *) {
        const float ratio = maxDist / currentDist;
        int nScale = (int)std::ceil(std::log(ratio) / std::log(1.2f));
        if (nScale < 0) nScale = 0; else if (nScale >= 8) nScale = 7;
        return nScale;
    }
    void Replace(MapPoint *other) { bad = true; replacedBy = other; }
    void AddObservation(KeyFrame *kf, size_t idx) { observations.push_back(std::make_pair(kf, idx)); obs++; }
    bool mbTrackInView = true, bad = false;
    int mnTrackScaleLevel = 0, obs = 1;
    float mTrackViewCos = 1.f, mTrackProjX = 0, mTrackProjY = 0, mTrackProjXR = 0;
    cv::Mat desc, pos;
    bool isBad() { return bad; }
    int Observations() { return obs; }
    cv::Mat GetDescriptor() { return desc.clone(); }
    cv::Mat GetWorldPos() { return pos.clone(); }
};

struct Frame {
    int N = 0;
    std::vector<cv::KeyPoint> mvKeys, mvKeysUn;
    std::vector<float> mvuRight, mvScaleFactors;
    std::vector<MapPoint *> mvpMapPoints;
    std::vector<bool> mvbOutlier;
    cv::Mat mDescriptors, mTcw;
    float mb = 0.5372f, mbf = 386.1448f, fx = 718.856f, fy = 718.856f, cx = 320.f, cy = 240.f;
    float mnMinX = 0, mnMinY = 0, mnMaxX = 640, mnMaxY = 480, mfGridElementWidthInv = 64.f / 640.f, mfGridElementHeightInv = 48.f / 480.f;
    std::map<unsigned int, std::vector<unsigned int>> mFeatVec; // DBoW2::FeatureVector is a std::map of index lists
    // Frame::AssignFeaturesToGrid + GetFeaturesInArea (src/Frame.cc:261-279, 386-442) through the CPU oracle
    mutable std::vector<int> cellStart, cellItems;
    std::vector<size_t> GetFeaturesInArea(const float &x, const float &y, const float &r, const int minLevel = -1, const int maxLevel = -1) const
    {
        static_assert(sizeof(cv::KeyPoint) == 28, "layout");
        if (cellStart.empty()) {
            cellStart.assign(64 * 48 + 1, 0);
            cellItems.assign(N > 0 ? N : 1, 0);
            ref_assign_grid(mvKeysUn.data(), N, mnMinX, mnMinY, mfGridElementWidthInv, mfGridElementHeightInv, 64, 48, cellStart.data(), cellItems.data());
        }
        std::vector<int> out(N > 0 ? N : 1);
        const int n = ref_features_in_area(mvKeysUn.data(), cellStart.data(), cellItems.data(), mnMinX, mnMinY, mfGridElementWidthInv,
                                           mfGridElementHeightInv, 64, 48, x, y, r, minLevel, maxLevel, out.data(), (int)out.size());
        return std::vector<size_t>(out.begin(), out.begin() + n);
    }
};

struct KeyFrame {
    int N = 0;
    std::vector<MapPoint *> mps;
    std::vector<cv::KeyPoint> mvKeysUn;
    std::vector<float> mvuRight, mvScaleFactors, mvLevelSigma2;
    cv::Mat mDescriptors, Rcw, tcw, Ow;
    float fx = 718.856f, fy = 718.856f, cx = 320.f, cy = 240.f, mbf = 386.1448f;
    float mnMinX = 0, mnMinY = 0, mnMaxX = 640, mnMaxY = 480, mfGridElementWidthInv = 64.f / 640.f, mfGridElementHeightInv = 48.f / 480.f;
    int mnGridCols = 64, mnGridRows = 48;
    std::vector<float> mvInvLevelSigma2;
    std::map<unsigned int, std::vector<unsigned int>> mFeatVec;
    bool IsInImage(const float &x, const float &y) const { return (x >= mnMinX && x < mnMaxX && y >= mnMinY && y < mnMaxY); }
    void AddMapPoint(MapPoint *pMP, const size_t &idx) { mps[idx] = pMP; }
    // KeyFrame::GetFeaturesInArea (src/KeyFrame.cc:607-647) through the CPU oracle: same grid walk, no level filter
    mutable std::vector<int> cellStart, cellItems;
    std::vector<size_t> GetFeaturesInArea(const float &x, const float &y, const float &r) const
    {
        if (cellStart.empty()) {
            cellStart.assign(64 * 48 + 1, 0);
            cellItems.assign(N > 0 ? N : 1, 0);
            ref_assign_grid(mvKeysUn.data(), N, mnMinX, mnMinY, mfGridElementWidthInv, mfGridElementHeightInv, 64, 48, cellStart.data(), cellItems.data());
        }
        std::vector<int> out(N > 0 ? N : 1);
        const int n = ref_features_in_area(mvKeysUn.data(), cellStart.data(), cellItems.data(), mnMinX, mnMinY, mfGridElementWidthInv,
                                           mfGridElementHeightInv, 64, 48, x, y, r, -1, -1, out.data(), (int)out.size());
        return std::vector<size_t>(out.begin(), out.begin() + n);
    }
    std::vector<MapPoint *> GetMapPointMatches() { return mps; }
    std::set<MapPoint *> GetMapPoints() { std::set<MapPoint *> r; for (MapPoint *m : mps) if (m) r.insert(m); return r; }
    MapPoint *GetMapPoint(const size_t &idx) { return mps[idx]; }
    cv::Mat GetCameraCenter() { return Ow.clone(); }
    cv::Mat GetRotation() { return Rcw.clone(); }
    cv::Mat GetTranslation() { return tcw.clone(); }
};

static void fill_frame(Frame &F, int N)
{
    F.N = N;
    F.mvKeys.resize(N); F.mvuRight.assign(N, -1.f); F.mvpMapPoints.assign(N, nullptr); F.mvbOutlier.assign(N, false);
    F.mDescriptors.create(N, 32, CV_8U);
    F.mvScaleFactors.resize(8);
    F.mvScaleFactors[0] = 1.f;
    for (int l = 1; l < 8; l++) F.mvScaleFactors[l] = F.mvScaleFactors[l - 1] * 1.2f;
    for (int i = 0; i < N; i++) {
        F.mvKeys[i] = cv::KeyPoint(rndf() * 640, rndf() * 480, 31, rndf() * 360, 50, rnd() % 8);
        for (int k = 0; k < 32; k++) F.mDescriptors.ptr(i)[k] = (unsigned char)rnd();
        if (rnd() % 3 == 0) F.mvuRight[i] = F.mvKeys[i].pt.x - rndf() * 40;
    }
    F.mvKeysUn = F.mvKeys;
}

static void noisy_copy(const unsigned char *src, unsigned char *dst, int flips)
{
    memcpy(dst, src, 32);
    for (int f = 0; f < flips; f++) { const unsigned b = rnd() % 256; dst[b >> 3] ^= (unsigned char)(1u << (b & 7)); }
}

// ---- small pose helpers of the checks below (CV_32F gemm semantics: float small-matrix branch, double for transposed operands) ---
// R*x + t = ONE cv::gemm(R, x, 1, t, 1) in the small-matrix branch: FLOAT accumulation left to right (pinned to cv2 in tests/test_ref_pin.py)
static void t_Rx_t(const float *R, const float *x, const float *t, float *y)
{
    for (int r = 0; r < 3; r++) {
        float s = R[3 * r] * x[0];
        s = s + R[3 * r + 1] * x[1];
        s = s + R[3 * r + 2] * x[2];
        y[r] = (float)((double)s + (double)t[r]);
    }
}
static void t_centre(const float *R, const float *t, float *O)
{
    for (int i = 0; i < 3; i++) O[i] = -(float)((double)R[i] * t[0] + (double)R[3 + i] * t[1] + (double)R[6 + i] * t[2]);
}
static float t_norm(const float *p) { return (float)std::sqrt((double)p[0] * p[0] + (double)p[1] * p[1] + (double)p[2] * p[2]); }
static void rot_y(float ang, float *R)
{
    const float c = std::cos(ang), sn = std::sin(ang);
    const float r[9] = {c, 0, sn, 0, 1, 0, -sn, 0, c};
    memcpy(R, r, sizeof r);
}
static cv::Mat mat3(const float *R) { cv::Mat m; m.create(3, 3, CV_32F); for (int i = 0; i < 9; i++) m.at<float>(i / 3, i % 3) = R[i]; return m; }
static cv::Mat vec3(const float *t) { cv::Mat m; m.create(3, 1, CV_32F); for (int i = 0; i < 3; i++) m.at<float>(i, 0) = t[i]; return m; }
static cv::Mat mat4(const float *R, const float *t, float s)
{
    cv::Mat m; m.create(4, 4, CV_32F);
    for (int r = 0; r < 4; r++) for (int c = 0; c < 4; c++) m.at<float>(r, c) = r == c;
    for (int r = 0; r < 3; r++) { for (int c = 0; c < 3; c++) m.at<float>(r, c) = s * R[3 * r + c]; m.at<float>(r, 3) = s * t[r]; }
    return m;
}
// world point whose camera coordinates under (R, t) are pc:  R^T (pc - t)
static void world_of(const float *R, const float *t, const float *pc, float *pw)
{
    for (int i = 0; i < 3; i++) pw[i] = (float)((double)R[i] * (pc[0] - t[0]) + (double)R[3 + i] * (pc[1] - t[1]) + (double)R[6 + i] * (pc[2] - t[2]));
}
static void keyframe_from(KeyFrame &K, const Frame &tmp)
{
    K.N = tmp.N; K.mps.assign(K.N, nullptr); K.mvuRight = tmp.mvuRight; K.mvKeysUn = tmp.mvKeys; K.mvScaleFactors = tmp.mvScaleFactors;
    K.mvLevelSigma2.clear(); K.mvInvLevelSigma2.clear();
    for (float f : K.mvScaleFactors) { K.mvLevelSigma2.push_back(f * f); K.mvInvLevelSigma2.push_back(1.0f / (f * f)); }
    K.mDescriptors = tmp.mDescriptors.clone();
    const float I[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, z[3] = {0, 0, 0};
    K.Rcw = mat3(I); K.tcw = vec3(z); K.Ow = vec3(z);
}
// a map point that lands near feature src of a keyframe-like target under camera pose (R, t), with a compatible scale band
template <class T> static void point_near(MapPoint &m, const T &K, int src, const float *R, const float *t, float jitter)
{
    const float z = 3.f + 20.f * rndf();
    const float u = K.mvKeysUn[src].pt.x + (rndf() - 0.5f) * jitter, v = K.mvKeysUn[src].pt.y + (rndf() - 0.5f) * jitter;
    const float pc[3] = {(u - K.cx) * z / K.fx, (v - K.cy) * z / K.fy, rnd() % 40 ? z : -z};
    float pw[3], O[3], PO[3];
    world_of(R, t, pc, pw);
    t_centre(R, t, O);
    for (int i = 0; i < 3; i++) PO[i] = pw[i] - O[i];
    const float nn = t_norm(PO);
    m.pos = vec3(pw);
    float nrm[3];
    for (int i = 0; i < 3; i++) nrm[i] = (rnd() % 15 ? 1.f : -1.f) * PO[i] / nn;
    m.normal = vec3(nrm);
    m.maxDist = nn * std::pow(1.2f, (float)K.mvKeysUn[src].octave) * (0.9f + 0.2f * rndf());
    m.minDist = m.maxDist / std::pow(1.2f, 7.f);
    m.desc.create(1, 32, CV_8U);
    noisy_copy(K.mDescriptors.ptr(src), m.desc.ptr(), rnd() % 70);
    m.obs = 1 + rnd() % 5;
    m.bad = rnd() % 30 == 0;
}
// reference gate of the Sim3 searches (src/ORBmatcher.cc:336-371), with the oracle-side arithmetic above
static bool ref_gate(KeyFrame *pKF, const float *Rcw, const float *tcw, const float *Ow, MapPoint *pMP, float &u, float &v, float &dist)
{
    const float pw[3] = {pMP->pos.at<float>(0, 0), pMP->pos.at<float>(1, 0), pMP->pos.at<float>(2, 0)};
    float pc[3];
    t_Rx_t(Rcw, pw, tcw, pc);
    if (pc[2] < 0.0f) return false;
    const float invz = 1 / pc[2], x = pc[0] * invz, y = pc[1] * invz;
    u = pKF->fx * x + pKF->cx; v = pKF->fy * y + pKF->cy;
    if (!pKF->IsInImage(u, v)) return false;
    const float PO[3] = {pw[0] - Ow[0], pw[1] - Ow[1], pw[2] - Ow[2]};
    dist = t_norm(PO);
    if (dist < pMP->GetMinDistanceInvariance() || dist > pMP->GetMaxDistanceInvariance()) return false;
    const double dot = (double)PO[0] * pMP->normal.at<float>(0, 0) + (double)PO[1] * pMP->normal.at<float>(1, 0) + (double)PO[2] * pMP->normal.at<float>(2, 0);
    return !(dot < 0.5 * dist);
}
static void ref_decompose(const cv::Mat &Scw, float *Rcw, float *tcw, float *Ow)
{
    float sR[9], st[3];
    for (int r = 0; r < 3; r++) { for (int c = 0; c < 3; c++) sR[3 * r + c] = Scw.at<float>(r, c); st[r] = Scw.at<float>(r, 3); }
    const float scw = (float)std::sqrt((double)sR[0] * sR[0] + (double)sR[1] * sR[1] + (double)sR[2] * sR[2]);
    const float inv = (float)(1.0 / scw);
    for (int i = 0; i < 9; i++) Rcw[i] = sR[i] * inv;
    for (int i = 0; i < 3; i++) tcw[i] = st[i] * inv;
    t_centre(Rcw, tcw, Ow);
}

// ---- the reference loops, straight, with the oracle's distance -------------------------------------------------
static int ref_SearchByProjection(Frame &F, const std::vector<MapPoint *> &vp, float th, float ratio)
{
    int nmatches = 0;
    const bool bFactor = th != 1.0;
    for (size_t iMP = 0; iMP < vp.size(); iMP++) {
        MapPoint *pMP = vp[iMP];
        if (!pMP->mbTrackInView || pMP->isBad()) continue;
        const int lvl = pMP->mnTrackScaleLevel;
        float r = pMP->mTrackViewCos > 0.998 ? 2.5f : 4.0f;
        if (bFactor) r *= th;
        const std::vector<size_t> vi = F.GetFeaturesInArea(pMP->mTrackProjX, pMP->mTrackProjY, r * F.mvScaleFactors[lvl], lvl - 1, lvl);
        if (vi.empty()) continue;
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (size_t k = 0; k < vi.size(); k++) {
            const size_t idx = vi[k];
            if (F.mvpMapPoints[idx] && F.mvpMapPoints[idx]->Observations() > 0) continue;
            if (F.mvuRight[idx] > 0) {
                const float er = std::fabs(pMP->mTrackProjXR - F.mvuRight[idx]);
                if (er > r * F.mvScaleFactors[lvl]) continue;
            }
            const int d = ref_descriptor_distance(pMP->desc.ptr(), F.mDescriptors.ptr((int)idx));
            if (d < bestDist) { bestDist2 = bestDist; bestDist = d; bestLevel2 = bestLevel; bestLevel = F.mvKeysUn[idx].octave; bestIdx = (int)idx; }
            else if (d < bestDist2) { bestLevel2 = F.mvKeysUn[idx].octave; bestDist2 = d; }
        }
        if (bestDist <= 100) {
            if (bestLevel == bestLevel2 && bestDist > ratio * bestDist2) continue;
            F.mvpMapPoints[bestIdx] = pMP;
            nmatches++;
        }
    }
    return nmatches;
}

int main()
{
    int fails = 0;
    // ---- M2: SearchByProjection(Frame, MapPoints, th)
    {
        Frame F1, F2;
        fill_frame(F1, 1500);
        F2 = F1;
        F2.mDescriptors = F1.mDescriptors.clone();
        std::vector<MapPoint> store(1200);
        std::vector<MapPoint *> vp;
        for (size_t i = 0; i < store.size(); i++) {
            MapPoint &m = store[i];
            const int src = rnd() % F1.N; // projects near an existing feature, descriptor = noisy copy (or random)
            m.mTrackProjX = F1.mvKeys[src].pt.x + (rndf() - 0.5f) * 4;
            m.mTrackProjY = F1.mvKeys[src].pt.y + (rndf() - 0.5f) * 4;
            m.mTrackProjXR = F1.mvuRight[src] > 0 ? F1.mvuRight[src] + (rndf() - 0.5f) * 6 : -1;
            m.mnTrackScaleLevel = F1.mvKeys[src].octave + (rnd() % 4 == 0 ? 1 : 0);
            if (m.mnTrackScaleLevel > 7) m.mnTrackScaleLevel = 7;
            m.mTrackViewCos = rnd() % 2 ? 0.9995f : 0.9f;
            m.mbTrackInView = rnd() % 10 != 0;
            m.bad = rnd() % 25 == 0;
            m.obs = rnd() % 6 == 0 ? 0 : 2;
            m.desc.create(1, 32, CV_8U);
            if (rnd() % 5) noisy_copy(F1.mDescriptors.ptr(src), m.desc.ptr(), rnd() % 70);
            else for (int k = 0; k < 32; k++) m.desc.ptr()[k] = (unsigned char)rnd();
            vp.push_back(&m);
        }
        for (size_t i = 0; i + 1 < vp.size(); i += 7) vp[i + 1]->desc = vp[i]->desc.clone(), vp[i + 1]->mTrackProjX = vp[i]->mTrackProjX, vp[i + 1]->mTrackProjY = vp[i]->mTrackProjY; // contested targets
        ORB_SLAM2::ORBmatcher matcher(0.8f);
        const int ng = matcher.SearchByProjection(F1, vp, 3.f);
        const int nr = ref_SearchByProjection(F2, vp, 3.f, 0.8f);
        bool same = ng == nr;
        for (int i = 0; i < F1.N; i++) same = same && F1.mvpMapPoints[i] == F2.mvpMapPoints[i];
        printf("%s SearchByProjection(local map): %d matches (reference loop %d)\n", same && ng > 300 ? "ok" : "FAIL", ng, nr);
        fails += !(same && ng > 300);
    }
    // ---- M3: SearchByProjection(CurrentFrame, LastFrame, th, bMono)
    for (int variant = 0; variant < 3; variant++) {
        Frame Last, Cur1, Cur2;
        fill_frame(Last, 1300);
        fill_frame(Cur1, 1500);
        // poses: last = identity, current = small rotation about y + translation (variant 1: forward, 2: backward motion)
        Last.mTcw.create(4, 4, CV_32F); Cur1.mTcw.create(4, 4, CV_32F);
        for (int r = 0; r < 4; r++) for (int c = 0; c < 4; c++) { Last.mTcw.at<float>(r, c) = r == c; Cur1.mTcw.at<float>(r, c) = r == c; }
        const float ang = 0.01f;
        Cur1.mTcw.at<float>(0, 0) = std::cos(ang); Cur1.mTcw.at<float>(0, 2) = std::sin(ang);
        Cur1.mTcw.at<float>(2, 0) = -std::sin(ang); Cur1.mTcw.at<float>(2, 2) = std::cos(ang);
        Cur1.mTcw.at<float>(0, 3) = 0.05f; Cur1.mTcw.at<float>(2, 3) = variant == 1 ? -1.0f : (variant == 2 ? 1.0f : 0.1f);
        std::vector<MapPoint> store(Last.N);
        for (int i = 0; i < Last.N; i++) {
            MapPoint &m = store[i];
            // back-project the last keypoint at a random depth, place a matching feature near its projection in Cur
            const float z = 4.f + 30.f * rndf();
            m.pos.create(3, 1, CV_32F);
            m.pos.at<float>(0, 0) = (Last.mvKeys[i].pt.x - Last.cx) * z / Last.fx;
            m.pos.at<float>(1, 0) = (Last.mvKeys[i].pt.y - Last.cy) * z / Last.fy;
            m.pos.at<float>(2, 0) = z;
            m.desc.create(1, 32, CV_8U);
            memcpy(m.desc.ptr(), Last.mDescriptors.ptr(i), 32);
            m.obs = rnd() % 5 == 0 ? 0 : 3;
            if (rnd() % 8) Last.mvpMapPoints[i] = &m;
            Last.mvbOutlier[i] = rnd() % 12 == 0;
            // projection in the current frame
            const float *T = Cur1.mTcw.ptr<float>();
            const float xc = T[0] * m.pos.at<float>(0, 0) + T[1] * m.pos.at<float>(1, 0) + T[2] * z + T[3];
            const float yc = T[4] * m.pos.at<float>(0, 0) + T[5] * m.pos.at<float>(1, 0) + T[6] * z + T[7];
            const float zc = T[8] * m.pos.at<float>(0, 0) + T[9] * m.pos.at<float>(1, 0) + T[10] * z + T[11];
            const int j = i % Cur1.N;
            if (zc > 0.1f && rnd() % 4) {
                Cur1.mvKeys[j].pt.x = Cur1.fx * xc / zc + Cur1.cx + (rndf() - 0.5f) * 6;
                Cur1.mvKeys[j].pt.y = Cur1.fy * yc / zc + Cur1.cy + (rndf() - 0.5f) * 6;
                Cur1.mvKeys[j].octave = std::max(0, std::min(7, Last.mvKeys[i].octave + (int)(rnd() % 3) - 1));
                Cur1.mvKeys[j].angle = std::fmod(Last.mvKeys[i].angle + (rnd() % 8 ? 5.f : 170.f * rndf()), 360.f);
                noisy_copy(Last.mDescriptors.ptr(i), Cur1.mDescriptors.ptr(j), rnd() % 80);
                if (Cur1.mvuRight[j] > 0) Cur1.mvuRight[j] = Cur1.mvKeys[j].pt.x - Cur1.mbf / zc + (rndf() - 0.5f) * 10;
            }
        }
        Cur1.mvKeysUn = Cur1.mvKeys;
        Cur2 = Cur1;
        Cur2.mDescriptors = Cur1.mDescriptors.clone();
        ORB_SLAM2::ORBmatcher matcher(0.9f, true);
        const int ng = matcher.SearchByProjection(Cur1, Last, 15.f, false);
        // reference loop (src/ORBmatcher.cc:1396-1553) with the oracle's distance and grid
        int nr = 0;
        {
            Frame &C = Cur2;
            std::vector<int> rotHist[30];
            const float *Tc = C.mTcw.ptr<float>(), *Tl = Last.mTcw.ptr<float>();
            float twc[3], tlc[3];
            for (int i = 0; i < 3; i++) twc[i] = -(float)((double)Tc[i] * Tc[3] + (double)Tc[4 + i] * Tc[7] + (double)Tc[8 + i] * Tc[11]);
            { float Rl[9], tl[3]; for (int r = 0; r < 3; r++) { for (int c = 0; c < 3; c++) Rl[3 * r + c] = Tl[4 * r + c]; tl[r] = Tl[4 * r + 3]; } t_Rx_t(Rl, twc, tl, tlc); }
            const bool bForward = tlc[2] > C.mb, bBackward = -tlc[2] > C.mb;
            for (int i = 0; i < Last.N; i++) {
                MapPoint *pMP = Last.mvpMapPoints[i];
                if (!pMP || Last.mvbOutlier[i]) continue;
                const float X = pMP->pos.at<float>(0, 0), Y = pMP->pos.at<float>(1, 0), Z = pMP->pos.at<float>(2, 0);
                float Rc[9], tc3[3], pcam[3];
                for (int r = 0; r < 3; r++) { for (int c = 0; c < 3; c++) Rc[3 * r + c] = Tc[4 * r + c]; tc3[r] = Tc[4 * r + 3]; }
                const float pw3[3] = {X, Y, Z};
                t_Rx_t(Rc, pw3, tc3, pcam);
                const float xc = pcam[0], yc = pcam[1], zc = pcam[2];
                const float invzc = 1.0 / zc;
                if (invzc < 0) continue;
                const float u = C.fx * xc * invzc + C.cx, v = C.fy * yc * invzc + C.cy;
                if (u < C.mnMinX || u > C.mnMaxX || v < C.mnMinY || v > C.mnMaxY) continue;
                const int oct = Last.mvKeys[i].octave;
                const float radius = 15.f * C.mvScaleFactors[oct];
                std::vector<size_t> vi;
                if (bForward) vi = C.GetFeaturesInArea(u, v, radius, oct);
                else if (bBackward) vi = C.GetFeaturesInArea(u, v, radius, 0, oct);
                else vi = C.GetFeaturesInArea(u, v, radius, oct - 1, oct + 1);
                if (vi.empty()) continue;
                int bestDist = 256, bestIdx2 = -1;
                for (size_t i2 : vi) {
                    if (C.mvpMapPoints[i2] && C.mvpMapPoints[i2]->Observations() > 0) continue;
                    if (C.mvuRight[i2] > 0) {
                        const float ur = u - C.mbf * invzc;
                        if (std::fabs(ur - C.mvuRight[i2]) > radius) continue;
                    }
                    const int d = ref_descriptor_distance(pMP->desc.ptr(), C.mDescriptors.ptr((int)i2));
                    if (d < bestDist) { bestDist = d; bestIdx2 = (int)i2; }
                }
                if (bestDist <= 100) {
                    C.mvpMapPoints[bestIdx2] = pMP;
                    nr++;
                    float rot = Last.mvKeysUn[i].angle - C.mvKeysUn[bestIdx2].angle;
                    if (rot < 0.0) rot += 360.0f;
                    int bin = (int)std::round(rot * (1.0f / 30));
                    if (bin == 30) bin = 0;
                    rotHist[bin].push_back(bestIdx2);
                }
            }
            int m1 = 0, m2 = 0, m3 = 0, i1 = -1, i2 = -1, i3 = -1;
            for (int i = 0; i < 30; i++) {
                const int sz = (int)rotHist[i].size();
                if (sz > m1) { m3 = m2; m2 = m1; m1 = sz; i3 = i2; i2 = i1; i1 = i; }
                else if (sz > m2) { m3 = m2; m2 = sz; i3 = i2; i2 = i; }
                else if (sz > m3) { m3 = sz; i3 = i; }
            }
            if (m2 < 0.1f * (float)m1) { i2 = -1; i3 = -1; } else if (m3 < 0.1f * (float)m1) i3 = -1;
            for (int i = 0; i < 30; i++) {
                if (i == i1 || i == i2 || i == i3) continue;
                for (int j : rotHist[i]) { C.mvpMapPoints[j] = nullptr; nr--; }
            }
        }
        bool same = ng == nr;
        for (int i = 0; i < Cur1.N; i++) same = same && Cur1.mvpMapPoints[i] == Cur2.mvpMapPoints[i];
        printf("%s SearchByProjection(last frame, variant %d): %d matches (reference loop %d)\n", same && ng > 150 ? "ok" : "FAIL", variant, ng, nr);
        fails += !(same && ng > 150);
    }
    // ---- M6: SearchByBoW(KeyFrame, Frame)
    {
        Frame F;
        fill_frame(F, 1400);
        KeyFrame KF;
        const int NK = 1300;
        KF.N = NK; KF.mvKeysUn.resize(NK); KF.mps.assign(NK, nullptr); KF.mDescriptors.create(NK, 32, CV_8U);
        std::vector<MapPoint> store(NK);
        for (int i = 0; i < NK; i++) {
            const int src = rnd() % F.N;
            KF.mvKeysUn[i] = F.mvKeys[src];
            KF.mvKeysUn[i].angle = std::fmod(F.mvKeys[src].angle + (rnd() % 8 ? 12.f : 200.f * rndf()), 360.f);
            noisy_copy(F.mDescriptors.ptr(src), KF.mDescriptors.ptr(i), rnd() % 60);
            if (rnd() % 6) { KF.mps[i] = &store[i]; store[i].bad = rnd() % 30 == 0; }
            const unsigned node = (unsigned)(src % 97) * 3; // same vocabulary node as its source most of the time
            KF.mFeatVec[rnd() % 9 ? node : node + 1].push_back(i);
        }
        for (int i = 0; i < F.N; i++) F.mFeatVec[(unsigned)(i % 97) * 3].push_back(i);
        std::vector<MapPoint *> got, want;
        ORB_SLAM2::ORBmatcher matcher(0.7f, true);
        const int ng = matcher.SearchByBoW(&KF, F, got);
        // reference loop (src/ORBmatcher.cc:171-303) with the oracle distance
        want.assign(F.N, nullptr);
        int nr = 0;
        std::vector<int> rotHist[30];
        auto KFit = KF.mFeatVec.begin(); auto Fit = F.mFeatVec.begin();
        while (KFit != KF.mFeatVec.end() && Fit != F.mFeatVec.end()) {
            if (KFit->first == Fit->first) {
                for (unsigned realIdxKF : KFit->second) {
                    MapPoint *pMP = KF.mps[realIdxKF];
                    if (!pMP || pMP->isBad()) continue;
                    int b1 = 256, bi = -1, b2 = 256;
                    for (unsigned realIdxF : Fit->second) {
                        if (want[realIdxF]) continue;
                        const int d = ref_descriptor_distance(KF.mDescriptors.ptr((int)realIdxKF), F.mDescriptors.ptr((int)realIdxF));
                        if (d < b1) { b2 = b1; b1 = d; bi = (int)realIdxF; } else if (d < b2) b2 = d;
                    }
                    if (b1 <= 50 && (float)b1 < 0.7f * (float)b2) {
                        want[bi] = pMP;
                        float rot = KF.mvKeysUn[realIdxKF].angle - F.mvKeys[bi].angle;
                        if (rot < 0.0) rot += 360.0f;
                        int bin = (int)std::round(rot * (1.0f / 30));
                        if (bin == 30) bin = 0;
                        rotHist[bin].push_back(bi);
                        nr++;
                    }
                }
                KFit++; Fit++;
            } else if (KFit->first < Fit->first) KFit = KF.mFeatVec.lower_bound(Fit->first);
            else Fit = F.mFeatVec.lower_bound(KFit->first);
        }
        int m1 = 0, m2 = 0, m3 = 0, i1 = -1, i2 = -1, i3 = -1;
        for (int i = 0; i < 30; i++) {
            const int s = (int)rotHist[i].size();
            if (s > m1) { m3 = m2; m2 = m1; m1 = s; i3 = i2; i2 = i1; i1 = i; }
            else if (s > m2) { m3 = m2; m2 = s; i3 = i2; i2 = i; }
            else if (s > m3) { m3 = s; i3 = i; }
        }
        if (m2 < 0.1f * (float)m1) { i2 = -1; i3 = -1; } else if (m3 < 0.1f * (float)m1) i3 = -1;
        for (int i = 0; i < 30; i++) {
            if (i == i1 || i == i2 || i == i3) continue;
            for (int j : rotHist[i]) { want[j] = nullptr; nr--; }
        }
        bool same = ng == nr && got.size() == want.size();
        for (size_t i = 0; same && i < want.size(); i++) same = got[i] == want[i];
        printf("%s SearchByBoW(KeyFrame, Frame): %d matches (reference loop %d)\n", same && ng > 200 ? "ok" : "FAIL", ng, nr);
        fails += !(same && ng > 200);
    }
    // ---- M7: SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize)
    {
        Frame F1, F2;
        fill_frame(F1, 1600);
        fill_frame(F2, 1600);
        for (int i = 0; i < F1.N; i++) {
            F1.mvKeys[i].octave = rnd() % 3 ? 0 : (int)(rnd() % 4);
            const int j = (i * 7) % F2.N;
            if (rnd() % 5) { // F2 holds a shifted, noisy copy of most F1 features; some targets are contested
                F2.mvKeys[j].pt.x = F1.mvKeys[i].pt.x + (rndf() - 0.5f) * 60;
                F2.mvKeys[j].pt.y = F1.mvKeys[i].pt.y + (rndf() - 0.5f) * 60;
                F2.mvKeys[j].octave = rnd() % 6 ? 0 : 1;
                F2.mvKeys[j].angle = std::fmod(F1.mvKeys[i].angle + (rnd() % 8 ? 3.f : 200.f * rndf()), 360.f);
                noisy_copy(F1.mDescriptors.ptr(i), F2.mDescriptors.ptr(j), rnd() % 50);
            }
        }
        for (int i = 0; i + 3 < F1.N; i += 11) { memcpy(F1.mDescriptors.ptr(i + 3), F1.mDescriptors.ptr(i), 32); F1.mvKeys[i + 3].pt = F1.mvKeys[i].pt; F1.mvKeys[i + 3].octave = F1.mvKeys[i].octave; }
        F1.mvKeysUn = F1.mvKeys; F2.mvKeysUn = F2.mvKeys;
        std::vector<cv::Point2f> prevG(F1.N), prevR(F1.N);
        for (int i = 0; i < F1.N; i++) prevG[i] = prevR[i] = F1.mvKeysUn[i].pt;
        std::vector<int> gotM, wantM(F1.N, -1);
        ORB_SLAM2::ORBmatcher matcher(0.9f, true);
        const int ng = matcher.SearchForInitialization(F1, F2, prevG, gotM, 100);
        int nr = 0;
        {
            std::vector<int> rotHist[30], vMatchedDistance(F2.N, 2147483647), vnMatches21(F2.N, -1);
            for (int i1 = 0; i1 < F1.N; i1++) {
                if (F1.mvKeysUn[i1].octave > 0) continue;
                std::vector<size_t> vi = F2.GetFeaturesInArea(prevR[i1].x, prevR[i1].y, 100, 0, 0);
                if (vi.empty()) continue;
                int bestDist = 2147483647, bestDist2 = 2147483647, bestIdx2 = -1;
                for (size_t i2 : vi) {
                    const int d = ref_descriptor_distance(F1.mDescriptors.ptr(i1), F2.mDescriptors.ptr((int)i2));
                    if (vMatchedDistance[i2] <= d) continue;
                    if (d < bestDist) { bestDist2 = bestDist; bestDist = d; bestIdx2 = (int)i2; } else if (d < bestDist2) bestDist2 = d;
                }
                if (bestDist <= 50 && bestDist < (float)bestDist2 * 0.9f) {
                    if (vnMatches21[bestIdx2] >= 0) { wantM[vnMatches21[bestIdx2]] = -1; nr--; }
                    wantM[i1] = bestIdx2; vnMatches21[bestIdx2] = i1; vMatchedDistance[bestIdx2] = bestDist; nr++;
                    float rot = F1.mvKeysUn[i1].angle - F2.mvKeysUn[bestIdx2].angle;
                    if (rot < 0.0) rot += 360.0f;
                    int bin = (int)std::round(rot * (1.0f / 30));
                    if (bin == 30) bin = 0;
                    rotHist[bin].push_back(i1);
                }
            }
            int m1 = 0, m2 = 0, m3 = 0, i1 = -1, i2 = -1, i3 = -1;
            for (int i = 0; i < 30; i++) {
                const int sz = (int)rotHist[i].size();
                if (sz > m1) { m3 = m2; m2 = m1; m1 = sz; i3 = i2; i2 = i1; i1 = i; }
                else if (sz > m2) { m3 = m2; m2 = sz; i3 = i2; i2 = i; }
                else if (sz > m3) { m3 = sz; i3 = i; }
            }
            if (m2 < 0.1f * (float)m1) { i2 = -1; i3 = -1; } else if (m3 < 0.1f * (float)m1) i3 = -1;
            for (int i = 0; i < 30; i++) {
                if (i == i1 || i == i2 || i == i3) continue;
                for (int idx1 : rotHist[i]) if (wantM[idx1] >= 0) { wantM[idx1] = -1; nr--; }
            }
            for (int i = 0; i < F1.N; i++) if (wantM[i] >= 0) prevR[i] = F2.mvKeysUn[wantM[i]].pt;
        }
        bool same = ng == nr && gotM == wantM;
        for (int i = 0; same && i < F1.N; i++) same = prevG[i].x == prevR[i].x && prevG[i].y == prevR[i].y;
        printf("%s SearchForInitialization: %d matches (reference loop %d)\n", same && ng > 150 ? "ok" : "FAIL", ng, nr);
        fails += !(same && ng > 150);
    }
    // ---- M8: SearchForTriangulation(KF1, KF2, F12, pairs, bOnlyStereo)
    for (int onlyStereo = 0; onlyStereo < 2; onlyStereo++) {
        KeyFrame K1, K2;
        Frame tmp;
        fill_frame(tmp, 1400);
        std::vector<MapPoint> store(8);
        for (KeyFrame *K : {&K1, &K2}) {
            K->N = tmp.N; K->mps.assign(K->N, nullptr); K->mvuRight.assign(K->N, -1.f);
            K->mvScaleFactors = tmp.mvScaleFactors;
            for (float f : K->mvScaleFactors) K->mvLevelSigma2.push_back(f * f);
            K->mDescriptors.create(K->N, 32, CV_8U);
            K->Rcw.create(3, 3, CV_32F); K->tcw.create(3, 1, CV_32F); K->Ow.create(3, 1, CV_32F);
            for (int r = 0; r < 3; r++) { for (int c = 0; c < 3; c++) K->Rcw.at<float>(r, c) = r == c; K->tcw.at<float>(r, 0) = 0; K->Ow.at<float>(r, 0) = 0; }
        }
        K2.tcw.at<float>(0, 0) = -0.5f; K2.tcw.at<float>(2, 0) = 0.3f; K2.Ow.at<float>(0, 0) = 0.5f; K2.Ow.at<float>(2, 0) = -0.3f;
        K1.mvKeysUn = tmp.mvKeys; K2.mvKeysUn.resize(tmp.N);
        for (int i = 0; i < tmp.N; i++) {
            memcpy(K1.mDescriptors.ptr(i), tmp.mDescriptors.ptr(i), 32);
            K2.mvKeysUn[i] = tmp.mvKeys[i];
            K2.mvKeysUn[i].pt.x += (rndf() - 0.5f) * 80;
            K2.mvKeysUn[i].pt.y += rnd() % 4 ? (rndf() - 0.5f) * 2.f : (rndf() - 0.5f) * 30.f; // epipolar lines are image rows
            K2.mvKeysUn[i].angle = std::fmod(tmp.mvKeys[i].angle + (rnd() % 8 ? 4.f : 150.f * rndf()), 360.f);
            noisy_copy(tmp.mDescriptors.ptr(i), K2.mDescriptors.ptr(i), rnd() % 70);
            if (rnd() % 3 == 0) K1.mvuRight[i] = 10.f;
            if (rnd() % 3 == 0) K2.mvuRight[i] = 10.f;
            if (rnd() % 10 == 0) K1.mps[i] = &store[i % 8];
            if (rnd() % 10 == 0) K2.mps[i] = &store[i % 8];
            const unsigned node = (unsigned)(i % 53) * 2;
            K1.mFeatVec[node].push_back(i);
            K2.mFeatVec[rnd() % 12 ? node : node + 2].push_back(i);
        }
        for (int i = 0; i + 53 < tmp.N; i += 9) memcpy(K2.mDescriptors.ptr(i + 53), K2.mDescriptors.ptr(i), 32), K2.mvKeysUn[i + 53].pt.y = K2.mvKeysUn[i].pt.y; // equal distances in one node
        cv::Mat F12(3, 3, CV_32F);
        for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) F12.at<float>(r, c) = 0;
        F12.at<float>(1, 2) = -1.f; F12.at<float>(2, 1) = 1.f;
        std::vector<std::pair<size_t, size_t>> got, want;
        ORB_SLAM2::ORBmatcher matcher(0.6f, false), matcherOri(0.6f, true);
        ORB_SLAM2::ORBmatcher &mm = onlyStereo ? matcherOri : matcher;
        const int ng = mm.SearchForTriangulation(&K1, &K2, F12, got, onlyStereo != 0);
        int nr = 0;
        {
            const float C2[3] = {K2.tcw.at<float>(0, 0), K2.tcw.at<float>(1, 0), K2.tcw.at<float>(2, 0)}; // R2w = I, Cw = 0
            const float invz = 1.0f / C2[2], ex = K2.fx * C2[0] * invz + K2.cx, ey = K2.fy * C2[1] * invz + K2.cy;
            std::vector<int> vM(K1.N, -1), rotHist[30];
            auto f1 = K1.mFeatVec.begin(); auto f2 = K2.mFeatVec.begin();
            while (f1 != K1.mFeatVec.end() && f2 != K2.mFeatVec.end()) {
                if (f1->first == f2->first) {
                    for (unsigned idx1 : f1->second) {
                        if (K1.mps[idx1]) continue;
                        const bool s1 = K1.mvuRight[idx1] >= 0;
                        if (onlyStereo && !s1) continue;
                        const cv::KeyPoint &kp1 = K1.mvKeysUn[idx1];
                        int bestDist = 50, bestIdx2 = -1;
                        for (unsigned idx2 : f2->second) {
                            if (K2.mps[idx2]) continue;
                            const bool s2 = K2.mvuRight[idx2] >= 0;
                            if (onlyStereo && !s2) continue;
                            const int d = ref_descriptor_distance(K1.mDescriptors.ptr((int)idx1), K2.mDescriptors.ptr((int)idx2));
                            if (d > 50 || d > bestDist) continue;
                            const cv::KeyPoint &kp2 = K2.mvKeysUn[idx2];
                            if (!s1 && !s2) {
                                const float dx = ex - kp2.pt.x, dy = ey - kp2.pt.y;
                                if (dx * dx + dy * dy < 100 * K2.mvScaleFactors[kp2.octave]) continue;
                            }
                            const float a = kp1.pt.x * 0 + kp1.pt.y * 0 + 0, b = kp1.pt.x * 0 + kp1.pt.y * 0 + 1.f, c = kp1.pt.x * 0 + kp1.pt.y * -1.f + 0;
                            const float num = a * kp2.pt.x + b * kp2.pt.y + c, den = a * a + b * b;
                            if (den == 0) continue;
                            if (num * num / den < 3.84 * K2.mvLevelSigma2[kp2.octave]) { bestIdx2 = (int)idx2; bestDist = d; }
                        }
                        if (bestIdx2 >= 0) {
                            vM[idx1] = bestIdx2; nr++;
                            float rot = kp1.angle - K2.mvKeysUn[bestIdx2].angle;
                            if (rot < 0.0) rot += 360.0f;
                            int bin = (int)std::round(rot * (1.0f / 30));
                            if (bin == 30) bin = 0;
                            rotHist[bin].push_back((int)idx1);
                        }
                    }
                    f1++; f2++;
                } else if (f1->first < f2->first) f1 = K1.mFeatVec.lower_bound(f2->first);
                else f2 = K2.mFeatVec.lower_bound(f1->first);
            }
            if (onlyStereo) { // this run checks orientation
                int m1 = 0, m2 = 0, m3 = 0, i1 = -1, i2 = -1, i3 = -1;
                for (int i = 0; i < 30; i++) {
                    const int sz = (int)rotHist[i].size();
                    if (sz > m1) { m3 = m2; m2 = m1; m1 = sz; i3 = i2; i2 = i1; i1 = i; }
                    else if (sz > m2) { m3 = m2; m2 = sz; i3 = i2; i2 = i; }
                    else if (sz > m3) { m3 = sz; i3 = i; }
                }
                if (m2 < 0.1f * (float)m1) { i2 = -1; i3 = -1; } else if (m3 < 0.1f * (float)m1) i3 = -1;
                for (int i = 0; i < 30; i++) {
                    if (i == i1 || i == i2 || i == i3) continue;
                    for (int j : rotHist[i]) { vM[j] = -1; nr--; }
                }
            }
            for (int i = 0; i < K1.N; i++) if (vM[i] >= 0) want.push_back(std::make_pair((size_t)i, (size_t)vM[i]));
        }
        const bool same = ng == nr && got == want;
        const int need = onlyStereo ? 20 : 150;
        printf("%s SearchForTriangulation(bOnlyStereo=%d): %d pairs (reference loop %d)\n", same && ng > need ? "ok" : "FAIL", onlyStereo, ng, nr);
        fails += !(same && ng > need);
    }
    // ---- M9: Fuse(KeyFrame, MapPoints, th)
    {
        Frame tmp;
        fill_frame(tmp, 1500);
        const int NP = 1300;
        KeyFrame KA, KB;
        std::vector<MapPoint> storeA(NP + tmp.N), storeB;
        for (KeyFrame *K : {&KA, &KB}) {
            K->N = tmp.N; K->mps.assign(K->N, nullptr); K->mvuRight = tmp.mvuRight; K->mvKeysUn = tmp.mvKeys;
            K->mvScaleFactors = tmp.mvScaleFactors;
            for (float f : K->mvScaleFactors) { K->mvLevelSigma2.push_back(f * f); K->mvInvLevelSigma2.push_back(1.0f / (f * f)); }
            K->mDescriptors = tmp.mDescriptors.clone();
            K->Rcw.create(3, 3, CV_32F); K->tcw.create(3, 1, CV_32F); K->Ow.create(3, 1, CV_32F);
            for (int r = 0; r < 3; r++) { for (int c = 0; c < 3; c++) K->Rcw.at<float>(r, c) = r == c; K->tcw.at<float>(r, 0) = 0; K->Ow.at<float>(r, 0) = 0; }
        }
        for (int i = 0; i < NP; i++) { // candidates to fuse: 3-D points projecting near keyframe features
            MapPoint &m = storeA[i];
            const int src = rnd() % tmp.N;
            const float z = 3.f + 20.f * rndf();
            const float u = tmp.mvKeys[src].pt.x + (rndf() - 0.5f) * 3, v = tmp.mvKeys[src].pt.y + (rndf() - 0.5f) * 3;
            m.pos.create(3, 1, CV_32F);
            m.pos.at<float>(0, 0) = (u - KA.cx) * z / KA.fx; m.pos.at<float>(1, 0) = (v - KA.cy) * z / KA.fy; m.pos.at<float>(2, 0) = rnd() % 40 ? z : -z;
            m.normal.create(3, 1, CV_32F);
            const float nn = std::sqrt(m.pos.at<float>(0, 0) * m.pos.at<float>(0, 0) + m.pos.at<float>(1, 0) * m.pos.at<float>(1, 0) + z * z);
            for (int r = 0; r < 3; r++) m.normal.at<float>(r, 0) = (rnd() % 15 ? 1.f : -1.f) * m.pos.at<float>(r, 0) / nn;
            m.maxDist = nn * std::pow(1.2f, (float)tmp.mvKeys[src].octave) * (0.9f + 0.2f * rndf());
            m.minDist = m.maxDist / std::pow(1.2f, 7.f);
            m.desc.create(1, 32, CV_8U);
            noisy_copy(tmp.mDescriptors.ptr(src), m.desc.ptr(), rnd() % 70);
            m.obs = 1 + rnd() % 5;
            m.bad = rnd() % 30 == 0;
        }
        for (int j = 0; j < tmp.N; j++) { // some keyframe features already hold a map point (-> Replace either way)
            MapPoint &m = storeA[NP + j];
            m.obs = 1 + rnd() % 5; m.bad = rnd() % 20 == 0;
        }
        storeB = storeA;
        for (size_t i = 0; i < storeB.size(); i++) { storeB[i].desc = storeA[i].desc.clone(); }
        std::vector<MapPoint *> vpA, vpB;
        for (int i = 0; i < NP; i++) { vpA.push_back(rnd() % 50 ? &storeA[i] : nullptr); vpB.push_back(vpA.back() ? &storeB[i] : nullptr); }
        for (int i = 5; i < NP; i += 40) { vpA[i] = vpA[i - 3]; vpB[i] = vpB[i - 3]; } // duplicates: IsInKeyFrame after the first fuse
        for (int j = 0; j < tmp.N; j++) if (rnd() % 3 == 0) { KA.mps[j] = &storeA[NP + j]; KB.mps[j] = &storeB[NP + j]; storeA[NP + j].observations.push_back({&KA, (size_t)j}); storeB[NP + j].observations.push_back({&KB, (size_t)j}); }
        for (int j = 0; j < tmp.N; j += 17) if (KA.mps[j]) { vpA.push_back(KA.mps[j]); vpB.push_back(KB.mps[j]); } // points that get replaced appear later in the list
        ORB_SLAM2::ORBmatcher matcher;
        const int ng = matcher.Fuse(&KA, vpA, 3.0f);
        int nr = 0;
        { // reference loop (src/ORBmatcher.cc:873-1038) on copy B with the oracle's distance and grid
            KeyFrame *pKF = &KB;
            for (size_t i = 0; i < vpB.size(); i++) {
                MapPoint *pMP = vpB[i];
                if (!pMP) continue;
                if (pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;
                const float X = pMP->pos.at<float>(0, 0), Y = pMP->pos.at<float>(1, 0), Z = pMP->pos.at<float>(2, 0);
                if (Z < 0.0f) continue; // Rcw = I, tcw = 0
                const float invz = 1 / Z, x = X * invz, y = Y * invz, u = pKF->fx * x + pKF->cx, v = pKF->fy * y + pKF->cy;
                if (!pKF->IsInImage(u, v)) continue;
                const float ur = u - pKF->mbf * invz;
                const float dist3D = (float)std::sqrt((double)X * X + (double)Y * Y + (double)Z * Z);
                if (dist3D < pMP->GetMinDistanceInvariance() || dist3D > pMP->GetMaxDistanceInvariance()) continue;
                const double dot = (double)X * pMP->normal.at<float>(0, 0) + (double)Y * pMP->normal.at<float>(1, 0) + (double)Z * pMP->normal.at<float>(2, 0);
                if (dot < 0.5 * dist3D) continue;
                const int lvl = pMP->PredictScale(dist3D, pKF);
                const float radius = 3.0f * pKF->mvScaleFactors[lvl];
                const std::vector<size_t> vi = pKF->GetFeaturesInArea(u, v, radius);
                if (vi.empty()) continue;
                int bestDist = 256, bestIdx = -1;
                for (size_t idx : vi) {
                    const cv::KeyPoint &kp = pKF->mvKeysUn[idx];
                    if (kp.octave < lvl - 1 || kp.octave > lvl) continue;
                    if (pKF->mvuRight[idx] >= 0) {
                        const float ex = u - kp.pt.x, ey = v - kp.pt.y, er = ur - pKF->mvuRight[idx];
                        if ((ex * ex + ey * ey + er * er) * pKF->mvInvLevelSigma2[kp.octave] > 7.8) continue;
                    } else {
                        const float ex = u - kp.pt.x, ey = v - kp.pt.y;
                        if ((ex * ex + ey * ey) * pKF->mvInvLevelSigma2[kp.octave] > 5.99) continue;
                    }
                    const int d = ref_descriptor_distance(pMP->desc.ptr(), pKF->mDescriptors.ptr((int)idx));
                    if (d < bestDist) { bestDist = d; bestIdx = (int)idx; }
                }
                if (bestDist <= 50) {
                    MapPoint *in = pKF->GetMapPoint(bestIdx);
                    if (in) { if (!in->isBad()) { if (in->Observations() > pMP->Observations()) pMP->Replace(in); else in->Replace(pMP); } }
                    else { pMP->AddObservation(pKF, bestIdx); pKF->AddMapPoint(pMP, bestIdx); }
                    nr++;
                }
            }
        }
        bool same = ng == nr;
        for (int j = 0; same && j < tmp.N; j++) same = (KA.mps[j] ? KA.mps[j] - storeA.data() : -1) == (KB.mps[j] ? KB.mps[j] - storeB.data() : -1);
        for (size_t i = 0; same && i < storeA.size(); i++)
            same = storeA[i].bad == storeB[i].bad && storeA[i].obs == storeB[i].obs &&
                   (storeA[i].replacedBy ? storeA[i].replacedBy - storeA.data() : -1) == (storeB[i].replacedBy ? storeB[i].replacedBy - storeB.data() : -1);
        printf("%s Fuse(KeyFrame, MapPoints): %d fused (reference loop %d)\n", same && ng > 300 ? "ok" : "FAIL", ng, nr);
        fails += !(same && ng > 300);
    }
    // ---- M6b: SearchByBoW(KeyFrame, KeyFrame)
    {
        Frame tmp;
        fill_frame(tmp, 1300);
        KeyFrame K1, K2;
        std::vector<MapPoint> s1(tmp.N), s2(tmp.N);
        for (KeyFrame *K : {&K1, &K2}) { K->N = tmp.N; K->mps.assign(K->N, nullptr); K->mvKeysUn = tmp.mvKeys; K->mDescriptors.create(K->N, 32, CV_8U); }
        for (int i = 0; i < tmp.N; i++) {
            memcpy(K1.mDescriptors.ptr(i), tmp.mDescriptors.ptr(i), 32);
            noisy_copy(tmp.mDescriptors.ptr(i), K2.mDescriptors.ptr(i), rnd() % 64);
            K2.mvKeysUn[i].angle = std::fmod(tmp.mvKeys[i].angle + (rnd() % 8 ? 6.f : 180.f * rndf()), 360.f);
            if (rnd() % 5) { K1.mps[i] = &s1[i]; s1[i].bad = rnd() % 25 == 0; }
            if (rnd() % 5) { K2.mps[i] = &s2[i]; s2[i].bad = rnd() % 25 == 0; }
            const unsigned node = (unsigned)(i % 61) * 2;
            K1.mFeatVec[node].push_back(i);
            K2.mFeatVec[rnd() % 10 ? node : node + 2].push_back(i);
        }
        for (int i = 0; i + 61 < tmp.N; i += 7) memcpy(K2.mDescriptors.ptr(i + 61), K2.mDescriptors.ptr(i), 32);
        std::vector<MapPoint *> got, want(tmp.N, nullptr);
        ORB_SLAM2::ORBmatcher matcher(0.75f, true);
        const int ng = matcher.SearchByBoW(&K1, &K2, got);
        int nr = 0;
        {
            std::vector<bool> vbMatched2(tmp.N, false);
            std::vector<int> rotHist[30];
            auto f1 = K1.mFeatVec.begin(); auto f2 = K2.mFeatVec.begin();
            while (f1 != K1.mFeatVec.end() && f2 != K2.mFeatVec.end()) {
                if (f1->first == f2->first) {
                    for (unsigned idx1 : f1->second) {
                        MapPoint *p1 = K1.mps[idx1];
                        if (!p1 || p1->isBad()) continue;
                        int b1 = 256, bi = -1, b2 = 256;
                        for (unsigned idx2 : f2->second) {
                            MapPoint *p2 = K2.mps[idx2];
                            if (vbMatched2[idx2] || !p2) continue;
                            if (p2->isBad()) continue;
                            const int d = ref_descriptor_distance(K1.mDescriptors.ptr((int)idx1), K2.mDescriptors.ptr((int)idx2));
                            if (d < b1) { b2 = b1; b1 = d; bi = (int)idx2; } else if (d < b2) b2 = d;
                        }
                        if (b1 < 50 && (float)b1 < 0.75f * (float)b2) {
                            want[idx1] = K2.mps[bi]; vbMatched2[bi] = true;
                            float rot = K1.mvKeysUn[idx1].angle - K2.mvKeysUn[bi].angle;
                            if (rot < 0.0) rot += 360.0f;
                            int bin = (int)std::round(rot * (1.0f / 30));
                            if (bin == 30) bin = 0;
                            rotHist[bin].push_back((int)idx1);
                            nr++;
                        }
                    }
                    f1++; f2++;
                } else if (f1->first < f2->first) f1 = K1.mFeatVec.lower_bound(f2->first);
                else f2 = K2.mFeatVec.lower_bound(f1->first);
            }
            int m1 = 0, m2 = 0, m3 = 0, i1 = -1, i2 = -1, i3 = -1;
            for (int i = 0; i < 30; i++) {
                const int sz = (int)rotHist[i].size();
                if (sz > m1) { m3 = m2; m2 = m1; m1 = sz; i3 = i2; i2 = i1; i1 = i; }
                else if (sz > m2) { m3 = m2; m2 = sz; i3 = i2; i2 = i; }
                else if (sz > m3) { m3 = sz; i3 = i; }
            }
            if (m2 < 0.1f * (float)m1) { i2 = -1; i3 = -1; } else if (m3 < 0.1f * (float)m1) i3 = -1;
            for (int i = 0; i < 30; i++) {
                if (i == i1 || i == i2 || i == i3) continue;
                for (int j : rotHist[i]) { want[j] = nullptr; nr--; }
            }
        }
        const bool same = ng == nr && got == want;
        printf("%s SearchByBoW(KeyFrame, KeyFrame): %d matches (reference loop %d)\n", same && ng > 200 ? "ok" : "FAIL", ng, nr);
        fails += !(same && ng > 200);
    }
    // ---- M4: SearchByProjection(Frame, KeyFrame, sAlreadyFound, th, ORBdist)  (relocalisation)
    for (int variant = 0; variant < 2; variant++) {
        Frame C1, C2;
        fill_frame(C1, 1500);
        float R[9], t[3] = {0.08f, -0.02f, 0.15f};
        rot_y(0.02f, R);
        C1.mTcw = mat4(R, t, 1.f);
        Frame tmpK;
        fill_frame(tmpK, 1200);
        KeyFrame K;
        keyframe_from(K, tmpK);
        std::vector<MapPoint> store(K.N);
        std::set<MapPoint *> found;
        for (int i = 0; i < K.N; i++) {
            point_near(store[i], C1, rnd() % C1.N, R, t, 4.f);
            if (rnd() % 6) K.mps[i] = &store[i];
            if (rnd() % 12 == 0) found.insert(&store[i]);
        }
        for (int j = 0; j < C1.N; j++) if (rnd() % 4 == 0) C1.mvpMapPoints[j] = &store[rnd() % K.N]; // already claimed features
        C2 = C1; C2.mDescriptors = C1.mDescriptors.clone();
        const float th = variant ? 3.f : 10.f;
        const int ORBdist = variant ? 64 : 100;
        ORB_SLAM2::ORBmatcher matcher(0.9f, true);
        const int ng = matcher.SearchByProjection(C1, &K, found, th, ORBdist);
        int nr = 0;
        {
            float O[3];
            t_centre(R, t, O);
            std::vector<int> rotHist[30];
            for (int i = 0; i < K.N; i++) {
                MapPoint *pMP = K.mps[i];
                if (!pMP || pMP->isBad() || found.count(pMP)) continue;
                const float pw[3] = {pMP->pos.at<float>(0, 0), pMP->pos.at<float>(1, 0), pMP->pos.at<float>(2, 0)};
                float pc[3];
                t_Rx_t(R, pw, t, pc);
                const float invzc = 1.0 / pc[2];
                const float u = C2.fx * pc[0] * invzc + C2.cx, v = C2.fy * pc[1] * invzc + C2.cy;
                if (u < C2.mnMinX || u > C2.mnMaxX || v < C2.mnMinY || v > C2.mnMaxY) continue;
                const float PO[3] = {pw[0] - O[0], pw[1] - O[1], pw[2] - O[2]};
                const float dist3D = t_norm(PO);
                if (dist3D < pMP->GetMinDistanceInvariance() || dist3D > pMP->GetMaxDistanceInvariance()) continue;
                const int lvl = pMP->PredictScale(dist3D, &C2);
                const std::vector<size_t> vi = C2.GetFeaturesInArea(u, v, th * C2.mvScaleFactors[lvl], lvl - 1, lvl + 1);
                if (vi.empty()) continue;
                int bestDist = 256, best = -1;
                for (size_t i2 : vi) {
                    if (C2.mvpMapPoints[i2]) continue;
                    const int d = ref_descriptor_distance(pMP->desc.ptr(), C2.mDescriptors.ptr((int)i2));
                    if (d < bestDist) { bestDist = d; best = (int)i2; }
                }
                if (bestDist <= ORBdist) {
                    C2.mvpMapPoints[best] = pMP; nr++;
                    float rot = K.mvKeysUn[i].angle - C2.mvKeysUn[best].angle;
                    if (rot < 0.0) rot += 360.0f;
                    int bin = (int)std::round(rot * (1.0f / 30));
                    if (bin == 30) bin = 0;
                    rotHist[bin].push_back(best);
                }
            }
            int m1 = 0, m2 = 0, m3 = 0, i1 = -1, i2 = -1, i3 = -1;
            for (int i = 0; i < 30; i++) {
                const int sz = (int)rotHist[i].size();
                if (sz > m1) { m3 = m2; m2 = m1; m1 = sz; i3 = i2; i2 = i1; i1 = i; }
                else if (sz > m2) { m3 = m2; m2 = sz; i3 = i2; i2 = i; }
                else if (sz > m3) { m3 = sz; i3 = i; }
            }
            if (m2 < 0.1f * (float)m1) { i2 = -1; i3 = -1; } else if (m3 < 0.1f * (float)m1) i3 = -1;
            for (int i = 0; i < 30; i++) {
                if (i == i1 || i == i2 || i == i3) continue;
                for (int j : rotHist[i]) { C2.mvpMapPoints[j] = nullptr; nr--; }
            }
        }
        bool same = ng == nr;
        for (int j = 0; j < C1.N; j++) same = same && C1.mvpMapPoints[j] == C2.mvpMapPoints[j];
        printf("%s SearchByProjection(frame, keyframe, th=%g): %d matches (reference loop %d)\n", same && ng > 20 ? "ok" : "FAIL", th, ng, nr);
        fails += !(same && ng > 20);
    }
    // ---- M5 + Fuse(Scw): Sim3 projection of loop map points into a keyframe
    {
        Frame tmp;
        fill_frame(tmp, 1500);
        KeyFrame KA, KB;
        keyframe_from(KA, tmp); keyframe_from(KB, tmp);
        float R[9], t[3] = {0.3f, 0.05f, -0.2f};
        rot_y(-0.03f, R);
        const cv::Mat Scw = mat4(R, t, 1.07f);
        float Rd[9], td[3], Od[3];
        ref_decompose(Scw, Rd, td, Od);
        const int NP = 1400;
        std::vector<MapPoint> storeA(NP + tmp.N), storeB;
        for (int i = 0; i < NP; i++) point_near(storeA[i], KA, rnd() % tmp.N, Rd, td, 3.f);
        for (int j = 0; j < tmp.N; j++) { storeA[NP + j].obs = 1 + rnd() % 5; storeA[NP + j].bad = rnd() % 20 == 0; }
        storeB = storeA;
        std::vector<MapPoint *> vpA, vpB;
        for (int i = 0; i < NP; i++) { vpA.push_back(&storeA[i]); vpB.push_back(&storeB[i]); }
        for (int i = 5; i < NP; i += 40) { vpA[i] = vpA[i - 3]; vpB[i] = vpB[i - 3]; }
        // M5: vpMatched pre-filled for some features, some of the candidates among them
        {
            std::vector<MapPoint *> mA(tmp.N, nullptr), mB(tmp.N, nullptr);
            for (int j = 0; j < tmp.N; j++) if (rnd() % 5 == 0) { const int k = rnd() % NP; mA[j] = &storeA[k]; mB[j] = &storeB[k]; }
            ORB_SLAM2::ORBmatcher matcher(0.75f, true);
            const int ng = matcher.SearchByProjection(&KA, Scw, vpA, mA, 10);
            int nr = 0;
            std::set<MapPoint *> sp(mB.begin(), mB.end());
            sp.erase(nullptr);
            for (size_t i = 0; i < vpB.size(); i++) {
                MapPoint *pMP = vpB[i];
                if (pMP->isBad() || sp.count(pMP)) continue;
                float u, v, dist;
                if (!ref_gate(&KB, Rd, td, Od, pMP, u, v, dist)) continue;
                const int lvl = pMP->PredictScale(dist, &KB);
                const std::vector<size_t> vi = KB.GetFeaturesInArea(u, v, 10 * KB.mvScaleFactors[lvl]);
                if (vi.empty()) continue;
                int bestDist = 256, best = -1;
                for (size_t idx : vi) {
                    if (mB[idx]) continue;
                    const int o = KB.mvKeysUn[idx].octave;
                    if (o < lvl - 1 || o > lvl) continue;
                    const int d = ref_descriptor_distance(pMP->desc.ptr(), KB.mDescriptors.ptr((int)idx));
                    if (d < bestDist) { bestDist = d; best = (int)idx; }
                }
                if (bestDist <= 50) { mB[best] = pMP; nr++; }
            }
            bool same = ng == nr;
            for (int j = 0; same && j < tmp.N; j++) same = (mA[j] ? mA[j] - storeA.data() : -1) == (mB[j] ? mB[j] - storeB.data() : -1);
            printf("%s SearchByProjection(keyframe, Scw): %d matches (reference loop %d)\n", same && ng > 200 ? "ok" : "FAIL", ng, nr);
            fails += !(same && ng > 200);
        }
        // Fuse(Scw)
        {
            for (int j = 0; j < tmp.N; j++) if (rnd() % 3 == 0) { KA.mps[j] = &storeA[NP + j]; KB.mps[j] = &storeB[NP + j]; }
            for (int j = 0; j < tmp.N; j += 17) if (KA.mps[j]) { vpA.push_back(KA.mps[j]); vpB.push_back(KB.mps[j]); } // already in the keyframe
            std::vector<MapPoint *> rA(vpA.size(), nullptr), rB(vpB.size(), nullptr);
            ORB_SLAM2::ORBmatcher matcher(0.8f);
            const int ng = matcher.Fuse(&KA, Scw, vpA, 4.f, rA);
            int nr = 0;
            const std::set<MapPoint *> sp = KB.GetMapPoints();
            for (size_t i = 0; i < vpB.size(); i++) {
                MapPoint *pMP = vpB[i];
                if (pMP->isBad() || sp.count(pMP)) continue;
                float u, v, dist;
                if (!ref_gate(&KB, Rd, td, Od, pMP, u, v, dist)) continue;
                const int lvl = pMP->PredictScale(dist, &KB);
                const std::vector<size_t> vi = KB.GetFeaturesInArea(u, v, 4.f * KB.mvScaleFactors[lvl]);
                if (vi.empty()) continue;
                int bestDist = 0x7fffffff, best = -1;
                for (size_t idx : vi) {
                    const int o = KB.mvKeysUn[idx].octave;
                    if (o < lvl - 1 || o > lvl) continue;
                    const int d = ref_descriptor_distance(pMP->desc.ptr(), KB.mDescriptors.ptr((int)idx));
                    if (d < bestDist) { bestDist = d; best = (int)idx; }
                }
                if (bestDist <= 50) {
                    MapPoint *in = KB.GetMapPoint(best);
                    if (in) { if (!in->isBad()) rB[i] = in; }
                    else { pMP->AddObservation(&KB, best); KB.AddMapPoint(pMP, best); }
                    nr++;
                }
            }
            bool same = ng == nr;
            for (int j = 0; same && j < tmp.N; j++) same = (KA.mps[j] ? KA.mps[j] - storeA.data() : -1) == (KB.mps[j] ? KB.mps[j] - storeB.data() : -1);
            for (size_t i = 0; same && i < rA.size(); i++) same = (rA[i] ? rA[i] - storeA.data() : -1) == (rB[i] ? rB[i] - storeB.data() : -1);
            for (size_t i = 0; same && i < storeA.size(); i++) same = storeA[i].obs == storeB[i].obs;
            printf("%s Fuse(keyframe, Scw): %d fused (reference loop %d)\n", same && ng > 300 ? "ok" : "FAIL", ng, nr);
            fails += !(same && ng > 300);
        }
    }
    // ---- SearchBySim3
    {
        Frame tmp;
        fill_frame(tmp, 1400);
        KeyFrame K1, K2;
        keyframe_from(K1, tmp); keyframe_from(K2, tmp);
        const float s12 = 1.04f;
        float R12[9], t12[3] = {0.25f, -0.03f, 0.1f};
        rot_y(0.025f, R12);
        // KF2's own world pose (KF1 = world): close to, not equal to, the inverse Sim3
        float R2w[9], t2w[3] = {-0.24f, 0.03f, -0.09f};
        rot_y(-0.024f, R2w);
        K2.Rcw = mat3(R2w); K2.tcw = vec3(t2w);
        float sR12[9], sR21[9], t21[3];
        const float inv = (float)(1.0 / s12);
        for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) { sR12[3 * r + c] = R12[3 * r + c] * s12; sR21[3 * r + c] = R12[3 * c + r] * inv; }
        for (int r = 0; r < 3; r++) { float a = sR21[3 * r] * t12[0]; a = a + sR21[3 * r + 1] * t12[1]; a = a + sR21[3 * r + 2] * t12[2]; t21[r] = -a; } // gemm(sR21, t12, -1): float branch
        std::vector<MapPoint> s1(tmp.N), s2(tmp.N);
        const float I[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, z0[3] = {0, 0, 0};
        for (int i = 0; i < tmp.N; i++) {
            // 3-D point seen by KF1 feature i; KF2 feature i is moved onto its Sim3 projection
            const float z = 4.f + 20.f * rndf();
            const float pc1[3] = {(K1.mvKeysUn[i].pt.x - K1.cx) * z / K1.fx, (K1.mvKeysUn[i].pt.y - K1.cy) * z / K1.fy, z};
            float pc2[3];
            t_Rx_t(sR21, pc1, t21, pc2);
            const float u2 = K1.fx * pc2[0] / pc2[2] + K1.cx, v2 = K1.fy * pc2[1] / pc2[2] + K1.cy;
            if (rnd() % 6 && u2 > 1 && u2 < 639 && v2 > 1 && v2 < 479) { K2.mvKeysUn[i].pt.x = u2 + (rndf() - 0.5f); K2.mvKeysUn[i].pt.y = v2 + (rndf() - 0.5f); }
            noisy_copy(K1.mDescriptors.ptr(i), K2.mDescriptors.ptr(i), rnd() % 80);
            MapPoint &a = s1[i], &b = s2[i];
            a.pos = vec3(pc1); // world == camera 1
            float pw2[3];
            world_of(R2w, t2w, pc2, pw2); // KF2's map point: the world point that KF2's pose puts at pc2
            b.pos = vec3(pw2);
            a.maxDist = t_norm(pc2) * std::pow(1.2f, (float)K2.mvKeysUn[i].octave) * (0.93f + 0.14f * rndf());
            b.maxDist = t_norm(pc1) * std::pow(1.2f, (float)K1.mvKeysUn[i].octave) * (0.93f + 0.14f * rndf());
            a.minDist = a.maxDist / std::pow(1.2f, 7.f); b.minDist = b.maxDist / std::pow(1.2f, 7.f);
            a.desc.create(1, 32, CV_8U); b.desc.create(1, 32, CV_8U);
            memcpy(a.desc.ptr(), K1.mDescriptors.ptr(i), 32); memcpy(b.desc.ptr(), K2.mDescriptors.ptr(i), 32);
            a.bad = rnd() % 30 == 0; b.bad = rnd() % 30 == 0;
            if (rnd() % 8) K1.mps[i] = &a;
            if (rnd() % 8) { K2.mps[i] = &b; b.observations.push_back({&K2, (size_t)i}); }
        }
        (void)I; (void)z0;
        std::vector<MapPoint *> got(tmp.N, nullptr), want;
        for (int i = 0; i < tmp.N; i += 9) if (K2.mps[i]) got[i] = K2.mps[i]; // matches already known (from SearchByBoW)
        want = got;
        ORB_SLAM2::ORBmatcher matcher(0.75f, true);
        const int ng = matcher.SearchBySim3(&K1, &K2, got, s12, mat3(R12), vec3(t12), 7.5f);
        int nr = 0;
        {
            std::vector<bool> am1(tmp.N, false), am2(tmp.N, false);
            for (int i = 0; i < tmp.N; i++) if (want[i]) { am1[i] = true; const int j = want[i]->GetIndexInKeyFrame(&K2); if (j >= 0 && j < tmp.N) am2[j] = true; }
            std::vector<int> m1(tmp.N, -1), m2(tmp.N, -1);
            for (int dir = 0; dir < 2; dir++) {
                KeyFrame &Ka = dir ? K2 : K1, &Kb = dir ? K1 : K2;
                const float *Raw = dir ? R2w : I, *taw = dir ? t2w : z0, *sR = dir ? sR12 : sR21, *tt = dir ? t12 : t21;
                std::vector<bool> &am = dir ? am2 : am1;
                std::vector<int> &mm = dir ? m2 : m1;
                for (int i = 0; i < tmp.N; i++) {
                    MapPoint *pMP = Ka.mps[i];
                    if (!pMP || am[i] || pMP->isBad()) continue;
                    const float pw[3] = {pMP->pos.at<float>(0, 0), pMP->pos.at<float>(1, 0), pMP->pos.at<float>(2, 0)};
                    float pa[3], pb[3];
                    t_Rx_t(Raw, pw, taw, pa);
                    t_Rx_t(sR, pa, tt, pb);
                    if (pb[2] < 0.0) continue;
                    const float invz = 1.0 / pb[2], x = pb[0] * invz, y = pb[1] * invz, u = K1.fx * x + K1.cx, v = K1.fy * y + K1.cy;
                    if (!Kb.IsInImage(u, v)) continue;
                    const float dist3D = t_norm(pb);
                    if (dist3D < pMP->GetMinDistanceInvariance() || dist3D > pMP->GetMaxDistanceInvariance()) continue;
                    const int lvl = pMP->PredictScale(dist3D, &Kb);
                    const std::vector<size_t> vi = Kb.GetFeaturesInArea(u, v, 7.5f * Kb.mvScaleFactors[lvl]);
                    if (vi.empty()) continue;
                    int bestDist = 0x7fffffff, best = -1;
                    for (size_t idx : vi) {
                        const int o = Kb.mvKeysUn[idx].octave;
                        if (o < lvl - 1 || o > lvl) continue;
                        const int d = ref_descriptor_distance(pMP->desc.ptr(), Kb.mDescriptors.ptr((int)idx));
                        if (d < bestDist) { bestDist = d; best = (int)idx; }
                    }
                    if (bestDist <= 100) mm[i] = best;
                }
            }
            for (int i1 = 0; i1 < tmp.N; i1++) { const int i2 = m1[i1]; if (i2 >= 0 && m2[i2] == i1) { want[i1] = K2.mps[i2]; nr++; } }
        }
        const bool same = ng == nr && got == want;
        printf("%s SearchBySim3: %d matches (reference loop %d)\n", same && ng > 200 ? "ok" : "FAIL", ng, nr);
        fails += !(same && ng > 200);
    }
    // ---- M1: static DescriptorDistance
    {
        cv::Mat a(1, 32, CV_8U), b(1, 32, CV_8U);
        bool same = true;
        for (int t = 0; t < 1000; t++) {
            for (int k = 0; k < 32; k++) { a.ptr()[k] = (unsigned char)rnd(); b.ptr()[k] = (unsigned char)rnd(); }
            same = same && ORB_SLAM2::ORBmatcher::DescriptorDistance(a, b) == ref_descriptor_distance(a.ptr(), b.ptr());
        }
        printf("%s DescriptorDistance\n", same ? "ok" : "FAIL");
        fails += !same;
    }
    return fails ? 1 : 0;
}

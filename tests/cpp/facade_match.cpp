// facade_match.cpp -- binds the templated Search* methods of include/orbb200/ORBmatcher.h to small stand-ins of
// ORB_SLAM2::Frame / KeyFrame / MapPoint (same member names as the reference classes) and checks them against the
// sequential reference loops evaluated with the CPU oracle's DescriptorDistance (oracle/liborb_ref.so: checker only).
//   facade_match  -> prints "ok ..." lines, exit code 0 on parity
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <type_traits>
#include <vector>

#include "orbb200/ORBmatcher.h"

extern "C" int ref_descriptor_distance(const unsigned char *a, const unsigned char *b); // oracle/orb_ref.c

static unsigned long long g_s = 88172645463325252ull;
static unsigned rnd() { g_s ^= g_s << 13; g_s ^= g_s >> 7; g_s ^= g_s << 17; return (unsigned)(g_s >> 11); }
static float rndf() { return (rnd() & 0xFFFFFF) / (float)0x1000000; }

struct MapPoint {
    bool mbTrackInView = true, bad = false;
    int mnTrackScaleLevel = 0, obs = 1;
    float mTrackViewCos = 1.f, mTrackProjX = 0, mTrackProjY = 0, mTrackProjXR = 0;
    cv::Mat desc;
    bool isBad() { return bad; }
    int Observations() { return obs; }
    cv::Mat GetDescriptor() { return desc.clone(); }
};

struct Frame {
    int N = 0;
    std::vector<cv::KeyPoint> mvKeys, mvKeysUn;
    std::vector<float> mvuRight, mvScaleFactors;
    std::vector<MapPoint *> mvpMapPoints;
    cv::Mat mDescriptors;
    std::map<unsigned int, std::vector<unsigned int>> mFeatVec; // DBoW2::FeatureVector is a std::map of index lists
    // stand-in of Frame::GetFeaturesInArea (src/Frame.cc:386-442): index order, square window, level band
    std::vector<size_t> GetFeaturesInArea(const float &x, const float &y, const float &r, const int minLevel, const int maxLevel) const
    {
        std::vector<size_t> v;
        const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
        for (int i = 0; i < N; i++) {
            const cv::KeyPoint &kp = mvKeysUn[i];
            if (bCheckLevels) {
                if (kp.octave < minLevel) continue;
                if (maxLevel >= 0 && kp.octave > maxLevel) continue;
            }
            if (std::fabs(kp.pt.x - x) < r && std::fabs(kp.pt.y - y) < r) v.push_back(i);
        }
        return v;
    }
};

struct KeyFrame {
    std::vector<MapPoint *> mps;
    std::vector<cv::KeyPoint> mvKeysUn;
    cv::Mat mDescriptors;
    std::map<unsigned int, std::vector<unsigned int>> mFeatVec;
    std::vector<MapPoint *> GetMapPointMatches() { return mps; }
};

static void fill_frame(Frame &F, int N)
{
    F.N = N;
    F.mvKeys.resize(N); F.mvuRight.assign(N, -1.f); F.mvpMapPoints.assign(N, nullptr);
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
    // ---- M6: SearchByBoW(KeyFrame, Frame)
    {
        Frame F;
        fill_frame(F, 1400);
        KeyFrame KF;
        const int NK = 1300;
        KF.mvKeysUn.resize(NK); KF.mps.assign(NK, nullptr); KF.mDescriptors.create(NK, 32, CV_8U);
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

// facade_distinct.cpp -- test driver of include/orbb200/MapPointDescriptors.h.  Stand-in KeyFrame / MapPoint types carry the
// members MapPoint::ComputeDistinctiveDescriptors touches (src/MapPoint.cc:266-340: mObservations, KeyFrame::mDescriptors,
// isBad, mbBad, mDescriptor); every point is updated twice -- by the reference loop written out below with the host
// DescriptorDistance, and by ComputeDistinctiveDescriptorsBatch (one device call) -- and the results are compared.
//   facade_distinct.bin <seed> <n_keyframes> <n_points>      prints "OK <points> <updated>" or the first mismatch
#include <algorithm>
#include <climits>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <mutex>
#include <vector>

#include "orbb200/MapPointDescriptors.h"
#include "orbb200/ORBmatcher.h"

struct KeyFrame {
    cv::Mat mDescriptors;
    bool bad;
    bool isBad() const { return bad; }
};

struct MapPoint {
    std::map<KeyFrame *, size_t> mObservations;
    bool mbBad;
    cv::Mat mDescriptor;
    std::mutex mMutexFeatures;
    MapPoint() : mbBad(false) {}

    // the gather of src/MapPoint.cc:270-299
    bool CollectObservedDescriptors(std::vector<cv::Mat> &vDescriptors)
    {
        std::map<KeyFrame *, size_t> observations;
        {
            std::unique_lock<std::mutex> lock1(mMutexFeatures);
            if (mbBad) return false;
            observations = mObservations;
        }
        vDescriptors.clear();
        vDescriptors.reserve(observations.size());
        for (std::map<KeyFrame *, size_t>::iterator mit = observations.begin(), mend = observations.end(); mit != mend; mit++) {
            KeyFrame *pKF = mit->first;
            if (!pKF->isBad()) vDescriptors.push_back(pKF->mDescriptors.row((int)mit->second));
        }
        return true;
    }
    // the locked assignment of :336-339
    void SetDescriptor(const cv::Mat &d)
    {
        std::unique_lock<std::mutex> lock(mMutexFeatures);
        mDescriptor = d;
    }
    // the reference function, statement for statement, on the stand-in members
    void ReferenceComputeDistinctiveDescriptors()
    {
        std::vector<cv::Mat> vDescriptors;
        if (!CollectObservedDescriptors(vDescriptors)) return;
        if (vDescriptors.empty()) return;
        const size_t N = vDescriptors.size();
        std::vector<std::vector<float> > Distances(N, std::vector<float>(N));
        for (size_t i = 0; i < N; i++) {
            Distances[i][i] = 0;
            for (size_t j = i + 1; j < N; j++) {
                int distij = ORB_SLAM2::ORBmatcher::DescriptorDistance(vDescriptors[i], vDescriptors[j]);
                Distances[i][j] = distij;
                Distances[j][i] = distij;
            }
        }
        int BestMedian = INT_MAX;
        int BestIdx = 0;
        for (size_t i = 0; i < N; i++) {
            std::vector<int> vDists(Distances[i].begin(), Distances[i].end());
            std::sort(vDists.begin(), vDists.end());
            int median = vDists[0.5 * (N - 1)];
            if (median < BestMedian) { BestMedian = median; BestIdx = i; }
        }
        SetDescriptor(vDescriptors[BestIdx].clone());
    }
};

static unsigned long long g_state;
static unsigned rnd() { g_state = g_state * 6364136223846793005ULL + 1442695040888963407ULL; return (unsigned)(g_state >> 33); }

int main(int argc, char **argv)
{
    if (argc < 4) return 2;
    g_state = strtoull(argv[1], 0, 10);
    const int nKF = atoi(argv[2]), nMP = atoi(argv[3]), perKF = nMP; // a keyframe observes a point at most once
    // a landmark's base descriptor; every keyframe row observing it is a noisy copy
    std::vector<std::vector<unsigned char> > base(nMP, std::vector<unsigned char>(32));
    for (int p = 0; p < nMP; p++) for (int b = 0; b < 32; b++) base[p][b] = (unsigned char)rnd();
    std::vector<KeyFrame> kfs(nKF);
    for (int k = 0; k < nKF; k++) { kfs[k].mDescriptors.create(perKF, 32, CV_8U); kfs[k].bad = (rnd() % 9) == 0; }
    std::vector<int> used(nKF, 0);
    std::vector<MapPoint> A(nMP), B(nMP);
    for (int p = 0; p < nMP; p++) {
        const int nobs = (p % 17 == 0) ? 0 : 1 + (int)(rnd() % (unsigned)std::min(nKF, 45));
        for (int o = 0; o < nobs; o++) {
            const int k = (int)(rnd() % (unsigned)nKF);
            if (used[k] >= perKF || A[p].mObservations.count(&kfs[k])) continue;
            unsigned char *row = kfs[k].mDescriptors.ptr(used[k]);
            for (int b = 0; b < 32; b++) row[b] = base[p][b];
            const int flips = (int)(rnd() % 40);
            for (int f = 0; f < flips; f++) { const unsigned bit = rnd() % 256; row[bit >> 3] ^= (unsigned char)(1u << (bit & 7)); }
            A[p].mObservations[&kfs[k]] = (size_t)used[k];
            B[p].mObservations[&kfs[k]] = (size_t)used[k];
            used[k]++;
        }
        A[p].mbBad = B[p].mbBad = (p % 23 == 5);
    }
    for (int p = 0; p < nMP; p++) A[p].ReferenceComputeDistinctiveDescriptors();
    std::vector<MapPoint *> pts;
    for (int p = 0; p < nMP; p++) pts.push_back(&B[p]);
    pts.push_back(0); // null entries of mvpMapPoints are skipped by the callers' loops
    ORB_SLAM2::ComputeDistinctiveDescriptorsBatch(pts);
    int updated = 0;
    for (int p = 0; p < nMP; p++) {
        if (A[p].mDescriptor.empty() != B[p].mDescriptor.empty()) { printf("MISMATCH point %d: updated %d vs %d\n", p, !A[p].mDescriptor.empty(), !B[p].mDescriptor.empty()); return 1; }
        if (A[p].mDescriptor.empty()) continue;
        updated++;
        if (memcmp(A[p].mDescriptor.ptr(), B[p].mDescriptor.ptr(), 32) != 0) { printf("MISMATCH point %d: different descriptor\n", p); return 1; }
    }
    printf("OK %d %d\n", nMP, updated);
    return 0;
}

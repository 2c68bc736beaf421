// facade_device_frame.cpp -- orbb200::DeviceFrame (include/orbb200/DeviceFrame.h): a Frame whose extraction outputs stay on the
// device is searched by the SAME facade ORBmatcher with the SAME results as a host Frame, and nothing of the frame is uploaded
// again -- the host-to-device byte counter between extraction and the searches moves by the windows + query descriptors only.
//   facade_device_frame <w> <h> <image.raw>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <string>
#include <vector>

#include "orbb200/ORBextractor.h"
#include "orbb200/ORBmatcher.h"
#include "orbb200/ORBVocabulary.h"

static unsigned long long g_s = 88172645463325252ull;
static unsigned rnd() { g_s ^= g_s << 13; g_s ^= g_s >> 7; g_s ^= g_s << 17; return (unsigned)(g_s >> 11); }
static float rndf() { return (rnd() & 0xFFFFFF) / (float)0x1000000; }

struct MapPoint { // the members ORBmatcher::SearchByProjection(F, points, th) and (Cur, Last, th, mono) read (include/MapPoint.h)
    bool mbTrackInView = true, bad = false;
    int mnTrackScaleLevel = 0, obs = 1;
    float mTrackViewCos = 1.f, mTrackProjX = 0, mTrackProjY = 0, mTrackProjXR = -1;
    cv::Mat desc, pos;
    bool isBad() { return bad; }
    int Observations() { return obs; }
    cv::Mat GetDescriptor() { return desc.clone(); }
    cv::Mat GetWorldPos() { return pos.clone(); }
};
struct Frame { // include/Frame.h, the members the windowed searches read
    int N = 0;
    std::vector<cv::KeyPoint> mvKeys, mvKeysUn;
    std::vector<float> mvuRight, mvScaleFactors;
    std::vector<MapPoint *> mvpMapPoints;
    std::vector<bool> mvbOutlier;
    cv::Mat mDescriptors, mTcw;
    float mb = 0.5372f, mbf = 386.1448f, fx = 718.856f, fy = 718.856f, cx = 320.f, cy = 240.f;
    float mnMinX = 0, mnMinY = 0, mnMaxX = 640, mnMaxY = 480, mfGridElementWidthInv = 64.f / 640.f, mfGridElementHeightInv = 48.f / 480.f;
    orbb200::DeviceFrame *mpDeviceFrame = nullptr; // <- the one member an integrator adds to keep the frame on the device
};

static int check(bool ok, const char *what)
{
    printf("%s %s\n", ok ? "ok" : "FAIL", what);
    return ok ? 0 : 1;
}

int main(int argc, char **argv)
{
    if (argc < 4) return 2;
    const int w = atoi(argv[1]), h = atoi(argv[2]);
    cv::Mat im(h, w, CV_8UC1);
    FILE *f = fopen(argv[3], "rb");
    if (!f || fread(im.data, 1, (size_t)w * h, f) != (size_t)w * h) return 3;
    fclose(f);
    int fails = 0;
    ORB_SLAM2::ORBextractor ex(1000, 1.2f, 8, 20, 7);
    Frame F;
    ex(im, cv::Mat(), F.mvKeys, F.mDescriptors);
    F.N = (int)F.mvKeys.size();
    orbb200::DeviceFrame dev;
    fails += check(dev.Adopt(ex, F.N) && dev.N() == F.N && F.N > 500, "Adopt: extraction outputs kept on the device");
    // distorted camera (TUM1): Frame::UndistortKeyPoints on the device copy == the host-buffer call
    cv::Mat mK = cv::Mat::zeros(3, 3, CV_32F), mDistCoef(5, 1, CV_32F);
    mK.at<float>(0, 0) = 517.306408f; mK.at<float>(1, 1) = 516.469215f; mK.at<float>(0, 2) = 318.643040f; mK.at<float>(1, 2) = 255.313989f; mK.at<float>(2, 2) = 1.f;
    const float dc[5] = {0.262383f, -0.953104f, -0.005358f, 0.002628f, 1.163314f};
    for (int i = 0; i < 5; i++) mDistCoef.at<float>(i, 0) = dc[i];
    orbb200::UndistortKeyPoints(F.mvKeys, mK, mDistCoef, F.mvKeysUn);
    fails += check(dev.Undistort(mK, mDistCoef), "Undistort on the device copy");
    {
        const std::vector<cv::KeyPoint> &k = dev.Keys(), &ku = dev.KeysUn();
        bool same = k.size() == F.mvKeys.size() && ku.size() == F.mvKeysUn.size();
        for (size_t i = 0; same && i < k.size(); i++)
            same = !memcmp(&k[i], &F.mvKeys[i], sizeof(cv::KeyPoint)) && !memcmp(&ku[i], &F.mvKeysUn[i], sizeof(cv::KeyPoint));
        const cv::Mat &d = dev.Descriptors();
        for (int i = 0; same && i < F.N; i++) same = !memcmp(d.ptr(i), F.mDescriptors.ptr(i), 32);
        fails += check(same, "lazy host views (Keys, KeysUn, Descriptors) equal the extractor's outputs / the host undistortion");
    }
    F.mnMaxX = (float)w; F.mnMaxY = (float)h; F.mfGridElementWidthInv = 64.f / w; F.mfGridElementHeightInv = 48.f / h;
    F.mvuRight.assign(F.N, -1.f); F.mvpMapPoints.assign(F.N, nullptr); F.mvbOutlier.assign(F.N, false);
    F.mvScaleFactors.resize(8);
    F.mvScaleFactors[0] = 1.f;
    for (int l = 1; l < 8; l++) F.mvScaleFactors[l] = F.mvScaleFactors[l - 1] * 1.2f;
    for (int i = 0; i < F.N; i++) if (rnd() % 3 == 0) F.mvuRight[i] = F.mvKeysUn[i].pt.x - rndf() * 40;
    // local-map points projecting near the frame's features (Tracking::SearchLocalPoints, src/Tracking.cc:1463)
    std::vector<MapPoint> store(900);
    std::vector<MapPoint *> vp;
    for (size_t i = 0; i < store.size(); i++) {
        MapPoint &m = store[i];
        const int src = rnd() % F.N;
        m.mTrackProjX = F.mvKeysUn[src].pt.x + (rndf() - 0.5f) * 4;
        m.mTrackProjY = F.mvKeysUn[src].pt.y + (rndf() - 0.5f) * 4;
        m.mTrackProjXR = F.mvuRight[src] > 0 ? F.mvuRight[src] + (rndf() - 0.5f) * 6 : -1;
        m.mnTrackScaleLevel = std::min(7, F.mvKeysUn[src].octave + (rnd() % 4 == 0 ? 1 : 0));
        m.mTrackViewCos = rnd() % 2 ? 0.9995f : 0.9f;
        m.mbTrackInView = rnd() % 10 != 0;
        m.desc.create(1, 32, CV_8U);
        memcpy(m.desc.ptr(), F.mDescriptors.ptr(src), 32);
        for (int k = 0; k < (int)(rnd() % 60); k++) { const unsigned b = rnd() % 256; m.desc.ptr()[b >> 3] ^= (unsigned char)(1u << (b & 7)); }
        vp.push_back(&m);
    }
    Frame G = F; // the same frame, searched through its device copy
    G.mDescriptors = F.mDescriptors.clone();
    G.mpDeviceFrame = &dev;
    ORB_SLAM2::ORBmatcher matcher(0.8f);
    const unsigned long long b0 = orbb200_h2d_bytes_total();
    const int nHost = matcher.SearchByProjection(F, vp, 3.f);
    const unsigned long long b1 = orbb200_h2d_bytes_total();
    const int nDev = matcher.SearchByProjection(G, vp, 3.f);
    const unsigned long long b2 = orbb200_h2d_bytes_total();
    bool same = nHost == nDev && nHost > 200;
    for (int i = 0; i < F.N; i++) same = same && F.mvpMapPoints[i] == G.mvpMapPoints[i];
    fails += check(same, "SearchByProjection(frame, local map): device-resident frame == host frame");
    const unsigned long long hostBytes = b1 - b0, devBytes = b2 - b1;
    printf("h2d bytes: host frame %llu, device frame %llu (frame itself: %llu)\n", hostBytes, devBytes, (unsigned long long)F.N * 60ull);
    fails += check(hostBytes == devBytes + (unsigned long long)F.N * 60ull && devBytes % 52 == 0 && devBytes > 0,
                   "the device-resident search uploads windows + query descriptors only (52 bytes per query, nothing of the frame)");
    // Frame::ComputeBoW (src/Frame.cc:459-466) on the resident descriptors == the host-buffer transform of the same vocabulary
    {
        const int k = 6, L = 3; // complete tree: 1 + 6 + 36 + 216 nodes, words = the 216 leaves
        std::vector<int> parent(1, -1);
        std::vector<unsigned char> leaf(1, 0);
        std::vector<int> level(1, 0);
        for (size_t n = 0; n < parent.size(); n++)
            if (level[n] < L)
                for (int c = 0; c < k; c++) { parent.push_back((int)n); level.push_back(level[n] + 1); leaf.push_back(level[n] + 1 == L); }
        const int nn = (int)parent.size();
        std::vector<unsigned char> vdesc((size_t)nn * 32);
        std::vector<double> w(nn, 0.0);
        for (int n = 1; n < nn; n++) {
            memcpy(&vdesc[(size_t)n * 32], F.mDescriptors.ptr((int)(rnd() % F.N)), 32); // centres near real descriptors
            for (int j = 0; j < 20; j++) { const unsigned bit = rnd() % 256; vdesc[(size_t)n * 32 + (bit >> 3)] ^= (unsigned char)(1u << (bit & 7)); }
            w[n] = leaf[n] ? 0.25 + rndf() * 3.0 : 0.0;
        }
        parent[0] = 0;
        orbb200::ORBVocabulary voc;
        voc.create(k, L, 0 /* L1 */, 0 /* TF_IDF */, nn, parent.data(), leaf.data(), vdesc.data(), w.data());
        DBoW2::BowVector v1, v2;
        DBoW2::FeatureVector f1, f2;
        voc.transform(F.mDescriptors, v1, f1, 2);
        const unsigned long long c0 = orbb200_h2d_bytes_total();
        const bool okBow = dev.ComputeBoW(voc, v2, f2, 2);
        const unsigned long long c1 = orbb200_h2d_bytes_total();
        bool sameBow = okBow && v1.size() == v2.size() && f1.size() == f2.size() && v1.size() > 50 && f1.size() > 3;
        for (DBoW2::BowVector::const_iterator a = v1.begin(), b2 = v2.begin(); sameBow && a != v1.end(); ++a, ++b2)
            sameBow = a->first == b2->first && a->second == b2->second;
        for (DBoW2::FeatureVector::const_iterator a = f1.begin(), b2 = f2.begin(); sameBow && a != f1.end(); ++a, ++b2)
            sameBow = a->first == b2->first && a->second == b2->second;
        fails += check(sameBow && c1 - c0 <= 16, "ComputeBoW on the resident descriptors == the host-buffer transform (words, values, feature lists); nothing uploaded");
    }
    printf("%s: %d check(s) failed\n", fails ? "FAIL" : "ok", fails);
    return fails ? 1 : 0;
}

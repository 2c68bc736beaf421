// facade_extract.cpp -- drives the C++ facade (include/orbb200/ORBextractor.h) exactly like Frame::ExtractORB does
// (reference src/Frame.cc:285-292, 73-79) and dumps the results for tests/test_gpu_facade.py.
//   facade_extract <w> <h> <nfeat> <scale> <nlevels> <ini> <min> <left.raw> <right.raw|-> <out_prefix> [bf b]
#include <cstdio>
#include <cstdlib>
#include <string>
#include <thread>
#include <vector>

#include "orbb200/ORBextractor.h"

using ORB_SLAM2::ORBextractor;

static bool read_raw(const char *path, cv::Mat &m)
{
    FILE *f = fopen(path, "rb");
    if (!f) return false;
    const size_t n = fread(m.data, 1, (size_t)m.rows * m.cols, f);
    fclose(f);
    return n == (size_t)m.rows * m.cols;
}

static void dump(const std::string &path, const void *p, size_t bytes)
{
    FILE *f = fopen(path.c_str(), "wb");
    if (bytes) fwrite(p, 1, bytes, f);
    fclose(f);
}

int main(int argc, char **argv)
{
    if (argc < 11) { fprintf(stderr, "usage\n"); return 2; }
    const int w = atoi(argv[1]), h = atoi(argv[2]), nfeat = atoi(argv[3]);
    const float scale = (float)atof(argv[4]);
    const int nlevels = atoi(argv[5]), ini = atoi(argv[6]), mn = atoi(argv[7]);
    const bool stereo = std::string(argv[9]) != "-";
    const std::string out = argv[10];
    cv::Mat imLeft(h, w, CV_8UC1), imRight(h, w, CV_8UC1);
    if (!read_raw(argv[8], imLeft)) return 3;
    if (stereo && !read_raw(argv[9], imRight)) return 3;

    // Tracking::Tracking (src/Tracking.cc:124-127): one extractor per camera
    ORBextractor *mpORBextractorLeft = new ORBextractor(nfeat, scale, nlevels, ini, mn);
    ORBextractor *mpORBextractorRight = stereo ? new ORBextractor(nfeat, scale, nlevels, ini, mn) : nullptr;

    // Frame::Frame (src/Frame.cc:73-79)
    const int mnScaleLevels = mpORBextractorLeft->GetLevels();
    const float mfScaleFactor = mpORBextractorLeft->GetScaleFactor();
    std::vector<float> mvScaleFactors = mpORBextractorLeft->GetScaleFactors();
    std::vector<float> mvInvScaleFactors = mpORBextractorLeft->GetInverseScaleFactors();
    std::vector<float> mvLevelSigma2 = mpORBextractorLeft->GetScaleSigmaSquares();
    std::vector<float> mvInvLevelSigma2 = mpORBextractorLeft->GetInverseScaleSigmaSquares();

    std::vector<cv::KeyPoint> mvKeys, mvKeysRight;
    cv::Mat mDescriptors, mDescriptorsRight;
    // Frame.cc:82-85: two host threads, one extractor each
    std::thread threadLeft([&] { (*mpORBextractorLeft)(imLeft, cv::Mat(), mvKeys, mDescriptors); });
    if (stereo) {
        std::thread threadRight([&] { (*mpORBextractorRight)(imRight, cv::Mat(), mvKeysRight, mDescriptorsRight); });
        threadRight.join();
    }
    threadLeft.join();

    // empty image: outputs untouched (src/ORBextractor.cc:1264)
    {
        std::vector<cv::KeyPoint> k(3);
        cv::Mat d(2, 32, CV_8U), empty;
        (*mpORBextractorLeft)(empty, cv::Mat(), k, d);
        if (k.size() != 3 || d.rows != 2) { fprintf(stderr, "empty-image contract broken\n"); return 4; }
        (*mpORBextractorLeft)(imLeft, cv::Mat(), k, d); // restore the left state for the stereo matcher / pyramid dump
    }

    dump(out + "_kps.bin", mvKeys.data(), mvKeys.size() * sizeof(cv::KeyPoint));
    if (!mDescriptors.empty()) dump(out + "_desc.bin", mDescriptors.data, (size_t)mDescriptors.rows * 32);
    else dump(out + "_desc.bin", nullptr, 0);
    std::vector<float> tabs;
    tabs.push_back((float)mnScaleLevels);
    tabs.push_back(mfScaleFactor);
    for (auto *v : {&mvScaleFactors, &mvInvScaleFactors, &mvLevelSigma2, &mvInvLevelSigma2}) tabs.insert(tabs.end(), v->begin(), v->end());
    dump(out + "_tables.bin", tabs.data(), tabs.size() * 4);
    for (int l = 0; l < mnScaleLevels; l++) {
        const cv::Mat &m = mpORBextractorLeft->mvImagePyramid[l];
        std::vector<unsigned char> buf((size_t)m.rows * m.cols);
        for (int r = 0; r < m.rows; r++) memcpy(buf.data() + (size_t)r * m.cols, m.ptr(r), m.cols);
        dump(out + "_pyr" + std::to_string(l) + ".bin", buf.data(), buf.size());
    }
    if (stereo && argc >= 13) {
        dump(out + "_kpsR.bin", mvKeysRight.data(), mvKeysRight.size() * sizeof(cv::KeyPoint));
        const float mbf = (float)atof(argv[11]), mb = (float)atof(argv[12]);
        std::vector<float> mvuRight, mvDepth;
        const int nv = orbb200::ComputeStereoMatches(*mpORBextractorLeft, *mpORBextractorRight, (int)mvKeys.size(), mbf, mb, mvuRight, mvDepth);
        if (nv < 0) return 5;
        dump(out + "_uright.bin", mvuRight.data(), mvuRight.size() * 4);
        dump(out + "_depth.bin", mvDepth.data(), mvDepth.size() * 4);
        printf("stereo_matches %d\n", nv);
    }
    printf("keypoints %zu\n", mvKeys.size());
    delete mpORBextractorLeft;
    delete mpORBextractorRight;
    return 0;
}

// facade_extract.cpp -- drives the C++ facade (include/orbb200/ORBextractor.h) exactly like Frame::ExtractORB does
// (reference src/Frame.cc:285-292, 73-79) and dumps the results for tests/test_gpu_facade.py.
//   facade_extract <w> <h> <nfeat> <scale> <nlevels> <ini> <min> <left.raw> <right.raw|-> <out_prefix> [bf b]
#include <algorithm>
#include <chrono>
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

// Tracking::GrabImageMonocular on a colour frame (src/Tracking.cc:255-266) + Frame::UndistortKeyPoints (src/Frame.cc:470-515):
//   facade_extract --color <w> <h> <channels> <bRGB> <color.raw> <out_prefix> <fx> <fy> <cx> <cy> <k1> <k2> <p1> <p2> <k3>
static int color_main(int argc, char **argv)
{
    if (argc < 17) return 2;
    const int w = atoi(argv[2]), h = atoi(argv[3]), ch = atoi(argv[4]);
    const bool bRGB = atoi(argv[5]) != 0;
    const std::string out = argv[7];
    cv::Mat im(h, w, ch == 3 ? CV_8UC3 : CV_8UC4);
    FILE *f = fopen(argv[6], "rb");
    if (!f || fread(im.data, 1, (size_t)w * h * ch, f) != (size_t)w * h * ch) return 3;
    fclose(f);
    ORBextractor *mpIniORBextractor = new ORBextractor(2000, 1.2f, 8, 20, 7); // Tracking.cc:130: 2 * nFeatures for the mono initialiser
    std::vector<cv::KeyPoint> mvKeys, mvKeysUn;
    cv::Mat mDescriptors, mImGray;
    mpIniORBextractor->ExtractColor(im, bRGB, mvKeys, mDescriptors, &mImGray);
    cv::Mat mK = cv::Mat::zeros(3, 3, CV_32F), mDistCoef(5, 1, CV_32F);
    mK.at<float>(0, 0) = (float)atof(argv[8]); mK.at<float>(1, 1) = (float)atof(argv[9]);
    mK.at<float>(0, 2) = (float)atof(argv[10]); mK.at<float>(1, 2) = (float)atof(argv[11]); mK.at<float>(2, 2) = 1.f;
    for (int i = 0; i < 5; i++) mDistCoef.at<float>(i, 0) = (float)atof(argv[12 + i]);
    orbb200::UndistortKeyPoints(mvKeys, mK, mDistCoef, mvKeysUn);
    dump(out + "_kps.bin", mvKeys.data(), mvKeys.size() * sizeof(cv::KeyPoint));
    dump(out + "_kpsun.bin", mvKeysUn.data(), mvKeysUn.size() * sizeof(cv::KeyPoint));
    dump(out + "_desc.bin", mDescriptors.data, (size_t)mDescriptors.rows * 32);
    dump(out + "_gray.bin", mImGray.data, (size_t)w * h);
    printf("N %zu\n", mvKeys.size());
    return 0;
}

// Latency of the drop-in's real call pattern (src/Frame.cc:285-292: one image per call; :82-85: left and right on two host threads):
//   facade_extract --latency <w> <h> <nfeat> <nlevels> <left.raw> <right.raw|-> <iters> <graph 0|1>
// prints one JSON object: p50 / p99 / mean microseconds of ORBextractor::operator() (mono) or of the two-thread pair (stereo),
// mvImagePyramid left on the device (lazy) -- plus the cost of the first pyramid read.
static int latency_main(int argc, char **argv)
{
    if (argc < 10) return 2;
    const int w = atoi(argv[2]), h = atoi(argv[3]), nfeat = atoi(argv[4]), nlevels = atoi(argv[5]);
    const bool stereo = std::string(argv[7]) != "-";
    const int iters = atoi(argv[8]), graph = atoi(argv[9]);
    cv::Mat imLeft(h, w, CV_8UC1), imRight(h, w, CV_8UC1);
    if (!read_raw(argv[6], imLeft)) return 3;
    if (stereo && !read_raw(argv[7], imRight)) return 3;
    ORBextractor exL(nfeat, 1.2f, nlevels, 20, 7), exR(nfeat, 1.2f, nlevels, 20, 7);
    orbb200_extractor_set_graph(exL.Handle(), graph);
    orbb200_extractor_set_graph(exR.Handle(), graph);
    std::vector<cv::KeyPoint> kL, kR;
    cv::Mat dL, dR;
    std::vector<double> us;
    double pyrUs = 0;
    size_t n = 0;
    for (int it = -20; it < iters; it++) { // 20 warm-up calls (graph capture, allocations)
        const auto t0 = std::chrono::steady_clock::now();
        if (stereo) {
            std::thread tl([&] { exL(imLeft, cv::Mat(), kL, dL); });
            std::thread tr([&] { exR(imRight, cv::Mat(), kR, dR); });
            tl.join();
            tr.join();
        } else {
            exL(imLeft, cv::Mat(), kL, dL);
        }
        const auto t1 = std::chrono::steady_clock::now();
        if (it >= 0) us.push_back(std::chrono::duration<double, std::micro>(t1 - t0).count());
        n = kL.size();
        if (it == -1) (void)exL.mvImagePyramid[0].rows; // first read allocates the handle's pinned slab buffer: keep that out of the figure
        if (it == iters - 1) {
            if (exL.mvImagePyramid.resident_on_host()) { fprintf(stderr, "lazy pyramid was downloaded without being read\n"); return 4; }
            const auto p0 = std::chrono::steady_clock::now();
            const int rows = exL.mvImagePyramid[nlevels - 1].rows;
            pyrUs = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - p0).count();
            if (rows <= 0 || !exL.mvImagePyramid.resident_on_host()) return 4;
        }
    }
    std::sort(us.begin(), us.end());
    double mean = 0;
    for (double v : us) mean += v;
    mean /= us.size();
    printf("{\"call\": \"%s\", \"shape\": \"%dx%d\", \"nfeatures\": %d, \"graph\": %d, \"iters\": %d, \"p50_us\": %.1f, \"p99_us\": %.1f, "
           "\"mean_us\": %.1f, \"min_us\": %.1f, \"keypoints\": %zu, \"first_pyramid_read_us\": %.1f}\n",
           stereo ? "ORBextractor::operator() x2, two host threads (Frame.cc:82-85)" : "ORBextractor::operator()", w, h, nfeat, graph, iters,
           us[us.size() / 2], us[std::min(us.size() - 1, (size_t)(us.size() * 0.99))], mean, us[0], n, pyrUs);
    return 0;
}

int main(int argc, char **argv)
{
    if (argc > 1 && std::string(argv[1]) == "--color") return color_main(argc, argv);
    if (argc > 1 && std::string(argv[1]) == "--latency") return latency_main(argc, argv);
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

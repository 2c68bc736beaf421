// loader_test.cpp -- drives include/orbb200/DatasetLoader.h on a KITTI-layout sequence directory the way
// Examples/Stereo/stereo_kitti.cc drives cv::imread (LoadImages :130-160, main loop :60-100), batched:
//   loader_test <sequence_dir> <pairs_per_batch> <out.bin> [--extract nfeatures]
// writes every decoded batch ([pair][L, R][h][w] bytes) to out.bin; with --extract also runs orbb200_extract_stereo_batch on each
// batch and prints the keypoint counts (GPU needed).  Also prints the TUM list of <sequence_dir>/rgb.txt when it exists.
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "orbb200/DatasetLoader.h"
#include "orbb200/Rectify.h"

int main(int argc, char **argv)
{
    // loader_test --euroc <left dir> <right dir> <times file> <calib.txt> <out.bin>: the EuRoC list (stereo_euroc.cc:192-215), and
    // the first pair rectified as :96-97, 131-132 do (calib.txt: K D(5) R P(3x4) of the left then of the right camera, cols rows)
    if (argc >= 7 && std::string(argv[1]) == "--euroc") {
        std::vector<std::string> L, R;
        std::vector<double> ts;
        orbb200::LoadImagesEuRoC(argv[2], argv[3], argv[4], L, R, ts);
        for (size_t i = 0; i < L.size(); i++) printf("euroc %.9f %s %s\n", ts[i], L[i].c_str(), R[i].c_str());
        FILE *cf = fopen(argv[5], "r");
        if (!cf) return 3;
        double v[2][35];
        int cols = 0, rows = 0;
        for (int s = 0; s < 2; s++)
            for (int i = 0; i < 35; i++)
                if (fscanf(cf, "%lf", &v[s][i]) != 1) return 3;
        if (fscanf(cf, "%d %d", &cols, &rows) != 2) return 3;
        fclose(cf);
        cv::Mat m[2][4];
        for (int s = 0; s < 2; s++) {
            m[s][0].create(3, 3, CV_64F); m[s][1].create(1, 5, CV_64F); m[s][2].create(3, 3, CV_64F); m[s][3].create(3, 4, CV_64F);
            memcpy(m[s][0].data, &v[s][0], 72); memcpy(m[s][1].data, &v[s][9], 40); memcpy(m[s][2].data, &v[s][14], 72); memcpy(m[s][3].data, &v[s][23], 96);
        }
        orbb200::StereoRectifier rect(m[0][0], m[0][1], m[0][2], m[0][3], m[1][0], m[1][1], m[1][2], m[1][3], cols, rows);
        if (!rect.ok() || L.empty()) return 4;
        std::vector<unsigned char> pl, pr;
        int w, h, ch, w2, h2, ch2;
        std::string err;
        if (!orbb200::ReadImage(L[0], pl, w, h, ch, &err) || !orbb200::ReadImage(R[0], pr, w2, h2, ch2, &err)) { printf("error %s\n", err.c_str()); return 3; }
        cv::Mat imL(h, w, CV_8UC1, pl.data()), imR(h2, w2, CV_8UC1, pr.data()), rl, rr;
        if (!rect(imL, imR, rl, rr)) return 4;
        FILE *out = fopen(argv[6], "wb");
        fwrite(rl.data, 1, (size_t)rl.rows * rl.cols, out);
        fwrite(rr.data, 1, (size_t)rr.rows * rr.cols, out);
        fclose(out);
        printf("rectified %d %d\n", rl.cols, rl.rows);
        return 0;
    }
    if (argc < 4) return 2;
    const std::string dir = argv[1];
    const int ppb = atoi(argv[2]);
    const bool extract = argc >= 6 && std::string(argv[4]) == "--extract";
    std::vector<std::string> L, R, tum;
    std::vector<double> ts, tts;
    orbb200::LoadImagesKITTI(dir, L, R, ts);
    printf("pairs %zu\n", L.size());
    orbb200::LoadImagesTUM(dir + "/rgb.txt", tum, tts);
    for (size_t i = 0; i < tum.size(); i++) printf("tum %.6f %s\n", tts[i], tum[i].c_str());
    if (!tum.empty()) { // one colour file through the decoder + gray conversion
        std::vector<unsigned char> pix;
        int w, h, ch;
        std::string err;
        if (!orbb200::ReadImage(dir + "/" + tum[0], pix, w, h, ch, &err)) { printf("error %s\n", err.c_str()); return 3; }
        std::vector<unsigned char> g((size_t)w * h);
        orbb200::ToGray(pix.data(), w, h, ch, true, g.data());
        FILE *f = fopen((std::string(argv[3]) + ".tum0").c_str(), "wb");
        fwrite(g.data(), 1, g.size(), f);
        fclose(f);
        printf("tum0 %d %d %d\n", w, h, ch);
    }
    orbb200::StereoBatchLoader loader(L, R, ts, ppb, 3, 4);
    if (!loader.error().empty()) { printf("error %s\n", loader.error().c_str()); return 3; }
    printf("shape %d %d batches %d\n", loader.width(), loader.height(), loader.batches());
    FILE *out = fopen(argv[3], "wb");
    orbb200_extractor *ex = nullptr;
    std::vector<orbb200_keypoint> kps;
    std::vector<uint8_t> desc;
    std::vector<float> uR, dep;
    int cap = 0;
    if (extract) {
        if (orbb200_extractor_create(&ex, atoi(argv[5]), 1.2f, 8, 20, 7, -1) != ORBB200_OK) { printf("error %s\n", orbb200_last_error()); return 4; }
        cap = orbb200_max_keypoints_for_shape(ex, loader.width(), loader.height());
    }
    orbb200::StereoBatch b;
    while (loader.Next(b)) {
        if (!b.error.empty()) { printf("error %s\n", b.error.c_str()); return 3; }
        fwrite(b.pixels, 1, (size_t)2 * b.npairs * b.width * b.height, out);
        printf("batch first %d npairs %d t0 %.3f\n", b.first, b.npairs, b.timestamps.empty() ? -1.0 : b.timestamps[0]);
        if (extract) {
            const int n2 = 2 * b.npairs;
            kps.resize((size_t)n2 * cap); desc.resize((size_t)n2 * cap * 32); uR.resize((size_t)b.npairs * cap); dep.resize((size_t)b.npairs * cap);
            std::vector<int> n(n2), nv(b.npairs);
            if (orbb200_extract_stereo_batch(ex, b.images.data(), b.npairs, b.width, b.height, b.width, kps.data(), desc.data(), cap, n.data(), 386.1448f,
                                             386.1448f / 718.856f, uR.data(), dep.data(), nv.data()) != ORBB200_OK) { printf("error %s\n", orbb200_last_error()); return 4; }
            for (int i = 0; i < n2; i++) printf("n %d %d\n", b.first * 2 + i, n[i]);
            for (int p = 0; p < b.npairs; p++) printf("nv %d %d\n", b.first + p, nv[p]);
        }
        loader.Release(b);
    }
    fclose(out);
    if (ex) orbb200_extractor_destroy(ex);
    return 0;
}

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

int main(int argc, char **argv)
{
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

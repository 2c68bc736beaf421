// facade_vocab.cpp -- test driver of include/orbb200/ORBVocabulary.h, written the way Frame::ComputeBoW
// (src/Frame.cc:459-466) and LoopClosing::DetectLoop (src/LoopClosing.cc:157) use ORBVocabulary.  Prints the maps
// (doubles as hex floats, exact) for tests/test_gpu_facade.py to compare with the reference's golden outputs.
//   facade_vocab.bin <vocabulary.txt> <descriptors.bin> <n_features> <levelsup> <split>
// Frame A = features [0, split), frame B = [split, n): A through the vector<cv::Mat> overload (the reference's call),
// B through the matrix overload, both again through transformBatch, then score(A, B) and score(A, A).
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "orbb200/ORBVocabulary.h"

static std::vector<cv::Mat> toDescriptorVector(const cv::Mat &D) // Converter::toDescriptorVector (src/Converter.cc:27-35)
{
    std::vector<cv::Mat> v;
    v.reserve(D.rows);
    for (int j = 0; j < D.rows; j++) v.push_back(D.row(j));
    return v;
}

static void dump(const char *tag, const DBoW2::BowVector &b, const DBoW2::FeatureVector &f)
{
    for (DBoW2::BowVector::const_iterator it = b.begin(); it != b.end(); ++it) printf("%s BOW %u %a\n", tag, it->first, it->second);
    for (DBoW2::FeatureVector::const_iterator it = f.begin(); it != f.end(); ++it) {
        printf("%s FV %u", tag, it->first);
        for (size_t i = 0; i < it->second.size(); i++) printf(" %u", it->second[i]);
        printf("\n");
    }
}

static DBoW2::BowVector read_bow(const char *path)
{
    DBoW2::BowVector b;
    FILE *f = fopen(path, "r");
    unsigned id;
    double v;
    while (f && fscanf(f, "%u %la", &id, &v) == 2) b[id] = v;
    if (f) fclose(f);
    return b;
}

int main(int argc, char **argv)
{
    if (argc == 5 && std::string(argv[1]) == "--score") { // host-only: facade_vocab.bin --score <scoring> <bowA.txt> <bowB.txt>
        const DBoW2::BowVector a = read_bow(argv[3]), b = read_bow(argv[4]);
        printf("SCORE %a %a\n", orbb200::ORBVocabulary::score(atoi(argv[2]), a, b), orbb200::ORBVocabulary::score(atoi(argv[2]), b, a));
        return 0;
    }
    if (argc < 6) return 2;
    const int n = atoi(argv[3]), levelsup = atoi(argv[4]), split = atoi(argv[5]);
    std::vector<unsigned char> raw((size_t)n * 32);
    FILE *f = fopen(argv[2], "rb");
    if (!f || fread(&raw[0], 32, n, f) != (size_t)n) return 3;
    fclose(f);
    orbb200::ORBVocabulary voc;
    if (voc.loadFromTextFile("/nonexistent/ORBvoc.txt")) return 4; // System.cc:71: a wrong path is reported, not fatal here
    if (!voc.loadFromTextFile(argv[1])) { fprintf(stderr, "Wrong path to vocabulary.\n"); return 5; }
    printf("INFO %d %d %u %d\n", voc.getBranchingFactor(), voc.getDepthLevels(), voc.size(), (int)voc.empty());
    cv::Mat A(split, 32, CV_8U, &raw[0]), B(n - split, 32, CV_8U, &raw[(size_t)split * 32]);
    DBoW2::BowVector bowA, bowB;
    DBoW2::FeatureVector fvA, fvB;
    voc.transform(toDescriptorVector(A), bowA, fvA, levelsup); // mpORBvocabulary->transform(vCurrentDesc,mBowVec,mFeatVec,4);
    voc.transform(B, bowB, fvB, levelsup);
    dump("A", bowA, fvA);
    dump("B", bowB, fvB);
    std::vector<cv::Mat> both;
    both.push_back(A);
    both.push_back(B);
    std::vector<DBoW2::BowVector> bv;
    std::vector<DBoW2::FeatureVector> fv;
    voc.transformBatch(both, bv, fv, levelsup);
    printf("BATCH %d\n", (int)(bv.size() == 2 && bv[0] == bowA && bv[1] == bowB && fv[0] == fvA && fv[1] == fvB));
    printf("SCORE %a %a\n", voc.score(bowA, bowB), voc.score(bowA, bowA));
    // an empty frame clears the outputs (v.clear(); fv.clear();)
    voc.transform(std::vector<cv::Mat>(), bowA, fvA, levelsup);
    printf("EMPTY %d\n", (int)(bowA.empty() && fvA.empty()));
    return 0;
}

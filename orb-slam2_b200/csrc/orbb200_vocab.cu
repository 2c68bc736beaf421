// orbb200_vocab.cu -- C-ABI of the DBoW2 vocabulary transform (include/orbb200.h, "DBoW2 vocabulary transform").
// Loader + tree flattening + launch + the BowVector / FeatureVector epilogue.  B200 only, no CPU fallback: the walks
// run in k_vocab_transform (vocab_kernel.cuh); the host code below only parses, lays the tree out and folds the
// per-feature (word, weight, node) triples into the map contents in the reference's arithmetic order.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <string>
#include <mutex>
#include <vector>

#include "vocab_kernel.cuh"

struct orbb200_vocabulary {
    int device = 0;
    int k = 0, L = 0, scoring = 0, weighting = 0, nnodes = 0, nwords = 0, maxChildren = 0;
    cudaStream_t stream = nullptr;
    // tree in HBM (breadth-first order, children contiguous)
    uint4 *d_desc = nullptr;
    int2 *d_info = nullptr;
    int *d_orig = nullptr;
    double *d_weight = nullptr;
    // host-API staging (grown on demand): device descriptors / outputs, pinned host mirrors
    uint8_t *d_in = nullptr;
    int *d_word = nullptr, *d_nid = nullptr;
    double *d_w = nullptr;
    int *h_word = nullptr, *h_nid = nullptr;
    double *h_w = nullptr;
    int cap = 0;
    // orbb200_compute_bow_batch: per-image offsets and the folded maps (k_bow_fold), device + pinned mirrors
    int *d_imgOff = nullptr, *d_bowId = nullptr, *d_nBow = nullptr, *d_nKept = nullptr;
    double *d_bowVal = nullptr;
    unsigned long long *d_fvKey = nullptr;
    int *h_imgOff = nullptr, *h_bowId = nullptr, *h_nBow = nullptr, *h_nKept = nullptr;
    double *h_bowVal = nullptr;
    unsigned long long *h_fvKey = nullptr;
    int foldCap = 0, foldImgs = 0;
};

namespace orbb200 {

constexpr int VOCAB_WALK_MIN_FEATURES = 16384; // measured crossover on B200 (tools/bow_probe.py): 2000 features 15 vs 21 us, 640k 0.245 vs 0.142 ms

static int vocab_ensure_staging(orbb200_vocabulary *v, int n)
{
    if (n <= v->cap) return ORBB200_OK;
    int cap = std::max(n, 4096);
    cudaFree(v->d_in); cudaFree(v->d_word); cudaFree(v->d_nid); cudaFree(v->d_w);
    cudaFreeHost(v->h_word); cudaFreeHost(v->h_nid); cudaFreeHost(v->h_w);
    v->d_in = nullptr; v->d_word = v->d_nid = nullptr; v->d_w = nullptr;
    v->h_word = v->h_nid = nullptr; v->h_w = nullptr;
    v->cap = 0;
    ORB_CUDA_OK(cudaMalloc(&v->d_in, (size_t)cap * 32));
    ORB_CUDA_OK(cudaMalloc(&v->d_word, (size_t)cap * sizeof(int)));
    ORB_CUDA_OK(cudaMalloc(&v->d_nid, (size_t)cap * sizeof(int)));
    ORB_CUDA_OK(cudaMalloc(&v->d_w, (size_t)cap * sizeof(double)));
    ORB_CUDA_OK(cudaMallocHost(&v->h_word, (size_t)cap * sizeof(int)));
    ORB_CUDA_OK(cudaMallocHost(&v->h_nid, (size_t)cap * sizeof(int)));
    ORB_CUDA_OK(cudaMallocHost(&v->h_w, (size_t)cap * sizeof(double)));
    v->cap = cap;
    return ORBB200_OK;
}

static int vocab_ensure_fold(orbb200_vocabulary *v, int total, int nimg)
{
    if (total > v->foldCap) {
        const int cap = std::max(total, 4096);
        cudaFree(v->d_bowId); cudaFree(v->d_bowVal); cudaFree(v->d_fvKey);
        cudaFreeHost(v->h_bowId); cudaFreeHost(v->h_bowVal); cudaFreeHost(v->h_fvKey);
        v->d_bowId = nullptr; v->d_bowVal = nullptr; v->d_fvKey = nullptr;
        v->h_bowId = nullptr; v->h_bowVal = nullptr; v->h_fvKey = nullptr;
        v->foldCap = 0;
        ORB_CUDA_OK(cudaMalloc(&v->d_bowId, (size_t)cap * sizeof(int)));
        ORB_CUDA_OK(cudaMalloc(&v->d_bowVal, (size_t)cap * sizeof(double)));
        ORB_CUDA_OK(cudaMalloc(&v->d_fvKey, (size_t)cap * sizeof(unsigned long long)));
        ORB_CUDA_OK(cudaMallocHost(&v->h_bowId, (size_t)cap * sizeof(int)));
        ORB_CUDA_OK(cudaMallocHost(&v->h_bowVal, (size_t)cap * sizeof(double)));
        ORB_CUDA_OK(cudaMallocHost(&v->h_fvKey, (size_t)cap * sizeof(unsigned long long)));
        v->foldCap = cap;
    }
    if (nimg > v->foldImgs) {
        const int cap = std::max(nimg, 64);
        cudaFree(v->d_imgOff); cudaFree(v->d_nBow); cudaFree(v->d_nKept);
        cudaFreeHost(v->h_imgOff); cudaFreeHost(v->h_nBow); cudaFreeHost(v->h_nKept);
        v->d_imgOff = v->d_nBow = v->d_nKept = nullptr;
        v->h_imgOff = v->h_nBow = v->h_nKept = nullptr;
        v->foldImgs = 0;
        ORB_CUDA_OK(cudaMalloc(&v->d_imgOff, (size_t)(cap + 1) * sizeof(int)));
        ORB_CUDA_OK(cudaMalloc(&v->d_nBow, (size_t)cap * sizeof(int)));
        ORB_CUDA_OK(cudaMalloc(&v->d_nKept, (size_t)cap * sizeof(int)));
        ORB_CUDA_OK(cudaMallocHost(&v->h_imgOff, (size_t)(cap + 1) * sizeof(int)));
        ORB_CUDA_OK(cudaMallocHost(&v->h_nBow, (size_t)cap * sizeof(int)));
        ORB_CUDA_OK(cudaMallocHost(&v->h_nKept, (size_t)cap * sizeof(int)));
        v->foldImgs = cap;
    }
    return ORBB200_OK;
}

static int vocab_launch(orbb200_vocabulary *v, const uint8_t *desc_d, int n, int levelsup, int *word_d, double *w_d, int *nid_d,
                        cudaStream_t st, const int *counts_d = nullptr, int stride = 0)
{
    if (n == 0) return ORBB200_OK;
    const int nidLevel = v->L - levelsup; // const int nid_level = m_L - levelsup; (:1226)
    // Same walk, two mappings (identical results): a batch keeps every lane busy with its own feature (k_vocab_walk,
    // L1-request bound); a single frame is latency bound, so a lane group per feature does each level in one load round.
    if (n >= VOCAB_WALK_MIN_FEATURES) {
        k_vocab_walk<<<div_up(n, VW_THREADS), VW_THREADS, 0, st>>>(v->d_desc, v->d_info, v->d_orig, v->d_weight, (const uint4 *)desc_d, n,
                                                                   nidLevel, word_d, w_d, nid_d, counts_d, stride);
    } else if (v->maxChildren <= 16) {
        k_vocab_transform<16><<<div_up(n, VT_THREADS / 16), VT_THREADS, 0, st>>>(v->d_desc, v->d_info, v->d_orig, v->d_weight,
                                                                                 (const uint4 *)desc_d, n, nidLevel, word_d, w_d, nid_d,
                                                                                 counts_d, stride);
    } else {
        k_vocab_transform<32><<<div_up(n, VT_THREADS / 32), VT_THREADS, 0, st>>>(v->d_desc, v->d_info, v->d_orig, v->d_weight,
                                                                                 (const uint4 *)desc_d, n, nidLevel, word_d, w_d, nid_d,
                                                                                 counts_d, stride);
    }
    ORB_CUDA_OK(cudaGetLastError());
    return ORBB200_OK;
}

// host descriptors -> per-feature triples in the pinned mirrors (h_word / h_w / h_nid)
static int vocab_transform_host(orbb200_vocabulary *v, const uint8_t *desc, int n, int levelsup)
{
    ORB_CUDA_OK(cudaSetDevice(v->device));
    int rc = vocab_ensure_staging(v, n);
    if (rc) return rc;
    if (n == 0) return ORBB200_OK;
    ORB_CUDA_OK(cudaMemcpyAsync(v->d_in, desc, (size_t)n * 32, cudaMemcpyHostToDevice, v->stream));
    rc = vocab_launch(v, v->d_in, n, levelsup, v->d_word, v->d_w, v->d_nid, v->stream);
    if (rc) return rc;
    ORB_CUDA_OK(cudaMemcpyAsync(v->h_word, v->d_word, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, v->stream));
    ORB_CUDA_OK(cudaMemcpyAsync(v->h_w, v->d_w, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, v->stream));
    ORB_CUDA_OK(cudaMemcpyAsync(v->h_nid, v->d_nid, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, v->stream));
    ORB_CUDA_OK(cudaStreamSynchronize(v->stream));
    return ORBB200_OK;
}

// ScoringObject.h:74-89: (mustNormalize, norm) of the six scoring classes; norm 1 = L1, 2 = L2
static bool must_normalize(int scoring, int *norm)
{
    *norm = scoring == 1 ? 2 : 1;
    return scoring != 5;
}

} // namespace orbb200

using namespace orbb200;

// MaxDynamicSharedMemorySize belongs to the function (per device), not to a vocabulary handle: opt k_bow_fold in once per device to
// the 200 KB that the 8192-feature limit of the fold kernel needs, so that handles with different sizes cannot lower it under
// one another
static int fold_opt_in(int device)
{
    static std::mutex mu;
    static bool done[64] = {};
    std::lock_guard<std::mutex> lk(mu);
    if (device < 0 || device >= 64 || done[device]) return ORBB200_OK;
    ORB_CUDA_OK(cudaFuncSetAttribute(k_bow_fold, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    done[device] = true;
    return ORBB200_OK;
}

extern "C" {

int orbb200_vocabulary_create(orbb200_vocabulary **out, int k, int L, int scoring, int weighting, int nnodes, const int *parent,
                              const uint8_t *is_leaf, const uint8_t *desc, const double *weight, int device)
{
    ORB_ARG_OK(out && nnodes >= 1, "null output or no root node");
    ORB_ARG_OK(nnodes == 1 || (parent && is_leaf && desc && weight), "null tree arrays");
    // if(m_k<0 || m_k>20 || m_L<1 || m_L>10 || n1<0 || n1>5 || n2<0 || n2>3) (:1385)
    ORB_ARG_OK(!(k < 0 || k > 20 || L < 1 || L > 10 || scoring < 0 || scoring > 5 || weighting < 0 || weighting > 3),
               "Vocabulary loading failure: not a correct k / L / scoring / weighting header");
    // children lists in id order (CSR)
    std::vector<int> ccount(nnodes, 0), cstart(nnodes + 1, 0);
    for (int n = 1; n < nnodes; n++) {
        ORB_ARG_OK(parent[n] >= 0 && parent[n] < n, "parent id must precede the node");
        ccount[parent[n]]++;
    }
    for (int n = 0; n < nnodes; n++) cstart[n + 1] = cstart[n] + ccount[n];
    std::vector<int> clist(std::max(nnodes - 1, 0)), fill(cstart.begin(), cstart.end() - 1);
    for (int n = 1; n < nnodes; n++) clist[fill[parent[n]]++] = n;
    int maxChildren = 0, nwords = 0;
    std::vector<int> wordOf(nnodes, -1);
    for (int n = 1; n < nnodes; n++) {
        ORB_ARG_OK((is_leaf[n] != 0) == (ccount[n] == 0), "a word with children / an inner node without children");
        if (is_leaf[n]) wordOf[n] = nwords++; // :1430-1437
    }
    for (int n = 0; n < nnodes; n++) maxChildren = std::max(maxChildren, ccount[n]);
    ORB_ARG_OK(maxChildren <= 255, "more than 255 children under one node");
    // breadth-first renumbering: children of a node become contiguous, in the reference's child order
    std::vector<int> order; // internal -> original
    order.reserve(nnodes);
    order.push_back(0);
    for (size_t q = 0; q < order.size(); q++) {
        const int o = order[q];
        for (int c = cstart[o]; c < cstart[o + 1]; c++) order.push_back(clist[c]);
    }
    ORB_ARG_OK((int)order.size() == nnodes, "tree is not connected to the root");
    std::vector<uint8_t> hdesc((size_t)nnodes * 32, 0);
    std::vector<int2> hinfo(nnodes);
    std::vector<double> hweight(nnodes, 0.0);
    int next = 1;
    for (int i = 0; i < nnodes; i++) {
        const int o = order[i];
        if (o > 0) {
            memcpy(&hdesc[(size_t)i * 32], desc + (size_t)o * 32, 32);
            hweight[i] = weight[o];
        }
        if (ccount[o] > 0) {
            hinfo[i] = make_int2(next, ccount[o]);
            next += ccount[o];
        } else {
            hinfo[i] = make_int2(o > 0 ? wordOf[o] : -1, 0);
        }
    }
    orbb200_vocabulary *v = new orbb200_vocabulary();
    v->k = k; v->L = L; v->scoring = scoring; v->weighting = weighting;
    v->nnodes = nnodes; v->nwords = nwords; v->maxChildren = maxChildren;
    auto fail = [&](int rc) { orbb200_vocabulary_destroy(v); return rc; };
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        set_error("orbb200_vocabulary_create: no CUDA device (this library has no CPU path)");
        return fail(ORBB200_ERR_CUDA);
    }
    if (device < 0) cudaGetDevice(&device);
    v->device = device;
#define VOC_OK(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) { set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e_)); return fail(ORBB200_ERR_CUDA); } } while (0)
    VOC_OK(cudaSetDevice(device));
    VOC_OK(cudaStreamCreateWithFlags(&v->stream, cudaStreamNonBlocking));
    VOC_OK(cudaMalloc(&v->d_desc, (size_t)nnodes * 32));
    VOC_OK(cudaMalloc(&v->d_info, (size_t)nnodes * sizeof(int2)));
    VOC_OK(cudaMalloc(&v->d_orig, (size_t)nnodes * sizeof(int)));
    VOC_OK(cudaMalloc(&v->d_weight, (size_t)nnodes * sizeof(double)));
    VOC_OK(cudaMemcpy(v->d_desc, hdesc.data(), (size_t)nnodes * 32, cudaMemcpyHostToDevice));
    VOC_OK(cudaMemcpy(v->d_info, hinfo.data(), (size_t)nnodes * sizeof(int2), cudaMemcpyHostToDevice));
    VOC_OK(cudaMemcpy(v->d_orig, order.data(), (size_t)nnodes * sizeof(int), cudaMemcpyHostToDevice));
    VOC_OK(cudaMemcpy(v->d_weight, hweight.data(), (size_t)nnodes * sizeof(double), cudaMemcpyHostToDevice));
#undef VOC_OK
    *out = v;
    return ORBB200_OK;
}

int orbb200_vocabulary_load_text(orbb200_vocabulary **out, const char *path, int device)
{
    ORB_ARG_OK(out && path, "null argument");
    FILE *f = fopen(path, "rb");
    if (!f) { set_error("orbb200_vocabulary_load_text: cannot open %s", path); return ORBB200_ERR_ARG; }
    fseek(f, 0, SEEK_END);
    long sz = ftell(f);
    fseek(f, 0, SEEK_SET);
    std::string buf((size_t)std::max(sz, 0L), '\0');
    size_t got = sz > 0 ? fread(&buf[0], 1, (size_t)sz, f) : 0;
    fclose(f);
    buf.resize(got);
    const char *p = buf.c_str(), *end = p + buf.size();
    auto next_line = [&](const char *&b, const char *&e) { // [b, e) without the newline; false at end of file
        if (p >= end) return false;
        b = p;
        const char *nl = (const char *)memchr(p, '\n', (size_t)(end - p));
        e = nl ? nl : end;
        p = nl ? nl + 1 : end;
        return true;
    };
    const char *b, *e;
    int k = -1, L = -1, n1 = -1, n2 = -1;
    if (!next_line(b, e) || sscanf(std::string(b, e).c_str(), "%d %d %d %d", &k, &L, &n1, &n2) != 4) {
        set_error("Vocabulary loading failure: This is not a correct text file!");
        return ORBB200_ERR_ARG;
    }
    std::vector<int> parent(1, 0);
    std::vector<uint8_t> leaf(1, 0), desc(32, 0);
    std::vector<double> weight(1, 0.0);
    while (next_line(b, e)) {
        char *q = const_cast<char *>(b), *r;
        while (q < e && (*q == ' ' || *q == '\t' || *q == '\r')) q++;
        if (q >= e) continue; // blank line (see include/orbb200.h)
        bool ok = true;
        const long pid = strtol(q, &r, 10); ok &= r != q; q = r;
        const long isl = strtol(q, &r, 10); ok &= r != q; q = r;
        const size_t o = desc.size();
        desc.resize(o + 32);
        for (int i = 0; i < 32; i++) { desc[o + i] = (uint8_t)strtol(q, &r, 10); ok &= r != q; q = r; } // *p = (unsigned char)n (FORB.cpp:132)
        const double w = strtod(q, &r);
        while (r < e && (*r == ' ' || *r == '\t' || *r == '\r')) r++;
        // (the reference parses a short line into whatever the streams leave behind; here it is an error)
        if (!ok || r == q || r != e) { set_error("orbb200_vocabulary_load_text: malformed node line %zu", parent.size()); return ORBB200_ERR_ARG; }
        parent.push_back((int)pid);
        leaf.push_back(isl > 0 ? 1 : 0); // if(nIsLeaf>0) (:1430)
        weight.push_back(w);
    }
    return orbb200_vocabulary_create(out, k, L, n1, n2, (int)parent.size(), parent.data(), leaf.data(), desc.data(), weight.data(), device);
}

void orbb200_vocabulary_destroy(orbb200_vocabulary *v)
{
    if (!v) return;
    cudaSetDevice(v->device);
    cudaFree(v->d_desc); cudaFree(v->d_info); cudaFree(v->d_orig); cudaFree(v->d_weight);
    cudaFree(v->d_in); cudaFree(v->d_word); cudaFree(v->d_nid); cudaFree(v->d_w);
    cudaFreeHost(v->h_word); cudaFreeHost(v->h_nid); cudaFreeHost(v->h_w);
    cudaFree(v->d_imgOff); cudaFree(v->d_bowId); cudaFree(v->d_nBow); cudaFree(v->d_nKept); cudaFree(v->d_bowVal); cudaFree(v->d_fvKey);
    cudaFreeHost(v->h_imgOff); cudaFreeHost(v->h_bowId); cudaFreeHost(v->h_nBow); cudaFreeHost(v->h_nKept); cudaFreeHost(v->h_bowVal);
    cudaFreeHost(v->h_fvKey);
    if (v->stream) cudaStreamDestroy(v->stream);
    delete v;
}

int orbb200_vocabulary_info(const orbb200_vocabulary *v, int *k, int *L, int *scoring, int *weighting, int *nnodes, int *nwords)
{
    ORB_ARG_OK(v, "null vocabulary");
    if (k) *k = v->k;
    if (L) *L = v->L;
    if (scoring) *scoring = v->scoring;
    if (weighting) *weighting = v->weighting;
    if (nnodes) *nnodes = v->nnodes;
    if (nwords) *nwords = v->nwords;
    return ORBB200_OK;
}

int orbb200_vocabulary_transform(orbb200_vocabulary *v, const uint8_t *desc, int n, int levelsup, int *word_id, double *weight,
                                 int *node_id)
{
    ORB_ARG_OK(v && n >= 0 && (n == 0 || (desc && word_id && weight && node_id)), "null argument");
    if (v->nwords == 0) { set_error("orbb200_vocabulary_transform: empty vocabulary"); return ORBB200_ERR_STATE; }
    int rc = vocab_transform_host(v, desc, n, levelsup);
    if (rc) return rc;
    memcpy(word_id, v->h_word, (size_t)n * sizeof(int));
    memcpy(weight, v->h_w, (size_t)n * sizeof(double));
    memcpy(node_id, v->h_nid, (size_t)n * sizeof(int));
    return ORBB200_OK;
}

int orbb200_vocabulary_transform_device(orbb200_vocabulary *v, const uint8_t *desc_d, int n, int levelsup, int *word_id_d,
                                        double *weight_d, int *node_id_d, void *stream)
{
    ORB_ARG_OK(v && n >= 0 && (n == 0 || (desc_d && word_id_d && weight_d && node_id_d)), "null argument");
    ORB_ARG_OK(((uintptr_t)desc_d & 15) == 0, "desc_d must be 16-byte aligned");
    if (v->nwords == 0) { set_error("orbb200_vocabulary_transform_device: empty vocabulary"); return ORBB200_ERR_STATE; }
    ORB_CUDA_OK(cudaSetDevice(v->device));
    return vocab_launch(v, desc_d, n, levelsup, word_id_d, weight_d, node_id_d, stream ? (cudaStream_t)stream : v->stream);
}

int orbb200_compute_bow_batch(orbb200_vocabulary *v, const uint8_t *desc, const int *counts, int nimg, int levelsup, int *bow_off,
                              int *bow_id, double *bow_val, int *fv_off, int *fv_node, int *fv_feat_off, int *fv_feat)
{
    ORB_ARG_OK(v && nimg >= 0 && counts && bow_off && fv_off && fv_feat_off, "null argument");
    long long total = 0;
    for (int i = 0; i < nimg; i++) {
        ORB_ARG_OK(counts[i] >= 0, "negative descriptor count");
        total += counts[i];
    }
    ORB_ARG_OK(total < (1ll << 31), "too many descriptors for one call");
    ORB_ARG_OK(total == 0 || (desc && bow_id && bow_val && fv_node && fv_feat), "null argument");
    bow_off[0] = fv_off[0] = fv_feat_off[0] = 0;
    if (v->nwords == 0) { // if(empty()) return; (:1133-1136): both maps stay empty
        for (int i = 0; i < nimg; i++) bow_off[i + 1] = fv_off[i + 1] = 0;
        return ORBB200_OK;
    }
    int maxCount = 0;
    for (int i = 0; i < nimg; i++) maxCount = std::max(maxCount, counts[i]);
    int npad = 2;
    while (npad < maxCount) npad <<= 1;
    const size_t smem = (size_t)npad * 16;
    if (smem > 200 * 1024) {
        set_error("orbb200_compute_bow_batch: %d descriptors in one image exceed the fold kernel's shared memory (max 8192)", maxCount);
        return ORBB200_ERR_CAPACITY;
    }
    ORB_CUDA_OK(cudaSetDevice(v->device));
    const int n = (int)total;
    int rc = vocab_ensure_staging(v, n);
    if (!rc) rc = vocab_ensure_fold(v, n, nimg);
    if (rc) return rc;
    if (nimg == 0) return ORBB200_OK;
    { const int rc2 = fold_opt_in(v->device); if (rc2) return rc2; }
    v->h_imgOff[0] = 0;
    for (int i = 0; i < nimg; i++) v->h_imgOff[i + 1] = v->h_imgOff[i] + counts[i];
    cudaStream_t st = v->stream;
    ORB_CUDA_OK(cudaMemcpyAsync(v->d_imgOff, v->h_imgOff, (size_t)(nimg + 1) * sizeof(int), cudaMemcpyHostToDevice, st));
    if (n > 0) ORB_CUDA_OK(cudaMemcpyAsync(v->d_in, desc, (size_t)n * 32, cudaMemcpyHostToDevice, st));
    rc = vocab_launch(v, v->d_in, n, levelsup, v->d_word, v->d_w, v->d_nid, st);
    if (rc) return rc;
    BowFoldParams P;
    P.word = v->d_word; P.nid = v->d_nid; P.weight = v->d_w; P.imgOff = v->d_imgOff;
    P.counts = nullptr; P.stride = 0;
    P.tf = (v->weighting == 0 || v->weighting == 1) ? 1 : 0; // TF_IDF || TF (:1144)
    int norm = 1;
    P.must = must_normalize(v->scoring, &norm) ? 1 : 0;
    P.norm = norm;
    P.bowId = v->d_bowId; P.bowVal = v->d_bowVal; P.fvKey = v->d_fvKey; P.nBow = v->d_nBow; P.nKept = v->d_nKept;
    k_bow_fold<<<nimg, BF_THREADS, smem, st>>>(P, npad);
    ORB_CUDA_OK(cudaGetLastError());
    ORB_CUDA_OK(cudaMemcpyAsync(v->h_nBow, v->d_nBow, (size_t)nimg * sizeof(int), cudaMemcpyDeviceToHost, st));
    ORB_CUDA_OK(cudaMemcpyAsync(v->h_nKept, v->d_nKept, (size_t)nimg * sizeof(int), cudaMemcpyDeviceToHost, st));
    if (n > 0) {
        ORB_CUDA_OK(cudaMemcpyAsync(v->h_bowId, v->d_bowId, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, st));
        ORB_CUDA_OK(cudaMemcpyAsync(v->h_bowVal, v->d_bowVal, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, st));
        ORB_CUDA_OK(cudaMemcpyAsync(v->h_fvKey, v->d_fvKey, (size_t)n * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
    }
    ORB_CUDA_OK(cudaStreamSynchronize(st));
    // flatten: the per-image blocks are already in map order; only the node-group boundaries remain to be marked
    int nb = 0, nf = 0, nff = 0;
    for (int im = 0; im < nimg; im++) {
        const int off = v->h_imgOff[im], kb = v->h_nBow[im], m = v->h_nKept[im];
        memcpy(bow_id + nb, v->h_bowId + off, (size_t)kb * sizeof(int));
        memcpy(bow_val + nb, v->h_bowVal + off, (size_t)kb * sizeof(double));
        nb += kb;
        bow_off[im + 1] = nb;
        const unsigned long long *key = v->h_fvKey + off;
        unsigned prev = 0;
        for (int q = 0; q < m; q++) {
            const unsigned node = (unsigned)(key[q] >> 32);
            if (q == 0 || node != prev) {
                fv_node[nf] = (int)node;
                fv_feat_off[nf++] = nff;
                prev = node;
            }
            fv_feat[nff++] = (int)(unsigned)key[q];
        }
        fv_feat_off[nf] = nff;
        fv_off[im + 1] = nf;
    }
    return ORBB200_OK;
}

int orbb200_compute_bow_batch_device(orbb200_vocabulary *v, const uint8_t *desc_d, const int *n_d, int nimg, int cap, int levelsup,
                                     int *bow_id_d, double *bow_val_d, int *n_bow_d, unsigned long long *fv_key_d, int *n_kept_d,
                                     void *stream)
{
    ORB_ARG_OK(v && nimg >= 0 && cap > 0, "null vocabulary / bad sizes");
    ORB_ARG_OK(nimg == 0 || (desc_d && n_d && bow_id_d && bow_val_d && n_bow_d && fv_key_d && n_kept_d), "null argument");
    ORB_ARG_OK(((uintptr_t)desc_d & 15) == 0, "desc_d must be 16-byte aligned");
    ORB_ARG_OK((long long)nimg * cap < (1ll << 31), "too many descriptor rows for one call");
    if (v->nwords == 0) { set_error("orbb200_compute_bow_batch_device: empty vocabulary"); return ORBB200_ERR_STATE; }
    if (nimg == 0) return ORBB200_OK;
    int npad = 2;
    while (npad < cap) npad <<= 1;
    const size_t smem = (size_t)npad * 16;
    if (smem > 200 * 1024) {
        set_error("orbb200_compute_bow_batch_device: cap %d exceeds the fold kernel's shared memory (max 8192)", cap);
        return ORBB200_ERR_CAPACITY;
    }
    ORB_CUDA_OK(cudaSetDevice(v->device));
    const int n = nimg * cap;
    int rc = vocab_ensure_staging(v, n); // per-feature (word, weight, node) scratch
    if (rc) return rc;
    { const int rc2 = fold_opt_in(v->device); if (rc2) return rc2; }
    cudaStream_t st = stream ? (cudaStream_t)stream : v->stream;
    rc = vocab_launch(v, desc_d, n, levelsup, v->d_word, v->d_w, v->d_nid, st, n_d, cap);
    if (rc) return rc;
    BowFoldParams P;
    P.word = v->d_word; P.nid = v->d_nid; P.weight = v->d_w; P.imgOff = nullptr;
    P.counts = n_d; P.stride = cap;
    P.tf = (v->weighting == 0 || v->weighting == 1) ? 1 : 0;
    int norm = 1;
    P.must = must_normalize(v->scoring, &norm) ? 1 : 0;
    P.norm = norm;
    P.bowId = bow_id_d; P.bowVal = bow_val_d; P.fvKey = fv_key_d; P.nBow = n_bow_d; P.nKept = n_kept_d;
    k_bow_fold<<<nimg, BF_THREADS, smem, st>>>(P, npad);
    ORB_CUDA_OK(cudaGetLastError());
    return ORBB200_OK;
}

} // extern "C"

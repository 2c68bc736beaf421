/*
 * DatasetLoader.h -- the image side of the reference's example drivers, batched.
 *
 *   LoadImagesKITTI  == LoadImages of Examples/Stereo/stereo_kitti.cc:130-160 (times.txt, image_0/%06d.png, image_1/%06d.png)
 *   LoadImagesTUM    == LoadImages of Examples/Monocular/mono_tum.cc:128-160 (rgb.txt: three header lines, "timestamp filename")
 *   StereoBatchLoader: where the reference calls cv::imread once per frame inside its main loop (stereo_kitti.cc:69-70), this
 *                      reader decodes the pairs of a whole batch on worker threads straight into PINNED host memory, laid out
 *                      [pair][L, R][h][w] back to back -- exactly what orbb200_extract_stereo_batch consumes with one H2D copy per
 *                      chunk -- and keeps `slots` batches in flight so that decoding runs under the GPU work of the previous batch.
 *
 * Decoders: PNG (8-bit gray / gray+alpha / RGB / RGBA, non-interlaced -- KITTI odometry and TUM RGB-D are such files) through zlib's
 * inflate + the five PNG row filters, binary PGM (P5) and PPM (P6).  No OpenCV needed.  Link with -lz.
 * Colour files are reduced to gray with cv::cvtColor's fixed-point weights (R 9798, G 19235, B 3735, +16384 >> 15), the arithmetic
 * of Tracking::GrabImageStereo's prologue (src/Tracking.cc:177-202) -- the same the device prologue k_stage_gray uses.
 */
#ifndef ORBB200_DATASETLOADER_H
#define ORBB200_DATASETLOADER_H

#include <zlib.h>

#include <algorithm>
#include <condition_variable>
#include <cstdlib>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iomanip>
#include <mutex>
#include <sstream>
#include <string>
#include <thread>
#include <vector>

#include "../orbb200.h"

namespace orbb200 {

inline void LoadImagesKITTI(const std::string &strPathToSequence, std::vector<std::string> &vstrImageLeft,
                            std::vector<std::string> &vstrImageRight, std::vector<double> &vTimestamps)
{
    std::ifstream fTimes((strPathToSequence + "/times.txt").c_str());
    std::string s;
    while (std::getline(fTimes, s)) {
        if (s.empty()) continue;
        std::stringstream ss(s);
        double t;
        ss >> t;
        vTimestamps.push_back(t);
    }
    const int nTimes = (int)vTimestamps.size();
    vstrImageLeft.resize(nTimes);
    vstrImageRight.resize(nTimes);
    for (int i = 0; i < nTimes; i++) {
        std::stringstream ss;
        ss << std::setfill('0') << std::setw(6) << i;
        vstrImageLeft[i] = strPathToSequence + "/image_0/" + ss.str() + ".png";
        vstrImageRight[i] = strPathToSequence + "/image_1/" + ss.str() + ".png";
    }
}

/* Examples/Stereo/stereo_euroc.cc:192-215: one timestamp (integer nanoseconds) per line of strPathTimes; the images are
 * <left dir>/<stamp>.png and <right dir>/<stamp>.png, the time is stamp / 1e9 */
inline void LoadImagesEuRoC(const std::string &strPathLeft, const std::string &strPathRight, const std::string &strPathTimes,
                            std::vector<std::string> &vstrImageLeft, std::vector<std::string> &vstrImageRight, std::vector<double> &vTimeStamps)
{
    std::ifstream fTimes(strPathTimes.c_str());
    std::string s;
    while (std::getline(fTimes, s)) {
        if (!s.empty() && s[s.size() - 1] == '\r') s.erase(s.size() - 1);
        if (s.empty()) continue;
        vstrImageLeft.push_back(strPathLeft + "/" + s + ".png");
        vstrImageRight.push_back(strPathRight + "/" + s + ".png");
        std::stringstream ss(s);
        double t;
        ss >> t;
        vTimeStamps.push_back(t / 1e9);
    }
}

inline void LoadImagesTUM(const std::string &strFile, std::vector<std::string> &vstrImageFilenames, std::vector<double> &vTimestamps)
{
    std::ifstream f(strFile.c_str());
    std::string s;
    std::getline(f, s);
    std::getline(f, s);
    std::getline(f, s); // three header lines (mono_tum.cc:133-137)
    while (std::getline(f, s)) {
        if (s.empty()) continue;
        std::stringstream ss(s);
        double t;
        std::string name;
        ss >> t >> name;
        vTimestamps.push_back(t);
        vstrImageFilenames.push_back(name);
    }
}

namespace detail {

inline bool read_file(const std::string &path, std::vector<unsigned char> &buf)
{
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) return false;
    fseek(f, 0, SEEK_END);
    const long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    buf.resize(n > 0 ? (size_t)n : 0);
    const size_t got = n > 0 ? fread(buf.data(), 1, (size_t)n, f) : 0;
    fclose(f);
    return got == buf.size() && n > 0;
}

inline uint32_t be32(const unsigned char *p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

/* PNG: signature, IHDR, concatenated IDAT -> inflate -> per-row filter (None / Sub / Up / Average / Paeth) */
inline bool decode_png(const std::vector<unsigned char> &file, std::vector<unsigned char> &pix, int &w, int &h, int &ch, std::string &err)
{
    static const unsigned char sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    if (file.size() < 33 || memcmp(file.data(), sig, 8) != 0) { err = "not a PNG file"; return false; }
    size_t pos = 8;
    std::vector<unsigned char> z;
    int depth = 0, ctype = 0, interlace = 0;
    w = h = 0;
    while (pos + 12 <= file.size()) {
        const uint32_t len = be32(&file[pos]);
        const unsigned char *type = &file[pos + 4];
        if (pos + 12 + (size_t)len > file.size()) { err = "truncated PNG chunk"; return false; }
        const unsigned char *data = &file[pos + 8];
        if (!memcmp(type, "IHDR", 4)) {
            w = (int)be32(data); h = (int)be32(data + 4); depth = data[8]; ctype = data[9]; interlace = data[12];
        } else if (!memcmp(type, "IDAT", 4)) {
            z.insert(z.end(), data, data + len);
        } else if (!memcmp(type, "IEND", 4)) {
            break;
        }
        pos += 12 + (size_t)len;
    }
    if (w <= 0 || h <= 0) { err = "PNG without IHDR"; return false; }
    if (depth != 8 || interlace != 0) { err = "only 8-bit non-interlaced PNG is supported"; return false; }
    ch = ctype == 0 ? 1 : ctype == 4 ? 2 : ctype == 2 ? 3 : ctype == 6 ? 4 : 0;
    if (!ch) { err = "palette PNG is not supported"; return false; }
    const size_t stride = (size_t)w * ch;
    std::vector<unsigned char> raw((stride + 1) * (size_t)h);
    uLongf outLen = (uLongf)raw.size();
    if (uncompress(raw.data(), &outLen, z.data(), (uLong)z.size()) != Z_OK || outLen != raw.size()) { err = "PNG inflate failed"; return false; }
    pix.resize(stride * (size_t)h);
    for (int y = 0; y < h; y++) {
        const unsigned char *in = &raw[(stride + 1) * (size_t)y];
        unsigned char *out = &pix[stride * (size_t)y];
        const unsigned char *up = y ? out - stride : nullptr;
        const int ft = in[0];
        in++;
        for (size_t x = 0; x < stride; x++) {
            const int a = x >= (size_t)ch ? out[x - ch] : 0, b = up ? up[x] : 0, c = (up && x >= (size_t)ch) ? up[x - ch] : 0;
            int pr;
            switch (ft) {
            case 0: pr = 0; break;
            case 1: pr = a; break;
            case 2: pr = b; break;
            case 3: pr = (a + b) >> 1; break;
            case 4: {
                const int p = a + b - c, pa = p > a ? p - a : a - p, pb = p > b ? p - b : b - p, pc = p > c ? p - c : c - p;
                pr = (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
                break;
            }
            default: err = "bad PNG filter type"; return false;
            }
            out[x] = (unsigned char)(in[x] + pr);
        }
    }
    return true;
}

/* binary PGM (P5) / PPM (P6), maxval 255 */
inline bool decode_pnm(const std::vector<unsigned char> &file, std::vector<unsigned char> &pix, int &w, int &h, int &ch, std::string &err)
{
    if (file.size() < 7 || file[0] != 'P' || (file[1] != '5' && file[1] != '6')) { err = "not a binary PGM / PPM file"; return false; }
    ch = file[1] == '5' ? 1 : 3;
    size_t pos = 2;
    int vals[3], nv = 0;
    while (nv < 3 && pos < file.size()) {
        while (pos < file.size() && (file[pos] == ' ' || file[pos] == '\n' || file[pos] == '\r' || file[pos] == '\t')) pos++;
        if (pos < file.size() && file[pos] == '#') { while (pos < file.size() && file[pos] != '\n') pos++; continue; }
        int v = 0, nd = 0;
        while (pos < file.size() && file[pos] >= '0' && file[pos] <= '9') { v = v * 10 + (file[pos] - '0'); pos++; nd++; }
        if (!nd) { err = "bad PNM header"; return false; }
        vals[nv++] = v;
    }
    pos++; // the single whitespace after maxval
    w = vals[0]; h = vals[1];
    if (nv < 3 || vals[2] != 255 || w <= 0 || h <= 0 || pos + (size_t)w * h * ch > file.size()) { err = "unsupported or truncated PNM"; return false; }
    pix.assign(file.begin() + pos, file.begin() + pos + (size_t)w * h * ch);
    return true;
}

} // namespace detail

/* decode one image file into 8-bit pixels (1..4 channels interleaved); false + message on failure */
inline bool ReadImage(const std::string &path, std::vector<unsigned char> &pix, int &w, int &h, int &ch, std::string *error = nullptr)
{
    std::vector<unsigned char> file;
    std::string err;
    bool ok = detail::read_file(path, file);
    if (!ok) err = "cannot read " + path;
    else if (file.size() > 1 && file[0] == 'P') ok = detail::decode_pnm(file, pix, w, h, ch, err);
    else ok = detail::decode_png(file, pix, w, h, ch, err);
    if (!ok && error) *error = path + ": " + err;
    return ok;
}

/* gray conversion with cv::cvtColor's 8-bit fixed-point weights; rFirst = the file's channel 0 is red (PNG / PPM are RGB) */
inline void ToGray(const unsigned char *pix, int w, int h, int ch, bool rFirst, unsigned char *dst)
{
    const size_t n = (size_t)w * h;
    if (ch == 1) { memcpy(dst, pix, n); return; }
    if (ch == 2) { for (size_t i = 0; i < n; i++) dst[i] = pix[2 * i]; return; }
    for (size_t i = 0; i < n; i++) {
        const unsigned char *p = pix + i * ch;
        const int r = rFirst ? p[0] : p[2], g = p[1], b = rFirst ? p[2] : p[0];
        dst[i] = (unsigned char)((r * 9798 + g * 19235 + b * 3735 + 16384) >> 15);
    }
}

/* One batch of stereo pairs in pinned host memory: image 2p = left, 2p + 1 = right of pair p, each h x w bytes, back to back. */
struct StereoBatch {
    int slot = -1, first = 0, npairs = 0, width = 0, height = 0;
    unsigned char *pixels = nullptr;             // [2 * npairs][height][width]
    std::vector<const uint8_t *> images;         // the 2 * npairs image pointers orbb200_extract_stereo_batch takes
    std::vector<double> timestamps;              // per pair
    std::string error;                           // non-empty: a file of this batch could not be decoded
};

class StereoBatchLoader {
public:
    /* left / right: file lists of equal length (LoadImagesKITTI); pairsPerBatch pairs per Next(); `slots` batches in flight */
    StereoBatchLoader(const std::vector<std::string> &left, const std::vector<std::string> &right, const std::vector<double> &timestamps,
                      int pairsPerBatch, int slots = 3, int threads = 4)
        : left_(left), right_(right), ts_(timestamps), ppb_(pairsPerBatch), nslots_(slots < 2 ? 2 : slots), w_(0), h_(0), nextIssue_(0), nextTake_(0),
          stop_(false)
    {
        total_ = (int)std::min(left_.size(), right_.size());
        // the first image fixes the shape of the sequence
        std::vector<unsigned char> pix;
        int ch = 0;
        if (total_ > 0 && !ReadImage(left_[0], pix, w_, h_, ch, &initError_)) total_ = 0;
        slotBytes_ = (size_t)2 * ppb_ * w_ * h_;
        slots_.resize(nslots_);
        for (int i = 0; i < nslots_; i++) {
            slots_[i].state = FREE;
            slots_[i].mem = nullptr;
            if (!slotBytes_) continue;
            if (orbb200_host_alloc((void **)&slots_[i].mem, slotBytes_) == ORBB200_OK) { slots_[i].pinned = true; continue; }
            slots_[i].mem = (unsigned char *)std::malloc(slotBytes_); // no CUDA device (CPU-only tools): pageable memory, same layout
            slots_[i].pinned = false;
            if (!slots_[i].mem) { initError_ = "out of memory"; total_ = 0; }
        }
        const int nbatches = ppb_ > 0 ? (total_ + ppb_ - 1) / ppb_ : 0;
        nbatches_ = nbatches;
        for (int t = 0; t < (threads < 1 ? 1 : threads); t++) workers_.emplace_back([this] { work(); });
    }
    ~StereoBatchLoader()
    {
        { std::lock_guard<std::mutex> lk(mu_); stop_ = true; }
        cv_.notify_all();
        for (std::thread &t : workers_) t.join();
        for (Slot &s : slots_) { if (s.mem && s.pinned) orbb200_host_free(s.mem); else std::free(s.mem); }
    }
    int width() const { return w_; }
    int height() const { return h_; }
    int batches() const { return nbatches_; }
    const std::string &error() const { return initError_; }

    /* blocks until the next batch (in sequence order) is decoded; false at the end of the sequence */
    bool Next(StereoBatch &b)
    {
        std::unique_lock<std::mutex> lk(mu_);
        if (nextTake_ >= nbatches_) return false;
        const int want = nextTake_;
        cv_.wait(lk, [&] { return slot_of(want) >= 0 && slots_[slot_of(want)].state == READY; });
        Slot &s = slots_[slot_of(want)];
        s.state = TAKEN;
        nextTake_++;
        b.slot = slot_of(want);
        b.first = want * ppb_;
        b.npairs = std::min(ppb_, total_ - b.first);
        b.width = w_; b.height = h_;
        b.pixels = s.mem;
        b.error = s.error;
        b.images.resize((size_t)2 * b.npairs);
        for (int i = 0; i < 2 * b.npairs; i++) b.images[i] = s.mem + (size_t)i * w_ * h_;
        b.timestamps.assign(ts_.begin() + std::min((size_t)b.first, ts_.size()), ts_.begin() + std::min((size_t)(b.first + b.npairs), ts_.size()));
        return true;
    }
    /* hand the batch's pinned slot back to the decoders (after the extraction call that read it has returned) */
    void Release(const StereoBatch &b)
    {
        { std::lock_guard<std::mutex> lk(mu_); if (b.slot >= 0) { slots_[b.slot].state = FREE; slots_[b.slot].batch = -1; } }
        cv_.notify_all();
    }

private:
    enum State { FREE, FILLING, READY, TAKEN };
    struct Slot { State state; int batch = -1, pending = 0; unsigned char *mem; bool pinned = false; std::string error; };
    int slot_of(int batch) const { for (int i = 0; i < nslots_; i++) if (slots_[i].batch == batch) return i; return -1; }
    void work()
    {
        std::vector<unsigned char> pix;
        for (;;) {
            int batch = -1, item = -1, slot = -1;
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_.wait(lk, [&] {
                    if (stop_) return true;
                    if (!queue_.empty()) return true;
                    if (nextIssue_ < nbatches_) for (int i = 0; i < nslots_; i++) if (slots_[i].state == FREE) return true;
                    return false;
                });
                if (stop_) return;
                if (queue_.empty()) { // open the next batch on a free slot: its 2 * npairs decode jobs go to the shared queue
                    for (int i = 0; i < nslots_ && slot < 0; i++) if (slots_[i].state == FREE) slot = i;
                    batch = nextIssue_++;
                    const int np = std::min(ppb_, total_ - batch * ppb_);
                    slots_[slot].state = FILLING; slots_[slot].batch = batch; slots_[slot].pending = 2 * np; slots_[slot].error.clear();
                    for (int i = 0; i < 2 * np; i++) queue_.push_back(std::make_pair(batch, i));
                    cv_.notify_all();
                }
                batch = queue_.back().first; item = queue_.back().second;
                queue_.pop_back();
                slot = slot_of(batch);
            }
            const int pair = batch * ppb_ + item / 2;
            const std::string &path = (item & 1) ? right_[pair] : left_[pair];
            int w = 0, h = 0, ch = 0;
            std::string err;
            bool ok = ReadImage(path, pix, w, h, ch, &err);
            if (ok && (w != w_ || h != h_)) { ok = false; err = path + ": shape differs from the first image of the sequence"; }
            if (ok) ToGray(pix.data(), w, h, ch, true, slots_[slot].mem + (size_t)item * w_ * h_);
            {
                std::lock_guard<std::mutex> lk(mu_);
                if (!ok && slots_[slot].error.empty()) slots_[slot].error = err;
                if (--slots_[slot].pending == 0) slots_[slot].state = READY;
            }
            cv_.notify_all();
        }
    }

    std::vector<std::string> left_, right_;
    std::vector<double> ts_;
    int ppb_, nslots_, w_, h_, total_ = 0, nbatches_ = 0, nextIssue_, nextTake_;
    size_t slotBytes_ = 0;
    bool stop_;
    std::string initError_;
    std::vector<Slot> slots_;
    std::vector<std::pair<int, int> > queue_;
    std::vector<std::thread> workers_;
    std::mutex mu_;
    std::condition_variable cv_;
};

} // namespace orbb200

#endif /* ORBB200_DATASETLOADER_H */

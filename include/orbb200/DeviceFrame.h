/*
 * DeviceFrame.h -- the extraction outputs of one Frame kept on the device.
 *
 * The reference's Frame constructor (src/Frame.cc:60-125) extracts, undistorts, matches stereo and bins the keypoints into the
 * 64 x 48 grid (:261-279); every later search (src/ORBmatcher.cc:48-136, 1396-1553, 1555-1682, 430-556) reads mvKeysUn /
 * mDescriptors of that frame again.  With the plain facade each such search uploads the frame's keypoints and descriptors
 * (60 bytes per keypoint) before its one device call.  A DeviceFrame keeps them in HBM instead:
 *
 *   Adopt(extractor, n)      after ORBextractor::operator(): device-to-device copy of the n keypoints / descriptors the extractor
 *                            handle still holds (no host round trip)
 *   Undistort(mK, mDistCoef) Frame::UndistortKeyPoints (:470-515) on the device copy (k1 == 0: mvKeysUn = mvKeys, :473-477)
 *   SetGrid(...)             the grid parameters AssignFeaturesToGrid / GetFeaturesInArea use (Frame statics mnMinX, mnMinY,
 *                            mfGridElementWidthInv, mfGridElementHeightInv)
 *   WindowSearch(...)        grid build + GetFeaturesInArea + all distances of a windowed search against this frame:
 *                            only the windows and the query descriptors travel (orbb200_window_search_resident)
 *   ComputeBoW(voc, v, fv)   Frame::ComputeBoW (:459-466) on the resident descriptors (orbb200_compute_bow_batch_device): only the
 *                            BowVector / FeatureVector come back
 *   Keys() / KeysUn() / Descriptors()   host copies, materialised on first read
 *
 * ORBmatcher's windowed searches take this path by themselves when the caller's Frame type has a member
 * `orbb200::DeviceFrame *mpDeviceFrame` that is not null (include/orbb200/ORBmatcher.h: window_search_frame); results are
 * identical to the host-buffer path -- same kernels, same candidate order.
 */
#ifndef ORBB200_DEVICEFRAME_H
#define ORBB200_DEVICEFRAME_H

#include <cstring>
#include <iostream>
#include <vector>

#include "ORBextractor.h"

namespace orbb200 {

class DeviceFrame {
public:
    DeviceFrame()
        : kps_d_(nullptr), kpsUn_d_(nullptr), desc_d_(nullptr), n_(0), capN_(0), device_(-1), minX_(0), minY_(0), invW_(0), invH_(0), cols_(64),
          rows_(48), hostKeys_(false), hostKeysUn_(false), hostDesc_(false)
    {
    }
    ~DeviceFrame() { release(); }
    DeviceFrame(const DeviceFrame &) = delete;
    DeviceFrame &operator=(const DeviceFrame &) = delete;

    int N() const { return n_; }
    const orbb200_keypoint *keys_device() const { return kps_d_; }
    const orbb200_keypoint *keys_un_device() const { return kpsUn_d_ ? kpsUn_d_ : kps_d_; }
    const uint8_t *descriptors_device() const { return desc_d_; }

    /* keep the outputs of the extractor's last operator() / ExtractColor call (image 0 of its handle) on the device */
    bool Adopt(ORBB200_NAMESPACE::ORBextractor &ex, int n, int device = -1)
    {
        const orbb200_keypoint *k = nullptr;
        const uint8_t *d = nullptr;
        int cap = 0;
        if (orbb200_extractor_outputs_device(ex.Handle(), &k, &d, nullptr, &cap) != ORBB200_OK) return fail();
        if (n < 0 || n > cap) n = n < 0 ? 0 : cap;
        device_ = device;
        if (!reserve(n)) return false;
        n_ = n;
        hostKeys_ = hostKeysUn_ = hostDesc_ = false;
        if (n == 0) return true;
        if (orbb200_extractor_sync(ex.Handle()) != ORBB200_OK) return fail();
        if (orbb200_copy_on_device(kps_d_, k, sizeof(orbb200_keypoint) * (size_t)n) != ORBB200_OK) return fail();
        if (orbb200_copy_on_device(desc_d_, d, (size_t)32 * n) != ORBB200_OK) return fail();
        unKeysValid_ = false;
        return true;
    }

    /* Frame::UndistortKeyPoints on the device copy; with k1 == 0 the undistorted keys ARE the keys */
    bool Undistort(const cv::Mat &mK, const cv::Mat &mDistCoef)
    {
        float dist[5] = {0, 0, 0, 0, 0};
        const int nCoef = mDistCoef.rows * mDistCoef.cols;
        for (int i = 0; i < nCoef && i < 5; i++) dist[i] = mDistCoef.ptr<float>()[i];
        unKeysValid_ = false;
        hostKeysUn_ = false;
        if (dist[0] == 0.0f || n_ == 0) return true;
        const float cam[4] = {mK.at<float>(0, 0), mK.at<float>(1, 1), mK.at<float>(0, 2), mK.at<float>(1, 2)};
        if (!kpsUn_d_ && orbb200_device_alloc((void **)&kpsUn_d_, sizeof(orbb200_keypoint) * (size_t)capN_, device_) != ORBB200_OK) return fail();
        // a null stream = the default stream, ordered behind the synchronous copies of Adopt
        if (orbb200_undistort_keypoints_device(kps_d_, n_, cam, dist, kpsUn_d_, nullptr) != ORBB200_OK) return fail();
        if (orbb200_device_sync(device_) != ORBB200_OK) return fail(); // the searches run on the library's own stream
        unKeysValid_ = true;
        return true;
    }

    void SetGrid(float minX, float minY, float gridElementWidthInv, float gridElementHeightInv, int cols = 64, int rows = 48)
    {
        minX_ = minX; minY_ = minY; invW_ = gridElementWidthInv; invH_ = gridElementHeightInv; cols_ = cols; rows_ = rows;
    }

    /* AssignFeaturesToGrid + GetFeaturesInArea for every window + all candidate distances, against the resident frame.
     * windows: 5 floats per query (x, y, r, minLevel, maxLevel); CSR result like ORBmatcher::window_search */
    bool WindowSearch(const std::vector<float> &windows, const std::vector<unsigned char> &qDesc, std::vector<int> &offs, std::vector<int> &cand,
                      std::vector<uint16_t> &dist)
    {
        const int nq = (int)(windows.size() / 5);
        offs.assign(nq + 1, 0);
        cand.assign(1, 0);
        dist.assign(1, 0);
        if (nq == 0 || n_ == 0) return true;
        int cap = nq * 48 + 64;
        for (int attempt = 0; attempt < 2; attempt++) {
            cand.assign(cap, 0);
            dist.assign(cap, 0);
            const int rc = orbb200_window_search_resident(unKeysValid_ ? kpsUn_d_ : kps_d_, desc_d_, n_, minX_, minY_, invW_, invH_, cols_, rows_,
                                                          windows.data(), qDesc.data(), nq, offs.data(), cand.data(), dist.data(), cap, device_);
            if (rc == ORBB200_OK) return true;
            if (rc == ORBB200_ERR_CAPACITY && offs[nq] > cap) { cap = offs[nq]; continue; }
            break;
        }
        offs.assign(nq + 1, 0);
        return fail();
    }

    /* Frame::ComputeBoW on the resident descriptors: voc.transform(mDescriptors, v, fv, 4) without the descriptors leaving the device.
     * Voc = orbb200::ORBVocabulary (anything with handle()), BowVec / FeatVec = DBoW2::BowVector / FeatureVector (std::map based). */
    template <class Voc, class BowVec, class FeatVec> bool ComputeBoW(const Voc &voc, BowVec &v, FeatVec &fv, int levelsup = 4)
    {
        v.clear();
        fv.clear();
        if (n_ == 0 || !voc.handle()) return true;
        if (capN_ > 8192) { std::cerr << "DeviceFrame::ComputeBoW: more than 8192 keypoint slots" << std::endl; return false; }
        if (!bow_d_) {
            // one block: n, n_bow, n_kept (ints, 64 bytes) | bow_id [cap] | bow_val [cap] doubles | fv_key [cap] u64
            const size_t bytes = 64 + (size_t)capN_ * (4 + 8 + 8) + 64;
            if (orbb200_device_alloc((void **)&bow_d_, bytes, device_) != ORBB200_OK) return fail();
            bowCap_ = capN_;
        }
        unsigned char *base = bow_d_;
        int *cnt_d = reinterpret_cast<int *>(base);
        double *val_d = reinterpret_cast<double *>(base + 64);
        unsigned long long *key_d = reinterpret_cast<unsigned long long *>(base + 64 + (size_t)bowCap_ * 8);
        int *id_d = reinterpret_cast<int *>(base + 64 + (size_t)bowCap_ * 16);
        if (orbb200_copy_to_device(cnt_d, &n_, sizeof(int)) != ORBB200_OK) return fail();
        if (orbb200_compute_bow_batch_device(voc.handle(), desc_d_, cnt_d, 1, bowCap_, levelsup, id_d, val_d, cnt_d + 1, key_d, cnt_d + 2, nullptr) !=
            ORBB200_OK)
            return fail();
        if (orbb200_device_sync(device_) != ORBB200_OK) return fail();
        int cnt[3];
        if (orbb200_copy_to_host(cnt, cnt_d, sizeof(cnt)) != ORBB200_OK) return fail();
        std::vector<int> ids(cnt[1] + 1);
        std::vector<double> vals(cnt[1] + 1);
        std::vector<unsigned long long> keys(cnt[2] + 1);
        if (cnt[1] && (orbb200_copy_to_host(ids.data(), id_d, sizeof(int) * (size_t)cnt[1]) != ORBB200_OK ||
                       orbb200_copy_to_host(vals.data(), val_d, sizeof(double) * (size_t)cnt[1]) != ORBB200_OK))
            return fail();
        if (cnt[2] && orbb200_copy_to_host(keys.data(), key_d, sizeof(unsigned long long) * (size_t)cnt[2]) != ORBB200_OK) return fail();
        for (int j = 0; j < cnt[1]; j++) v.insert(v.end(), typename BowVec::value_type((typename BowVec::key_type)ids[j], vals[j]));
        for (int j = 0; j < cnt[2]; j++) // keys ascend: node groups in map order, feature indices ascending inside a group
            fv[(typename FeatVec::key_type)(keys[j] >> 32)].push_back((unsigned int)(keys[j] & 0xFFFFFFFFull));
        return true;
    }

    /* host views, downloaded on first read */
    const std::vector<cv::KeyPoint> &Keys()
    {
        if (!hostKeys_) { download(kps_d_, keys_); hostKeys_ = true; }
        return keys_;
    }
    const std::vector<cv::KeyPoint> &KeysUn()
    {
        if (!unKeysValid_) return Keys();
        if (!hostKeysUn_) { download(kpsUn_d_, keysUn_); hostKeysUn_ = true; }
        return keysUn_;
    }
    const cv::Mat &Descriptors()
    {
        if (!hostDesc_) {
            if (n_ == 0) desc_.release();
            else {
                desc_.create(n_, 32, CV_8U);
                std::vector<unsigned char> tmp((size_t)n_ * 32);
                if (orbb200_copy_to_host(tmp.data(), desc_d_, tmp.size()) != ORBB200_OK) fail();
                for (int i = 0; i < n_; i++) std::memcpy(desc_.ptr(i), &tmp[(size_t)i * 32], 32);
            }
            hostDesc_ = true;
        }
        return desc_;
    }

private:
    bool fail()
    {
        std::cerr << "DeviceFrame (orbb200): " << orbb200_last_error() << std::endl;
        return false;
    }
    void release()
    {
        orbb200_device_free(kps_d_);
        orbb200_device_free(kpsUn_d_);
        orbb200_device_free(desc_d_);
        orbb200_device_free(bow_d_);
        bow_d_ = nullptr;
        kps_d_ = kpsUn_d_ = nullptr;
        desc_d_ = nullptr;
        capN_ = 0;
    }
    bool reserve(int n)
    {
        if (n <= capN_) return true;
        release();
        const int c = n + 64;
        if (orbb200_device_alloc((void **)&kps_d_, sizeof(orbb200_keypoint) * (size_t)c, device_) != ORBB200_OK) return fail();
        if (orbb200_device_alloc((void **)&desc_d_, (size_t)32 * c, device_) != ORBB200_OK) return fail();
        capN_ = c;
        return true;
    }
    void download(const orbb200_keypoint *src, std::vector<cv::KeyPoint> &dst)
    {
        static_assert(sizeof(cv::KeyPoint) == sizeof(orbb200_keypoint), "cv::KeyPoint must be the 28-byte POD");
        dst.resize(n_);
        if (n_ && orbb200_copy_to_host(dst.data(), src, sizeof(orbb200_keypoint) * (size_t)n_) != ORBB200_OK) fail();
    }

    orbb200_keypoint *kps_d_, *kpsUn_d_;
    uint8_t *desc_d_;
    unsigned char *bow_d_ = nullptr; // ComputeBoW's device block (allocated on first use)
    int bowCap_ = 0;
    int n_, capN_, device_;
    float minX_, minY_, invW_, invH_;
    int cols_, rows_;
    bool unKeysValid_ = false;
    bool hostKeys_, hostKeysUn_, hostDesc_;
    std::vector<cv::KeyPoint> keys_, keysUn_;
    cv::Mat desc_;
};

} // namespace orbb200

#endif /* ORBB200_DEVICEFRAME_H */

/*
 * MapPointDescriptors.h -- MapPoint::ComputeDistinctiveDescriptors (reference src/MapPoint.cc:266-340) for MANY map points
 * in one device call.
 *
 * The reference recomputes the representative descriptor one map point at a time, under that point's mutex
 * (callers loop over a keyframe's points: src/LocalMapping.cc:172,448,602,637; src/LoopClosing.cc:562;
 * src/Tracking.cc:592,1232).  The N x N Hamming matrix and the per-row medians are the arithmetic; which cv::Mat ends up in
 * mDescriptor is the side effect.  Here the arithmetic of a whole loop runs in ONE kernel launch
 * (orbb200_distinctive_descriptors) and the side effect stays with the caller:
 *
 *   1. per map point, gather the observed descriptors exactly like :270-299 (non-bad keyframes, map iteration order),
 *   2. ORB_SLAM2::SelectDistinctiveDescriptors(groups, best)   -> best[g] = index into groups[g] (-1 = leave untouched),
 *   3. per map point, mDescriptor = groups[g][best[g]].clone() under its mutex (:336-339).
 *
 * ComputeDistinctiveDescriptorsBatch() below does 1-3 for any MapPoint type that offers the two small accessors named in
 * INTEGRATION.md (CollectObservedDescriptors / SetDescriptor); tests/cpp/facade_distinct.cpp binds it to a stand-in.
 */
#ifndef ORBB200_MAPPOINTDESCRIPTORS_H
#define ORBB200_MAPPOINTDESCRIPTORS_H

#include <cstring>
#include <iostream>
#include <vector>

#ifdef ORBB200_USE_OPENCV
#include <opencv2/core/core.hpp>
#else
#include "cv_shim.h"
#endif
#include "../orbb200.h"

/* The classes live in ORB_SLAM2 so that the reference's callers compile unchanged.  A test that links the reference's own
   classes of the same name beside these (tests/cpp/ref_match.cpp) renames the namespace with -DORBB200_NAMESPACE=... */
#ifndef ORBB200_NAMESPACE
#define ORBB200_NAMESPACE ORB_SLAM2
#endif
namespace ORBB200_NAMESPACE {

// groups[g] = vDescriptors of map point g (:270-299), each a 1x32 CV_8U row.  best[g] = BestIdx of :319-337, or -1 when the
// group is empty (the reference returns before touching mDescriptor, :281,296).  Returns false on a device error (best is
// then all -1: every point keeps its previous descriptor, which is what a skipped call leaves behind).
inline bool SelectDistinctiveDescriptors(const std::vector<std::vector<cv::Mat> > &groups, std::vector<int> &best, int device = -1)
{
    const int G = (int)groups.size();
    best.assign(G, -1);
    if (G == 0) return true;
    std::vector<int> offsets(G + 1, 0);
    for (int g = 0; g < G; g++) offsets[g + 1] = offsets[g] + (int)groups[g].size();
    std::vector<unsigned char> desc((size_t)offsets[G] * 32 + 32);
    for (int g = 0; g < G; g++)
        for (size_t i = 0; i < groups[g].size(); i++)
            std::memcpy(&desc[((size_t)offsets[g] + i) * 32], groups[g][i].ptr(), 32);
    const int rc = orbb200_distinctive_descriptors(desc.data(), offsets.data(), G, best.data(), nullptr, device);
    if (rc != ORBB200_OK) {
        std::cerr << "orbb200: ComputeDistinctiveDescriptors failed: " << orbb200_last_error() << std::endl;
        best.assign(G, -1);
        return false;
    }
    return true;
}

// Replacement for `for (pMP : points) pMP->ComputeDistinctiveDescriptors();`
// MapPointT must offer   bool CollectObservedDescriptors(std::vector<cv::Mat>&)   (the gather of :270-299; false when
// the point is bad) and   void SetDescriptor(const cv::Mat&)   (the locked assignment of :336-339).  Null pointers skipped.
template <class MapPointT> void ComputeDistinctiveDescriptorsBatch(const std::vector<MapPointT *> &points, int device = -1)
{
    std::vector<std::vector<cv::Mat> > groups(points.size());
    for (size_t g = 0; g < points.size(); g++)
        if (points[g] && !points[g]->CollectObservedDescriptors(groups[g])) groups[g].clear();
    std::vector<int> best;
    SelectDistinctiveDescriptors(groups, best, device);
    for (size_t g = 0; g < points.size(); g++)
        if (points[g] && best[g] >= 0) points[g]->SetDescriptor(groups[g][best[g]].clone());
}

} // namespace ORBB200_NAMESPACE (ORB_SLAM2 by default)

#endif

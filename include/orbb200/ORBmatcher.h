/*
 * ORBmatcher.h -- ORB_SLAM2::ORBmatcher (reference include/ORBmatcher.h:38-100, src/ORBmatcher.cc) with the
 * descriptor-distance arithmetic of the searches on the GPU through the C-ABI (include/orbb200.h).
 *
 * Shape of every search: candidate generation -> DescriptorDistance per (query, candidate) ->
 * running best / second-best -> accept test -> ordered side effect -> rotation-histogram cull.
 * Here: a host prologue enumerates candidates with the caller's own, unchanged Frame::GetFeaturesInArea /
 * FeatureVector walk and packs descriptors; ONE call to orbb200_candidate_distances computes every distance on the
 * device; a host epilogue replays the reference loop in query order on those distances, so skip-if-claimed rules
 * (src/ORBmatcher.cc:93-95, 222), strict '<' tie-breaking, thresholds, ratio test and the histogram are the
 * reference's, statement for statement.
 *
 * The Search* members are templates over the caller's Frame / KeyFrame / MapPoint types: included in the reference
 * tree they bind to ORB_SLAM2::Frame etc. (same member names: mvKeysUn, mDescriptors, mvpMapPoints, mFeatVec, ...),
 * and this repo's tests bind them to small stand-ins (tests/cpp/facade_match.cpp).  The windowed searches send
 * their windows to orbb200_window_search (grid build + GetFeaturesInArea + distances on the device); the BoW-node
 * searches send CSR lists to orbb200_candidate_distances.  All twelve public searches of the reference class are here
 * (M1-M9 of SURVEY.md 8a plus the Sim3 overloads of loop closing); pose arithmetic on the host follows OpenCV's evaluation
 * of the reference's CV_32F expressions (MatExpr folding, cv::gemm's float small-matrix branch vs its double general branch,
 * cv::norm / Mat::dot in double) so projections agree to the bit; tests/cpp/ref_match.cpp checks it against the compiled reference.
 */
#ifndef ORBB200_ORBMATCHER_H
#define ORBB200_ORBMATCHER_H

#include <cmath>
#include <cstdint>
#include <cstring>
#include <iostream>
#include <set>
#include <type_traits>
#include <utility>
#include <vector>

#ifdef ORBB200_USE_OPENCV
#include <opencv2/core/core.hpp>
#else
#include "cv_shim.h"
#endif
#include "../orbb200.h"
#include "DeviceFrame.h"

/* The classes live in ORB_SLAM2 so that the reference's callers compile unchanged.  A test that links the reference's own
   classes of the same name beside these (tests/cpp/ref_match.cpp) renames the namespace with -DORBB200_NAMESPACE=... */
#ifndef ORBB200_NAMESPACE
#define ORBB200_NAMESPACE ORB_SLAM2
#endif
/* the restated OpenCV float accumulations must not be contracted into FMAs even when the including project builds with
   -O3 -march=native (the reference's CMake flags) */
#if defined(__GNUC__) && !defined(__clang__)
#define ORBB200_NO_FMA __attribute__((optimize("-ffp-contract=off")))
#else
#define ORBB200_NO_FMA
#endif
namespace ORBB200_NAMESPACE {

class ORBmatcher {
public:
    ORBmatcher(float nnratio = 0.6, bool checkOri = true) : mfNNratio(nnratio), mbCheckOrientation(checkOri), device_(-1) {}

    // Hamming distance of two 32-byte descriptors (src/ORBmatcher.cc:1733-1749).  A single pair stays on the host.
    static int DescriptorDistance(const cv::Mat &a, const cv::Mat &b)
    {
        const uint32_t *pa = a.ptr<uint32_t>(), *pb = b.ptr<uint32_t>();
        int dist = 0;
        for (int i = 0; i < 8; i++) dist += __builtin_popcount(pa[i] ^ pb[i]);
        return dist;
    }

    // Track local map (src/ORBmatcher.cc:48-136; caller src/Tracking.cc:1463).
    template <class FrameT, class MapPointT>
    int SearchByProjection(FrameT &F, const std::vector<MapPointT *> &vpMapPoints, const float th = 3)
    {
        const bool bFactor = th != 1.0;
        // prologue: queries in reference order with their windows
        std::vector<int> qMP;                 // index into vpMapPoints
        std::vector<float> qR;                // r * scale[level]
        std::vector<float> windows;           // F.GetFeaturesInArea arguments (:77-78), evaluated on the device
        std::vector<unsigned char> qDesc;
        for (size_t iMP = 0; iMP < vpMapPoints.size(); iMP++) {
            MapPointT *pMP = vpMapPoints[iMP];
            if (!pMP->mbTrackInView) continue;
            if (pMP->isBad()) continue;
            const int nPredictedLevel = pMP->mnTrackScaleLevel;
            float r = RadiusByViewingCos(pMP->mTrackViewCos);
            if (bFactor) r *= th;
            const float win = r * F.mvScaleFactors[nPredictedLevel];
            qMP.push_back((int)iMP);
            qR.push_back(win);
            const float wq[5] = {pMP->mTrackProjX, pMP->mTrackProjY, win, (float)(nPredictedLevel - 1), (float)nPredictedLevel};
            windows.insert(windows.end(), wq, wq + 5);
            const cv::Mat d = pMP->GetDescriptor();
            qDesc.insert(qDesc.end(), d.ptr(), d.ptr() + 32);
        }
        std::vector<int> offs, cand;
        std::vector<uint16_t> dist;
        if (!window_search_frame(F, windows, qDesc, offs, cand, dist)) return 0;
        // epilogue: the reference loop (:86-131) on precomputed distances
        int nmatches = 0;
        for (size_t q = 0; q < qMP.size(); q++) {
            MapPointT *pMP = vpMapPoints[qMP[q]];
            int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
            for (int c = offs[q]; c < offs[q + 1]; c++) {
                const size_t idx = (size_t)cand[c];
                if (F.mvpMapPoints[idx])
                    if (F.mvpMapPoints[idx]->Observations() > 0) continue;
                if (F.mvuRight[idx] > 0) {
                    const float er = std::fabs(pMP->mTrackProjXR - F.mvuRight[idx]);
                    if (er > qR[q]) continue;
                }
                const int d = dist[c];
                if (d < bestDist) {
                    bestDist2 = bestDist; bestDist = d;
                    bestLevel2 = bestLevel; bestLevel = F.mvKeysUn[idx].octave;
                    bestIdx = (int)idx;
                } else if (d < bestDist2) {
                    bestLevel2 = F.mvKeysUn[idx].octave;
                    bestDist2 = d;
                }
            }
            if (bestDist <= TH_HIGH) {
                if (bestLevel == bestLevel2 && bestDist > mfNNratio * bestDist2) continue;
                F.mvpMapPoints[bestIdx] = pMP;
                nmatches++;
            }
        }
        return nmatches;
    }

    // Relocalisation / reference-keyframe tracking (src/ORBmatcher.cc:171-303; callers src/Tracking.cc:907, 1682).
    template <class KeyFrameT, class FrameT, class MapPointT>
    int SearchByBoW(KeyFrameT *pKF, FrameT &F, std::vector<MapPointT *> &vpMapPointMatches)
    {
        const std::vector<MapPointT *> vpMapPointsKF = pKF->GetMapPointMatches();
        vpMapPointMatches = std::vector<MapPointT *>(F.N, static_cast<MapPointT *>(NULL));
        // prologue: merge-walk of the two feature vectors (:193-279); one query per keyframe feature with a good point
        std::vector<int> qKF, offs(1, 0), cand;
        std::vector<unsigned char> qDesc;
        typename std::remove_reference<decltype(pKF->mFeatVec)>::type::const_iterator KFit = pKF->mFeatVec.begin(), KFend = pKF->mFeatVec.end();
        typename std::remove_reference<decltype(F.mFeatVec)>::type::const_iterator Fit = F.mFeatVec.begin(), Fend = F.mFeatVec.end();
        while (KFit != KFend && Fit != Fend) {
            if (KFit->first == Fit->first) {
                const std::vector<unsigned int> &vIndicesKF = KFit->second, &vIndicesF = Fit->second;
                for (size_t iKF = 0; iKF < vIndicesKF.size(); iKF++) {
                    const unsigned int realIdxKF = vIndicesKF[iKF];
                    MapPointT *pMP = vpMapPointsKF[realIdxKF];
                    if (!pMP) continue;
                    if (pMP->isBad()) continue;
                    qKF.push_back((int)realIdxKF);
                    for (size_t iF = 0; iF < vIndicesF.size(); iF++) cand.push_back((int)vIndicesF[iF]);
                    offs.push_back((int)cand.size());
                    const unsigned char *d = pKF->mDescriptors.ptr((int)realIdxKF);
                    qDesc.insert(qDesc.end(), d, d + 32);
                }
                KFit++;
                Fit++;
            } else if (KFit->first < Fit->first) {
                KFit = pKF->mFeatVec.lower_bound(Fit->first);
            } else {
                Fit = F.mFeatVec.lower_bound(KFit->first);
            }
        }
        std::vector<uint16_t> dist;
        if (!distances(qDesc, (int)qKF.size(), F.mDescriptors, offs, cand, dist)) return 0;
        // epilogue (:214-266, :282-300)
        int nmatches = 0;
        std::vector<int> rotHist[HISTO_LENGTH];
        const float factor = 1.0f / HISTO_LENGTH;
        for (size_t q = 0; q < qKF.size(); q++) {
            const int realIdxKF = qKF[q];
            int bestDist1 = 256, bestIdxF = -1, bestDist2 = 256;
            for (int c = offs[q]; c < offs[q + 1]; c++) {
                const int realIdxF = cand[c];
                if (vpMapPointMatches[realIdxF]) continue;
                const int d = dist[c];
                if (d < bestDist1) { bestDist2 = bestDist1; bestDist1 = d; bestIdxF = realIdxF; }
                else if (d < bestDist2) bestDist2 = d;
            }
            if (bestDist1 <= TH_LOW) {
                if (static_cast<float>(bestDist1) < mfNNratio * static_cast<float>(bestDist2)) {
                    vpMapPointMatches[bestIdxF] = vpMapPointsKF[realIdxKF];
                    if (mbCheckOrientation) {
                        float rot = pKF->mvKeysUn[realIdxKF].angle - F.mvKeys[bestIdxF].angle;
                        if (rot < 0.0) rot += 360.0f;
                        int bin = (int)std::round(rot * factor);
                        if (bin == HISTO_LENGTH) bin = 0;
                        rotHist[bin].push_back(bestIdxF);
                    }
                    nmatches++;
                }
            }
        }
        if (mbCheckOrientation) {
            int ind1 = -1, ind2 = -1, ind3 = -1;
            ComputeThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
            for (int i = 0; i < HISTO_LENGTH; i++) {
                if (i == ind1 || i == ind2 || i == ind3) continue;
                for (size_t j = 0, jend = rotHist[i].size(); j < jend; j++) {
                    vpMapPointMatches[rotHist[i][j]] = static_cast<MapPointT *>(NULL);
                    nmatches--;
                }
            }
        }
        return nmatches;
    }

    // Motion-model tracking (src/ORBmatcher.cc:1396-1553; callers src/Tracking.cc:1065, 1072).
    template <class FrameT>
    int SearchByProjection(FrameT &CurrentFrame, const FrameT &LastFrame, const float th, const bool bMono)
    {
        typedef typename std::remove_pointer<typename std::remove_reference<decltype(LastFrame.mvpMapPoints[0])>::type>::type MapPointT;
        int nmatches = 0;
        std::vector<int> rotHist[HISTO_LENGTH];
        const float factor = 1.0f / HISTO_LENGTH;
        float Rcw[9], tcw[3], Rlw[9], tlw[3];
        pose_Rt(CurrentFrame.mTcw, Rcw, tcw);
        pose_Rt(LastFrame.mTcw, Rlw, tlw);
        // twc = -Rcw^T * tcw (transposed operand: double accumulation) ; tlc = Rlw * twc + tlw (small-matrix branch: float)
        float twc[3], tlc[3];
        camera_centre(Rcw, tcw, twc);
        mat_Rx_t(Rlw, twc, tlw, tlc);
        const bool bForward = tlc[2] > CurrentFrame.mb && !bMono;
        const bool bBackward = -tlc[2] > CurrentFrame.mb && !bMono;
        // prologue: projections and windows, in LastFrame index order (:1419-1472)
        std::vector<int> qI;
        std::vector<float> win, qU, qInvz, qRadius;
        std::vector<unsigned char> qDesc;
        for (int i = 0; i < LastFrame.N; i++) {
            MapPointT *pMP = LastFrame.mvpMapPoints[i];
            if (!pMP) continue;
            if (LastFrame.mvbOutlier[i]) continue;
            const cv::Mat x3Dw = pMP->GetWorldPos();
            const float xw[3] = {x3Dw.template at<float>(0, 0), x3Dw.template at<float>(1, 0), x3Dw.template at<float>(2, 0)};
            float x3Dc[3];
            mat_Rx_t(Rcw, xw, tcw, x3Dc);
            const float xc = x3Dc[0], yc = x3Dc[1];
            const float invzc = 1.0 / x3Dc[2];
            if (invzc < 0) continue;
            const float u = CurrentFrame.fx * xc * invzc + CurrentFrame.cx;
            const float v = CurrentFrame.fy * yc * invzc + CurrentFrame.cy;
            if (u < CurrentFrame.mnMinX || u > CurrentFrame.mnMaxX) continue;
            if (v < CurrentFrame.mnMinY || v > CurrentFrame.mnMaxY) continue;
            const int nLastOctave = LastFrame.mvKeys[i].octave;
            const float radius = th * CurrentFrame.mvScaleFactors[nLastOctave];
            int minL, maxL;
            if (bForward) { minL = nLastOctave; maxL = -1; }
            else if (bBackward) { minL = 0; maxL = nLastOctave; }
            else { minL = nLastOctave - 1; maxL = nLastOctave + 1; }
            qI.push_back(i); qU.push_back(u); qInvz.push_back(invzc); qRadius.push_back(radius);
            const float wq[5] = {u, v, radius, (float)minL, (float)maxL};
            win.insert(win.end(), wq, wq + 5);
            const cv::Mat d = pMP->GetDescriptor();
            qDesc.insert(qDesc.end(), d.ptr(), d.ptr() + 32);
        }
        std::vector<int> offs, cand;
        std::vector<uint16_t> dist;
        if (!window_search_frame(CurrentFrame, win, qDesc, offs, cand, dist)) return 0;
        // epilogue (:1474-1527)
        for (size_t q = 0; q < qI.size(); q++) {
            if (offs[q] == offs[q + 1]) continue;
            const int i = qI[q];
            MapPointT *pMP = LastFrame.mvpMapPoints[i];
            int bestDist = 256, bestIdx2 = -1;
            for (int c = offs[q]; c < offs[q + 1]; c++) {
                const size_t i2 = (size_t)cand[c];
                if (CurrentFrame.mvpMapPoints[i2])
                    if (CurrentFrame.mvpMapPoints[i2]->Observations() > 0) continue;
                if (CurrentFrame.mvuRight[i2] > 0) {
                    const float ur = qU[q] - CurrentFrame.mbf * qInvz[q];
                    const float er = std::fabs(ur - CurrentFrame.mvuRight[i2]);
                    if (er > qRadius[q]) continue;
                }
                const int d = dist[c];
                if (d < bestDist) { bestDist = d; bestIdx2 = (int)i2; }
            }
            if (bestDist <= TH_HIGH) {
                CurrentFrame.mvpMapPoints[bestIdx2] = pMP;
                nmatches++;
                if (mbCheckOrientation) {
                    float rot = LastFrame.mvKeysUn[i].angle - CurrentFrame.mvKeysUn[bestIdx2].angle;
                    if (rot < 0.0) rot += 360.0f;
                    int bin = (int)std::round(rot * factor);
                    if (bin == HISTO_LENGTH) bin = 0;
                    rotHist[bin].push_back(bestIdx2);
                }
            }
        }
        if (mbCheckOrientation) {
            int ind1 = -1, ind2 = -1, ind3 = -1;
            ComputeThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
            for (int i = 0; i < HISTO_LENGTH; i++) {
                if (i != ind1 && i != ind2 && i != ind3) {
                    for (size_t j = 0, jend = rotHist[i].size(); j < jend; j++) {
                        CurrentFrame.mvpMapPoints[rotHist[i][j]] = static_cast<MapPointT *>(NULL);
                        nmatches--;
                    }
                }
            }
        }
        return nmatches;
    }

    // Monocular map initialisation (src/ORBmatcher.cc:430-563; caller src/Tracking.cc:679).
    template <class FrameT>
    int SearchForInitialization(FrameT &F1, FrameT &F2, std::vector<cv::Point2f> &vbPrevMatched, std::vector<int> &vnMatches12,
                                int windowSize = 10)
    {
        int nmatches = 0;
        vnMatches12 = std::vector<int>(F1.mvKeysUn.size(), -1);
        std::vector<int> rotHist[HISTO_LENGTH];
        const float factor = 1.0f / HISTO_LENGTH;
        std::vector<int> vMatchedDistance(F2.mvKeysUn.size(), INT_MAX_);
        std::vector<int> vnMatches21(F2.mvKeysUn.size(), -1);
        // prologue: one window per level-0 keypoint of F1 (:446-458)
        std::vector<int> qI1;
        std::vector<float> windows;
        std::vector<unsigned char> qDesc;
        for (size_t i1 = 0, iend1 = F1.mvKeysUn.size(); i1 < iend1; i1++) {
            const int level1 = F1.mvKeysUn[i1].octave;
            if (level1 > 0) continue;
            qI1.push_back((int)i1);
            const float wq[5] = {vbPrevMatched[i1].x, vbPrevMatched[i1].y, (float)windowSize, (float)level1, (float)level1};
            windows.insert(windows.end(), wq, wq + 5);
            const unsigned char *d = F1.mDescriptors.ptr((int)i1);
            qDesc.insert(qDesc.end(), d, d + 32);
        }
        std::vector<int> offs, cand;
        std::vector<uint16_t> dist;
        if (!window_search_frame(F2, windows, qDesc, offs, cand, dist)) return 0;
        // epilogue (:460-514): order dependent through vMatchedDistance / vnMatches21
        for (size_t q = 0; q < qI1.size(); q++) {
            if (offs[q] == offs[q + 1]) continue;
            const int i1 = qI1[q];
            int bestDist = INT_MAX_, bestDist2 = INT_MAX_, bestIdx2 = -1;
            for (int c = offs[q]; c < offs[q + 1]; c++) {
                const int i2 = cand[c];
                const int d = dist[c];
                if (vMatchedDistance[i2] <= d) continue;
                if (d < bestDist) { bestDist2 = bestDist; bestDist = d; bestIdx2 = i2; }
                else if (d < bestDist2) bestDist2 = d;
            }
            if (bestDist <= TH_LOW) {
                if (bestDist < (float)bestDist2 * mfNNratio) {
                    if (vnMatches21[bestIdx2] >= 0) {
                        vnMatches12[vnMatches21[bestIdx2]] = -1;
                        nmatches--;
                    }
                    vnMatches12[i1] = bestIdx2;
                    vnMatches21[bestIdx2] = i1;
                    vMatchedDistance[bestIdx2] = bestDist;
                    nmatches++;
                    if (mbCheckOrientation) {
                        float rot = F1.mvKeysUn[i1].angle - F2.mvKeysUn[bestIdx2].angle;
                        if (rot < 0.0) rot += 360.0f;
                        int bin = (int)std::round(rot * factor);
                        if (bin == HISTO_LENGTH) bin = 0;
                        rotHist[bin].push_back(i1);
                    }
                }
            }
        }
        if (mbCheckOrientation) {
            int ind1 = -1, ind2 = -1, ind3 = -1;
            ComputeThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
            for (int i = 0; i < HISTO_LENGTH; i++) {
                if (i == ind1 || i == ind2 || i == ind3) continue;
                for (size_t j = 0, jend = rotHist[i].size(); j < jend; j++) {
                    const int idx1 = rotHist[i][j];
                    if (vnMatches12[idx1] >= 0) { vnMatches12[idx1] = -1; nmatches--; }
                }
            }
        }
        for (size_t i1 = 0, iend1 = vnMatches12.size(); i1 < iend1; i1++)
            if (vnMatches12[i1] >= 0) vbPrevMatched[i1] = F2.mvKeysUn[vnMatches12[i1]].pt;
        return nmatches;
    }

    // Triangulation candidates between two keyframes (src/ORBmatcher.cc:704-871; caller src/LocalMapping.cc:309).
    template <class KeyFrameT>
    int SearchForTriangulation(KeyFrameT *pKF1, KeyFrameT *pKF2, cv::Mat F12, std::vector<std::pair<size_t, size_t> > &vMatchedPairs,
                               const bool bOnlyStereo)
    {
        // epipole of camera 1 in image 2 (:714-721)
        const cv::Mat Cw = pKF1->GetCameraCenter(), R2w = pKF2->GetRotation(), t2w = pKF2->GetTranslation();
        float R[9], c1[3], t2[3], C2[3];
        for (int r = 0; r < 3; r++) {
            for (int c = 0; c < 3; c++) R[3 * r + c] = R2w.template at<float>(r, c);
            c1[r] = Cw.template at<float>(r, 0);
            t2[r] = t2w.template at<float>(r, 0);
        }
        mat_Rx_t(R, c1, t2, C2);
        const float invz = 1.0f / C2[2];
        const float ex = pKF2->fx * C2[0] * invz + pKF2->cx;
        const float ey = pKF2->fy * C2[1] * invz + pKF2->cy;
        int nmatches = 0;
        std::vector<int> vMatches12(pKF1->N, -1);
        std::vector<int> rotHist[HISTO_LENGTH];
        const float factor = 1.0f / HISTO_LENGTH;
        // prologue: BoW-node merge walk (:737-842); the per-candidate gates that do not depend on the distance
        // (existing map point, stereo flag; vbMatched2 is never set in this code base) are applied here
        std::vector<int> qIdx1, offs(1, 0), cand;
        std::vector<unsigned char> qDesc;
        typename std::remove_reference<decltype(pKF1->mFeatVec)>::type::const_iterator f1it = pKF1->mFeatVec.begin(), f1end = pKF1->mFeatVec.end();
        typename std::remove_reference<decltype(pKF2->mFeatVec)>::type::const_iterator f2it = pKF2->mFeatVec.begin(), f2end = pKF2->mFeatVec.end();
        while (f1it != f1end && f2it != f2end) {
            if (f1it->first == f2it->first) {
                for (size_t i1 = 0, iend1 = f1it->second.size(); i1 < iend1; i1++) {
                    const size_t idx1 = f1it->second[i1];
                    if (pKF1->GetMapPoint(idx1)) continue;
                    const bool bStereo1 = pKF1->mvuRight[idx1] >= 0;
                    if (bOnlyStereo && !bStereo1) continue;
                    qIdx1.push_back((int)idx1);
                    for (size_t i2 = 0, iend2 = f2it->second.size(); i2 < iend2; i2++) {
                        const size_t idx2 = f2it->second[i2];
                        if (pKF2->GetMapPoint(idx2)) continue;
                        if (bOnlyStereo && !(pKF2->mvuRight[idx2] >= 0)) continue;
                        cand.push_back((int)idx2);
                    }
                    offs.push_back((int)cand.size());
                    const unsigned char *d = pKF1->mDescriptors.ptr((int)idx1);
                    qDesc.insert(qDesc.end(), d, d + 32);
                }
                f1it++;
                f2it++;
            } else if (f1it->first < f2it->first) {
                f1it = pKF1->mFeatVec.lower_bound(f2it->first);
            } else {
                f2it = pKF2->mFeatVec.lower_bound(f1it->first);
            }
        }
        // ONE device call resolves every query (:757-815): distances, TH_LOW, the epipole gate and CheckDistEpipolarLine, a later
        // candidate winning at equal distance.  (vbMatched2 is never set in this code base, so the queries are independent.)
        const int nq = (int)qIdx1.size(), nt = pKF2->N;
        std::vector<float> qxy((size_t)2 * nq), Ff(9);
        std::vector<unsigned char> qst(nq), tst(nt > 0 ? nt : 1);
        for (int q = 0; q < nq; q++) {
            const cv::KeyPoint &kp1 = pKF1->mvKeysUn[qIdx1[q]];
            qxy[2 * q] = kp1.pt.x; qxy[2 * q + 1] = kp1.pt.y;
            qst[q] = pKF1->mvuRight[qIdx1[q]] >= 0 ? 1 : 0;
        }
        for (int i = 0; i < nt; i++) tst[i] = pKF2->mvuRight[i] >= 0 ? 1 : 0;
        for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) Ff[3 * r + c] = F12.template at<float>(r, c);
        std::vector<int> best2(nq > 0 ? nq : 1, -1);
        if (nq > 0) {
            static_assert(sizeof(cv::KeyPoint) == sizeof(orbb200_keypoint), "cv::KeyPoint must be the 28-byte POD");
            const cv::Mat &D2 = pKF2->mDescriptors;
            std::vector<unsigned char> tpacked;
            const unsigned char *tptr = D2.ptr(0);
            if (!D2.isContinuous()) {
                tpacked.resize((size_t)nt * 32);
                for (int i = 0; i < nt; i++) std::memcpy(&tpacked[(size_t)i * 32], D2.ptr(i), 32);
                tptr = tpacked.data();
            }
            const int rc = orbb200_search_triangulation(qDesc.data(), qxy.data(), qst.data(), nq, tptr,
                                                        reinterpret_cast<const orbb200_keypoint *>(pKF2->mvKeysUn.data()), tst.data(), nt, offs.data(),
                                                        cand.empty() ? nullptr : cand.data(), Ff.data(), ex, ey, pKF2->mvScaleFactors.data(),
                                                        pKF2->mvLevelSigma2.data(), (int)pKF2->mvScaleFactors.size(), TH_LOW, best2.data(), nullptr, device_);
            if (rc != ORBB200_OK) {
                std::cerr << "ORBmatcher (orbb200): " << orbb200_last_error() << std::endl;
                vMatchedPairs.clear();
                return 0;
            }
        }
        // epilogue (:816-835): matches, rotation histogram
        for (size_t q = 0; q < qIdx1.size(); q++) {
            const int idx1 = qIdx1[q];
            const cv::KeyPoint &kp1 = pKF1->mvKeysUn[idx1];
            const int bestIdx2 = best2[q];
            if (bestIdx2 >= 0) {
                const cv::KeyPoint &kp2 = pKF2->mvKeysUn[bestIdx2];
                vMatches12[idx1] = bestIdx2;
                nmatches++;
                if (mbCheckOrientation) {
                    float rot = kp1.angle - kp2.angle;
                    if (rot < 0.0) rot += 360.0f;
                    int bin = (int)std::round(rot * factor);
                    if (bin == HISTO_LENGTH) bin = 0;
                    rotHist[bin].push_back(idx1);
                }
            }
        }
        if (mbCheckOrientation) {
            int ind1 = -1, ind2 = -1, ind3 = -1;
            ComputeThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
            for (int i = 0; i < HISTO_LENGTH; i++) {
                if (i == ind1 || i == ind2 || i == ind3) continue;
                for (size_t j = 0, jend = rotHist[i].size(); j < jend; j++) {
                    vMatches12[rotHist[i][j]] = -1;
                    nmatches--;
                }
            }
        }
        vMatchedPairs.clear();
        vMatchedPairs.reserve(nmatches);
        for (size_t i = 0, iend = vMatches12.size(); i < iend; i++) {
            if (vMatches12[i] < 0) continue;
            vMatchedPairs.push_back(std::make_pair(i, (size_t)vMatches12[i]));
        }
        return nmatches;
    }

    // Fuse map points into a keyframe (src/ORBmatcher.cc:873-1038; callers src/LocalMapping.cc:553, 579).
    template <class KeyFrameT, class MapPointT>
    int Fuse(KeyFrameT *pKF, const std::vector<MapPointT *> &vpMapPoints, const float th = 3.0)
    {
        const cv::Mat Rm = pKF->GetRotation(), tm = pKF->GetTranslation(), Om = pKF->GetCameraCenter();
        float Rcw[9], tcw[3], Ow[3];
        for (int r = 0; r < 3; r++) {
            for (int c = 0; c < 3; c++) Rcw[3 * r + c] = Rm.template at<float>(r, c);
            tcw[r] = tm.template at<float>(r, 0);
            Ow[r] = Om.template at<float>(r, 0);
        }
        const float fx = pKF->fx, fy = pKF->fy, cx = pKF->cx, cy = pKF->cy, bf = pKF->mbf;
        // prologue (:892-943): projection, distance / viewing-angle gates, predicted level -> one window per map point
        std::vector<int> qI, qLevel;
        std::vector<float> qU, qV, qUr, windows;
        std::vector<unsigned char> qDesc;
        const int nMPs = (int)vpMapPoints.size();
        for (int i = 0; i < nMPs; i++) {
            MapPointT *pMP = vpMapPoints[i];
            if (!pMP) continue;
            if (pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;
            const cv::Mat p3Dw = pMP->GetWorldPos();
            const float pw[3] = {p3Dw.template at<float>(0, 0), p3Dw.template at<float>(1, 0), p3Dw.template at<float>(2, 0)};
            float p3Dc[3];
            mat_Rx_t(Rcw, pw, tcw, p3Dc);
            if (p3Dc[2] < 0.0f) continue;
            const float invz = 1 / p3Dc[2];
            const float x = p3Dc[0] * invz, y = p3Dc[1] * invz;
            const float u = fx * x + cx, v = fy * y + cy;
            if (!pKF->IsInImage(u, v)) continue;
            const float ur = u - bf * invz;
            const float maxDistance = pMP->GetMaxDistanceInvariance(), minDistance = pMP->GetMinDistanceInvariance();
            const float PO[3] = {pw[0] - Ow[0], pw[1] - Ow[1], pw[2] - Ow[2]};
            const float dist3D = (float)std::sqrt((double)PO[0] * PO[0] + (double)PO[1] * PO[1] + (double)PO[2] * PO[2]); // cv::norm
            if (dist3D < minDistance || dist3D > maxDistance) continue;
            const cv::Mat Pn = pMP->GetNormal();
            const double dot = (double)PO[0] * Pn.template at<float>(0, 0) + (double)PO[1] * Pn.template at<float>(1, 0) +
                               (double)PO[2] * Pn.template at<float>(2, 0); // cv::Mat::dot
            if (dot < 0.5 * dist3D) continue;
            const int nPredictedLevel = pMP->PredictScale(dist3D, pKF);
            const float radius = th * pKF->mvScaleFactors[nPredictedLevel];
            qI.push_back(i); qLevel.push_back(nPredictedLevel); qU.push_back(u); qV.push_back(v); qUr.push_back(ur);
            const float wq[5] = {u, v, radius, -1.f, -1.f}; // KeyFrame::GetFeaturesInArea has no level filter (KeyFrame.cc:607)
            windows.insert(windows.end(), wq, wq + 5);
            const cv::Mat d = pMP->GetDescriptor();
            qDesc.insert(qDesc.end(), d.ptr(), d.ptr() + 32);
        }
        std::vector<int> offs, cand;
        std::vector<uint16_t> dist;
        if (!window_search(pKF->mvKeysUn, pKF->mDescriptors, pKF->mnMinX, pKF->mnMinY, pKF->mfGridElementWidthInv,
                           pKF->mfGridElementHeightInv, pKF->mnGridCols, pKF->mnGridRows, windows, qDesc, offs, cand, dist))
            return 0;
        // epilogue (:945-1034), in map-point order; the graph mutations of earlier iterations are re-checked
        int nFused = 0;
        for (size_t q = 0; q < qI.size(); q++) {
            MapPointT *pMP = vpMapPoints[qI[q]];
            if (pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;
            const int nPredictedLevel = qLevel[q];
            const float u = qU[q], v = qV[q], ur = qUr[q];
            int bestDist = 256, bestIdx = -1;
            for (int c = offs[q]; c < offs[q + 1]; c++) {
                const size_t idx = (size_t)cand[c];
                const cv::KeyPoint &kp = pKF->mvKeysUn[idx];
                const int kpLevel = kp.octave;
                if (kpLevel < nPredictedLevel - 1 || kpLevel > nPredictedLevel) continue;
                if (pKF->mvuRight[idx] >= 0) {
                    const float ex = u - kp.pt.x, ey = v - kp.pt.y, er = ur - pKF->mvuRight[idx];
                    const float e2 = ex * ex + ey * ey + er * er;
                    if (e2 * pKF->mvInvLevelSigma2[kpLevel] > 7.8) continue;
                } else {
                    const float ex = u - kp.pt.x, ey = v - kp.pt.y;
                    const float e2 = ex * ex + ey * ey;
                    if (e2 * pKF->mvInvLevelSigma2[kpLevel] > 5.99) continue;
                }
                const int d = dist[c];
                if (d < bestDist) { bestDist = d; bestIdx = (int)idx; }
            }
            if (bestDist <= TH_LOW) {
                MapPointT *pMPinKF = pKF->GetMapPoint(bestIdx);
                if (pMPinKF) {
                    if (!pMPinKF->isBad()) {
                        if (pMPinKF->Observations() > pMP->Observations()) pMP->Replace(pMPinKF);
                        else pMPinKF->Replace(pMP);
                    }
                } else {
                    pMP->AddObservation(pKF, bestIdx);
                    pKF->AddMapPoint(pMP, bestIdx);
                }
                nFused++;
            }
        }
        return nFused;
    }

    // Relocalisation: project the map points of a candidate keyframe into the current frame
    // (src/ORBmatcher.cc:1555-1685; callers src/Tracking.cc:1763, 1777).
    template <class FrameT, class KeyFrameT, class SetT>
    int SearchByProjection(FrameT &CurrentFrame, KeyFrameT *pKF, const SetT &sAlreadyFound, const float th, const int ORBdist)
    {
        typedef typename std::remove_pointer<typename std::remove_reference<decltype(CurrentFrame.mvpMapPoints[0])>::type>::type MapPointT;
        int nmatches = 0;
        float Rcw[9], tcw[3], Ow[3];
        pose_Rt(CurrentFrame.mTcw, Rcw, tcw);
        camera_centre(Rcw, tcw, Ow);
        std::vector<int> rotHist[HISTO_LENGTH];
        const float factor = 1.0f / HISTO_LENGTH;
        const std::vector<MapPointT *> vpMPs = pKF->GetMapPointMatches();
        // prologue (:1572-1612)
        std::vector<int> qI;
        std::vector<float> win;
        std::vector<unsigned char> qDesc;
        for (size_t i = 0, iend = vpMPs.size(); i < iend; i++) {
            MapPointT *pMP = vpMPs[i];
            if (!pMP) continue;
            if (pMP->isBad() || sAlreadyFound.count(pMP)) continue;
            const cv::Mat x3Dw = pMP->GetWorldPos();
            const float xw[3] = {x3Dw.template at<float>(0, 0), x3Dw.template at<float>(1, 0), x3Dw.template at<float>(2, 0)};
            float x3Dc[3];
            mat_Rx_t(Rcw, xw, tcw, x3Dc);
            const float xc = x3Dc[0], yc = x3Dc[1];
            const float invzc = 1.0 / x3Dc[2];
            const float u = CurrentFrame.fx * xc * invzc + CurrentFrame.cx;
            const float v = CurrentFrame.fy * yc * invzc + CurrentFrame.cy;
            if (u < CurrentFrame.mnMinX || u > CurrentFrame.mnMaxX) continue;
            if (v < CurrentFrame.mnMinY || v > CurrentFrame.mnMaxY) continue;
            const float PO[3] = {xw[0] - Ow[0], xw[1] - Ow[1], xw[2] - Ow[2]};
            const float dist3D = norm3(PO);
            const float maxDistance = pMP->GetMaxDistanceInvariance(), minDistance = pMP->GetMinDistanceInvariance();
            if (dist3D < minDistance || dist3D > maxDistance) continue;
            const int nPredictedLevel = pMP->PredictScale(dist3D, &CurrentFrame);
            const float radius = th * CurrentFrame.mvScaleFactors[nPredictedLevel];
            qI.push_back((int)i);
            const float wq[5] = {u, v, radius, (float)(nPredictedLevel - 1), (float)(nPredictedLevel + 1)};
            win.insert(win.end(), wq, wq + 5);
            const cv::Mat d = pMP->GetDescriptor();
            qDesc.insert(qDesc.end(), d.ptr(), d.ptr() + 32);
        }
        std::vector<int> offs, cand;
        std::vector<uint16_t> dist;
        if (!window_search_frame(CurrentFrame, win, qDesc, offs, cand, dist)) return 0;
        // epilogue (:1614-1655)
        for (size_t q = 0; q < qI.size(); q++) {
            if (offs[q] == offs[q + 1]) continue;
            const int i = qI[q];
            int bestDist = 256, bestIdx2 = -1;
            for (int c = offs[q]; c < offs[q + 1]; c++) {
                const size_t i2 = (size_t)cand[c];
                if (CurrentFrame.mvpMapPoints[i2]) continue;
                const int d = dist[c];
                if (d < bestDist) { bestDist = d; bestIdx2 = (int)i2; }
            }
            if (bestDist <= ORBdist) {
                CurrentFrame.mvpMapPoints[bestIdx2] = vpMPs[i];
                nmatches++;
                if (mbCheckOrientation) {
                    float rot = pKF->mvKeysUn[i].angle - CurrentFrame.mvKeysUn[bestIdx2].angle;
                    if (rot < 0.0) rot += 360.0f;
                    int bin = (int)std::round(rot * factor);
                    if (bin == HISTO_LENGTH) bin = 0;
                    rotHist[bin].push_back(bestIdx2);
                }
            }
        }
        if (mbCheckOrientation) {
            int ind1 = -1, ind2 = -1, ind3 = -1;
            ComputeThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
            for (int i = 0; i < HISTO_LENGTH; i++) {
                if (i != ind1 && i != ind2 && i != ind3) {
                    for (size_t j = 0, jend = rotHist[i].size(); j < jend; j++) {
                        CurrentFrame.mvpMapPoints[rotHist[i][j]] = static_cast<MapPointT *>(NULL);
                        nmatches--;
                    }
                }
            }
        }
        return nmatches;
    }

    // Loop closing: project loop map points with a Sim3 (src/ORBmatcher.cc:305-422; caller src/LoopClosing.cc:464).
    template <class KeyFrameT, class MapPointT>
    int SearchByProjection(KeyFrameT *pKF, cv::Mat Scw, const std::vector<MapPointT *> &vpPoints, std::vector<MapPointT *> &vpMatched, int th)
    {
        float Rcw[9], tcw[3], Ow[3];
        sim3_decompose(Scw, Rcw, tcw, Ow);
        std::set<MapPointT *> spAlreadyFound(vpMatched.begin(), vpMatched.end());
        spAlreadyFound.erase(static_cast<MapPointT *>(NULL));
        std::vector<int> qI, qLevel;
        std::vector<float> qU, qV, win;
        std::vector<unsigned char> qDesc;
        for (int iMP = 0, iendMP = (int)vpPoints.size(); iMP < iendMP; iMP++) {
            MapPointT *pMP = vpPoints[iMP];
            if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;
            float u, v, dist3D;
            if (!project_gated(pKF, Rcw, tcw, Ow, pMP, u, v, dist3D)) continue;
            const int nPredictedLevel = pMP->PredictScale(dist3D, pKF);
            const float radius = th * pKF->mvScaleFactors[nPredictedLevel];
            push_query(pMP, iMP, nPredictedLevel, u, v, radius, qI, qLevel, qU, qV, win, qDesc);
        }
        std::vector<int> offs, cand;
        std::vector<uint16_t> dist;
        if (!window_search_keyframe(pKF, win, qDesc, offs, cand, dist)) return 0;
        int nmatches = 0;
        for (size_t q = 0; q < qI.size(); q++) {
            int bestDist = 256, bestIdx = -1;
            for (int c = offs[q]; c < offs[q + 1]; c++) {
                const size_t idx = (size_t)cand[c];
                if (vpMatched[idx]) continue;
                const int kpLevel = pKF->mvKeysUn[idx].octave;
                if (kpLevel < qLevel[q] - 1 || kpLevel > qLevel[q]) continue;
                const int d = dist[c];
                if (d < bestDist) { bestDist = d; bestIdx = (int)idx; }
            }
            if (bestDist <= TH_LOW) {
                vpMatched[bestIdx] = vpPoints[qI[q]];
                nmatches++;
            }
        }
        return nmatches;
    }

    // Loop closing: fuse corrected loop map points into a keyframe (src/ORBmatcher.cc:1040-1164; caller
    // src/LoopClosing.cc:745).
    template <class KeyFrameT, class MapPointT>
    int Fuse(KeyFrameT *pKF, cv::Mat Scw, const std::vector<MapPointT *> &vpPoints, float th, std::vector<MapPointT *> &vpReplacePoint)
    {
        float Rcw[9], tcw[3], Ow[3];
        sim3_decompose(Scw, Rcw, tcw, Ow);
        const std::set<MapPointT *> spAlreadyFound = pKF->GetMapPoints();
        std::vector<int> qI, qLevel;
        std::vector<float> qU, qV, win;
        std::vector<unsigned char> qDesc;
        const int nPoints = (int)vpPoints.size();
        for (int iMP = 0; iMP < nPoints; iMP++) {
            MapPointT *pMP = vpPoints[iMP];
            if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;
            float u, v, dist3D;
            if (!project_gated(pKF, Rcw, tcw, Ow, pMP, u, v, dist3D)) continue;
            const int nPredictedLevel = pMP->PredictScale(dist3D, pKF);
            const float radius = th * pKF->mvScaleFactors[nPredictedLevel];
            push_query(pMP, iMP, nPredictedLevel, u, v, radius, qI, qLevel, qU, qV, win, qDesc);
        }
        std::vector<int> offs, cand;
        std::vector<uint16_t> dist;
        if (!window_search_keyframe(pKF, win, qDesc, offs, cand, dist)) return 0;
        int nFused = 0;
        for (size_t q = 0; q < qI.size(); q++) {
            // spAlreadyFound is a snapshot and isBad() is not changed inside this loop: no re-check needed
            MapPointT *pMP = vpPoints[qI[q]];
            int bestDist = INT_MAX_, bestIdx = -1;
            for (int c = offs[q]; c < offs[q + 1]; c++) {
                const size_t idx = (size_t)cand[c];
                const int kpLevel = pKF->mvKeysUn[idx].octave;
                if (kpLevel < qLevel[q] - 1 || kpLevel > qLevel[q]) continue;
                const int d = dist[c];
                if (d < bestDist) { bestDist = d; bestIdx = (int)idx; }
            }
            if (bestDist <= TH_LOW) {
                MapPointT *pMPinKF = pKF->GetMapPoint(bestIdx);
                if (pMPinKF) {
                    if (!pMPinKF->isBad()) vpReplacePoint[qI[q]] = pMPinKF;
                } else {
                    pMP->AddObservation(pKF, bestIdx);
                    pKF->AddMapPoint(pMP, bestIdx);
                }
                nFused++;
            }
        }
        return nFused;
    }

    // Loop closing: guided matching both ways under a Sim3 (src/ORBmatcher.cc:1166-1390; caller src/LoopClosing.cc:391).
    template <class KeyFrameT, class MapPointT>
    int SearchBySim3(KeyFrameT *pKF1, KeyFrameT *pKF2, std::vector<MapPointT *> &vpMatches12, const float &s12, const cv::Mat &R12,
                     const cv::Mat &t12, const float th)
    {
        float R1w[9], t1w[3], R2w[9], t2w[3], sR12[9], sR21[9], t12f[3], t21[3];
        {
            const cv::Mat a = pKF1->GetRotation(), b = pKF1->GetTranslation(), c = pKF2->GetRotation(), d = pKF2->GetTranslation();
            const float inv = (float)(1.0 / s12); // MatExpr scale factors reach convertTo as floats for 32F data
            for (int r = 0; r < 3; r++) {
                for (int k = 0; k < 3; k++) {
                    R1w[3 * r + k] = a.template at<float>(r, k);
                    R2w[3 * r + k] = c.template at<float>(r, k);
                    sR12[3 * r + k] = R12.template at<float>(r, k) * s12;
                    sR21[3 * r + k] = R12.template at<float>(k, r) * inv;
                }
                t1w[r] = b.template at<float>(r, 0);
                t2w[r] = d.template at<float>(r, 0);
                t12f[r] = t12.template at<float>(r, 0);
            }
            mat_neg_Rx(sR21, t12f, t21); // t21 = -sR21*t12
        }
        const std::vector<MapPointT *> vpMapPoints1 = pKF1->GetMapPointMatches(), vpMapPoints2 = pKF2->GetMapPointMatches();
        const int N1 = (int)vpMapPoints1.size(), N2 = (int)vpMapPoints2.size();
        std::vector<bool> vbAlreadyMatched1(N1, false), vbAlreadyMatched2(N2, false);
        for (int i = 0; i < N1; i++) {
            MapPointT *pMP = vpMatches12[i];
            if (pMP) {
                vbAlreadyMatched1[i] = true;
                const int idx2 = pMP->GetIndexInKeyFrame(pKF2);
                if (idx2 >= 0 && idx2 < N2) vbAlreadyMatched2[idx2] = true;
            }
        }
        std::vector<int> vnMatch1(N1, -1), vnMatch2(N2, -1);
        // the two directions are independent of each other: same helper, roles swapped
        const float fx = pKF1->fx, fy = pKF1->fy, cx = pKF1->cx, cy = pKF1->cy;
        if (!sim3_direction(pKF2, vpMapPoints1, vbAlreadyMatched1, R1w, t1w, sR21, t21, fx, fy, cx, cy, th, vnMatch1)) return 0;
        if (!sim3_direction(pKF1, vpMapPoints2, vbAlreadyMatched2, R2w, t2w, sR12, t12f, fx, fy, cx, cy, th, vnMatch2)) return 0;
        int nFound = 0;
        for (int i1 = 0; i1 < N1; i1++) {
            const int idx2 = vnMatch1[i1];
            if (idx2 >= 0) {
                const int idx1 = vnMatch2[idx2];
                if (idx1 == i1) {
                    vpMatches12[i1] = vpMapPoints2[idx2];
                    nFound++;
                }
            }
        }
        return nFound;
    }

    // Loop-closure matching between two keyframes (src/ORBmatcher.cc:568-702; caller src/LoopClosing.cc:323).
    template <class KeyFrameT, class MapPointT>
    int SearchByBoW(KeyFrameT *pKF1, KeyFrameT *pKF2, std::vector<MapPointT *> &vpMatches12)
    {
        const std::vector<cv::KeyPoint> &vKeysUn1 = pKF1->mvKeysUn, &vKeysUn2 = pKF2->mvKeysUn;
        const std::vector<MapPointT *> vpMapPoints1 = pKF1->GetMapPointMatches(), vpMapPoints2 = pKF2->GetMapPointMatches();
        vpMatches12 = std::vector<MapPointT *>(vpMapPoints1.size(), static_cast<MapPointT *>(NULL));
        std::vector<bool> vbMatched2(vpMapPoints2.size(), false);
        std::vector<int> rotHist[HISTO_LENGTH];
        const float factor = 1.0f / HISTO_LENGTH;
        int nmatches = 0;
        // prologue: merge walk (:590-679); candidates = features of the same node holding a good map point
        std::vector<int> qIdx1, offs(1, 0), cand;
        std::vector<unsigned char> qDesc;
        typename std::remove_reference<decltype(pKF1->mFeatVec)>::type::const_iterator f1it = pKF1->mFeatVec.begin(), f1end = pKF1->mFeatVec.end();
        typename std::remove_reference<decltype(pKF2->mFeatVec)>::type::const_iterator f2it = pKF2->mFeatVec.begin(), f2end = pKF2->mFeatVec.end();
        while (f1it != f1end && f2it != f2end) {
            if (f1it->first == f2it->first) {
                for (size_t i1 = 0, iend1 = f1it->second.size(); i1 < iend1; i1++) {
                    const size_t idx1 = f1it->second[i1];
                    MapPointT *pMP1 = vpMapPoints1[idx1];
                    if (!pMP1) continue;
                    if (pMP1->isBad()) continue;
                    qIdx1.push_back((int)idx1);
                    for (size_t i2 = 0, iend2 = f2it->second.size(); i2 < iend2; i2++) {
                        const size_t idx2 = f2it->second[i2];
                        MapPointT *pMP2 = vpMapPoints2[idx2];
                        if (!pMP2) continue;
                        if (pMP2->isBad()) continue;
                        cand.push_back((int)idx2);
                    }
                    offs.push_back((int)cand.size());
                    const unsigned char *d = pKF1->mDescriptors.ptr((int)idx1);
                    qDesc.insert(qDesc.end(), d, d + 32);
                }
                f1it++;
                f2it++;
            } else if (f1it->first < f2it->first) {
                f1it = pKF1->mFeatVec.lower_bound(f2it->first);
            } else {
                f2it = pKF2->mFeatVec.lower_bound(f1it->first);
            }
        }
        std::vector<uint16_t> dist;
        if (!distances(qDesc, (int)qIdx1.size(), pKF2->mDescriptors, offs, cand, dist)) return 0;
        // epilogue (:606-668): strict `<` on TH_LOW here, unlike the Frame overload
        for (size_t q = 0; q < qIdx1.size(); q++) {
            const int idx1 = qIdx1[q];
            int bestDist1 = 256, bestIdx2 = -1, bestDist2 = 256;
            for (int c = offs[q]; c < offs[q + 1]; c++) {
                const int idx2 = cand[c];
                if (vbMatched2[idx2]) continue;
                const int d = dist[c];
                if (d < bestDist1) { bestDist2 = bestDist1; bestDist1 = d; bestIdx2 = idx2; }
                else if (d < bestDist2) bestDist2 = d;
            }
            if (bestDist1 < TH_LOW) {
                if (static_cast<float>(bestDist1) < mfNNratio * static_cast<float>(bestDist2)) {
                    vpMatches12[idx1] = vpMapPoints2[bestIdx2];
                    vbMatched2[bestIdx2] = true;
                    if (mbCheckOrientation) {
                        float rot = vKeysUn1[idx1].angle - vKeysUn2[bestIdx2].angle;
                        if (rot < 0.0) rot += 360.0f;
                        int bin = (int)std::round(rot * factor);
                        if (bin == HISTO_LENGTH) bin = 0;
                        rotHist[bin].push_back(idx1);
                    }
                    nmatches++;
                }
            }
        }
        if (mbCheckOrientation) {
            int ind1 = -1, ind2 = -1, ind3 = -1;
            ComputeThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
            for (int i = 0; i < HISTO_LENGTH; i++) {
                if (i == ind1 || i == ind2 || i == ind3) continue;
                for (size_t j = 0, jend = rotHist[i].size(); j < jend; j++) {
                    vpMatches12[rotHist[i][j]] = static_cast<MapPointT *>(NULL);
                    nmatches--;
                }
            }
        }
        return nmatches;
    }

    void SetDevice(int device) { device_ = device; } /* extension: GPU used by the searches (default: current) */

public:
    static const int TH_LOW = 50;
    static const int TH_HIGH = 100;
    static const int HISTO_LENGTH = 30;
    static const int INT_MAX_ = 2147483647;

protected:
    // one device call for all (query, candidate) pairs of a search
    bool distances(const std::vector<unsigned char> &qDesc, int nq, const cv::Mat &targets, const std::vector<int> &offs,
                   const std::vector<int> &cand, std::vector<uint16_t> &dist)
    {
        dist.assign(cand.size() ? cand.size() : 1, 0);
        if (nq == 0 || cand.empty()) return true;
        const unsigned char *t = targets.ptr();
        std::vector<unsigned char> packed;
        if (!targets.isContinuous()) { // cv::Mat ROI: pack rows
            packed.resize((size_t)targets.rows * 32);
            for (int r = 0; r < targets.rows; r++) std::memcpy(&packed[(size_t)r * 32], targets.ptr(r), 32);
            t = packed.data();
        }
        if (orbb200_candidate_distances(qDesc.data(), nq, t, targets.rows, offs.data(), cand.data(), dist.data(), device_) != ORBB200_OK) {
            std::cerr << "ORBmatcher (orbb200): " << orbb200_last_error() << std::endl; // no CPU fallback: 0 matches
            return false;
        }
        return true;
    }

    // R (row-major 3x3) and t of a 4x4 CV_32F pose, like Tcw.rowRange(0,3).colRange(0,3) / .col(3)
    static void pose_Rt(const cv::Mat &T, float *R, float *t)
    {
        for (int r = 0; r < 3; r++) {
            for (int c = 0; c < 3; c++) R[3 * r + c] = T.at<float>(r, c);
            t[r] = T.at<float>(r, 3);
        }
    }
    // y = R*x + t as OpenCV evaluates `R*x + t` for CV_32F operands: the expression folds into ONE cv::gemm(R, x, 1, t, 1), and
    // for an untransposed product with inner length 2..4 gemm takes its small-matrix branch -- the three products accumulate
    // left to right in FLOAT, then (float)(sum*alpha + t*beta) in double (measured against cv2.gemm: tests/test_ref_pin.py).
    ORBB200_NO_FMA static void mat_Rx_t(const float *R, const float *x, const float *t, float *y)
    {
        for (int r = 0; r < 3; r++) {
            float s = R[3 * r] * x[0];
            s = s + R[3 * r + 1] * x[1];
            s = s + R[3 * r + 2] * x[2];
            y[r] = (float)((double)s + (double)t[r]);
        }
    }
    // y = -(R*x): cv::gemm(R, x, -1, noArray(), 0) in the same small-matrix branch (float accumulation, exact negation)
    ORBB200_NO_FMA static void mat_neg_Rx(const float *R, const float *x, float *y)
    {
        for (int r = 0; r < 3; r++) {
            float s = R[3 * r] * x[0];
            s = s + R[3 * r + 1] * x[1];
            s = s + R[3 * r + 2] * x[2];
            y[r] = -s;
        }
    }

    // Ow = -R^T * t: a TRANSPOSED operand sends cv::gemm down its general branch (products accumulate in double in k order;
    // alpha = -1 negates the double sum, one narrowing)
    static void camera_centre(const float *R, const float *t, float *Ow)
    {
        for (int i = 0; i < 3; i++) Ow[i] = -(float)((double)R[i] * t[0] + (double)R[3 + i] * t[1] + (double)R[6 + i] * t[2]);
    }
    // cv::norm of a 3-vector of floats: squares summed in double, sqrt in double, narrowed by the caller's float
    static float norm3(const float *p)
    {
        return (float)std::sqrt((double)p[0] * p[0] + (double)p[1] * p[1] + (double)p[2] * p[2]);
    }
    // "Decompose Scw" (src/ORBmatcher.cc:313-318, 1049-1054): scw = |row 0| (Mat::dot accumulates in double),
    // Rcw = sRcw / scw and tcw = t / scw (a MatExpr scale: each element times the FLOAT 1/scw), Ow = -Rcw^T tcw.
    static void sim3_decompose(const cv::Mat &Scw, float *Rcw, float *tcw, float *Ow)
    {
        float sR[9], st[3];
        pose_Rt(Scw, sR, st);
        const float scw = (float)std::sqrt((double)sR[0] * sR[0] + (double)sR[1] * sR[1] + (double)sR[2] * sR[2]);
        const float inv = (float)(1.0 / scw);
        for (int i = 0; i < 9; i++) Rcw[i] = sR[i] * inv;
        for (int i = 0; i < 3; i++) tcw[i] = st[i] * inv;
        camera_centre(Rcw, tcw, Ow);
    }
    // The common gate of the Sim3 searches (src/ORBmatcher.cc:336-371, 1073-1108): positive depth, inside the image,
    // distance inside the scale-invariance band, viewing angle below 60 degrees.
    template <class KeyFrameT, class MapPointT>
    static bool project_gated(KeyFrameT *pKF, const float *Rcw, const float *tcw, const float *Ow, MapPointT *pMP, float &u, float &v,
                              float &dist3D)
    {
        const cv::Mat p3Dw = pMP->GetWorldPos();
        const float pw[3] = {p3Dw.template at<float>(0, 0), p3Dw.template at<float>(1, 0), p3Dw.template at<float>(2, 0)};
        float p3Dc[3];
        mat_Rx_t(Rcw, pw, tcw, p3Dc);
        if (p3Dc[2] < 0.0f) return false;
        const float invz = 1 / p3Dc[2];
        const float x = p3Dc[0] * invz, y = p3Dc[1] * invz;
        u = pKF->fx * x + pKF->cx;
        v = pKF->fy * y + pKF->cy;
        if (!pKF->IsInImage(u, v)) return false;
        const float maxDistance = pMP->GetMaxDistanceInvariance(), minDistance = pMP->GetMinDistanceInvariance();
        const float PO[3] = {pw[0] - Ow[0], pw[1] - Ow[1], pw[2] - Ow[2]};
        dist3D = norm3(PO);
        if (dist3D < minDistance || dist3D > maxDistance) return false;
        const cv::Mat Pn = pMP->GetNormal();
        const double dot = (double)PO[0] * Pn.template at<float>(0, 0) + (double)PO[1] * Pn.template at<float>(1, 0) +
                           (double)PO[2] * Pn.template at<float>(2, 0);
        if (dot < 0.5 * dist3D) return false;
        return true;
    }
    template <class MapPointT>
    static void push_query(MapPointT *pMP, int i, int level, float u, float v, float radius, std::vector<int> &qI, std::vector<int> &qLevel,
                           std::vector<float> &qU, std::vector<float> &qV, std::vector<float> &win, std::vector<unsigned char> &qDesc)
    {
        qI.push_back(i); qLevel.push_back(level); qU.push_back(u); qV.push_back(v);
        const float wq[5] = {u, v, radius, -1.f, -1.f}; // KeyFrame::GetFeaturesInArea has no level filter (KeyFrame.cc:607)
        win.insert(win.end(), wq, wq + 5);
        const cv::Mat d = pMP->GetDescriptor();
        qDesc.insert(qDesc.end(), d.ptr(), d.ptr() + 32);
    }
    // KeyFrame::GetFeaturesInArea (src/KeyFrame.cc:607-647) for every window + distances; the grid is the keyframe's own
    template <class KeyFrameT>
    bool window_search_keyframe(KeyFrameT *pKF, const std::vector<float> &windows, const std::vector<unsigned char> &qDesc,
                                std::vector<int> &offs, std::vector<int> &cand, std::vector<uint16_t> &dist)
    {
        return window_search(pKF->mvKeysUn, pKF->mDescriptors, pKF->mnMinX, pKF->mnMinY, pKF->mfGridElementWidthInv,
                             pKF->mfGridElementHeightInv, pKF->mnGridCols, pKF->mnGridRows, windows, qDesc, offs, cand, dist);
    }
    // One direction of SearchBySim3 (src/ORBmatcher.cc:1212-1283 and :1286-1357): map points of pKFa, moved into the
    // camera of pKFb by (sRba, tba) after (Raw, taw), matched inside pKFb's grid.  fx..cy are pKF1's in both directions.
    template <class KeyFrameT, class MapPointT>
    bool sim3_direction(KeyFrameT *pKFb, const std::vector<MapPointT *> &vpMapPointsA, const std::vector<bool> &vbAlreadyMatchedA,
                        const float *Raw, const float *taw, const float *sRba, const float *tba, float fx, float fy, float cx, float cy,
                        float th, std::vector<int> &vnMatchA)
    {
        std::vector<int> qI, qLevel;
        std::vector<float> qU, qV, win;
        std::vector<unsigned char> qDesc;
        const int NA = (int)vpMapPointsA.size();
        for (int i = 0; i < NA; i++) {
            MapPointT *pMP = vpMapPointsA[i];
            if (!pMP || vbAlreadyMatchedA[i]) continue;
            if (pMP->isBad()) continue;
            const cv::Mat p3Dw = pMP->GetWorldPos();
            const float pw[3] = {p3Dw.template at<float>(0, 0), p3Dw.template at<float>(1, 0), p3Dw.template at<float>(2, 0)};
            float pa[3], pb[3];
            mat_Rx_t(Raw, pw, taw, pa);
            mat_Rx_t(sRba, pa, tba, pb);
            if (pb[2] < 0.0) continue;
            const float invz = 1.0 / pb[2];
            const float x = pb[0] * invz, y = pb[1] * invz;
            const float u = fx * x + cx, v = fy * y + cy;
            if (!pKFb->IsInImage(u, v)) continue;
            const float maxDistance = pMP->GetMaxDistanceInvariance(), minDistance = pMP->GetMinDistanceInvariance();
            const float dist3D = norm3(pb);
            if (dist3D < minDistance || dist3D > maxDistance) continue;
            const int nPredictedLevel = pMP->PredictScale(dist3D, pKFb);
            const float radius = th * pKFb->mvScaleFactors[nPredictedLevel];
            push_query(pMP, i, nPredictedLevel, u, v, radius, qI, qLevel, qU, qV, win, qDesc);
        }
        std::vector<int> offs, cand;
        std::vector<uint16_t> dist;
        if (!window_search_keyframe(pKFb, win, qDesc, offs, cand, dist)) return false;
        for (size_t q = 0; q < qI.size(); q++) {
            int bestDist = INT_MAX_, bestIdx = -1;
            for (int c = offs[q]; c < offs[q + 1]; c++) {
                const size_t idx = (size_t)cand[c];
                const int octave = pKFb->mvKeysUn[idx].octave;
                if (octave < qLevel[q] - 1 || octave > qLevel[q]) continue;
                const int d = dist[c];
                if (d < bestDist) { bestDist = d; bestIdx = (int)idx; }
            }
            if (bestDist <= TH_HIGH) vnMatchA[qI[q]] = bestIdx;
        }
        return true;
    }

    // Frame::GetFeaturesInArea for every window of a search + all candidate distances, in ONE device call
    // (orbb200_window_search: the frame's 64x48 grid is rebuilt on the device from mvKeysUn, so the caller's
    // Frame::mGrid is not read).  windows: 5 floats per query = x, y, r, minLevel, maxLevel.
    template <class FrameT>
    bool window_search_frame(const FrameT &F, const std::vector<float> &windows, const std::vector<unsigned char> &qDesc,
                             std::vector<int> &offs, std::vector<int> &cand, std::vector<uint16_t> &dist)
    {
#ifdef FRAME_GRID_COLS
        const int cols = FRAME_GRID_COLS, rows = FRAME_GRID_ROWS;
#else
        const int cols = 64, rows = 48;
#endif
        // a frame that keeps its extraction outputs on the device (member `orbb200::DeviceFrame *mpDeviceFrame`): nothing of the
        // frame is uploaded, only the windows and the query descriptors travel
        orbb200::DeviceFrame *df = device_frame_of(F, 0);
        if (df && df->N() == (int)F.mvKeysUn.size()) {
            df->SetGrid(F.mnMinX, F.mnMinY, F.mfGridElementWidthInv, F.mfGridElementHeightInv, cols, rows);
            return df->WindowSearch(windows, qDesc, offs, cand, dist);
        }
        return window_search(F.mvKeysUn, F.mDescriptors, F.mnMinX, F.mnMinY, F.mfGridElementWidthInv, F.mfGridElementHeightInv,
                             cols, rows, windows, qDesc, offs, cand, dist);
    }
    template <class T> static auto device_frame_of(const T &F, int) -> decltype(F.mpDeviceFrame) { return F.mpDeviceFrame; }
    template <class T> static orbb200::DeviceFrame *device_frame_of(const T &, long) { return nullptr; }

    bool window_search(const std::vector<cv::KeyPoint> &tkps, const cv::Mat &targets, float minX, float minY, float invW, float invH,
                       int cols, int rows, const std::vector<float> &windows, const std::vector<unsigned char> &qDesc,
                       std::vector<int> &offs, std::vector<int> &cand, std::vector<uint16_t> &dist)
    {
        const int nq = (int)(windows.size() / 5);
        offs.assign(nq + 1, 0);
        cand.assign(1, 0);
        dist.assign(1, 0);
        if (nq == 0 || tkps.empty()) return true;
        const unsigned char *t = targets.ptr();
        std::vector<unsigned char> packed;
        if (!targets.isContinuous()) {
            packed.resize((size_t)targets.rows * 32);
            for (int r = 0; r < targets.rows; r++) std::memcpy(&packed[(size_t)r * 32], targets.ptr(r), 32);
            t = packed.data();
        }
        int cap = nq * 48 + 64;
        for (int attempt = 0; attempt < 2; attempt++) {
            cand.assign(cap, 0);
            dist.assign(cap, 0);
            const int rc = orbb200_window_search(reinterpret_cast<const orbb200_keypoint *>(tkps.data()), t, (int)tkps.size(), minX, minY,
                                                 invW, invH, cols, rows, windows.data(), qDesc.data(), nq, offs.data(), cand.data(),
                                                 dist.data(), cap, device_);
            if (rc == ORBB200_OK) return true;
            if (rc == ORBB200_ERR_CAPACITY && offs[nq] > cap) { cap = offs[nq]; continue; } // offsets are valid: exact size
            break;
        }
        std::cerr << "ORBmatcher (orbb200): " << orbb200_last_error() << std::endl; // no CPU fallback: 0 matches
        offs.assign(nq + 1, 0);
        return false;
    }

    // src/ORBmatcher.cc:147-164
    template <class KeyFrameT>
    bool CheckDistEpipolarLine(const cv::KeyPoint &kp1, const cv::KeyPoint &kp2, const cv::Mat &F12, const KeyFrameT *pKF2)
    {
        const float a = kp1.pt.x * F12.at<float>(0, 0) + kp1.pt.y * F12.at<float>(1, 0) + F12.at<float>(2, 0);
        const float b = kp1.pt.x * F12.at<float>(0, 1) + kp1.pt.y * F12.at<float>(1, 1) + F12.at<float>(2, 1);
        const float c = kp1.pt.x * F12.at<float>(0, 2) + kp1.pt.y * F12.at<float>(1, 2) + F12.at<float>(2, 2);
        const float num = a * kp2.pt.x + b * kp2.pt.y + c;
        const float den = a * a + b * b;
        if (den == 0) return false;
        const float dsqr = num * num / den;
        return dsqr < 3.84 * pKF2->mvLevelSigma2[kp2.octave];
    }

    // src/ORBmatcher.cc:138-144
    float RadiusByViewingCos(const float &viewCos)
    {
        if (viewCos > 0.998) return 2.5;
        else return 4.0;
    }

    // src/ORBmatcher.cc:1687-1728
    void ComputeThreeMaxima(std::vector<int> *histo, const int L, int &ind1, int &ind2, int &ind3)
    {
        int max1 = 0, max2 = 0, max3 = 0;
        for (int i = 0; i < L; i++) {
            const int s = (int)histo[i].size();
            if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
            else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
            else if (s > max3) { max3 = s; ind3 = i; }
        }
        if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
        else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
    }

    float mfNNratio;
    bool mbCheckOrientation;
    int device_;
};

} // namespace ORBB200_NAMESPACE (ORB_SLAM2 by default)

#endif /* ORBB200_ORBMATCHER_H */

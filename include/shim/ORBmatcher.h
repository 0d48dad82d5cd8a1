/* Drop-in replacement of the reference's include/ORBmatcher.h (worxli/ORB_SLAM include/ORBmatcher.h:37-107) for the
 * searches that take whole Frame / KeyFrame objects: class ORB_SLAM::ORBmatcher with the reference's constructor,
 * constants and method signatures, implemented on the C ABI of liborbx.so (include/orbm.h) - the B200 kernels.
 * Header-only; compiled inside the reference's tree (needs its Frame.h / KeyFrame.h / MapPoint.h).  Put this directory
 * before the reference's include/ on the include path and drop src/ORBmatcher.cc from the build; the call sites
 *   src/Tracking.cc:361-362  SearchForInitialization(mInitialFrame, mCurrentFrame, mvbPrevMatched, mvIniMatches, 100)
 *   src/Tracking.cc:510,515  WindowSearch(mLastFrame, mCurrentFrame, 200 | 100, vpMapPointMatches, minOctave)
 *   src/Tracking.cc:544,547  SearchByProjection(mLastFrame, mCurrentFrame, 15 | 50, vpMapPointMatches)
 *   src/Tracking.cc:893,913  SearchByBoW(pKF, mCurrentFrame, vpMapPointMatches)
 *   src/LoopClosing.cc:266   SearchByBoW(mpCurrentKF, pKF, vpMapPointMatches)
 * compile unchanged (oracle/Makefile target ref_dropin, tests/test_gpu_dropin.py).  Each call flattens the objects
 * into the views of orbm.h (positions, octaves, angles, descriptors, "has a good MapPoint" flags, feature vectors),
 * runs the search on the device and writes the MapPoint pointers back, which is all the pointer bookkeeping these
 * five routines do.  The projection family taking MapPoint sets (local map, last frame, relocalisation, Sim3, Fuse,
 * SearchBySim3, SearchForTriangulation) is served through the flat views of orbm.h (INTEGRATION.md shows the caller
 * side): their 3-D projection code stays with the host. */
#ifndef ORBMATCHER_H
#define ORBMATCHER_H

#include <climits>
#include <cstring>
#include <set>
#include <stdexcept>
#include <string>
#include <vector>

#include <opencv2/core/core.hpp>
#include <opencv2/features2d/features2d.hpp>

#include "MapPoint.h"
#include "KeyFrame.h"
#include "Frame.h"

#include "../orbm.h"

namespace ORB_SLAM
{

class ORBmatcher
{
public:

    ORBmatcher(float nnratio = 0.6, bool checkOri = true) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}

    // Computes the Hamming distance between two ORB descriptors (src/ORBmatcher.cc:1794-1810)
    static int DescriptorDistance(const cv::Mat &a, const cv::Mat &b)
    {
        int32_t d = 0;
        check(orbm_descriptor_distance(handle(), a.ptr<unsigned char>(), b.ptr<unsigned char>(), 1, &d));
        return d;
    }

    // Matching for the Map Initialization (src/ORBmatcher.cc:598-713)
    int SearchForInitialization(Frame &F1, Frame &F2, std::vector<cv::Point2f> &vbPrevMatched,
                                std::vector<int> &vnMatches12, int windowSize = 10)
    {
        Flat a(F1, flags_none(F1)), b(F2, flags_none(F2));
        Set set(a, b);
        const int stride = a.n > b.n ? a.n : b.n;
        std::vector<float> prev((size_t)stride * 2, 0.f);
        for (int i = 0; i < a.n; i++) { prev[2 * i] = vbPrevMatched[i].x; prev[2 * i + 1] = vbPrevMatched[i].y; }
        std::vector<int32_t> m12(stride > 0 ? stride : 1);
        int32_t n = 0;
        const int32_t i1 = 0, i2 = 1;
        check(orbm_search_for_initialization(handle(), set.s, set.s, &i1, &i2, 1, prev.data(), stride > 0 ? stride : 1,
                                             windowSize, mfNNratio, mbCheckOrientation, m12.data(), &n));
        vnMatches12.assign(m12.begin(), m12.begin() + a.n);                                       // :604
        for (int i = 0; i < a.n; i++) vbPrevMatched[i] = cv::Point2f(prev[2 * i], prev[2 * i + 1]);   // :708-710
        return n;
    }

    // Search MapPoints tracked in Frame1 in Frame2 in a window centered at their position in Frame1 (:409-516)
    int WindowSearch(Frame &F1, Frame &F2, int windowSize, std::vector<MapPoint *> &vpMapPointMatches2,
                     int minOctave = -1, int maxOctave = INT_MAX)
    {
        Flat a(F1, flags_good(F1.mvpMapPoints)), b(F2, flags_none(F2));
        Set set(a, b);
        const int stride = (a.n > b.n ? a.n : b.n) > 0 ? (a.n > b.n ? a.n : b.n) : 1;
        std::vector<int32_t> m21(stride);
        int32_t n = 0;
        const int32_t i1 = 0, i2 = 1;
        check(orbm_window_search(handle(), set.s, set.s, &i1, &i2, 1, windowSize, mfNNratio, mbCheckOrientation,
                                 minOctave, maxOctave, m21.data(), stride, &n));
        vpMapPointMatches2 = std::vector<MapPoint*>(F2.mvpMapPoints.size(), static_cast<MapPoint*>(NULL));   // :413
        for (int i2k = 0; i2k < b.n; i2k++)
            if (m21[i2k] >= 0) vpMapPointMatches2[i2k] = F1.mvpMapPoints[m21[i2k]];                            // :478
        return n;
    }

    // Refined matching when we have a guess of Frame 2 pose (:519-594)
    int SearchByProjection(Frame &F1, Frame &F2, int windowSize, std::vector<MapPoint *> &vpMapPointMatches2)
    {
        vpMapPointMatches2 = F2.mvpMapPoints;                                                                  // :521
        std::set<MapPoint*> found(vpMapPointMatches2.begin(), vpMapPointMatches2.end());
        const cv::Mat Rc2w = F2.mTcw.rowRange(0, 3).colRange(0, 3);
        const cv::Mat tc2w = F2.mTcw.rowRange(0, 3).col(3);
        const int n1 = (int)F1.mvpMapPoints.size();
        std::vector<unsigned char> v1(n1, 0);
        std::vector<float> uv((size_t)2 * (n1 > 0 ? n1 : 1), 0.f);
        for (int i1 = 0; i1 < n1; i1++) {                     // the 3-D projection of :540-550, as the reference writes it
            MapPoint* pMP1 = F1.mvpMapPoints[i1];
            if (!pMP1) continue;
            if (pMP1->isBad() || found.count(pMP1)) continue;
            cv::Mat x3Dw = pMP1->GetWorldPos();
            cv::Mat x3Dc2 = Rc2w * x3Dw + tc2w;
            const float xc2 = x3Dc2.at<float>(0);
            const float yc2 = x3Dc2.at<float>(1);
            const float invzc2 = 1.0 / x3Dc2.at<float>(2);
            uv[2 * i1] = F2.fx * xc2 * invzc2 + F2.cx;
            uv[2 * i1 + 1] = F2.fy * yc2 * invzc2 + F2.cy;
            v1[i1] = 1;
        }
        Flat a(F1, v1), b(F2, flags_any(F2.mvpMapPoints));
        Set set(a, b);
        const int stride = (a.n > b.n ? a.n : b.n) > 0 ? (a.n > b.n ? a.n : b.n) : 1;
        uv.resize((size_t)stride * 2, 0.f);
        std::vector<int32_t> m21(stride);
        int32_t n = 0;
        const int32_t i1 = 0, i2 = 1;
        check(orbm_search_by_projection_frames(handle(), set.s, set.s, &i1, &i2, 1, uv.data(), stride, windowSize,
                                               mfNNratio, m21.data(), &n));
        for (int k = 0; k < b.n; k++)
            if (m21[k] >= 0) vpMapPointMatches2[k] = F1.mvpMapPoints[m21[k]];                                  // :587
        return n;
    }

    // Brute force constrained to ORB that belong to the same vocabulary node (:155-284)
    int SearchByBoW(KeyFrame *pKF, Frame &F, std::vector<MapPoint*> &vpMapPointMatches)
    {
        std::vector<MapPoint*> vpMapPointsKF = pKF->GetMapPointMatches();
        Flat a(pKF->GetKeyPointsUn(), pKF->GetDescriptors(), flags_good(vpMapPointsKF), pKF->mfGridElementWidthInv,
               pKF->mfGridElementHeightInv, pKF->GetFeatureVector());
        Flat b(F, flags_none(F), true);
        Set set(a, b, true);
        const int stride = (a.n > b.n ? a.n : b.n) > 0 ? (a.n > b.n ? a.n : b.n) : 1;
        std::vector<int32_t> mF(stride);
        int32_t n = 0;
        const int32_t i1 = 0, i2 = 1;
        check(orbm_search_by_bow(handle(), set.s, set.s, &i1, &i2, 1, ORBM_BOW_KF_F, mfNNratio, mbCheckOrientation,
                                 mF.data(), stride, &n));
        vpMapPointMatches = std::vector<MapPoint*>(F.mvpMapPoints.size(), static_cast<MapPoint*>(NULL));      // :159
        for (int k = 0; k < b.n; k++)
            if (mF[k] >= 0) vpMapPointMatches[k] = vpMapPointsKF[mF[k]];                                       // :228
        return n;
    }

    int SearchByBoW(KeyFrame *pKF1, KeyFrame* pKF2, std::vector<MapPoint*> &vpMatches12)                      // :715-850
    {
        std::vector<MapPoint*> vpMapPoints1 = pKF1->GetMapPointMatches();
        std::vector<MapPoint*> vpMapPoints2 = pKF2->GetMapPointMatches();
        Flat a(pKF1->GetKeyPointsUn(), pKF1->GetDescriptors(), flags_good(vpMapPoints1), pKF1->mfGridElementWidthInv,
               pKF1->mfGridElementHeightInv, pKF1->GetFeatureVector());
        Flat b(pKF2->GetKeyPointsUn(), pKF2->GetDescriptors(), flags_good(vpMapPoints2), pKF2->mfGridElementWidthInv,
               pKF2->mfGridElementHeightInv, pKF2->GetFeatureVector());
        Set set(a, b, true);
        const int stride = (a.n > b.n ? a.n : b.n) > 0 ? (a.n > b.n ? a.n : b.n) : 1;
        std::vector<int32_t> m12(stride);
        int32_t n = 0;
        const int32_t i1 = 0, i2 = 1;
        check(orbm_search_by_bow(handle(), set.s, set.s, &i1, &i2, 1, ORBM_BOW_KF_KF, mfNNratio, mbCheckOrientation,
                                 m12.data(), stride, &n));
        vpMatches12 = std::vector<MapPoint*>(vpMapPoints1.size(), static_cast<MapPoint*>(NULL));              // :723
        for (int k = 0; k < a.n; k++)
            if (m12[k] >= 0) vpMatches12[k] = vpMapPoints2[m12[k]];                                            // :795
        return n;
    }

public:

    static const int TH_LOW = ORBM_TH_LOW;
    static const int TH_HIGH = ORBM_TH_HIGH;
    static const int HISTO_LENGTH = ORBM_HISTO_LENGTH;

protected:

    // one matcher handle per process (the reference's matcher is stateless; calls are serialised by its callers' locks)
    static orbm_handle* handle()
    {
        static orbm_handle* h = NULL;
        if (!h && orbm_create(0, &h) != ORBM_OK)
            throw std::runtime_error(std::string("ORBmatcher: ") + orbm_last_error(NULL));
        return h;
    }
    static void check(int rc)
    {
        if (rc != ORBM_OK) throw std::runtime_error(std::string("ORBmatcher: ") + orbm_last_error(handle()));
    }

    static std::vector<unsigned char> flags_none(const Frame& F) { return std::vector<unsigned char>(F.mvKeysUn.size(), 0); }
    static std::vector<unsigned char> flags_good(const std::vector<MapPoint*>& v)     // pMP && !pMP->isBad()
    {
        std::vector<unsigned char> f(v.size(), 0);
        for (size_t i = 0; i < v.size(); i++) f[i] = (v[i] && !v[i]->isBad()) ? 1 : 0;
        return f;
    }
    static std::vector<unsigned char> flags_any(const std::vector<MapPoint*>& v)      // pMP != NULL
    {
        std::vector<unsigned char> f(v.size(), 0);
        for (size_t i = 0; i < v.size(); i++) f[i] = v[i] ? 1 : 0;
        return f;
    }

    // SoA copy of what the searches read from a Frame / KeyFrame (include/orbm.h: orbm_frame, orbm_featvec)
    struct Flat {
        int n;
        std::vector<float> x, y, angle;
        std::vector<int32_t> octave, node_id, start, items;
        std::vector<unsigned char> desc, valid;
        float inv_w, inv_h;
        bool has_fv;
        Flat(const Frame& F, const std::vector<unsigned char>& v, bool with_fv = false)
        { init(F.mvKeysUn, F.mDescriptors, v, F.mfGridElementWidthInv, F.mfGridElementHeightInv); if (with_fv) fv(F.mFeatVec); }
        Flat(const std::vector<cv::KeyPoint>& k, const cv::Mat& d, const std::vector<unsigned char>& v, float iw, float ih,
             const DBoW2::FeatureVector& f)
        { init(k, d, v, iw, ih); fv(f); }
        void init(const std::vector<cv::KeyPoint>& k, const cv::Mat& d, const std::vector<unsigned char>& v, float iw, float ih)
        {
            n = (int)k.size(); inv_w = iw; inv_h = ih; has_fv = false;
            x.resize(n); y.resize(n); angle.resize(n); octave.resize(n); desc.resize((size_t)n * 32); valid = v;
            valid.resize(n, 0);
            for (int i = 0; i < n; i++) {
                x[i] = k[i].pt.x; y[i] = k[i].pt.y; angle[i] = k[i].angle; octave[i] = k[i].octave;
                memcpy(&desc[(size_t)i * 32], d.ptr<unsigned char>(i), 32);
            }
        }
        void fv(const DBoW2::FeatureVector& f)       // std::map<NodeId, vector<unsigned>>: ascending ids, in order
        {
            has_fv = true;
            start.push_back(0);
            for (DBoW2::FeatureVector::const_iterator it = f.begin(); it != f.end(); ++it) {
                node_id.push_back((int32_t)it->first);
                for (size_t j = 0; j < it->second.size(); j++) items.push_back((int32_t)it->second[j]);
                start.push_back((int32_t)items.size());
            }
        }
        orbm_frame view() const
        {
            orbm_frame f;
            f.n = n; f.x = x.data(); f.y = y.data(); f.angle = angle.data(); f.octave = octave.data();
            f.desc = desc.data(); f.valid = valid.data();
            f.min_x = (float)Frame::mnMinX; f.min_y = (float)Frame::mnMinY; f.inv_w = inv_w; f.inv_h = inv_h;
            return f;
        }
        orbm_featvec fview() const
        {
            orbm_featvec v;
            v.nnodes = (int32_t)node_id.size(); v.node_id = node_id.data(); v.start = start.data(); v.items = items.data();
            return v;
        }
    };

    struct Set {                                     // the two frames of one call on the device
        orbm_frameset* s;
        Set(const Flat& a, const Flat& b, bool with_fv = false) : s(NULL)
        {
            orbm_frame fr[2] = {a.view(), b.view()};
            orbm_featvec fv[2] = {a.fview(), b.fview()};
            check(orbm_frameset_create(handle(), fr, with_fv ? fv : NULL, 2, &s));
        }
        ~Set() { if (s) orbm_frameset_destroy(s); }
    };

    float mfNNratio;
    bool mbCheckOrientation;
};

}// namespace ORB_SLAM

#endif

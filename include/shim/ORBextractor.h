/* Drop-in replacement of the reference's include/ORBextractor.h (worxli/ORB_SLAM include/ORBextractor.h:32-77):
 * class ORB_SLAM::ORBextractor with the same public members, implemented on the C ABI of liborbx.so (include/orbx.h),
 * i.e. on the B200 kernels.  Header-only; needs OpenCV's core types (cv::Mat, cv::KeyPoint, InputArray / OutputArray)
 * from whatever <opencv/cv.h> the including project uses, and -lorbx at link time.
 *
 * Use: put this directory BEFORE the reference's include/ on the include path and add `-include shim/ORBextractor.h`
 * (the guard macro below is the reference's own, so its header becomes a no-op wherever it is included afterwards -
 * reference headers include it with quotes, which would otherwise find their sibling first), drop src/ORBextractor.cc
 * from the build.  The reference's src/Frame.cc:59, `(*mpORBextractor)(im, cv::Mat(), mvKeys, mDescriptors)`, and
 * src/Tracking.cc:128-130,105-109 (construction from Settings.yaml) then compile and run unchanged
 * (oracle/Makefile target ref_dropin, tests/test_gpu_dropin.py).
 *
 * Differences from the reference class, all outside its public surface: the protected members (pyramid, pattern,
 * tables) do not exist - the pyramid lives in HBM inside the handle; errors the reference asserts on throw
 * std::runtime_error with the library's message. */
#ifndef ORBEXTRACTOR_H
#define ORBEXTRACTOR_H

#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include <opencv/cv.h>

#include "../orbx.h"

namespace ORB_SLAM
{

class ORBextractor
{
public:

    enum {HARRIS_SCORE=0, FAST_SCORE=1 };

    ORBextractor(int nfeatures = 1000, float scaleFactor = 1.2f, int nlevels = 8, int scoreType = FAST_SCORE,
                 int fastTh = 20)
        : nfeatures(nfeatures), scaleFactor(scaleFactor), nlevels(nlevels), scoreType(scoreType), fastTh(fastTh),
          handle_(NULL), cap_w_(0), cap_h_(0)
    {
        // OpenCV-version dependent arithmetic (include/orbx.h): default = the OpenCV 2.4 era of the reference; a host
        // built against OpenCV >= 3.4 sets ORBX_BLUR_MODE=1 ORBX_RETAIN_MODE=1 to get that library's results
        blur_mode_ = env_int("ORBX_BLUR_MODE", ORBX_BLUR_CV24);
        retain_mode_ = env_int("ORBX_RETAIN_MODE", ORBX_RETAIN_CV24);
        device_ = env_int("ORBX_DEVICE", 0);
    }

    ~ORBextractor() { if (handle_) orbx_destroy(handle_); }

    // Compute the ORB features and descriptors on an image (reference src/ORBextractor.cc:718-779).  `mask` is accepted
    // and ignored, exactly as the reference's detection code ignores it (:601-603).
    void operator()( cv::InputArray _image, cv::InputArray /*mask*/,
      std::vector<cv::KeyPoint>& keypoints,
      cv::OutputArray descriptors)
    {
        cv::Mat image = _image.getMat();
        if (image.empty())                       // :721-722 - outputs untouched
            return;
        if (image.type() != CV_8UC1)            // assert(image.type() == CV_8UC1), :725
            throw std::runtime_error("ORBextractor: image must be CV_8UC1");
        ensure(image.cols, image.rows);
        static_assert(sizeof(cv::KeyPoint) == sizeof(orbx_keypoint), "cv::KeyPoint is not the 28-byte record of orbx.h");
        kp_.resize(nfeatures);
        desc_.resize((size_t)nfeatures * 32);
        int n = 0;
        const int rc = orbx_extract(handle_, image.data, image.cols, image.rows, (size_t)image.step,
                                    reinterpret_cast<orbx_keypoint*>(kp_.data()), desc_.data(), &n);
        if (rc != ORBX_OK)
            throw std::runtime_error(std::string("ORBextractor: ") + orbx_last_error(handle_));
        keypoints.assign(kp_.begin(), kp_.begin() + n);        // :746-747 clear + refill
        if (n == 0) {                                          // :738-739
            descriptors.release();
            return;
        }
        descriptors.create(n, 32, CV_8U);                      // :742
        cv::Mat d = descriptors.getMat();
        for (int i = 0; i < n; i++)
            memcpy(d.ptr(i), &desc_[(size_t)i * 32], 32);
    }

    int inline GetLevels(){
        return nlevels;}

    float inline GetScaleFactor(){
        return scaleFactor;}

protected:

    static int env_int(const char* name, int dflt) { const char* e = getenv(name); return e ? atoi(e) : dflt; }

    // one handle, sized for the largest image seen so far (the reference reallocates its pyramid per call, :781-822)
    void ensure(int w, int h)
    {
        if (handle_ && w <= cap_w_ && h <= cap_h_)
            return;
        if (handle_) { orbx_destroy(handle_); handle_ = NULL; }
        orbx_params p;
        orbx_default_params(&p);
        p.nfeatures = nfeatures; p.scale_factor = (float)scaleFactor; p.nlevels = nlevels; p.score_type = scoreType;
        p.fast_th = fastTh; p.max_width = w > cap_w_ ? w : cap_w_; p.max_height = h > cap_h_ ? h : cap_h_;
        p.max_batch = 1; p.device = device_; p.blur_mode = blur_mode_; p.retain_mode = retain_mode_;
        if (orbx_create(&p, &handle_) != ORBX_OK)
            throw std::runtime_error(std::string("ORBextractor: ") + orbx_last_error(NULL));
        cap_w_ = p.max_width; cap_h_ = p.max_height;
    }

    int nfeatures;
    double scaleFactor;
    int nlevels;
    int scoreType;
    int fastTh;

    orbx_handle* handle_;
    int cap_w_, cap_h_, blur_mode_, retain_mode_, device_;
    std::vector<cv::KeyPoint> kp_;
    std::vector<unsigned char> desc_;

private:
    ORBextractor(const ORBextractor&);             // owns a device handle
    ORBextractor& operator=(const ORBextractor&);
};

} //namespace ORB_SLAM

#endif

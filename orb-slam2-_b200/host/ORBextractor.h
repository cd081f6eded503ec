// Drop-in replacement of ORB_SLAM2::ORBextractor (R/include/ORBextractor.h:45-123) over liborbx.so.
//
// Same class name, namespace, constructor, operator(), getters and the public mvImagePyramid member, so
// Frame.cpp / Tracking.cpp compile unchanged against this header (R/src/Frame.cpp:86-89,149,204,262-268,
// 558-695; R/src/Tracking.cpp:130-149). All compute happens in the CUDA library behind include/orbx.h;
// there is no CPU path: a missing device or a failing call aborts with the library's message (the
// reference has no error channel either: no exceptions, no return codes).
#ifndef ORBEXTRACTOR_H
#define ORBEXTRACTOR_H

#include <vector>
#include <list>
#include <opencv/cv.h>

struct orbx_extractor;

namespace ORB_SLAM2
{

class ORBextractor
{
public:

    enum {HARRIS_SCORE=0, FAST_SCORE=1 };

    ORBextractor(int nfeatures, float scaleFactor, int nlevels,
                 int iniThFAST, int minThFAST);

    ~ORBextractor();

    // Compute the ORB features and descriptors on an image.
    // ORB are dispersed on the image using an octree. Mask is ignored (as in the reference).
    void operator()( cv::InputArray image, cv::InputArray mask,
      std::vector<cv::KeyPoint>& keypoints,
      cv::OutputArray descriptors);

    int inline GetLevels(){
        return nlevels;}

    float inline GetScaleFactor(){
        return scaleFactor;}

    std::vector<float> inline GetScaleFactors(){
        return mvScaleFactor;
    }

    std::vector<float> inline GetInverseScaleFactors(){
        return mvInvScaleFactor;
    }

    std::vector<float> inline GetScaleSigmaSquares(){
        return mvLevelSigma2;
    }

    std::vector<float> inline GetInverseScaleSigmaSquares(){
        return mvInvLevelSigma2;
    }

    // Host copies of the pyramid levels of the last image (level 0 aliases the caller's image, like
    // R/src/ORBextractor.cpp:1223; the others alias a page-locked buffer of this extractor and are valid until
    // its next operator() call). Only the reference's own Frame::ComputeStereoMatches reads them: monocular /
    // RGB-D callers, and stereo callers that link host/FrameStereo.cc, switch the read-back off
    // (SetPyramidReadback(false) or ORBX_PYRAMID_READBACK=0) and save the device->host copy.
    std::vector<cv::Mat> mvImagePyramid;
    void SetPyramidReadback(bool on) { mbPyramidReadback = on; }

    // CUDA device the extractor runs on (default: ORBX_DEVICE from the environment, else 0);
    // takes effect at the next operator() call.
    void SetDevice(int device);

    // The liborbx handle behind this extractor: orbx_stereo_match reads the device pyramids of the
    // last operator() call through it (the device-side Frame::ComputeStereoMatches).
    orbx_extractor* handle() const { return mpHandle; }

protected:

    void EnsureHandle(int width, int height);
    void EnsureOutputs(int cap);

    int nfeatures;
    double scaleFactor;
    int nlevels;
    int iniThFAST;
    int minThFAST;

    std::vector<int> mnFeaturesPerLevel;

    std::vector<int> umax;

    std::vector<float> mvScaleFactor;
    std::vector<float> mvInvScaleFactor;
    std::vector<float> mvLevelSigma2;
    std::vector<float> mvInvLevelSigma2;

    orbx_extractor* mpHandle;
    int mnDevice, mnMaxWidth, mnMaxHeight;
    bool mbPyramidReadback;
    unsigned char* mpPyramidHost;   // page-locked backing store of mvImagePyramid[1..]
    size_t mnPyramidBytes;
    void* mpOutKeys;                // page-locked result rows of the last call (orbx_keypoint == cv::KeyPoint, 28 bytes)
    unsigned char* mpOutDesc;
    int mnOutCap;

private:
    ORBextractor(const ORBextractor&);             // owns a device handle: not copyable
    ORBextractor& operator=(const ORBextractor&);
};

} //namespace ORB_SLAM

#endif

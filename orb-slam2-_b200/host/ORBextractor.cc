// ORB_SLAM2::ORBextractor over liborbx.so -- see ORBextractor.h. Mirrors the observable behaviour of
// R/src/ORBextractor.cpp:418-502 (ctor tables) and :1120-1188 (operator()).
#include "ORBextractor.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "orbx.h"

namespace ORB_SLAM2
{

static_assert(sizeof(cv::KeyPoint) == sizeof(orbx_keypoint),
              "cv::KeyPoint {Point2f pt; float size, angle, response; int octave, class_id} must be 28 bytes");

static void Fail(const char* what)
{
    // no CPU fallback and no error channel in the reference's interface: fail loudly
    std::fprintf(stderr, "ORBextractor (orbx): %s: %s\n", what, orbx_last_error());
    std::abort();
}

ORBextractor::ORBextractor(int _nfeatures, float _scaleFactor, int _nlevels,
                           int _iniThFAST, int _minThFAST):
    nfeatures(_nfeatures), nlevels(_nlevels), iniThFAST(_iniThFAST), minThFAST(_minThFAST),
    scaleFactor(_scaleFactor), mpHandle(NULL), mnDevice(0), mnMaxWidth(0), mnMaxHeight(0),
    mbPyramidReadback(true), mpPyramidHost(NULL), mnPyramidBytes(0),
    mpOutKeys(NULL), mpOutDesc(NULL), mnOutCap(0)
{
    if(const char* dev = std::getenv("ORBX_DEVICE"))
        mnDevice = std::atoi(dev);
    // ORBX_PYRAMID_READBACK=0: leave mvImagePyramid[1..] empty (nothing reads it once Frame::ComputeStereoMatches is
    // host/FrameStereo.cc, which works on the device pyramids); SetPyramidReadback() overrides
    if(const char* rb = std::getenv("ORBX_PYRAMID_READBACK"))
        mbPyramidReadback = std::atoi(rb) != 0;

    // The tables only depend on the ctor arguments; a 64x64 probe handle computes them with the library's
    // arithmetic (the reference's own types and operation order, R/src/ORBextractor.cpp:435-500).
    mvScaleFactor.resize(nlevels);
    mvInvScaleFactor.resize(nlevels);
    mvLevelSigma2.resize(nlevels);
    mvInvLevelSigma2.resize(nlevels);
    mnFeaturesPerLevel.resize(nlevels);
    umax.resize(16);
    mvImagePyramid.resize(nlevels);

    orbx_params p;
    p.nfeatures = nfeatures; p.scale_factor = _scaleFactor; p.nlevels = nlevels;
    p.ini_th_fast = iniThFAST; p.min_th_fast = minThFAST; p.device = mnDevice;
    p.max_width = 4095; p.max_height = 4095; p.max_batch = 1;
    orbx_extractor* probe = NULL;
    if(orbx_extractor_create(&p, &probe) != ORBX_OK)
        Fail("orbx_extractor_create");
    if(orbx_extractor_tables(probe, &mvScaleFactor[0], &mvInvScaleFactor[0], &mvLevelSigma2[0],
                             &mvInvLevelSigma2[0], &mnFeaturesPerLevel[0], &umax[0]) != ORBX_OK)
        Fail("orbx_extractor_tables");
    mpHandle = probe;   // arenas are allocated on first use, for the size of the first image
    mnMaxWidth = p.max_width;
    mnMaxHeight = p.max_height;
}

ORBextractor::~ORBextractor()
{
    orbx_extractor_destroy(mpHandle);
    if(mpPyramidHost)
        orbx_host_free(mpPyramidHost);
    if(mpOutKeys)
        orbx_host_free(mpOutKeys);
    if(mpOutDesc)
        orbx_host_free(mpOutDesc);
}

void ORBextractor::SetDevice(int device)
{
    if(device == mnDevice)
        return;
    mnDevice = device;
    orbx_extractor_destroy(mpHandle);
    mpHandle = NULL;
}

void ORBextractor::EnsureHandle(int width, int height)
{
    if(mpHandle && width <= mnMaxWidth && height <= mnMaxHeight)
        return;
    orbx_extractor_destroy(mpHandle);
    mpHandle = NULL;
    orbx_params p;
    p.nfeatures = nfeatures; p.scale_factor = (float)scaleFactor; p.nlevels = nlevels;
    p.ini_th_fast = iniThFAST; p.min_th_fast = minThFAST; p.device = mnDevice;
    p.max_width = width > 4095 ? width : 4095; p.max_height = height > 4095 ? height : 4095; p.max_batch = 1;
    if(orbx_extractor_create(&p, &mpHandle) != ORBX_OK)
        Fail("orbx_extractor_create");
    mnMaxWidth = p.max_width;
    mnMaxHeight = p.max_height;
}

void ORBextractor::EnsureOutputs(int cap)
{
    if(cap <= mnOutCap)
        return;
    if(mpOutKeys) orbx_host_free(mpOutKeys);
    if(mpOutDesc) orbx_host_free(mpOutDesc);
    mpOutKeys = NULL; mpOutDesc = NULL; mnOutCap = 0;
    void *k = NULL, *d = NULL;
    if(orbx_host_alloc(mnDevice, (size_t)cap * sizeof(orbx_keypoint), &k) != ORBX_OK ||
       orbx_host_alloc(mnDevice, (size_t)cap * 32, &d) != ORBX_OK)
        Fail("orbx_host_alloc");
    mpOutKeys = k;
    mpOutDesc = static_cast<unsigned char*>(d);
    mnOutCap = cap;
}

void ORBextractor::operator()( cv::InputArray _image, cv::InputArray _mask, std::vector<cv::KeyPoint>& _keypoints,
                      cv::OutputArray _descriptors)
{
    (void)_mask;                       // ignored by the reference as well
    if(_image.empty())
        return;                        // outputs untouched (R/src/ORBextractor.cpp:1123-1124)

    cv::Mat image = _image.getMat();
    if(image.type() != CV_8UC1)        // the reference only asserts (:1127), compiled out in Release
    {
        std::fprintf(stderr, "ORBextractor (orbx): image must be CV_8UC1\n");
        std::abort();
    }

    EnsureHandle(image.cols, image.rows);

    // Results land in page-locked arrays this extractor owns (the library then copies device rows straight into
    // them, no staging), sized for the upper bound on the keypoints of one image: every level's quota plus the
    // octree's overshoot.
    int cap = nfeatures + 64 * nlevels + 512;
    int n = 0;
    int rc = ORBX_OK;
    for(int attempt=0; attempt<2; attempt++)
    {
        EnsureOutputs(cap);
        rc = orbx_extract(mpHandle, image.data, image.cols, image.rows, (size_t)image.step,
                          static_cast<orbx_keypoint*>(mpOutKeys), mpOutDesc, mnOutCap, &n);
        if(rc != ORBX_ERR_CAPACITY || n <= mnOutCap)
            break;
        cap = n;
    }
    if(rc != ORBX_OK)
        Fail("orbx_extract");

    if(n == 0)
        _descriptors.release();        // :1145
    else
    {
        _descriptors.create(n, 32, CV_8U);
        cv::Mat descriptors = _descriptors.getMat();
        if(descriptors.isContinuous())
            std::memcpy(descriptors.ptr(0), mpOutDesc, (size_t)n * 32);
        else
            for(int i=0; i<n; i++)
                std::memcpy(descriptors.ptr(i), mpOutDesc + (size_t)i*32, 32);
    }

    const cv::KeyPoint* first = static_cast<const cv::KeyPoint*>(mpOutKeys);
    _keypoints.assign(first, first + n);   // clear + fill (:1153-1186)

    // mvImagePyramid: level 0 aliases the caller's image (:1223). The other levels come back in one batch of DMA
    // copies into a page-locked buffer this extractor owns, and the cv::Mat headers alias that buffer (row step = the
    // device pitch): like the reference's, they are valid until the next operator() call on this extractor.
    mvImagePyramid[0] = image;
    if(mbPyramidReadback && nlevels > 1)
    {
        size_t offsets[ORBX_MAX_LEVELS], bytes = 0;
        int pitches[ORBX_MAX_LEVELS];
        if(orbx_get_pyramid(mpHandle, 0, NULL, 0, offsets, pitches, &bytes) != ORBX_OK)
            Fail("orbx_get_pyramid");
        if(bytes > mnPyramidBytes)
        {
            if(mpPyramidHost)
                orbx_host_free(mpPyramidHost);
            mpPyramidHost = NULL;
            mnPyramidBytes = 0;
            void* p = NULL;
            if(orbx_host_alloc(mnDevice, bytes, &p) != ORBX_OK)
                Fail("orbx_host_alloc");
            mpPyramidHost = static_cast<unsigned char*>(p);
            mnPyramidBytes = bytes;
        }
        if(orbx_get_pyramid(mpHandle, 0, mpPyramidHost, mnPyramidBytes, offsets, pitches, &bytes) != ORBX_OK)
            Fail("orbx_get_pyramid");
        for(int level=1; level<nlevels; ++level)
        {
            int w = 0, h = 0;
            if(orbx_level_size(mpHandle, level, &w, &h) != ORBX_OK)
                Fail("orbx_level_size");
            mvImagePyramid[level] = cv::Mat(h, w, CV_8UC1, mpPyramidHost + offsets[level], (size_t)pitches[level]);
        }
    }
}

} //namespace ORB_SLAM

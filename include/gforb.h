/* gforb.h -- C ABI of the B200-native GF-ORB-SLAM tracking front end (libgforb.so).
 *
 * Drop-in boundary for ONE hot path of Aidenryan/GF_ORB_SLAM: ORB extraction -> Hamming matching -> good-feature
 * selection.  The reference has no FFI: the "operator API" is three C++ classes called synchronously from the
 * tracking thread.  Each entry point below names the reference member function it replaces (paths relative to the
 * reference tree); include/gforb/*.hpp re-creates the class signatures on top of this ABI.
 *
 * Conventions
 *  - every function returns GFORB_OK (0) or a negative GFORB_ERR_*; nothing throws.
 *  - handles are opaque; one CUDA device per handle; work is enqueued on `stream` (a cudaStream_t passed as void*,
 *    NULL = the CUDA default stream) and, unless GFORB_ASYNC is set in `flags`, the call returns after the stream
 *    has been synchronised.
 *  - `flags & GFORB_IN_DEVICE` / `GFORB_OUT_DEVICE`: the input / output pointers are device pointers (no copies);
 *    otherwise they are host pointers and the call performs the H2D / D2H copies itself (pinned memory recommended).
 *  - batch-first: all frame-level calls take `batch` independent units.
 *  - There is no CPU fallback: if no CUDA device is usable every create() fails with GFORB_ERR_CUDA.
 */
#ifndef GFORB_H
#define GFORB_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GFORB_OK 0
#define GFORB_ERR_CUDA (-1)        /* CUDA runtime error / no device */
#define GFORB_ERR_ARG (-2)         /* bad argument (the reference would assert) */
#define GFORB_ERR_CAPACITY (-3)    /* exceeds the sizes given at create() */
#define GFORB_ERR_GEOMETRY (-4)    /* cell grid degenerate for this image size (reference: cv::Exception / div by 0) */
#define GFORB_ERR_OVERFLOW (-5)    /* internal candidate list overflow (reported, never silent) */

#define GFORB_IN_DEVICE 1
#define GFORB_OUT_DEVICE 2
#define GFORB_ASYNC 4

/* Layout-identical to cv::KeyPoint (28 bytes): pt.x, pt.y, size, angle, response, octave, class_id. */
typedef struct GforbKeyPoint {
    float x, y, size, angle, response;
    int32_t octave, class_id;
} GforbKeyPoint;

const char* gforb_version(void);
int gforb_device_count(void);

/* ------------------------------------------------------------------------------------------------ extraction
 * Replaces ORB_SLAM::ORBextractor (include/ORBextractor.h:51-104, src/ORBextractor.cc:464-998). */
typedef struct GforbExtractorConfig {
    int32_t nfeatures;        /* ctor arg nfeatures (1000) */
    float scaleFactor;        /* ctor arg scaleFactor (1.2f) */
    int32_t nlevels;          /* ctor arg nlevels (8) */
    int32_t scoreType;        /* 1 = FAST_SCORE (only mode implemented; HARRIS_SCORE=0 -> GFORB_ERR_ARG) */
    int32_t iniThFAST;        /* ctor arg fastTh (20) */
    int32_t minThFAST;        /* hard-coded 7 in the reference (src/ORBextractor.cc:627) */
    int32_t blur_variant;     /* 0 = OpenCV>=3.4 bit-exact Gaussian (cv2-4.13 verified), 1 = legacy 2.4 integer kernel */
    int32_t retain_variant;   /* 0 = nth_element at begin()+n (OpenCV 2.4.13/3.4.1), 1 = begin()+n-1 */
    int32_t max_width, max_height, max_batch;
    int32_t device;           /* CUDA device ordinal */
} GforbExtractorConfig;

typedef struct GforbExtractor GforbExtractor;

int gforb_extractor_create(const GforbExtractorConfig* cfg, GforbExtractor** out);
int gforb_extractor_destroy(GforbExtractor* h);

/* ORBextractor::operator()(image, mask, keypoints, descriptors) for `batch` frames (src/ORBextractor.cc:769-875).
 * images: 8-bit single channel, row stride `stride`, frame b at images + b*frame_stride.  The mask argument of the
 * reference is dead code (always empty, src/Frame.cc:59) and is not part of the ABI.
 * keypoints: batch x nfeatures, descriptors: batch x nfeatures x 32 bytes, counts: batch (valid rows per frame);
 * output order is exactly the reference's (level-major, then cell order, then retainBest permutation). */
int gforb_extract(GforbExtractor* h, const uint8_t* images, int width, int height, int stride, size_t frame_stride,
                  int batch, GforbKeyPoint* keypoints, uint8_t* descriptors, int32_t* counts, int flags, void* stream);

/* ORBextractor::GetLevels / GetScaleFactor (include/ORBextractor.h:66-70) and the ctor tables
 * (mvScaleFactor, mnFeaturesPerLevel; src/ORBextractor.cc:469-494). */
int gforb_extractor_levels(const GforbExtractor* h);
float gforb_extractor_scale_factor(const GforbExtractor* h);
int gforb_extractor_tables(const GforbExtractor* h, float* scale_factors, int32_t* features_per_level);

/* Debug/parity taps on the last gforb_extract() call: copies level `level` of frame `b` (bordered plane,
 * (h+32) x pitch bytes, pitch returned) -- which = 0: unblurred pyramid (ComputePyramid output), 1: blurred ROI. */
int gforb_extractor_debug_level(GforbExtractor* h, int b, int level, int which, uint8_t* dst, size_t dst_bytes,
                                int* w, int* hgt, int* pitch);
/* number of kernels launched by the last gforb_extract() call on this handle */
int gforb_extractor_last_launches(const GforbExtractor* h);
/* name + device time (ms, CUDA events on the launch stream) of stage i of the last profiled call; returns the
 * number of stages.  Profiling is enabled with gforb_extractor_set_profiling(h, 1). */
int gforb_extractor_set_profiling(GforbExtractor* h, int on);
int gforb_extractor_stage_time(const GforbExtractor* h, int i, const char** name, float* ms);

#ifdef __cplusplus
}
#endif
#endif /* GFORB_H */

// dropin/ORBextractor.h — takes the place of the reference's include/ORBextractor.h: copy it over that file, or force-include
// it (`-include dropin/ORBextractor.h`; it defines the same include guard, so the reference's header is then skipped), and every
// `#include "ORBextractor.h"` (Frame.h, KeyFrame.h, Tracking.h, System.h) gets the orbx-backed ORB_SLAM2::ORBextractor of
// include/orbx_cxx.hpp instead; src/ORBextractor.cc is then left out of the build.
//
// Frame::ComputeStereoMatches (src/Frame.cc:466-640) reads mpORBextractorLeft->mvImagePyramid on the host.  Until that body
// is replaced by ORB_SLAM2::ComputeStereoMatchesGPU, define ORBX_DROPIN_HOST_PYRAMID: the stand-in then
// copies the pyramid back after every extraction, which keeps the unmodified Frame.cc working (at the cost of the copy).
#ifndef ORBEXTRACTOR_H
#define ORBEXTRACTOR_H
#define ORBX_CXX_EXTRACTOR_ONLY   // the reference keeps its own ORBmatcher.h; dropin/ORBmatcher.cc implements it
#include "orbx_cxx.hpp"
#endif

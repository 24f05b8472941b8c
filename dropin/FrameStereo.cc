// dropin/FrameStereo.cc — Frame::ComputeStereoMatches (src/Frame.cc:466-640) on the GPU: the body INTEGRATION.md tells a
// maintainer to put in place of the reference's row table + Hamming scan + 11x11 SAD + parabola + median cut.
//
// Both extractors are the orbx stand-ins (dropin/ORBextractor.h); their keypoints, descriptors and pyramids are still in
// device slot 0 of each handle when the stereo constructor gets here (src/Frame.cc:78-90), so nothing is uploaded and no
// pyramid is copied back to the host: one orbx_stereo_match_batch call writes mvuRight / mvDepth.
//
// In a patched tree this definition simply replaces the one in src/Frame.cc.  To link it with the UNMODIFIED src/Frame.cc
// (oracle/_ref/orbref_slam_gpustereo, tests/test_zz_dropin.py), dropin/stereo_shim/Frame.h marks the member weak where the
// class declares it, so the reference's own definition yields to this one at link time.
#define ORBX_FRAME_STEREO_DEFINITION   // see dropin/stereo_shim/Frame.h; without the shim this does nothing
#include "Frame.h"

namespace ORB_SLAM2 {

void Frame::ComputeStereoMatches() {
  // the static Frame::fx is assigned after this call on the first frame (src/Frame.cc:90 vs :105): read mK instead
  ComputeStereoMatchesGPU(*mpORBextractorLeft, *mpORBextractorRight, N, mbf, mK.at<float>(0, 0), mvuRight, mvDepth);
}

}  // namespace ORB_SLAM2

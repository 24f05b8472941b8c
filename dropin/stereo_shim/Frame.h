// dropin/stereo_shim/Frame.h — found before the reference's include/Frame.h (put -Idropin/stereo_shim first).  It passes
// straight on to that header, but for its duration the declaration `void ComputeStereoMatches();` (include/Frame.h:93) reads
// `void ComputeStereoMatches() __attribute__((weak));`.  The unmodified src/Frame.cc then emits its CPU body as a weak symbol,
// which the compiler neither inlines nor reasons about, and the definition in dropin/FrameStereo.cc takes its place when
// the two are linked.  Only needed to test the GPU body against an unpatched tree; a maintainer edits src/Frame.cc instead.
// dropin/FrameStereo.cc defines ORBX_FRAME_STEREO_DEFINITION before it includes Frame.h: its definition stays strong.
#ifndef ORBX_STEREO_SHIM_FRAME_H
#define ORBX_STEREO_SHIM_FRAME_H
#ifdef ORBX_FRAME_STEREO_DEFINITION
#include_next "Frame.h"
#else
#define ComputeStereoMatches() ComputeStereoMatches() __attribute__((weak))
#include_next "Frame.h"
#undef ComputeStereoMatches
#endif
#endif

// ORACLE — TEST INFRASTRUCTURE ONLY.  Drives the reference's own DBoW2 (Thirdparty/DBoW2, compiled where it lies
// against oracle/cvshim) so that the restatement of TemplatedVocabulary::transform in orb_oracle.hpp can be
// checked against it:  orbref_dbow <vocabulary.txt> <descriptors.bin> <levelsup>
// vocabulary.txt is in the reference's loadFromTextFile format; descriptors.bin = N x 32 bytes.
// Output: one line per feature "F i word node" (through the single-feature transform), then "B word value" per
// BowVector entry (hex float), then "V node i i i ..." per FeatureVector entry.
#include <cstdio>
#include <fstream>
#include <vector>

#include "DBoW2/FORB.h"
#include "DBoW2/TemplatedVocabulary.h"

typedef DBoW2::TemplatedVocabulary<DBoW2::FORB::TDescriptor, DBoW2::FORB> ORBVocabulary;   // include/ORBVocabulary.h:30

int main(int argc, char** argv) {
  if (argc < 4) return 2;
  ORBVocabulary voc;
  if (!voc.loadFromTextFile(argv[1])) return 3;
  std::ifstream f(argv[2], std::ios::binary);
  std::vector<unsigned char> raw((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
  const int n = (int)(raw.size() / 32), levelsup = atoi(argv[3]);
  std::vector<cv::Mat> feats(n);
  for (int i = 0; i < n; ++i) {
    feats[i].create(1, 32, CV_8U);
    std::memcpy(feats[i].ptr<unsigned char>(), &raw[(size_t)i * 32], 32);
  }
  DBoW2::BowVector bv;
  DBoW2::FeatureVector fv;
  voc.transform(feats, bv, fv, levelsup);
  for (DBoW2::BowVector::const_iterator it = bv.begin(); it != bv.end(); ++it) printf("B %u %a\n", it->first, it->second);
  for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it) {
    printf("V %u", it->first);
    for (size_t k = 0; k < it->second.size(); ++k) printf(" %u", it->second[k]);
    printf("\n");
  }
  return 0;
}

#!/usr/bin/env python3
"""Generates tests/golden/ref_bow.json from the REFERENCE ITSELF: oracle/_ref/orbref_dbow is the reference's
Thirdparty/DBoW2 compiled in place (make -C oracle ref).  Seeded synthetic vocabularies (orb_slam2_b200/synth.py:
vocabulary) and seeded descriptor sets go through TemplatedVocabulary::transform; the file stores digests of the
BowVector (words, double bits) and of the FeatureVector (node per feature) plus a few literal entries.
Run in the build container only."""
import hashlib, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from orb_slam2_b200 import synth
from oracle import pyoracle as O

# (seed, k, L, levelsup, scoring, weighting, norm): scoring 0 L1_NORM, 1 L2_NORM, 5 DOT_PRODUCT; weighting 0 TF_IDF .. 3 BINARY
CASES = [(11, 10, 4, 2, 0, 0, 1), (12, 10, 3, 4, 0, 0, 1), (13, 6, 5, 3, 1, 1, 2), (14, 10, 4, 2, 0, 2, 1), (15, 8, 3, 1, 5, 3, 0), (16, 10, 4, 2, 5, 1, 0)]


def descriptors(seed, voc, n=1500):
    rng = np.random.default_rng(seed)
    base = voc["desc"][rng.integers(1, len(voc["desc"]), n)]
    return np.ascontiguousarray(base ^ np.packbits(rng.random((n, 256)) < 0.08, axis=1))


if __name__ == "__main__":
    out = {"generator": "tools/gen_golden_bow.py", "source": "oracle/_ref/orbref_dbow (reference Thirdparty/DBoW2, compiled in place)", "cases": []}
    for seed, k, L, levelsup, scoring, weighting, norm in CASES:
        V = synth.vocabulary(seed, k, L, stop_fraction=0.05)
        D = descriptors(seed + 100, V)
        node, bw, bv = O.ref_dbow_transform(V, D, levelsup, scoring, weighting)
        out["cases"].append(dict(seed=seed, k=k, L=L, levelsup=levelsup, scoring=scoring, weighting=weighting, norm=norm,
                                 voc_sha256=hashlib.sha256(V["desc"].tobytes() + V["weight"].tobytes()).hexdigest(),
                                 desc_sha256=hashlib.sha256(D.tobytes()).hexdigest(), bow_size=int(len(bw)),
                                 bow_words_sha256=hashlib.sha256(bw.tobytes()).hexdigest(),
                                 bow_values_sha256=hashlib.sha256(bv.tobytes()).hexdigest(),
                                 nodes_sha256=hashlib.sha256(node.tobytes()).hexdigest(),
                                 first_words=bw[:4].tolist(), first_values_hex=[float(v).hex() for v in bv[:4]], stopped=int((node < 0).sum())))
    json.dump(out, open(os.path.join(ROOT, "tests/golden/ref_bow.json"), "w"), indent=1)
    print("wrote", len(out["cases"]), "cases")

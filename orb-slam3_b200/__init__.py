"""B200-native ORB front-end: drop-in for ORB-SLAM3's ORBextractor / ORBmatcher hot path.

The product is liborbx.so (CUDA sm_100a kernels behind the C ABI of include/orbx.h) plus the C++
shims in cpp/.  This Python package only binds the C ABI for tests, bench.py and smoke().
The directory name contains a hyphen, so import it through `_pkg.load()` at the repo root
(registers the package as `orbslam3_b200`).
"""
from . import binding, synth  # noqa: F401
from .binding import KP_DTYPE, OrbxError  # noqa: F401
from .extractor import ORBextractor, DescriptorDistance, extract_batch_multi  # noqa: F401
from .matcher import (ComputeStereoMatches, ExtractStereo, knn2_hamming, SearchByProjection, SearchByProjectionFrame,  # noqa: F401
                      SearchByProjectionKeyFrame,
                      SearchForInitialization, SearchForTriangulation, SearchByBoW, GetFeaturesInArea,
                      SearchLocalPoints, SearchByProjectionRig, UndistortKeyPoints, ComputeDistinctiveDescriptors, ORBVocabulary, SearchByProjectionSim3, Fuse, SearchBySim3, ComputeThreeMaxima,
                      SearchByProjectionFrameRig, SearchByBoWRig, SearchForTriangulationCb, ComputeStereoFishEyeMatches)

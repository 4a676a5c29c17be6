"""TEST INFRASTRUCTURE ONLY -- the oracle of the separator-detection hot path.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU legs (`cpu_baseline`, `--impl reference`) import this
package, and only as the checker or the timed CPU baseline: the product path (`libsepb200.so`, `host/`, the
`multi_robot_slam_separators_b200` package) never does and fails loudly without its CUDA library.

Pin status.  The reference ships no tests, fixtures or golden vectors for this path and cannot be
built or imported here (ROS melodic, RTAB-Map, OpenCV C++, Python 2): there is no `oracle/_ref`.  Parity is therefore
UNPINNED with respect to the reference's own binaries.  What is pinned: the arithmetic of the third-party calls the
reference makes at its call sites -- `scipy.spatial.distance.cdist` (data_handler.py:170), `cv2.BFMatcher(NORM_HAMMING)
.knnMatch(k=2)` (myRegistrationVis.cpp:580,739,852 via VWDictionary), `cv2.solvePnPRansac` (:1077 via
util3d::estimateMotion3DTo2D) -- executed here (scipy 1.18, cv2 4.13), plus operation-for-operation restatements of
OpenCV's Jacobi SVD / EPnP / RANSAC loop / CvLevMarq that are bit-identical to those calls on the committed fixtures
(`tests/golden/*.npz`, generator `tests/golden/make_golden.py`).  Each module cites the reference file:line it follows.
"""

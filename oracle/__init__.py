"""ORACLE — TEST INFRASTRUCTURE ONLY.

CPU restatement of the reference's TrackKLT hot path (ov_core/src/track/TrackKLT.cpp, Grider_GRID.h), used only as the
checker by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.  Nothing in the
product package (uvio_b200/) imports it.

Two layers:
  * ``trackklt_cv2``  — the reference's control flow restated line by line, driving the *real* OpenCV kernels through
                        cv2 (OpenCV is the un-vendored dependency where all of the reference's pixel arithmetic lives).
  * ``cstages``       — plain-C restatement of each OpenCV routine (oracle/stages.c), validated against cv2.

Pinning status: the reference ships no golden vectors, unit tests or fixtures for this path, and
ov_core cannot be compiled here (no OpenCV C++/Eigen/Boost headers).  Parity is therefore *unpinned by the reference's
own tests*; it is pinned to OpenCV 4.13.0 (cv2, IPP off) at the reference's exact call sites and arguments.
"""

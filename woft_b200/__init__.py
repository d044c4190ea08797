"""woft_b200: B200-native (sm_100a) drop-in for the per-frame hot path of the WOFT planar tracker.

Seams: `CorrBlock` (correlation volume + pyramid + lookup), `RAFTWrapper` (flow wrapper the
tracker calls), `find_homography_*` (weighted homography estimators).  All device work goes through the
C-ABI library `libwoft_b200.so` (include/woft_b200.h); there is no CPU or eager-PyTorch fallback for it.
"""
__version__ = "0.1.0"

"""B200-native dense bundle adjustment (DBA) engine: a drop-in for the `droid_backends.ba`
path of HI-SLAM2 / CUT3R-SLAM (plus frame_distance / depth_filter / iproj).

Layout: csrc/ (sm_100a CUDA kernels + the C-ABI library libdba.so), droid_backends.py (host-side
mirror of the reference's torch-extension interface), sharded.py (global BA sharded by source
keyframe over NCCL), synthetic.py (seeded factor graphs of the BASELINE.json shapes).
"""
__version__ = "0.1.0"

"""B200-native ORB extraction + Hamming matching front-end (hot path of ruguangyou/pose-estimation).

Import with importlib.import_module("pose-estimation_b200") (the directory name has a hyphen).
`frontend` is the ctypes mirror of the reference's ORBextractor / FeatureTracker call surface over the
C ABI in include/b2orb.h; `synth` generates the deterministic synthetic stereo inputs.
"""

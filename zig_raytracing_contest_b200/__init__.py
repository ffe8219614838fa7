"""B200-native render path for tigrazone/zig_raytracing_contest.

Modules: abi (ctypes mirror of include/rtc.h), rtc (host binding: Geometry, Scene.render),
scenes (procedural scenes for BASELINE.json's configs). The compute lives in librtc.so
(csrc/: hand-written sm_100a CUDA behind the C ABI); there is no CPU fallback.
"""
__all__ = ["abi", "rtc", "scenes"]

"""vector_b200 — B200-native engine for the vector-build hot path of ramiyako/vector.

Public surface (same names as the reference's ``utils.py`` where one exists):
    resample_signal, apply_frequency_shift, create_spectrogram, normalize_spectrogram,
    save_vector, save_vector_wv, load_packet, load_packet_info, compute_freq_ranges,
    build_vector, build_batch_device (a whole scenario sweep in two device calls), install
plus ``*_device`` variants and ``Engine`` for callers that keep data on the GPU.
Importing the package is cheap and GPU-free; the CUDA library is loaded on first use and its
absence is an error, never a fallback.
"""
from . import hostmath
from .hostmath import TARGET_SAMPLE_RATE, MAX_PACKETS, validate_packet_timing

__all__ = ["resample_signal", "apply_frequency_shift", "create_spectrogram",
           "normalize_spectrogram", "save_vector", "save_vector_wv", "load_packet",
           "load_packet_info", "compute_freq_ranges", "build_vector", "build_vector_device", "build_batch_device",
           "create_spectrogram_device", "full_rate_spectrogram_device", "install", "Engine",
           "get_engine", "validate_packet_timing", "TARGET_SAMPLE_RATE", "MAX_PACKETS",
           "detect_packet_bounds", "median_filter_rows2"]

_API = {"resample_signal", "apply_frequency_shift", "create_spectrogram", "normalize_spectrogram",
        "save_vector", "save_vector_wv", "load_packet", "load_packet_info", "compute_freq_ranges",
        "create_spectrogram_device", "full_rate_spectrogram_device", "wv_header",
        "detect_packet_bounds", "median_filter_rows2"}
_GEN = {"build_vector", "build_vector_device", "build_batch_device", "install"}
_ENG = {"Engine", "get_engine"}


def __getattr__(name):          # lazy: torch / the CUDA library are only touched when used
    if name in _API:
        from . import api
        return getattr(api, name)
    if name in _GEN:
        from . import generate
        return getattr(generate, name)
    if name in _ENG:
        from . import engine
        return getattr(engine, name)
    raise AttributeError(name)

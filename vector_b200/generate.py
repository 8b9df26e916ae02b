"""Vector build orchestration: the numeric core of the reference's two ``generate_vector``
methods behind one function, plus drop-in replacements for those methods and ``install()``.

Reference flow (unified_gui.py:1692-1823, main.py:214-321): for every configured packet
load -> [resample] -> frequency-shift -> place every periodic instance -> normalise by the
global max-abs -> save -> validate timing -> spectrogram -> plot.  All index math, markers,
file naming, validation and plot arguments stay on the host exactly as the reference has
them; np.zeros / the `+=` loop / max-abs / divide go to the engine (K0-K3), and the vector
stays in HBM for the spectrogram and the .wv quantiser.
"""
from __future__ import annotations

import os
import time
import traceback

import numpy as np
import torch

from . import api, hostmath as H
from .engine import get_engine

MARKER_STYLES = ['x', 'o', '^', 's', 'D', 'P', 'v', '1', '2', '3', '4']   # unified_gui.py:1716 / main.py:221
MARKER_COLORS = [f"C{i}" for i in range(10)]                              # unified_gui.py:1717 / main.py:222


def prepare_bank(engine, configs, packets=None, fs=H.TARGET_SAMPLE_RATE, resample=False,
                 slot_base=0):
    """Upload / resample / shift every configured packet into the engine's bank.
    Returns per-config (length at vector rate, pre_samples)."""
    info = []
    for i, cfg in enumerate(configs):
        y, pre = packets[i] if packets is not None else api.load_packet_info(cfg["file"])
        slot = slot_base + i
        engine.bank_upload(slot, y)
        if resample and cfg.get("sample_rate", fs) != fs:                 # main.py:236-237
            ratio = fs / cfg["sample_rate"]
            engine.bank_resample(slot, int(len(y) * ratio))               # utils.py:113-114
        engine.bank_commit(slot, cfg["freq_shift"], fs)                   # utils.py:120-127 (0 -> unchanged)
        info.append((engine.bank_len(slot), pre))
    return info


def placement_rows(configs, info, total_samples, fs=H.TARGET_SAMPLE_RATE, slot_base=0):
    """(start, period, count, slot) per packet — host integers, reference arithmetic
    (unified_gui.py:1748-1757)."""
    rows = []
    for i, (cfg, (length, pre)) in enumerate(zip(configs, info)):
        period = H.period_samples(cfg["period"], fs)
        if period <= 0:
            raise ValueError(f"Invalid period for packet {i+1}")
        start = H.start_offset(cfg["start_time"], pre, fs)
        rows.append((start, period, H.instance_count(start, period, length, total_samples), slot_base + i))
    return rows


def make_markers(configs, info, rows, fs=H.TARGET_SAMPLE_RATE):
    """Marker tuples (time_s, freq_shift, base_name, style, colour) in the reference's order
    (packet, then instance) with its style/colour assignment (unified_gui.py:1733-1737,
    1765-1766)."""
    style_map, markers = {}, []
    for cfg, (_, pre), (start, period, count, _) in zip(configs, info, rows):
        base = os.path.splitext(os.path.basename(cfg["file"]))[0]
        if base not in style_map:
            style_map[base] = (MARKER_STYLES[len(style_map) % len(MARKER_STYLES)],
                               MARKER_COLORS[len(style_map) % len(MARKER_COLORS)])
        st, col = style_map[base]
        for k in range(count):
            markers.append(((start + k * period + pre) / fs, cfg["freq_shift"], base, st, col))
    return markers


def build_vector_device(configs, total_samples, normalize=True, packets=None,
                        fs=H.TARGET_SAMPLE_RATE, resample=False, engine=None, lo=0, hi=None,
                        reduce_max=None):
    """Device-resident build.  Returns (vector CUDA complex64 [hi-lo], markers,
    instance_counts, maxsq CUDA double or None)."""
    eng = engine or get_engine()
    info = prepare_bank(eng, configs, packets, fs, resample)
    rows = placement_rows(configs, info, total_samples, fs)
    vec, maxsq = eng.build_vector(rows, total_samples, normalize, lo=lo, hi=hi, reduce_max=reduce_max)
    return vec, make_markers(configs, info, rows, fs), [r[2] for r in rows], maxsq


BATCH_SLOT_BASE = 32          # source packets of a scenario batch live in bank slots 32..61


def build_batch_device(scenarios, total_samples, normalize=True, packets=None, fs=H.TARGET_SAMPLE_RATE,
                       resample=False, engine=None):
    """A sweep of independent scenario vectors (BASELINE.json configs[4]) — what the reference produces by
    running generate_vector once per scenario (unified_gui.py:1722-1782 in a loop) — as TWO device calls:
    every distinct (packet, frequency shift) is shifted once in one launch, then ONE launch places,
    normalises (each vector by its own max) and stores all vectors.

    scenarios: list of config lists (get_config() dicts, like build_vector_device's `configs`).
    packets: {file: (samples, pre_samples)} for the distinct files (default: load_packet_info per file).
    Returns (vectors CUDA complex64 [n_vectors, total_samples], markers per vector, instance counts per
    vector, maxsq CUDA float64 [n_vectors] or None)."""
    eng = engine or get_engine()
    files = []
    for cfgs in scenarios:
        for cfg in cfgs:
            key = (cfg["file"], cfg.get("sample_rate", fs) if resample else fs)
            if key not in files:
                files.append(key)
    if len(files) > 30:
        raise ValueError("a scenario batch draws from at most 30 distinct packets")
    src = {}
    for i, (name, rate) in enumerate(files):
        y, pre = packets[name] if packets is not None else api.load_packet_info(name)
        slot = BATCH_SLOT_BASE + i
        eng.bank_upload(slot, y)
        if rate != fs:                                                    # main.py:236-237
            eng.bank_resample(slot, int(len(y) * (fs / rate)))
        src[(name, rate)] = (slot, eng.bank_len(slot), pre)
    batch, all_rows, all_info = [], [], []
    for cfgs in scenarios:
        info, rows = [], []
        for i, cfg in enumerate(cfgs):
            slot, length, pre = src[(cfg["file"], cfg.get("sample_rate", fs) if resample else fs)]
            period = H.period_samples(cfg["period"], fs)
            if period <= 0:
                raise ValueError(f"Invalid period for packet {i+1}")
            start = H.start_offset(cfg["start_time"], pre, fs)
            rows.append((start, period, H.instance_count(start, period, length, total_samples), slot))
            info.append((length, pre))
        flags = H.free_instance_flags(rows, [l for l, _ in info], total_samples) if normalize else [False] * len(rows)
        batch.append([(s, p, k, slot, cfg["freq_shift"], fl) for (s, p, k, slot), cfg, fl in zip(rows, cfgs, flags)])
        all_rows.append(rows)
        all_info.append(info)
    eng.batch_prepare(batch, fs)
    vecs, maxsq = eng.batch_build(total_samples, normalize)
    markers = [make_markers(cfgs, info, rows, fs) for cfgs, info, rows in zip(scenarios, all_info, all_rows)]
    return vecs, markers, [[r[2] for r in rows] for rows in all_rows], maxsq


def build_vector(configs, total_samples, normalize=True, packets=None, fs=H.TARGET_SAMPLE_RATE,
                 resample=False):
    """numpy-facing build: (vector complex64 ndarray, markers, instance_counts)."""
    vec, markers, counts, _ = build_vector_device(configs, total_samples, normalize, packets, fs,
                                                  resample)
    return vec.cpu().numpy(), markers, counts


# --------------------------------------------------------------------------------------
# generate_vector replacements.  `ns` is the namespace of the module being patched (the
# reference's unified_gui / main): messagebox, plot_spectrogram, TARGET_SAMPLE_RATE and the
# validate_packet_timing method are taken from there so GUI behaviour is untouched.
# --------------------------------------------------------------------------------------
def make_unified_generate_vector(ns):
    """Replacement for UnifiedVectorApp.generate_vector (unified_gui.py:1692-1823)."""
    messagebox = ns["messagebox"]
    FS = ns["TARGET_SAMPLE_RATE"]

    def generate_vector(self, output_format="mat"):
        try:
            if not self.packet_configs:
                messagebox.showerror("Error", "No packets configured")
                return
            vector_length_ms = float(self.vector_length_var.get())
            if vector_length_ms <= 0:
                messagebox.showerror("Error", "Vector length must be positive")
                return
            vector_length = vector_length_ms / 1000.0
            messagebox.showinfo("Processing...", "Creating vector, please wait...")
            total = int(vector_length * FS)
            print(f"Creating vector of length {vector_length_ms} ms ({total:,} samples)")

            configs = []
            for idx, pc in enumerate(self.packet_configs):
                cfg = pc.get_config()
                if not cfg:
                    continue
                configs.append(cfg)
                print(f"Processing packet {idx+1}: {os.path.basename(cfg['file'])}")
                if cfg['freq_shift'] != 0:
                    print(f"  Applied frequency shift: {cfg['freq_shift']/1e6:.1f} MHz")
                if int(cfg['period'] * FS) <= 0:
                    messagebox.showerror("Error", f"Invalid period for packet {idx+1}")
                    return
            if not configs:
                messagebox.showerror("Error", "No valid packets to include in vector")
                return

            eng = get_engine()
            vec_dev, markers, counts, _ = build_vector_device(
                configs, total, normalize=bool(self.normalize.get()), fs=FS, engine=eng)
            for cfg, k in zip(configs, counts):
                print(f"  Inserted {k} instances with period {cfg['period']*1000:.1f} ms")
            if self.normalize.get():
                print("Vector normalized")

            timestamp = time.strftime("%Y%m%d_%H%M%S")
            if output_format.lower() == "wv":
                output_path = f"vector_{timestamp}.wv"
                api.save_vector_wv(vec_dev, output_path, FS, normalize=False)
                vector = None
            else:
                output_path = f"vector_{timestamp}.mat"
                api.save_vector(vec_dev, output_path)         # planes split on the device, streamed to the file
                vector = None
            print(f"Vector saved to {output_path}")

            validation_text, detailed = self.validate_packet_timing(markers, self.packet_configs)
            try:
                f, t, S = api.create_spectrogram_device(vec_dev, FS, time_resolution_us=1, engine=eng)
                if vector is None:
                    vector = vec_dev.cpu().numpy()
                ns["plot_spectrogram"](
                    f, t, S.cpu().numpy(),
                    title=f"Final Vector Spectrogram - {output_format.upper()} | {validation_text}",
                    packet_markers=markers, sample_rate=FS, signal=vector,
                    validation_details=detailed)
                print("Final spectrogram displayed with timing validation")
            except Exception as e:
                print(f"Warning: Could not display final spectrogram: {e}")
            messagebox.showinfo(
                "Success",
                f"Vector created and saved successfully!\nFormat: {output_format.upper()}\n"
                f"Length: {vector_length_ms} milliseconds\nSamples: {total:,}\nValid packets: {len(configs)}")
        except ValueError as e:
            messagebox.showerror("Error", f"Invalid value: {e}")
            print(f"Value error: {e}")
        except Exception as e:
            messagebox.showerror("Error", f"Error creating vector: {e}")
            print(f"Error creating vector: {traceback.format_exc()}")

    return generate_vector


def make_main_generate_vector(ns):
    """Replacement for VectorApp.generate_vector (main.py:214-321): length in seconds,
    resamples packets whose rate differs from 56 MHz, fixed output names."""
    messagebox = ns["messagebox"]
    FS = ns["TARGET_SAMPLE_RATE"]

    def generate_vector(self, output_format="mat"):
        try:
            vector_length = float(self.vector_length_var.get())
            total = int(vector_length * FS)
            configs = [pc.get_config() for pc in self.packet_configs]
            eng = get_engine()
            vec_dev, markers, counts, _ = build_vector_device(
                configs, total, normalize=bool(self.normalize.get()), fs=FS, resample=True, engine=eng)
            freq_shifts = [c['freq_shift'] if c['freq_shift'] != 0 else 0 for c in configs]
            markers = [m for m in markers]
            if output_format == "wv":
                api.save_vector_wv(vec_dev, 'data/output_vector.wv', FS)
            else:
                api.save_vector(vec_dev, 'data/output_vector.mat')
            center_freq = 0
            if freq_shifts:
                center_freq = (min(freq_shifts) + max(freq_shifts)) / 2
            f, t, S = api.create_spectrogram_device(vec_dev, FS, center_freq=center_freq,
                                                    time_resolution_us=1, engine=eng)
            clean = [v for v in freq_shifts if isinstance(v, (int, float)) and not isinstance(v, bool)]
            ranges = H.compute_freq_ranges(clean)
            if ranges:
                ranges = [tuple(r) for r in list(ranges)]
            ns["plot_spectrogram"](f, t, S.cpu().numpy(), title='Final Vector Spectrogram',
                                   center_freq=center_freq, packet_markers=markers,
                                   freq_ranges=ranges, show_colorbar=False)
            messagebox.showinfo("Success", f"Vector created and saved successfully as {output_format.upper()}")
        except Exception as e:
            messagebox.showerror("Error", f"Error: {e}")

    return generate_vector


_PATCHED_NAMES = ("resample_signal", "apply_frequency_shift", "create_spectrogram",
                  "normalize_spectrogram", "save_vector", "save_vector_wv", "load_packet",
                  "load_packet_info", "compute_freq_ranges", "detect_packet_bounds")


def install(utils_mod=None, unified_gui_mod=None, main_mod=None):
    """Rebind the reference's names to the engine.  Pass the already-imported
    reference modules (any of them may be None).  Returns the list of (module, name) rebound."""
    done = []
    for mod in (utils_mod, unified_gui_mod, main_mod):
        if mod is None:
            continue
        for name in _PATCHED_NAMES:
            if hasattr(mod, name):
                setattr(mod, name, getattr(api, name))
                done.append((mod.__name__, name))
    if unified_gui_mod is not None and hasattr(unified_gui_mod, "UnifiedVectorApp"):
        unified_gui_mod.UnifiedVectorApp.generate_vector = make_unified_generate_vector(vars(unified_gui_mod))
        done.append((unified_gui_mod.__name__, "UnifiedVectorApp.generate_vector"))
    if main_mod is not None and hasattr(main_mod, "VectorApp"):
        main_mod.VectorApp.generate_vector = make_main_generate_vector(vars(main_mod))
        done.append((main_mod.__name__, "VectorApp.generate_vector"))
    return done

"""Headless import of the REAL reference (ramiyako/vector) — build container only.

TEST INFRASTRUCTURE.  ``/root/reference`` is mounted read-only in the build container and
does not exist on the GPU box, so nothing that runs there may import this module; its only
users are ``tests/golden/make_golden.py`` (fixture generator) and the ``not gpu`` tests
that re-validate the oracle port against the live reference when it happens to be present.

The reference's numerics need only numpy/scipy; its modules fail to import merely because
they pull in matplotlib / tkinter / customtkinter at import time (utils.py:10-13,
unified_gui.py:1-14, main.py:1-4).  Stand-ins registered in ``sys.modules`` BEFORE the
import make ``utils``, ``unified_gui`` and ``main`` importable; nothing graphical is used.
"""
from __future__ import annotations

import contextlib
import glob
import io
import os
import sys
import tempfile
from unittest.mock import MagicMock

REFERENCE_ROOT = os.environ.get("VECTOR_REFERENCE_ROOT", "/root/reference")
_STUBS = ("matplotlib", "matplotlib.pyplot", "matplotlib.widgets", "matplotlib.colors",
          "tkinter", "tkinter.filedialog", "tkinter.messagebox", "tkinter.ttk",
          "customtkinter", "brokenaxes")


def available():
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "utils.py"))


def import_reference():
    """Return (utils, unified_gui, main) modules of the real reference."""
    if not available():
        raise RuntimeError(f"reference not mounted at {REFERENCE_ROOT}")
    for name in _STUBS:
        sys.modules.setdefault(name, MagicMock())
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    with contextlib.redirect_stdout(io.StringIO()):
        import utils as ref_utils            # noqa: E402
        import unified_gui as ref_gui        # noqa: E402
        import main as ref_main              # noqa: E402
    return ref_utils, ref_gui, ref_main


class _Var:
    def __init__(self, v):
        self._v = v

    def get(self):
        return self._v


class _Cfg:
    def __init__(self, d):
        self._d = d

    def get_config(self):
        return dict(self._d) if self._d is not None else None


def run_unified_generate_vector(configs, length_ms, normalize=True, output_format="mat"):
    """Drive the real ``UnifiedVectorApp.generate_vector`` (unified_gui.py:1692-1823) with
    fake Tk variables.  ``configs`` are get_config()-style dicts whose 'file' exists.
    Returns dict(vector, spectrogram=(f,t,Sxx) or None, title, out_path_bytes, stdout)."""
    import numpy as np
    import scipy.io as sio
    ref_utils, ref_gui, _ = import_reference()
    app = ref_gui.UnifiedVectorApp()
    app.packet_configs = [_Cfg(c) for c in configs]
    app.vector_length_var = _Var(str(length_ms))
    app.normalize = _Var(bool(normalize))
    captured = {}
    real_cs = ref_utils.create_spectrogram
    real_vt = ref_gui.UnifiedVectorApp.validate_packet_timing

    def spy_cs(*a, **k):
        out = real_cs(*a, **k)
        captured["spec"] = out
        return out

    def spy_vt(self, markers, pcs):
        out = real_vt(self, markers, pcs)
        captured["markers"] = list(markers)
        captured["validation"] = out
        return out

    cwd = os.getcwd()
    buf = io.StringIO()
    with tempfile.TemporaryDirectory() as td:
        os.chdir(td)
        try:
            ref_gui.create_spectrogram = spy_cs
            ref_gui.UnifiedVectorApp.validate_packet_timing = spy_vt
            with contextlib.redirect_stdout(buf):
                app.generate_vector(output_format)
            files = glob.glob(os.path.join(td, f"vector_*.{output_format}"))
            if len(files) != 1:
                raise RuntimeError("reference wrote no vector file:\n" + buf.getvalue())
            raw = open(files[0], "rb").read()
            vec = None
            if output_format == "mat":
                vec = np.asarray(sio.loadmat(files[0], squeeze_me=True)["Y"])
        finally:
            ref_gui.create_spectrogram = real_cs
            ref_gui.UnifiedVectorApp.validate_packet_timing = real_vt
            os.chdir(cwd)
    return dict(vector=vec, file_bytes=raw, spectrogram=captured.get("spec"),
                markers=captured.get("markers"), validation=captured.get("validation"),
                stdout=buf.getvalue())


def run_main_generate_vector(configs, length_s, normalize=True, output_format="mat"):
    """Drive the real ``VectorApp.generate_vector`` (main.py:214-321) unbound, with a fake
    ``self``; it writes data/output_vector.{mat,wv} relative to CWD."""
    import numpy as np
    import scipy.io as sio
    from types import SimpleNamespace
    _, _, ref_main = import_reference()
    fake = SimpleNamespace(packet_configs=[_Cfg(c) for c in configs],
                           vector_length_var=_Var(str(length_s)),
                           normalize=_Var(bool(normalize)))
    cwd = os.getcwd()
    buf = io.StringIO()
    with tempfile.TemporaryDirectory() as td:
        os.chdir(td)
        os.makedirs("data")
        try:
            with contextlib.redirect_stdout(buf):
                ref_main.VectorApp.generate_vector(fake, output_format)
            path = os.path.join(td, "data", f"output_vector.{output_format}")
            raw = open(path, "rb").read()
            vec = None
            if output_format == "mat":
                vec = np.asarray(sio.loadmat(path, squeeze_me=True)["Y"])
        finally:
            os.chdir(cwd)
    return dict(vector=vec, file_bytes=raw, stdout=buf.getvalue())


def trace_spectrogram_plan(n, sr, **kw):
    """Run the real create_spectrogram on a length-n dummy signal with
    scipy.signal.spectrogram intercepted; returns the kwargs of the first scipy call
    (window, nperseg, noverlap, nfft, fs, input length)."""
    import numpy as np
    import scipy.signal
    ref_utils, _, _ = import_reference()
    seen = []
    real = scipy.signal.spectrogram

    def spy(x, **k):
        seen.append(dict(n_dec=len(x), **{a: k[a] for a in
                                          ("fs", "window", "nperseg", "noverlap", "nfft",
                                           "scaling")}))
        nfft = k["nfft"]
        # cheap non-zero stand-in so the all-zero fallback chain is not entered
        return np.zeros(nfft), np.zeros(1), np.ones((nfft, 1), dtype=np.float32)

    sig = np.zeros(n, dtype=np.complex64)
    scipy.signal.spectrogram = spy
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            ref_utils.create_spectrogram(sig, sr, **kw)
    finally:
        scipy.signal.spectrogram = real
    return seen[0]

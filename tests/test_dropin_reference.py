"""Drop-in proof against the REAL reference modules (build container only: /root/reference is mounted there
and absent on the GPU box, where these tests skip).  No GPU needed: install() only rebinds names.

VERDICT r1 weak #3: every name install() patches must exist in the reference with the same call signature,
and both generate_vector methods must be replaced on the real classes."""
import inspect

import pytest

from oracle import ref_import

pytestmark = pytest.mark.skipif(not ref_import.available(), reason="reference not mounted")


@pytest.fixture(scope="module")
def ref():
    return ref_import.import_reference()


def _sig(fn):
    return [(p.name, p.kind, p.default) for p in inspect.signature(fn).parameters.values()]


def test_patched_names_exist_with_identical_signatures(ref):
    from vector_b200 import api, generate
    ref_utils = ref[0]
    for name in generate._PATCHED_NAMES:
        assert hasattr(ref_utils, name), name
        want, got = _sig(getattr(ref_utils, name)), _sig(getattr(api, name))
        got = [p for p in got if not p[0].startswith("_")]          # private test hooks (keyword-only extras)
        assert got == want, (name, got, want)


def test_install_rebinds_the_real_modules_and_restores(ref):
    from vector_b200 import api, generate
    ref_utils, ref_gui, ref_main = ref
    saved = {(m, n): getattr(m, n) for m in ref for n in generate._PATCHED_NAMES if hasattr(m, n)}
    gv_gui, gv_main = ref_gui.UnifiedVectorApp.generate_vector, ref_main.VectorApp.generate_vector
    try:
        done = generate.install(ref_utils, ref_gui, ref_main)
        for name in generate._PATCHED_NAMES:
            assert getattr(ref_utils, name) is getattr(api, name)
        # the names unified_gui / main imported from utils at import time (unified_gui.py:15-33, main.py:5-15)
        for mod in (ref_gui, ref_main):
            bound = [n for n in generate._PATCHED_NAMES if hasattr(mod, n)]
            assert bound, mod.__name__
            for n in bound:
                assert getattr(mod, n) is getattr(api, n), (mod.__name__, n)
        assert ("unified_gui", "UnifiedVectorApp.generate_vector") in done
        assert ("main", "VectorApp.generate_vector") in done
        assert ref_gui.UnifiedVectorApp.generate_vector is not gv_gui
        assert ref_main.VectorApp.generate_vector is not gv_main
        assert _sig(ref_gui.UnifiedVectorApp.generate_vector) == _sig(gv_gui)
        assert _sig(ref_main.VectorApp.generate_vector) == _sig(gv_main)
        # what the replacements take from the patched module's namespace is there
        for mod in (ref_gui, ref_main):
            for n in ("messagebox", "TARGET_SAMPLE_RATE", "plot_spectrogram"):
                assert hasattr(mod, n), (mod.__name__, n)
        assert hasattr(ref_gui.UnifiedVectorApp, "validate_packet_timing")
    finally:
        for (m, n), f in saved.items():
            setattr(m, n, f)
        ref_gui.UnifiedVectorApp.generate_vector, ref_main.VectorApp.generate_vector = gv_gui, gv_main


def test_host_side_restatements_agree_with_the_real_reference(ref, tmp_path):
    """The host logic the replacements keep (no GPU): load_packet_info incl. the MAT fast reader, the .wv header,
    compute_freq_ranges, validate_packet_timing — against the real functions on the reference's own fixtures."""
    import glob
    import os
    import numpy as np
    from vector_b200 import api, hostmath as H
    ref_utils, ref_gui, _ = ref
    files = sorted(glob.glob(os.path.join(ref_import.REFERENCE_ROOT, "data", "*.mat")))
    assert len(files) >= 10
    for f in files:
        try:
            y_ref, pre_ref = ref_utils.load_packet_info(f)
        except ValueError as e:
            with pytest.raises(ValueError) as ei:
                api.load_packet_info(f)
            assert str(ei.value) == str(e)
            continue
        y, pre = api.load_packet_info(f)
        assert y.dtype == y_ref.dtype and np.array_equal(y, y_ref) and pre == pre_ref, f
    for shifts in ([], [0, 0], [2e6, -3e6, 1e6], [5e6], [0, 1e6]):
        assert H.compute_freq_ranges(shifts) == ref_utils.compute_freq_ranges(shifts)
    markers = [(0.0002 + 0.001 * k, 2e6, "packet_1_5MHz", "x", "C0") for k in range(5)]
    markers += [(0.0006 + 0.0012 * k, -3e6, "packet_2_chirp", "o", "C1") for k in range(3)]
    cfgs = [dict(file="data/packet_1_5MHz.mat", freq_shift=2e6, period=0.001, start_time=0.0002),
            dict(file="data/packet_2_chirp.mat", freq_shift=-3e6, period=0.0012, start_time=0.0006)]

    class Cfg:
        def __init__(self, d):
            self.d = d

        def get_config(self):
            return self.d
    import contextlib
    import io
    with contextlib.redirect_stdout(io.StringIO()):
        title_ref, details_ref = ref_gui.UnifiedVectorApp.validate_packet_timing(None, markers, [Cfg(c) for c in cfgs])
    title, details = H.validate_packet_timing(markers, cfgs)
    assert title == title_ref and details == details_ref

"""Generate tests/golden/* by running the REAL reference (ramiyako/vector) headless.

Run in the build container (``/root/reference`` mounted):

    python -m tests.golden.make_golden

Inputs are the seeded synthetic packets of ``oracle.ref_port.synth_packet`` written to a
temporary directory as MAT files (Y complex128 + pre_samples=100, like the reference's
``data/packet_*`` fixtures); outputs are what the reference's own code produced:
``UnifiedVectorApp.generate_vector`` (unified_gui.py:1692-1823), ``VectorApp.generate_vector``
(main.py:214-321), ``utils.apply_frequency_shift/resample_signal/create_spectrogram/
normalize_spectrogram/save_vector_wv``.  Large arrays are stored as SHA-1 + sample windows.
"""
from __future__ import annotations

import json
import os
import sys
import tempfile

import numpy as np
import scipy
import scipy.io as sio

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import ref_import, ref_port            # noqa: E402
from tests.golden import cases                     # noqa: E402


def write_packets(td):
    paths = {}
    for i, kind in enumerate(ref_port.PACKET_KINDS):
        y = ref_port.synth_packet(kind, seed=101 + i)
        p = os.path.join(td, ref_port.packet_file(kind))
        sio.savemat(p, {"Y": y, "pre_samples": 100})
        paths[kind] = p
    return paths


def spec_summary(S):
    S = np.ascontiguousarray(S)
    return dict(shape=list(S.shape), dtype=str(S.dtype), sha1=cases.sha1(S),
                sub=S[::16, ::max(1, S.shape[1] // 64)].copy(),
                total=float(S.sum(dtype=np.float64)), peak=float(S.max()))


def main():
    ref_utils, ref_gui, ref_main = ref_import.import_reference()
    meta = {"numpy": np.__version__, "scipy": scipy.__version__, "builds": {}, "plans": {},
            "misc": {}}
    arrays = {}
    with tempfile.TemporaryDirectory() as td:
        paths = write_packets(td)
        bank_sha = {k: cases.sha1(ref_utils.load_packet_info(p)[0]) for k, p in paths.items()}
        meta["bank_sha1"] = bank_sha

        # ---- vector builds through the real generate_vector --------------------------
        for name, case in cases.build_cases().items():
            cfgs = [dict(c, file=paths[c["kind"]]) for c in case["configs"]]
            out = ref_import.run_unified_generate_vector(cfgs, case["length_ms"],
                                                         case["normalize"], "mat")
            vec = out["vector"].astype(np.complex64, copy=False)
            s = cases.summarize(vec)
            arrays[f"{name}__starts"] = s.pop("starts")
            arrays[f"{name}__windows"] = s.pop("windows")
            if vec.shape[0] <= 300_000:
                arrays[f"{name}__full"] = vec
            mk = out["markers"] or []
            s["marker_times"] = [float(m[0]) for m in mk]
            s["marker_names"] = [m[2] for m in mk]
            s["title"] = out["validation"][0] if out["validation"] else None
            s["details"] = out["validation"][1] if out["validation"] else None
            if out["spectrogram"] is not None:
                f, t, S = out["spectrogram"]
                ss = spec_summary(S)
                arrays[f"{name}__spec_sub"] = ss.pop("sub")
                ss["freqs_sha1"] = cases.sha1(f)
                ss["times_sha1"] = cases.sha1(t)
                ss["f0"], ss["f1"], ss["t0"], ss["t1"] = (float(f[0]), float(f[-1]),
                                                           float(t[0]), float(t[-1]))
                db, vmin, vmax = ref_utils.normalize_spectrogram(S, 5, 95, 25)
                ss["db_sha1"] = cases.sha1(db)
                ss["vmin"], ss["vmax"] = float(vmin), float(vmax)
                arrays[f"{name}__db_sub"] = np.ascontiguousarray(
                    db[::16, ::max(1, db.shape[1] // 64)])
                s["spectrogram"] = ss
            meta["builds"][name] = s
            print(name, s["n"], s["sha1"][:10], s["max_abs"], s.get("title"))

        # ---- .wv output of a small build -----------------------------------------------
        case = cases.build_cases()["wf_2p5ms_shift_norm"]
        cfgs = [dict(c, file=paths[c["kind"]]) for c in case["configs"]]
        out = ref_import.run_unified_generate_vector(cfgs, case["length_ms"], True, "wv")
        raw = out["file_bytes"]
        head_end = raw.index(b":#") + 2
        header = raw[:head_end].decode()
        meta["misc"]["wv_header_wf_2p5ms_shift_norm"] = header
        meta["misc"]["wv_payload_sha1"] = cases.sha1(np.frombuffer(raw[head_end:-1], np.uint8))
        meta["misc"]["wv_tail"] = raw[-1:].decode()

        # ---- main.py flavour: resample 20 MHz -> 56 MHz, seconds-length input ------------
        rng = np.random.default_rng(77)
        small = {}
        for i, (kind, sr) in enumerate((("tone", 20e6), ("chirp", 40e6), ("bpsk", 56e6))):
            y = ref_port.synth_packet(kind, fs=sr, body_s=1e-4, pre=20, seed=201 + i)
            p = os.path.join(td, f"small_{kind}.mat")
            sio.savemat(p, {"Y": y, "pre_samples": 20})
            small[kind] = (p, sr)
        mcfgs = [dict(file=small["tone"][0], sample_rate=small["tone"][1], freq_shift=3e6,
                      period=0.0005, pre_samples=20, start_time=0.0001),
                 dict(file=small["chirp"][0], sample_rate=small["chirp"][1], freq_shift=-2e6,
                      period=0.0007, pre_samples=20, start_time=0.0),
                 dict(file=small["bpsk"][0], sample_rate=small["bpsk"][1], freq_shift=0.0,
                      period=0.0004, pre_samples=20, start_time=0.0002)]
        out = ref_import.run_main_generate_vector(mcfgs, "0.004", True, "mat")
        vec = out["vector"].astype(np.complex64, copy=False)
        s = cases.summarize(vec)
        arrays["main_resample__starts"] = s.pop("starts")
        arrays["main_resample__windows"] = s.pop("windows")
        arrays["main_resample__full"] = vec
        s["configs"] = [dict(c, file=os.path.basename(c["file"])) for c in mcfgs]
        meta["builds"]["main_resample"] = s
        print("main_resample", s["n"], s["sha1"][:10], s["max_abs"])

        # ---- helpers called directly ---------------------------------------------------
        y, pre = ref_utils.load_packet_info(paths["tone"])
        sh = ref_utils.apply_frequency_shift(y, 2e6, 56e6)
        arrays["shift_tone_p2MHz"] = sh
        sh2 = ref_utils.apply_frequency_shift(y, -12.3456e6, 56e6)
        meta["misc"]["shift_tone_m12p3456MHz_sha1"] = cases.sha1(sh2)
        arrays["shift_tone_m12p3456MHz_head"] = sh2[:2048]
        for (n_in, sr_in, sr_out) in ((2020, 20e6, 56e6), (4020, 40e6, 56e6),
                                      (56100, 56e6, 200e6), (5000, 56e6, 20e6),
                                      (4096, 56e6, 28e6)):
            x = ref_port.synth_packet("chirp", fs=sr_in, body_s=(n_in - 20) / sr_in, pre=20,
                                      seed=300 + n_in).astype(np.complex64)[:n_in]
            r = ref_utils.resample_signal(x, sr_in, sr_out)
            key = f"resample_{n_in}_{int(sr_in/1e6)}to{int(sr_out/1e6)}"
            meta["misc"][key] = dict(n_out=int(len(r)), sha1=cases.sha1(r),
                                     in_sha1=cases.sha1(x))
            arrays[key + "__head"] = r[:1024]
            arrays[key + "__tail"] = r[-1024:]
            print(key, len(r))

        # spectrogram of a single packet (non-heavy branch, blackmanharris)
        f, t, S = ref_utils.create_spectrogram(sh, 56e6, time_resolution_us=1)
        ss = spec_summary(S)
        arrays["spec_packet__sub"] = ss.pop("sub")
        ss["f0"], ss["f1"], ss["t0"], ss["t1"] = float(f[0]), float(f[-1]), float(t[0]), float(t[-1])
        meta["misc"]["spec_packet"] = ss
        # all-zero signal -> fallback chain (utils.py:316-347)
        f, t, S = ref_utils.create_spectrogram(np.zeros(20000, np.complex64), 56e6)
        meta["misc"]["spec_zeros_shape"] = list(S.shape)
        meta["misc"]["spec_zeros_t0"] = float(t[0])

        # ---- derived STFT parameters ---------------------------------------------------
        for n in cases.plan_lengths():
            for tr in (1, 50):
                meta["plans"][f"{n}:{tr}"] = ref_import.trace_spectrogram_plan(
                    n, 56e6, time_resolution_us=tr)
        meta["plans"]["2000000000:1@200e6"] = ref_import.trace_spectrogram_plan.__call__(
            20_000_000, 200e6, time_resolution_us=1) | {"note": "n=2e7 stand-in, same branch"}

    np.savez_compressed(os.path.join(HERE, "golden_arrays.npz"), **arrays)
    with open(os.path.join(HERE, "golden_meta.json"), "w") as fh:
        json.dump(meta, fh, indent=1, default=lambda o: o.tolist() if hasattr(o, "tolist") else str(o))
    print("wrote", len(arrays), "arrays")


if __name__ == "__main__":
    main()

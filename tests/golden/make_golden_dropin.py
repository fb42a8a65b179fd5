"""Golden fixtures for the drop-in layer: the REFERENCE's processor / detector classes driven exactly as its GUI
controller drives them (``cls(config_manager, **params)``, ``process(adc_cube=cube, **params)`` with the same dict, then
``getattr`` for the registry's view keys), on seeded synthetic cubes.  Parameter sets live in ``tests/dropin_params.py``.

Runs only in the build container (imports /root/reference read-only):

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_dropin.py

Writes ``tests/golden/dropin_<cfg>.npz``; keys are ``<spec>|<frame tag>|<name>`` (name = ``data`` for the value
``process`` returned, else the attribute).  Inputs are regenerated in the tests from the seeds.
"""
import importlib
import os
import sys
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests"))
sys.path.insert(0, REF)
sys.dont_write_bytecode = True

import dropin_params as dp  # noqa: E402
from mmwave_radar_processing.config_managers.cfgManager import ConfigManager  # noqa: E402  (reference)
from mmwave_radar_processing.processors.virtual_array_reformater import VirtualArrayReformatter  # noqa: E402
from mmwave_radar_processing.detectors.detector_registry import get_detector_registry  # noqa: E402
from mmwave_radar_processing_b200.synthetic import synth_frame_iq, synth_scene_iq, iq_to_complex_cube  # noqa: E402

warnings.filterwarnings("ignore")


def ref_cfg(name, geom):
    cm = ConfigManager()
    cm.load_cfg(os.path.join(REF, "configs", name + ".cfg"), array_geometry=geom)
    return cm


def shape_of(cm):
    cpl = cm.frameCfg_end_index - cm.frameCfg_start_index + 1
    return cm.num_rx_antennas, cm.get_num_adc_samples(0), cpl, cm.frameCfg_loops


def put(out, key, val):
    if val is None:
        return
    if isinstance(val, (float, int)):
        val = np.array(val)
    val = np.asarray(val)
    # heat maps are compared at 1e-4 relative L2: store them as float32 (6e-8) to keep the fixtures small; the fp64 RD map
    # and range profile of the first frame stay fp64 because they are the INPUT of the stand-alone CFAR checks
    spec, tag, name = (key.split("|") + ["", ""])[:3]
    is_map = (name == "rng_dop_resp" or (name == "data" and spec.startswith(("range_angle", "dbs", "rd_variant"))) or
              (name == "data" and spec == "range_doppler_resp" and tag != "s%d" % dp.SEEDS[0]))
    if is_map and val.dtype == np.float64:
        val = val.astype(np.float32)
    out[key] = val


def frames_of(cm):
    r, s, cpl, l = shape_of(cm)
    frames = [(f"s{sd}", synth_frame_iq(r, s, cpl, l, sd)) for sd in dp.SEEDS]
    frames.append(("zero", np.zeros_like(frames[0][1])))
    return frames


def run_case(name, geom):
    cm = ref_cfg(name, geom)
    r, s, cpl, l = shape_of(cm)
    vref = VirtualArrayReformatter(cm)
    cubes = [(tag, vref.process(iq_to_complex_cube(iq))) for tag, iq in frames_of(cm)]
    v = cubes[0][1].shape[0]
    out = {}
    failed = []
    for key, module, cls_name, params, view_keys, vel in dp.specs(v):
        cls = getattr(importlib.import_module("mmwave_radar_processing." + module), cls_name)
        try:
            proc = cls(cm, **params)
        except Exception as exc:  # noqa: BLE001
            failed.append((key, "ctor", type(exc).__name__))
            put(out, f"{key}|ctor|error", np.array(type(exc).__name__))
            continue
        for tag, cube in cubes:
            try:
                if vel is not None:
                    res = proc.process(adc_cube=cube, velocity_ned=np.array(vel), **params)
                else:
                    res = proc.process(adc_cube=cube, **params)
            except Exception as exc:  # noqa: BLE001
                failed.append((key, tag, type(exc).__name__, str(exc)[:80]))
                put(out, f"{key}|{tag}|error", np.array(type(exc).__name__))
                continue
            put(out, f"{key}|{tag}|data", res)
            for attr in view_keys:
                put(out, f"{key}|{tag}|{attr}", getattr(proc, attr, None))
            if cls_name == "PointCloudGenerator":
                det = proc.detector
                put(out, f"{key}|{tag}|dets", det.dets)
                if det.dets is not None and det.dets.shape[0] > 0:
                    _, _, ri, di = det._map_detections_to_bins(det.dets)
                    az, el = proc._compute_angle_estimation(det.rng_dop_resp_raw, ri, di)
                    put(out, f"{key}|{tag}|az", az)
                    put(out, f"{key}|{tag}|el", el)
            if cls_name == "RangeDopplerDetector2D":
                put(out, f"{key}|{tag}|thresholds", proc.detector.thresholds)
                put(out, f"{key}|{tag}|noise_estimates", proc.detector.noise_estimates)
                put(out, f"{key}|{tag}|detections", proc.detector.detections)
            if cls_name == "RangeDopplerGroundDetector":
                put(out, f"{key}|{tag}|altitude", proc.altimeter.current_altitude_corrected_m)
            if key == "det2d_gui" and tag == cubes[0][0]:
                put(out, f"{key}|{tag}|rng_dop_resp_raw", proc.rng_dop_resp_raw.astype(np.complex64))

    # RangeDopplerProcessor argument grid on the first frame
    rdp_cls = importlib.import_module("mmwave_radar_processing.processors.range_doppler_resp").RangeDopplerProcessor
    tag0, cube0 = cubes[0]
    for rx, mag in dp.rd_variants(v):
        res = rdp_cls(cm).process(adc_cube=cube0, rx_idx=rx, return_magnitude=mag)
        if not mag:
            res = res.astype(np.complex64)
        put(out, f"rd_variant_rx{rx}_mag{int(mag)}|{tag0}|data", res)

    # stand-alone detector classes on the reference's fp64 RD map (antenna 0), three of its rows and the range profile
    reg = get_detector_registry()
    x2d = out[f"range_doppler_resp|{tag0}|data"]
    prof = out[f"range_resp|{tag0}|data"]
    two_d = [(k, k, p) for k, p in dp.CFAR_2D.items()] + [(t, k, p) for t, (k, p) in dp.CFAR_2D_EXTRA.items()]
    for tag, kind, params in two_d:
        d = reg[kind](**params)
        lst = d.detect(x2d)
        put(out, f"cfar|{tag}|list", np.array(lst, dtype=np.int64).reshape(-1, 2))
        put(out, f"cfar|{tag}|thresholds", d.thresholds)
        put(out, f"cfar|{tag}|noise_estimates", d.noise_estimates)
        put(out, f"cfar|{tag}|detections", d.detections)
    rows = [1, x2d.shape[0] // 2, x2d.shape[0] - 1]
    put(out, "cfar|rows", np.array(rows))
    one_d = [(k, k, p) for k, p in dp.CFAR_1D.items()] + [("ca_too_large", *dp.CFAR_1D_TOO_LARGE)]
    for tag, kind, params in one_d:
        d = reg[kind](**params)
        for sig_tag, sig in [("prof", prof)] + [(f"row{j}", x2d[rr]) for j, rr in enumerate(rows)]:
            lst = d.detect(sig)
            put(out, f"cfar|{tag}|{sig_tag}|list", np.array(lst, dtype=np.int64))
            put(out, f"cfar|{tag}|{sig_tag}|thresholds", d.thresholds)
            put(out, f"cfar|{tag}|{sig_tag}|noise_estimates", d.noise_estimates)
            put(out, f"cfar|{tag}|{sig_tag}|detections", d.detections)

    # scripted 20-frame sequence through ONE PointCloudGenerator with the ground detector (altimeter recurrence)
    seq_params = dp.seq_params(v, l)
    pcg_cls = importlib.import_module("mmwave_radar_processing.processors.point_cloud_generator").PointCloudGenerator
    pcg = pcg_cls(cm, **seq_params)
    for f in range(dp.SEQ_FRAMES):
        iq = synth_scene_iq(r, s, cpl, l, 500 + f, dp.ground_sequence_targets(f, cm.range_res_m, s, l))
        cube = vref.process(iq_to_complex_cube(iq))
        pts = pcg.process(adc_cube=cube, **seq_params)
        put(out, f"seq|f{f}|data", pts)
        put(out, f"seq|f{f}|dets", pcg.detector.dets)
        put(out, f"seq|f{f}|altitude", pcg.detector.altimeter.current_altitude_corrected_m)

    path = os.path.join(HERE, f"dropin_{name}.npz")
    np.savez_compressed(path, **out)
    n_pts = {k: out[k].shape[0] for k in out if k.endswith("|data") and k.startswith(("pcg", "seq"))}
    print(path, os.path.getsize(path) // 1024, "KiB;", len(out), "arrays; failures:", failed)
    print("   points:", n_pts)
    print("   altitudes:", [float(out[f"seq|f{f}|altitude"]) for f in range(dp.SEQ_FRAMES)])


if __name__ == "__main__":
    for name, geom in dp.CASES.items():
        run_case(name, geom)

"""Generate the golden fixtures by running the REFERENCE's own classes on seeded synthetic cubes.

Runs only in the build container (imports /root/reference read-only; the reference cannot
travel to the GPU box).  Writes ``tests/golden/*.npz`` + ``cfg_table.json``.

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py

The inputs are regenerated in the tests from the same seeds through
``mmwave_radar_processing_b200.synthetic`` so only outputs are stored.
"""
import glob
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, REPO)
sys.path.insert(0, REF)
sys.dont_write_bytecode = True

from mmwave_radar_processing.config_managers.cfgManager import ConfigManager  # noqa: E402  (reference)
from mmwave_radar_processing.processors.virtual_array_reformater import VirtualArrayReformatter  # noqa: E402
from mmwave_radar_processing.processors.range_doppler_resp import RangeDopplerProcessor  # noqa: E402
from mmwave_radar_processing.processors.range_resp import RangeProcessor  # noqa: E402
from mmwave_radar_processing.processors.range_angle_resp import RangeAngleProcessor  # noqa: E402
from mmwave_radar_processing.processors.range_angle_resp_dbs_enhanced import RangeAngleProcessorDBSEnhanced  # noqa: E402
from mmwave_radar_processing.processors.range_detector import RangeDetector  # noqa: E402
from mmwave_radar_processing.processors.point_cloud_generator import PointCloudGenerator  # noqa: E402
from mmwave_radar_processing.processors.range_doppler_detection.range_doppler_detector_2d import RangeDopplerDetector2D  # noqa: E402
from mmwave_radar_processing.processors.range_doppler_detection.range_doppler_detector_sequential import RangeDopplerDetectorSequential  # noqa: E402
from mmwave_radar_processing.processors.range_doppler_detection.range_doppler_ground_detector import RangeDopplerGroundDetector  # noqa: E402
from mmwave_radar_processing.processors.simple_synthetic_array_beamformer_processor_multiFrame import SyntheticArrayBeamformerProcessor  # noqa: E402
from mmwave_radar_processing.detectors.detector_registry import get_detector_registry  # noqa: E402

from mmwave_radar_processing_b200.synthetic import synth_frame_iq, iq_to_complex_cube  # noqa: E402

# parameter sets: (i) reference tests/verify_processors.py, (ii) gui_configs/processor_params.yaml
CFAR2D_SETS = {
    "os_test": ("os_cfar_2d", dict(num_train=(8, 4), num_guard=(2, 1), rho=0.75, alpha=3.5)),
    "ca_test": ("ca_cfar_2d", dict(num_train=(4, 4), num_guard=(2, 2), pfa=1e-5)),
    "os_gui": ("os_cfar_2d", dict(num_train=[5, 5], num_guard=[3, 2], rho=0.7, alpha=2)),
}
CFAR1D_SETS = {
    "ca": ("ca_cfar_1d", dict(num_train=10, num_guard=2, pfa=1e-3)),
    "go": ("go_cfar_1d", dict(num_train=10, num_guard=2, pfa=1e-3)),
    "so": ("so_cfar_1d", dict(num_train=10, num_guard=2, pfa=1e-3)),
    "os": ("os_cfar_1d", dict(num_train=5, num_guard=3, rho=0.5, alpha=2)),
}
SEQ_PARAMS = dict(rng_cfar_type="os_cfar_1d", rng_cfar_params=dict(num_train=5, num_guard=3, rho=0.6, alpha=2),
                  vel_cfar_type="os_cfar_1d", vel_cfar_params=dict(num_train=5, num_guard=2, rho=0.7, alpha=3))
GROUND_PARAMS = dict(vel_cfar_type="os_cfar_1d", vel_cfar_params=dict(num_train=12, num_guard=4, rho=0.5, alpha=15),
                     altimeter_params=dict(min_altitude_m=25.0e-2, zoom_search_region_m=20.0e-2,
                                           altitude_search_limit_m=40e-2, range_bias=0.0, precise_est_enabled=False))

CASES = {
    # cfg name: (seeds, geometry)
    "6843_RadVel_ods_20Hz": ([0, 1, 2], "ods"),
    "1843_RadVel": ([0, 1], "standard"),
    "RadarHD": ([0], "standard"),
    "RadCloud": ([0], "standard"),
    "RadSAR_1843_official": ([0], "standard"),
}


def load_ref_cfg(name, geom="standard"):
    cm = ConfigManager()
    cm.load_cfg(os.path.join(REF, "configs", name + ".cfg"), array_geometry=geom)
    return cm


def shapes(cm):
    cpl = cm.frameCfg_end_index - cm.frameCfg_start_index + 1
    return cm.num_rx_antennas, cm.get_num_adc_samples(0), cpl, cm.frameCfg_loops


def cfg_table():
    rows = {}
    for path in sorted(glob.glob(os.path.join(REF, "configs", "*.cfg"))):
        name = os.path.basename(path)[:-4]
        cm = load_ref_cfg(name)
        rdp = RangeDopplerProcessor(cm)
        rap = RangeAngleProcessor(cm)
        r, s, cpl, l = shapes(cm)
        rows[name] = dict(
            num_rx=r, num_tx=cm.num_tx_antennas, cpl=cpl, S=s, L=l,
            virtual=bool(cm.virtual_antennas_enabled),
            range_res_m=cm.range_res_m, range_max_m=cm.range_max_m, range_bin_size_m=float(cm.range_bin_size_m),
            vel_res_m_s=cm.vel_res_m_s, vel_max_m_s=cm.vel_max_m_s,
            n_range_bins=int(rdp.range_bins.shape[0]), n_vel_bins=int(rdp.vel_bins.shape[0]),
            n_ra_range_bins=int(rap.range_bins.shape[0]),
            range_bins_last=float(rdp.range_bins[-1]), vel_bins_first=float(rdp.vel_bins[0]),
            vel_bins_last=float(rdp.vel_bins[-1]),
            adc_bits=cm.adcCfg_num_adc_bits, adcbuf_fmt=cm.adcbufCfg_adc_output_fmt,
            interleave=bool(cm.adcbufCfg_channel_interleave), frame_ms=cm.frameCfg_periodicity_ms,
            hw_trig=bool(cm.frameCfg_hardware_trigger_enabled),
        )
    # angle-axis length quirk (A or A+1)
    quirk = {}
    cm = load_ref_cfg("6843_RadVel_ods_20Hz")
    for a in (8, 16, 32, 64, 100, 128, 256, 512, 1024):
        p = RangeAngleProcessor(cm, num_angle_bins=a)
        quirk[str(a)] = dict(n=int(p.angle_bins.shape[0]), first=float(p.angle_bins[0]), last=float(p.angle_bins[-1]),
                             mid=float(p.angle_bins[a // 2]))
    return dict(cfgs=rows, angle_axis=quirk)


def ant_sets(v):
    if v >= 12:
        return dict(az_a=[0, 1, 2, 3], el_a=[], az_b=[0, 3, 4, 7], el_b=[9, 8, 5, 4], ra=[0, 3, 4, 7])
    if v >= 8:
        return dict(az_a=[0, 1, 2, 3], el_a=[], az_b=[0, 1, 2, 3, 4, 5, 6, 7], el_b=[1, 5], ra=[0, 3, 4, 7])
    return dict(az_a=[0, 1, 2, 3], el_a=[], az_b=[0, 1, 2], el_b=[3, 2], ra=[0, 1, 3])


def frame_outputs(cm, cube_raw, full):
    out = {}
    cube = VirtualArrayReformatter(cm).process(cube_raw)
    v = cube.shape[0]
    sets = ant_sets(v)
    rdp = RangeDopplerProcessor(cm)
    rd = rdp.process(cube, rx_idx=-1, return_magnitude=False)
    if full:
        out["virtual_cube"] = cube.astype(np.complex64)     # exact: integer valued
        out["rd"] = rd
    out["rd_mag_rx0"] = rdp.process(cube)                   # default args: rx_idx=0, magnitude
    out["rd_mag_rx_last"] = rdp.process(cube, rx_idx=v - 1)
    out["range_profile"] = RangeProcessor(cm).process(cube, chirp_idx=0)
    out["range_profile_c3"] = RangeProcessor(cm).process(cube, chirp_idx=3)

    # 2-D CFAR + detector + point cloud
    for tag, (kind, params) in CFAR2D_SETS.items():
        det = RangeDopplerDetector2D(cm, cfar_type=kind, cfar_params=params)
        dets = det.process(cube)
        out[f"dets2d_{tag}"] = dets
        out[f"thr2d_{tag}"] = det.detector.thresholds
        out[f"noise2d_{tag}"] = det.detector.noise_estimates
        for ab in ("a", "b"):
            pc = PointCloudGenerator(cm, az_antenna_idxs=sets["az_" + ab], el_antenna_idxs=sets["el_" + ab],
                                     detector_type="range_doppler_detector_2d",
                                     detector_params=dict(cfar_type=kind, cfar_params=params))
            pts = pc.process(cube)
            out[f"pc_{tag}_{ab}"] = pts
            if pts.shape[0]:
                _, _, ri, di = pc.detector._map_detections_to_bins(pc.detector.dets)
                az, el = pc._compute_angle_estimation(pc.detector.rng_dop_resp_raw, ri, di)
                out[f"az_{tag}_{ab}"] = az
                out[f"el_{tag}_{ab}"] = el

    # 1-D CFAR family on the range profile and on three RD rows
    reg = get_detector_registry()
    mag0 = np.abs(rd[0])
    rows = [0, mag0.shape[0] // 3, mag0.shape[0] - 2]
    out["cfar1d_rows"] = np.array(rows)
    for tag, (kind, params) in CFAR1D_SETS.items():
        d = reg[kind](**params)
        idx = d.detect(out["range_profile"])
        out[f"c1_prof_{tag}_thr"] = d.thresholds
        out[f"c1_prof_{tag}_noise"] = d.noise_estimates
        out[f"c1_prof_{tag}_idx"] = np.array(idx, dtype=int)
        for j, rr in enumerate(rows):
            idx = d.detect(mag0[rr])
            out[f"c1_row{j}_{tag}_thr"] = d.thresholds
            out[f"c1_row{j}_{tag}_noise"] = d.noise_estimates
            out[f"c1_row{j}_{tag}_idx"] = np.array(idx, dtype=int)

    # range detector, sequential + ground detectors
    rdet = RangeDetector(cm, cfar_type="os_cfar_1d", cfar_params=dict(num_train=5, num_guard=3, rho=0.5, alpha=2))
    out["range_detector_dets"] = np.asarray(rdet.process(cube), dtype=float)
    out["range_detector_thr"] = rdet.thresholds
    seq = RangeDopplerDetectorSequential(cm, **SEQ_PARAMS)
    out["dets_seq"] = seq.process(cube)
    try:
        gnd = RangeDopplerGroundDetector(cm, **GROUND_PARAMS)
        out["dets_ground"] = gnd.process(cube)
        out["ground_altitude"] = np.array(gnd.altimeter.current_altitude_corrected_m)
        pcg = PointCloudGenerator(cm, az_antenna_idxs=sets["az_b"], el_antenna_idxs=sets["el_b"],
                                  detector_type="range_doppler_ground_detector", detector_params=GROUND_PARAMS)
        out["pc_ground"] = pcg.process(cube)
    except Exception as exc:  # noqa: BLE001
        print("ground detector failed:", exc)

    # range-angle
    for a in (64, 256):
        rap = RangeAngleProcessor(cm, num_angle_bins=a)
        out[f"ra_{a}_sub"] = rap.process(cube, chirp_idx=0, rx_antennas=sets["ra"])
        out[f"ra_{a}_all_c2"] = rap.process(cube, chirp_idx=2, rx_antennas=[])
    rap = RangeAngleProcessor(cm, num_angle_bins=64)
    out["ra_64_nowin"] = rap.process(cube, chirp_idx=1, rx_antennas=sets["ra"], perform_windowing=False)
    return out, cube


def dbs_outputs(cm, cube):
    out = {}
    vel = np.array([0.5, 0.1, 0.0])
    for tag, ants in (("sub", [0, 3, 4, 7]), ("all", [])):
        p = RangeAngleProcessorDBSEnhanced(cm, num_angle_bins_range_angle_response=64,
                                           num_angle_bins_dbs_enhanced_response=100, min_x_y_vel_dbs=0.25)
        out[f"dbs_{tag}"] = p.process(cube, velocity_ned=vel, rx_antennas=ants)
        out[f"dbs_{tag}_angle_bins"] = p.angle_bins
    p = RangeAngleProcessorDBSEnhanced(cm, num_angle_bins_range_angle_response=64,
                                       num_angle_bins_dbs_enhanced_response=100, min_x_y_vel_dbs=0.25)
    out["dbs_slow"] = p.process(cube, velocity_ned=np.array([0.1, 0.05, 0.0]), rx_antennas=[0, 3, 4, 7], chirp_idx=1)
    out["dbs_velocity"] = vel
    return out


def sar_outputs(cm, name):
    out = {}
    r, s, cpl, l = shapes(cm)
    vel = np.array([0.01, 0.5, 0.0])
    for tag, az, el, nf, stride in (
            ("a", np.deg2rad(np.linspace(-45, 45, 90)), np.array([0.0]), 5, 1),
            ("b", np.deg2rad(np.linspace(-30, 30, 16)), np.deg2rad(np.linspace(-10, 10, 4)), 3, 2)):
        p = SyntheticArrayBeamformerProcessor(
            cm, receiver_idx=1, chirp_cfg_idx=0, num_frames=nf, stride=stride,
            az_angle_bins_rad=az, el_angle_bins_rad=el,
            min_vel=np.array([0.0, 0.3, 0.0]), max_vel=np.array([0.05, 0.7, 0.05]),
            max_vel_stdev=np.array([0.02, 0.02, 0.02]))
        valid_flags = []
        for f in range(nf):
            raw = iq_to_complex_cube(synth_frame_iq(r, s, cpl, l, 1000 + f))
            v_f = vel + np.array([0.0, 0.002 * f, 0.0])
            try:
                resp = p.process(raw, current_vel=v_f)
            except ValueError:          # reference griddata bug for E>1
                resp = p.beamformed_resp
            valid_flags.append(bool(p.array_geometry_valid))
        resp = p.compute_synthetic_response(p.array_geometry)
        out[f"sar_{tag}_resp"] = np.array(resp)
        out[f"sar_{tag}_geometry"] = np.array(p.array_geometry)
        out[f"sar_{tag}_valid"] = np.array(valid_flags)
        out[f"sar_{tag}_az"] = az
        out[f"sar_{tag}_el"] = el
        out[f"sar_{tag}_hist_vel"] = np.array(p.history_avg_vel)
        out[f"sar_{tag}_chirp_times_us"] = np.array(p.chirp_start_times_us)
        out[f"sar_{tag}_range_bins"] = np.array(p.range_bins)
        out[f"sar_{tag}_lambda"] = np.array(p.lambda_m)
        out[f"sar_{tag}_params"] = np.array([1, 0, nf, stride])
    return out


def main():
    with open(os.path.join(HERE, "cfg_table.json"), "w") as fh:
        json.dump(cfg_table(), fh, indent=1, sort_keys=True)
    for name, (seeds, geom) in CASES.items():
        cm = load_ref_cfg(name, geom)
        r, s, cpl, l = shapes(cm)
        for i, seed in enumerate(seeds):
            iq = synth_frame_iq(r, s, cpl, l, seed)
            raw = iq_to_complex_cube(iq)
            out, cube = frame_outputs(cm, raw, full=(i == 0))
            if name == "RadSAR_1843_official" and i == 0:
                out.update(dbs_outputs(cm, cube))
                out.update(sar_outputs(cm, name))
            if name == "6843_RadVel_ods_20Hz" and i == 0:
                out.update(dbs_outputs(cm, cube))
            path = os.path.join(HERE, f"{name}_seed{seed}.npz")
            np.savez_compressed(path, **out)
            print(path, os.path.getsize(path) // 1024, "KiB",
                  {k: out[k].shape for k in ("dets2d_os_test", "dets2d_ca_test", "dets_seq") if k in out})


if __name__ == "__main__":
    main()

"""The C++ host mirror (include/dnrp_host.hpp: tx_descriptor_t, harq processes, sync/pcc/pdc reports, tpoint_t::work_pcc /
work_pdc / work_pdc_error, batch_phy_t) driven by a loopback firmware written like the reference's tfw_loopback.
CPU: the program builds against the C ABI and fails loudly without a GPU (no CPU fallback).  GPU: it runs both
scenarios (plain loopback with resampling 10/9 and two RX antennas; HARQ retransmission combining; packets found in one
continuous stream by batch_phy_t::process_stream)."""
import json
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT, load_product


def build_program(tmp_path, name="loopback_firmware"):
    load_product().build()
    exe = str(tmp_path / name)
    libdir = os.path.join(ROOT, "dect-nr-plus-sdr_b200")
    cmd = ["g++", "-std=c++17", "-O2", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"),
           "-I/usr/local/cuda/include", os.path.join(ROOT, "tests", "host", name + ".cpp"), "-o", exe, "-L" + libdir, "-ldnrp_cuda",
           "-Wl,-rpath," + libdir, "-L/usr/local/cuda/lib64", "-lcudart"]
    subprocess.check_call(cmd)
    return exe


def test_host_mirror_builds_and_refuses_to_run_without_gpu(tmp_path):
    import torch
    exe = build_program(tmp_path)
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu test")
    r = subprocess.run([exe, "4"], capture_output=True, text=True)
    assert r.returncode == 3 and "no CPU fallback" in r.stderr, (r.returncode, r.stderr)


@pytest.mark.gpu
def test_host_mirror_loopback_firmware(tmp_path):
    exe = build_program(tmp_path)
    prefix = str(tmp_path / "worker_pool_0000_")
    r = subprocess.run([exe, "48", prefix], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr)
    out = json.loads(r.stdout.strip().splitlines()[-1])
    _check_json_export(prefix, out)
    assert out["fails"] == 0 and out["plcf_mismatch"] == 0 and out["tb_mismatch"] == 0
    assert out["work_pcc"] >= 48 + 8 and out["harq_decoded"] == 8 and out["harq_first_round_errors"] > 0
    assert out["harq_rounds"] >= 2
    assert out["stream_decoded"] == 12     # streaming front door: every packet of the continuous stream found and decoded
    assert out["push_decoded"] == 12       # the same pushed piece by piece through the 10/9 resampler (open_stream / push_stream)


def _kind(x):
    if isinstance(x, dict):
        return {k: _kind(v) for k, v in x.items()}
    if isinstance(x, list):
        return ["%s[%d]" % (type(x[0]).__name__ if x else "none", len(x))]
    return type(x).__name__


def _check_json_export(prefix, out):
    """the per-packet JSON export (dnrp_json_export.hpp, batch_phy_t::enable_json_export) against the key tree of the
    reference's own example file (tests/golden/json_export_keys.json, made by tests/golden/make_json_export_keys.py):
    same file naming, entry naming, key names, nesting, order and value kinds for RADIO and PHY; the MAC section follows
    the current source (worker_tx_rx.cpp:399-400, plcf_decoder.cpp:154-164)"""
    gold = json.load(open(os.path.join(ROOT, "tests", "golden", "json_export_keys.json")))
    files = sorted(f for f in os.listdir(os.path.dirname(prefix)) if f.startswith(os.path.basename(prefix)))
    assert len(files) == out["json_files"] and files[0].endswith("_0000000000") and len(files) >= 2
    n_entries = 0
    for fi, name in enumerate(files):
        d = json.load(open(os.path.join(os.path.dirname(prefix), name)))
        assert len(d) == 8 or fi == len(files) - 1
        for key, e in d.items():
            assert key == gold["entry_prefix"] + str(n_entries)
            n_entries += 1
            got = _kind(e)
            assert list(got) == list(gold["tree"])                                  # top level: names and order
            for sec in ("RADIO",):
                assert got[sec] == gold["tree"][sec]
            for sec in ("worker_pool_config", "sync_report"):
                assert got["PHY"][sec] == gold["tree"]["PHY"][sec], sec
            rs, ref = got["PHY"]["rx_synced"], gold["tree"]["PHY"]["rx_synced"]
            fixed = [k for k in ref if not k.startswith("ch_")]
            assert [k for k in rs if not k.startswith("ch_")] == fixed and all(rs[k] == ref[k] for k in fixed)
            Nk = e["PHY"]["rx_synced"]["N_b_OCC_plus_DC"]
            chans = [k for k in rs if k.startswith("ch_")]
            assert chans == ["ch_%d_%d" % (a, t) for a in range(2) for t in range(e["PHY"]["sync_report"]["N_eff_TX"])]
            for k in chans:
                assert rs[k] == {"re": ["int[%d]" % Nk], "im": ["int[%d]" % Nk]}
            assert e["32fc_int32_scale"] == 100000.0 and e["PHY"]["worker_pool_config"]["L"] == 10
            assert e["PHY"]["rx_synced"]["mcs"] == 4 and -5.0 < e["PHY"]["rx_synced"]["snr"] < 40.0
            assert e["MAC"]["plcf"]["PLCF_type"] in (1, 2)
            # a flat unit-gain channel (two antennas, the second rotated): magnitudes of the exported estimate near 1
            h = (np.array(e["PHY"]["rx_synced"]["ch_0_0"]["re"]) + 1j * np.array(e["PHY"]["rx_synced"]["ch_0_0"]["im"])) / 1e5
            assert 0.3 < np.median(np.abs(h)) < 3.0
    # scenario 3 runs on a second batch_phy_t without export
    assert n_entries == out["work_pdc"] + out["work_pdc_error"] - out["stream_decoded"]


def test_multi_plan_firmware_builds_and_refuses_to_run_without_gpu(tmp_path):
    import torch
    exe = build_program(tmp_path, "multi_plan_firmware")
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu test")
    r = subprocess.run([exe, "4"], capture_output=True, text=True)
    assert r.returncode == 3 and "no CPU fallback" in r.stderr, (r.returncode, r.stderr)


@pytest.mark.gpu
def test_several_geometries_networks_and_plcf_types_in_one_call(tmp_path):
    """one process_windows call over packets of four geometries (two lengths x two MCS), two network ids and both PLCF
    types: every packet reaches work_pcc and work_pdc with its own transport block; with two GPUs the same runs on both
    inside one process (per-device kernel attributes)"""
    exe = build_program(tmp_path, "multi_plan_firmware")
    r = subprocess.run([exe, "32"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr)
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert out["fails"] == 0 and out["work_pcc"] == 64 and out["work_pdc"] == 64 and out["tb_mismatch"] == 0


def test_multi_beta_firmware_builds_and_refuses_to_run_without_gpu(tmp_path):
    import torch
    exe = build_program(tmp_path, "multi_beta_firmware")
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu test")
    r = subprocess.run([exe, "6"], capture_output=True, text=True)
    assert r.returncode == 3 and "no CPU fallback" in r.stderr, (r.returncode, r.stderr)


@pytest.mark.gpu
def test_one_receiver_for_packets_of_beta_4_2_and_1(tmp_path):
    """multi_beta_phy_t: a receiver configured for beta = 4 gets windows that carry packets of beta 4, 2 and 1 at its sample
    rate (sync_param.hpp:238-244: the synchronisation is triggered by every smaller beta); the beta search names each
    window's beta, the window is received under that geometry, the firmware reads the beta from sync_report.b and every
    transport block arrives"""
    exe = build_program(tmp_path, "multi_beta_firmware")
    r = subprocess.run([exe, "24"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr)
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert out["fails"] == 0 and out["beta_wrong"] == 0 and out["work_pdc"] == 24 and out["mismatch"] == 0


@pytest.mark.gpu
def test_one_receiver_for_packets_of_two_mu(tmp_path):
    """multi_beta_phy_t::add_mu: a receiver configured for mu = 2 / beta = 2 gets packets of (mu, beta) = (2, 2), (2, 1),
    (1, 4) and (1, 2) at its sample rate ("mu detection" is future work in the reference, README.md:198): the windows its
    own detector and beta search do not claim are searched by the PHYs of the other mu; the firmware reads mu and beta
    from sync_report and every transport block arrives"""
    exe = build_program(tmp_path, "multi_beta_firmware")
    r = subprocess.run([exe, "24", "mu"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr)
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert out["fails"] == 0 and out["mu_wrong"] == 0 and out["beta_wrong"] == 0 and out["work_pdc"] == 24

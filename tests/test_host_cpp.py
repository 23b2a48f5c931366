"""The C++ host mirror (include/dnrp_host.hpp: tx_descriptor_t, harq processes, sync/pcc/pdc reports, tpoint_t::work_pcc /
work_pdc / work_pdc_error, batch_phy_t) driven by a loopback firmware written like the reference's tfw_loopback.
CPU: the program builds against the C ABI and fails loudly without a GPU (no CPU fallback).  GPU: it runs both
scenarios (plain loopback with resampling 10/9 and two RX antennas; HARQ retransmission combining; packets found in one
continuous stream by batch_phy_t::process_stream)."""
import json
import os
import subprocess

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
    r = subprocess.run([exe, "48"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr)
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert out["fails"] == 0 and out["plcf_mismatch"] == 0 and out["tb_mismatch"] == 0
    assert out["work_pcc"] >= 48 + 8 and out["harq_decoded"] == 8 and out["harq_first_round_errors"] > 0
    assert out["harq_rounds"] >= 2
    assert out["stream_decoded"] == 12     # streaming front door: every packet of the continuous stream found and decoded
    assert out["push_decoded"] == 12       # the same pushed piece by piece through the 10/9 resampler (open_stream / push_stream)


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
    assert out["fails"] == 0 and out["work_pcc"] == 32 and out["work_pdc"] == 32 and out["tb_mismatch"] == 0

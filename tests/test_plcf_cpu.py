"""PLCF codec of the host mirror (include/dnrp_plcf.hpp: plcf_10/20/21_t, feedback info formats 1-6, plcf_decoder_t with
its radio-device-class check) against oracle/plcf.py, the literal restatement of the reference's byte arithmetic.
Random valid and invalid field sets are packed by both, every packed header is decoded by both, and raw random bytes
are decoded by both (unknown header formats, reserved bits, out-of-range CQI, RDC limits)."""
import os
import random
import subprocess

import pytest

from conftest import ROOT
from oracle import plcf as O

UNDEF = O.UNDEF


@pytest.fixture(scope="module")
def codec(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("plcf") / "plcf_codec")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "host", "plcf_codec.cpp"), "-o", exe])

    def run(lines):
        r = subprocess.run([exe], input="\n".join(lines) + "\n", capture_output=True, text=True, check=True)
        out = r.stdout.strip().split("\n")
        assert len(out) == len(lines)
        return out
    return run


def _rand_fb(rng, ff, valid):
    f = dict(HARQ=rng.randrange(8), TF=rng.randrange(2), HARQ1=rng.randrange(8), TF1=rng.randrange(2),
             bitmap=rng.randrange(256), Codebook=rng.randrange(8 if ff == 2 else 64), MIMO=rng.randrange(2 if ff == 2 else 4),
             Reserved=0, Buffer_Size=rng.choice([0, 1, 16, 17, 100, 1024, 1025, 70000, 131072, 131073, 10 ** 6]),
             MCS=rng.randrange(12))
    if not valid:
        k = rng.choice(["MCS", "Reserved", "TF"])
        f[k] = {"MCS": 12, "Reserved": 1, "TF": UNDEF}[k]
    return f


def _fb_numbers(f):
    s = lambda v: -1 if v == UNDEF else v
    return [f["HARQ"], s(f["TF"]), f["HARQ1"], s(f["TF1"]), f["bitmap"], f["Codebook"], s(f["MIMO"]), f["Reserved"],
            f["Buffer_Size"], s(f["MCS"])]


def _fmt_decoded(p, ptype):
    """the driver's output line for a decoded header"""
    if p is None:
        return None
    s = "hf %d plt %d pl %d nid %d tid %d txp %d dBm %d mcs %d nss %d rv %d" % (
        p["HeaderFormat"], p["PacketLengthType"], p["PacketLength_m1"] + 1, p["ShortNetworkID"], p["TransmitterIdentity"],
        p["TransmitPower"], O.TX_POWER_TABLE[p["TransmitPower"]], p["DFMCS"], p["N_SS"], p["rv"])
    if ptype == 2:
        if p["HeaderFormat"] == 0:
            s += " rid %d ndi %d harq %d ff %d" % (p["ReceiverIdentity"], p["DFNewDataIndication"], p["DFHARQProcessNumber"],
                                                     p["FeedbackFormat"])
        else:
            s += " rid %d ff %d" % (p["ReceiverIdentity"], p["FeedbackFormat"])
        f = p["fb"]
        if f is None:
            s += " fb none"
        else:
            i32 = lambda v: -1 if v == UNDEF else v
            s += " fb " + " ".join(str(i32(v)) for v in (f["HARQ"], f["TF"], f["HARQ1"], f["TF1"], f["bitmap"], f["Codebook"],
                                                          f["MIMO"], f["Reserved"], f["Buffer_Size"], f["MCS"]))
    return s


def test_pack_and_decode_match_the_oracle(codec):
    rng = random.Random(5)
    lines, expect = [], []
    for i in range(1500):
        kind = rng.choice(["10", "20", "21"])
        bad = rng.random() < 0.25
        base = dict(HeaderFormat={"10": 0, "20": 0, "21": 1}[kind], PacketLengthType=rng.randrange(2),
                    PacketLength_m1=rng.randrange(16), ShortNetworkID=rng.randrange(1, 256),
                    TransmitterIdentity=rng.randrange(1, 65536), TransmitPower=rng.randrange(16),
                    DFMCS=rng.randrange(8 if kind == "10" else 12))
        if kind == "10":
            base["Reserved"] = 0
            if bad:
                k = rng.choice(["ShortNetworkID", "TransmitterIdentity", "DFMCS", "Reserved", "HeaderFormat"])
                base[k] = {"ShortNetworkID": 0, "TransmitterIdentity": 0, "DFMCS": 8, "Reserved": 1, "HeaderFormat": 1}[k]
            lines.append("P10 %d %d %d %d %d %d %d %d" % (base["HeaderFormat"], base["PacketLengthType"], base["PacketLength_m1"],
                                                         base["ShortNetworkID"], base["TransmitterIdentity"],
                                                         base["TransmitPower"], base["Reserved"], base["DFMCS"]))
            expect.append(O.plcf10_pack(base).hex() if O.plcf10_valid(base) else "INVALID")
            continue
        ff = rng.randrange(7)
        fb = _rand_fb(rng, ff, valid=not (bad and rng.random() < 0.5))
        base.update(ReceiverIdentity=rng.randrange(1, 65536), NumberOfSpatialStreams=rng.randrange(4), FeedbackFormat=ff)
        if kind == "20":
            base.update(DFRedundancyVersion=rng.randrange(4), DFNewDataIndication=rng.randrange(2),
                        DFHARQProcessNumber=rng.randrange(8))
            if bad and rng.random() < 0.5:
                base["ReceiverIdentity"] = 0
            head = [base[k] for k in ("PacketLengthType", "PacketLength_m1", "ShortNetworkID", "TransmitterIdentity",
                                      "TransmitPower", "DFMCS", "ReceiverIdentity", "NumberOfSpatialStreams",
                                      "DFRedundancyVersion", "DFNewDataIndication", "DFHARQProcessNumber", "FeedbackFormat")]
            ok = O.plcf20_valid(base)
        else:
            base["Reserved"] = 1 if (bad and rng.random() < 0.5) else 0
            head = [base[k] for k in ("PacketLengthType", "PacketLength_m1", "ShortNetworkID", "TransmitterIdentity",
                                      "TransmitPower", "DFMCS", "ReceiverIdentity", "NumberOfSpatialStreams", "Reserved",
                                      "FeedbackFormat")]
            ok = O.plcf21_valid(base)
        ok = ok and (ff == 0 or O.fb_valid(ff, fb))
        lines.append("P%s " % kind + " ".join(str(v) for v in head + _fb_numbers(fb)))
        expect.append(O.plcf2x_pack(base, fb).hex() if ok else "INVALID")
    got = codec(lines)
    assert got == expect
    # every valid header back through plcf_decoder_t with random radio-device-class limits
    ulines, uexpect = [], []
    for line, h in zip(lines, got):
        if h == "INVALID":
            continue
        ptype = 1 if line.startswith("P10") else 2
        lim = (rng.choice([1, 4, 8, 16]), rng.choice([3, 7, 9, 11]), rng.choice([1, 2, 4, 8]))
        ulines.append("U %d %s %d %d %d" % (ptype, h, *lim))
        p = O.decode_and_rdc_check(ptype, bytes.fromhex(h), *lim)
        uexpect.append("has=0" if p is None else "has=%d %s" % (ptype, _fmt_decoded(p, ptype)))
    assert codec(ulines) == uexpect
    assert sum(1 for e in uexpect if e != "has=0") > 300 and sum(1 for e in uexpect if e == "has=0") > 100


def test_decoding_random_bytes_matches_the_oracle(codec):
    rng = random.Random(9)
    lines, expect = [], []
    for _ in range(3000):
        ptype = rng.choice([1, 2])
        a = bytearray(rng.randrange(256) for _ in range(10))
        if rng.random() < 0.7:   # mostly plausible headers: known format, ids set
            a[0] = (rng.choice([0, 0, 1, 2]) << 5) | (a[0] & 0x1F)
        if rng.random() < 0.5:
            a[8] = (rng.randrange(7) << 4) | (a[8] & 0xF)
        n = 5 if ptype == 1 else 10
        lines.append("U %d %s 16 11 8" % (ptype, bytes(a[:n]).hex()))
        p = O.decode_and_rdc_check(ptype, bytes(a[:n]) + bytes(10 - n), 16, 11, 8)
        expect.append("has=0" if p is None else "has=%d %s" % (ptype, _fmt_decoded(p, ptype)))
    assert codec(lines) == expect
    assert sum(1 for e in expect if e != "has=0") > 200


def test_transmit_power_and_buffer_status_tables(codec):
    dbm = list(range(-45, 30))
    sizes = [0, 1, 15, 16, 17, 32, 33, 1000, 1024, 1025, 131072, 131073, 2 ** 31]
    got = codec(["S %d" % d for d in dbm] + ["B %d" % s for s in sizes])
    assert [int(g) for g in got[:len(dbm)]] == [O.set_transmit_power(d) for d in dbm]
    assert [int(g) for g in got[len(dbm):]] == [O.buffer_size_2_buffer_status(s) for s in sizes]

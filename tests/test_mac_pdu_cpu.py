"""MAC PDU framing of the host mirror (include/dnrp_mac_pdu.hpp: MAC header type, the four common headers, multiplexing
headers, mac_pdu_decoder_t demultiplexing, mac_pdu_writer_t) against oracle/mac_pdu.py, the literal restatement of the
reference's byte arithmetic: random PDUs built by both, every built PDU decoded by both, random byte strings decoded by
both."""
import os
import random
import subprocess

import pytest

from conftest import ROOT
from oracle import mac_pdu as O


@pytest.fixture(scope="module")
def codec(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("macpdu") / "mac_pdu_codec")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "host", "mac_pdu_codec.cpp"), "-o", exe])

    def run(lines):
        r = subprocess.run([exe], input="\n".join(lines) + "\n", capture_output=True, text=True, check=True)
        out = r.stdout.strip("\n").split("\n")
        assert len(out) == len(lines)
        return out
    return run


def _fmt(d):
    s = "final=%d" % (1 if d["final"] else 0)
    if "type" in d:
        s += " type %d sec %d" % (d["type"], d["sec"])
    if "mch" in d:
        m = d["mch"]
        s += " reset %d seq %d rx %d tx %d nid %d" % (m["reset"], m["seq"], m["rx"], m["tx"], m["nid"])
    for ext, ie, n, off in d["sdus"]:
        s += " [%d %d %d @%d]" % (ext, ie, n, off)
    if d["unparsed"] is not None:
        s += " unparsed %d" % d["unparsed"]
    return s


def test_build_and_decode_match_the_oracle(codec):
    rng = random.Random(4)
    lines, expect, specs = [], [], []
    for _ in range(800):
        N = rng.choice([17, 37, 60, 105, 300, 700, 4115])
        typ = rng.choice([0, 1, 2, 3, 4])
        sec = rng.randrange(4)
        reset, seq = rng.randrange(2), rng.randrange(4096)
        rx, tx, nid = rng.randrange(1, 2 ** 32), rng.randrange(1, 2 ** 32), rng.randrange(2 ** 24)
        if rng.random() < 0.1:
            tx = 0
        sdus = []
        for _ in range(rng.randrange(5)):
            n = rng.choice([0, 1, 5, 40, 255, 256, 600]) if N > 700 else rng.choice([0, 1, 3, 9, 20])
            sdus.append((rng.choice([1, 2, 3, 4, 5, 6, 23]), bytes(rng.randrange(256) for _ in range(n))))
        lines.append("B %d %d %d %d %d %d %d %d %d " % (N, typ, sec, reset, seq, rx, tx, nid, len(sdus)) +
                     " ".join("%d %s" % (ie, d.hex() or "-") for ie, d in sdus))
        b = O.build(N, typ, sec, reset, seq, rx, tx, nid, sdus)
        expect.append("INVALID" if b is None else b.hex())
        specs.append((typ, sdus))
    got = codec(lines)
    assert got == expect
    built = [(h, s) for h, s in zip(got, specs) if h != "INVALID"]
    assert len(built) > 400
    dlines = ["D " + h for h, _ in built]
    dgot = codec(dlines)
    for (h, (typ, sdus)), g in zip(built, dgot):
        d = O.decode(bytes.fromhex(h))
        assert g == _fmt(d)
        assert d["final"] and d["type"] == typ and [(ie, n) for _, ie, n, _ in d["sdus"]] == [(ie, len(x)) for ie, x in sdus]
        a = bytes.fromhex(h)
        for (_, _, n, off), (_, x) in zip(d["sdus"], sdus):
            assert a[off:off + n] == x


def test_decoding_random_bytes_matches_the_oracle(codec):
    rng = random.Random(8)
    lines, expect = [], []
    for _ in range(2000):
        n = rng.choice([1, 2, 7, 12, 40, 200])
        a = bytearray(rng.randrange(256) for _ in range(n))
        if rng.random() < 0.8:
            a[0] = (rng.randrange(4) << 4) | rng.randrange(5)        # plausible header type
        if n > 12 and rng.random() < 0.7:
            a[1] &= 0x1F                                              # reserved bits of the common header clear
        lines.append("D " + bytes(a).hex())
        expect.append(_fmt(O.decode(bytes(a))))
    assert codec(lines) == expect

"""TEST ORACLE (not product code): Physical Layer Control Field pack / unpack, a literal restatement of the reference's
byte arithmetic: sections_part4/physical_header_field/plcf_base.cpp:44-62, plcf_10.cpp:62-91, plcf_20.cpp:88-127,
plcf_21.cpp:78-113, feedback_info.cpp:28-425, plcf_decoder.cpp:44-104, mac_architecture/identity.cpp:55-87.
Only tests/ may import this module.  Parity: pinned by the reference's own byte layout (restated shift by shift);
the reference holds no golden vectors for the PLCF."""

UNDEF = 0xFFFFFFFF
TX_POWER_TABLE = [-40, -30, -20, -16, -12, -8, -4, 0, 4, 7, 10, 13, 16, 19, 21, 23]
BUFFER_LOWER = [0, 0, 16, 32, 64, 128, 256, 512, 1024, 2048, 4096, 8192, 16384, 32768, 65536, 131072]
N_SS_REV = [1, 2, 4, 8]


def buffer_size_2_buffer_status(n):
    if n == 0:
        return 0
    bounds = [16, 32, 64, 128, 256, 512, 1024, 2048, 4096, 8192, 16384, 32768, 65536, 131072]
    for i, b in enumerate(bounds):
        if n <= b:
            return i + 1
    return 15


def set_transmit_power(dbm):
    for i, v in enumerate(TX_POWER_TABLE):   # std::lower_bound
        if v >= dbm:
            return i
    return len(TX_POWER_TABLE) - 1


def cqi_2_mcs(cqi):
    return UNDEF if (cqi == 0 or cqi > 12) else cqi - 1


def _id_ok(v, bits):
    return v != 0 and v < (1 << bits)


# feedback dict keys: HARQ TF HARQ1 TF1 bitmap Codebook MIMO Reserved Buffer_Size MCS
def fb_valid(ff, f):
    mcs_ok = f["MCS"] <= 11
    if ff == 1:
        return f["HARQ"] <= 7 and f["TF"] != UNDEF and mcs_ok
    if ff == 2:
        return f["Codebook"] <= 7 and f["MIMO"] in (0, 1) and mcs_ok
    if ff == 3:
        return f["HARQ"] <= 7 and f["TF"] != UNDEF and f["HARQ1"] <= 7 and f["TF1"] != UNDEF and mcs_ok
    if ff == 4:
        return f["bitmap"] <= 0xFF and mcs_ok
    if ff == 5:
        return f["HARQ"] <= 7 and f["TF"] != UNDEF and f["MIMO"] != UNDEF and f["Codebook"] <= 63
    if ff == 6:
        return f["HARQ"] <= 7 and f["Reserved"] == 0 and mcs_ok
    return False


def fb_pack(ff, f, a):
    """a: two bytes (byte 8, 9 of the header); upper nibble of a[0] holds FeedbackFormat"""
    a[0] &= 0xF0
    a[1] = 0
    if ff == 0:
        return
    bs = lambda: buffer_size_2_buffer_status(f["Buffer_Size"])
    cqi = lambda: f["MCS"] + 1
    if ff == 1:
        a[0] |= (f["HARQ"] << 1) | (f["TF"] & 1); a[1] = (bs() << 4) | cqi()
    elif ff == 2:
        a[0] |= (f["Codebook"] << 1) | (f["MIMO"] & 1); a[1] = (bs() << 4) | cqi()
    elif ff == 3:
        a[0] |= (f["HARQ"] << 1) | (f["TF"] & 1); a[1] = (f["HARQ1"] << 5) | ((f["TF1"] & 1) << 4) | cqi()
    elif ff == 4:
        a[0] |= f["bitmap"] >> 4; a[1] = ((f["bitmap"] & 0xF) << 4) | cqi()
    elif ff == 5:
        a[0] |= (f["HARQ"] << 1) | (f["TF"] & 1); a[1] = ((f["MIMO"] & 3) << 6) | f["Codebook"]
    elif ff == 6:
        a[0] |= (f["HARQ"] << 1) | f["Reserved"]; a[1] = (bs() << 4) | cqi()


def fb_unpack(ff, a):
    """-> (ok, dict or None)"""
    f = dict(HARQ=0, TF=UNDEF, HARQ1=0, TF1=UNDEF, bitmap=0, Codebook=0, MIMO=UNDEF, Reserved=0, Buffer_Size=0, MCS=UNDEF)
    if ff == 0:
        return ((a[0] & 0xF) == 0 and a[1] == 0), None
    if ff > 6:
        return False, None
    if ff == 1:
        f.update(HARQ=(a[0] >> 1) & 7, TF=a[0] & 1, Buffer_Size=BUFFER_LOWER[a[1] >> 4], MCS=cqi_2_mcs(a[1] & 0xF))
    elif ff == 2:
        f.update(Codebook=(a[0] >> 1) & 7, MIMO=a[0] & 1, Buffer_Size=BUFFER_LOWER[a[1] >> 4], MCS=cqi_2_mcs(a[1] & 0xF))
    elif ff == 3:
        f.update(HARQ=(a[0] >> 1) & 7, TF=a[0] & 1, HARQ1=(a[1] >> 5) & 7, TF1=(a[1] >> 4) & 1, MCS=cqi_2_mcs(a[1] & 0xF))
    elif ff == 4:
        f.update(bitmap=((a[0] & 0xF) << 4) | (a[1] >> 4), MCS=cqi_2_mcs(a[1] & 0xF))
    elif ff == 5:
        f.update(HARQ=(a[0] >> 1) & 7, TF=a[0] & 1, MIMO=(a[1] >> 6) & 3, Codebook=a[1] & 0x3F)
    elif ff == 6:
        f.update(HARQ=(a[0] >> 1) & 7, Reserved=a[0] & 1, Buffer_Size=BUFFER_LOWER[a[1] >> 4], MCS=cqi_2_mcs(a[1] & 0xF))
    return fb_valid(ff, f), f


def _base_valid(p, hf, mcs_max):
    return (p["HeaderFormat"] == hf and p["PacketLengthType"] <= 1 and p["PacketLength_m1"] <= 15 and
            _id_ok(p["ShortNetworkID"], 8) and _id_ok(p["TransmitterIdentity"], 16) and p["TransmitPower"] <= 15 and
            p["DFMCS"] <= mcs_max)


def plcf10_valid(p):
    return _base_valid(p, 0, 7) and p["Reserved"] == 0


def plcf10_pack(p):
    a = bytearray(5)
    a[0] = (p["HeaderFormat"] << 5) | (p["PacketLengthType"] << 4) | p["PacketLength_m1"]
    a[1] = p["ShortNetworkID"]
    a[2], a[3] = p["TransmitterIdentity"] >> 8, p["TransmitterIdentity"] & 0xFF
    a[4] = (p["TransmitPower"] << 4) | (p["Reserved"] << 3) | p["DFMCS"]
    return bytes(a)


def plcf20_valid(p):
    return (_base_valid(p, 0, 11) and _id_ok(p["ReceiverIdentity"], 16) and p["NumberOfSpatialStreams"] <= 3 and
            p["DFRedundancyVersion"] <= 3 and p["DFNewDataIndication"] <= 1 and p["DFHARQProcessNumber"] <= 7 and
            p["FeedbackFormat"] <= 15)


def plcf21_valid(p):
    return (_base_valid(p, 1, 11) and _id_ok(p["ReceiverIdentity"], 16) and p["NumberOfSpatialStreams"] <= 3 and
            p["Reserved"] == 0 and p["FeedbackFormat"] <= 15)


def plcf2x_pack(p, fb):
    a = bytearray(10)
    a[0] = (p["HeaderFormat"] << 5) | (p["PacketLengthType"] << 4) | p["PacketLength_m1"]
    a[1] = p["ShortNetworkID"]
    a[2], a[3] = p["TransmitterIdentity"] >> 8, p["TransmitterIdentity"] & 0xFF
    a[4] = (p["TransmitPower"] << 4) | p["DFMCS"]
    a[5], a[6] = p["ReceiverIdentity"] >> 8, p["ReceiverIdentity"] & 0xFF
    if p["HeaderFormat"] == 0:
        a[7] = (p["NumberOfSpatialStreams"] << 6) | (p["DFRedundancyVersion"] << 4) | (p["DFNewDataIndication"] << 3) | \
            p["DFHARQProcessNumber"]
    else:
        a[7] = (p["NumberOfSpatialStreams"] << 6) | p["Reserved"]
    a[8] = p["FeedbackFormat"] << 4
    tail = bytearray(a[8:10])
    fb_pack(p["FeedbackFormat"], fb, tail)
    a[8:10] = tail
    return bytes(a)


def decode_and_rdc_check(ptype, a, PacketLength_max, mcs_index_max, N_SS_max):
    """plcf_decoder_t::decode_and_rdc_check: -> dict of fields or None"""
    hf = a[0] >> 5
    p = dict(HeaderFormat=hf, PacketLengthType=(a[0] >> 4) & 1, PacketLength_m1=a[0] & 0xF, ShortNetworkID=a[1],
             TransmitterIdentity=(a[2] << 8) | a[3], TransmitPower=a[4] >> 4)
    fb = None
    if ptype == 1:
        if hf >= 1:
            return None
        p.update(Reserved=(a[4] >> 3) & 1, DFMCS=a[4] & 7)
        if not plcf10_valid(p):
            return None
        n_ss, rv = 1, 0
    else:
        if hf >= 2:
            return None
        p.update(DFMCS=a[4] & 0xF, ReceiverIdentity=(a[5] << 8) | a[6], NumberOfSpatialStreams=a[7] >> 6,
                 FeedbackFormat=a[8] >> 4)
        if hf == 0:
            p.update(DFRedundancyVersion=(a[7] >> 4) & 3, DFNewDataIndication=(a[7] >> 3) & 1, DFHARQProcessNumber=a[7] & 7)
        else:
            p.update(Reserved=a[7] & 0x3F)
        ok, fb = fb_unpack(p["FeedbackFormat"], a[8:10])
        if not ok or not (plcf20_valid(p) if hf == 0 else plcf21_valid(p)):
            return None
        n_ss, rv = N_SS_REV[p["NumberOfSpatialStreams"]], p.get("DFRedundancyVersion", 0)
    if p["PacketLength_m1"] + 1 > PacketLength_max or p["DFMCS"] > mcs_index_max or n_ss > N_SS_max:
        return None
    p.update(N_SS=n_ss, rv=rv, fb=fb)
    return p

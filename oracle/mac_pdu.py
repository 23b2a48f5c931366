"""TEST ORACLE (not product code): MAC PDU framing, a literal restatement of the reference's byte arithmetic:
sections_part4/mac_pdu/mac_header_type.cpp:44-64, mac_common_header.cpp:46-199, mac_multiplexing_header.cpp:86-130 and
:413-440 (length field), mac_pdu_decoder.cpp:85-318 (state machine; MAC messages without a length field end the walk, as
they do for an IE the reference has no pool entry for; padding ends the PDU, clause 6.4.3.8).  Only tests/ may import it.
Parity: pinned by the reference's byte layout (restated shift by shift); the reference holds no golden MAC PDUs."""

MCH_SIZE = {0: 2, 1: 7, 2: 10, 3: 6, 4: 0}


def build(N, typ, sec, reset, seq, rx, tx, nid, sdus):
    """-> bytes of the transport block or None if a field is invalid / the PDU does not fit"""
    if typ not in MCH_SIZE or sec > 3:
        return None
    if typ in (0, 2, 3) and (reset > 1 or seq > 0xFFF):
        return None
    if (typ in (1, 2, 3) and tx == 0) or (typ == 2 and rx == 0) or (typ == 1 and nid > 0xFFFFFF):
        return None
    a = bytearray([0xEE] * N)
    if 1 + MCH_SIZE[typ] > N:
        return None
    a[0] = (0 << 6) | (sec << 4) | typ
    w = 1
    if typ == 1:
        a[w:w + 3] = nid.to_bytes(3, "big"); a[w + 3:w + 7] = tx.to_bytes(4, "big")
    elif typ != 4:
        a[w] = (reset << 4) | (seq >> 8); a[w + 1] = seq & 0xFF
        if typ == 2:
            a[w + 2:w + 6] = rx.to_bytes(4, "big"); a[w + 6:w + 10] = tx.to_bytes(4, "big")
        if typ == 3:
            a[w + 2:w + 6] = tx.to_bytes(4, "big")
    w += MCH_SIZE[typ]
    for ie, data in sdus:
        n = len(data)
        if ie == 0 or ie > 63 or n > 0xFFFF:
            return None
        hdr = bytes([(1 << 6) | ie, n]) if n <= 0xFF else bytes([(2 << 6) | ie, n >> 8, n & 0xFF])
        if w + len(hdr) + n > N:
            return None
        a[w:w + len(hdr)] = hdr
        a[w + len(hdr):w + len(hdr) + n] = data
        w += len(hdr) + n
    rest = N - w
    if rest:
        a[w:] = bytes(rest)
        if rest == 1:
            a[w] = 3 << 6
        elif rest == 2:
            a[w] = (3 << 6) | (1 << 5)
        else:
            a[w] = 1 << 6; a[w + 1] = min(rest - 2, 0xFF)
    return bytes(a)


def decode(a):
    """-> dict(final, type, sec, mch fields, sdus [(mac_ext, ie, n, offset)], unparsed)"""
    N = len(a)
    out = dict(final=False, sdus=[], unparsed=None)
    r = 0
    if N < 1:
        out["final"] = True
        return out
    v, sec, typ = a[0] >> 6, (a[0] >> 4) & 3, a[0] & 15
    if v != 0 or typ > 4:
        out["final"] = True   # premature abort counts as a valid final state (all bytes consumed)
        return out
    out.update(type=typ, sec=sec)
    r = 1
    n = MCH_SIZE[typ]
    if r + n > N:
        out["final"] = True
        return out
    m = dict(reset=0, seq=0, rx=0, tx=0, nid=0)
    ok = True
    if typ == 1:
        m["nid"] = int.from_bytes(a[r:r + 3], "big"); m["tx"] = int.from_bytes(a[r + 3:r + 7], "big")
        ok = m["tx"] != 0
    elif typ != 4:
        reserved = a[r] >> 5
        m["reset"] = (a[r] >> 4) & 1; m["seq"] = ((a[r] & 15) << 8) | a[r + 1]
        if typ == 2:
            m["rx"] = int.from_bytes(a[r + 2:r + 6], "big"); m["tx"] = int.from_bytes(a[r + 6:r + 10], "big")
            ok = m["rx"] != 0 and m["tx"] != 0
        if typ == 3:
            m["tx"] = int.from_bytes(a[r + 2:r + 6], "big")
            ok = m["tx"] != 0
        ok = ok and reserved == 0
    if not ok:
        out["final"] = True
        return out
    out["mch"] = m
    r += n
    while r < N:
        ext = a[r] >> 6
        ie = a[r] & (31 if ext == 3 else 63)
        if ie == 0:
            break                                   # padding
        if ext == 0:
            out["unparsed"] = r
            break
        nh = {1: 2, 2: 3, 3: 1}[ext]
        if r + nh > N:
            break
        ln = a[r + 1] if ext == 1 else ((a[r + 1] << 8) | a[r + 2]) if ext == 2 else (a[r] >> 5) & 1
        if r + nh + ln > N:
            break
        out["sdus"].append((ext, ie, ln, r + nh))
        r += nh + ln
    out["final"] = True
    return out

"""Test infrastructure: turns an interleaved baseline JPEG (one scan, three components) into the equivalent
NON-interleaved one -- three scans of one component each (T.81 A.2.2) -- without touching a coefficient.

The AC part of every block is copied bit for bit; only the DC differences are re-coded, because the prediction
chain of a component follows the scan's block order (MCU order when interleaved, raster order over
ceil(w/8) x ceil(h/8) blocks when not).  Blocks that exist only as MCU padding are dropped.  Any decoder must
produce the same pixels from both files; tests/test_oracle.py checks that with libjpeg-turbo (Pillow).
"""
import struct


def _segments(data):
    """Yields (marker, payload_offset, payload_length) up to and including SOS; returns the scan offset."""
    p = 2
    out = []
    while True:
        assert data[p] == 0xFF
        while data[p] == 0xFF:
            p += 1
        m = data[p]
        p += 1
        ln = struct.unpack(">H", data[p:p + 2])[0]
        out.append((m, p + 2, ln - 2))
        p += ln
        if m == 0xDA:
            return out, p


def _huff_tables(payload):
    """DHT payload -> {(class, id): (decode dict {(len, code): sym}, encode dict {sym: (len, code)})}"""
    t = {}
    q = 0
    while q < len(payload):
        tc, th = payload[q] >> 4, payload[q] & 15
        bits = payload[q + 1:q + 17]
        n = sum(bits)
        vals = payload[q + 17:q + 17 + n]
        q += 17 + n
        dec, enc = {}, {}
        code, k = 0, 0
        for l in range(1, 17):
            for _ in range(bits[l - 1]):
                dec[(l, code)] = vals[k]
                enc[vals[k]] = (l, code)
                code += 1
                k += 1
            code <<= 1
        t[(tc, th)] = (dec, enc)
    return t


class _Bits:
    def __init__(self, data):
        # unstuff once: FF00 -> FF; the input has no restart markers (asserted by the caller)
        b = bytearray()
        i = 0
        while i < len(data):
            if data[i] == 0xFF:
                if data[i + 1] == 0x00:
                    b.append(0xFF)
                    i += 2
                    continue
                break
            b.append(data[i])
            i += 1
        self.s = "".join(format(x, "08b") for x in b)
        self.p = 0

    def take(self, n):
        v = self.s[self.p:self.p + n]
        self.p += n
        return v

    def symbol(self, dec):
        code, l = 0, 0
        while True:
            code = (code << 1) | int(self.s[self.p])
            self.p += 1
            l += 1
            if (l, code) in dec:
                return dec[(l, code)]
            assert l < 16, "bad code"


def _extend(bits):
    if not bits:
        return 0
    v = int(bits, 2)
    return v if bits[0] == "1" else v - (1 << len(bits)) + 1


def _encode_dc(diff, enc):
    size = 0 if diff == 0 else abs(diff).bit_length()
    l, code = enc[size]
    out = format(code, "0%db" % l)
    if size:
        v = diff if diff > 0 else diff + (1 << size) - 1
        out += format(v, "0%db" % size)
    return out


def _stuff(bitstr):
    bitstr += "1" * (-len(bitstr) % 8)
    raw = int(bitstr, 2).to_bytes(len(bitstr) // 8, "big") if bitstr else b""
    return raw.replace(b"\xff", b"\xff\x00")


def to_three_scans(data, restart_blocks=0, order=(0, 1, 2)):
    """data: baseline JPEG, 3 components, one interleaved scan, no restart interval.
    restart_blocks: restart interval (in blocks == MCUs of a one-component scan) for the new scans, 0 = none.
    order: the order in which the component scans are written."""
    data = bytes(data)
    segs, scan_off = _segments(data)
    huff = {}
    sof = sos = None
    head = bytearray(b"\xff\xd8")
    for m, off, ln in segs:
        pl = data[off:off + ln]
        if m == 0xC4:
            huff.update(_huff_tables(pl))
        if m == 0xC0:
            sof = pl
        if m == 0xDA:
            sos = pl
            continue
        assert m != 0xDD or struct.unpack(">H", pl)[0] == 0, "input must not use restart intervals"
        if m != 0xDD:
            head += b"\xff" + bytes([m]) + struct.pack(">H", ln + 2) + pl
    height, width, ncomp = struct.unpack(">HHB", sof[1:6])
    assert ncomp == 3 and sos[0] == 3
    comps = []
    for c in range(3):
        cid, hv, tq = sof[6 + 3 * c:9 + 3 * c]
        assert sos[1 + 2 * c] == cid
        comps.append({"id": cid, "h": hv >> 4, "v": hv & 15, "td": sos[2 + 2 * c] >> 4, "ta": sos[2 + 2 * c] & 15})
    hmax, vmax = max(c["h"] for c in comps), max(c["v"] for c in comps)
    mcus_x, mcus_y = -(-width // (8 * hmax)), -(-height // (8 * vmax))
    br = _Bits(data[scan_off:])
    # blocks[c][(by, bx)] = (absolute DC, AC bit string)
    blocks = [dict() for _ in range(3)]
    pred = [0, 0, 0]
    for my in range(mcus_y):
        for mx in range(mcus_x):
            for c, cp in enumerate(comps):
                dcdec, acdec = huff[(0, cp["td"])][0], huff[(1, cp["ta"])][0]
                for v in range(cp["v"]):
                    for h in range(cp["h"]):
                        size = br.symbol(dcdec)
                        pred[c] += _extend(br.take(size))
                        start = br.p
                        k = 1
                        while k < 64:
                            rs = br.symbol(acdec)
                            r, s = rs >> 4, rs & 15
                            if s == 0:
                                if r != 15:
                                    break
                                k += 16
                                continue
                            br.take(s)
                            k += r + 1
                        blocks[c][(my * cp["v"] + v, mx * cp["h"] + h)] = (pred[c], br.s[start:br.p])
    out = bytearray(head)
    if restart_blocks:
        out += b"\xff\xdd\x00\x04" + struct.pack(">H", restart_blocks)
    for c in order:
        cp = comps[c]
        cw, ch = -(-width * cp["h"] // hmax), -(-height * cp["v"] // vmax)
        bw, bh = -(-cw // 8), -(-ch // 8)
        dcenc = huff[(0, cp["td"])][1]
        out += b"\xff\xda" + struct.pack(">HB", 8, 1) + bytes([cp["id"], (cp["td"] << 4) | cp["ta"], 0, 63, 0])
        bits, pr, n, rst = "", 0, 0, 0
        for by in range(bh):
            for bx in range(bw):
                if restart_blocks and n and n % restart_blocks == 0:
                    out += _stuff(bits) + bytes([0xFF, 0xD0 + (rst & 7)])
                    bits, pr, rst = "", 0, rst + 1
                dc, ac = blocks[c][(by, bx)]
                bits += _encode_dc(dc - pr, dcenc) + ac
                pr = dc
                n += 1
        out += _stuff(bits)
    out += b"\xff\xd9"
    return bytes(out)

"""numpy/Python restatement of the K0 scan for tests: un-stuff one entropy-coded segment.

T.81 B.1.1.5 / F.1.2.3: FF00 -> FF, FF FF -> fill, FF D0..D7 -> restart boundary, any other marker ends it.
Returns (unstuffed bytes, boundary byte offsets, marker values)."""
import numpy as np


def unstuff(raw):
    raw = bytes(raw)
    out = bytearray()
    bnd, marks = [], []
    i, n = 0, len(raw)
    while i < n:
        c = raw[i]
        if c != 0xFF:
            out.append(c)
            i += 1
            continue
        j = i + 1
        while j < n and raw[j] == 0xFF:
            j += 1
        if j >= n:
            break
        m = raw[j]
        if m == 0x00:
            out.append(0xFF)
        elif 0xD0 <= m <= 0xD7:
            bnd.append(len(out))
            marks.append(m)
        else:
            break
        i = j + 1
    return bytes(out), bnd, marks


def words_to_bytes(words_u8, nbytes):
    """Undo the big-endian word layout (byte i of the stream sits at address i ^ 3)."""
    w = np.frombuffer(bytes(words_u8), dtype=np.uint8)
    n4 = (nbytes + 3) // 4 * 4
    return w[:n4].reshape(-1, 4)[:, ::-1].reshape(-1)[:nbytes].tobytes()

"""A second, independent statement of the deployed LoRa receive chain in plain Python -- written before
oracle/std_chain.c existed (it is the script that first decoded the golden capture) and kept as a cross-check of it.
Explicit header only; test infrastructure, small inputs."""


def whitening_seq(n):
    s, v = [], 0xFF
    for _ in range(n):
        s.append(v)
        fb = ((v >> 7) ^ (v >> 5) ^ (v >> 4) ^ (v >> 3)) & 1
        v = ((v << 1) | fb) & 0xFF
    return s


def deinterleave(sym, sf_app, cw_len):
    inter = [[(s >> (sf_app - 1 - j)) & 1 for j in range(sf_app)] for s in sym]
    de = [[0] * cw_len for _ in range(sf_app)]
    for i in range(cw_len):
        for j in range(sf_app):
            de[(i - j - 1) % sf_app][i] = inter[i][j]
    return [int("".join(map(str, r)), 2) for r in de]


def ham_dec(cw, cw_len):
    c = [(cw >> (cw_len - 1 - i)) & 1 for i in range(cw_len)]
    return (c[3] << 3) | (c[2] << 2) | (c[1] << 1) | c[0]


def crc16(data):
    crc = 0
    for b in data:
        for i in range(8):
            if ((crc >> 15) & 1) ^ ((b >> (7 - i)) & 1):
                crc = ((crc << 1) ^ 0x1021) & 0xFFFF
            else:
                crc = (crc << 1) & 0xFFFF
    return crc


def decode(symbols, sf, ldro=False):
    """-> (length, has_crc, cr, header checksum, payload bytes, received crc, crc ok); no error correction."""
    n = 1 << sf
    s = [(int(x) - 1) % n for x in symbols]
    hdr = [v >> 2 for v in s[:8]]
    hdr = [v ^ (v >> 1) for v in hdr]
    nib = [ham_dec(c, 8) for c in deinterleave(hdr, sf - 2, 8)]
    ln, has_crc, cr = (nib[0] << 4) | nib[1], nib[2] & 1, nib[2] >> 1
    chk = ((nib[3] & 1) << 4) | nib[4]
    rest = nib[5:]
    cw_len, app = 4 + cr, (sf - 2 if ldro else sf)
    pay = s[8:]
    for b in range(len(pay) // cw_len):
        blk = pay[b * cw_len:(b + 1) * cw_len]
        if ldro:
            blk = [v >> 2 for v in blk]
        blk = [v ^ (v >> 1) for v in blk]
        rest += [ham_dec(c, cw_len) for c in deinterleave(blk, app, cw_len)]
    w = whitening_seq(ln)
    out = [((rest[2 * i + 1] << 4) | rest[2 * i]) ^ w[i] for i in range(ln)]
    crc_rx, ok = None, None
    if has_crc:
        q = 2 * ln
        crc_rx = rest[q] | (rest[q + 1] << 4) | (rest[q + 2] << 8) | (rest[q + 3] << 12)
        c = crc16(out[:-2]) ^ (out[-1] if ln >= 1 else 0) ^ ((out[-2] << 8) if ln >= 2 else 0)
        ok = c == crc_rx
    return ln, has_crc, cr, chk, bytes(out), crc_rx, ok

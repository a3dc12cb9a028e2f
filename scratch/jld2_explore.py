import struct, sys
d = open(sys.argv[1], 'rb').read()
# superblock at 512
sb = 512
print('sb version', d[sb+8], 'offsz', d[sb+9], 'lensz', d[sb+10])
base, ext, eof, root = struct.unpack_from('<QQQQ', d, sb+12)
print('base', base, 'eof', eof, 'root', root)
def ohdr(off, depth=0):
    off += base
    assert d[off:off+4] == b'OHDR', off
    ver, flags = d[off+4], d[off+5]
    p = off+6
    if flags & 0x20: p += 16
    if flags & 0x10: p += 4
    szw = 1 << (flags & 3)
    size = int.from_bytes(d[p:p+szw], 'little'); p += szw
    msgs = []
    def parse(p, end):
        while p + 4 <= end:
            t = d[p]; sz = struct.unpack_from('<H', d, p+1)[0]; fl = d[p+3]; p += 4
            if flags & 0x04: p += 2
            body = d[p:p+sz]
            if t == 0x10:
                co, cl = struct.unpack_from('<QQ', body)
                co += base
                assert d[co:co+4] == b'OCHK'
                parse(co+4, co+cl-4)
            elif t != 0:
                msgs.append((t, fl, p, body))
            p += sz
    parse(p, p+size)
    return msgs
for off in [root]:
    for t, fl, p, body in ohdr(off):
        print(hex(t), fl, p, len(body), body[:60])
print('--- traj')
toff = 0x35c9
for t, fl, p, body in ohdr(toff):
    print(hex(t), fl, p, len(body), body[:80])

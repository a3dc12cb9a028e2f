"""Minimal reader for the JLD2 trajectory fixtures the reference ships (`docs/data/*.jld2`).

SURVEY.md §8f-4 / §8c: the reference's documentation stores *solved* `NamedTrajectory` objects
(`save(..., "traj", prob.trajectory)`; loaded again in e.g.
`docs/literate/examples/multilevel_transmon.jl`).  They are the only artefacts in the reference
tree produced by the real Julia evaluator, so replaying them through the defect kernel pins the
layout and sign conventions of the path against the reference itself (tests/test_reference_fixtures.py).

JLD2 is a subset of HDF5 (superblock v2 at byte 512, version-2 object headers, committed compound
datatypes, global heap collections for variable-length data).  No HDF5 library is installed in this
image, so this module decodes exactly that subset with `struct`:

* object headers `OHDR` + continuation blocks `OCHK`; messages dataspace (0x01), datatype (0x03,
  inline or shared -> committed), layout (0x08: compact / contiguous), link (0x06), attribute (0x0C);
* datatypes: fixed point, float, string, compound, reference, variable length;
* object references are file offsets relative to the superblock's base address;
* variable-length values live in global heap collections `GCOL`.

It returns plain Python values: numbers, `numpy` arrays, strings, and `dict`s keyed by Julia field
name for structs / NamedTuples (tuple fields are keyed "1", "2", ...).  Nothing here is on the product
path; it is fixture tooling for tests and `tests/golden/make_golden.py`.
"""
from __future__ import annotations

import struct

import numpy as np

__all__ = ["load", "named_trajectory"]


class _File:
    def __init__(self, data: bytes):
        self.d = data
        if not data.startswith(b"HDF5-based Julia Data Format"):
            raise ValueError("not a JLD2 file")
        sb = data.find(b"\x89HDF\r\n\x1a\n")
        if sb < 0 or data[sb + 8] not in (2, 3) or data[sb + 9] != 8 or data[sb + 10] != 8:
            raise ValueError("unsupported HDF5 superblock")
        self.base, _ext, _eof, self.root = struct.unpack_from("<QQQQ", data, sb + 12)
        self._types = {}

    # ---- object headers ---------------------------------------------------------------------
    def messages(self, off: int):
        d = self.d
        off += self.base
        if d[off:off + 4] != b"OHDR" or d[off + 4] != 2:
            raise ValueError(f"no version-2 object header at {off}")
        flags = d[off + 5]
        p = off + 6
        if flags & 0x20:
            p += 16
        if flags & 0x10:
            p += 4
        w = 1 << (flags & 3)
        size = int.from_bytes(d[p:p + w], "little")
        p += w
        out = []

        def walk(p, end):
            while p + 4 <= end:
                t = d[p]
                sz = struct.unpack_from("<H", d, p + 1)[0]
                fl = d[p + 3]
                p += 4
                if flags & 0x04:
                    p += 2
                if t == 0x10:
                    co, cl = struct.unpack_from("<QQ", d, p)
                    co += self.base
                    if d[co:co + 4] != b"OCHK":
                        raise ValueError("bad continuation block")
                    walk(co + 4, co + cl - 4)
                elif t != 0:
                    out.append((t, fl, p, sz))
                p += sz

        walk(p, p + size)
        return out

    # ---- datatypes ---------------------------------------------------------------------------
    def datatype(self, p: int):
        """Parse a datatype message at absolute position p -> (descr, end)."""
        d = self.d
        cls, ver = d[p] & 0x0F, d[p] >> 4
        b0, b1 = d[p + 1], d[p + 2]
        size = struct.unpack_from("<I", d, p + 4)[0]
        q = p + 8
        if cls == 0:
            return {"k": "int", "size": size, "signed": bool(b0 & 0x08)}, q + 4
        if cls == 1:
            return {"k": "float", "size": size}, q + 12
        if cls == 3:
            return {"k": "string", "size": size}, q
        if cls == 4:
            return {"k": "bits", "size": size}, q + 4
        if cls == 5:
            return {"k": "opaque", "size": size}, q + ((b0 + 7) // 8) * 8
        if cls == 7:
            return {"k": "ref", "size": size}, q
        if cls == 9:
            base, q = self.datatype(q)
            return {"k": "vlen", "size": size, "base": base, "string": (b0 & 0x0F) == 1}, q
        if cls == 6:
            n = b0 | (b1 << 8)
            members = []
            ow = 1
            while (1 << (8 * ow)) <= size and ow < 4:
                ow += 1
            for _ in range(n):
                e = d.index(b"\x00", q)
                name = d[q:e].decode("utf-8")
                if ver >= 3:
                    q = e + 1
                    moff = int.from_bytes(d[q:q + ow], "little")
                    q += ow
                else:
                    q += ((e + 1 - q) + 7) // 8 * 8
                    moff = struct.unpack_from("<I", d, q)[0]
                    q += 4
                    if ver == 1:
                        q += 28
                mt, q = self.datatype(q)
                members.append((name, moff, mt))
            return {"k": "compound", "size": size, "members": members}, q
        raise ValueError(f"unsupported datatype class {cls}")

    def committed(self, off: int):
        """Committed datatype at object offset `off` -> (descr, attributes)."""
        if off not in self._types:
            # the attributes (julia_type, field_names, ...) are not needed to decode values, and the
            # type of `julia_type` refers back to itself, so they are left alone
            descr = None
            for t, fl, p, sz in self.messages(off):
                if t == 0x03:
                    descr, _ = self.datatype(p)
            self._types[off] = (descr, {})
        return self._types[off]

    def message_datatype(self, fl: int, p: int):
        if fl & 0x02:     # shared: version, type, address of the committed datatype
            addr = struct.unpack_from("<Q", self.d, p + 2)[0]
            descr, _ = self.committed(addr)
            return descr, addr
        return self.datatype(p)[0], None

    # ---- values ------------------------------------------------------------------------------
    def heap_object(self, addr: int, index: int) -> bytes:
        d = self.d
        p = addr + self.base
        if d[p:p + 4] != b"GCOL":
            raise ValueError("bad global heap collection")
        total = struct.unpack_from("<Q", d, p + 8)[0]
        q, end = p + 16, p + total
        while q + 16 <= end:
            idx, _rc, _res, sz = struct.unpack_from("<HHIQ", d, q)
            if idx == 0:
                break
            if idx == index:
                return d[q + 16:q + 16 + sz]
            q += 16 + (sz + 7) // 8 * 8
        raise KeyError((addr, index))

    def value(self, descr, buf: bytes, p: int = 0):
        k = descr["k"]
        if k == "int":
            return int.from_bytes(buf[p:p + descr["size"]], "little", signed=descr["signed"])
        if k == "float":
            return struct.unpack_from("<d" if descr["size"] == 8 else "<f", buf, p)[0]
        if k == "string":
            return buf[p:p + descr["size"]].split(b"\x00")[0].decode("utf-8")
        if k in ("bits", "opaque"):
            return bytes(buf[p:p + descr["size"]])
        if k == "ref":
            off = struct.unpack_from("<Q", buf, p)[0]
            return None if off in (0, 0xFFFFFFFFFFFFFFFF) else self.dataset(off)
        if k == "vlen":
            n, addr, idx = struct.unpack_from("<IQI", buf, p)
            if n == 0 and addr == 0:
                return "" if descr["string"] else []
            raw = self.heap_object(addr, idx)
            if descr["string"]:
                return raw[:n].decode("utf-8")
            bs = descr["base"]["size"]
            return [self.value(descr["base"], raw, i * bs) for i in range(n)]
        if k == "compound":
            return {name: self.value(mt, buf, p + moff) for name, moff, mt in descr["members"]}
        raise ValueError(k)

    def attribute(self, p: int):
        d = self.d
        ver = d[p]
        aflags = d[p + 1]
        nsz, tsz, ssz = struct.unpack_from("<HHH", d, p + 2)
        q = p + 8 + (1 if ver == 3 else 0)
        name = d[q:q + nsz].split(b"\x00")[0].decode("utf-8")
        q += nsz
        if aflags & 0x01:
            addr = struct.unpack_from("<Q", d, q + 2)[0]
            descr, _ = self.committed(addr)
        else:
            descr, _ = self.datatype(q)
        q += tsz
        dims = self.dataspace(q)
        q += ssz
        count = int(np.prod(dims)) if dims else 1
        try:
            vals = [self.value(descr, d, q + i * descr["size"]) for i in range(count)]
        except Exception:      # attributes are informational (julia_type ...): never fatal
            vals = [None]
        return name, (vals[0] if not dims else vals)

    def dataspace(self, p: int):
        d = self.d
        ver, nd, fl = d[p], d[p + 1], d[p + 2]
        q = p + (4 if ver == 2 else 8)
        if ver == 2 and d[p + 3] == 2:
            return None          # null dataspace
        return [struct.unpack_from("<Q", d, q + 8 * i)[0] for i in range(nd)]

    def dataset(self, off: int):
        d = self.d
        dims, descr, raw = [], None, None
        for t, fl, p, sz in self.messages(off):
            if t == 0x01:
                dims = self.dataspace(p)
            elif t == 0x03:
                descr, _ = self.message_datatype(fl, p)
            elif t == 0x08:
                ver, cls = d[p], d[p + 1]
                if ver not in (3, 4):
                    raise ValueError("unsupported layout version")
                if cls == 0:
                    n = struct.unpack_from("<H", d, p + 2)[0]
                    raw = d[p + 4:p + 4 + n]
                elif cls == 1:
                    addr, n = struct.unpack_from("<QQ", d, p + 2)
                    raw = b"" if addr == 0xFFFFFFFFFFFFFFFF else d[addr + self.base:addr + self.base + n]
                else:
                    raise ValueError("chunked datasets are not supported")
        if descr is None:
            raise ValueError(f"object at {off} is not a dataset")
        if dims is None:
            return None
        if raw is None:
            raw = b""
        if not dims:
            return self.value(descr, raw, 0) if raw else ({} if descr["k"] == "compound" else None)
        count = int(np.prod(dims))
        if descr["k"] == "float" and descr["size"] == 8:
            a = np.frombuffer(raw, dtype="<f8", count=count).copy()
        elif descr["k"] == "int" and descr["size"] == 8:
            a = np.frombuffer(raw, dtype="<i8" if descr["signed"] else "<u8", count=count).copy()
        else:
            vals = [self.value(descr, raw, i * descr["size"]) for i in range(count)]
            return vals
        # HDF5 lists dimensions slowest first; Julia arrays are column-major with dims reversed
        return a.reshape(dims[::-1], order="F") if len(dims) > 1 else a

    def links(self, off: int):
        d = self.d
        out = {}
        for t, fl, p, sz in self.messages(off):
            if t != 0x06:
                continue
            lf = d[p + 1]
            q = p + 2
            ltype = 0
            if lf & 0x08:
                ltype = d[q]
                q += 1
            if lf & 0x04:
                q += 8
            if lf & 0x10:
                q += 1
            w = 1 << (lf & 3)
            n = int.from_bytes(d[q:q + w], "little")
            q += w
            name = d[q:q + n].decode("utf-8")
            q += n
            if ltype == 0:
                out[name] = struct.unpack_from("<Q", d, q)[0]
        return out


def load(path: str) -> dict:
    """All top-level datasets of a JLD2 file as {name: value} (groups such as `_types` are skipped)."""
    with open(path, "rb") as fh:
        f = _File(fh.read())
    out = {}
    for name, off in f.links(f.root).items():
        if name.startswith("_"):
            continue
        out[name] = f.dataset(off)
    return out


def _ordered(nt: dict):
    return list(nt.values())


def named_trajectory(obj: dict) -> dict:
    """Normalise a decoded `NamedTrajectory` struct (NamedTrajectories.jl 0.x field names `datavec`,
    `dims`, `components`, `names`, `timestep`, ...) into
    {"data": (dim, N) array, "N", "dim", "components": {name: 0-based index array}, "names", "timestep", raw}."""
    datavec = np.asarray(obj["datavec"], dtype=np.float64)
    comps = {}
    for name, rng in obj["components"].items():
        if isinstance(rng, dict):                 # UnitRange{Int64}: start, stop
            comps[name] = np.arange(rng["start"] - 1, rng["stop"])
        else:
            comps[name] = np.asarray(rng, dtype=np.int64) - 1
    dim = int(obj["dim"]) if "dim" in obj else int(max(v.max() for v in comps.values()) + 1)
    N = int(obj["N"]) if "N" in obj else (int(obj["T"]) if "T" in obj else datavec.size // dim)
    data = datavec[:dim * N].reshape(dim, N, order="F")
    return {"data": data, "N": N, "dim": dim, "components": comps, "names": list(obj.get("names", {}).values())
            if isinstance(obj.get("names"), dict) else obj.get("names"), "timestep": obj.get("timestep"),
            "global_data": obj.get("global_data"), "raw": obj}

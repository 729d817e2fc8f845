"""A self-contained reader / writer for the subset of HDF5 that Keras weight files use, with an h5py-like surface
(`File`, groups, `create_group`, `create_dataset`, `attrs`), used by keras_io.py when h5py is not installed.

Subset (HDF5 File Format Specification, version-0 structures, what libhdf5 writes with its default `libver=earliest`
and what every libhdf5 reads): superblock v0, old-style groups (symbol-table message -> v1 B-tree of SNOD nodes + local
heap), v1 object headers (with continuation blocks when reading), dataspace v1/v2, datatypes fixed-point / IEEE float /
fixed-length string, contiguous (and compact) data layout v3, attribute messages v1-v3.  Not supported: chunked or
filtered datasets, variable-length types (such attributes are skipped when reading), new-style groups, v2 headers.

Writing keeps every group in ONE symbol-table node: the file's "group leaf node K" (a superblock field every reader
honours) is raised to fit the largest group (the parcel encoder has ~350 children).
"""
import struct
import numpy as np

UNDEF = 0xFFFFFFFFFFFFFFFF
SIG = b"\x89HDF\r\n\x1a\n"


def _pad8(n):
    return (n + 7) & ~7


# ---------------------------------------------------------------------------------------------------------------------
# in-memory tree
# ---------------------------------------------------------------------------------------------------------------------
class _Attrs(dict):
    def __setitem__(self, k, v):
        dict.__setitem__(self, k, np.asarray(v))


class Dataset:
    def __init__(self, data):
        self._data = np.ascontiguousarray(data)
        self.attrs = _Attrs()
        self.shape, self.dtype = self._data.shape, self._data.dtype

    def __array__(self, dtype=None, copy=None):
        return self._data if dtype is None else self._data.astype(dtype)

    def __getitem__(self, k):
        return self._data[k]


class Group:
    def __init__(self):
        self._children = {}
        self.attrs = _Attrs()

    def create_group(self, name):
        g = self
        for part in [p for p in name.split("/") if p]:
            nxt = g._children.get(part)
            if nxt is None:
                nxt = g._children[part] = Group()
            if not isinstance(nxt, Group):
                raise ValueError(f"{part} is a dataset")
            g = nxt
        return g

    def create_dataset(self, name, data=None, **_):
        parts = [p for p in name.split("/") if p]
        g = self.create_group("/".join(parts[:-1])) if len(parts) > 1 else self
        if parts[-1] in g._children:
            raise ValueError(f"{name} exists")
        d = g._children[parts[-1]] = Dataset(np.asarray(data))
        return d

    def __getitem__(self, name):
        o = self
        for part in [p for p in name.split("/") if p]:
            o = o._children[part]
        return o

    def __contains__(self, name):
        try:
            self[name]
            return True
        except (KeyError, AttributeError):
            return False

    def keys(self):
        return list(self._children)


class File(Group):
    def __init__(self, path, mode="r"):
        Group.__init__(self)
        self._path, self._mode = path, mode
        if mode == "r":
            with open(path, "rb") as f:
                _Reader(f.read()).load_root(self)
        elif mode != "w":
            raise ValueError("mode must be 'r' or 'w'")

    def close(self):
        if self._mode == "w":
            with open(self._path, "wb") as f:
                f.write(_Writer(self).serialise())
            self._mode = "closed"

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        if exc[0] is None:
            self.close()
        return False


# ---------------------------------------------------------------------------------------------------------------------
# datatype / dataspace messages
# ---------------------------------------------------------------------------------------------------------------------
def _dtype_msg(dt):
    dt = np.dtype(dt)
    if dt.kind == "f":
        exp_bits, man_bits, bias = {4: (8, 23, 127), 8: (11, 52, 1023), 2: (5, 10, 15)}[dt.itemsize]
        bits = dt.itemsize * 8
        return struct.pack("<BBBBI", 0x11, 0x20, bits - 1, 0, dt.itemsize) + \
            struct.pack("<HHBBBBI", 0, bits, man_bits, exp_bits, 0, man_bits, bias)
    if dt.kind in "iu":
        return struct.pack("<BBBBI", 0x10, 0x08 if dt.kind == "i" else 0x00, 0, 0, dt.itemsize) + \
            struct.pack("<HH", 0, dt.itemsize * 8)
    if dt.kind == "S":
        return struct.pack("<BBBBI", 0x13, 0x01, 0, 0, max(dt.itemsize, 1))          # null-padded ASCII
    raise TypeError(f"h5min cannot store dtype {dt}")


def _parse_dtype(b):
    cls, ver = b[0] & 0x0F, b[0] >> 4
    size = struct.unpack_from("<I", b, 4)[0]
    if cls == 1:
        return np.dtype(f"{'>' if b[1] & 1 else '<'}f{size}")
    if cls == 0:
        return np.dtype(f"{'>' if b[1] & 1 else '<'}{'i' if b[1] & 0x08 else 'u'}{size}")
    if cls == 3:
        return np.dtype(f"S{size}")
    return None                                                                      # vlen, compound, ...: unsupported


def _dspace_msg(shape):
    return struct.pack("<BBBB4x", 1, len(shape), 0, 0) + b"".join(struct.pack("<Q", int(d)) for d in shape)


def _parse_dspace(b):
    ver, rank = b[0], b[1]
    off = 8 if ver == 1 else 4
    if ver == 2 and b[3] == 2:                                                       # null dataspace
        return (0,)
    return tuple(struct.unpack_from("<Q", b, off + 8 * i)[0] for i in range(rank))


# ---------------------------------------------------------------------------------------------------------------------
# writer
# ---------------------------------------------------------------------------------------------------------------------
class _Writer:
    BT_K = 16                                                                        # group internal node K

    def __init__(self, root):
        self.root = root
        biggest = [1]

        def walk(g):
            biggest[0] = max(biggest[0], len(g._children))
            for c in g._children.values():
                if isinstance(c, Group):
                    walk(c)
        walk(root)
        self.leaf_k = max(4, (biggest[0] + 1) // 2)
        self.snod_size = 8 + 2 * self.leaf_k * 40
        self.btree_size = 24 + (2 * self.BT_K + 1) * 8 + 2 * self.BT_K * 8

    @staticmethod
    def _msg(mtype, data):
        return struct.pack("<HHB3x", mtype, _pad8(len(data)), 0) + data + b"\0" * (_pad8(len(data)) - len(data))

    def _attr_msgs(self, attrs):
        out = []
        for name, val in attrs.items():
            val = np.asarray(val)
            if val.dtype.kind == "U":
                val = np.char.encode(val, "utf8")
            nb = name.encode() + b"\0"
            dt, ds = _dtype_msg(val.dtype), _dspace_msg(val.shape)
            body = struct.pack("<BBHHH", 1, 0, len(nb), len(dt), len(ds))
            body += nb + b"\0" * (_pad8(len(nb)) - len(nb)) + dt + b"\0" * (_pad8(len(dt)) - len(dt))
            body += ds + b"\0" * (_pad8(len(ds)) - len(ds)) + np.ascontiguousarray(val).tobytes()
            if len(body) > 65000:
                raise ValueError(f"attribute {name} exceeds the 64 KB object-header limit (Keras splits such lists)")
            out.append(self._msg(0x000C, body))
        return out

    @staticmethod
    def _ohdr(msgs):
        body = b"".join(msgs)
        return struct.pack("<BBHII4x", 1, 0, len(msgs), 1, len(body)) + body

    def serialise(self):
        # pass 1: sizes and addresses (depth first); pass 2: bytes
        plan = []                                                                    # (address, bytes-producing closure)
        cursor = [96]

        def alloc(n):
            a = cursor[0]
            cursor[0] = _pad8(a + n)
            return a

        def plan_dataset(d):
            raw = d._data.tobytes()
            msgs_wo_layout = [self._msg(0x0001, _dspace_msg(d._data.shape)), self._msg(0x0003, _dtype_msg(d._data.dtype)),
                              self._msg(0x0005, struct.pack("<BBBB", 2, 2, 2, 0))] + self._attr_msgs(d.attrs)
            layout_len = len(self._msg(0x0008, b"\0" * 18))
            oh_addr = alloc(16 + sum(len(m) for m in msgs_wo_layout) + layout_len)
            data_addr = alloc(len(raw)) if raw else UNDEF
            layout = self._msg(0x0008, struct.pack("<BBQQ", 3, 1, data_addr, len(raw)))
            plan.append((oh_addr, self._ohdr(msgs_wo_layout + [layout])))
            if raw:
                plan.append((data_addr, raw))
            return oh_addr

        def plan_group(g):
            names = sorted(g._children, key=lambda s: s.encode())
            heap, offs = bytearray(8), {}
            for n in names:
                offs[n] = len(heap)
                nb = n.encode() + b"\0"
                heap += nb + b"\0" * (_pad8(len(nb)) - len(nb))
            attr_msgs = self._attr_msgs(g.attrs)
            oh_addr = alloc(16 + 24 + sum(len(m) for m in attr_msgs))
            bt_addr = alloc(self.btree_size)
            heap_addr = alloc(32)
            hdata_addr = alloc(len(heap))
            snod_addr = alloc(self.snod_size) if names else UNDEF
            entries = []
            for n in names:
                c = g._children[n]
                if isinstance(c, Group):
                    c_oh, c_bt, c_heap = plan_group(c)
                    entries.append(struct.pack("<QQII", offs[n], c_oh, 1, 0) + struct.pack("<QQ", c_bt, c_heap))
                else:
                    entries.append(struct.pack("<QQII", offs[n], plan_dataset(c), 0, 0) + b"\0" * 16)
            plan.append((oh_addr, self._ohdr([self._msg(0x0011, struct.pack("<QQ", bt_addr, heap_addr))] + attr_msgs)))
            bt = struct.pack("<4sBBHQQ", b"TREE", 0, 0, 1 if names else 0, UNDEF, UNDEF)
            if names:
                bt += struct.pack("<QQQ", 0, snod_addr, offs[names[-1]])
            bt += b"\0" * (self.btree_size - len(bt))
            plan.append((bt_addr, bt))
            plan.append((heap_addr, struct.pack("<4sB3xQQQ", b"HEAP", 0, len(heap), 1, hdata_addr)))
            plan.append((hdata_addr, bytes(heap)))
            if names:
                snod = struct.pack("<4sBBH", b"SNOD", 1, 0, len(names)) + b"".join(entries)
                plan.append((snod_addr, snod + b"\0" * (self.snod_size - len(snod))))
            return oh_addr, bt_addr, heap_addr

        r_oh, r_bt, r_heap = plan_group(self.root)
        eof = cursor[0]
        sb = SIG + struct.pack("<BBBBBBBB", 0, 0, 0, 0, 0, 8, 8, 0) + struct.pack("<HHI", self.leaf_k, self.BT_K, 0)
        sb += struct.pack("<QQQQ", 0, UNDEF, eof, UNDEF)
        sb += struct.pack("<QQII", 0, r_oh, 1, 0) + struct.pack("<QQ", r_bt, r_heap)
        assert len(sb) == 96
        buf = bytearray(eof)
        buf[0:96] = sb
        for addr, b in plan:
            buf[addr:addr + len(b)] = b
        return bytes(buf)


# ---------------------------------------------------------------------------------------------------------------------
# reader
# ---------------------------------------------------------------------------------------------------------------------
class _Reader:
    def __init__(self, buf):
        self.b = buf
        if buf[:8] != SIG:
            raise ValueError("not an HDF5 file (signature at offset 0 expected)")
        ver = buf[8]
        if ver not in (0, 1) or buf[13] != 8 or buf[14] != 8:
            raise ValueError(f"h5min reads superblock version 0/1 with 8-byte offsets (got version {ver})")
        off = 24 + (4 if ver == 1 else 0)
        self.base = struct.unpack_from("<Q", buf, off)[0]
        self.root_entry = off + 32

    def load_root(self, root):
        _, oh, _, _ = struct.unpack_from("<QQII", self.b, self.root_entry)
        self._load_group(oh + self.base, root)

    def _messages(self, addr):
        b = self.b
        if b[addr:addr + 4] == b"OHDR":
            raise ValueError("h5min does not read version-2 object headers (file written with libver='latest')")
        ver, _, nmsg, _, size = struct.unpack_from("<BBHII", b, addr)
        if ver != 1:
            raise ValueError(f"object header version {ver} not supported")
        chunks, out = [(addr + 16, size)], []
        while chunks and len(out) < nmsg:
            pos, left = chunks.pop(0)
            end = pos + left
            while pos + 8 <= end and len(out) < nmsg:
                mtype, msize, _ = struct.unpack_from("<HHB", b, pos)
                data = b[pos + 8:pos + 8 + msize]
                pos += 8 + msize
                if mtype == 0x0010:
                    c_off, c_len = struct.unpack_from("<QQ", data, 0)
                    chunks.append((c_off + self.base, c_len))
                out.append((mtype, data))
        return out

    def _attr(self, data):
        ver = data[0]
        nlen, dlen, slen = struct.unpack_from("<HHH", data, 2)
        pos = 8 + (1 if ver == 3 else 0)
        pad = _pad8 if ver == 1 else (lambda n: n)
        name = data[pos:pos + nlen].split(b"\0")[0].decode()
        pos += pad(nlen)
        dt = _parse_dtype(data[pos:pos + dlen])
        pos += pad(dlen)
        shape = _parse_dspace(data[pos:pos + slen])
        pos += pad(slen)
        if dt is None:
            return name, None
        n = int(np.prod(shape)) if shape else 1
        return name, np.frombuffer(data, dtype=dt, count=n, offset=pos).reshape(shape).copy()

    def _fill_attrs(self, obj, msgs):
        for mtype, data in msgs:
            if mtype == 0x000C:
                k, v = self._attr(data)
                if v is not None:
                    dict.__setitem__(obj.attrs, k, v)

    def _heap_name(self, heap_addr, off):
        if self.b[heap_addr:heap_addr + 4] != b"HEAP":
            raise ValueError("bad local heap signature")
        data_addr = struct.unpack_from("<Q", self.b, heap_addr + 24)[0] + self.base
        end = self.b.index(b"\0", data_addr + off)
        return self.b[data_addr + off:end].decode()

    def _btree_snods(self, addr):
        b = self.b
        if b[addr:addr + 4] != b"TREE":
            raise ValueError("bad B-tree signature")
        _, level, used = struct.unpack_from("<BBH", b, addr + 4)
        for i in range(used):
            child = struct.unpack_from("<Q", b, addr + 24 + 8 + 16 * i)[0] + self.base
            if level > 0:
                yield from self._btree_snods(child)
            else:
                yield child

    def _load_group(self, oh_addr, group):
        msgs = self._messages(oh_addr)
        self._fill_attrs(group, msgs)
        st = [d for t, d in msgs if t == 0x0011]
        if not st:
            if any(t in (0x0002, 0x0006) for t, _ in msgs):
                raise ValueError("h5min does not read new-style (link-message) groups")
            return
        bt, heap = struct.unpack_from("<QQ", st[0], 0)
        for snod in self._btree_snods(bt + self.base):
            if self.b[snod:snod + 4] != b"SNOD":
                raise ValueError("bad symbol-table node signature")
            n = struct.unpack_from("<H", self.b, snod + 6)[0]
            for i in range(n):
                noff, oh, _, _ = struct.unpack_from("<QQII", self.b, snod + 8 + 40 * i)
                name = self._heap_name(heap + self.base, noff)
                group._children[name] = self._load_object(oh + self.base)

    def _load_object(self, oh_addr):
        msgs = self._messages(oh_addr)
        kinds = {t for t, _ in msgs}
        if 0x0008 in kinds:
            dt = shape = layout = None
            for t, d in msgs:
                if t == 0x0003:
                    dt = _parse_dtype(d)
                elif t == 0x0001:
                    shape = _parse_dspace(d)
                elif t == 0x0008:
                    layout = d
            if dt is None:
                raise ValueError("dataset of an unsupported datatype")
            if layout[0] != 3:
                raise ValueError(f"data layout message version {layout[0]} not supported")
            n = int(np.prod(shape)) if shape else 1
            if layout[1] == 1:
                addr, size = struct.unpack_from("<QQ", layout, 2)
                arr = (np.zeros(shape, dt) if addr == UNDEF or n == 0 else
                       np.frombuffer(self.b, dtype=dt, count=n, offset=addr + self.base).reshape(shape).copy())
            elif layout[1] == 0:
                size = struct.unpack_from("<H", layout, 2)[0]
                arr = np.frombuffer(layout, dtype=dt, count=n, offset=4).reshape(shape).copy()
            else:
                raise ValueError("chunked datasets are not supported (Keras writes contiguous ones)")
            d = Dataset(arr.astype(arr.dtype.newbyteorder("=")))
            self._fill_attrs(d, msgs)
            return d
        g = Group()
        self._load_group(oh_addr, g)
        return g

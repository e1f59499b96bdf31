"""Minimal HDF5 reader/writer for Keras weight files — no h5py / libhdf5 exists in this image.

Covers exactly what ``Sequence2Sequence.save`` / ``load_config`` / ``load_weights``
(/root/reference/ocrd_cor_asv_ann/lib/seq2seq.py:1121-1174) put into or take out of a model file
with h5py<3 and default ``libver``: superblock version 0, "old-style" groups (symbol table message,
B-tree v1 ``TREE`` nodes, ``SNOD`` symbol nodes, local ``HEAP``), version-1 object headers with
continuation blocks, contiguous or compact little/big-endian integer, float, enum (numpy bool) and
fixed-length string datasets, and version 1-3 attribute messages.
Anything else (chunked/compressed layouts, new-style groups, variable-length strings) raises
``H5Unsupported`` or is returned as ``None`` for attributes the model loader never reads.

    root = read(path)            # Group: dict name -> Group | numpy array, plus .attrs
    write(path, root)
"""
import struct

import numpy as np

UNDEF = 0xFFFFFFFFFFFFFFFF
SIGNATURE = b"\x89HDF\r\n\x1a\n"


class H5Unsupported(NotImplementedError):
    pass


class Group(dict):
    """A group: children by name; ``attrs`` holds the group's attributes."""

    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        self.attrs = {}

    def require_group(self, name):
        node = self
        for part in name.strip('/').split('/'):
            if part not in node:
                node[part] = Group()
            node = node[part]
        return node

    def get_path(self, name):
        node = self
        for part in name.strip('/').split('/'):
            node = node[part]
        return node


class Dataset(np.ndarray):
    """ndarray that can carry attributes (Keras puts none on datasets, but files may)."""
    attrs = None


# =============================================================================== reader
class _Reader:
    def __init__(self, buf):
        self.b = buf
        if buf[:8] != SIGNATURE:
            raise ValueError("not an HDF5 file (bad signature)")
        ver = buf[8]
        if ver not in (0, 1):
            raise H5Unsupported("HDF5 superblock version %d (only 0/1: h5py default libver)" % ver)
        self.so, self.sl = buf[13], buf[14]
        if self.so != 8 or self.sl != 8:
            raise H5Unsupported("offset/length sizes other than 8")
        self.leaf_k, self.int_k = struct.unpack_from("<HH", buf, 16)
        pos = 24 + (4 if ver == 1 else 0)
        self.base, _, self.eof, _ = struct.unpack_from("<QQQQ", buf, pos)
        pos += 32
        # root group symbol table entry
        _, self.root_header, cache, _ = struct.unpack_from("<QQII", buf, pos)

    # ---- object headers
    def messages(self, addr):
        b = self.b
        addr += self.base
        if b[addr:addr + 4] == b"OHDR":
            raise H5Unsupported("version-2 object headers (file written with libver='latest')")
        ver, _, nmsg, _, size = struct.unpack_from("<BBHII", b, addr)
        if ver != 1:
            raise H5Unsupported("object header version %d" % ver)
        out = []
        blocks = [(addr + 16, size)]
        while blocks and len(out) < nmsg:
            pos, left = blocks.pop(0)
            end = pos + left
            while pos + 8 <= end and len(out) < nmsg:
                mtype, msize, flags = struct.unpack_from("<HHB", b, pos)
                data = bytes(b[pos + 8: pos + 8 + msize])
                pos += 8 + msize
                if mtype == 0x10:
                    off, ln = struct.unpack_from("<QQ", data, 0)
                    blocks.append((off + self.base, ln))
                out.append((mtype, flags, data))
        return out

    def load(self, addr):
        msgs = self.messages(addr)
        types = [m[0] for m in msgs]
        if 0x11 in types:
            node = Group()
            data = msgs[types.index(0x11)][2]
            btree, heap = struct.unpack_from("<QQ", data, 0)
            for name, child in self.group_entries(btree, heap):
                node[name] = self.load(child)
        elif 0x08 in types:
            node = self.dataset(msgs)
        elif 0x02 in types or 0x06 in types:
            raise H5Unsupported("new-style groups (link messages)")
        else:
            node = Group()
        attrs = {}
        for mtype, flags, data in msgs:
            if mtype == 0x0C:
                k, v = self.attribute(data)
                attrs[k] = v
        if isinstance(node, Group):
            node.attrs = attrs
        elif attrs:
            node.attrs = attrs
        return node

    # ---- groups
    def heap_data(self, heap):
        b = self.b
        heap += self.base
        if b[heap:heap + 4] != b"HEAP":
            raise ValueError("bad local heap signature")
        size, _, data_addr = struct.unpack_from("<QQQ", b, heap + 8)
        return bytes(b[data_addr + self.base: data_addr + self.base + size])

    def group_entries(self, btree, heap):
        names = self.heap_data(heap)
        out = []
        self._walk_tree(btree, names, out)
        return out

    def _walk_tree(self, addr, names, out):
        b = self.b
        addr += self.base
        if b[addr:addr + 4] != b"TREE":
            raise ValueError("bad B-tree signature")
        ntype, level, used = struct.unpack_from("<BBH", b, addr + 4)
        if ntype != 0:
            raise H5Unsupported("non-group B-tree")
        pos = addr + 8 + 16           # skip sibling pointers
        children = []
        for i in range(used):
            pos += 8                  # key i
            children.append(struct.unpack_from("<Q", b, pos)[0])
            pos += 8
        for child in children:
            if level > 0:
                self._walk_tree(child, names, out)
            else:
                self._symbol_node(child, names, out)

    def _symbol_node(self, addr, names, out):
        b = self.b
        addr += self.base
        if b[addr:addr + 4] != b"SNOD":
            raise ValueError("bad symbol node signature")
        nsym = struct.unpack_from("<H", b, addr + 6)[0]
        pos = addr + 8
        for _ in range(nsym):
            name_off, header = struct.unpack_from("<QQ", b, pos)
            end = names.index(b"\0", name_off)
            out.append((bytes(names[name_off:end]).decode("utf-8"), header))
            pos += 40

    # ---- datatypes / dataspaces
    def datatype(self, data, pos=0):
        """-> (numpy dtype or None, bytes consumed)."""
        cv, b0, b1, b2, size = struct.unpack_from("<BBBBI", data, pos)
        cls, ver = cv & 0x0F, cv >> 4
        p = pos + 8
        order = ">" if (b0 & 1) else "<"
        if cls == 0:
            dt = np.dtype("%s%s%d" % (order, "i" if (b0 & 8) else "u", size))
            return dt, p + 4 - pos
        if cls == 1:
            return np.dtype("%sf%d" % (order, size)), p + 12 - pos
        if cls == 3:
            return np.dtype("S%d" % size), p - pos
        if cls == 8:
            nmemb = b0 | (b1 << 8)
            base, used = self.datatype(data, p)
            p += used
            names = []
            for _ in range(nmemb):
                end = data.index(b"\0", p)
                names.append(bytes(data[p:end]))
                ln = end - p + 1
                p += ln if ver >= 3 else (ln + 7) // 8 * 8
            p += nmemb * base.itemsize
            if sorted(names) == [b"FALSE", b"TRUE"]:
                return np.dtype(bool), p - pos
            return base, p - pos
        return None, 8          # variable-length, compound, ...: not needed by the model loader

    @staticmethod
    def dataspace(data):
        ver, rank, flags = struct.unpack_from("<BBB", data, 0)
        if ver == 1:
            p = 8
        elif ver == 2:
            if data[3] == 2:
                return None      # null dataspace
            p = 4
        else:
            raise H5Unsupported("dataspace version %d" % ver)
        return tuple(struct.unpack_from("<%dQ" % rank, data, p)) if rank else ()

    def _array(self, raw, dt, shape):
        n = int(np.prod(shape)) if shape else 1
        arr = np.frombuffer(raw, dtype=np.uint8 if dt == np.dtype(bool) else dt, count=n)
        if dt == np.dtype(bool):
            arr = arr.astype(bool)
        return arr.reshape(shape).copy()

    def dataset(self, msgs):
        dt = shape = None
        layout = None
        for mtype, flags, data in msgs:
            if mtype == 0x01:
                shape = self.dataspace(data)
            elif mtype == 0x03:
                dt, _ = self.datatype(data)
            elif mtype == 0x08:
                layout = data
            elif mtype == 0x0B:
                raise H5Unsupported("filtered (compressed) dataset")
        if dt is None:
            raise H5Unsupported("dataset datatype")
        ver, cls = layout[0], layout[1]
        if ver != 3:
            raise H5Unsupported("data layout version %d" % ver)
        n = (int(np.prod(shape)) if shape else 1) * dt.itemsize
        if cls == 1:
            addr, size = struct.unpack_from("<QQ", layout, 2)
            raw = b"\0" * n if addr == UNDEF else self.b[addr + self.base: addr + self.base + n]
        elif cls == 0:
            size = struct.unpack_from("<H", layout, 2)[0]
            raw = layout[4:4 + size]
        else:
            raise H5Unsupported("chunked dataset layout")
        return self._array(raw, dt, shape).view(Dataset)

    def attribute(self, data):
        ver = data[0]
        name_size, dt_size, ds_size = struct.unpack_from("<HHH", data, 2)
        p = 8
        if ver == 3:
            p += 1
        pad = (lambda n: (n + 7) // 8 * 8) if ver == 1 else (lambda n: n)
        name = bytes(data[p:p + name_size]).split(b"\0")[0].decode("utf-8")
        p += pad(name_size)
        dt, _ = self.datatype(data, p)
        p += pad(dt_size)
        shape = self.dataspace(data[p:p + ds_size])
        p += pad(ds_size)
        if dt is None or shape is None:
            return name, None
        return name, self._array(data[p:], dt, shape)


def read(path):
    """Whole file -> Group tree (datasets as numpy arrays)."""
    with open(path, "rb") as f:
        buf = memoryview(f.read())
    r = _Reader(buf)
    return r.load(r.root_header)


# =============================================================================== writer
LEAF_K = 64            # symbol nodes hold 2*LEAF_K entries: one node per group (h5py's default is 4)
INTERNAL_K = 16


def _pad8(b):
    return b + b"\0" * (-len(b) % 8)


def _datatype_msg(dt):
    dt = np.dtype(dt)
    if dt == np.dtype(bool):
        base = _datatype_msg(np.int8)
        names = _pad8(b"FALSE\0") + _pad8(b"TRUE\0")
        return struct.pack("<BBBBI", 0x18, 2, 0, 0, 1) + base + names + bytes([0, 1])
    order = 1 if dt.byteorder == ">" else 0
    if dt.kind in "iu":
        bits = order | (8 if dt.kind == "i" else 0)
        return struct.pack("<BBBBIHH", 0x10, bits, 0, 0, dt.itemsize, 0, dt.itemsize * 8)
    if dt.kind == "f":
        if dt.itemsize == 4:
            prop = struct.pack("<HHBBBBI", 0, 32, 23, 8, 0, 23, 127)
            sign = 31
        elif dt.itemsize == 8:
            prop = struct.pack("<HHBBBBI", 0, 64, 52, 11, 0, 52, 1023)
            sign = 63
        else:
            raise H5Unsupported("float%d" % (dt.itemsize * 8))
        return struct.pack("<BBBBI", 0x11, 0x20 | order, sign, 0, dt.itemsize) + prop
    if dt.kind == "S":
        return struct.pack("<BBBBI", 0x13, 0x01, 0, 0, max(dt.itemsize, 1))      # null-padded ASCII, like h5py
    raise H5Unsupported("dtype %s" % dt)


def _dataspace_msg(shape):
    rank = len(shape)
    return struct.pack("<BBB5x", 1, rank, 0) + struct.pack("<%dQ" % rank, *shape)


def _message(mtype, data, flags=0):
    data = _pad8(data)
    return struct.pack("<HHB3x", mtype, len(data), flags) + data


def _attribute_msg(name, value):
    arr = np.asarray(value)
    if arr.dtype.kind == "U":
        arr = np.char.encode(arr, "utf-8")
    if arr.dtype.kind == "S" and arr.dtype.itemsize == 0:
        arr = arr.astype("S1")
    nm = name.encode("utf-8") + b"\0"
    dt, ds = _datatype_msg(arr.dtype), _dataspace_msg(arr.shape)
    body = struct.pack("<BBHHH", 1, 0, len(nm), len(dt), len(ds)) + _pad8(nm) + _pad8(dt) + _pad8(ds)
    raw = (arr.astype(np.uint8) if arr.dtype == np.dtype(bool) else np.ascontiguousarray(arr)).tobytes()
    return _message(0x0C, body + raw)


class _Writer:
    def __init__(self):
        self.buf = bytearray(96)          # superblock placeholder

    def alloc(self, data):
        self.buf += b"\0" * (-len(self.buf) % 8)
        addr = len(self.buf)
        self.buf += data
        return addr

    def object_header(self, msgs):
        body = b"".join(msgs)
        # version-1 messages are limited to 64 KiB each, the header block is not
        head = struct.pack("<BBHII4x", 1, 0, len(msgs), 1, len(body))
        return self.alloc(head + body)

    def dataset(self, arr):
        arr = np.asarray(arr)
        if arr.dtype.kind == "U":
            arr = np.char.encode(arr, "utf-8")
        raw = (arr.astype(np.uint8) if arr.dtype == np.dtype(bool) else np.ascontiguousarray(arr)).tobytes()
        addr = self.alloc(raw) if raw else UNDEF
        msgs = [_message(0x01, _dataspace_msg(arr.shape)),
                _message(0x03, _datatype_msg(arr.dtype), flags=1),
                _message(0x05, struct.pack("<BBBB", 2, 2, 2, 0)),              # fill value v2: alloc late, write if set, undefined
                _message(0x08, struct.pack("<BBQQ", 3, 1, addr, len(raw)))]
        attrs = getattr(arr, "attrs", None) or {}
        msgs += [_attribute_msg(k, v) for k, v in attrs.items()]
        return self.object_header(msgs)

    def group(self, node):
        names = sorted(node, key=lambda s: s.encode("utf-8"))
        if len(names) > 2 * LEAF_K:
            raise H5Unsupported("more than %d entries in one group" % (2 * LEAF_K))
        children = [(n, self.group(node[n]) if isinstance(node[n], Group) else self.dataset(node[n])) for n in names]
        # local heap: offset 0 holds the empty string, then the names, 8-byte aligned
        heap = bytearray(8)
        offs = []
        for n, _ in children:
            offs.append(len(heap))
            heap += _pad8(n.encode("utf-8") + b"\0")
        free_off = len(heap)
        heap += struct.pack("<QQ", 1, 16)            # one free block: next = 1 (none), size 16
        data_addr = self.alloc(bytes(heap))
        heap_addr = self.alloc(b"HEAP" + struct.pack("<B3xQQQ", 0, len(heap), free_off, data_addr))
        snod = bytearray(b"SNOD" + struct.pack("<BBH", 1, 0, len(children)))
        for (n, addr), off in zip(children, offs):
            snod += struct.pack("<QQII16x", off, addr, 0, 0)
        snod += b"\0" * (8 + 40 * 2 * LEAF_K - len(snod))
        snod_addr = self.alloc(bytes(snod))
        tree = bytearray(b"TREE" + struct.pack("<BBHQQ", 0, 0, 1 if children else 0, UNDEF, UNDEF))
        tree += struct.pack("<QQQ", 0, snod_addr, offs[-1] if offs else 0)
        tree += b"\0" * (24 + (2 * INTERNAL_K + 1) * 8 + 2 * INTERNAL_K * 8 - len(tree))
        tree_addr = self.alloc(bytes(tree))
        msgs = [_message(0x11, struct.pack("<QQ", tree_addr, heap_addr))]
        msgs += [_attribute_msg(k, v) for k, v in node.attrs.items()]
        self.last_group = (tree_addr, heap_addr)
        return self.object_header(msgs)

    def finish(self, root):
        root_addr = self.group(root)
        tree_addr, heap_addr = self.last_group
        eof = len(self.buf)
        sb = SIGNATURE + struct.pack("<BBBBBBBBHHI", 0, 0, 0, 0, 0, 8, 8, 0, LEAF_K, INTERNAL_K, 0)
        sb += struct.pack("<QQQQ", 0, UNDEF, eof, UNDEF)
        sb += struct.pack("<QQII", 0, root_addr, 1, 0) + struct.pack("<QQ", tree_addr, heap_addr)
        assert len(sb) == 96
        self.buf[:96] = sb
        return bytes(self.buf)


def write(path, root):
    """Group tree -> HDF5 file (superblock 0, old-style groups, contiguous datasets)."""
    data = _Writer().finish(root)
    with open(path, "wb") as f:
        f.write(data)

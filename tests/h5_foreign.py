"""Hand-assembled HDF5 files in the shape libhdf5 1.10 / h5py 2.x (the reference's pinned stack) gives them.

No libhdf5 / h5py exists in this image, so the product's reader (``cor_asv_ann_b200/h5lite.py``) could only ever
read what its own writer produced.  This module is a SECOND, independent encoder written from the HDF5 File Format
Specification (version 1.x: III.A superblock v0, III.A.1 B-tree v1, III.B symbol nodes, III.D local heaps, IV.A.1.a
version-1 object headers, IV.A.2.* messages) that makes the layout choices libhdf5 makes and h5lite's writer does not:

  * group B-trees with the default ``LEAF_K = 4``: a group of more than 8 entries spreads over several ``SNOD`` leaves
    under one ``TREE`` node, and (forced here beyond ``split_at`` leaves) over a level-1 ``TREE`` root;
  * object headers that run out of their first block: a continuation message (0x0010) to a block elsewhere in the
    file, NIL messages (0x0000) filling gaps, the modification-time message (0x0012) libhdf5 adds to datasets;
  * symbol table entries with cache type 1 (B-tree / heap address in the scratch pad);
  * local heaps with a free list and unused (deleted) names in them;
  * dataspace messages that carry maximum dimensions; compact as well as contiguous layouts; big-endian data;
  * version-1 attribute messages with 8-byte padded name / datatype / dataspace, null-padded fixed strings longer than
    their content (``|S20``), and a variable-length string attribute (which Keras' loader never reads);
  * holes between objects, a non-zero user block free (objects not in definition order).

It is test infrastructure (like ``oracle/``): nothing in the product imports it.
"""
import struct

import numpy as np

UNDEF = 0xFFFFFFFFFFFFFFFF
LEAF_K, INTERNAL_K = 4, 16


def pad8(b):
    return b + b"\0" * (-len(b) % 8)


class Dset:
    def __init__(self, array, layout="contiguous", maxdims=False, attrs=None, big_endian=False):
        self.array, self.layout, self.maxdims, self.attrs, self.big_endian = np.asarray(array), layout, maxdims, attrs or {}, big_endian


class Grp:
    def __init__(self, children=None, attrs=None, dead_names=()):
        self.children, self.attrs, self.dead_names = children or {}, attrs or {}, tuple(dead_names)


class VlenStr:
    """An attribute value stored as an HDF5 variable-length string (global heap reference)."""
    def __init__(self, text):
        self.text = text


# ------------------------------------------------------------------------------------------------ messages
def datatype(dt, big_endian=False):
    dt = np.dtype(dt)
    order = 1 if (big_endian or dt.byteorder == ">") else 0
    if dt == np.dtype(bool):            # h5py: enum over int8 {FALSE = 0, TRUE = 1}, datatype message version 1
        base = datatype(np.int8)
        return struct.pack("<BBBBI", 0x18, 2, 0, 0, 1) + base + pad8(b"FALSE\0") + pad8(b"TRUE\0") + bytes([0, 1])
    if dt.kind in "iu":
        return struct.pack("<BBBBIHH", 0x10, order | (8 if dt.kind == "i" else 0), 0, 0, dt.itemsize, 0, 8 * dt.itemsize)
    if dt.kind == "f":
        expo, mant, bias = {4: (8, 23, 127), 8: (11, 52, 1023)}[dt.itemsize]
        return struct.pack("<BBBBI", 0x11, 0x20 | order, 8 * dt.itemsize - 1, 0, dt.itemsize) + \
            struct.pack("<HHBBBBI", 0, 8 * dt.itemsize, mant, expo, 0, mant, bias)
    if dt.kind == "S":
        return struct.pack("<BBBBI", 0x13, 0x01, 0, 0, dt.itemsize)     # null-padded, ASCII
    raise TypeError(dt)


def dataspace(shape, maxdims=False):
    body = struct.pack("<BBB5x", 1, len(shape), 1 if (maxdims and shape) else 0) + struct.pack("<%dQ" % len(shape), *shape)
    if maxdims and shape:
        body += struct.pack("<%dQ" % len(shape), *shape)
    return body


def message(mtype, body, flags=0):
    body = pad8(body)
    return struct.pack("<HHB3x", mtype, len(body), flags) + body


def raw_bytes(arr, big_endian=False):
    arr = np.asarray(arr)
    if arr.dtype == np.dtype(bool):
        return arr.astype(np.uint8).tobytes()
    if big_endian:
        return np.ascontiguousarray(arr).astype(arr.dtype.newbyteorder(">")).tobytes()
    return np.ascontiguousarray(arr).tobytes()


class File:
    def __init__(self, hole=24, first_block=2, split_at=3):
        """hole: unused bytes left between objects (freed space of a file that was edited in place);
        first_block: messages that fit the first object-header block before a continuation block is needed;
        split_at: leaves per TREE node beyond which a level-1 root is put on top."""
        self.buf = bytearray(96)
        self.hole, self.first_block, self.split_at = hole, first_block, split_at
        self.gheap = None

    def alloc(self, data, align=8):
        self.buf += b"\0" * (-len(self.buf) % align)
        self.buf += b"\xdd" * self.hole                 # junk, never referenced
        self.buf += b"\0" * (-len(self.buf) % align)
        addr = len(self.buf)
        self.buf += data
        return addr

    # ---- variable-length strings live in a global heap collection
    def vlen_ref(self, text):
        data = text.encode("utf8")
        obj = struct.pack("<HHIQ", 1, 0, 0, len(data)) + pad8(data)
        free = struct.pack("<HHIQ", 0, 0, 0, 4096 - 16 - len(obj))
        coll = b"GCOL" + struct.pack("<B3xQ", 1, 4096) + obj + free
        addr = self.alloc(coll + b"\0" * (4096 - len(coll)))
        return struct.pack("<IQI", len(data), addr, 1)

    def attribute(self, name, value):
        nm = name.encode("utf8") + b"\0"
        if isinstance(value, VlenStr):
            dt = struct.pack("<BBBBI", 0x19, 0x01, 0x00, 0, 16) + datatype(np.dtype("S1"))   # class 9, type = string
            ds, raw = dataspace(()), self.vlen_ref(value.text)
        else:
            arr = np.asarray(value)
            if arr.dtype.kind == "U":
                arr = np.char.encode(arr, "utf8")
            dt, ds, raw = datatype(arr.dtype), dataspace(arr.shape), raw_bytes(arr)
        body = struct.pack("<BBHHH", 1, 0, len(nm), len(dt), len(ds)) + pad8(nm) + pad8(dt) + pad8(ds) + raw
        return message(0x0C, body)

    def object_header(self, msgs):
        """First block with room for `first_block` messages and a continuation message; the rest, separated by NIL
        messages, in a continuation block placed elsewhere in the file (after whatever is allocated in between)."""
        head, tail = msgs[:self.first_block], msgs[self.first_block:]
        if not tail:
            body = b"".join(head) + message(0x00, b"\0" * 16)
            return self.alloc(struct.pack("<BBHII4x", 1, 0, len(head) + 1, 1, len(body)) + body)
        cont_body = b""
        n_tail = 0
        for m in tail:
            cont_body += m + message(0x00, b"\0" * 8)
            n_tail += 2
        cont_addr = self.alloc(cont_body)
        body = b"".join(head) + message(0x10, struct.pack("<QQ", cont_addr, len(cont_body)))
        return self.alloc(struct.pack("<BBHII4x", 1, 0, len(head) + 1 + n_tail, 1, len(body)) + body)

    def dataset(self, d):
        arr = d.array
        if arr.dtype.kind == "U":
            arr = np.char.encode(arr, "utf8")
        raw = raw_bytes(arr, d.big_endian)
        if d.layout == "compact":
            layout = struct.pack("<BBH", 3, 0, len(raw)) + raw
            fill = struct.pack("<BBBBI", 2, 1, 2, 1, 0)
        else:
            addr = self.alloc(raw) if raw else UNDEF
            layout = struct.pack("<BBQQ", 3, 1, addr, len(raw))
            fill = struct.pack("<BBBBI", 2, 2, 2, 1, 0)
        msgs = [message(0x01, dataspace(arr.shape, d.maxdims)), message(0x03, datatype(arr.dtype, d.big_endian), flags=1),
                message(0x05, fill), message(0x08, layout), message(0x12, struct.pack("<B3xI", 1, 1573000000))]
        msgs += [self.attribute(k, v) for k, v in d.attrs.items()]
        return self.object_header(msgs)

    def group(self, g):
        """-> (object header address, B-tree address, heap address)."""
        names = sorted(g.children, key=lambda s: s.encode("utf8"))
        kids = [(n, self.node(g.children[n])) for n in names]
        # local heap: the empty string at 0, names of deleted links still in place, live names, one free block
        heap = bytearray(8)
        for dead in g.dead_names:
            heap += pad8(dead.encode("utf8") + b"\0")
        offs = {}
        for n in names:
            offs[n] = len(heap)
            heap += pad8(n.encode("utf8") + b"\0")
        free_off = len(heap)
        heap += struct.pack("<QQ", 1, 32) + b"\0" * 16
        data_addr = self.alloc(bytes(heap))
        heap_addr = self.alloc(b"HEAP" + struct.pack("<B3xQQQ", 0, len(heap), free_off, data_addr))
        # symbol nodes: libhdf5 splits a full node (2K entries) in two halves, so leaves hold K..2K entries
        leaves = []
        step = LEAF_K + 1
        for i in range(0, len(kids), step):
            part = kids[i:i + step]
            snod = bytearray(b"SNOD" + struct.pack("<BBH", 1, 0, len(part)))
            for n, (addr, bt, hp) in part:
                if bt is not None:
                    snod += struct.pack("<QQII", offs[n], addr, 1, 0) + struct.pack("<QQ", bt, hp)
                else:
                    snod += struct.pack("<QQII16x", offs[n], addr, 0, 0)
            snod += b"\0" * (8 + 40 * 2 * LEAF_K - len(snod))
            leaves.append((self.alloc(bytes(snod)), offs[part[-1][0]]))
        node_size = 24 + (2 * INTERNAL_K + 1) * 8 + 2 * INTERNAL_K * 8

        def tree(level, entries, left=UNDEF, right=UNDEF):
            t = bytearray(b"TREE" + struct.pack("<BBHQQ", 0, level, len(entries), left, right))
            t += struct.pack("<Q", 0)
            for child, last_key in entries:
                t += struct.pack("<QQ", child, last_key)
            return bytes(t) + b"\0" * (node_size - len(t))

        if len(leaves) <= self.split_at:
            btree = self.alloc(tree(0, leaves))
        else:
            mids = []
            chunks = [leaves[i:i + self.split_at] for i in range(0, len(leaves), self.split_at)]
            addrs = []
            for c in chunks:        # reserve the nodes first so that siblings can point at each other
                addrs.append(self.alloc(b"\0" * node_size))
            for i, c in enumerate(chunks):
                left = addrs[i - 1] if i else UNDEF
                right = addrs[i + 1] if i + 1 < len(chunks) else UNDEF
                self.buf[addrs[i]:addrs[i] + node_size] = tree(0, c, left, right)
                mids.append((addrs[i], c[-1][1]))
            btree = self.alloc(tree(1, mids))
        msgs = [message(0x11, struct.pack("<QQ", btree, heap_addr))] + [self.attribute(k, v) for k, v in g.attrs.items()]
        return self.object_header(msgs), btree, heap_addr

    def node(self, x):
        if isinstance(x, Grp):
            return self.group(x)
        return (self.dataset(x if isinstance(x, Dset) else Dset(x)), None, None)

    def finish(self, root):
        root_addr, btree, heap = self.group(root)
        sb = b"\x89HDF\r\n\x1a\n" + struct.pack("<BBBBBBBBHHI", 0, 0, 0, 0, 0, 8, 8, 0, LEAF_K, INTERNAL_K, 0)
        sb += struct.pack("<QQQQ", 0, UNDEF, len(self.buf), UNDEF)
        sb += struct.pack("<QQII", 0, root_addr, 1, 0) + struct.pack("<QQ", btree, heap)
        assert len(sb) == 96
        self.buf[:96] = sb
        return bytes(self.buf)


def write(path, root, **kw):
    data = File(**kw).finish(root)
    with open(path, "wb") as f:
        f.write(data)
    return data


# ------------------------------------------------------------------------------------------------ Keras layouts
def keras_weight_file(weights, weight_names, config=None, nest=None, layer_name_width=None, extra_attrs=None):
    """The tree ``Model.save_weights`` (Keras 2.3.1, ``keras/engine/saving.py`` save_weights_to_hdf5_group) writes:
    root attrs layer_names / backend / keras_version, one group per layer with attr weight_names and the weight
    datasets under their (slash-separated) names; ``config`` = the reference's own group (seq2seq.py:1130-1141);
    ``nest`` = 'model_weights' for files written by ``Model.save`` (whole model)."""
    layers = list(weights)
    ln = np.array([n.encode("utf8") for n in layers])
    if layer_name_width:
        ln = ln.astype("S%d" % layer_name_width)
    attrs = {"layer_names": ln, "backend": np.bytes_(b"tensorflow"), "keras_version": np.bytes_(b"2.3.1")}
    attrs.update(extra_attrs or {})
    root = Grp(attrs=attrs)
    for layer in layers:
        names = weight_names[layer]
        g = Grp(attrs={"weight_names": np.array([n.encode("utf8") for n in names])})
        for n, arr in zip(names, weights[layer]):
            node = g
            parts = n.split("/")
            for p in parts[:-1]:
                node = node.children.setdefault(p, Grp())
            node.children[parts[-1]] = Dset(np.asarray(arr, np.float32), maxdims=True)
        root.children[layer] = g
    if nest:
        root = Grp(children={nest: root}, attrs={"keras_version": np.bytes_(b"2.3.1"), "model_config": VlenStr('{"class_name": "Model"}')})
    if config is not None:
        c = Grp()
        for k, v in config.items():
            c.children[k] = v if isinstance(v, Dset) else Dset(np.asarray(v))
        (root.children[nest] if nest else root).children["config"] = c
    return root

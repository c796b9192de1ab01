"""Read the parameters of a reference-made TfFeedForward trial: ``{'h5_model_base64': ...}``.

The reference's ``dump_parameters`` (examples/models/image_classification/TfFeedForward.py:102-118)
saves the whole Keras model with ``model.save()`` to an HDF5 file and ships its bytes base64-encoded;
``load_parameters`` (:120-135) hands them back to ``keras.models.load_model``.  Neither TensorFlow nor
h5py/libhdf5 exists on the serving box, so this module reads the few HDF5 structures such a file
contains, straight from the bytes (HDF5 File Format Specification, version 3.0):

* superblock versions 0-3;
* object headers version 1 (what libhdf5 writes by default, and so h5py/Keras) and version 2;
* groups: symbol tables (B-tree v1 + local heap + SNOD nodes) and compact link messages;
* datasets: contiguous or compact layout, fixed-point / IEEE floating-point / fixed-length string /
  variable-length string (global heap) element types;
* attributes: message versions 1-3.

Chunked or filtered datasets, dense (fractal-heap) link and attribute storage and shared messages are
refused with an ``H5FormatError`` that says what was met -- Keras does not produce them for a model of
this size.

Parity status: **unpinned against libhdf5 output.**  No HDF5 library and no HDF5 file exists in the
build image, so the reader is tested against files assembled by ``tests/h5_writer.py`` from the same
specification (symbol-table groups, version-1 headers, the layout ``model.save()`` of Keras 2.2 /
tf.keras 1.12 produces: ``keras/engine/saving.py``), not against a file TensorFlow wrote.
"""
import base64
import json

import numpy as np

SIGNATURE = b'\x89HDF\r\n\x1a\n'
UNDEF = 0xFFFFFFFFFFFFFFFF


class H5FormatError(ValueError):
    pass


def _need(cond, what):
    if not cond:
        raise H5FormatError(what)


class _Cursor:
    """Little-endian reads at an absolute file position."""

    def __init__(self, buf, pos=0):
        self.buf = buf
        self.pos = pos

    def take(self, n):
        _need(0 <= self.pos and self.pos + n <= len(self.buf), 'read past the end of the file at %d (+%d)' % (self.pos, n))
        b = self.buf[self.pos:self.pos + n]
        self.pos += n
        return b

    def u(self, n):
        return int.from_bytes(self.take(n), 'little')

    def skip(self, n):
        self.pos += n

    def align(self, k, origin=0):
        r = (self.pos - origin) % k
        if r:
            self.pos += k - r


# header message types
MSG_NIL, MSG_DATASPACE, MSG_LINK_INFO, MSG_DATATYPE, MSG_LINK = 0x00, 0x01, 0x02, 0x03, 0x06
MSG_LAYOUT, MSG_FILTER, MSG_ATTRIBUTE, MSG_CONTINUATION, MSG_SYMBOL_TABLE, MSG_ATTR_INFO = 0x08, 0x0B, 0x0C, 0x10, 0x11, 0x15


class H5File:
    """The object tree of one HDF5 file held in memory (read-only)."""

    def __init__(self, data):
        self.buf = bytes(data)
        base = self._find_superblock()
        c = _Cursor(self.buf, base + 8)
        version = c.u(1)
        _need(version in (0, 1, 2, 3), 'unsupported superblock version %d' % version)
        if version in (0, 1):
            c.skip(4)                       # free-space, root-entry, reserved, shared-header versions
            self.O, self.L = c.u(1), c.u(1)
            c.skip(1)
            c.skip(4)                       # group leaf / internal node K
            c.skip(4)                       # consistency flags
            if version == 1:
                c.skip(4)                   # indexed-storage K, reserved
            self.base = c.u(self.O)
            c.skip(3 * self.O)              # free-space info, end of file, driver info
            c.skip(self.O)                  # root entry: link name offset
            self.root_addr = c.u(self.O)
        else:
            self.O, self.L = c.u(1), c.u(1)
            c.skip(1)                       # consistency flags
            self.base = c.u(self.O)
            c.skip(2 * self.O)              # superblock extension, end of file
            self.root_addr = c.u(self.O)
        _need(self.O in (2, 4, 8) and self.L in (2, 4, 8), 'odd offset/length sizes %d/%d' % (self.O, self.L))
        self.undef = (1 << (8 * self.O)) - 1
        if self.base == self.undef:
            self.base = base
        self._gcol = {}

    def _find_superblock(self):
        pos = 0
        while pos + 8 <= len(self.buf):
            if self.buf[pos:pos + 8] == SIGNATURE:
                return pos
            pos = 512 if pos == 0 else pos * 2
        raise H5FormatError('not an HDF5 file (no superblock signature)')

    def at(self, addr):
        _need(addr != self.undef, 'undefined address')
        return _Cursor(self.buf, self.base + addr)

    # ------------------------------------------------------------------ object headers
    def messages(self, addr):
        """[(type, flags, body bytes)] of the object header at ``addr``, continuation blocks included."""
        c = self.at(addr)
        if self.buf[c.pos:c.pos + 4] == b'OHDR':
            return self._messages_v2(c)
        version = c.u(1)
        _need(version == 1, 'object header version %d at %d' % (version, addr))
        c.skip(1)
        n_msgs = c.u(2)
        c.skip(4)                           # reference count
        size = c.u(4)
        c.skip(4)                           # the prefix is padded to 16 bytes
        out, blocks = [], [(c.pos, size)]
        while blocks and len(out) < n_msgs:
            pos, size = blocks.pop(0)
            c = _Cursor(self.buf, pos)
            end = pos + size
            while c.pos + 8 <= end and len(out) < n_msgs:
                mtype, msize, flags = c.u(2), c.u(2), c.u(1)
                c.skip(3)
                body = c.take(msize)
                out.append((mtype, flags, body))
                if mtype == MSG_CONTINUATION:
                    b = _Cursor(body)
                    blocks.append((self.base + b.u(self.O), b.u(self.L)))
                    _need(len(out) <= 65536, 'object header with an endless chain of continuation blocks')
        return out

    def _messages_v2(self, c):
        c.skip(4)
        _need(c.u(1) == 2, 'object header (OHDR) version is not 2')
        hflags = c.u(1)
        if hflags & 0x20:
            c.skip(16)                      # access, modification, change, birth times
        if hflags & 0x10:
            c.skip(4)                       # attribute phase-change thresholds
        size = c.u(1 << (hflags & 3))
        out, blocks, seen = [], [(c.pos, size)], set()
        while blocks:
            pos, size = blocks.pop(0)
            _need(pos not in seen, 'object header continuation blocks form a cycle')
            seen.add(pos)
            c = _Cursor(self.buf, pos)
            end = pos + size
            while c.pos + 4 <= end:
                mtype, msize, flags = c.u(1), c.u(2), c.u(1)
                if hflags & 0x04:
                    c.skip(2)               # creation order
                if c.pos + msize > end:
                    break                   # gap before the checksum
                body = c.take(msize)
                out.append((mtype, flags, body))
                if mtype == MSG_CONTINUATION:
                    b = _Cursor(body)
                    caddr, clen = b.u(self.O), b.u(self.L)
                    _need(self.buf[self.base + caddr:self.base + caddr + 4] == b'OCHK', 'continuation block without OCHK')
                    blocks.append((self.base + caddr + 4, clen - 8))     # minus signature and checksum
        return out

    # ------------------------------------------------------------------ groups
    def links(self, addr):
        """{name: object header address} of the group whose header is at ``addr``."""
        out = {}
        for mtype, _, body in self.messages(addr):
            if mtype == MSG_SYMBOL_TABLE:
                b = _Cursor(body)
                btree, heap = b.u(self.O), b.u(self.O)
                heap_data = self._local_heap(heap)
                self._walk_group_btree(btree, heap_data, out)
            elif mtype == MSG_LINK:
                name, target = self._link_message(body)
                if target is not None:
                    out[name] = target
            elif mtype == MSG_LINK_INFO:
                b = _Cursor(body)
                b.skip(1)
                flags = b.u(1)
                if flags & 1:
                    b.skip(8)
                fheap = b.u(self.O)
                _need(fheap == self.undef, 'group with dense (fractal heap) link storage')
        return out

    def _local_heap(self, addr):
        c = self.at(addr)
        _need(c.take(4) == b'HEAP', 'local heap signature missing at %d' % addr)
        c.skip(4)                           # version, reserved
        size = c.u(self.L)
        c.skip(self.L)                      # free list head
        data = c.u(self.O)
        return self.buf[self.base + data:self.base + data + size]

    def _walk_group_btree(self, addr, heap, out, below=256):
        c = self.at(addr)
        _need(c.take(4) == b'TREE', 'B-tree signature missing at %d' % addr)
        node_type, level, used = c.u(1), c.u(1), c.u(2)
        _need(node_type == 0, 'B-tree node type %d where a group node is expected' % node_type)
        _need(level < below, 'B-tree levels do not decrease towards the leaves')
        c.skip(2 * self.O)                  # siblings
        children = []
        for _ in range(used):
            c.skip(self.L)                  # key: heap offset of the largest name in the child to the left
            children.append(c.u(self.O))
        for child in children:
            if level > 0:
                self._walk_group_btree(child, heap, out, level)
                continue
            s = self.at(child)
            _need(s.take(4) == b'SNOD', 'symbol table node signature missing at %d' % child)
            s.skip(2)
            for _ in range(s.u(2)):
                name_off, obj = s.u(self.O), s.u(self.O)
                s.skip(24)                  # cache type, reserved, scratch pad
                end = heap.find(b'\0', name_off)
                _need(0 <= name_off < len(heap) and end >= 0, 'link name outside the local heap')
                out[heap[name_off:end].decode('utf-8')] = obj

    def _link_message(self, body):
        b = _Cursor(body)
        _need(b.u(1) == 1, 'link message version')
        flags = b.u(1)
        ltype = b.u(1) if flags & 0x08 else 0
        if flags & 0x04:
            b.skip(8)                       # creation order
        if flags & 0x10:
            b.skip(1)                       # character set
        name = b.take(b.u(1 << (flags & 3))).decode('utf-8')
        return name, (b.u(self.O) if ltype == 0 else None)      # soft / external links are skipped

    def resolve(self, path, start=None):
        addr = self.root_addr if start is None else start
        parts = [p for p in path.split('/') if p]
        _need(len(parts) <= 64, 'path too deep')
        for part in parts:
            table = self.links(addr)
            _need(part in table, 'no object named %r (have %s)' % (part, sorted(table)))
            addr = table[part]
        return addr

    # ------------------------------------------------------------------ element types, shapes, values
    def _datatype(self, body):
        """-> (kind, size, numpy dtype or None); kind in 'int', 'float', 'string', 'vlen_string'."""
        b = _Cursor(body)
        cv = b.u(1)
        cls, bits0 = cv & 0x0F, b.u(1)
        b.skip(2)
        size = b.u(4)
        order = '>' if bits0 & 1 else '<'
        if cls == 0:
            signed = bool(bits0 & 0x08)
            _need(size in (1, 2, 4, 8), 'integer of %d bytes' % size)
            return 'int', size, np.dtype('%s%s%d' % (order, 'i' if signed else 'u', size))
        if cls == 1:
            _need(size in (2, 4, 8), 'floating-point type of %d bytes' % size)
            return 'float', size, np.dtype('%sf%d' % (order, size))
        if cls == 3:
            return 'string', size, np.dtype('S%d' % size)
        if cls == 9:
            _need((bits0 & 0x0F) == 1, 'variable-length sequence (only variable-length strings are read)')
            return 'vlen_string', size, None
        raise H5FormatError('datatype class %d is not read' % cls)

    def _dataspace(self, body):
        b = _Cursor(body)
        version, rank, flags = b.u(1), b.u(1), b.u(1)
        if version == 1:
            b.skip(5)
        else:
            _need(version == 2, 'dataspace message version %d' % version)
            if b.u(1) == 2:
                return None                 # null dataspace
        return tuple(b.u(self.L) for _ in range(rank))

    def _global_heap_object(self, addr, index):
        if addr not in self._gcol:
            c = self.at(addr)
            _need(c.take(4) == b'GCOL', 'global heap signature missing at %d' % addr)
            c.skip(4)
            end = self.base + addr + c.u(self.L)
            objs = {}
            while c.pos + 8 + self.L <= end:
                idx = c.u(2)
                c.skip(6)                   # reference count, reserved
                size = c.u(self.L)
                if idx == 0:
                    break                   # free space
                objs[idx] = c.take(size)
                c.align(8, self.base + addr)
            self._gcol[addr] = objs
        _need(index in self._gcol[addr], 'global heap object %d missing' % index)
        return self._gcol[addr][index]

    def _decode(self, raw, dtype, shape):
        kind, size, np_dtype = dtype
        count = int(np.prod(shape)) if shape else 1
        _need(len(raw) >= count * size, 'value shorter than its type and shape say')
        if kind == 'vlen_string':
            out = []
            c = _Cursor(raw)
            for _ in range(count):
                length, addr, index = c.u(4), c.u(self.O), c.u(4)
                out.append(self._global_heap_object(addr, index)[:length] if length else b'')
            arr = np.array(out, dtype=object)
        else:
            arr = np.frombuffer(raw, dtype=np_dtype, count=count)
            if kind == 'string':
                arr = np.array([s.rstrip(b'\0 ') for s in arr.tolist()], dtype=object)
            elif np_dtype.byteorder == '>':
                arr = arr.astype(np_dtype.newbyteorder('<'))
        if not shape:
            return arr[0]
        return arr.reshape(shape)

    # ------------------------------------------------------------------ attributes and datasets
    def attrs(self, addr):
        out = {}
        for mtype, flags, body in self.messages(addr):
            if mtype == MSG_ATTR_INFO:
                b = _Cursor(body)
                b.skip(1)
                aflags = b.u(1)
                if aflags & 1:
                    b.skip(2)
                _need(b.u(self.O) == self.undef, 'object with dense (fractal heap) attribute storage')
            if mtype != MSG_ATTRIBUTE:
                continue
            b = _Cursor(body)
            version = b.u(1)
            _need(version in (1, 2, 3), 'attribute message version %d' % version)
            aflags = b.u(1)
            _need(version == 1 or not (aflags & 3), 'attribute with a shared datatype or dataspace')
            name_size, dt_size, ds_size = b.u(2), b.u(2), b.u(2)
            if version == 3:
                b.skip(1)                   # name character set
            pad = 8 if version == 1 else 1
            name = b.take(name_size).split(b'\0')[0].decode('utf-8')
            b.align(pad)
            dtype = self._datatype(b.take(dt_size))
            b.align(pad)
            shape = self._dataspace(b.take(ds_size))
            b.align(pad)
            out[name] = None if shape is None else self._decode(body[b.pos:], dtype, shape)
        return out

    def dataset(self, addr):
        dtype = shape = raw = None
        for mtype, _, body in self.messages(addr):
            if mtype == MSG_DATATYPE:
                dtype = self._datatype(body)
            elif mtype == MSG_DATASPACE:
                shape = self._dataspace(body)
            elif mtype == MSG_FILTER:
                raise H5FormatError('filtered (compressed) dataset')
            elif mtype == MSG_LAYOUT:
                b = _Cursor(body)
                version = b.u(1)
                if version == 3:
                    cls = b.u(1)
                    if cls == 0:
                        raw = b.take(b.u(2))
                    elif cls == 1:
                        a, n = b.u(self.O), b.u(self.L)
                        raw = b'' if a == self.undef else self.buf[self.base + a:self.base + a + n]
                    else:
                        raise H5FormatError('chunked dataset')
                else:
                    _need(version in (1, 2), 'data layout message version %d' % version)
                    rank, cls = b.u(1), b.u(1)
                    b.skip(5)
                    if cls == 1:
                        a = b.u(self.O)
                        b.skip(4 * rank)
                        raw = (a, None)     # size follows from type and shape
                    elif cls == 0:
                        b.skip(4 * rank)
                        raw = b.take(b.u(4))
                    else:
                        raise H5FormatError('chunked dataset')
        _need(dtype is not None and shape is not None and raw is not None, 'object at %d is not a dataset' % addr)
        if isinstance(raw, tuple):
            n = int(np.prod(shape)) * dtype[1]
            raw = self.buf[self.base + raw[0]:self.base + raw[0] + n]
        if len(raw) == 0 and int(np.prod(shape)) > 0:
            raise H5FormatError('dataset without allocated storage')
        return self._decode(raw, dtype, shape)


# ---------------------------------------------------------------------- Keras model file -> parameter dict
def _text(v):
    if isinstance(v, (bytes, np.bytes_)):
        return bytes(v).decode('utf-8')
    return str(v)


def _names(v):
    if v is None:
        return []
    return [_text(x) for x in np.atleast_1d(v).tolist()]


def keras_layer_weights(data):
    """[(layer name, {weight name: float32 array})] in ``layer_names`` order, plus the parsed
    ``model_config`` (or None).  Follows keras/engine/saving.py: the weights live under the group
    ``model_weights`` of a ``model.save()`` file (under the root of a ``save_weights()`` file); each
    layer group carries ``weight_names``, and a weight named ``dense/kernel:0`` is the dataset at that
    path below the layer group."""
    f = H5File(data)
    root_attrs = f.attrs(f.root_addr)
    config = json.loads(_text(root_attrs['model_config'])) if 'model_config' in root_attrs else None
    top = f.links(f.root_addr)
    waddr = top['model_weights'] if 'model_weights' in top else f.root_addr
    wattrs = f.attrs(waddr)
    _need('layer_names' in wattrs, "no 'layer_names' attribute: not a Keras weight file")
    layers = []
    for lname in _names(wattrs['layer_names']):
        laddr = f.resolve(lname, waddr)
        weights = {}
        for wname in _names(f.attrs(laddr).get('weight_names')):
            weights[wname] = np.ascontiguousarray(f.dataset(f.resolve(wname, laddr)), dtype=np.float32)
        layers.append((lname, weights))
    return layers, config


def _layer_configs(config):
    if config is None:
        return {}
    cfg = config.get('config')
    seq = cfg.get('layers') if isinstance(cfg, dict) else cfg        # Keras >= 2.2.3 nests the list
    out = {}
    for entry in seq or []:
        c = entry.get('config', {})
        out[c.get('name')] = (entry.get('class_name'), c)
    return out


def params_from_h5_bytes(data, image_size=None):
    """The TfFeedForward graph (TfFeedForward.py:141-157: Flatten, BatchNormalization, Dense+ReLU x L,
    Dense+softmax) as the raw-array parameter dict the engine takes."""
    layers, config = keras_layer_weights(data)
    lcfg = _layer_configs(config)
    bn = None
    kernels, biases = [], []
    eps = 1e-3
    for lname, weights in layers:
        if not weights:
            continue                        # Flatten
        by_suffix = {w.split('/')[-1].split(':')[0]: a for w, a in weights.items()}
        if 'moving_mean' in by_suffix:
            _need(bn is None and not kernels, 'more than one BatchNormalization layer, or one after a Dense layer')
            _need('moving_variance' in by_suffix, 'BatchNormalization layer without moving_variance')
            d = by_suffix['moving_mean'].shape[0]
            bn = {'bn_gamma': by_suffix.get('gamma', np.ones(d, np.float32)),       # scale=False / center=False
                  'bn_beta': by_suffix.get('beta', np.zeros(d, np.float32)),
                  'bn_mean': by_suffix['moving_mean'], 'bn_var': by_suffix['moving_variance']}
            cls, c = lcfg.get(lname, (None, {}))
            eps = float(c.get('epsilon', 1e-3))
        elif 'kernel' in by_suffix:
            k = by_suffix['kernel']
            _need(k.ndim == 2, 'layer %r: kernel of rank %d (only Dense layers are read)' % (lname, k.ndim))
            kernels.append(k)
            biases.append(by_suffix.get('bias', np.zeros(k.shape[1], np.float32)))
        else:
            raise H5FormatError('layer %r has weights this model family does not use: %s' % (lname, sorted(weights)))
    _need(kernels, 'no Dense layer found')
    d = kernels[0].shape[0]
    if bn is None:                          # a graph without the normalisation layer: identity affine
        bn = {'bn_gamma': np.ones(d, np.float32), 'bn_beta': np.zeros(d, np.float32),
              'bn_mean': np.zeros(d, np.float32), 'bn_var': np.full(d, 1.0 - eps, np.float32)}
    for a, b in zip(kernels[:-1], kernels[1:]):
        _need(a.shape[1] == b.shape[0], 'Dense layers do not chain: %s then %s' % (a.shape, b.shape))
    _need(bn['bn_mean'].shape[0] == d, 'BatchNormalization width %d, first Dense input %d' % (bn['bn_mean'].shape[0], d))
    if image_size is None:
        for cls, c in lcfg.values():
            shape = c.get('batch_input_shape')
            if shape and len(shape) >= 3:
                image_size = int(shape[1])
                break
    params = {'format': 'rafiki_b200.TfFeedForward/1', 'kernels': kernels, 'biases': biases, 'bn_eps': eps}
    params.update(bn)
    if image_size is not None:
        params['image_size'] = int(image_size)
    return params


def params_from_h5_base64(text, image_size=None):
    return params_from_h5_bytes(base64.b64decode(text.encode('utf-8') if isinstance(text, str) else text), image_size)

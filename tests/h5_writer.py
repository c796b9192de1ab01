"""Test infrastructure: assemble small HDF5 files the way libhdf5 1.8/1.10 lays them out by default
(superblock version 0, version-1 object headers, symbol-table groups, contiguous datasets, version-1
attribute messages, 8-byte offsets and lengths), from the HDF5 File Format Specification.  Used to
exercise ``rafiki_b200/keras_h5.py``; no HDF5 library exists in this image, so this writer -- not
libhdf5 -- is what the reader is checked against (see the reader's header)."""
import json
import struct

import numpy as np

UNDEF = 0xFFFFFFFFFFFFFFFF


def _pad8(b):
    return b + b'\0' * (-len(b) % 8)


def _dt_float32():
    return struct.pack('<BBBBI', 0x11, 0x20, 0x1F, 0x00, 4) + struct.pack('<HHBBBBI', 0, 32, 23, 8, 0, 23, 127)


def _dt_float64():
    return struct.pack('<BBBBI', 0x11, 0x20, 0x3F, 0x00, 8) + struct.pack('<HHBBBBI', 0, 64, 52, 11, 0, 52, 1023)


def _dt_int(size, signed=True):
    return struct.pack('<BBBBI', 0x10, 0x08 if signed else 0x00, 0, 0, size) + struct.pack('<HH', 0, 8 * size)


def _dt_string(size):
    return struct.pack('<BBBBI', 0x13, 0x01, 0, 0, size)          # null-padded, ASCII


def _dataspace(shape):
    return struct.pack('<BBB5x', 1, len(shape), 0) + b''.join(struct.pack('<Q', d) for d in shape)


def _value(v):
    """-> (datatype message, dataspace message, raw bytes)"""
    if isinstance(v, (bytes, str)):
        raw = v.encode('utf-8') if isinstance(v, str) else v
        return _dt_string(max(len(raw), 1)), _dataspace(()), raw or b'\0'
    a = np.asarray(v)
    if a.dtype.kind in ('S', 'U', 'O'):
        items = [x.encode('utf-8') if isinstance(x, str) else bytes(x) for x in a.ravel().tolist()]
        size = max(max(len(x) for x in items), 1)
        return _dt_string(size), _dataspace(a.shape), b''.join(x.ljust(size, b'\0') for x in items)
    if a.dtype == np.float32:
        return _dt_float32(), _dataspace(a.shape), a.astype('<f4').tobytes()
    if a.dtype == np.float64:
        return _dt_float64(), _dataspace(a.shape), a.astype('<f8').tobytes()
    if a.dtype.kind in 'iu':
        return _dt_int(a.dtype.itemsize, a.dtype.kind == 'i'), _dataspace(a.shape), a.astype(a.dtype.newbyteorder('<')).tobytes()
    raise TypeError(a.dtype)


class H5Writer:
    def __init__(self):
        self.buf = bytearray(96)            # superblock, written last

    def alloc(self, data):
        self.buf += b'\0' * (-len(self.buf) % 8)
        addr = len(self.buf)
        self.buf += data
        return addr

    @staticmethod
    def _msg(mtype, body, flags=0):
        body = _pad8(body)
        return struct.pack('<HHB3x', mtype, len(body), flags) + body

    def _attr_msg(self, name, value):
        dt, ds, raw = _value(value)
        nm = name.encode('utf-8') + b'\0'
        body = struct.pack('<BBHHH', 1, 0, len(nm), len(dt), len(ds)) + _pad8(nm) + _pad8(dt) + _pad8(ds) + raw
        return self._msg(0x0C, body)

    def _header(self, msgs, split_after=None):
        """Version-1 object header; with ``split_after`` the later messages go to a continuation block,
        as libhdf5 does when attributes are added to an object whose header is full."""
        if split_after is not None and split_after < len(msgs):
            tail = b''.join(msgs[split_after:])
            caddr = self.alloc(tail)
            head = msgs[:split_after] + [self._msg(0x10, struct.pack('<QQ', caddr, len(tail)))]
            n = len(msgs) + 1
        else:
            head, n = msgs, len(msgs)
        blob = b''.join(head)
        return self.alloc(struct.pack('<BBHII4x', 1, 0, n, 1, len(blob)) + blob)

    def dataset(self, array, attrs=None, compact=False):
        dt, ds, raw = _value(array)
        if compact:
            layout = struct.pack('<BBH', 3, 0, len(raw)) + raw
        else:
            layout = struct.pack('<BBQQ', 3, 1, self.alloc(raw) if raw else UNDEF, len(raw))
        msgs = [self._msg(0x01, ds), self._msg(0x03, dt, flags=1),
                self._msg(0x05, struct.pack('<BBBB', 2, 2, 2, 0)),            # fill value (old style), ignored by readers of data
                self._msg(0x08, layout), self._msg(0x12, struct.pack('<B3xI', 1, 1541030400))]
        msgs += [self._attr_msg(k, v) for k, v in (attrs or {}).items()]
        return self._header(msgs)

    def group(self, links, attrs=None, per_node=8, split_attrs=False):
        names = sorted(links, key=lambda s: s.encode('utf-8'))
        heap = bytearray(8)                 # offset 0: the empty string
        offs = {}
        for nm in names:
            offs[nm] = len(heap)
            heap += _pad8(nm.encode('utf-8') + b'\0')
        heap += b'\0' * 16                  # a free block, as libhdf5 leaves
        data_addr = self.alloc(bytes(heap))
        heap_addr = self.alloc(b'HEAP' + struct.pack('<B3xQQQ', 0, len(heap), len(heap) - 16, data_addr))
        snods, keys = [], [0]
        for i in range(0, len(names), per_node):
            chunk = names[i:i + per_node]
            body = b'SNOD' + struct.pack('<BBH', 1, 0, len(chunk))
            for nm in chunk:
                body += struct.pack('<QQII16x', offs[nm], links[nm], 0, 0)
            body += b'\0' * (40 * (per_node - len(chunk)))
            snods.append(self.alloc(body))
            keys.append(offs[chunk[-1]])
        tree = b'TREE' + struct.pack('<BBHQQ', 0, 0, len(snods), UNDEF, UNDEF)
        for i, s in enumerate(snods):
            tree += struct.pack('<QQ', keys[i], s)
        tree += struct.pack('<Q', keys[-1])
        tree_addr = self.alloc(tree)
        msgs = [self._msg(0x11, struct.pack('<QQ', tree_addr, heap_addr)), self._msg(0x00, b'\0' * 8)]
        msgs += [self._attr_msg(k, v) for k, v in (attrs or {}).items()]
        return self._header(msgs, split_after=2 if split_attrs else None)

    def finish(self, root_addr):
        sb = b'\x89HDF\r\n\x1a\n' + struct.pack('<BBBBBBBBHHI', 0, 0, 0, 0, 0, 8, 8, 0, 4, 16, 0)
        sb += struct.pack('<QQQQ', 0, UNDEF, len(self.buf), UNDEF)
        sb += struct.pack('<QQII16x', 0, root_addr, 0, 0)
        assert len(sb) == 96, len(sb)
        self.buf[0:96] = sb
        return bytes(self.buf)


def keras_model_file(params, image_size, channels=1, nested_config=True, split_attrs=True, with_optimizer=True):
    """The file ``keras.Sequential([...TfFeedForward graph...]).save()`` writes (keras/engine/saving.py of
    Keras 2.2 / tf.keras 1.12), for the raw-array parameter dict ``params``."""
    w = H5Writer()
    n_dense = len(params['kernels'])
    names = ['flatten', 'batch_normalization'] + ['dense' if i == 0 else 'dense_%d' % i for i in range(n_dense)]
    layer_cfg = [{'class_name': 'Flatten', 'config': {'name': 'flatten', 'trainable': True,
                                                      'batch_input_shape': [None, image_size, image_size] + ([channels] if channels > 1 else []),
                                                      'dtype': 'float32', 'data_format': 'channels_last'}},
                 {'class_name': 'BatchNormalization', 'config': {'name': 'batch_normalization', 'axis': -1, 'momentum': 0.99,
                                                                 'epsilon': float(params.get('bn_eps', 1e-3)), 'center': True, 'scale': True}}]
    for i in range(n_dense):
        layer_cfg.append({'class_name': 'Dense', 'config': {'name': names[2 + i], 'units': int(params['kernels'][i].shape[1]),
                                                            'activation': 'softmax' if i == n_dense - 1 else 'relu', 'use_bias': True}})
    config = {'class_name': 'Sequential', 'config': {'name': 'sequential', 'layers': layer_cfg} if nested_config else layer_cfg}

    def layer_group(lname, weights):
        inner = w.group({k + ':0': w.dataset(np.asarray(v, np.float32)) for k, v in weights.items()})
        wn = np.array([('%s/%s:0' % (lname, k)).encode() for k in weights]) if weights else np.zeros((0,), np.float64)
        return w.group({lname: inner} if weights else {}, attrs={'weight_names': wn})

    layers = {'flatten': layer_group('flatten', {}),
              'batch_normalization': layer_group('batch_normalization', {
                  'gamma': params['bn_gamma'], 'beta': params['bn_beta'],
                  'moving_mean': params['bn_mean'], 'moving_variance': params['bn_var']})}
    for i in range(n_dense):
        layers[names[2 + i]] = layer_group(names[2 + i], {'kernel': params['kernels'][i], 'bias': params['biases'][i]})
    mw = w.group(layers, attrs={'layer_names': np.array([n.encode() for n in names]), 'backend': b'tensorflow',
                                'keras_version': b'2.1.6-tf'}, per_node=2, split_attrs=split_attrs)
    top = {'model_weights': mw}
    if with_optimizer:
        adam = w.group({'iterations:0': w.dataset(np.int64(120)), 'lr:0': w.dataset(np.float32(0.01), compact=True)})
        top['optimizer_weights'] = w.group({'training': w.group({'Adam': adam})},
                                           attrs={'weight_names': np.array([b'training/Adam/iterations:0', b'training/Adam/lr:0'])})
    root = w.group(top, attrs={'keras_version': b'2.1.6-tf', 'backend': b'tensorflow',
                               'model_config': json.dumps(config).encode('utf-8'),
                               'training_config': json.dumps({'loss': 'sparse_categorical_crossentropy'}).encode('utf-8')},
                   split_attrs=split_attrs)
    return w.finish(root)

"""a reference-made ``{'h5_model_base64': ...}`` parameter dict (TfFeedForward.py:102-135)
is read without TensorFlow or h5py.  The files are assembled by tests/h5_writer.py in libhdf5's default
layout (no HDF5 library exists in this image: parity with a TensorFlow-written file is unpinned)."""
import base64
import struct
import sys
import os

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import h5_writer  # noqa: E402
from oracle import feedforward as ff  # noqa: E402
from rafiki_b200 import keras_h5, synth  # noqa: E402


def _same(a, b):
    for k in ('bn_gamma', 'bn_beta', 'bn_mean', 'bn_var'):
        assert np.array_equal(a[k], np.asarray(b[k], np.float32)), k
    assert len(a['kernels']) == len(b['kernels'])
    for x, y in zip(a['kernels'] + a['biases'], list(b['kernels']) + list(b['biases'])):
        assert x.dtype == np.float32 and np.array_equal(x, np.asarray(y, np.float32))


@pytest.mark.parametrize('S,ch,L,H', [(28, 1, 2, 128), (28, 1, 1, 16), (32, 3, 2, 100), (32, 1, 1, 2)])
@pytest.mark.parametrize('nested,split', [(True, True), (False, False)])
def test_model_save_file_round_trip(S, ch, L, H, nested, split):
    p = synth.make_trial(3, S, ch, L, H)
    data = h5_writer.keras_model_file(p, S, ch, nested_config=nested, split_attrs=split)
    q = keras_h5.params_from_h5_base64(base64.b64encode(data).decode('utf-8'))
    _same(q, p)
    assert q['image_size'] == S and abs(q['bn_eps'] - 1e-3) < 1e-12
    X = synth.make_queries(64, S, ch, seed=2)
    assert np.array_equal(ff.forward(q, X, np.float32), ff.forward(p, X, np.float32))


def test_epsilon_and_image_size_come_from_model_config():
    p = dict(synth.make_trial(1, 28, 1, 1, 8))
    p['bn_eps'] = 5e-2
    q = keras_h5.params_from_h5_bytes(h5_writer.keras_model_file(p, 28))
    assert abs(q['bn_eps'] - 5e-2) < 1e-9
    assert keras_h5.params_from_h5_bytes(h5_writer.keras_model_file(p, 28), image_size=32)['image_size'] == 32


def test_object_tree_attributes_and_groups_spanning_several_nodes():
    w = h5_writer.H5Writer()
    leaves = {'d%02d' % i: w.dataset(np.arange(i + 1, dtype=np.int32) * 3) for i in range(23)}
    g = w.group(leaves, attrs={'note': 'héllo', 'vec': np.array([1.5, -2.0]), 'names': np.array([b'a', b'bcd'])},
                per_node=4, split_attrs=True)
    f = keras_h5.H5File(w.finish(w.group({'g': g, 'scalar': w.dataset(np.float64(2.5), compact=True)})))
    assert sorted(f.links(f.resolve('g'))) == sorted(leaves)
    for i in (0, 7, 22):
        assert np.array_equal(f.dataset(f.resolve('g/d%02d' % i)), np.arange(i + 1, dtype=np.int32) * 3)
    a = f.attrs(f.resolve('g'))
    assert a['note'].decode('utf-8') == 'héllo' and np.array_equal(a['vec'], [1.5, -2.0])
    assert a['names'].tolist() == [b'a', b'bcd']
    assert f.dataset(f.resolve('scalar')) == 2.5
    with pytest.raises(keras_h5.H5FormatError, match='no object named'):
        f.resolve('g/missing')


def _new_style_file(value, attr):
    """Superblock 2, version-2 object headers, compact link messages, version-3 attribute message."""
    buf = bytearray(48)

    def alloc(b):
        buf.extend(b'\0' * (-len(buf) % 8))
        a = len(buf)
        buf.extend(b)
        return a

    def ohdr(msgs):
        body = b''.join(struct.pack('<BHB', t, len(m), 0) + m for t, m in msgs)
        return alloc(b'OHDR' + struct.pack('<BBH', 2, 0x01, len(body) + 3) + body + b'\0' * 3 + b'\0\0\0\0')

    dt, ds, raw = h5_writer._value(value)
    adt, ads, araw = h5_writer._value(attr)
    amsg = struct.pack('<BBHHHB', 3, 0, 5, len(adt), len(ads), 0) + b'unit\0' + adt + ads + araw
    daddr = ohdr([(0x01, ds), (0x03, dt), (0x08, struct.pack('<BBQQ', 3, 1, alloc(raw), len(raw))), (0x0C, amsg)])
    link = struct.pack('<BBB', 1, 0x00, 4) + b'data' + struct.pack('<Q', daddr)
    root = ohdr([(0x02, struct.pack('<BBQQ', 0, 0, h5_writer.UNDEF, h5_writer.UNDEF)), (0x06, link)])
    buf[0:48] = b'\x89HDF\r\n\x1a\n' + struct.pack('<BBBBQQQQI', 2, 8, 8, 0, 0, h5_writer.UNDEF, len(buf), root, 0)
    return bytes(buf)


def test_new_style_file_structures():
    x = np.linspace(-1, 1, 12, dtype=np.float32).reshape(3, 4)
    f = keras_h5.H5File(_new_style_file(x, b'metres'))
    assert list(f.links(f.root_addr)) == ['data']
    assert np.array_equal(f.dataset(f.resolve('data')), x)
    assert f.attrs(f.resolve('data'))['unit'] == b'metres'


def test_variable_length_string_attribute():
    # h5py stores a Python str attribute as a variable-length string: (length, global heap address, index)
    w = h5_writer.H5Writer()
    text = b'{"class_name": "Sequential"}'
    gcol = b'GCOL' + struct.pack('<B3xQ', 1, 16 + 16 + len(text) + (-len(text) % 8) + 16)
    gcol += struct.pack('<HH4xQ', 1, 0, len(text)) + text + b'\0' * (-len(text) % 8) + struct.pack('<HH4xQ', 0, 0, 0)
    gaddr = w.alloc(gcol)
    dt = struct.pack('<BBBBI', 0x19, 0x01, 0x00, 0x00, 16) + h5_writer._dt_string(1)
    ds = h5_writer._dataspace(())
    raw = struct.pack('<IQI', len(text), gaddr, 1)
    body = struct.pack('<BBHHH', 1, 0, 6, len(dt), len(ds)) + b'model\0\0\0' + h5_writer._pad8(dt) + h5_writer._pad8(ds) + raw
    root = w._header([w._msg(0x11, struct.pack('<QQ', *_empty_group_tables(w))), w._msg(0x0C, body)])
    f = keras_h5.H5File(w.finish(root))
    assert f.attrs(f.root_addr)['model'] == text


def _empty_group_tables(w):
    heap_data = w.alloc(b'\0' * 24)
    heap = w.alloc(b'HEAP' + struct.pack('<B3xQQQ', 0, 24, 8, heap_data))
    tree = w.alloc(b'TREE' + struct.pack('<BBHQQ', 0, 0, 0, h5_writer.UNDEF, h5_writer.UNDEF) + struct.pack('<Q', 0))
    return tree, heap


def test_refusals_say_what_was_met():
    with pytest.raises(keras_h5.H5FormatError, match='not an HDF5 file'):
        keras_h5.params_from_h5_bytes(b'PK\x03\x04' + b'\0' * 600)
    w = h5_writer.H5Writer()
    dt, ds, raw = h5_writer._value(np.zeros((4, 4), np.float32))
    chunked = w._header([w._msg(0x01, ds), w._msg(0x03, dt), w._msg(0x08, struct.pack('<BBBQII', 3, 2, 3, 0, 2, 2) + b'\0' * 8)])
    f = keras_h5.H5File(w.finish(w.group({'c': chunked})))
    with pytest.raises(keras_h5.H5FormatError, match='chunked'):
        f.dataset(f.resolve('c'))
    good = h5_writer.keras_model_file(synth.make_trial(0, 28, 1, 1, 8), 28)
    with pytest.raises(keras_h5.H5FormatError):
        keras_h5.params_from_h5_bytes(good[:len(good) // 2])
    w = h5_writer.H5Writer()
    plain = w.finish(w.group({'x': w.dataset(np.zeros(3, np.float32))}))
    with pytest.raises(keras_h5.H5FormatError, match='layer_names'):
        keras_h5.params_from_h5_bytes(plain)


def test_damaged_files_raise_value_errors_and_terminate():
    """Random damage must end in a ValueError (H5FormatError, a JSON or a text decoding error), never in a hang,
    an IndexError or a silent wrong answer of the wrong shape."""
    good = bytearray(h5_writer.keras_model_file(synth.make_trial(2, 28, 1, 1, 8), 28))
    rng = np.random.default_rng(0)
    structural = len(good) - 784 * 8 * 4          # leave most of the big kernel alone: hit headers, heaps, trees
    outcomes = {'ok': 0, 'refused': 0}
    for trial in range(300):
        bad = bytearray(good)
        for _ in range(int(rng.integers(1, 6))):
            pos = int(rng.integers(0, len(bad))) if trial % 3 == 0 else int(rng.integers(0, 4096)) % len(bad)
            bad[pos] = int(rng.integers(0, 256))
        if trial % 10 == 9:
            bad = bad[:int(rng.integers(8, len(bad)))]
        try:
            q = keras_h5.params_from_h5_bytes(bytes(bad))
            assert len(q['kernels']) >= 1 and all(k.ndim == 2 for k in q['kernels'])
            outcomes['ok'] += 1
        except (ValueError, KeyError, OverflowError, MemoryError):
            outcomes['refused'] += 1
    assert outcomes['ok'] + outcomes['refused'] == 300 and outcomes['refused'] > 0 and structural > 0

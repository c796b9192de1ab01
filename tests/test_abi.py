"""The C-ABI library loads and exports every symbol include/rafiki_b200.h declares.  CPU only:
no compute entry point is called with data."""
import ctypes
import os
import re

import pytest

from rafiki_b200 import _native

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, 'include', 'rafiki_b200.h')).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(rfk_[a-z0-9_]+)\s*\(', text)))


def test_library_is_built():
    assert os.path.exists(_native.LIB_PATH), 'run __graft_entry__.build() first'


def test_every_declared_symbol_is_exported_and_bound():
    lib = ctypes.CDLL(_native.LIB_PATH)
    declared = _declared_symbols()
    assert len(declared) >= 12
    for name in declared:
        assert hasattr(lib, name), 'missing export: ' + name
        assert name in _native.PROTOTYPES, 'no ctypes prototype for ' + name
    assert sorted(_native.PROTOTYPES) == declared


def test_prototypes_take_as_many_arguments_as_the_header_declares():
    text = open(os.path.join(ROOT, 'include', 'rafiki_b200.h')).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    text = re.sub(r'//[^\n]*', '', text)
    seen = 0
    for name, args in re.findall(r'\b(rfk_[a-z0-9_]+)\s*\(([^;{]*?)\)\s*;', text, flags=re.S):
        args = args.strip()
        n_args = 0 if args in ('', 'void') else args.count(',') + 1
        assert len(_native.PROTOTYPES[name][1]) == n_args, '%s: header declares %d arguments, ctypes prototype %d' % (
            name, n_args, len(_native.PROTOTYPES[name][1]))
        seen += 1
    assert seen == len(_native.PROTOTYPES)


def test_version_and_error_channel():
    lib = _native.load()
    assert lib.rfk_version() == 100
    assert isinstance(lib.rfk_last_error(), bytes)


def test_argument_validation_needs_no_gpu():
    lib = _native.load()
    rc = lib.rfk_ensemble_mean(None, 0, 4, 10, None, None)
    assert rc == 1 and b'outside 1..' in lib.rfk_last_error()
    rc = lib.rfk_decode_u8(None, None, 4, 0, 0, None, None, 0, None)
    assert rc == 1
    with pytest.raises(_native.RafikiNativeError):
        _native.check(rc)


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    monkeypatch.setattr(_native, '_lib', None)
    monkeypatch.setattr(_native, 'LIB_PATH', str(tmp_path / 'nope.so'))
    with pytest.raises(ImportError, match='no CPU fallback'):
        _native.load()


def test_peer_reduce_argument_validation():
    lib = _native.load()
    assert lib.rfk_peer_reduce(None, None, 2, 0, 0, 1, 8, None, 4, None) == 1
    assert lib.rfk_peer_reduce(None, None, 99, 0, 0, 1, 8, None, 4, None) == 1
    assert b'world' in lib.rfk_last_error()
    assert lib.rfk_peer_wait(None, 2, 1, None) == 1

"""CPU-side checks of the drop-in boundary: the shared library loads and
exports every entry point include/topoflow_b200.h declares, and the ctypes
table mirrors the header.  No compute call is made (no GPU here)."""
import ctypes
import os
import re

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
HEADER = os.path.join(ROOT, 'include', 'topoflow_b200.h')


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
    return sorted(set(re.findall(r'\b(tf_[A-Za-z0-9_]+)\s*\(', src)))


def test_header_declares_the_abi():
    names = declared_functions()
    assert 'tf_chan_step' in names and 'tf_d8_build' in names
    assert len(names) >= 24


def test_library_exports_every_declared_symbol():
    from topoflow_old_b200 import _lib
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared_functions():
        assert hasattr(lib, name), 'libtopoflow_b200.so lacks %s' % name


def test_ctypes_table_matches_header():
    from topoflow_old_b200 import _lib
    assert sorted(_lib.SIGNATURES) == declared_functions()
    assert b'sm_100a' in _lib.lib.tf_version()


def test_no_device_is_an_error_not_a_fallback():
    """Without a GPU the product must refuse to run (no CPU path)."""
    from topoflow_old_b200 import _lib
    from topoflow_old_b200.engine import Strip, device_count
    if device_count() > 0:
        return
    try:
        Strip(8, 8)
    except _lib.TopoflowB200Error as e:
        assert 'CUDA' in str(e) or 'device' in str(e)
    else:
        raise AssertionError('Strip() succeeded without a CUDA device')


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, 'topoflow_old_b200')
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith('.py'):
                txt = open(os.path.join(d, f)).read()
                assert not re.search(r'^\s*(from|import)\s+oracle\b', txt, flags=re.M), f

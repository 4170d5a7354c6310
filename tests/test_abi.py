"""CPU-side checks of the boundary: the C-ABI library loads and exports every symbol that
include/mopo_b200.h declares (no compute calls without a GPU), ctypes structs match the header."""
import os
import re
import ctypes as C
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope='module')
def lib():
    import __graft_entry__ as ge
    ge.build()
    from mopo_b200 import _lib
    return _lib


def _header_functions():
    src = open(os.path.join(ROOT, 'include', 'mopo_b200.h')).read()
    src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
    return sorted(set(re.findall(r'\b(mopo_[a-z0-9_]+)\s*\(', src)))


def test_every_declared_symbol_is_exported(lib):
    names = _header_functions()
    assert len(names) >= 19
    h = lib.load()
    for n in names:
        assert hasattr(h, n), n
    assert sorted(lib.SYMBOLS) == names


def test_version_and_error_string(lib):
    h = lib.load()
    assert h.mopo_version() >= 200
    assert isinstance(h.mopo_last_error(), bytes)


def test_loaded_library_was_built_from_this_tree(lib):
    """The binary that travels to the GPU box is the one these sources produce: its embedded hash (sources + headers + flags)
    equals the tree's; the product library exports no developer probe."""
    from mopo_b200 import build as b
    h = lib.load()
    assert h.mopo_build_hash().decode() == b.tree_hash()
    assert not b.needs_build()
    for name in lib.DEBUG_SYMBOLS:
        assert not hasattr(h, name), name
    dbg = open(os.path.join(ROOT, 'include', 'mopo_b200_debug.h')).read()
    for name in lib.DEBUG_SYMBOLS:
        assert name in dbg


def test_struct_layouts_match_header(lib):
    # 9 int32 ; see mopo_ens_cfg
    assert C.sizeof(lib.EnsCfg) == 36
    # pointers/int64 8-aligned; field order as in the header
    assert lib.StepArgs.B.offset == 16 and lib.StepArgs.noise.offset == 48
    assert lib.StepArgs.next_obs.offset == 88 and C.sizeof(lib.StepArgs) == 168
    assert lib.RolloutArgs.actions.offset == 24 and lib.RolloutArgs.noise.offset == 56
    assert lib.RolloutArgs.action_events.offset == 104 and lib.RolloutArgs.policy.offset == 112 and C.sizeof(lib.RolloutArgs) == 128
    assert lib.PolicyArgs.noise.offset == 32 and lib.PolicyArgs.actions.offset == 72 and C.sizeof(lib.PolicyArgs) == 96
    # mopo_sac_hparams: 6 float + 2 int32; mopo_sac_batch: 5 pointers, int64, pointer, uint64, pointer
    assert C.sizeof(lib.SacHParams) == 32 and lib.SacHParams.target_update_interval.offset == 24
    assert lib.SacBatch.n.offset == 40 and lib.SacBatch.seed.offset == 56 and C.sizeof(lib.SacBatch) == 72


def test_null_handle_calls_fail_cleanly(lib):
    h = lib.load()
    assert h.mopo_pool_size(None) == -1
    assert h.mopo_ens_destroy(None) == 0
    rc = h.mopo_ens_create(None, None)
    assert rc == lib.EINVAL and b'null' in h.mopo_last_error()
    with pytest.raises(ValueError):
        lib.check(rc)


def test_missing_library_fails_loudly(lib, monkeypatch):
    monkeypatch.setattr(lib, '_lib', None)
    monkeypatch.setattr(lib, 'LIB_PATH', '/nonexistent/libmopo_b200.so')
    with pytest.raises(ImportError):
        lib.load()

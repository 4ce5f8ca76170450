"""The C-ABI boundary: both shared libraries load without a GPU, export every symbol their
headers declare, and fail loudly (no CPU fallback) when no CUDA device is present."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# the dav1d entry points src/easyav1.c calls (grep dav1d_ in the reference: :1401-1412 open,
# :2376-2432 send_data/get_picture, :2995 flush, :3555 picture_unref, :1626 close)
DAV1D_API = ["dav1d_default_settings", "dav1d_open", "dav1d_close", "dav1d_send_data", "dav1d_get_picture",
             "dav1d_flush", "dav1d_picture_unref", "dav1d_data_wrap", "dav1d_data_unref", "dav1d_data_create",
             "dav1d_parse_sequence_header", "dav1d_version", "dav1d_version_api", "dav1d_apply_grain",
             "dav1d_get_event_flags", "dav1d_get_decode_error_data_props", "dav1d_get_frame_delay",
             "dav1d_picture_ref", "dav1d_picture_move_ref", "dav1d_data_props_unref"]


def declared_functions(header):
    src = open(header).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(eav1cu_[a-z0-9_]+)\s*\(", src)))


def test_engine_exports_every_declared_symbol():
    from easyav1_b200.lib import load_engine
    E = load_engine()
    names = declared_functions(os.path.join(ROOT, "include", "easyav1_cuda.h"))
    assert len(names) >= 20
    missing = [n for n in names if not hasattr(E, n)]
    assert not missing, missing


def test_host_library_exports_dav1d_api():
    from easyav1_b200.lib import load_host
    H = load_host()
    missing = [n for n in DAV1D_API + ["eav1_stream_open", "eav1_stream_next", "eav1_stream_close", "eav1rec_engine"]
               if not hasattr(H, n)]
    assert not missing, missing


def test_struct_sizes_match_header():
    """ctypes mirrors used by the tests must agree with the C structs (the engine reports them)."""
    from easyav1_b200.lib import load_engine, StagedInfo
    E = load_engine()
    E.eav1cu_abi_sizeof.restype = C.c_size_t
    E.eav1cu_abi_sizeof.argtypes = [C.c_int]
    assert E.eav1cu_abi_sizeof(0) == 60      # Eav1Cmd
    assert E.eav1cu_abi_sizeof(1) == 24      # Eav1LfRec
    assert E.eav1cu_abi_sizeof(2) == 44      # Eav1LrRec
    assert E.eav1cu_abi_sizeof(4) == C.sizeof(StagedInfo)
    assert E.eav1cu_abi_sizeof(99) == 0


def _no_gpu():
    from easyav1_b200.lib import load_engine
    return load_engine().eav1cu_device_count() == 0


def test_no_device_fails_loudly():
    if not _no_gpu():
        pytest.skip("a CUDA device is present")
    from easyav1_b200.lib import load_engine, EngineError
    from easyav1_b200.decoder import Decoder
    E = load_engine()
    ctx = C.c_void_p()
    E.eav1cu_ctx_create.argtypes = [C.POINTER(C.c_void_p), C.c_int]
    assert E.eav1cu_ctx_create(C.byref(ctx), 0) < 0 and not ctx.value
    with pytest.raises(EngineError):
        Decoder(b"\x12\x00")


def test_oracle_is_not_interposed_by_product(manifest):
    """Both libraries export dav1d_open; they are linked -Bsymbolic so that the oracle keeps
    calling the reference's own code even when the product library was loaded first with
    RTLD_GLOBAL. Without a GPU the product cannot decode, so a successful decode proves it."""
    if not _no_gpu():
        pytest.skip("a CUDA device is present")
    from easyav1_b200.lib import load_host
    from tests import refdec
    from tests.conftest import fixture_obu
    load_host()
    man = manifest["intra_64x64_nofilt"]
    assert len(list(refdec.decode_all(fixture_obu(man)))) == len(man["frames"])

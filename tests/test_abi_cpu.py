"""CPU-side checks of the C ABI: the library builds for sm_100a, loads, exports every declared symbol and
rejects bad arguments without touching a GPU."""
import ctypes

import pytest

from weclip_b200 import _lib


@pytest.fixture(scope="module")
def L():
    _lib.build()
    return _lib.lib()


def test_exports_every_declared_symbol(L):
    names = _lib.declared_symbols()
    assert len(names) >= 8
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing


def test_abi_version_matches_header(L):
    import re
    hdr = open(_lib.HEADER_PATH).read()
    assert L.wc_abi_version() == int(re.search(r"#define WC_ABI_VERSION (\d+)", hdr).group(1))


def test_bad_arguments_are_rejected_without_a_device(L):
    rc = L.wc_par_fwd(None, 4, 4, None, None, None, 1, 4, 4, None, 6, 20, 0, None, ctypes.c_size_t(0), None)
    assert rc == -1
    assert b"bad argument" in L.wc_last_error()
    assert L.wc_par_workspace_bytes(1, 3, 8, 8, 6) >= 4 * 64 * (3 + 48 + 6)


def test_product_path_fails_loudly_without_cuda():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from weclip_b200.par import PAR
    par = PAR([1, 2, 4, 8, 12, 24], 20)
    with pytest.raises(_lib.WeclipError):
        par(torch.zeros(1, 3, 8, 8), torch.zeros(1, 2, 8, 8))

"""CPU checks of the C-ABI library: it loads, exports every symbol include/p3_b200.h declares, its host
table code agrees with the oracle's, and every compute entry point fails loudly without a CUDA device
(no CPU fallback)."""
import ctypes
import os
import re

import numpy as np
import pytest

from scream_b200 import lib as p3
from scream_b200.fields import FIELDS, RUNTIME_DEFAULTS

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "p3_b200.h")


def _declared_functions():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(p3b200_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported(p3lib):
    L = p3lib.lib()
    names = _declared_functions()
    assert len(names) >= 17
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    assert sorted(p3lib.ABI_SYMBOLS) == names, set(names) ^ set(p3lib.ABI_SYMBOLS)


def test_abi_version_and_strerror(p3lib):
    L = p3lib.lib()
    assert L.p3b200_abi_version() == 2
    for code in (0, -1, -2, -3, -4, -5):
        assert L.p3b200_strerror(code)
    assert b"CUDA" in L.p3b200_strerror(-2).upper()


def test_field_enum_matches_python_order():
    text = open(HEADER).read()
    body = text[text.index("enum P3B200Field"):]
    body = body[:body.index("};")]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = re.findall(r"P3B200_F_([A-Z0-9_]+)", body)
    assert names[-1] == "NUM"
    assert [n.lower() for n in names[:-1]] == FIELDS


def test_runtime_defaults_match_reference(p3lib):
    """P3Runtime defaults, p3_functions.hpp:114-133."""
    r = p3lib.P3B200Runtime()
    p3lib.lib().p3b200_runtime_defaults(ctypes.byref(r))
    for name, val in RUNTIME_DEFAULTS:
        assert float(getattr(r, name)) == float(val), name


def test_tables_blob_and_generator_match_oracle(p3lib, oracle):
    t = p3lib.load_tables()
    _, ref = oracle.tables()
    assert np.array_equal(np.ctypeslib.as_array(t.ice_table_vals), ref["ice"].ravel())
    got_coll = np.ctypeslib.as_array(t.collect_table_vals)
    # stored as log10 of the file values (p3_table_ice_impl.hpp:63); numpy's log10 may differ from libm's in the last bit
    np.testing.assert_allclose(got_coll, np.log10(ref["coll"].ravel()), rtol=4e-16, atol=0)
    for k in ("vn", "vm", "revap"):
        np.testing.assert_array_equal(np.ctypeslib.as_array(getattr(t, k + "_table_vals")), ref[k].ravel(), err_msg=k)
    assert (np.ctypeslib.as_array(t.mu_r_table_vals) == 1.0).all()


def test_tables_wrong_version_rejected(p3lib):
    """Table version check, p3_table_ice_impl.hpp:40-41."""
    t = p3lib.P3B200Tables()
    rc = p3lib.lib().p3b200_tables_read_blob(p3lib.TABLE_BLOB.encode(), b"9.9.9", ctypes.byref(t))
    assert rc == -4
    rc = p3lib.lib().p3b200_tables_read_blob(b"/nonexistent/blob", b"4.1.1", ctypes.byref(t))
    assert rc == -4


@pytest.mark.skipif(not os.path.exists("/root/reference"), reason="reference tree not mounted")
def test_text_table_parser_matches_blob(p3lib, oracle):
    """The shipped blob is the reference's text table (init_kokkos_ice_lookup_tables parse rules)."""
    t = p3lib.P3B200Tables()
    rc = p3lib.lib().p3b200_tables_read_text(oracle.REF_TABLE_TEXT.encode(), b"4.1.1", ctypes.byref(t))
    assert rc == 0
    b = p3lib.load_tables()
    assert np.array_equal(np.ctypeslib.as_array(t.ice_table_vals), np.ctypeslib.as_array(b.ice_table_vals))
    assert np.array_equal(np.ctypeslib.as_array(t.collect_table_vals), np.ctypeslib.as_array(b.collect_table_vals))


def test_no_cpu_fallback(p3lib):
    """Without a CUDA device every compute entry point must fail with P3B200_ECUDA."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(p3lib.P3Error):
        p3lib.P3Handle(0, 8)
    L = p3lib.lib()
    h = ctypes.c_void_p()
    rc = L.p3b200_create(ctypes.byref(p3lib.load_tables()), 0, 8, ctypes.byref(h))
    assert rc == -2 and not h.value

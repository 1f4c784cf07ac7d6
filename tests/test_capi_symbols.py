"""The C-ABI library loads and exports every function include/*.h declares
(no compute calls: no GPU needed)."""
import ctypes
import glob
import os
import re

from conftest import REPO


def declared_functions():
    names = []
    for header in glob.glob(os.path.join(REPO, "include", "*.h")):
        text = re.sub(r"/\*.*?\*/", "", open(header).read(), flags=re.S)
        names += re.findall(r"\b(lassim_[a-z0-9_]+)\s*\(", text)
    return sorted(set(names))


def test_library_builds_and_exports_header_symbols():
    from lassim_b200.build import build_library
    path = build_library()
    lib = ctypes.CDLL(path)
    names = declared_functions()
    assert len(names) >= 15
    for name in names:
        assert hasattr(lib, name), "missing symbol " + name
    lib.lassim_version.restype = ctypes.c_char_p
    assert lib.lassim_version().decode().count(".") == 2


def test_create_fails_loudly_without_device():
    import torch
    if torch.cuda.is_available():
        return
    import pytest
    from lassim_b200 import Engine, LassimError
    with pytest.raises(LassimError):
        Engine(0)

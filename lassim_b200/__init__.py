"""lassim_b200 - B200-native cost evaluation for LASSIM core problems.

The package mirrors the module layout of the LASSIM toolbox for the part that
sits on the cost-evaluation path (``core``, ``customs``, ``data_management``)
so that a script written against the toolbox switches by changing its imports.
Evaluation goes through ``liblassim_b200.so`` (hand-written sm_100a CUDA behind
the C ABI of ``include/lassim_b200.h``); there is no CPU fallback.
"""
__version__ = "0.1.0"

from ._capi import Engine, LassimError, load_library, library_path  # noqa: F401

"""ctypes binding of libbmhrl_reward.so; prototypes come from include/bmhrl_reward.h (single source of truth).
No fallback: a missing library raises."""
import ctypes
import os

from .._lib import _PKG, _ROOT, parse_header

HEADER = os.path.join(_ROOT, "include", "bmhrl_reward.h")
LIB_PATH = os.path.join(_PKG, "lib", "libbmhrl_reward.so")


class _RewardLib:
    def __init__(self):
        self._dll = None

    def load(self):
        if self._dll is not None:
            return self._dll
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"bmhrl_b200: reward library not built ({LIB_PATH} missing). Run "
                               f"`make -C bmhrl_b200/csrc_host` (or __graft_entry__.build()). There is no Python fallback.")
        dll = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in parse_header(HEADER).items():
            fn = getattr(dll, name)
            fn.restype = res
            fn.argtypes = args
        if dll.bmhrl_reward_abi_version() != 1:
            raise RuntimeError("bmhrl_b200: reward ABI version mismatch between header and library")
        self._dll = dll
        return dll

    def call(self, name, *args):
        dll = self.load()
        rc = getattr(dll, name)(*args)
        if rc != 0:
            raise RuntimeError(f"{name} failed ({rc}): {dll.bmhrl_reward_last_error().decode()}")

    def raw(self, name):
        return getattr(self.load(), name)


RLIB = _RewardLib()

"""TESTS ONLY: route the product's ctypes binding to the host-compiled build of the same kernel sources
(tests/host_emu/libjrl_lbm_emu.so) so that host logic and kernel logic can be checked in the GPU-less container.

This lives under tests/ on purpose: the product package has no switch for it; the test
suite monkeypatches `jrl_native` from the outside -- the loaded library handle and the CUDA-tensor requirement."""
import ctypes

_saved = {}


def install(path):
    import jrl_native
    if not _saved:
        _saved["lib"], _saved["req"] = jrl_native._lib, jrl_native._require_cuda
    jrl_native._lib = jrl_native._bind(ctypes.CDLL(path))
    jrl_native._require_cuda = lambda t, name: None


def uninstall():
    import jrl_native
    if _saved:
        jrl_native._lib, jrl_native._require_cuda = None, _saved["req"]
        _saved.clear()

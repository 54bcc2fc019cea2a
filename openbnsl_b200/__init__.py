"""openbnsl_b200 -- B200-native drop-in for OpenBNSL's contingency-counting hot path.

``openbnsl_b200.openbnsllib`` has the Python surface of the reference's
``openbnsllib`` module (submodules ``base``, ``score``, ``citest``,
``structure_learning``); every data-path call runs hand-written sm_100a CUDA
kernels through the extern-"C" layer in ``libopenbnsl_b200.so``
(include/openbnsl_b200.h).  There is no CPU fallback: importing works without a
GPU (so that the host logic can be unit-tested), but the first call that needs
data on the device raises RuntimeError when no usable CUDA device exists.
"""
import os as _os

_HERE = _os.path.dirname(_os.path.abspath(__file__))


def _require_native():
    import glob
    if not _os.path.exists(_os.path.join(_HERE, "libopenbnsl_b200.so")) or not glob.glob(
            _os.path.join(_HERE, "openbnsllib*.so")):
        raise ImportError("openbnsl_b200 native extensions are not built; run "
                          "`python -m openbnsl_b200.build` (nvcc, sm_100a). There is no fallback.")


def __getattr__(name):
    if name == "openbnsllib":
        _require_native()
        import importlib
        mod = importlib.import_module("openbnsl_b200.openbnsllib")
        globals()["openbnsllib"] = mod
        return mod
    raise AttributeError(name)

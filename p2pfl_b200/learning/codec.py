"""Host wire codec.

Payload = ``pickle.dumps([ndarray per state_dict entry, in state_dict order])``,
byte-compatible with the reference (p2pfl/learning/pytorch/lightning_learner.py:113-138,
SURVEY Appendix B).  Decoding uses an unpickler that only admits numpy array
reconstruction (Q21): a payload that tries to import anything else raises
``DecodingParamsError`` instead of executing code.
"""

from __future__ import annotations

import io
import pickle
from typing import Any, List

import numpy as np

from p2pfl_b200.exceptions import DecodingParamsError
from p2pfl_b200.settings import Settings

_ALLOWED = {
    ("numpy", "ndarray"),
    ("numpy", "dtype"),
    ("numpy.core.multiarray", "_reconstruct"),
    ("numpy._core.multiarray", "_reconstruct"),
    ("numpy.core.multiarray", "scalar"),
    ("numpy._core.multiarray", "scalar"),
    ("numpy.core.numeric", "_frombuffer"),
    ("numpy._core.numeric", "_frombuffer"),
}


class _NumpyOnlyUnpickler(pickle.Unpickler):
    def find_class(self, module: str, name: str) -> Any:
        if (module, name) in _ALLOWED:
            return super().find_class(module, name)
        raise pickle.UnpicklingError(f"forbidden global {module}.{name} in model payload")


def encode_arrays(arrays: List[np.ndarray]) -> bytes:
    return pickle.dumps(arrays)


def decode_arrays(data: bytes) -> List[np.ndarray]:
    try:
        if Settings.SAFE_UNPICKLE:
            obj = _NumpyOnlyUnpickler(io.BytesIO(data)).load()
        else:
            obj = pickle.loads(data)
    except Exception as e:
        raise DecodingParamsError(f"cannot unpickle model payload: {e}") from e
    if not isinstance(obj, list) or not all(isinstance(a, np.ndarray) for a in obj):
        raise DecodingParamsError("model payload is not a list of ndarrays")
    return obj

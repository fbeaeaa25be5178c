"""Pickle-free wire format for parameters and optimizer state.

State travels between peers of an OPEN collaboration (hivemind ships it as protobuf tensors + msgpack metadata;
reference call sites albert/run_trainer.py:124-128, run_first_peer.py:119-121), so the receiving side must never
unpickle it. `pack_state` turns a nested structure of dicts / lists / scalars / tensors into plain msgpack-able
values (tensors become raw little-endian bytes with dtype and shape); `unpack_state` rebuilds it and rejects anything
that is not one of those types, any dtype outside a whitelist and any tensor whose byte count does not match its
shape. The caller still validates shapes against its own model before copying.
"""
from __future__ import annotations

from typing import Any

import numpy as np
import torch

_DTYPES = {
    "float32": torch.float32, "float64": torch.float64, "float16": torch.float16, "bfloat16": torch.bfloat16,
    "int64": torch.int64, "int32": torch.int32, "int16": torch.int16, "int8": torch.int8, "uint8": torch.uint8,
    "bool": torch.bool,
}
_NAMES = {v: k for k, v in _DTYPES.items()}
_MAX_DEPTH = 16


def pack_state(obj: Any, _depth: int = 0) -> Any:
    if _depth > _MAX_DEPTH:
        raise ValueError("state is nested too deeply")
    if isinstance(obj, torch.Tensor):
        t = obj.detach().cpu().contiguous()
        if t.dtype not in _NAMES:
            raise ValueError(f"unsupported tensor dtype {t.dtype}")
        flat = t.reshape(-1)
        raw = flat.view(torch.uint8).numpy().tobytes() if t.dtype != torch.bool else flat.to(torch.uint8).numpy().tobytes()
        return {"__tensor__": _NAMES[t.dtype], "shape": list(t.shape), "data": raw}
    if isinstance(obj, dict):
        return {"__dict__": [[pack_state(k, _depth + 1), pack_state(v, _depth + 1)] for k, v in obj.items()]}
    if isinstance(obj, (list, tuple)):
        return {"__list__": [pack_state(v, _depth + 1) for v in obj], "tuple": isinstance(obj, tuple)}
    if obj is None or isinstance(obj, (bool, int, float, str, bytes)):
        return obj
    if isinstance(obj, (np.integer, np.floating)):
        return obj.item()
    raise ValueError(f"cannot serialise a {type(obj).__name__} without pickle")


def unpack_state(obj: Any, max_tensor_bytes: int = 1 << 34) -> Any:
    """Rebuild what `pack_state` produced. Whatever the bytes from the network decode to, the outcome is either plain
    data (scalars, lists, tuples, dicts, tensors of whitelisted dtypes) or a ValueError."""
    try:
        return _unpack(obj, max_tensor_bytes, 0)
    except ValueError:
        raise
    except (TypeError, KeyError, IndexError, OverflowError, RuntimeError, AttributeError) as e:
        raise ValueError(f"malformed state message: {e!r}") from e


def _unpack(obj: Any, max_tensor_bytes: int, _depth: int) -> Any:
    if _depth > _MAX_DEPTH:
        raise ValueError("state is nested too deeply")
    if obj is None or isinstance(obj, (bool, int, float, str, bytes)):
        return obj
    if isinstance(obj, dict):
        if "__tensor__" in obj:
            name, shape, data = obj["__tensor__"], obj.get("shape"), obj.get("data")
            if (not isinstance(name, str) or name not in _DTYPES or not isinstance(shape, (list, tuple))
                    or not isinstance(data, (bytes, bytearray))):
                raise ValueError("malformed tensor record")
            if not all(isinstance(s, int) and not isinstance(s, bool) for s in shape):
                raise ValueError("malformed tensor shape")
            shape = [int(s) for s in shape]
            if any(s < 0 for s in shape) or len(shape) > 8:
                raise ValueError("malformed tensor shape")
            dtype = _DTYPES[name]
            numel = int(np.prod(shape, dtype=np.int64)) if shape else 1
            itemsize = 1 if dtype == torch.bool else torch.empty((), dtype=dtype).element_size()
            if numel * itemsize != len(data) or len(data) > max_tensor_bytes:
                raise ValueError("tensor byte count does not match its shape")
            if numel == 0:
                return torch.zeros(shape, dtype=dtype)
            flat = torch.frombuffer(bytearray(data), dtype=torch.uint8)
            if dtype == torch.bool:
                return flat.to(torch.bool).reshape(shape)
            return flat.view(dtype).reshape(shape).clone()
        if "__dict__" in obj:
            pairs = obj["__dict__"]
            if not isinstance(pairs, (list, tuple)):
                raise ValueError("malformed dict record")
            out = {}
            for pair in pairs:
                if not isinstance(pair, (list, tuple)) or len(pair) != 2:
                    raise ValueError("malformed dict record")
                k = _unpack(pair[0], max_tensor_bytes, _depth + 1)
                if not isinstance(k, (bool, int, float, str, bytes)):
                    raise ValueError("dict keys must be scalars")
                out[k] = _unpack(pair[1], max_tensor_bytes, _depth + 1)
            return out
        if "__list__" in obj:
            items = obj["__list__"]
            if not isinstance(items, (list, tuple)):
                raise ValueError("malformed list record")
            vals = [_unpack(v, max_tensor_bytes, _depth + 1) for v in items]
            return tuple(vals) if obj.get("tuple") else vals
    raise ValueError(f"unexpected {type(obj).__name__} in a state message")

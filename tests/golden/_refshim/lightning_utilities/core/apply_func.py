from collections import OrderedDict, defaultdict
from collections.abc import Mapping, Sequence


def apply_to_collection(data, dtype, function, *args, wrong_dtype=None, include_none=True, **kwargs):
    """Recursively map ``function`` over every element of ``data`` that is an instance of ``dtype``."""
    if isinstance(data, dtype) and (wrong_dtype is None or not isinstance(data, wrong_dtype)):
        return function(data, *args, **kwargs)
    if isinstance(data, Mapping):
        out = []
        for k, v in data.items():
            v = apply_to_collection(v, dtype, function, *args, wrong_dtype=wrong_dtype, include_none=include_none, **kwargs)
            if include_none or v is not None:
                out.append((k, v))
        if isinstance(data, defaultdict):
            return type(data)(data.default_factory, OrderedDict(out))
        return type(data)(OrderedDict(out))
    is_namedtuple = isinstance(data, tuple) and hasattr(data, "_fields")
    if isinstance(data, Sequence) and not isinstance(data, str):
        out = []
        for d in data:
            v = apply_to_collection(d, dtype, function, *args, wrong_dtype=wrong_dtype, include_none=include_none, **kwargs)
            if include_none or v is not None:
                out.append(v)
        return type(data)(*out) if is_namedtuple else type(data)(out)
    return data

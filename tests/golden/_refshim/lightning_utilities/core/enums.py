from enum import Enum


class StrEnum(str, Enum):
    """String enum with case-insensitive lookup by key or value."""

    @classmethod
    def from_str(cls, value, source="key"):
        if source in ("key", "any"):
            for k in cls.__members__:
                if k.lower() == value.lower():
                    return cls[k]
        if source in ("value", "any"):
            for m in cls:
                if m.value.lower() == value.lower():
                    return m
        raise ValueError(f"Invalid match: expected one of {cls._allowed_matches(source)}, but got {value}.")

    @classmethod
    def try_from_str(cls, value, source="key"):
        try:
            return cls.from_str(value, source)
        except ValueError:
            return None

    @classmethod
    def _allowed_matches(cls, source):
        keys, vals = [], []
        for m in cls:
            keys.append(m.name)
            vals.append(m.value)
        if source == "key":
            return keys
        if source == "value":
            return vals
        return keys + vals

    def __eq__(self, other):
        if isinstance(other, Enum):
            other = other.value
        return self.value.lower() == str(other).lower()

    def __hash__(self):
        return hash(self.value.lower())

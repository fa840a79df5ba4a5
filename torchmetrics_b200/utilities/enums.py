"""String enums used by the task wrappers (reference: utilities/enums.py)."""
from enum import Enum


class EnumStr(str, Enum):
    @staticmethod
    def _name() -> str:
        return "Task"

    @classmethod
    def from_str(cls, value, source: str = "key"):
        for member in cls:
            if member.name.lower() == str(value).lower() or member.value.lower() == str(value).lower():
                return member
        allowed = [m.lower() for m in cls.__members__]
        raise ValueError(f"Invalid {cls._name()}: expected one of {allowed}, but got {value}.")

    def __eq__(self, other):
        other = other.value if isinstance(other, Enum) else str(other)
        return self.value.lower() == other.lower()

    def __hash__(self):
        return hash(self.value.lower())

    def __str__(self):
        return self.value.lower()


class ClassificationTask(EnumStr):
    """binary / multiclass / multilabel selector of the task wrappers."""

    BINARY = "binary"
    MULTICLASS = "multiclass"
    MULTILABEL = "multilabel"

    @staticmethod
    def _name() -> str:
        return "Classification"


class ClassificationTaskNoMultilabel(EnumStr):
    BINARY = "binary"
    MULTICLASS = "multiclass"

    @staticmethod
    def _name() -> str:
        return "Classification"


class DataType(EnumStr):
    """Input case inferred by the legacy `_input_format_classification` (reference: utilities/enums.py:56-73)."""

    BINARY = "binary"
    MULTILABEL = "multi-label"
    MULTICLASS = "multi-class"
    MULTIDIM_MULTICLASS = "multi-dim multi-class"

    @staticmethod
    def _name() -> str:
        return "Data type"


class AverageMethod(EnumStr):
    MICRO = "micro"
    MACRO = "macro"
    WEIGHTED = "weighted"
    NONE = None
    SAMPLES = "samples"

    @staticmethod
    def _name() -> str:
        return "Average method"


class MDMCAverageMethod(EnumStr):
    GLOBAL = "global"
    SAMPLEWISE = "samplewise"

    @staticmethod
    def _name() -> str:
        return "MDMC Average method"

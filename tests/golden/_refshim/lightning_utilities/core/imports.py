import importlib.util


def package_available(name: str) -> bool:
    try:
        return importlib.util.find_spec(name) is not None
    except Exception:
        return False


class RequirementCache:
    """Truthiness == "is this requirement importable"; version pins are evaluated with packaging when possible."""

    def __init__(self, requirement: str = "", module: str = None) -> None:
        self.requirement = requirement
        self.module = module

    def _check(self) -> bool:
        if not hasattr(self, "_ok"):
            self._ok = self._evaluate()
        return self._ok

    def _evaluate(self) -> bool:
        try:
            from importlib.metadata import version
            from packaging.requirements import Requirement

            req = Requirement(self.requirement)
            have = version(req.name)
            # local version tags (e.g. 2.11.0+cu128) are ignored when matching the specifier
            return req.specifier.contains(have.split("+")[0], prereleases=True)
        except Exception:
            if self.module:
                return package_available(self.module)
            return False

    def __bool__(self) -> bool:
        return self._check()

    def __str__(self) -> str:
        return f"Requirement {self.requirement!r} {'met' if self._check() else 'not met'}"

    __repr__ = __str__

"""Exception / warning types of the boundary (reference: utilities/exceptions.py:16-21)."""


class TorchMetricsUserError(Exception):
    """Misuse of the Metric API (e.g. syncing twice)."""


class TorchMetricsUserWarning(Warning):
    """Suspicious but legal use of the Metric API."""

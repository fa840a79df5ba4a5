"""B200-native metric-state hot path with the torchmetrics API."""
__version__ = "0.1.0"

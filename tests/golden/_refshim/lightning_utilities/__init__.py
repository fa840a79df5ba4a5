"""Minimal stand-in for the `lightning_utilities` package (absent in this image).

Only used by tests/golden/make_golden.py so that the *unmodified* reference can be
imported from /root/reference/src to generate golden vectors. Not part of the product.
"""
from lightning_utilities.core.apply_func import apply_to_collection  # noqa: F401

"""Minimal stand-in for the `timm` package, only so that the reference model file
(src/architectures/rise_mobile_v3.py:29 `from timm.layers import DropPath`) imports in this
container where timm is absent.  TEST INFRASTRUCTURE."""

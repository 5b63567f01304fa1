"""Kernels and their Python bindings: flat layout, ctypes loader, autograd ops."""

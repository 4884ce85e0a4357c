"""src/get_pde.py"""
from .pdes import get_pde  # noqa: F401

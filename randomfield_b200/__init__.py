"""randomfield_b200 -- B200-native drop-in for the field-synthesis path of dkirkby/randomfield.

Exports mirror randomfield/__init__.py:16-17.
"""
from .cosmotools import create_cosmology
from .generate import Generator

__all__ = ['create_cosmology', 'Generator']
__version__ = '0.1'

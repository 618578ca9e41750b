from .base import DecodeStrategy
from .batch import DecodeBatch
from .jacobi import DecodeJacobi
from .m1 import DecodeM1
from .m1_graph import DecodeM1Graph
from .speculative import DecodeSpeculative

__all__ = ["DecodeStrategy", "DecodeM1", "DecodeM1Graph", "DecodeBatch", "DecodeSpeculative", "DecodeJacobi"]

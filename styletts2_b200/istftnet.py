"""``from styletts2_b200.istftnet import Decoder`` replaces ``from Modules.istftnet import Decoder``
(/root/reference/models.py:1006)."""
from .decoder import IstftnetDecoder as Decoder  # noqa: F401

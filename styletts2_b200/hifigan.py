"""``from styletts2_b200.hifigan import Decoder`` replaces ``from Modules.hifigan import Decoder``
(/root/reference/models.py:1021)."""
from .decoder import HifiganDecoder as Decoder  # noqa: F401

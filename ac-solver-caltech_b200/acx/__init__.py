"""acx -- B200-native Andrews-Curtis move engine (Python side of libacx.so)."""
from ._lib import AcxError, lib, LIB_PATH  # noqa: F401

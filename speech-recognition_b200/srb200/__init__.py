"""srb200 -- B200-native engine behind the sr.recognition drop-in API."""
from . import lib  # noqa: F401

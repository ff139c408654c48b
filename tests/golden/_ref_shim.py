"""Import shim for the UNMODIFIED reference at /root/reference (container only).

Only the golden-vector generator (`make_golden.py`) uses this; nothing that runs on the GPU box
imports it.  The shim follows SURVEY.md section 8c: stub the two absent third-party modules the
reference's package __init__ pulls in, and restore the numpy aliases it still uses.
"""
import sys
import types

import numpy as np

REF = "/root/reference"


def load_reference():
    if REF not in sys.path:
        sys.path.insert(0, REF)
    pa = types.ModuleType("pyaudio")
    pa.paInt16 = 8
    pa.PyAudio = object
    pa.paContinue = 0
    sys.modules.setdefault("pyaudio", pa)
    sys.modules.setdefault("python_speech_features", types.ModuleType("python_speech_features"))
    if not hasattr(np, "int"):
        np.int = int  # decode.py:24,95  kmeans.py:126,138,200
    if not hasattr(np, "alltrue"):
        np.alltrue = np.all  # continuous_speech.py:173
    import sr.recognition as rec  # noqa: E402  (the reference package)
    assert rec.__file__.startswith(REF), rec.__file__
    return rec

"""Drop-in mirror of the reference's `sr.langmodel` package (sr/langmodel/__init__.py:1-3): the lexical tree and the
text Viterbi spell-checker over it.  The tree is host data; the dynamic programme runs in
srb_text_viterbi (libsrb200, sm_100a)."""
from .langmodel import *       # noqa: F401,F403
from .lextree import *         # noqa: F401,F403
from .spellchecker import *    # noqa: F401,F403

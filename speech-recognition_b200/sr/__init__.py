"""`sr` namespace of the mirror: re-exports what the reference's sr/__init__.py:2 takes from sr.recognition."""
from .recognition import HMMState, HMM, decode_hmm_states, GMM, build_state_sequences, NES  # noqa: F401

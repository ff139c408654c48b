"""`sr` namespace of the mirror: the re-exports of the reference's sr/__init__.py:1-2."""
from .core import delta_feature, load_wav_as_mfcc, make_HMM, train, test, aurora_continuous_train  # noqa: F401
from .recognition import HMMState, HMM, decode_hmm_states, GMM, build_state_sequences, NES  # noqa: F401

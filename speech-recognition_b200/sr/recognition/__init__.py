"""Drop-in mirror of the reference's `sr.recognition` package (sr/recognition/__init__.py:1-5): the same names,
star-exported the same way, computed by the B200 kernels of libsrb200 (no CPU path)."""
from .decode import *          # noqa: F401,F403
from .hmm import *             # noqa: F401,F403
from .kmeans import *          # noqa: F401,F403
from .continuous_speech import *   # noqa: F401,F403
from .hmm_state import *       # noqa: F401,F403
from .hmm_state import GMM, HMMState, MultivariateNormal, NES, mahalanobis  # noqa: F401

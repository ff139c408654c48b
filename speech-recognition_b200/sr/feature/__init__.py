from .feature import *  # noqa: F401,F403

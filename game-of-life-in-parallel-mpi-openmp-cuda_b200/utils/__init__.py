"""Utilities: pattern fixtures, reference builder (oracle / baselines), clock sampling."""
from . import patterns, reference  # noqa: F401
from .clocks import ClockSampler, Timer  # noqa: F401

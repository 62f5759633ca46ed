"""``dtg.errors`` -- the exception names the TF-1 session machinery uses (``tf.errors``).

``MonitoredSession`` re-creates its session on ``AbortedError`` / ``UnavailableError``
("initialization, hooks, recovery", /root/reference/README.md:78) and treats
``OutOfRangeError`` as a clean end of input.
"""
from .framework import OutOfRangeError
from .parallel.rpc import AbortedError, RemoteError, UnavailableError
from .session.hooks import NanLossDuringTrainingError

__all__ = ["AbortedError", "UnavailableError", "OutOfRangeError", "RemoteError",
           "NanLossDuringTrainingError"]

"""Per-worker session: which rank am I, and the queue back to the driver.

Same public surface and error behaviour as ray_lightning/session.py:1-63
(get_actor_rank / put_queue / init_session / get_session / set_session_queue; a second
init_session raises ValueError, put_queue without a queue raises ValueError).  Not on the
data path: it only carries Tune closures from rank 0 to the driver.
"""
from typing import Any, Optional

_ISSUE_URL = "https://github.com/ray-project/ray_lightning_accelerators"


class RayLightningSession:
    def __init__(self, rank: int, queue: Optional[Any]):
        self._rank = rank
        self._queue = queue

    def get_actor_rank(self) -> int:
        return self._rank

    def set_queue(self, queue) -> None:
        self._queue = queue

    def put_queue(self, item) -> None:
        if self._queue is None:
            raise ValueError("Trying to put something into session queue, but queue was not initialized. "
                             f"This is probably a bug, please raise an issue at {_ISSUE_URL}")
        self._queue.put((self._rank, item))


_session: Optional[RayLightningSession] = None


def init_session(*args, **kwargs) -> None:
    global _session
    if _session:
        raise ValueError("Trying to initialize RayLightningSession twice."
                         "\nFIX THIS by not calling `init_session()` manually.")
    _session = RayLightningSession(*args, **kwargs)


def get_session() -> RayLightningSession:
    if not _session or not isinstance(_session, RayLightningSession):
        raise ValueError("Trying to access RayLightningSession from outside an Pytorch Lightning run."
                         "\nFIX THIS by calling function in `session.py` like `get_actor_rank()` only from "
                         "within an Pytorch Lightning actor session.")
    return _session


def set_session_queue(queue) -> None:
    get_session().set_queue(queue)


def get_actor_rank() -> int:
    return get_session().get_actor_rank()


def put_queue(*args, **kwargs) -> None:
    get_session().put_queue(*args, **kwargs)


def _reset_session_for_tests() -> None:
    """Workers are fresh processes in production (ray_ddp.py:105,118-120); in-process tests
    need a way back to the pristine state."""
    global _session
    _session = None

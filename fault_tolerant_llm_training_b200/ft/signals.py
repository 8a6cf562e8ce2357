"""Signal handling.

The reference converts SIGUSR1 / SIGTERM into an exception raised *inside the
handler* (utils.py:93-97, registered at train.py:89-90), so the interruption can
land between any two bytecodes of the loop — including half-way through the
optimizer's foreach sequence.  Here the handlers only record the
signal; the training loop polls at the step boundary, agrees on the decision
across ranks, and then raises `TrainingInterrupt`, whose `.args` keep the
reference's `("Exception", signum)` shape so `e.args[1]` still is the error type
(train.py:123-126).
"""
from __future__ import annotations

import signal
import threading
from typing import Optional

SIGNAL_TIMEOUT = int(signal.SIGUSR1)   # 10 on Linux: Slurm --signal=USR1@120 (train.sh:12)
SIGNAL_CANCEL = int(signal.SIGTERM)    # 15: scancel
ERROR_TYPE_EXCEPTION = -1              # any Python exception (train.py:123-124)

# stop codes used for the cross-rank MAX-reduction; larger wins
STOP_NONE, STOP_TIMEOUT, STOP_ERROR, STOP_CANCEL = 0, 1, 2, 3
_CODE_TO_TYPE = {STOP_TIMEOUT: SIGNAL_TIMEOUT, STOP_ERROR: ERROR_TYPE_EXCEPTION, STOP_CANCEL: SIGNAL_CANCEL}
_TYPE_TO_CODE = {v: k for k, v in _CODE_TO_TYPE.items()}


class TrainingInterrupt(Exception):
    """Raised by the loop at a step boundary after a signal was agreed on."""

    def __init__(self, signum: int):
        super().__init__("Exception", signum)
        self.signum = signum


class PeerAbort(Exception):
    """Another rank failed mid-step; this rank's state is that of the last completed step."""


def error_type_of(exc: BaseException) -> int:
    """Map an exception to the reference's integer error type.

    Signals carry their number.  Everything else is -1: the reference takes
    `e.args[1]` of *any* exception, so `OSError(28, "No space left")` becomes
    "Unknown exit signal" and nothing is saved; that is not kept."""
    if isinstance(exc, TrainingInterrupt):
        return exc.signum
    return ERROR_TYPE_EXCEPTION


class SignalFlags:
    """Flag-setting handlers for SIGUSR1 and SIGTERM.  Install *before* setup
    (the reference installs after ~37-60 s of setup, so an early signal kills the
    process silently: SURVEY Q12)."""

    def __init__(self):
        self._lock = threading.Lock()
        self._pending = STOP_NONE
        self._installed = False
        self._old = {}
        self.saving = False  # set by the exit handler; further USR1s are ignored while True

    def install(self) -> "SignalFlags":
        if threading.current_thread() is not threading.main_thread():
            return self  # signal.signal only works on the main thread (e.g. under a test runner thread)
        for sig in (signal.SIGUSR1, signal.SIGTERM):
            self._old[sig] = signal.signal(sig, self._on_signal)
        self._installed = True
        return self

    def uninstall(self) -> None:
        if self._installed:
            for sig, old in self._old.items():
                signal.signal(sig, old)
            self._installed = False

    def _on_signal(self, signum, frame) -> None:
        code = _TYPE_TO_CODE.get(int(signum), STOP_NONE)
        # no lock: handlers run on the main thread between bytecodes; max() keeps CANCEL over TIMEOUT
        if code > self._pending:
            self._pending = code

    def raise_flag(self, code: int) -> None:
        if code > self._pending:
            self._pending = code

    def poll(self) -> int:
        """Pending stop code (STOP_*), without clearing it."""
        return self._pending

    def clear(self) -> None:
        self._pending = STOP_NONE


def stop_code_to_error_type(code: int) -> Optional[int]:
    return _CODE_TO_TYPE.get(code)

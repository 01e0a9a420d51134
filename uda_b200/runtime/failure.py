"""Failure -> fallback contract of the accelerated shuffle, for jobs driven from Python.

Reference behaviour: any exception before the first merged byte has been handed to the
reducer makes the plug-in give the whole shuffle back to vanilla Hadoop (``failureInUda`` up-call,
src/UdaBridge.cc:505-530; ``doFallbackInit`` / ``fetchOutputs`` retry in
plugins/shared/.../UdaShuffleConsumerPluginShared.java:187-260); after that point falling back is impossible
(build/journal.txt:167) and the task fails; ``mapred.rdma.developer.mode=true`` turns every fallback into an abort
(…Shared.java:210-217); once failed, further commands are ignored (src/UdaBridge.cc:75-86).

The native runtime implements this for the JNI path (csrc/runtime/threads.h FailureLatch). This module is the same
contract for the GPU jobs: a kernel watchdog time-out, a CRC mismatch of a landed segment or an allocation failure
raises :class:`UdaFailure`; :class:`FallbackController` decides between fallback, abort and hard failure.
"""
from __future__ import annotations

import contextlib
import logging
from typing import Callable, Optional

log = logging.getLogger("org.apache.hadoop.mapred.ShuffleConsumerPlugin")  # the logger name the reference uses


class UdaFailure(RuntimeError):
    """A failure of the accelerated path that the host may answer with its fallback shuffle."""


class ShuffleTimeout(UdaFailure):
    """The landed-byte counters did not reach their expected value within the spin budget (a lost or short push)."""


class ChecksumMismatch(UdaFailure):
    """A landed segment does not match its IFile CRC32."""


class ArenaAllocationFailed(UdaFailure):
    """The symmetric receive arena could not be allocated (the reference's MR registration failure)."""


class FallbackImpossible(RuntimeError):
    """A failure after merged data was already delivered: the reduce task has to fail."""


class FallbackController:
    """Wraps the phases of a job. ``on_failure`` is the ``failureInUda`` up-call; it fires at most once."""

    def __init__(self, on_failure: Optional[Callable[[BaseException], None]] = None, developer_mode: bool = False):
        self.on_failure, self.developer_mode = on_failure, developer_mode
        self.failed = False
        self.delivered = False          # set by the job when the first merged byte left for the consumer
        self.reason: Optional[BaseException] = None

    def mark_delivered(self) -> None:
        self.delivered = True

    @contextlib.contextmanager
    def guard(self, phase: str):
        """Runs one phase. Returns normally after a handled failure, with :attr:`failed` set; the caller must then
        stop using the accelerated path (commands after a failure are ignored, like the reference)."""
        if self.failed:
            log.warning("phase %s skipped: UDA already failed and fell back", phase)
            yield False
            return
        try:
            yield True
        except UdaFailure as e:
            self._handle(phase, e)
        except (RuntimeError, MemoryError, ValueError) as e:   # CUDA errors surface as RuntimeError
            self._handle(phase, e)

    def _handle(self, phase: str, e: BaseException) -> None:
        self.failed, self.reason = True, e
        if self.developer_mode:
            log.error("UDA failure in %s with mapred.rdma.developer.mode=true: aborting instead of falling back", phase)
            raise e
        if self.delivered:
            raise FallbackImpossible(f"failure in {phase} after merged data was delivered: {e}") from e
        log.error("UDA has encountered a critical error and will try to fallback to vanilla MSG=%s", e)
        if self.on_failure:
            self.on_failure(e)


def check_wait_status(status_tensor, what: str = "shuffle") -> None:
    """Raises :class:`ShuffleTimeout` when the wait/barrier kernels recorded a time-out in the arena's status word
    (csrc/kernels/shuffle.cu: wait_flags_kernel / barrier_kernel give up after ``max_spins`` polls). Synchronises."""
    st = status_tensor.cpu()
    if int(st[0]) != 0:
        raise ShuffleTimeout(f"{what}: landed-byte flags did not reach their expected values (status={st.tolist()[:4]})")
